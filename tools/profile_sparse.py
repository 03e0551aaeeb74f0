"""Small driver for profiling the sparse-frame kernel and the extraction sweep:
S-real 512^3 (fr3_office keyframe poses), 6 frames, then one surface extraction."""
import importlib, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("semantic-slam_b200")
synth = pkg.synth
dims, vs = (512, 512, 512), 0.01
o2, fr2 = synth.s_real(dims, vs, n_frames=6, poses=synth.load_pose_fixture()[::32])
cfg = pkg.make_config(dims, vs, o2)
with pkg.Volume(cfg) as v:
    for d, c, T in fr2:
        v.integrate(d, c, T, sync=False)
    v.sync()
    print("written last", v.stats().voxels_written, "points", len(v.extract_points(pkg.PRED_ABS_LT, 0.5, 1.5)[2]))
