"""Small driver for profiling the sparse-frame kernel and the extraction sweep.
usage: python tools/profile_sparse.py [real|kitti]
  real : S-real 512^3 (fr3_office keyframe poses), 6 frames, then one surface extraction
  kitti: S-kitti 1241x376 into 1024x256x1024 @ 10 cm, max_depth 50 m, 6 frames"""
import importlib, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("semantic-slam_b200")
synth = pkg.synth
scene = sys.argv[1] if len(sys.argv) > 1 else "real"
if scene == "kitti":
    dims, vs = (1024, 256, 1024), 0.1
    o2, fr2 = synth.s_kitti(n_frames=6, voxel_size=vs, dims=dims)
    cfg = pkg.make_config(dims, vs, o2, K=synth.KITTI_K, im_hw=synth.KITTI_HW, max_depth=50.0)
else:
    dims, vs = (512, 512, 512), 0.01
    o2, fr2 = synth.s_real(dims, vs, n_frames=6, poses=synth.load_pose_fixture()[::32])
    cfg = pkg.make_config(dims, vs, o2)
with pkg.Volume(cfg) as v:
    for d, c, T in fr2:
        v.integrate(d, c, T, sync=False)
    v.sync()
    print("written last", v.stats().voxels_written, "points", len(v.extract_points(pkg.PRED_ABS_LT, 0.5, 1.5)[2]))
