"""Where does the sparse integrate kernel spend its work?  (dev library: tsdf_debug_counters)
usage: python tools/sparse_counters.py [real|kitti]"""
import ctypes as C, importlib, json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("semantic-slam_b200")
pkg.capi.select(dev=True)
synth, lib = pkg.synth, pkg.load()
scene = sys.argv[1] if len(sys.argv) > 1 else "real"
if scene == "kitti":
    dims, vs = (1024, 256, 1024), 0.1
    o2, fr2 = synth.s_kitti(n_frames=4, voxel_size=vs, dims=dims)
    cfg = pkg.make_config(dims, vs, o2, K=synth.KITTI_K, im_hw=synth.KITTI_HW, max_depth=50.0)
else:
    dims, vs = (512, 512, 512), 0.01
    o2, fr2 = synth.s_real(dims, vs, n_frames=12, poses=synth.load_pose_fixture()[::16])
    cfg = pkg.make_config(dims, vs, o2)
names = ["blocks", "blocks_skipped", "segments", "segments_culled", "runs_projected", "runs_gathered", "runs_updated", "-"]
with pkg.Volume(cfg) as v:
    for d, c, T in fr2:
        v.integrate(d, c, T, sync=False)
    v.sync()
    pkg.capi.check(lib.tsdf_debug_counters(v.h, 1, None))
    written = 0
    for d, c, T in fr2:
        v.integrate(d, c, T)
        written += v.stats().voxels_written
    out = (C.c_uint64 * 8)()
    pkg.capi.check(lib.tsdf_debug_counters(v.h, 0, out))
    n = len(fr2)
    r = {k: out[i] / n for i, k in enumerate(names[:7])}
    r["voxels_written"] = written / n
    r["voxels"] = int(np.prod(dims))
    r["scene"] = scene
    print(json.dumps(r))
