"""Timing + correctness probe of the extraction path at C2 size (512^3)."""
import importlib, sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("semantic-slam_b200")
synth = pkg.synth
dims, vs = (512, 512, 512), 0.01
o2, fr2 = synth.s_real(dims, vs, n_frames=4, poses=synth.load_pose_fixture()[::48])
cfg = pkg.make_config(dims, vs, o2)
with pkg.Volume(cfg) as v:
    for d, c, T in fr2:
        v.integrate(d, c, T, sync=False)
    v.sync()
    for k, (tt, wt) in enumerate([(0.5, 1.25), (0.5, 1.5), (0.5, 1.5), (0.4, 1.5)]):
        t0 = time.perf_counter(); n = v.extract_count(pkg.PRED_ABS_LT, tt, wt); t1 = time.perf_counter()
        xyz, rgb, idx = v.extract_points(pkg.PRED_ABS_LT, tt, wt); t2 = time.perf_counter()
        print(f"pass {k}: count {n} in {(t1-t0)*1e3:.2f} ms, points in {(t2-t1)*1e3:.2f} ms", flush=True)
    t, w, c = v.download()
    sel = np.flatnonzero((np.abs(t) < np.float32(0.4)) & (w > np.float32(1.5)))
    print("index set equal:", np.array_equal(sel, idx), len(sel), len(idx))
