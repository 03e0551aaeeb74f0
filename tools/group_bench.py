"""Throughput of the single-process multi-GPU handle (tsdf_group): the bench workload (1024^3, S-full +
sphere, pinned host frames) cut into z-slabs over N devices, driven by ONE host thread.
usage: python tools/group_bench.py <n_devices> [steps]   -> one JSON line (host wall clock around the loop,
synchronised on both sides; compare with bench.py's one-process-per-GPU numbers)."""
import ctypes as C, importlib, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("semantic-slam_b200")
synth, lib, check = pkg.synth, pkg.load(), pkg.capi.check
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 40
DIM, VS = int(os.environ.get("TSDF_BENCH_DIM", "1024")), 0.002
dims = (DIM,) * 3
origin, frames = synth.s_full(dims, VS * 1024 / DIM, n_frames=4, sphere=True)
cfg = pkg.make_config(dims, VS * 1024 / DIM, origin)
hd, hr = [], []
for d, rgb, _ in frames:   # pinned host frames
    pd, pr = C.c_void_p(), C.c_void_p()
    check(lib.tsdf_host_alloc(C.byref(pd), d.nbytes)); check(lib.tsdf_host_alloc(C.byref(pr), rgb.nbytes))
    C.memmove(pd, d.ctypes.data, d.nbytes); C.memmove(pr, rgb.ctypes.data, rgb.nbytes)
    hd.append(pd.value); hr.append(pr.value)
with pkg.Group(cfg, list(range(n))) as g:
    for i in range(6):
        g.integrate(hd[i % 4], hr[i % 4], frames[i % 4][2], sync=False)
    g.sync()
    wins = []
    for _ in range(5):
        t0 = time.perf_counter()
        for i in range(steps):
            g.integrate(hd[i % 4], hr[i % 4], frames[i % 4][2], sync=False)
        g.sync()
        wins.append((time.perf_counter() - t0) * 1e3 / steps)
    st = g.stats()
    t0 = time.perf_counter()
    pool = pkg.PinnedPool()
    g.extract_points(pkg.PRED_ABS_LT, 0.999, 0.5, pool=pool)
    t0 = time.perf_counter()
    pts = g.extract_points(pkg.PRED_ABS_LT, 0.999, 0.5, pool=pool)
    rb_ms = (time.perf_counter() - t0) * 1e3
    ms = float(np.median(wins))
    print(json.dumps({"what": "tsdf_group, one process / one host thread", "n_devices": n, "peer_direct": g.peer_access(),
                      "grid": list(dims), "steps": steps, "ms_per_step_median": ms, "windows": wins,
                      "Gvox_per_s": DIM ** 3 / ms / 1e6, "coverage": st.voxels_written / DIM ** 3,
                      "readback_ms_cached_mark": rb_ms, "points": len(pts[2])}))
