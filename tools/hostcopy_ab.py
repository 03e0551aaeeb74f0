"""host_copy (pageable -> pinned staging on several threads) against a plain memcpy, on this box's host.
usage: python tools/hostcopy_ab.py   (TSDF_B200_COPY_THREADS picks the helper count, default 3)"""
import ctypes as C, importlib, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("semantic-slam_b200")
lib = pkg.load()
f = getattr(lib, "_Z9host_copyPvPKvm")
f.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
f.restype = None
out = {"threads_env": os.environ.get("TSDF_B200_COPY_THREADS"), "cpu_count": os.cpu_count()}
rng = np.random.default_rng(0)
for n in (1_228_800, 2_150_400, 3_266_312):
    a = rng.integers(0, 255, n, dtype=np.uint8)
    p = C.c_void_p()
    pkg.capi.check(lib.tsdf_host_alloc(C.byref(p), n))
    for name, fn in (("pool", lambda: f(p, a.ctypes.data, n)), ("memcpy", lambda: C.memmove(p, a.ctypes.data, n))):
        fn()
        ts = []
        for _ in range(200):
            t0 = time.perf_counter(); fn(); ts.append(time.perf_counter() - t0)
        out[f"{n}_{name}_ms"] = round(float(np.median(ts)) * 1e3, 4)
    lib.tsdf_host_free(p)
print(json.dumps(out))
