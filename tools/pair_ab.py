"""A/B of the two-frames-per-pass kernel (k_integrate2) against one launch per frame.
usage: python tools/pair_ab.py <scene> <dim> [k]
  scene: full (S-full, every voxel written) | sphere (bench workload) | slab (one of 8 z-slabs of S-full)
The k frames of a tsdf_integrate_multi_async call are resident in HBM (the call's staging cost is
not what is compared); variant bit 14 disables pairing.  Round-robin, R rounds; median / min ms per frame."""
import ctypes as C, importlib, json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("semantic-slam_b200")
pkg.capi.select(dev=True)   # A/B variants live in libtsdf_b200_dev.so only
synth, lib, check = pkg.synth, pkg.load(), pkg.capi.check
scene = sys.argv[1] if len(sys.argv) > 1 else "full"
dim = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
k = int(sys.argv[3]) if len(sys.argv) > 3 else 8
R = int(os.environ.get("AB_ROUNDS", "7"))
dims, vs = (dim,) * 3, 2.048 / dim
origin, frames = synth.s_full(dims, vs, n_frames=4, sphere=(scene == "sphere"))
kw = dict(z_range=(dim // 2, dim // 2 + dim // 8)) if scene == "slab" else {}
cfg = pkg.make_config(dims, vs, origin, **kw)
depth = np.ascontiguousarray(np.stack([frames[i % 4][0] for i in range(k)]))
rgb = np.ascontiguousarray(np.stack([frames[i % 4][1] for i in range(k)]))
T = np.ascontiguousarray(np.stack([frames[i % 4][2].reshape(16) for i in range(k)]).astype(np.float32))
dd, dr = C.c_void_p(), C.c_void_p()
check(lib.tsdf_device_alloc(0, C.byref(dd), depth.nbytes)); check(lib.tsdf_device_alloc(0, C.byref(dr), rgb.nbytes))
check(lib.tsdf_memcpy_h2d(0, dd, depth.ctypes.data, depth.nbytes)); check(lib.tsdf_memcpy_h2d(0, dr, rgb.ctypes.data, rgb.nbytes))
variants = {"pairs": 0, "singles": 1 << 14}
times = {n: [] for n in variants}
with pkg.Volume(cfg) as vol:
    for rnd in range(R + 1):
        for name, var in variants.items():
            vol.set_variant(var)
            vol.timer_start()
            for _ in range(2):
                check(lib.tsdf_integrate_multi_async(vol.h, k, dd, dr, None, T.ctypes.data, np.float32(1.0)))
            ms = vol.timer_stop()
            if rnd:
                times[name].append(ms / (2 * k))
    cover = vol.stats().voxels_written / (vol.n_local * k)
lib.tsdf_device_free(0, dd); lib.tsdf_device_free(0, dr)
print(json.dumps({"scene": scene, "dim": dim, "k": k, "coverage": round(cover, 4), "rounds": R,
                  "ms_per_frame": {n: {"median": round(float(np.median(t)), 4), "min": round(float(min(t)), 4)} for n, t in times.items()}}))
