"""Mesh readback timing: the sweep (k_mc_mark* + two scans + counts to the host)
and the emit kernels, device time by CUDA events on the library's compute stream, the 8-wide
sweep against the 4-wide one (variant bit 18) and the scalar one (bit 16), round-robin.
usage: python tools/mesh_ab.py <scene: real|sphere> <dim> [cu]
  real   = S-real fused from 8 fr3_office keyframes (C2-like); sphere = analytic sphere uploaded
  cu     = class-TSDF storage order (x fast) instead of the plugin's (z fast)"""
import importlib, sys, os, json, ctypes as C
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("semantic-slam_b200")
pkg.capi.select(dev=True)   # A/B variants live in libtsdf_b200_dev.so only
synth = pkg.synth
scene = sys.argv[1] if len(sys.argv) > 1 else "real"
dim = int(sys.argv[2]) if len(sys.argv) > 2 else 512
sem = pkg.SEM_TSDFCU if "cu" in sys.argv[3:] else pkg.SEM_FUSIONPY
R = int(os.environ.get("AB_ROUNDS", "7"))
dims, vs = (dim,) * 3, 5.12 / dim
if scene == "real":
    origin, frames = synth.s_real(dims, vs, n_frames=8, poses=synth.load_pose_fixture()[::24])
else:
    origin, frames = (0.0, 0.0, 0.0), []
cfg = pkg.make_config(dims, vs, origin, semantics=sem, with_color=(sem == pkg.SEM_FUSIONPY),
                      max_depth=6.0 if sem == pkg.SEM_TSDFCU else None)
lib = pkg.load()
VARIANTS = ((0, "default"), (262144, "sweep_4_wide"), (65536, "scalar_sweep"))
res = {v: {"sweep": [], "emit": []} for v, _ in VARIANTS}
with pkg.Volume(cfg) as vol:
    if frames:
        for d, c, T in frames:
            vol.integrate(d, c if sem == pkg.SEM_FUSIONPY else None, T)
    else:
        ax = np.arange(dim, dtype=np.float32)
        r = np.sqrt((ax[:, None, None] - 0.47 * dim) ** 2 + (ax[None, :, None] - 0.52 * dim) ** 2 +
                    (ax[None, None, :] - 0.5 * dim) ** 2)
        vol.upload(tsdf=np.clip((r - 0.31 * dim) / 5.0, -1, 1).astype(np.float32))
    nv, nf = vol.mesh_count()
    dv = C.c_void_p(); dn = C.c_void_p(); dc = C.c_void_p(); df = C.c_void_p()
    for ptr_, nbytes in ((dv, nv * 12), (dn, nv * 12), (dc, nv * 3), (df, nf * 12)):
        pkg.capi.check(lib.tsdf_device_alloc(cfg.device, C.byref(ptr_), max(nbytes, 16)))
    a, b = C.c_int64(0), C.c_int64(0)
    for rnd in range(R + 1):
        for var, _ in VARIANTS:
            vol.set_variant(var)
            vol.upload()                      # no arrays: only invalidates the cached sweep
            vol.timer_start()
            assert vol.mesh_count() == (nv, nf)
            t_sweep = vol.timer_stop()
            vol.timer_start()
            pkg.capi.check(lib.tsdf_extract_mesh_device(vol.h, dv, dn, dc, nv, df, nf, C.byref(a), C.byref(b)))
            t_emit = vol.timer_stop()
            if rnd:
                res[var]["sweep"].append(t_sweep)
                res[var]["emit"].append(t_emit)
    for ptr_ in (dv, dn, dc, df):
        lib.tsdf_device_free(cfg.device, ptr_)
    n = vol.n_local
out = {"scene": scene, "dim": dim, "order": "x-fast" if sem == pkg.SEM_TSDFCU else "z-fast", "vertices": nv,
       "faces": nf, "rounds": R}
for var, name in VARIANTS:
    sw, em = float(np.median(res[var]["sweep"])), float(np.median(res[var]["emit"]))
    out[name] = {"sweep_ms": round(sw, 4), "sweep_GBps_algorithmic": round(5.5 * n / sw / 1e6, 1),
                 "emit_ms": round(em, 4)}
print(json.dumps(out))
