"""A/B timing of kernel variants on the S-full workload (evidence for DESIGN.md choices).
usage: python tools/variants.py [dim] [variant ...]   (variant bits: 1 literal math, 2/4 prefetch L2/L1,
8/16 rows-per-block x64/x32)"""
import importlib, sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("semantic-slam_b200")
dim = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
variants = [int(v) for v in sys.argv[2:]] or [0, 2, 4]
vs = 2.048 / dim
origin, frames = pkg.synth.s_full((dim,) * 3, vs, n_frames=2)
cfg = pkg.make_config((dim,) * 3, vs, origin)
out = {}
with pkg.Volume(cfg) as vol:
    for var in variants:
        vol.reset()
        vol.set_variant(var)
        d, c, T = frames[0]
        vol.integrate(d, c, T)
        vol.bench_kernel(3)
        ms = min(vol.bench_kernel(10) for _ in range(3))
        gb = (24 * vol.n_local + 16 * d.size) / 1e9
        out[var] = {"ms": round(ms, 4), "GBps": round(gb / ms * 1e3, 1), "written": vol.stats().voxels_written}
        print(var, out[var], flush=True)
print(json.dumps(out))
