"""A/B timing of kernel variants (evidence for DESIGN.md choices).
usage: python tools/variants.py <scene> <dim> [variant ...]
  scene: slab  = z-slab [dim/2, dim/2+dim/8) of S-full (one rank of 8)
         full  = S-full (every voxel written; the roofline workload)
         cu    = S-full in class-TSDF semantics (x-fast, depth + weight only; 200 = the reference's default volume)
         real  = S-real with fr3_office keyframe poses (C2-like, ~7 % coverage)
         kitti = S-kitti 1241x376 into (dim, dim/4, dim) at 102.4 m / dim voxels (C3-like)
  variant bits: 1 literal math, 2/4 prefetch L2/L1, 8/16/256/512 rows-per-block x64/x32/x8/x4, 32 prefetch off,
                64 segment cull off, 128 cull + staged probe, 2048 wave balancing off
Variants are timed round-robin (R rounds of 5 launches each per frame) so that clock / power
drift hits all of them alike; the report is the median and the minimum over rounds."""
import importlib, sys, os, json
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("semantic-slam_b200")
pkg.capi.select(dev=True)   # A/B variants live in libtsdf_b200_dev.so only
synth = pkg.synth
scene = sys.argv[1] if len(sys.argv) > 1 else "full"
dim = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
variants = [int(v) for v in sys.argv[3:]] or [0, 64]
R = int(os.environ.get("AB_ROUNDS", "7"))
kw = {}
if scene == "slab":      # one of 8 z-slabs of the S-full grid (what each rank of an 8-GPU run integrates)
    dims, vs = (dim,) * 3, 2.048 / dim
    origin, frames = synth.s_full(dims, vs, n_frames=2)
    kw = dict(z_range=(dim // 2, dim // 2 + dim // 8))
elif scene == "full":
    dims, vs = (dim,) * 3, 2.048 / dim
    origin, frames = synth.s_full(dims, vs, n_frames=2)
elif scene == "cu":
    dims, vs = (dim,) * 3, 2.048 / dim
    origin, frames = synth.s_full(dims, vs, n_frames=2)
    kw = dict(semantics=pkg.SEM_TSDFCU, with_color=False, max_depth=6.0)
elif scene == "real":
    dims, vs = (dim,) * 3, 5.12 / dim
    origin, frames = synth.s_real(dims, vs, n_frames=4, poses=synth.load_pose_fixture()[::48])
else:
    dims, vs = (dim, dim // 4, dim), 102.4 / dim
    origin, frames = synth.s_kitti(n_frames=4, voxel_size=vs, dims=dims)
    kw = dict(K=synth.KITTI_K, im_hw=synth.KITTI_HW, max_depth=50.0)
cfg = pkg.make_config(dims, vs, origin, **kw)
times = {v: [] for v in variants}
cover = 0.0
with pkg.Volume(cfg) as vol:
    for rnd in range(R + 1):               # round 0 = warm-up, dropped
        for var in variants:
            tot = 0.0
            for d, c, T in frames:
                vol.set_variant(var)
                vol.integrate(d, c, T)      # uploads this frame; bench_kernel re-launches on it
                tot += vol.bench_kernel(5)
                cover = vol.stats().voxels_written / vol.n_local
            if rnd:
                times[var].append(tot / len(frames))
out = {}
for var in variants:
    t = np.array(times[var])
    out[var] = {"median_ms": round(float(np.median(t)), 4), "min_ms": round(float(t.min()), 4),
                "Gvox_visited_per_s": round(vol.n_local / float(np.median(t)) / 1e6, 1)}
print(json.dumps({"scene": scene, "dim": dim, "coverage_last": round(cover, 4), "rounds": R, "variants": out}))
