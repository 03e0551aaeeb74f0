"""Generates the committed golden vectors.  Run in the build container (needs oracle/_ref,
i.e. /root/reference at build time):  python tests/golden/make_golden.py

  tsdfcu_small.npz   inputs + outputs of the REFERENCE's own source lines
                     (oracle/_ref/libtsdf_ref.so = src/tsdf.cu:15-60,170-218,253-403 compiled
                     as CPU C++, -O2 -ffp-contract=off): 3 frames of 60x80 depth into a
                     40x36x32 grid.  Pins oracle/tsdf_oracle.c (SEM_TSDFCU) and the CUDA path.
  fusionpy_small.npz same inputs + colour through oracle/tsdf_oracle.c (SEM_FUSIONPY).  The
                     third-party upstream is unavailable offline: regression fixture only
                     ("parity unpinned").
"""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.oracle import Oracle, RefLift, SEM_TSDFCU, SEM_FUSIONPY  # noqa: E402

synth = importlib.import_module("semantic-slam_b200.synth")
HERE = os.path.dirname(os.path.abspath(__file__))

dims, vs = (40, 36, 32), 0.05
hw = (60, 80)
K = synth.TUM3_K.copy()
K[[0, 2, 4, 5]] /= 8.0  # 640x480 -> 80x60
rng = np.random.default_rng(0xB200)
origin, frames = synth.s_real(dims, vs, n_frames=3, K=K, hw=hw, seed=0xB200)
base = synth.pose(synth.rot_xyz(0.02, -0.03, 0.01), (0.05, -0.02, 0.03))
trunc = np.float32(vs) * np.float32(5)

o, r = Oracle(), RefLift()
n = dims[0] * dims[1] * dims[2]
t, w = np.ones(n, np.float32), np.zeros(n, np.float32)
c2bs = []
binv = r.invert(base)
for d, rgb, T in frames:
    c2b = r.multiply(binv, T)          # src/tsdf.cu:142
    c2bs.append(c2b)
    r.integrate(K, c2b, d, dims, origin, vs, trunc, t, w)
ply = os.path.join(HERE, "_tmp.ply")
r.save_ply(ply, dims, vs, origin, t, w)
ply_bytes = np.frombuffer(open(ply, "rb").read(), np.uint8)
os.remove(ply)
np.savez_compressed(os.path.join(HERE, "tsdfcu_small.npz"), dims=dims, vs=vs, hw=hw, K=K, origin=origin,
                    base2world=base, depth=np.stack([f[0] for f in frames]),
                    rgb=np.stack([f[1] for f in frames]), T_wc=np.stack([f[2] for f in frames]),
                    cam2base=np.stack(c2bs), tsdf=t, weight=w, ply=ply_bytes)

p = o.params(dims, vs, trunc, origin, hw, K, np.inf, SEM_FUSIONPY)
t2, w2, c2 = o.new_volume(p, True)
for d, rgb, T in frames:
    o.integrate(p, T.reshape(16), d, o.pack_color(rgb), 0.75, t2, w2, c2)
xyz, rgbp, idx = o.extract(p, t2, w2, c2)
np.savez_compressed(os.path.join(HERE, "fusionpy_small.npz"), tsdf=t2, weight=w2, color=c2, idx=idx,
                    xyz=xyz, rgb=rgbp, obs_weight=0.75)
print("tsdfcu: written voxels", int((w > 0).sum()), "ply bytes", ply_bytes.size, "| fusionpy points", len(idx))
