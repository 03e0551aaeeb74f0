"""Writes tests/golden/mesh_small.npz: a seeded tsdf + colour volume (a bumpy sphere cut by a plane,
[x][y][z] order) and the mesh the CPU oracle (oracle/tsdf_oracle.c: oracle_mesh) gives for it.
Regression fixture: the mesh definition is this repo's (upstream's skimage is absent — parity
unpinned); the properties in tests/test_mesh.py are what validates it.
usage: python tests/golden/make_mesh_golden.py"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle.oracle import Oracle, SEM_FUSIONPY  # noqa: E402

dims, vs, origin = (22, 18, 20), 0.02, (-0.2, 0.1, 0.5)
rng = np.random.default_rng(0xB200 + 6)
X, Y, Z = np.meshgrid(*[np.arange(d, dtype=np.float64) for d in dims], indexing="ij")
t = np.sqrt((X - 10.4) ** 2 + (Y - 8.7) ** 2 + (Z - 9.2) ** 2) - 6.3 + 0.4 * np.sin(X) * np.cos(1.3 * Y)
t = np.minimum(t, 15.5 - Z) / 5.0
t = np.clip(t, -1, 1).astype(np.float32)
color = rng.integers(0, 1 << 24, dims).astype(np.float32)
o = Oracle()
p = o.params(dims, vs, 5 * vs, origin, (480, 640), np.eye(3, dtype=np.float32), 6.0, SEM_FUSIONPY)
v, f, n, c = o.mesh(p, t.reshape(-1), color.reshape(-1))
np.savez_compressed(os.path.join(HERE, "mesh_small.npz"), tsdf=t, color=color, origin=np.array(origin, np.float32),
                    verts=v, faces=f, norms=n, colors=c)
print("verts", len(v), "faces", len(f))
