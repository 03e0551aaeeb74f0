import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.oracle import Oracle, SEM_FUSIONPY
pkg = importlib.import_module("semantic-slam_b200")
synth = pkg.synth
o = Oracle()
dims, vs, md = (56, 40, 96), 0.04, 40.0
origin, frames = synth.s_corridor(dims, vs, n_frames=8, seed=5)
def run(devs, pattern):
    p = o.params(dims, vs, np.float32(vs) * np.float32(5), origin, frames[0][0].shape, synth.TUM3_K, md, SEM_FUSIONPY)
    t, w, c = o.new_volume(p, True)
    cfg = pkg.make_config(dims, vs, origin, semantics=SEM_FUSIONPY, with_color=True, max_depth=md)
    with pkg.Group(cfg, devs) as g:
        for i, (d, rgb, T) in enumerate(frames):
            o.integrate(p, T.reshape(16), d, o.pack_color(rgb), 1.0, t, w, c)
            g.integrate(d, rgb, T, sync=pattern(i))
        g.sync()
        gt, gw, gc = g.download()
        bad = (gw != w).reshape(dims)
        per = [int(bad[:, :, a:b].sum()) for a, b in [pkg.slab_range(dims[2], len(devs), r) for r in range(len(devs))]]
        print(devs, "bad per slab", per, "gw sum", float(gw.sum()), "w sum", float(w.sum()), "max gw", gw.max(), flush=True)
for rep in range(2):
    run([0, 1], lambda i: False)
    run([0, 1], lambda i: i % 3 == 2)
    run([0, 0, 0], lambda i: i % 3 == 2)
    run([0, 1], lambda i: i % 3 == 2)
    run([1, 0, 1], lambda i: False)
    os.environ["TSDF_B200_GROUP_NO_PEER"] = "1"
    run([0, 1], lambda i: False)
    del os.environ["TSDF_B200_GROUP_NO_PEER"]
