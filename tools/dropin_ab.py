"""Drop-in call latency / throughput from PAGEABLE host memory (what the reference hands over):
  tsdf200 : class-TSDF 200^3 volume, blocking tsdf_integrate per frame (src/Object.cpp:164 pattern), median ms per call
  c2      : 512^3 S-real, tsdf_integrate_async stream of 200 frames, frames/s
  c5      : 256 instance volumes, boxes + crops, keyframes/s
usage: python tools/dropin_ab.py   (TSDF_B200_COPY_THREADS selects the staging-copy helpers)"""
import importlib, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("semantic-slam_b200")
synth = pkg.synth
out = {"copy_threads": os.environ.get("TSDF_B200_COPY_THREADS", "default")}
o_r, fr_r = synth.s_full((200,) * 3, 2.048 / 200, n_frames=2)
cfg = pkg.make_config((200,) * 3, 2.048 / 200, o_r, semantics=pkg.SEM_TSDFCU, with_color=False, max_depth=1e9)
with pkg.Volume(cfg) as v:
    for i in range(20):
        v.integrate(fr_r[i % 2][0], None, fr_r[i % 2][2])
    ts = []
    for i in range(300):
        t0 = time.perf_counter(); v.integrate(fr_r[i % 2][0], None, fr_r[i % 2][2]); ts.append(time.perf_counter() - t0)
    out["tsdf200_blocking_ms"] = round(float(np.median(ts)) * 1e3, 4)
dims, vs = (512, 512, 512), 0.01
o2, fr2 = synth.s_real(dims, vs, n_frames=12, poses=synth.load_pose_fixture()[::16])
with pkg.Volume(pkg.make_config(dims, vs, o2)) as v2:
    for i in range(12):
        v2.integrate(*fr2[i][:2], fr2[i][2], sync=False)
    v2.sync()
    t0 = time.perf_counter()
    for i in range(300):
        d, c, T = fr2[i % 12]
        v2.integrate(d, c, T, sync=False)
    v2.sync()
    out["c2_fps"] = round(300 / (time.perf_counter() - t0))
n_inst = 256
_, fr5 = synth.s_real((64, 64, 64), 0.08, n_frames=4, seed=91)
masks, boxes5 = synth.instance_masks(n_inst, seed=92)
base = fr5[0][2]
cfgs = [pkg.make_config((128,) * 3, 0.004, (0.0, 0.0, 0.5 + 0.001 * i), semantics=pkg.SEM_TSDFCU, with_color=False, base2world=base,
                        max_depth=30.0, vol_id=i) for i in range(n_inst)]
act = np.arange(n_inst, dtype=np.int32)
bx, crops = pkg.Batch.crop_masks(masks, boxes5)
with pkg.Batch(cfgs) as b:
    for i in range(4):
        b.integrate_boxes(fr5[i % 4][0], None, act, bx, crops, fr5[i % 4][2])
    b.sync()
    t0 = time.perf_counter()
    for i in range(40):
        b.integrate_boxes(fr5[i % 4][0], None, act, bx, crops, fr5[i % 4][2])
    b.sync()
    out["c5_keyframes_per_s"] = round(40 / (time.perf_counter() - t0))
print(json.dumps(out))
