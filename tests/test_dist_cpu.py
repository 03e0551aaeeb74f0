"""Host-side logic of the N>1 path on CPU: world_size-2 gloo (control plane only — the data
plane is NCCL on the GPU box, tests/test_multigpu.py)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import importlib
    pkg = importlib.import_module("semantic-slam_b200")
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # 1. the 128-byte id made on rank 0 reaches every rank unchanged
        uid = pkg.dist.exchange_uid(lambda: bytes(range(128)), rank, dist)
        assert uid == bytes(range(128))
        # 2. pose metadata broadcast
        poses = np.arange(32, dtype=np.float32).reshape(2, 16) if rank == 0 else np.zeros((2, 16), np.float32)
        got = pkg.dist.broadcast_array(poses, rank, dist)
        assert np.array_equal(got, np.arange(32, dtype=np.float32).reshape(2, 16))
        # 3. every rank derives the same slab plan and the slabs tile the grid
        plan = pkg.dist.plan_slabs(1024, world)
        assert plan[0][0] == 0 and plan[-1][1] == 1024 and all(a[1] == b[0] for a, b in zip(plan, plan[1:]))
        mine = pkg.slab_range(1024, world, rank)
        assert mine == plan[rank]
        # 4. merge-by-index of per-rank point lists (z-fast slabs interleave)
        rng = np.random.default_rng(5)
        idx_all = np.sort(rng.choice(10000, 500, replace=False)).astype(np.int64)
        parts = []
        for r in range(world):
            sel = idx_all[idx_all % world == r]
            parts.append((np.stack([sel] * 3, 1).astype(np.float32), np.zeros((len(sel), 3), np.uint8), sel))
        xyz, rgb, idx = pkg.dist.merge_by_index(parts)
        assert np.array_equal(idx, idx_all) and np.array_equal(xyz[:, 0], idx_all.astype(np.float32))
        # 5. instance volumes shard by instance: same plan on every rank, every instance owned once,
        #    loads within one largest volume of each other; a keyframe's active list, boxes and
        #    concatenated mask crops split into per-rank shares that together cover it exactly
        sizes = [int(x) for x in rng.integers(16, 129, 37) ** 3]
        owner = pkg.dist.plan_instances(sizes, world)
        t = torch.tensor(owner)
        dist.broadcast(t, src=0)
        assert t.tolist() == owner and set(owner) <= set(range(world))
        loads = [sum(sz for sz, o in zip(sizes, owner) if o == r) for r in range(world)]
        assert max(loads) - min(loads) <= max(sizes)
        active = [int(x) for x in rng.permutation(37)[:20]]
        boxes = rng.integers(1, 40, (20, 4)).astype(np.int32)
        crops = np.concatenate([np.full(int(b[2]) * int(b[3]), g, np.uint8) for g, b in zip(active, boxes)])
        seen = []
        for r in range(world):
            mine = [i for i, o in enumerate(owner) if o == r]
            local = {g: k for k, g in enumerate(mine)}
            loc, bx, cr, _ = pkg.dist.share_of_active(local, active, boxes, crops)
            off = 0
            for k, b in zip(loc, bx):
                g = mine[int(k)]
                n = int(b[2]) * int(b[3])
                assert (cr[off:off + n] == g).all() and np.array_equal(b, boxes[active.index(g)])
                off += n
                seen.append(g)
            assert off == len(cr)
        assert sorted(seen) == sorted(active)
        q.put((rank, "ok"))
    except Exception as e:  # pragma: no cover
        q.put((rank, repr(e)))
    finally:
        dist.destroy_process_group()


def test_control_plane_world2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, 29533, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, "ok"), (1, "ok")], res
