"""Worker for the multi-GPU parity test: launched by torchrun with N ranks, one GPU each.
Every rank integrates its z-slab from frames broadcast by rank 0 over NCCL, checks the slab
against the oracle's slab, and the root checks the gathered point cloud against the oracle's
whole-grid extraction.  Exits non-zero on any mismatch."""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    from oracle.oracle import Oracle, SEM_TSDFCU, SEM_FUSIONPY
    pkg = importlib.import_module("semantic-slam_b200")
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    o = Oracle()
    synth = pkg.synth
    # >= 64 slices per rank (a C3 slab is 128): the slab kernel runs whole blocks, not an 8-slice sliver
    dims, vs = (56, 40, max(72, 64 * world)), 0.04
    origin, frames = synth.s_corridor(dims, vs, n_frames=4, seed=77)   # same seed => same data on all ranks
    MD = 40.0   # the corridor is up to 20 m long: class-TSDF's default max_depth of 6 m would cut it
    for sem, color in ((SEM_FUSIONPY, True), (SEM_TSDFCU, False)):
        cfg = pkg.make_config(dims, vs, origin, semantics=sem, with_color=color, max_depth=MD)
        uid = pkg.dist.exchange_uid(pkg.dist.make_nccl_uid, rank, dist, dev)
        sv = pkg.dist.ShardedVolume(cfg, rank, world, local, uid)
        zr = (sv.vol.cfg.z_begin, sv.vol.cfg.z_end)
        assert zr == pkg.slab_range(dims[2], world, rank)
        # 10 distributed frames (the 6-slot ring wraps) with a plain single-handle integrate of a
        # frame every rank holds in between: the distribution has its own ring and sequence number
        seq = [frames[i % 4] for i in range(10)]
        for i, (d, rgb, T) in enumerate(seq):
            if i == 4:
                sv.vol.integrate(d, rgb if color else None, T, sync=False)
            elif rank == 0:   # only the root holds the images; poses are control-plane metadata
                sv.integrate_bcast(d, rgb if color else None, T)
            else:
                sv.integrate_bcast(None, None, T)
        sv.sync()
        frames_run = seq
        want_nccl = os.environ.get("TSDF_B200_BCAST") == "nccl"
        assert not (want_nccl and sv.peer_pull())
        mode = "peer-pull" if sv.peer_pull() else "nccl"
        gt, gw, gc = sv.vol.download()
        p = o.params(dims, vs, cfg.trunc_margin, origin, synth.TUM3_HW, synth.TUM3_K, cfg.max_depth, sem, zr)
        t, w, c = o.new_volume(p, color)
        for d, rgb, T in frames_run:
            o.integrate(p, T.reshape(16), d, o.pack_color(rgb) if color else None, 1.0, t, w, c)
        assert w.max() >= 2, f"rank {rank}: slab {zr} saw nothing (scene does not reach it)"
        assert np.array_equal(gw, w), f"rank {rank}: weights differ"
        assert np.array_equal(gt.view(np.uint32), t.view(np.uint32)), f"rank {rank}: tsdf differs"
        if color:
            assert np.array_equal(gc, c), f"rank {rank}: colour differs"
        xyz, rgbp, idx, n = sv.gather_points()
        pool = pkg.PinnedPool()
        xyz2, rgb2, idx2, n2 = sv.gather_points(pool=pool)      # straight into page-locked outputs
        mesh = sv.gather_mesh() if color else None              # whole-grid mesh at the root
        if rank == 0:
            pw = o.params(dims, vs, cfg.trunc_margin, origin, synth.TUM3_HW, synth.TUM3_K, cfg.max_depth, sem)
            tw, ww, cw = o.new_volume(pw, color)
            for d, rgb, T in frames_run:
                o.integrate(pw, T.reshape(16), d, o.pack_color(rgb) if color else None, 1.0, tw, ww, cw)
            oxyz, orgb, oidx = o.extract(pw, tw, ww, cw)
            assert n == len(oidx) > 1000 and np.array_equal(idx, oidx), "gathered index set differs"
            assert np.array_equal(xyz.view(np.uint32), oxyz.view(np.uint32))
            assert n2 == n and np.array_equal(idx2, oidx) and np.array_equal(xyz2.view(np.uint32), oxyz.view(np.uint32))
            if color:
                assert np.array_equal(rgbp, orgb) and np.array_equal(rgb2, orgb)
                om = o.mesh(pw, tw, cw)
                assert len(om[1]) > 1000 and np.array_equal(mesh[1], om[1]), "gathered mesh faces differ"
                assert np.array_equal(mesh[0].view(np.uint32), om[0].view(np.uint32)) and np.array_equal(mesh[3], om[3])
                assert np.array_equal(mesh[2].view(np.uint32), om[2].view(np.uint32))
        pool.close()
        sv.close()
        dist.barrier()
    # ---- instance volumes: sharded by instance, no data-path collective ----
    n_inst, idims, ivs = 7, (32, 32, 32), 0.02
    _, fr = synth.s_real((64, 64, 64), 0.08, n_frames=3, seed=93)
    masks, boxes = synth.instance_masks(n_inst, seed=94)
    origins = [o.instance_origin(fr[0][0], masks[i], synth.TUM3_K) for i in range(n_inst)]
    cfgs = [pkg.make_config(idims, ivs, origins[i], semantics=SEM_FUSIONPY, with_color=True, vol_id=i)
            for i in range(n_inst)]
    sb = pkg.dist.ShardedBatch(cfgs, rank, world, local)
    active_sets = [list(range(n_inst)), [3, 1, 6], [4, 0, 2, 5]]
    written = 0
    for (d, rgb, T), act in zip(fr, active_sets):
        bx, crops = pkg.Batch.crop_masks(masks[act], boxes[act])
        if sb.integrate_boxes(d, rgb, act, bx, crops, T, sync=True):
            written += sb.voxels_written()
    tot = torch.tensor([written], dtype=torch.int64, device=dev)
    dist.all_reduce(tot)
    want = 0
    for i in range(n_inst):
        p = o.params(idims, ivs, np.float32(ivs) * np.float32(5), origins[i], synth.TUM3_HW, synth.TUM3_K, np.inf, SEM_FUSIONPY)
        t, w, c = o.new_volume(p, True)
        for (d, rgb, T), act in zip(fr, active_sets):
            if i in act:
                md = (d * (masks[i] >= 128)).astype(np.float32)
                want += o.integrate(p, T.reshape(16), md, o.pack_color(rgb), 1.0, t, w, c)
        if sb.owner[i] == rank:
            gt, gw, gc = sb.download(i)
            assert np.array_equal(gw, w) and np.array_equal(gt.view(np.uint32), t.view(np.uint32)) and np.array_equal(gc, c), \
                f"rank {rank}: instance {i} differs"
    assert int(tot.item()) == want, (int(tot.item()), want)
    assert sorted(i for r in range(world) for i, ow in enumerate(sb.owner) if ow == r) == list(range(n_inst))
    sb.close()
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        print("dist_worker ok", world, mode)


if __name__ == "__main__":
    main()
