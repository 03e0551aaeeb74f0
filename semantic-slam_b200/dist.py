"""One-process-per-GPU plumbing for z-slab sharded volumes.

torch.distributed is used ONLY as the control plane (rendezvous, shipping the 128-byte NCCL
unique id and the per-frame pose metadata); the data plane — frame broadcast and point
gather — is NCCL inside libtsdf_b200.so (csrc/comm.cu).  Host-side logic here is covered by
world_size-2 gloo tests on CPU (tests/test_dist_cpu.py).
"""
import ctypes as C

import numpy as np

from . import capi
from .capi import check, ptr, f32
from .volume import Volume, make_config, slab_range


def exchange_uid(make_uid, rank, dist_mod, device=None):
    """Rank 0 calls make_uid() -> 128 bytes; everyone receives them via torch.distributed."""
    import torch
    buf = torch.zeros(128, dtype=torch.uint8)
    if rank == 0:
        uid = make_uid()
        assert len(uid) == 128
        buf = torch.frombuffer(bytearray(uid), dtype=torch.uint8).clone()
    if device is not None:
        buf = buf.to(device)
    dist_mod.broadcast(buf, src=0)
    return bytes(buf.cpu().numpy().tobytes())


def broadcast_array(arr, rank, dist_mod, device=None):
    """Control-plane broadcast of a small float32 array (poses, intrinsics) from rank 0."""
    import torch
    t = torch.from_numpy(np.ascontiguousarray(arr, np.float32).copy())
    if device is not None:
        t = t.to(device)
    dist_mod.broadcast(t, src=0)
    return t.cpu().numpy()


def plan_slabs(dim_z, world):
    """[(z_begin, z_end)] per rank; contiguous, 4-slice aligned, covering [0, dim_z)."""
    return [slab_range(dim_z, world, r) for r in range(world)]


def merge_by_index(parts):
    """parts: list of (xyz, rgb, idx) with each idx ascending -> merged ascending by idx.
    Host-side mirror of the k-way merge the library does at the root (test helper)."""
    xyz = np.concatenate([p[0] for p in parts])
    rgb = np.concatenate([p[1] for p in parts])
    idx = np.concatenate([p[2] for p in parts])
    order = np.argsort(idx, kind="stable")
    return xyz[order], rgb[order], idx[order]


def make_nccl_uid():
    uid = (C.c_uint8 * 128)()
    check(capi.load().tsdf_comm_unique_id(uid))
    return bytes(uid)


class ShardedVolume:
    """This rank's z-slab of a grid + its NCCL communicator."""

    def __init__(self, cfg, rank, world, device, uid):
        self.rank, self.world = rank, world
        zr = slab_range(cfg.dim_z, world, rank)
        cfg.z_begin, cfg.z_end = zr
        cfg.device = device
        self.vol = Volume(cfg)
        self.lib = capi.load()
        self.comm = C.c_void_p()
        ub = (C.c_uint8 * 128).from_buffer_copy(uid)
        check(self.lib.tsdf_comm_create(ub, world, rank, device, C.byref(self.comm)))
        self._keep = []

    def integrate_bcast(self, depth, rgb, T_wc, K=None, obs_weight=1.0, root=0):
        T = f32(T_wc, 16)
        Kc = None if K is None else f32(K, 9)
        self._keep.append((depth, rgb, T, Kc))
        check(self.lib.tsdf_integrate_bcast_async(self.vol.h, self.comm, root, ptr(depth), ptr(rgb),
                                                  ptr(Kc), ptr(T), np.float32(obs_weight)))

    def peer_pull(self):
        """True when frames are distributed by the fused peer-pull prep (else ncclBroadcast)."""
        a = C.c_int(0)
        check(self.lib.tsdf_comm_peer_pull(self.comm, C.byref(a)))
        return bool(a.value)

    def distribution(self):
        return ("peer-pull fused into the prep kernel (NVLink, cudaIpc)" if self.peer_pull() else "ncclBroadcast")

    def sync(self):
        self.vol.sync()
        self._keep.clear()

    def gather_mesh(self, root=0):
        """Collective: verts, faces, norms, colors of the WHOLE grid at the root (None elsewhere)."""
        nv, nf = C.c_int64(0), C.c_int64(0)
        args = (self.vol.h, self.comm, root)
        check(self.lib.tsdf_gather_mesh(*args, None, None, None, 0, None, 0, C.byref(nv), C.byref(nf)))
        if self.rank != root:
            check(self.lib.tsdf_gather_mesh(*args, None, None, None, 0, None, 0, C.byref(nv), C.byref(nf)))
            return None
        nv, nf = nv.value, nf.value
        verts = np.zeros((max(nv, 1), 3), np.float32)
        norms = np.zeros((max(nv, 1), 3), np.float32)
        cols = np.zeros((max(nv, 1), 3), np.uint8)
        faces = np.zeros((max(nf, 1), 3), np.int32)
        cv, cf = C.c_int64(0), C.c_int64(0)
        check(self.lib.tsdf_gather_mesh(*args, ptr(verts), ptr(norms), ptr(cols), nv, ptr(faces), nf, C.byref(cv), C.byref(cf)))
        return verts[:nv], faces[:nf], norms[:nv], cols[:nv]

    def gather_points(self, pred=capi.PRED_NONZERO, tsdf_thresh=1.2, weight_thresh=0.9, root=0, cap=None, pool=None):
        n = C.c_int64(0)
        if cap is None:  # collective sizing call first
            check(self.lib.tsdf_gather_points(self.vol.h, self.comm, root, pred, np.float32(tsdf_thresh),
                                              np.float32(weight_thresh), None, None, None, -1, C.byref(n)))
            cap = n.value
        if self.rank != root:
            check(self.lib.tsdf_gather_points(self.vol.h, self.comm, root, pred, np.float32(tsdf_thresh),
                                              np.float32(weight_thresh), None, None, None, 0, C.byref(n)))
            return None, None, None, n.value
        cap = int(cap)
        if pool is not None:
            xyz = pool.array("xyz", (max(cap, 1), 3), np.float32)
            rgb = pool.array("rgb", (max(cap, 1), 3), np.uint8)
            idx = pool.array("idx", (max(cap, 1),), np.int64)
        else:
            xyz = np.zeros((max(cap, 1), 3), np.float32)
            rgb = np.zeros((max(cap, 1), 3), np.uint8)
            idx = np.zeros(max(cap, 1), np.int64)
        check(self.lib.tsdf_gather_points(self.vol.h, self.comm, root, pred, np.float32(tsdf_thresh),
                                          np.float32(weight_thresh), ptr(xyz), ptr(rgb), ptr(idx), cap, C.byref(n)))
        return xyz[:n.value], rgb[:n.value], idx[:n.value], n.value

    def close(self):
        if self.comm:
            self.lib.tsdf_comm_destroy(self.comm)
            self.comm = None
        self.vol.close()


def plan_instances(sizes, world):
    """Instance volumes (BASELINE config 5) shard by instance, not by z: largest
    first, each to the rank with the fewest voxels so far (ties: lowest rank).  Deterministic, so
    every rank computes the same plan without communication.  Returns owner[i] for every instance."""
    load = [0] * world
    owner = [0] * len(sizes)
    for i in sorted(range(len(sizes)), key=lambda k: (-int(sizes[k]), k)):
        r = min(range(world), key=lambda k: (load[k], k))
        owner[i] = r
        load[r] += int(sizes[i])
    return owner


def share_of_active(local, active, boxes=None, crops=None, masks=None):
    """One rank's share of a keyframe's active list.  local: {global instance -> index in this
    rank's batch}.  Returns (local indices, boxes, concatenated crops, full masks) of the active
    instances this rank owns, in the order of `active`."""
    keep = [k for k, g in enumerate(active) if g in local]
    if len(keep) == len(active) and boxes is not None and masks is None:   # this rank owns them all: nothing to cut
        return (np.array([local[g] for g in active], np.int32), np.ascontiguousarray(boxes, np.int32).reshape(-1, 4),
                np.ascontiguousarray(crops, np.uint8), None)
    loc = np.array([local[active[k]] for k in keep], np.int32)
    bx = cr = mk = None
    if boxes is not None:
        b = np.asarray(boxes, np.int32).reshape(-1, 4)
        off = np.concatenate([[0], np.cumsum(b[:, 2].astype(np.int64) * b[:, 3])])
        bx = np.ascontiguousarray(b[keep])
        parts = [np.asarray(crops)[off[k]:off[k + 1]] for k in keep]
        cr = np.ascontiguousarray(np.concatenate(parts) if parts else np.zeros(0, np.uint8), np.uint8)
    if masks is not None:
        mk = np.ascontiguousarray(np.asarray(masks)[keep])
    return loc, bx, cr, mk


class ShardedBatch:
    """Per-object volumes of one keyframe spread over the ranks by instance.  No data-path
    collective: a keyframe's depth / colour / masks are on every rank's host (in the off-line
    replay of examples/label_instance_rgbd.cpp:79-110 every rank reads the same files), each rank
    launches one batched integrate over the active instances it owns."""

    def __init__(self, cfgs, rank, world, device):
        from .volume import Batch
        self.rank, self.world = rank, world
        sizes = [c.dim_x * c.dim_y * c.dim_z for c in cfgs]
        self.owner = plan_instances(sizes, world)
        self.mine = [i for i, r in enumerate(self.owner) if r == rank]
        self.local = {g: k for k, g in enumerate(self.mine)}
        for i in self.mine:
            cfgs[i].device = int(device)
        self.batch = Batch([cfgs[i] for i in self.mine]) if self.mine else None

    def split_active(self, active, boxes=None, crops=None, masks=None):
        return share_of_active(self.local, active, boxes, crops, masks)

    def integrate_boxes(self, depth, rgb, active, boxes, crops, T_wc, K=None, obs_weight=1.0, sync=False):
        loc, bx, cr, _ = self.split_active(active, boxes, crops)
        if len(loc):
            self.batch.integrate_boxes(depth, rgb, loc, bx, cr, T_wc, K, obs_weight, sync)
        return len(loc)

    def integrate(self, depth, rgb, active, masks, T_wc, K=None, obs_weight=1.0, sync=False):
        loc, _, _, mk = self.split_active(active, masks=masks)
        if len(loc):
            self.batch.integrate(depth, rgb, loc, mk, T_wc, K, obs_weight, sync)
        return len(loc)

    def sync(self):
        if self.batch is not None:
            self.batch.sync()

    def voxels_written(self):
        """This rank's count for its last launch (sum over ranks = the keyframe's total)."""
        return self.batch.stats().voxels_written if self.batch is not None else 0

    def download(self, i):
        """Instance i (global index); only on its owner."""
        assert i in self.local, f"instance {i} lives on rank {self.owner[i]}"
        return self.batch.download(self.local[i])

    def close(self):
        if self.batch is not None:
            self.batch.close()
            self.batch = None
