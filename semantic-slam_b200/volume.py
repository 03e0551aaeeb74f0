"""Host-side mirror of the reference's operator surface for the TSDF hot path, over the C ABI.

  Volume      thin RAII wrapper of a `tsdf_vol*` handle (one grid or one z-slab on one GPU)
  TSDF        mirrors class TSDF (reference include/tsdf.hpp:22-98): TSDF(h, w, id,
              base2world, origin), Integrate(depth, cam2world), host mirrors
              voxel_grid_TSDF / voxel_grid_weight, and the destructor's .ply/.bin dump
              (src/tsdf.cu:98-133) as close()
  TSDFVolume  mirrors the plugin surface reached through class TSDFfusion
              (src/TSDFfusion.py.in:21-53 -> fusion.TSDFVolume): TSDFVolume(vol_bnds,
              voxel_size), integrate(color_im, depth_im, cam_intr, cam_pose, obs_weight),
              get_volume(), get_point_cloud()

All compute happens in libtsdf_b200.so on the GPU; nothing here computes a voxel.
"""
import ctypes as C

import numpy as np

from . import capi
from .capi import (Config, Stats, check, f32, ptr, SEM_TSDFCU, SEM_FUSIONPY, PRED_NONZERO,
                   PRED_ABS_LT, DEPTH_F32, DEPTH_U16)


def default_config():
    cfg = Config()
    capi.load().tsdf_config_default(C.byref(cfg))
    return cfg


def config_from_bounds(vol_bnds, voxel_size):
    cfg = Config()
    b = (C.c_double * 6)(*np.asarray(vol_bnds, np.float64).reshape(6))
    check(capi.load().tsdf_config_from_bounds(b, float(voxel_size), C.byref(cfg)))
    return cfg


def make_config(dims, voxel_size, origin, semantics=SEM_FUSIONPY, with_color=True, trunc=None,
                im_hw=(480, 640), K=None, max_depth=None, base2world=None, device=0, z_range=None,
                vol_id=0):
    cfg = default_config()
    cfg.dim_x, cfg.dim_y, cfg.dim_z = [int(d) for d in dims]
    cfg.voxel_size = np.float32(voxel_size)
    cfg.trunc_margin = np.float32(trunc) if trunc is not None else np.float32(np.float32(voxel_size) * np.float32(5))
    cfg.origin[:] = [float(np.float32(o)) for o in origin]
    cfg.im_h, cfg.im_w = int(im_hw[0]), int(im_hw[1])
    if K is not None:
        cfg.K[:] = [float(k) for k in np.asarray(K, np.float32).reshape(9)]
    if max_depth is None:
        max_depth = 6.0 if semantics == SEM_TSDFCU else np.inf
    cfg.max_depth = np.float32(max_depth)
    cfg.semantics = int(semantics)
    cfg.with_color = int(bool(with_color))
    if base2world is not None:
        cfg.base2world[:] = [float(x) for x in np.asarray(base2world, np.float32).reshape(16)]
    cfg.device = int(device)
    if z_range is not None:
        cfg.z_begin, cfg.z_end = int(z_range[0]), int(z_range[1])
    cfg.id = int(vol_id)
    return cfg


def slab_range(dim_z, nranks, rank):
    a, b = C.c_int(0), C.c_int(0)
    capi.load().tsdf_slab_range(int(dim_z), int(nranks), int(rank), C.byref(a), C.byref(b))
    return a.value, b.value


def device_count():
    n = C.c_int(0)
    rc = capi.load().tsdf_device_count(C.byref(n))
    return n.value if rc == capi.OK else 0


class PinnedPool:
    """Page-locked numpy arrays from tsdf_host_alloc, grown on demand and kept: readback calls copy
    straight into them (cudaMemcpyAsync, no staging), uploads from them skip the staging copy."""

    def __init__(self):
        self.lib = capi.load()
        self._bufs = {}   # name -> (ptr, nbytes)

    def array(self, name, shape, dtype):
        dtype = np.dtype(dtype)
        need = max(1, int(np.prod(shape)) * dtype.itemsize)
        ptr, cap = self._bufs.get(name, (None, 0))
        if cap < need:
            if ptr:
                self.lib.tsdf_host_free(ptr)
            p = C.c_void_p()
            cap = need + need // 4
            check(self.lib.tsdf_host_alloc(C.byref(p), cap))
            ptr = p.value
            self._bufs[name] = (ptr, cap)
        raw = np.ctypeslib.as_array((C.c_uint8 * need).from_address(ptr))
        return raw.view(dtype)[:int(np.prod(shape))].reshape(shape)

    def close(self):
        for ptr, _ in self._bufs.values():
            self.lib.tsdf_host_free(ptr)
        self._bufs = {}


class Volume:
    """One `tsdf_vol*`.  Arrays returned by download() are in the semantics' reference order
    (TSDFCU: [z][y][x]; FUSIONPY: [x][y][z]) restricted to this handle's z-slab."""

    def __init__(self, cfg):
        self.lib = capi.load()
        self.h = C.c_void_p()
        check(self.lib.tsdf_create(C.byref(cfg), C.byref(self.h)))
        self.cfg = Config()
        check(self.lib.tsdf_get_config(self.h, C.byref(self.cfg)))
        self._keep = []  # keeps borrowed input arrays alive until sync()

    # -- lifecycle --------------------------------------------------------------------
    def close(self):
        if getattr(self, "h", None) is not None and self.h:
            self.lib.tsdf_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def reset(self):
        check(self.lib.tsdf_reset(self.h))

    # -- shape helpers ------------------------------------------------------------------
    @property
    def dims(self):
        return (self.cfg.dim_x, self.cfg.dim_y, self.cfg.dim_z)

    @property
    def dz_local(self):
        return self.cfg.z_end - self.cfg.z_begin

    @property
    def n_local(self):
        return self.cfg.dim_x * self.cfg.dim_y * self.dz_local

    def local_shape(self):
        if self.cfg.semantics == SEM_TSDFCU:
            return (self.dz_local, self.cfg.dim_y, self.cfg.dim_x)
        return (self.cfg.dim_x, self.cfg.dim_y, self.dz_local)

    # -- integration -------------------------------------------------------------------
    def _frame_args(self, depth, rgb, K, T_wc, dtype=np.float32):
        H, W = self.cfg.im_h, self.cfg.im_w
        if not isinstance(depth, (int, np.integer)):
            depth = np.ascontiguousarray(depth, dtype)
            assert depth.size == H * W, "depth shape mismatch"
        if rgb is not None and not isinstance(rgb, (int, np.integer)):
            rgb = np.ascontiguousarray(rgb, np.uint8)
            assert rgb.size == H * W * 3, "rgb shape mismatch"
        Kc = None if K is None else f32(K, 9)
        T = f32(T_wc, 16)
        self._keep.append((depth, rgb, Kc, T))
        return depth, rgb, Kc, T

    def integrate(self, depth, rgb, T_wc, K=None, obs_weight=1.0, sync=True):
        depth, rgb, Kc, T = self._frame_args(depth, rgb, K, T_wc)
        fn = self.lib.tsdf_integrate if sync else self.lib.tsdf_integrate_async
        check(fn(self.h, ptr(depth), ptr(rgb), ptr(Kc), ptr(T), np.float32(obs_weight)))
        if sync:
            self._keep.clear()

    def integrate_raw(self, depth_u16, depth_scale, rgb, T_wc, K=None, obs_weight=1.0, subsample=(1, 1), fmt=DEPTH_U16):
        """fmt: DEPTH_U16 / DEPTH_F32 raw depth, DEPTH_RANGE_* lidar range images."""
        dt = np.uint16 if fmt in (DEPTH_U16, capi.DEPTH_RANGE_U16) else np.float32
        depth, rgb, Kc, T = self._frame_args(depth_u16, rgb, K, T_wc, dtype=dt)
        check(self.lib.tsdf_integrate_raw_async(self.h, ptr(depth), int(fmt), np.float32(depth_scale),
                                                int(subsample[0]), int(subsample[1]), ptr(rgb), ptr(Kc),
                                                ptr(T), np.float32(obs_weight)))
        self.sync()

    def integrate_multi(self, depths, rgbs, T_wcs, K=None, obs_weight=1.0, sync=True):
        k = len(T_wcs)
        d = np.ascontiguousarray(np.stack([np.asarray(x, np.float32) for x in depths]))
        c = None if rgbs is None else np.ascontiguousarray(np.stack([np.asarray(x, np.uint8) for x in rgbs]))
        T = np.ascontiguousarray(np.stack([f32(t, 16) for t in T_wcs]))
        Kc = None if K is None else f32(K, 9)
        self._keep.append((d, c, T, Kc))
        check(self.lib.tsdf_integrate_multi_async(self.h, k, ptr(d), ptr(c), ptr(Kc), ptr(T), np.float32(obs_weight)))
        if sync:
            self.sync()

    def sync(self):
        check(self.lib.tsdf_sync(self.h))
        self._keep.clear()

    def stats(self):
        s = Stats()
        check(self.lib.tsdf_get_stats(self.h, C.byref(s)))
        self._keep.clear()
        return s

    def bench_kernel(self, iters):
        ms = C.c_float(0)
        check(self.lib.tsdf_bench_kernel(self.h, int(iters), C.byref(ms)))
        return ms.value

    def timer_start(self):
        check(self.lib.tsdf_timer_start(self.h))

    def timer_stop(self):
        ms = C.c_float(0)
        check(self.lib.tsdf_timer_stop(self.h, C.byref(ms)))
        self._keep.clear()
        return ms.value

    def kernel_timing(self, enable=True):
        check(self.lib.tsdf_kernel_timing(self.h, int(enable)))

    def kernel_timing_read(self, reset=True):
        s, n = C.c_double(0), C.c_int64(0)
        check(self.lib.tsdf_kernel_timing_read(self.h, C.byref(s), C.byref(n), int(reset)))
        return s.value, n.value

    def trace_enable(self, enable=True):
        check(self.lib.tsdf_trace_enable(self.h, int(enable)))

    def trace_read(self):
        """(n_frames, 4) ms relative to the first traced event: copy begin, frame ready, kernel begin, end."""
        n = C.c_int(0)
        check(self.lib.tsdf_trace_read(self.h, None, 0, C.byref(n)))
        out = np.zeros((max(n.value, 1), 4), np.float32)
        check(self.lib.tsdf_trace_read(self.h, ptr(out), n.value, C.byref(n)))
        return out[:n.value]

    def set_variant(self, variant):
        check(self.lib.tsdf_set_variant(self.h, int(variant)))

    # -- readback ------------------------------------------------------------------------
    def download(self, want_color=None):
        n = self.n_local
        t = np.empty(n, np.float32)
        w = np.empty(n, np.float32)
        has_c = bool(self.cfg.with_color) if want_color is None else want_color
        c = np.empty(n, np.float32) if has_c else None
        check(self.lib.tsdf_download(self.h, ptr(t), ptr(w), ptr(c), 0))
        return t, w, c

    def download_range(self, first_voxel, n_voxels, want_color=None):
        """Voxels [first_voxel, first_voxel + n_voxels) of the local slab in storage order."""
        t = np.empty(n_voxels, np.float32)
        w = np.empty(n_voxels, np.float32)
        has_c = bool(self.cfg.with_color) if want_color is None else want_color
        c = np.empty(n_voxels, np.float32) if has_c else None
        check(self.lib.tsdf_download_range(self.h, int(first_voxel), int(n_voxels), ptr(t), ptr(w), ptr(c)))
        return t, w, c

    def download_into(self, tsdf, weight, color):
        """Scatter this slab into GLOBAL host arrays (global_layout=1)."""
        check(self.lib.tsdf_download(self.h, ptr(tsdf), ptr(weight), ptr(color), 1))

    def upload(self, tsdf=None, weight=None, color=None, global_layout=False):
        a = [None if x is None else f32(x) for x in (tsdf, weight, color)]
        check(self.lib.tsdf_upload(self.h, ptr(a[0]), ptr(a[1]), ptr(a[2]), int(global_layout)))

    def extract_count(self, pred=PRED_NONZERO, tsdf_thresh=1.2, weight_thresh=0.9):
        n = C.c_int64(0)
        check(self.lib.tsdf_extract_points(self.h, pred, np.float32(tsdf_thresh), np.float32(weight_thresh),
                                           None, None, None, 0, C.byref(n)))
        return n.value

    def extract_points(self, pred=PRED_NONZERO, tsdf_thresh=1.2, weight_thresh=0.9, pool=None, out=None):
        """pool: a PinnedPool -> the outputs are views of page-locked memory owned by the pool (valid
        until the next call with the same pool): the D2H copies land in them directly.
        out: (xyz, rgb, idx) arrays of a previous call with the same point count to write into."""
        n = self.extract_count(pred, tsdf_thresh, weight_thresh)
        if out is not None:
            xyz, rgb, idx = out
            assert len(idx) == n and xyz.shape == (n, 3) and rgb.shape == (n, 3)
        elif pool is not None:
            xyz = pool.array("xyz", (max(n, 1), 3), np.float32)
            rgb = pool.array("rgb", (max(n, 1), 3), np.uint8)
            idx = pool.array("idx", (max(n, 1),), np.int64)
        else:
            xyz = np.zeros((max(n, 1), 3), np.float32)
            rgb = np.zeros((max(n, 1), 3), np.uint8)
            idx = np.zeros(max(n, 1), np.int64)
        cnt = C.c_int64(0)
        check(self.lib.tsdf_extract_points(self.h, pred, np.float32(tsdf_thresh), np.float32(weight_thresh),
                                           ptr(xyz), ptr(rgb), ptr(idx), n, C.byref(cnt)))
        assert cnt.value == n
        return xyz[:n], rgb[:n], idx[:n]

    def mesh_count(self):
        nv, nf = C.c_int64(0), C.c_int64(0)
        check(self.lib.tsdf_extract_mesh(self.h, None, None, None, 0, None, 0, C.byref(nv), C.byref(nf)))
        return nv.value, nf.value

    def extract_mesh(self, pool=None, out=None):
        """Level-0 marching cubes on the device: verts (nv,3) f32, faces (nf,3) i32, norms (nv,3) f32,
        colors (nv,3) u8 — the tuple order of fusion.TSDFVolume.get_mesh()."""
        nv, nf = self.mesh_count()
        if out is not None:
            verts, faces, norms, cols = out
            assert verts.shape == (nv, 3) and faces.shape == (nf, 3)
        elif pool is not None:
            verts = pool.array("verts", (max(nv, 1), 3), np.float32)
            norms = pool.array("norms", (max(nv, 1), 3), np.float32)
            cols = pool.array("cols", (max(nv, 1), 3), np.uint8)
            faces = pool.array("faces", (max(nf, 1), 3), np.int32)
        else:
            verts = np.zeros((max(nv, 1), 3), np.float32)
            norms = np.zeros((max(nv, 1), 3), np.float32)
            cols = np.zeros((max(nv, 1), 3), np.uint8)
            faces = np.zeros((max(nf, 1), 3), np.int32)
        cv, cf = C.c_int64(0), C.c_int64(0)
        check(self.lib.tsdf_extract_mesh(self.h, ptr(verts), ptr(norms), ptr(cols), nv, ptr(faces), nf,
                                         C.byref(cv), C.byref(cf)))
        assert (cv.value, cf.value) == (nv, nf)
        return verts[:nv], faces[:nf], norms[:nv], cols[:nv]

    def save_mesh_ply(self, path):
        check(self.lib.tsdf_save_mesh_ply(self.h, path.encode()))

    def save_ply(self, path, pred=PRED_NONZERO, tsdf_thresh=1.2, weight_thresh=0.9):
        check(self.lib.tsdf_save_ply(self.h, path.encode(), pred, np.float32(tsdf_thresh), np.float32(weight_thresh)))

    def save_bin(self, path):
        check(self.lib.tsdf_save_bin(self.h, path.encode()))

    def load_bin(self, path):
        check(self.lib.tsdf_load_bin(self.h, path.encode()))


class Group:
    """One `tsdf_group*`: a whole grid as z-slabs over a device list, ONE process (the C++ host's
    multi-GPU handle).  Arrays and point / mesh readbacks are those of one whole-grid Volume."""

    def __init__(self, cfg, devices):
        self.lib = capi.load()
        self.h = C.c_void_p()
        devs = (C.c_int32 * len(devices))(*[int(d) for d in devices])
        check(self.lib.tsdf_group_create(C.byref(cfg), devs, len(devices), C.byref(self.h)))
        self.cfg, self.n = cfg, len(devices)
        self._keep = []

    def close(self):
        if getattr(self, "h", None) is not None and self.h:
            self.lib.tsdf_group_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    @property
    def n_voxels(self):
        return self.cfg.dim_x * self.cfg.dim_y * self.cfg.dim_z

    def peer_access(self):
        n = C.c_int(0)
        check(self.lib.tsdf_group_peer_access(self.h, C.byref(n)))
        return n.value

    def integrate(self, depth, rgb, T_wc, K=None, obs_weight=1.0, sync=True):
        if not isinstance(depth, (int, np.integer)):
            depth = np.ascontiguousarray(depth, np.float32)
        if rgb is not None and not isinstance(rgb, (int, np.integer)):
            rgb = np.ascontiguousarray(rgb, np.uint8)
        Kc = None if K is None else f32(K, 9)
        T = f32(T_wc, 16)
        self._keep.append((depth, rgb, Kc, T))
        fn = self.lib.tsdf_group_integrate if sync else self.lib.tsdf_group_integrate_async
        check(fn(self.h, ptr(depth), ptr(rgb), ptr(Kc), ptr(T), np.float32(obs_weight)))
        if sync:
            self._keep.clear()

    def sync(self):
        check(self.lib.tsdf_group_sync(self.h))
        self._keep.clear()

    def reset(self):
        check(self.lib.tsdf_group_reset(self.h))

    def stats(self):
        s = Stats()
        check(self.lib.tsdf_group_get_stats(self.h, C.byref(s)))
        self._keep.clear()
        return s

    def download(self):
        n = self.n_voxels
        t, w = np.empty(n, np.float32), np.empty(n, np.float32)
        c = np.empty(n, np.float32) if self.cfg.with_color else None
        check(self.lib.tsdf_group_download(self.h, ptr(t), ptr(w), ptr(c)))
        return t, w, c

    def extract_points(self, pred=PRED_NONZERO, tsdf_thresh=1.2, weight_thresh=0.9, pool=None):
        n = C.c_int64(0)
        args = (self.h, pred, np.float32(tsdf_thresh), np.float32(weight_thresh))
        check(self.lib.tsdf_group_extract_points(*args, None, None, None, 0, C.byref(n)))
        n = n.value
        mk = (lambda name, shp, dt: pool.array(name, shp, dt)) if pool is not None else (lambda name, shp, dt: np.zeros(shp, dt))
        xyz, rgb, idx = mk("xyz", (max(n, 1), 3), np.float32), mk("rgb", (max(n, 1), 3), np.uint8), mk("idx", (max(n, 1),), np.int64)
        cnt = C.c_int64(0)
        check(self.lib.tsdf_group_extract_points(*args, ptr(xyz), ptr(rgb), ptr(idx), n, C.byref(cnt)))
        assert cnt.value == n
        return xyz[:n], rgb[:n], idx[:n]

    def extract_mesh(self):
        nv, nf = C.c_int64(0), C.c_int64(0)
        check(self.lib.tsdf_group_extract_mesh(self.h, None, None, None, 0, None, 0, C.byref(nv), C.byref(nf)))
        nv, nf = nv.value, nf.value
        verts = np.zeros((max(nv, 1), 3), np.float32)
        norms = np.zeros((max(nv, 1), 3), np.float32)
        cols = np.zeros((max(nv, 1), 3), np.uint8)
        faces = np.zeros((max(nf, 1), 3), np.int32)
        cv, cf = C.c_int64(0), C.c_int64(0)
        check(self.lib.tsdf_group_extract_mesh(self.h, ptr(verts), ptr(norms), ptr(cols), nv, ptr(faces), nf,
                                               C.byref(cv), C.byref(cf)))
        return verts[:nv], faces[:nf], norms[:nv], cols[:nv]

    def release_scratch(self):
        check(self.lib.tsdf_group_release_scratch(self.h))

    def save_ply(self, path, pred=PRED_NONZERO, tsdf_thresh=1.2, weight_thresh=0.9):
        check(self.lib.tsdf_group_save_ply(self.h, path.encode(), pred, np.float32(tsdf_thresh), np.float32(weight_thresh)))

    def save_bin(self, path):
        check(self.lib.tsdf_group_save_bin(self.h, path.encode()))

    def save_mesh_ply(self, path):
        check(self.lib.tsdf_group_save_mesh_ply(self.h, path.encode()))


class Batch:
    """N per-instance volumes that share one frame per call (`tsdf_batch*`): BASELINE config 5, the
    Engine::Run mask loop (reference src/Engine.cpp:172-233) as ONE upload + ONE launch."""

    def __init__(self, cfgs):
        self.lib = capi.load()
        self.n = len(cfgs)
        arr = (Config * self.n)(*cfgs)
        self.h = C.c_void_p()
        check(self.lib.tsdf_batch_create(arr, self.n, C.byref(self.h)))
        self.dims = [(c.dim_x, c.dim_y, c.dim_z) for c in cfgs]
        self.with_color = [bool(c.with_color) for c in cfgs]
        self._keep = []

    def close(self):
        if getattr(self, "h", None) is not None and self.h:
            self.lib.tsdf_batch_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def integrate(self, depth, rgb, active, masks, T_wc, K=None, obs_weight=1.0, sync=False):
        """masks: len(active) x H x W uint8 (0/255) or None; depth/rgb/masks may be device pointers."""
        a = np.ascontiguousarray(active, np.int32)
        T = f32(T_wc, 16)
        Kc = None if K is None else f32(K, 9)
        self._keep.append((depth, rgb, a, masks, T, Kc))
        check(self.lib.tsdf_batch_integrate_async(self.h, ptr(depth), ptr(rgb), ptr(a), len(a), ptr(masks),
                                                  ptr(Kc), ptr(T), np.float32(obs_weight)))
        if sync:
            self.sync()

    def integrate_boxes(self, depth, rgb, active, boxes, crops, T_wc, K=None, obs_weight=1.0, sync=False):
        """boxes: len(active) x 4 int32 {x, y, w, h}; crops: the masks inside their boxes, concatenated."""
        a = np.ascontiguousarray(active, np.int32)
        bx = np.ascontiguousarray(boxes, np.int32).reshape(-1, 4)
        T = f32(T_wc, 16)
        Kc = None if K is None else f32(K, 9)
        self._keep.append((depth, rgb, a, bx, crops, T, Kc))
        check(self.lib.tsdf_batch_integrate_boxes_async(self.h, ptr(depth), ptr(rgb), ptr(a), len(a), ptr(bx),
                                                        ptr(crops), ptr(Kc), ptr(T), np.float32(obs_weight)))
        if sync:
            self.sync()

    @staticmethod
    def crop_masks(masks, boxes):
        """Full-frame masks + boxes {x, y, w, h} -> (clipped boxes, concatenated crops)."""
        H, W = masks.shape[1:]
        bx = np.array(boxes, np.int32).reshape(-1, 4).copy()
        x0, y0 = np.clip(bx[:, 0], 0, W), np.clip(bx[:, 1], 0, H)
        x1, y1 = np.clip(bx[:, 0] + bx[:, 2], 0, W), np.clip(bx[:, 1] + bx[:, 3], 0, H)
        bx = np.stack([x0, y0, x1 - x0, y1 - y0], 1).astype(np.int32)
        crops = np.concatenate([masks[i, b[1]:b[1] + b[3], b[0]:b[0] + b[2]].reshape(-1) for i, b in enumerate(bx)])
        return bx, np.ascontiguousarray(crops, np.uint8)

    def sync(self):
        check(self.lib.tsdf_batch_sync(self.h))
        self._keep.clear()

    def stats(self):
        s = Stats()
        check(self.lib.tsdf_batch_get_stats(self.h, C.byref(s)))
        self._keep.clear()
        return s

    def download(self, i):
        n = int(np.prod(self.dims[i]))
        t, w = np.empty(n, np.float32), np.empty(n, np.float32)
        c = np.empty(n, np.float32) if self.with_color[i] else None
        check(self.lib.tsdf_batch_download(self.h, int(i), ptr(t), ptr(w), ptr(c)))
        return t, w, c

    def extract_mesh(self, i):
        """verts, faces, norms, colors of instance i (tsdf_batch_extract_mesh)."""
        nv, nf = C.c_int64(0), C.c_int64(0)
        check(self.lib.tsdf_batch_extract_mesh(self.h, int(i), None, None, None, 0, None, 0, C.byref(nv), C.byref(nf)))
        verts = np.zeros((max(nv.value, 1), 3), np.float32)
        norms = np.zeros((max(nv.value, 1), 3), np.float32)
        cols = np.zeros((max(nv.value, 1), 3), np.uint8)
        faces = np.zeros((max(nf.value, 1), 3), np.int32)
        check(self.lib.tsdf_batch_extract_mesh(self.h, int(i), ptr(verts), ptr(norms), ptr(cols), nv.value, ptr(faces),
                                               nf.value, C.byref(nv), C.byref(nf)))
        return verts[:nv.value], faces[:nf.value], norms[:nv.value], cols[:nv.value]


class TSDF:
    """Mirror of the reference's `class TSDF` (include/tsdf.hpp:22-98).

    TSDF(h, w, id, base2world_vec, origin): 200^3 grid, 4 mm voxels, trunc = 5 voxels, TUM3
    intrinsics (include/tsdf.hpp:63-67,96).  Integrate(depth_im, cam2world_vec) as
    src/tsdf.cu:135-168.  The reference refreshes its public host mirrors only in the
    destructor (src/tsdf.cu:101-104); here `refresh()` does it on demand and `close()`
    reproduces the destructor: download, tsdf<id>.ply, tsdf<id>.bin (src/tsdf.cu:98-133)."""

    def __init__(self, h, w, id, base2world_vec, origin, device=0, dims=(200, 200, 200),
                 voxel_size=0.004):
        cfg = make_config(dims, voxel_size, origin, semantics=SEM_TSDFCU, with_color=False,
                          im_hw=(h, w), base2world=base2world_vec, device=device, vol_id=id)
        self.vol = Volume(cfg)
        self.mnId = int(id)
        n = self.vol.n_local
        self.voxel_grid_TSDF = np.ones(n, np.float32)    # src/tsdf.cu:79-80
        self.voxel_grid_weight = np.zeros(n, np.float32)  # src/tsdf.cu:81

    def Integrate(self, depth_im, cam2world_vec):
        self.vol.integrate(depth_im, None, cam2world_vec, obs_weight=1.0, sync=False)

    def refresh(self):
        t, w, _ = self.vol.download(want_color=False)
        self.voxel_grid_TSDF[:] = t
        self.voxel_grid_weight[:] = w

    def close(self, out_dir=None):
        if self.vol is None:
            return
        self.refresh()
        if out_dir is not None:
            self.vol.save_ply(f"{out_dir}/tsdf{self.mnId}.ply", PRED_NONZERO, 1.2, 0.9)  # src/tsdf.cu:109-112
            self.vol.save_bin(f"{out_dir}/tsdf{self.mnId}.bin")                           # src/tsdf.cu:116-132
        self.vol.close()
        self.vol = None


class TSDFVolume:
    """Mirror of the plugin surface (`fusion.TSDFVolume` as driven by src/TSDFfusion.py.in).

    vol_bnds: 3x2 [[xmin,xmax],[ymin,ymax],[zmin,zmax]] in metres; dims = ceil(extent/voxel_size),
    trunc = 5*voxel_size, origin = mins (src/TSDFfusion.py.in:23-29)."""

    def __init__(self, vol_bnds, voxel_size, device=0, im_hw=(480, 640), z_range=None):
        cfg = config_from_bounds(vol_bnds, voxel_size)
        cfg.device = int(device)
        cfg.im_h, cfg.im_w = int(im_hw[0]), int(im_hw[1])
        if z_range is not None:
            cfg.z_begin, cfg.z_end = int(z_range[0]), int(z_range[1])
        self.vol = Volume(cfg)
        self._vol_dim = np.array(self.vol.dims)
        self._voxel_size = float(voxel_size)
        self._trunc_margin = float(self.vol.cfg.trunc_margin)
        self._vol_origin = np.array(self.vol.cfg.origin[:], np.float32)

    def integrate(self, color_im, depth_im, cam_intr, cam_pose, obs_weight=1.0):
        """src/TSDFfusion.py.in:43 — integrate(color HxWx3 u8, depth HxW m, K 3x3, T_wc 4x4)."""
        self.vol.integrate(depth_im, color_im, cam_pose, K=cam_intr, obs_weight=obs_weight, sync=False)

    def get_volume(self):
        t, w, c = self.vol.download()
        shp = self.vol.local_shape()
        return t.reshape(shp), c.reshape(shp)

    def get_weights(self):
        return self.vol.download()[1].reshape(self.vol.local_shape())

    def get_point_cloud(self, tsdf_thresh=1.2, weight_thresh=0.9, pred=PRED_NONZERO):
        """Headless replacement of the mesh/viewer readback (src/TSDFfusion.py.in:48-53;
        viewers are unavailable on this box): xyz + rgb of the kept voxels."""
        xyz, rgb, _ = self.vol.extract_points(pred, tsdf_thresh, weight_thresh)
        return xyz, rgb

    def get_mesh(self):
        """src/TSDFfusion.py.in:52 — verts, faces, norms, colors = tsdf_vol.get_mesh(); marching cubes
        runs on the device (csrc/mesh.cu), vertex/face order is this library's."""
        return self.vol.extract_mesh()

    def close(self):
        self.vol.close()


def meshwrite(filename, verts, faces, norms, colors):
    """fusion.meshwrite (src/TSDFfusion.py.in:53): ascii PLY of a mesh already on the host.  Pure file
    formatting, no voxel is computed here; Volume.save_mesh_ply writes the same bytes from the C ABI."""
    with open(filename, "w") as f:
        f.write("ply\nformat ascii 1.0\nelement vertex %d\n" % verts.shape[0])
        f.write("property float x\nproperty float y\nproperty float z\n")
        f.write("property float nx\nproperty float ny\nproperty float nz\n")
        f.write("property uchar red\nproperty uchar green\nproperty uchar blue\n")
        f.write("element face %d\nproperty list uchar int vertex_index\nend_header\n" % faces.shape[0])
        for i in range(verts.shape[0]):
            f.write("%f %f %f %f %f %f %d %d %d\n" % (verts[i, 0], verts[i, 1], verts[i, 2], norms[i, 0],
                                                     norms[i, 1], norms[i, 2], colors[i, 0], colors[i, 1],
                                                     colors[i, 2]))
        for i in range(faces.shape[0]):
            f.write("3 %d %d %d\n" % (faces[i, 0], faces[i, 1], faces[i, 2]))
