"""ctypes front-end of the CPU oracle (oracle/tsdf_oracle.c) and of the reference lift
(oracle/_ref/libtsdf_ref*.so = the reference's own src/tsdf.cu lines compiled as CPU C++).

TEST INFRASTRUCTURE ONLY.  Importers allowed: tests/, __graft_entry__.smoke(), and
bench.py's cpu_baseline / --impl reference legs.  The product package never imports this.
Nothing here reads /root/reference at run time; only `make -C oracle` (build time, in the
build container) does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SEM_TSDFCU, SEM_FUSIONPY = 0, 1
PRED_NONZERO, PRED_ABS_LT = 0, 1

_f32p = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")


class OracleParams(C.Structure):
    _fields_ = [
        ("dim_x", C.c_int32), ("dim_y", C.c_int32), ("dim_z", C.c_int32),
        ("voxel_size", C.c_float), ("trunc_margin", C.c_float),
        ("origin", C.c_float * 3),
        ("im_h", C.c_int32), ("im_w", C.c_int32),
        ("K", C.c_float * 9),
        ("max_depth", C.c_float),
        ("semantics", C.c_int32),
        ("z_begin", C.c_int32), ("z_end", C.c_int32),
    ]


def build(force=False):
    """Compile liboracle*.so (and oracle/_ref when /root/reference is present)."""
    need = force or not os.path.exists(os.path.join(HERE, "liboracle.so")) \
        or not os.path.exists(os.path.join(HERE, "liboracle_omp.so")) \
        or os.path.getmtime(os.path.join(HERE, "liboracle.so")) < max(
            os.path.getmtime(os.path.join(HERE, "tsdf_oracle.c")), os.path.getmtime(os.path.join(HERE, "mc_table.inc")))
    if need:
        subprocess.check_call(["make", "-C", HERE, "liboracle.so", "liboracle_omp.so"],
                              stdout=subprocess.DEVNULL)
    if os.path.exists("/root/reference/src/tsdf.cu") and (
            force or not os.path.exists(os.path.join(HERE, "_ref", "libtsdf_ref.so"))):
        subprocess.check_call(["make", "-C", HERE, "ref"], stdout=subprocess.DEVNULL)


def _as_ptr(a, ct):
    return None if a is None else a.ctypes.data_as(C.POINTER(ct))


class Oracle:
    """The C restatement.  omp=True loads the OpenMP build (same arithmetic, all cores)."""

    def __init__(self, omp=False):
        build()
        self.lib = C.CDLL(os.path.join(HERE, "liboracle_omp.so" if omp else "liboracle.so"))
        L = self.lib
        L.oracle_integrate.restype = C.c_int64
        L.oracle_integrate.argtypes = [C.POINTER(OracleParams), _f32p, _f32p, C.c_void_p, C.c_float,
                                       _f32p, _f32p, C.c_void_p]
        L.oracle_integrate_range.restype = C.c_int64
        L.oracle_integrate_range.argtypes = [C.POINTER(OracleParams), _f32p, _f32p, C.c_void_p, C.c_float,
                                             C.c_int64, C.c_int64, _f32p, _f32p, C.c_void_p]
        L.oracle_count_trunc_ties.restype = C.c_int64
        L.oracle_count_trunc_ties.argtypes = [C.POINTER(OracleParams), _f32p, _f32p]
        L.oracle_extract.restype = C.c_int64
        L.oracle_extract.argtypes = [C.POINTER(OracleParams), _f32p, _f32p, C.c_void_p, C.c_int,
                                     C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p,
                                     C.c_int64]
        L.oracle_compose_cam2base.restype = C.c_int
        L.oracle_compose_cam2base.argtypes = [_f32p, _f32p, _f32p]
        L.oracle_invert_matrix.restype = C.c_int
        L.oracle_invert_matrix.argtypes = [_f32p, _f32p]
        L.oracle_multiply_matrix.argtypes = [_f32p, _f32p, _f32p]
        L.oracle_pack_color.argtypes = [_u8p, C.c_int64, _f32p]
        L.oracle_instance_origin.argtypes = [_f32p, C.c_void_p, C.c_int, C.c_int, _f32p, _f32p]
        L.oracle_range_to_depth.argtypes = [_f32p, C.c_int, C.c_int, _f32p, C.c_float, _f32p]
        L.oracle_range_to_depth.restype = None
        L.oracle_save_bin.restype = C.c_int
        L.oracle_save_bin.argtypes = [C.c_char_p, C.POINTER(OracleParams), _f32p]
        L.oracle_mesh.restype = C.c_int64
        L.oracle_mesh.argtypes = [C.POINTER(OracleParams), _f32p, C.c_void_p, C.c_void_p, C.c_void_p,
                                  C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.POINTER(C.c_int64)]
        L.oracle_save_mesh_ply.restype = C.c_int
        L.oracle_save_mesh_ply.argtypes = [C.c_char_p, _f32p, _f32p, _u8p, C.c_int64, C.c_void_p, C.c_int64]
        L.oracle_save_ply.restype = C.c_int
        L.oracle_save_ply.argtypes = [C.c_char_p, C.POINTER(OracleParams), _f32p, _f32p, C.c_int,
                                      C.c_float, C.c_float]

    def set_threads(self, n):
        """OpenMP threads of the omp build, set explicitly; returns what the runtime will use."""
        self.lib.oracle_set_threads(int(n))
        return int(self.lib.oracle_max_threads())

    @staticmethod
    def params(dims, voxel_size, trunc, origin, im_hw, K, max_depth, semantics, z_range=None):
        p = OracleParams()
        p.dim_x, p.dim_y, p.dim_z = [int(d) for d in dims]
        p.voxel_size = np.float32(voxel_size)
        p.trunc_margin = np.float32(trunc)
        p.origin[:] = [float(np.float32(o)) for o in origin]
        p.im_h, p.im_w = int(im_hw[0]), int(im_hw[1])
        p.K[:] = [float(k) for k in np.asarray(K, np.float32).reshape(9)]
        p.max_depth = np.float32(max_depth)
        p.semantics = int(semantics)
        z0, z1 = z_range if z_range else (0, int(dims[2]))
        p.z_begin, p.z_end = int(z0), int(z1)
        return p

    @staticmethod
    def n_local(p):
        return int(p.dim_x) * int(p.dim_y) * (int(p.z_end) - int(p.z_begin))

    def new_volume(self, p, color=False):
        n = self.n_local(p)
        t = np.ones(n, np.float32)
        w = np.zeros(n, np.float32)
        c = np.zeros(n, np.float32) if color else None
        return t, w, c

    def compose(self, base2world, cam2world):
        out = np.zeros(16, np.float32)
        ok = self.lib.oracle_compose_cam2base(np.ascontiguousarray(base2world, np.float32).reshape(16),
                                              np.ascontiguousarray(cam2world, np.float32).reshape(16), out)
        if not ok:
            raise ValueError("singular base2world")
        return out

    def invert(self, m):
        out = np.zeros(16, np.float32)
        ok = self.lib.oracle_invert_matrix(np.ascontiguousarray(m, np.float32).reshape(16), out)
        return out if ok else None

    def multiply(self, a, b):
        out = np.zeros(16, np.float32)
        self.lib.oracle_multiply_matrix(np.ascontiguousarray(a, np.float32).reshape(16),
                                        np.ascontiguousarray(b, np.float32).reshape(16), out)
        return out

    def pack_color(self, rgb):
        rgb = np.ascontiguousarray(rgb, np.uint8)
        n = rgb.size // 3
        out = np.zeros(n, np.float32)
        self.lib.oracle_pack_color(rgb.reshape(-1), n, out)
        return out

    def integrate(self, p, cam2base, depth, color_packed, obs_weight, tsdf, weight, color):
        cp = None if color_packed is None else color_packed.ctypes.data
        cv = None if color is None else color.ctypes.data
        return int(self.lib.oracle_integrate(C.byref(p), np.ascontiguousarray(cam2base, np.float32).reshape(16),
                                             np.ascontiguousarray(depth, np.float32).reshape(-1), cp,
                                             np.float32(obs_weight), tsdf, weight, cv))

    def count_trunc_ties(self, p, cam2base, depth):
        """Voxels that reach src/tsdf.cu:49 with diff == -trunc exactly (where `<` and `<=` differ)."""
        return int(self.lib.oracle_count_trunc_ties(C.byref(p), np.ascontiguousarray(cam2base, np.float32).reshape(16),
                                                    np.ascontiguousarray(depth, np.float32).reshape(-1)))

    def integrate_range(self, p, cam2base, depth, color_packed, obs_weight, first, n, tsdf, weight, color):
        """As integrate, but only the voxels [first, first + n) of the slab's storage order
        (arrays of n voxels)."""
        cp = None if color_packed is None else color_packed.ctypes.data
        cv = None if color is None else color.ctypes.data
        return int(self.lib.oracle_integrate_range(C.byref(p), np.ascontiguousarray(cam2base, np.float32).reshape(16),
                                                   np.ascontiguousarray(depth, np.float32).reshape(-1), cp,
                                                   np.float32(obs_weight), int(first), int(n), tsdf, weight, cv))

    def extract(self, p, tsdf, weight, color, pred=PRED_NONZERO, tsdf_thresh=1.2, weight_thresh=0.9):
        cv = None if color is None else color.ctypes.data
        n = int(self.lib.oracle_extract(C.byref(p), tsdf, weight, cv, pred, np.float32(tsdf_thresh),
                                        np.float32(weight_thresh), None, None, None, 0))
        xyz = np.zeros((max(n, 1), 3), np.float32)
        rgb = np.zeros((max(n, 1), 3), np.uint8)
        idx = np.zeros(max(n, 1), np.int64)
        self.lib.oracle_extract(C.byref(p), tsdf, weight, cv, pred, np.float32(tsdf_thresh),
                                np.float32(weight_thresh), xyz.ctypes.data, rgb.ctypes.data,
                                idx.ctypes.data, n)
        return xyz[:n], rgb[:n], idx[:n]

    def mesh(self, p, tsdf, color):
        """Level-0 marching cubes of the slab's tsdf array (restated get_mesh(), see tsdf_oracle.c):
        verts (nv,3) f32, faces (nf,3) i32, norms (nv,3) f32, colors (nv,3) u8."""
        cv = None if color is None else color.ctypes.data
        nf = C.c_int64(0)
        nv = int(self.lib.oracle_mesh(C.byref(p), tsdf, cv, None, None, None, 0, None, 0, C.byref(nf)))
        verts = np.zeros((max(nv, 1), 3), np.float32)
        norms = np.zeros((max(nv, 1), 3), np.float32)
        cols = np.zeros((max(nv, 1), 3), np.uint8)
        faces = np.zeros((max(nf.value, 1), 3), np.int32)
        self.lib.oracle_mesh(C.byref(p), tsdf, cv, verts.ctypes.data, norms.ctypes.data, cols.ctypes.data,
                             nv, faces.ctypes.data, nf.value, C.byref(nf))
        return verts[:nv], faces[:nf.value], norms[:nv], cols[:nv]

    def save_mesh_ply(self, path, verts, faces, norms, colors):
        return bool(self.lib.oracle_save_mesh_ply(
            path.encode(), np.ascontiguousarray(verts, np.float32).reshape(-1),
            np.ascontiguousarray(norms, np.float32).reshape(-1),
            np.ascontiguousarray(colors, np.uint8).reshape(-1), len(verts),
            np.ascontiguousarray(faces, np.int32).ctypes.data, len(faces)))

    def instance_origin(self, depth, mask, K):
        H, W = depth.shape
        out = np.zeros(3, np.float32)
        m = None if mask is None else np.ascontiguousarray(mask, np.uint8).ctypes.data
        self.lib.oracle_instance_origin(np.ascontiguousarray(depth, np.float32).reshape(-1), m, H, W,
                                        np.ascontiguousarray(K, np.float32).reshape(9), out)
        return out

    def range_to_depth(self, rng, K, scale=1.0):
        """Lidar range image -> camera-z depth (reference src/DoN.cpp:86-100, src/Object.cpp:611-619)."""
        H, W = rng.shape
        out = np.empty(H * W, np.float32)
        self.lib.oracle_range_to_depth(np.ascontiguousarray(rng, np.float32).reshape(-1), H, W,
                                       np.ascontiguousarray(K, np.float32).reshape(-1), np.float32(scale), out)
        return out.reshape(H, W)

    def save_bin(self, path, p, tsdf):
        return bool(self.lib.oracle_save_bin(path.encode(), C.byref(p), tsdf))

    def save_ply(self, path, p, tsdf, weight, pred=PRED_NONZERO, tsdf_thresh=1.2, weight_thresh=0.9):
        return bool(self.lib.oracle_save_ply(path.encode(), C.byref(p), tsdf, weight, pred,
                                             np.float32(tsdf_thresh), np.float32(weight_thresh)))


def ref_available(kind="strict"):
    name = {"strict": "libtsdf_ref.so", "omp": "libtsdf_ref_omp.so", "gpu": "libtsdf_ref_gpu.so"}[kind]
    return os.path.exists(os.path.join(HERE, "_ref", name))


class RefLift:
    """The reference's own src/tsdf.cu lines (15-60, 170-218, 253-403) compiled as CPU C++."""

    def __init__(self, omp=False):
        path = os.path.join(HERE, "_ref", "libtsdf_ref_omp.so" if omp else "libtsdf_ref.so")
        self.lib = C.CDLL(path)
        L = self.lib
        L.ref_integrate.argtypes = [_f32p, _f32p, _f32p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                    C.c_float, C.c_float, C.c_float, C.c_float, C.c_float, _f32p, _f32p]
        L.ref_multiply_matrix.argtypes = [_f32p, _f32p, _f32p]
        L.ref_invert_matrix.restype = C.c_int
        L.ref_invert_matrix.argtypes = [_f32p, _f32p]
        L.ref_save_ply.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float,
                                   C.c_float, C.c_float, _f32p, _f32p, C.c_float, C.c_float]

    def set_threads(self, n):
        self.lib.ref_set_threads(int(n))
        return int(self.lib.ref_max_threads())

    def integrate(self, K, cam2base, depth, dims, origin, vs, trunc, tsdf, weight):
        H, W = depth.shape
        self.lib.ref_integrate(np.ascontiguousarray(K, np.float32).reshape(9),
                               np.ascontiguousarray(cam2base, np.float32).reshape(16),
                               np.ascontiguousarray(depth, np.float32).reshape(-1), H, W,
                               int(dims[0]), int(dims[1]), int(dims[2]),
                               np.float32(origin[0]), np.float32(origin[1]), np.float32(origin[2]),
                               np.float32(vs), np.float32(trunc), tsdf, weight)

    def invert(self, m):
        out = np.zeros(16, np.float32)
        ok = self.lib.ref_invert_matrix(np.ascontiguousarray(m, np.float32).reshape(16), out)
        return out if ok else None

    def multiply(self, a, b):
        out = np.zeros(16, np.float32)
        self.lib.ref_multiply_matrix(np.ascontiguousarray(a, np.float32).reshape(16),
                                     np.ascontiguousarray(b, np.float32).reshape(16), out)
        return out

    def save_ply(self, path, dims, vs, origin, tsdf, weight, tsdf_thresh=1.2, weight_thresh=0.9):
        self.lib.ref_save_ply(path.encode(), int(dims[0]), int(dims[1]), int(dims[2]), np.float32(vs),
                              np.float32(origin[0]), np.float32(origin[1]), np.float32(origin[2]),
                              tsdf, weight, np.float32(tsdf_thresh), np.float32(weight_thresh))


class RefGpu:
    """The reference's own __global__ GpuIntegrate (src/tsdf.cu:15-60) compiled by nvcc for sm_100a
    with nvcc's defaults (mul+add contracted to FMA, as the reference's CMake would build it) and
    driven by the call sequence of src/tsdf.cu:161-167 (oracle/ref_gpu_epilogue.inc).  Checker and
    "bar to beat" only; needs dim_y <= 1024 (one thread per y, `<<<dim_z, dim_y>>>`)."""

    def __init__(self, dims, hw, K):
        self.lib = C.CDLL(os.path.join(HERE, "_ref", "libtsdf_ref_gpu.so"))
        L = self.lib
        L.ref_gpu_create.restype = C.c_void_p
        L.ref_gpu_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _f32p]
        L.ref_gpu_integrate.restype = C.c_float
        L.ref_gpu_integrate.argtypes = [C.c_void_p, _f32p, _f32p, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float]
        L.ref_gpu_download.restype = C.c_int
        L.ref_gpu_download.argtypes = [C.c_void_p, _f32p, _f32p]
        L.ref_gpu_destroy.argtypes = [C.c_void_p]
        assert dims[1] <= 1024
        self.dims = tuple(int(d) for d in dims)
        self.h = L.ref_gpu_create(self.dims[0], self.dims[1], self.dims[2], int(hw[0]), int(hw[1]),
                                  np.ascontiguousarray(K, np.float32).reshape(9))
        if not self.h:
            raise RuntimeError("ref_gpu_create failed (no device / out of memory)")

    def integrate(self, cam2base, depth, origin, vs, trunc):
        """Returns the CUDA-event ms of the reference's upload + launch."""
        ms = self.lib.ref_gpu_integrate(self.h, np.ascontiguousarray(cam2base, np.float32).reshape(16),
                                        np.ascontiguousarray(depth, np.float32).reshape(-1), np.float32(origin[0]),
                                        np.float32(origin[1]), np.float32(origin[2]), np.float32(vs), np.float32(trunc))
        if ms < 0:
            raise RuntimeError("reference kernel failed")
        return float(ms)

    def download(self):
        n = int(np.prod(self.dims))
        t, w = np.empty(n, np.float32), np.empty(n, np.float32)
        assert self.lib.ref_gpu_download(self.h, t, w) == 0
        return t, w

    def close(self):
        if self.h:
            self.lib.ref_gpu_destroy(self.h)
            self.h = None
