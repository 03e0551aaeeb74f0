"""ctypes binding of libtsdf_b200.so (include/tsdf_b200.h) — the same stub a Python host would
write; tests/ and bench.py call the product exclusively through it.

There is no CPU fallback: if the shared library is missing `load()` raises, and every compute
entry point returns TSDF_E_NODEVICE on a box without a GPU.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# TSDF_B200_LIB: A/B tooling only (tools/variants.py against an older build of the same ABI)
LIB_PATH = os.environ.get("TSDF_B200_LIB") or os.path.join(HERE, "libtsdf_b200.so")
# the same sources built with -DTSDF_B200_DEV: A/B kernel variants, tsdf_bench_kernel, the exact-math
# self-test.  tools/ and the tests that force a code path select it; the product never does.
DEV_LIB_PATH = os.environ.get("TSDF_B200_DEV_LIB") or os.path.join(HERE, "libtsdf_b200_dev.so")

OK, E_INVALID, E_CUDA, E_NOMEM, E_CAPACITY, E_IO, E_NCCL, E_SINGULAR, E_NODEVICE = range(9)
SEM_TSDFCU, SEM_FUSIONPY = 0, 1
PRED_NONZERO, PRED_ABS_LT = 0, 1
DEPTH_F32, DEPTH_U16, DEPTH_RANGE_F32, DEPTH_RANGE_U16 = 0, 1, 2, 3


class Config(C.Structure):
    _fields_ = [
        ("dim_x", C.c_int32), ("dim_y", C.c_int32), ("dim_z", C.c_int32),
        ("voxel_size", C.c_float), ("trunc_margin", C.c_float),
        ("origin", C.c_float * 3),
        ("base2world", C.c_float * 16),
        ("im_h", C.c_int32), ("im_w", C.c_int32),
        ("K", C.c_float * 9),
        ("max_depth", C.c_float),
        ("semantics", C.c_int32),
        ("with_color", C.c_int32),
        ("device", C.c_int32),
        ("z_begin", C.c_int32), ("z_end", C.c_int32),
        ("id", C.c_int32),
        ("reserved", C.c_int32 * 7),
    ]


class Stats(C.Structure):
    _fields_ = [
        ("voxels_visited", C.c_int64), ("voxels_written", C.c_int64), ("frames", C.c_int64),
        ("last_kernel_ms", C.c_float), ("last_upload_ms", C.c_float),
        ("bytes_algorithmic", C.c_int64),
    ]


# name -> (restype, argtypes); the list is also what tests/test_abi.py checks against the header
_VP, _I, _F, _I64 = C.c_void_p, C.c_int, C.c_float, C.c_int64
SIGNATURES = {
    "tsdf_version": (_I, []),
    "tsdf_last_error": (C.c_char_p, []),
    "tsdf_device_count": (_I, [C.POINTER(_I)]),
    "tsdf_config_default": (None, [C.POINTER(Config)]),
    "tsdf_config_from_bounds": (_I, [C.POINTER(C.c_double), C.c_double, C.POINTER(Config)]),
    "tsdf_create": (_I, [C.POINTER(Config), C.POINTER(_VP)]),
    "tsdf_destroy": (_I, [_VP]),
    "tsdf_reset": (_I, [_VP]),
    "tsdf_get_config": (_I, [_VP, C.POINTER(Config)]),
    "tsdf_integrate": (_I, [_VP, _VP, _VP, _VP, _VP, _F]),
    "tsdf_integrate_async": (_I, [_VP, _VP, _VP, _VP, _VP, _F]),
    "tsdf_integrate_raw_async": (_I, [_VP, _VP, _I, _F, _I, _I, _VP, _VP, _VP, _F]),
    "tsdf_integrate_multi_async": (_I, [_VP, _I, _VP, _VP, _VP, _VP, _F]),
    "tsdf_sync": (_I, [_VP]),
    "tsdf_get_stats": (_I, [_VP, C.POINTER(Stats)]),
    "tsdf_download": (_I, [_VP, _VP, _VP, _VP, _I]),
    "tsdf_download_range": (_I, [_VP, _I64, _I64, _VP, _VP, _VP]),
    "tsdf_upload": (_I, [_VP, _VP, _VP, _VP, _I]),
    "tsdf_extract_points": (_I, [_VP, _I, _F, _F, _VP, _VP, _VP, _I64, C.POINTER(_I64)]),
    "tsdf_extract_points_device": (_I, [_VP, _I, _F, _F, _VP, _VP, _VP, _I64, C.POINTER(_I64)]),
    "tsdf_extract_points_static": (_I, [C.POINTER(Config), _VP, _VP, _VP, _I, _F, _F, _VP, _VP, _VP, _I64, C.POINTER(_I64), _VP]),
    "tsdf_extract_mesh": (_I, [_VP, _VP, _VP, _VP, _I64, _VP, _I64, C.POINTER(_I64), C.POINTER(_I64)]),
    "tsdf_extract_mesh_device": (_I, [_VP, _VP, _VP, _VP, _I64, _VP, _I64, C.POINTER(_I64), C.POINTER(_I64)]),
    "tsdf_save_mesh_ply": (_I, [_VP, C.c_char_p]),
    "tsdf_save_ply": (_I, [_VP, C.c_char_p, _I, _F, _F]),
    "tsdf_save_bin": (_I, [_VP, C.c_char_p]),
    "tsdf_load_bin": (_I, [_VP, C.c_char_p]),
    "tsdf_instance_origin": (_I, [_I, _VP, _VP, _I, _I, _VP, _VP]),
    "tsdf_batch_create": (_I, [C.POINTER(Config), _I, C.POINTER(_VP)]),
    "tsdf_batch_destroy": (_I, [_VP]),
    "tsdf_batch_size": (_I, [_VP, C.POINTER(_I)]),
    "tsdf_batch_integrate_async": (_I, [_VP, _VP, _VP, _VP, _I, _VP, _VP, _VP, _F]),
    "tsdf_batch_integrate_boxes_async": (_I, [_VP, _VP, _VP, _VP, _I, _VP, _VP, _VP, _VP, _F]),
    "tsdf_batch_sync": (_I, [_VP]),
    "tsdf_batch_get_stats": (_I, [_VP, C.POINTER(Stats)]),
    "tsdf_batch_download": (_I, [_VP, _I, _VP, _VP, _VP]),
    "tsdf_batch_extract_points": (_I, [_VP, _I, _I, _F, _F, _VP, _VP, _VP, _I64, C.POINTER(_I64)]),
    "tsdf_batch_extract_mesh": (_I, [_VP, _I, _VP, _VP, _VP, _I64, _VP, _I64, C.POINTER(_I64), C.POINTER(_I64)]),
    "tsdf_comm_unique_id": (_I, [_VP]),
    "tsdf_comm_create": (_I, [_VP, _I, _I, _I, C.POINTER(_VP)]),
    "tsdf_comm_destroy": (_I, [_VP]),
    "tsdf_comm_peer_pull": (_I, [_VP, C.POINTER(_I)]),
    "tsdf_slab_range": (None, [_I, _I, _I, C.POINTER(_I), C.POINTER(_I)]),
    "tsdf_integrate_bcast_async": (_I, [_VP, _VP, _I, _VP, _VP, _VP, _VP, _F]),
    "tsdf_gather_points": (_I, [_VP, _VP, _I, _I, _F, _F, _VP, _VP, _VP, _I64, C.POINTER(_I64)]),
    "tsdf_gather_mesh": (_I, [_VP, _VP, _I, _VP, _VP, _VP, _I64, _VP, _I64, C.POINTER(_I64), C.POINTER(_I64)]),
    "tsdf_group_create": (_I, [C.POINTER(Config), C.POINTER(C.c_int32), _I, C.POINTER(_VP)]),
    "tsdf_group_destroy": (_I, [_VP]),
    "tsdf_group_size": (_I, [_VP, C.POINTER(_I)]),
    "tsdf_group_slab": (_I, [_VP, _I, C.POINTER(_VP)]),
    "tsdf_group_peer_access": (_I, [_VP, C.POINTER(_I)]),
    "tsdf_group_integrate": (_I, [_VP, _VP, _VP, _VP, _VP, _F]),
    "tsdf_group_integrate_async": (_I, [_VP, _VP, _VP, _VP, _VP, _F]),
    "tsdf_group_sync": (_I, [_VP]),
    "tsdf_group_reset": (_I, [_VP]),
    "tsdf_group_get_stats": (_I, [_VP, C.POINTER(Stats)]),
    "tsdf_group_download": (_I, [_VP, _VP, _VP, _VP]),
    "tsdf_group_extract_points": (_I, [_VP, _I, _F, _F, _VP, _VP, _VP, _I64, C.POINTER(_I64)]),
    "tsdf_group_extract_mesh": (_I, [_VP, _VP, _VP, _VP, _I64, _VP, _I64, C.POINTER(_I64), C.POINTER(_I64)]),
    "tsdf_group_release_scratch": (_I, [_VP]),
    "tsdf_group_save_ply": (_I, [_VP, C.c_char_p, _I, _F, _F]),
    "tsdf_group_save_bin": (_I, [_VP, C.c_char_p]),
    "tsdf_group_save_mesh_ply": (_I, [_VP, C.c_char_p]),
    "tsdf_host_alloc": (_I, [C.POINTER(_VP), C.c_size_t]),
    "tsdf_host_free": (_I, [_VP]),
    "tsdf_device_alloc": (_I, [_I, C.POINTER(_VP), C.c_size_t]),
    "tsdf_device_free": (_I, [_I, _VP]),
    "tsdf_memcpy_h2d": (_I, [_I, _VP, _VP, C.c_size_t]),
    "tsdf_memcpy_d2h": (_I, [_I, _VP, _VP, C.c_size_t]),
    "tsdf_multiply_matrix": (None, [_VP, _VP, _VP]),
    "tsdf_invert_matrix": (_I, [_VP, _VP]),
    "tsdf_read_bundle_poses": (_I, [C.c_char_p, _VP, _I, C.POINTER(_I)]),
    "tsdf_read_trajectory_kitti": (_I, [C.c_char_p, _VP, _I, C.POINTER(_I)]),
    "tsdf_read_frame_ids": (_I, [C.c_char_p, _VP, _I, C.POINTER(_I)]),
    "tsdf_read_associations": (_I, [C.c_char_p, _VP, _VP, _VP, _VP, _I, _I, C.POINTER(_I)]),
    "tsdf_timer_start": (_I, [_VP]),
    "tsdf_timer_stop": (_I, [_VP, C.POINTER(_F)]),
    "tsdf_launch_count": (_I, [C.POINTER(_I64)]),
    "tsdf_trace_enable": (_I, [_VP, _I]),
    "tsdf_trace_read": (_I, [_VP, _VP, _I, C.POINTER(_I)]),
    "tsdf_kernel_timing": (_I, [_VP, _I]),
    "tsdf_kernel_timing_read": (_I, [_VP, C.POINTER(C.c_double), C.POINTER(_I64), _I]),
}
# exported by libtsdf_b200_dev.so only (include/tsdf_b200.h, #ifdef TSDF_B200_DEV)
DEV_SIGNATURES = {
    "tsdf_bench_kernel": (_I, [_VP, _I, C.POINTER(_F)]),
    "tsdf_set_variant": (_I, [_VP, _I]),
    "tsdf_selftest_exactmath": (_I, [_I, _I64, C.c_uint64, C.POINTER(_I64)]),
    "tsdf_debug_counters": (_I, [_VP, _I, _VP]),
}

_libs = {}
_dev_selected = False


def _open(path, sigs):
    if not os.path.exists(path):
        raise RuntimeError(
            f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(make -C semantic-slam_b200/csrc). There is no CPU fallback.")
    lib = C.CDLL(path)   # RTLD_LOCAL; both libraries are linked -Bsymbolic and may coexist in one process
    for name, (res, args) in sigs.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return lib


def select(dev):
    """Chooses which library load() returns from now on (objects already created keep theirs)."""
    global _dev_selected
    _dev_selected = bool(dev)


def load(dev=None):
    """Loads the CUDA extension; raises loudly when it has not been built.  dev=None: the library
    chosen with select() (the product library unless a test / tool selected the dev build)."""
    dev = _dev_selected if dev is None else bool(dev)
    if dev not in _libs:
        _libs[dev] = _open(DEV_LIB_PATH, {**SIGNATURES, **DEV_SIGNATURES}) if dev else _open(LIB_PATH, SIGNATURES)
    return _libs[dev]


class TsdfError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"tsdf_b200 error {code}: {msg}")
        self.code = code


def check(rc):
    if rc != OK:
        raise TsdfError(rc, load().tsdf_last_error().decode(errors="replace"))


def ptr(a):
    """Host numpy array / int device pointer / None -> c_void_p-compatible value."""
    if a is None:
        return None
    if isinstance(a, (int, np.integer)):
        return int(a)
    assert a.flags["C_CONTIGUOUS"], "array must be C-contiguous"
    return a.ctypes.data


def f32(a, n=None):
    a = np.ascontiguousarray(a, np.float32).reshape(-1)
    assert n is None or a.size == n
    return a
