"""GPU test of tsdf_extract_points_static — the point extraction for hosts that cannot allocate.

Some hosts of this library may not (or cannot) let it call cudaMalloc while it runs: a mapping
loop with a frame deadline (cudaMalloc / cudaFree stall the whole device), a process whose device
memory already belongs to its own pool or to torch's caching allocator, or a caller whose voxel
arrays are its own buffers and who therefore has no tsdf_vol handle.  For them the library keeps the
compaction scratch in a fixed array inside its device image and works on caller-owned buffers only.

The test plays that caller: it owns every device buffer (allocated up front, before the calls under
test), calls the new entry point through the C ABI and requires bit-identical output to the existing
handle-based entry point, tsdf_extract_points_device, on the same voxels — on several calls in a row,
because all calls share the one scratch array and it keeps the previous call's contents (a big grid
followed by a small one, both semantics, both predicates, a z-slab).  Bit-exact: this is index /
compaction work, the tolerance is zero.

This file is named test_zz_* so that pytest collects it last: with `-x` a failure here cannot hide
the other tests.  It passes alone in a fresh process and after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

STATIC_MAX_VOXELS = 1 << 27


class DeviceBuf:
    """A device buffer owned by the test (the "caller's memory")."""

    def __init__(self, lib, device, nbytes):
        self.lib, self.device, self.nbytes = lib, device, max(int(nbytes), 16)
        self.p = C.c_void_p()
        assert lib.tsdf_device_alloc(device, C.byref(self.p), self.nbytes) == 0

    def put(self, a):
        a = np.ascontiguousarray(a)
        assert a.nbytes <= self.nbytes
        assert self.lib.tsdf_memcpy_h2d(self.device, self.p, a.ctypes.data, a.nbytes) == 0

    def get(self, shape, dtype):
        out = np.empty(shape, dtype)
        if out.nbytes:
            assert self.lib.tsdf_memcpy_d2h(self.device, out.ctypes.data, self.p, out.nbytes) == 0
        return out

    def free(self):
        if self.p:
            self.lib.tsdf_device_free(self.device, self.p)
            self.p = C.c_void_p()


def _handle_path(pkg, vol, pred, tt, wt, bufs):
    """The existing entry point: extraction on a tsdf_vol handle into device buffers."""
    lib = pkg.load()
    n = C.c_int64(0)
    args = (vol.h, pred, np.float32(tt), np.float32(wt))
    pkg.capi.check(lib.tsdf_extract_points_device(*args, None, None, None, 0, C.byref(n)))
    cnt = C.c_int64(0)
    pkg.capi.check(lib.tsdf_extract_points_device(*args, bufs["xyz"].p, bufs["rgb"].p, bufs["idx"].p, n.value, C.byref(cnt)))
    vol.sync()
    assert cnt.value == n.value
    k = n.value
    return bufs["xyz"].get((k, 3), np.float32), bufs["rgb"].get((k, 3), np.uint8), bufs["idx"].get((k,), np.int64)


def _static_path(pkg, cfg, d_t, d_w, d_c, pred, tt, wt, bufs):
    """The entry point under test: caller-owned voxel arrays, scratch inside the library image."""
    lib = pkg.load()
    n = C.c_int64(-1)
    args = (C.byref(cfg), d_t.p, d_w.p, d_c.p if d_c is not None else None, pred, np.float32(tt), np.float32(wt))
    pkg.capi.check(lib.tsdf_extract_points_static(*args, None, None, None, 0, C.byref(n), None))   # sizing call
    if n.value > 0:   # too small a capacity is reported, nothing is written
        short = C.c_int64(0)
        rc = lib.tsdf_extract_points_static(*args, bufs["xyz"].p, bufs["rgb"].p, bufs["idx"].p, n.value - 1, C.byref(short), None)
        assert rc == pkg.capi.E_CAPACITY and short.value == n.value
    cnt = C.c_int64(0)
    pkg.capi.check(lib.tsdf_extract_points_static(*args, bufs["xyz"].p, bufs["rgb"].p, bufs["idx"].p, n.value, C.byref(cnt), None))
    assert cnt.value == n.value
    k = n.value
    return bufs["xyz"].get((k, 3), np.float32), bufs["rgb"].get((k, 3), np.uint8), bufs["idx"].get((k,), np.int64)


CASES = [
    # dims, semantics, colour, predicate, tsdf_thresh, weight_thresh, z_range
    ((96, 80, 100), "FUSIONPY", True, "NONZERO", 1.2, 0.9, None),    # big one first: leaves a full scratch behind
    ((40, 36, 28), "TSDFCU", False, "NONZERO", 1.2, 0.9, None),      # small one right after it, other layout
    ((64, 48, 63), "TSDFCU", False, "ABS_LT", 0.2, 0.0, None),       # ragged: n % 4 != 0 tail path of k_mark
    ((64, 48, 64), "FUSIONPY", True, "ABS_LT", 0.3, 1.5, (16, 48)),  # a z-slab: global indices, z offset
]


def _run_case(pkg, synth, dims, sem, color, pred, tt, wt, zr, vs=0.02, n_frames=3, dev=0):
    lib = pkg.load()
    sem = getattr(pkg, "SEM_" + sem)
    pred = getattr(pkg, "PRED_" + pred)
    origin, frames = synth.s_real(dims, vs, n_frames=n_frames)
    cfg = pkg.make_config(dims, vs, origin, semantics=sem, with_color=color, device=dev, z_range=zr)
    with pkg.Volume(cfg) as vol:
        for d, rgb, T in frames:
            vol.integrate(d, rgb if color else None, T)
        t, w, c = vol.download()
        n = t.size
        cap = max(vol.extract_count(pred, tt, wt), 1)
        # the caller's memory, all of it allocated BEFORE the calls under test
        d_t, d_w = DeviceBuf(lib, dev, 4 * n), DeviceBuf(lib, dev, 4 * n)
        d_c = DeviceBuf(lib, dev, 4 * n) if color else None
        bufs = {"xyz": DeviceBuf(lib, dev, 12 * cap), "rgb": DeviceBuf(lib, dev, 3 * cap), "idx": DeviceBuf(lib, dev, 8 * cap)}
        try:
            d_t.put(t); d_w.put(w)
            if color:
                d_c.put(c)
            del t, w, c
            want = _handle_path(pkg, vol, pred, tt, wt, bufs)
            # poison the outputs so that a call that writes nothing cannot pass
            for b in bufs.values():
                b.put(np.full(b.nbytes, 0xA5, np.uint8))
            for rep in range(2):   # two calls in a row on the same inputs, sharing the scratch
                got = _static_path(pkg, vol.cfg, d_t, d_w, d_c, pred, tt, wt, bufs)
                assert len(got[2]) == len(want[2]), (dims, rep, len(got[2]), len(want[2]))
                assert np.array_equal(got[2], want[2]), "indices differ"
                assert np.array_equal(got[0].view(np.uint32), want[0].view(np.uint32)), "xyz differs"
                assert np.array_equal(got[1], want[1]), "rgb differs"
            return len(want[2])
        finally:
            for b in [d_t, d_w, d_c, *bufs.values()]:
                if b is not None:
                    b.free()


def test_static_scratch_extraction_matches_handle_extraction(pkg, synth):
    results = [_run_case(pkg, synth, *case) for case in CASES]
    assert results[0] > 1000 and min(results) > 0, results   # every case extracted a real surface


def test_static_scratch_at_its_maximum_size(pkg, synth):
    """Exactly 2^27 voxels (512^3), the largest slab the entry point takes: every byte of the static
    array's ballot region is in use."""
    n = _run_case(pkg, synth, (512, 512, 512), "TSDFCU", False, "ABS_LT", 0.25, 0.0, None, vs=0.01, n_frames=1)
    assert n > 100000, n


def test_static_scratch_limit_is_enforced(pkg):
    """A slab beyond 2^27 voxels is refused before anything is launched (the pointers are never
    dereferenced), with a message that names the limit."""
    lib = pkg.load()
    cfg = pkg.make_config((1024, 512, 257), 0.02, (0, 0, 0), semantics=pkg.SEM_TSDFCU, with_color=False)
    assert cfg.dim_x * cfg.dim_y * cfg.dim_z > STATIC_MAX_VOXELS
    n = C.c_int64(0)
    dummy = DeviceBuf(lib, 0, 256)
    try:
        rc = lib.tsdf_extract_points_static(C.byref(cfg), dummy.p, dummy.p, None, pkg.PRED_NONZERO, np.float32(1.2),
                                            np.float32(0.9), None, None, None, 0, C.byref(n), None)
        assert rc == pkg.capi.E_INVALID
        msg = lib.tsdf_last_error().decode()
        assert "limit" in msg and str(STATIC_MAX_VOXELS) in msg, msg
    finally:
        dummy.free()
