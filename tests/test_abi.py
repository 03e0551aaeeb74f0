"""CPU tests of the drop-in boundary: the library loads, exports every symbol declared in
include/tsdf_b200.h, the host-only entry points behave, and compute entry points fail LOUDLY
(no CPU fallback) when there is no device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT, bits


def header_functions(dev):
    """Functions declared in the header; dev=True: only those inside #ifdef TSDF_B200_DEV, else the rest."""
    src = open(os.path.join(ROOT, "include", "tsdf_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    m = re.search(r"#ifdef TSDF_B200_DEV(.*?)#endif", src, flags=re.S)
    part = m.group(1) if dev else src.replace(m.group(0), "")
    return sorted(set(re.findall(r"\b(tsdf_[a-z0-9_]+)\s*\(", part)))


def test_library_exports_every_declared_symbol(pkg):
    lib = pkg.load(dev=False)
    names = header_functions(dev=False)
    assert len(names) >= 65
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/tsdf_b200.h but not exported"
    assert set(names) == set(pkg.capi.SIGNATURES), "ctypes stub and header disagree"
    assert lib.tsdf_version() == 100


def test_dev_hooks_are_not_in_the_product_library(pkg):
    """A/B variants, tsdf_bench_kernel and the exact-math self-test ship only in libtsdf_b200_dev.so."""
    dev_names = header_functions(dev=True)
    assert set(dev_names) == set(pkg.capi.DEV_SIGNATURES) and len(dev_names) == 4
    prod, dev = pkg.load(dev=False), pkg.load(dev=True)
    for n in dev_names:
        assert not hasattr(prod, n), f"{n} exported by the product library"
        assert hasattr(dev, n)
    for n in header_functions(dev=False):
        assert hasattr(dev, n)


def test_config_struct_layout(pkg):
    assert C.sizeof(pkg.capi.Config) == 4 * (3 + 2 + 3 + 16 + 2 + 9 + 1 + 3 + 2 + 1 + 7)
    cfg = pkg.default_config()
    assert (cfg.dim_x, cfg.dim_y, cfg.dim_z) == (200, 200, 200)          # include/tsdf.hpp:65-67
    assert cfg.voxel_size == np.float32(0.004)                            # include/tsdf.hpp:63
    assert cfg.trunc_margin == np.float32(0.004) * np.float32(5)          # include/tsdf.hpp:64
    assert list(cfg.K)[:6] == [np.float32(v) for v in (535.4, 0, 320.1, 0, 539.2, 247.6)]
    assert cfg.max_depth == 6.0 and (cfg.im_h, cfg.im_w) == (480, 640)


def test_config_from_bounds_follows_fusion_rule(pkg):
    cfg = pkg.config_from_bounds([[0, 10], [0, 10], [0, 10]], 0.02)       # src/TSDFfusion.py.in:23-29
    assert (cfg.dim_x, cfg.dim_y, cfg.dim_z) == (500, 500, 500)
    assert cfg.trunc_margin == np.float32(0.02 * 5) and cfg.semantics == pkg.SEM_FUSIONPY
    cfg = pkg.config_from_bounds([[-1.03, 2.0], [0.5, 0.77], [3, 3.001]], 0.05)
    ext = np.array([3.03, 0.27, 0.001])
    assert [cfg.dim_x, cfg.dim_y, cfg.dim_z] == list(np.ceil(ext / 0.05).astype(int))
    assert list(cfg.origin) == [np.float32(-1.03), np.float32(0.5), np.float32(3)]
    with pytest.raises(pkg.TsdfError):
        pkg.config_from_bounds([[0, 0], [0, 1], [0, 1]], 0.05)


def test_host_matrix_helpers_match_oracle(pkg, oracle, synth):
    lib = pkg.load()
    rng = np.random.default_rng(2)
    for i in range(500):
        A = synth.pose(synth.rot_xyz(*rng.uniform(-3, 3, 3)), rng.uniform(-5, 5, 3)).reshape(16)
        B = rng.normal(size=16).astype(np.float32)
        out = np.zeros(16, np.float32)
        lib.tsdf_multiply_matrix(A.ctypes.data, B.ctypes.data, out.ctypes.data)
        assert np.array_equal(bits(out), bits(oracle.multiply(A, B)))
        M = A if i % 2 else B
        rc = lib.tsdf_invert_matrix(M.ctypes.data, out.ctypes.data)
        ref = oracle.invert(M)
        assert (rc != 0) == (ref is None)
        if ref is not None:
            assert np.array_equal(bits(out), bits(ref))
    Z = np.zeros(16, np.float32)
    assert lib.tsdf_invert_matrix(Z.ctypes.data, out.ctypes.data) == pkg.capi.E_SINGULAR


def test_read_bundle_poses_matches_fixture(pkg, tmp_path):
    """bundle.txt ingest (reference examples/label_instance_rgbd.cpp:289-330): the first cameras of
    the reference's saved fr3_office run against the committed Twc fixture made from the same file."""
    import ctypes as C
    lib = pkg.load()
    path = os.path.join(ROOT, "tests", "golden", "bundle_head.txt").encode()
    n = C.c_int(0)
    assert lib.tsdf_read_bundle_poses(path, None, 0, C.byref(n)) == 0 and n.value == 4
    T = np.zeros((4, 16), np.float32)
    assert lib.tsdf_read_bundle_poses(path, T.ctypes.data_as(C.c_void_p), 2, C.byref(n)) == pkg.capi.E_CAPACITY
    assert lib.tsdf_read_bundle_poses(path, T.ctypes.data_as(C.c_void_p), 4, C.byref(n)) == 0
    ref = np.load(os.path.join(ROOT, "tests", "golden", "fr3_office_Twc.npy"))[:4]
    assert np.array_equal(T.reshape(4, 4, 4)[:, :3, :3], ref.reshape(4, 4, 4)[:, :3, :3])   # transpose is exact
    assert np.max(np.abs(T - ref)) < 1e-6 and (T[:, 12:] == [0, 0, 0, 1]).all()
    assert lib.tsdf_read_bundle_poses(str(tmp_path / "missing.txt").encode(), None, 0, C.byref(n)) == pkg.capi.E_IO
    (tmp_path / "bad.txt").write_text("2 0\n0 0 0\n1 0 0\n0 1\n")
    assert lib.tsdf_read_bundle_poses(str(tmp_path / "bad.txt").encode(), T.ctypes.data_as(C.c_void_p), 4, C.byref(n)) == pkg.capi.E_IO


def test_read_associations_matches_fixture(pkg, tmp_path):
    """associations.txt ingest (reference LoadImageNames, examples/label_instance_rgbd.cpp:239-268):
    the first lines of the reference's saved run; empty lines skipped, malformed ones rejected."""
    import ctypes as C
    lib = pkg.load()
    src = os.path.join(ROOT, "tests", "golden", "associations_head.txt")
    text = open(src).read()
    p2 = tmp_path / "assoc.txt"
    p2.write_text(text.replace("\n", "\n\n", 1) + "\n")          # an empty line inside and at the end
    n = C.c_int(0)
    assert lib.tsdf_read_associations(str(p2).encode(), None, None, None, None, 0, 0, C.byref(n)) == 0 and n.value == 4
    stride = 64
    tr, td = np.zeros(4), np.zeros(4)
    a, b = C.create_string_buffer(4 * stride), C.create_string_buffer(4 * stride)
    args = (tr.ctypes.data_as(C.c_void_p), td.ctypes.data_as(C.c_void_p), a, b)
    assert lib.tsdf_read_associations(str(p2).encode(), *args, stride, 2, C.byref(n)) == pkg.capi.E_CAPACITY
    assert lib.tsdf_read_associations(str(p2).encode(), *args, 8, 4, C.byref(n)) == pkg.capi.E_CAPACITY   # names do not fit
    assert lib.tsdf_read_associations(str(p2).encode(), *args, stride, 4, C.byref(n)) == 0 and n.value == 4
    want = [ln.split() for ln in text.splitlines()]
    for i, (t0, f0, t1, f1) in enumerate(want):
        assert a.raw[i * stride:].split(b"\0")[0].decode() == f0 and b.raw[i * stride:].split(b"\0")[0].decode() == f1
        assert tr[i] == float(t0) and td[i] == float(t1)
    assert lib.tsdf_read_associations(str(tmp_path / "none.txt").encode(), None, None, None, None, 0, 0, C.byref(n)) == pkg.capi.E_IO
    (tmp_path / "bad.txt").write_text("1.0 rgb/a.png 1.0\n")
    assert lib.tsdf_read_associations(str(tmp_path / "bad.txt").encode(), None, None, None, None, 0, 0, C.byref(n)) == pkg.capi.E_IO


def test_slab_ranges_partition_the_grid(pkg):
    for dz in (1024, 256, 200, 130, 7):
        for n in (1, 2, 3, 4, 8):
            r = [pkg.slab_range(dz, n, k) for k in range(n)]
            assert r[0][0] == 0 and r[-1][1] == dz
            for a, b in zip(r, r[1:]):
                assert a[1] == b[0]
            assert all(z0 % 4 == 0 for z0, z1 in r if z1 > z0)
    assert [pkg.slab_range(1024, 8, k) for k in range(8)] == [(128 * k, 128 * k + 128) for k in range(8)]


def test_invalid_configs_are_rejected(pkg):
    cfg = pkg.default_config()
    cfg.dim_x = 0
    with pytest.raises(pkg.TsdfError) as e:
        pkg.Volume(cfg)
    assert e.value.code in (pkg.capi.E_INVALID, pkg.capi.E_NODEVICE)


def test_no_cpu_fallback(pkg):
    """Without a device every compute entry point must fail loudly."""
    if pkg.device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(pkg.TsdfError) as e:
        pkg.Volume(pkg.default_config())
    assert e.value.code == pkg.capi.E_NODEVICE
    out = np.zeros(3, np.float32)
    d = np.ones((4, 4), np.float32)
    K = pkg.synth.TUM3_K
    rc = pkg.load().tsdf_instance_origin(0, d.ctypes.data, None, 4, 4, K.ctypes.data, out.ctypes.data)
    assert rc == pkg.capi.E_NODEVICE


def test_product_does_not_import_oracle():
    """The product path must never route through oracle/ (test infrastructure)."""
    pk = os.path.join(ROOT, "semantic-slam_b200")
    for dirpath, _, files in os.walk(pk):
        for f in files:
            path = os.path.join(dirpath, f)
            if f.endswith(".py"):
                txt = open(path).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", txt, flags=re.M), f"{f} imports oracle"
                assert "liboracle" not in txt and "_ref" not in txt, f"{f} loads oracle artefacts"
            elif f.endswith((".cu", ".cuh", ".cpp", ".h", ".hpp")) or f == "Makefile":
                txt = open(path, errors="replace").read()
                assert not re.search(r"#include\s+[\"<][^\">]*oracle", txt), f"{f} includes oracle code"
                assert "liboracle" not in txt and "libtsdf_ref" not in txt, f"{f} links oracle artefacts"


def test_read_trajectory_kitti_and_frame_ids(pkg, synth, tmp_path):
    """CameraTrajectory.txt as SaveTrajectoryKITTI writes it (src/ORB_SLAM2/System.cc:502-505: `fixed`,
    setprecision(9), the 3x4 [Rwc | twc] row by row) and frame_id.txt (:1104-1121, one integer per line)."""
    lib = pkg.load()
    rng = np.random.default_rng(8)
    poses = [synth.pose(synth.rot_xyz(*rng.uniform(-1, 1, 3)), rng.uniform(-50, 50, 3)).reshape(4, 4) for _ in range(7)]
    path = tmp_path / "CameraTrajectory.txt"
    with open(path, "w") as f:
        for T in poses:
            f.write(" ".join("%.9f" % float(v) for v in T[:3].reshape(-1)) + "\n")
        f.write("\n")
    n = C.c_int(0)
    p = str(path).encode()
    assert lib.tsdf_read_trajectory_kitti(p, None, 0, C.byref(n)) == 0 and n.value == 7
    T = np.zeros((7, 16), np.float32)
    assert lib.tsdf_read_trajectory_kitti(p, T.ctypes.data_as(C.c_void_p), 3, C.byref(n)) == pkg.capi.E_CAPACITY
    assert lib.tsdf_read_trajectory_kitti(p, T.ctypes.data_as(C.c_void_p), 7, C.byref(n)) == 0 and n.value == 7
    for got, want in zip(T.reshape(7, 4, 4), poses):
        # "%.9f" of a float32 parses back to the same float32 for |x| < 2^6 or so; translations up to 50 m
        # keep 9 decimals >> the 2^-18 spacing of float32 there
        assert np.array_equal(got[:3], np.float32([[float("%.9f" % float(v)) for v in r] for r in want[:3]]))
        assert list(got[3]) == [0, 0, 0, 1]
    (tmp_path / "bad.txt").write_text("1 0 0 0 0 1 0 0 0 0 1\n")
    assert lib.tsdf_read_trajectory_kitti(str(tmp_path / "bad.txt").encode(), None, 0, C.byref(n)) == pkg.capi.E_IO
    assert "11 values" in lib.tsdf_last_error().decode()
    assert lib.tsdf_read_trajectory_kitti(str(tmp_path / "missing.txt").encode(), None, 0, C.byref(n)) == pkg.capi.E_IO

    ids = [0, 3, 17, 18, 4540]
    (tmp_path / "frame_id.txt").write_text("".join("%d\n" % i for i in ids))
    q = str(tmp_path / "frame_id.txt").encode()
    assert lib.tsdf_read_frame_ids(q, None, 0, C.byref(n)) == 0 and n.value == 5
    out = np.zeros(5, np.int64)
    assert lib.tsdf_read_frame_ids(q, out.ctypes.data_as(C.c_void_p), 2, C.byref(n)) == pkg.capi.E_CAPACITY
    assert lib.tsdf_read_frame_ids(q, out.ctypes.data_as(C.c_void_p), 5, C.byref(n)) == 0 and list(out) == ids
    (tmp_path / "bad_ids.txt").write_text("12\nx\n")
    assert lib.tsdf_read_frame_ids(str(tmp_path / "bad_ids.txt").encode(), None, 0, C.byref(n)) == pkg.capi.E_IO
