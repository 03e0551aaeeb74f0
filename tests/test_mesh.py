"""Mesh readback (plugin SaveMesh -> get_mesh() + meshwrite, reference src/TSDFfusion.py.in:48-53).

CPU part: the oracle's marching cubes (oracle_mesh) and the generated case table are checked through
properties that do not depend on any implementation — every mesh edge is shared by exactly two
triangles with opposite directions (watertight, consistently oriented, 2-manifold), Euler
characteristic 2 for a sphere, triangle normals and vertex normals point to free space, vertices
sit on the zero crossing — and against the committed fixture tests/golden/mesh_small.npz.
GPU part: tsdf_extract_mesh against the oracle, bit for bit (positions, normals, colours, face
indices), on random volumes in both storage orders, ragged dims, slabs, a fused scene, and the PLY
bytes.  skimage (upstream's marching cubes) is absent: vertex/face ORDER parity is unpinned."""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import GOLDEN, ROOT, bits
from oracle.oracle import SEM_FUSIONPY, SEM_TSDFCU

VS = 0.02
K_ID = np.eye(3, dtype=np.float32)


def to_storage(t_xyz, sem):
    """[x][y][z] numpy array -> the semantics' storage order, flat."""
    if sem == SEM_FUSIONPY:
        return np.ascontiguousarray(t_xyz, np.float32).reshape(-1)
    return np.ascontiguousarray(t_xyz.transpose(2, 1, 0), np.float32).reshape(-1)


def oracle_mesh(oracle, t_xyz, sem, color=None, origin=(0, 0, 0), z_range=None):
    dims = t_xyz.shape if z_range is None else (t_xyz.shape[0], t_xyz.shape[1], z_range[2])
    p = oracle.params(dims, VS, 5 * VS, origin, (480, 640), K_ID, 6.0, sem,
                      z_range=None if z_range is None else z_range[:2])
    return oracle.mesh(p, to_storage(t_xyz, sem), None if color is None else to_storage(color, sem))


def edge_audit(v, f):
    """-> (#directed edges seen twice, boolean mask of directed edges without their reverse, edges)"""
    e = np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]]).astype(np.int64)
    key = e[:, 0] * len(v) + e[:, 1]
    rkey = e[:, 1] * len(v) + e[:, 0]
    _, cnt = np.unique(key, return_counts=True)
    return int((cnt > 1).sum()), ~np.isin(rkey, key), e


def sphere(dims, c, r):
    X, Y, Z = np.meshgrid(*[np.arange(d, dtype=np.float64) for d in dims], indexing="ij")
    return ((np.sqrt((X - c[0]) ** 2 + (Y - c[1]) ** 2 + (Z - c[2]) ** 2) - r) / 5.0).astype(np.float32)


def test_case_table_is_the_generators_output():
    out = subprocess.check_output([sys.executable, os.path.join(ROOT, "tools", "gen_mc_table.py"), "--print"]).decode()
    for rel in ("oracle/mc_table.inc", "semantic-slam_b200/csrc/mc_table.inc"):
        assert open(os.path.join(ROOT, rel)).read() == out, rel + " is stale: run tools/gen_mc_table.py"


@pytest.mark.parametrize("sem", [SEM_FUSIONPY, SEM_TSDFCU])
def test_oracle_sphere_is_closed_oriented_genus0(oracle, sem):
    dims, c, r = (24, 20, 28), (11.3, 9.7, 13.1), 7.4
    t = sphere(dims, c, r)
    v, f, n, _ = oracle_mesh(oracle, t, sem)
    dup, missing, e = edge_audit(v, f)
    assert dup == 0 and missing.sum() == 0                 # every edge: two triangles, opposite directions
    assert np.isin(np.arange(len(v)), f).all()             # no unused vertex
    n_edges = len(np.unique(np.minimum(e[:, 0], e[:, 1]) * len(v) + np.maximum(e[:, 0], e[:, 1])))
    assert len(v) - n_edges + len(f) == 2                  # Euler characteristic of a sphere
    cw = np.array(c) * VS
    tri = v[f]
    fn = np.cross(tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0])
    assert ((fn * (tri.mean(1) - cw)).sum(1) > 0).all()    # wound towards free space (tsdf > 0)
    assert ((n * (v - cw)).sum(1) > 0).all() and np.abs(np.linalg.norm(n, axis=1) - 1).max() < 1e-6
    assert np.abs(np.linalg.norm(v - cw, axis=1) - r * VS).max() < 0.05 * VS   # on the zero crossing


@pytest.mark.parametrize("sem", [SEM_FUSIONPY, SEM_TSDFCU])
def test_oracle_noise_volume_is_manifold_away_from_the_boundary(oracle, sem):
    """Worst case for the case table: every one of the 256 configurations, ambiguous faces included."""
    dims = (17, 13, 19)
    t = np.random.default_rng(0xB200).standard_normal(dims).astype(np.float32)
    v, f, _, _ = oracle_mesh(oracle, t, sem)
    dup, missing, e = edge_audit(v, f)
    assert dup == 0
    vv = v / VS
    for a, b in e[missing]:                                # open edges lie in a face of the grid
        on = [abs(vv[a][k] - lim) < 1e-4 and abs(vv[b][k] - lim) < 1e-4 for k in range(3) for lim in (0, dims[k] - 1)]
        assert any(on)
    # one vertex per sign-crossing grid edge
    ins = t < 0
    want = sum(int((np.take(ins, range(0, dims[k] - 1), k) != np.take(ins, range(1, dims[k]), k)).sum()) for k in range(3))
    assert len(v) == want


def test_oracle_layouts_give_the_same_mesh(oracle):
    t = np.random.default_rng(5).standard_normal((9, 11, 7)).astype(np.float32)
    col = np.random.default_rng(6).integers(0, 1 << 24, t.shape).astype(np.float32)
    va, fa, na, ca = oracle_mesh(oracle, t, SEM_FUSIONPY, col)
    vb, fb, nb, cb = oracle_mesh(oracle, t, SEM_TSDFCU, col)
    ka = np.lexsort(va.T[::-1])
    kb = np.lexsort(vb.T[::-1])
    assert np.array_equal(bits(va[ka]), bits(vb[kb])) and np.array_equal(bits(na[ka]), bits(nb[kb]))
    assert np.array_equal(ca[ka], cb[kb])
    ra, rb = np.empty(len(ka), np.int64), np.empty(len(kb), np.int64)
    ra[ka] = np.arange(len(ka))
    rb[kb] = np.arange(len(kb))
    canon = lambda f, r: sorted(tuple(np.roll(t3, -int(np.argmin(t3)))) for t3 in r[f])
    assert canon(fa, ra) == canon(fb, rb)


def test_oracle_colour_and_golden(oracle):
    g = np.load(os.path.join(GOLDEN, "mesh_small.npz"))
    v, f, n, c = oracle_mesh(oracle, g["tsdf"], SEM_FUSIONPY, g["color"], origin=tuple(g["origin"]))
    assert np.array_equal(bits(v), bits(g["verts"])) and np.array_equal(f, g["faces"])
    assert np.array_equal(bits(n), bits(g["norms"])) and np.array_equal(c, g["colors"])
    # colour = voxel at round-half-even of the vertex position (upstream: color_vol[np.round(verts)])
    ijk = np.round((v.astype(np.float64) - g["origin"]) / VS).astype(int)
    cc = g["color"][ijk[:, 0], ijk[:, 1], ijk[:, 2]].astype(np.uint32)
    # positions that are within rounding of x.5 may round either way in float64; compare the rest
    frac = np.abs(((v.astype(np.float64) - g["origin"]) / VS) % 1 - 0.5).min(1)
    ok = frac > 1e-3
    unpack = lambda x: np.stack([x & 255, (x >> 8) & 255, (x >> 16) & 255], 1).astype(np.uint8)
    assert ok.mean() > 0.5
    assert np.array_equal(c[ok], unpack(cc)[ok])
    # the fixture's plane z = 15.5 gives exact ties: half to even -> voxel 16, never 15
    q = (v.astype(np.float64) - g["origin"]) / VS
    tie = (np.abs(q[:, 2] - 15.5) < 1e-5) & (np.abs(q[:, 0] - np.round(q[:, 0])) < 1e-5) & (np.abs(q[:, 1] - np.round(q[:, 1])) < 1e-5)
    assert tie.sum() > 100
    ct = g["color"][np.round(q[tie, 0]).astype(int), np.round(q[tie, 1]).astype(int), 16].astype(np.uint32)
    assert np.array_equal(c[tie], unpack(ct))


def test_meshwrite_matches_oracle_writer(pkg, oracle, tmp_path):
    g = np.load(os.path.join(GOLDEN, "mesh_small.npz"))
    a, b = str(tmp_path / "a.ply"), str(tmp_path / "b.ply")
    pkg.meshwrite(a, g["verts"], g["faces"], g["norms"], g["colors"])
    assert oracle.save_mesh_ply(b, g["verts"], g["faces"], g["norms"], g["colors"])
    assert open(a, "rb").read() == open(b, "rb").read()
    head = open(a).read().split("end_header")[0].split("\n")
    assert head[:3] == ["ply", "format ascii 1.0", "element vertex %d" % len(g["verts"])]


# ---- GPU parity ------------------------------------------------------------------------------------
def gpu_mesh(pkg, t_xyz, sem, color=None, origin=(0, 0, 0), z_range=None):
    dims = t_xyz.shape
    cfg = pkg.make_config(dims if z_range is None else (dims[0], dims[1], z_range[2]), VS, origin, semantics=sem,
                          with_color=color is not None, z_range=None if z_range is None else z_range[:2])
    with pkg.Volume(cfg) as vol:
        vol.upload(tsdf=to_storage(t_xyz, sem), color=None if color is None else to_storage(color, sem))
        nv, nf = vol.mesh_count()
        out = vol.extract_mesh()
        assert (len(out[0]), len(out[1])) == (nv, nf)
        return out


def assert_same_mesh(a, b):
    (va, fa, na, ca), (vb, fb, nb, cb) = a, b
    assert (len(va), len(fa)) == (len(vb), len(fb)), ((len(va), len(fa)), (len(vb), len(fb)))
    assert np.array_equal(fa, fb), "face indices differ"
    assert np.array_equal(bits(va), bits(vb)), "vertex positions differ"
    assert np.array_equal(bits(na), bits(nb)), "normals differ"
    assert np.array_equal(ca, cb), "colours differ"


@pytest.mark.gpu
@pytest.mark.parametrize("sem", [SEM_FUSIONPY, SEM_TSDFCU])
@pytest.mark.parametrize("dims", [(17, 13, 19), (64, 48, 40), (5, 70, 33), (1, 9, 9), (3, 2, 2), (40, 36, 132)])
def test_gpu_mesh_equals_oracle_on_noise(pkg, oracle, sem, dims):
    rng = np.random.default_rng(sum(dims) + sem)
    t = rng.standard_normal(dims).astype(np.float32)
    t[rng.random(dims) < 0.05] = 0.0                       # exact zeros and -0.0 count as outside
    t[rng.random(dims) < 0.02] = -0.0
    col = rng.integers(0, 1 << 24, dims).astype(np.float32)
    assert_same_mesh(gpu_mesh(pkg, t, sem, col, origin=(-0.31, 0.2, 1.7)),
                     oracle_mesh(oracle, t, sem, col, origin=(-0.31, 0.2, 1.7)))


@pytest.mark.gpu
def test_gpu_mesh_sphere_and_empty(pkg, oracle):
    t = sphere((96, 80, 72), (47.2, 40.1, 33.3), 25.6)
    got = gpu_mesh(pkg, t, SEM_FUSIONPY)
    assert_same_mesh(got, oracle_mesh(oracle, t, SEM_FUSIONPY))
    dup, missing, _ = edge_audit(got[0], got[1])
    assert dup == 0 and missing.sum() == 0
    v, f, n, c = gpu_mesh(pkg, np.ones((32, 32, 32), np.float32), SEM_FUSIONPY)   # a fresh volume: no surface
    assert len(v) == 0 and len(f) == 0


@pytest.mark.gpu
def test_gpu_mesh_of_a_z_slab(pkg, oracle):
    dims = (20, 16, 48)
    t = np.random.default_rng(11).standard_normal(dims).astype(np.float32)
    for sem in (SEM_FUSIONPY, SEM_TSDFCU):
        for z0, z1 in ((0, 16), (16, 48)):
            sub = np.ascontiguousarray(t[:, :, z0:z1])
            assert_same_mesh(gpu_mesh(pkg, sub, sem, z_range=(z0, z1, dims[2])),
                             oracle_mesh(oracle, sub, sem, z_range=(z0, z1, dims[2])))


@pytest.mark.gpu
def test_gpu_mesh_of_a_fused_scene_and_ply(pkg, oracle, synth, tmp_path):
    dims = (96, 80, 96)
    origin, frames = synth.s_real(dims, VS, n_frames=4)
    cfg = pkg.make_config(dims, VS, origin, semantics=pkg.SEM_FUSIONPY, with_color=True)
    with pkg.Volume(cfg) as vol:
        for d, rgb, T in frames:
            vol.integrate(d, rgb, T)
        t, w, c = vol.download()
        got = vol.extract_mesh()
        assert vol.mesh_count() == (len(got[0]), len(got[1]))          # cached sweep
        vol.integrate(*frames[0][:2], frames[0][2])                    # invalidates it
        t2, _, c2 = vol.download()
        got2 = vol.extract_mesh()
        vol.save_mesh_ply(str(tmp_path / "gpu.ply"))
    p = oracle.params(dims, VS, cfg.trunc_margin, origin, synth.TUM3_HW, synth.TUM3_K, cfg.max_depth, SEM_FUSIONPY)
    want = oracle.mesh(p, t, c)
    assert len(want[0]) > 1000 and len(want[1]) > 1000
    assert_same_mesh(got, want)
    assert not np.array_equal(t, t2)
    assert_same_mesh(got2, oracle.mesh(p, t2, c2))
    # the PLY written by the library is the oracle writer's bytes for the same mesh
    assert oracle.save_mesh_ply(str(tmp_path / "cpu.ply"), got2[0], got2[1], got2[2], got2[3])
    assert open(tmp_path / "gpu.ply", "rb").read() == open(tmp_path / "cpu.ply", "rb").read()


@pytest.mark.gpu
def test_gpu_mesh_capacity_and_device_outputs(pkg):
    import ctypes as C
    t = sphere((40, 40, 40), (20, 20, 20), 12.0)
    cfg = pkg.make_config(t.shape, VS, (0, 0, 0), semantics=pkg.SEM_FUSIONPY, with_color=False)
    lib = pkg.load()
    with pkg.Volume(cfg) as vol:
        vol.upload(tsdf=to_storage(t, SEM_FUSIONPY))
        nv, nf = vol.mesh_count()
        verts = np.zeros((nv, 3), np.float32)
        faces = np.zeros((nf, 3), np.int32)
        a, b = C.c_int64(0), C.c_int64(0)
        rc = lib.tsdf_extract_mesh(vol.h, verts.ctypes.data, None, None, nv - 1, None, 0, C.byref(a), C.byref(b))
        assert rc == pkg.capi.E_CAPACITY and (a.value, b.value) == (nv, nf)
        rc = lib.tsdf_extract_mesh(vol.h, None, None, None, 0, faces.ctypes.data, nf - 1, C.byref(a), C.byref(b))
        assert rc == pkg.capi.E_CAPACITY
        # faces only, then vertices only: each array is optional
        assert lib.tsdf_extract_mesh(vol.h, None, None, None, 0, faces.ctypes.data, nf, C.byref(a), C.byref(b)) == 0
        assert lib.tsdf_extract_mesh(vol.h, verts.ctypes.data, None, None, nv, None, 0, C.byref(a), C.byref(b)) == 0
        full = vol.extract_mesh()
        assert np.array_equal(full[1], faces) and np.array_equal(bits(full[0]), bits(verts))
        assert full[3].max() == 0                                      # no colour array: zeros


@pytest.mark.gpu
@pytest.mark.parametrize("sem", [SEM_FUSIONPY, SEM_TSDFCU])
def test_gpu_mesh_scalar_and_vector_sweeps_agree(pkg, oracle, sem, dev_lib):
    """Fast axes that are a multiple of 8 / of 4 take k_mc_mark8 / k_mc_mark4 (a thread owns 8 / 4 voxels);
    variant bit 16 forces the scalar sweep, bit 18 the 4-wide one.  Both must give the oracle's mesh — also where a 128-voxel
    chunk straddles rows (fast axis 36) and where the volume ends inside a tile."""
    for dims in ((36, 20, 36), (128, 12, 256), (8, 8, 8), (4, 4, 4), (132, 9, 68)):
        rng = np.random.default_rng(sum(dims))
        t = rng.standard_normal(dims).astype(np.float32)
        col = rng.integers(0, 1 << 24, dims).astype(np.float32)
        want = oracle_mesh(oracle, t, sem, col)
        cfg = pkg.make_config(dims, VS, (0, 0, 0), semantics=sem, with_color=True)
        with pkg.Volume(cfg) as vol:
            vol.upload(tsdf=to_storage(t, sem), color=to_storage(col, sem))
            assert_same_mesh(vol.extract_mesh(), want)
            for variant in (65536, 262144):                # scalar sweep; 4-wide sweep
                vol.set_variant(variant)
                vol.upload()                               # invalidate the cached sweep
                assert_same_mesh(vol.extract_mesh(), want)


@pytest.mark.gpu
def test_gpu_scans_over_many_tiles(pkg, oracle):
    """42 M voxels: 20480 mesh tiles and 5120 extraction tiles, i.e. several 4096-entry iterations of the
    one-block scan (register carry, prefetched counts, double-buffered warp totals) behind both the
    point extraction and the mesh; sparse content so that the oracle's outputs stay small."""
    dims = (512, 512, 160)
    rng = np.random.default_rng(0xB200 + 9)
    t = np.ones(dims, np.float32)
    hit = rng.random(dims) < 0.01
    t[hit] = rng.standard_normal(int(hit.sum())).astype(np.float32)
    w = hit.astype(np.float32) * 2
    cfg = pkg.make_config(dims, VS, (0, 0, 0), semantics=pkg.SEM_FUSIONPY, with_color=False)
    p = oracle.params(dims, VS, 5 * VS, (0, 0, 0), (480, 640), K_ID, np.inf, SEM_FUSIONPY)
    with pkg.Volume(cfg) as vol:
        vol.upload(tsdf=t.reshape(-1), weight=w.reshape(-1))
        gxyz, _, gidx = vol.extract_points()
        got = vol.extract_mesh()
    oxyz, _, oidx = oracle.extract(p, t.reshape(-1), w.reshape(-1), None)
    assert len(oidx) > 300000 and np.array_equal(gidx, oidx) and np.array_equal(bits(gxyz), bits(oxyz))
    want = oracle.mesh(p, t.reshape(-1), None)
    assert len(want[0]) > 1000000
    assert_same_mesh(got, want)
