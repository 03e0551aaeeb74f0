"""CPU tests of the oracle: pinned against the golden vectors made from the reference's own
source lines (tests/golden/make_golden.py), and against oracle/_ref itself when present."""
import os

import numpy as np
import pytest

from conftest import bits
from oracle.oracle import Oracle, RefLift, ref_available, SEM_TSDFCU, SEM_FUSIONPY, PRED_ABS_LT


def run_cu(o, g):
    dims, vs, hw, K = tuple(g["dims"]), float(g["vs"]), tuple(g["hw"]), g["K"]
    trunc = np.float32(vs) * np.float32(5)
    p = o.params(dims, vs, trunc, g["origin"], hw, K, 6.0, SEM_TSDFCU)
    t, w, _ = o.new_volume(p)
    for d, T, c2b in zip(g["depth"], g["T_wc"], g["cam2base"]):
        mine = o.compose(g["base2world"], T)
        assert np.array_equal(bits(mine), bits(c2b)), "pose composition differs from the reference"
        o.integrate(p, mine, d, None, 1.0, t, w, None)
    return p, t, w


def test_oracle_matches_reference_golden(oracle, golden_cu, tmp_path):
    p, t, w = run_cu(oracle, golden_cu)
    assert np.array_equal(w, golden_cu["weight"])
    assert np.array_equal(bits(t), bits(golden_cu["tsdf"]))
    path = str(tmp_path / "o.ply")
    assert oracle.save_ply(path, p, t, w)
    assert np.array_equal(np.frombuffer(open(path, "rb").read(), np.uint8), golden_cu["ply"])


def test_oracle_fusionpy_regression(oracle, golden_cu, golden_py):
    g = golden_cu
    dims, vs, hw, K = tuple(g["dims"]), float(g["vs"]), tuple(g["hw"]), g["K"]
    p = oracle.params(dims, vs, np.float32(vs) * np.float32(5), g["origin"], hw, K, np.inf, SEM_FUSIONPY)
    t, w, c = oracle.new_volume(p, True)
    for d, rgb, T in zip(g["depth"], g["rgb"], g["T_wc"]):
        oracle.integrate(p, T.reshape(16), d, oracle.pack_color(rgb), float(golden_py["obs_weight"]), t, w, c)
    assert np.array_equal(w, golden_py["weight"]) and np.array_equal(bits(t), bits(golden_py["tsdf"]))
    assert np.array_equal(c, golden_py["color"])
    xyz, rgb, idx = oracle.extract(p, t, w, c)
    assert np.array_equal(idx, golden_py["idx"]) and np.array_equal(rgb, golden_py["rgb"])


@pytest.mark.skipif(not ref_available(), reason="oracle/_ref not built (needs /root/reference at build time)")
def test_oracle_vs_reference_lift_random(oracle, synth):
    r = RefLift()
    rng = np.random.default_rng(7)
    for i in range(300):
        T = synth.pose(synth.rot_xyz(*rng.uniform(-3, 3, 3)), rng.uniform(-5, 5, 3))
        if i % 3 == 0:
            T = rng.normal(size=(4, 4)).astype(np.float32)
        a, b = oracle.invert(T), r.invert(T)
        assert (a is None and b is None) or np.array_equal(bits(a), bits(b))
        T2 = rng.normal(size=(4, 4)).astype(np.float32)
        assert np.array_equal(bits(oracle.multiply(T, T2)), bits(r.multiply(T, T2)))
    dims, vs = (50, 44, 38), 0.03          # odd sizes on purpose
    origin, frames = synth.s_real(dims, vs, n_frames=4, seed=11)
    p = oracle.params(dims, vs, 5 * vs, origin, synth.TUM3_HW, synth.TUM3_K, 6.0, SEM_TSDFCU)
    t, w, _ = oracle.new_volume(p)
    t2, w2 = np.ones_like(t), np.zeros_like(w)
    base = frames[1][2]
    for d, _, T in frames:
        c2b = oracle.compose(base, T)
        oracle.integrate(p, c2b, d, None, 1.0, t, w, None)
        r.integrate(synth.TUM3_K, c2b, d, dims, origin, vs, 5 * vs, t2, w2)
    assert w.max() >= 2 and np.array_equal(w, w2) and np.array_equal(bits(t), bits(t2))


def test_oracle_slabs_equal_whole(oracle, synth):
    dims, vs = (32, 24, 40), 0.04
    origin, frames = synth.s_real(dims, vs, n_frames=2, seed=3)
    for sem in (SEM_TSDFCU, SEM_FUSIONPY):
        p = oracle.params(dims, vs, 5 * vs, origin, synth.TUM3_HW, synth.TUM3_K, 6.0, sem)
        t, w, c = oracle.new_volume(p, True)
        for d, rgb, T in frames:
            oracle.integrate(p, T.reshape(16), d, oracle.pack_color(rgb), 1.0, t, w, c)
        shape = (dims[2], dims[1], dims[0]) if sem == SEM_TSDFCU else dims
        zax = 0 if sem == SEM_TSDFCU else 2
        parts = []
        for z0, z1 in ((0, 12), (12, 28), (28, 40)):
            ps = oracle.params(dims, vs, 5 * vs, origin, synth.TUM3_HW, synth.TUM3_K, 6.0, sem, (z0, z1))
            ts, ws, cs = oracle.new_volume(ps, True)
            for d, rgb, T in frames:
                oracle.integrate(ps, T.reshape(16), d, oracle.pack_color(rgb), 1.0, ts, ws, cs)
            lshape = list(shape)
            lshape[zax] = z1 - z0
            parts.append(ws.reshape(lshape))
        assert np.array_equal(np.concatenate(parts, axis=zax), w.reshape(shape))


def test_oracle_extract_predicates_and_origin(oracle, synth):
    dims, vs = (24, 24, 24), 0.05
    origin, frames = synth.s_real(dims, vs, n_frames=2, seed=5)
    p = oracle.params(dims, vs, 5 * vs, origin, synth.TUM3_HW, synth.TUM3_K, 6.0, SEM_TSDFCU)
    t, w, _ = oracle.new_volume(p)
    for d, _, T in frames:
        oracle.integrate(p, T.reshape(16), d, None, 1.0, t, w, None)
    _, _, idx = oracle.extract(p, t, w, None)
    assert np.array_equal(idx, np.flatnonzero((np.abs(t) != 0) & (w > np.float32(0.9))))
    _, _, idx2 = oracle.extract(p, t, w, None, PRED_ABS_LT, 0.5, 0.9)
    assert np.array_equal(idx2, np.flatnonzero((np.abs(t) < np.float32(0.5)) & (w > np.float32(0.9))))
    d = frames[0][0]
    mask = np.zeros(d.shape, np.uint8)
    mask[100:200, 300:400] = 255
    og = oracle.instance_origin(d, mask, synth.TUM3_K)
    v, u = np.mgrid[0:d.shape[0], 0:d.shape[1]]
    z = np.where(mask >= 128, d, 0).astype(np.float32)
    ok = z > 0
    K = synth.TUM3_K
    x = ((u.astype(np.float32) - K[2]) * z * (np.float32(1) / K[0]))[ok]
    y = ((v.astype(np.float32) - K[5]) * z * (np.float32(1) / K[4]))[ok]
    assert np.allclose(og, [x.min(), y.min(), z[ok].min()], rtol=0, atol=0)


def test_oracle_range_to_depth_known_answers(oracle, synth):
    """Lidar range -> depth (reference src/DoN.cpp:86-100): on the principal ray d == range; a
    pixel whose ray has x = 3/4, y = 1 has rim = sqrt(9/16 + 1 + 1) = 1.6007810593582121."""
    K = np.array([400, 0, 100, 0, 300, 50, 0, 0, 1], np.float32)
    rng = np.zeros((120, 500), np.float32)
    rng[50, 100] = 7.0           # principal point: x = y = 0
    rng[350 % 120, 0] = 0.0
    rng[110, 400] = 4.0          # x = 300/400, y = 60/300 -> rim = sqrt(0.5625 + 0.04 + 1)
    d = oracle.range_to_depth(rng, K, scale=0.5)
    assert d[50, 100] == np.float32(3.5)
    x, y = np.float32(300) / np.float32(400), np.float32(60) / np.float32(300)
    rim = np.sqrt(np.float32(x * x + y * y + np.float32(1)), dtype=np.float32)
    assert d[110, 400] == np.float32(np.float32(4.0 * 0.5) / rim)
    assert d[0, 0] == 0 and np.count_nonzero(d) == 2
    # against float64 geometry: depth = range * cos(angle to the optical axis)
    rr = np.full((120, 500), 10.0, np.float32)
    dd = oracle.range_to_depth(rr, K, 1.0)
    cc, rw = np.meshgrid(np.arange(500), np.arange(120))
    ref = 10.0 / np.sqrt(((cc - 100.0) / 400.0) ** 2 + ((rw - 50.0) / 300.0) ** 2 + 1.0)
    assert np.max(np.abs(dd - ref)) < 5e-6


def test_numpy_float64_restatement_agrees_with_oracle(oracle, synth):
    """oracle/fusion_numpy.py (the float64 'reference CPU TSDFfusion path' of SURVEY §8d, parity
    unpinned) against the fp32 C oracle in SEM_FUSIONPY: same update counts up to pixel/truncation
    flips from float64 vs float32 rounding, tsdf within 1e-5 and colour within one
    step per channel (tie rounding) wherever the weights agree and no flip touched the voxel."""
    from oracle.fusion_numpy import TSDFVolumeF64
    dims, vs = (48, 40, 44), 0.03
    origin, frames = synth.s_real(dims, vs, n_frames=3, seed=5)
    o64 = np.asarray(origin, np.float32).astype(np.float64)
    bnds = np.stack([o64, o64 + np.array(dims) * float(np.float32(vs))], axis=1)
    v = TSDFVolumeF64(bnds, float(np.float32(vs)))
    assert tuple(v.dims) == dims
    p = oracle.params(dims, vs, np.float32(vs) * np.float32(5), origin, synth.TUM3_HW, synth.TUM3_K, np.inf, SEM_FUSIONPY)
    t, w, c = oracle.new_volume(p, True)
    n64 = n32 = 0
    for d, rgb, T in frames:
        n32 += oracle.integrate(p, T.reshape(16), d, oracle.pack_color(rgb), 1.0, t, w, c)
        n64 += v.integrate(rgb, d, synth.TUM3_K, T.reshape(4, 4), 1.0)
    assert n32 > 10000 and abs(n64 - n32) <= 2e-3 * n32
    same = v.weight == w
    assert same.mean() > 0.995
    # a voxel whose weights agree can still have sampled a neighbouring pixel in one frame: bounded fraction
    dt = np.abs(v.tsdf[same] - t[same])
    assert (dt > 1e-5).mean() < 5e-3
    # colour: the float64 CPU mode rounds ties to even (np.round), the kernel arithmetic away from zero
    # (roundf), and (o*w0 + n)/2 ties often: per channel the two may differ by one step, not more
    clean = same & (np.abs(v.tsdf - t) <= 1e-5)
    a, b = v.color[clean].astype(np.int64), c[clean].astype(np.int64)
    step = np.maximum.reduce([np.abs((a >> s & 255) - (b >> s & 255)) for s in (0, 8, 16)])
    assert (step > 1).mean() < 5e-3 and (step == 0).mean() > 0.9


def test_oracle_range_equals_whole(oracle, synth):
    """oracle_integrate_range (a window of the linear storage order) against oracle_integrate on
    the same slab, both semantics, whole grid and a z-slab."""
    dims, vs = (40, 36, 44), 0.04
    origin, frames = synth.s_real(dims, vs, n_frames=2, seed=3)
    for sem in (SEM_TSDFCU, SEM_FUSIONPY):
        for zr in (None, (8, 30)):
            p = oracle.params(dims, vs, 5 * vs, origin, synth.TUM3_HW, synth.TUM3_K, 6.0, sem, zr)
            t, w, c = oracle.new_volume(p, True)
            n = oracle.n_local(p)
            first, cnt = n // 3, n // 2
            t2, w2, c2 = np.ones(cnt, np.float32), np.zeros(cnt, np.float32), np.zeros(cnt, np.float32)
            total = part = 0
            for d, rgb, T in frames:
                total += oracle.integrate(p, T.reshape(16), d, oracle.pack_color(rgb), 1.0, t, w, c)
                part += oracle.integrate_range(p, T.reshape(16), d, oracle.pack_color(rgb), 1.0, first, cnt, t2, w2, c2)
            s = slice(first, first + cnt)
            assert 0 < part < total and part == int(w[s].sum())
            assert np.array_equal(bits(t[s]), bits(t2)) and np.array_equal(w[s], w2) and np.array_equal(c[s], c2)


# ---- SEM_FUSIONPY geometry / weight / tsdf half pinned against the reference's own lines ---------
# With colour off, obs_weight = 1 and max_depth = 6 the plugin semantics differs from
# src/tsdf.cu:15-60 only in the storage order ([x][y][z] vs [z][y][x]) and in the truncation
# compare (`<` vs `<=`, src/tsdf.cu:49), and the compare matters only for a voxel whose
# diff == -trunc exactly.  On inputs with zero such ties (asserted) the two must agree bit for bit
# after a transpose.  Only the colour blend and the mesh stay unpinned (third-party, un-vendored).
def fusionpy_geometry(oracle, dims, vs, origin, hw, K, frames_c2b, trunc=None):
    trunc = np.float32(vs) * np.float32(5) if trunc is None else trunc
    p = oracle.params(dims, vs, trunc, origin, hw, K, 6.0, SEM_FUSIONPY)
    t, w, _ = oracle.new_volume(p)
    ties = 0
    for d, c2b in frames_c2b:
        ties += oracle.count_trunc_ties(p, c2b, d)
        oracle.integrate(p, c2b, d, None, 1.0, t, w, None)
    to_zyx = lambda a: np.ascontiguousarray(a.reshape(dims).transpose(2, 1, 0)).reshape(-1)  # noqa: E731
    return to_zyx(t), to_zyx(w), ties


def test_fusionpy_geometry_matches_reference_golden(oracle, golden_cu):
    g = golden_cu
    dims, vs, hw = tuple(int(x) for x in g["dims"]), float(g["vs"]), tuple(int(x) for x in g["hw"])
    t, w, ties = fusionpy_geometry(oracle, dims, vs, g["origin"], hw, g["K"], list(zip(g["depth"], g["cam2base"])))
    assert ties == 0, "golden input has diff == -trunc ties: `<` vs `<=` would matter"
    assert w.max() >= 2 and np.array_equal(w, g["weight"]) and np.array_equal(bits(t), bits(g["tsdf"]))


@pytest.mark.skipif(not ref_available(), reason="oracle/_ref not built (needs /root/reference at build time)")
def test_fusionpy_geometry_vs_reference_lift_random(oracle, synth):
    r = RefLift()
    rng = np.random.default_rng(23)
    for k, (dims, vs) in enumerate([((50, 44, 38), 0.03), ((36, 52, 64), 0.025), ((64, 64, 64), 0.02)]):
        origin, frames = synth.s_real(dims, vs, n_frames=4, seed=40 + k)
        base = synth.pose(synth.rot_xyz(*rng.uniform(-0.05, 0.05, 3)), rng.uniform(-0.1, 0.1, 3))
        fc = [(d, oracle.compose(base, T)) for d, _, T in frames]
        t, w, ties = fusionpy_geometry(oracle, dims, vs, origin, synth.TUM3_HW, synth.TUM3_K, fc)
        t2 = np.ones(int(np.prod(dims)), np.float32)
        w2 = np.zeros_like(t2)
        for d, c2b in fc:
            r.integrate(synth.TUM3_K, c2b, d, dims, origin, vs, np.float32(vs) * np.float32(5), t2, w2)
        assert ties == 0
        assert w2.max() >= 2 and np.array_equal(w, w2) and np.array_equal(bits(t), bits(t2))


def test_trunc_tie_counter_sees_a_planted_tie(oracle, synth):
    """The tie counter itself: a depth image made so that one voxel's diff is exactly -trunc."""
    dims, vs = (8, 8, 8), np.float32(0.125)
    trunc = np.float32(0.5)
    K = np.array([100, 0, 32, 0, 100, 32, 0, 0, 1], np.float32)
    T = np.eye(4, dtype=np.float32)
    origin = np.array([-0.5, -0.5, 1.0], np.float32)   # voxel (4,4,k) sits on the optical axis at z = 1 + k/8
    depth = np.full((64, 64), 1.0, np.float32)          # diff = 1 - (1 + k/8) == -0.5 exactly at k = 4
    for sem in (SEM_TSDFCU, SEM_FUSIONPY):
        p = oracle.params(dims, vs, trunc, origin, (64, 64), K, 6.0, sem)
        assert oracle.count_trunc_ties(p, T.reshape(16), depth) > 0
        t, w, _ = oracle.new_volume(p)
        oracle.integrate(p, T.reshape(16), depth, None, 1.0, t, w, None)
        idx = 4 * 64 + 4 * 8 + 4   # (4,4,4) in either order
        assert w[idx] == (0.0 if sem == SEM_TSDFCU else 1.0)   # `<=` skips the tie, `<` keeps it


def test_openmp_build_equals_strict_build(oracle, synth):
    """liboracle_omp.so (-O3 -march=x86-64-v3 -fopenmp, still -ffp-contract=off: the checker of the
    full-size GPU tests and the cpu_baseline port) computes the same bits as the strict build."""
    omp = Oracle(omp=True)
    dims, vs = (72, 56, 64), 0.03
    origin, frames = synth.s_real(dims, vs, n_frames=3, seed=17)
    for sem, color in ((SEM_TSDFCU, False), (SEM_FUSIONPY, True)):
        p = oracle.params(dims, vs, 5 * vs, origin, synth.TUM3_HW, synth.TUM3_K, 6.0, sem)
        a, b = oracle.new_volume(p, color), omp.new_volume(p, color)
        for d, rgb, T in frames:
            pc = oracle.pack_color(rgb) if color else None
            assert oracle.integrate(p, T.reshape(16), d, pc, 0.5, *a) == omp.integrate(p, T.reshape(16), d, pc, 0.5, *b)
        assert np.array_equal(bits(a[0]), bits(b[0])) and np.array_equal(a[1], b[1])
        assert not color or np.array_equal(a[2], b[2])
