"""Direct oracle parity at the BASELINE.json sizes: the CUDA path through the C
ABI against the OpenMP build of oracle/tsdf_oracle.c (same arithmetic as the strict build, checked
in tests/test_oracle.py) on the FULL grids of configs 1, 2 and one 128-slice slab of config 3, and
the geometry / weight / tsdf half of the plugin semantics against the reference's OWN lines
(tests/golden/tsdfcu_small.npz, made from oracle/_ref; oracle/_ref itself where it travelled).
Bar: weights, N_written and extracted index sets bit-exact; tsdf and colour bit-exact (tolerance
1e-5 abs written in TOL, never needed)."""
import numpy as np
import pytest

from conftest import bits
from oracle.oracle import Oracle, RefLift, ref_available, SEM_TSDFCU, SEM_FUSIONPY, PRED_NONZERO, PRED_ABS_LT

pytestmark = pytest.mark.gpu
TOL = 1e-5


@pytest.fixture(scope="module")
def omp():
    return Oracle(omp=True)


def compare(vol, t, w, c):
    gt, gw, gc = vol.download()
    assert np.array_equal(gw, w), f"weights differ at {np.flatnonzero(gw != w)[:5]}"
    assert np.max(np.abs(gt - t), initial=0) <= TOL
    assert np.array_equal(bits(gt), bits(t)), "tsdf not bit-exact"
    if c is not None:
        assert np.max(np.abs(gc - c), initial=0) <= TOL and np.array_equal(gc, c), "colour differs"
    return gt, gw, gc


def to_zyx(a, dims):
    return np.ascontiguousarray(a.reshape(dims).transpose(2, 1, 0)).reshape(-1)


def test_golden_fusionpy_geometry_from_reference_lines(pkg, oracle, golden_cu):
    """SEM_FUSIONPY with colour off, obs_weight = 1, max_depth = 6 on the reference-made golden
    input: equal to the reference's own output after the [x][y][z] -> [z][y][x] transpose (the
    input has no diff == -trunc tie, asserted, so `<` vs `<=` cannot matter)."""
    g = golden_cu
    dims, vs, hw = tuple(int(x) for x in g["dims"]), float(g["vs"]), tuple(int(x) for x in g["hw"])
    p = oracle.params(dims, vs, np.float32(vs) * np.float32(5), g["origin"], hw, g["K"], 6.0, SEM_FUSIONPY)
    assert sum(oracle.count_trunc_ties(p, c2b, d) for d, c2b in zip(g["depth"], g["cam2base"])) == 0
    cfg = pkg.make_config(dims, vs, g["origin"], semantics=SEM_FUSIONPY, with_color=False, im_hw=hw, K=g["K"],
                          max_depth=6.0, base2world=g["base2world"])
    with pkg.Volume(cfg) as vol:
        for d, T in zip(g["depth"], g["T_wc"]):
            vol.integrate(d, None, T, obs_weight=1.0)
        t, w, _ = vol.download()
    assert np.array_equal(to_zyx(w, dims), g["weight"])
    assert np.array_equal(bits(to_zyx(t, dims)), bits(g["tsdf"]))


@pytest.mark.parametrize("sem,color", [(SEM_TSDFCU, False), (SEM_FUSIONPY, True), (SEM_FUSIONPY, False)])
def test_c1_full_size_vs_oracle(pkg, omp, synth, sem, color):
    """BASELINE config 1: one 640x480 TUM3 frame into the full 256^3 grid at 2 cm (S-real room:
    partial coverage, holes, noise), both semantics.  The colour-off plugin run is also compared
    with the reference's own lines (oracle/_ref) where that library is present."""
    dims, vs = (256, 256, 256), 0.02
    origin, frames = synth.s_real(dims, vs, n_frames=1, seed=301)
    d, rgb, T = frames[0]
    md = 6.0
    cfg = pkg.make_config(dims, vs, origin, semantics=sem, with_color=color, max_depth=md)
    p = omp.params(dims, vs, cfg.trunc_margin, origin, synth.TUM3_HW, synth.TUM3_K, md, sem)
    t, w, c = omp.new_volume(p, color)
    n = omp.integrate(p, T.reshape(16), d, omp.pack_color(rgb) if color else None, 1.0, t, w, c)
    with pkg.Volume(cfg) as vol:
        vol.integrate(d, rgb if color else None, T)
        assert vol.stats().voxels_written == n and 0 < n < vol.n_local
        compare(vol, t, w, c)
        for pred, tt, wt in ((PRED_NONZERO, 1.2, 0.9), (PRED_ABS_LT, 0.5, 0.5)):
            gx, gr, gi = vol.extract_points(pred, tt, wt)
            ox, orr, oi = omp.extract(p, t, w, c, pred, tt, wt)
            assert len(oi) > 1000 and np.array_equal(gi, oi) and np.array_equal(bits(gx), bits(ox))
            if color:
                assert np.array_equal(gr, orr)
    if not color and ref_available("omp"):
        r = RefLift(omp=True)
        t2, w2 = np.ones(int(np.prod(dims)), np.float32), np.zeros(int(np.prod(dims)), np.float32)
        r.integrate(synth.TUM3_K, T.reshape(16), d, dims, origin, vs, cfg.trunc_margin, t2, w2)
        if sem == SEM_FUSIONPY:
            assert omp.count_trunc_ties(p, T.reshape(16), d) == 0
            t, w = to_zyx(t, dims), to_zyx(w, dims)
        assert np.array_equal(w, w2) and np.array_equal(bits(t), bits(t2)), "differs from the reference's own lines"


def test_c2_full_size_vs_oracle(pkg, omp, synth):
    """BASELINE config 2: fr3_office keyframe poses (committed fixture), S-real frames with colour
    into the full 512^3 grid at 1 cm, 5 keyframes; all three arrays, N_written per frame and the
    extracted index set / xyz / rgb against the oracle."""
    dims, vs = (512, 512, 512), 0.01
    poses = synth.load_pose_fixture()[::40]
    origin, frames = synth.s_real(dims, vs, n_frames=5, poses=poses)
    cfg = pkg.make_config(dims, vs, origin, semantics=SEM_FUSIONPY, with_color=True)
    p = omp.params(dims, vs, cfg.trunc_margin, origin, synth.TUM3_HW, synth.TUM3_K, cfg.max_depth, SEM_FUSIONPY)
    t, w, c = omp.new_volume(p, True)
    with pkg.Volume(cfg) as vol:
        for d, rgb, T in frames:
            n = omp.integrate(p, T.reshape(16), d, omp.pack_color(rgb), 1.0, t, w, c)
            vol.integrate(d, rgb, T)
            assert vol.stats().voxels_written == n > 0
        assert w.max() >= 3
        compare(vol, t, w, c)
        gx, gr, gi = vol.extract_points(PRED_ABS_LT, 0.5, 1.5)
        ox, orr, oi = omp.extract(p, t, w, c, PRED_ABS_LT, 0.5, 1.5)
        assert len(oi) > 50000 and np.array_equal(gi, oi)
        assert np.array_equal(bits(gx), bits(ox)) and np.array_equal(gr, orr)


@pytest.mark.parametrize("zr", [(128, 256), (0, 128)])
def test_c3_slab_vs_oracle(pkg, omp, synth, zr):
    """BASELINE config 3: KITTI-shaped 1241x376 stereo-depth frames into one 128-slice z-slab (what
    each of 8 GPUs owns) of the 1024x256x1024 outdoor grid at 10 cm, max_depth 50 m, colour."""
    dims, vs = (1024, 256, 1024), 0.1
    origin, frames = synth.s_kitti(n_frames=3, voxel_size=vs, dims=dims)
    cfg = pkg.make_config(dims, vs, origin, semantics=SEM_FUSIONPY, with_color=True, K=synth.KITTI_K,
                          im_hw=synth.KITTI_HW, max_depth=50.0, z_range=zr)
    p = omp.params(dims, vs, cfg.trunc_margin, origin, synth.KITTI_HW, synth.KITTI_K, 50.0, SEM_FUSIONPY, zr)
    t, w, c = omp.new_volume(p, True)
    with pkg.Volume(cfg) as vol:
        total = 0
        for d, rgb, T in frames:
            n = omp.integrate(p, T.reshape(16), d, omp.pack_color(rgb), 1.0, t, w, c)
            vol.integrate(d, rgb, T)
            assert vol.stats().voxels_written == n
            total += n
        assert total > 10000
        compare(vol, t, w, c)
        gx, gr, gi = vol.extract_points(PRED_ABS_LT, 0.8, 0.5)
        ox, orr, oi = omp.extract(p, t, w, c, PRED_ABS_LT, 0.8, 0.5)
        assert len(oi) > 1000 and np.array_equal(gi, oi)
        assert np.array_equal(bits(gx), bits(ox)) and np.array_equal(gr, orr)
