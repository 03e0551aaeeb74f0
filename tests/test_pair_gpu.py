"""Oracle parity of the two-frames-per-pass dense kernel (k_integrate2, csrc/integrate.cu) — the path
tsdf_integrate_multi_async takes when the slab lies inside both frames' frusta (the
off-line keyframe replay of examples/label_instance_rgbd.cpp:79-110 hands the library many frames per
volume).  Every test (1) compares the volume bit for bit with oracle.integrate run frame by frame
(weights, tsdf, colour, N_written) and (2) PROVES the pair kernel ran, from the library's own launch
counter: a k-frame call must cost k prep + ceil(k/2) integrate + ceil(k/2) publish launches; if pairing
silently did not happen the integrate count is k and the test fails.
Bar: bit-exact (TOL written for the record, never needed)."""
import ctypes as C

import numpy as np
import pytest

from conftest import bits
from oracle.oracle import Oracle, SEM_TSDFCU, SEM_FUSIONPY

pytestmark = pytest.mark.gpu
TOL = 1e-5


@pytest.fixture(scope="module")
def omp():
    return Oracle(omp=True)


def launches(pkg):
    n = C.c_int64(0)
    pkg.capi.check(pkg.load().tsdf_launch_count(C.byref(n)))
    return n.value


def expect_paired(k):
    return k + 2 * ((k + 1) // 2)        # preps + integrates (pairs, odd tail alone) + publishes


def expect_unpaired(k):
    return 2 * k + (k + 1) // 2


def compare(vol, t, w, c):
    gt, gw, gc = vol.download()
    assert np.array_equal(gw, w), f"weights differ at {np.flatnonzero(gw != w)[:5]}"
    assert np.max(np.abs(gt - t), initial=0) <= TOL and np.array_equal(bits(gt), bits(t)), "tsdf not bit-exact"
    if c is not None:
        assert np.array_equal(gc, c), "colour differs"


def run_multi(pkg, omp, synth, dims, vs, frames, origin, sem, color, obs, zr, k, paired=True):
    md = 6.0 if sem == SEM_TSDFCU else np.inf
    cfg = pkg.make_config(dims, vs, origin, semantics=sem, with_color=color, z_range=zr, max_depth=md)
    p = omp.params(dims, vs, cfg.trunc_margin, origin, synth.TUM3_HW, synth.TUM3_K, md, sem, zr)
    t, w, c = omp.new_volume(p, color)
    fr = frames[:k]
    want = 0
    for d, rgb, T in fr:
        want += omp.integrate(p, T.reshape(16), d, omp.pack_color(rgb) if color else None, obs, t, w, c)
    with pkg.Volume(cfg) as vol:
        # one warm call so that the slots exist (their one-time set-up launches nothing, but keep the count clean)
        vol.integrate_multi([f[0] for f in fr], [f[1] for f in fr] if color else None, [f[2] for f in fr], obs_weight=obs)
        vol.reset()
        l0 = launches(pkg)
        vol.integrate_multi([f[0] for f in fr], [f[1] for f in fr] if color else None, [f[2] for f in fr], obs_weight=obs)
        l1 = launches(pkg)
        assert vol.stats().voxels_written == want and want > 0
        compare(vol, t, w, c)
    got = l1 - l0
    if k >= 2:
        assert expect_paired(k) != expect_unpaired(k)
    assert got == (expect_paired(k) if paired else expect_unpaired(k)), \
        f"k={k}: {got} launches, paired form costs {expect_paired(k)}, unpaired {expect_unpaired(k)}"


@pytest.fixture(scope="module")
def dense_scene(synth):
    dims, vs = (128, 96, 128), 0.012
    origin, frames = synth.s_full(dims, vs, n_frames=8, sphere=True)
    return dims, vs, origin, frames


@pytest.mark.parametrize("k", [2, 3, 8])
@pytest.mark.parametrize("sem,color,obs", [(SEM_FUSIONPY, True, 1.0), (SEM_FUSIONPY, True, 0.5),
                                           (SEM_FUSIONPY, False, 1.0), (SEM_TSDFCU, False, 1.0)])
def test_pair_kernel_vs_oracle_whole_grid(pkg, omp, synth, dense_scene, sem, color, obs, k):
    """Dense frames (S-full + sphere: every voxel in view, a real surface, a shadow), k = 2, 3 (odd tail),
    8; both semantics; colour on/off; obs_weight 1 (integer-weight blend, WMODE 1) and 0.5 (general blend)."""
    dims, vs, origin, frames = dense_scene
    run_multi(pkg, omp, synth, dims, vs, frames, origin, sem, color, obs, None, k)


@pytest.mark.parametrize("zr", [(0, 32), (96, 128)])
def test_pair_kernel_vs_oracle_z_slab(pkg, omp, synth, dense_scene, zr):
    """One z-slab of the same grid (32-voxel rows in the z-fast layout; the far slab holds the sphere)."""
    dims, vs, origin, frames = dense_scene
    run_multi(pkg, omp, synth, dims, vs, frames, origin, SEM_FUSIONPY, True, 1.0, zr, 5)


def test_sparse_frames_are_not_paired(pkg, omp, synth):
    """The negative of the launch-count proof: S-real frames do not contain the slab, so every frame is
    its own launch — the counter tells the two forms apart."""
    dims, vs = (64, 48, 64), 0.03
    origin, frames = synth.s_real(dims, vs, n_frames=4, seed=77)
    run_multi(pkg, omp, synth, dims, vs, frames, origin, SEM_FUSIONPY, True, 1.0, None, 4, paired=False)


def test_bench_workload_slab_vs_oracle(pkg, omp, synth):
    """The bench workload itself (C4: 1024^3 @ 2 mm, S-full + sphere, colour): the last 128-slice slab
    (the one the sphere is in; 134 M voxels, 1024-voxel rows would be the whole grid's — this is the
    8-GPU slab shape) against the OpenMP oracle, three frames: one single launch, then one pair."""
    dims, vs, zr = (1024, 1024, 1024), 0.002, (896, 1024)
    origin, frames = synth.s_full(dims, vs, n_frames=3, sphere=True)
    cfg = pkg.make_config(dims, vs, origin, semantics=SEM_FUSIONPY, with_color=True, z_range=zr)
    p = omp.params(dims, vs, cfg.trunc_margin, origin, synth.TUM3_HW, synth.TUM3_K, np.inf, SEM_FUSIONPY, zr)
    t, w, c = omp.new_volume(p, True)
    with pkg.Volume(cfg) as vol:
        d, rgb, T = frames[0]
        n0 = omp.integrate(p, T.reshape(16), d, omp.pack_color(rgb), 1.0, t, w, c)
        vol.integrate(d, rgb, T)
        assert vol.stats().voxels_written == n0
        l0 = launches(pkg)
        vol.integrate_multi([f[0] for f in frames[1:]], [f[1] for f in frames[1:]], [f[2] for f in frames[1:]])
        assert launches(pkg) - l0 == expect_paired(2)
        n12 = sum(omp.integrate(p, T.reshape(16), d, omp.pack_color(rgb), 1.0, t, w, c) for d, rgb, T in frames[1:])
        assert vol.stats().voxels_written == n12
        assert 0.9 * vol.n_local < n0 < vol.n_local          # dense, but the sphere's inside/shadow is skipped
        compare(vol, t, w, c)


def test_inside_kernels_equal_bounds_checked_kernels(pkg, omp, synth, dense_scene, dev_lib):
    """When the host proves that every voxel projects strictly inside the image (slab_inside_with_margin,
    csrc/integrate.cu) the dense kernels drop the image-bounds tests of src/tsdf.cu:39,43 (CULL = 3 /
    k_integrate2<INSIDE>).  Variant bit 15 keeps the tests: both forms must give the same bits, single
    launches and pairs, and equal the oracle."""
    dims, vs, origin, frames = dense_scene
    cfg = pkg.make_config(dims, vs, origin, semantics=SEM_FUSIONPY, with_color=True)
    p = omp.params(dims, vs, cfg.trunc_margin, origin, synth.TUM3_HW, synth.TUM3_K, np.inf, SEM_FUSIONPY)
    t, w, c = omp.new_volume(p, True)
    for d, rgb, T in frames[:5]:
        omp.integrate(p, T.reshape(16), d, omp.pack_color(rgb), 1.0, t, w, c)
    out = {}
    for variant in (0, 1 << 15):
        with pkg.Volume(cfg) as vol:
            vol.set_variant(variant)
            vol.integrate(*frames[0][:2], frames[0][2])                                    # single launch
            vol.integrate_multi([f[0] for f in frames[1:5]], [f[1] for f in frames[1:5]], [f[2] for f in frames[1:5]])  # two pairs
            compare(vol, t, w, c)
            out[variant] = vol.download()
    for a, b in zip(out[0], out[1 << 15]):
        assert np.array_equal(bits(a), bits(b))
