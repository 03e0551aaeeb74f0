"""The conservative segment cull of the integrate kernel (csrc/integrate.cu box_is_skipped) must
never change a result: it may only skip 128-voxel segments in which the reference's own tests
(src/tsdf.cu:39 behind the camera, :43 outside the image, :49 beyond the surface) reject every
voxel.  Checked against the oracle and against the same kernel with the cull switched off
(variant bit 6), on poses chosen to put segment ends on both sides of every frustum plane."""
import numpy as np
import pytest

from conftest import bits
from oracle.oracle import SEM_TSDFCU, SEM_FUSIONPY
from test_parity_gpu import oracle_run, assert_volume_equal

pytestmark = [pytest.mark.gpu, pytest.mark.usefixtures("dev_lib")]   # variants: libtsdf_b200_dev.so
NO_CULL, FORCE_CULL = 64, 128   # tsdf_set_variant bits (default: chosen per launch by the host)


def run_gpu(pkg, cfg, frames, color, variant, fused=False):
    with pkg.Volume(cfg) as vol:
        vol.set_variant(variant)
        written = []
        if fused:
            vol.integrate_multi([f[0] for f in frames], [f[1] for f in frames] if color else None,
                                [f[2] for f in frames])
        else:
            for d, rgb, T in frames:
                vol.integrate(d, rgb if color else None, T)
                written.append(vol.stats().voxels_written)
        return vol.download() + (written,)


def random_poses(synth, rng, center, radius, n):
    """Cameras around (and inside) the volume looking in arbitrary directions: most of the grid is
    outside the frustum, rows cross its planes at every angle, some cameras face away."""
    out = []
    for i in range(n):
        eye = center + rng.normal(size=3) * radius * (0.2 if i % 3 == 0 else 1.0)
        target = center + rng.normal(size=3) * radius * (0.3 if i % 4 else 3.0)
        T = synth.look_at(eye, target)
        T[:3, :3] = T[:3, :3] @ synth.rot_xyz(0, 0, rng.uniform(-3.1, 3.1))[:3, :3]  # roll
        out.append(np.ascontiguousarray(T, np.float32))
    return out


@pytest.mark.parametrize("sem,color", [(SEM_FUSIONPY, True), (SEM_TSDFCU, False)])
@pytest.mark.parametrize("dims", [(160, 24, 288), (40, 36, 132), (300, 20, 24)])
def test_cull_never_changes_results(pkg, oracle, synth, sem, color, dims):
    vs = 0.02
    rng = np.random.default_rng(dims[0] * 7 + sem)
    ext = np.array(dims, np.float32) * vs
    origin = np.array([-ext[0] / 2, -ext[1] / 2, 0.5], np.float32)
    center = origin + ext / 2
    scene = synth.room_scene(center, ext.max() * 0.45).sphere(tuple(center), float(ext.min()) * 0.3)
    poses = random_poses(synth, rng, center, float(ext.max()) * 0.6, 10)
    frames = []
    for T in poses:
        d, rgb = synth.render(scene, synth.TUM3_K, synth.TUM3_HW, T, rng, noise_sigma=0.002, hole_frac=0.05)
        frames.append((d, rgb, T))
    cfg = pkg.make_config(dims, vs, origin, semantics=sem, with_color=color)
    p, t, w, c, ow = oracle_run(oracle, dims, vs, origin, frames, sem, color, max_depth=cfg.max_depth)
    assert 0 < sum(ow) < 0.5 * len(frames) * np.prod(dims)   # sparse: the cull has work to do
    gt, gw, gc, written = run_gpu(pkg, cfg, frames, color, FORCE_CULL)
    assert written == ow
    assert_volume_equal(gt, gw, gc, t, w, c)
    if color:
        assert np.array_equal(gc, c)
    nt, nw, nc, nwritten = run_gpu(pkg, cfg, frames, color, NO_CULL)
    assert nwritten == ow and np.array_equal(bits(nt), bits(gt)) and np.array_equal(nw, gw)
    ft, fw, fc, _ = run_gpu(pkg, cfg, frames[:8], color, FORCE_CULL, fused=True)
    p8, t8, w8, c8, _ = oracle_run(oracle, dims, vs, origin, frames[:8], sem, color, max_depth=cfg.max_depth)
    assert_volume_equal(ft, fw, fc, t8, w8, c8)
    # dense kernels forced: consecutive frames go through the two-frames-per-pass kernel (k_integrate2,
    # launch_integrate pairs under variant 64), which here meets frustum edges, cameras facing away,
    # holes and surfaces — frames its host-side "slab inside both frusta" test would never hand it
    pt, pw, pc, _ = run_gpu(pkg, cfg, frames[:7], color, NO_CULL, fused=True)
    p7, t7, w7, c7, _ = oracle_run(oracle, dims, vs, origin, frames[:7], sem, color, max_depth=cfg.max_depth)
    assert_volume_equal(pt, pw, pc, t7, w7, c7)
    if color:
        assert np.array_equal(pc, c7)


def test_depth_cull_wall_in_front(pkg, oracle, synth):
    """A wall close to the camera hides almost the whole grid: everything farther than the
    frame's largest depth + trunc is skipped (src/tsdf.cu:49); one NaN pixel must disable that
    shortcut without changing the result (the reference integrates NaN depth)."""
    dims, vs = (64, 48, 256), 0.02
    origin = np.array([-0.64, -0.48, 0.4], np.float32)
    T = synth.pose()
    rng = np.random.default_rng(3)
    scene = synth.Scene().plane((0, 0, 1), 1.0)
    d, rgb = synth.render(scene, synth.TUM3_K, synth.TUM3_HW, T, rng, noise_sigma=0.003)
    d_nan = d.copy()
    d_nan[200, 300] = np.nan
    for depth in (d, d_nan):
        frames = [(depth, rgb, T)]
        cfg = pkg.make_config(dims, vs, origin, semantics=SEM_FUSIONPY, with_color=True)
        p, t, w, c, ow = oracle_run(oracle, dims, vs, origin, frames, SEM_FUSIONPY, True)
        gt, gw, gc, written = run_gpu(pkg, cfg, frames, True, FORCE_CULL)
        assert written == ow and 0 < ow[0] < 0.2 * np.prod(dims)
        assert np.array_equal(gw, w) and np.array_equal(bits(gt), bits(t)) and np.array_equal(bits(gc), bits(c))


def test_kitti_shaped_config3(pkg, oracle, synth):
    """BASELINE config 3 in small: 1241x376 stereo-depth frames (KITTI00-02 intrinsics, depth
    quantised like bf/disparity, max_depth 50 m) into an outdoor grid of the same 4:1:4 aspect,
    whole grid and 8 z-slabs."""
    dims, vs = (128, 32, 128), 0.8
    origin, frames = synth.s_kitti(n_frames=3, voxel_size=vs, dims=dims)
    H, W = synth.KITTI_HW
    cfg = pkg.make_config(dims, vs, origin, semantics=SEM_FUSIONPY, with_color=True, K=synth.KITTI_K, im_hw=(H, W),
                          max_depth=50.0)
    p, t, w, c, ow = oracle_run(oracle, dims, vs, origin, frames, SEM_FUSIONPY, True, K=synth.KITTI_K, hw=(H, W),
                                max_depth=50.0)
    assert sum(ow) > 0
    gt, gw, gc, written = run_gpu(pkg, cfg, frames, True, FORCE_CULL)
    assert written == ow
    assert_volume_equal(gt, gw, gc, t, w, c)
    t3, w3, c3 = (a.reshape(dims) for a in (t, w, c))
    for r in range(8):
        zr = pkg.slab_range(dims[2], 8, r)
        cfg_r = pkg.make_config(dims, vs, origin, semantics=SEM_FUSIONPY, with_color=True, K=synth.KITTI_K,
                                im_hw=(H, W), max_depth=50.0, z_range=zr)
        st, sw, sc, _ = run_gpu(pkg, cfg_r, frames, True, 0)
        sl = (slice(None), slice(None), slice(*zr))
        assert np.array_equal(sw, w3[sl].reshape(-1)) and np.array_equal(bits(st), bits(t3[sl].reshape(-1)))
        assert np.array_equal(sc, c3[sl].reshape(-1))


STAGED = 1 << 20   # tsdf_set_variant bit 20: staged image tile (TMA box in shared memory) in the sparse kernel


@pytest.mark.parametrize("sem", [SEM_FUSIONPY, SEM_TSDFCU])
def test_staged_image_tile_reads_the_same_pixels(pkg, oracle, synth, sem):
    """The staged variant of the sparse kernel (cp.async.bulk.tensor.2d box of the prepared frame in
    shared memory when a block's pixel rectangle fits 64 x 32 pixels) must gather the same bits as the
    __ldg path (src/tsdf.cu:45: same pixel index -> same value).  Two geometries: a small far volume
    whose block tiles all fit the box (straddling the sphere's surface and its shadow), and random
    poses around a larger volume where almost none fits (the fallback inside the same kernel)."""
    big = (512, 512, 512)
    _, frames = synth.s_full(big, 0.004, n_frames=3, sphere=True)
    z_near = float(_[2])
    r = 0.15 * 2.048
    dims, vs = (128, 128, 128), 0.001
    origin = np.array([-0.064 + 0.11, -0.064 - 0.05, z_near + 2.048 - 1.5 * r - 0.064], np.float32)
    cases = [(dims, vs, origin, frames)]
    d2, v2 = (160, 24, 288), 0.02
    rng = np.random.default_rng(77)
    ext = np.array(d2, np.float32) * v2
    o2 = np.array([-ext[0] / 2, -ext[1] / 2, 0.5], np.float32)
    center = o2 + ext / 2
    scene = synth.room_scene(center, ext.max() * 0.45).sphere(tuple(center), float(ext.min()) * 0.3)
    fr2 = []
    for T in random_poses(synth, rng, center, float(ext.max()) * 0.6, 6):
        d, rgb = synth.render(scene, synth.TUM3_K, synth.TUM3_HW, T, rng, noise_sigma=0.002, hole_frac=0.05)
        fr2.append((d, rgb, T))
    cases.append((d2, v2, o2, fr2))
    for dm, vsz, org, frs in cases:
        cfg = pkg.make_config(dm, vsz, org, semantics=sem, with_color=True, max_depth=20.0)
        p, t, w, c, ow = oracle_run(oracle, dm, vsz, org, frs, sem, True, max_depth=20.0)
        assert sum(ow) > 1000
        a = run_gpu(pkg, cfg, frs, True, FORCE_CULL)
        b = run_gpu(pkg, cfg, frs, True, FORCE_CULL | STAGED)
        assert a[3] == b[3] == ow
        for x, y in zip(a[:3], b[:3]):
            assert np.array_equal(bits(x), bits(y))
        assert_volume_equal(b[0], b[1], b[2], t, w, c)
    # the far volume must not be trivially free space: the sphere's surface crosses it
    dm, vsz, org, frs = cases[0]
    p, t, w, c, ow = oracle_run(oracle, dm, vsz, org, frs, sem, True, max_depth=20.0)
    assert (t < 0.5).any() and (w == 0).any() and (w == 3).any()
