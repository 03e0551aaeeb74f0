"""GPU parity tests proper: the CUDA path (through the C ABI) against the CPU oracle on the same
seeded inputs, against the committed golden vectors, and — at BASELINE sizes — through
size-independent properties.  Bar (BASELINE.json north_star): weights and extracted index
sets bit-exact; tsdf and colour within 1e-5 abs.  The implementation is in fact bit-exact on
tsdf too, and the tests assert that (any non-zero diff is a bug signal)."""
import os

import numpy as np
import pytest

from conftest import bits
from oracle.oracle import SEM_TSDFCU, SEM_FUSIONPY, PRED_NONZERO, PRED_ABS_LT

pytestmark = pytest.mark.gpu
TOL = 1e-5  # north_star tolerance for tsdf / colour


def oracle_run(oracle, dims, vs, origin, frames, sem, color, obs=1.0, base=None, K=None, hw=None,
               max_depth=None, z_range=None, trunc=None):
    K = K if K is not None else np.array([535.4, 0, 320.1, 0, 539.2, 247.6, 0, 0, 1], np.float32)
    hw = hw or frames[0][0].shape
    if max_depth is None:
        max_depth = 6.0 if sem == SEM_TSDFCU else np.inf
    trunc = trunc if trunc is not None else np.float32(vs) * np.float32(5)
    p = oracle.params(dims, vs, trunc, origin, hw, K, max_depth, sem, z_range)
    t, w, c = oracle.new_volume(p, color)
    written = []
    for d, rgb, T in frames:
        c2b = T.reshape(16) if base is None else oracle.compose(base, T)
        written.append(oracle.integrate(p, c2b, d, oracle.pack_color(rgb) if color else None, obs, t, w, c))
    return p, t, w, c, written


def assert_volume_equal(gt, gw, gc, t, w, c):
    assert np.array_equal(gw, w), f"weights differ at {np.flatnonzero(gw != w)[:5]}"
    assert np.max(np.abs(gt - t), initial=0) <= TOL
    assert np.array_equal(bits(gt), bits(t)), "tsdf not bit-exact"
    if c is not None:
        assert np.max(np.abs(gc - c), initial=0) <= TOL, "colour differs"


@pytest.mark.parametrize("sem,color", [(SEM_TSDFCU, False), (SEM_FUSIONPY, True)])
@pytest.mark.parametrize("dims", [(1100, 6, 10), (10, 6, 1100), (1027, 5, 3), (3, 5, 1027), (2052, 4, 4), (4, 4, 2052)])
def test_long_and_ragged_fast_axis(pkg, oracle, synth, sem, color, dims):
    """Rows longer than one block's 1024 voxels (several blocks along the fast axis) and row
    lengths that are not a multiple of 4 (scalar path, partial last warp), in both storage
    orders: S-full frames (every voxel written) and the same frames with the camera moved
    sideways and turned so that part of the long axis leaves the image."""
    vs = 0.004 if max(dims) < 2000 else 0.002     # keeps the wall behind the volume within max_depth
    origin, frames = synth.s_full(dims, vs, n_frames=2)
    turned = [(d, rgb, (T.reshape(4, 4) @ synth.pose(synth.rot_xyz(0.02 * (i + 1), 0.2 - 0.45 * i, 0.05),
                                                     (1.1 - 2.4 * i, 0.05, 0.0))).astype(np.float32))
              for i, (d, rgb, T) in enumerate(frames)]
    frames = frames + turned
    p, t, w, c, written = oracle_run(oracle, dims, vs, origin, frames, sem, color, max_depth=6.0)
    cfg = pkg.make_config(dims, vs, origin, semantics=sem, with_color=color, max_depth=6.0)
    with pkg.Volume(cfg) as vol:
        for (d, rgb, T), ow in zip(frames, written):
            vol.integrate(d, rgb if color else None, T)
            assert vol.stats().voxels_written == ow
        gt, gw, gc = vol.download()
    n = int(np.prod(dims))
    assert written[0] == n and any(0 < x < n for x in written[2:]), written
    assert_volume_equal(gt, gw, gc, t, w, c)


@pytest.mark.parametrize("sem,color", [(SEM_TSDFCU, False), (SEM_FUSIONPY, True)])
@pytest.mark.parametrize("dims", [(48, 40, 1), (1, 40, 48), (48, 40, 4), (4, 40, 48), (2, 2, 2), (1, 1, 1), (5, 1, 7)])
def test_degenerate_axes(pkg, oracle, synth, sem, color, dims):
    """Axes of 1-4 voxels (a fast axis shorter than one vector, a single row, a single voxel): the
    thread-to-row mapping must keep one pass of rows inside the block's row table."""
    vs = 0.05
    origin, frames = synth.s_full(dims, vs, n_frames=2)
    frames = frames + [(d, rgb, (T.reshape(4, 4) @ synth.pose(synth.rot_xyz(0.03, 0.25 - 0.5 * i, 0.1),
                                                              (0.6 - 1.2 * i, 0.1, 0.0))).astype(np.float32))
                       for i, (d, rgb, T) in enumerate(frames)]
    p, t, w, c, written = oracle_run(oracle, dims, vs, origin, frames, sem, color, max_depth=6.0)
    cfg = pkg.make_config(dims, vs, origin, semantics=sem, with_color=color, max_depth=6.0)
    with pkg.Volume(cfg) as vol:
        for (d, rgb, T), ow in zip(frames, written):
            vol.integrate(d, rgb if color else None, T)
            assert vol.stats().voxels_written == ow
        gt, gw, gc = vol.download()
        n_pts = vol.extract_count(PRED_NONZERO, 1.2, 0.9)
    assert written[0] == int(np.prod(dims)) and n_pts == int(((t != 0) & (w > 0.9)).sum())
    assert_volume_equal(gt, gw, gc, t, w, c)


@pytest.mark.parametrize("sem,color", [(SEM_TSDFCU, False), (SEM_FUSIONPY, True), (SEM_FUSIONPY, False)])
@pytest.mark.parametrize("dims", [(64, 48, 64), (50, 44, 38), (33, 17, 21), (128, 8, 260)])
def test_integrate_matches_oracle(pkg, oracle, synth, sem, color, dims):
    vs = 0.03
    origin, frames = synth.s_real(dims, vs, n_frames=4, seed=21)
    base = frames[1][2] if sem == SEM_TSDFCU else None
    obs = 1.0 if sem == SEM_TSDFCU else 0.5
    p, t, w, c, written = oracle_run(oracle, dims, vs, origin, frames, sem, color, obs, base)
    cfg = pkg.make_config(dims, vs, origin, semantics=sem, with_color=color, base2world=base)
    with pkg.Volume(cfg) as vol:
        for (d, rgb, T), ow in zip(frames, written):
            vol.integrate(d, rgb if color else None, T, obs_weight=obs)
            st = vol.stats()
            assert st.voxels_written == ow and st.voxels_visited == vol.n_local
        gt, gw, gc = vol.download()
    assert sum(written) > 0
    assert_volume_equal(gt, gw, gc, t, w, c)


def test_golden_vectors_from_reference_lines(pkg, golden_cu, tmp_path):
    """CUDA path vs outputs of the reference's own src/tsdf.cu lines (tests/golden)."""
    g = golden_cu
    dims, vs, hw = tuple(int(x) for x in g["dims"]), float(g["vs"]), tuple(int(x) for x in g["hw"])
    cfg = pkg.make_config(dims, vs, g["origin"], semantics=SEM_TSDFCU, with_color=False, im_hw=hw,
                          K=g["K"], base2world=g["base2world"], vol_id=7)
    with pkg.Volume(cfg) as vol:
        for d, T in zip(g["depth"], g["T_wc"]):
            vol.integrate(d, None, T)
        gt, gw, _ = vol.download()
        ply, binf = str(tmp_path / "tsdf7.ply"), str(tmp_path / "tsdf7.bin")
        vol.save_ply(ply, PRED_NONZERO, 1.2, 0.9)
        vol.save_bin(binf)
    assert np.array_equal(gw, g["weight"]) and np.array_equal(bits(gt), bits(g["tsdf"]))
    assert np.array_equal(np.frombuffer(open(ply, "rb").read(), np.uint8), g["ply"])  # byte-compatible PLY
    raw = np.fromfile(binf, np.float32)
    assert list(raw[:3]) == list(map(float, dims)) and np.array_equal(bits(raw[8:]), bits(g["tsdf"]))
    assert raw[6] == np.float32(vs) and raw[7] == np.float32(vs) * np.float32(5)


def test_golden_fusionpy_regression(pkg, golden_cu, golden_py):
    g = golden_cu
    dims, vs, hw = tuple(int(x) for x in g["dims"]), float(g["vs"]), tuple(int(x) for x in g["hw"])
    cfg = pkg.make_config(dims, vs, g["origin"], semantics=SEM_FUSIONPY, with_color=True, im_hw=hw, K=g["K"])
    with pkg.Volume(cfg) as vol:
        for d, rgb, T in zip(g["depth"], g["rgb"], g["T_wc"]):
            vol.integrate(d, rgb, T, obs_weight=float(golden_py["obs_weight"]))
        gt, gw, gc = vol.download()
        xyz, rgb, idx = vol.extract_points()
    assert_volume_equal(gt, gw, gc, golden_py["tsdf"], golden_py["weight"], golden_py["color"])
    assert np.array_equal(idx, golden_py["idx"]) and np.array_equal(rgb, golden_py["rgb"])
    assert np.array_equal(bits(xyz), bits(golden_py["xyz"]))


@pytest.mark.parametrize("sem", [SEM_TSDFCU, SEM_FUSIONPY])
def test_extraction_index_sets_bit_exact(pkg, oracle, synth, sem):
    dims, vs = (72, 40, 56), 0.03
    color = sem == SEM_FUSIONPY
    origin, frames = synth.s_real(dims, vs, n_frames=5, seed=31)
    p, t, w, c, _ = oracle_run(oracle, dims, vs, origin, frames, sem, color)
    cfg = pkg.make_config(dims, vs, origin, semantics=sem, with_color=color)
    with pkg.Volume(cfg) as vol:
        for d, rgb, T in frames:
            vol.integrate(d, rgb if color else None, T, sync=False)
        for pred, tt, wt in ((PRED_NONZERO, 1.2, 0.9), (PRED_ABS_LT, 0.4, 1.5), (PRED_ABS_LT, 0.0, 0.0)):
            oxyz, orgb, oidx = oracle.extract(p, t, w, c, pred, tt, wt)
            assert vol.extract_count(pred, tt, wt) == len(oidx)
            xyz, rgb, idx = vol.extract_points(pred, tt, wt)
            assert np.array_equal(idx, oidx)
            assert np.array_equal(bits(xyz), bits(oxyz))
            if color:
                assert np.array_equal(rgb, orgb)
        # capacity protocol
        import ctypes as C
        n = C.c_int64(0)
        buf = np.zeros(3, np.float32)
        rc = vol.lib.tsdf_extract_points(vol.h, PRED_NONZERO, np.float32(1.2), np.float32(0.9),
                                         buf.ctypes.data, None, None, 1, C.byref(n))
        assert rc == pkg.capi.E_CAPACITY and n.value == len(oracle.extract(p, t, w, c)[2])


def test_empty_and_degenerate_frames(pkg, oracle, synth):
    dims, vs = (32, 32, 32), 0.05
    origin = np.array([-0.8, -0.8, 1.0], np.float32)
    H, W = 480, 640
    I = np.eye(4, dtype=np.float32)
    zeros = np.zeros((H, W), np.float32)
    far = np.full((H, W), 7.5, np.float32)            # > max_depth 6 (src/tsdf.cu:46)
    neg = np.full((H, W), -1.0, np.float32)
    behind = synth.pose(synth.rot_xyz(0, np.pi, 0), (0, 0, 0))  # volume behind the camera (:39)
    wall = np.full((H, W), 2.0, np.float32)
    wall[::7, ::5] = 0.0
    nanny = wall.copy()
    nanny[150:330, 220:420] = np.nan                    # NaN depth is integrated by the reference
    rgb = np.zeros((H, W, 3), np.uint8)
    for sem in (SEM_TSDFCU, SEM_FUSIONPY):
        frames = [(zeros, rgb, I), (far, rgb, I), (neg, rgb, I), (wall, rgb, behind), (wall, rgb, I), (nanny, rgb, I)]
        p, t, w, c, written = oracle_run(oracle, dims, vs, origin, frames, sem, False, max_depth=6.0)
        assert written[0] == written[2] == written[3] == 0 and written[4] > 0
        cfg = pkg.make_config(dims, vs, origin, semantics=sem, with_color=False, max_depth=6.0)
        with pkg.Volume(cfg) as vol:
            for (d, _, T), ow in zip(frames, written):
                vol.integrate(d, None, T)
                assert vol.stats().voxels_written == ow
            gt, gw, _ = vol.download()
        # a NaN sample passes both validity tests of src/tsdf.cu:46 and fmin(1, NaN) = 1, so the
        # voxels behind the NaN patch ARE updated (weight 2 after the two wall frames)
        assert np.array_equal(gw, w) and w.max() == 2
        assert np.array_equal(bits(gt), bits(t))


def test_reset_and_upload_roundtrip(pkg, synth):
    dims, vs = (40, 40, 40), 0.04
    origin, frames = synth.s_real(dims, vs, n_frames=2, seed=41)
    cfg = pkg.make_config(dims, vs, origin)
    with pkg.Volume(cfg) as vol:
        for d, rgb, T in frames:
            vol.integrate(d, rgb, T)
        a = vol.download()
        vol.reset()
        t0, w0, c0 = vol.download()
        assert (t0 == 1).all() and (w0 == 0).all() and (c0 == 0).all()   # src/tsdf.cu:79-81
        vol.upload(*a)
        b = vol.download()
        assert all(np.array_equal(bits(x), bits(y)) for x, y in zip(a, b))
        vol.reset()
        for d, rgb, T in frames:
            vol.integrate(d, rgb, T)
        b = vol.download()
        assert all(np.array_equal(bits(x), bits(y)) for x, y in zip(a, b)), "not deterministic"


@pytest.mark.parametrize("sem", [SEM_TSDFCU, SEM_FUSIONPY])
def test_fused_multi_frame_equals_sequential(pkg, oracle, synth, sem):
    dims, vs = (48, 40, 64), 0.035
    color = sem == SEM_FUSIONPY
    origin, frames = synth.s_real(dims, vs, n_frames=11, seed=51)   # 8 + 2 + 1 fused launches
    p, t, w, c, written = oracle_run(oracle, dims, vs, origin, frames, sem, color, 0.5 if color else 1.0)
    cfg = pkg.make_config(dims, vs, origin, semantics=sem, with_color=color)
    with pkg.Volume(cfg) as vol:
        vol.integrate_multi([f[0] for f in frames], [f[1] for f in frames] if color else None,
                            [f[2] for f in frames], obs_weight=0.5 if color else 1.0)
        gt, gw, gc = vol.download()
    assert_volume_equal(gt, gw, gc, t, w, c)


@pytest.mark.parametrize("sem", [SEM_TSDFCU, SEM_FUSIONPY])
def test_z_slabs_equal_whole_grid(pkg, oracle, synth, sem):
    """Multi-GPU sharding emulated on one GPU: 3 slab handles == 1 whole-grid handle == oracle."""
    dims, vs = (40, 32, 72), 0.04
    color = sem == SEM_FUSIONPY
    origin, frames = synth.s_real(dims, vs, n_frames=3, seed=61)
    p, t, w, c, _ = oracle_run(oracle, dims, vs, origin, frames, sem, color)
    n = dims[0] * dims[1] * dims[2]
    T_, W_, C_ = np.zeros(n, np.float32), np.zeros(n, np.float32), (np.zeros(n, np.float32) if color else None)
    pts = []
    for r in range(3):
        zr = pkg.slab_range(dims[2], 3, r)
        cfg = pkg.make_config(dims, vs, origin, semantics=sem, with_color=color, z_range=zr)
        with pkg.Volume(cfg) as vol:
            for d, rgb, T in frames:
                vol.integrate(d, rgb if color else None, T)
            vol.download_into(T_, W_, C_)
            pts.append(vol.extract_points())
    assert_volume_equal(T_, W_, C_, t, w, c)
    oxyz, orgb, oidx = oracle.extract(p, t, w, c)
    idx = np.concatenate([q[2] for q in pts])
    xyz = np.concatenate([q[0] for q in pts])
    order = np.argsort(idx, kind="stable")
    assert np.array_equal(idx[order], oidx) and np.array_equal(bits(xyz[order]), bits(oxyz))


def test_raw_u16_depth_and_subsampling(pkg, oracle, synth):
    """Depth pre-processing fused into the upload (examples/label_instance_rgbd.cpp:89-100)."""
    dims, vs = (48, 48, 48), 0.04
    origin, frames = synth.s_real(dims, vs, n_frames=2, seed=71)
    scale = np.float32(1.0) / np.float32(5000.0)
    raw = [np.clip(np.round(f[0] * 5000.0), 0, 65535).astype(np.uint16) for f in frames]
    sub = []
    for r in raw:
        d = r.astype(np.float32) * scale
        m = np.zeros_like(d)
        m[::4, ::3] = d[::4, ::3]
        sub.append(m)
    fr = [(s, f[1], f[2]) for s, f in zip(sub, frames)]
    p, t, w, c, _ = oracle_run(oracle, dims, vs, origin, fr, SEM_FUSIONPY, True)
    cfg = pkg.make_config(dims, vs, origin)
    with pkg.Volume(cfg) as vol:
        for r, f in zip(raw, frames):
            vol.integrate_raw(r, scale, f[1], f[2], subsample=(3, 4))
        gt, gw, gc = vol.download()
    assert_volume_equal(gt, gw, gc, t, w, c)


def test_lidar_range_input_mode(pkg, oracle, synth):
    """TSDF_DEPTH_RANGE_*: the frame holds ranges along the ray; the prep kernel converts them with
    the reference's formula (src/DoN.cpp:86-100, src/Object.cpp:611-619) bit for bit."""
    dims, vs = (64, 48, 64), 0.04
    origin, frames = synth.s_real(dims, vs, n_frames=3, seed=55)
    K = synth.TUM3_K
    H, W = synth.TUM3_HW
    cc, rr = np.meshgrid(np.arange(W, dtype=np.float64), np.arange(H, dtype=np.float64))
    rim = np.sqrt(((cc - K[2]) / K[0]) ** 2 + ((rr - K[5]) / K[4]) ** 2 + 1.0)
    ranges = [(d.astype(np.float64) * rim).astype(np.float32) for d, _, _ in frames]       # metres
    ranges_u16 = [np.clip(np.round(r.astype(np.float64) * 500.0), 0, 65535).astype(np.uint16) for r in ranges]
    for fmt, imgs, scale in ((pkg.capi.DEPTH_RANGE_F32, ranges, 1.0), (pkg.capi.DEPTH_RANGE_U16, ranges_u16, 1.0 / 500.0)):
        conv = [oracle.range_to_depth(np.asarray(im, np.float32), K, np.float32(scale)) for im in imgs]
        oframes = [(c, f[1], f[2]) for c, f in zip(conv, frames)]
        p, t, w, c, ow = oracle_run(oracle, dims, vs, origin, oframes, SEM_FUSIONPY, True)
        cfg = pkg.make_config(dims, vs, origin, semantics=SEM_FUSIONPY, with_color=True)
        with pkg.Volume(cfg) as vol:
            for im, (_, rgb, T) in zip(imgs, frames):
                vol.integrate_raw(im, scale, rgb, T, fmt=fmt)
            gt, gw, gc = vol.download()
        assert sum(ow) > 0
        assert_volume_equal(gt, gw, gc, t, w, c)


def test_instance_origin_on_device(pkg, oracle, synth):
    import ctypes as C
    _, frames = synth.s_real((32, 32, 32), 0.05, n_frames=1, seed=81)
    d = frames[0][0]
    masks, _ = synth.instance_masks(3)
    for m in (None, masks[0], masks[2]):
        out = np.zeros(3, np.float32)
        rc = pkg.load().tsdf_instance_origin(0, d.ctypes.data, None if m is None else m.ctypes.data,
                                             d.shape[0], d.shape[1], synth.TUM3_K.ctypes.data, out.ctypes.data)
        assert rc == 0
        assert np.array_equal(bits(out), bits(oracle.instance_origin(d, m, synth.TUM3_K)))


def test_reference_class_mirrors(pkg, oracle, synth, tmp_path):
    """TSDF(h,w,id,base2world,origin).Integrate(depth, cam2world) and
    TSDFVolume(vol_bnds, voxel_size).integrate(color, depth, K, pose, obs_weight)."""
    dims, vs = (200, 200, 200), 0.004      # the reference's hard-coded grid (include/tsdf.hpp:63-67)
    L = 200 * vs
    scene = synth.Scene().sphere((0, 0, 1.2), 0.25).plane((0, 0, 1), 1.6)
    rng = np.random.default_rng(5)
    poses = [synth.pose(synth.rot_xyz(0.01 * i, -0.02 * i, 0), (0.01 * i, 0, 0)) for i in range(3)]
    frames = [synth.render(scene, synth.TUM3_K, synth.TUM3_HW, T, rng) + (T,) for T in poses]
    origin = np.array([-L / 2, -L / 2, 0.8], np.float32)
    base = poses[0]
    p, t, w, _, _ = oracle_run(oracle, dims, vs, origin, frames, SEM_TSDFCU, False, base=base)
    obj = pkg.TSDF(480, 640, 3, base.reshape(16), origin)
    for d, _, T in frames:
        obj.Integrate(d, T.reshape(16))
    assert (obj.voxel_grid_weight == 0).all()          # mirrors are stale until readback (src/tsdf.cu:101-104)
    obj.close(str(tmp_path))
    assert np.array_equal(obj.voxel_grid_weight, w) and np.array_equal(bits(obj.voxel_grid_TSDF), bits(t))
    assert oracle.save_ply(str(tmp_path / "o.ply"), p, t, w)
    assert open(tmp_path / "o.ply", "rb").read() == open(tmp_path / "tsdf3.ply", "rb").read()
    assert oracle.save_bin(str(tmp_path / "o.bin"), p, t)
    assert open(tmp_path / "o.bin", "rb").read() == open(tmp_path / "tsdf3.bin", "rb").read()

    bnds = np.array([[-0.6, 0.6], [-0.5, 0.5], [0.8, 1.8]])
    tv = pkg.TSDFVolume(bnds, 0.01)
    d3 = tuple(tv._vol_dim)
    assert d3 == tuple(np.ceil((bnds[:, 1] - bnds[:, 0]) / 0.01).astype(int))
    for d, rgb, T in frames:
        tv.integrate(rgb, d, synth.TUM3_K.reshape(3, 3), T, obs_weight=1.0)
    tsdf_vol, color_vol = tv.get_volume()
    p2, t2, w2, c2, _ = oracle_run(oracle, d3, 0.01, bnds[:, 0], frames, SEM_FUSIONPY, True,
                                   trunc=np.float32(0.01 * 5))
    assert np.array_equal(bits(tsdf_vol.reshape(-1)), bits(t2)) and np.array_equal(color_vol.reshape(-1), c2)
    xyz, rgb = tv.get_point_cloud()
    oxyz, orgb, _ = oracle.extract(p2, t2, w2, c2)
    assert np.array_equal(bits(xyz), bits(oxyz)) and np.array_equal(rgb, orgb)
    tv.close()


@pytest.mark.parametrize("dims,vs", [((32, 32, 32), 0.02), ((128, 24, 40), 0.006), ((136, 16, 24), 0.006)])
def test_batched_instances_match_per_instance_oracle(pkg, oracle, synth, dims, vs):
    """Config 5 in small: N instance volumes, one launch per frame, masked depth semantics of
    src/Engine.cpp:192-193 and per-instance base/origin of src/Object.cpp:23-49.  Fast axis 32: the
    plain sparse body; 128 / 136 (ragged second window): the quarter work-list kernel
    (k_integrate_q_batch), what BASELINE config 5's 128^3 instance volumes run."""
    import ctypes as C
    n_inst = 6
    _, frames = synth.s_real((64, 64, 64), 0.08, n_frames=3, seed=91)
    masks, _ = synth.instance_masks(n_inst, seed=92)
    base = frames[0][2]
    d0 = frames[0][0]
    cfgs = (pkg.capi.Config * n_inst)()
    origins = []
    for i in range(n_inst):
        og = oracle.instance_origin(d0, masks[i], synth.TUM3_K)
        origins.append(og)
        cfgs[i] = pkg.make_config(dims, vs, og, semantics=SEM_TSDFCU, with_color=False, base2world=base,
                                  max_depth=30.0, vol_id=i)
    lib = pkg.load()
    hb = C.c_void_p()
    pkg.capi.check(lib.tsdf_batch_create(cfgs, n_inst, C.byref(hb)))
    try:
        active_sets = [list(range(n_inst)), [0, 2, 5], [4, 1]]
        for (d, rgb, T), act in zip(frames, active_sets):
            a = np.array(act, np.int32)
            m = np.ascontiguousarray(masks[act])
            Tc = np.ascontiguousarray(T, np.float32)
            pkg.capi.check(lib.tsdf_batch_integrate_async(hb, d.ctypes.data, None, a.ctypes.data, len(act),
                                                          m.ctypes.data, None, Tc.ctypes.data, np.float32(1.0)))
            pkg.capi.check(lib.tsdf_batch_sync(hb))
        for i in range(n_inst):
            p = oracle.params(dims, vs, np.float32(vs) * np.float32(5), origins[i], synth.TUM3_HW, synth.TUM3_K,
                              30.0, SEM_TSDFCU)
            t, w, _ = oracle.new_volume(p)
            for (d, rgb, T), act in zip(frames, active_sets):
                if i in act:
                    md = d * (masks[i] >= 128)
                    oracle.integrate(p, oracle.compose(base, T), md.astype(np.float32), None, 1.0, t, w, None)
            gt, gw = np.zeros_like(t), np.zeros_like(w)
            pkg.capi.check(lib.tsdf_batch_download(hb, i, gt.ctypes.data, gw.ctypes.data, None))
            assert np.array_equal(gw, w) and np.array_equal(bits(gt), bits(t)), f"instance {i}"
    finally:
        lib.tsdf_batch_destroy(hb)


def test_batched_instances_cropped_masks_equal_full_masks(pkg, oracle, synth):
    """tsdf_batch_integrate_boxes_async (masks cropped to their Mask-RCNN boxes, reference
    src/Engine.cpp:127,197) gives the volumes of the full-frame mask call, colour included, and
    matches the per-instance oracle."""
    n_inst, dims, vs = 5, (32, 32, 32), 0.02
    _, frames = synth.s_real((64, 64, 64), 0.08, n_frames=3, seed=93)
    masks, boxes = synth.instance_masks(n_inst, seed=94)
    origins = [oracle.instance_origin(frames[0][0], masks[i], synth.TUM3_K) for i in range(n_inst)]
    cfgs = [pkg.make_config(dims, vs, origins[i], semantics=SEM_FUSIONPY, with_color=True, vol_id=i)
            for i in range(n_inst)]
    active_sets = [[0, 1, 2, 3, 4], [3, 1], [4, 0, 2]]
    with pkg.Batch(cfgs) as full, pkg.Batch(cfgs) as crop:
        for (d, rgb, T), act in zip(frames, active_sets):
            full.integrate(d, rgb, act, np.ascontiguousarray(masks[act]), T, sync=True)
            bx, crops = pkg.Batch.crop_masks(masks[act], boxes[act])
            assert crops.size < 0.2 * masks[act].size
            crop.integrate_boxes(d, rgb, act, bx, crops, T, sync=True)
        assert full.stats().voxels_written == crop.stats().voxels_written
        for i in range(n_inst):
            ft, fw, fc = full.download(i)
            ct, cw, cc = crop.download(i)
            assert np.array_equal(bits(ft), bits(ct)) and np.array_equal(fw, cw) and np.array_equal(fc, cc)
            p = oracle.params(dims, vs, np.float32(vs) * np.float32(5), origins[i], synth.TUM3_HW, synth.TUM3_K,
                              np.inf, SEM_FUSIONPY)
            t, w, c = oracle.new_volume(p, True)
            for (d, rgb, T), act in zip(frames, active_sets):
                if i in act:
                    md = (d * (masks[i] >= 128)).astype(np.float32)
                    oracle.integrate(p, T.reshape(16), md, oracle.pack_color(rgb), 1.0, t, w, c)
            assert_volume_equal(ct, cw, cc, t, w, c)
            # per-object mesh (tsdf_batch_extract_mesh) of the same instance volume
            gm, om = crop.extract_mesh(i), oracle.mesh(p, t, c)
            assert np.array_equal(gm[1], om[1]) and np.array_equal(bits(gm[0]), bits(om[0]))
            assert np.array_equal(bits(gm[2]), bits(om[2])) and np.array_equal(gm[3], om[3])


def test_full_size_properties_c1(pkg, synth):
    """BASELINE config 1 size (256^3, 2 cm) through size-independent properties: S-full writes
    every voxel every frame (N_written == N_vox, weights == frame count everywhere, tsdf stays
    exactly 1 in free space), weights are linear in obs_weight, a reset + replay is
    deterministic, and extraction returns exactly the predicate's index set, ascending."""
    dims, vs = (256, 256, 256), 0.02
    origin, frames = synth.s_full(dims, vs, n_frames=3)
    cfg = pkg.make_config(dims, vs, origin, semantics=SEM_FUSIONPY, with_color=True)
    with pkg.Volume(cfg) as vol:
        for d, rgb, T in frames:
            vol.integrate(d, rgb, T, obs_weight=1.0)
            assert vol.stats().voxels_written == vol.n_local
        t, w, c = vol.download()
        assert (w == 3.0).all() and (t == 1.0).all()
        assert (c >= 0).all() and (c < 2 ** 24).all() and (c == np.floor(c)).all()
        d, rgb, T = frames[0]
        vol.integrate(d, rgb, T, obs_weight=0.5)
        assert (vol.download()[1] == 3.5).all()
        # realistic scene at the same size: extraction == numpy selection, strictly ascending
        vol.reset()
        o2, fr2 = synth.s_real(dims, vs, n_frames=3, seed=101)
    cfg = pkg.make_config(dims, vs, o2, semantics=SEM_FUSIONPY, with_color=True)
    with pkg.Volume(cfg) as vol:
        for d, rgb, T in fr2:
            vol.integrate(d, rgb, T, sync=False)
        t, w, c = vol.download()
        for pred, tt, wt in ((PRED_ABS_LT, 0.5, 1.5), (PRED_NONZERO, 0.0, 0.9)):
            xyz, rgb, idx = vol.extract_points(pred, tt, wt)
            sel = ((np.abs(t) < np.float32(tt)) if pred == PRED_ABS_LT else (np.abs(t) != 0)) & (w > np.float32(wt))
            assert len(idx) > 0 and (np.diff(idx) > 0).all()
            assert np.array_equal(idx, np.flatnonzero(sel))
            cc = c[idx].astype(np.int64)
            assert np.array_equal(rgb, np.stack([cc & 255, (cc >> 8) & 255, cc >> 16], 1).astype(np.uint8))


def test_full_size_properties_c2(pkg, synth, dev_lib):
    """BASELINE config 2 size (512^3, 1 cm, fr3_office keyframe poses) through size-independent
    properties: k frames fused into one pass == k single passes (bit for bit, all three arrays),
    the segment cull never changes a result, and N_written adds up."""
    dims, vs = (512, 512, 512), 0.01
    poses = synth.load_pose_fixture()[::24]
    origin, frames = synth.s_real(dims, vs, n_frames=8, poses=poses)
    cfg = pkg.make_config(dims, vs, origin, semantics=SEM_FUSIONPY, with_color=True)
    with pkg.Volume(cfg) as vol:
        total = 0
        for d, rgb, T in frames:
            vol.integrate(d, rgb, T)
            total += vol.stats().voxels_written
        seq = vol.download()
        n_seq = vol.extract_count(PRED_ABS_LT, 0.5, 1.5)
        vol.reset()
        vol.integrate_multi([f[0] for f in frames], [f[1] for f in frames], [f[2] for f in frames])
        assert vol.stats().voxels_written == total
        fused = vol.download()
        vol.reset()
        vol.set_variant(64)                       # cull off
        for d, rgb, T in frames:
            vol.integrate(d, rgb, T, sync=False)
        plain = vol.download()
        assert vol.extract_count(PRED_ABS_LT, 0.5, 1.5) == n_seq
    assert 0 < total < 8 * np.prod(dims)
    for a, b, c in zip(seq, fused, plain):
        assert np.array_equal(bits(a), bits(b)) and np.array_equal(bits(a), bits(c))


def test_full_size_properties_c4(pkg, synth):
    """BASELINE config 4 size (1024^3, the bench workload) without downloading 13 GB: S-full
    writes every voxel every frame, so after k frames the extraction predicate
    `weight > k - 0.5` selects all N voxels and `weight > k + 0.5` none; an 8-way z-slab of the
    same grid writes exactly its share."""
    dims, vs = (1024, 1024, 1024), 0.002
    origin, frames = synth.s_full(dims, vs, n_frames=3)
    n = int(np.prod(dims))
    cfg = pkg.make_config(dims, vs, origin, semantics=SEM_FUSIONPY, with_color=True)
    with pkg.Volume(cfg) as vol:
        for k, (d, rgb, T) in enumerate(frames, 1):
            vol.integrate(d, rgb, T)
            assert vol.stats().voxels_written == n
            assert vol.extract_count(PRED_NONZERO, 0.0, k - 0.5) == n     # tsdf == 1 != 0 everywhere
            assert vol.extract_count(PRED_NONZERO, 0.0, k + 0.5) == 0
        assert vol.extract_count(PRED_ABS_LT, 1.0, 0.5) == 0              # |tsdf| < 1 nowhere: free space
    zr = pkg.slab_range(dims[2], 8, 5)
    cfg = pkg.make_config(dims, vs, origin, semantics=SEM_FUSIONPY, with_color=True, z_range=zr)
    with pkg.Volume(cfg) as slab:
        slab.integrate(*frames[0][:2], frames[0][2])
        assert slab.stats().voxels_written == dims[0] * dims[1] * (zr[1] - zr[0]) == slab.n_local


def test_against_reference_gpu_kernel(pkg, oracle, synth):
    """"Correctness is checked against the reference's own integration on identical synthetic
    inputs" (north_star), literally: the reference's __global__ GpuIntegrate lines compiled by nvcc
    for sm_100a with nvcc's defaults run beside this library on the reference's own 200^3 @ 4 mm
    grid.  nvcc contracts the reference's a*b+c into FMA, so a pixel or truncation
    decision can flip in the last ulp: the no-FMA source-order semantics is canonical here (and
    this library equals the oracle bit for bit); against the FMA build a vanishing fraction of
    voxels may pick the neighbouring pixel or flip the truncation test (different weight, or a
    different depth sample), everywhere else tsdf agrees to 1e-5."""
    from oracle.oracle import RefGpu, ref_available
    if not ref_available("gpu"):
        pytest.skip("oracle/_ref/libtsdf_ref_gpu.so not built (needs /root/reference at build time)")
    dims, vs = (200, 200, 200), 0.004
    L = 200 * vs
    scene = synth.Scene().sphere((0, 0, 1.2), 0.25).plane((0, 0, 1), 1.6)
    rng = np.random.default_rng(11)
    poses = [synth.pose(synth.rot_xyz(0.013 * i, -0.021 * i, 0.004 * i), (0.011 * i, -0.003 * i, 0.002 * i)) for i in range(4)]
    frames = [synth.render(scene, synth.TUM3_K, synth.TUM3_HW, T, rng, noise_sigma=0.002) + (T,) for T in poses]
    origin = np.array([-L / 2, -L / 2, 0.8], np.float32)
    base = poses[0]
    trunc = np.float32(vs) * np.float32(5)
    ref = RefGpu(dims, synth.TUM3_HW, synth.TUM3_K)
    cfg = pkg.make_config(dims, vs, origin, semantics=SEM_TSDFCU, with_color=False, base2world=base)
    with pkg.Volume(cfg) as vol:
        for d, _, T in frames:
            ref.integrate(oracle.compose(base, T), d, origin, vs, trunc)
            vol.integrate(d, None, T)
        gt, gw, _ = vol.download()
    rt, rw = ref.download()
    ref.close()
    assert rw.max() == len(frames) and gw.max() == len(frames)
    differ = np.count_nonzero(rw != gw)
    assert differ <= 1e-4 * gw.size, f"{differ} of {gw.size} weights differ from the reference's FMA build"
    dt = np.abs(rt - gt)
    flipped = np.count_nonzero(dt > TOL)
    assert flipped <= 1e-3 * gw.size, f"{flipped} of {gw.size} tsdf values differ by more than {TOL}"
    print(f"reference GPU kernel (FMA build): {differ} / {gw.size} weights differ, {flipped} tsdf values differ by "
          f"> {TOL} (pixel / truncation flips), median |diff| {np.median(dt[gw > 0]):.2e}")


@pytest.mark.parametrize("sem", [SEM_TSDFCU, SEM_FUSIONPY])
def test_more_than_2_31_voxels(pkg, oracle, synth, sem, dev_lib):
    """Maximum sizes: a 1280 x 1280 x 1536 grid (2.52 G voxels > 2^31; 20-30 GB of HBM) — every
    linear index, byte offset and extraction index must be 64-bit.  The volume is too large to
    mirror on the host, so windows of the storage order (tsdf_download_range) are compared bit for
    bit with the oracle restricted to the same windows (oracle_integrate_range): the first and
    last 2^20 voxels, the window straddling voxel 2^31, and windows around written voxels found
    through the extraction's own int64 indices."""
    dims, vs = (1280, 1280, 1536), 0.004
    color = sem == SEM_FUSIONPY
    n = int(np.prod(dims))
    assert n > 2 ** 31
    # two views from near the centre, towards the +z and the +x wall of the room: the surfaces with
    # the largest linear indices in either storage order
    poses = [synth.look_at((0.1, 0.0, -1.0), (0.0, 0.2, 3.0)), synth.look_at((-1.0, 0.1, 0.0), (3.0, 0.2, 0.1))]
    origin, frames = synth.s_real(dims, vs, n_frames=2, seed=11, poses=poses)
    cfg = pkg.make_config(dims, vs, origin, semantics=sem, with_color=color, max_depth=6.0)
    p = oracle.params(dims, vs, np.float32(vs) * np.float32(5), origin, synth.TUM3_HW, synth.TUM3_K, 6.0, sem)
    W = 1 << 20
    with pkg.Volume(cfg) as vol:
        assert vol.n_local == n
        written = 0
        for d, rgb, T in frames:
            vol.integrate(d, rgb if color else None, T)
            written += vol.stats().voxels_written
        assert written > 10 ** 6
        xyz, _, idx = vol.extract_points(PRED_ABS_LT, 0.999, 0.5)
        assert len(idx) > 10 ** 5 and np.all(np.diff(idx) > 0) and idx[0] >= 0 and idx[-1] < n
        assert idx[-1] > 2 ** 31, "scene must reach past voxel 2^31"
        hi = int(idx[np.searchsorted(idx, 2 ** 31)])
        firsts = [0, 2 ** 31 - W // 2, n - W, max(0, int(idx[0]) - W // 2), min(n - W, hi - W // 2),
                  min(n - W, int(idx[-1]) - W // 2)]
        touched = 0
        for first in firsts:
            t, w, c = np.ones(W, np.float32), np.zeros(W, np.float32), (np.zeros(W, np.float32) if color else None)
            for d, rgb, T in frames:
                touched += oracle.integrate_range(p, T.reshape(16), d, oracle.pack_color(rgb) if color else None,
                                                  1.0, first, W, t, w, c)
            gt, gw, gc = vol.download_range(first, W)
            assert np.array_equal(gw, w) and np.array_equal(bits(gt), bits(t)), first
            if color:
                assert np.array_equal(gc, c), first
            # the extraction's indices inside the window are exactly the oracle's surface voxels
            sel = np.flatnonzero((np.abs(t) < np.float32(0.999)) & (w > 0.5)) + first
            a, b = np.searchsorted(idx, first), np.searchsorted(idx, first + W)
            assert np.array_equal(idx[a:b], sel), first
            if len(sel):   # coordinates of the last one (src/tsdf.cu:201-208 / the plugin's order)
                g = int(sel[-1])
                if sem == SEM_TSDFCU:
                    z, r = divmod(g, dims[0] * dims[1]); y, x = divmod(r, dims[0])
                else:
                    x, r = divmod(g, dims[1] * dims[2]); y, z = divmod(r, dims[2])
                o32, v32 = np.asarray(origin, np.float32), np.float32(vs)
                want = np.array([o32[0] + np.float32(x) * v32, o32[1] + np.float32(y) * v32, o32[2] + np.float32(z) * v32], np.float32)
                assert np.array_equal(bits(xyz[b - 1]), bits(want))
        assert touched > 10 ** 4
        with pytest.raises(Exception):
            vol.download_range(n - 10, 11)
        # mesh readback past voxel 2^31 (64-bit coordinate decode, tile offsets and record indices): the
        # 8-wide, 4-wide and scalar sweeps agree, every vertex lies on its grid edge inside the volume
        counts = []
        for variant in (0, 262144, 65536):
            vol.set_variant(variant)
            vol.upload()                                  # no arrays: invalidates the cached sweep
            counts.append(vol.mesh_count())
        vol.set_variant(0)
        assert counts[0] == counts[1] == counts[2] and counts[0][0] > 10 ** 5 and counts[0][1] > 10 ** 5
        mv, mf, mn, _ = vol.extract_mesh()
        assert mf.min() >= 0 and mf.max() == len(mv) - 1
        q = (mv.astype(np.float64) - np.asarray(origin, np.float32).astype(np.float64)) / float(np.float32(vs))
        assert q.min() > -1e-3 and np.all(q.max(0) < np.array(dims) - 1 + 1e-3)
        off_grid = np.abs(q - np.round(q)) > 2e-3         # exactly one coordinate of a vertex is fractional (or none)
        assert off_grid.sum(1).max() <= 1
        # the three vertices of a triangle sit on edges of ONE cell (a wrong 64-bit rank would point far away);
        # the edge audit itself (every edge shared by two triangles) runs at sizes the host can sort, tests/test_mesh.py
        tri = q[mf[:: max(1, len(mf) // 200000)]]
        assert (tri.max(1) - tri.min(1)).max() <= 1 + 4e-3
        assert np.all(np.floor(tri.min(1) + 2e-3) + 1 >= tri.max(1) - 2e-3)


def test_handles_driven_from_different_host_threads(pkg, oracle, synth):
    """Threading contract of the boundary: each handle is single-writer, different
    handles may be driven from different host threads (the reference's mask loop,
    src/Engine.cpp:172, is annotated `omp parallel for`).  Four threads, four volumes of
    different shapes and semantics, interleaved async launches (ctypes drops the GIL during the
    calls): every volume must equal its own oracle run, and the per-thread error slots must not
    leak into each other."""
    import threading
    cases = [((64, 48, 64), SEM_FUSIONPY, True, 21), ((50, 44, 38), SEM_TSDFCU, False, 22),
             ((200, 40, 36), SEM_FUSIONPY, False, 23), ((36, 40, 132), SEM_TSDFCU, False, 24)]
    jobs = []
    for dims, sem, color, seed in cases:
        origin, frames = synth.s_real(dims, 0.03, n_frames=6, seed=seed)
        p, t, w, c, written = oracle_run(oracle, dims, 0.03, origin, frames, sem, color)
        jobs.append((pkg.make_config(dims, 0.03, origin, semantics=sem, with_color=color), frames, color, (t, w, c), written))
    out, errs = [None] * len(jobs), []
    start = threading.Barrier(len(jobs))

    def run(k):
        cfg, frames, color, _, written = jobs[k]
        try:
            with pkg.Volume(cfg) as vol:
                start.wait()
                for rep in range(3):                      # same frames three times: reset in between
                    vol.reset()
                    for (d, rgb, T), ow in zip(frames, written):
                        vol.integrate(d, rgb if color else None, T, sync=False)
                    assert vol.stats().voxels_written == written[-1]
                    if k == 0 and rep == 1:               # a failing call on this thread only
                        with pytest.raises(Exception):
                            vol.download_range(vol.n_local, 1)
                        assert "range" in pkg.load().tsdf_last_error().decode()
                out[k] = vol.download()
        except BaseException as e:   # noqa: BLE001
            errs.append((k, repr(e)))

    threads = [threading.Thread(target=run, args=(k,)) for k in range(len(jobs))]
    for th in threads:
        th.start()
    for th in threads:
        th.join(timeout=240)
    assert not errs, errs
    for k, (_, _, _, (t, w, c), _) in enumerate(jobs):
        assert out[k] is not None
        assert_volume_equal(out[k][0], out[k][1], out[k][2], t, w, c)
