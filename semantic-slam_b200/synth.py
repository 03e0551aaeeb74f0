"""Seeded synthetic RGB-D inputs for the parity tests and the benchmark.

Host-side numpy only: analytic scenes (planes, spheres, boxes) ray-cast through a pinhole
camera into fp32 depth (metres, camera z) + 8-bit colour.  No file or network access except
`load_pose_fixture`, which reads the committed tests/golden pose fixture.

Intrinsics come from the reference's configs: TUM3 (config/TUM3.yaml:8-11, also hard-coded
in include/tsdf.hpp:96) and KITTI00-02 (config/KITTI00-02.yaml:8-11,25).
"""
import os

import numpy as np

TUM3_K = np.array([535.4, 0, 320.1, 0, 539.2, 247.6, 0, 0, 1], np.float32)
TUM3_HW = (480, 640)
KITTI_K = np.array([718.856, 0, 607.1928, 0, 718.856, 185.2157, 0, 0, 1], np.float32)
KITTI_HW = (376, 1241)
KITTI_BF = 386.1448
SEED_BASE = 0xB200


def rot_xyz(rx, ry, rz):
    cx, sx, cy, sy, cz, sz = np.cos(rx), np.sin(rx), np.cos(ry), np.sin(ry), np.cos(rz), np.sin(rz)
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


def pose(R=None, t=(0, 0, 0)):
    T = np.eye(4)
    if R is not None:
        T[:3, :3] = R
    T[:3, 3] = t
    return T.astype(np.float32)


def look_at(eye, target, up=(0, -1, 0)):
    """Camera-to-world pose with +z towards target, +y down (image convention)."""
    eye, target, up = np.asarray(eye, float), np.asarray(target, float), np.asarray(up, float)
    z = target - eye
    z /= np.linalg.norm(z)
    x = np.cross(-up, z)  # y down => x = z cross (down) ... keep right-handed
    if np.linalg.norm(x) < 1e-9:
        x = np.array([1.0, 0, 0])
    x /= np.linalg.norm(x)
    y = np.cross(z, x)
    R = np.stack([x, y, z], axis=1)
    return pose(R, eye)


class Scene:
    """Union of analytic primitives; `cast` returns camera-z depth of the nearest hit."""

    def __init__(self):
        self.planes = []   # (n[3], d) : n.x = d
        self.spheres = []  # (c[3], r)
        self.boxes = []    # (lo[3], hi[3])

    def plane(self, n, d):
        self.planes.append((np.asarray(n, float), float(d)))
        return self

    def sphere(self, c, r):
        self.spheres.append((np.asarray(c, float), float(r)))
        return self

    def box(self, lo, hi):
        self.boxes.append((np.asarray(lo, float), np.asarray(hi, float)))
        return self

    def cast(self, K, hw, T_wc):
        H, W = hw
        K = np.asarray(K, float).reshape(3, 3)
        u, v = np.meshgrid(np.arange(W, dtype=float), np.arange(H, dtype=float))
        d_cam = np.stack([(u - K[0, 2]) / K[0, 0], (v - K[1, 2]) / K[1, 1], np.ones_like(u)], -1)
        T = np.asarray(T_wc, float).reshape(4, 4)
        D = d_cam @ T[:3, :3].T
        o = T[:3, 3]
        best = np.full((H, W), np.inf)
        for n, d in self.planes:
            den = D @ n
            with np.errstate(divide="ignore", invalid="ignore"):
                t = (d - o @ n) / den
            t[~np.isfinite(t) | (t <= 1e-6)] = np.inf
            best = np.minimum(best, t)
        for c, r in self.spheres:
            oc = o - c
            a = np.sum(D * D, -1)
            b = 2 * (D @ oc)
            cc = oc @ oc - r * r
            disc = b * b - 4 * a * cc
            with np.errstate(invalid="ignore"):
                t = (-b - np.sqrt(disc)) / (2 * a)
            t[~(disc >= 0) | ~(t > 1e-6)] = np.inf
            best = np.minimum(best, t)
        for lo, hi in self.boxes:
            with np.errstate(divide="ignore", invalid="ignore"):
                t1 = (lo - o) / D
                t2 = (hi - o) / D
            tn = np.nanmax(np.minimum(t1, t2), -1)
            tf = np.nanmin(np.maximum(t1, t2), -1)
            t = np.where((tn <= tf) & (tn > 1e-6), tn, np.inf)
            best = np.minimum(best, t)
        hit = o[None, None, :] + best[..., None] * D
        return best, hit


def texture(hit, rng, noise=24):
    """Procedural 8-bit colour from world hit points + seeded noise (exercises rounding)."""
    h = np.nan_to_num(hit, nan=0.0, posinf=0.0, neginf=0.0)
    base = np.stack([(np.floor(h[..., 0] * 4) + np.floor(h[..., 2] * 4)) % 2 * 120 + 60,
                     (np.sin(h[..., 1] * 7.0) * 0.5 + 0.5) * 200 + 20,
                     (np.cos(h[..., 0] * 3.0 + h[..., 2] * 5.0) * 0.5 + 0.5) * 220 + 10], -1)
    n = rng.integers(-noise, noise + 1, size=base.shape)
    return np.clip(base + n, 0, 255).astype(np.uint8)


def render(scene, K, hw, T_wc, rng, noise_sigma=0.0, hole_frac=0.0, max_range=np.inf):
    depth, hit = scene.cast(K, hw, T_wc)
    rgb = texture(hit, rng)
    depth = np.where(np.isfinite(depth) & (depth < max_range), depth, 0.0)
    if noise_sigma > 0:
        depth = np.where(depth > 0, depth + rng.normal(0, noise_sigma, depth.shape), 0.0)
    if hole_frac > 0:
        depth = np.where(rng.random(depth.shape) < hole_frac, 0.0, depth)
    return np.ascontiguousarray(depth, np.float32), np.ascontiguousarray(rgb, np.uint8)


# --------------------------------------------------------------------------------------
# S-full: every voxel of the grid is inside the frustum and in front of the wall, so
# N_written == N_vox each frame (the roofline case).
# --------------------------------------------------------------------------------------
def s_full(dims, voxel_size, n_frames=4, seed=SEED_BASE, K=TUM3_K, hw=TUM3_HW, jitter=True, sphere=False):
    """Returns (origin[3], frames) with frames = list of (depth, rgb, T_wc).
    The volume is an axis-aligned box in the world frame, camera looks along +z.
    sphere=True puts a sphere of radius 0.15 * min(L) (0.31 m at the 2.048 m bench volume) against the
    far face of the volume: a surface for the readback leg; its interior and (short) shadow are not
    written."""
    rng = np.random.default_rng(seed)
    Kf = np.asarray(K, float)
    H, W = hw
    L = np.array(dims, float) * voxel_size
    # tightest half-angles of the image (tan), shrunk 12% so that small pose jitter keeps coverage
    tx = min(Kf[2], W - 1 - Kf[2]) / Kf[0] * 0.88
    ty = min(Kf[5], H - 1 - Kf[5]) / Kf[4] * 0.88
    z_near = max(L[0] / 2 / tx, L[1] / 2 / ty) * 1.02 + 0.05
    origin = np.array([-L[0] / 2, -L[1] / 2, z_near], np.float32)
    wall_z = z_near + L[2] + 0.25
    scene = Scene().plane((0, 0, 1), wall_z)
    if sphere:
        r = 0.15 * L.min()
        scene.sphere((0.0, 0.0, z_near + L[2] - 0.5 * r), r)   # near the far face: a short shadow
    frames = []
    for i in range(n_frames):
        if jitter and i > 0:
            R = rot_xyz(*(rng.uniform(-0.004, 0.004, 3)))
            t = rng.uniform(-0.004, 0.004, 3)
        else:
            R, t = np.eye(3), np.zeros(3)
        T = pose(R, t)
        depth, rgb = render(scene, K, hw, T, rng, noise_sigma=0.002)
        frames.append((depth, rgb, T))
    return origin, frames


# --------------------------------------------------------------------------------------
# S-real: sphere on a floor inside a room, 5% holes, 5 mm noise; poses either orbiting or
# from the fr3_office keyframe fixture.
# --------------------------------------------------------------------------------------
def room_scene(center, half):
    c, h = np.asarray(center, float), float(half)
    s = Scene()
    s.plane((0, 1, 0), c[1] + 0.45 * h)          # floor (y is down in camera convention)
    s.plane((0, 0, 1), c[2] + 0.95 * h).plane((0, 0, 1), c[2] - 0.95 * h)
    s.plane((1, 0, 0), c[0] + 0.95 * h).plane((1, 0, 0), c[0] - 0.95 * h)
    s.plane((0, 1, 0), c[1] - 0.9 * h)           # ceiling
    s.sphere((c[0], c[1] + 0.45 * h - 0.6, c[2]), 0.6)
    s.box((c[0] + 0.8, c[1] + 0.45 * h - 0.7, c[2] - 0.5), (c[0] + 1.4, c[1] + 0.45 * h, c[2] + 0.3))
    return s


def s_real(dims, voxel_size, n_frames=8, seed=SEED_BASE + 1, K=TUM3_K, hw=TUM3_HW, poses=None,
           max_range=np.inf):
    """Volume centred on the scene; returns (origin, frames)."""
    rng = np.random.default_rng(seed)
    L = np.array(dims, float) * voxel_size
    center = np.zeros(3)
    origin = (center - L / 2).astype(np.float32)
    scene = room_scene(center, L.min() / 2)
    frames = []
    for i in range(n_frames):
        if poses is not None:
            T = np.asarray(poses[i % len(poses)], np.float32).reshape(4, 4)
        else:
            a = 2 * np.pi * i / max(n_frames, 1)
            r = 0.55 * L.min() / 2
            eye = center + np.array([r * np.cos(a), -0.15 * L[1] / 2 + 0.05 * np.sin(3 * a), r * np.sin(a)])
            T = look_at(eye, center + np.array([0, 0.1 * L[1] / 2, 0]))
        depth, rgb = render(scene, K, hw, T, rng, noise_sigma=0.005, hole_frac=0.05, max_range=max_range)
        frames.append((depth, rgb, T))
    return origin, frames


def load_pose_fixture(scale=1.0):
    """194 fr3_office keyframe poses (Twc, row-major 4x4) from the committed fixture
    tests/golden/fr3_office_Twc.npy (made by tests/golden/make_pose_fixture.py from the
    reference's result/rgbd/bundle.txt).  Translations are re-centred on their mean."""
    here = os.path.dirname(os.path.abspath(__file__))
    path = os.path.join(here, "..", "tests", "golden", "fr3_office_Twc.npy")
    T = np.load(path).astype(np.float64).reshape(-1, 4, 4)
    T[:, :3, 3] = (T[:, :3, 3] - T[:, :3, 3].mean(0)) * scale
    return T.astype(np.float32)


# --------------------------------------------------------------------------------------
# S-kitti (config 3): ground plane 1.65 m below the camera + box obstacles, stereo-like
# depth quantisation d = bf / round16(bf / z).
# --------------------------------------------------------------------------------------
def s_kitti(n_frames=4, seed=SEED_BASE + 2, voxel_size=0.1, dims=(1024, 256, 1024)):
    rng = np.random.default_rng(seed)
    L = np.array(dims, float) * voxel_size
    origin = np.array([-L[0] / 2, -L[1] + 3.0, -2.0], np.float32)  # y down: ground at +1.65
    scene = Scene().plane((0, 1, 0), 1.65)
    for k in range(24):
        x = rng.uniform(-20, 20)
        z = rng.uniform(8, 90)
        w, h, d = rng.uniform(1.5, 4), rng.uniform(1.2, 3.5), rng.uniform(1.5, 5)
        scene.box((x, 1.65 - h, z), (x + w, 1.65, z + d))
    frames = []
    for i in range(n_frames):
        T = pose(rot_xyz(0, 0.002 * i, 0), (0.02 * i, 0, 1.0 * i))
        depth, rgb = render(scene, KITTI_K, KITTI_HW, T, rng, max_range=50.0)
        with np.errstate(divide="ignore", invalid="ignore"):
            disp = np.round(KITTI_BF / depth * 16.0) / 16.0
            dq = np.where((depth > 0) & (disp > 0), KITTI_BF / disp, 0.0)
        frames.append((np.ascontiguousarray(dq, np.float32), rgb, T))
    return origin, frames


# --------------------------------------------------------------------------------------
# S-corridor: a long volume seen end-on (floor, side walls, an end wall and boxes all the way
# along z), so that EVERY z-slab of a sharded grid holds written voxels — the multi-GPU parity
# worker's scene (a room in the middle of a long grid leaves the outer slabs empty).
# --------------------------------------------------------------------------------------
def s_corridor(dims, voxel_size, n_frames=4, seed=SEED_BASE + 3, K=TUM3_K, hw=TUM3_HW):
    rng = np.random.default_rng(seed)
    L = np.array(dims, float) * voxel_size
    origin = np.array([-L[0] / 2, -L[1] / 2, 0.5], np.float32)
    scene = Scene().plane((0, 1, 0), 0.35 * L[1]).plane((1, 0, 0), 0.47 * L[0]).plane((1, 0, 0), -0.47 * L[0])
    scene.plane((0, 0, 1), 0.5 + 0.985 * L[2])
    n_box = max(6, int(L[2] / 1.2))
    for k in range(n_box):
        z = 0.5 + (k + rng.uniform(0.1, 0.6)) * L[2] / n_box
        x = rng.uniform(-0.4, 0.2) * L[0]
        w, h, d = rng.uniform(0.1, 0.3) * L[0], rng.uniform(0.15, 0.5) * L[1], rng.uniform(0.2, 0.6)
        scene.box((x, 0.35 * L[1] - h, z), (x + w, 0.35 * L[1], z + d))
    frames = []
    for i in range(n_frames):
        T = pose(rot_xyz(0.01 * i, 0.015 * (i - 1), 0.005 * i), (0.03 * i, -0.02 * i, -1.0 + 0.35 * i))
        depth, rgb = render(scene, K, hw, T, rng, noise_sigma=0.004, hole_frac=0.03)
        frames.append((depth, rgb, T))
    return origin, frames


# --------------------------------------------------------------------------------------
# S-inst (config 5): N seeded elliptical instance masks (0/255) with boxes.
# --------------------------------------------------------------------------------------
def instance_masks(n, hw=TUM3_HW, seed=SEED_BASE + 4):
    rng = np.random.default_rng(seed)
    H, W = hw
    masks = np.zeros((n, H, W), np.uint8)
    boxes = np.zeros((n, 4), np.int32)  # x, y, w, h
    v, u = np.mgrid[0:H, 0:W]
    for i in range(n):
        a, b = rng.integers(12, 70), rng.integers(12, 60)
        cx, cy = rng.integers(a, W - a), rng.integers(b, H - b)
        m = ((u - cx) / a) ** 2 + ((v - cy) / b) ** 2 <= 1.0
        masks[i][m] = 255
        boxes[i] = (cx - a, cy - b, 2 * a + 1, 2 * b + 1)
    return masks, boxes
