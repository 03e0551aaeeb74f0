"""How often could an image tile be STAGED in shared memory for the sparse integrate kernel?  (CPU only.)
For the C2 (S-real 512^3, fr3_office poses) and C3 (S-kitti 1024x256x1024) workloads: the pixel rectangle the
8 corners of every block tile project to (what the kernel's block-level hull test bounds), for the kernel's
tile shape (1 x 128 rows x 128 voxels along z) and for smaller hypothetical ones, against one TMA box of
64 x 32 pixels (32 KB of float4) and against an any-shape budget of 8192 pixels (128 KB).
usage: python tools/stage_fit.py   -> one JSON line per (workload, tile shape)"""
import importlib, json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
synth = importlib.import_module("semantic-slam_b200.synth")


def analyse(name, dims, vs, origin, frames, K, hw, tile):
    fx, cx, fy, cy = K[0], K[2], K[4], K[5]
    H, W = hw
    tx, ty, tz = tile
    res = []
    for d, _, T in frames:
        T = T.reshape(4, 4).astype(np.float64)
        R, t = T[:3, :3], T[:3, 3]
        dmax = d.max()
        nx, ny, nz = dims[0] // tx, dims[1] // ty, dims[2] // tz
        ix, iy, iz = np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij")
        big = np.full(ix.shape, 1e9)
        umin, vmin, zmin, umax, vmax = big.copy(), big.copy(), big.copy(), -big, -big
        for a in (0, 1):
            for b in (0, 1):
                for c in (0, 1):
                    p = np.stack([origin[0] + (ix * tx + a * (tx - 1)) * vs, origin[1] + (iy * ty + b * (ty - 1)) * vs,
                                  origin[2] + (iz * tz + c * (tz - 1)) * vs], -1)
                    pc = (p - t) @ R
                    z = pc[..., 2]
                    with np.errstate(divide="ignore", invalid="ignore"):
                        u, v = fx * pc[..., 0] / z + cx, fy * pc[..., 1] / z + cy
                    umin, umax = np.minimum(umin, u), np.maximum(umax, u)
                    vmin, vmax = np.minimum(vmin, v), np.maximum(vmax, v)
                    zmin = np.minimum(zmin, z)
        inview = (zmin > 0.05) & (umax >= 0) & (umin <= W - 1) & (vmax >= 0) & (vmin <= H - 1) & (zmin < dmax + 5 * vs)
        w = np.clip(umax, 0, W - 1) - np.clip(umin, 0, W - 1) + 1
        h = np.clip(vmax, 0, H - 1) - np.clip(vmin, 0, H - 1) + 1
        res.append((inview.sum(), (inview & (w <= 64) & (h <= 32)).sum(), (inview & (w * h <= 8192)).sum(),
                    np.median(w[inview]), np.median(h[inview])))
    r = np.array(res, float)
    print(json.dumps({"workload": name, "tile_rows_x_rows_x_voxels": list(tile), "blocks_in_view_per_frame": round(r[:, 0].mean()),
                      "fit_one_64x32_box_pct": round(100 * r[:, 1].sum() / r[:, 0].sum(), 1),
                      "fit_8192_px_any_shape_pct": round(100 * r[:, 2].sum() / r[:, 0].sum(), 1),
                      "median_rect_px": [round(r[:, 3].mean()), round(r[:, 4].mean())]}))


dims, vs = (512, 512, 512), 0.01
o2, fr2 = synth.s_real(dims, vs, n_frames=12, poses=synth.load_pose_fixture()[::16])
for tile in ((1, 128, 128), (8, 16, 128), (4, 8, 128), (2, 4, 128)):
    analyse("C2", dims, vs, o2, fr2, synth.TUM3_K, synth.TUM3_HW, tile)
dims, vs = (1024, 256, 1024), 0.1
o3, fr3 = synth.s_kitti(n_frames=4, voxel_size=vs, dims=dims)
for tile in ((1, 128, 128), (8, 16, 128), (4, 8, 128)):
    analyse("C3", dims, vs, o3, fr3, synth.KITTI_K, synth.KITTI_HW, tile)
