"""TEST / MEASUREMENT INFRASTRUCTURE ONLY — never imported by the product path.

float64 numpy restatement of the CPU mode of the third-party volume class the reference's plugin
drives (`fusion.TSDFVolume`, module andyzeng/tsdf-fusion-python, NOT vendored and without a
pinned version: only `sys.path.append` in reference src/TSDFfusion.py.in:14 and the calls at
:21-29 (constructor from bounds) and :43 (integrate(color, depth, K, pose, obs_weight=1.)).
PARITY UNPINNED: restated from the published algorithm, upstream source unavailable offline.

It is the "reference CPU TSDFfusion path" of SURVEY.md §8(d): one vectorised pass over all
voxels per frame, voxel coordinates re-materialised every frame, arithmetic in float64 with the
volumes stored as float32.  tests/test_oracle.py checks it against oracle/tsdf_oracle.c
(SEM_FUSIONPY, fp32): the two differ only where float64 vs float32 rounding moves a projected
pixel across x.5 or a depth difference across the truncation boundary, a bounded fraction.
bench.py times it on C1 (256^3) as `cpu_baseline.numpy_float64_c1`.
"""
import numpy as np


class TSDFVolumeF64:
    """Same constructor rule as reference src/TSDFfusion.py.in:21-29 relies on:
    dims = ceil((max - min) / voxel_size), origin = mins, trunc = 5 * voxel_size; [x][y][z] C order."""

    def __init__(self, vol_bnds, voxel_size):
        b = np.asarray(vol_bnds, np.float64).reshape(3, 2)
        self.voxel_size = float(voxel_size)
        self.trunc = 5.0 * self.voxel_size
        self.dims = np.ceil((b[:, 1] - b[:, 0]) / self.voxel_size).astype(np.int64)
        self.origin = b[:, 0].astype(np.float32).astype(np.float64)
        n = int(np.prod(self.dims))
        self.tsdf = np.ones(n, np.float32)
        self.weight = np.zeros(n, np.float32)
        self.color = np.zeros(n, np.float32)

    def integrate(self, color_im, depth_im, cam_intr, cam_pose, obs_weight=1.0):
        H, W = depth_im.shape
        K = np.asarray(cam_intr, np.float64).reshape(3, 3)
        packed = (color_im[..., 2].astype(np.float64) * 65536.0 + color_im[..., 1].astype(np.float64) * 256.0
                  + color_im[..., 0].astype(np.float64))
        # every voxel centre, in world then camera coordinates (3 x N float64, rebuilt per frame)
        gx, gy, gz = np.meshgrid(*[np.arange(int(d)) for d in self.dims], indexing="ij")
        vox = np.stack([gx.ravel(), gy.ravel(), gz.ravel()]).astype(np.float64)
        world = self.origin[:, None] + vox * self.voxel_size
        Tcw = np.linalg.inv(np.asarray(cam_pose, np.float64).reshape(4, 4))
        cam = Tcw[:3, :3] @ world + Tcw[:3, 3:4]
        z = cam[2]
        with np.errstate(divide="ignore", invalid="ignore"):
            px = np.round(K[0, 0] * (cam[0] / z) + K[0, 2])
            py = np.round(K[1, 1] * (cam[1] / z) + K[1, 2])
        ok = (z > 0) & (px >= 0) & (px < W) & (py >= 0) & (py < H)
        idx = np.flatnonzero(ok)
        pix = py[idx].astype(np.int64) * W + px[idx].astype(np.int64)
        d = depth_im.reshape(-1).astype(np.float64)[pix]
        diff = d - z[idx]
        keep = (d > 0) & (diff >= -self.trunc)
        idx, pix, diff = idx[keep], pix[keep], diff[keep]
        dist = np.minimum(1.0, diff / self.trunc)
        w0 = self.weight[idx].astype(np.float64)
        w1 = w0 + obs_weight
        self.tsdf[idx] = ((self.tsdf[idx].astype(np.float64) * w0 + dist) / w1).astype(np.float32)
        self.weight[idx] = w1.astype(np.float32)
        old = self.color[idx].astype(np.float64)
        new = packed.reshape(-1)[pix]
        out = np.zeros_like(old)
        for scale in (65536.0, 256.0, 1.0):       # B, G, R of the float-packed colour
            o = np.floor(old / scale)
            n = np.floor(new / scale)
            old -= o * scale
            new -= n * scale
            out += np.minimum(np.round((o * w0 + n) / w1), 255.0) * scale
        self.color[idx] = out.astype(np.float32)
        return int(idx.size)
