"""Seeded ray sets shared by the oracle tests, the golden-vector generator and the GPU parity tests."""
import numpy as np


def mixed_rays(dims, n, seed):
    """Random incoherent rays around/inside a world of `dims` voxels plus the special cases the
    reference code branches on: axis-parallel directions, integer starts, starts on
    the far faces, starts outside the grid."""
    rng = np.random.default_rng(seed)
    dims = np.asarray(dims, np.float32)
    o = (rng.uniform(-0.3, 1.3, (n, 3)) * dims).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    k = max(n // 100, 1)
    d[0 * k:2 * k, 0] = 0            # dx == 0
    d[2 * k:4 * k, 1] = 0            # dy == 0
    d[4 * k:5 * k, :2] = 0           # +-z axis rays
    d[5 * k:6 * k, 1:] = 0           # +-x axis rays
    o[6 * k:8 * k] = np.floor(o[6 * k:8 * k])          # integer starts
    o[8 * k:9 * k, 1] = dims[1]                          # exactly on the far y face
    o[9 * k:10 * k, 0] = dims[0]                         # exactly on the far x face
    o[10 * k:11 * k] = np.clip(o[10 * k:11 * k], 0, dims - 1)  # inside
    o[11 * k:12 * k, 2] = 0                              # on the near z face
    return o, d


def camera_rays(origin, fwd, up, right, W, H, tan_half):
    """Primary directions of Renderer.cu:44-59 in float32 (un-normalised is fine: Raytrace normalises)."""
    x = np.arange(W, dtype=np.float32)
    y = np.arange(H, dtype=np.float32)
    u = (x / np.float32(W)) * np.float32(2) - np.float32(1)
    v = (y / np.float32(H)) * np.float32(2) - np.float32(1)
    aspect = np.float32(W) / np.float32(H)
    a = u * (np.float32(tan_half) * aspect)
    b = v * np.float32(tan_half)
    fwd, up, right = (np.asarray(t, np.float32) for t in (fwd, up, right))
    d = fwd[None, None, :] + a[None, :, None] * right[None, None, :] + b[:, None, None] * up[None, None, :]
    d = d.reshape(-1, 3).astype(np.float32)
    o = np.broadcast_to(np.asarray(origin, np.float32), d.shape).copy()
    return o, d
