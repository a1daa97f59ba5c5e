"""Parity at the benchmark's own scale (BASELINE config 3): the 8192x512x8192 world and the 3840x2160
frame the headline number is measured on.  What this scale exercises and the small worlds do not: the 2^26-entry record
grid, the 1.3 GB brick pool, 31-bit padded cell indices, rays that run into the kMaxSteps cap of
VolumeRaytracer.cuh:235 on the grazing camera, corner-to-corner traversals, and the build's early-out far from the
origin of the noise field.

The oracle cannot build this world (2^35 voxels of 32-octave noise), so the checks are:
  * >= 4096 sampled bricks (incl. the far corner) re-derived voxel by voxel with the oracle's full 32-octave noise:
    mask, bounds and both top-level bits equal (VoxelWorldBuilder.cu:10-35, VolumeRaytracer.cuh:379-516);
  * sampled rows of the 4K frame for both config-3 cameras and a corner-to-corner view, marched by the oracle on the
    downloaded brickmap: step plane bit-exact, written pixels identical, RGB within 1/255
    (VolumeRaytracer.cu:354-525, Renderer.cu:90-276).
"""
import numpy as np
import pytest

import oracle_lib as ol

pytestmark = pytest.mark.gpu

DIMS, FACTOR = (8192, 512, 8192), 8
W, H = 3840, 2160
# (a) overview, (b) grazing; plus a view from one corner along the diagonal, whose traversals can
# reach the step cap before they reach the far border (1024 + 64 + 1024 coarse cells)
CAMERAS = {
    "overview": dict(xz=(4096, 7936), lift=128, euler=(-0.5, 0.0, 0.0)),
    "grazing": dict(xz=(4096, 4096), lift=0, euler=(-0.05, 0.8, 0.0)),
    "diagonal": dict(xz=(8180, 8180), lift=0, euler=(-0.02, 0.7853982, 0.0)),
}


@pytest.fixture(scope="module")
def w2(vx):
    w = vx.World.from_noise(*DIMS, FACTOR)
    yield w
    w.close()


@pytest.fixture(scope="module")
def w2_oracle(w2):
    top, bw, b6 = w2.download()
    return ol.OracleWorld.from_brickmap(top, bw, b6, *DIMS, FACTOR)


def test_sampled_bricks_equal_full_octave_noise(vx, w2):
    rng = np.random.default_rng(20262)
    cx, cy, cz = (d // FACTOR for d in DIMS)
    n = 4096
    c = np.stack([rng.integers(0, cx, n), rng.integers(0, cy, n), rng.integers(0, cz, n)], 1).astype(np.uint32)
    c[:512, 0] = rng.integers(1000, cx, 512); c[:512, 2] = rng.integers(1000, cz, 512)   # far corner: x, z > 8000
    c[512:520] = [[0, 0, 0], [cx - 1, cy - 1, cz - 1], [cx - 1, 0, 0], [0, cy - 1, 0], [0, 0, cz - 1], [cx - 1, 0, cz - 1],
                  [cx - 1, cy - 1, 0], [0, cy - 1, cz - 1]]
    # most random bricks are all-solid or all-empty: put half of the sample on the terrain surface (found with vertical rays)
    rt = vx.VoxelRaytracer3D(w2, 4096)
    k = 2048
    px = rng.integers(0, DIMS[0], k) + 0.5
    pz = rng.integers(0, DIMS[2], k) + 0.5
    o = np.stack([px, np.full(k, DIMS[1] + 10.0), pz], 1).astype(np.float32)
    r = rt.raytrace(o, np.tile(np.array([[0, -1, 0]], np.float32), (k, 1)))
    rt.close()
    surf = np.nonzero(r["hit"])[0]
    assert surf.size > k // 2
    c[1024:1024 + surf.size] = (r["voxel"][surf] // FACTOR).astype(np.uint32)
    occ, bw, b6 = w2.download_bricks(c)
    assert np.array_equal(occ & 1, occ >> 1), "traversal-grid and reference-indexed top-level bits disagree"
    partial = 0
    for i in range(n):
        bits = ol.populate_dense(FACTOR, FACTOR, FACTOR, int(c[i, 0]) * FACTOR, int(c[i, 1]) * FACTOR, int(c[i, 2]) * FACTOR, nthreads=1)
        assert np.array_equal(bits, bw[i]), "brick %s: mask differs from the 32-octave noise" % c[i]
        v = np.unpackbits(bits.view(np.uint8), bitorder="little").reshape(8, 8, 8)  # [z][y][x]
        assert bool(occ[i] & 1) == bool(v.any())
        if v.any():
            zs, ys, xs = np.nonzero(v)
            want = [xs.min(), ys.min(), zs.min(), xs.max(), ys.max(), zs.max()]
            partial += int(not v.all())
        else:
            want = [0, 0, 0, -1, -1, -1]
        assert np.array_equal(b6[i], np.array(want, np.float32)), "brick %s: bounds differ" % c[i]
    assert partial > 500, "the sample must contain surface bricks (got %d)" % partial


def camera_position(vx, world, cam):
    """Camera height rule of SURVEY.md §8d."""
    rt = vx.VoxelRaytracer3D(world, 16)
    x, z = cam["xz"]
    r = rt.raytrace(np.array([[x + 0.5, DIMS[1] + 10.0, z + 0.5]], np.float32), np.array([[0.0, -1.0, 0.0]], np.float32))
    rt.close()
    top = int(r["voxel"][0, 1]) if r["hit"][0] else -1
    return (float(x), float((top + 16 if top < DIMS[1] - 1 else DIMS[1] + 32) + cam["lift"]), float(z))


@pytest.mark.parametrize("name", list(CAMERAS))
def test_4k_frame_rows_match_oracle(vx, w2, w2_oracle, name):
    import torch
    cam = CAMERAS[name]
    pos = camera_position(vx, w2, cam)
    tan_half = vx.device_tanf(np.float32(np.float32(90.0 * 3.1415 / 180.0) / np.float32(2.0)))
    vcam = vx.make_camera(pos, cam["euler"], 90.0)
    ocam = ol.make_camera(pos, cam["euler"], 90.0, tan_half=tan_half)
    rt = vx.VoxelRaytracer3D(w2, 1)
    frame = torch.full((H, W, 4), 7, dtype=torch.uint8, device="cuda")
    steps = torch.full((H, W), -1, dtype=torch.int32, device="cuda")
    rt.render_device(vcam, vx.default_env(), vx.RenderOpts(0, 0, 0, 1, 1, vx.NORM_HOST, 0, 0, 0, 0, None), frame.data_ptr(), W, H,
                     d_steps=steps.data_ptr())
    torch.cuda.synchronize()
    got, gsteps = frame.cpu().numpy(), steps.cpu().numpy()
    rt.close()
    stride, first = 48, 17
    ref = np.full((H, W, 4), 7, np.uint8)
    rsteps = np.full((H, W), -1, np.int32)
    w2_oracle.render_sample_planes(ocam, ol.default_env(), ol.RenderOpts(0, 0, 0, 1, 1, ol.NORM_HOST), W, H, first, H, stride, ref, rsteps)
    rows = np.arange(first, H, stride)
    assert np.array_equal(gsteps[rows], rsteps[rows]), "%s: step planes differ at %d pixels" % (name, int((gsteps[rows] != rsteps[rows]).sum()))
    diff = np.abs(got[rows].astype(np.int16) - ref[rows].astype(np.int16))
    assert diff.max() <= 1, "%s: max channel difference %d" % (name, diff.max())
    print("%s: %d rows, max steps %d, mean %.1f" % (name, rows.size, int(gsteps[rows].max()), float(gsteps[rows].mean())))
