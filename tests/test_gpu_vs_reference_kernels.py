"""GPU parity against the REFERENCE'S OWN KERNELS (PopulateVoxels, dispatch, screenDispatch) built
unmodified for sm_100a by oracle/Makefile (oracle/_ref/ref_gpu_nofma: -fmad=false, the parity
authority of SURVEY.md §8c).  Bit-exact: world bits/bounds, hit positions, normals, step counts and
the shipped debug-view frame."""
import os
import subprocess

import numpy as np
import pytest

import oracle_lib as ol
from raysets import mixed_rays

pytestmark = pytest.mark.gpu

REF = os.path.join(ol.ORACLE_DIR, "_ref", "ref_gpu_nofma")
needs_ref = pytest.mark.skipif(not os.path.exists(REF), reason="oracle/_ref/ref_gpu_nofma not built")

DIMS, FACTOR = (256, 128, 256), 8


def run_ref(args, tmp_path):
    out = subprocess.run([REF] + [str(a) for a in args], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    return out.stdout


@pytest.fixture(scope="module")
def my_world(vx):
    w = vx.World.from_noise(*DIMS, FACTOR)
    yield w
    w.close()


@needs_ref
def test_world_equals_reference_populate_and_lowres(vx, my_world, tmp_path):
    """PopulateVoxels (device noise incl. saturating float->uint) + GenerateLowresVoxelBuffer."""
    path = tmp_path / "world.bin"
    run_ref(["build", *DIMS, FACTOR, path], tmp_path)
    raw = np.fromfile(path, np.uint8)
    count, wpb = (int(v) for v in raw[:16].view(np.uint64))
    off = 16
    ntop = (count + 31) // 32
    top = raw[off:off + ntop * 4].view(np.uint32); off += ntop * 4
    bw = raw[off:off + count * wpb * 4].view(np.uint32); off += count * wpb * 4
    b6 = raw[off:off + count * 24].view(np.float32).reshape(-1, 6)
    mt, mb, m6 = my_world.download()
    assert np.array_equal(mt, top)
    assert np.array_equal(mb, bw)
    assert np.array_equal(m6, b6)


@needs_ref
def test_trace_equals_reference_dispatch_kernel(vx, my_world, tmp_path):
    n = 200000
    o, d = mixed_rays(DIMS, n, seed=2024)
    rays = tmp_path / "rays.bin"
    with open(rays, "wb") as f:
        f.write(np.uint64(n).tobytes()); f.write(o.tobytes()); f.write(d.tobytes())
    out = tmp_path / "hits.bin"
    run_ref(["trace", *DIMS, FACTOR, rays, out], tmp_path)
    raw = np.fromfile(out, np.uint8)
    assert int(raw[:8].view(np.uint64)[0]) == n
    pos = raw[8:8 + 12 * n].view(np.float32).reshape(-1, 3)
    nrm = raw[8 + 12 * n:8 + 24 * n].view(np.float32).reshape(-1, 3)
    steps = raw[8 + 24 * n:8 + 28 * n].view(np.int32)
    rt = vx.VoxelRaytracer3D(my_world, n)
    got = rt.raytrace(o, d, norm_mode=vx.NORM_DEVICE)
    assert np.array_equal(got["steps"], steps)
    assert np.array_equal(got["pos"].view(np.uint32), pos.view(np.uint32))
    assert np.array_equal(got["normal"], nrm)
    assert np.array_equal(got["hit"].astype(bool), np.isfinite(pos).all(1))
    rt.close()


@needs_ref
@pytest.mark.parametrize("view", [((128, 150, 300), (-0.35, 0.0, 0.0)), ((40, 60, 40), (-0.2, 2.2, 0.0))])
def test_frame_equals_reference_screen_dispatch(vx, my_world, tmp_path, view):
    """As shipped: checkerboard on, DEBUG_VIEW quadrants, no shadow ray, crosshair; frame number 0."""
    pos, eul = view
    W, H = 320, 180
    out = tmp_path / "frame.bgra"
    run_ref(["render", *DIMS, FACTOR, W, H, *pos, *eul, 1, out], tmp_path)
    ref = np.fromfile(out, np.uint8).reshape(H, W, 4)
    rt = vx.VoxelRaytracer3D(my_world, 1)
    cam = vx.make_camera(pos, eul, 90.0)
    opts = vx.RenderOpts(1, 0, 1, 0, 1, vx.NORM_DEVICE, 0, 0, 0, 0, None)
    got = rt.render_host(cam, vx.default_env(), opts, W, H, bgra=np.full((H, W, 4), 255, np.uint8))
    assert (ref != 255).any(), "reference frame is empty"
    assert np.array_equal(got, ref), "differing pixels: %d" % int((got != ref).any(2).sum())
    rt.close()


REF_ORTHO = os.path.join(ol.ORACLE_DIR, "_ref", "ref_gpu_nofma_ortho")


@pytest.mark.skipif(not os.path.exists(REF_ORTHO), reason="oracle/_ref/ref_gpu_nofma_ortho not built")
@pytest.mark.parametrize("view", [((128, 150, 300), (-0.6, 0.0, 0.0), (100.0, 60.0)), ((40, 100, 40), (-0.9, 2.2, 0.0), (10.0, 10.0))])
def test_ortho_frame_equals_reference_ortho_build(vx, my_world, tmp_path, view):
    """The reference built with -DORTHO (getRayDirectionOrtho, Renderer.cu:61-70,202-204): as shipped otherwise
    (checkerboard, DEBUG_VIEW quadrants, crosshair); vx_camera.ortho_w / ortho_h select the same camera here."""
    pos, eul, size = view
    W, H = 320, 180
    out = tmp_path / "frame.bgra"
    res = subprocess.run([REF_ORTHO] + [str(a) for a in ["render", *DIMS, FACTOR, W, H, *pos, *eul, 1, out, *size]],
                         capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stderr[-2000:]
    ref = np.fromfile(out, np.uint8).reshape(H, W, 4)
    rt = vx.VoxelRaytracer3D(my_world, 1)
    cam = vx.make_camera(pos, eul, 90.0, ortho=size)
    opts = vx.RenderOpts(1, 0, 1, 0, 1, vx.NORM_DEVICE, 0, 0, 0, 0, None)
    got = rt.render_host(cam, vx.default_env(), opts, W, H, bgra=np.full((H, W, 4), 255, np.uint8))
    assert (ref != 255).any(), "reference frame is empty"
    assert np.array_equal(got, ref), "differing pixels: %d" % int((got != ref).any(2).sum())
    persp = rt.render_host(vx.make_camera(pos, eul, 90.0), vx.default_env(), opts, W, H, bgra=np.full((H, W, 4), 255, np.uint8))
    assert not np.array_equal(persp, ref)
    rt.close()
