"""vx_render_static (include/vxrt.h): vx_render for callers that may not allocate — the frame's scratch (ray counters and the
queue of primary hits waiting for their sun ray) is a fixed file-scope device array of the library.  No counterpart in the
reference.  Collected last on purpose.

Checks, all through the C ABI: the pixels and the step plane equal vx_render's byte for byte (same kernels; bit-exact is the
tolerance of every frame test here) on two calls in a row and again after a different frame has used the array in between;
frames beyond the limit are refused with a clear message; the call leaves the device's free memory untouched and can be
captured into a CUDA graph (a cudaMalloc inside a capture fails)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

DIMS, ORIGIN = (128, 128, 128), (4032, 336, 7872)   # a hilly window of the noise field (the one smoke() renders)


@pytest.fixture(scope="module")
def scene(vx):
    world = vx.World.from_noise(*DIMS, 8, origin=ORIGIN)
    rt = vx.VoxelRaytracer3D(world, 16)
    tan_half = vx.device_tanf(np.float32(np.float32(90.0 * 3.1415 / 180.0) / np.float32(2.0)))  # Renderer.cu:50-52, once at set-up
    yield world, rt, tan_half
    rt.close()
    world.close()


def cameras(vx, tan_half):
    return [vx.make_camera((64.0, 120.0, 128.0), (-0.5, 0.2, 0.0), 90.0, tan_half=tan_half),   # smoke()'s view
            vx.make_camera((64.0, 120.0, 128.0), (-0.7, -0.3, 0.0), 90.0, tan_half=tan_half)]


def render_both(vx, scene, cam, opts, W, H, fill):
    import torch
    world, rt, _ = scene
    out = []
    for call in (lambda b, s: rt.render_device(cam, vx.default_env(), opts, b.data_ptr(), W, H, d_steps=s.data_ptr()),
                 lambda b, s: world.render_static(cam, vx.default_env(), opts, b.data_ptr(), W, H, d_steps=s.data_ptr())):
        frame = torch.full((H, W, 4), fill, dtype=torch.uint8, device="cuda")
        steps = torch.full((H, W), -1, dtype=torch.int32, device="cuda")
        call(frame, steps)
        torch.cuda.synchronize()
        out.append((frame.cpu().numpy(), steps.cpu().numpy()))
    return out


@pytest.mark.parametrize("W,H", [(256, 144), (640, 360)])
def test_static_frame_equals_vx_render_twice_in_a_row(vx, scene, W, H):
    opts = vx.RenderOpts(0, 0, 0, 1, 1, vx.NORM_DEVICE, 0, 0, 0, 0, None)
    cam_a, cam_b = cameras(vx, scene[2])
    for i, cam in enumerate([cam_a, cam_a, cam_b, cam_a]):   # same frame twice, another one, the first again
        (want, wsteps), (got, gsteps) = render_both(vx, scene, cam, opts, W, H, fill=7 + i)
        assert i > 0 or int((wsteps > 0).sum()) > W * H // 4, "the frame barely marches: not a test"
        assert np.array_equal(gsteps, wsteps), "call %d: step planes differ at %d pixels" % (i, int((gsteps != wsteps).sum()))
        assert np.array_equal(got, want), "call %d: %d bytes differ" % (i, int((got != want).sum()))
    shadowed = want[..., :3].astype(np.int32).sum(2)
    assert len(np.unique(shadowed)) > 16   # a shaded, shadowed frame, not a constant


def test_static_modes_match(vx, scene):
    """Checkerboard + debug view (the reference's shipped mode), host normalisation, a band share and a tile."""
    W, H = 320, 200
    cam = cameras(vx, scene[2])[0]
    for opts in (vx.RenderOpts(1, 3, 1, 0, 1, vx.NORM_DEVICE, 0, 0, 0, 0, None),
                 vx.RenderOpts(1, 4, 0, 1, 1, vx.NORM_HOST, 0, 0, 0, 0, None),
                 vx.RenderOpts(0, 0, 0, 1, 0, vx.NORM_DEVICE, 0, 0, 0, 0, None, 3, 1),
                 vx.RenderOpts(0, 0, 0, 1, 1, vx.NORM_DEVICE, 32, 16, 96, 64, None)):
        (want, wsteps), (got, gsteps) = render_both(vx, scene, cam, opts, W, H, fill=200)
        assert np.array_equal(got, want) and np.array_equal(gsteps, wsteps)


def test_static_limit_and_arguments(vx, scene):
    import torch
    world, _, tan_half = scene
    cam = cameras(vx, tan_half)[0]
    opts = vx.RenderOpts(0, 0, 0, 1, 1, vx.NORM_DEVICE, 0, 0, 0, 0, None)
    W, H = 1280, 720
    frame = torch.full((H, W, 4), 9, dtype=torch.uint8, device="cuda")
    with pytest.raises(vx.VxError) as e:
        world.render_static(cam, vx.default_env(), opts, frame.data_ptr(), W, H)
    assert "VX_STATIC_MAX_RAY_IDS" in str(e.value) and str(W // 16 * (H // 8) * 128) in str(e.value)
    torch.cuda.synchronize()
    assert int(frame.min()) == 9 and int(frame.max()) == 9, "a refused call must not touch the frame"
    # ... but the same frame fits in four band shares
    rt = scene[1]
    want = torch.zeros_like(frame)
    rt.render_device(cam, vx.default_env(), opts, want.data_ptr(), W, H)
    for k in range(4):
        world.render_static(cam, vx.default_env(), vx.RenderOpts(0, 0, 0, 1, 1, vx.NORM_DEVICE, 0, 0, 0, 0, None, 4, k), frame.data_ptr(), W, H)
    torch.cuda.synchronize()
    assert torch.equal(frame, want)
    no_tan = vx.make_camera((64.0, 120.0, 128.0), (-0.5, 0.2, 0.0), 90.0)
    with pytest.raises(vx.VxError, match="tan_half"):
        world.render_static(no_tan, vx.default_env(), opts, frame.data_ptr(), 64, 64)
    stats = torch.zeros(8, dtype=torch.int64, device="cuda")
    with pytest.raises(vx.VxError, match="d_stats"):
        world.render_static(cam, vx.default_env(), vx.RenderOpts(0, 0, 0, 1, 1, vx.NORM_DEVICE, 0, 0, 0, 0, stats.data_ptr()), frame.data_ptr(), 64, 64)


def test_static_call_allocates_nothing(vx, scene):
    """Free device memory is the same before and after, and the call can be captured into a CUDA graph and replayed
    (cudaMalloc inside a global-mode capture is an error, and would invalidate the capture)."""
    import torch
    world, rt, tan_half = scene
    W, H = 640, 360
    cam_a, cam_b = cameras(vx, tan_half)
    opts = vx.RenderOpts(0, 0, 0, 1, 1, vx.NORM_DEVICE, 0, 0, 0, 0, None)
    frame = torch.zeros((H, W, 4), dtype=torch.uint8, device="cuda")
    want = torch.zeros_like(frame)
    world.render_static(cam_a, vx.default_env(), opts, frame.data_ptr(), W, H)   # first call on the device: symbol look-up
    rt.render_device(cam_b, vx.default_env(), opts, want.data_ptr(), W, H)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for _ in range(3):
        world.render_static(cam_a, vx.default_env(), opts, frame.data_ptr(), W, H)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0
    side = torch.cuda.Stream()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        graph.capture_begin(capture_error_mode="global")
        try:
            world.render_static(cam_b, vx.default_env(), opts, frame.data_ptr(), W, H, stream=side.cuda_stream)
        finally:
            graph.capture_end()
    frame.zero_()
    torch.cuda.synchronize()
    for _ in range(2):
        graph.replay()
    torch.cuda.synchronize()
    assert torch.equal(frame, want)
