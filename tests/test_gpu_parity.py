"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle and the committed
golden vectors of the reference.  Integer/index planes and float bit patterns must be identical."""
import numpy as np
import pytest

import oracle_lib as ol
import worlds
from raysets import camera_rays, mixed_rays

pytestmark = pytest.mark.gpu

INT_PLANES = ("hit", "steps", "voxel", "brick", "coarse_steps", "fine_steps")


def assert_same_hits(got, ref, counters=False):
    for k in INT_PLANES:
        bad = np.nonzero((got[k] != ref[k]).reshape(len(ref["hit"]), -1).any(1))[0]
        assert bad.size == 0, "plane %s differs at rays %s" % (k, bad[:8])
    assert np.array_equal(got["pos"].view(np.uint32), ref["pos"].view(np.uint32)), "hit positions differ"
    assert np.array_equal(got["normal"], ref["normal"]), "normals differ"
    if counters:
        assert np.array_equal(got["counters"], ref["counters"]), "work counters differ"


@pytest.fixture(scope="module")
def gpu_worlds(vx):
    cache = {}

    def get(dims, origin, factor, source="noise"):
        key = (dims, origin, factor, source)
        if key not in cache:
            if source == "noise":
                cache[key] = vx.World.from_noise(*dims, factor, origin=origin)
            else:
                cache[key] = vx.World.from_dense_bits(worlds.dense(dims, origin), *dims, factor)
        return cache[key]
    yield get
    for w in cache.values():
        w.close()


@pytest.mark.parametrize("name", worlds.GOLDEN_CASES)
def test_world_build_matches_reference(vx, gpu_worlds, name):
    """Device noise build == GenerateLowresVoxelBuffer(PopulateVoxels) of the reference (golden CRCs)
    and == the oracle, word for word."""
    g = worlds.golden_case(name)
    w = gpu_worlds(g["dims"], g["origin"], g["factor"])
    top, bw, b6 = w.download()
    assert np.array_equal(top, g["top"])
    assert worlds.crc(bw) == g["brick_crc"]
    assert worlds.crc(b6) == g["bounds_crc"]
    ow = worlds.oracle_world(g["dims"], g["origin"], g["factor"])
    assert np.array_equal(bw, ow.brick_bits) and np.array_equal(b6, ow.bounds)
    info = w.info
    assert info.occupied_bricks == sum(bin(int(x)).count("1") for x in top)
    assert info.solid_voxels == sum(bin(int(x)).count("1") for x in bw)


def test_early_out_build_equals_full_octave_build(vx):
    dims, org = (128, 64, 128), (4096, 64, 2048)
    a = vx.World.from_noise(*dims, 8, origin=org)
    b = vx.World.from_noise(*dims, 8, origin=org, full_octaves=True)
    for p, q in zip(a.download(), b.download()):
        assert np.array_equal(p, q)
    a.close(); b.close()


def test_world_from_dense_bits_equals_noise_build(vx, gpu_worlds):
    dims, org = (64, 64, 64), (100, 40, 900)
    a = gpu_worlds(dims, org, 8, "noise").download()
    b = gpu_worlds(dims, org, 8, "dense").download()
    for p, q in zip(a, b):
        assert np.array_equal(p, q)


@pytest.mark.parametrize("name", worlds.GOLDEN_CASES)
def test_trace_matches_reference_golden_vectors(vx, gpu_worlds, name):
    g = worlds.golden_case(name)
    w = gpu_worlds(g["dims"], g["origin"], g["factor"])
    o, d = mixed_rays(g["dims"], g["nrays"], g["seed"])
    rt = vx.VoxelRaytracer3D(w, g["nrays"])
    r = rt.raytrace(o, d, norm_mode=vx.NORM_HOST)
    assert np.array_equal(r["hit"], g["hit"])
    assert np.array_equal(r["steps"], g["steps"])
    assert np.array_equal(r["pos"].view(np.uint32), g["pos"])
    assert np.array_equal(r["normal"], g["normal"])
    rt.close()


@pytest.mark.parametrize("case", [((64, 64, 64), (0, 0, 0), 8, 200000), ((128, 64, 128), (512, 0, 256), 8, 200000),
                                  ((128, 128, 128), (100, 40, 900), 16, 100000), ((256, 256, 256), (0, 0, 0), 32, 50000)])
def test_trace_matches_oracle_all_planes(vx, gpu_worlds, case):
    dims, org, factor, n = case
    w = gpu_worlds(dims, org, factor)
    ow = worlds.oracle_world(dims, org, factor)
    o, d = mixed_rays(dims, n, seed=321)
    rt = vx.VoxelRaytracer3D(w, n)
    got = rt.raytrace(o, d, norm_mode=vx.NORM_HOST, counters=True)
    ref = ow.trace(o, d, norm_mode=ol.NORM_HOST)
    assert_same_hits(got, ref, counters=True)
    # reference host post-processing (VolumeRaytracer.cu:601-614)
    assert np.array_equal(got["valid"], ref["hit"])
    rt.close()


def test_device_normalisation_path_matches_oracle(vx, gpu_worlds):
    """Default mode (device rsqrtf, what the reference kernels do): the oracle is fed the
    device-normalised directions and must then agree bit for bit (SURVEY H1 ii)."""
    dims, org = (128, 64, 128), (512, 0, 256)
    w = gpu_worlds(dims, org, 8)
    ow = worlds.oracle_world(dims, org, 8)
    n = 100000
    o, d = mixed_rays(dims, n, seed=77)
    rt = vx.VoxelRaytracer3D(w, n)
    got = rt.raytrace(o, d, norm_mode=vx.NORM_DEVICE)
    dn = vx.device_normalize(d)
    ref = ow.trace(o, dn, norm_mode=ol.NORM_NONE)
    assert_same_hits(got, ref)
    got2 = rt.raytrace(o, dn, norm_mode=vx.NORM_NONE)
    assert_same_hits(got2, ref)
    rt.close()


def test_camera_rays_and_step_cap(vx, gpu_worlds):
    dims, org = (128, 64, 128), (512, 0, 256)
    w = gpu_worlds(dims, org, 8)
    ow = worlds.oracle_world(dims, org, 8)
    f, u, r = vx.get_directions((-0.45, 0.3, 0.0))
    o, d = camera_rays((64, 90, 150), f, u, r, 320, 180, 1.0)
    rt = vx.VoxelRaytracer3D(w, o.shape[0])
    got = rt.raytrace(o, d, norm_mode=vx.NORM_HOST, counters=True)
    ref = ow.trace(o, d, norm_mode=ol.NORM_HOST)
    assert_same_hits(got, ref, counters=True)
    assert got["hit"].mean() > 0.2
    got8 = rt.raytrace(o, d, norm_mode=vx.NORM_HOST, max_steps=8)
    ref8 = ow.trace(o, d, norm_mode=ol.NORM_HOST, max_steps=8)
    assert_same_hits(got8, ref8)
    rt.close()


def test_empty_batch_and_capacity(vx, gpu_worlds):
    w = gpu_worlds((64, 64, 64), (0, 0, 0), 8)
    rt = vx.VoxelRaytracer3D(w, 16)
    r = rt.raytrace(np.zeros((0, 3), np.float32), np.zeros((0, 3), np.float32))
    assert r["hit"].size == 0
    with pytest.raises(vx.VxError):
        rt.raytrace(np.zeros((17, 3), np.float32), np.ones((17, 3), np.float32))
    rt.close()


@pytest.mark.parametrize("mode", [dict(checkerboard=1, debug_view=1, shadows=0, frame=0), dict(checkerboard=1, debug_view=1, shadows=0, frame=1),
                                  dict(checkerboard=0, debug_view=0, shadows=1, frame=0), dict(checkerboard=1, debug_view=0, shadows=1, frame=3),
                                  dict(checkerboard=0, debug_view=0, shadows=0, frame=0)])
def test_render_matches_oracle(vx, gpu_worlds, mode):
    """screenDispatch pixel logic: identical written-pixel set; RGB within 1/255 (powf/tanf are
    library functions: tolerance stated by north_star); step plane bit-exact."""
    dims, org = (128, 64, 128), (512, 0, 256)
    w = gpu_worlds(dims, org, 8)
    ow = worlds.oracle_world(dims, org, 8)
    W, H = 256, 144
    tan_half = vx.device_tanf(np.float32(np.float32(90.0 * 3.1415 / 180.0) / np.float32(2.0)))
    euler, cam_pos = (-0.45, 0.3, 0.0), (64, 90, 150)
    cam = vx.make_camera(cam_pos, euler, 90.0)
    ocam = ol.make_camera(cam_pos, euler, 90.0, tan_half=tan_half)
    opts = vx.RenderOpts(mode["checkerboard"], mode["frame"], mode["debug_view"], mode["shadows"], 1, vx.NORM_HOST, 0, 0, 0, 0, None)
    oopts = ol.RenderOpts(mode["checkerboard"], mode["frame"], mode["debug_view"], mode["shadows"], 1, ol.NORM_HOST)
    rt = vx.VoxelRaytracer3D(w, 1)
    seed = np.full((H, W, 4), 7, np.uint8)  # unwritten pixels must keep this value
    got = rt.render_host(cam, vx.default_env(), opts, W, H, bgra=seed.copy())
    ref, _ = ow.render(ocam, ol.default_env(), oopts, W, H, bgra=seed.copy())
    written_g = (got != 7).any(2) | (got[..., 3] == 255)
    written_r = (ref != 7).any(2) | (ref[..., 3] == 255)
    assert np.array_equal(written_g, written_r)
    diff = np.abs(got.astype(np.int16) - ref.astype(np.int16))
    assert diff.max() <= 1, "max channel difference %d" % diff.max()
    assert (diff > 0).mean() < 0.01
    rt.close()


@pytest.mark.parametrize("mode", [dict(checkerboard=0, debug_view=0, shadows=1), dict(checkerboard=1, debug_view=1, shadows=0),
                                  dict(checkerboard=0, debug_view=0, shadows=0)])
def test_ortho_render_matches_oracle(vx, gpu_worlds, mode):
    """Orthographic camera (getRayDirectionOrtho, Renderer.cu:61-70): per-pixel origins, rays along fwd; shading and the
    debug distance use the pixel's own origin.  Same criteria as the perspective test; dense and fused shading agree."""
    dims, org = (128, 64, 128), (512, 0, 256)
    w = gpu_worlds(dims, org, 8)
    ow = worlds.oracle_world(dims, org, 8)
    W, H = 256, 144
    euler, cam_pos, size = (-0.7, 0.3, 0.0), (64, 110, 120), (40.0, 30.0)
    cam = vx.make_camera(cam_pos, euler, 90.0, ortho=size)
    ocam = ol.make_camera(cam_pos, euler, 90.0, tan_half=1.0, ortho=size)
    opts = vx.RenderOpts(mode["checkerboard"], 2, mode["debug_view"], mode["shadows"], 1, vx.NORM_HOST, 0, 0, 0, 0, None)
    oopts = ol.RenderOpts(mode["checkerboard"], 2, mode["debug_view"], mode["shadows"], 1, ol.NORM_HOST)
    rt = vx.VoxelRaytracer3D(w, 1)
    seed = np.full((H, W, 4), 7, np.uint8)
    ref, rsteps = ow.render(ocam, ol.default_env(), oopts, W, H, bgra=seed.copy())
    for fused in (0, 1):
        rt.set_fused_shade(fused)
        got = rt.render_host(cam, vx.default_env(), opts, W, H, bgra=seed.copy())
        assert np.array_equal((got != 7).any(2) | (got[..., 3] == 255), (ref != 7).any(2) | (ref[..., 3] == 255))
        diff = np.abs(got.astype(np.int16) - ref.astype(np.int16))
        assert diff.max() <= 1, "max channel difference %d" % diff.max()
        assert (diff > 0).mean() < 0.01
    hit_share = ((ref[..., :3] != 7).any(2)).mean()
    assert hit_share > 0.3
    rt.close()


def test_full_scale_frame_matches_oracle(vx):
    """BASELINE config 2 scale: 1024x256x1024 world built on the device, handed to the oracle in the
    reference's form; whole shaded + shadowed frame: identical written pixels, RGB within 1/255, and the
    per-pixel step plane (primary + shadow DDA steps) bit-exact."""
    import torch
    X, Y, Z = 1024, 256, 1024
    w = vx.World.from_noise(X, Y, Z, 8)
    top, bw, b6 = w.download()
    ow = ol.OracleWorld.from_brickmap(top, bw, b6, X, Y, Z, 8)
    W, H = 960, 540
    tan_half = vx.device_tanf(np.float32(np.float32(90.0 * 3.1415 / 180.0) / np.float32(2.0)))
    rt = vx.VoxelRaytracer3D(w, 1)
    probe = vx.VoxelRaytracer3D(w, 4)
    for (px, pz), eul in [((512, 960), (-0.45, 0.0, 0.0)), ((300, 300), (-0.1, 0.8, 0.0))]:
        r = probe.raytrace(np.array([[px + 0.5, Y + 10.0, pz + 0.5]], np.float32), np.array([[0, -1, 0]], np.float32))
        top_y = int(r["voxel"][0, 1]) if r["hit"][0] else -1
        pos = (px, (top_y + 16 if top_y < Y - 1 else Y + 32), pz)  # camera height rule of SURVEY.md 8d
        cam = vx.make_camera(pos, eul, 90.0)
        ocam = ol.make_camera(pos, eul, 90.0, tan_half=tan_half)
        frame = torch.full((H, W, 4), 7, dtype=torch.uint8, device="cuda")
        steps = torch.full((H, W), -1, dtype=torch.int32, device="cuda")
        opts = vx.RenderOpts(0, 0, 0, 1, 1, vx.NORM_HOST, 0, 0, 0, 0, None)
        rt.render_device(cam, vx.default_env(), opts, frame.data_ptr(), W, H, d_steps=steps.data_ptr())
        torch.cuda.synchronize()
        got, gsteps = frame.cpu().numpy(), steps.cpu().numpy()
        ref, rsteps = ow.render(ocam, ol.default_env(), ol.RenderOpts(0, 0, 0, 1, 1, ol.NORM_HOST), W, H,
                                bgra=np.full((H, W, 4), 7, np.uint8))
        assert np.array_equal(gsteps, rsteps), "step planes differ at %d pixels" % int((gsteps != rsteps).sum())
        diff = np.abs(got.astype(np.int16) - ref.astype(np.int16))
        assert diff.max() <= 1
        assert gsteps.mean() > 5
    probe.close(); rt.close(); w.close()


def test_bands_and_tiles_compose_the_full_frame(vx, gpu_worlds):
    """Multi-GPU partition, device side: rendering the bands of every rank (vx_render_opts.band_count/band_index), or a set
    of tile rectangles, into one buffer gives the same bytes and steps as rendering the whole frame at once."""
    import torch
    dims, org = (128, 64, 128), (512, 0, 256)
    w = gpu_worlds(dims, org, 8)
    W, H = 200, 117  # not a multiple of the 16x8 macro tile
    cam = vx.make_camera((64, 90, 150), (-0.45, 0.3, 0.0), 90.0)
    rt = vx.VoxelRaytracer3D(w, 1)

    def render(**kw):
        frame = torch.full((H, W, 4), 7, dtype=torch.uint8, device="cuda")
        steps = torch.full((H, W), -1, dtype=torch.int32, device="cuda")
        for o in kw["passes"]:
            rt.render_device(cam, vx.default_env(), o, frame.data_ptr(), W, H, d_steps=steps.data_ptr())
        torch.cuda.synchronize()
        return frame.cpu().numpy(), steps.cpu().numpy()

    for cb in (0, 1):
        whole = render(passes=[vx.RenderOpts(cb, 2, 0, 1, 1, vx.NORM_DEVICE, 0, 0, 0, 0, None)])
        for n in (2, 3, 8):
            banded = render(passes=[vx.RenderOpts(cb, 2, 0, 1, 1, vx.NORM_DEVICE, 0, 0, 0, 0, None, n, r) for r in range(n)])
            assert np.array_equal(whole[0], banded[0]) and np.array_equal(whole[1], banded[1]), "bands x%d, checkerboard %d" % (n, cb)
        rows = H // 2 if cb else H
        tiles = [(x0, y0, 96, 40) for y0 in range(0, rows, 40) for x0 in range(0, 224, 96)]  # thread-space rectangles
        tiled = render(passes=[vx.RenderOpts(cb, 2, 0, 1, 1, vx.NORM_DEVICE, x0, y0, tw, th, None) for (x0, y0, tw, th) in tiles])
        assert np.array_equal(whole[0], tiled[0]) and np.array_equal(whole[1], tiled[1]), "tiles, checkerboard %d" % cb
    # shading inside the sun-ray march (the variant chosen for a frame in another GPU's memory) writes the same frame
    whole = render(passes=[vx.RenderOpts(0, 2, 0, 1, 1, vx.NORM_DEVICE, 0, 0, 0, 0, None)])
    for norm in (vx.NORM_DEVICE, vx.NORM_HOST):
        rt.set_fused_shade(0)
        dense = render(passes=[vx.RenderOpts(0, 2, 0, 1, 1, norm, 0, 0, 0, 0, None)])
        rt.set_fused_shade(1)
        fused = render(passes=[vx.RenderOpts(0, 2, 0, 1, 1, norm, 0, 0, 0, 0, None)])
        assert np.array_equal(dense[0], fused[0]) and np.array_equal(dense[1], fused[1]), "fused shade, norm %d" % norm
    rt.set_fused_shade(0)
    with pytest.raises(vx.VxError):
        rt.set_fused_shade(2)
    # row chunks x bands, the passes FramePartition.render_and_gather issues on N ranks (bench.py, N > 1)
    from voxelengine_b200.partition import FramePartition
    whole = render(passes=[vx.RenderOpts(0, 2, 0, 1, 1, vx.NORM_DEVICE, 0, 0, 0, 0, None)])
    for n in (2, 3):
        chunks = FramePartition(W, H, n, 0).row_chunks(3)
        passes = [vx.RenderOpts(0, 2, 0, 1, 1, vx.NORM_DEVICE, 0, y0, 224, h, None, n, r) for (y0, h) in chunks for r in range(n)]
        chunked = render(passes=passes)
        assert np.array_equal(whole[0], chunked[0]) and np.array_equal(whole[1], chunked[1]), "chunks x bands x%d" % n
    rt.close()


from bounce_ref import host_normalize as _host_normalize, noise_hash as _noise_hash, sample_rays  # noqa: E402


@pytest.mark.parametrize("checkerboard,frame", [(0, 0), (1, 3)])
def test_bounce_rays_match_reference_sample_loop(vx, gpu_worlds, checkerboard, frame):
    """BASELINE config 5: the AO sample ray of Renderer.cu:121-165 for every primary hit.  Generation is checked against
    a numpy restatement (hash, randomFloat, hemisphere flip, origin offset) fed with the primary hits of the batch API;
    the incoherent rays are then traced on the GPU and by the oracle: every plane identical."""
    import torch
    from raysets import camera_rays
    dims, org = (128, 64, 128), (512, 0, 256)
    w = gpu_worlds(dims, org, 8)
    ow = worlds.oracle_world(dims, org, 8)
    W, H = 256, 144
    tan_half = vx.device_tanf(np.float32(np.float32(90.0 * 3.1415 / 180.0) / np.float32(2.0)))
    euler, cam_pos = (-0.45, 0.3, 0.0), (64, 90, 150)
    cam = vx.make_camera(cam_pos, euler, 90.0)
    rt = vx.VoxelRaytracer3D(w, W * H)
    opts = vx.RenderOpts(checkerboard, frame, 0, 1, 1, vx.NORM_HOST, 0, 0, 0, 0, None)
    d_o = torch.zeros(W * H * 3, dtype=torch.float32, device="cuda")
    d_d = torch.zeros(W * H * 3, dtype=torch.float32, device="cuda")
    d_p = torch.zeros(W * H, dtype=torch.int32, device="cuda")
    d_n = torch.zeros(1, dtype=torch.int64, device="cuda")
    rt.bounce_rays(cam, opts, W, H, d_o.data_ptr(), d_d.data_ptr(), d_p.data_ptr(), d_n.data_ptr())
    torch.cuda.synchronize()
    n = int(d_n.item())
    order = np.argsort(d_p[:n].cpu().numpy())
    pix = d_p[:n].cpu().numpy()[order]
    bo = d_o[:3 * n].cpu().numpy().reshape(-1, 3)[order]
    bd = d_d[:3 * n].cpu().numpy().reshape(-1, 3)[order]
    assert n > 1000 and len(np.unique(pix)) == n

    # expected: primary rays of the written pixels through the batch API, then the sample loop in numpy
    fwd, up, right = [np.array(v, np.float32) for v in (cam.fwd, cam.up, cam.right)]
    po, pd = camera_rays(cam_pos, fwd, up, right, W, H, tan_half)
    prim = rt.raytrace(po, _host_normalize(pd), norm_mode=vx.NORM_HOST)
    ys, xs = np.divmod(np.arange(W * H), W)
    if checkerboard:  # thread (tx, ty) -> pixel (x, 2 ty + (x even) + (frame even)); other pixels are not rendered
        adj = (xs % 2 == 0).astype(np.int64) + (1 if frame % 2 == 0 else 0)
        rendered = ((ys - adj) >= 0) & ((ys - adj) % 2 == 0) & ((ys - adj) // 2 < H // 2)
        tys = (ys - adj) // 2
    else:
        rendered, tys = np.ones(W * H, bool), ys
    sel = np.nonzero(rendered & prim["hit"].astype(bool))[0]
    assert np.array_equal(sel, pix), "bounce rays must exist exactly for the rendered primary hits"
    nrm = (-prim["normal"][sel]).astype(np.float32)
    eo, sd = sample_rays(xs[sel], tys[sel], W, frame, prim["pos"][sel], nrm)
    assert np.array_equal(bd.view(np.uint32), sd.view(np.uint32)), "sample directions differ"
    assert np.array_equal(bo.view(np.uint32), eo.view(np.uint32)), "sample origins differ"

    got = rt.raytrace(bo, bd, norm_mode=vx.NORM_HOST, counters=True)
    ref = ow.trace(bo, bd, norm_mode=ol.NORM_HOST)
    assert_same_hits(got, ref, counters=True)
    got8 = rt.raytrace(bo, bd, norm_mode=vx.NORM_HOST, max_steps=8)  # the cap the reference passes (Renderer.cu:143)
    ref8 = ow.trace(bo, bd, norm_mode=ol.NORM_HOST, max_steps=8)
    assert_same_hits(got8, ref8)
    rt.close()


def test_render_host_sliced_path_equals_device_render(vx, gpu_worlds):
    """vx_render_host renders large plain frames in slices to overlap the copy to the host; the bytes must be those of
    one vx_render into device memory."""
    import torch
    dims, org = (128, 64, 128), (512, 0, 256)
    w = gpu_worlds(dims, org, 8)
    W, H = 320, 300  # >= 256 rows: sliced
    cam = vx.make_camera((64, 90, 150), (-0.45, 0.3, 0.0), 90.0)
    for dbg, sh in ((0, 1), (1, 0)):
        rt = vx.VoxelRaytracer3D(w, 1)  # fresh device frame: pixels a mode does not write keep the caller's bytes
        opts = vx.RenderOpts(0, 0, dbg, sh, 1, vx.NORM_DEVICE, 0, 0, 0, 0, None)
        frame = torch.full((H, W, 4), 7, dtype=torch.uint8, device="cuda")
        rt.render_device(cam, vx.default_env(), opts, frame.data_ptr(), W, H)
        torch.cuda.synchronize()
        got = rt.render_host(cam, vx.default_env(), opts, W, H, bgra=np.full((H, W, 4), 7, np.uint8))
        assert np.array_equal(got, frame.cpu().numpy())
        rt.close()


def test_batch_call_pipeline_planes_and_ragged_chunks(vx, gpu_worlds):
    """vx_trace with host arrays runs as a pipeline over 256 Ki-ray chunks through a page-locked ring: a batch that is not
    a whole number of chunks, every plane and subsets of planes, must equal the device-pointer call and the oracle."""
    import torch
    dims, org = (128, 64, 128), (512, 0, 256)
    w = gpu_worlds(dims, org, 8)
    ow = worlds.oracle_world(dims, org, 8)
    n = 3 * 256 * 1024 + 12345   # four chunks, the last one ragged; more chunks than ring slots
    o, d = mixed_rays(dims, n, seed=4242)
    rt = vx.VoxelRaytracer3D(w, n)
    full = rt.raytrace(o, d, norm_mode=vx.NORM_HOST, counters=True)
    ref = ow.trace(o, d, norm_mode=ol.NORM_HOST)
    assert_same_hits(full, ref, counters=True)
    sub = rt.raytrace(o, d, norm_mode=vx.NORM_HOST, planes=rt.REFERENCE_PLANES)
    assert sorted(sub) == sorted(rt.REFERENCE_PLANES)
    for k in rt.REFERENCE_PLANES:
        assert np.array_equal(sub[k].view(np.uint8), full[k].view(np.uint8)), k
    one = rt.raytrace(o, d, norm_mode=vx.NORM_HOST, planes=("brick",), out=dict(brick=np.full(n, 7, np.uint32)))
    assert np.array_equal(one["brick"], full["brick"])
    # a second call reuses ring, threads and streams; a small batch takes the plain path
    again = rt.raytrace(o[:1000], d[:1000], norm_mode=vx.NORM_HOST)
    for k in INT_PLANES:
        assert np.array_equal(again[k], full[k][:1000])
    rt.close()


@pytest.mark.parametrize("case", [((256, 128, 256), (512, 0, 256), 8, 3), ((128, 128, 256), (100, 40, 900), 16, 2), ((64, 64, 64), (0, 0, 0), 8, 4)])
def test_partitioned_build_equals_single_device_build(vx, case):
    """BASELINE configs[4]: the world built slab by slab (vx_slab_build_noise, here all slabs on one device, one after the
    other) and assembled from the concatenated parts (vx_world_from_parts) is word for word the world of one
    vx_world_build_noise — top-level bits, masks, bounds, counters — and traces the same (more builders than z-tiles
    leaves some slabs empty: the last case)."""
    import torch
    dims, org, factor, n = case
    whole = vx.World.from_noise(*dims, factor, origin=org)
    nb = (dims[0] // factor) * (dims[1] // factor) * (dims[2] // factor)
    wpb = factor ** 3 // 32
    slabs = [vx.Slab(*dims, factor, r, n, origin=org) for r in range(n)]
    infos = [s.info for s in slabs]
    assert [int(i.brick_begin) for i in infos] == [sum(int(j.brick_count) for j in infos[:r]) for r in range(n)]
    assert sum(int(i.brick_count) for i in infos) == nb
    occupied = sum(int(i.occupied_bricks) for i in infos)
    top = torch.zeros(nb // 32, dtype=torch.int32, device="cuda")
    bounds = torch.zeros(nb, dtype=torch.int32, device="cuda")
    pool = torch.zeros(max(occupied, 1) * wpb, dtype=torch.int32, device="cuda")
    slot = 0
    for s, i in zip(slabs, infos):
        b0, bn, occ = int(i.brick_begin), int(i.brick_count), int(i.occupied_bricks)
        if bn:
            s.export(top[b0 // 32:].data_ptr(), bounds[b0:].data_ptr(), pool[slot * wpb:].data_ptr() if occ else 0)
        slot += occ
        s.close()
    torch.cuda.synchronize()
    parts = vx.World.from_parts(*dims, factor, top.data_ptr(), bounds.data_ptr(), pool.data_ptr(), occupied, sum(int(i.solid_voxels) for i in infos))
    for a, b in zip(whole.download(), parts.download()):
        assert np.array_equal(a, b)
    assert (whole.info.occupied_bricks, whole.info.solid_voxels) == (parts.info.occupied_bricks, parts.info.solid_voxels)
    o, d = mixed_rays(dims, 50000, seed=5)
    r1, r2 = vx.VoxelRaytracer3D(whole, 50000), vx.VoxelRaytracer3D(parts, 50000)
    assert_same_hits(r1.raytrace(o, d), r2.raytrace(o, d))
    with pytest.raises(vx.VxError):   # a pool that does not belong to these top-level bits is refused
        vx.World.from_parts(*dims, factor, top.data_ptr(), bounds.data_ptr(), pool.data_ptr(), occupied + 1, 0)
    r1.close(); r2.close(); whole.close(); parts.close()


def test_chain_shortcuts_equal_the_ieee_operations(vx):
    """The two shortcuts of the per-brick chain (vx_march.cuh): tmax_div(a, d) must be a / d for every float a and every
    finite non-zero d (zero numerators are answered without the division), next_toward(x, down) must be
    nextafterf(x, -+inf) for every float (VolumeRaytracer.cu:203-205, :452-460)."""
    import ctypes as C
    rng = np.random.default_rng(77)
    special = np.array([0.0, -0.0, 1.0, -1.0, 8.0, 0.125, 1e-45, -1e-45, 1.1754944e-38, -1.1754944e-38, 5e-39, 3.4028235e38, -3.4028235e38,
                        np.inf, -np.inf, np.nan, 1e-30, -1e-30, 1e30, 0.99999994, 1.0000001, 511.99997, 512.0], np.float32)
    bits = rng.integers(0, 2**32, 200000, dtype=np.uint64).astype(np.uint32).view(np.float32)   # every exponent, NaNs and denormals included
    near = (rng.uniform(-1030, 1030, 100000)).astype(np.float32)
    a = np.concatenate([np.repeat(special, special.size), bits, near, np.zeros(1000, np.float32), -np.zeros(1000, np.float32)])
    d = np.concatenate([np.tile(special, special.size), np.roll(bits, 1), rng.uniform(-1, 1, 100000).astype(np.float32),
                        rng.uniform(-1, 1, 2000).astype(np.float32)])
    ok = np.isfinite(d) & (d != 0)
    d = np.where(ok, d, np.float32(0.57735026)).astype(np.float32)
    n = a.size
    out = np.zeros((6, n), np.float32)
    L = vx.lib()
    L.vx_debug_chain_math.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int]
    vx._check(L.vx_debug_chain_math(a.ctypes.data, d.ctypes.data, n, out.ctypes.data, 0))
    div, dev_div, up, down, dev_up, dev_down = out

    def same(got, want, what):
        nan = np.isnan(want)
        assert np.array_equal(np.isnan(got), nan), what + ": NaNs differ"
        bad = np.nonzero(~nan & (got.view(np.uint32) != want.view(np.uint32)))[0]
        assert bad.size == 0, "%s: %d differ, first a=%r d=%r got=%r want=%r" % (what, bad.size, a[bad[0]], d[bad[0]], got[bad[0]], want[bad[0]])

    # against the device's own operations (what the reference's kernels execute) ...
    same(div, dev_div, "tmax_div vs div.rn")
    same(up, dev_up, "next_toward(up) vs nextafterf")
    same(down, dev_down, "next_toward(down) vs nextafterf")
    # ... and against the host's (what the reference computes when its source runs on a CPU)
    with np.errstate(all="ignore"):
        same(div, (a / d).astype(np.float32), "tmax_div vs IEEE division")
        same(up, np.nextafter(a, np.float32(np.inf)), "next_toward(up) vs libm")
        same(down, np.nextafter(a, np.float32(-np.inf)), "next_toward(down) vs libm")
    assert int((a == 0).sum()) > 2000 and int((np.abs(d) < 1.1754944e-38).sum()) > 50   # zero numerators and denormal divisors were in the sample
