"""Pins the CPU oracle: (1) against golden vectors produced by the REFERENCE compiled for the host
(tests/golden/gen_golden.py), (2) directly against that host build when oracle/_ref is present."""
import ctypes as C
import os

import numpy as np
import pytest

import oracle_lib as ol
import worlds
from raysets import mixed_rays


@pytest.mark.parametrize("name", worlds.GOLDEN_CASES)
def test_oracle_matches_reference_golden_vectors(name):
    g = worlds.golden_case(name)
    bits = worlds.dense(g["dims"], g["origin"])
    assert worlds.crc(bits) == g["dense_crc"]
    w = worlds.oracle_world(g["dims"], g["origin"], g["factor"])
    # GenerateLowresVoxelBuffer outputs
    assert np.array_equal(w.top, g["top"])
    assert worlds.crc(w.brick_bits) == g["brick_crc"]
    assert worlds.crc(w.bounds) == g["bounds_crc"]
    # GPUDDA::Raytrace outputs: bit-exact
    o, d = mixed_rays(g["dims"], g["nrays"], g["seed"])
    r = w.trace(o, d, norm_mode=ol.NORM_HOST)
    assert np.array_equal(r["hit"], g["hit"])
    assert np.array_equal(r["steps"], g["steps"])
    assert np.array_equal(r["pos"].view(np.uint32), g["pos"])
    assert np.array_equal(r["normal"], g["normal"])
    assert np.array_equal(r["steps"], r["coarse_steps"] + r["fine_steps"])


def test_noise_matches_reference_golden_vectors():
    g = np.load(os.path.join(worlds.GOLDEN, "noise.npz"))
    L = ol.lib()
    s = np.float32(0.005)
    vals = np.array([L.vxo_terrain_noise(np.float32(x) * s, np.float32(y) * s, np.float32(z) * s, ol.F2U_X86)
                     for x, y, z in g["pts"]], np.float32)
    assert np.array_equal(vals.view(np.uint32), g["vals"])
    hashes = np.array([L.vxo_hash(int(v)) for v in range(0, 4000 * 7919, 7919)], np.uint32)
    assert np.array_equal(hashes, g["hashes"])


def test_float_to_uint_modes_differ_only_by_saturation():
    """cuda_noise.cuh:118-121: the lattice sum exceeds 2^32 at high octaves; CUDA saturates, x86 wraps."""
    L = ol.lib()
    a = L.vxo_perlin(1.0e6, 3.0, 5.0e6, 1.0, 38 * 27389482, ol.F2U_CUDA)
    b = L.vxo_perlin(1.0e6, 3.0, 5.0e6, 1.0, 38 * 27389482, ol.F2U_X86)
    assert np.isfinite(a) and np.isfinite(b)
    # low octaves / small coordinates: both conversions agree
    for p in [(0.5, 0.25, 0.75), (3.2, 1.1, 7.7), (40.9, 2.5, 40.9)]:
        assert L.vxo_perlin(*p, 1.0, 38 * 27389482, ol.F2U_CUDA) == L.vxo_perlin(*p, 1.0, 38 * 27389482, ol.F2U_X86)


def test_tiled_index_matches_definition():
    L = ol.lib()
    rng = np.random.default_rng(0)
    W, H = 128, 64
    for _ in range(200):
        x, y, z = int(rng.integers(0, W)), int(rng.integers(0, H)), int(rng.integers(0, 256))
        tile = (x // 8) + (y // 8) * (W // 8) + (z // 8) * (W // 8) * (H // 8)
        assert L.vxo_sample_index(x, y, z, W, H) == tile * 512 + (x % 8) + 8 * (y % 8) + 64 * (z % 8)


def test_voxel_rule_bottom_layer_is_solid():
    """y == 0 is always solid: !(0 > max(t, 0)) — VoxelWorldBuilder.cu:23-34."""
    L = ol.lib()
    for x, z in [(0, 0), (17, 900), (8000, 4000)]:
        assert L.vxo_voxel_solid(x, 0, z, ol.F2U_CUDA) == 1


# ---- direct comparison with the reference's host build (present in the build container and on the
# ---- GPU box, where oracle/_ref travels; skipped if absent) ------------------------------------
ref = ol.ref_host()
needs_ref = pytest.mark.skipif(ref is None, reason="oracle/_ref/libvxref_host.so not built")


@needs_ref
def test_slab_test_against_reference():
    rng = np.random.default_rng(3)
    L = ol.lib()
    for i in range(3000):
        s = rng.uniform(-4, 12, 3).astype(np.float32)
        d = rng.normal(size=3).astype(np.float32)
        if i % 7 == 0:
            d[rng.integers(0, 3)] = 0
        lo = rng.uniform(0, 8, 3).astype(np.float32)
        hi = (lo + rng.uniform(0, 4, 3)).astype(np.float32)
        if i % 11 == 0:
            s = np.floor(s)
            lo = np.floor(lo)
        pa, na, pb, nb = (np.zeros(3, np.float32) for _ in range(4))
        ha = L.vxo_ray_aabb(s, d, lo, hi, pa, na)
        hb = ref.vxref_ray_aabb(s, d, lo, hi, pb, nb)
        assert ha == hb
        if ha:
            assert np.array_equal(pa.view(np.uint32), pb.view(np.uint32)) and np.array_equal(na, nb)


@needs_ref
@pytest.mark.parametrize("case", [((64, 64, 64), (0, 0, 0), 8), ((128, 128, 128), (100, 40, 900), 16)])
def test_march_against_reference_fresh_rays(case):
    dims, org, factor = case
    bits = worlds.dense(dims, org)
    w = worlds.oracle_world(dims, org, factor)
    rw = ref.vxref_build(bits, *dims, factor)
    n = 30000
    o, d = mixed_rays(dims, n, seed=99)
    r = w.trace(o, d, norm_mode=ol.NORM_HOST)
    hit = np.zeros(n, np.uint8)
    pos = np.zeros((n, 3), np.float32)
    nrm = np.zeros((n, 3), np.float32)
    st = np.zeros(n, np.int32)
    ref.vxref_trace(rw, o, d, n, 2048, hit.ctypes.data, pos, nrm, st.ctypes.data, 0)
    assert np.array_equal(hit, r["hit"]) and np.array_equal(st, r["steps"])
    assert np.array_equal(pos.view(np.uint32), r["pos"].view(np.uint32)) and np.array_equal(nrm, r["normal"])


@needs_ref
def test_step_cap_against_reference():
    """Raytrace(maxSteps = 8, ...) as the AO loop uses (Renderer.cu:141)."""
    dims, org = (64, 64, 64), (0, 0, 0)
    w = worlds.oracle_world(dims, org, 8)
    rw = ref.vxref_build(worlds.dense(dims, org), *dims, 8)
    n = 5000
    o, d = mixed_rays(dims, n, seed=5)
    r = w.trace(o, d, norm_mode=ol.NORM_HOST, max_steps=8)
    hit = np.zeros(n, np.uint8)
    pos = np.zeros((n, 3), np.float32)
    nrm = np.zeros((n, 3), np.float32)
    st = np.zeros(n, np.int32)
    ref.vxref_trace(rw, o, d, n, 8, hit.ctypes.data, pos, nrm, st.ctypes.data, 0)
    assert np.array_equal(hit, r["hit"]) and np.array_equal(st, r["steps"])


def test_ortho_camera_is_the_batch_trace_of_its_rays():
    """getRayDirectionOrtho (Renderer.cu:61-70): the oracle's orthographic frame has, pixel for pixel, the step counts of
    its own batch march over rays built here independently: origin + right*(u*2-1)*size.x*ratio + up*(v*2-1)*size.y in
    float32, left to right, direction = fwd.  (The GPU test against the reference's -DORTHO build pins the camera.)"""
    dims, org = (128, 64, 128), (512, 0, 256)
    w = worlds.oracle_world(dims, org, 8)
    W, H, size = 96, 54, (40.0, 30.0)
    pos, eul = (64, 110, 120), (-0.7, 0.3, 0.0)
    cam = ol.make_camera(pos, eul, 90.0, tan_half=1.0, ortho=size)
    opts = ol.RenderOpts(0, 0, 1, 0, 0, ol.NORM_HOST)
    _, steps = w.render(cam, ol.default_env(), opts, W, H)
    f, u, r = ol.get_directions(eul)
    f32 = np.float32
    ys, xs = np.mgrid[0:H, 0:W]
    a = (xs.astype(f32) / f32(W) * f32(2) - f32(1)).astype(f32)
    b = (ys.astype(f32) / f32(H) * f32(2) - f32(1)).astype(f32)
    ratio = f32(W) / f32(H)
    o = np.empty((H, W, 3), f32)
    for k in range(3):
        t = ((r[k] * a).astype(f32) * f32(size[0])).astype(f32) * ratio
        t2 = (u[k] * b).astype(f32) * f32(size[1])
        o[..., k] = (f32(pos[k]) + t.astype(f32)).astype(f32) + t2.astype(f32)
    d = np.broadcast_to(f, (H * W, 3)).copy()
    res = w.trace(o.reshape(-1, 3), d, norm_mode=ol.NORM_HOST)
    assert np.array_equal(res["steps"].reshape(H, W), steps)
    assert 0.2 < res["hit"].mean() < 1.0
    persp = w.render(ol.make_camera(pos, eul, 90.0, tan_half=1.0), ol.default_env(), opts, W, H)[1]
    assert not np.array_equal(persp, steps)
