"""CPU pins against the REFERENCE'S OWN code compiled into oracle/_ref (oracle/Makefile): the pieces of the path whose
oracle form was only a restatement in round 1.

  * the bounce-ray generator (BASELINE config 5; Renderer.cu:121-165): cudaNoise::randomFloat (cuda_noise.cuh:66-71) and
    helper_math.h's normalize (:1325-1329), reflect (:1427-1430) and dot (:1264-1267), as the reference's sample loop
    combines them, against tests/bounce_ref.py — which tests/test_gpu_parity.py in turn compares with the CUDA generator;
  * GPUDDA::Raytrace on a brickmap handed over in the reference's form without its dense grid (vxref_from_brickmap: how
    bench.py's reference arm runs the reference's own march on the 8192x512x8192 world) against the same march on the
    world GenerateLowresVoxelBuffer built, and against the oracle;
  * the oracle's row-sample entry points against its own whole-frame render.
"""
import numpy as np
import pytest

import bounce_ref
import oracle_lib as ol
import worlds
from raysets import mixed_rays

needs_ref = pytest.mark.skipif(ol.ref_host() is None or ol.ref_noise() is None or not hasattr(ol.ref_host(), "vxref_from_brickmap"),
                               reason="oracle/_ref not built (reference sources absent)")


@needs_ref
def test_random_float_is_the_references():
    R = ol.ref_noise()
    rng = np.random.default_rng(5)
    seeds = np.concatenate([np.arange(0, 4096, dtype=np.uint32), rng.integers(0, 2 ** 32, 20000, dtype=np.uint64).astype(np.uint32),
                            np.array([0xffffffff, 0x80000000, 0x7fffffff], np.uint32)])
    got = bounce_ref.random_float(seeds)
    ref = np.array([R.vxref_random_float(int(s)) for s in seeds], np.float32)
    assert np.array_equal(got.view(np.uint32), ref.view(np.uint32))
    assert np.array_equal(bounce_ref.noise_hash(seeds), np.array([R.vxref_hash(int(s)) for s in seeds], np.uint32))


@needs_ref
def test_sample_ray_generator_is_the_references_sample_loop():
    """Renderer.cu:128-142 evaluated with the reference's own randomFloat / normalize / dot / reflect, one hit at a time."""
    H, N = ol.ref_host(), ol.ref_noise()
    rng = np.random.default_rng(11)
    n, W, frame = 3000, 3840, 3
    tx = rng.integers(0, W, n); ty = rng.integers(0, 2160, n)
    pos = (rng.uniform(0, 8192, (n, 3))).astype(np.float32)
    axes = np.eye(3, dtype=np.float32)
    nrm = (axes[rng.integers(0, 3, n)] * rng.choice(np.array([-1.0, 1.0], np.float32), n)[:, None]).astype(np.float32)
    eo, sd = bounce_ref.sample_rays(tx, ty, W, frame, pos, nrm)
    out, tmp = np.zeros(3, np.float32), np.zeros(3, np.float32)
    for i in range(n):
        seed = np.int32(ty[i] * W + tx[i])
        with np.errstate(over="ignore"):
            si = np.int32(seed + np.int32(0 * 1000) + np.int32((frame + 1) * 1000))
            s = [np.uint32(np.int32(si * np.int32(k))) for k in (1, 10, 100)]
        v = np.array([np.float32(N.vxref_random_float(int(x))) * np.float32(2) - np.float32(1) for x in s], np.float32)
        H.vxref_normalize(v, out)
        d = out.copy()
        if H.vxref_dot(d, nrm[i]) < 0:
            H.vxref_reflect(d, nrm[i], tmp)
            d = tmp.copy()
        assert np.array_equal(d.view(np.uint32), sd[i].view(np.uint32)), "sample %d: direction differs" % i
        o = (pos[i] + (nrm[i] * np.float32(0.01)).astype(np.float32)).astype(np.float32)   # position + normal * 0.01f
        assert np.array_equal(o.view(np.uint32), eo[i].view(np.uint32))


@needs_ref
@pytest.mark.parametrize("case", [((128, 64, 128), (512, 0, 256), 8), ((128, 128, 128), (100, 40, 900), 16)])
def test_reference_march_on_a_brickmap_view(case):
    dims, org, factor = case
    ow = worlds.oracle_world(dims, org, factor)
    o, d = mixed_rays(dims, 30000, seed=99)
    view = ol.RefWorld(ow.top, ow.brick_bits, ow.bounds, *dims, factor)
    got = view.trace(o, d)
    ref = ow.trace(o, d, norm_mode=ol.NORM_HOST)
    assert np.array_equal(got["hit"], ref["hit"]) and np.array_equal(got["steps"], ref["steps"])
    assert np.array_equal(got["pos"].view(np.uint32), ref["pos"].view(np.uint32)) and np.array_equal(got["normal"], ref["normal"])
    # and the reference's march on the world its own GenerateLowresVoxelBuffer built
    L = ol.ref_host()
    h = L.vxref_build(worlds.dense(dims, org), *dims, factor)
    n = o.shape[0]
    hit, steps = np.zeros(n, np.uint8), np.zeros(n, np.int32)
    pos, nrm = np.zeros((n, 3), np.float32), np.zeros((n, 3), np.float32)
    L.vxref_trace(h, o, d, n, 2048, hit.ctypes.data, pos, nrm, steps.ctypes.data, 0)
    assert np.array_equal(hit, got["hit"]) and np.array_equal(steps, got["steps"]) and np.array_equal(pos.view(np.uint32), got["pos"].view(np.uint32))


@pytest.mark.parametrize("mode", [dict(cb=0, shadows=1), dict(cb=1, shadows=1), dict(cb=0, shadows=0)])
def test_row_sample_entry_points_equal_the_whole_frame(mode):
    dims, org = (128, 64, 128), (512, 0, 256)
    ow = worlds.oracle_world(dims, org, 8)
    W, H = 256, 144
    cam = ol.make_camera((64, 90, 150), (-0.45, 0.3, 0.0), 90.0)
    opts = ol.RenderOpts(mode["cb"], 2, 0, mode["shadows"], 1, ol.NORM_HOST)
    full, fsteps = ow.render(cam, ol.default_env(), opts, W, H, bgra=np.full((H, W, 4), 7, np.uint8))
    b, s = np.full((H, W, 4), 7, np.uint8), np.zeros((H, W), np.int32)
    rows = H // 2 if mode["cb"] else H
    ow.render_sample_planes(cam, ol.default_env(), opts, W, H, 0, rows, 1, b, s)
    assert np.array_equal(b, full) and np.array_equal(s, fsteps)
    st = ow.render_sample(cam, ol.default_env(), opts, W, H, 3, rows, 5)
    ro, rd = ow.frame_rays(cam, ol.default_env(), opts, W, H, 3, rows, 5)
    assert len(ro) == int(st[0] + st[2]) and len(ro) > 1000
    r = ow.trace(ro, rd, norm_mode=ol.NORM_HOST)
    assert int(r["steps"].sum()) == int(st[4]) and int(r["hit"].sum()) == int(st[1] + st[3])
