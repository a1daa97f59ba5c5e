"""numpy restatement of the reference's ambient-occlusion sample ray (Renderer.cu:121-165, sample i = 0) — the bounce
rays of BASELINE config 5.  test_reference_pins.py pins every function here against the reference's own code
(cudaNoise::randomFloat of cuda_noise.cuh:66-71; normalize / reflect / dot of helper_math.h) compiled into oracle/_ref;
test_gpu_parity.py compares the CUDA generator (vx_bounce_rays) with it bit for bit."""
import numpy as np


def host_normalize(v):
    """normalize() with CUDA's host rsqrtf: (float)(1.0 / sqrt((double)dot)), dot in float32 left to right."""
    v = v.astype(np.float32)
    d = (v[:, 0] * v[:, 0] + v[:, 1] * v[:, 1]).astype(np.float32) + (v[:, 2] * v[:, 2]).astype(np.float32)
    inv = (1.0 / np.sqrt(d.astype(np.float64))).astype(np.float32)
    return (v * inv[:, None]).astype(np.float32)


def noise_hash(seed):
    s = seed.astype(np.uint32)
    with np.errstate(over="ignore"):
        s = (s + np.uint32(0x7ed55d16)) + (s << np.uint32(12))
        s = (s ^ np.uint32(0xc761c23c)) ^ (s >> np.uint32(19))
        s = (s + np.uint32(0x165667b1)) + (s << np.uint32(5))
        s = (s + np.uint32(0xd3a2646c)) ^ (s << np.uint32(9))
        s = (s + np.uint32(0xfd7046c5)) + (s << np.uint32(3))
        s = (s ^ np.uint32(0xb55a4f09)) ^ (s >> np.uint32(16))
    return s


def random_float(seed_u32):
    """cudaNoise::randomFloat: (float)hash(seed) / (float)0xffffffff."""
    return (noise_hash(seed_u32).astype(np.float32) / np.float32(0xffffffff)).astype(np.float32)


def dot(a, b):
    return ((a[:, 0] * b[:, 0] + a[:, 1] * b[:, 1]).astype(np.float32) + a[:, 2] * b[:, 2]).astype(np.float32)


def reflect(i, n):
    """helper_math.h:1427-1430: i - 2.0f * n * dot(n, i)."""
    dn = dot(n, i)
    return (i - ((np.float32(2) * n).astype(np.float32) * dn[:, None]).astype(np.float32)).astype(np.float32)


def sample_dirs(si):
    """Direction of sample seed si (int32 array) before the hemisphere flip (Renderer.cu:131-134)."""
    with np.errstate(over="ignore"):
        r = [random_float((si * np.int32(k)).astype(np.int32).view(np.uint32)) for k in (1, 10, 100)]
    return host_normalize(np.stack([v * np.float32(2) - np.float32(1) for v in r], 1))


def sample_rays(tx, ty, W, frame_number, pos, normal):
    """(origins, dirs) of the sample rays for hits at thread coordinates (tx, ty) with shading normal `normal`."""
    with np.errstate(over="ignore"):
        si = (ty * W + tx + (frame_number + 1) * 1000).astype(np.int32)
    sd = sample_dirs(si)
    flip = dot(sd, normal) < 0
    sd = np.where(flip[:, None], reflect(sd, normal), sd)
    origins = (pos + (normal * np.float32(0.01)).astype(np.float32)).astype(np.float32)
    return origins, sd
