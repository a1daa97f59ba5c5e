"""Generates tests/golden/*.npz from the REFERENCE ITSELF compiled for the host
(oracle/_ref/libvxref_host.so + libvxref_noise.so, built by oracle/Makefile from /root/reference).
Run in the build container only:  python tests/golden/gen_golden.py
Inputs are regenerated from seeds at test time (tests/raysets.py); the files hold the reference's outputs:
  trace_*.npz : hit, steps, pos (bit pattern), normal of GPUDDA::Raytrace for seeded rays on a
                seeded noise window; plus top/brick/bounds CRCs of GenerateLowresVoxelBuffer
  noise.npz   : cudaNoise::repeaterPerlin values (x86 float->uint semantics) at seeded points
"""
import os
import sys
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_lib as ol  # noqa: E402
from raysets import mixed_rays  # noqa: E402

CASES = [
    # name, dims, window origin, factor, rays, seed
    ("w64", (64, 64, 64), (0, 0, 0), 8, 6000, 11),
    ("w128x64", (128, 64, 128), (512, 0, 256), 8, 6000, 12),
    ("w64sparse", (64, 64, 64), (100, 40, 900), 8, 6000, 13),
    ("w256f32", (256, 256, 256), (0, 0, 0), 32, 4000, 14),
]


def main():
    R, N = ol.ref_host(), ol.ref_noise()
    assert R is not None and N is not None, "build oracle/_ref first (make -C oracle all)"
    for name, dims, org, factor, nr, seed in CASES:
        X, Y, Z = dims
        # dense grid through the oracle's populate (pinned separately by noise.npz and on the GPU box)
        bits = ol.populate_dense(X, Y, Z, *org)
        rw = R.vxref_build(bits, X, Y, Z, factor)
        nb = (X // factor) * (Y // factor) * (Z // factor)
        top = np.zeros((nb + 31) // 32, np.uint32)
        bw = np.zeros(nb * factor ** 3 // 32, np.uint32)
        b6 = np.zeros(nb * 6, np.float32)
        R.vxref_export(rw, top, bw, b6)
        o, d = mixed_rays(dims, nr, seed)
        hit = np.zeros(nr, np.uint8)
        pos = np.zeros((nr, 3), np.float32)
        nrm = np.zeros((nr, 3), np.float32)
        st = np.zeros(nr, np.int32)
        R.vxref_trace(rw, o, d, nr, 2048, hit.ctypes.data, pos, nrm, st.ctypes.data, 0)
        np.savez_compressed(os.path.join(HERE, "trace_%s.npz" % name), dims=np.array(dims), origin=np.array(org),
                            factor=factor, nrays=nr, seed=seed, dense_crc=zlib.crc32(bits.tobytes()),
                            top=top, brick_crc=zlib.crc32(bw.tobytes()), bounds_crc=zlib.crc32(b6.tobytes()),
                            hit=np.packbits(hit), steps=st.astype(np.int16), pos=pos.view(np.uint32), normal=nrm.astype(np.int8))
        print(name, "hits", int(hit.sum()), "mean steps", st.mean())
    rng = np.random.default_rng(5)
    pts = np.stack([rng.integers(0, 8192, 4000), rng.integers(0, 512, 4000), rng.integers(0, 8192, 4000)], 1)
    vals = np.array([N.vxref_terrain_noise(np.float32(x) * np.float32(0.005), np.float32(y) * np.float32(0.005),
                                           np.float32(z) * np.float32(0.005)) for x, y, z in pts], np.float32)
    hashes = np.array([N.vxref_hash(int(s)) for s in range(0, 4000 * 7919, 7919)], np.uint32)
    np.savez_compressed(os.path.join(HERE, "noise.npz"), pts=pts.astype(np.int32), vals=vals.view(np.uint32), hashes=hashes)
    print("noise ok")


if __name__ == "__main__":
    main()
