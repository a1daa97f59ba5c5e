"""Cached seeded worlds for the tests (dense grids come from the oracle's noise restatement)."""
import functools
import os
import zlib

import numpy as np

import oracle_lib as ol

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@functools.lru_cache(maxsize=None)
def dense(dims, origin):
    return ol.populate_dense(*dims, *origin)


@functools.lru_cache(maxsize=None)
def oracle_world(dims, origin, factor):
    return ol.OracleWorld.from_dense(dense(dims, origin), *dims, factor)


def golden_case(name):
    g = np.load(os.path.join(GOLDEN, "trace_%s.npz" % name))
    dims = tuple(int(v) for v in g["dims"])
    origin = tuple(int(v) for v in g["origin"])
    n = int(g["nrays"])
    return dict(dims=dims, origin=origin, factor=int(g["factor"]), nrays=n, seed=int(g["seed"]),
                dense_crc=int(g["dense_crc"]), top=g["top"], brick_crc=int(g["brick_crc"]), bounds_crc=int(g["bounds_crc"]),
                hit=np.unpackbits(g["hit"])[:n], steps=g["steps"].astype(np.int32), pos=g["pos"], normal=g["normal"].astype(np.float32))


GOLDEN_CASES = ["w64", "w128x64", "w64sparse", "w256f32"]


def crc(a):
    return zlib.crc32(np.ascontiguousarray(a).tobytes())
