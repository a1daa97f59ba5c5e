"""Editing a built world in place and storing it (vx_world_edit / vx_world_save / vx_world_load).
The criterion is the reference's own build: after any sequence of edits the device brickmap must equal, word for word,
what GenerateLowresVoxelBuffer (the oracle's restatement of it) builds from the dense grid with the same bits flipped
through the reference's tiled index (GetSampleIndex, VolumeRaytracer.cuh:111-131), and rays must see the edited world."""
import numpy as np
import pytest

import oracle_lib as ol
import worlds
from raysets import mixed_rays

pytestmark = pytest.mark.gpu


def tiled_index(x, y, z, X, Y):
    tW, tH = X // 8, Y // 8
    tile = (x // 8) + (y // 8) * tW + (z // 8) * tW * tH
    return tile.astype(np.int64) * 512 + (x % 8) + 8 * (y % 8) + 64 * (z % 8)


def flip(bits, vox, dims, solid):
    idx = tiled_index(vox[:, 0].astype(np.int64), vox[:, 1].astype(np.int64), vox[:, 2].astype(np.int64), dims[0], dims[1])
    words, masks = idx >> 5, (np.uint32(1) << (idx & 31).astype(np.uint32))
    if solid:
        np.bitwise_or.at(bits, words, masks)
    else:
        np.bitwise_and.at(bits, words, ~masks)


def assert_same_world(w, bits, dims, factor):
    ow = ol.OracleWorld.from_dense(bits, *dims, factor)
    top, bw, b6 = w.download()
    assert np.array_equal(top, ow.top)
    assert np.array_equal(bw, ow.brick_bits)
    assert np.array_equal(b6, ow.bounds)
    info = w.info
    assert info.solid_voxels == int(np.unpackbits(bits.view(np.uint8)).sum())
    assert info.occupied_bricks == int(np.unpackbits(top.view(np.uint8)).sum())
    return ow


@pytest.mark.parametrize("case", [((64, 64, 64), (0, 0, 0), 8), ((128, 64, 128), (512, 0, 256), 8), ((128, 128, 128), (100, 40, 900), 16),
                                  ((256, 256, 256), (0, 0, 0), 32)])
def test_edits_equal_a_rebuild_from_the_edited_dense_grid(vx, case):
    dims, org, factor = case
    rng = np.random.default_rng(11)
    bits = worlds.dense(dims, org).copy()
    w = vx.World.from_dense_bits(bits, *dims, factor)
    n = 4000
    # 1. scattered new voxels (mostly in empty bricks: new pool slots), with duplicates and out-of-range coordinates
    vox = rng.integers(0, [dims[0], dims[1], dims[2]], size=(n, 3)).astype(np.uint32)
    vox[:50] = vox[50:100]
    junk = np.array([[dims[0], 0, 0], [0, dims[1] + 7, 0], [2 ** 31, 1, 1]], np.uint32)
    w.edit(np.vstack([vox, junk]), solid=True)
    flip(bits, vox, dims, True)
    assert_same_world(w, bits, dims, factor)
    # 2. dig: clear a box that empties whole bricks and shaves others, plus clearing voxels that were never set
    lo = np.array([dims[0] // 4, 0, dims[2] // 4]); hi = np.minimum(lo + np.array([factor * 2 + 3, dims[1], factor + 5]), dims)
    gx, gy, gz = np.mgrid[lo[0]:hi[0], lo[1]:hi[1], lo[2]:hi[2]]
    box = np.stack([gx.ravel(), gy.ravel(), gz.ravel()], 1).astype(np.uint32)
    w.edit(box, solid=False)
    flip(bits, box, dims, False)
    assert_same_world(w, bits, dims, factor)
    # 3. build again inside the emptied bricks (their slots are reused), and a second scatter
    w.edit(box[::7], solid=True)
    flip(bits, box[::7], dims, True)
    vox2 = rng.integers(0, [dims[0], dims[1], dims[2]], size=(n, 3)).astype(np.uint32)
    w.edit(vox2, solid=False)
    flip(bits, vox2, dims, False)
    ow = assert_same_world(w, bits, dims, factor)
    # rays see the edited world, all planes
    o, d = mixed_rays(dims, 50000, seed=5)
    rt = vx.VoxelRaytracer3D(w, o.shape[0])
    got = rt.raytrace(o, d, norm_mode=vx.NORM_HOST)
    ref = ow.trace(o, d, norm_mode=ol.NORM_HOST)
    for k in ("hit", "steps", "voxel", "brick"):
        assert np.array_equal(got[k], ref[k]), k
    assert np.array_equal(got["pos"].view(np.uint32), ref["pos"].view(np.uint32))
    rt.close(); w.close()


def test_edit_an_empty_world_and_empty_it_again(vx):
    dims = (64, 64, 64)
    bits = np.zeros(64 * 64 * 64 // 32, np.uint32)
    w = vx.World.from_dense_bits(bits, *dims, 8)
    assert w.info.occupied_bricks == 0
    vox = np.array([[0, 0, 0], [63, 63, 63], [8, 8, 8], [8, 8, 9]], np.uint32)
    w.edit(vox)
    flip(bits, vox, dims, True)
    assert_same_world(w, bits, dims, 8)
    assert w.info.occupied_bricks == 3 and w.info.solid_voxels == 4
    w.edit(vox, solid=False)
    flip(bits, vox, dims, False)
    assert_same_world(w, bits, dims, 8)
    assert w.info.occupied_bricks == 0 and w.info.solid_voxels == 0
    w.edit(np.zeros((0, 3), np.uint32))
    w.close()


def test_save_and_load_round_trip(vx, tmp_path):
    dims, org = (128, 64, 128), (512, 0, 256)
    w = vx.World.from_noise(*dims, 8, origin=org)
    w.edit(np.array([[3, 60, 3], [100, 63, 77]], np.uint32))
    path = str(tmp_path / "world.vxb")
    w.save(path)
    w2 = vx.World.load(path)
    for a, b in zip(w.download(), w2.download()):
        assert np.array_equal(a, b)
    i1, i2 = w.info, w2.info
    assert (i1.occupied_bricks, i1.solid_voxels, tuple(i1.dims), i1.factor) == (i2.occupied_bricks, i2.solid_voxels, tuple(i2.dims), i2.factor)
    o, d = mixed_rays(dims, 20000, seed=9)
    r1 = vx.VoxelRaytracer3D(w, o.shape[0]); r2 = vx.VoxelRaytracer3D(w2, o.shape[0])
    a, b = r1.raytrace(o, d), r2.raytrace(o, d)
    for k in ("hit", "steps", "voxel", "brick"):
        assert np.array_equal(a[k], b[k])
    # a loaded world can be edited (the pool grows from its saved size)
    more = np.array([[1, 1, 1], [120, 62, 120], [64, 63, 64]], np.uint32)
    w.edit(more); w2.edit(more)
    for p, q in zip(w.download(), w2.download()):
        assert np.array_equal(p, q)
    r1.close(); r2.close(); w.close(); w2.close()
    # damaged files are refused
    raw = bytearray(open(path, "rb").read())
    raw[len(raw) // 2] ^= 0x40
    bad = tmp_path / "bad.vxb"
    bad.write_bytes(bytes(raw))
    with pytest.raises(vx.VxError):
        vx.World.load(str(bad))
    (tmp_path / "short.vxb").write_bytes(bytes(raw[: len(raw) // 3]))
    with pytest.raises(vx.VxError):
        vx.World.load(str(tmp_path / "short.vxb"))
    (tmp_path / "junk.vxb").write_bytes(b"not a world")
    with pytest.raises(vx.VxError):
        vx.World.load(str(tmp_path / "junk.vxb"))
