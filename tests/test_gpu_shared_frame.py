"""Shared frame of the multi-GPU partition (SharedFrame, vx_shared_*): two processes map one frame buffer over CUDA IPC,
each renders its bands into it through the C ABI, and the result is byte-identical to one process rendering the whole
frame.  Both processes use cuda:0 (the test box has one GPU); on a multi-GPU node the same calls go over NVLink."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
W, H = 328, 203
DIMS = (128, 64, 128)


def _worker(rank, world_size, port, out_path):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    import voxelengine_b200 as vx
    from voxelengine_b200.partition import FramePartition, SharedFrame

    torch.cuda.set_device(0)
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world_size)
    world = vx.World.from_noise(*DIMS, 8, device=0)
    rt = vx.VoxelRaytracer3D(world, 1)
    cam = vx.make_camera((64, 90, 150), (-0.45, 0.3, 0.0), 90.0)
    part = FramePartition(W, H, world_size, rank)
    sf = SharedFrame(part, 0)
    bc, bi = part.render_opts()
    assert not sf.remote  # every rank of this test is on cuda:0
    sf.attach(rt)
    if rank == world_size - 1:
        rt.set_fused_shade(1)  # what a rank on another GPU does by itself (the frame is peer memory there)
    for frame_no in range(2):
        rt.render_device(cam, vx.default_env(), vx.RenderOpts(0, frame_no, 0, 1, 1, vx.NORM_DEVICE, 0, 0, 0, 0, None, bc, bi), sf.ptr, W, H)
        sf.fence()
    if rank == 0:
        host = torch.empty(H * W * 4, dtype=torch.uint8).pin_memory()
        sf.to_host(host, H * W * 4)
        whole = torch.zeros(H * W * 4, dtype=torch.uint8, device="cuda")
        rt.render_device(cam, vx.default_env(), vx.RenderOpts(0, 1, 0, 1, 1, vx.NORM_DEVICE, 0, 0, 0, 0, None), whole.data_ptr(), W, H)
        torch.cuda.synchronize()
        np.save(out_path, np.stack([host.numpy(), whole.cpu().numpy()]))
    sf.close()
    # the same frame for a consumer on the host: every rank copies its own bands into one shared, pinned host frame
    from voxelengine_b200.partition import HostFrame
    hf = HostFrame(part, 0)
    mine = torch.full((part.frame_bytes(),), 9, dtype=torch.uint8, device="cuda")
    rt.set_fused_shade(0)
    rt.render_device(cam, vx.default_env(), vx.RenderOpts(0, 1, 0, 1, 1, vx.NORM_DEVICE, 0, 0, 0, 0, None, bc, bi), mine.data_ptr(), W, H)
    hf.store(mine.data_ptr())
    hf.fence()
    if rank == 0:
        np.save(out_path + ".host.npy", np.array(hf.array).reshape(-1))
    hf.fence()
    # the overlapped variant: two row ranges, the first one's copy on a second stream
    hf.array[:] = 3
    hf.fence()
    cols = (W + 31) // 32 * 32

    def render_rows(y0, rows):
        rt.render_device(cam, vx.default_env(), vx.RenderOpts(0, 1, 0, 1, 1, vx.NORM_DEVICE, 0, y0, cols, rows, None, bc, bi), mine.data_ptr(), W, H)
    hf.render_and_store(render_rows, mine.data_ptr(), 0.6)
    hf.fence()
    if rank == 0:
        np.save(out_path + ".host2.npy", np.array(hf.array).reshape(-1))
    hf.fence()
    hf.close()
    rt.close()
    dist.destroy_process_group()


@pytest.mark.parametrize("world_size", [2, 3])
def test_ranks_render_into_rank0_frame(tmp_path, world_size):
    import torch.multiprocessing as mp
    out = str(tmp_path / "frames.npy")
    port = 29650 + world_size
    mp.spawn(_worker, args=(world_size, port, out), nprocs=world_size, join=True)
    got, want = np.load(out)
    assert want.any()
    assert np.array_equal(got, want)
    assert np.array_equal(np.load(out + ".host.npy"), want)  # H = 203: the last band is ragged
    assert np.array_equal(np.load(out + ".host2.npy"), want)


def test_shared_handle_validation():
    import voxelengine_b200 as vx
    with pytest.raises(ValueError):
        vx.shared_open(b"short")
    ptr, handle = vx.shared_alloc(4096, 0)
    assert ptr and len(handle) == vx.SHARED_HANDLE_BYTES
    vx.shared_free(ptr, 0)
