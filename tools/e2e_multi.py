"""Phases of the multi-GPU host-frame path (torchrun): render only, + copy of the bands to the shared host frame, + rank
barrier.  python -m torch.distributed.run --nproc-per-node N tools/e2e_multi.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import voxelengine_b200 as vx, bench
from voxelengine_b200.partition import FramePartition, HostFrame

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dev = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=torch.device("cuda", dev))
wl = bench.WORKLOADS["w2-8k"]
W, H = wl["W"], wl["H"]
w = vx.World.from_noise(*wl["dims"], 8, device=dev)
rt = vx.VoxelRaytracer3D(w, 16)
cam, _ = bench.camera_for(vx, rt, wl)
env = vx.default_env()
part = FramePartition(W, H, world, rank)
bc, bi = part.render_opts()
hf = HostFrame(part, dev)
local = part.alloc_frame("cuda")
stream = torch.cuda.current_stream().cuda_stream
o = vx.RenderOpts(0, 0, 0, 1, 1, vx.NORM_DEVICE, 0, 0, 0, 0, None, bc, bi)
cols = (W + 31) // 32 * 32
opts = {}
def rows(y0, h):
    oo = opts.setdefault((y0, h), vx.RenderOpts(0, 0, 0, 1, 1, vx.NORM_DEVICE, 0, y0, cols, h, None, bc, bi))
    rt.render_device(cam, env, oo, local.data_ptr(), W, H, stream=stream)
K = 20
def timed(name, fn):
    for _ in range(3): fn()
    torch.cuda.synchronize(); dist.barrier()
    t0 = time.perf_counter()
    for _ in range(K): fn()
    torch.cuda.synchronize(); dist.barrier()
    dt = (time.perf_counter() - t0) / K * 1e3
    t = torch.tensor([dt], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0: print("%-28s %.3f ms" % (name, t.item()), flush=True)
def render(): rt.render_device(cam, env, o, local.data_ptr(), W, H, stream=stream); torch.cuda.synchronize()
def render_copy(): rt.render_device(cam, env, o, local.data_ptr(), W, H, stream=stream); hf.store(local.data_ptr(), stream); torch.cuda.synchronize()
def copy_only(): hf.store(local.data_ptr(), stream); torch.cuda.synchronize()
def barrier_only(): dist.barrier()
def full(): rt.render_device(cam, env, o, local.data_ptr(), W, H, stream=stream); hf.store(local.data_ptr(), stream); hf.fence()
def split(): hf.render_and_store(rows, local.data_ptr(), 0.8); hf.fence()
def split_nobarrier(): hf.render_and_store(rows, local.data_ptr(), 0.8); torch.cuda.synchronize()
for name, fn in [("render+sync", render), ("copy only", copy_only), ("render+copy+sync", render_copy), ("barrier only", barrier_only),
                 ("render+copy+fence", full), ("split 0.8 + sync", split_nobarrier), ("split 0.8 + fence", split)]:
    timed(name, fn)
hf.close(); dist.destroy_process_group()
