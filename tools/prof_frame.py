"""Renders a few frames of a bench workload (for ncu captures): python tools/prof_frame.py [workload] [frames]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import voxelengine_b200 as vx
import bench

name = sys.argv[1] if len(sys.argv) > 1 else "w2-4k"
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 3
wl = bench.WORKLOADS[name]
world = vx.World.from_noise(*wl["dims"], 8)
rt = vx.VoxelRaytracer3D(world, 16)
cam, pos = bench.camera_for(vx, rt, wl["dims"], bench.CAMERAS[os.environ["VX_CAMERA"]] if os.environ.get("VX_CAMERA") else bench.camera_spec(wl))
W, H = wl["W"], wl["H"]
buf = torch.zeros(H * W * 4, dtype=torch.uint8, device="cuda")
opts = vx.RenderOpts(0, 0, 0, 1, 1, vx.NORM_DEVICE, 0, 0, 0, 0, None)
ts = []
for _ in range(frames):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); rt.render_device(cam, vx.default_env(), opts, buf.data_ptr(), W, H, stream=torch.cuda.current_stream().cuda_stream); e1.record()
    torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
import zlib
crc = zlib.crc32(buf.cpu().numpy().tobytes())
times = ""
if os.environ.get("VX_KERNEL_TIMES"):  # per-kernel device times of a few more frames (events between the launches)
    rt.set_timing(True)
    acc = [0.0] * 4
    for i in range(6):
        rt.render_device(cam, vx.default_env(), opts, buf.data_ptr(), W, H, stream=torch.cuda.current_stream().cuda_stream)
        acc = [a + b for a, b in zip(acc, rt.frame_times())]
    times = " ".join("%s=%.3f" % (n, a / 6) for n, a in zip(rt.KERNELS, acc))
print(name, pos, "frame ms:", ["%.3f" % t for t in ts], "crc32=%08x" % crc, times)
