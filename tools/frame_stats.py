"""Per-frame ray/work counters of a bench workload: python tools/frame_stats.py [workload]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import voxelengine_b200 as vx
import bench

name = sys.argv[1] if len(sys.argv) > 1 else "w2-4k"
wl = bench.WORKLOADS[name]
world = vx.World.from_noise(*wl["dims"], 8)
rt = vx.VoxelRaytracer3D(world, 16)
cam, pos = bench.camera_for(vx, rt, wl["dims"], bench.CAMERAS[os.environ["VX_CAMERA"]] if os.environ.get("VX_CAMERA") else bench.camera_spec(wl))
W, H = wl["W"], wl["H"]
buf = torch.zeros(H * W * 4, dtype=torch.uint8, device="cuda")
for shadows in (0, 1):
    stats = torch.zeros(8, dtype=torch.int64, device="cuda")
    opts = vx.RenderOpts(0, 0, 0, shadows, 1, vx.NORM_DEVICE, 0, 0, 0, 0, stats.data_ptr())
    rt.render_device(cam, vx.default_env(), opts, buf.data_ptr(), W, H, stream=torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    s = stats.tolist()
    print("shadows=%d primary rays %d hits %d | shadow rays %d hits %d | steps %d C %d K %d E %d" % (shadows, *s))
