import os, sys, time, ctypes as C
sys.path.insert(0, os.getcwd())
import torch, voxelengine_b200 as vx, bench
wl = bench.WORKLOADS["w2-4k"]
world = vx.World.from_noise(*wl["dims"], 8)
rt = vx.VoxelRaytracer3D(world, 16)
cam, pos = bench.camera_for(vx, rt, wl)
W, H = wl["W"], wl["H"]
host = torch.empty(H * W * 4, dtype=torch.uint8).pin_memory()
o = vx.RenderOpts(0, 0, 0, 1, 1, vx.NORM_DEVICE, 0, 0, 0, 0, None)
env = vx.default_env()
for _ in range(5):
    vx._check(vx.lib().vx_render_host(rt.h, C.byref(cam), C.byref(env), C.byref(o), host.data_ptr(), W, H))
t0 = time.perf_counter()
K = 40
for _ in range(K):
    vx._check(vx.lib().vx_render_host(rt.h, C.byref(cam), C.byref(env), C.byref(o), host.data_ptr(), W, H))
dt = (time.perf_counter() - t0) / K
print("slices", os.environ.get("VX_HOST_SLICES"), "split", os.environ.get("VX_HOST_SPLIT"), "ms %.3f" % (dt * 1e3))
