"""Per-kernel frame times for every tuning variant: python tools/sweep_kernels.py [workload]"""
import glob, os, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
code = '''
import sys, os
sys.path.insert(0, %r)
import torch, voxelengine_b200 as vx, bench
wl = bench.WORKLOADS[%r]
world = vx.World.from_noise(*wl["dims"], 8)
rt = vx.VoxelRaytracer3D(world, 16)
cam, pos = bench.camera_for(vx, rt, wl)
W, H = wl["W"], wl["H"]
buf = torch.zeros(H * W * 4, dtype=torch.uint8, device="cuda")
opts = vx.RenderOpts(0, 0, 0, 1, 1, vx.NORM_DEVICE, 0, 0, 0, 0, None)
rt.set_timing(True)
acc = [0.0] * 4
for i in range(9):
    rt.render_device(cam, vx.default_env(), opts, buf.data_ptr(), W, H, stream=torch.cuda.current_stream().cuda_stream)
    t = rt.frame_times()
    if i >= 3: acc = [a + b for a, b in zip(acc, t)]
print(" ".join("%%s=%%.3f" %% (n, a / 6) for n, a in zip(rt.KERNELS, acc)), "sum=%%.3f" %% (sum(acc) / 6))
''' % (root, sys.argv[1] if len(sys.argv) > 1 else "w2-4k")
for so in sorted(glob.glob(os.path.join(root, "voxelengine_b200/lib/variants/libvxrt_*.so"))):
    out = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, VXRT_LIB=so), capture_output=True, text=True)
    print(os.path.basename(so), out.stdout.strip().splitlines()[-1] if out.stdout.strip() else out.stderr[-300:])
