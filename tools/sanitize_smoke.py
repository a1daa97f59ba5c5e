"""Small end-to-end pass for compute-sanitizer: world build, batch rays, shaded frame (all four kernels), bounce rays.
  compute-sanitizer --tool memcheck python tools/sanitize_smoke.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import voxelengine_b200 as vx

w = vx.World.from_noise(128, 64, 128, 8, origin=(512, 0, 256))
rt = vx.VoxelRaytracer3D(w, 20000)
rng = np.random.default_rng(3)
o = (rng.uniform(-0.3, 1.3, (20000, 3)) * np.array([128, 64, 128])).astype(np.float32)
d = rng.normal(size=(20000, 3)).astype(np.float32)
r = rt.raytrace(o, d, counters=True)
W, H = 160, 96
cam = vx.make_camera((64, 90, 150), (-0.45, 0.3, 0.0), 90.0)
for cb, dbg, sh in ((0, 0, 1), (1, 1, 0), (1, 0, 1)):
    opts = vx.RenderOpts(cb, 1, dbg, sh, 1, vx.NORM_DEVICE, 0, 0, 0, 0, None)
    img = rt.render_host(cam, vx.default_env(), opts, W, H)
big = rt.render_host(cam, vx.default_env(), vx.RenderOpts(0, 0, 0, 1, 1, vx.NORM_DEVICE, 0, 0, 0, 0, None), 192, 272)  # sliced path
d_o = torch.zeros(W * H * 3, dtype=torch.float32, device="cuda"); d_d = torch.zeros_like(d_o)
d_p = torch.zeros(W * H, dtype=torch.int32, device="cuda"); d_n = torch.zeros(1, dtype=torch.int64, device="cuda")
rt.bounce_rays(cam, vx.RenderOpts(0, 0, 0, 1, 1, vx.NORM_DEVICE, 0, 0, 0, 0, None), W, H, d_o.data_ptr(), d_d.data_ptr(), d_p.data_ptr(), d_n.data_ptr())
torch.cuda.synchronize()
print("ok: hits %d, frame mean %.1f, bounce rays %d" % (int(r["hit"].sum()), float(img.mean()), int(d_n.item())))
rt.close(); w.close()
