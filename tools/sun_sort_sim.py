"""How much lane time the sun rays of a frame could gain from being re-ordered before shadow_rays draws them 32 at a time.
python tools/sun_sort_sim.py [workload] [camera]

Per pixel: sun-ray steps = step plane of the shadowed frame - step plane of the same frame without shadows.  The hits are taken
in the order the frame queue holds them (8x4 tiles, four to a 16x8 macro tile, row-major), cut into warps of 32, and a warp's
cost is its longest ray (no refill: an upper bound of what divergence costs; the kernel refills in groups of 8).  Orders
compared: queue order; sorted by sun-ray steps inside windows of 256 / 1024 / 4096 entries (a perfect predictor); sorted by
the height of the hit inside the same windows (the predictor shadow_setup could afford)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import voxelengine_b200 as vx
import bench

name = sys.argv[1] if len(sys.argv) > 1 else "w2-4k"
wl = bench.WORKLOADS[name]
world = vx.World.from_noise(*wl["dims"], 8)
rt = vx.VoxelRaytracer3D(world, 16)
camname = sys.argv[2] if len(sys.argv) > 2 else wl["camera"]
cam, pos = bench.camera_for(vx, rt, wl["dims"], bench.CAMERAS[camname])
W, H = wl["W"], wl["H"]
buf = torch.zeros(H * W * 4, dtype=torch.uint8, device="cuda")
planes = []
for shadows in (1, 0):
    st = torch.full((H, W), -1, dtype=torch.int32, device="cuda")
    rt.render_device(cam, vx.default_env(), vx.RenderOpts(0, 0, 0, shadows, 0, vx.NORM_DEVICE, 0, 0, 0, 0, None), buf.data_ptr(), W, H, d_steps=st.data_ptr())
    torch.cuda.synchronize()
    planes.append(st.cpu().numpy().astype(np.int64))
sun = planes[0] - planes[1]
# hit mask and hit height from the bounce-ray generator (origin = hit position + normal * 0.01)
n_max = W * H
d_o = torch.zeros(n_max * 3, dtype=torch.float32, device="cuda"); d_d = torch.zeros(n_max * 3, dtype=torch.float32, device="cuda")
d_p = torch.zeros(n_max, dtype=torch.int32, device="cuda"); d_n = torch.zeros(1, dtype=torch.int64, device="cuda")
rt.bounce_rays(cam, vx.RenderOpts(0, 0, 0, 1, 0, vx.NORM_DEVICE, 0, 0, 0, 0, None), W, H, d_o.data_ptr(), d_d.data_ptr(), d_p.data_ptr(), d_n.data_ptr())
torch.cuda.synchronize()
n = int(d_n.item())
pix = d_p[:n].cpu().numpy().astype(np.int64)
hy = d_o[:3 * n].cpu().numpy().reshape(n, 3)[:, 1]
height = np.full(H * W, np.nan); height[pix] = hy
hit = np.zeros(H * W, bool); hit[pix] = True
# queue order: macro tiles 16x8 row-major, inside: four 8x4 tiles, inside: row-major
ys, xs = np.mgrid[0:H, 0:W]
mrow, mcol = ys // 8, xs // 16
k = ((ys % 8) // 4) * 2 + ((xs % 16) // 8)
j = (ys % 4) * 8 + (xs % 8)
key = ((mrow * ((W + 15) // 16) + mcol) * 4 + k) * 32 + j
order = np.argsort(key.ravel(), kind="stable")
order = order[hit.ravel()[order]]
s = sun.ravel()[order].astype(np.float64)
hgt = height[order]
print("sun rays %d, steps mean %.1f, median %.0f, p90 %.0f, max %.0f; rays with <= 4 steps: %.1f %%" % (
    s.size, s.mean(), np.median(s), np.percentile(s, 90), s.max(), 100.0 * (s <= 4).mean()))


def warp_cost(v):
    m = v.size // 32 * 32
    w = v[:m].reshape(-1, 32)
    return w.max(1).sum() * 32, w.sum()


def report(label, v):
    c, u = warp_cost(v)
    print("%-44s lane use %.3f   warp-steps %.3e" % (label, u / c, c / 32))


report("queue order", s)
for win in (256, 1024, 4096, 65536):
    m = s.size // win * win
    a = s[:m].reshape(-1, win); h = hgt[:m].reshape(-1, win)
    report("sorted by steps, windows of %d" % win, np.sort(a, 1).ravel())
    idx = np.argsort(h, 1, kind="stable")
    report("sorted by hit height, windows of %d" % win, np.take_along_axis(a, idx, 1).ravel())
    q = np.floor(h / 16.0)
    idx = np.argsort(q, 1, kind="stable")
    report("sorted by hit height / 16, windows of %d" % win, np.take_along_axis(a, idx, 1).ravel())
