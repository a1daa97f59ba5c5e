"""Why does a rank's 1/N share of a 4K frame take more than 1/N of the frame's time?  Offline model of primary_rays' tile
scheduling: per 8x4 tile the cost is taken as the largest step count among its rays (a warp keeps a tile until its last ray
is done) plus a constant; tiles are handed in id order to 5328 persistent warps (list scheduling).  Prints the makespan over
the ideal (sum / workers) for the whole frame and for band sets r of N.   python tools/tile_sim.py [workload] [N]"""
import heapq, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, voxelengine_b200 as vx, bench
name = sys.argv[1] if len(sys.argv) > 1 else "w2-4k"
N = int(sys.argv[2]) if len(sys.argv) > 2 else 8
wl = bench.WORKLOADS[name]
world = vx.World.from_noise(*wl["dims"], 8)
rt = vx.VoxelRaytracer3D(world, 16)
cam, pos = bench.camera_for(vx, rt, wl["dims"], bench.camera_spec(wl))
W, H = wl["W"], wl["H"]
buf = torch.zeros(H * W * 4, dtype=torch.uint8, device="cuda")
steps = torch.zeros(H * W, dtype=torch.int32, device="cuda")
rt.render_device(cam, vx.default_env(), vx.RenderOpts(0, 0, 0, 0, 1, vx.NORM_DEVICE, 0, 0, 0, 0, None), buf.data_ptr(), W, H, d_steps=steps.data_ptr())
torch.cuda.synchronize()
s = steps.cpu().numpy().reshape(H, W)
th, tw = H // 4, W // 8
t = s[:th * 4, :tw * 8].reshape(th, 4, tw, 8)
tmax, tsum = t.max(axis=(1, 3)).astype(np.float64), t.sum(axis=(1, 3)).astype(np.float64)
rows = s.mean(axis=1)
print("steps per ray by image row decile:", [round(float(rows[int(H * k / 10):int(H * (k + 1) / 10)].mean()), 1) for k in range(10)])
print("tile max steps: mean %.1f, p50 %.0f, p90 %.0f, p99 %.0f, max %.0f;  tile mean-of-rays %.1f -> lanes busy in a tile %.2f" % (
    tmax.mean(), np.percentile(tmax, 50), np.percentile(tmax, 90), np.percentile(tmax, 99), tmax.max(), tsum.mean() / 32, tsum.sum() / (32 * tmax.sum())))
def makespan(costs, workers=5328):
    h = [0.0] * workers
    heapq.heapify(h)
    for c in costs:
        heapq.heappush(h, heapq.heappop(h) + c)
    return max(h)
for const in (0.0, 20.0, 60.0):
    cost = tmax + const
    # ids: 16x8 macro tiles (four 8x4 tiles) row-major inside a rank's 8-row bands
    def order(r, n):
        out = []
        for band in range(r, H // 8, n):
            for mc in range(W // 16):
                for k in range(4):
                    out.append(cost[band * 2 + (k >> 1), mc * 2 + (k & 1)])
        return out
    full = order(0, 1)
    line = "const %3.0f: whole frame makespan/ideal %.3f" % (const, makespan(full) / (sum(full) / 5328))
    for r in (0, N // 2, N - 1):
        c = order(r, N)
        line += " | band set %d/%d: %.3f (ideal share of the frame %.3f, modelled %.3f)" % (r, N, makespan(c) / (sum(c) / 5328), 1.0 / N, makespan(c) / makespan(full))
    print(line)
