"""Timeline of primary_rays' tiles for one rank's share of a banded frame: python tools/tile_trace.py [workload] [N] [r]
(needs VX_TILE_TRACE=1).  Prints how many warps hold a tile over the kernel's duration and the longest tiles."""
import ctypes as C, os, sys
os.environ["VX_TILE_TRACE"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, voxelengine_b200 as vx, bench
name = sys.argv[1] if len(sys.argv) > 1 else "w2-4k"
N = int(sys.argv[2]) if len(sys.argv) > 2 else 8
r = int(sys.argv[3]) if len(sys.argv) > 3 else 0
wl = bench.WORKLOADS[name]
world = vx.World.from_noise(*wl["dims"], 8)
rt = vx.VoxelRaytracer3D(world, 16)
cam, pos = bench.camera_for(vx, rt, wl["dims"], bench.camera_spec(wl))
W, H = wl["W"], wl["H"]
buf = torch.zeros(H * W * 4 + 8 * W * 4, dtype=torch.uint8, device="cuda")
opts = vx.RenderOpts(0, 0, 0, 0, 1, vx.NORM_DEVICE, 0, 0, 0, 0, None, N if N > 1 else 0, r if N > 1 else 0)  # primaries only
rt.set_timing(True)
for i in range(4):
    rt.render_device(cam, vx.default_env(), opts, buf.data_ptr(), W, H)
    ms = rt.frame_times()[0]
L = vx.lib()
L.vx_debug_tile_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]
cap = W * H // 32 + 1024
cost, start, n = np.zeros(cap, np.uint32), np.zeros(cap, np.uint32), C.c_size_t(0)
vx._check(L.vx_debug_tile_trace(rt.h, cost.ctypes.data, start.ctypes.data, cap, C.byref(n)))
n = n.value
cost, start = cost[:n].astype(np.float64) / 1965.0, (start[:n] - start[:n].min()).astype(np.uint32).astype(np.float64) / 1000.0  # both in microseconds
end = start + cost
T = end.max()
print("kernel %.3f ms by events; %d tiles; first draw -> last finish %.1f us; sum of tile times %.0f us over 5328 warps = %.1f us per warp" % (ms, n, T, cost.sum(), cost.sum() / 5328))
print("tile time us: mean %.1f p50 %.1f p90 %.1f p99 %.1f max %.1f" % (cost.mean(), np.percentile(cost, 50), np.percentile(cost, 90), np.percentile(cost, 99), cost.max()))
edges = np.linspace(0, T, 21)
busy = [float(((start < b) & (end > a)).sum()) for a, b in zip(edges[:-1], edges[1:])]
print("warps holding a tile per 5%% of the duration:", [int(v) for v in busy])
late = np.argsort(end)[-8:]
print("last tiles to finish (start us, time us):", [(round(float(start[i]), 1), round(float(cost[i]), 1)) for i in late])
