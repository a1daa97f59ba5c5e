"""Scratch timing probe (not the bench): world build and frame times at the BASELINE configs."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import voxelengine_b200 as vx

def column_top(rt, x, z, Y):
    r = rt.raytrace(np.array([[x + 0.5, Y + 10.0, z + 0.5]], np.float32), np.array([[0, -1, 0]], np.float32))
    return int(r["voxel"][0, 1]) if r["hit"][0] else -1

def frame(rt, cam, W, H, shadows, stats=False, reps=5, checkerboard=0, debug=0):
    buf = torch.zeros(H * W * 4, dtype=torch.uint8, device="cuda")
    st = torch.zeros(8, dtype=torch.int64, device="cuda")
    opts = vx.RenderOpts(checkerboard, 0, debug, shadows, 1, vx.NORM_DEVICE, 0, 0, 0, 0, st.data_ptr() if stats else None)
    env = vx.default_env()
    s = torch.cuda.current_stream().cuda_stream
    rt.render_device(cam, env, opts, buf.data_ptr(), W, H, stream=s)
    torch.cuda.synchronize()
    if stats:
        return st.cpu().numpy()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); rt.render_device(cam, env, opts, buf.data_ptr(), W, H, stream=s); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts), float(np.median(ts))

print("L2-resident read GB/s (32 MiB):", vx.measure_read_bandwidth(32 << 20, 50))
print("HBM read GB/s (4 GiB):", vx.measure_read_bandwidth(4 << 30, 2))
for name, dims, W, H, xz, eul, extra in [("W1", (1024, 256, 1024), 1920, 1080, (512, 960), (-0.45, 0, 0), 0),
                                         ("W2a", (8192, 512, 8192), 3840, 2160, (4096, 7936), (-0.5, 0, 0), 128),
                                         ("W2b", (8192, 512, 8192), 3840, 2160, (4096, 4096), (-0.05, 0.8, 0), 0)]:
    if name != "W2b":
        t0 = time.time(); world = vx.World.from_noise(*dims, 8); t1 = time.time()
        i = world.info
        print(name, "build ms", i.build_ms, "wall", t1 - t0, "occupied", i.occupied_bricks, "of", i.bricks, "solid", i.solid_voxels, "pool MB", i.bytes_pool / 1e6)
        rt = vx.VoxelRaytracer3D(world, 16)
    top = column_top(rt, xz[0], xz[1], dims[1])
    camy = (top + 16 if top < dims[1] - 1 else dims[1] + 32) + extra
    cam = vx.make_camera((xz[0], camy, xz[1]), eul)
    print(name, "column top", top, "cam y", camy)
    for sh in (0, 1):
        st = frame(rt, cam, W, H, sh, stats=True)
        t = frame(rt, cam, W, H, sh)
        rays = st[0] + st[2]
        byts = 4 * st[5] + 8 * st[6] + 64 * st[7] + 4 * W * H
        print(name, "shadows", sh, "ms min/med", t, "stats", st.tolist(), "Mrays/s", rays / t[0] / 1e3, "steps/ray", st[4] / rays, "alg GB/s", byts / t[0] / 1e6)
