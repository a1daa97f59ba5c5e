"""Kernel times of ONE rank's share of a banded frame on one GPU (what limits strong scaling of a small frame):
python tools/band_scaling.py [workload] [bands N] — renders band set r of N for a few r and prints the per-kernel times,
for the resident-block counts given in VX_BLOCKS_PER_SM_LIST (comma separated; each needs its own process, so this script
re-runs itself)."""
import os, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, root)
if os.environ.get("VX_BAND_CHILD") != "1":
    for b in os.environ.get("VX_BLOCKS_PER_SM_LIST", "0").split(","):
        env = dict(os.environ, VX_BAND_CHILD="1")
        if int(b) > 0:
            env["VX_BLOCKS_PER_SM"] = b
        print("== blocks per SM cap:", b, flush=True)
        subprocess.run([sys.executable, __file__] + sys.argv[1:], env=env)
    sys.exit(0)
import torch, voxelengine_b200 as vx, bench
name = sys.argv[1] if len(sys.argv) > 1 else "w2-4k"
N = int(sys.argv[2]) if len(sys.argv) > 2 else 8
wl = bench.WORKLOADS[name]
world = vx.World.from_noise(*wl["dims"], 8)
rt = vx.VoxelRaytracer3D(world, 16)
cam, pos = bench.camera_for(vx, rt, wl["dims"], bench.camera_spec(wl))
W, H = wl["W"], wl["H"]
buf = torch.zeros(H * W * 4 + 8 * W * 4, dtype=torch.uint8, device="cuda")
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
stream = torch.cuda.current_stream().cuda_stream
rt.set_timing(True)
for r in sorted({0, N // 2, N - 1}):
    opts = vx.RenderOpts(0, 0, 0, 1, 1, vx.NORM_DEVICE, 0, 0, 0, 0, None, N if N > 1 else 0, r if N > 1 else 0)
    acc = [0.0] * 4
    tot = 0.0
    for i in range(12):
        if os.environ.get('VX_NO_FLUSH') != '1':
            flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); rt.render_device(cam, vx.default_env(), opts, buf.data_ptr(), W, H, stream=stream); e1.record(); e1.synchronize()
        t = rt.frame_times()
        if i >= 4:
            acc = [a + b for a, b in zip(acc, t)]; tot += e0.elapsed_time(e1)
    print("bands %d/%d: " % (r, N) + " ".join("%s=%.3f" % (n, a / 8) for n, a in zip(rt.KERNELS, acc)), "sum=%.3f step=%.3f" % (sum(acc) / 8, tot / 8), flush=True)
# contiguous row blocks of the same size, for comparison with the interleaved band sets (cache locality vs balance)
if os.environ.get("VX_BAND_CONTIG", "1") == "1" and N > 1:
    rows = H // N // 8 * 8
    cols = (W + 31) // 32 * 32
    tot_p = 0.0
    for r in range(N):
        opts = vx.RenderOpts(0, 0, 0, 1, 1, vx.NORM_DEVICE, 0, r * rows, cols, rows, None, 0, 0)
        acc = [0.0] * 4
        for i in range(10):
            if os.environ.get('VX_NO_FLUSH') != '1':
                flush.fill_(1)
            rt.render_device(cam, vx.default_env(), opts, buf.data_ptr(), W, H, stream=stream)
            t = rt.frame_times()
            if i >= 4:
                acc = [a + b for a, b in zip(acc, t)]
        tot_p += acc[0] / 6
        print("rows %4d..%4d: " % (r * rows, (r + 1) * rows) + " ".join("%s=%.3f" % (n, a / 6) for n, a in zip(rt.KERNELS, acc)), "sum=%.3f" % (sum(acc) / 6), flush=True)
    print("sum of primary_rays over the %d blocks: %.3f" % (N, tot_p))
