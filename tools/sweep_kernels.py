"""Per-kernel frame times and frame CRCs for the built library and every tuning variant (tools/build_variants.sh):
python tools/sweep_kernels.py [workload] [names...]   (names: substrings of variant file names; default all)
One line per library and camera: kernel times in ms (mean of 6 frames after 3 warm-up frames, CUDA events between the
launches, L2 flushed before every frame), the frame's CRC32 (must not change between variants) and the bounce-ray pass."""
import glob, os, subprocess, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
code = '''
import sys, os, zlib
sys.path.insert(0, %r)
import torch, voxelengine_b200 as vx, bench
wl = bench.WORKLOADS[%r]
world = vx.World.from_noise(*wl["dims"], 8)
rt = vx.VoxelRaytracer3D(world, 16)
W, H = wl["W"], wl["H"]
buf = torch.zeros(H * W * 4, dtype=torch.uint8, device="cuda")
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
opts = vx.RenderOpts(0, 0, 0, 1, 1, vx.NORM_DEVICE, 0, 0, 0, 0, None)
stream = torch.cuda.current_stream().cuda_stream
rt.set_timing(True)
for name in (["overview", "grazing", "diagonal"] if isinstance(wl["camera"], str) else ["camera"]):
    cam, pos = bench.camera_for(vx, rt, wl["dims"], bench.CAMERAS[name] if name in bench.CAMERAS else wl["camera"])
    acc = [0.0] * 4
    for i in range(9):
        flush.fill_(1)
        rt.render_device(cam, vx.default_env(), opts, buf.data_ptr(), W, H, stream=stream)
        t = rt.frame_times()
        if i >= 3: acc = [a + b for a, b in zip(acc, t)]
    crc = zlib.crc32(buf.cpu().numpy().tobytes())
    if hasattr(vx.lib(), "vx_debug_fine_stats"):
        import ctypes
        st = (ctypes.c_uint64 * 8)()
        vx.lib().vx_debug_fine_stats(st, 0)
        print("RESULT %%-8s fine stats (9 frames): started %%d, out of range %%d, two far faces %%d, far face not first %%d, test inconclusive %%d, far-face starts %%d" %% (name, *list(st)[:6]))
    print("RESULT %%-8s " %% name + " ".join("%%s=%%.3f" %% (n, a / 6) for n, a in zip(rt.KERNELS, acc)), "sum=%%.3f crc=%%08x" %% (sum(acc) / 6, crc))
    if name == "overview" and os.environ.get("VX_SWEEP_BOUNCE", "1") != "0":
        n_max = W * H
        d_o = torch.zeros(n_max * 3, dtype=torch.float32, device="cuda"); d_d = torch.zeros(n_max * 3, dtype=torch.float32, device="cuda")
        d_p = torch.zeros(n_max, dtype=torch.int32, device="cuda"); d_n = torch.zeros(1, dtype=torch.int64, device="cuda")
        rt.set_timing(False)
        rt.bounce_rays(cam, opts, W, H, d_o.data_ptr(), d_d.data_ptr(), d_p.data_ptr(), d_n.data_ptr(), stream=stream)
        torch.cuda.synchronize()
        n = int(d_n.item())
        hit = torch.zeros(n, dtype=torch.uint8, device="cuda"); steps = torch.zeros(n, dtype=torch.int32, device="cuda")
        planes = dict(hit=hit.data_ptr(), steps=steps.data_ptr())
        ms = []
        for i in range(9):
            flush.fill_(1)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); rt.raytrace_device(d_o.data_ptr(), d_d.data_ptr(), n, planes, stream=stream); e1.record(); torch.cuda.synchronize()
            if i >= 3: ms.append(e0.elapsed_time(e1))
        print("RESULT bounce   rays=%%d ms=%%.3f steps=%%d hits=%%d" %% (n, sum(ms) / len(ms), int(steps.sum().item()), int(hit.sum().item())))
        rt.set_timing(True)
''' % (root, sys.argv[1] if len(sys.argv) > 1 else "w2-4k")
want = sys.argv[2:]
libs = [os.path.join(root, "voxelengine_b200/lib/libvxrt.so")] + sorted(glob.glob(os.path.join(root, "voxelengine_b200/lib/variants/libvxrt_*.so")))
for so in libs:
    if want and not any(w in os.path.basename(so) for w in want) and "variants" in so:
        continue
    out = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, VXRT_LIB=so), capture_output=True, text=True)
    print("==", os.path.relpath(so, root), flush=True)
    res = [l[7:] for l in out.stdout.splitlines() if l.startswith("RESULT ")]
    print("\n".join(res) if res else out.stderr[-600:], flush=True)
