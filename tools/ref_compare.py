"""Times the REFERENCE'S OWN kernels (oracle/_ref/ref_gpu: all VoxelRT/*.cu unmodified, default flags, headless driver) and
ours on the same box, same world, same camera, same render mode: BASELINE.md O1, "the kernel to beat".

  python tools/ref_compare.py [X Y Z] [W H]        default 1024 256 1024 at 1920x1080 (BASELINE config 2)

The mode is the reference as shipped (Renderer.cu:4-5: DEBUG_VIEW, checkerboard, no sun ray), the only one its source
can run; frame 0 of both must be identical byte for byte.  Also times the batch API on the same primary rays.
Prints one JSON line."""
import json, os, subprocess, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import voxelengine_b200 as vx

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
a = [int(v) for v in sys.argv[1:]]
X, Y, Z = a[:3] if len(a) >= 3 else (1024, 256, 1024)
W, H = a[3:5] if len(a) >= 5 else (1920, 1080)
FRAMES = 8
ref_bin = os.path.join(ROOT, "oracle", "_ref", "ref_gpu")
world = vx.World.from_noise(X, Y, Z, 8)
# camera: the reference always builds the window at the origin of the noise field, which is mostly flat; look at its
# highest terrain (a 64x64 grid of vertical probe rays finds it) from 200 voxels further along +z, 40 above it
g = 64
xs, zs = np.meshgrid((np.arange(g) + 0.5) * X / g, (np.arange(g) + 0.5) * Z / g, indexing="ij")
po = np.stack([xs.ravel(), np.full(g * g, Y + 10.0), zs.ravel()], 1).astype(np.float32)
probe = vx.VoxelRaytracer3D(world, g * g)
r = probe.raytrace(po, np.tile(np.array([[0, -1, 0]], np.float32), (g * g, 1)))
hgt = np.where(r["hit"].astype(bool), r["voxel"][:, 1], -1)
k = int(np.argmax(hgt))
pos = (float(int(po[k, 0])), float(min(int(hgt[k]) + 40, Y + 32)), float(min(int(po[k, 2]) + 200, Z - 8)))
eul = (-0.45, 0.0, 0.0)
probe.close()

out = {"world": [X, Y, Z], "frame": [W, H], "camera": pos, "euler": eul, "mode": "as shipped: checkerboard, DEBUG_VIEW, no sun ray",
       "terrain_max_height": int(hgt.max()), "probe_hit_fraction": float((hgt > 0).mean())}
with tempfile.TemporaryDirectory() as tmp:
    path = os.path.join(tmp, "ref.bgra")
    t0 = time.perf_counter()
    p = subprocess.run([ref_bin, "render", X, Y, Z, 8, W, H, *pos, *eul, FRAMES, path] and
                       [str(v) for v in [ref_bin, "render", X, Y, Z, 8, W, H, *pos, *eul, FRAMES, path]], capture_output=True, text=True, timeout=3000)
    out["reference_process_s"] = time.perf_counter() - t0
    if p.returncode != 0:
        print(json.dumps({"error": p.stderr[-500:]})); sys.exit(1)
    kv = dict(l.split("=", 1) for l in p.stdout.splitlines() if "=" in l)
    ref_ms = [float(v) for k, v in kv.items() if k.startswith("ref_frame_ms")]
    out["reference_frame_ms"] = ref_ms
    out["reference_startup_ms"] = {k: float(kv[k]) for k in ("ref_voxel_generation_ms", "ref_buffer_generation_ms", "ref_upload_ms") if k in kv}
    # parity authority: the same sources built with -fmad=false; one frame
    nofma = os.path.join(ROOT, "oracle", "_ref", "ref_gpu_nofma")
    p2 = subprocess.run([str(v) for v in [nofma, "render", X, Y, Z, 8, W, H, *pos, *eul, 1, path]], capture_output=True, text=True, timeout=3000)
    if p2.returncode != 0:
        print(json.dumps({"error": p2.stderr[-500:]})); sys.exit(1)
    ref0 = np.fromfile(path, np.uint8).reshape(H, W, 4)

rt = vx.VoxelRaytracer3D(world, W * H)
cam = vx.make_camera(pos, eul, 90.0)
buf = torch.full((H, W, 4), 255, dtype=torch.uint8, device="cuda")
ours = []
first = None
for i in range(FRAMES):
    o = vx.RenderOpts(1, i, 1, 0, 1, vx.NORM_DEVICE, 0, 0, 0, 0, None)  # the reference's frame counter advances every call
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    rt.render_device(cam, vx.default_env(), o, buf.data_ptr(), W, H, stream=torch.cuda.current_stream().cuda_stream)
    e1.record(); torch.cuda.synchronize()
    ours.append(e0.elapsed_time(e1))
    if i == 0:
        first = buf.cpu().numpy().copy()
out["ours_frame_ms"] = ours
out["frame0_identical_to_nofma_build"] = bool(np.array_equal(first, ref0))
out["frame0_differing_pixels"] = int((first != ref0).any(2).sum())
out["ours_world_build_ms"] = world.info.build_ms
med = lambda v: float(np.median(v[2:]))
out["speedup_frame"] = med(ref_ms) / med(ours)
print(json.dumps(out))
