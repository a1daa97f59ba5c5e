"""Times the REFERENCE'S OWN kernels (oracle/_ref: all VoxelRT/*.cu compiled from /root/reference by oracle/Makefile, headless
driver) and ours on the same box, same world, same camera, same render mode: BASELINE.md O1, "the kernel to beat".

  python tools/ref_compare.py [X Y Z] [W H]        default 1024 256 1024 at 1920x1080 (BASELINE config 2)

Two modes: the reference as shipped (Renderer.cu:4-5: DEBUG_VIEW, checkerboard, no sun ray; ref_gpu / ref_gpu_nofma), and
the shaded + shadowed branch its source holds but compiles out (ref_gpu_shaded / ref_gpu_nofma_shaded, see
oracle/Makefile).  Timing uses the default-flag builds (what a user of the reference gets), parity the -fmad=false builds
(SURVEY.md §8c).  Prints one JSON line; bench.py imports compare() for its `ref_kernel` object."""
import json, os, subprocess, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np

REF_DIR = os.path.join(ROOT, "oracle", "_ref")
FRAMES = 8


def available():
    return all(os.path.exists(os.path.join(REF_DIR, b)) for b in ("ref_gpu", "ref_gpu_nofma"))


def _run_ref(binary, X, Y, Z, W, H, pos, eul, frames, path, last=None):
    env = dict(os.environ)
    if last:
        env["VXREF_DUMP_LAST"] = last
    p = subprocess.run([str(v) for v in [os.path.join(REF_DIR, binary), "render", X, Y, Z, 8, W, H, *pos, *eul, frames, path]],
                       capture_output=True, text=True, timeout=3000, env=env)
    if p.returncode != 0:
        raise RuntimeError("%s failed: %s" % (binary, p.stderr[-500:]))
    kv = dict(l.split("=", 1) for l in p.stdout.splitlines() if "=" in l)
    ms = [float(v) for k, v in kv.items() if k.startswith("ref_frame_ms")]
    startup = {k: float(kv[k]) for k in ("ref_voxel_generation_ms", "ref_buffer_generation_ms", "ref_upload_ms") if k in kv}
    return ms, startup


def compare(X=1024, Y=256, Z=1024, W=1920, H=1080, device=0):
    import torch
    import voxelengine_b200 as vx
    world = vx.World.from_noise(X, Y, Z, 8, device=device)
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
    out = {"world": [X, Y, Z], "frame": [W, H], "camera": pos, "euler": eul, "terrain_max_height": int(hgt.max()),
           "ours_world_build_ms": world.info.build_ms, "frames_timed": FRAMES - 2,
           "what": "the reference's screenDispatch (32x1 blocks, one thread per pixel) vs our four-kernel frame on the same B200, same world, camera and mode; "
                   "median of frames 2..%d, CUDA events; checkerboard frames (half the pixels per frame), as the reference renders" % (FRAMES - 1)}
    rt = vx.VoxelRaytracer3D(world, 16)
    cam = vx.make_camera(pos, eul, 90.0)
    stream = torch.cuda.current_stream().cuda_stream
    med = lambda v: float(np.median(v[2:]))
    modes = {"shipped": dict(fast="ref_gpu", exact="ref_gpu_nofma", debug=1, shadows=0,
                             desc="as shipped: checkerboard, DEBUG_VIEW quadrants, no sun ray (Renderer.cu:4-5,102)"),
             "shaded": dict(fast="ref_gpu_shaded", exact="ref_gpu_nofma_shaded", debug=0, shadows=1,
                            desc="calculateColor + Tonemap with the sun-shadow Raytrace of Renderer.cu:102 restored, checkerboard")}
    with tempfile.TemporaryDirectory() as tmp:
        path = os.path.join(tmp, "ref.bgra")
        for name, m in modes.items():
            if not all(os.path.exists(os.path.join(REF_DIR, m[b])) for b in ("fast", "exact")):
                continue
            ref_ms, startup = _run_ref(m["fast"], X, Y, Z, W, H, pos, eul, FRAMES, path)
            _run_ref(m["exact"], X, Y, Z, W, H, pos, eul, 1, path)
            ref0 = np.fromfile(path, np.uint8).reshape(H, W, 4)
            buf = torch.full((H, W, 4), 255, dtype=torch.uint8, device="cuda")
            ours, first = [], None
            for i in range(FRAMES):
                o = vx.RenderOpts(1, i, m["debug"], m["shadows"], 1, vx.NORM_DEVICE, 0, 0, 0, 0, None)  # the reference's frame counter advances every call
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                rt.render_device(cam, vx.default_env(), o, buf.data_ptr(), W, H, stream=stream)
                e1.record(); torch.cuda.synchronize()
                ours.append(e0.elapsed_time(e1))
                if i == 0:
                    first = buf.cpu().numpy().copy()
            diff = np.abs(first.astype(np.int16) - ref0.astype(np.int16))
            out[name] = {"mode": m["desc"], "reference_frame_ms": med(ref_ms), "ours_frame_ms": med(ours), "speedup": med(ref_ms) / med(ours),
                         "frame0_identical_to_fmad_false_build": bool(np.array_equal(first, ref0)),
                         "frame0_differing_pixels": int((diff > 0).any(2).sum()), "frame0_max_channel_diff": int(diff.max()),
                         "reference_startup_ms": startup}
    rt.close(); world.close()
    return out


if __name__ == "__main__":
    a = [int(v) for v in sys.argv[1:]]
    dims = a[:3] if len(a) >= 3 else (1024, 256, 1024)
    frame = a[3:5] if len(a) >= 5 else (1920, 1080)
    print(json.dumps(compare(*dims, *frame)))
