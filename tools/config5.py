"""BASELINE config 5 on one GPU: 1-spp incoherent bounce rays of the 4K frame (vx_bounce_rays -> vx_trace_device) and
world-build throughput.  python tools/config5.py [workload]   -> one JSON line"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import voxelengine_b200 as vx
import bench

name = sys.argv[1] if len(sys.argv) > 1 else "w2-4k"
wl = bench.WORKLOADS[name]
X, Y, Z = wl["dims"]
builds = []
for _ in range(2):
    world = vx.World.from_noise(X, Y, Z, 8)
    builds.append(world.info.build_ms)
    if _ == 0:
        world.close()
small = vx.World.from_noise(1024, 256, 1024, 8)
w1_ms = small.info.build_ms
small.close()
W, H = wl["W"], wl["H"]
rt = vx.VoxelRaytracer3D(world, 16)
cam, pos = bench.camera_for(vx, rt, wl["dims"], bench.CAMERAS[os.environ["VX_CAMERA"]] if os.environ.get("VX_CAMERA") else bench.camera_spec(wl))
opts = vx.RenderOpts(0, 0, 0, 1, 1, vx.NORM_DEVICE, 0, 0, 0, 0, None)
d_o = torch.zeros(W * H * 3, dtype=torch.float32, device="cuda")
d_d = torch.zeros(W * H * 3, dtype=torch.float32, device="cuda")
d_p = torch.zeros(W * H, dtype=torch.int32, device="cuda")
d_n = torch.zeros(1, dtype=torch.int64, device="cuda")
stream = torch.cuda.current_stream().cuda_stream
rt.bounce_rays(cam, opts, W, H, d_o.data_ptr(), d_d.data_ptr(), d_p.data_ptr(), d_n.data_ptr(), stream=stream)
torch.cuda.synchronize()
n = int(d_n.item())
hit = torch.zeros(n, dtype=torch.uint8, device="cuda")
steps = torch.zeros(n, dtype=torch.int32, device="cuda")
posb = torch.zeros(n * 3, dtype=torch.float32, device="cuda")
nrm = torch.zeros(n * 3, dtype=torch.float32, device="cuda")
cnt = torch.zeros(n * 3, dtype=torch.int32, device="cuda")
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
planes = dict(hit=hit.data_ptr(), pos=posb.data_ptr(), normal=nrm.data_ptr(), steps=steps.data_ptr())
rt.raytrace_device(d_o.data_ptr(), d_d.data_ptr(), n, dict(planes, counters=cnt.data_ptr()), stream=stream)  # counting variant, untimed
torch.cuda.synchronize()
c = cnt.view(n, 3).sum(0).tolist()
ms = []
for i in range(13):
    flush.fill_(1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); rt.raytrace_device(d_o.data_ptr(), d_d.data_ptr(), n, planes, stream=stream); e1.record()
    torch.cuda.synchronize()
    if i >= 3:
        ms.append(e0.elapsed_time(e1))
t = sum(ms) / len(ms)
alg = 4 * c[0] + 8 * c[1] + 64 * c[2] + 52 * n
print(json.dumps({"workload": name + " bounce: 1 spp AO sample rays of the primary hits (Renderer.cu:121-165), full MAX_STEPS cap",
                  "rays": n, "hits": int(hit.sum().item()), "steps_per_ray": float(steps.sum().item()) / n, "ms": t, "mrays_per_s": n / t / 1e3,
                  "C_K_E_per_ray": [v / n for v in c], "algorithmic_bytes": alg, "achieved_gbs": alg / t / 1e6,
                  "world_build": {"w2_ms": builds, "w2_gvoxels_per_s": X * Y * Z / min(builds) / 1e6,
                                  "w1_ms": w1_ms, "w1_gvoxels_per_s": 1024 * 256 * 1024 / w1_ms / 1e6}}))
