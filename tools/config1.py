"""BASELINE config 1: CPU DDA traversal of a 256^3 dense bit-array grid, 1920x1080 primary rays, fixed camera, headless
 — the reference's OWN code on the host cores next to ours on the GPU, same world, same rays.

What runs on the CPU is oracle/_ref/libvxref_host.so: the reference's VolumeRaytracer.cu compiled unmodified for the host
(oracle/Makefile), i.e. GenerateLowresVoxelBuffer + GPUDDA::Raytrace as written, on all host threads ("kind": "reference"),
and — for the single-level variant DDATestCpp stood for — its DDARayTraversal over the dense grid (one thread: the wrapper
loops).  The CPU restatement (oracle/, "port") is timed beside it.  The world is the reference generator's 256^3 terrain
(oracle's CPU PopulateVoxels, checked against the reference's cuda_noise.cuh elsewhere); the camera sits outside the grid on
the +z side, so every ray takes the clip path (VolumeRaytracer.cu:369-381).

On a GPU box the same rays go through vx_trace (host-normalisation mode, which is what the reference computes when it runs
on a CPU) and every plane must equal the reference's output bit for bit.

python tools/config1.py            -> one JSON line (GPU part only when a CUDA device is present)
"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import oracle_lib as ol
from raysets import camera_rays

X = Y = Z = 256
W, H = 1920, 1080
FACTOR = 8
CAM_XZ, EULER = (128.0, 320.0), (-0.35, 0.0, 0.0)


def camera_height(bits):
    """16 above the highest solid voxel of the column nearest to the camera (the camera itself is outside the grid in z)."""
    x, z = int(CAM_XZ[0]), Z - 1
    top = -1
    for y in range(Y):
        i = ((x // 8) + (y // 8) * (X // 8) + (z // 8) * (X // 8) * (Y // 8)) * 512 + (x % 8) + (y % 8) * 8 + (z % 8) * 64
        if (int(bits[i >> 5]) >> (i & 31)) & 1:
            top = y
    return float(top + 16) if top < Y - 1 else float(Y + 32)


def main():
    ref = ol.ref_host()
    threads = os.cpu_count() or 1
    t0 = time.perf_counter()
    bits = ol.populate_dense(X, Y, Z)
    t_noise = time.perf_counter() - t0
    cam_pos = (CAM_XZ[0], camera_height(bits), CAM_XZ[1])
    fwd, up, right = ol.get_directions(EULER)
    tan_half = float(np.tan(np.float32(np.float32(90.0 * 3.1415 / 180.0) / np.float32(2.0))))
    o, d = camera_rays(cam_pos, fwd, up, right, W, H, tan_half)
    n = o.shape[0]
    line = {"config": "1: 256^3 noise terrain (dense bit array -> 32^3 bricks of 8^3), %dx%d primary rays, camera %s euler %s, "
                      "every ray enters through the grid clip" % (W, H, list(cam_pos), list(EULER)),
            "rays": n, "host_threads": threads, "cpu_noise_build_s": t_noise}

    # ---- the oracle port (always present) ----
    w = ol.OracleWorld.from_dense(bits, X, Y, Z, FACTOR)
    w.trace(o[:1000], d[:1000])
    t0 = time.perf_counter()
    r = w.trace(o, d, norm_mode=ol.NORM_HOST)
    t_port = time.perf_counter() - t0
    t0 = time.perf_counter()
    dense_hit, dense_steps = w.trace_dense(o, d, norm_mode=ol.NORM_HOST)
    t_port_dense = time.perf_counter() - t0
    line["cpu_port"] = {"two_level_mrays_s": n / t_port / 1e6, "single_level_dense_mrays_s": n / t_port_dense / 1e6, "threads": threads,
                        "hits": int(r["hit"].sum()), "steps_per_ray": float(r["steps"].mean()), "dense_steps_per_ray": float(dense_steps.mean())}

    # ---- the reference's own code on the host ----
    hit = np.zeros(n, np.uint8); pos = np.zeros((n, 3), np.float32); nrm = np.zeros((n, 3), np.float32); st = np.zeros(n, np.int32)
    if ref is not None:
        t0 = time.perf_counter()
        rw = ref.vxref_build(bits, X, Y, Z, FACTOR)
        t_build = time.perf_counter() - t0
        ref.vxref_trace(rw, o[:1000].copy(), d[:1000].copy(), 1000, 2048, hit.ctypes.data, pos, nrm, st.ctypes.data, 0)
        t0 = time.perf_counter()
        ref.vxref_trace(rw, o, d, n, 2048, hit.ctypes.data, pos, nrm, st.ctypes.data, 0)
        t_ref = time.perf_counter() - t0
        same = bool(np.array_equal(hit, r["hit"]) and np.array_equal(st, r["steps"]) and
                    np.array_equal(pos.view(np.uint32), r["pos"].view(np.uint32)) and np.array_equal(nrm, r["normal"]))
        line["cpu_reference"] = {"kind": "reference", "what": "GPUDDA::Raytrace of VolumeRaytracer.cu compiled for the host, all threads",
                                 "two_level_mrays_s": n / t_ref / 1e6, "threads": threads, "GenerateLowresVoxelBuffer_s": t_build,
                                 "hits": int(hit.sum()), "steps_per_ray": float(st.mean()), "port_equals_reference_bit_for_bit": same}
        if not same:
            line["error"] = "oracle port and reference disagree"
    else:
        line["cpu_reference"] = None

    # ---- ours, when a GPU is here ----
    try:
        import torch
        have_gpu = torch.cuda.is_available()
    except Exception:
        have_gpu = False
    if have_gpu:
        import voxelengine_b200 as vx
        t0 = time.perf_counter()
        world = vx.World.from_noise(X, Y, Z, FACTOR)
        t_gpu_build = time.perf_counter() - t0
        rt = vx.VoxelRaytracer3D(world, n)
        g = rt.raytrace(o, d, norm_mode=vx.NORM_HOST)           # host buffers in, host planes out (the e2e call)
        ms_fresh = []
        for _ in range(3):  # every plane the library has (8 + valid / distance), into freshly allocated numpy arrays: the page faults
            t0 = time.perf_counter()  # of 130 MB of new result arrays are most of this figure
            g = rt.raytrace(o, d, norm_mode=vx.NORM_HOST)
            ms_fresh.append((time.perf_counter() - t0) * 1e3)
        # the call VoxelRaytracer3D::Raytrace(vector, vector) stands for: the reference's result planes, result arrays re-used
        res = rt.raytrace(o, d, norm_mode=vx.NORM_HOST, planes=vx.VoxelRaytracer3D.REFERENCE_PLANES)
        ms = []
        for _ in range(6):
            t0 = time.perf_counter()
            res = rt.raytrace(o, d, norm_mode=vx.NORM_HOST, planes=vx.VoxelRaytracer3D.REFERENCE_PLANES, out=res)
            ms.append((time.perf_counter() - t0) * 1e3)
        d_o, d_d = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
        planes = {k: torch.zeros(s, dtype=t, device="cuda") for k, s, t in (("hit", n, torch.uint8), ("pos", 3 * n, torch.float32),
                                                                              ("normal", 3 * n, torch.float32), ("steps", n, torch.int32))}
        ptrs = {k: v.data_ptr() for k, v in planes.items()}
        stream = torch.cuda.current_stream().cuda_stream
        dev = []
        for i in range(8):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); rt.raytrace_device(d_o.data_ptr(), d_d.data_ptr(), n, ptrs, norm_mode=vx.NORM_HOST, stream=stream); e1.record()
            torch.cuda.synchronize()
            if i >= 3:
                dev.append(e0.elapsed_time(e1))
        want = (hit, st, pos, nrm) if ref is not None else (r["hit"], r["steps"], r["pos"], r["normal"])
        same = bool(np.array_equal(g["hit"], want[0]) and np.array_equal(g["steps"], want[1]) and
                    np.array_equal(g["pos"].view(np.uint32), want[2].view(np.uint32)) and np.array_equal(g["normal"], want[3]))
        line["gpu"] = {"device_mrays_s": n / (sum(dev) / len(dev)) / 1e3, "device_ms": sum(dev) / len(dev),
                       "host_api_mrays_s": n / min(ms) / 1e3, "host_api_ms": min(ms), "host_api_planes": list(vx.VoxelRaytracer3D.REFERENCE_PLANES),
                       "host_api_all_planes_fresh_arrays_ms": min(ms_fresh), "world_build_s": t_gpu_build,
                       "equals_cpu_%s_bit_for_bit" % ("reference" if ref is not None else "port"): same}
        if not same:
            line["error"] = "GPU planes differ from the CPU result"
    print(json.dumps(line))
    return 1 if "error" in line else 0


if __name__ == "__main__":
    sys.exit(main())
