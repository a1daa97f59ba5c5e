#!/usr/bin/env python
"""bench.py — headline benchmark of the brickmap ray tracer (contract: task brief, "Measurement").

A step = one frame of the hot path: every pixel's primary ray through the two-level DDA plus, for
every primary hit, the sun shadow ray and shading (BASELINE.json configs[2]/[3]).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload w2-4k|w2-8k|w1-1080p]

N = 1   : 8192x512x8192 noise world, 8^3 bricks, 3840x2160, camera "overview" (SURVEY.md §8d config 3a).
N > 1   : same world replicated on every GPU, 7680x4320 frame split into 8-row bands dealt round-robin
          across ranks; rank 0 owns the frame and the other ranks' kernels store their pixels into it over NVLink
          peer memory, closed by a fence (config 4; --exchange gather: NCCL gather instead): strong scaling of one frame.
--impl reference : the CPU restatement of the reference's path (oracle/, kind "port": the reference has no
          CPU implementation of the 3-D march and its GPU build lives in oracle/_ref for parity only) on all
          host cores, each step a bounded row sample of the same frame.
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: dims, frame, camera (x,z), extra height, euler
    "w2-4k": dict(dims=(8192, 512, 8192), W=3840, H=2160, xz=(4096, 7936), lift=128, euler=(-0.5, 0.0, 0.0),
                  desc="8192x512x8192 noise brickmap (8^3 bricks), 3840x2160 primary + sun-shadow rays, overview camera"),
    "w2-8k": dict(dims=(8192, 512, 8192), W=7680, H=4320, xz=(4096, 7936), lift=128, euler=(-0.5, 0.0, 0.0),
                  desc="8192x512x8192 noise brickmap (8^3 bricks), 7680x4320 primary + sun-shadow rays, 8-row bands "
                       "round-robin over ranks, frame assembled on rank 0"),
    "w1-1080p": dict(dims=(1024, 256, 1024), W=1920, H=1080, xz=(512, 960), lift=0, euler=(-0.45, 0.0, 0.0),
                     desc="1024x256x1024 noise brickmap (8^3 bricks), 1920x1080 primary + sun-shadow rays"),
}
FACTOR = 8
L2_FLUSH_BYTES = 512 << 20


def read_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """SM clock and throttle reasons sampled WHILE the timed region runs: NVML from a thread every ~2 ms (the timed
    region of this bench lasts tens of milliseconds), with nvidia-smi -lms as the fall-back when NVML is not usable."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    REASONS = ((0x8, "hw_slowdown"), (0x40, "hw_thermal_slowdown"), (0x20, "sw_thermal_slowdown"), (0x4, "sw_power_cap"))

    def __init__(self, index):
        import threading
        self.path = tempfile.mktemp(prefix="clocks_", suffix=".csv")
        self.proc = None
        self.sm, self.mx, self.bits, self.halt, self.thread = [], [], 0, False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            try:
                import torch
                uuid = str(torch.cuda.get_device_properties(index).uuid)
                h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid if not uuid.startswith("GPU-") else uuid).encode())
            except Exception:
                h = pynvml.nvmlDeviceGetHandleByIndex(index)
            reasons = getattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons", None) or pynvml.nvmlDeviceGetCurrentClocksThrottleReasons
            mx = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))

            def loop():
                failures = 0
                while not self.halt and failures < 1000:
                    try:
                        self.sm.append(float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)))
                        self.mx.append(mx)
                    except Exception:
                        failures += 1
                    try:
                        self.bits |= int(reasons(h))
                    except Exception:
                        failures += 1
                    time.sleep(0.002)
            self.thread = threading.Thread(target=loop, daemon=True)
            self.thread.start()
        except Exception:
            self.thread = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "200"], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "source": None}
        self.halt = True
        if self.thread is not None:
            self.thread.join(timeout=2)
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        try:
            for line in open(self.path):
                c = [v.strip() for v in line.split(",")]
                if len(c) < 7:
                    continue
                sm.append(float(c[0])); mx.append(float(c[1]))
                for n, v in zip(names, c[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            os.unlink(self.path)
        except Exception:
            pass
        if self.sm:  # NVML samples, all of them taken inside the timed region
            out.update(sm_mhz=float(np.median(self.sm)), sm_max_mhz=float(max(self.mx)), samples=len(self.sm), source="nvml, every 2 ms")
            reasons |= {n for bit, n in self.REASONS if self.bits & bit}
        elif sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), samples=len(sm), source="nvidia-smi -lms 200")
        out["reasons"] = sorted(reasons)
        return out


def camera_for(vx, rt, wl):
    """Camera height rule of SURVEY.md §8d: 16 above the highest solid voxel of the camera's column
    (found with one vertical ray), Y+32 when the column is solid to the top; plus the workload's lift."""
    X, Y, Z = wl["dims"]
    x, z = wl["xz"]
    r = rt.raytrace(np.array([[x + 0.5, Y + 10.0, z + 0.5]], np.float32), np.array([[0.0, -1.0, 0.0]], np.float32))
    top = int(r["voxel"][0, 1]) if r["hit"][0] else -1
    y = (top + 16 if top < Y - 1 else Y + 32) + wl["lift"]
    return vx.make_camera((x, y, z), wl["euler"], 90.0), (float(x), float(y), float(z))


def cpu_sample(ol, ow, wl, cam_pos, tan_half, target_s, threads):
    """Times the CPU oracle on a bounded row sample of the workload frame; returns (Mrays/s, rays, seconds, stride)."""
    W, H = wl["W"], wl["H"]
    cam = ol.make_camera(cam_pos, wl["euler"], 90.0, tan_half=tan_half)
    env = ol.default_env()
    opts = ol.RenderOpts(0, 0, 0, 1, 1, ol.NORM_HOST)
    # calibrate on 1/256 of the rows, then size the sample for about target_s seconds
    stride = 256
    t0 = time.perf_counter()
    st = ow.render_sample(cam, env, opts, W, H, 0, H, stride, threads)
    dt = time.perf_counter() - t0
    rows_now = (H + stride - 1) // stride
    rows_want = max(rows_now, min(H, int(rows_now * target_s / max(dt, 1e-3))))
    stride = max(1, H // rows_want)
    return cam, env, opts, stride


def run_reference(args, wl, wl_name):
    """--impl reference: the CPU port of the reference's march, all host threads, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol
    import voxelengine_b200 as vx
    threads = os.cpu_count() or 1
    X, Y, Z = wl["dims"]
    # the world is the benchmark's INPUT: it is generated on the GPU (the CPU cannot generate 2^35 voxels of
    # 32-octave noise in minutes) and handed to the CPU code in the reference's own form
    world = vx.World.from_noise(X, Y, Z, FACTOR, device=0)
    rt = vx.VoxelRaytracer3D(world, 16)
    _, cam_pos = camera_for(vx, rt, wl)
    fov = np.float32(np.float32(90.0 * 3.1415 / 180.0) / np.float32(2.0))
    tan_half = vx.device_tanf(fov)
    top, bw, b6 = world.download()
    rt.close(); world.close()
    ow = ol.OracleWorld.from_brickmap(top, bw, b6, X, Y, Z, FACTOR)
    cam, env, opts, stride = cpu_sample(ol, ow, wl, cam_pos, tan_half, 2.0, threads)
    W, H = wl["W"], wl["H"]
    for _ in range(args.warmup):
        ow.render_sample(cam, env, opts, W, H, 0, H, stride, threads)
    rays = 0
    t0 = time.perf_counter()
    for k in range(args.steps):
        st = ow.render_sample(cam, env, opts, W, H, k % stride, H, stride, threads)
        rays += int(st[0] + st[2])
    dt = time.perf_counter() - t0
    val = rays / dt / 1e6
    sample = "every %d-th pixel row of the %dx%d frame per step (%d rays/step), same world/camera as the GPU arm" % (
        stride, W, H, rays // max(args.steps, 1))
    line = {"impl": "reference", "metric": "Mrays/s", "value": val, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt / max(args.steps, 1) * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl_name + ": " + wl["desc"], "camera": cam_pos, "timing": "host wall clock, CPU only"},
            "cpu_baseline": {"value": val, "unit": "Mrays/s", "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def run_ours(args, wl, wl_name):
    import torch
    import torch.distributed as dist
    import voxelengine_b200 as vx
    import ctypes as C

    rank = int(os.environ.get("RANK", "0"))
    world_size = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world_size > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = local_rank
    X, Y, Z = wl["dims"]
    W, H = wl["W"], wl["H"]
    world = vx.World.from_noise(X, Y, Z, FACTOR, device=dev)
    info = world.info
    rt = vx.VoxelRaytracer3D(world, 16)
    if "VX_FUSED_SHADE" in os.environ:  # experiments only; the default picks by where the frame lives
        rt.set_fused_shade(int(os.environ["VX_FUSED_SHADE"]))
    cam, cam_pos = camera_for(vx, rt, wl)
    env = vx.default_env()
    stream = torch.cuda.current_stream().cuda_stream
    from voxelengine_b200.partition import FramePartition, SharedFrame
    part = FramePartition(W, H, world_size, rank)   # 8-row bands dealt round-robin over the ranks
    shared = None
    if world_size > 1 and args.exchange == "shared":
        try:
            shared = SharedFrame(part, dev)
        except RuntimeError as e:  # raised on every rank together: fall back to the NCCL gather
            if rank == 0:
                print("bench: shared frame unavailable (%s); using the NCCL gather" % e, file=sys.stderr)
    frame = part.alloc_frame("cuda") if shared is None else None
    if shared is not None and "VX_FUSED_SHADE" not in os.environ:
        shared.attach(rt)
    frame_ptr = shared.ptr if shared else frame.data_ptr()
    stats = torch.zeros(8, dtype=torch.int64, device="cuda")
    flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device="cuda")
    host = torch.empty(H * W * 4, dtype=torch.uint8).pin_memory() if world_size == 1 else None
    band_count, band_index = part.render_opts()

    def opts(stats_ptr=None):
        return vx.RenderOpts(0, 0, 0, 1, 1, vx.NORM_DEVICE, 0, 0, 0, 0, stats_ptr, band_count, band_index)

    def render(o):
        rt.render_device(cam, env, o, frame_ptr, W, H, stream=stream)

    def gather():
        part.gather(frame)

    # N > 1: the frame is rendered in row chunks and each chunk's bands start their way to rank 0 while the next chunk
    # renders (FramePartition.render_and_gather); kernel times are summed over the chunks
    chunk_ms = [0.0, 0.0, 0.0, 0.0]
    chunk_opts = {}
    cols = (W + 31) // 32 * 32

    def render_rows(y0, h):
        o = chunk_opts.get((y0, h))
        if o is None:
            o = chunk_opts[(y0, h)] = vx.RenderOpts(0, 0, 0, 1, 1, vx.NORM_DEVICE, 0, y0, cols, h, None, band_count, band_index)
        render(o)
        if timing_chunks[0]:
            for i, v in enumerate(rt.frame_times()):
                chunk_ms[i] += v
    timing_chunks = [False]

    def frame_step():
        if world_size == 1:
            render(o_timed)
        elif shared is not None:
            # every rank's kernels store their pixels into rank 0's frame over NVLink; the fence is the only collective
            render(o_timed)
            shared.fence()
        elif args.gather_chunks <= 1:
            render(o_timed)
            gather()
        else:
            part.render_and_gather(frame, render_rows, args.gather_chunks)

    # counters of exactly this frame (untimed counting variants of the kernels): first the primary rays alone
    # (shadows off), then the whole frame
    o_prim = opts(stats.data_ptr())
    o_prim.shadows = 0
    render(o_prim)
    torch.cuda.synchronize()
    st_prim = stats.clone()
    stats.zero_()
    render(opts(stats.data_ptr()))
    torch.cuda.synchronize()
    st = stats.clone()
    if world_size > 1:
        dist.all_reduce(st)
    st = st.cpu().numpy()
    rays = int(st[0] + st[2])
    alg_bytes_rank = int(4 * stats[5].item() + 8 * stats[6].item() + 64 * stats[7].item() + 4 * int(stats[0].item()))
    # per kernel: primary_rays reads its cells, writes the miss pixels (4 B) or queues the hit (20 B); shadow_rays reads its
    # lane record (48 B) and cells and writes its result (4 B)
    p_hits, p_rays = int(st_prim[1].item()), int(st_prim[0].item())
    alg_primary = int(4 * st_prim[5].item() + 8 * st_prim[6].item() + 64 * st_prim[7].item() + 4 * (p_rays - p_hits) + 20 * p_hits)
    s_rays = int(stats[2].item())
    alg_shadow = int(4 * (stats[5] - st_prim[5]).item() + 8 * (stats[6] - st_prim[6]).item() + 64 * (stats[7] - st_prim[7]).item() + 52 * s_rays)
    rt.set_timing(True)
    o_timed = opts(None)
    for _ in range(max(args.warmup, 3)):
        flush.fill_(1)
        frame_step()
    torch.cuda.synchronize()

    # ---- timed region: K steps, device time by CUDA events on the launching stream -------------
    sampler = ClockSampler(dev) if rank == 0 else None
    if world_size > 1:
        dist.barrier()
    torch.cuda.synchronize()
    wall0 = time.perf_counter()
    k_ms, step_ms = 0.0, 0.0
    kern_ms = [0.0, 0.0, 0.0, 0.0]
    for _ in range(args.steps):
        flush.fill_(1)  # L2 flush between timed iterations (512 MiB > 126 MB L2)
        e0, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(2))
        e0.record()
        frame_step()
        e2.record()
        e2.synchronize()
        step_ms += e0.elapsed_time(e2)
        if world_size == 1 or shared is not None or args.gather_chunks <= 1:
            kern_ms = [a + b for a, b in zip(kern_ms, rt.frame_times())]
    torch.cuda.synchronize()
    if world_size > 1:
        dist.barrier()
    wall = time.perf_counter() - wall0
    if world_size > 1 and shared is None and args.gather_chunks > 1:
        # per-kernel times of the chunked frame: an untimed extra pass (reading the events waits for each chunk)
        timing_chunks[0] = True
        for _ in range(args.steps):
            flush.fill_(1)
            frame_step()
        torch.cuda.synchronize()
        timing_chunks[0] = False
        kern_ms = list(chunk_ms)
    k_ms = sum(kern_ms[:4])
    rt.set_timing(False)
    t = torch.tensor([step_ms, k_ms] + kern_ms, dtype=torch.float64, device="cuda")
    if world_size > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    step_ms_max, k_ms_max = (float(v) for v in t.cpu()[:2])
    kern_ms = [float(v) for v in t.cpu()[2:]]
    clocks = sampler.stop() if sampler else None

    exchange = "none" if world_size == 1 else ("peer stores into rank 0's frame (CUDA IPC over NVLink) + fence" if shared else "torch.distributed gather to rank 0")
    # ---- e2e: the reference-facing call with HOST buffers (RenderScreen + cudaMemcpy of main.cu:165-167) ----
    e2e_rays_s = None
    d2h = W * H * 4
    h2d = C.sizeof(vx.Camera) + C.sizeof(vx.Env) + C.sizeof(vx.RenderOpts)
    hostframe = None
    if world_size > 1:
        # N > 1: one pinned host frame shared by the ranks' processes; every rank renders its bands into its own device
        # frame and copies them down its own PCIe link (HostFrame) — the GPUs exchange nothing on this path
        from voxelengine_b200.partition import HostFrame
        if shared is not None:
            shared.close()
            shared = None
        rt.set_fused_shade(0)
        local = frame if frame is not None else part.alloc_frame("cuda")
        try:
            hostframe = HostFrame(part, dev)
        except RuntimeError as e:  # every rank together: gather to rank 0, which copies the whole frame down
            if rank == 0:
                print("bench: shared host frame unavailable (%s); rank 0 copies the gathered frame" % e, file=sys.stderr)
            frame = local
            host = torch.empty(H * W * 4, dtype=torch.uint8).pin_memory() if rank == 0 else None

        e2e_opts = {}

        def e2e_rows(y0, h):
            o = e2e_opts.get((y0, h))
            if o is None:
                o = e2e_opts[(y0, h)] = vx.RenderOpts(0, 0, 0, 1, 1, vx.NORM_DEVICE, 0, y0, cols, h, None, band_count, band_index)
            rt.render_device(cam, env, o, local.data_ptr(), W, H, stream=stream)

        def e2e_step():
            if hostframe is None:
                rt.render_device(cam, env, o_timed, local.data_ptr(), W, H, stream=stream)
                part.gather(local)
                if rank == 0:
                    host.copy_(local[: H * W * 4], non_blocking=True)
                torch.cuda.synchronize()
                return
            if args.host_split >= 1.0:
                rt.render_device(cam, env, o_timed, local.data_ptr(), W, H, stream=stream)
                hostframe.store(local.data_ptr(), stream)
            else:
                hostframe.render_and_store(e2e_rows, local.data_ptr(), args.host_split)
            hostframe.fence()
        e2e_step()
        dist.barrier()
    else:
        # untimed first call: allocates the tracer's device frame and copy stream and seeds the frame from the host buffer
        vx._check(vx.lib().vx_render_host(rt.h, C.byref(cam), C.byref(env), C.byref(o_timed), host.data_ptr(), W, H))
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        if world_size == 1:
            vx._check(vx.lib().vx_render_host(rt.h, C.byref(cam), C.byref(env), C.byref(o_timed), host.data_ptr(), W, H))
        else:
            e2e_step()
    if world_size > 1:
        dist.barrier()
    e2e_s = time.perf_counter() - t0
    e2e_rays_s = rays * args.steps / e2e_s / 1e6
    if hostframe is not None:
        hostframe.close()
    chunks = len(part.row_chunks(args.gather_chunks)) if world_size > 1 and shared is None and args.gather_chunks > 1 else 1
    return dict(chunks=chunks, exchange=exchange, vx=vx, rank=rank, world_size=world_size, rays=rays, st=st, alg_bytes_rank=alg_bytes_rank, step_ms=step_ms_max,
                k_ms=k_ms_max, kern_ms=kern_ms, alg_primary=alg_primary, alg_shadow=alg_shadow, wall=wall, clocks=clocks, e2e=e2e_rays_s, h2d=h2d, d2h=d2h, info=info, cam_pos=cam_pos, world=world,
                rt=rt, dev=dev)


def cpu_baseline_leg(res, wl):
    """N=1, rank 0: the oracle (CPU port of the reference's march) on a bounded sample of the same frame."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol
    vx = res["vx"]
    threads = os.cpu_count() or 1
    X, Y, Z = wl["dims"]
    fov = np.float32(np.float32(90.0 * 3.1415 / 180.0) / np.float32(2.0))
    tan_half = vx.device_tanf(fov, res["dev"])
    top, bw, b6 = res["world"].download()
    ow = ol.OracleWorld.from_brickmap(top, bw, b6, X, Y, Z, FACTOR)
    cam, env, opts, stride = cpu_sample(ol, ow, wl, res["cam_pos"], tan_half, 12.0, threads)
    # about 10 s of CPU work: the row sample, repeated when even the whole frame is quicker than that
    rays, passes, st = 0, 0, None
    t0 = time.perf_counter()
    while passes == 0 or (time.perf_counter() - t0 < 10.0 and passes < 64):
        st = ow.render_sample(cam, env, opts, wl["W"], wl["H"], (stride // 2 + passes) % stride, wl["H"], stride, threads)
        rays += int(st[0] + st[2])
        passes += 1
    dt = time.perf_counter() - t0
    return {"value": rays / dt / 1e6, "unit": "Mrays/s", "cores": threads, "kind": "port",
            "sample": "every %d-th pixel row of the %dx%d frame, %d passes (%d rays, %.1f s), same world and camera" % (
                stride, wl["W"], wl["H"], passes, rays, dt),
            "steps_per_ray": float(st[4]) / max(int(st[0] + st[2]), 1)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, choices=list(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--exchange", default=os.environ.get("VX_EXCHANGE", "shared"), choices=["shared", "gather"],
                    help="N > 1: how the bands reach rank 0 (shared: peer stores into rank 0's frame; gather: NCCL gather)")
    ap.add_argument("--host-split", type=float, default=float(os.environ.get("VX_HOST_SPLIT_N", "0.8")),
                    help="N > 1, e2e: share of the rows rendered before the first copy to the host starts (1 = no overlap)")
    ap.add_argument("--gather-chunks", type=int, default=int(os.environ.get("VX_GATHER_CHUNKS", "1")),
                    help="N > 1: row chunks of the frame whose gather overlaps the next chunk's render (1 = render, then gather)")
    args = ap.parse_args()
    world_size = int(os.environ.get("WORLD_SIZE", "1"))
    wl_name = args.workload or ("w2-4k" if max(args.gpus, world_size) == 1 else "w2-8k")
    wl = WORKLOADS[wl_name]
    if args.impl == "reference":
        run_reference(args, wl, wl_name)
        return
    res = run_ours(args, wl, wl_name)
    if res["rank"] != 0:
        import torch.distributed as dist
        dist.destroy_process_group()
        return
    vx = res["vx"]
    K = args.steps
    ms_step = res["step_ms"] / K
    ms_kernel = res["k_ms"] / K
    value = res["rays"] / (ms_step * 1e-3) / 1e6
    hbm_peak, peak_src = read_peaks()
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get(wl_name)
    except Exception:
        pass
    l2_peak = vx.measure_read_bandwidth(32 << 20, 50, res["dev"])
    names = vx.VoxelRaytracer3D.KERNELS
    kern = {n: res["kern_ms"][i] / K for i, n in enumerate(names)}

    def roof(kernel, alg_bytes, formula):
        ms = kern[kernel]
        a = alg_bytes / (ms * 1e-3) / 1e9
        return {"bound": "hbm", "achieved": a, "peak": hbm_peak, "unit": "GB/s", "frac": a / hbm_peak,
                "traffic": (traffic or {}).get(kernel) if isinstance(traffic, dict) else None, "peak_source": peak_src,
                "kernel": "vx::" + kernel, "kernel_ms": ms, "algorithmic_bytes_per_launch": alg_bytes, "formula": formula,
                "l2_peak_gbs_measured": l2_peak, "frac_of_l2": a / l2_peak}
    st = res["st"]
    line = {
        "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": res["world_size"], "steps": K, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl_name + ": " + wl["desc"], "camera": res["cam_pos"], "euler": wl["euler"], "fov_deg": 90,
                   "rays_per_step": res["rays"], "primary_rays": int(st[0]), "shadow_rays": int(st[2]), "steps_per_ray": float(st[4]) / res["rays"],
                   "parallelism": "image bands x%d, world replicated" % res["world_size"],
                   "frame_chunks": res["chunks"], "exchange": res["exchange"], "l2": "flushed between timed steps (512 MiB write)",
                   "world": {"occupied_bricks": int(res["info"].occupied_bricks), "bricks": int(res["info"].bricks),
                             "pool_mb": res["info"].bytes_pool / 1e6, "build_ms": res["info"].build_ms}},
        "clocks": res["clocks"],
        "e2e": {"value": res["e2e"], "unit": "Mrays/s", "h2d_bytes_per_step": res["h2d"], "d2h_bytes_per_step": res["d2h"]},
        "gpu_launches": 4 * K * res["chunks"],
        "kernels_ms": dict(kern, frame=ms_kernel),
        # the dominant kernel first; the sun-ray march, the other half of the frame, beside it
        "roofline": roof("primary_rays", res["alg_primary"], "4*C + 8*K + 64*E of the primary rays + 4 B per miss pixel + 20 B per queued hit (SURVEY.md 8d)"),
        "roofline_shadow_rays": roof("shadow_rays", res["alg_shadow"], "4*C + 8*K + 64*E of the sun rays + 48 B lane record + 4 B result per ray"),
        "frame_algorithmic_bytes": res["alg_bytes_rank"],
        "wall_s_timed_region": res["wall"],
    }
    if res["world_size"] == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline_leg(res, wl)
    print(json.dumps(line), flush=True)
    if res["world_size"] > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
