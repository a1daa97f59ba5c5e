#!/usr/bin/env python
"""bench.py — headline benchmark of the brickmap ray tracer (contract: task brief, "Measurement").

A step = one frame of the hot path: every pixel's primary ray through the two-level DDA plus, for
every primary hit, the sun shadow ray and shading (BASELINE.json configs[2]).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload w2-4k|w2-8k|w1-1080p]

Every N : 8192x512x8192 noise world (replicated per GPU), 3840x2160, camera (a) "overview" of SURVEY.md §8d config 3
          — BASELINE.json's metric is quoted "at 4K ... 1/2/4/8 B200", so `value` is the SAME frame at every N
          (strong scaling, like for like).  N > 1: the frame's 8-row bands are dealt round-robin over the ranks; rank 0
          owns the frame and the other ranks' kernels store their pixels into it over NVLink peer memory, closed by a
          one-word fence (--exchange gather: NCCL gather instead).
          Beside it, in the same line: camera (b) "grazing" (`grazing`, N = 1), the 7680x4320 frame of configs[3]
          (`config4_8k`, N > 1) with its own single-GPU base measured in the same run (`scaling_base`), bounce rays and
          world-build throughput of configs[4] (`config5`), the parity of the timed frame against the CPU oracle at
          this scale (`parity`, N = 1) and the identity of the assembled multi-GPU frame (`frame_identical`, N > 1).
--impl reference : the reference's own march (GPUDDA::Raytrace compiled for the host, oracle/_ref; the CPU port in
          oracle/ when that is absent) on all host cores over the rays of a bounded row sample of the same frame.
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
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

DIMS_W2 = (8192, 512, 8192)
CAMERAS = {
    "overview": dict(xz=(4096, 7936), lift=128, euler=(-0.5, 0.0, 0.0)),
    "grazing": dict(xz=(4096, 4096), lift=0, euler=(-0.05, 0.8, 0.0)),
    # not a config of BASELINE.json: from one corner along the diagonal, traversals that can meet the step cap before the far
    # border (tests/test_gpu_w2_scale.py checks it against the oracle, tools/sweep_kernels.py times it)
    "diagonal": dict(xz=(8180, 8180), lift=0, euler=(-0.02, 0.7853982, 0.0)),
}
WORKLOADS = {
    "w2-4k": dict(dims=DIMS_W2, W=3840, H=2160, camera="overview",
                  desc="8192x512x8192 noise brickmap (8^3 bricks), 3840x2160 primary + sun-shadow rays, overview camera"),
    "w2-8k": dict(dims=DIMS_W2, W=7680, H=4320, camera="overview",
                  desc="8192x512x8192 noise brickmap (8^3 bricks), 7680x4320 primary + sun-shadow rays, overview camera"),
    "w1-1080p": dict(dims=(1024, 256, 1024), W=1920, H=1080, camera=dict(xz=(512, 960), lift=0, euler=(-0.45, 0.0, 0.0)),
                     desc="1024x256x1024 noise brickmap (8^3 bricks), 1920x1080 primary + sun-shadow rays"),
}
FACTOR = 8
L2_FLUSH_BYTES = 512 << 20
NOISE_OPS_PER_VOXEL = 12800  # 32 octaves x ~400 scalar operations


def read_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)", float(p.get("sm_max_mhz", 1965.0))
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)", 1965.0


class ClockSampler:
    """SM clock and throttle reasons sampled WHILE a timed region runs: NVML from a thread every ~2 ms (the timed
    regions of this bench last tens of milliseconds), with nvidia-smi -lms as the fall-back when NVML is not usable."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    REASONS = ((0x8, "hw_slowdown"), (0x40, "hw_thermal_slowdown"), (0x20, "sw_thermal_slowdown"), (0x4, "sw_power_cap"))

    def __init__(self, index, smi=True):
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
        if smi:
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


def camera_spec(wl):
    return CAMERAS[wl["camera"]] if isinstance(wl["camera"], str) else wl["camera"]


def camera_for(vx, rt, dims, cam):
    """Camera height rule of SURVEY.md §8d: 16 above the highest solid voxel of the camera's column
    (found with one vertical ray), Y+32 when the column is solid to the top; plus the camera's lift."""
    X, Y, Z = dims
    x, z = cam["xz"]
    r = rt.raytrace(np.array([[x + 0.5, Y + 10.0, z + 0.5]], np.float32), np.array([[0.0, -1.0, 0.0]], np.float32))
    top = int(r["voxel"][0, 1]) if r["hit"][0] else -1
    y = (top + 16 if top < Y - 1 else Y + 32) + cam["lift"]
    return vx.make_camera((x, y, z), cam["euler"], 90.0), (float(x), float(y), float(z))


def device_tan_half(vx, dev=0):
    fov = np.float32(np.float32(90.0 * 3.1415 / 180.0) / np.float32(2.0))
    return vx.device_tanf(fov, dev)


def world_check(vx, ol, world, dims, n=256, seed=1):
    """"Same world" checked, not assumed: CRC of the top-level bits, and n bricks (half of them on the terrain surface)
    re-derived voxel by voxel from the oracle's 32-octave noise."""
    top = world.download_top()
    rng = np.random.default_rng(seed)
    cx, cy, cz = (d // FACTOR for d in dims)
    c = np.stack([rng.integers(0, cx, n), rng.integers(0, cy, n), rng.integers(0, cz, n)], 1).astype(np.uint32)
    rt = vx.VoxelRaytracer3D(world, n)
    k = n // 2
    o = np.stack([rng.integers(0, dims[0], k) + 0.5, np.full(k, dims[1] + 10.0), rng.integers(0, dims[2], k) + 0.5], 1).astype(np.float32)
    r = rt.raytrace(o, np.tile(np.array([[0, -1, 0]], np.float32), (k, 1)))
    rt.close()
    surf = np.nonzero(r["hit"])[0]
    c[:surf.size] = (r["voxel"][surf] // FACTOR).astype(np.uint32)
    occ, bw, _ = world.download_bricks(c)
    bad = 0
    for i in range(n):
        bits = ol.populate_dense(FACTOR, FACTOR, FACTOR, int(c[i, 0]) * FACTOR, int(c[i, 1]) * FACTOR, int(c[i, 2]) * FACTOR, nthreads=1)
        bad += int(not np.array_equal(bits, bw[i]) or bool(occ[i] & 1) != bool(bits.any()))
    return {"top_bits_crc32": "%08x" % zlib.crc32(top.tobytes()), "bricks_rederived_from_oracle_noise": n, "bricks_differing": bad}


def pick_stride(ol, ow, cam, env, opts, W, H, target_s, threads):
    """Row stride of a sample of the frame that costs the oracle about target_s seconds on `threads` cores."""
    stride = 256
    t0 = time.perf_counter()
    ow.render_sample(cam, env, opts, W, H, 0, H, stride, threads)
    dt = time.perf_counter() - t0
    rows_now = (H + stride - 1) // stride
    rows_want = max(rows_now, min(H, int(rows_now * target_s / max(dt, 1e-3))))
    return max(1, H // rows_want)


# ---------------------------------------------------------------------------------------------------------------------
# --impl reference
# ---------------------------------------------------------------------------------------------------------------------
def run_reference(args, wl, wl_name):
    """The reference's CPU implementation of the path on the host cores.  The reference has no 3-D CPU march of its own
    (SURVEY F1); what runs here is its GPUDDA::Raytrace source compiled for the host (oracle/_ref/libvxref_host.so, kind
    "reference") over the rays screenDispatch would hand it for a row sample of the frame — or, where /root/reference was
    not available to build that, the CPU port in oracle/ (kind "port")."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol
    import voxelengine_b200 as vx
    threads = os.cpu_count() or 1
    X, Y, Z = wl["dims"]
    W, H = wl["W"], wl["H"]
    # the world is the benchmark's INPUT: it is generated on the GPU (a CPU cannot generate 2^35 voxels of 32-octave
    # noise in minutes), checked against the oracle's noise on sampled bricks, and handed over in the reference's own form
    world = vx.World.from_noise(X, Y, Z, FACTOR, device=0)
    rt = vx.VoxelRaytracer3D(world, 16)
    _, cam_pos = camera_for(vx, rt, wl["dims"], camera_spec(wl))
    tan_half = device_tan_half(vx)
    check = world_check(vx, ol, world, wl["dims"])
    top, bw, b6 = world.download()
    rt.close(); world.close()
    ow = ol.OracleWorld.from_brickmap(top, bw, b6, X, Y, Z, FACTOR)
    cam = ol.make_camera(cam_pos, camera_spec(wl)["euler"], 90.0, tan_half=tan_half)
    env = ol.default_env()
    opts = ol.RenderOpts(0, 0, 0, 1, 1, ol.NORM_HOST)
    try:
        ref = ol.RefWorld(top, bw, b6, X, Y, Z, FACTOR)
    except Exception:
        ref = None
    kind = "reference" if ref is not None else "port"
    # the port is ~1.5x faster than the reference's own code: size the sample for ~2 s per step of the slower one
    stride = pick_stride(ol, ow, cam, env, opts, W, H, 1.2 if ref is not None else 2.0, threads)
    if ref is not None:
        samples = [ow.frame_rays(cam, env, opts, W, H, k % stride, H, stride) for k in range(min(stride, args.steps + args.warmup))]

        def step(k):
            o, d = samples[k % len(samples)]
            ref.trace(o, d, 2048, threads)
            return o.shape[0]
    else:
        def step(k):
            st = ow.render_sample(cam, env, opts, W, H, k % stride, H, stride, threads)
            return int(st[0] + st[2])
    for k in range(args.warmup):
        step(k)
    rays = 0
    t0 = time.perf_counter()
    for k in range(args.steps):
        rays += step(args.warmup + k)
    dt = time.perf_counter() - t0
    val = rays / dt / 1e6
    sample = "every %d-th pixel row of the %dx%d frame per step (%d rays/step: primaries + one sun ray per hit), same world/camera as the GPU arm" % (
        stride, W, H, rays // max(args.steps, 1))
    what = ("the reference's GPUDDA::Raytrace (VolumeRaytracer.cu:354-525) compiled for the host, one std::thread per core"
            if ref is not None else "CPU port of the reference's march (oracle/)")
    line = {"impl": "reference", "metric": "Mrays/s", "value": val, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt / max(args.steps, 1) * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl_name + ": " + wl["desc"], "camera": cam_pos, "timing": "host wall clock, CPU only", "what": what,
                       "world": check},
            "cpu_baseline": {"value": val, "unit": "Mrays/s", "cores": threads, "kind": kind, "sample": sample},
            "e2e": {"value": val, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


# ---------------------------------------------------------------------------------------------------------------------
# ours
# ---------------------------------------------------------------------------------------------------------------------
class Bench:
    def __init__(self, args):
        import torch
        import torch.distributed as dist
        import voxelengine_b200 as vx
        self.torch, self.dist, self.vx, self.args = torch, dist, vx, args
        self.rank = int(os.environ.get("RANK", "0"))
        self.world_size = int(os.environ.get("WORLD_SIZE", "1"))
        self.dev = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback")
        torch.cuda.set_device(self.dev)
        if self.world_size > 1:
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.dev))
        self.stream = torch.cuda.current_stream().cuda_stream
        self.flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device="cuda")
        self.stats = torch.zeros(8, dtype=torch.int64, device="cuda")
        self.env = vx.default_env()
        self.launches = 0

    def barrier(self):
        if self.world_size > 1:
            self.dist.barrier()

    def opts(self, bands=(0, 0), stats=False, norm=None, shadows=1, rows=None, W=0):
        vx = self.vx
        y0, h = rows if rows else (0, 0)
        return vx.RenderOpts(0, 0, 0, shadows, 1, vx.NORM_DEVICE if norm is None else norm, 0, y0, (W + 31) // 32 * 32 if rows else 0, h,
                             self.stats.data_ptr() if stats else None, bands[0], bands[1])

    def count_frame(self, rt, cam, W, H, bands):
        """Counters of exactly this frame from the counting variants of the kernels (untimed): primary pass alone, then all."""
        torch = self.torch
        tmp = torch.empty(W * H * 4 + 8 * W * 4, dtype=torch.uint8, device="cuda")
        out = []
        for shadows in (0, 1):
            self.stats.zero_()
            rt.render_device(cam, self.env, self.opts(bands, stats=True, shadows=shadows), tmp.data_ptr(), W, H, stream=self.stream)
            torch.cuda.synchronize()
            out.append(self.stats.clone())
        return out  # [primary only, whole frame] for this rank's bands

    def timed(self, step, K, warm, rt=None, solo=False):
        """K timed steps (CUDA events on the launching stream, L2 flushed before each), after `warm` untimed ones.
        Returns (sum of step ms, per-kernel ms sums or None).  solo: only this rank works (no barriers)."""
        torch = self.torch
        for _ in range(warm):
            self.flush.fill_(1)
            step()
        torch.cuda.synchronize()
        if not solo:
            self.barrier()
        torch.cuda.synchronize()
        total, kern = 0.0, [0.0, 0.0, 0.0, 0.0]
        for _ in range(K):
            self.flush.fill_(1)  # L2 flush between timed iterations (512 MiB write > 126 MB L2)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            step()
            e1.record()
            e1.synchronize()
            total += e0.elapsed_time(e1)
            if rt is not None:
                kern = [a + b for a, b in zip(kern, rt.frame_times())]
        torch.cuda.synchronize()
        if not solo:
            self.barrier()
        return total, kern

    def max_over_ranks(self, values):
        t = self.torch.tensor(values, dtype=self.torch.float64, device="cuda")
        if self.world_size > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return [float(v) for v in t.cpu()]

    def gather_ranks(self, values):
        t = self.torch.tensor(values, dtype=self.torch.float64, device="cuda")
        if self.world_size == 1:
            return [[float(v) for v in t.cpu()]]
        out = [self.torch.zeros_like(t) for _ in range(self.world_size)]
        self.dist.all_gather(out, t)
        return [[float(v) for v in o.cpu()] for o in out]


def measure_single(B, rt, cam, W, H, K, warm):
    """One GPU, whole frame into a local buffer: counters, timed steps, per-kernel times."""
    torch = B.torch
    frame = torch.zeros(W * H * 4 + 8 * W * 4, dtype=torch.uint8, device="cuda")
    st_prim, st_all = B.count_frame(rt, cam, W, H, (0, 0))
    o = B.opts()
    rt.set_timing(True)
    total, kern = B.timed(lambda: rt.render_device(cam, B.env, o, frame.data_ptr(), W, H, stream=B.stream), K, warm, rt, solo=True)
    rt.set_timing(False)
    B.launches += 4 * K
    return dict(frame=frame, st_prim=st_prim.cpu().numpy(), st=st_all.cpu().numpy(), ms=total / K, kern=[k / K for k in kern])


def frame_summary(m, names):
    st = m["st"]
    rays = int(st[0] + st[2])
    return {"ms_per_step": m["ms"], "value": rays / (m["ms"] * 1e-3) / 1e6, "unit": "Mrays/s", "rays_per_step": rays, "primary_rays": int(st[0]),
            "primary_hits": int(st[1]), "shadow_rays": int(st[2]), "shadow_hits": int(st[3]), "steps_per_ray": float(st[4]) / max(rays, 1),
            "kernels_ms": dict({n: m["kern"][i] for i, n in enumerate(names)}, frame=sum(m["kern"]))}


def rooflines(m, names, l2_peak, l2_clocks, hbm_peak, peak_src, traffic):
    """bytes(ray) = 4 C + 8 K + 64 E + P, P = 4 B per pixel for the primary rays of the screen path and 0
    for sun rays.  The records the four kernels of a frame hand to each other are reported beside it, not inside it."""
    sp, st = m["st_prim"], m["st"]
    p_rays, p_hits, s_rays = int(sp[0]), int(sp[1]), int(st[2])
    alg = {"primary_rays": int(4 * sp[5] + 8 * sp[6] + 64 * sp[7] + 4 * p_rays),
           "shadow_rays": int(4 * (st[5] - sp[5]) + 8 * (st[6] - sp[6]) + 64 * (st[7] - sp[7]))}
    hand = {"primary_rays": 16 * p_hits, "shadow_rays": 52 * s_rays}
    formula = {"primary_rays": "4*C + 8*K + 64*E of the primary rays + 4 B per pixel (SURVEY.md 8d)",
               "shadow_rays": "4*C + 8*K + 64*E of the sun rays, P = 0 (SURVEY.md 8d); C, K, E include the first brick, which shadow_setup marches"}
    out = {}
    for k in ("primary_rays", "shadow_rays"):
        ms = m["kern"][names.index(k)]
        a = alg[k] / (ms * 1e-3) / 1e9
        out[k] = {"bound": "l2", "achieved": a, "peak": l2_peak, "unit": "GB/s", "frac": a / l2_peak,
                  "traffic": (traffic or {}).get(k) if isinstance(traffic, dict) else None,
                  "peak_source": "measured in this run: all-SM ld.global.cg read of a 32 MiB (L2-resident) buffer, best of 4", "peak_clocks": l2_clocks,
                  "hbm_peak": hbm_peak, "hbm_peak_source": peak_src, "frac_of_hbm": a / hbm_peak,
                  "kernel": "vx::" + k, "kernel_ms": ms, "algorithmic_bytes_per_launch": alg[k], "formula": formula[k],
                  "handover_bytes": hand[k],
                  "handover": "hit queue entries written (16 B per primary hit)" if k == "primary_rays" else "lane record read + result written (48 + 4 B per sun ray)"}
    frame_bytes = alg["primary_rays"] + alg["shadow_rays"]
    return out, frame_bytes


def parity_leg(B, ol, ow, rt, cam_specs, W, H, threads, target_s):
    """The timed frame against the CPU oracle at its own scale: one untimed GPU frame per camera with the host definition of
    normalize (what the reference computes on a CPU) and a step plane, compared on a row sample the oracle walks in about
    target_s seconds: step plane bit-exact, same pixels written, RGB within 1/255."""
    torch, vx = B.torch, B.vx
    tan_half = device_tan_half(vx, B.dev)
    out = {}
    for name, (cam, pos, euler) in cam_specs.items():
        frame = torch.full((H, W, 4), 7, dtype=torch.uint8, device="cuda")
        steps = torch.full((H, W), -1, dtype=torch.int32, device="cuda")
        rt.render_device(cam, B.env, B.opts(norm=vx.NORM_HOST), frame.data_ptr(), W, H, d_steps=steps.data_ptr(), stream=B.stream)
        torch.cuda.synchronize()
        got, gsteps = frame.cpu().numpy(), steps.cpu().numpy()
        del frame, steps
        ocam = ol.make_camera(pos, euler, 90.0, tan_half=tan_half)
        oopts = ol.RenderOpts(0, 0, 0, 1, 1, ol.NORM_HOST)
        stride = pick_stride(ol, ow, ocam, ol.default_env(), oopts, W, H, target_s, threads)
        first = stride // 2
        ref = np.full((H, W, 4), 7, np.uint8)
        rsteps = np.full((H, W), -1, np.int32)
        t0 = time.perf_counter()
        ow.render_sample_planes(ocam, ol.default_env(), oopts, W, H, first, H, stride, ref, rsteps, threads)
        dt = time.perf_counter() - t0
        rows = np.arange(first, H, stride)
        diff = np.abs(got[rows].astype(np.int16) - ref[rows].astype(np.int16))
        out[name] = {"workload": "8192x512x8192, %dx%d" % (W, H), "camera": pos, "euler": euler, "rows_compared": int(rows.size),
                     "pixels_compared": int(rows.size * W), "steps_equal": bool(np.array_equal(gsteps[rows], rsteps[rows])),
                     "written_pixels_equal": bool(np.array_equal((got[rows] != 7).any(2), (ref[rows] != 7).any(2))),
                     "rgb_max_diff": int(diff.max()), "rgb_pixels_differing": int((diff > 0).any(2).sum()),
                     "max_steps_in_sample": int(rsteps[rows].max()), "oracle_seconds": dt,
                     "normalize": "host definition of rsqrtf on both sides (VX_NORM_HOST); the timed frames use the device rsqrtf, as the reference's kernels do"}
        ok = out[name]["steps_equal"] and out[name]["written_pixels_equal"] and out[name]["rgb_max_diff"] <= 1
        out[name]["ok"] = bool(ok)
    return out


def cpu_baseline_leg(B, ol, ow, refw, cam_pos, euler, W, H, threads):
    """N = 1, rank 0: the reference's own Raytrace compiled for the host (kind "reference") and the CPU port (oracle) on a
    bounded row sample of the same frame, about 10 s each."""
    vx = B.vx
    tan_half = device_tan_half(vx, B.dev)
    cam = ol.make_camera(cam_pos, euler, 90.0, tan_half=tan_half)
    env, opts = ol.default_env(), ol.RenderOpts(0, 0, 0, 1, 1, ol.NORM_HOST)
    stride = pick_stride(ol, ow, cam, env, opts, W, H, 8.0, threads)
    rays, passes, st = 0, 0, None
    t0 = time.perf_counter()
    while passes == 0 or (time.perf_counter() - t0 < 8.0 and passes < 64):
        st = ow.render_sample(cam, env, opts, W, H, (stride // 2 + passes) % stride, H, stride, threads)
        rays += int(st[0] + st[2])
        passes += 1
    dt = time.perf_counter() - t0
    port = rays / dt / 1e6
    out = {"value": port, "unit": "Mrays/s", "cores": threads, "kind": "port",
           "sample": "every %d-th pixel row of the %dx%d frame, %d passes (%d rays, %.1f s), same world and camera" % (stride, W, H, passes, rays, dt),
           "steps_per_ray": float(st[4]) / max(int(st[0] + st[2]), 1)}
    if refw is not None:
        o, d = ow.frame_rays(cam, env, opts, W, H, stride // 2, H, stride)
        n, reps = o.shape[0], 0
        t0 = time.perf_counter()
        while reps == 0 or (time.perf_counter() - t0 < 8.0 and reps < 64):
            refw.trace(o, d, 2048, threads)
            reps += 1
        dt = time.perf_counter() - t0
        out.update(value=n * reps / dt / 1e6, kind="reference", port_value=port,
                   sample="the %d rays (primaries + one sun ray per hit) of every %d-th pixel row of the %dx%d frame, %d passes (%.1f s), through the "
                          "reference's GPUDDA::Raytrace compiled for the host (oracle/_ref/libvxref_host.so); port_value = the oracle's renderer on the same rows"
                          % (n, stride, W, H, reps, dt))
    return out


def e2e_batch(B, rt_world, cam, W, H):
    """The host-array batch call (VoxelRaytracer3D::Raytrace(vector, vector), VolumeRaytracer.cu:574-618): the 1920x1080
    primary rays of the overview camera as host arrays through vx_trace, the planes the reference returns coming back."""
    vx = B.vx
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from raysets import camera_rays
    w, h = 1920, 1080
    o, d = camera_rays(tuple(cam.origin), np.array(cam.fwd, np.float32), np.array(cam.up, np.float32), np.array(cam.right, np.float32), w, h,
                       device_tan_half(vx, B.dev))
    n = o.shape[0]
    rt = vx.VoxelRaytracer3D(rt_world, n)
    planes = rt.REFERENCE_PLANES
    out = rt.raytrace(o, d, planes=planes)           # first call builds the staging ring (untimed)
    out = rt.raytrace(o, d, planes=planes, out=out)
    reps = 5
    t0 = time.perf_counter()
    for _ in range(reps):
        rt.raytrace(o, d, planes=planes, out=out)
    dt = (time.perf_counter() - t0) / reps
    rt.close()
    B.launches += reps * ((n + 262143) // 262144)
    return {"value": n / dt / 1e6, "unit": "Mrays/s", "rays": n, "ms_per_call": dt * 1e3, "h2d_bytes_per_call": n * 24, "d2h_bytes_per_call": n * 29,
            "planes": list(planes), "hit_rate": float(out["hit"].mean()),
            "what": "vx_trace with pageable host arrays (numpy): chunked H2D / march / D2H pipeline through a page-locked ring"}


def config5(B, rt, world, cam, W, H, K):
    """BASELINE configs[4]: the 1-spp diffuse-bounce (AO sample) ray of every primary hit this rank owns, traced with the full
    step cap; rays stay on the GPU that owns the pixel.  Returns this rank's (rays, ms per pass)."""
    torch, vx = B.torch, B.vx
    bands = (B.world_size, B.rank) if B.world_size > 1 else (0, 0)
    n_max = W * H // max(B.world_size, 1) + 16 * W
    d_o = torch.zeros(n_max * 3, dtype=torch.float32, device="cuda")
    d_d = torch.zeros(n_max * 3, dtype=torch.float32, device="cuda")
    d_p = torch.zeros(n_max, dtype=torch.int32, device="cuda")
    d_n = torch.zeros(1, dtype=torch.int64, device="cuda")
    rt.bounce_rays(cam, B.opts(bands), W, H, d_o.data_ptr(), d_d.data_ptr(), d_p.data_ptr(), d_n.data_ptr(), stream=B.stream)
    torch.cuda.synchronize()
    n = int(d_n.item())
    hit = torch.zeros(n, dtype=torch.uint8, device="cuda")
    steps = torch.zeros(n, dtype=torch.int32, device="cuda")
    planes = dict(hit=hit.data_ptr(), steps=steps.data_ptr())

    def step():
        rt.raytrace_device(d_o.data_ptr(), d_d.data_ptr(), n, planes, stream=B.stream)
    total, _ = B.timed(step, K, 3)
    B.launches += K
    return n, total / K, float(steps.float().mean().item()), float(hit.float().mean().item())


def run_ours(args, wl, wl_name):
    import ctypes as C
    B = Bench(args)
    torch, dist, vx = B.torch, B.dist, B.vx
    from voxelengine_b200.partition import FramePartition, SharedFrame, HostFrame
    N, rank = B.world_size, B.rank
    X, Y, Z = wl["dims"]
    W, H = wl["W"], wl["H"]
    K, warm = args.steps, max(args.warmup, 3)
    world = vx.World.from_noise(X, Y, Z, FACTOR, device=B.dev)
    info = world.info
    rt = vx.VoxelRaytracer3D(world, 16)
    cam_spec = camera_spec(wl)
    cam, cam_pos = camera_for(vx, rt, wl["dims"], cam_spec)
    names = list(vx.VoxelRaytracer3D.KERNELS)
    hbm_peak, peak_src, sm_max = read_peaks()
    line = {}

    # ---------------- N = 1 ----------------
    if N == 1:
        sampler = ClockSampler(B.dev)
        wall0 = time.perf_counter()
        m = measure_single(B, rt, cam, W, H, K, warm)
        wall = time.perf_counter() - wall0
        clocks = sampler.stop()
        s = frame_summary(m, names)
        # e2e: the reference-facing call with a HOST frame (RenderScreen + cudaMemcpy of main.cu:165-167)
        host = torch.empty(H * W * 4, dtype=torch.uint8).pin_memory()
        o_timed = B.opts()
        vx._check(vx.lib().vx_render_host(rt.h, C.byref(cam), C.byref(B.env), C.byref(o_timed), host.data_ptr(), W, H))
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(K):
            vx._check(vx.lib().vx_render_host(rt.h, C.byref(cam), C.byref(B.env), C.byref(o_timed), host.data_ptr(), W, H))
        e2e_s = time.perf_counter() - t0
        B.launches += 8 * K  # two slices of four kernels
        host_equal = bool(torch.equal(host.view(-1)[: W * H * 4], m["frame"][: W * H * 4].cpu()))
        if not host_equal:
            raise SystemExit("bench: the host frame of vx_render_host differs from the device frame of vx_render")
        sampler = ClockSampler(B.dev, smi=False)
        l2_peak = vx.measure_read_bandwidth(32 << 20, 50, B.dev)
        l2_clocks = sampler.stop()
        traffic = None
        try:
            with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
                traffic = json.load(f).get(wl_name)
        except Exception:
            pass
        roofs, frame_bytes = rooflines(m, names, l2_peak, {k: l2_clocks[k] for k in ("sm_mhz", "sm_max_mhz", "reasons", "samples")}, hbm_peak, peak_src, traffic)
        line.update({
            "metric": "Mrays/s", "value": s["value"], "unit": "Mrays/s", "n_gpus": 1, "steps": K, "warmup": warm, "ms_per_step": s["ms_per_step"],
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl_name + ": " + wl["desc"], "camera": cam_pos, "euler": cam_spec["euler"], "fov_deg": 90,
                       "rays_per_step": s["rays_per_step"], "primary_rays": s["primary_rays"], "shadow_rays": s["shadow_rays"],
                       "steps_per_ray": s["steps_per_ray"], "parallelism": "one GPU", "exchange": "none",
                       "l2": "flushed between timed steps (512 MiB write)", "normalize": "device rsqrtf (VX_NORM_DEVICE), as the reference's kernels",
                       "world": {"occupied_bricks": int(info.occupied_bricks), "bricks": int(info.bricks), "pool_mb": info.bytes_pool / 1e6,
                                 "build_ms": info.build_ms}},
            "clocks": clocks,
            "e2e": {"value": s["rays_per_step"] * K / e2e_s / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": C.sizeof(vx.Camera) + C.sizeof(vx.Env) + C.sizeof(vx.RenderOpts),
                    "d2h_bytes_per_step": W * H * 4, "ms_per_step": e2e_s / K * 1e3, "host_frame_equals_device_frame": host_equal,
                    "what": "vx_render_host: frame rendered and copied into pinned host memory, host wall clock"},
            "kernels_ms": s["kernels_ms"],
            "roofline": roofs["primary_rays"], "roofline_shadow_rays": roofs["shadow_rays"],
            "frame_algorithmic_bytes": frame_bytes,
            "frame_roofline": {"achieved": frame_bytes / (s["ms_per_step"] * 1e-3) / 1e9, "unit": "GB/s",
                               "frac_of_l2": frame_bytes / (s["ms_per_step"] * 1e-3) / 1e9 / l2_peak, "frac_of_hbm": frame_bytes / (s["ms_per_step"] * 1e-3) / 1e9 / hbm_peak},
            "wall_s_timed_region": wall,
        })
        del host
        # the world-build kernel (noise_tiles, PopulateVoxels' replacement): ALU roofline of SURVEY.md 8d
        alu_peak = 148 * 128 * sm_max * 1e6
        voxels = float(X) * Y * Z
        full = vx.World.from_noise(1024, 256, 1024, FACTOR, device=B.dev, full_octaves=True)
        fi = full.info
        full.close()
        line["roofline_build"] = {
            "bound": "alu", "kernel": "vx::noise_tiles", "unit": "Gop/s", "peak": alu_peak / 1e9, "peak_source": "148 SMs x 128 lanes x %.0f MHz (MEASURED_PEAKS.json sm_max_mhz)" % sm_max,
            "ops_per_voxel": NOISE_OPS_PER_VOXEL, "formula": "32 octaves x ~400 scalar operations per voxel (SURVEY.md 8d)",
            "full_octaves": {"world": "1024x256x1024", "kernel_ms": fi.tiles_ms, "gvoxels_s": 1024.0 * 256 * 1024 / (fi.tiles_ms * 1e-3) / 1e9,
                             "achieved": NOISE_OPS_PER_VOXEL * 1024.0 * 256 * 1024 / (fi.tiles_ms * 1e-3) / 1e9,
                             "frac": NOISE_OPS_PER_VOXEL * 1024.0 * 256 * 1024 / (fi.tiles_ms * 1e-3) / alu_peak},
            "early_out": {"world": "8192x512x8192", "kernel_ms": info.tiles_ms, "build_ms": info.build_ms, "gvoxels_s": voxels / (info.tiles_ms * 1e-3) / 1e9,
                          "effective": NOISE_OPS_PER_VOXEL * voxels / (info.tiles_ms * 1e-3) / 1e9,
                          "effective_frac": NOISE_OPS_PER_VOXEL * voxels / (info.tiles_ms * 1e-3) / alu_peak,
                          "note": "exact early-out: the un-skipped operation count is kept as the numerator (SURVEY.md 8d), so > 1 means work was skipped"}}
        # camera (b) "grazing": the worst case for coarse steps, same world, same frame size
        cam_b, pos_b = camera_for(vx, rt, wl["dims"], CAMERAS["grazing"])
        if wl["camera"] == "overview":
            mb = measure_single(B, rt, cam_b, W, H, K, warm)
            sb = frame_summary(mb, names)
            rb, fb = rooflines(mb, names, l2_peak, None, hbm_peak, peak_src, None)
            line["grazing"] = dict(sb, camera=pos_b, euler=CAMERAS["grazing"]["euler"],
                                   roofline_frac_of_l2={k: rb[k]["frac"] for k in rb}, roofline_frac_of_hbm={k: rb[k]["frac_of_hbm"] for k in rb})
            del mb
        line["e2e_batch"] = e2e_batch(B, world, cam, W, H)
        # configs[4]: bounce rays
        nb, ms_b, steps_b, hit_b = config5(B, rt, world, cam, W, H, max(K // 2, 3))
        line["config5"] = {"bounce": {"value": nb / (ms_b * 1e-3) / 1e6, "unit": "Mrays/s", "rays": nb, "ms": ms_b, "steps_per_ray": steps_b, "hit_rate": hit_b,
                                      "what": "1-spp AO sample ray (Renderer.cu:121-165) of every primary hit of the frame, full step cap, trace_rays kernel"},
                           "build": {"gvoxels_s": voxels / (info.build_ms * 1e-3) / 1e9, "build_ms": info.build_ms, "world": "8192x512x8192, replicated build"}}
        # "the kernel to beat" (BASELINE.md O1): the reference's own kernels, compiled from its sources into oracle/_ref, on this GPU
        if not args.no_ref_kernel:
            try:
                sys.path.insert(0, os.path.join(ROOT, "tools"))
                import ref_compare
                if ref_compare.available():
                    sampler = ClockSampler(B.dev, smi=False)
                    rk = ref_compare.compare(device=B.dev)
                    rk["clocks"] = sampler.stop()
                    line["ref_kernel"] = rk
                    B.launches += 4 * 16
                else:
                    line["ref_kernel"] = {"unavailable": "oracle/_ref/ref_gpu not built (reference sources absent at build time)"}
            except Exception as e:  # a diagnostic beside the headline: never takes the bench down
                line["ref_kernel"] = {"error": repr(e)[:300]}
        if not args.no_cpu_baseline:
            sys.path.insert(0, os.path.join(ROOT, "tests"))
            import oracle_lib as ol
            threads = os.cpu_count() or 1
            line["config"]["world"]["check"] = world_check(vx, ol, world, wl["dims"])
            top, bw, b6 = world.download()
            ow = ol.OracleWorld.from_brickmap(top, bw, b6, X, Y, Z, FACTOR)
            specs = {wl["camera"] if isinstance(wl["camera"], str) else "camera": (cam, cam_pos, cam_spec["euler"])}
            if wl["camera"] == "overview":
                specs["grazing"] = (cam_b, pos_b, CAMERAS["grazing"]["euler"])
            line["parity"] = parity_leg(B, ol, ow, rt, specs, W, H, threads, 6.0)
            try:
                refw = ol.RefWorld(top, bw, b6, X, Y, Z, FACTOR)
            except Exception:
                refw = None
            line["cpu_baseline"] = cpu_baseline_leg(B, ol, ow, refw, cam_pos, cam_spec["euler"], W, H, threads)
            if not all(p["ok"] for p in line["parity"].values()):
                print(json.dumps({"parity": line["parity"]}), file=sys.stderr)
                raise SystemExit("bench: the timed frame does not match the CPU oracle (see parity above); no value printed")
        line["gpu_launches"] = B.launches
        emit(line)
        return

    # ---------------- N > 1 ----------------
    def banded(wname):
        """The frame of workload `wname` banded over the N ranks into rank 0's HBM: timed, verified, with its single-GPU base."""
        w = WORKLOADS[wname]
        Wn, Hn = w["W"], w["H"]
        part = FramePartition(Wn, Hn, N, rank)
        bands = part.render_opts()
        shared = None
        if args.exchange == "shared":
            try:
                shared = SharedFrame(part, B.dev)
            except RuntimeError as e:  # raised on every rank together: fall back to the NCCL gather
                if rank == 0:
                    print("bench: shared frame unavailable (%s); using the NCCL gather" % e, file=sys.stderr)
        local = part.alloc_frame("cuda")
        if shared is not None:
            if "VX_FUSED_SHADE" in os.environ:
                rt.set_fused_shade(int(os.environ["VX_FUSED_SHADE"]))
            else:
                shared.attach(rt)
        target = shared.ptr if shared is not None else local.data_ptr()
        o = B.opts(bands)

        def step():
            rt.render_device(cam, B.env, o, target, Wn, Hn, stream=B.stream)
            if shared is not None:
                shared.fence()      # every rank's pixel stores are in rank 0's frame; the only collective
            else:
                part.gather(local)
        st_prim, st_all = B.count_frame(rt, cam, Wn, Hn, bands)
        st = st_all.clone()
        dist.all_reduce(st)
        st = st.cpu().numpy()
        rays = int(st[0] + st[2])
        rt.set_timing(True)
        sampler = ClockSampler(B.dev) if rank == 0 else None
        total, kern = B.timed(step, K, warm, rt)
        clocks = sampler.stop() if sampler else None
        rt.set_timing(False)
        B.launches += 4 * K
        step_ms = B.max_over_ranks([total])[0] / K
        per_rank = B.gather_ranks([k / K for k in kern] + [total / K])
        # ---- the assembled frame against rank 0's own whole-frame render ----
        nbytes = Wn * Hn * 4
        if rank == 0:
            if shared is not None:
                vx.device_fill(shared.ptr, 0, part.frame_bytes(), B.dev)
            else:
                local.zero_()
        torch.cuda.synchronize()
        B.barrier()
        step()
        torch.cuda.synchronize()
        B.barrier()
        identical, crc = None, None
        base_ms, base_kern = None, None
        if rank == 0:
            whole = torch.zeros(part.frame_bytes(), dtype=torch.uint8, device="cuda")
            rt.set_fused_shade(0)
            rt.render_device(cam, B.env, B.opts(), whole.data_ptr(), Wn, Hn, stream=B.stream)
            torch.cuda.synchronize()
            if shared is not None:
                got = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
                shared.to_host(got, nbytes)
            else:
                got = local[:nbytes].cpu()
            ref = whole[:nbytes].cpu()
            identical = bool(torch.equal(got, ref))
            crc = "%08x" % zlib.crc32(ref.numpy().tobytes())
            # ---- like-for-like base: the same frame on rank 0 alone, measured now, same flush, same clocks ----
            rt.set_timing(True)
            o1 = B.opts()
            total1, kern1 = B.timed(lambda: rt.render_device(cam, B.env, o1, whole.data_ptr(), Wn, Hn, stream=B.stream), K, warm, rt, solo=True)
            rt.set_timing(False)
            B.launches += 4 * K
            base_ms, base_kern = total1 / K, [k / K for k in kern1]
            del whole
        B.barrier()
        flag = torch.tensor([1 if (identical is None or identical) else 0], dtype=torch.int32, device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) == 0:
            raise SystemExit("bench: the frame assembled from %d ranks differs from the single-GPU render of %s; no value printed" % (N, wname))
        # ---- e2e: the frame delivered to HOST memory, every rank copying its own bands down its own PCIe link ----
        if shared is not None:
            shared.close()
        rt.set_fused_shade(0)
        hostframe = None
        try:
            hostframe = HostFrame(part, B.dev)
        except RuntimeError as e:
            if rank == 0:
                print("bench: shared host frame unavailable (%s); rank 0 copies the gathered frame" % e, file=sys.stderr)
        host = torch.empty(nbytes, dtype=torch.uint8).pin_memory() if (hostframe is None and rank == 0) else None
        e2e_opts = {}

        def e2e_rows(y0, h):
            oo = e2e_opts.get((y0, h))
            if oo is None:
                oo = e2e_opts[(y0, h)] = B.opts(bands, rows=(y0, h), W=Wn)
            rt.render_device(cam, B.env, oo, local.data_ptr(), Wn, Hn, stream=B.stream)

        def e2e_step():
            if hostframe is None:
                rt.render_device(cam, B.env, o, local.data_ptr(), Wn, Hn, stream=B.stream)
                part.gather(local)
                if rank == 0:
                    host.copy_(local[:nbytes], non_blocking=True)
                torch.cuda.synchronize()
                return
            if args.host_split >= 1.0:
                rt.render_device(cam, B.env, o, local.data_ptr(), Wn, Hn, stream=B.stream)
                hostframe.store(local.data_ptr(), B.stream)
            else:
                hostframe.render_and_store(e2e_rows, local.data_ptr(), args.host_split)
            hostframe.fence()
        if hostframe is not None and rank == 0:
            hostframe.array[:] = 0
        B.barrier()
        e2e_step()
        B.barrier()
        host_identical = None
        if rank == 0:
            arr = hostframe.array.reshape(-1) if hostframe is not None else host.numpy()
            host_identical = bool(zlib.crc32(arr[:nbytes].tobytes()) == int(crc, 16))
        torch.cuda.synchronize()
        B.barrier()
        t0 = time.perf_counter()
        for _ in range(K):
            e2e_step()
        B.barrier()
        e2e_s = time.perf_counter() - t0
        B.launches += (8 if args.host_split < 1.0 and hostframe is not None else 4) * K
        # the same with two host frames: the ranks start frame k + 1 before anybody waits for frame k (a consumer that reads
        # one frame while the next one renders); every frame is still delivered whole, and the last pair is checked
        piped_s, piped_identical = None, None
        if hostframe is not None:
            second, local2 = None, part.alloc_frame("cuda")
            try:
                second = HostFrame(part, B.dev)
            except RuntimeError:
                second = None
            if second is not None:
                frames, locals_ = [hostframe, second], [local, local2]
                if rank == 0:
                    hostframe.array[:] = 0; second.array[:] = 0
                B.barrier()

                def piped(k, pending):
                    hf, loc = frames[k % 2], locals_[k % 2]

                    def rows(y0, h):
                        oo = e2e_opts.get((y0, h))
                        if oo is None:
                            oo = e2e_opts[(y0, h)] = B.opts(bands, rows=(y0, h), W=Wn)
                        rt.render_device(cam, B.env, oo, loc.data_ptr(), Wn, Hn, stream=B.stream)
                    if args.host_split >= 1.0:
                        rt.render_device(cam, B.env, o, loc.data_ptr(), Wn, Hn, stream=B.stream)
                        hf.store(loc.data_ptr(), B.stream)
                        ev = hf.mark(torch.cuda.current_stream())
                    else:
                        hf.render_and_store(rows, loc.data_ptr(), args.host_split)
                        ev = hf.mark()
                    if pending is not None:
                        pending[0].wait(pending[1])
                    return (hf, ev)
                pend = None
                for k in range(2):
                    pend = piped(k, pend)
                pend[0].wait(pend[1])
                B.barrier()
                t0 = time.perf_counter()
                pend = None
                for k in range(K):
                    pend = piped(k, pend)
                pend[0].wait(pend[1])
                B.barrier()
                piped_s = time.perf_counter() - t0
                B.launches += (8 if args.host_split < 1.0 else 4) * K
                if rank == 0:
                    piped_identical = all(zlib.crc32(hf.array.reshape(-1)[:nbytes].tobytes()) == int(crc, 16) for hf in frames)
                second.close()
        if hostframe is not None:
            hostframe.close()
        flag = torch.tensor([1 if (host_identical is None or host_identical) else 0], dtype=torch.int32, device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) == 0:
            raise SystemExit("bench: the HOST frame assembled from %d ranks differs from the single-GPU render of %s; no value printed" % (N, wname))
        out = None
        if rank == 0:
            kn = names + ["step"]
            out = {"workload": wname + ": " + w["desc"], "ms_per_step": step_ms, "value": rays / (step_ms * 1e-3) / 1e6, "unit": "Mrays/s",
                   "rays_per_step": rays, "primary_rays": int(st[0]), "shadow_rays": int(st[2]), "steps_per_ray": float(st[4]) / rays,
                   "frame_identical": identical, "host_frame_identical": host_identical, "frame_crc32": crc,
                   "exchange": ("peer stores into rank 0's frame (CUDA IPC over NVLink) + one-word all-reduce fence" if args.exchange == "shared" and shared is not None
                                else "torch.distributed gather to rank 0"),
                   "kernels_ms_per_rank": {"min": {n: min(r[i] for r in per_rank) for i, n in enumerate(kn)},
                                           "max": {n: max(r[i] for r in per_rank) for i, n in enumerate(kn)}},
                   "scaling_base": {"workload": wname, "n1_ms": base_ms, "n1_value": rays / (base_ms * 1e-3) / 1e6,
                                    "n1_kernels_ms": {n: base_kern[i] for i, n in enumerate(names)},
                                    "how": "rank 0 alone renders the whole frame in this same run (other ranks wait at a barrier): same world, camera, L2 flush, clocks"},
                   "speedup_like_for_like": base_ms / step_ms, "efficiency_like_for_like": base_ms / step_ms / N,
                   "e2e": {"value": rays * K / e2e_s / 1e6, "unit": "Mrays/s", "ms_per_step": e2e_s / K * 1e3, "d2h_bytes_per_step": nbytes,
                           "h2d_bytes_per_step": (C.sizeof(vx.Camera) + C.sizeof(vx.Env) + C.sizeof(vx.RenderOpts)) * N,
                           "what": "every rank copies its bands into one page-locked host frame shared by the ranks' processes (HostFrame); each frame is waited for before the next one starts",
                           "two_frames_in_flight": None if piped_s is None else {
                               "value": rays * K / piped_s / 1e6, "unit": "Mrays/s", "ms_per_step": piped_s / K * 1e3, "host_frames_identical": piped_identical,
                               "what": "two host frames: frame k + 1 is launched before frame k is waited for, so the inter-process barrier overlaps rendering"}},
                   "clocks": clocks}
        return out

    wall0 = time.perf_counter()
    main_name = wl_name
    res = banded(main_name)
    other = None
    if args.workload is None:   # default run: configs[3] (8K) beside the 4K headline
        other = banded("w2-8k")
    # configs[4] on N ranks: bounce rays of this rank's pixels + replicated build time (max over ranks)
    nb, ms_b, steps_b, hit_b = config5(B, rt, world, cam, W, H, max(K // 2, 3))
    tot = torch.tensor([nb], dtype=torch.int64, device="cuda")
    dist.all_reduce(tot)
    # configs[4], world build: every rank building its own replica (what the frames above ran on) against the z-slab build,
    # where each rank runs the voxel kernel on 1/N of the bricks and the parts are all-gathered over NVLink
    from voxelengine_b200.partition import build_world_partitioned
    torch.cuda.synchronize()
    B.barrier()
    t0 = time.perf_counter()
    pworld, ptimes = build_world_partitioned(X, Y, Z, FACTOR, device=B.dev)
    torch.cuda.synchronize()
    B.barrier()
    part_ms = (time.perf_counter() - t0) * 1e3
    rng = np.random.default_rng(77)
    cc = np.stack([rng.integers(0, X // FACTOR, 8192), rng.integers(0, Y // FACTOR, 8192), rng.integers(0, Z // FACTOR, 8192)], 1).astype(np.uint32)
    same = (zlib.crc32(pworld.download_top().tobytes()) == zlib.crc32(world.download_top().tobytes())
            and (pworld.info.occupied_bricks, pworld.info.solid_voxels) == (info.occupied_bricks, info.solid_voxels)
            and all(np.array_equal(a, b) for a, b in zip(pworld.download_bricks(cc), world.download_bricks(cc))))
    prt = vx.VoxelRaytracer3D(pworld, 16)
    chk = torch.zeros(W * H * 4 + 8 * W * 4, dtype=torch.uint8, device="cuda")
    prt.render_device(cam, B.env, B.opts(), chk.data_ptr(), W, H, stream=B.stream)
    torch.cuda.synchronize()
    pcrc = "%08x" % zlib.crc32(chk[: W * H * 4].cpu().numpy().tobytes())
    prt.close(); pworld.close(); del chk
    crc_ok = rank != 0 or pcrc == res["frame_crc32"]   # the frame rendered on the partitioned world == the verified frame of `value`
    flag = torch.tensor([1 if (same and crc_ok) else 0], dtype=torch.int32, device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    part_ok = bool(int(flag.item()))
    ms_b_max, build_max, part_max, slab_max, exch_max, asm_max = B.max_over_ranks([ms_b, info.build_ms, part_ms, ptimes["slab"], ptimes["exchange"], ptimes["assemble"]])
    wall = time.perf_counter() - wall0
    if rank == 0:
        line.update({
            "metric": "Mrays/s", "value": res["value"], "unit": "Mrays/s", "n_gpus": N, "steps": K, "warmup": warm, "ms_per_step": res["ms_per_step"],
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": res["workload"], "camera": cam_pos, "euler": cam_spec["euler"], "fov_deg": 90, "rays_per_step": res["rays_per_step"],
                       "primary_rays": res["primary_rays"], "shadow_rays": res["shadow_rays"], "steps_per_ray": res["steps_per_ray"],
                       "parallelism": "8-row image bands dealt round-robin over %d GPUs, world replicated" % N, "exchange": res["exchange"],
                       "l2": "flushed between timed steps (512 MiB write)",
                       "world": {"occupied_bricks": int(info.occupied_bricks), "bricks": int(info.bricks), "pool_mb": info.bytes_pool / 1e6, "build_ms": info.build_ms}},
            "clocks": res["clocks"], "e2e": res["e2e"],
            "frame_identical": res["frame_identical"], "host_frame_identical": res["host_frame_identical"], "frame_crc32": res["frame_crc32"],
            "scaling_base": res["scaling_base"], "speedup_like_for_like": res["speedup_like_for_like"], "efficiency_like_for_like": res["efficiency_like_for_like"],
            "kernels_ms_per_rank": res["kernels_ms_per_rank"],
            "strong_4k": {k: res[k] for k in ("ms_per_step", "value", "scaling_base", "speedup_like_for_like", "efficiency_like_for_like", "frame_identical")} if main_name == "w2-4k" else None,
            "config4_8k": other,
            "config5": {"bounce": {"value": int(tot.item()) / (ms_b_max * 1e-3) / 1e6, "unit": "Mrays/s", "rays": int(tot.item()), "ms": ms_b_max,
                                   "what": "1-spp AO sample rays of each rank's own pixels (no ray exchange), max over ranks"},
                        "build": {"build_ms_replicated": build_max, "gvoxels_s_replicated": float(X) * Y * Z / (build_max * 1e-3) / 1e9,
                                  "build_ms_partitioned": part_max, "gvoxels_s_partitioned": float(X) * Y * Z / (part_max * 1e-3) / 1e9,
                                  "partitioned_phases_ms": {"slab_kernels": slab_max, "exchange": exch_max, "assemble": asm_max},
                                  "partitioned_world_identical": part_ok,
                                  "what": "replicated: every rank builds the whole world (device time, max over ranks); partitioned: z-slab per rank + "
                                          "all-gather of top-level bits, bounds and mask pools over NCCL + assembly on every rank (host wall clock between "
                                          "barriers, max over ranks); identical = same top-level bits, counters, 8192 sampled bricks and the same 4K frame"}},
            "gpu_launches": B.launches, "wall_s_timed_region": wall,
        })
        emit(line)
    dist.destroy_process_group()


_RESULT_FD = None


def claim_stdout():
    """stdout carries exactly ONE JSON line.  Libraries write there too (NCCL prints its version banner to stdout at NCCL_DEBUG=WARN
    as well as =VERSION, from every rank, and all ranks share torchrun's stdout), so file descriptor 1 is pointed at stderr for
    the whole run and the line is written to the original descriptor by emit()."""
    global _RESULT_FD
    if _RESULT_FD is None:
        sys.stdout.flush()
        _RESULT_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_RESULT_FD, data)


def main():
    claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, choices=list(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true", help="N = 1: skip the CPU legs (parity at scale, cpu_baseline)")
    ap.add_argument("--no-ref-kernel", action="store_true", help="N = 1: skip timing the reference's own kernels (oracle/_ref) beside ours")
    ap.add_argument("--exchange", default=os.environ.get("VX_EXCHANGE", "shared"), choices=["shared", "gather"],
                    help="N > 1: how the bands reach rank 0 (shared: peer stores into rank 0's frame; gather: NCCL gather)")
    ap.add_argument("--host-split", type=float, default=float(os.environ.get("VX_HOST_SPLIT_N", "0.8")),
                    help="N > 1, e2e: share of the rows rendered before the first copy to the host starts (1 = no overlap)")
    args = ap.parse_args()
    wl_name = args.workload or "w2-4k"
    wl = WORKLOADS[wl_name]
    if args.impl == "reference":
        run_reference(args, wl, wl_name)
        return
    run_ours(args, wl, wl_name)


if __name__ == "__main__":
    main()
