"""One rank's share of a frame rendered as K interleaved band sets on K streams (K tracers), against the same share on one stream:
does the next stage of one part fill the GPU while the other part's last tiles finish?
python tools/two_stream_frame.py [workload] [ranks N] [parts K ...]

The share is band set 0 of N (N = 1: the whole frame).  With K parts, part k renders band set k * N of K * N — together exactly
the bands of set 0 of N — on its own stream with its own tracer (a tracer owns one frame's queue).  Times are CUDA events on a
third stream that waits for all parts, L2 flushed before every frame; the frame is compared with the one-stream render."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import voxelengine_b200 as vx
import bench

name = sys.argv[1] if len(sys.argv) > 1 else "w2-4k"
N = int(sys.argv[2]) if len(sys.argv) > 2 else 1
parts = [int(v) for v in sys.argv[3:]] or [1, 2, 3, 4]
wl = bench.WORKLOADS[name]
world = vx.World.from_noise(*wl["dims"], 8)
tracers = [vx.VoxelRaytracer3D(world, 16) for _ in range(max(parts))]
cam, pos = bench.camera_for(vx, tracers[0], wl["dims"], bench.camera_spec(wl))
W, H = wl["W"], wl["H"]
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
streams = [torch.cuda.Stream() for _ in range(max(parts))]
main = torch.cuda.current_stream()
ref = None
for K in parts:
    buf = torch.zeros(H * W * 4, dtype=torch.uint8, device="cuda")
    ms = []
    for i in range(14):
        flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(main)
        done = []
        for k in range(K):
            streams[k].wait_event(e0)
            total, index = K * N, k * N
            opts = vx.RenderOpts(0, 0, 0, 1, 1, vx.NORM_DEVICE, 0, 0, 0, 0, None, total if total > 1 else 0, index if total > 1 else 0)
            tracers[k].render_device(cam, vx.default_env(), opts, buf.data_ptr(), W, H, stream=streams[k].cuda_stream)
            ev = torch.cuda.Event(); ev.record(streams[k]); done.append(ev)
        for ev in done:
            main.wait_event(ev)
        e1.record(main); e1.synchronize()
        if i >= 4:
            ms.append(e0.elapsed_time(e1))
    got = buf.cpu()
    if ref is None:
        ref = got
    print("share 1/%d in %d part(s): %.3f ms (min %.3f)  identical to one stream: %s" % (N, K, sum(ms) / len(ms), min(ms), bool(torch.equal(got, ref))), flush=True)
