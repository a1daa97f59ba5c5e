"""Multi-GPU frame partition: the frame's 8-row bands are dealt round-robin to the ranks, every rank
renders its bands into their true positions of a full-size frame (vx_render_opts.band_count / band_index), and the
bands are gathered to rank 0.  Rays never leave their GPU: the world is replicated, the gather is the only exchange.

Host-side only (torch tensors + torch.distributed); the same code runs over NCCL on CUDA tensors and over gloo on CPU
tensors, which is how tests/test_partition.py covers it without a GPU.

Two exchanges are offered.  SharedFrame (the default of bench.py on GPUs): rank 0 owns the frame, the other ranks map
it over CUDA IPC and their frame kernels store pixels straight into rank 0's HBM over NVLink; the only collective left
is a completion fence.  FramePartition.gather: every rank renders into its own frame and the bands travel in one
torch.distributed gather (works on any backend).
"""
import torch
import torch.distributed as dist

BAND_ROWS = 8  # rows per band: one 16x8 macro tile row of the primary-ray kernel


class FramePartition:
    """`checkerboard`: the frames are rendered with vx_render_opts.checkerboard = 1 (the reference's default,
    Renderer.cu:186-194).  A band is 8 rows of THREADS; with the checkerboard mapping y = 2 * ty + {0, 1, 2} it covers
    pixel rows 16 b .. 16 b + 17, overlapping the next band's, so a rank's pixels are not whole pixel rows and cannot be
    moved by row copies: gather(), render_and_gather() and HostFrame refuse such a partition.  SharedFrame (every rank's
    kernels store their own pixels into the one frame) is exact for checkerboard frames and is the path to use."""

    def __init__(self, W, H, world_size, rank, bytes_per_pixel=4, checkerboard=False):
        if not (0 <= rank < world_size):
            raise ValueError("rank %d outside world of %d" % (rank, world_size))
        self.checkerboard = bool(checkerboard)
        self.W, self.H, self.world_size, self.rank = W, H, world_size, rank
        self.bands = (H + BAND_ROWS - 1) // BAND_ROWS
        self.band_bytes = BAND_ROWS * W * bytes_per_pixel
        self.per_rank = (self.bands + world_size - 1) // world_size  # bands per rank, padded to the largest share
        self._send = None
        self._recv = None

    # ---- which bands / rows belong to whom ------------------------------------------------------
    def bands_of(self, rank):
        return range(rank, self.bands, self.world_size)

    def rows_of(self, rank):
        rows = []
        for b in self.bands_of(rank):
            rows.extend(range(b * BAND_ROWS, min((b + 1) * BAND_ROWS, self.H)))
        return rows

    def frame_bytes(self):
        """Bytes of a frame buffer padded to whole bands (what alloc_frame allocates)."""
        return self.bands * self.band_bytes

    def alloc_frame(self, device):
        return torch.zeros(self.frame_bytes(), dtype=torch.uint8, device=device)

    def render_opts(self):
        """(band_count, band_index) for vx_render_opts; (0, 0) renders the whole frame."""
        return (self.world_size, self.rank) if self.world_size > 1 else (0, 0)

    def row_chunks(self, n_chunks):
        """Split the frame's rows into up to n_chunks ranges (y0, rows), each a whole number of band rounds
        (BAND_ROWS * world_size rows) so that inside every chunk the bands are still dealt rank, rank + N, ..."""
        rnd = BAND_ROWS * self.world_size
        rounds = (self.H + rnd - 1) // rnd
        per = max(1, (rounds + n_chunks - 1) // max(n_chunks, 1))
        out, y = [], 0
        while y < self.H:
            h = min(per * rnd, self.H - y)
            out.append((y, h))
            y += per * rnd
        return out

    def render_and_gather(self, frame, render_rows, n_chunks=4, group=None):
        """Pipelined frame: for every row chunk call render_rows(y0, rows) (which must enqueue this rank's render of
        those rows on the current stream / do it synchronously on CPU), then start the gather of that chunk's bands
        asynchronously, so the transfer of one chunk overlaps the render of the next.  Collective."""
        self._require_row_bands("render_and_gather")
        if self.world_size == 1:
            render_rows(0, self.H)
            return frame
        fb = self._views(frame)
        chunks = self.row_chunks(n_chunks)
        if self._send is None or self._send.device != frame.device:
            self._send = torch.zeros(self.per_rank, self.band_bytes, dtype=torch.uint8, device=frame.device)
            self._recv = ([torch.zeros_like(self._send) for _ in range(self.world_size)] if self.rank == 0 else None)
        pending = []
        for (y0, h) in chunks:
            render_rows(y0, h)
            b0, b1 = y0 // BAND_ROWS, (y0 + h + BAND_ROWS - 1) // BAND_ROWS          # global bands of the chunk
            s0, s1 = b0 // self.world_size, (b1 + self.world_size - 1) // self.world_size  # their slots in the compact buffers
            mine = fb[b0:b1][self.rank::self.world_size]                             # b0 is a multiple of world_size
            n = mine.shape[0]
            send = self._send[s0:s0 + max(s1 - s0, 1)]
            send[:n].copy_(mine)
            recv = [r[s0:s0 + send.shape[0]] for r in self._recv] if self.rank == 0 else None
            work = dist.gather(send, recv, dst=0, group=group, async_op=True)
            pending.append((work, b0, b1, recv))
        for work, b0, b1, recv in pending:
            work.wait()
            if self.rank == 0:
                for r in range(1, self.world_size):
                    dst = fb[b0:b1][r::self.world_size]
                    dst.copy_(recv[r][: dst.shape[0]])
        return frame

    # ---- gather ------------------------------------------------------------------------------------
    def _views(self, frame):
        return frame.view(self.bands, self.band_bytes)

    def _require_row_bands(self, what):
        if self.checkerboard and self.world_size > 1:
            raise ValueError("%s moves whole pixel rows, but the bands of a checkerboard frame interleave "
                             "(thread row ty writes pixel rows 2 ty .. 2 ty + 2): use SharedFrame, or render without checkerboard" % what)

    def gather(self, frame, group=None):
        """Collect every rank's bands into rank 0's `frame` (a flat uint8 tensor of frame_bytes()).  Collective."""
        self._require_row_bands("gather")
        if self.world_size == 1:
            return frame
        fb = self._views(frame)
        if self._send is None or self._send.device != frame.device:
            self._send = torch.zeros(self.per_rank, self.band_bytes, dtype=torch.uint8, device=frame.device)
            self._recv = ([torch.zeros_like(self._send) for _ in range(self.world_size)] if self.rank == 0 else None)
        mine = fb[self.rank::self.world_size]
        self._send[: mine.shape[0]].copy_(mine)
        dist.gather(self._send, self._recv, dst=0, group=group)
        if self.rank == 0:
            for r in range(1, self.world_size):
                dst = fb[r::self.world_size]
                dst.copy_(self._recv[r][: dst.shape[0]])
        return frame


def _all_ok(err, world_size, group, what):
    """Collective: raise on EVERY rank if `err` is set on any of them, so that the callers can fall back together."""
    errs = [err]
    if world_size > 1:
        errs = [None] * world_size
        dist.all_gather_object(errs, None if err is None else repr(err), group=group)
    bad = [(r, e) for r, e in enumerate(errs) if e is not None]
    if bad:
        raise RuntimeError("%s failed on rank %d: %s" % (what, bad[0][0], bad[0][1]))


class SharedFrame:
    """The frame of a multi-GPU partition, living in rank 0's HBM and mapped by every other rank (vx_shared_alloc /
    vx_shared_open, include/vxrt.h).  `ptr` is what each rank passes to render_device as d_bgra, with
    FramePartition.render_opts() selecting its bands; after `fence()` rank 0's buffer holds the whole frame.

    Collective constructor.  The handle travels through broadcast_object_list, so any backend does; every process must
    have its CUDA device selected (`device` is the CUDA ordinal it renders on)."""

    def __init__(self, part, device, group=None):
        import voxelengine_b200 as vx
        self._vx, self.part, self.device, self.group = vx, part, device, group
        self.nbytes = part.frame_bytes()
        self.owner = part.rank == 0
        box = [None, None]
        my_gpu = str(getattr(torch.cuda.get_device_properties(device), "uuid", device))
        if self.owner:
            self.ptr, handle = vx.shared_alloc(self.nbytes, device)
            box = [handle, my_gpu]
        if part.world_size > 1:
            dist.broadcast_object_list(box, src=0, group=group)
        err = None
        if not self.owner:
            try:
                self.ptr = vx.shared_open(box[0], device)
            except Exception as e:  # e.g. no peer access between the two GPUs
                self.ptr, err = None, e
        try:
            _all_ok(err, part.world_size, group, "mapping the shared frame")
        except RuntimeError:
            if self.owner:
                vx.shared_free(self.ptr, device)
            elif self.ptr is not None:
                vx.shared_close(self.ptr, device)
            self.ptr = None
            raise
        self.remote = box[1] != my_gpu   # the frame is in another GPU's memory: this rank's pixel stores cross NVLink
        self._token = None

    def attach(self, tracer):
        """Tell a VoxelRaytracer3D of this rank where the frame lives: a rank writing into another GPU's memory shades
        inside the sun-ray march (vx_tracer_set_fused_shade), which spreads its NVLink stores over the frame time
        instead of all writers bursting into rank 0 together.  Measured on 8K frames (B200, NVSwitch): with 2 ranks the
        dense pass is still faster (4.90 vs 5.00 ms), with 4 the fused one (2.63 vs 2.71 ms), with 8 clearly (1.46 vs
        1.61 ms) — the burst costs about the same 0.2 ms however few pixels each writer has."""
        tracer.set_fused_shade(1 if self.remote and self.part.world_size >= 4 else 0)

    def fence(self):
        """All ranks' pixel stores issued so far are in rank 0's buffer once this returns (NCCL: once the current
        stream passes this point — the all-reduce is stream-ordered after each rank's kernels, and a kernel's peer
        stores are performed before it completes)."""
        if self.part.world_size == 1:
            return
        if dist.get_backend(self.group) == "nccl":
            if self._token is None:
                self._token = torch.zeros(1, dtype=torch.int32, device=torch.device("cuda", self.device))
            dist.all_reduce(self._token, group=self.group)
        else:
            torch.cuda.synchronize(self.device)
            dist.barrier(group=self.group)

    def to_host(self, host, nbytes=None, stream=0):
        """Rank 0: copy the first nbytes of the frame into a (pinned) host uint8 tensor and wait for it."""
        self._vx.device_to_host(host.data_ptr(), self.ptr, self.nbytes if nbytes is None else nbytes, self.device, stream)

    def close(self):
        if self.ptr is None:
            return
        if self.part.world_size > 1:      # nobody unmaps or frees while another rank may still be writing
            torch.cuda.synchronize(self.device)
            dist.barrier(group=self.group)
        if not self.owner:
            self._vx.shared_close(self.ptr, self.device)
        if self.part.world_size > 1:
            dist.barrier(group=self.group)
        if self.owner:
            self._vx.shared_free(self.ptr, self.device)
        self.ptr = None


class HostFrame:
    """The frame in host memory for a consumer on the CPU side (the reference copies its frame down every frame,
    main.cu:165-167), shared by all ranks' processes: a file in /dev/shm that every rank maps and page-locks
    (vx_host_register).  Each rank renders its bands into its own device frame and `store(d_frame_ptr)` copies exactly
    those bands into place (vx_bands_to_host) — N PCIe links working at once and nothing crossing between the GPUs.
    After `fence()` the host frame is whole; `array` is its numpy view (H x W x 4 bytes, B,G,R,A).

    Collective constructor; any backend (the name travels through broadcast_object_list).  Single node only."""

    def __init__(self, part, device, group=None):
        import mmap
        import os
        import ctypes
        import numpy as np
        import voxelengine_b200 as vx
        part._require_row_bands("HostFrame")
        self._vx, self.part, self.device, self.group = vx, part, device, group
        self.nbytes = part.frame_bytes()
        box = [None]
        if part.rank == 0:
            box[0] = "/dev/shm/vxrt_frame_%d_%s" % (os.getpid(), os.urandom(4).hex())
            fd = os.open(box[0], os.O_CREAT | os.O_EXCL | os.O_RDWR, 0o600)
            os.ftruncate(fd, self.nbytes)
        if part.world_size > 1:
            dist.broadcast_object_list(box, src=0, group=group)
        self.path = box[0]
        if part.rank != 0:
            fd = os.open(self.path, os.O_RDWR)
        self._map = mmap.mmap(fd, self.nbytes)
        os.close(fd)
        self._view = (ctypes.c_uint8 * self.nbytes).from_buffer(self._map)
        self.ptr = ctypes.addressof(self._view)
        err = None
        try:
            vx.host_register(self.ptr, self.nbytes)
        except Exception as e:
            err = e
        try:
            _all_ok(err, part.world_size, group, "page-locking the shared host frame")
        except RuntimeError:
            if err is None:
                vx.host_unregister(self.ptr)
            if part.rank == 0:
                os.unlink(self.path)
            self._view = None
            self._map.close()
            self._map = None
            raise
        self.array = np.frombuffer(self._map, np.uint8, part.H * part.W * 4).reshape(part.H, part.W, 4)
        self._copy_stream = None
        if part.world_size > 1:
            dist.barrier(group=group)   # everyone has it mapped: the name can go
        if part.rank == 0:
            os.unlink(self.path)

    def store(self, d_frame_ptr, stream=0, y0=0, rows=None):
        """Enqueue the copy of this rank's bands from its device frame into the host frame (on `stream`); with y0 / rows
        only the bands inside that row range (y0 a multiple of 8 * world_size, so the dealing of bands is unchanged)."""
        bc, bi = self.part.render_opts()
        rows = self.part.H - y0 if rows is None else rows
        if y0 % (BAND_ROWS * max(bc, 1)):
            raise ValueError("y0 must be a multiple of %d" % (BAND_ROWS * max(bc, 1)))
        off = y0 * self.part.W * 4
        self._vx.bands_to_host(self.ptr + off, d_frame_ptr + off, self.part.W, rows, bc, bi, self.device, stream)

    def render_and_store(self, render_rows, d_frame_ptr, split=0.8):
        """One frame to the host with the copy overlapped: render_rows(y0, rows) renders this rank's bands of the first
        `split` of the rows (on the current stream), their copy runs on a second stream while the rest renders, and
        only the last part's copy is left after the last kernel (same idea and measured split as vx_render_host)."""
        n = max(self.part.world_size, 1) * BAND_ROWS
        r0 = min(self.part.H, max(n, int(self.part.H * split) // n * n))
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(self.device)
            self._events = [torch.cuda.Event(), torch.cuda.Event()]
        for k, (y0, rows) in enumerate(((0, r0), (r0, self.part.H - r0))):
            if rows <= 0:
                continue
            render_rows(y0, rows)
            self._events[k].record()
            self._copy_stream.wait_event(self._events[k])
            self.store(d_frame_ptr, self._copy_stream.cuda_stream, y0, rows)

    def fence(self):
        """Every rank's copies enqueued so far have landed when this returns on every rank."""
        torch.cuda.synchronize(self.device)
        if self.part.world_size > 1:
            dist.barrier(group=self.group)

    def mark(self, stream=None):
        """An event after the copies enqueued so far (on the copy stream of render_and_store, or on `stream` for store()):
        wait(mark()) is fence() for exactly this frame, so a consumer that keeps two host frames can let the ranks start frame
        k + 1 before it waits for frame k — the inter-process barrier then overlaps rendering instead of adding to it."""
        ev = torch.cuda.Event()
        ev.record(self._copy_stream if (stream is None and self._copy_stream is not None) else stream)
        return ev

    def wait(self, event):
        event.synchronize()
        if self.part.world_size > 1:
            dist.barrier(group=self.group)

    def close(self):
        if self._map is None:
            return
        torch.cuda.synchronize(self.device)
        self._vx.host_unregister(self.ptr)
        self.array = None
        self._view = None
        try:
            self._map.close()
        except BufferError:   # a caller still holds a view of `array`; the mapping goes with the last reference
            pass
        self._map = None


def build_world_partitioned(X, Y, Z, factor=8, device=0, origin=(0, 0, 0), group=None, rank=None, world_size=None):
    """The noise world built by all ranks together (BASELINE configs[4]): rank r runs the voxel kernel on its z-slab
    (vx_slab_build_noise), the slabs' top-level bits, packed bounds and compact mask pools are all-gathered (plain bulk
    collectives: NCCL over NVLink on GPUs), and every rank assembles the same world from the parts (vx_world_from_parts) —
    word for word the world vx_world_build_noise builds on one device.  Collective.  Returns (World, timings dict in ms:
    slab = this rank's voxel kernel + brick analysis, exchange = the collectives, assemble = vx_world_from_parts)."""
    import time
    import voxelengine_b200 as vx
    rank = dist.get_rank(group) if rank is None else rank
    world_size = dist.get_world_size(group) if world_size is None else world_size
    dev = torch.device("cuda", device)
    slab = vx.Slab(X, Y, Z, factor, rank, world_size, origin=origin, device=device)
    info = slab.info
    wpb = int(info.words_per_brick)
    nb = (X // factor) * (Y // factor) * (Z // factor)
    mine = torch.tensor([int(info.brick_begin), int(info.brick_count), int(info.occupied_bricks), int(info.solid_voxels)], dtype=torch.int64, device=dev)
    parts = [torch.zeros_like(mine) for _ in range(world_size)]
    torch.cuda.synchronize(device)
    t0 = time.perf_counter()
    if world_size > 1:
        dist.all_gather(parts, mine, group=group)
    else:
        parts = [mine]
    parts = [[int(v) for v in p.cpu()] for p in parts]
    occupied = sum(p[2] for p in parts)
    solid = sum(p[3] for p in parts)
    top = torch.empty(nb // 32, dtype=torch.int32, device=dev)
    bounds = torch.empty(nb, dtype=torch.int32, device=dev)
    pool = torch.empty(max(occupied, 1) * wpb, dtype=torch.int32, device=dev)
    slot0 = [sum(p[2] for p in parts[:r]) for r in range(world_size)]
    views = [(top[p[0] // 32:(p[0] + p[1]) // 32], bounds[p[0]:p[0] + p[1]], pool[slot0[r] * wpb:(slot0[r] + p[2]) * wpb])
             for r, p in enumerate(parts)]
    slab.export(views[rank][0].data_ptr(), views[rank][1].data_ptr(), views[rank][2].data_ptr() if parts[rank][2] else 0,
                stream=torch.cuda.current_stream(device).cuda_stream)
    if world_size > 1:
        # every part goes from its builder to everybody: one broadcast per part and builder, straight into place
        for r in range(world_size):
            for v in views[r]:
                if v.numel():
                    dist.broadcast(v, src=dist.get_global_rank(group, r) if group is not None else r, group=group)
    torch.cuda.synchronize(device)
    t1 = time.perf_counter()
    slab.close()
    world = vx.World.from_parts(X, Y, Z, factor, top.data_ptr(), bounds.data_ptr(), pool.data_ptr(), occupied, solid, device=device)
    torch.cuda.synchronize(device)
    t2 = time.perf_counter()
    return world, {"slab": float(info.build_ms), "exchange": (t1 - t0) * 1e3, "assemble": (t2 - t1) * 1e3,
                   "slab_bricks": int(info.brick_count), "slab_occupied": int(info.occupied_bricks)}
