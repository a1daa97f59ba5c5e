"""Multi-GPU frame partition: the frame's 8-row bands are dealt round-robin to the ranks, every rank
renders its bands into their true positions of a full-size frame (vx_render_opts.band_count / band_index), and the
bands are gathered to rank 0.  Rays never leave their GPU: the world is replicated, the gather is the only exchange.

Host-side only (torch tensors + torch.distributed); the same code runs over NCCL on CUDA tensors and over gloo on CPU
tensors, which is how tests/test_partition.py covers it without a GPU.
"""
import torch
import torch.distributed as dist

BAND_ROWS = 8  # rows per band: one 16x8 macro tile row of the primary-ray kernel


class FramePartition:
    def __init__(self, W, H, world_size, rank, bytes_per_pixel=4):
        if not (0 <= rank < world_size):
            raise ValueError("rank %d outside world of %d" % (rank, world_size))
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

    # ---- gather ------------------------------------------------------------------------------------
    def _views(self, frame):
        return frame.view(self.bands, self.band_bytes)

    def gather(self, frame, group=None):
        """Collect every rank's bands into rank 0's `frame` (a flat uint8 tensor of frame_bytes()).  Collective."""
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
