"""Row-band sharding of one frame over the ranks of one node.

The reference renders a frame on one Vulkan device; its ray-gen launches are per-pixel independent
(src/shaders/lighting_passes/refrence.rgen:18-47: the seed is a function of the pixel and the
frame only), so the frame splits by rows with no exchange step inside the pass.  A rank owns every
`world`-th band of `band_rows` rows (interleaved so that sky/floor imbalance spreads evenly); the scene
and acceleration structure are replicated.  The only collective is the gather of the finished bands.

Pure host logic: nothing here touches CUDA, so the plan and the gather layout are covered by
world_size-2 `gloo` tests on CPU (tests/test_multi_rank_cpu.py) and used as-is over NCCL by bench.py.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Tuple

DEFAULT_BAND_ROWS = 16


@dataclass(frozen=True)
class BandPlan:
    height: int
    band_rows: int
    world: int

    def __post_init__(self):
        if self.height <= 0 or self.band_rows <= 0 or self.world <= 0:
            raise ValueError("BandPlan: height, band_rows and world must be positive")

    @property
    def num_bands(self) -> int:
        return (self.height + self.band_rows - 1) // self.band_rows

    def owner(self, y: int) -> int:
        """Rank that renders row y — the same rule the kernels apply (rt2_launch_params.band_*)."""
        return (y // self.band_rows) % self.world

    def bands(self, rank: int) -> List[Tuple[int, int]]:
        """[(row0, row1)) bands of `rank`, top to bottom; the last band of the frame may be ragged."""
        if not 0 <= rank < self.world:
            raise ValueError(f"rank {rank} outside world {self.world}")
        out = []
        for b in range(rank, self.num_bands, self.world):
            y0 = b * self.band_rows
            out.append((y0, min(y0 + self.band_rows, self.height)))
        return out

    def rows(self, rank: int) -> int:
        return sum(y1 - y0 for (y0, y1) in self.bands(rank))

    @property
    def max_rows(self) -> int:
        """Slot size (in rows) of one rank in the gathered buffer: every rank sends this many rows, padded."""
        return max(self.rows(r) for r in range(self.world))

    def pack_offsets(self, rank: int) -> List[Tuple[int, int, int]]:
        """(row0, row1, first packed row) for each band of `rank` in its contiguous send buffer."""
        out, off = [], 0
        for (y0, y1) in self.bands(rank):
            out.append((y0, y1, off))
            off += y1 - y0
        return out

    def scatter_index(self):
        """For the gathered [world, max_rows] slot layout: list of (frame_row0, frame_row1, rank, packed_row0)."""
        out = []
        for r in range(self.world):
            for (y0, y1, off) in self.pack_offsets(r):
                out.append((y0, y1, r, off))
        out.sort()
        return out


def pack_bands(plan: BandPlan, rank: int, frame, send):
    """Copy `rank`'s rows of `frame` ([H, ...]) into `send` ([max_rows, ...]); padding rows are left as they are."""
    for (y0, y1, off) in plan.pack_offsets(rank):
        send[off:off + (y1 - y0)] = frame[y0:y1]
    return send


def unpack_gathered(plan: BandPlan, gathered, frame):
    """Inverse of the gather: `gathered` is [world * max_rows, ...] (rank-major slots) -> `frame` [H, ...]."""
    m = plan.max_rows
    for (y0, y1, r, off) in plan.scatter_index():
        frame[y0:y1] = gathered[r * m + off:r * m + off + (y1 - y0)]
    return frame


def gather_frame(plan: BandPlan, rank: int, send, recv, dist, async_op: bool = False):
    """All-gather every rank's packed bands (`send`: [max_rows, ...]) into `recv` ([world*max_rows, ...]).
    `dist` is torch.distributed (NCCL on the GPU box, gloo in the CPU tests).  With async_op the collective's work handle is
    returned instead of `recv`: the caller's stream does not wait for the gather (and so not for the slowest rank) until it
    calls handle.wait() — used with double-buffered send/recv to let the gather of frame n overlap the render of frame n+1."""
    if plan.world == 1:
        recv.copy_(send)
        return None if async_op else recv
    if dist.get_backend() == "gloo":                 # gloo has no all_gather_into_tensor
        chunks = list(recv.view(plan.world, -1).unbind(0))
        h = dist.all_gather(chunks, send.reshape(-1), async_op=async_op)
    else:
        h = dist.all_gather_into_tensor(recv.view(-1), send.reshape(-1), async_op=async_op)
    return h if async_op else recv


# ---- in-frame exchange of per-pixel buffers (K6 / K7 under sharding; the host-side mirror of k_exchange_pack / k_exchange_unpack,
# raytracer2_b200/csrc/rt2_group.cu) --------------------------------------------------------------------------------------------------
def exchange_entry_index(width: int, height: int, layout: int, block_row_pitch: int = 0):
    """Entry index of pixel (x, y) for every pixel, as an [H, W] int64 array.  layout 0: row-major; layout 1: the reservoirs'
    16 x 16 block-linear order (RTXDI_ReservoirPositionToPointer, rtxdi/RtxdiHelpers.hlsli:75-88)."""
    import numpy as np
    y, x = np.mgrid[0:height, 0:width].astype(np.int64)
    if layout == 0:
        return y * width + x
    return (y >> 4) * block_row_pitch + (x >> 4) * 256 + (y & 15) * 16 + (x & 15)


def exchange_pack(plan: BandPlan, rank: int, buf, width: int, layout: int, block_row_pitch: int = 0):
    """`buf`: [entries, words] array addressed by exchange_entry_index.  Returns this rank's pixels' entries, [max_rows * W, words]
    (bands top to bottom, padded with zeros)."""
    import numpy as np
    idx = exchange_entry_index(width, plan.height, layout, block_row_pitch)
    out = np.zeros((plan.max_rows * width, buf.shape[1]), buf.dtype)
    for (y0, y1, off) in plan.pack_offsets(rank):
        out[off * width:(off + y1 - y0) * width] = buf[idx[y0:y1].ravel()]
    return out


def exchange_unpack(plan: BandPlan, rank: int, gathered, buf, width: int, layout: int, block_row_pitch: int = 0):
    """`gathered`: [world, max_rows * W, words].  Writes every OTHER rank's entries into `buf` in place; this rank's stay as they are."""
    idx = exchange_entry_index(width, plan.height, layout, block_row_pitch)
    for (y0, y1, r, off) in plan.scatter_index():
        if r != rank:
            buf[idx[y0:y1].ravel()] = gathered[r, off * width:(off + y1 - y0) * width]
    return buf
