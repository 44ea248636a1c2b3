"""Pixel-sharded single-frame mode of the dense-depth fitter (BASELINE.json configs[3]).

The image is cut into one band of rows per GPU. Every GPU holds the whole mesh and graph, warps it redundantly, rasterises and
assembles its own band, and the per-node partial sums of J^T J and J^T r are all-reduced once per Gauss-Newton iteration; the ARAP
term, the arrowhead solve and the node update are replicated. The sums live in 103-bit fixed point (two int64 words per entry,
csrc/fixed_accum.cuh) and are exchanged as int64: integer addition is associative, so every rank holds bit-identical sums whatever
order NCCL reduces in, they equal the single-GPU sums bit for bit (bands are multiples of the 8-row assembly tile, so every warp sees
the same 32 pixels as in the unsharded fit), and the replicated solves -- which use no atomics -- keep the node state identical on
all ranks with no broadcast. At 640x480 one band is a few microseconds of work against an all-reduce of tens of microseconds: the
mode is built for 4K frames and reported as measured.
"""
import os

import torch
import torch.distributed as dist

# the sharded Gauss-Newton loop is replayed from a CUDA graph (NCCL collective inside) from the second fit of a key on; NNRT_B200_SHARD_GRAPH=0 keeps it eager
GRAPH_CAPTURE = os.environ.get("NNRT_B200_SHARD_GRAPH", "1") != "0"

# Default split: stripes of this many rows dealt round-robin to the ranks (stripe s -> rank s % world_size). Contiguous bands
# (stripe_rows=None in shard_pixels) leave the outer GPUs idle when the surface covers the middle of the frame: at 3840x2160 on 8 GPUs
# the busiest band's data term took 0.179 ms against 0.150 ms for a rank's stripes (tests/measure/pixel_shard_stage_times.py). 32 rows =
# four assembly tiles: fine enough to balance 8 GPUs (64-row stripes no longer do), coarse enough that a face of a 4K frame rarely meets
# two stripes (8- and 16-row stripes queue more band items per face: 0.160 / 0.159 ms).
STRIPE_ROWS = 32

# How the ranks' sums meet: "peer" = the kernel that consumes them reads every rank's buffer over NVLink peer memory (CUDA IPC) and adds
# the words itself, no library collective on the iteration; "nccl" = int64 all-reduce between the two halves of the iteration.
COLLECTIVE = os.environ.get("NNRT_B200_SHARD_COLLECTIVE", "peer")


def stripe_rows_of(rank, world_size, height, stripe_rows=STRIPE_ROWS):
    """The image rows of `rank`'s stripes as a list of [begin, end) ranges (host-side statement of csrc/fitter_kernels.cuh::row_owned)."""
    if not (0 <= rank < world_size):
        raise ValueError(f"rank {rank} outside world of {world_size}")
    if stripe_rows <= 0 or stripe_rows % 8:
        raise ValueError(f"stripe_rows must be a positive multiple of 8, got {stripe_rows}")
    return [(b, min(height, b + stripe_rows)) for b in range(rank * stripe_rows, height, world_size * stripe_rows)]


def pixel_rows(rank, world_size, height, align=8):
    """Rows [begin, end) of `rank`'s band: `height` split as evenly as the assembly tile height (8 rows) allows; the last band takes
    the remainder. Every row belongs to exactly one rank."""
    if not (0 <= rank < world_size):
        raise ValueError(f"rank {rank} outside world of {world_size}")
    if world_size > max(1, height // align):
        raise ValueError(f"{world_size} bands of at least {align} rows do not fit an image of height {height}")
    per = (height // world_size) // align * align
    begin = rank * per
    end = height if rank == world_size - 1 else begin + per
    return begin, end


class DeviceBuffer:
    """Zero-copy torch view of an int64 device buffer owned by the C library (CUDA array interface)."""

    def __init__(self, pointer, count):
        self.__cuda_array_interface__ = {"shape": (int(count),), "typestr": "<i8", "data": (int(pointer), False), "version": 2}

    def tensor(self, device):
        return torch.as_tensor(self, device=device)


def all_reduce_sum(tensor, group=None):
    """Sums the accumulator buffer over the ranks of `group` in place (NCCL on GPUs; a single rank is a no-op)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(tensor, op=dist.ReduceOp.SUM, group=group)
    return tensor
