"""Frame-batch (sequence-parallel) sharding of the dense-depth fitter across GPUs (BASELINE.json configs[4]).

Independent RGB-D sequences are the unit: frames inside a sequence are serial (each fit warm-starts from the previous one), sequences
never exchange data, so there is NO data-path collective. One process per GPU; torch.distributed is used only to agree on timing
(max over ranks) and to gather per-rank results. Works with the nccl backend on GPUs and with gloo on CPU (tests).
"""
import torch
import torch.distributed as dist


def shard_sequences(sequence_count, rank, world_size):
    """Round-robin assignment: sequence s runs on rank s % world_size (64 sequences over 1/2/4/8 GPUs divide evenly)."""
    if not (0 <= rank < world_size):
        raise ValueError(f"rank {rank} outside world of {world_size}")
    return list(range(rank, sequence_count, world_size))


def max_over_ranks(values, device="cpu"):
    """Element-wise max of a list of floats over all ranks (the timed region's duration is the slowest rank's)."""
    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(v) for v in t]


def sum_over_ranks(values, device="cpu"):
    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return [float(v) for v in t]


def frames_per_second(frames_this_rank, elapsed_ms_this_rank, device="cpu"):
    """Whole-job throughput: all ranks' frames over the slowest rank's time."""
    total = sum_over_ranks([frames_this_rank], device)[0]
    slowest = max_over_ranks([elapsed_ms_this_rank], device)[0]
    return total / (slowest * 1e-3), total, slowest


def run_sequences(sequence_ids, frames_per_sequence, fit_frame):
    """Runs `fit_frame(sequence_id, frame_index)` for every frame of every local sequence, sequence by sequence (frames of one sequence
    are serial). Returns the number of frames processed."""
    count = 0
    for seq in sequence_ids:
        for frame in range(frames_per_sequence):
            fit_frame(seq, frame)
            count += 1
    return count


class SequenceStreams:
    """One CUDA stream per sequence resident on this GPU. A fit is a chain of dependent launches that ends in a single-CTA corner
    solve, so one sequence leaves most SMs idle; independent sequences (configs[4]) fill them when each has its own fitter, warp field
    and stream. Nothing is shared between sequences, so no cross-stream ordering is needed beyond fork() / join() with the caller's
    stream."""

    def __init__(self, count, device=None):
        if count < 1:
            raise ValueError("at least one sequence")
        if not torch.cuda.is_available():
            raise RuntimeError("SequenceStreams needs a CUDA device; there is no CPU fallback")
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.streams = [torch.cuda.Stream(self.device) for _ in range(count)]

    def __len__(self):
        return len(self.streams)

    def fork(self):
        """Every sequence stream waits for the work already enqueued on the caller's current stream."""
        event = torch.cuda.current_stream(self.device).record_event()
        for stream in self.streams:
            stream.wait_event(event)

    def join(self):
        """The caller's current stream waits for every sequence stream (no host synchronisation)."""
        current = torch.cuda.current_stream(self.device)
        for stream in self.streams:
            current.wait_event(stream.record_event())

    def run(self, sequence, fn, *args, **kwargs):
        """Calls fn(*args, **kwargs) with sequence's stream current, so everything it enqueues lands on that stream."""
        with torch.cuda.stream(self.streams[sequence]):
            return fn(*args, **kwargs)

    def synchronize(self, sequence=None):
        for stream in (self.streams if sequence is None else [self.streams[sequence]]):
            stream.synchronize()
