"""Multi-GPU plumbing: one process per GPU, the environment batch sharded in contiguous blocks, NO data-path
collective (environments never exchange state: nothing crosses envs in collision_avoidance_env.py).  The only
collective of the design is the all-reduce of the per-step episode statistics ([8] floats), issued on a side
stream so that it overlaps the next step's kernels.  torch.distributed is the plumbing (NCCL over NVLink on
GPUs, gloo in the CPU tests)."""
import torch
import torch.distributed as dist


def shard_bounds(num_envs, world_size, rank):
    """Contiguous block [lo, hi) of environments owned by ``rank``; blocks differ by at most one env."""
    lo = (num_envs * rank) // world_size
    hi = (num_envs * (rank + 1)) // world_size
    return lo, hi


def shard_scenario(arrays, world_size, rank):
    """Slice every per-environment array (leading dim E) of a scenario tuple to this rank's block; entries that are
    None or not per-environment (e.g. shared obstacle rectangles [n,4]) pass through."""
    E = len(arrays[0])
    lo, hi = shard_bounds(E, world_size, rank)
    out = []
    for a in arrays:
        out.append(a[lo:hi] if a is not None and hasattr(a, "__len__") and len(a) == E else a)
    return tuple(out)


class StatsReducer:
    """all_reduce(SUM) of the [8] episode-statistics vector across ranks, off the critical path."""

    def __init__(self, device=None):
        self.world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
        self.device = device
        self.stream = torch.cuda.Stream(device=device) if (device is not None and torch.device(device).type == "cuda") else None
        self.total = None

    def reduce(self, stats):
        """stats: [8] float tensor written by casim_step on the current stream.  Returns the tensor that will hold
        the global sums once ``wait()`` returns."""
        if self.stream is not None:
            # the snapshot is taken on the CURRENT stream, i.e. ordered before the next casim_step overwrites `stats`;
            # only the collective runs on the side stream (it overlaps the next step's kernels)
            self.total = stats.clone()
            self.stream.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(self.stream):
                self.total.record_stream(self.stream)
                if self.world > 1:
                    dist.all_reduce(self.total, op=dist.ReduceOp.SUM)
        else:
            self.total = stats.clone()
            if self.world > 1:
                dist.all_reduce(self.total, op=dist.ReduceOp.SUM)
        return self.total

    def reduce_row(self, row, ready_event=None):
        """In-place all_reduce(SUM) of one statistics row on the side stream, which first waits for ``ready_event`` (the
        event casim_set_stats_event records after the step kernel) or, without one, for the current stream.  The row
        must not be rewritten before ``wait()`` (bench.py gives every step of a CUDA graph its own row)."""
        if self.stream is None:
            if self.world > 1:
                dist.all_reduce(row, op=dist.ReduceOp.SUM)
            return row
        if ready_event is not None:
            self.stream.wait_event(ready_event)
        else:
            self.stream.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(self.stream):
            if self.world > 1:
                dist.all_reduce(row, op=dist.ReduceOp.SUM)
        self.total = row
        return row

    def wait(self):
        if self.stream is not None:
            torch.cuda.current_stream(self.device).wait_stream(self.stream)
        return self.total
