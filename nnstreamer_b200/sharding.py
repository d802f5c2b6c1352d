"""Multi-GPU plumbing for the stream-sharded path.

The tensor_transform path shards by stream: pipeline branch (stream) s runs on GPU s mod G for
the life of its element, and no data crosses GPUs -- so there is NO collective on the data
path.  torch.distributed is used only by the benchmark driver: a barrier around the timed
region and a MAX over ranks of the measured time.  Works with the `gloo` backend on CPU (tests)
and `nccl` on GPUs (bench.py under torchrun).
"""


def stream_to_gpu(stream_index, n_gpus):
    """Pipeline branch -> GPU (fixed for the life of the element via its device-id)."""
    if n_gpus < 1:
        raise ValueError("n_gpus must be >= 1")
    return stream_index % n_gpus


def streams_of_rank(n_streams, world_size, rank):
    """The streams rank `rank` (one process per GPU) drives."""
    return [s for s in range(n_streams) if stream_to_gpu(s, world_size) == rank]


def barrier(dist):
    if dist is not None:
        dist.barrier()


def max_over_ranks(value, dist, device="cpu"):
    """MAX all-reduce of a Python float; identity without a process group."""
    if dist is None:
        return float(value)
    import torch
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(value, dist, device="cpu"):
    if dist is None:
        return float(value)
    import torch
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def gather_over_ranks(values, dist, device="cpu"):
    """All-gather of a short list of floats: returns one list per rank, in rank order."""
    vals = [float(v) for v in values]
    if dist is None:
        return [vals]
    import torch
    t = torch.tensor(vals, dtype=torch.float64, device=device)
    outs = [torch.zeros_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(outs, t)
    return [[float(x) for x in o.tolist()] for o in outs]


def aggregate_throughput(units_per_rank, seconds_this_rank, dist, device="cpu"):
    """Whole-job throughput: all units processed by all ranks / slowest rank's time."""
    total_units = sum_over_ranks(units_per_rank, dist, device)
    slowest = max_over_ranks(seconds_this_rank, dist, device)
    return total_units / slowest if slowest > 0 else 0.0
