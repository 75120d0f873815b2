"""Host-side sharding of independent sequences over ranks (replicas only, no
data-path collective).  One process per GPU; a rank owns the sequences `shard_sequences` gives it and a
handle on its own device.  torch.distributed is used for exactly two things: the barrier that brackets a
timed region and the max-over-ranks of the device time (`max_over_ranks`) — NCCL on the GPU box, gloo in
the CPU tests."""


def shard_sequences(n_sequences, world, rank):
    """Contiguous, balanced partition of sequence ids 0..n-1: the first n % world ranks take one extra."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("rank outside 0..world-1")
    base, extra = divmod(n_sequences, world)
    lo = rank * base + min(rank, extra)
    return list(range(lo, lo + base + (1 if rank < extra else 0)))


def sequence_seed(base_seed, sequence_id):
    """Seed of an independent synthetic sequence (bench.py: rank r owns sequence r at N ranks)."""
    return base_seed + sequence_id


def max_over_ranks(values, world, device="cpu"):
    """Element-wise MAX of a list of floats over all ranks (identity when world == 1)."""
    if world == 1:
        return [float(v) for v in values]
    import torch
    import torch.distributed as dist
    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(v) for v in t.cpu()]


def aggregate_throughput(units_per_rank, ms_max, world):
    """Whole-job units/s under weak scaling: every rank processed `units_per_rank` in at most ms_max."""
    return units_per_rank * world / (ms_max * 1e-3)
