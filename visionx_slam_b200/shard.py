"""Host-side sharding of independent sequences over ranks (replicas only, no
data-path collective).  One process per GPU; a rank owns the sequences `shard_sequences` gives it and a
handle on its own device.  torch.distributed is used for exactly two things: the barrier that brackets a
timed region and the max-over-ranks of the device time (`max_over_ranks`) — both on CPU tensors over gloo, on
the GPU box as in the CPU tests: the path has no collective, so no NCCL communicator is ever created."""
import os


def init_host_group(world):
    """gloo process group from the torchrun environment (RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT); a no-op
    for a single rank."""
    if world <= 1:
        return
    import torch.distributed as dist
    if not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("gloo")


def host_barrier(world):
    if world > 1:
        import torch.distributed as dist
        dist.barrier()


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


def max_over_ranks(values, world, device="cpu"):  # device stays "cpu": gloo
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
