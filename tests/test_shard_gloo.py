"""N>1 host logic on CPU: two gloo ranks shard independent sequences, time a fake region and reduce it the
way bench.py does (barrier, max over ranks, whole-job aggregate).  No GPU, no data-path collective."""
import os
import socket
import sys

import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from visionx_slam_b200.shard import aggregate_throughput, max_over_ranks, sequence_seed, shard_sequences  # noqa: E402


def test_shard_partition_properties():
    for n in (0, 1, 7, 8, 16, 129):
        for world in (1, 2, 3, 4, 8):
            parts = [shard_sequences(n, world, r) for r in range(world)]
            assert sorted(sum(parts, [])) == list(range(n))  # disjoint cover
            assert max(map(len, parts)) - min(map(len, parts)) <= 1  # balanced
    assert sequence_seed(2026, 3) == 2029


def _worker(rank, world, port, q):
    import numpy as np
    import torch.distributed as dist
    from visionx_slam_b200.synth import synth_sequence
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    mine = shard_sequences(5, world, rank)
    # every rank builds its own sequences; different ranks get different frames
    digest = [int(np.asarray(synth_sequence(64, 48, 2, seed=sequence_seed(2026, s))).sum()) for s in mine]
    dist.barrier()
    ms = [10.0 + 5.0 * rank, 20.0 - 3.0 * rank]  # pretend device times (resident, e2e)
    mx = max_over_ranks(ms, world)
    dist.barrier()
    q.put((rank, mine, digest, mx))
    dist.destroy_process_group()


def test_two_gloo_ranks_shard_and_reduce():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, s0, d0, m0), (r1, s1, d1, m1) = got
    assert s0 == [0, 1, 2] and s1 == [3, 4]
    assert len(set(d0 + d1)) == 5  # five distinct sequences
    assert m0 == m1 == [15.0, 20.0]  # max over ranks, identical everywhere
    assert aggregate_throughput(128 * 20, m0[0], 2) == 128 * 20 * 2 / 15e-3
