"""The N>1 path of bench.py (replicas only): rank -> sequence assignment, barrier and
max-over-ranks / sum reduction, exercised with a world-size-2 gloo group on CPU."""
import os
import socket

import torch.multiprocessing as mp

from lidarslam_b200 import shard


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    assert shard.rank_info() == (rank, world, rank)
    local_ms = [100.0 + 50.0 * rank, 7.0 - rank]           # rank 1 is slower on the first timer, faster on the second
    dist.barrier()
    ms, cnt = shard.reduce_job(local_ms, [10 + rank])
    q.put((rank, ms, cnt, shard.sequence_seed(rank)))
    dist.destroy_process_group()


def test_world_size_2_reduce():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    out = sorted(q.get(timeout=120) for _ in ps)
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, ms, cnt, seed in out:
        assert ms == [150.0, 7.0] and cnt == [21.0]          # MAX over ranks, SUM of counters: identical on every rank
        assert seed == 1234 + rank
    assert shard.job_throughput(20, 2, 0.150) == 2 * 20 / 0.150


def test_sequence_assignment():
    a = shard.assign_sequences(8, 4)
    assert a == [[0, 4], [1, 5], [2, 6], [3, 7]]
    assert sorted(s for r in shard.assign_sequences(8, 3) for s in r) == list(range(8))
    assert shard.assign_sequences(8, 8) == [[i] for i in range(8)]
    assert shard.reduce_job([3.0], [2.0]) == ([3.0], [2.0])   # no process group: identity
