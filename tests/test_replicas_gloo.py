"""N > 1 host logic (replicas only, no data-path collective) over world-size-2
gloo on CPU: stream partition, least-loaded routing, sum/max aggregation of a timed region."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from nano_qwen3tts_vllm_b200 import replicas


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = replicas.partition_streams(list(range(11)), world, rank)
    ms = 5.0 + rank                      # rank 1 is the slow one
    units, t = replicas.aggregate(len(mine) * 10, ms, dist)
    dist.barrier()
    q.put((rank, mine, units, t))
    dist.destroy_process_group()


def test_partition_and_router():
    ids = [f"s{i}" for i in range(7)]
    parts = [replicas.partition_streams(ids, 3, r) for r in range(3)]
    assert sorted(sum(parts, [])) == sorted(ids)
    assert parts[0] == ["s0", "s3", "s6"]
    assert max(map(len, parts)) - min(map(len, parts)) <= 1
    assert replicas.least_loaded([3, 1, 1]) == 1
    assert replicas.aggregate(5, 2.0) == (5.0, 2.0)


def test_two_replicas_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (_, m0, u0, t0), (_, m1, u1, t1) = res
    assert m0 == [0, 2, 4, 6, 8, 10] and m1 == [1, 3, 5, 7, 9]
    assert u0 == u1 == 110.0             # all ranks' units
    assert t0 == t1 == 6.0               # max over ranks
