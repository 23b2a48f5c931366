"""Host logic of the multi-GPU bench path on CPU: two processes (gloo, 127.0.0.1) run bench.py's reduction exactly as
the ranks of a torchrun launch do: MAX of the step times, SUM of the counters, whole-job value from both."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import bench
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    # rank r: P packets, P - r failures, its own step time
    P, bits = 1000, 32920
    ok = P - rank
    times, counters = bench.reduce_over_ranks([10.0 + 2.0 * rank, 30.0 - rank], [P, ok, ok * bits, ok, P, P], world, "cpu")
    q.put((rank, times, counters, bench.whole_job_value(counters[2], times[0]), bench.rank_seed(1000, rank)))
    dist.barrier()
    dist.destroy_process_group()


def test_reduce_over_two_ranks():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, t0, c0, v0, s0), (r1, t1, c1, v1, s1) = res
    assert t0 == t1 == [12.0, 30.0]                      # MAX over ranks
    assert c0 == c1 == [2000, 1999, 1999 * 32920, 1999, 2000, 2000]   # SUM over ranks
    assert v0 == v1 == pytest.approx(1999 * 32920 / 12.0e-3 / 1e6)
    assert s0 != s1                                       # disjoint synthetic streams


def test_single_rank_is_identity():
    sys.path.insert(0, ROOT)
    import bench
    t, c = bench.reduce_over_ranks([3.5], [7, 8], 1, "cpu")
    assert t == [3.5] and c == [7, 8]
