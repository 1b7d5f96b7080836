"""N>1 host logic on CPU: world_size-2 gloo run of the rank plumbing bench.py uses (barrier, max-over-ranks
time, whole-job rate, unit sharding) plus the runner's component -> device assignment."""
import os
import socket
import sys

import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    sys.path.insert(0, ROOT)
    from lattisense_b200.dist import Ranks
    r = Ranks(backend="gloo")
    r.barrier()
    slow = r.max(0.010 * (rank + 1))               # rank 1 is the slow one
    rate = r.whole_job_rate(1024, 0.010 * (rank + 1))
    lo, hi = r.shard(1027)
    r.barrier()
    q.put((rank, slow, rate, lo, hi))
    r.close()


def test_world_size_2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(i, 2, port, q)) for i in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(abs(r[1] - 0.020) < 1e-12 for r in res)           # max over ranks
    assert all(abs(r[2] - 2048 / 0.020) < 1e-6 for r in res)     # total units / slowest time
    assert (res[0][3], res[0][4], res[1][3], res[1][4]) == (0, 514, 514, 1027)


def test_single_process_defaults():
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        os.environ.pop(k, None)
    sys.path.insert(0, ROOT)
    from lattisense_b200.dist import Ranks
    r = Ranks(backend="gloo")
    assert (r.rank, r.world) == (0, 1)
    assert r.max(3.5) == 3.5 and r.shard(10) == (0, 10)
    r.barrier()
