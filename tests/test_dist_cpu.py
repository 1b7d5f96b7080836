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


# ---- limb-sharded key-switch: host-side layout logic over a real world_size-2 gloo all-gather ----
def _shard_worker(rank, world, port, q):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    sys.path.insert(0, ROOT)
    import numpy as np
    import torch
    import torch.distributed as dist
    from lattisense_b200.dist import Ranks, ShardedKeySwitch, shard_layout, shard_poly
    r = Ranks(backend="gloo")
    level, k, n = 4, 3, 8
    lay = shard_layout(level, k, world, rank)
    poly = np.arange((level + 1) * n, dtype=np.int64).reshape(level + 1, n)        # limb i holds i*n .. i*n+n-1
    own = shard_poly(poly, lay["own_q"])
    local = torch.zeros((lay["cq"], n), dtype=torch.int64)
    local[:own.shape[0]] = torch.from_numpy(own)
    out = torch.zeros((world, lay["cq"], n), dtype=torch.int64)
    ShardedKeySwitch._nccl_gather(local, out)   # the very call the GPU path makes (backend-agnostic form)
    flat = out.reshape(world * lay["cq"], n).numpy()
    ok = all(np.array_equal(flat[(i % world) * lay["cq"] + i // world], poly[i]) for i in range(level + 1))
    q.put((rank, lay, ok))
    r.close()


def test_shard_layout_and_gather_slots_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_shard_worker, args=(i, 2, port, q)) for i in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res[0][1] == {"own_q": [0, 2, 4], "own_p": [0, 2], "cq": 3, "cp": 2}
    assert res[1][1] == {"own_q": [1, 3], "own_p": [1], "cq": 3, "cp": 2}
    assert res[0][2] and res[1][2]       # limb i sits at gathered slot (i % G) * cq + i // G, as lsb_shard.cu assumes


def test_shard_helpers_roundtrip():
    import numpy as np
    sys.path.insert(0, ROOT)
    from lattisense_b200.dist import shard_evk, shard_layout, shard_poly, unshard_ct
    level, k, world, n = 6, 3, 4, 4
    ct = np.arange(2 * (level + 1) * n, dtype=np.uint64).reshape(2, level + 1, n)
    parts = [shard_poly(ct, shard_layout(level, k, world, r)["own_q"]) for r in range(world)]
    assert np.array_equal(unshard_ct(parts, level, world), ct)
    evk = np.arange(3 * 2 * (level + 1 + k) * n, dtype=np.uint64).reshape(3, 2, level + 1 + k, n)
    lay = shard_layout(level, k, world, 1)
    e = shard_evk(evk, level, k, lay["own_q"], lay["own_p"])
    assert e.shape == (3, 2, len(lay["own_q"]) + len(lay["own_p"]), n)
    assert np.array_equal(e[:, :, 0], evk[:, :, 1]) and np.array_equal(e[:, :, -1], evk[:, :, level + 1 + 1])
