"""World-size-2 gloo test of the N>1 host logic (shard ranges + hit-list gather), on CPU.
Each rank produces the hit list of ITS id range with the oracle (stand-in for the GPU engine,
which is exercised by the gpu tests); the gather must reproduce the single-rank list exactly."""
import os
import socket
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port_no, n, seed, gens, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from lifeapi_b200 import HIT_DTYPE, gather_hits, shard_range
    from oracle.loader import Port
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port_no)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        port = Port()
        lo, hi = shard_range(n, rank, world)
        ids, pops, ps, hx = port.survey(lo, hi - lo, seed, 0, gens, hi - lo)
        hits = np.zeros(len(ids), dtype=HIT_DTYPE)
        hits["id"], hits["g_first_false"] = ids, pops
        merged, sums, xors = gather_hits(hits, [ps, hx])
        if rank == 0:
            q.put((merged.tobytes(), sums[0], xors[1]))
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_shard_range_partitions():
    from lifeapi_b200 import shard_range
    for n in (0, 1, 7, 1 << 20, (1 << 24) + 3):
        for w in (1, 2, 3, 4, 8):
            r = [shard_range(n, k, w) for k in range(w)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[k][1] == r[k + 1][0] for k in range(w - 1))
            assert max(b - a for a, b in r) - min(b - a for a, b in r) <= 1
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


def test_gather_hits_world2_gloo():
    from conftest import SEED
    from lifeapi_b200 import HIT_DTYPE
    from oracle.loader import Port
    n, gens, world = 96, 200, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port_no = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port_no, n, SEED, gens, q)) for r in range(world)]
    for p in procs:
        p.start()
    raw, pop_sum, hash_xor = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    merged = np.frombuffer(raw, dtype=HIT_DTYPE)
    ids, pops, ps, hx = Port().survey(0, n, SEED, 0, gens, n)
    assert (merged["id"] == ids).all() and (merged["g_first_false"] == pops).all()
    assert pop_sum == ps and hash_xor == hx


def test_bind_host_to_device_is_optional(monkeypatch):
    """The NUMA binding of one-process-per-GPU runs is an optimisation: without NVML (this container) or when
    switched off it reports why and leaves the affinity alone."""
    from lifeapi_b200 import bind_host_to_device
    before = os.sched_getaffinity(0)
    monkeypatch.setenv("LIFEB200_NO_NUMA_BIND", "1")
    assert bind_host_to_device(0).startswith("off")
    monkeypatch.delenv("LIFEB200_NO_NUMA_BIND")
    what = bind_host_to_device(0)
    assert isinstance(what, str) and what
    if what.startswith("unavailable"):
        assert os.sched_getaffinity(0) == before
    else:  # a box with NVML: restore every core for the rest of the session
        os.sched_setaffinity(0, before)
