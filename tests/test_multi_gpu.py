"""GPU parity tests of the ordered hit compaction, the device-resident survey / enumeration calls and the
multi-GPU groups of the C ABI (lifeb200_multi_*, csrc/multi.cu).  On a one-GPU box the group has world = 1
(no NCCL needed); with two or more GPUs the same calls run sharded and must return the same bytes."""
import ctypes as C

import numpy as np
import pytest

from conftest import SEED, golden

pytestmark = pytest.mark.gpu


def _device_count():
    from lifeapi_b200 import load_library
    n = C.c_int()
    load_library().lifeb200_device_count(C.byref(n))
    return n.value


def test_hits_leave_the_device_in_id_order(engine, port):
    """No host sort anywhere: ragged sizes around the 2048-id compaction tile, dense and sparse hit lists."""
    from lifeapi_b200 import HIT_DTYPE
    for (first, n, gens) in [(0, 2047, 2), (5, 2048, 2), (1 << 35, 2049, 2), (11, 6000, 16), (0, 1, 2), (123, 4097, 300)]:
        d_hits = engine.malloc(max(n, 1) * 16)
        try:
            nh, ps, hx = engine.soup_survey_dev(first, n, SEED, 0, gens, d_hits, n)
            got = np.zeros(nh, dtype=HIT_DTYPE)
            if nh:
                engine.d2h(got, d_hits)
                engine.sync()
        finally:
            engine.free(d_hits)
        ids, pops, ps_o, hx_o = port.survey(first, n, SEED, 0, gens, n)
        assert nh == len(ids) and (got["id"] == ids).all() and (got["g_first_false"] == pops).all(), (first, n, gens)
        assert (got["g_recovered"] == 0).all() and (ps, hx) == (ps_o, hx_o)


def test_overflow_reports_count_and_keeps_the_sorted_prefix(engine, port):
    from lifeapi_b200 import E_OVERFLOW, HIT_DTYPE, LifeB200Error
    first, n, gens, cap = 7, 5000, 2, 100
    ids, pops, _, _ = port.survey(first, n, SEED, 0, gens, n)
    assert len(ids) > cap
    d_hits = engine.malloc(cap * 16)
    try:
        with pytest.raises(LifeB200Error) as ei:
            engine.soup_survey_dev(first, n, SEED, 0, gens, d_hits, cap)
        assert ei.value.code == E_OVERFLOW and str(len(ids)) in str(ei.value)
        got = np.zeros(cap, dtype=HIT_DTYPE)
        engine.d2h(got, d_hits)
        engine.sync()
        assert (got["id"] == ids[:cap]).all() and (got["g_first_false"] == pops[:cap]).all()
    finally:
        engine.free(d_hits)


def test_enumeration_device_resident(engine):
    from lifeapi_b200 import HIT_DTYPE
    c = golden("collisions.npz")
    n_cat = int(c["enum_n_cat"])
    total = n_cat * 32 * 4096
    cap = 1 << 16
    d_hits = engine.malloc(cap * 16)
    try:
        for rep in range(2):  # the second call finds the tables cached in the context
            nh, skipped = engine.enumerate_collisions_dev(c["catalysts"][:n_cat], c["gliders"], 0, total, int(c["enum_gens"]),
                                                          int(c["enum_interval"]), d_hits, cap)
            got = np.zeros(nh, dtype=HIT_DTYPE)
            engine.d2h(got, d_hits)
            engine.sync()
            assert skipped == int(c["enum_skipped"])
            assert (got["id"] == c["enum_ids"]).all() and (got["g_first_false"] == c["enum_first_false"]).all()
            assert (got["g_recovered"] == c["enum_recovered"]).all()
        # a different table set must not be served from the cache
        nh2, _ = engine.enumerate_collisions_dev(c["catalysts"][1:n_cat], c["gliders"], 0, 32 * 4096, int(c["enum_gens"]),
                                                 int(c["enum_interval"]), d_hits, cap)
        got2 = np.zeros(nh2, dtype=HIT_DTYPE)
        engine.d2h(got2, d_hits)
        engine.sync()
        sel = (c["enum_ids"] >= 32 * 4096) & (c["enum_ids"] < 2 * 32 * 4096)
        assert (got2["id"] + 32 * 4096 == c["enum_ids"][sel]).all() and (got2["g_first_false"] == c["enum_first_false"][sel]).all()
    finally:
        engine.free(d_hits)


@pytest.fixture(scope="module")
def groups():
    """world-1 group always; a world-k group over every visible GPU when there are several."""
    from lifeapi_b200 import MultiEngine
    out = [MultiEngine([0])]
    k = _device_count()
    if k >= 2:
        out.append(MultiEngine(list(range(k))))
    yield out
    for g in out:
        g.close()


def test_multi_group_info(groups):
    for g in groups:
        info = g.info()
        assert info["world"] == info["n_local"] >= 1 and info["first_rank"] == 0
        if info["world"] > 1:
            assert info["nccl_version"] >= 21800
        r = [g.shard(100, 1003, k) for k in range(info["world"])]
        assert r[0][0] == 100 and r[-1][1] == 1103 and all(r[k][1] == r[k + 1][0] for k in range(len(r) - 1))


def test_multi_survey_equals_single_engine(groups, engine, port):
    for g in groups:
        g.reserve(4099, 4099)
        for mode in ((0, 1) if g.info()["world"] > 1 else (0,)):
            g.set_gather(mode)
            for (first, n, dist, gens) in [(0, 4099, 0, 96), (1 << 33, 1000, 1, 130), (3, 5, 0, 20), (9, 0, 0, 4)]:
                hits, ps, hx = g.soup_survey(first, n, SEED, dist, gens, cap=max(n, 1))
                h1, ps1, hx1 = engine.soup_survey(first, n, SEED, dist, gens, cap=max(n, 1))
                assert hits.tobytes() == h1.tobytes() and (ps, hx) == (ps1, hx1), (g.info(), mode, first, n)
                assert g.last_gather_ms() >= 0.0
        g.set_gather(0)
    ids, pops, ps_o, hx_o = port.survey(0, 4099, SEED, 0, 96, 4099)
    hits, ps, hx = groups[-1].soup_survey(0, 4099, SEED, 0, 96, cap=4099)
    assert (hits["id"] == ids).all() and (hits["g_first_false"] == pops).all() and (ps, hx) == (ps_o, hx_o)


def test_multi_enumeration_equals_golden(groups):
    from lifeapi_b200 import E_OVERFLOW, LifeB200Error
    c = golden("collisions.npz")
    n_cat = int(c["enum_n_cat"])
    total = n_cat * 32 * 4096
    for g in groups:
        hits, skipped = g.enumerate_collisions(c["catalysts"][:n_cat], c["gliders"], 0, total, int(c["enum_gens"]),
                                               int(c["enum_interval"]), cap=1 << 16)
        assert skipped == int(c["enum_skipped"])
        assert (hits["id"] == c["enum_ids"]).all() and (hits["g_first_false"] == c["enum_first_false"]).all()
        assert (hits["g_recovered"] == c["enum_recovered"]).all()
        with pytest.raises(LifeB200Error) as ei:
            g.enumerate_collisions(c["catalysts"][:n_cat], c["gliders"], 0, total, int(c["enum_gens"]),
                                   int(c["enum_interval"]), cap=10)
        assert ei.value.code == E_OVERFLOW


def test_multi_step_batch_host(groups, port):
    soups = port.soup(0, 1031, SEED, 0)
    want = port.step_batch(soups, 40)
    for g in groups:
        got = g.step_batch_host(soups, 40)
        assert (got == want).all()
        e0 = g.engine(0)  # a borrowed per-device context works like any other
        assert (e0.step_batch_host(soups[:33], 3) == port.step_batch(soups[:33], 3)).all()


def test_multi_bad_arguments():
    from lifeapi_b200 import E_ARG, LifeB200Error, MultiEngine
    with pytest.raises(LifeB200Error) as ei:
        MultiEngine([0, 0])
    assert ei.value.code == E_ARG
    with pytest.raises(LifeB200Error) as ei:
        MultiEngine([99])
    assert ei.value.code == E_ARG
    g = MultiEngine(device=0, rank=0, world=1)  # a one-rank "process per GPU" group needs no unique id
    try:
        hits, ps, hx = g.soup_survey(0, 100, SEED, 0, 8, cap=100)
        assert len(hits) > 0
        with pytest.raises(LifeB200Error):
            g.soup_survey(0, 100, SEED, 0, 1, cap=100)
    finally:
        g.close()


def _rank_worker(rank, world, uid, first, n, gens, q):
    import os
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from lifeapi_b200 import MultiEngine
    g = MultiEngine(device=rank, rank=rank, world=world, unique_id=uid)
    try:
        hits, ps, hx = g.soup_survey(first, n, SEED, 0, gens, cap=n)
        q.put((rank, hits.tobytes() if rank == 0 else b"", ps, hx))
    finally:
        g.close()


def test_multi_one_process_per_gpu(engine):
    """lifeb200_multi_create_rank: two processes, one GPU each, the unique id handed over by the parent; rank 0
    receives the merged list over NCCL, every rank the counters."""
    if _device_count() < 2:
        pytest.skip("needs two GPUs")
    import multiprocessing as mp
    from lifeapi_b200 import MultiEngine
    uid = MultiEngine.unique_id()
    first, n, gens = 1000, 5001, 64
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_rank_worker, args=(r, 2, uid, first, n, gens, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=300) for _ in range(2))
    for p in procs:
        p.join(60)
    h1, ps1, hx1 = engine.soup_survey(first, n, SEED, 0, gens, cap=n)
    assert res[0][1] == h1.tobytes()
    assert all((r[2], r[3]) == (ps1, hx1) for r in res)
