"""lifeb200_step_match_static_dev: the allocation-free variant of lifeb200_step_match_dev (static __device__ scratch,
no counterpart in the reference) must return the same hit records as the existing entry point, bit for bit, on two
calls in a row -- the second sees the first call's leftovers in the shared scratch.  Collected last on purpose."""
import numpy as np
import pytest

from conftest import golden

pytestmark = pytest.mark.gpu

STATIC_MAX = 262144  # LIFEB200_STEP_MATCH_STATIC_MAX (include/lifeapi_b200.h)


def _placements(port, c, ids):
    cats, gl = c["catalysts"], c["gliders"]
    pats = np.empty((len(ids), 64), dtype=np.uint64)
    tidx = np.empty(len(ids), dtype=np.uint32)
    for k, i in enumerate(ids):
        i = int(i)
        dy, dx, cg = i & 63, (i >> 6) & 63, i >> 12
        tidx[k] = cg // 32
        pats[k] = cats[cg // 32] | port.moved(gl[cg % 32], dx, dy)
    return pats, tidx


def _run(engine, fn, pats, targets, tidx, gens, interval, first_id, cap):
    """Device buffers come from the caller (here: lifeb200_malloc, outside the call under test)."""
    from lifeapi_b200 import HIT_DTYPE
    n = len(pats)
    d_p, d_t = engine.malloc(max(pats.nbytes, 16)), engine.malloc(targets.nbytes)
    d_i, d_h, d_n = engine.malloc(max(tidx.nbytes, 16)), engine.malloc(max(cap, 1) * 16), engine.malloc(8)
    try:
        if n:
            engine.h2d(d_p, pats)
            engine.h2d(d_i, tidx)
        engine.h2d(d_t, targets)
        engine.memset(d_h, 0xEE, max(cap, 1) * 16)
        fn(d_p, d_t, d_i, n, first_id, gens, interval, d_h, cap, d_n)
        cnt = np.zeros(1, dtype=np.uint64)
        engine.d2h(cnt, d_n)
        engine.sync()
        hits = np.zeros(min(int(cnt[0]), cap), dtype=HIT_DTYPE)
        if len(hits):
            engine.d2h(hits, d_h)
            engine.sync()
        return int(cnt[0]), hits
    finally:
        for d in (d_p, d_t, d_i, d_h, d_n):
            engine.free(d)


def test_static_scratch_equals_existing_entry_point_twice(engine, port):
    c = golden("collisions.npz")
    targets = np.ascontiguousarray(c["targets"], dtype=np.uint64)
    rng = np.random.default_rng(11)
    # first call: 3000 placements (hits + random ones); second call: a SHORTER, different batch with another first_id,
    # so stale dense words and tile counts of call 1 lie right behind the range call 2 reads
    ids1 = np.concatenate([rng.choice(c["enum_ids"], 400, replace=False), rng.integers(0, 8 * 32 * 4096, size=2600).astype(np.uint64)])
    rng.shuffle(ids1)
    ids2 = np.concatenate([rng.choice(c["enum_ids"], 150, replace=False), rng.integers(0, 8 * 32 * 4096, size=1083).astype(np.uint64)])
    rng.shuffle(ids2)
    for ids, gens, interval, first_id in ((ids1, 128, 8, 77), (ids2, 96, 5, 1 << 40)):
        pats, tidx = _placements(port, c, ids)
        n_ref, ref = _run(engine, engine.step_match_dev, pats, targets, tidx, gens, interval, first_id, len(ids))
        n_got, got = _run(engine, engine.step_match_static_dev, pats, targets, tidx, gens, interval, first_id, len(ids))
        assert n_ref > 50
        assert n_got == n_ref
        assert got.tobytes() == ref.tobytes()  # integer records: bit-exact, same order
        # a sample against the oracle too, so that "equal" cannot mean "equally wrong"
        for h in got[:: max(1, len(got) // 16)]:
            r = port.step_match(pats[int(h["id"]) - first_id], targets[tidx[int(h["id"]) - first_id]], gens, interval)
            assert r and (int(h["g_first_false"]), int(h["g_recovered"])) == (r[0], r[1])


def test_static_scratch_lane_group_kernel_and_small_cap(engine, port):
    """The lane-group kernel (pinned lanes) through the static scratch; cap smaller than the hit count."""
    c = golden("collisions.npz")
    targets = np.ascontiguousarray(c["targets"], dtype=np.uint64)
    ids = np.random.default_rng(5).choice(c["enum_ids"], 300, replace=False)
    pats, tidx = _placements(port, c, ids)
    engine.set_lanes_per_pattern(8)
    try:
        n_ref, ref = _run(engine, engine.step_match_dev, pats, targets, tidx, 128, 8, 0, 40)
        n_got, got = _run(engine, engine.step_match_static_dev, pats, targets, tidx, 128, 8, 0, 40)
    finally:
        engine.set_lanes_per_pattern(0)
    assert n_got == n_ref > 40 and len(got) == 40 and got.tobytes() == ref.tobytes()
    n_got, got = _run(engine, engine.step_match_static_dev, pats[:0], targets, tidx[:0], 128, 8, 0, 4)
    assert n_got == 0 and len(got) == 0


def test_static_scratch_rejects_sizes_beyond_the_maximum(engine):
    from lifeapi_b200 import E_ARG, LifeB200Error
    d = engine.malloc(64)
    try:
        with pytest.raises(LifeB200Error) as e:
            engine.step_match_static_dev(d, d, None, STATIC_MAX + 1, 0, 16, 2, d, 1, d)  # rejected before any access
        assert e.value.code == E_ARG and "LIFEB200_STEP_MATCH_STATIC_MAX" in str(e.value)
    finally:
        engine.free(d)


def test_static_scratch_is_shared_by_every_context_of_the_device(engine, port):
    """The array belongs to the device, not to a lifeb200_ctx: a second context on the same GPU (its own stream) uses the
    same address, taken once; calls made one after the other from different contexts agree with the existing entry point."""
    from lifeapi_b200 import Engine
    c = golden("collisions.npz")
    targets = np.ascontiguousarray(c["targets"], dtype=np.uint64)
    ids = np.random.default_rng(23).choice(c["enum_ids"], 257, replace=False)
    pats, tidx = _placements(port, c, ids)
    n_ref, ref = _run(engine, engine.step_match_dev, pats, targets, tidx, 128, 8, 5, len(ids))
    other = Engine(0)
    try:
        n_b, got_b = _run(other, other.step_match_static_dev, pats, targets, tidx, 128, 8, 5, len(ids))
    finally:
        other.close()
    n_a, got_a = _run(engine, engine.step_match_static_dev, pats, targets, tidx, 128, 8, 5, len(ids))
    assert n_a == n_b == n_ref > 100
    assert got_a.tobytes() == ref.tobytes() and got_b.tobytes() == ref.tobytes()
