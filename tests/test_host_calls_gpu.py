"""GPU parity tests of the round-2 host-pointer entry points (the single-object members of the drop-in header go
through them) and of the shifted Contains batch, against the unmodified reference (oracle/_ref) and the C port."""
import numpy as np
import pytest

from conftest import SEED, golden
from stable_inputs import stable_records

pytestmark = pytest.mark.gpu


def test_weld_history_interaction_host_vs_reference(engine, ref, port):
    n = 1500  # more than one pipeline chunk would need 16 Ki records; ragged small batches exercise the plan edges
    raw = port.soup(40, 4 * n, SEED, 0).reshape(n, 4, 64)
    recs = raw.copy()
    recs[:, 1] &= recs[:, 2] & recs[:, 3] & ~recs[:, 0]
    recs[:, 2] &= recs[:, 3] & ~recs[:, 0]
    recs[:, 3] &= ~recs[:, 0]
    for gens in (1, 7, 40):
        assert (engine.weld_step_batch_host(recs, gens) == ref.weld_step_batch(recs, gens, 0)).all(), gens
    assert (engine.weld_step_batch_host(recs[:0], 3).size == 0)
    assert (engine.interaction_counts_host(raw[:, 0]) == ref.interaction_counts_batch(raw[:, 0], 0)).all()
    hist = np.zeros((n, 4, 64), dtype=np.uint64)
    hist[:, 0] = raw[:, 0]
    hist[:, 1] = raw[:, 0]
    hist[:, 2] = raw[:, 1]
    out = engine.history_step_batch_host(hist, 9)
    cur, acc = raw[:, 0].copy(), raw[:, 0].copy()
    for _ in range(9):
        cur = port.step_batch(cur, 1)
        acc |= cur
    assert (out[:, 0] == cur).all() and (out[:, 1] == acc).all() and (out[:, 2] == raw[:, 1]).all()


def test_vulnerable_and_refined_host_equal_device_path(engine, port):
    recs = stable_records(257, 21)
    recs, _ = engine.stable_pass_host(recs, 5)  # propagated records: option planes are populated
    assert (engine.stable_vulnerable_host(recs) == engine.stable_vulnerable(recs)).all()
    assert (engine.stable_vulnerable_host(recs) == port.stable_vulnerable(recs)).all()
    soups = port.soup(7, 257, SEED, 0)
    con = soups & ~recs[:, 1]
    cun = recs[:, 1] & np.roll(soups, 1, axis=0)
    con &= ~cun
    for variant in (0, 1):
        a = engine.unknown_step_refined_host(recs, con, cun, variant)
        b = port.refined_step(recs, con, cun, variant)
        assert all((x == y).all() for x, y in zip(a, b)), variant


def test_step_match_host_equals_device_path(engine, port):
    c = golden("collisions.npz")
    rng = np.random.default_rng(11)
    ids = np.concatenate([rng.choice(c["enum_ids"], 200, replace=False), rng.integers(0, 8 * 32 * 4096, size=600).astype(np.uint64)])
    cats, gl = c["catalysts"], c["gliders"]
    pats = np.empty((len(ids), 64), dtype=np.uint64)
    tidx = np.empty(len(ids), dtype=np.uint32)
    for k, i in enumerate(ids):
        i = int(i)
        tidx[k] = (i >> 12) // 32
        pats[k] = cats[(i >> 12) // 32] | port.moved(gl[(i >> 12) % 32], (i >> 6) & 63, i & 63)
    a = engine.step_match_host(pats, c["targets"], tidx, 128, 8, first_id=77)
    b = engine.step_match(pats, c["targets"], tidx, 128, 8, first_id=77)
    assert len(a) > 100 and a.tobytes() == b.tobytes()
    want = [(77 + p, *r) for p in range(len(ids)) if (r := port.step_match(pats[p], c["targets"][tidx[p]], 128, 8))]
    assert [(int(h["id"]), int(h["g_first_false"]), int(h["g_recovered"])) for h in a] == want
    one = engine.step_match_host(pats[tidx == 0], c["targets"][:1], None, 128, 8)
    assert len(one) == sum(1 for p in np.nonzero(tidx == 0)[0] if port.step_match(pats[p], c["targets"][0], 128, 8))


def test_contains_target_shifted_vs_reference(engine, ref, port):
    """LifeState::Contains(const LifeTarget&, dx, dy) (LifeTarget.hpp:38-42): a catalyst planted at a seeded offset in
    every soup is found at that offset and (mostly) nowhere else; every verdict equals the reference's."""
    eater = port.parse("2b2o$bobo$bo$2o!")
    target = ref.target_from_state(eater).reshape(2, 64)
    n = 600
    soups = port.soup(900, n, SEED, 0) & port.soup(5000, n, SEED, 0) & port.soup(9000, n, SEED, 0)
    rng = np.random.default_rng(2)
    offs = rng.integers(0, 64, size=(n, 2))
    states = np.empty_like(soups)
    for p in range(n):
        placed = port.moved(eater, int(offs[p, 0]), int(offs[p, 1]))
        states[p] = (soups[p] & ~port.zoi(placed)) | placed
    for (dx, dy) in [(0, 0), (5, 60), (63, 1), (-3, 70), (int(offs[0, 0]), int(offs[0, 1]))]:
        got = engine.contains_target_batch_host(states, target, dx, dy)
        want = np.array([ref.contains_target_shifted(states[p], target, dx, dy) for p in range(n)])
        assert (got == want).all(), (dx, dy)
    assert engine.contains_target_batch_host(states[:1], target, int(offs[0, 0]), int(offs[0, 1]))[0]
    # one target per state (moved to where the catalyst was planted): Contains at offset (0, 0) for all of them
    per = np.stack([np.stack([port.moved(target[0], int(o[0]), int(o[1])), port.moved(target[1], int(o[0]), int(o[1]))]) for o in offs])
    assert engine.contains_target_batch_host(states, per, 0, 0).all()
    # resident form
    d_s, d_t, d_o = engine.malloc(states.nbytes), engine.malloc(target.nbytes), engine.malloc(n)
    try:
        engine.h2d(d_s, states); engine.h2d(d_t, np.ascontiguousarray(target))
        engine.contains_target_batch_dev(d_s, n, d_t, 1, 5, 60, d_o)
        out = np.zeros(n, dtype=np.uint8)
        engine.d2h(out, d_o)
        engine.sync()
        assert (out.astype(bool) == engine.contains_target_batch_host(states, target, 5, 60)).all()
    finally:
        for p in (d_s, d_t, d_o):
            engine.free(p)


def test_partial_overlap_rejected_and_in_place_allowed(engine, port):
    from lifeapi_b200 import E_ARG, LifeB200Error
    soups = port.soup(0, 64, SEED, 0)
    d = engine.malloc(2 * soups.nbytes)
    try:
        engine.h2d(d, soups)
        engine.step_batch_dev(d, d, 64, 5)  # in place: allowed
        out = np.empty_like(soups)
        engine.d2h(out, d)
        engine.sync()
        assert (out == port.step_batch(soups, 5)).all()
        with pytest.raises(LifeB200Error) as ei:
            engine.step_batch_dev(d, d + 512, 64, 1)  # shifted by one pattern: partial overlap
        assert ei.value.code == E_ARG
    finally:
        engine.free(d)


def test_neighbour_count_arithmetic_batches_vs_reference(engine, ref, port):
    """NeighbourCount::Add / operator+ / Subtract / WithExactly (NeighbourCount.hpp:71-102) on batches: real count
    records (0..9 per cell) and arbitrary planes (every carry pattern), with and without a carry-in plane."""
    n = 777
    soups = port.soup(11, 2 * n, SEED, 0)
    ca, cb = engine.neighbour_count_host(soups[:n]), engine.neighbour_count_host(soups[n:])
    ra, rb = port.soup(5000, 4 * n, SEED, 0).reshape(n, 4, 64), port.soup(9000, 4 * n, SEED, 0).reshape(n, 4, 64)
    carry = port.soup(123, n, SEED, 0)
    for a, b in ((ca, cb), (ra, rb)):
        assert (engine.nc_arith("add", a, b) == ref.nc_add_batch(a, b)).all()
        assert (engine.nc_arith("add", a, b, incarry=carry) == ref.nc_add_batch(a, b, carry)).all()
        assert (engine.nc_arith("subtract", a, b) == ref.nc_subtract_batch(a, b)).all()
        for k in (0, 3, 4, 9, 15):
            assert (engine.nc_arith("with_exactly", a, k=k) == ref.nc_with_exactly_batch(a, k)).all(), k
    z = engine.nc_arith("subtract", ca, ca)
    assert not z.any()  # x - x == 0 on every plane
    assert engine.nc_arith("add", ca[:0], cb[:0]).size == 0


def test_weld_interaction_counts_vs_reference(engine, ref, port):
    """LifeWeld::InteractionCounts (LifeWeld.hpp:193-215): welded still lifes (some cells frozen) and dense random records."""
    n = 515
    raw = port.soup(31, 4 * n, SEED, 0).reshape(n, 4, 64)
    recs = raw.copy()
    recs[:, 0] &= raw[:, 1] & raw[:, 2]          # sparser states: isolated cells and small clumps, real interaction counts
    recs[:, 1] &= recs[:, 0] & raw[:, 3]         # a share of the live cells is frozen ...
    recs[:, 2] = np.roll(recs[:, 1], 1, axis=0) & ~recs[:, 0]   # ... and some dead ones carry frozen counts too
    recs[:, 3] &= raw[:, 2] & raw[:, 1]
    for r in (recs, raw):
        assert (engine.weld_interaction_counts_host(r) == ref.weld_interaction_counts_batch(r, 0)).all()
    plain = recs.copy()
    plain[:, 1:] = 0  # no frozen cells: LifeState::InteractionCounts masked by the state's own ZOI (a no-op on these planes)
    got = engine.weld_interaction_counts_host(plain)
    assert (got == engine.interaction_counts_host(plain[:, 0])[:, :3]).all()


def test_weld_interaction_offsets_vs_reference(engine, ref, port):
    """LifeWeld::InteractionOffsets (LifeWeld.hpp:217-262): one welded catalyst against many small objects (shared
    operand), per-problem pairs, and welds without frozen cells."""
    n = 300
    eater = port.parse("2b2o$bobo$bo$2o!", 20, 20)
    cat = np.zeros((1, 4, 64), dtype=np.uint64)
    cat[0, 0] = eater
    cat[0, 3] = port.parse("2o!", 22, 23)           # two stator cells frozen ...
    cat[0, 2] = port.parse("o!", 19, 22)            # ... and a frozen count on a dead neighbour
    box = np.zeros(64, dtype=np.uint64)
    box[:8] = np.uint64(0xFF)
    objs = np.zeros((n, 4, 64), dtype=np.uint64)
    objs[:, 0] = port.soup(4000, n, SEED, 0) & port.soup(8000, n, SEED, 0) & box
    objs[::2, 3] = objs[::2, 0] & port.soup(12000, n, SEED, 0)[::2]   # half of the objects have frozen cells too
    got = engine.weld_interaction_offsets_host(cat, objs)
    assert (got == ref.weld_interaction_offsets_batch(cat, objs, 0)).all()
    got2 = engine.weld_interaction_offsets_host(objs, np.roll(objs, 1, axis=0))
    assert (got2 == ref.weld_interaction_offsets_batch(objs, np.roll(objs, 1, axis=0), 0)).all()
    # no frozen cells anywhere: still the reference's member (its second and third terms mask the OTHER weld's operand
    # with the first weld's zone, so this is NOT LifeState::InteractionOffsets -- the quirk is part of the parity)
    plain = objs.copy()
    plain[:, 1:] = 0
    pc = cat.copy()
    pc[:, 1:] = 0
    assert (engine.weld_interaction_offsets_host(pc, plain) == ref.weld_interaction_offsets_batch(pc, plain, 0)).all()


def test_host_register_caller_owned_buffer(engine, port):
    """lifeb200_host_register: a caller-owned (numpy) batch is page-locked in place and stepped through the host
    pipeline; same bytes as the pageable path, several pipeline chunks."""
    n = 70_001
    soups = port.soup(0, n, SEED, 0)
    want = engine.step_batch_host(soups, 12)
    mine = soups.copy()
    out = np.empty_like(mine)
    engine.host_register(mine)
    engine.host_register(out)
    try:
        engine.step_batch_host(mine, 12, out=out)
        assert (out == want).all()
        engine.step_batch_host(out, 0, out=out)  # gens = 0 on a registered buffer, in place
        assert (out == want).all()
    finally:
        engine.host_unregister(mine)
        engine.host_unregister(out)
    assert (want[:64] == port.step_batch(soups[:64], 12)).all()
