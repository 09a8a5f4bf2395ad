"""GPU parity tests of the fused step + LifeTarget match, collision enumeration and soup survey."""
import numpy as np
import pytest

from conftest import SEED, golden

pytestmark = pytest.mark.gpu


def placements(port, c, ids):
    """pattern = catalyst | projectile.Moved(dx, dy) for enumeration ids (include/lifeapi_b200.h)."""
    cats, gl = c["catalysts"], c["gliders"]
    pats = np.empty((len(ids), 64), dtype=np.uint64)
    tidx = np.empty(len(ids), dtype=np.uint32)
    for k, i in enumerate(ids):
        i = int(i)
        dy, dx, cg = i & 63, (i >> 6) & 63, i >> 12
        tidx[k] = cg // 32
        pats[k] = cats[cg // 32] | port.moved(gl[cg % 32], dx, dy)
    return pats, tidx


def test_step_match_vs_oracle(engine, port):
    c = golden("collisions.npz")
    targets = c["targets"]
    rng = np.random.default_rng(3)
    # 300 placements the reference found to be hits + 1200 random ones (mostly misses / overlaps)
    ids = np.concatenate([rng.choice(c["enum_ids"], 300, replace=False),
                          rng.integers(0, 8 * 32 * 4096, size=1200).astype(np.uint64)])
    rng.shuffle(ids)
    pats, tidx = placements(port, c, ids)
    n = len(ids)
    # lanes pinned -> step_match_kernel (layout 2); default / layout 8 -> step_match_kernel_x (4 patterns per lane
    # group, per-pattern targets packed on the device)
    for lpp, layout, gens, interval in ((4, 0, 128, 8), (8, 0, 160, 6), (32, 0, 90, 3), (0, 0, 128, 8), (0, 8, 20, 3), (0, 2, 128, 8)):
        engine.set_lanes_per_pattern(lpp)
        engine.set_step_layout(layout)
        try:
            hits = engine.step_match(pats, targets, tidx, gens=gens, stable_interval=interval, first_id=1000)
        finally:
            engine.set_lanes_per_pattern(0)
            engine.set_step_layout(0)
        want = []
        for p in range(n):
            r = port.step_match(pats[p], targets[tidx[p]], gens, interval)
            if r:
                want.append((1000 + p, r[0], r[1]))
        assert len(want) > (100 if gens > 50 else 0)
        assert [(int(h["id"]), int(h["g_first_false"]), int(h["g_recovered"])) for h in hits] == want, lpp


def test_step_match_single_target_and_overflow(engine, port):
    from lifeapi_b200 import E_ARG, E_OVERFLOW, LifeB200Error
    c = golden("collisions.npz")
    tgt = c["targets"][0]
    first = c["enum_ids"][c["enum_ids"] < 32 * 4096]  # catalyst 0 only
    ids = np.concatenate([first[:40], np.arange(5000, 5360, dtype=np.uint64)])
    pats, _ = placements(port, c, ids)
    hits = engine.step_match(pats, tgt[None], None, 128, 8)
    want = [k for k in range(len(ids)) if port.step_match(pats[k], tgt, 128, 8)]
    assert list(hits["id"]) == want and len(want) >= 40
    with pytest.raises(LifeB200Error) as ei:
        engine.step_match(pats, tgt[None], None, 128, 8, cap=1)
    assert ei.value.code == E_OVERFLOW
    with pytest.raises(LifeB200Error) as ei:
        engine.step_match(pats, tgt[None], None, 128, 0)
    assert ei.value.code == E_ARG
    assert len(engine.step_match(pats[:0], tgt[None], None, 10, 2)) == 0
    for n in (1, 2, 3, 5, 7):  # ragged around the 4-pattern group of step_match_kernel_x
        got = engine.step_match(pats[:n], tgt[None], None, 128, 8)
        assert list(got["id"]) == [k for k in range(n) if port.step_match(pats[k], tgt, 128, 8)]


def test_enumeration_vs_reference_golden(engine):
    """The golden list was produced by reference calls only (ZOI, Moved, |, Step, Contains)."""
    from lifeapi_b200 import E_OVERFLOW, LifeB200Error
    c = golden("collisions.npz")
    n_cat = int(c["enum_n_cat"])
    total = n_cat * 32 * 4096
    hits, skipped = engine.enumerate_collisions(c["catalysts"][:n_cat], c["gliders"], 0, total, int(c["enum_gens"]),
                                                int(c["enum_interval"]), cap=1 << 16)
    assert skipped == int(c["enum_skipped"])
    assert (hits["id"] == c["enum_ids"]).all()
    assert (hits["g_first_false"] == c["enum_first_false"]).all()
    assert (hits["g_recovered"] == c["enum_recovered"]).all()
    # shards of the id space concatenate to the same list (multi-GPU partitioning)
    from lifeapi_b200 import merge_hits, shard_range
    parts, sk = [], 0
    for r in range(3):
        lo, hi = shard_range(total, r, 3)
        h, s = engine.enumerate_collisions(c["catalysts"][:n_cat], c["gliders"], lo, hi, int(c["enum_gens"]),
                                           int(c["enum_interval"]), cap=1 << 16)
        parts.append(h)
        sk += s
    assert sk == skipped and (merge_hits(parts) == hits).all()
    with pytest.raises(LifeB200Error) as ei:
        engine.enumerate_collisions(c["catalysts"][:n_cat], c["gliders"], 0, total, int(c["enum_gens"]),
                                    int(c["enum_interval"]), cap=10)
    assert ei.value.code == E_OVERFLOW


def test_enumeration_vs_oracle_other_catalysts(engine, port):
    """Ragged id ranges (neither end a multiple of the 4-placement group), every kernel: lanes per pattern
    pinned -> enumerate_kernel; default with >= 48 gens or layout 8 -> enumerate_kernel_x (4 placements per
    lane group); layout 2 -> enumerate_kernel again."""
    c = golden("collisions.npz")
    cats = c["catalysts"]
    lo, hi = 4 * 32 * 4096 + 7, 4 * 32 * 4096 + 3 * 4096 + 11  # beehive, ragged range over 3+ glider variants
    want = {}
    for gens, interval in ((100, 5), (20, 3)):
        want[gens] = port.enumerate(cats, c["gliders"], lo, hi, gens, interval, 1 << 14)
    for lpp, layout, gens, interval in ((4, 0, 100, 5), (16, 0, 100, 5), (0, 0, 100, 5), (0, 8, 100, 5), (0, 2, 100, 5),
                                        (0, 8, 20, 3), (0, 0, 20, 3)):
        engine.set_lanes_per_pattern(lpp)
        engine.set_step_layout(layout)
        try:
            hits, skipped = engine.enumerate_collisions(cats, c["gliders"], lo, hi, gens, interval, cap=1 << 14)
        finally:
            engine.set_lanes_per_pattern(0)
            engine.set_step_layout(0)
        ids, ff, rec, sk = want[gens]
        assert skipped == sk, (lpp, layout, gens)
        assert (hits["id"] == ids).all() and (hits["g_first_false"] == ff).all() and (hits["g_recovered"] == rec).all()
    # tiny ranges inside one group, across a group edge, across a catalyst edge, and the last ids of the space
    total = len(cats) * 32 * 4096
    for a, b in ((5, 6), (6, 9), (32 * 4096 - 2, 32 * 4096 + 3), (total - 5, total + 100), (total, total + 4)):
        engine.set_step_layout(8)
        try:
            hits, skipped = engine.enumerate_collisions(cats, c["gliders"], a, b, 60, 4, cap=64)
        finally:
            engine.set_step_layout(0)
        ids, ff, rec, sk = port.enumerate(cats, c["gliders"], a, min(b, total), 60, 4, 64)
        assert skipped == sk and (hits["id"] == ids).all() and (hits["g_first_false"] == ff).all()


def test_soup_survey_vs_oracle(engine, port):
    for (first, n, dist, gens, lpp) in [(0, 333, 0, 300, 8), (1 << 33, 1000, 1, 130, 2), (17, 64, 0, 2, 32), (5, 257, 0, 64, 16)]:
        engine.set_lanes_per_pattern(lpp)
        try:
            hits, ps, hx = engine.soup_survey(first, n, SEED, dist, gens, cap=n)
        finally:
            engine.set_lanes_per_pattern(0)
        ids, pops, ps_o, hx_o = port.survey(first, n, SEED, dist, gens, n)
        assert (hits["id"] == ids).all() and (hits["g_first_false"] == pops).all()
        assert (ps, hx) == (ps_o, hx_o)


def test_soup_survey_default_path_four_soups_per_group(engine, port):
    """gens >= 96 with no pinned lane grouping takes survey_kernel_x (4 soups bit-interleaved per lane group):
    ragged counts around the group size, both distributions, huge first ids."""
    for (first, n, dist, gens) in [(0, 1, 0, 96), (3, 2, 1, 100), (9, 3, 0, 97), (1 << 40, 5, 0, 128), (77, 1023, 0, 200),
                                   (0, 4099, 1, 96)]:
        hits, ps, hx = engine.soup_survey(first, n, SEED, dist, gens, cap=n)
        ids, pops, ps_o, hx_o = port.survey(first, n, SEED, dist, gens, n)
        assert (hits["id"] == ids).all() and (hits["g_first_false"] == pops).all(), (first, n, gens)
        assert (ps, hx) == (ps_o, hx_o), (first, n, gens)


def test_fold_matches_oracle(engine, port):
    soups = port.soup(50, 999, SEED)
    d = engine.malloc(soups.nbytes)
    try:
        engine.h2d(d, soups)
        assert engine.fold_dev(d, 999, 50) == port.fold_batch(soups, 50)
    finally:
        engine.free(d)


def _survey_on_reference(ref, port, ids, gens):
    """(active, pop) of the soups `ids` after `gens` generations, by the reference's Step / == / GetPop on all threads."""
    ids = np.asarray(ids, dtype=np.uint64)
    if len(ids) and int(ids[-1]) - int(ids[0]) == len(ids) - 1:  # one contiguous run
        soups = port.soup(int(ids[0]), len(ids), SEED, 0)
    else:
        soups = np.concatenate([port.soup(int(i), 1, SEED, 0) for i in ids])
    _, act, pop = ref.survey_states_mt(soups, gens, 0)
    return act.astype(bool), pop


def test_config5_real_shape_sample_and_prefix_vs_reference(engine, ref, port):
    """BASELINE config 5 as SURVEY 8(d) words it: 16 Mi on-device soups x 4096 generations; a 1/1024 sample of the
    ids checked fully on the CPU (active flag and population of every sampled soup) and the whole hit list of a
    64 Ki prefix; plus the shard edges a multi-GPU run cuts at (first ids 2^21 k) and the last id 2^24 - 1."""
    total, gens = 1 << 24, 4096
    hits, pop_sum, hash_xor = engine.soup_survey(0, total, SEED, 0, gens, cap=total)
    ids = hits["id"]
    assert (np.diff(ids.astype(np.int64)) > 0).all() and int(ids[-1]) < total
    hit_pop = dict(zip(ids[ids % 1024 == 0].tolist(), hits["g_first_false"][ids % 1024 == 0].tolist()))
    # (1) 1/1024 sample: 16 Ki soups, every one stepped 4096 generations by the reference
    sample = np.arange(0, total, 1024, dtype=np.uint64)
    act, pop = _survey_on_reference(ref, port, sample, gens)
    assert set(sample[act].tolist()) == set(hit_pop)
    assert all(hit_pop[int(i)] == int(p) for i, p in zip(sample[act], pop[act]))
    # (2) 64 Ki prefix: the complete hit list (count, ids, populations)
    prefix = np.arange(0, 1 << 16, dtype=np.uint64)
    act, pop = _survey_on_reference(ref, port, prefix, gens)
    sel = ids < (1 << 16)
    assert int(act.sum()) == int(sel.sum()) and (ids[sel] == prefix[act]).all() and (hits["g_first_false"][sel] == pop[act]).all()
    # (3) shard edges: a rank that starts at 2^21 k reports the same records for its first ids as the full run,
    #     and the last 4096 ids of the space agree with the reference
    for k in range(1, 8):
        lo = k << 21
        h, _, _ = engine.soup_survey(lo, 8192, SEED, 0, gens, cap=8192)
        sel = (ids >= lo) & (ids < lo + 8192)
        assert h.tobytes() == hits[sel].tobytes(), k
    tail = np.arange(total - 4096, total, dtype=np.uint64)
    act, pop = _survey_on_reference(ref, port, tail, gens)
    sel = ids >= total - 4096
    assert (ids[sel] == tail[act]).all() and (hits["g_first_false"][sel] == pop[act]).all()
    # (4) checksum of checksums: the eight shards' counters add / xor up to the single call's
    ps, hx = 0, 0
    for k in range(0, 8, 4):  # two of the eight shards re-run (the other six are covered by the merged list above)
        _, a, b = engine.soup_survey(k << 21, 1 << 21, SEED, 0, gens, cap=1 << 21)
        ps, hx = ps + a, hx ^ b
    assert 0 < ps < pop_sum


def test_config4_full_size_vs_reference(engine, ref):
    """BASELINE config 4 at its full size: all 4 Mi placements (32 catalysts x 32 glider variants x 4096 offsets, <= 256
    generations, stable_interval 8) -- EVERY hit record and the skipped count against the reference's own
    Moved / ZOI / Step / Contains loop on all host threads (about half a minute of CPU)."""
    import os
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from bench_extras import D4, GLIDER, STILL_LIFES, _pattern
    names = list(STILL_LIFES)
    cats = np.stack([_pattern(STILL_LIFES[names[i % 8]], 28, 28, D4[(i // 8) * 2]) for i in range(32)])
    g0 = np.stack([_pattern(GLIDER, 8, 8, tf) for tf in D4])
    phases = [g0]
    for _ in range(3):
        phases.append(engine.step_batch_host(phases[-1], 1))
    gliders = np.ascontiguousarray(np.stack(phases, axis=1).reshape(32, 64))
    total = 32 * 32 * 4096
    hits, skipped = engine.enumerate_collisions(cats, gliders, 0, total, 256, 8, 1 << 20)
    ids, ff, rec, sk = ref.enumerate_mt(cats, gliders, 0, total, 256, 8, 1 << 20, 0)
    assert skipped == sk and len(hits) == len(ids) > 1000
    assert (hits["id"] == ids).all() and (hits["g_first_false"] == ff).all() and (hits["g_recovered"] == rec).all()
