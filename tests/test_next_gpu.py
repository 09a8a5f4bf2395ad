"""GPU parity tests of the SURVEY 8(f) rows: LifeWeld::Step, InteractionCountsAndNext, GetHash (XXH3)."""
import numpy as np
import pytest

from conftest import SEED, golden

pytestmark = pytest.mark.gpu


def test_next_rows_golden(engine):
    g = golden("next_rows.npz")
    for i, k in enumerate(g["weld_gens"]):
        assert (engine.weld_step_batch(g["weld"], int(k)) == g["weld_out"][i]).all()
    assert (engine.interaction_counts(g["inter_in"]) == g["inter_out"]).all()
    assert (engine.hash_batch_host(g["hash_in"]) == g["hash_out"]).all()


def test_hash_known_answers_and_sizes(engine, port):
    rp = golden("rpentomino.npz")
    assert (engine.hash_batch_host(rp["states"]) == rp["xxh3"]).all()  # SURVEY 8c known answers
    for n in (0, 1, 3, 31, 33, 4099):
        s = port.soup(50, n, SEED, 0)
        got = engine.hash_batch_host(s)
        assert [int(h) for h in got] == [port.hash(x) for x in s]
    # 70 000 patterns cross a chunk boundary of the host pipeline; spot-check + XOR fold
    s = port.soup(0, 70000, SEED, 1)
    got = engine.hash_batch_host(s)
    for i in (0, 65535, 65536, 69999):
        assert int(got[i]) == port.hash(s[i])


def test_step_then_hash_matches_survey_xor(engine):
    """XOR of GetHash over 8 stored mt19937_64(42) soups after 1024 gens: resident step + resident hash."""
    g = golden("soups.npz")
    soups = np.ascontiguousarray(g["mt42_first8"])
    d, dh = engine.malloc(soups.nbytes), engine.malloc(64)
    try:
        engine.h2d(d, soups)
        engine.step_batch_dev(d, d, 8, 1024)
        engine.hash_batch_dev(d, dh, 8)
        out, h = np.empty_like(soups), np.empty(8, dtype=np.uint64)
        engine.d2h(out, d)
        engine.d2h(h, dh)
        engine.sync()
    finally:
        engine.free(d)
        engine.free(dh)
    assert (out == g["mt42_first8_final"]).all()


@pytest.mark.parametrize("lpp", (0, 8, 16, 32))
def test_weld_and_interaction_vs_oracle(engine, port, lpp):
    engine.set_lanes_per_pattern(lpp)
    try:
        rng = np.random.default_rng(17)
        for n in (1, 5, 33, 257):
            s = port.soup(400, n, SEED, 0)
            w = np.zeros((n, 4, 64), dtype=np.uint64)
            w[:, 0] = s
            w[:, 1:] = rng.integers(0, 1 << 63, size=(n, 3, 64), dtype=np.uint64) & rng.integers(0, 1 << 63, size=(n, 3, 64), dtype=np.uint64)
            for gens in (1, 6):
                assert (engine.weld_step_batch(w, gens) == port.weld_step(w, gens)).all(), (lpp, n, gens)
            assert (engine.interaction_counts(s) == port.interaction_counts(s)).all(), (lpp, n)
        # frozen = 0 reduces to LifeState::Step; the `next` plane of InteractionCountsAndNext is Step too
        s = port.soup(0, 64, SEED, 0)
        w = np.zeros((64, 4, 64), dtype=np.uint64)
        w[:, 0] = s
        assert (engine.weld_step_batch(w, 11)[:, 0] == port.step_batch(s, 11)).all()
        assert (engine.interaction_counts(s)[:, 3] == port.step_batch(s, 1)).all()
    finally:
        engine.set_lanes_per_pattern(0)


def test_weld_step_every_kernel_vs_reference(engine, ref, port):
    """LifeWeld::Step through both kernels -- 4 records per lane group in the row-class layout (default from 24
    generations, or layout 8 pinned) and the lane-group kernel (layout 1 / pinned lanes) -- on dense frozen planes
    (every carry of the 3-bit add exercised), on sparse ones, and on records with no frozen cell (the plain-step
    fast path); ragged batch sizes around the 4-record group."""
    raw = port.soup(77, 4 * 1030, SEED, 0).reshape(1030, 4, 64)
    dense = raw.copy()
    sparse = raw.copy()
    sparse[:, 1] &= sparse[:, 2] & sparse[:, 3] & ~sparse[:, 0]
    sparse[:, 2] &= sparse[:, 3] & ~sparse[:, 0]
    sparse[:, 3] &= ~sparse[:, 0]
    plain = raw.copy()
    plain[:, 1:] = 0
    mixed = sparse.copy()
    mixed[::3, 1:] = 0  # warps with and without frozen cells side by side
    for name, recs in (("dense", dense), ("sparse", sparse), ("plain", plain), ("mixed", mixed)):
        for n in (1, 3, 4, 5, 1030):
            for layout, lpp, gens in ((0, 0, 40), (8, 0, 3), (1, 0, 40), (0, 16, 25)):
                engine.set_step_layout(layout)
                engine.set_lanes_per_pattern(lpp)
                try:
                    got = engine.weld_step_batch(recs[:n], gens)
                finally:
                    engine.set_step_layout(0)
                    engine.set_lanes_per_pattern(0)
                want = ref.weld_step_batch(recs[:n], gens, 0)
                assert (got == want).all(), (name, n, layout, lpp, gens)
    assert (engine.weld_step_batch(plain[:64], 30)[:, 0] == port.step_batch(plain[:64, 0], 30)).all()  # frozen = 0 is Step
