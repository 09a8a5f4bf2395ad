"""GPU parity tests of the generated-circuit kernels (three-valued step, LifeHistory step,
LifeStable passes, refined steps), through the C ABI.  Oracle = oracle/stable_oracle.c (cube
tables of the reference's committed fragments), pinned against the unmodified reference in
test_oracle.py; golden = outputs of the reference's own LifeStable members.  Bar: bit-exact."""
import numpy as np
import pytest

from conftest import SEED, golden
from stable_inputs import stable_records, three_valued_patterns

pytestmark = pytest.mark.gpu


def test_unknown_step_golden(engine):
    g = golden("unknown_step.npz")
    for i, k in enumerate(g["gens"]):
        s, u = engine.unknown_step_batch_host(g["state"], g["unknown"], int(k))
        assert (s == g["next_state"][i]).all() and (u == g["next_unknown"][i]).all()


@pytest.mark.parametrize("layout", (0, 1, 2))
@pytest.mark.parametrize("lpp", (0, 4, 8, 16, 32))
@pytest.mark.parametrize("n", (1, 5, 33, 1000))
def test_unknown_step_vs_oracle(engine, port, lpp, n, layout):
    engine.set_lanes_per_pattern(lpp)
    engine.set_step_layout(layout)
    try:
        st, un = three_valued_patterns(n, 100 + n, port.soup(9000, n, SEED, 0), port.zoi)
        for gens in (0, 1, 7, 41):
            want = port.unknown_step_sop_batch(st, un, gens)
            got = engine.unknown_step_batch_host(st, un, gens)
            assert (got[0] == want[0]).all() and (got[1] == want[1]).all(), (lpp, n, gens)
    finally:
        engine.set_lanes_per_pattern(0)
        engine.set_step_layout(0)


@pytest.mark.parametrize("layout", (1, 2))
def test_unknown_step_two_select_path_equals_the_cover(engine, port, layout):
    """The default path steps disjoint planes through two B3/S23 selects (csrc/life_three.cuh) and overlapping
    ones through the fragment text; lifeb200_set_three_valued_path(1) forces the text for everything.  Both
    must agree bit for bit: dense unknowns, sparse unknowns, no unknowns, everything unknown, and a batch
    where every 7th pattern overlaps (so warps holding both kinds take the literal path as a whole)."""
    n = 4096
    rng = np.random.default_rng(31)
    soups = port.soup(0, n, SEED, 0)
    r = rng.integers(0, 1 << 63, size=(3, n, 64), dtype=np.uint64) << np.uint64(1)
    un = r[0].copy()
    un[n // 4: n // 2] &= r[1][n // 4: n // 2] & r[2][n // 4: n // 2]
    un[n // 2: 3 * n // 4] = 0
    un[-1] = ~np.uint64(0)
    st = soups & ~un
    mixed_st = st.copy()
    mixed_st[::7] |= un[::7] & r[1][::7]
    engine.set_step_layout(layout)
    try:
        for s_in, u_in in ((st, un), (mixed_st, un)):
            for gens in (1, 2, 16, 50):
                engine.set_three_valued_path(1)
                want = engine.unknown_step_batch_host(s_in, u_in, gens)
                engine.set_three_valued_path(0)
                got = engine.unknown_step_batch_host(s_in, u_in, gens)
                assert (got[0] == want[0]).all() and (got[1] == want[1]).all(), (layout, gens)
        sel = np.arange(0, n, 61)
        want = port.unknown_step_sop_batch(mixed_st[sel], un[sel], 16)
        got = engine.unknown_step_batch_host(mixed_st[sel], un[sel], 16)
        assert (got[0] == want[0]).all() and (got[1] == want[1]).all()
    finally:
        engine.set_three_valued_path(0)
        engine.set_step_layout(0)
    from lifeapi_b200 import LifeB200Error
    with pytest.raises(LifeB200Error):
        engine.set_three_valued_path(2)


def test_unknown_step_dont_care_rows_and_life_limit(engine, port):
    """Overlapping state/unknown planes hit the cover's don't-care rows: the kernel must follow the
    fragment text, not the semantics.  With no unknown cells the step is plain Life."""
    rng = np.random.default_rng(8)
    r = rng.integers(0, 1 << 63, size=(2, 200, 64), dtype=np.uint64) << np.uint64(1)
    want = port.unknown_step_sop_batch(r[0], r[1], 4)
    got = engine.unknown_step_batch_host(r[0], r[1], 4)
    assert (got[0] == want[0]).all() and (got[1] == want[1]).all()
    soups = port.soup(0, 64, SEED, 0)
    s, u = engine.unknown_step_batch_host(soups, np.zeros_like(soups), 20)
    assert (s == port.step_batch(soups, 20)).all() and not u.any()


def test_unknown_step_dev_in_place_and_config3_properties(engine, port):
    """Resident path at a larger size (64 Ki patterns): in place == out of place, a sample equals the
    oracle, and a pattern with every cell unknown stays unknown (size-independent properties)."""
    n = 1 << 16
    base = port.soup(0, n, SEED, 0)
    box = np.zeros(64, dtype=np.uint64)
    box[16:48] = np.uint64(0x0000FFFFFFFF0000)
    st = base & box
    un = (np.roll(base, 1, axis=0) & np.roll(base, 2, axis=0) & np.roll(base, 3, axis=0)) & box
    st &= ~un
    un[7] = ~np.uint64(0)
    st[7] = 0
    nbytes = st.nbytes
    ds, du, ds2, du2 = (engine.malloc(nbytes) for _ in range(4))
    try:
        engine.h2d(ds, st); engine.h2d(du, un)
        engine.unknown_step_batch_dev(ds, du, ds2, du2, n, 16)
        engine.unknown_step_batch_dev(ds, du, ds, du, n, 16)
        a, b, c, d = (np.empty_like(st) for _ in range(4))
        engine.d2h(a, ds); engine.d2h(b, du); engine.d2h(c, ds2); engine.d2h(d, du2)
        engine.sync()
    finally:
        for p in (ds, du, ds2, du2):
            engine.free(p)
    assert (a == c).all() and (b == d).all()
    sel = np.arange(0, n, 257)
    want = port.unknown_step_sop_batch(st[sel], un[sel], 16)
    assert (a[sel] == want[0]).all() and (b[sel] == want[1]).all()
    assert (b[7] == ~np.uint64(0)).all() and not a[7].any()
    assert not (a & b).any()  # ON and UNKNOWN stay disjoint on reachable inputs


def test_history_step(engine, port):
    soups = port.soup(300, 37, SEED, 0)
    rec = np.zeros((37, 4, 64), dtype=np.uint64)
    rec[:, 0] = soups
    rec[:, 1] = soups
    rec[:, 2] = port.soup(900, 37, SEED, 0)  # marked / original must come back untouched
    rec[:, 3] = soups
    out = engine.history_step_batch(rec, 25)
    cur, hist = soups.copy(), soups.copy()
    for _ in range(25):
        cur = port.step_batch(cur, 1)
        hist |= cur
    assert (out[:, 0] == cur).all() and (out[:, 1] == hist).all()
    assert (out[:, 2] == rec[:, 2]).all() and (out[:, 3] == rec[:, 3]).all()


def test_stable_passes_golden(engine):
    g = golden("stable.npz")
    for which in range(8):
        out, flags = engine.stable_pass_host(g["records"], which)
        assert (flags == g["flags"][which]).all(), which
        assert (out.reshape(g["passes"][which].shape) == g["passes"][which]).all(), which
    assert (engine.stable_vulnerable(g["vulnerable_in"]) == g["vulnerable"]).all()
    for v, key in ((0, "refined_m"), (1, "refined_old")):
        got = engine.unknown_step_refined(g["refined_records"], g["refined_on"], g["refined_unknown"], v)
        assert (np.stack(got) == g[key]).all(), key


def test_stable_passes_chained_vs_oracle(engine, port):
    rng = np.random.default_rng(2718)
    clean = stable_records(300, 4242)
    dirty = rng.integers(0, 1 << 63, size=(41, 10, 64), dtype=np.uint64) << np.uint64(1)
    half = rng.integers(0, 1 << 63, size=(40, 10, 64), dtype=np.uint64)
    half[:, 2:] &= rng.integers(0, 1 << 63, size=(40, 8, 64), dtype=np.uint64) & rng.integers(0, 1 << 63, size=(40, 8, 64), dtype=np.uint64)
    half[:, 0] &= ~half[:, 1]
    cur = np.concatenate([clean, dirty, half])
    seen = set()
    for which in (1, 2, 3, 0, 4, 6, 7, 5, 5):
        want, wf = port.stable_pass_batch(cur, which)
        got, gf = engine.stable_pass_host(cur, which)
        assert (gf == wf).all(), which
        assert (got.reshape(want.shape) == want).all(), which
        seen |= set(int(f) for f in wf)
        cur = want
    assert {0, 1, 3} <= seen  # inconsistent, consistent-unchanged and consistent-changed all exercised
    assert (engine.stable_vulnerable(cur) == port.stable_vulnerable(cur)).all()
    con = rng.integers(0, 1 << 63, size=(cur.shape[0], 64), dtype=np.uint64)
    for v in (0, 1):
        got = engine.unknown_step_refined(cur, con, cur[:, 1], v)
        want = port.refined_step(cur, con, cur[:, 1], v)
        assert all((x == y).all() for x, y in zip(got, want)), v


def test_stable_propagate_properties_at_size(engine, port):
    """16 Ki records (80 MiB) through the chunked host path: Propagate is idempotent (a second call
    reports consistent & unchanged and leaves the record alone), never un-decides a cell, and a
    sample equals the oracle."""
    base = stable_records(256, 11)
    recs = np.tile(base, (64, 1, 1))
    recs[:, 0] = np.roll(recs[:, 0], 3, axis=1)  # translate copies so chunks differ
    recs[:, 1] = np.roll(recs[:, 1], 3, axis=1)
    out, flags = engine.stable_pass_host(recs, 5)
    out = out.reshape(recs.shape)
    ok = (flags & 1) == 1
    assert ok.any()
    again, f2 = engine.stable_pass_host(out[ok], 5)
    assert (f2 == 1).all() and (again.reshape(out[ok].shape) == out[ok]).all()
    assert not (out[ok][:, 1] & ~recs[ok][:, 1]).any()  # unknown only shrinks
    sel = np.arange(0, recs.shape[0], 97)
    want, wf = port.stable_pass_batch(recs[sel], 5)
    assert (flags[sel] == wf).all() and (out[sel] == want).all()


def test_stable_argument_errors(engine):
    from lifeapi_b200 import LifeB200Error
    rec = np.zeros((1, 10, 64), dtype=np.uint64)
    with pytest.raises(LifeB200Error):
        engine.stable_pass_host(rec, 8)
    out, flags = engine.stable_pass_host(np.zeros((0, 10, 64), dtype=np.uint64), 5)
    assert out.size == 0 and flags.size == 0


def test_config3_full_size_every_plane_vs_reference(engine, ref):
    """BASELINE config 3 at its full size: 256 Ki partial patterns x 16 three-valued generations, BOTH planes of
    EVERY pattern against the reference harness (NeighbourCount + the committed unknown_step.hpp text, all
    host threads).  Inputs as bench_extras.config3_unknown_step builds them."""
    n, gens = 1 << 18, 16
    d = engine.malloc(n * 512)
    base = np.empty((n, 64), dtype=np.uint64)
    try:
        engine.soup_dev(d, 0, n, SEED, 0)
        engine.d2h(base, d)
        engine.sync()
    finally:
        engine.free(d)
    box = np.zeros(64, dtype=np.uint64)
    box[16:48] = np.uint64(0x0000FFFFFFFF0000)
    un = (np.roll(base, 1, axis=0) & np.roll(base, 2, axis=0) & np.roll(base, 3, axis=0)) & box
    st = (base & box) & ~un
    got_s, got_u = engine.unknown_step_batch_host(st, un, gens)
    want_s, want_u = ref.unknown_step_batch(st, un, gens, 0)
    assert (got_s == want_s).all() and (got_u == want_u).all()
    assert got_u.any() and got_s.any()
