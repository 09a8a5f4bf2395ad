"""GPU parity tests of the SURVEY 8(f) rank-4 rows: Convolve, MatchLive / MatchLiveAndDead / Match,
Match(LifeTarget), InteractionOffsets -- through the C ABI, bit-exact (==) against the golden vectors of
the unmodified reference and against the oracle on seeded inputs."""
import numpy as np
import pytest

from conftest import SEED, golden
from lifeapi_b200 import MATCH_BOUNDARY, MATCH_LIVE, MATCH_LIVE_AND_DEAD, LifeB200Error

pytestmark = pytest.mark.gpu


def sparse(rng, n, k):
    """n planes of density 2^-k"""
    out = rng.integers(0, 1 << 63, size=(n, 64), dtype=np.uint64) << np.uint64(1) | rng.integers(0, 2, size=(n, 64), dtype=np.uint64)
    for _ in range(k - 1):
        out &= rng.integers(0, 1 << 63, size=(n, 64), dtype=np.uint64) << np.uint64(1) | rng.integers(0, 2, size=(n, 64), dtype=np.uint64)
    return out


def test_convolve_golden(engine):
    g = golden("locate.npz")
    a, b = g["a"], g["b"]
    assert (engine.convolve_batch_host(a, b) == g["convolve"]).all()
    assert (engine.convolve_batch_host(b, a) == g["convolve"]).all()          # commutes
    assert (engine.convolve_batch_host(a[0], b) == g["convolve_bcast"]).all()  # one operand shared by all
    assert (engine.convolve_batch_host(b, a[:1]) == g["convolve_bcast"]).all()


def test_match_golden(engine, port):
    g = golden("locate.npz")
    s, l, d, t = g["states"], g["live"], g["dead"], g["targets"]
    assert (engine.match_batch_host(s, l, None, MATCH_LIVE) == g["match_live"]).all()
    assert (engine.match_batch_host(s, l, d, MATCH_LIVE_AND_DEAD) == g["match_live_and_dead"]).all()
    assert (engine.match_batch_host(s, l, None, MATCH_BOUNDARY) == g["match_state"]).all()
    assert (engine.match_target_batch_host(s, t) == g["match_target"]).all()
    assert (engine.match_batch_host(s, l[0], None, MATCH_BOUNDARY) == g["match_state_bcast"]).all()
    assert (engine.match_target_batch_host(s[3], t[3]) == g["match_target"][3]).all()
    # dense x dense operands (the a/b pairs) with dead = b.Mirrored(): match_on_ab = a.MatchLiveAndDead(b, b.Mirrored())
    a, b = g["a"], g["b"]
    mir = np.stack([port.mirrored(x) for x in b])
    assert (mir == np.stack([port.mirrored(port.mirrored(port.mirrored(x))) for x in b])).all()  # an involution
    assert (engine.match_batch_host(a, b, mir, MATCH_LIVE_AND_DEAD) == g["match_on_ab"]).all()
    assert not engine.convolve_batch_host(np.zeros(64, dtype=np.uint64), b).any()  # empty (*) x = empty


def test_interaction_offsets_golden(engine):
    g = golden("locate.npz")
    a, b = g["a"], g["b"]
    assert (engine.interaction_offsets_host(a, b) == g["offsets"]).all()
    assert (engine.interaction_offsets_host(a[11], b) == g["offsets_bcast"]).all()


@pytest.mark.parametrize("n", [0, 1, 7, 8, 9, 255, 1000])
def test_locate_vs_oracle_ragged_sizes(engine, port, n):
    rng = np.random.default_rng(1000 + n)
    a = sparse(rng, n, 1 + n % 6)
    b = sparse(rng, n, 6 - n % 6)
    conv = engine.convolve_batch_host(a, b)
    offs = engine.interaction_offsets_host(a, b)
    mtc = engine.match_batch_host(a, b, None, MATCH_BOUNDARY)
    assert conv.shape == (n, 64)
    step = max(1, n // 40)
    for i in range(0, n, step):
        assert (conv[i] == port.convolve(a[i], b[i])).all()
        assert (offs[i] == port.interaction_offsets(a[i], b[i])).all()
        assert (mtc[i] == port.match_state(a[i], b[i])).all()


def test_convolve_is_a_dilation_and_distributes(engine, port):
    """Size-independent properties on a large batch: a (*) single cell at (dx, dy) = a.Moved(dx, dy);
    a (*) (b | c) = (a (*) b) | (a (*) c); a (*) block-3x3-at-origin = a.ZOI()."""
    n = 20000  # crosses a chunk boundary of the host pipeline
    soups = port.soup(0, n, SEED, 0)
    a = soups & np.roll(soups, 1, axis=0) & np.roll(soups, 2, axis=0)  # density 1/8
    cell = np.zeros(64, dtype=np.uint64)
    cell[5] = np.uint64(1) << np.uint64(61)
    moved = engine.convolve_batch_host(a, cell)
    for i in (0, 1, 9999, n - 1):
        assert (moved[i] == port.moved(a[i], 5, 61)).all()
    block = np.zeros(64, dtype=np.uint64)
    block[[63, 0, 1]] = np.uint64(0x8000000000000003)  # rows -1, 0, 1 of columns -1, 0, 1
    zoi = engine.convolve_batch_host(block, a)
    for i in (0, 7, 12345, n - 1):
        assert (zoi[i] == port.zoi(a[i])).all()
    b, c = np.roll(a, 3, axis=0)[:2000] & soups[:2000], np.roll(a, 5, axis=0)[:2000] & soups[1:2001]
    lhs = engine.convolve_batch_host(a[:2000], b | c)
    rhs = engine.convolve_batch_host(a[:2000], b) | engine.convolve_batch_host(a[:2000], c)
    assert (lhs == rhs).all()


def test_match_finds_every_planted_catalyst(engine, port):
    """Plant a catalyst at a seeded offset in each of 4096 sparse soups; Match(catalyst) must be ON at that
    offset, and every ON offset must satisfy the definition (checked on a sample with the oracle)."""
    cats = golden("collisions.npz")["catalysts"]
    n = 4096
    rng = np.random.default_rng(5)
    states = sparse(rng, n, 4)
    offs = rng.integers(0, 64, size=(n, 2))
    which = rng.integers(0, len(cats), size=n)
    lives = np.empty((n, 64), dtype=np.uint64)
    for i in range(n):
        placed = port.moved(cats[which[i]], int(offs[i, 0]), int(offs[i, 1]))
        states[i] = (states[i] & ~port.zoi(placed)) | placed
        lives[i] = cats[which[i]]
    got = engine.match_batch_host(states, lives, None, MATCH_BOUNDARY)
    for i in range(n):
        x, y = int(offs[i, 0]), int(offs[i, 1])
        assert (int(got[i, x]) >> y) & 1, i
    for i in range(0, n, 97):
        assert (got[i] == port.match_state(states[i], lives[i])).all()


def test_locate_resident_path_and_errors(engine, port):
    g = golden("locate.npz")
    a, b = np.ascontiguousarray(g["a"]), np.ascontiguousarray(g["b"])
    n = len(a)
    da, db, do = engine.malloc(a.nbytes), engine.malloc(b.nbytes), engine.malloc(a.nbytes)
    try:
        engine.h2d(da, a)
        engine.h2d(db, b)
        out = np.empty_like(a)
        engine.convolve_batch_dev(da, n, db, n, do, n)
        engine.d2h(out, do)
        engine.sync()
        assert (out == g["convolve"]).all()
        engine.interaction_offsets_dev(da, n, db, n, do, n)
        engine.d2h(out, do)
        engine.sync()
        assert (out == g["offsets"]).all()
        engine.match_batch_dev(da, n, db, None, n, MATCH_LIVE, do)
        engine.d2h(out, do)
        engine.sync()
        for i in range(n):
            assert (out[i] == port.match_live(a[i], b[i])).all()
        with pytest.raises(LifeB200Error):
            engine.convolve_batch_dev(da, 3, db, n, do, n)  # count must be 1 or n
        with pytest.raises(LifeB200Error):
            engine.match_batch_dev(da, n, db, None, n, MATCH_LIVE_AND_DEAD, do)  # needs a dead plane
        with pytest.raises(LifeB200Error):
            engine.match_batch_dev(da, n, db, None, n, 7, do)
    finally:
        for p in (da, db, do):
            engine.free(p)
