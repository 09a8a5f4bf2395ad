"""GPU parity tests of the stepping path, all through the C ABI (lifeapi_b200.capi.Engine is a
1:1 ctypes stub).  Oracle = oracle/ C restatement, pinned in test_oracle.py; golden = outputs of
the unmodified reference.  Bar: bit-exact (integer work) -- every comparison is ==."""
import numpy as np
import pytest

from conftest import SEED, golden

pytestmark = pytest.mark.gpu

LPPS = (1, 2, 4, 8, 16, 32)


def dev_step(engine, soups, gens, out_of_place=False):
    soups = np.ascontiguousarray(soups, dtype=np.uint64)
    n = soups.size // 64
    d_in = engine.malloc(max(soups.nbytes, 16))
    d_out = engine.malloc(max(soups.nbytes, 16)) if out_of_place else d_in
    try:
        engine.h2d(d_in, soups)
        engine.step_batch_dev(d_in, d_out, n, gens)
        out = np.empty_like(soups)
        engine.d2h(out, d_out)
        engine.sync()
        return out
    finally:
        engine.free(d_in)
        if out_of_place:
            engine.free(d_out)


def test_config1_rpentomino_every_generation(engine, port):
    """BASELINE config 1: R-pentomino, 1000 x Step(), full state compared at EVERY generation."""
    g = golden("rpentomino.npz")
    want = {int(k): g["states"][i] for i, k in enumerate(g["gens"])}
    cur = want[0].copy()
    d = engine.malloc(512)
    try:
        engine.h2d(d, cur)
        got = np.empty(64, dtype=np.uint64)
        for gen in range(1, 1001):
            engine.step_batch_dev(d, d, 1, 1)
            engine.d2h(got, d)
            engine.sync()
            cur = port.step(cur)
            assert (got == cur).all(), f"generation {gen}"
            if gen in want:
                assert (got == want[gen]).all(), f"golden generation {gen}"
    finally:
        engine.free(d)
    # and Step(n) in one launch
    assert (engine.step_batch_host(want[0], 1000) == want[1000]).all()
    assert (engine.step_batch_host(want[0], 1024) == want[1024]).all()


@pytest.mark.parametrize("lpp", LPPS)
def test_golden_soups_all_lane_groupings(engine, lpp):
    s = golden("soups.npz")
    engine.set_lanes_per_pattern(lpp)
    try:
        for i, k in enumerate(s["gens"]):
            assert (dev_step(engine, s["soups0"], int(k)) == s["stepped0"][i]).all(), (lpp, int(k))
            assert (dev_step(engine, s["soups1"], int(k), out_of_place=True) == s["stepped1"][i]).all(), (lpp, int(k))
        assert (dev_step(engine, s["mt1"][None], 1024)[0] == s["mt1_final"]).all()
        assert (dev_step(engine, s["mt42_first8"], 1024) == s["mt42_first8_final"]).all()
    finally:
        engine.set_lanes_per_pattern(0)


@pytest.mark.parametrize("lpp", LPPS)
@pytest.mark.parametrize("n", (1, 3, 31, 33, 1000, 4099))
def test_ragged_batches_vs_oracle(engine, port, lpp, n):
    """StepAltTest shape: seeded random states, intermediate generations G in {1,2,3,64}."""
    engine.set_lanes_per_pattern(lpp)
    try:
        soups = port.soup(77, n, SEED, n & 1)
        for g in (1, 2, 3, 64):
            assert (dev_step(engine, soups, g) == port.step_batch(soups, g)).all(), (lpp, n, g)
    finally:
        engine.set_lanes_per_pattern(0)


def test_edge_cases(engine, port):
    from lifeapi_b200 import E_ARG, LifeB200Error
    soups = port.soup(5, 9, SEED)
    # gens = 0 is the identity, in place and out of place
    assert (dev_step(engine, soups, 0) == soups).all()
    assert (dev_step(engine, soups, 0, out_of_place=True) == soups).all()
    # n = 0 is a no-op (null pointers allowed)
    engine.step_batch_dev(None, None, 0, 5)
    assert engine.step_batch_host(np.zeros((0, 64), dtype=np.uint64), 3).shape == (0, 64)
    # empty and full universes
    z = np.zeros((2, 64), dtype=np.uint64)
    z[1] = ~np.uint64(0)
    assert not dev_step(engine, z, 1).any()  # all-on dies of overcrowding on the torus
    # misaligned device pointer is rejected, not silently mis-read
    d = engine.malloc(1024 + 16)
    try:
        with pytest.raises(LifeB200Error) as ei:
            engine.step_batch_dev(d + 8, d + 8, 1, 1)
        assert ei.value.code == E_ARG
        with pytest.raises(LifeB200Error):
            engine.set_lanes_per_pattern(3)
    finally:
        engine.free(d)
    # torus wrap: a glider crossing both seams keeps period-4 translation (LifeAPI.hpp:1202-1212)
    gl = np.zeros(64, dtype=np.uint64)
    for x, y in [(1, 0), (2, 1), (0, 2), (1, 2), (2, 2)]:
        gl[(61 + x) & 63] |= np.uint64(1) << np.uint64((61 + y) & 63)
    assert (engine.step_batch_host(gl, 4) == port.moved(gl, 1, 1)).all()
    assert (engine.step_batch_host(gl, 256) == gl).all()  # 64 cells in 256 generations: back home


def test_host_path_pageable_and_pinned(engine, port):
    n = 70_000  # > one 64 Ki chunk: exercises the 3-slot pipeline and the ragged tail chunk
    soups = port.soup(1 << 20, n, SEED)
    want = port.step_batch(soups, 8)
    assert (engine.step_batch_host(soups, 8) == want).all()
    pin_in = engine.host_alloc((n, 64))
    pin_out = engine.host_alloc((n, 64))
    try:
        pin_in[...] = soups
        engine.step_batch_host(pin_in, 8, out=pin_out)
        assert (pin_out == want).all()
    finally:
        engine.host_free(pin_in)
        engine.host_free(pin_out)


def test_soup_source_matches_oracle(engine, port):
    for dist in (0, 1):
        n = 1027
        d = engine.malloc(n * 512)
        try:
            engine.soup_dev(d, 99, n, SEED, dist)
            got = np.empty((n, 64), dtype=np.uint64)
            engine.d2h(got, d)
            engine.sync()
        finally:
            engine.free(d)
        assert (got == port.soup(99, n, SEED, dist)).all()


def test_neighbour_count_planes(engine, port):
    p = golden("planes.npz")
    got = engine.neighbour_count_host(p["soups"])
    assert (got == p["ncount"]).all()
    soups = port.soup(31337, 777, SEED) & port.soup(4242, 777, SEED)
    got = engine.neighbour_count_host(soups)
    for i in range(0, 777, 7):
        assert (got[i] == port.neighbour_count(soups[i])).all()
    for lpp in (4, 16, 32):
        engine.set_lanes_per_pattern(lpp)
        try:
            assert (engine.neighbour_count_host(soups) == got).all()
        finally:
            engine.set_lanes_per_pattern(0)


def test_config2_scaled_down_vs_oracle(engine, port):
    """BASELINE config 2 at a size the oracle finishes in seconds: 8 Ki soups x 1024 generations,
    ALL final states compared, both soup distributions."""
    n = 8192
    for dist in (0, 1):
        soups = port.soup(dist * 10_000_000, n, SEED, dist)
        assert (dev_step(engine, soups, 1024) == port.step_batch(soups, 1024)).all()


def test_config2_full_size_properties(engine, port):
    """BASELINE config 2 at FULL size (1 Mi soups x 1024 generations) through size-independent
    properties: (a) the soup survey kernel (soups built in registers) and soup_dev + step_batch_dev +
    fold_dev agree on the checksum-of-checksums and the population sum; (b) Step(1024) ==
    Step(512) o Step(512); (c) a 1/1024 sample of the final states equals the oracle bit for bit."""
    n, gens = 1 << 20, 1024
    d = engine.malloc(n * 512)
    try:
        engine.soup_dev(d, 0, n, SEED, 0)
        engine.step_batch_dev(d, d, n, gens)
        pop_a, hash_a = engine.fold_dev(d, n, 0)
        sample_idx = np.arange(0, n, 1024)
        sample = np.empty((len(sample_idx), 64), dtype=np.uint64)
        for k, i in enumerate(sample_idx):
            engine.d2h(sample[k], d + int(i) * 512)
        engine.sync()
        engine.soup_dev(d, 0, n, SEED, 0)
        engine.step_batch_dev(d, d, n, 512)
        engine.step_batch_dev(d, d, n, 512)
        pop_b, hash_b = engine.fold_dev(d, n, 0)
    finally:
        engine.free(d)
    hits, pop_c, hash_c = engine.soup_survey(0, n, SEED, 0, gens, n)
    assert (pop_a, hash_a) == (pop_b, hash_b) == (pop_c, hash_c)
    ids = sample_idx.astype(np.uint64)
    soups = np.concatenate([port.soup(int(i), 1, SEED, 0) for i in ids])
    want = port.step_batch(soups, gens)
    assert (sample == want).all()
    # the survey's active list restricted to the sample agrees with the oracle's predicate
    before = port.step_batch(soups, gens - 2)
    active = {int(i) for i, a, b in zip(ids, before, want) if not (a == b).all()}
    assert active == set(int(h) for h in hits["id"] if int(h) % 1024 == 0)


def test_config2_full_size_every_final_state_vs_reference(engine, port):
    """BASELINE config 2 at FULL size, SURVEY 8d's check as stated: ALL 1 Mi final states after 1024
    generations equal the CPU's bit for bit.  The CPU side is the unmodified reference's Step(n)
    (oracle/_ref) spread over every host thread, else the C port; about 10 s on the GPU box's 16 threads."""
    from oracle.loader import Ref
    cpu = Ref.load() or port
    n, gens = 1 << 20, 1024
    soups = port.soup(0, n, SEED, 0)
    got = engine.step_batch_host(soups, gens)
    want = cpu.step_batch(soups, gens, 0)
    bad = np.flatnonzero((got != want).any(axis=1))
    assert bad.size == 0, f"{bad.size} of {n} final states differ, first id {bad[:4]}"


def test_translation_equivariance(engine, port):
    """Size-independent property: Step commutes with Moved (torus symmetry)."""
    soups = port.soup(900, 64, SEED)
    moved = np.stack([port.moved(s, 13, -27) for s in soups])
    a = dev_step(engine, moved, 100)
    b = np.stack([port.moved(s, 13, -27) for s in dev_step(engine, soups, 100)])
    assert (a == b).all()


@pytest.mark.parametrize("layout", [1, 2, 4, 8])
def test_step_every_register_layout(engine, port, layout):
    """The register layout of the step kernel (reference, even/odd rows, 2 or 4 patterns bit-interleaved per
    register) never shows: same bits for ragged batch sizes around the 2- and 4-pattern group size, both soup
    distributions, odd generation counts, in place and out of place."""
    engine.set_step_layout(layout)
    try:
        for n, gens, dist in ((1, 1, 0), (2, 3, 1), (3, 17, 0), (5, 64, 1), (63, 33, 0), (130, 7, 0), (1027, 100, 0)):
            soups = port.soup(900 + n, n, SEED, dist)
            assert (engine.step_batch_host(soups, gens) == port.step_batch(soups, gens)).all(), (layout, n, gens)
        g = golden("rpentomino.npz")
        want = {int(k): g["states"][i] for i, k in enumerate(g["gens"])}
        for gen in (1, 2, 3, 10, 64, 100, 256, 1000, 1024):
            assert (engine.step_batch_host(want[0][None, :], gen)[0] == want[gen]).all(), (layout, gen)
        soups = np.ascontiguousarray(port.soup(5, 77, SEED, 0))
        d = engine.malloc(soups.nbytes)
        try:
            engine.h2d(d, soups)
            engine.step_batch_dev(d, d, 77, 129)  # in place
            out = np.empty_like(soups)
            engine.d2h(out, d)
            engine.sync()
        finally:
            engine.free(d)
        assert (out == port.step_batch(soups, 129)).all()
    finally:
        engine.set_step_layout(0)
