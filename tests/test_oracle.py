"""CPU tests: pin the oracle (C restatement) against the reference's known answers, the golden
vectors produced by the unmodified reference, and -- when oracle/_ref is present -- the reference
itself, differentially (the pattern of tests/StepAltTest.cpp:5-13: seeded random states, bit-exact ==)."""
import numpy as np
import pytest

from conftest import SEED, golden

# SURVEY.md 8(c) / BASELINE.md 5: reference Step() on LifeState::Parse("b2o$2o$bo!")
KNOWN = {1: (6, 0x0549DE903160DDAF), 2: (7, 0x5ACAA23574BB6D97), 10: (11, 0x4D9E729E7A379063),
         100: (121, 0x03517B63993A38C3), 256: (76, 0x50ABDA84234A962D), 500: (247, 0x6559805396198161),
         1000: (113, 0xDC35D1573487DC91), 1024: (113, 0xDC35D1573487DC91)}


def test_golden_matches_survey_known_answers():
    g = golden("rpentomino.npz")
    gens = list(g["gens"])
    for gen, (pop, h) in KNOWN.items():
        i = gens.index(gen)
        assert int(g["pops"][i]) == pop
        assert int(g["xxh3"][i]) == h
    s = golden("soups.npz")
    assert int(s["mt1_pop0"]) == 2014 and int(s["mt1_pop"]) == 340
    assert int(s["mt1_xxh3"]) == 0x77B09695A0065A59
    assert int(s["mt42_xor_xxh3"]) == 0xA0384537038C7533


def test_port_rpentomino_every_generation(port):
    g = golden("rpentomino.npz")
    want = {int(k): g["states"][i] for i, k in enumerate(g["gens"])}
    cur = want[0].copy()
    assert port.pop(cur) == 5
    for gen in range(1, 1025):
        nxt = port.step(cur)
        assert (nxt == port.step(cur, "alt")).all()
        if gen <= 64 or gen % 97 == 0:
            assert (nxt == port.step(cur, "naive")).all()
        cur = nxt
        if gen in want:
            assert (cur == want[gen]).all(), f"gen {gen}"
            assert port.pop(cur) == int(g["pops"][list(g["gens"]).index(gen)])


def test_port_soups_golden(port):
    s = golden("soups.npz")
    assert (port.soup(0, 16, int(s["seed"]), 0) == s["soups0"]).all()
    assert (port.soup(0, 16, int(s["seed"]), 1) == s["soups1"]).all()
    # reference-shaped distribution: bits 63,62 clear, bit 61 set (LifeAPI.hpp:21-22)
    assert (s["soups1"] >> np.uint64(62) == 0).all() and ((s["soups1"] >> np.uint64(61)) & np.uint64(1) == 1).all()
    for i, k in enumerate(s["gens"]):
        assert (port.step_batch(s["soups0"], int(k)) == s["stepped0"][i]).all()
        assert (port.step_batch(s["soups1"], int(k), threads=1) == s["stepped1"][i]).all()
    assert (port.step_batch(s["mt1"][None], 1024)[0] == s["mt1_final"]).all()
    assert (port.step_batch(s["mt42_first8"], 1024) == s["mt42_first8_final"]).all()


def test_port_planes_golden(port):
    p = golden("planes.npz")
    for i in range(8):
        assert (port.neighbour_count(p["soups"][i]) == p["ncount"][i]).all()
    for i in range(len(p["zoi_in"])):
        assert (port.zoi(p["zoi_in"][i]) == p["zoi"][i]).all()
    for (x, y), want in zip(p["moves"], p["moved"]):
        assert (port.moved(p["soups"][0], int(x), int(y)) == want).all()


def test_port_neighbour_count_is_the_3x3_sum(port):
    s = port.soup(7, 1, SEED)[0]
    planes = port.neighbour_count(s)  # bit3, bit2, bit1, bit0
    bits = ((s[:, None] >> np.arange(64, dtype=np.uint64)[None, :]) & np.uint64(1)).astype(np.int64)  # [x, y]
    total = sum(np.roll(np.roll(bits, dx, 0), dy, 1) for dx in (-1, 0, 1) for dy in (-1, 0, 1))
    got = sum(((planes[3 - k][:, None] >> np.arange(64, dtype=np.uint64)[None, :]) & np.uint64(1)).astype(np.int64) << k
              for k in range(4))
    assert (got == total).all()


def test_port_collisions_golden(port):
    c = golden("collisions.npz")
    for cat, tgt in zip(c["catalysts"], c["targets"]):
        assert (port.target_from_state(cat) == tgt).all()
        assert port.contains_target(cat, tgt)
    n_cat = int(c["enum_n_cat"])
    # a slice of the id space (the whole space is checked on the GPU): first catalyst, 3 glider variants
    lo, hi = 0, 3 * 4096
    ids, ff, rec, skipped = port.enumerate(c["catalysts"][:n_cat], c["gliders"], lo, hi, int(c["enum_gens"]),
                                           int(c["enum_interval"]), 1 << 16)
    sel = (c["enum_ids"] >= lo) & (c["enum_ids"] < hi)
    assert (ids == c["enum_ids"][sel]).all()
    assert (ff == c["enum_first_false"][sel]).all() and (rec == c["enum_recovered"][sel]).all()


def test_port_unknown_step_golden(port):
    u = golden("unknown_step.npz")
    assert not (u["state"] & u["unknown"]).any()
    for i, k in enumerate(u["gens"]):
        s, k_ = port.unknown_step_batch(u["state"], u["unknown"], int(k))
        assert (s == u["next_state"][i]).all(), f"state after {k} gens"
        assert (k_ == u["next_unknown"][i]).all(), f"unknown after {k} gens"
        assert not (s & k_).any()


def test_port_unknown_step_reduces_to_life_without_unknowns(port):
    s = port.soup(3, 4, SEED)
    z = np.zeros_like(s)
    a, b = port.unknown_step_batch(s, z, 5)
    assert (a == port.step_batch(s, 5)).all() and not b.any()


# ---- differential against the compiled reference (this container only) -------------------------

def test_port_vs_reference_step_random(port, ref):
    """tests/StepAltTest.cpp shape: many random states, Step == StepAlt == port, bit-exact."""
    n = 2000
    soups = np.concatenate([port.soup(10_000, n // 2, SEED, 0), port.soup(20_000, n // 2, SEED, 1)])
    for g in (1, 2, 7):
        assert (port.step_batch(soups, g) == ref.step_batch(soups, g)).all()
    for s in soups[:200]:
        r = ref.step(s)
        assert (r == ref.step(s, "alt")).all() and (r == port.step(s)).all() and (r == port.step(s, "alt")).all()


def test_port_vs_reference_planes_and_targets(port, ref):
    soups = port.soup(500, 64, SEED, 0)
    sparse = soups & port.soup(600, 64, SEED, 0) & port.soup(700, 64, SEED, 0)
    for s in list(soups[:32]) + list(sparse[:32]):
        assert (port.neighbour_count(s) == ref.neighbour_count(s)).all()
        assert (port.zoi(s) == ref.zoi(s)).all()
        t = ref.target_from_state(s)
        assert (port.target_from_state(s) == t).all()
        for other in (s, ref.step(s), soups[0]):
            assert port.contains_target(other, t) == ref.contains_target(other, t)
        assert port.pop(s) == ref.pop(s)
    for x, y in [(0, 0), (5, -7), (-64, 64), (100, 3), (-1, -1)]:
        assert (port.moved(soups[1], x, y) == ref.moved(soups[1], x, y)).all()


def test_port_vs_reference_step_match_and_enumerate(port, ref):
    c = golden("collisions.npz")
    cats, gl = c["catalysts"], c["gliders"]
    lo, hi = 5 * 4096 + 100, 5 * 4096 + 1500
    a = port.enumerate(cats[:3], gl, lo, hi, 96, 4, 4096)
    b = ref.enumerate(cats[:3], gl, lo, hi, 96, 4, 4096)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)
    t = ref.target_from_state(cats[0])
    for gi in range(0, 32, 5):
        for dx, dy in [(20, 20), (24, 22), (22, 25), (30, 18)]:
            s = cats[0] | ref.moved(gl[gi], dx, dy)
            assert port.step_match(s, t, 200, 6) == ref.step_match(s, t, 200, 6)


def test_port_vs_reference_unknown_step(port, ref):
    rng = np.random.default_rng(7)
    n = 64
    state = port.soup(4000, n, SEED, 0) & port.soup(5000, n, SEED, 0)
    unk = rng.integers(0, 1 << 63, size=(n, 64), dtype=np.uint64) & rng.integers(0, 1 << 63, size=(n, 64), dtype=np.uint64) \
        & rng.integers(0, 1 << 63, size=(n, 64), dtype=np.uint64)
    unk[: n // 2] &= np.array([ref.zoi(s) for s in state[: n // 2]])
    state &= ~unk
    for g in (1, 3):
        ps, pu = port.unknown_step_batch(state, unk, g)
        rs, ru = ref.unknown_step_batch(state, unk, g)
        assert (ps == rs).all() and (pu == ru).all()


def test_port_survey_matches_reference_predicates(port, ref):
    n, gens = 48, 300
    ids, pops, ps, hx = port.survey(123, n, SEED, 0, gens, n)
    soups = port.soup(123, n, SEED, 0)
    final, act, pop = ref.survey_states(soups, gens)
    assert (np.nonzero(act)[0] + 123 == ids).all()
    assert (pop[act.astype(bool)] == pops).all()
    assert int(pop.sum()) == ps
    assert port.fold_batch(final, 123) == (ps, hx)


# ---- generated circuits: emitter, cube tables, LifeStable passes ------------------------------
def _lop3(lut, a, b, c):
    r = 0
    for k in range(8):
        if lut >> k & 1:
            r |= (a if k & 4 else ~a) & (b if k & 2 else ~b) & (c if k & 1 else ~c)
    return r & 1


def test_rule_luts_exhaustive():
    """The 7-LUT B3/S23 select in csrc/life_core.cuh (rule32) equals the rule on all 2^7 inputs."""
    import os
    import re
    from conftest import ROOT
    src = open(os.path.join(ROOT, "lifeapi_b200", "csrc", "life_core.cuh")).read()
    body = src[src.index("uint32_t rule32("):]
    body = body[:body.index("\n}\n")]
    consts = {"kNotAllEq": 0x7E, "kMaj": 0xE8, "kXor3": 0x96}
    luts = [consts.get(m, None) if not m.startswith("0x") else int(m, 16) for m in re.findall(r"lop3<(\w+)>", body)]
    assert luts == [0x7E, 0xE8, 0x96, 0xE8, 0x16, 0x6D, 0x09]
    for r in range(128):
        a, u0, c0, d0, u1, c1, d1 = [(r >> i) & 1 for i in range(7)]
        s = u0 + c0 + d0 + 2 * (u1 + c1 + d1)
        xa, xb = _lop3(0x7E, u0, c0, d0), _lop3(0xE8, u0, c0, d0)
        ya, yb = _lop3(0x96, u1, c1, d1), _lop3(0xE8, u1, c1, d1)
        g1 = _lop3(0x16, xb, xa, a)
        g2 = _lop3(0x6D, g1, yb, a)
        assert _lop3(0x09, g2, ya, xa) == int(s == 3 or (a and s == 4))


def test_emitted_circuits_equal_the_fragments_on_every_input():
    """circuits.json -> LUT networks == SOP cubes on all 2^k inputs; generated files are in sync."""
    import os
    from conftest import ROOT
    from lifeapi_b200.bitslicing import emit_cuda as E
    items = E.load_json()
    assert [c.name for c, _ in items] == list(E.CIRCUITS)
    for c, net in items:
        E.verify(c, net)
        assert E.synthesize(c).to_json() == net.to_json()
        assert c.inputs == E.CIRCUITS[c.name]["inputs"] and c.outputs == E.CIRCUITS[c.name]["outputs"]
    assert E.render_cuda(items) == open(E.CUH_PATH).read()
    from oracle import make_tables
    assert make_tables.render_tables([c for c, _ in items]) == open(make_tables.TABLES_PATH).read()
    ref_dir = "/root/reference/bitslicing"
    if os.path.isdir(ref_dir):  # the committed cubes are exactly what the reference's text says
        for c, _ in items:
            spec = E.CIRCUITS[c.name]
            parsed = E.parse_hpp(c.name, open(os.path.join(ref_dir, spec["file"])).read(), spec["inputs"], spec["outputs"],
                                 f"bitslicing/{spec['file']}")
            assert parsed.to_json() == c.to_json()


def test_emitter_espresso_hook_round_trip():
    from lifeapi_b200.bitslicing import emit_cuda as E
    c = E.from_espresso("toy", ["1-0 10", "-11 11", "0-- 01"], "#.phase 10", ["a", "b", "c"], ["x", "y"])
    assert c.inverted == [False, True] and len(c.terms) == 3
    net = E.synthesize(c)
    E.verify(c, net)
    text = "x |= a & (~c);\n{ uint64_t temp = b & c; x |= temp; y |= temp; }\ny |= (~a);\ny = ~y;\n"
    assert E.parse_hpp("toy", text, ["a", "b", "c"], ["x", "y"]).terms == c.terms


def test_port_cube_tables_vs_json(port):
    from lifeapi_b200.bitslicing import emit_cuda as E
    rng = np.random.default_rng(5)
    for which, (c, net) in enumerate(E.load_json()):
        words = rng.integers(0, 1 << 63, size=len(c.inputs), dtype=np.uint64) << np.uint64(1) | rng.integers(0, 2, size=len(c.inputs), dtype=np.uint64)
        got = port.circuit_eval(which, words)
        full = (1 << 64) - 1
        want = E.eval_net(net, [int(w) for w in words], full)
        assert [int(g) for g in got] == want, c.name


def test_port_stable_golden(port):
    g = golden("stable.npz")
    for which in range(8):
        out, flags = port.stable_pass_batch(g["records"], which)
        assert (out == g["passes"][which]).all(), which
        assert (flags == g["flags"][which]).all(), which
    assert (port.stable_vulnerable(g["vulnerable_in"]) == g["vulnerable"]).all()
    for v, key in ((0, "refined_m"), (1, "refined_old")):
        got = port.refined_step(g["refined_records"], g["refined_on"], g["refined_unknown"], v)
        assert (np.stack(got) == g[key]).all()


def test_port_unknown_step_sop_golden(port):
    g = golden("unknown_step.npz")
    for i, k in enumerate(g["gens"]):
        s, u = port.unknown_step_sop_batch(g["state"], g["unknown"], int(k))
        assert (s == g["next_state"][i]).all() and (u == g["next_unknown"][i]).all()


def test_port_vs_reference_stable_passes(port, ref):
    from stable_inputs import stable_records
    rng = np.random.default_rng(99)
    clean = stable_records(48, 31337)
    dirty = rng.integers(0, 1 << 63, size=(16, 10, 64), dtype=np.uint64) << np.uint64(1)
    recs = np.concatenate([clean, dirty])
    cur = recs
    for which in (1, 2, 3, 0, 4, 6, 7, 5, 5):  # chained: later passes see realistic intermediate records
        a, fa = port.stable_pass_batch(cur, which)
        b, fb = ref.stable_pass_batch(cur, which)
        assert (a == b).all() and (fa == fb).all(), which
        cur = b
    assert (port.stable_vulnerable(cur) == ref.stable_vulnerable(cur)).all()
    con = rng.integers(0, 1 << 63, size=(64, 64), dtype=np.uint64)
    for v in (0, 1):
        assert all((x == y).all() for x, y in zip(port.refined_step(cur, con, recs[:, 1], v), ref.refined_step(cur, con, recs[:, 1], v)))
    for s in clean[:8, 0]:
        assert (port.zoi_hollow(s) == ref.zoi_hollow(s)).all()


def test_port_unknown_step_sop_vs_reference_and_semantics(port, ref):
    from stable_inputs import three_valued_patterns
    st, un = three_valued_patterns(64, 77, port.soup(5000, 64, SEED, 0), port.zoi)
    a = port.unknown_step_sop_batch(st, un, 6)
    b = ref.unknown_step_batch(st, un, 6)
    c = port.unknown_step_batch(st, un, 6)  # per-cell semantic restatement (unknown_step.py:3-38)
    assert (a[0] == b[0]).all() and (a[1] == b[1]).all()
    assert (a[0] == c[0]).all() and (a[1] == c[1]).all()
    # overlapping state/unknown planes exercise the cover's don't-care rows: SOP port == reference text
    rng = np.random.default_rng(4)
    r = rng.integers(0, 1 << 63, size=(2, 32, 64), dtype=np.uint64)
    a, b = port.unknown_step_sop_batch(r[0], r[1], 3), ref.unknown_step_batch(r[0], r[1], 3)
    assert (a[0] == b[0]).all() and (a[1] == b[1]).all()


def test_weld_rule_luts_exhaustive():
    """weld_rule in csrc/next.cu (LifeWeld::Step in layout 8): two full adders over the column counts, then the 6-gate
    tail from exact synthesis.  Checked against LifeWeld.hpp:169-186 -- next = [T == 3] | cell & [T == 4] with
    T = (3x3 sum + frozen) mod 8 -- on every (left, mid, right) column count, cell and frozen value a record can produce."""
    import os
    import re
    from conftest import ROOT
    src = open(os.path.join(ROOT, "lifeapi_b200", "csrc", "next.cu")).read()
    body = src[src.index("uint32_t weld_rule("):src.index("weld_step_kernel_x(uint64_t* records")]
    assert re.findall(r"lop3<(\w+)>", body) == ["kXor3", "kMaj", "kXor3", "kMaj", "kXor3", "kMaj", "0xDA", "0x98", "kXor3", "0x48"]
    checked = 0
    for l in range(4):
        for m in range(4):
            for r in range(4):
                for cell in (0, 1):
                    if (cell and m == 0) or (not cell and m == 3):
                        continue
                    s0, k0 = _lop3(0x96, l & 1, m & 1, r & 1), _lop3(0xE8, l & 1, m & 1, r & 1)
                    s1, k1 = _lop3(0x96, l >> 1, m >> 1, r >> 1), _lop3(0xE8, l >> 1, m >> 1, r >> 1)
                    for f in range(8):
                        f0, f1, f2 = f & 1, (f >> 1) & 1, f >> 2
                        g0 = _lop3(0x96, k0, s1, f1)
                        g1 = _lop3(0xE8, k0, s1, f1)
                        g2 = _lop3(0xDA, s0, cell, f0)
                        g3 = _lop3(0x98, cell, g0, g2)
                        g4 = _lop3(0x96, k1, f2, g1)
                        t = (l + m + r + f) & 7
                        assert _lop3(0x48, g2, g3, g4) == int(t == 3 or (t == 4 and cell)), (l, m, r, cell, f)
                        checked += 1
    assert checked == 64 * 2 * 8 - 2 * 16 * 8


def test_pair_shared_rule_luts_exhaustive():
    """pair_code / pair_select in csrc/life_core.cuh: the code of (a, b) serves both columns of a pair.  The
    select relies on the count's don't-cares (the cell is the middle bit of its own column's triple: cell = 1
    implies count >= 1, count = 3 implies cell = 1), so it is checked on every input a column can produce,
    once with the cell belonging to column a (left use) and once to column b (right use)."""
    import os
    import re
    from conftest import ROOT
    src = open(os.path.join(ROOT, "lifeapi_b200", "csrc", "life_core.cuh")).read()
    body = src[src.index("PairCode pair_code("):src.index("template <int LPP>\n__device__ __forceinline__ Cnt halo")]
    body = body.replace("kMaj", "0xE8").replace("kXor3", "0x96")
    assert [int(m, 16) for m in re.findall(r"lop3<(0x\w+)>", body)] == [0xB2, 0xD4, 0xE8, 0xE8, 0x16, 0x96, 0xC8]
    checked = 0
    for ca in range(4):
        for cb in range(4):
            a0, a1, b0, b1 = ca & 1, ca >> 1, cb & 1, cb >> 1
            p0 = _lop3(0xB2, a1, b0, b1)
            p1 = _lop3(0xD4, a1, b1, p0)
            p2 = _lop3(0xE8, a0, b0, p0)
            for co in range(4):
                for cell in (0, 1):
                    for own in (ca, cb):
                        if (cell == 1 and own == 0) or (cell == 0 and own == 3):
                            continue  # no column has a live centre and an empty triple, or a full triple and a dead centre
                        t0 = _lop3(0xE8, co & 1, p0, p2)
                        t1 = _lop3(0x16, co >> 1, p1, t0)
                        t2 = _lop3(0x96, co & 1, a0, b0)
                        n = ca + cb + co
                        assert _lop3(0xC8, cell, t1, t2) == int(n == 3 or (cell and n == 4))
                        checked += 1
    assert checked == 192


def test_three_valued_fast_path_exhaustive():
    """csrc/life_three.cuh: two B3/S23 selects (on state and on state | unknown) + 3 combining LUTs must equal
    the reference's unknown_step.hpp cover on EVERY input with disjoint state / unknown planes.  All
    neighbourhoods are enumerated column-wise: the centre column as raw three-valued cells (27 triples), the
    two other columns as (count of ON, count of ON-or-UNKNOWN) pairs, once with the centre column as the left
    and once as the right member of the column pair.  Bit-parallel: one Python int bit per case."""
    import os
    import re
    from conftest import ROOT
    from lifeapi_b200.bitslicing import emit_cuda as E
    src = open(os.path.join(ROOT, "lifeapi_b200", "csrc", "life_three.cuh")).read()
    body = src[src.index("PairCode4 pair_code4("):src.index("template <int LPP, bool IL>")]
    body = body.replace("kMaj", "0xE8").replace("kOr3", "0xFE")
    luts = [int(m, 16) for m in re.findall(r"lop3<(0x\w+)>", body)]
    assert luts == [0x56, 0xBE, 0xE4, 0x92, 0xB2, 0x48, 0xE8, 0xFE, 0x46, 0x92, 0xE8, 0xE2, 0xFE, 0x16, 0x08, 0x47, 0xF8]
    code_l, left_l, right_l, comb_l = luts[0:4], luts[4:9], luts[9:14], luts[14:17]

    cases = []  # (own column cells [(s,u)]*3, other pair member (cs, ct), far column (cs, ct), own_is_left)
    others = [(cs, ct) for cs in range(4) for ct in range(cs, 4)]
    for cells in __import__("itertools").product(((0, 0), (1, 0), (0, 1)), repeat=3):
        for mate in others:
            for far in others:
                for own_is_left in (True, False):
                    cases.append((cells, mate, far, own_is_left))
    n = len(cases)
    assert n == 27 * 10 * 10 * 2
    full = (1 << n) - 1

    def col(fn):
        v = 0
        for i, cse in enumerate(cases):
            if fn(cse):
                v |= 1 << i
        return v

    def lop3(lut, a, b, c):
        r = 0
        for k in range(8):
            if lut >> k & 1:
                r |= (a if k & 4 else ~a) & (b if k & 2 else ~b) & (c if k & 1 else ~c)
        return r & full

    def plane_signals(own_cnt, mate_cnt, far_cnt, cell):
        """own_cnt / mate_cnt / far_cnt: case -> 0..3; returns (alive, A) columns of this plane."""
        # pair (a, b): a = left member.  own is a when own_is_left else b
        a = lambda c: own_cnt(c) if c[3] else mate_cnt(c)
        b = lambda c: mate_cnt(c) if c[3] else own_cnt(c)
        a0, a1 = col(lambda c: a(c) & 1), col(lambda c: a(c) >> 1)
        b0, b1 = col(lambda c: b(c) & 1), col(lambda c: b(c) >> 1)
        c0, c1 = col(lambda c: far_cnt(c) & 1), col(lambda c: far_cnt(c) >> 1)
        p0 = lop3(code_l[0], a0, b0, b1)
        p1 = lop3(code_l[1], a1, b1, p0)
        p2 = lop3(code_l[2], a1, b1, p0)
        p3 = lop3(code_l[3], a0, b0, p1)
        l0 = lop3(left_l[0], c1, cell, p2)
        l1 = lop3(left_l[1], c1, p1, p2)
        l2 = lop3(left_l[2], c0, p3, l0)
        la, lalive = lop3(left_l[3], p2, l1, l2), lop3(left_l[4], p2, l1, l2)
        r0 = lop3(right_l[0], c1, cell, p2)
        r1 = lop3(right_l[1], c0, p3, r0)
        r2 = lop3(right_l[2], c1, p1, p2)
        ra, ralive = lop3(right_l[3], p2, r1, r2), lop3(right_l[4], p2, r1, r2)
        is_left = col(lambda c: c[3])
        return (lalive & is_left) | (ralive & ~is_left & full), (la & is_left) | (ra & ~is_left & full)

    s_own = lambda c: sum(x[0] for x in c[0])
    t_own = lambda c: sum(x[0] | x[1] for x in c[0])
    cell_s, cell_u = col(lambda c: c[0][1][0]), col(lambda c: c[0][1][1])
    lo_alive, lo_a = plane_signals(s_own, lambda c: c[1][0], lambda c: c[2][0], cell_s)
    hi_alive, hi_a = plane_signals(t_own, lambda c: c[1][1], lambda c: c[2][1], cell_s | cell_u)
    next_on = lop3(comb_l[0], cell_u, lo_alive, hi_alive)
    x = lop3(comb_l[1], lo_a, lo_alive, hi_alive)
    next_unknown = lop3(comb_l[2], cell_u, hi_a, x)

    cover = dict((c.name, c) for c, _ in E.load_json())["unknown_step"]
    on = lambda c: s_own(c) + c[1][0] + c[2][0]
    unk = lambda c: (t_own(c) - s_own(c)) + (c[1][1] - c[1][0]) + (c[2][1] - c[2][0])
    named = {"current_unknown": cell_u, "current_on": cell_s}
    for k in range(4):
        named[f"on{k}"] = col(lambda c, k=k: on(c) >> k & 1)
        named[f"unk{k}"] = col(lambda c, k=k: unk(c) >> k & 1)
    want = dict(zip(cover.outputs, E.eval_sop(cover, [named[i] for i in cover.inputs], full)))
    assert next_on == want["naive_next_on"]
    assert next_unknown == want["naive_next_unknown"]
    assert next_on & next_unknown == 0 and next_on != 0 and next_unknown != 0


# ---- 8f next rows ---------------------------------------------------------------------------------
def test_port_next_rows_golden(port):
    g = golden("next_rows.npz")
    for i, k in enumerate(g["weld_gens"]):
        assert (port.weld_step(g["weld"], int(k)) == g["weld_out"][i]).all()
    assert (port.interaction_counts(g["inter_in"]) == g["inter_out"]).all()
    assert [port.hash(x) for x in g["hash_in"]] == [int(h) for h in g["hash_out"]]
    # XXH3 restatement against the survey's known answers (R-pentomino generations)
    rp = golden("rpentomino.npz")
    assert [port.hash(s) for s in rp["states"]] == [int(h) for h in rp["xxh3"]]
    assert port.hash(golden("soups.npz")["mt1_final"]) == 0x77B09695A0065A59


def test_port_vs_reference_next_rows(port, ref):
    rng = np.random.default_rng(2)
    s = port.soup(123, 40, SEED, 1)
    w = np.zeros((40, 4, 64), dtype=np.uint64)
    w[:, 0] = s
    w[:, 1:] = rng.integers(0, 1 << 63, size=(40, 3, 64), dtype=np.uint64) << np.uint64(1)  # dense, arbitrary frozen planes
    assert (port.weld_step(w, 4) == ref.weld_step(w, 4)).all()
    assert (port.interaction_counts(s) == ref.interaction_counts(s)).all()
    assert all(port.hash(x) == ref.hash(x) for x in s)


# ---- Mirrored / Convolve / Match / InteractionOffsets / RLE ------------------
def test_port_locate_golden(port):
    g = golden("locate.npz")
    a, b = g["a"], g["b"]
    for i in range(len(a)):
        assert (port.mirrored(a[i]) == g["mirrored"][i]).all()
        assert (port.convolve(a[i], b[i]) == g["convolve"][i]).all()
        assert (port.convolve(b[i], a[i]) == g["convolve"][i]).all()  # the Minkowski sum commutes
        assert (port.convolve(a[0], b[i]) == g["convolve_bcast"][i]).all()
        assert (port.interaction_offsets(a[i], b[i]) == g["offsets"][i]).all()
        assert (port.match_live_and_dead(a[i], b[i], port.mirrored(b[i])) == g["match_on_ab"][i]).all()
    for i in range(len(g["states"])):
        s, l, d = g["states"][i], g["live"][i], g["dead"][i]
        assert (port.match_live(s, l) == g["match_live"][i]).all()
        assert (port.match_live_and_dead(s, l, d) == g["match_live_and_dead"][i]).all()
        assert (port.match_state(s, l) == g["match_state"][i]).all()
        assert (port.match_target(s, g["targets"][i]) == g["match_target"][i]).all()
        assert port.pop(g["match_state"][i]) >= 1  # the catalyst pasted into the soup is found
    for s, text in zip(g["rle_in"], g["rle"]):
        assert port.rle(s) == str(text)
        assert (port.parse(str(text)) == port.moved(s, 32, 32)).all()  # RLE is written from cell (32, 32): a half-torus shift
    for h, text in zip(g["history_in"], g["history_rle"]):
        assert port.history_rle(h) == str(text)


def test_port_match_is_the_definition(port):
    """Match by brute force: offset (x, y) is ON iff every live cell lands on an ON cell and every dead cell
    on an OFF cell (what MatchLiveAndDead's two convolutions compute, LifeAPI.hpp:432-435)."""
    g = golden("locate.npz")
    s, l, d = g["states"][1], g["live"][1], g["dead"][1]

    def cells(p):
        return [(x, y) for x in range(64) for y in range(64) if (int(p[x]) >> y) & 1]
    lc, dc = cells(l), cells(d)
    got = port.match_live_and_dead(s, l, d)
    for x in range(0, 64, 5):
        for y in range(64):
            want = all((int(s[(x + i) & 63]) >> ((y + j) & 63)) & 1 for i, j in lc) and \
                not any((int(s[(x + i) & 63]) >> ((y + j) & 63)) & 1 for i, j in dc)
            assert bool((int(got[x]) >> y) & 1) == want


def test_port_vs_reference_locate(port, ref):
    rng = np.random.default_rng(4242)
    for trial in range(24):
        k = 1 + trial % 7
        a = rng.integers(0, 1 << 63, size=(k, 64), dtype=np.uint64)
        b = rng.integers(0, 1 << 63, size=(8 - k, 64), dtype=np.uint64)
        a = np.bitwise_and.reduce(a, axis=0) << np.uint64(trial % 2)
        b = np.bitwise_and.reduce(b, axis=0)
        assert (port.mirrored(a) == ref.mirrored(a)).all()
        assert (port.convolve(a, b) == ref.convolve(a, b)).all()
        assert (port.interaction_offsets(a, b) == ref.interaction_offsets(a, b)).all()
        assert (port.match_live(a, b) == ref.match_live(a, b)).all()
        assert (port.match_state(a, b) == ref.match_state(a, b)).all()
        assert (port.match_live_and_dead(a, b, port.mirrored(a)) == ref.match_live_and_dead(a, b, port.mirrored(a))).all()
        assert port.rle(a) == ref.rle(a)
        h = np.stack([a, b, a & b, port.mirrored(b)])
        assert port.history_rle(h) == ref.history_rle(h)


def test_library_rle_and_parse_golden(port):
    """The product's host-side RLE writer / parser (lifeb200_rle ...; no GPU involved) against the reference's strings."""
    from lifeapi_b200 import history_parse_rle, history_rle, parse_rle, rle
    g = golden("locate.npz")
    for s, text in zip(g["rle_in"], g["rle"]):
        assert rle(s) == str(text)
        assert (parse_rle(str(text)) == port.moved(s, 32, 32)).all()
    for h, text in zip(g["history_in"], g["history_rle"]):
        assert history_rle(h) == str(text)
        back = history_parse_rle(str(text))  # 'F' cells (combinations StateToChar has no letter for) do not come back
        assert not (back[0] & ~h[0]).any()  # every parsed state cell is a state cell of the record
    assert rle(np.zeros(64, dtype=np.uint64)) == "!"
    header = "x = 3, y = 3, rule = B3/S23\nbo$2bo$\n3o!\n"
    assert (parse_rle(header) == port.parse(header)).all() and port.pop(parse_rle(header)) == 5
    cat = golden("collisions.npz")
    for c in cat["catalysts"]:
        assert (parse_rle(rle(c)) == port.moved(c, 32, 32)).all()


def test_library_rle_vs_reference_random(ref):
    from lifeapi_b200 import history_parse_rle, history_rle, parse_rle, rle
    rng = np.random.default_rng(99)
    for trial in range(40):
        planes = rng.integers(0, 1 << 63, size=(4, 1 + trial % 5, 64), dtype=np.uint64)
        planes = np.bitwise_and.reduce(planes, axis=1) << np.uint64(trial % 2)
        assert rle(planes[0]) == ref.rle(planes[0])
        text = ref.history_rle(planes)
        assert history_rle(planes) == text
        assert (history_parse_rle(text) == ref.history_parse(text)).all()
        assert (parse_rle(ref.rle(planes[1])) == ref.moved(planes[1], 32, 32)).all()
    messy = "#C comment\r\nx = 3\r\n3o 2b o$$o!ignored"
    assert (parse_rle(messy) == ref.parse(messy)).all()
