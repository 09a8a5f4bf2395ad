"""Generate tests/golden/*.npz from the UNMODIFIED reference compiled in this container.

    python tests/golden/make_golden.py        (needs /root/reference; run `make -C oracle ref` first)

Everything here is produced by calls into oracle/_ref (reference member functions only); the
script itself contains no Life logic apart from the seeded input generators.  The committed .npz
files are what travels to the GPU box, where /root/reference does not exist.
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle.loader import Port, Ref, build_ref  # noqa: E402
from stable_inputs import stable_records  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
SEED = 0x4C696665415049

CATALYSTS = {  # parsed with LifeState::Parse (not ConstantParse: SURVEY.md 0.1)
    "eater1": "2b2o$bobo$bo$2o!",
    "block": "2o$2o!",
    "boat": "2o$obo$bo!",
    "tub": "bo$obo$bo!",
    "beehive": "b2o$o2bo$b2o!",
    "loaf": "b2o$o2bo$bobo$2bo!",
    "ship": "2o$obo$b2o!",
    "pond": "b2o$o2bo$o2bo$b2o!",
}
GLIDER = "bo$2bo$3o!"


def mt19937_64(seed: int, count: int) -> np.ndarray:
    """std::mt19937_64 (for the survey's mt19937_64(1) known-answer soup)."""
    nn, mm = 312, 156
    mask = (1 << 64) - 1
    mt = [0] * nn
    mt[0] = seed & mask
    for i in range(1, nn):
        mt[i] = (6364136223846793005 * (mt[i - 1] ^ (mt[i - 1] >> 62)) + i) & mask
    out = []
    idx = nn
    for _ in range(count):
        if idx >= nn:
            for i in range(nn):
                x = (mt[i] & 0xFFFFFFFF80000000) | (mt[(i + 1) % nn] & 0x7FFFFFFF)
                xa = x >> 1
                if x & 1:
                    xa ^= 0xB5026F5AA96619E9
                mt[i] = mt[(i + mm) % nn] ^ xa
            idx = 0
        y = mt[idx]
        idx += 1
        y ^= (y >> 29) & 0x5555555555555555
        y ^= (y << 17) & 0x71D67FFFEDA60000
        y ^= (y << 37) & 0xFFF7EEE000000000
        y ^= y >> 43
        out.append(y & mask)
    return np.array(out, dtype=np.uint64)


def glider_variants(ref: Ref) -> np.ndarray:
    """8 orientations (the D4 images of the 3x3 glider box) x 4 phases (reference Step)."""
    cells = [(1, 0), (2, 1), (0, 2), (1, 2), (2, 2)]  # bo$2bo$3o!
    maps = [lambda x, y: (x, y), lambda x, y: (2 - x, y), lambda x, y: (x, 2 - y), lambda x, y: (2 - x, 2 - y),
            lambda x, y: (y, x), lambda x, y: (2 - y, x), lambda x, y: (y, 2 - x), lambda x, y: (2 - y, 2 - x)]
    out = []
    for m in maps:
        s = np.zeros(64, dtype=np.uint64)
        for (x, y) in cells:
            tx, ty = m(x, y)
            s[8 + tx] |= np.uint64(1) << np.uint64(8 + ty)  # LifeState::Set(x, y): state[x] |= 1 << y
        for _ in range(4):
            out.append(s.copy())
            s = ref.step(s)
    return np.stack(out)


def unknown_free_noise(rng, n: int) -> np.ndarray:
    """A few random live cells per pattern (perturbs the still-life field so live != stable counts)."""
    out = np.zeros((n, 64), dtype=np.uint64)
    for p in range(n):
        for _ in range(6):
            out[p, int(rng.integers(64))] |= np.uint64(1) << np.uint64(int(rng.integers(64)))
    return out


def density_planes(rng, n: int, density_log2: int) -> np.ndarray:
    """n random planes with cell density 2^-density_log2 (AND of independent random words)."""
    out = rng.integers(0, 1 << 63, size=(n, 64), dtype=np.uint64) << np.uint64(1) | rng.integers(0, 2, size=(n, 64), dtype=np.uint64)
    for _ in range(density_log2 - 1):
        out &= rng.integers(0, 1 << 63, size=(n, 64), dtype=np.uint64) << np.uint64(1) | rng.integers(0, 2, size=(n, 64), dtype=np.uint64)
    return out


def locate_inputs(ref: Ref, port: Port):
    """Operand pairs for Convolve / Match / InteractionOffsets: random planes of several densities, full and
    empty planes, runs that wrap around bit 63 -> 0, and real objects (catalysts, gliders) placed on soups."""
    rng = np.random.default_rng(777)
    cats = np.stack([ref.parse(r, 28, 28) for r in CATALYSTS.values()])
    gl = glider_variants(ref)
    a = [density_planes(rng, 4, 6), density_planes(rng, 4, 1), density_planes(rng, 3, 3), cats[:4]]
    b = [density_planes(rng, 4, 7), density_planes(rng, 4, 5), density_planes(rng, 3, 1), gl[[0, 5, 18, 31]]]
    a, b = np.concatenate(a), np.concatenate(b)
    full, empty = ~np.zeros((1, 64), dtype=np.uint64), np.zeros((1, 64), dtype=np.uint64)
    wrap = np.zeros((1, 64), dtype=np.uint64)
    wrap[0, 0] = np.uint64(0xF00000000000000F)   # a run of ON bits through the 63 -> 0 seam
    wrap[0, 63] = np.uint64(0x8000000000000001)
    wrap[0, 31] = np.uint64(0x00FFFFFFFFFFFF00)  # a run longer than 32 (the reference caps runs at 32)
    a = np.concatenate([a, full, empty, wrap, full, wrap])
    b = np.concatenate([b, gl[:1], cats[:1], cats[1:2], full, wrap])
    # states that really contain the objects looked for: a soup with catalysts pasted in at known offsets
    states = density_planes(rng, 8, 3)
    live = cats[:8].copy()
    for i in range(8):
        dx, dy = int(rng.integers(64)), int(rng.integers(64))
        zone = ref.zoi(ref.moved(cats[i], dx, dy))
        states[i] = (states[i] & ~zone) | ref.moved(cats[i], dx, dy)
    return a, b, states, live


def make_locate(ref: Ref, port: Port) -> None:
    a, b, states, live = locate_inputs(ref, port)
    rng = np.random.default_rng(778)
    dead = np.stack([ref.zoi(x) & ~x for x in live])
    dead[::2] &= density_planes(rng, 4, 1)  # partial dead masks as well as the full boundary
    targets = np.stack([np.concatenate([l, d]) for l, d in zip(live, dead)])
    hist = np.stack([np.stack([states[i], states[i] | a[i], density_planes(rng, 1, 4)[0], density_planes(rng, 1, 6)[0]])
                     for i in range(4)])
    rle_in = np.concatenate([a[[0, 4, 11, 15, 16, 17]], states[:2]])
    np.savez_compressed(
        os.path.join(OUT, "locate.npz"), a=a, b=b,
        mirrored=np.stack([ref.mirrored(x) for x in a]),
        convolve=np.stack([ref.convolve(x, y) for x, y in zip(a, b)]),
        convolve_bcast=np.stack([ref.convolve(a[0], y) for y in b]),
        offsets=np.stack([ref.interaction_offsets(x, y) for x, y in zip(a, b)]),
        offsets_bcast=np.stack([ref.interaction_offsets(a[11], y) for y in b]),
        states=states, live=live, dead=dead, targets=targets,
        match_live=np.stack([ref.match_live(s, l) for s, l in zip(states, live)]),
        match_live_and_dead=np.stack([ref.match_live_and_dead(s, l, d) for s, l, d in zip(states, live, dead)]),
        match_state=np.stack([ref.match_state(s, l) for s, l in zip(states, live)]),
        match_target=np.stack([ref.match_target(s, t) for s, t in zip(states, targets)]),
        match_state_bcast=np.stack([ref.match_state(s, live[0]) for s in states]),
        match_on_ab=np.stack([ref.match_live_and_dead(x, y, ref.mirrored(y)) for x, y in zip(a, b)]),
        rle_in=rle_in, rle=np.array([ref.rle(x) for x in rle_in]),
        history_in=hist, history_rle=np.array([ref.history_rle(h) for h in hist]))
    found = [int(ref.pop(m)) for m in np.stack([ref.match_state(s, l) for s, l in zip(states, live)])]
    print("locate: Match(live) offsets found per state:", found)


def main() -> None:
    build_ref()
    ref = Ref.load()
    assert ref is not None, "oracle/_ref missing"
    port = Port()
    if sys.argv[1:] == ["locate"]:  # only the 8f rank-4 vectors (the others are unchanged)
        make_locate(ref, port)
        return

    # ---- config 1: R-pentomino known answers -------------------------------------------------
    rp = ref.parse("b2o$2o$bo!")
    gens = [0, 1, 2, 3, 10, 64, 100, 256, 500, 1000, 1024]
    states, pops, hashes = [], [], []
    cur, g = rp.copy(), 0
    for target in gens:
        while g < target:
            cur = ref.step(cur)
            g += 1
        states.append(cur.copy())
        pops.append(ref.pop(cur))
        hashes.append(ref.hash(cur))
    np.savez_compressed(os.path.join(OUT, "rpentomino.npz"), gens=np.array(gens), states=np.stack(states),
                        pops=np.array(pops, dtype=np.uint32), xxh3=np.array(hashes, dtype=np.uint64))

    # ---- survey's mt19937_64(1) soup --------------------------------------------------------
    mt1 = mt19937_64(1, 64)
    mt1_final = ref.step_batch(mt1[None, :], 1024, 1)[0]
    # ---- seeded soups (both distributions), several generation counts ------------------------
    soups0 = port.soup(0, 16, SEED, 0)
    soups1 = port.soup(0, 16, SEED, 1)
    gsel = [1, 2, 3, 64, 257, 1024]
    st0 = np.stack([ref.step_batch(soups0, k, 1) for k in gsel])
    st1 = np.stack([ref.step_batch(soups1, k, 1) for k in gsel])
    # XOR of GetHash over 8192 soups of one mt19937_64(42) stream after 1024 gens
    mt42 = mt19937_64(42, 8192 * 64).reshape(8192, 64)
    fin42 = ref.step_batch(mt42, 1024, 0)
    x = 0
    for p in range(8192):
        x ^= ref.hash(fin42[p])
    np.savez_compressed(os.path.join(OUT, "soups.npz"), seed=np.uint64(SEED), gens=np.array(gsel), soups0=soups0, soups1=soups1,
                        stepped0=st0, stepped1=st1, mt1=mt1, mt1_final=mt1_final, mt1_pop0=np.uint32(ref.pop(mt1)),
                        mt1_pop=np.uint32(ref.pop(mt1_final)), mt1_xxh3=np.uint64(ref.hash(mt1_final)),
                        mt42_xor_xxh3=np.uint64(x), mt42_first8=mt42[:8], mt42_first8_final=fin42[:8])

    # ---- NeighbourCount planes / ZOI / Moved / LifeTarget --------------------------------------
    nc = np.stack([ref.neighbour_count(s) for s in soups0[:8]])
    nc_alt = np.stack([ref.neighbour_count(s, "ref_count_neighbourhood") for s in soups0[:8]])
    assert (nc == nc_alt).all()
    zoi = np.stack([ref.zoi(s) for s in soups0[:4]] + [ref.zoi(rp)])
    moves = [(0, 0), (1, 0), (0, 1), (-3, 5), (63, 63), (17, -20), (64, 64), (-65, 130)]
    moved = np.stack([ref.moved(soups0[0], x_, y_) for x_, y_ in moves])
    np.savez_compressed(os.path.join(OUT, "planes.npz"), soups=soups0[:8], ncount=nc, zoi_in=np.vstack([soups0[:4], rp[None]]),
                        zoi=zoi, moves=np.array(moves), moved=moved)

    # ---- fused step+match and collision enumeration -------------------------------------------
    cats = np.stack([ref.parse(r, 28, 28) for r in CATALYSTS.values()])
    targets = np.stack([ref.target_from_state(c) for c in cats])
    gl = glider_variants(ref)
    # enumeration over 2 catalysts x 32 glider variants x 4096 offsets, 128 generations
    sub_cats = cats[:2]
    ids, ff, rec, skipped = ref.enumerate(sub_cats, gl, 0, 2 * 32 * 4096, 128, 8, 1 << 18)
    np.savez_compressed(os.path.join(OUT, "collisions.npz"), catalyst_names=np.array(list(CATALYSTS.keys())), catalysts=cats,
                        targets=targets, gliders=gl, enum_n_cat=np.uint32(2), enum_gens=np.uint32(128),
                        enum_interval=np.uint32(8), enum_ids=ids, enum_first_false=ff, enum_recovered=rec,
                        enum_skipped=np.uint64(skipped))
    print(f"collisions: {len(ids)} hits, {skipped} skipped of {2 * 32 * 4096}")

    # ---- three-valued step through bitslicing/unknown_step.hpp ---------------------------------
    rng = np.random.default_rng(12345)
    n3 = 16
    base = port.soup(1000, n3, SEED, 0)
    box = np.zeros(64, dtype=np.uint64)
    box[16:48] = np.uint64(0x0000FFFFFFFF0000)
    state = base & box
    unk = np.zeros_like(state)
    for p in range(n3):
        z = ref.zoi(state[p])
        r = rng.integers(0, 1 << 63, size=(3, 64), dtype=np.uint64)
        unk[p] = r[0] & r[1] & r[2] & z  # density 1/8 inside the ZOI
    state &= ~unk
    g3 = [1, 2, 16]
    res = [ref.unknown_step_batch(state, unk, k, 1) for k in g3]
    np.savez_compressed(os.path.join(OUT, "unknown_step.npz"), state=state, unknown=unk, gens=np.array(g3),
                        next_state=np.stack([r[0] for r in res]), next_unknown=np.stack([r[1] for r in res]))

    # ---- LifeStable passes, Vulnerable, refined steps (reference LifeStable.hpp members) -------
    clean = stable_records(10, 2024)
    dirty = rng.integers(0, 1 << 63, size=(6, 10, 64), dtype=np.uint64) << np.uint64(1)
    dirty[3:, 2:] &= rng.integers(0, 1 << 63, size=(3, 8, 64), dtype=np.uint64) & rng.integers(0, 1 << 63, size=(3, 8, 64), dtype=np.uint64)
    dirty[3:, 0] &= ~dirty[3:, 1]
    recs = np.concatenate([clean, dirty])
    outs, flags = [], []
    for which in range(8):
        o, f = ref.stable_pass_batch(recs, which, 1)
        outs.append(o)
        flags.append(f)
    propagated, _ = ref.stable_pass_batch(recs, 5, 1)
    vuln = ref.stable_vulnerable(np.concatenate([recs, propagated]))
    cur_on = np.concatenate([clean[:, 0], rng.integers(0, 1 << 63, size=(6, 64), dtype=np.uint64)])
    cur_on[:10] = ref.step_batch(cur_on[:10] | unknown_free_noise(rng, 10), 1, 1)
    cur_unk = recs[:, 1].copy()
    refined = [np.stack(ref.refined_step(propagated, cur_on, cur_unk, v)) for v in (0, 1)]
    np.savez_compressed(os.path.join(OUT, "stable.npz"), records=recs, passes=np.stack(outs), flags=np.stack(flags),
                        vulnerable_in=np.concatenate([recs, propagated]), vulnerable=vuln, refined_records=propagated,
                        refined_on=cur_on, refined_unknown=cur_unk, refined_m=refined[0], refined_old=refined[1])
    print("stable flags per pass:", [np.bincount(f, minlength=4).tolist() for f in flags])

    # ---- 8f next rows: LifeWeld::Step, InteractionCountsAndNext, GetHash --------------------------
    wsoups = port.soup(7000, 12, SEED, 0)
    weld = np.zeros((12, 4, 64), dtype=np.uint64)
    weld[:, 0] = wsoups
    fr = rng.integers(0, 1 << 63, size=(3, 12, 3, 64), dtype=np.uint64)
    weld[:8, 1:] = fr[0, :8] & fr[1, :8] & fr[2, :8] & ~wsoups[:8, None, :]  # sparse frozen counts on dead cells
    wg = [1, 2, 9]
    np.savez_compressed(os.path.join(OUT, "next_rows.npz"), weld=weld, weld_gens=np.array(wg),
                        weld_out=np.stack([ref.weld_step(weld, k) for k in wg]),
                        inter_in=wsoups, inter_out=ref.interaction_counts(wsoups),
                        hash_in=np.concatenate([wsoups, np.zeros((1, 64), dtype=np.uint64), ~np.zeros((1, 64), dtype=np.uint64)]),
                        hash_out=np.array([ref.hash(x) for x in np.concatenate(
                            [wsoups, np.zeros((1, 64), dtype=np.uint64), ~np.zeros((1, 64), dtype=np.uint64)])], dtype=np.uint64))

    # ---- 8f rank 4: Mirrored / Convolve / Match / InteractionOffsets / RLE -----------------------
    make_locate(ref, port)

    for f in sorted(os.listdir(OUT)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(OUT, f)), "bytes")


if __name__ == "__main__":
    main()
