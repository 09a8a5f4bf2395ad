"""Side measurements printed in bench.py's `other_configs` (N = 1 only): BASELINE configs 3, 4, 5 and
the LifeStable Propagate pass.  Not bench lines of their own: the headline stays config 2.  Parity
of every kernel used here is the job of tests/ (-m gpu); the CPU numbers beside them come from the
reference compiled into oracle/_ref (bench.py's cpu_baseline leg), on bounded samples."""
from __future__ import annotations

import time

import numpy as np

SEED = 0x4C696665415049

STILL_LIFES = {  # (x, y) cells
    "block": [(0, 0), (1, 0), (0, 1), (1, 1)],
    "tub": [(1, 0), (0, 1), (2, 1), (1, 2)],
    "boat": [(0, 0), (1, 0), (0, 1), (2, 1), (1, 2)],
    "ship": [(0, 0), (1, 0), (0, 1), (2, 1), (1, 2), (2, 2)],
    "beehive": [(1, 0), (2, 0), (0, 1), (3, 1), (1, 2), (2, 2)],
    "loaf": [(1, 0), (2, 0), (0, 1), (3, 1), (1, 2), (3, 2), (2, 3)],
    "pond": [(1, 0), (2, 0), (0, 1), (3, 1), (0, 2), (3, 2), (1, 3), (2, 3)],
    "eater1": [(0, 0), (1, 0), (0, 1), (2, 1), (2, 2), (2, 3), (3, 3)],
}
GLIDER = [(1, 0), (2, 1), (0, 2), (1, 2), (2, 2)]
D4 = [lambda x, y: (x, y), lambda x, y: (3 - x, y), lambda x, y: (x, 3 - y), lambda x, y: (3 - x, 3 - y),
      lambda x, y: (y, x), lambda x, y: (3 - y, x), lambda x, y: (y, 3 - x), lambda x, y: (3 - y, 3 - x)]


def _pattern(cells, ox, oy, tf=None):
    s = np.zeros(64, dtype=np.uint64)
    for (x, y) in cells:
        if tf is not None:
            x, y = tf(x, y)
        s[(ox + x) & 63] |= np.uint64(1) << np.uint64((oy + y) & 63)
    return s


def _time(eng, fn, reps=3):
    fn()
    eng.sync()
    eng.timer_start()
    for _ in range(reps):
        fn()
    return eng.timer_stop() / reps


def _cpu_ref():
    try:
        from oracle.loader import Ref
        return Ref.load()
    except Exception:  # noqa: BLE001
        return None


def config3_unknown_step(eng, p_int):
    """256 Ki partial patterns x 16 three-valued generations (bitslicing/unknown_step.hpp)."""
    n, gens = 1 << 18, 16
    d_s, d_u, d_t, d_o = eng.malloc(n * 512), eng.malloc(n * 512), eng.malloc(n * 512), eng.malloc(n * 512)
    try:
        # state = soup masked to the centred 32x32 box; unknown = AND of three other soups (density 1/8) in the box
        eng.soup_dev(d_s, 0, n, SEED, 0)
        host = np.empty((n, 64), dtype=np.uint64)
        eng.d2h(host, d_s)
        eng.sync()
        box = np.zeros(64, dtype=np.uint64)
        box[16:48] = np.uint64(0x0000FFFFFFFF0000)
        unk = (np.roll(host, 1, axis=0) & np.roll(host, 2, axis=0) & np.roll(host, 3, axis=0)) & box
        st = (host & box) & ~unk
        eng.h2d(d_s, st)
        eng.h2d(d_u, unk)
        eng.sync()
        ms = _time(eng, lambda: eng.unknown_step_batch_dev(d_s, d_u, d_t, d_o, n, gens))
        eng.set_three_valued_path(1)
        ms_lit = _time(eng, lambda: eng.unknown_step_batch_dev(d_s, d_u, d_t, d_o, n, gens))
        eng.set_three_valued_path(0)
        long_gens = 256
        ms_long = _time(eng, lambda: eng.unknown_step_batch_dev(d_s, d_u, d_t, d_o, n, long_gens))
        d_o2 = d_o
        cells = n * 4096.0 * gens
        out = {"workload": f"config3: {n} partial patterns x {gens} three-valued generations", "ms": ms,
               "value": cells / (ms * 1e-3) / 1e9, "unit": "Gcell/s",
               "int_pipe_frac": cells * (40.0 / 32.0) / (ms * 1e-3) / p_int,
               "algorithmic": "40 int-pipe instr per 32 cells per generation: two NeighbourCounts + the 18-LUT cover (DESIGN.md 3.3)",
               "executed": "26 per 32 cells per generation (22 LOP3 + 4 SHF): two B3/S23 selects on state and state|unknown, "
                           "csrc/life_three.cuh; planes are disjoint, so every warp takes that path",
               "pipe_utilisation": cells * (26.0 / 32.0) / (ms * 1e-3) / p_int,
               "literal_cover": {"ms": ms_lit, "value": cells / (ms_lit * 1e-3) / 1e9, "unit": "Gcell/s",
                                 "what": "same launch forced through the fragment text (lifeb200_set_three_valued_path 1)"},
               "long_run": {"generations": long_gens, "ms": ms_long, "value": n * 4096.0 * long_gens / (ms_long * 1e-3) / 1e9,
                            "unit": "Gcell/s", "executed": "24 per 32 cells per generation (row-interleaved registers)",
                            "pipe_utilisation": n * 4096.0 * long_gens * (24.0 / 32.0) / (ms_long * 1e-3) / p_int}}
        ref = _cpu_ref()
        if ref is not None:
            t0 = time.perf_counter()
            want_s, want_u = ref.unknown_step_batch(st, unk, gens, 0)
            dt = time.perf_counter() - t0
            out["cpu_baseline"] = {"value": n * 4096.0 * gens / dt / 1e9, "unit": "Gcell/s", "cores": ref.threads(),
                                   "kind": "reference", "sample": f"all {n} patterns x {gens} gens, {dt:.2f} s, {ref.threads()} threads"}
            eng.unknown_step_batch_dev(d_s, d_u, d_t, d_o2, n, gens)
            got_s, got_u = np.empty_like(st), np.empty_like(unk)
            eng.d2h(got_s, d_t)
            eng.d2h(got_u, d_o2)
            eng.sync()
            out["parity"] = {"checked": int(n), "equal": bool((got_s == want_s).all() and (got_u == want_u).all()), "vs": "reference",
                             "what": "both planes of every pattern against NeighbourCount + the committed unknown_step.hpp text"}
        return out
    finally:
        for p in (d_s, d_u, d_t, d_o):
            eng.free(p)


def config4_enumeration(eng, p_int):
    """32 catalysts x 32 glider variants x 4096 offsets = 4 Mi placements, <= 256 gens, fused Contains + hit compaction."""
    names = list(STILL_LIFES)
    cats = np.stack([_pattern(STILL_LIFES[names[i % 8]], 28, 28, D4[(i // 8) * 2]) for i in range(32)])
    g0 = np.stack([_pattern(GLIDER, 8, 8, tf) for tf in D4])
    phases = [g0]
    for _ in range(3):
        phases.append(eng.step_batch_host(phases[-1], 1))  # glider phases from the product path itself
    gliders = np.ascontiguousarray(np.stack(phases, axis=1).reshape(32, 64))
    total = 32 * 32 * 4096
    gens, interval = 256, 8
    t0 = time.perf_counter()
    hits, skipped = eng.enumerate_collisions(cats, gliders, 0, total, gens, interval, 1 << 20)
    dt = time.perf_counter() - t0
    t0 = time.perf_counter()
    hits, skipped = eng.enumerate_collisions(cats, gliders, 0, total, gens, interval, 1 << 20)
    dt = min(dt, time.perf_counter() - t0)
    out = {"workload": f"config4: {total} placements (32 catalysts x 32 glider variants x 4096 offsets), <= {gens} gens, "
                       f"stable_interval {interval}", "s": dt, "placements_per_s": total / dt, "hits": int(len(hits)),
           "skipped_overlapping": int(skipped), "timing": "wall clock around the host-facing C-ABI call (tables up, hits down, sorted)"}
    ref = _cpu_ref()
    if ref is not None:
        # 1/16 of the id space, spread over all 32 catalysts: for every catalyst the 2 glider variants c % 16 and
        # (c + 8) % 16 + 16, all 4096 offsets of each, the reference's own loop on every host thread
        per_cat = 32 * 4096
        ranges = []
        for c in range(32):
            for g in (c % 16, (c + 8) % 16 + 16):
                lo = c * per_cat + g * 4096
                ranges.append((lo, lo + 4096))
        checked, equal, n_ref_hits, dt_cpu = 0, True, 0, 0.0
        for lo, hi in ranges:
            t0 = time.perf_counter()
            ids, ff, rec, sk = ref.enumerate_mt(cats, gliders, lo, hi, gens, interval, hi - lo, 0)
            dt_cpu += time.perf_counter() - t0
            sub = hits[(hits["id"] >= lo) & (hits["id"] < hi)]
            equal = equal and len(sub) == len(ids) and bool((sub["id"] == ids).all() and (sub["g_first_false"] == ff).all()
                                                             and (sub["g_recovered"] == rec).all())
            checked += hi - lo
            n_ref_hits += len(ids)
        out["cpu_baseline"] = {"placements_per_s": checked / dt_cpu, "cores": ref.threads(), "kind": "reference",
                               "sample": f"{checked} placement ids = 1/16 of the space, 2 glider variants of every one of the 32 "
                                         f"catalysts, {dt_cpu:.2f} s on {ref.threads()} threads"}
        out["parity"] = {"checked": int(checked), "equal": bool(equal), "vs": "reference", "reference_hits": int(n_ref_hits),
                         "what": "hit records (id, g_first_false, g_recovered) of the sampled id ranges against the reference enumeration"}
    return out


def step_match_explicit(eng, p_int):
    """The explicit-pattern API (lifeb200_step_match_dev): 256 Ki placements built on the host as catalyst | glider.Moved,
    one LifeTarget per catalyst, 256 generations; 4 patterns per lane group (step_match_kernel_x) against the
    lane-group kernel (layout 2 pinned)."""
    names = list(STILL_LIFES)
    cats = np.stack([_pattern(STILL_LIFES[names[i % 8]], 28, 28, D4[(i // 8) * 2]) for i in range(32)])
    n, gens, interval = 1 << 18, 256, 8
    rng = np.random.default_rng(4)
    c_idx = rng.integers(0, 32, size=n)
    gl = np.stack([_pattern(GLIDER, int(x), int(y), D4[int(t)]) for x, y, t in zip(rng.integers(0, 64, 512), rng.integers(0, 64, 512), rng.integers(0, 8, 512))])
    pats = cats[c_idx] | gl[rng.integers(0, 512, size=n)]
    tg = np.zeros((32, 2, 64), dtype=np.uint64)
    for c in range(32):
        z = np.zeros(64, dtype=np.uint64)
        for dx in (-1, 0, 1):
            col = np.roll(cats[c], dx)
            z |= col | (col << np.uint64(1)) | (col >> np.uint64(63)) | (col >> np.uint64(1)) | (col << np.uint64(63))
        tg[c, 0], tg[c, 1] = cats[c], z & ~cats[c]
    d_p, d_t, d_i, d_h, d_n = eng.malloc(pats.nbytes), eng.malloc(tg.nbytes), eng.malloc(n * 4), eng.malloc(n * 16), eng.malloc(8)
    try:
        eng.h2d(d_p, np.ascontiguousarray(pats))
        eng.h2d(d_t, tg)
        eng.h2d(d_i, c_idx.astype(np.uint32))
        eng.sync()
        out = {"workload": f"step_match_dev: {n} explicit placements x <= {gens} gens, one LifeTarget per catalyst, stable_interval {interval}"}
        res = {}
        for name, layout in (("four_per_group", 0), ("lane_group_layout2", 2)):
            eng.set_step_layout(layout)
            try:
                ms = _time(eng, lambda: eng.step_match_dev(d_p, d_t, d_i, n, 0, gens, interval, d_h, n, d_n))
                cnt = np.zeros(1, dtype=np.uint64)
                eng.d2h(cnt, d_n)
                eng.sync()
                hits = np.zeros(int(cnt[0]), dtype=[("id", "<u8"), ("a", "<u4"), ("b", "<u4")])
                if len(hits):
                    eng.d2h(hits, d_h)
                    eng.sync()
                res[name] = hits
                out[name] = {"ms": ms, "placements_per_s": n / (ms * 1e-3), "hits": int(cnt[0])}
            finally:
                eng.set_step_layout(0)
        out["kernels_agree"] = bool(res["four_per_group"].tobytes() == res["lane_group_layout2"].tobytes())
        return out
    finally:
        for p in (d_p, d_t, d_i, d_h, d_n):
            eng.free(p)


def config5_survey(eng, p_int):
    """Soups generated on device, 4096 gens, activity test + hit compaction (1/16 of the 16 Mi-soup config per call)."""
    n, gens = 1 << 20, 4096
    dt = float("inf")
    for _ in range(2):  # the first call also loads the kernel image; report the faster of two, like config 4
        t0 = time.perf_counter()
        hits, pop_sum, hash_xor = eng.soup_survey(0, n, SEED, 0, gens, n)
        dt = min(dt, time.perf_counter() - t0)
    cells = n * 4096.0 * gens
    return {"workload": f"config5 (1/16 scale): {n} on-device soups x {gens} gens, active iff state(G) != state(G-2)", "s": dt,
            "value": cells / dt / 1e9, "unit": "Gcell/s", "active": int(len(hits)), "pop_sum": int(pop_sum),
            "hash_xor": f"{int(hash_xor):016x}", "int_pipe_frac_executed": cells * (7.75 / 32.0) / dt / p_int,
            "timing": "wall clock around the host-facing call, hits downloaded and sorted; faster of two calls"}


def stable_propagate(eng, p_int):
    """LifeStable::Propagate on 64 Ki records (320 MiB resident)."""
    import sys
    import os
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "tests"))
    from stable_inputs import stable_records
    base = stable_records(512, 5)
    n = 1 << 16
    recs = np.ascontiguousarray(np.tile(base, (n // 512, 1, 1)))
    d, d_f, d_w = eng.malloc(recs.nbytes), eng.malloc(n), eng.malloc(recs.nbytes)
    try:
        eng.h2d(d, recs)
        eng.sync()

        def once():
            eng._check(eng.lib.lifeb200_memset(eng._h, d_f, 0, n))
            eng._check(_d2d(eng, d_w, d, recs.nbytes))
            eng.stable_pass_dev(d_w, n, 5, d_f)
        ms = _time(eng, once)
        ms_copy = _time(eng, lambda: eng._check(_d2d(eng, d_w, d, recs.nbytes)))
        flags = np.zeros(n, dtype=np.uint8)
        eng.d2h(flags, d_f)
        eng.sync()
        out = {"workload": f"LifeStable::Propagate to the fixed point on {n} records (5120 B each)", "ms": ms - ms_copy,
               "records_per_s": n / ((ms - ms_copy) * 1e-3), "consistent": int((flags & 1).sum()),
               "hbm_gbs": 2 * recs.nbytes / ((ms - ms_copy) * 1e-3) / 1e9}
        ref = _cpu_ref()
        if ref is not None:
            m = 1 << 15
            t0 = time.perf_counter()
            ref.stable_pass_batch(recs[:m], 5, 0)
            dt = time.perf_counter() - t0
            out["cpu_baseline"] = {"records_per_s": m / dt, "cores": ref.threads(), "kind": "reference",
                                   "sample": f"{m} records, {dt:.2f} s (includes one copy of the sample)"}
        return out
    finally:
        for p in (d, d_f, d_w):
            eng.free(p)


def locate_rows(eng, p_int):
    """SURVEY 8f rank 4 -- the steps either side of the stepping path, resident batches:
    Match(LifeTarget) of one catalyst in 256 Ki evolved soups, and InteractionOffsets of one catalyst
    against 256 Ki glider-sized objects (a soup masked to an 8x8 box)."""
    n = 1 << 18
    d_s, d_o, d_t = eng.malloc(n * 512), eng.malloc(n * 512), eng.malloc(1024)
    try:
        eng.soup_dev(d_s, 0, n, SEED, 0)
        eng.step_batch_dev(d_s, d_s, n, 256)  # evolved soups: the ash a search would look through
        eater = _pattern(STILL_LIFES["eater1"], 0, 0)
        zoi = np.zeros(64, dtype=np.uint64)
        for dx in (-1, 0, 1):
            for dy in (-1, 0, 1):
                zoi |= _pattern(STILL_LIFES["eater1"], dx, dy)
        target = np.concatenate([eater, zoi & ~eater])  # LifeTarget(eater): wanted, unwanted = boundary
        eng.h2d(d_t, target)
        eng.sync()
        ms = _time(eng, lambda: eng.match_target_batch_dev(d_s, n, d_t, 1, d_o))
        found = np.empty((n, 64), dtype=np.uint64)
        eng.d2h(found, d_o)
        eng.sync()
        out = {"match_target": {"workload": f"Match(LifeTarget(eater1)) over {n} soups evolved 256 gens", "ms": ms,
                                "states_per_s": n / (ms * 1e-3), "hbm_gbs": 2 * n * 512 / (ms * 1e-3) / 1e9,
                                "soups_with_an_eater": int((found != 0).any(axis=1).sum())}}
        ref = _cpu_ref()
        host = None
        if ref is not None:
            m = 1 << 12
            host = np.empty((n, 64), dtype=np.uint64)
            eng.d2h(host, d_s)
            eng.sync()
            t0 = time.perf_counter()
            want = ref.match_target_batch(host[:m], target, 0)
            dt = time.perf_counter() - t0
            out["match_target"]["cpu_baseline"] = {"states_per_s": m / dt, "cores": ref.threads(), "kind": "reference",
                                                   "sample": f"{m} states, {dt:.2f} s", "equal_on_sample": bool((want == found[:m]).all())}
        # InteractionOffsets: catalyst (broadcast) x small objects
        eng.soup_dev(d_s, 1 << 24, n, SEED, 0)
        objs = np.empty((n, 64), dtype=np.uint64)
        eng.d2h(objs, d_s)
        eng.sync()
        box = np.zeros(64, dtype=np.uint64)
        box[:8] = np.uint64(0xFF)
        objs &= box
        eng.h2d(d_s, objs)
        d_c = eng.malloc(512)
        try:
            eng.h2d(d_c, eater)
            eng.sync()
            ms = _time(eng, lambda: eng.interaction_offsets_dev(d_c, 1, d_s, n, d_o, n))
            out["interaction_offsets"] = {"workload": f"eater1.InteractionOffsets(object) over {n} 8x8 objects", "ms": ms,
                                          "pairs_per_s": n / (ms * 1e-3), "hbm_gbs": 2 * n * 512 / (ms * 1e-3) / 1e9}
            if ref is not None:
                m = 1 << 12
                got = np.empty((n, 64), dtype=np.uint64)
                eng.d2h(got, d_o)
                eng.sync()
                t0 = time.perf_counter()
                want = ref.interaction_offsets_batch(eater, objs[:m], 0)
                dt = time.perf_counter() - t0
                out["interaction_offsets"]["cpu_baseline"] = {"pairs_per_s": m / dt, "cores": ref.threads(), "kind": "reference",
                                                              "sample": f"{m} pairs, {dt:.2f} s",
                                                              "equal_on_sample": bool((want == got[:m]).all())}
        finally:
            eng.free(d_c)
        return out
    finally:
        for p in (d_s, d_o, d_t):
            eng.free(p)


def next_rows(eng, p_int):
    """SURVEY 8f ranks 1-3 on resident batches: LifeWeld::Step (64 generations in registers), GetHash (XXH3-64
    of every state) and InteractionCountsAndNext (four planes per pattern)."""
    out = {}
    ref = _cpu_ref()
    n, gens = 1 << 17, 64
    d = eng.malloc(n * 2048)
    try:
        eng.soup_dev(d, 0, 4 * n, SEED, 0)  # record = (state, frozen2, frozen1, frozen0), 4 x 512 B
        recs = np.empty((n, 4, 64), dtype=np.uint64)
        eng.d2h(recs, d)
        eng.sync()
        recs[:, 1] &= recs[:, 2] & recs[:, 3]   # frozen counts: sparse planes (1/8, 1/4 and 1/2 density)
        recs[:, 2] &= recs[:, 3]
        eng.h2d(d, recs)
        eng.sync()
        ms = _time(eng, lambda: eng._check(eng.lib.lifeb200_weld_step_batch_dev(eng._h, d, n, gens)))
        out["weld_step"] = {"workload": f"LifeWeld::Step x {gens} generations on {n} records (2048 B each), in place, dense random frozen planes", "ms": ms,
                            "value": n * 4096.0 * gens / (ms * 1e-3) / 1e9, "unit": "Gcell/s",
                            "executed": "12.25 int-pipe instr per 32 cells per generation (weld_step_kernel_x, profiles/r02_sass_mix.txt) "
                                        "+ about 55 per 32 cells per launch for the five plane conversions",
                            "pipe_utilisation": n * 4096.0 * (gens * 12.25 + 55.0) / 32.0 / (ms * 1e-3) / p_int}
        if ref is not None:
            m = 1 << 14
            t0 = time.perf_counter()
            want = ref.weld_step_batch(recs[:m], gens, 0)
            dt = time.perf_counter() - t0
            got = np.empty((m, 4, 64), dtype=np.uint64)
            eng.h2d(d, recs)  # the timing loop stepped the records in place: start again from the inputs
            eng._check(eng.lib.lifeb200_weld_step_batch_dev(eng._h, d, n, gens))
            eng.d2h(got, d)
            eng.sync()
            out["weld_step"]["cpu_baseline"] = {"value": m * 4096.0 * gens / dt / 1e9, "unit": "Gcell/s", "cores": ref.threads(),
                                                "kind": "reference", "sample": f"{m} records x {gens} gens, {dt:.2f} s, {ref.threads()} threads",
                                                "equal_on_sample": bool((got == want).all())}
    finally:
        eng.free(d)
    n = 1 << 20
    d_s, d_h = eng.malloc(n * 512), eng.malloc(n * 8)
    try:
        eng.soup_dev(d_s, 0, n, SEED, 0)
        ms = _time(eng, lambda: eng.hash_batch_dev(d_s, d_h, n))
        out["get_hash"] = {"workload": f"GetHash (XXH3-64 of the 512-byte state) over {n} resident soups", "ms": ms,
                           "states_per_s": n / (ms * 1e-3), "hbm_gbs": n * 520 / (ms * 1e-3) / 1e9}
        if ref is not None:
            m = 1 << 18
            host = np.empty((m, 64), dtype=np.uint64)
            eng.d2h(host, d_s)
            got = np.empty(m, dtype=np.uint64)
            eng.d2h(got, d_h)
            eng.sync()
            t0 = time.perf_counter()
            want = ref.hash_batch(host, 0)
            dt = time.perf_counter() - t0
            out["get_hash"]["cpu_baseline"] = {"states_per_s": m / dt, "cores": ref.threads(), "kind": "reference",
                                               "sample": f"{m} states, {dt:.3f} s, {ref.threads()} threads (native loop over GetHash)",
                                               "equal_on_sample": bool((want == got).all())}
    finally:
        eng.free(d_s)
        eng.free(d_h)
    n = 1 << 18
    d_s, d_p = eng.malloc(n * 512), eng.malloc(n * 2048)
    try:
        eng.soup_dev(d_s, 0, n, SEED, 0)
        ms = _time(eng, lambda: eng._check(eng.lib.lifeb200_interaction_counts_dev(eng._h, d_s, d_p, n)))
        out["interaction_counts"] = {"workload": f"InteractionCountsAndNext over {n} resident soups (4 planes out)", "ms": ms,
                                     "states_per_s": n / (ms * 1e-3), "hbm_gbs": n * 2560 / (ms * 1e-3) / 1e9}
        if ref is not None:
            m = 1 << 16
            host = np.empty((m, 64), dtype=np.uint64)
            eng.d2h(host, d_s)
            got = np.empty((m, 4, 64), dtype=np.uint64)
            eng.d2h(got, d_p)
            eng.sync()
            t0 = time.perf_counter()
            want = ref.interaction_counts_batch(host, 0)
            dt = time.perf_counter() - t0
            out["interaction_counts"]["cpu_baseline"] = {"states_per_s": m / dt, "cores": ref.threads(), "kind": "reference",
                                                         "sample": f"{m} states, {dt:.3f} s, {ref.threads()} threads",
                                                         "equal_on_sample": bool((want == got).all())}
    finally:
        eng.free(d_s)
        eng.free(d_p)
    return out


def _d2d(eng, dst, src, nbytes):
    # device-to-device copy through the step entry point's gens == 0 path (a plain cudaMemcpyAsync D2D)
    return eng.lib.lifeb200_step_batch_dev(eng._h, src, dst, nbytes // 512, 0)


def run_extras(eng, p_int):
    out = {}
    for name, fn in (("config3", config3_unknown_step), ("config4", config4_enumeration), ("step_match", step_match_explicit), ("config5", config5_survey),
                     ("stable_propagate", stable_propagate), ("next_rows", next_rows), ("locate", locate_rows)):
        try:
            out[name] = fn(eng, p_int)
        except Exception as e:  # noqa: BLE001  (a side measurement must not kill the headline)
            out[name] = {"error": repr(e)}
    return out
