"""Seeded input generators shared by tests/golden/make_golden.py and the tests (no Life logic).

LifeStable records (10 planes x 64 words, LifeStable.hpp:39-53) describing a stable background
with unknown cells: a field of well-separated still lifes, some rectangles of which are marked
unknown (state &= ~unknown, LifeStable.hpp:320-335 invariant), all option planes open (0).
"""
from __future__ import annotations

import numpy as np

STILL_LIFES = {  # (x, y) cells, bounding box at most 4x4
    "block": [(0, 0), (1, 0), (0, 1), (1, 1)],
    "tub": [(1, 0), (0, 1), (2, 1), (1, 2)],
    "boat": [(0, 0), (1, 0), (0, 1), (2, 1), (1, 2)],
    "ship": [(0, 0), (1, 0), (0, 1), (2, 1), (1, 2), (2, 2)],
    "beehive": [(1, 0), (2, 0), (0, 1), (3, 1), (1, 2), (2, 2)],
    "loaf": [(1, 0), (2, 0), (0, 1), (3, 1), (1, 2), (3, 2), (2, 3)],
    "pond": [(1, 0), (2, 0), (0, 1), (3, 1), (0, 2), (3, 2), (1, 3), (2, 3)],
    "empty": [],
}


def _set(plane: np.ndarray, x: int, y: int) -> None:
    plane[x & 63] |= np.uint64(1) << np.uint64(y & 63)  # LifeState::Set: state[x] |= 1 << y


def still_life_field(rng: np.random.Generator) -> np.ndarray:
    """One 64x64 torus of 8x8 tiles, each holding at most one still life with a 1-cell margin."""
    names = list(STILL_LIFES)
    plane = np.zeros(64, dtype=np.uint64)
    for tx in range(8):
        for ty in range(8):
            cells = STILL_LIFES[names[int(rng.integers(len(names)))]]
            if not cells:
                continue
            w = max(c[0] for c in cells) + 1
            h = max(c[1] for c in cells) + 1
            ox = int(rng.integers(1, 8 - w))
            oy = int(rng.integers(1, 8 - h))
            for (x, y) in cells:
                _set(plane, tx * 8 + ox + x, ty * 8 + oy + y)
    return plane


def unknown_blobs(rng: np.random.Generator, count: int, max_size: int) -> np.ndarray:
    plane = np.zeros(64, dtype=np.uint64)
    for _ in range(count):
        x0, y0 = int(rng.integers(64)), int(rng.integers(64))
        w, h = int(rng.integers(1, max_size + 1)), int(rng.integers(1, max_size + 1))
        for x in range(w):
            for y in range(h):
                _set(plane, x0 + x, y0 + y)
    return plane


def stable_records(n: int, seed: int) -> np.ndarray:
    """(n, 10, 64) records: state, unknown, 8 open option planes."""
    rng = np.random.default_rng(seed)
    rec = np.zeros((n, 10, 64), dtype=np.uint64)
    for i in range(n):
        field = still_life_field(rng)
        unk = unknown_blobs(rng, int(rng.integers(0, 12)), int(rng.integers(1, 7)))
        rec[i, 0] = field & ~unk
        rec[i, 1] = unk
    return rec


def three_valued_patterns(n: int, seed: int, soups: np.ndarray, zoi_fn) -> tuple[np.ndarray, np.ndarray]:
    """soup masked to the centred 32x32 box, unknown at density 1/8 inside its ZOI."""
    rng = np.random.default_rng(seed)
    box = np.zeros(64, dtype=np.uint64)
    box[16:48] = np.uint64(0x0000FFFFFFFF0000)
    state = np.ascontiguousarray(soups[:n] & box)
    unk = np.zeros_like(state)
    for p in range(n):
        r = rng.integers(0, 1 << 63, size=(3, 64), dtype=np.uint64) << np.uint64(1) | rng.integers(0, 2, size=(3, 64), dtype=np.uint64)
        unk[p] = r[0] & r[1] & r[2] & zoi_fn(state[p])
    state &= ~unk
    return state, unk
