"""oracle/loader.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

ctypes loaders for the two checkers:
  * Port  : oracle/life_oracle.c + circuits_oracle.c (C restatement; built on demand with gcc)
  * Ref   : oracle/_ref/libliferef_{v2,v3,v4}.so = the UNMODIFIED reference headers compiled in
            the build container (prebuilt binary; None if absent).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
Nothing under lifeapi_b200/ does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(HERE, "_build", "liblifeoracle.so")
REF_DIR = os.path.join(HERE, "_ref")

_u64p = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")
_u32p = np.ctypeslib.ndpointer(dtype=np.uint32, flags="C_CONTIGUOUS")
_u8p = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")


def aligned_copy(a: np.ndarray, align: int = 64) -> np.ndarray:
    """C-contiguous uint64 copy whose data pointer is `align`-byte aligned (LifeState is alignas(64))."""
    a = np.asarray(a, dtype=np.uint64)
    raw = np.empty(a.size * 8 + align, dtype=np.uint8)
    off = (-raw.ctypes.data) % align
    out = raw[off:off + a.size * 8].view(np.uint64).reshape(a.shape)
    out[...] = a
    return out


def build_port(force: bool = False) -> str:
    srcs = [os.path.join(HERE, f) for f in ("life_oracle.c", "circuits_oracle.c", "stable_oracle.c", "next_oracle.c", "convolve_oracle.c", "circuit_tables.h")]
    if force or not os.path.exists(PORT_SO) or any(os.path.getmtime(s) > os.path.getmtime(PORT_SO) for s in srcs):
        subprocess.run(["make", "-C", HERE, "port"], check=True, stdout=subprocess.DEVNULL)
    return PORT_SO


def build_ref() -> bool:
    """Compile the reference where it lies (only possible where /root/reference exists)."""
    if not os.path.isdir("/root/reference"):
        return False
    subprocess.run(["make", "-j3", "-C", HERE, "ref"], check=True, stdout=subprocess.DEVNULL)
    return True


def _cpu_flags() -> set[str]:
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("flags"):
                return set(line.split(":", 1)[1].split())
    except OSError:
        pass
    return set()


def ref_path() -> str | None:
    flags = _cpu_flags()
    order = []
    if {"avx512f", "avx512bw", "avx512vl", "avx512dq", "avx512cd"} <= flags:
        order.append("v4")
    if {"avx2", "bmi2", "fma"} <= flags:
        order.append("v3")
    order.append("v2")
    for lvl in order:
        p = os.path.join(REF_DIR, f"libliferef_{lvl}.so")
        if os.path.exists(p):
            return p
    return None


def _sig(lib, name, res, args):
    fn = getattr(lib, name)
    fn.restype = res
    fn.argtypes = args
    return fn


class Port:
    """The C restatement (kind = "port")."""

    kind = "port"

    def __init__(self):
        lib = C.CDLL(build_port())
        self.lib = lib
        _sig(lib, "lo_step", None, [_u64p])
        _sig(lib, "lo_step_alt", None, [_u64p])
        _sig(lib, "lo_step_naive", None, [_u64p])
        _sig(lib, "lo_step_n", None, [_u64p, C.c_uint])
        _sig(lib, "lo_step_batch", None, [_u64p, C.c_size_t, C.c_uint, C.c_int])
        _sig(lib, "lo_max_threads", C.c_int, [])
        _sig(lib, "lo_neighbour_count", None, [_u64p, _u64p])
        _sig(lib, "lo_nc_add", None, [_u64p, _u64p, _u64p, _u64p])
        _sig(lib, "lo_nc_subtract", None, [_u64p, _u64p, _u64p])
        _sig(lib, "lo_nc_with_exactly", None, [_u64p, C.c_uint, _u64p])
        _sig(lib, "lo_zoi", None, [_u64p, _u64p])
        _sig(lib, "lo_target_from_state", None, [_u64p, _u64p])
        _sig(lib, "lo_contains_target", C.c_int, [_u64p, _u64p])
        _sig(lib, "lo_moved", None, [_u64p, C.c_int, C.c_int, _u64p])
        _sig(lib, "lo_pop", C.c_uint, [_u64p])
        _sig(lib, "lo_soup", None, [_u64p, C.c_uint64, C.c_size_t, C.c_uint64, C.c_int])
        _sig(lib, "lo_step_match", C.c_int, [_u64p, _u64p, C.c_uint, C.c_uint, C.POINTER(C.c_uint), C.POINTER(C.c_uint)])
        _sig(lib, "lo_unknown_step", None, [_u64p, _u64p])
        _sig(lib, "lo_unknown_step_batch", None, [_u64p, _u64p, C.c_size_t, C.c_uint, C.c_int])
        _sig(lib, "lo_fold_batch", None, [_u64p, C.c_size_t, C.c_uint64, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)])
        _sig(lib, "lo_enumerate", C.c_size_t, [_u64p, C.c_uint, _u64p, C.c_uint, C.c_uint64, C.c_uint64, C.c_uint, C.c_uint,
                                               _u64p, _u32p, _u32p, C.c_size_t, C.POINTER(C.c_uint64)])
        _sig(lib, "lo_survey", C.c_size_t, [C.c_uint64, C.c_size_t, C.c_uint64, C.c_int, C.c_uint, _u64p, _u32p, C.c_size_t,
                                            C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)])
        _sig(lib, "lo_circuit_eval", C.c_int, [C.c_int, _u64p, _u64p])
        _sig(lib, "lo_zoi_hollow", None, [_u64p, _u64p])
        _sig(lib, "lo_unknown_step_sop_batch", None, [_u64p, _u64p, C.c_size_t, C.c_uint, C.c_int])
        _sig(lib, "lo_stable_pass", C.c_uint, [_u64p, C.c_int])
        _sig(lib, "lo_stable_pass_batch", None, [_u64p, C.c_size_t, C.c_int, _u8p, C.c_int])
        _sig(lib, "lo_stable_vulnerable", None, [_u64p, _u64p])
        _sig(lib, "lo_weld_step_n", None, [_u64p, C.c_uint])
        _sig(lib, "lo_interaction_counts", None, [_u64p, _u64p])
        _sig(lib, "lo_xxh3_512", C.c_uint64, [_u64p])
        _sig(lib, "lo_refined_step", None, [_u64p, _u64p, _u64p, C.c_int, _u64p, _u64p, _u64p])
        _sig(lib, "lo_mirrored", None, [_u64p, _u64p])
        _sig(lib, "lo_convolve", None, [_u64p, _u64p, _u64p])
        _sig(lib, "lo_match_live", None, [_u64p, _u64p, _u64p])
        _sig(lib, "lo_match_live_and_dead", None, [_u64p, _u64p, _u64p, _u64p])
        _sig(lib, "lo_match_state", None, [_u64p, _u64p, _u64p])
        _sig(lib, "lo_match_target", None, [_u64p, _u64p, _u64p])
        _sig(lib, "lo_interaction_offsets", None, [_u64p, _u64p, _u64p])
        _sig(lib, "lo_rle", C.c_size_t, [_u64p, C.c_char_p, C.c_size_t])
        _sig(lib, "lo_history_rle", C.c_size_t, [_u64p, C.c_char_p, C.c_size_t])
        _sig(lib, "lo_parse", None, [C.c_char_p, _u64p])

    # -- Convolve / Match / InteractionOffsets / RLE -------------------------
    _PFX = "lo_"

    def _plane_fn(self, name, *planes):
        out = np.empty(64, dtype=np.uint64)
        getattr(self.lib, self._PFX + name)(*[np.ascontiguousarray(p, dtype=np.uint64).reshape(-1) for p in planes], out)
        return out

    def mirrored(self, s):
        return self._plane_fn("mirrored", s)

    def convolve(self, a, b):
        return self._plane_fn("convolve", a, b)

    def match_live(self, s, live):
        return self._plane_fn("match_live", s, live)

    def match_live_and_dead(self, s, live, dead):
        return self._plane_fn("match_live_and_dead", s, live, dead)

    def match_state(self, s, live):
        return self._plane_fn("match_state", s, live)

    def match_target(self, s, target):
        return self._plane_fn("match_target", s, target)

    def interaction_offsets(self, a, b):
        return self._plane_fn("interaction_offsets", a, b)

    def _rle(self, name, planes) -> str:
        buf = C.create_string_buffer(1 << 16)
        n = getattr(self.lib, self._PFX + name)(np.ascontiguousarray(planes, dtype=np.uint64).reshape(-1), buf, len(buf))
        assert n < len(buf)
        return buf.value.decode()

    def rle(self, s) -> str:
        return self._rle("rle", s)

    def history_rle(self, rec) -> str:
        return self._rle("history_rle", rec)

    def parse(self, rle: str, dx: int = 0, dy: int = 0) -> np.ndarray:
        out = np.empty(64, dtype=np.uint64)
        self.lib.lo_parse(rle.encode(), out)
        return self.moved(out, dx, dy) if (dx or dy) else out

    # -- helpers returning fresh arrays ---------------------------------------------------------
    def threads(self) -> int:
        return self.lib.lo_max_threads()

    def soup(self, first_id: int, n: int, seed: int, dist: int = 0) -> np.ndarray:
        out = np.empty((n, 64), dtype=np.uint64)
        self.lib.lo_soup(out, first_id, n, seed, dist)
        return out

    def step_batch(self, states: np.ndarray, gens: int, threads: int = 0) -> np.ndarray:
        out = np.ascontiguousarray(states, dtype=np.uint64).copy()
        self.lib.lo_step_batch(out.reshape(-1), out.size // 64, gens, threads)
        return out

    def step(self, s: np.ndarray, kind: str = "step") -> np.ndarray:
        out = np.ascontiguousarray(s, dtype=np.uint64).copy()
        getattr(self.lib, {"step": "lo_step", "alt": "lo_step_alt", "naive": "lo_step_naive"}[kind])(out)
        return out

    def neighbour_count(self, s: np.ndarray) -> np.ndarray:
        out = np.empty((4, 64), dtype=np.uint64)
        self.lib.lo_neighbour_count(np.ascontiguousarray(s, dtype=np.uint64), out.reshape(-1))
        return out

    def zoi(self, s):
        out = np.empty(64, dtype=np.uint64)
        self.lib.lo_zoi(np.ascontiguousarray(s, dtype=np.uint64), out)
        return out

    def moved(self, s, x, y):
        out = np.empty(64, dtype=np.uint64)
        self.lib.lo_moved(np.ascontiguousarray(s, dtype=np.uint64), x, y, out)
        return out

    def target_from_state(self, s):
        out = np.empty(128, dtype=np.uint64)
        self.lib.lo_target_from_state(np.ascontiguousarray(s, dtype=np.uint64), out)
        return out

    def contains_target(self, s, target) -> bool:
        return bool(self.lib.lo_contains_target(np.ascontiguousarray(s, dtype=np.uint64),
                                                np.ascontiguousarray(target, dtype=np.uint64).reshape(-1)))

    def pop(self, s) -> int:
        return self.lib.lo_pop(np.ascontiguousarray(s, dtype=np.uint64))

    def step_match(self, s, target, gens, stable_interval):
        st = np.ascontiguousarray(s, dtype=np.uint64).copy()
        a, b = C.c_uint(), C.c_uint()
        hit = self.lib.lo_step_match(st, np.ascontiguousarray(target, dtype=np.uint64).reshape(-1), gens, stable_interval,
                                     C.byref(a), C.byref(b))
        return (a.value, b.value) if hit else None

    def unknown_step_batch(self, state, unknown, gens, threads: int = 0):
        s = np.ascontiguousarray(state, dtype=np.uint64).copy()
        u = np.ascontiguousarray(unknown, dtype=np.uint64).copy()
        self.lib.lo_unknown_step_batch(s.reshape(-1), u.reshape(-1), s.size // 64, gens, threads)
        return s, u

    def fold_batch(self, states, first_id: int = 0):
        st = np.ascontiguousarray(states, dtype=np.uint64).reshape(-1)
        ps, hx = C.c_uint64(), C.c_uint64()
        self.lib.lo_fold_batch(st, st.size // 64, first_id, C.byref(ps), C.byref(hx))
        return ps.value, hx.value

    def enumerate(self, catalysts, projectiles, id_begin, id_end, gens, stable_interval, cap, fn="lo_enumerate", lib=None):
        cat = np.ascontiguousarray(catalysts, dtype=np.uint64).reshape(-1)
        pro = np.ascontiguousarray(projectiles, dtype=np.uint64).reshape(-1)
        ids = np.zeros(max(cap, 1), dtype=np.uint64)
        ff = np.zeros(max(cap, 1), dtype=np.uint32)
        rec = np.zeros(max(cap, 1), dtype=np.uint32)
        ns = C.c_uint64()
        nh = getattr(lib or self.lib, fn)(cat, cat.size // 64, pro, pro.size // 64, id_begin, id_end, gens,
                                          stable_interval, ids, ff, rec, cap, C.byref(ns))
        nh = min(nh, cap)
        return ids[:nh], ff[:nh], rec[:nh], ns.value

    def circuit_eval(self, which: int, inputs: np.ndarray) -> np.ndarray:
        """inputs: (n_in,) uint64 words -> (n_out,) uint64 words through the cube table of fragment `which`."""
        out = np.zeros(16, dtype=np.uint64)
        k = self.lib.lo_circuit_eval(which, np.ascontiguousarray(inputs, dtype=np.uint64), out)
        if k < 0:
            raise ValueError(which)
        return out[:k]

    def zoi_hollow(self, s):
        out = np.empty(64, dtype=np.uint64)
        self.lib.lo_zoi_hollow(np.ascontiguousarray(s, dtype=np.uint64), out)
        return out

    def unknown_step_sop_batch(self, state, unknown, gens, threads: int = 0):
        s = np.ascontiguousarray(state, dtype=np.uint64).copy()
        u = np.ascontiguousarray(unknown, dtype=np.uint64).copy()
        self.lib.lo_unknown_step_sop_batch(s.reshape(-1), u.reshape(-1), s.size // 64, gens, threads)
        return s, u

    def stable_pass_batch(self, planes, which: int, threads: int = 0, fn="lo_stable_pass_batch", lib=None):
        """planes: (n, 10, 64) uint64 LifeStable records -> (new planes, flags uint8: bit0 consistent, bit1 changed)."""
        p = aligned_copy(np.asarray(planes, dtype=np.uint64).reshape(-1, 10, 64))
        flags = np.zeros(p.shape[0], dtype=np.uint8)
        getattr(lib or self.lib, fn)(p.reshape(-1), p.shape[0], which, flags, threads)
        return p, flags

    def stable_vulnerable(self, planes, fn="lo_stable_vulnerable", lib=None):
        p = aligned_copy(np.asarray(planes, dtype=np.uint64).reshape(-1, 10, 64))
        out = np.empty((p.shape[0], 64), dtype=np.uint64)
        for i in range(p.shape[0]):
            getattr(lib or self.lib, fn)(p[i].reshape(-1), out[i])
        return out

    def refined_step(self, planes, cur_on, cur_unk, variant: int, fn="lo_refined_step", lib=None):
        p = aligned_copy(np.asarray(planes, dtype=np.uint64).reshape(-1, 10, 64))
        con = aligned_copy(np.asarray(cur_on, dtype=np.uint64).reshape(-1, 64))
        cun = aligned_copy(np.asarray(cur_unk, dtype=np.uint64).reshape(-1, 64))
        outs = [np.empty_like(con) for _ in range(3)]
        for i in range(p.shape[0]):
            getattr(lib or self.lib, fn)(p[i].reshape(-1), con[i], cun[i], variant, outs[0][i], outs[1][i], outs[2][i])
        return tuple(outs)

    def weld_step(self, records, gens: int, fn="lo_weld_step_n", lib=None):
        r = aligned_copy(np.asarray(records, dtype=np.uint64).reshape(-1, 4, 64))
        for i in range(r.shape[0]):
            getattr(lib or self.lib, fn)(r[i].reshape(-1), gens)
        return r

    def interaction_counts(self, patterns, fn="lo_interaction_counts", lib=None):
        p = aligned_copy(np.asarray(patterns, dtype=np.uint64).reshape(-1, 64))
        out = aligned_copy(np.zeros((p.shape[0], 4, 64), dtype=np.uint64))
        for i in range(p.shape[0]):
            getattr(lib or self.lib, fn)(p[i], out[i].reshape(-1))
        return out

    def hash(self, s) -> int:
        return self.lib.lo_xxh3_512(np.ascontiguousarray(s, dtype=np.uint64))

    def survey(self, first_id, n, seed, dist, gens, cap):
        ids = np.zeros(max(cap, 1), dtype=np.uint64)
        pops = np.zeros(max(cap, 1), dtype=np.uint32)
        ps, hx = C.c_uint64(), C.c_uint64()
        nh = self.lib.lo_survey(first_id, n, seed, dist, gens, ids, pops, cap, C.byref(ps), C.byref(hx))
        nh = min(nh, cap)
        return ids[:nh], pops[:nh], ps.value, hx.value


class Ref:
    """The unmodified reference compiled here (kind = "reference"); Ref.load() is None if absent."""

    kind = "reference"

    @staticmethod
    def load() -> "Ref | None":
        p = ref_path()
        return Ref(p) if p else None

    def __init__(self, path: str):
        lib = C.CDLL(path)
        self.lib = lib
        self.path = path
        _sig(lib, "ref_isa", C.c_char_p, [])
        _sig(lib, "ref_step_batch", None, [_u64p, C.c_size_t, C.c_uint, C.c_int])
        _sig(lib, "ref_hardware_threads", C.c_int, [])
        _sig(lib, "ref_step", None, [_u64p])
        _sig(lib, "ref_step_alt", None, [_u64p])
        _sig(lib, "ref_step_n", None, [_u64p, C.c_uint])
        _sig(lib, "ref_hash", C.c_uint64, [_u64p])
        _sig(lib, "ref_pop", C.c_uint, [_u64p])
        _sig(lib, "ref_parse", None, [C.c_char_p, _u64p])
        _sig(lib, "ref_parse_at", None, [C.c_char_p, C.c_int, C.c_int, _u64p])
        _sig(lib, "ref_moved", None, [_u64p, C.c_int, C.c_int, _u64p])
        _sig(lib, "ref_zoi", None, [_u64p, _u64p])
        _sig(lib, "ref_neighbour_count", None, [_u64p, _u64p])
        _sig(lib, "ref_count_neighbourhood", None, [_u64p, _u64p])
        _sig(lib, "ref_nc_add", None, [_u64p, _u64p, _u64p, _u64p])
        _sig(lib, "ref_nc_subtract", None, [_u64p, _u64p, _u64p])
        _sig(lib, "ref_nc_with_exactly", None, [_u64p, C.c_uint, _u64p])
        _sig(lib, "ref_target_from_state", None, [_u64p, _u64p])
        _sig(lib, "ref_contains_target", C.c_int, [_u64p, _u64p])
        _sig(lib, "ref_step_match", C.c_int, [_u64p, _u64p, C.c_uint, C.c_uint, C.POINTER(C.c_uint), C.POINTER(C.c_uint)])
        _sig(lib, "ref_unknown_step", None, [_u64p, _u64p])
        _sig(lib, "ref_unknown_step_batch", None, [_u64p, _u64p, C.c_size_t, C.c_uint, C.c_int])
        _sig(lib, "ref_enumerate", C.c_size_t, [_u64p, C.c_uint, _u64p, C.c_uint, C.c_uint64, C.c_uint64, C.c_uint, C.c_uint,
                                                _u64p, _u32p, _u32p, C.c_size_t, C.POINTER(C.c_uint64)])
        _sig(lib, "ref_survey_states", None, [_u64p, C.c_size_t, C.c_uint, _u8p, _u32p])
        _sig(lib, "ref_stable_pass", C.c_uint, [_u64p, C.c_int])
        _sig(lib, "ref_stable_pass_batch", None, [_u64p, C.c_size_t, C.c_int, _u8p, C.c_int])
        _sig(lib, "ref_stable_vulnerable", None, [_u64p, _u64p])
        _sig(lib, "ref_zoi_hollow", None, [_u64p, _u64p])
        _sig(lib, "ref_weld_step_n", None, [_u64p, C.c_uint])
        _sig(lib, "ref_interaction_counts", None, [_u64p, _u64p])
        _sig(lib, "ref_refined_step", None, [_u64p, _u64p, _u64p, C.c_int, _u64p, _u64p, _u64p])
        _sig(lib, "ref_mirrored", None, [_u64p, _u64p])
        _sig(lib, "ref_convolve", None, [_u64p, _u64p, _u64p])
        _sig(lib, "ref_match_live", None, [_u64p, _u64p, _u64p])
        _sig(lib, "ref_match_live_and_dead", None, [_u64p, _u64p, _u64p, _u64p])
        _sig(lib, "ref_match_state", None, [_u64p, _u64p, _u64p])
        _sig(lib, "ref_match_target", None, [_u64p, _u64p, _u64p])
        _sig(lib, "ref_interaction_offsets", None, [_u64p, _u64p, _u64p])
        _sig(lib, "ref_rle", C.c_size_t, [_u64p, C.c_char_p, C.c_size_t])
        _sig(lib, "ref_history_rle", C.c_size_t, [_u64p, C.c_char_p, C.c_size_t])
        _sig(lib, "ref_history_parse", None, [C.c_char_p, _u64p])
        _sig(lib, "ref_match_target_batch", None, [_u64p, C.c_size_t, _u64p, _u64p, C.c_int])
        _sig(lib, "ref_interaction_offsets_batch", None, [_u64p, _u64p, C.c_size_t, _u64p, C.c_int])
        # native all-thread batch loops (oracle/ref_batch.cpp)
        _sig(lib, "ref_hash_batch", None, [_u64p, C.c_size_t, _u64p, C.c_int])
        _sig(lib, "ref_weld_step_batch", None, [_u64p, C.c_size_t, C.c_uint, C.c_int])
        _sig(lib, "ref_interaction_counts_batch", None, [_u64p, C.c_size_t, _u64p, C.c_int])
        _sig(lib, "ref_weld_interaction_counts_batch", None, [_u64p, C.c_size_t, _u64p, C.c_int])
        _sig(lib, "ref_weld_interaction_offsets_batch", None, [_u64p, C.c_size_t, _u64p, C.c_size_t, C.c_size_t, _u64p, C.c_int])
        _sig(lib, "ref_nc_add_batch", None, [_u64p, _u64p, C.c_void_p, C.c_size_t, _u64p, C.c_int])
        _sig(lib, "ref_nc_subtract_batch", None, [_u64p, _u64p, C.c_size_t, _u64p, C.c_int])
        _sig(lib, "ref_nc_with_exactly_batch", None, [_u64p, C.c_size_t, C.c_uint, _u64p, C.c_int])
        _sig(lib, "ref_contains_target_shifted", C.c_int, [_u64p, _u64p, C.c_int, C.c_int])
        _sig(lib, "ref_enumerate_mt", C.c_size_t, [_u64p, C.c_uint, _u64p, C.c_uint, C.c_uint64, C.c_uint64, C.c_uint, C.c_uint,
                                                   _u64p, _u32p, _u32p, C.c_size_t, C.POINTER(C.c_uint64), C.c_int])
        _sig(lib, "ref_survey_states_mt", None, [_u64p, C.c_size_t, C.c_uint, _u8p, _u32p, C.c_int])

    _PFX = "ref_"
    _plane_fn = Port._plane_fn
    mirrored = Port.mirrored
    convolve = Port.convolve
    match_live = Port.match_live
    match_live_and_dead = Port.match_live_and_dead
    match_state = Port.match_state
    match_target = Port.match_target
    interaction_offsets = Port.interaction_offsets
    _rle = Port._rle
    rle = Port.rle
    history_rle = Port.history_rle

    def history_parse(self, rle: str) -> np.ndarray:
        out = np.empty((4, 64), dtype=np.uint64)
        self.lib.ref_history_parse(rle.encode(), out.reshape(-1))
        return out

    def match_target_batch(self, states, target, threads: int = 0) -> np.ndarray:
        st = aligned_copy(np.asarray(states, dtype=np.uint64).reshape(-1, 64))
        out = aligned_copy(np.zeros_like(st))
        self.lib.ref_match_target_batch(st.reshape(-1), st.shape[0], aligned_copy(np.asarray(target).reshape(-1)), out.reshape(-1), threads)
        return out

    def interaction_offsets_batch(self, a, bs, threads: int = 0) -> np.ndarray:
        b = aligned_copy(np.asarray(bs, dtype=np.uint64).reshape(-1, 64))
        out = aligned_copy(np.zeros_like(b))
        self.lib.ref_interaction_offsets_batch(aligned_copy(np.asarray(a).reshape(-1)), b.reshape(-1), b.shape[0], out.reshape(-1), threads)
        return out

    # -- all-thread batch loops over the reference's own members (oracle/ref_batch.cpp) ----------
    def hash_batch(self, states, threads: int = 0) -> np.ndarray:
        st = aligned_copy(np.asarray(states, dtype=np.uint64).reshape(-1, 64))
        out = np.empty(st.shape[0], dtype=np.uint64)
        self.lib.ref_hash_batch(st.reshape(-1), st.shape[0], out, threads)
        return out

    def weld_step_batch(self, records, gens: int, threads: int = 0) -> np.ndarray:
        r = aligned_copy(np.asarray(records, dtype=np.uint64).reshape(-1, 4, 64))
        self.lib.ref_weld_step_batch(r.reshape(-1), r.shape[0], gens, threads)
        return r

    def interaction_counts_batch(self, patterns, threads: int = 0) -> np.ndarray:
        p = aligned_copy(np.asarray(patterns, dtype=np.uint64).reshape(-1, 64))
        out = aligned_copy(np.zeros((p.shape[0], 4, 64), dtype=np.uint64))
        self.lib.ref_interaction_counts_batch(p.reshape(-1), p.shape[0], out.reshape(-1), threads)
        return out

    def weld_interaction_counts_batch(self, records, threads: int = 0) -> np.ndarray:
        r = aligned_copy(np.asarray(records, dtype=np.uint64).reshape(-1, 4, 64))
        out = aligned_copy(np.zeros((r.shape[0], 3, 64), dtype=np.uint64))
        self.lib.ref_weld_interaction_counts_batch(r.reshape(-1), r.shape[0], out.reshape(-1), threads)
        return out

    def weld_interaction_offsets_batch(self, a, b, threads: int = 0) -> np.ndarray:
        a = aligned_copy(np.asarray(a, dtype=np.uint64).reshape(-1, 4, 64))
        b = aligned_copy(np.asarray(b, dtype=np.uint64).reshape(-1, 4, 64))
        n = max(a.shape[0], b.shape[0])
        out = aligned_copy(np.zeros((n, 64), dtype=np.uint64))
        self.lib.ref_weld_interaction_offsets_batch(a.reshape(-1), a.shape[0], b.reshape(-1), b.shape[0], n, out.reshape(-1), threads)
        return out

    def nc_add_batch(self, a, b, incarry=None, threads: int = 0) -> np.ndarray:
        a = aligned_copy(np.asarray(a, dtype=np.uint64).reshape(-1, 4, 64))
        b = aligned_copy(np.asarray(b, dtype=np.uint64).reshape(-1, 4, 64))
        out = aligned_copy(np.zeros_like(a))
        ic = None
        if incarry is not None:
            ic = aligned_copy(np.asarray(incarry, dtype=np.uint64).reshape(-1, 64))
        self.lib.ref_nc_add_batch(a.reshape(-1), b.reshape(-1), ic.ctypes.data if ic is not None else None, a.shape[0],
                                  out.reshape(-1), threads)
        return out

    def nc_subtract_batch(self, a, b, threads: int = 0) -> np.ndarray:
        a = aligned_copy(np.asarray(a, dtype=np.uint64).reshape(-1, 4, 64))
        b = aligned_copy(np.asarray(b, dtype=np.uint64).reshape(-1, 4, 64))
        out = aligned_copy(np.zeros_like(a))
        self.lib.ref_nc_subtract_batch(a.reshape(-1), b.reshape(-1), a.shape[0], out.reshape(-1), threads)
        return out

    def nc_with_exactly_batch(self, a, k: int, threads: int = 0) -> np.ndarray:
        a = aligned_copy(np.asarray(a, dtype=np.uint64).reshape(-1, 4, 64))
        out = aligned_copy(np.zeros((a.shape[0], 64), dtype=np.uint64))
        self.lib.ref_nc_with_exactly_batch(a.reshape(-1), a.shape[0], k, out.reshape(-1), threads)
        return out

    def contains_target_shifted(self, s, target, dx: int, dy: int) -> bool:
        return bool(self.lib.ref_contains_target_shifted(aligned_copy(np.asarray(s, dtype=np.uint64).reshape(-1)),
                                                         aligned_copy(np.asarray(target, dtype=np.uint64).reshape(-1)), dx, dy))

    def enumerate_mt(self, catalysts, projectiles, id_begin, id_end, gens, stable_interval, cap, threads: int = 0):
        cat = aligned_copy(np.asarray(catalysts, dtype=np.uint64).reshape(-1))
        pro = aligned_copy(np.asarray(projectiles, dtype=np.uint64).reshape(-1))
        ids = np.zeros(max(cap, 1), dtype=np.uint64)
        ff = np.zeros(max(cap, 1), dtype=np.uint32)
        rec = np.zeros(max(cap, 1), dtype=np.uint32)
        ns = C.c_uint64()
        nh = self.lib.ref_enumerate_mt(cat, cat.size // 64, pro, pro.size // 64, id_begin, id_end, gens, stable_interval,
                                       ids, ff, rec, cap, C.byref(ns), threads)
        nh = min(nh, cap)
        return ids[:nh], ff[:nh], rec[:nh], ns.value

    def survey_states_mt(self, states, gens, threads: int = 0):
        st = aligned_copy(np.asarray(states, dtype=np.uint64).reshape(-1, 64))
        n = st.shape[0]
        act = np.zeros(n, dtype=np.uint8)
        pop = np.zeros(n, dtype=np.uint32)
        self.lib.ref_survey_states_mt(st.reshape(-1), n, gens, act, pop, threads)
        return st, act, pop

    def isa(self) -> str:
        return self.lib.ref_isa().decode()

    def threads(self) -> int:
        return self.lib.ref_hardware_threads()

    def parse(self, rle: str, dx: int = 0, dy: int = 0) -> np.ndarray:
        out = np.empty(64, dtype=np.uint64)
        self.lib.ref_parse_at(rle.encode(), dx, dy, out)
        return out

    def step_batch(self, states: np.ndarray, gens: int, threads: int = 0) -> np.ndarray:
        out = aligned_copy(states)
        self.lib.ref_step_batch(out.reshape(-1), out.size // 64, gens, threads)
        return out

    def step(self, s, kind="step"):
        out = np.ascontiguousarray(s, dtype=np.uint64).copy()
        getattr(self.lib, {"step": "ref_step", "alt": "ref_step_alt"}[kind])(out)
        return out

    def hash(self, s) -> int:
        return self.lib.ref_hash(np.ascontiguousarray(s, dtype=np.uint64))

    def pop(self, s) -> int:
        return self.lib.ref_pop(np.ascontiguousarray(s, dtype=np.uint64))

    def neighbour_count(self, s, fn="ref_neighbour_count"):
        out = np.empty((4, 64), dtype=np.uint64)
        getattr(self.lib, fn)(np.ascontiguousarray(s, dtype=np.uint64), out.reshape(-1))
        return out

    def zoi(self, s):
        out = np.empty(64, dtype=np.uint64)
        self.lib.ref_zoi(np.ascontiguousarray(s, dtype=np.uint64), out)
        return out

    def moved(self, s, x, y):
        out = np.empty(64, dtype=np.uint64)
        self.lib.ref_moved(np.ascontiguousarray(s, dtype=np.uint64), x, y, out)
        return out

    def target_from_state(self, s):
        out = np.empty(128, dtype=np.uint64)
        self.lib.ref_target_from_state(np.ascontiguousarray(s, dtype=np.uint64), out)
        return out

    def contains_target(self, s, target) -> bool:
        return bool(self.lib.ref_contains_target(np.ascontiguousarray(s, dtype=np.uint64),
                                                 np.ascontiguousarray(target, dtype=np.uint64).reshape(-1)))

    def step_match(self, s, target, gens, stable_interval):
        st = np.ascontiguousarray(s, dtype=np.uint64).copy()
        a, b = C.c_uint(), C.c_uint()
        hit = self.lib.ref_step_match(st, np.ascontiguousarray(target, dtype=np.uint64).reshape(-1), gens, stable_interval,
                                      C.byref(a), C.byref(b))
        return (a.value, b.value) if hit else None

    def unknown_step_batch(self, state, unknown, gens, threads: int = 0):
        s = np.ascontiguousarray(state, dtype=np.uint64).copy()
        u = np.ascontiguousarray(unknown, dtype=np.uint64).copy()
        self.lib.ref_unknown_step_batch(s.reshape(-1), u.reshape(-1), s.size // 64, gens, threads)
        return s, u

    def enumerate(self, catalysts, projectiles, id_begin, id_end, gens, stable_interval, cap):
        return Port.enumerate(self, catalysts, projectiles, id_begin, id_end, gens, stable_interval, cap,
                              fn="ref_enumerate", lib=self.lib)

    def stable_pass_batch(self, planes, which: int, threads: int = 0):
        return Port.stable_pass_batch(self, planes, which, threads, fn="ref_stable_pass_batch", lib=self.lib)

    def stable_vulnerable(self, planes):
        return Port.stable_vulnerable(self, planes, fn="ref_stable_vulnerable", lib=self.lib)

    def weld_step(self, records, gens: int):
        return Port.weld_step(self, records, gens, fn="ref_weld_step_n", lib=self.lib)

    def interaction_counts(self, patterns):
        return Port.interaction_counts(self, patterns, fn="ref_interaction_counts", lib=self.lib)

    def refined_step(self, planes, cur_on, cur_unk, variant: int):
        return Port.refined_step(self, planes, cur_on, cur_unk, variant, fn="ref_refined_step", lib=self.lib)

    def zoi_hollow(self, s):
        out = np.empty(64, dtype=np.uint64)
        self.lib.ref_zoi_hollow(np.ascontiguousarray(s, dtype=np.uint64), out)
        return out

    def survey_states(self, states, gens):
        st = np.ascontiguousarray(states, dtype=np.uint64).copy()
        n = st.size // 64
        act = np.zeros(n, dtype=np.uint8)
        pop = np.zeros(n, dtype=np.uint32)
        self.lib.ref_survey_states(st.reshape(-1), n, gens, act, pop)
        return st, act, pop
