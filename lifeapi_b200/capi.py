"""ctypes binding of the C ABI in include/lifeapi_b200.h.

This is the Python-side stub a maintainer would use to drive the library (INTEGRATION.md shows
the C++ one).  It adds nothing: every method is one C call plus error translation.  There is no
fallback of any kind -- if the shared library is missing, loading raises; if no sm_100 GPU is
present, `Engine()` raises `LifeB200Error`.
"""
from __future__ import annotations

import ctypes as C
import os
import re

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_lib", "liblifeapi_b200.so")
HEADER_PATH = os.path.join(HERE, "..", "include", "lifeapi_b200.h")

WORDS = 64
PATTERN_BYTES = 512

E_OK, E_CUDA, E_ARG, E_OVERFLOW, E_NODEVICE, E_NCCL, E_HOST = 0, 1, 2, 3, 4, 5, 6
UNIQUE_ID_BYTES = 128

STABLE_WORDS = 640
STABLE_BYTES = 5120
(STABLE_PROPAGATE_SIMPLE_STEP, STABLE_SYNCHRONISE, STABLE_UPDATE_OPTIONS, STABLE_SIGNAL_NEIGHBOURS, STABLE_PROPAGATE_STEP,
 STABLE_PROPAGATE, STABLE_STABILISE_OPTIONS, STABLE_PROPAGATE_SIMPLE) = range(8)
STABLE_CONSISTENT, STABLE_CHANGED = 1, 2


class LifeB200Error(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"lifeapi_b200 error {code}: {msg}")
        self.code = code


class Hit(C.Structure):
    _fields_ = [("id", C.c_uint64), ("g_first_false", C.c_uint32), ("g_recovered", C.c_uint32)]


HIT_DTYPE = np.dtype([("id", "<u8"), ("g_first_false", "<u4"), ("g_recovered", "<u4")])

_u64p = C.POINTER(C.c_uint64)
_vp = C.c_void_p

# name -> (restype, argtypes); must list every function declared in the header
SIGNATURES = {
    "lifeb200_abi_version": (C.c_int, []),
    "lifeb200_last_error": (C.c_char_p, []),
    "lifeb200_device_count": (C.c_int, [C.POINTER(C.c_int)]),
    "lifeb200_create": (C.c_int, [C.c_int, C.POINTER(_vp)]),
    "lifeb200_destroy": (C.c_int, [_vp]),
    "lifeb200_set_stream": (C.c_int, [_vp, _vp]),
    "lifeb200_sync": (C.c_int, [_vp]),
    "lifeb200_device_info": (C.c_int, [_vp, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int),
                                       C.POINTER(C.c_int), C.POINTER(C.c_size_t)]),
    "lifeb200_set_lanes_per_pattern": (C.c_int, [_vp, C.c_int]),
    "lifeb200_set_step_layout": (C.c_int, [_vp, C.c_int]),
    "lifeb200_set_three_valued_path": (C.c_int, [_vp, C.c_int]),
    "lifeb200_malloc": (C.c_int, [_vp, C.POINTER(_vp), C.c_size_t]),
    "lifeb200_free": (C.c_int, [_vp, _vp]),
    "lifeb200_host_alloc": (C.c_int, [_vp, C.POINTER(_vp), C.c_size_t]),
    "lifeb200_host_free": (C.c_int, [_vp, _vp]),
    "lifeb200_host_register": (C.c_int, [_vp, _vp, C.c_size_t]),
    "lifeb200_host_unregister": (C.c_int, [_vp, _vp]),
    "lifeb200_memcpy_h2d": (C.c_int, [_vp, _vp, _vp, C.c_size_t]),
    "lifeb200_memcpy_d2h": (C.c_int, [_vp, _vp, _vp, C.c_size_t]),
    "lifeb200_memset": (C.c_int, [_vp, _vp, C.c_int, C.c_size_t]),
    "lifeb200_timer_start": (C.c_int, [_vp]),
    "lifeb200_timer_stop": (C.c_int, [_vp, C.POINTER(C.c_float)]),
    "lifeb200_launch_count": (C.c_int, [_vp, C.POINTER(C.c_uint64)]),
    "lifeb200_step_batch_dev": (C.c_int, [_vp, _vp, _vp, C.c_size_t, C.c_uint]),
    "lifeb200_step_batch_host": (C.c_int, [_vp, _vp, _vp, C.c_size_t, C.c_uint]),
    "lifeb200_soup_dev": (C.c_int, [_vp, _vp, C.c_uint64, C.c_size_t, C.c_uint64, C.c_int]),
    "lifeb200_neighbour_count_dev": (C.c_int, [_vp, _vp, _vp, C.c_size_t]),
    "lifeb200_neighbour_count_host": (C.c_int, [_vp, _vp, _vp, C.c_size_t]),
    "lifeb200_step_match_dev": (C.c_int, [_vp, _vp, _vp, _vp, C.c_size_t, C.c_uint64, C.c_uint, C.c_uint, _vp,
                                          C.c_size_t, _vp]),
    "lifeb200_step_match_static_dev": (C.c_int, [_vp, _vp, _vp, _vp, C.c_size_t, C.c_uint64, C.c_uint, C.c_uint, _vp,
                                                 C.c_size_t, _vp]),
    "lifeb200_step_match_host": (C.c_int, [_vp, _vp, _vp, _vp, C.c_size_t, C.c_size_t, C.c_uint64, C.c_uint, C.c_uint, _vp,
                                           C.c_size_t, _u64p]),
    "lifeb200_contains_target_batch_dev": (C.c_int, [_vp, _vp, C.c_size_t, _vp, C.c_size_t, C.c_int, C.c_int, _vp]),
    "lifeb200_contains_target_batch_host": (C.c_int, [_vp, _vp, C.c_size_t, _vp, C.c_size_t, C.c_int, C.c_int, _vp]),
    "lifeb200_history_step_batch_host": (C.c_int, [_vp, _vp, C.c_size_t, C.c_uint]),
    "lifeb200_stable_vulnerable_host": (C.c_int, [_vp, _vp, _vp, C.c_size_t]),
    "lifeb200_unknown_step_refined_host": (C.c_int, [_vp, _vp, _vp, _vp, C.c_size_t, C.c_int, _vp, _vp, _vp]),
    "lifeb200_weld_step_batch_host": (C.c_int, [_vp, _vp, C.c_size_t, C.c_uint]),
    "lifeb200_interaction_counts_host": (C.c_int, [_vp, _vp, _vp, C.c_size_t]),
    "lifeb200_nc_add_dev": (C.c_int, [_vp, _vp, _vp, _vp, _vp, C.c_size_t]),
    "lifeb200_nc_subtract_dev": (C.c_int, [_vp, _vp, _vp, _vp, C.c_size_t]),
    "lifeb200_nc_with_exactly_dev": (C.c_int, [_vp, _vp, C.c_uint, _vp, C.c_size_t]),
    "lifeb200_weld_interaction_counts_dev": (C.c_int, [_vp, _vp, _vp, C.c_size_t]),
    "lifeb200_weld_interaction_counts_host": (C.c_int, [_vp, _vp, _vp, C.c_size_t]),
    "lifeb200_enumerate_collisions": (C.c_int, [_vp, _vp, C.c_uint, _vp, C.c_uint, C.c_uint64, C.c_uint64, C.c_uint,
                                                C.c_uint, _vp, C.c_size_t, _u64p, _u64p]),
    "lifeb200_soup_survey": (C.c_int, [_vp, C.c_uint64, C.c_size_t, C.c_uint64, C.c_int, C.c_uint, _vp, C.c_size_t,
                                       _u64p, _u64p, _u64p]),
    "lifeb200_fold_dev": (C.c_int, [_vp, _vp, C.c_size_t, C.c_uint64, _u64p, _u64p]),
    "lifeb200_enumerate_collisions_dev": (C.c_int, [_vp, _vp, C.c_uint, _vp, C.c_uint, C.c_uint64, C.c_uint64, C.c_uint,
                                                    C.c_uint, _vp, C.c_size_t, _u64p, _u64p]),
    "lifeb200_soup_survey_dev": (C.c_int, [_vp, C.c_uint64, C.c_size_t, C.c_uint64, C.c_int, C.c_uint, _vp, C.c_size_t,
                                           _u64p, _u64p, _u64p]),
    "lifeb200_multi_create": (C.c_int, [C.POINTER(C.c_int), C.c_int, C.POINTER(_vp)]),
    "lifeb200_multi_unique_id": (C.c_int, [C.c_char_p]),
    "lifeb200_multi_create_rank": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_char_p, C.POINTER(_vp)]),
    "lifeb200_multi_destroy": (C.c_int, [_vp]),
    "lifeb200_multi_info": (C.c_int, [_vp, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "lifeb200_multi_ctx": (C.c_int, [_vp, C.c_int, C.POINTER(_vp)]),
    "lifeb200_multi_sync": (C.c_int, [_vp]),
    "lifeb200_multi_shard": (C.c_int, [_vp, C.c_uint64, C.c_uint64, C.c_int, _u64p, _u64p]),
    "lifeb200_multi_set_gather": (C.c_int, [_vp, C.c_int]),
    "lifeb200_multi_last_gather_ms": (C.c_int, [_vp, C.POINTER(C.c_float)]),
    "lifeb200_multi_reserve": (C.c_int, [_vp, C.c_uint64, C.c_size_t]),
    "lifeb200_multi_soup_survey": (C.c_int, [_vp, C.c_uint64, C.c_uint64, C.c_uint64, C.c_int, C.c_uint, _vp, C.c_size_t,
                                             _u64p, _u64p, _u64p]),
    "lifeb200_multi_enumerate_collisions": (C.c_int, [_vp, _vp, C.c_uint, _vp, C.c_uint, C.c_uint64, C.c_uint64, C.c_uint,
                                                      C.c_uint, _vp, C.c_size_t, _u64p, _u64p]),
    "lifeb200_multi_step_batch_host": (C.c_int, [_vp, _vp, _vp, C.c_size_t, C.c_uint]),
    "lifeb200_unknown_step_batch_dev": (C.c_int, [_vp, _vp, _vp, _vp, _vp, C.c_size_t, C.c_uint]),
    "lifeb200_unknown_step_batch_host": (C.c_int, [_vp, _vp, _vp, _vp, _vp, C.c_size_t, C.c_uint]),
    "lifeb200_history_step_batch_dev": (C.c_int, [_vp, _vp, C.c_size_t, C.c_uint]),
    "lifeb200_stable_pass_dev": (C.c_int, [_vp, _vp, C.c_size_t, C.c_int, _vp]),
    "lifeb200_stable_pass_host": (C.c_int, [_vp, _vp, C.c_size_t, C.c_int, _vp]),
    "lifeb200_stable_vulnerable_dev": (C.c_int, [_vp, _vp, _vp, C.c_size_t]),
    "lifeb200_unknown_step_refined_dev": (C.c_int, [_vp, _vp, _vp, _vp, C.c_size_t, C.c_int, _vp, _vp, _vp]),
    "lifeb200_weld_step_batch_dev": (C.c_int, [_vp, _vp, C.c_size_t, C.c_uint]),
    "lifeb200_interaction_counts_dev": (C.c_int, [_vp, _vp, _vp, C.c_size_t]),
    "lifeb200_hash_batch_dev": (C.c_int, [_vp, _vp, _vp, C.c_size_t]),
    "lifeb200_hash_batch_host": (C.c_int, [_vp, _vp, _vp, C.c_size_t]),
    "lifeb200_measure_int_peak": (C.c_int, [_vp, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "lifeb200_convolve_batch_dev": (C.c_int, [_vp, _vp, C.c_size_t, _vp, C.c_size_t, _vp, C.c_size_t]),
    "lifeb200_convolve_batch_host": (C.c_int, [_vp, _vp, C.c_size_t, _vp, C.c_size_t, _vp, C.c_size_t]),
    "lifeb200_match_batch_dev": (C.c_int, [_vp, _vp, C.c_size_t, _vp, _vp, C.c_size_t, C.c_int, _vp]),
    "lifeb200_match_batch_host": (C.c_int, [_vp, _vp, C.c_size_t, _vp, _vp, C.c_size_t, C.c_int, _vp]),
    "lifeb200_match_target_batch_dev": (C.c_int, [_vp, _vp, C.c_size_t, _vp, C.c_size_t, _vp]),
    "lifeb200_match_target_batch_host": (C.c_int, [_vp, _vp, C.c_size_t, _vp, C.c_size_t, _vp]),
    "lifeb200_interaction_offsets_dev": (C.c_int, [_vp, _vp, C.c_size_t, _vp, C.c_size_t, _vp, C.c_size_t]),
    "lifeb200_interaction_offsets_host": (C.c_int, [_vp, _vp, C.c_size_t, _vp, C.c_size_t, _vp, C.c_size_t]),
    "lifeb200_weld_interaction_offsets_dev": (C.c_int, [_vp, _vp, C.c_size_t, _vp, C.c_size_t, _vp, C.c_size_t]),
    "lifeb200_weld_interaction_offsets_host": (C.c_int, [_vp, _vp, C.c_size_t, _vp, C.c_size_t, _vp, C.c_size_t]),
    "lifeb200_rle": (C.c_int, [_vp, C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t)]),
    "lifeb200_history_rle": (C.c_int, [_vp, C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t)]),
    "lifeb200_parse_rle": (C.c_int, [C.c_char_p, _vp]),
    "lifeb200_history_parse_rle": (C.c_int, [C.c_char_p, _vp]),
}

MATCH_LIVE_AND_DEAD, MATCH_LIVE, MATCH_BOUNDARY = 0, 1, 2


def declared_functions(header_path: str = HEADER_PATH) -> list[str]:
    """Names of every function the public header declares."""
    text = open(header_path).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(lifeb200_[a-z0-9_]+)\s*\(", text)))


_lib = None


def load_library(path: str = LIB_PATH) -> C.CDLL:
    """dlopen the in-tree library; raises OSError if it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(path):
            raise OSError(f"{path} not built: run `python -m lifeapi_b200.build` (needs nvcc)")
        lib = C.CDLL(path)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def _ptr(a) -> int:
    """Raw address of a numpy array, int (device pointer) or None."""
    if a is None:
        return 0
    if isinstance(a, (int, np.integer)):
        return int(a)
    if isinstance(a, np.ndarray):
        if not a.flags["C_CONTIGUOUS"]:
            raise ValueError("array must be C-contiguous")
        return a.ctypes.data
    if hasattr(a, "data_ptr"):  # torch tensor (device or host)
        return int(a.data_ptr())
    raise TypeError(type(a))


class Engine:
    """One context == one GPU + one stream (lifeb200_create)."""

    def __init__(self, device: int = 0, _borrowed=None):
        self.lib = load_library()
        self._h = None
        self._owned = _borrowed is None
        if _borrowed is not None:  # a context owned by a MultiEngine (lifeb200_multi_ctx)
            self._h = _borrowed
        else:
            h = _vp()
            self._check(self.lib.lifeb200_create(device, C.byref(h)))
            self._h = h
        self.device = device

    def _check(self, rc: int) -> None:
        if rc != 0:
            raise LifeB200Error(rc, self.lib.lifeb200_last_error().decode())

    def close(self) -> None:
        if self._h is not None:
            if self._owned:
                self.lib.lifeb200_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- management --------------------------------------------------------------------------
    def set_stream(self, cuda_stream: int | None) -> None:
        self._check(self.lib.lifeb200_set_stream(self._h, cuda_stream or None))

    def sync(self) -> None:
        self._check(self.lib.lifeb200_sync(self._h))

    def set_lanes_per_pattern(self, lpp: int) -> None:
        self._check(self.lib.lifeb200_set_lanes_per_pattern(self._h, lpp))

    def set_step_layout(self, layout: int) -> None:
        self._check(self.lib.lifeb200_set_step_layout(self._h, layout))

    def set_three_valued_path(self, path: int) -> None:
        self._check(self.lib.lifeb200_set_three_valued_path(self._h, path))

    def device_info(self) -> dict:
        sm, ma, mi, khz, mem = C.c_int(), C.c_int(), C.c_int(), C.c_int(), C.c_size_t()
        self._check(self.lib.lifeb200_device_info(self._h, C.byref(sm), C.byref(ma), C.byref(mi), C.byref(khz), C.byref(mem)))
        return {"sm_count": sm.value, "cc": (ma.value, mi.value), "clock_khz": khz.value, "total_mem": mem.value}

    def launch_count(self) -> int:
        v = C.c_uint64()
        self._check(self.lib.lifeb200_launch_count(self._h, C.byref(v)))
        return v.value

    def timer_start(self) -> None:
        self._check(self.lib.lifeb200_timer_start(self._h))

    def timer_stop(self) -> float:
        ms = C.c_float()
        self._check(self.lib.lifeb200_timer_stop(self._h, C.byref(ms)))
        return ms.value

    def measure_int_peak(self) -> tuple[float, float]:
        r, mhz = C.c_double(), C.c_double()
        self._check(self.lib.lifeb200_measure_int_peak(self._h, C.byref(r), C.byref(mhz)))
        return r.value, mhz.value

    # -- memory ------------------------------------------------------------------------------
    def malloc(self, nbytes: int) -> int:
        p = _vp()
        self._check(self.lib.lifeb200_malloc(self._h, C.byref(p), nbytes))
        return p.value or 0

    def free(self, dptr: int) -> None:
        self._check(self.lib.lifeb200_free(self._h, dptr))

    def host_alloc(self, shape, dtype=np.uint64) -> np.ndarray:
        """Pinned host array (lifeb200_host_alloc); freed with host_free(arr)."""
        dtype = np.dtype(dtype)
        n = int(np.prod(shape)) * dtype.itemsize
        p = _vp()
        self._check(self.lib.lifeb200_host_alloc(self._h, C.byref(p), max(n, 1)))
        buf = (C.c_char * max(n, 1)).from_address(p.value)
        arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
        arr.flags.writeable = True
        return arr

    def host_free(self, arr: np.ndarray) -> None:
        self._check(self.lib.lifeb200_host_free(self._h, arr.ctypes.data))

    def host_register(self, arr: np.ndarray) -> None:
        """Page-lock a caller-owned numpy array (lifeb200_host_register); pair with host_unregister(arr)."""
        self._check(self.lib.lifeb200_host_register(self._h, arr.ctypes.data, arr.nbytes))

    def host_unregister(self, arr: np.ndarray) -> None:
        self._check(self.lib.lifeb200_host_unregister(self._h, arr.ctypes.data))

    def h2d(self, dptr: int, src: np.ndarray) -> None:
        self._check(self.lib.lifeb200_memcpy_h2d(self._h, dptr, _ptr(src), src.nbytes))

    def d2h(self, dst: np.ndarray, dptr: int) -> None:
        self._check(self.lib.lifeb200_memcpy_d2h(self._h, _ptr(dst), dptr, dst.nbytes))

    def memset(self, dptr: int, value: int, nbytes: int) -> None:
        self._check(self.lib.lifeb200_memset(self._h, dptr, value, nbytes))

    # -- compute -----------------------------------------------------------------------------
    def step_batch_dev(self, d_in, d_out, n: int, gens: int) -> None:
        self._check(self.lib.lifeb200_step_batch_dev(self._h, _ptr(d_in), _ptr(d_out), n, gens))

    def step_batch_host(self, src: np.ndarray, gens: int, out: np.ndarray | None = None) -> np.ndarray:
        src = np.ascontiguousarray(src, dtype=np.uint64)
        n = src.size // WORDS
        if out is None:
            out = np.empty_like(src)
        self._check(self.lib.lifeb200_step_batch_host(self._h, _ptr(src), _ptr(out), n, gens))
        return out

    def soup_dev(self, d_out, first_id: int, n: int, seed: int, dist: int = 0) -> None:
        self._check(self.lib.lifeb200_soup_dev(self._h, _ptr(d_out), first_id, n, seed, dist))

    def neighbour_count_dev(self, d_in, d_planes, n: int) -> None:
        self._check(self.lib.lifeb200_neighbour_count_dev(self._h, _ptr(d_in), _ptr(d_planes), n))

    def neighbour_count_host(self, src: np.ndarray) -> np.ndarray:
        src = np.ascontiguousarray(src, dtype=np.uint64)
        n = src.size // WORDS
        out = np.empty((n, 4, WORDS), dtype=np.uint64)
        self._check(self.lib.lifeb200_neighbour_count_host(self._h, _ptr(src), _ptr(out), n))
        return out

    def step_match_dev(self, d_patterns, d_targets, d_target_index, n, first_id, gens, stable_interval, d_hits, cap,
                       d_n_hits) -> None:
        self._check(self.lib.lifeb200_step_match_dev(self._h, _ptr(d_patterns), _ptr(d_targets), _ptr(d_target_index), n,
                                                     first_id, gens, stable_interval, _ptr(d_hits), cap, _ptr(d_n_hits)))

    def step_match_static_dev(self, d_patterns, d_targets, d_target_index, n, first_id, gens, stable_interval, d_hits, cap,
                              d_n_hits) -> None:
        """lifeb200_step_match_static_dev: the same call on the library's static device scratch (no allocation)."""
        self._check(self.lib.lifeb200_step_match_static_dev(self._h, _ptr(d_patterns), _ptr(d_targets), _ptr(d_target_index),
                                                            n, first_id, gens, stable_interval, _ptr(d_hits), cap,
                                                            _ptr(d_n_hits)))

    def step_match(self, patterns: np.ndarray, targets: np.ndarray, target_index: np.ndarray | None, gens: int,
                   stable_interval: int, first_id: int = 0, cap: int | None = None) -> np.ndarray:
        """Convenience: host arrays in, hit records (in id order, as the device leaves them) out; *_dev entry points only."""
        patterns = np.ascontiguousarray(patterns, dtype=np.uint64)
        targets = np.ascontiguousarray(targets, dtype=np.uint64)
        n = patterns.size // WORDS
        cap = n if cap is None else cap
        d_p = self.malloc(max(patterns.nbytes, 16))
        d_t = self.malloc(max(targets.nbytes, 16))
        d_i = 0
        d_h = self.malloc(max(cap, 1) * 16)
        d_n = self.malloc(8)
        try:
            self.h2d(d_p, patterns)
            self.h2d(d_t, targets)
            if target_index is not None:
                target_index = np.ascontiguousarray(target_index, dtype=np.uint32)
                d_i = self.malloc(max(target_index.nbytes, 16))
                self.h2d(d_i, target_index)
            self.step_match_dev(d_p, d_t, d_i or None, n, first_id, gens, stable_interval, d_h, cap, d_n)
            cnt = np.zeros(1, dtype=np.uint64)
            self.d2h(cnt, d_n)
            self.sync()
            found = int(cnt[0])
            if found > cap:
                raise LifeB200Error(E_OVERFLOW, f"{found} hits, capacity {cap}")
            hits = np.zeros(found, dtype=HIT_DTYPE)
            if found:
                self.d2h(hits, d_h)
                self.sync()
            return hits  # already in id order: the device compaction is ordered (csrc/match.cu)
        finally:
            for p in (d_p, d_t, d_i, d_h, d_n):
                if p:
                    self.free(p)

    def enumerate_collisions(self, catalysts: np.ndarray, projectiles: np.ndarray, id_begin: int, id_end: int, gens: int,
                             stable_interval: int, cap: int) -> tuple[np.ndarray, int]:
        catalysts = np.ascontiguousarray(catalysts, dtype=np.uint64)
        projectiles = np.ascontiguousarray(projectiles, dtype=np.uint64)
        hits = np.zeros(max(cap, 1), dtype=HIT_DTYPE)
        nh, ns = C.c_uint64(), C.c_uint64()
        self._check(self.lib.lifeb200_enumerate_collisions(
            self._h, _ptr(catalysts), catalysts.size // WORDS, _ptr(projectiles), projectiles.size // WORDS, id_begin,
            id_end, gens, stable_interval, _ptr(hits), cap, C.byref(nh), C.byref(ns)))
        return hits[: nh.value].copy(), ns.value

    def soup_survey(self, first_id: int, n: int, seed: int, dist: int, gens: int, cap: int):
        hits = np.zeros(max(cap, 1), dtype=HIT_DTYPE)
        nh, ps, hx = C.c_uint64(), C.c_uint64(), C.c_uint64()
        self._check(self.lib.lifeb200_soup_survey(self._h, first_id, n, seed, dist, gens, _ptr(hits), cap, C.byref(nh),
                                                  C.byref(ps), C.byref(hx)))
        return hits[: nh.value].copy(), ps.value, hx.value

    def soup_survey_dev(self, first_id: int, n: int, seed: int, dist: int, gens: int, d_hits, cap: int):
        """Hits stay on the device (sorted by id) in d_hits; returns (n_hits, pop_sum, hash_xor)."""
        nh, ps, hx = C.c_uint64(), C.c_uint64(), C.c_uint64()
        self._check(self.lib.lifeb200_soup_survey_dev(self._h, first_id, n, seed, dist, gens, _ptr(d_hits), cap, C.byref(nh),
                                                      C.byref(ps), C.byref(hx)))
        return nh.value, ps.value, hx.value

    def enumerate_collisions_dev(self, catalysts: np.ndarray, projectiles: np.ndarray, id_begin: int, id_end: int, gens: int,
                                 stable_interval: int, d_hits, cap: int) -> tuple[int, int]:
        catalysts = np.ascontiguousarray(catalysts, dtype=np.uint64)
        projectiles = np.ascontiguousarray(projectiles, dtype=np.uint64)
        nh, ns = C.c_uint64(), C.c_uint64()
        self._check(self.lib.lifeb200_enumerate_collisions_dev(
            self._h, _ptr(catalysts), catalysts.size // WORDS, _ptr(projectiles), projectiles.size // WORDS, id_begin,
            id_end, gens, stable_interval, _ptr(d_hits), cap, C.byref(nh), C.byref(ns)))
        return nh.value, ns.value

    def fold_dev(self, d_patterns, n: int, first_id: int = 0) -> tuple[int, int]:
        ps, hx = C.c_uint64(), C.c_uint64()
        self._check(self.lib.lifeb200_fold_dev(self._h, _ptr(d_patterns), n, first_id, C.byref(ps), C.byref(hx)))
        return ps.value, hx.value

    # -- generated circuits: three-valued step, LifeHistory, LifeStable passes -----------------------
    def unknown_step_batch_dev(self, d_state_in, d_unknown_in, d_state_out, d_unknown_out, n: int, gens: int) -> None:
        self._check(self.lib.lifeb200_unknown_step_batch_dev(self._h, _ptr(d_state_in), _ptr(d_unknown_in), _ptr(d_state_out),
                                                             _ptr(d_unknown_out), n, gens))

    def unknown_step_batch_host(self, state: np.ndarray, unknown: np.ndarray, gens: int):
        state = np.ascontiguousarray(state, dtype=np.uint64)
        unknown = np.ascontiguousarray(unknown, dtype=np.uint64)
        if state.shape != unknown.shape:
            raise ValueError("state / unknown shapes differ")
        so, uo = np.empty_like(state), np.empty_like(unknown)
        self._check(self.lib.lifeb200_unknown_step_batch_host(self._h, _ptr(state), _ptr(unknown), _ptr(so), _ptr(uo),
                                                              state.size // WORDS, gens))
        return so, uo

    def history_step_batch_dev(self, d_records, n: int, gens: int) -> None:
        self._check(self.lib.lifeb200_history_step_batch_dev(self._h, _ptr(d_records), n, gens))

    def history_step_batch(self, records: np.ndarray, gens: int) -> np.ndarray:
        """Convenience: (n, 4, 64) LifeHistory records, host in / host out via the *_dev entry point."""
        records = np.ascontiguousarray(records, dtype=np.uint64)
        n = records.size // (4 * WORDS)
        out = np.empty_like(records)
        d = self.malloc(max(records.nbytes, 16))
        try:
            self.h2d(d, records)
            self.history_step_batch_dev(d, n, gens)
            self.d2h(out, d)
            self.sync()
        finally:
            self.free(d)
        return out

    def stable_pass_dev(self, d_records, n: int, which: int, d_flags=None) -> None:
        self._check(self.lib.lifeb200_stable_pass_dev(self._h, _ptr(d_records), n, which, _ptr(d_flags)))

    def stable_pass_host(self, records: np.ndarray, which: int):
        """(n, 10, 64) LifeStable records -> (updated records, flags uint8)."""
        rec = np.ascontiguousarray(records, dtype=np.uint64).copy()
        n = rec.size // STABLE_WORDS
        flags = np.zeros(n, dtype=np.uint8)
        self._check(self.lib.lifeb200_stable_pass_host(self._h, _ptr(rec), n, which, _ptr(flags)))
        return rec, flags

    def stable_vulnerable(self, records: np.ndarray) -> np.ndarray:
        records = np.ascontiguousarray(records, dtype=np.uint64)
        n = records.size // STABLE_WORDS
        out = np.empty((n, WORDS), dtype=np.uint64)
        d, do = self.malloc(max(records.nbytes, 16)), self.malloc(max(out.nbytes, 16))
        try:
            self.h2d(d, records)
            self._check(self.lib.lifeb200_stable_vulnerable_dev(self._h, d, do, n))
            self.d2h(out, do)
            self.sync()
        finally:
            self.free(d)
            self.free(do)
        return out

    def unknown_step_refined(self, records: np.ndarray, current_on: np.ndarray, current_unknown: np.ndarray, variant: int = 0):
        records = np.ascontiguousarray(records, dtype=np.uint64)
        current_on = np.ascontiguousarray(current_on, dtype=np.uint64)
        current_unknown = np.ascontiguousarray(current_unknown, dtype=np.uint64)
        n = records.size // STABLE_WORDS
        outs = [np.empty((n, WORDS), dtype=np.uint64) for _ in range(3)]
        bufs = [self.malloc(max(a.nbytes, 16)) for a in (records, current_on, current_unknown, *outs)]
        try:
            for b, a in zip(bufs[:3], (records, current_on, current_unknown)):
                self.h2d(b, a)
            self._check(self.lib.lifeb200_unknown_step_refined_dev(self._h, bufs[0], bufs[1], bufs[2], n, variant, bufs[3],
                                                                   bufs[4], bufs[5]))
            for b, a in zip(bufs[3:], outs):
                self.d2h(a, b)
            self.sync()
        finally:
            for b in bufs:
                self.free(b)
        return tuple(outs)

    # -- 8f next rows: LifeWeld::Step, InteractionCountsAndNext, GetHash -------------------------------
    def _roundtrip(self, src: np.ndarray, out_shape, out_dtype, call) -> np.ndarray:
        src = np.ascontiguousarray(src)
        out = np.empty(out_shape, dtype=out_dtype)
        d_in, d_out = self.malloc(max(src.nbytes, 16)), self.malloc(max(out.nbytes, 16))
        try:
            self.h2d(d_in, src)
            call(d_in, d_out)
            if out.nbytes:
                self.d2h(out, d_out)
            self.sync()
        finally:
            self.free(d_in)
            self.free(d_out)
        return out

    def weld_step_batch(self, records: np.ndarray, gens: int) -> np.ndarray:
        """(n, 4, 64) LifeWeld records (state, frozen2, frozen1, frozen0) stepped `gens` generations."""
        records = np.ascontiguousarray(records, dtype=np.uint64)
        n = records.size // (4 * WORDS)
        out = np.empty_like(records)
        d = self.malloc(max(records.nbytes, 16))
        try:
            self.h2d(d, records)
            self._check(self.lib.lifeb200_weld_step_batch_dev(self._h, d, n, gens))
            self.d2h(out, d)
            self.sync()
        finally:
            self.free(d)
        return out

    def interaction_counts(self, patterns: np.ndarray) -> np.ndarray:
        patterns = np.ascontiguousarray(patterns, dtype=np.uint64)
        n = patterns.size // WORDS
        return self._roundtrip(patterns, (n, 4, WORDS), np.uint64,
                               lambda di, do: self._check(self.lib.lifeb200_interaction_counts_dev(self._h, di, do, n)))

    def weld_step_batch_host(self, records: np.ndarray, gens: int) -> np.ndarray:
        rec = np.ascontiguousarray(records, dtype=np.uint64).copy()
        self._check(self.lib.lifeb200_weld_step_batch_host(self._h, _ptr(rec), rec.size // (4 * WORDS), gens))
        return rec

    def history_step_batch_host(self, records: np.ndarray, gens: int) -> np.ndarray:
        rec = np.ascontiguousarray(records, dtype=np.uint64).copy()
        self._check(self.lib.lifeb200_history_step_batch_host(self._h, _ptr(rec), rec.size // (4 * WORDS), gens))
        return rec

    def interaction_counts_host(self, patterns: np.ndarray) -> np.ndarray:
        patterns = np.ascontiguousarray(patterns, dtype=np.uint64)
        n = patterns.size // WORDS
        out = np.empty((n, 4, WORDS), dtype=np.uint64)
        self._check(self.lib.lifeb200_interaction_counts_host(self._h, _ptr(patterns), _ptr(out), n))
        return out

    def stable_vulnerable_host(self, records: np.ndarray) -> np.ndarray:
        records = np.ascontiguousarray(records, dtype=np.uint64)
        n = records.size // STABLE_WORDS
        out = np.empty((n, WORDS), dtype=np.uint64)
        self._check(self.lib.lifeb200_stable_vulnerable_host(self._h, _ptr(records), _ptr(out), n))
        return out

    def unknown_step_refined_host(self, records: np.ndarray, current_on: np.ndarray, current_unknown: np.ndarray, variant: int = 0):
        records = np.ascontiguousarray(records, dtype=np.uint64)
        current_on = np.ascontiguousarray(current_on, dtype=np.uint64)
        current_unknown = np.ascontiguousarray(current_unknown, dtype=np.uint64)
        n = records.size // STABLE_WORDS
        outs = [np.empty((n, WORDS), dtype=np.uint64) for _ in range(3)]
        self._check(self.lib.lifeb200_unknown_step_refined_host(self._h, _ptr(records), _ptr(current_on), _ptr(current_unknown), n,
                                                                variant, _ptr(outs[0]), _ptr(outs[1]), _ptr(outs[2])))
        return tuple(outs)

    def step_match_host(self, patterns: np.ndarray, targets: np.ndarray, target_index: np.ndarray | None, gens: int,
                        stable_interval: int, first_id: int = 0, cap: int | None = None) -> np.ndarray:
        patterns = np.ascontiguousarray(patterns, dtype=np.uint64)
        targets = np.ascontiguousarray(targets, dtype=np.uint64)
        n, nt = patterns.size // WORDS, targets.size // (2 * WORDS)
        cap = n if cap is None else cap
        if target_index is not None:
            target_index = np.ascontiguousarray(target_index, dtype=np.uint32)
        hits = np.zeros(max(cap, 1), dtype=HIT_DTYPE)
        nh = C.c_uint64()
        self._check(self.lib.lifeb200_step_match_host(self._h, _ptr(patterns), _ptr(targets), _ptr(target_index), n, nt, first_id,
                                                      gens, stable_interval, _ptr(hits), cap, C.byref(nh)))
        return hits[: nh.value].copy()

    def contains_target_batch_host(self, states: np.ndarray, targets: np.ndarray, dx: int, dy: int) -> np.ndarray:
        """LifeState::Contains(const LifeTarget&, dx, dy) for every state; targets: one (2, 64) record or one per state."""
        (states, n), (targets, nt) = self._planes(states), self._planes(targets, 2 * WORDS)
        out = np.zeros(n, dtype=np.uint8)
        self._check(self.lib.lifeb200_contains_target_batch_host(self._h, _ptr(states), n, _ptr(targets), nt, dx, dy, _ptr(out)))
        return out.astype(bool)

    def contains_target_batch_dev(self, d_states, n: int, d_targets, n_targets: int, dx: int, dy: int, d_out) -> None:
        self._check(self.lib.lifeb200_contains_target_batch_dev(self._h, _ptr(d_states), n, _ptr(d_targets), n_targets, dx, dy, _ptr(d_out)))

    def weld_interaction_counts_host(self, records: np.ndarray) -> np.ndarray:
        records = np.ascontiguousarray(records, dtype=np.uint64)
        n = records.size // (4 * WORDS)
        out = np.empty((n, 3, WORDS), dtype=np.uint64)
        self._check(self.lib.lifeb200_weld_interaction_counts_host(self._h, _ptr(records), _ptr(out), n))
        return out

    def nc_arith(self, op: str, a: np.ndarray, b: np.ndarray | None = None, incarry: np.ndarray | None = None, k: int = 0) -> np.ndarray:
        """NeighbourCount batch arithmetic through the *_dev entry points: op in add / subtract / with_exactly."""
        a = np.ascontiguousarray(a, dtype=np.uint64)
        n = a.size // (4 * WORDS)
        out = np.empty((n, WORDS) if op == "with_exactly" else (n, 4, WORDS), dtype=np.uint64)
        bufs = {"a": self.malloc(max(a.nbytes, 16)), "o": self.malloc(max(out.nbytes, 16))}
        try:
            self.h2d(bufs["a"], a)
            for name, arr in (("b", b), ("c", incarry)):
                if arr is not None:
                    arr = np.ascontiguousarray(arr, dtype=np.uint64)
                    bufs[name] = self.malloc(max(arr.nbytes, 16))
                    self.h2d(bufs[name], arr)
            if op == "add":
                self._check(self.lib.lifeb200_nc_add_dev(self._h, bufs["a"], bufs["b"], bufs.get("c"), bufs["o"], n))
            elif op == "subtract":
                self._check(self.lib.lifeb200_nc_subtract_dev(self._h, bufs["a"], bufs["b"], bufs["o"], n))
            elif op == "with_exactly":
                self._check(self.lib.lifeb200_nc_with_exactly_dev(self._h, bufs["a"], k, bufs["o"], n))
            else:
                raise ValueError(op)
            if out.nbytes:
                self.d2h(out, bufs["o"])
            self.sync()
        finally:
            for p in bufs.values():
                self.free(p)
        return out

    def hash_batch_dev(self, d_patterns, d_hashes, n: int) -> None:
        self._check(self.lib.lifeb200_hash_batch_dev(self._h, _ptr(d_patterns), _ptr(d_hashes), n))

    def hash_batch_host(self, patterns: np.ndarray) -> np.ndarray:
        patterns = np.ascontiguousarray(patterns, dtype=np.uint64)
        n = patterns.size // WORDS
        out = np.empty(n, dtype=np.uint64)
        self._check(self.lib.lifeb200_hash_batch_host(self._h, _ptr(patterns), _ptr(out), n))
        return out

    # -- 8f rank 4: Convolve / Match / InteractionOffsets (host arrays through the *_host entry points) ----
    @staticmethod
    def _planes(a, words=WORDS):
        a = np.ascontiguousarray(a, dtype=np.uint64).reshape(-1, words)
        return a, a.shape[0]

    def convolve_batch_dev(self, d_a, n_a: int, d_b, n_b: int, d_out, n: int) -> None:
        self._check(self.lib.lifeb200_convolve_batch_dev(self._h, _ptr(d_a), n_a, _ptr(d_b), n_b, _ptr(d_out), n))

    def convolve_batch_host(self, a: np.ndarray, b: np.ndarray) -> np.ndarray:
        """a, b: (n or 1, 64) planes -> (n, 64); an operand of one plane is shared by all problems."""
        (a, na), (b, nb) = self._planes(a), self._planes(b)
        n = max(na, nb)
        out = np.empty((n, WORDS), dtype=np.uint64)
        self._check(self.lib.lifeb200_convolve_batch_host(self._h, _ptr(a), na, _ptr(b), nb, _ptr(out), n))
        return out

    def match_batch_dev(self, d_states, n: int, d_live, d_dead, n_t: int, mode: int, d_out) -> None:
        self._check(self.lib.lifeb200_match_batch_dev(self._h, _ptr(d_states), n, _ptr(d_live), _ptr(d_dead), n_t, mode, _ptr(d_out)))

    def match_batch_host(self, states: np.ndarray, live: np.ndarray, dead: np.ndarray | None = None,
                         mode: int = MATCH_LIVE_AND_DEAD) -> np.ndarray:
        (states, n), (live, nt) = self._planes(states), self._planes(live)
        if dead is not None:
            dead, nd = self._planes(dead)
            if nd != nt:
                raise ValueError("live / dead counts differ")
        out = np.empty((n, WORDS), dtype=np.uint64)
        self._check(self.lib.lifeb200_match_batch_host(self._h, _ptr(states), n, _ptr(live), _ptr(dead), nt, mode, _ptr(out)))
        return out

    def match_target_batch_dev(self, d_states, n: int, d_targets, n_targets: int, d_out) -> None:
        self._check(self.lib.lifeb200_match_target_batch_dev(self._h, _ptr(d_states), n, _ptr(d_targets), n_targets, _ptr(d_out)))

    def match_target_batch_host(self, states: np.ndarray, targets: np.ndarray) -> np.ndarray:
        (states, n), (targets, nt) = self._planes(states), self._planes(targets, 2 * WORDS)
        out = np.empty((n, WORDS), dtype=np.uint64)
        self._check(self.lib.lifeb200_match_target_batch_host(self._h, _ptr(states), n, _ptr(targets), nt, _ptr(out)))
        return out

    def weld_interaction_offsets_host(self, a: np.ndarray, b: np.ndarray) -> np.ndarray:
        """a, b: (n or 1, 4, 64) LifeWeld records -> (n, 64) offsets plane."""
        (a, na), (b, nb) = self._planes(a, 4 * WORDS), self._planes(b, 4 * WORDS)
        n = max(na, nb)
        out = np.empty((n, WORDS), dtype=np.uint64)
        self._check(self.lib.lifeb200_weld_interaction_offsets_host(self._h, _ptr(a), na, _ptr(b), nb, _ptr(out), n))
        return out

    def interaction_offsets_dev(self, d_a, n_a: int, d_b, n_b: int, d_out, n: int) -> None:
        self._check(self.lib.lifeb200_interaction_offsets_dev(self._h, _ptr(d_a), n_a, _ptr(d_b), n_b, _ptr(d_out), n))

    def interaction_offsets_host(self, a: np.ndarray, b: np.ndarray) -> np.ndarray:
        (a, na), (b, nb) = self._planes(a), self._planes(b)
        n = max(na, nb)
        out = np.empty((n, WORDS), dtype=np.uint64)
        self._check(self.lib.lifeb200_interaction_offsets_host(self._h, _ptr(a), na, _ptr(b), nb, _ptr(out), n))
        return out


def _prefer_bundled_nccl() -> None:
    """If this interpreter has a pip-bundled NCCL (nvidia-nccl-cu12, what torch links), point the library at that file
    unless the caller chose one: a process must not end up with two different libnccl.so.2 (the first one loaded wins
    the soname, and a later `import torch` against an older system NCCL fails on missing symbols).  No torch import."""
    import sys
    if os.environ.get("LIFEB200_NCCL_LIB"):
        return
    for base in sys.path:
        cand = os.path.join(base, "nvidia", "nccl", "lib", "libnccl.so.2")
        if os.path.exists(cand):
            os.environ["LIFEB200_NCCL_LIB"] = cand
            return


class MultiEngine:
    """A multi-GPU group (lifeb200_multi_*): every rank in this process (devices=[...]) or one rank per process
    (rank / world / unique_id, the id made by MultiEngine.unique_id() on rank 0 and distributed by the caller)."""

    def __init__(self, devices=None, *, device: int | None = None, rank: int | None = None, world: int | None = None,
                 unique_id: bytes | None = None):
        self.lib = load_library()
        self._h = None
        _prefer_bundled_nccl()
        h = _vp()
        if rank is None:
            arr = None
            n = 0
            if devices is not None:
                devices = list(devices)
                n = len(devices)
                arr = (C.c_int * n)(*devices)
            self._check(self.lib.lifeb200_multi_create(arr, n, C.byref(h)))
        else:
            self._check(self.lib.lifeb200_multi_create_rank(device if device is not None else 0, rank, world, unique_id, C.byref(h)))
        self._h = h

    @staticmethod
    def unique_id() -> bytes:
        lib = load_library()
        _prefer_bundled_nccl()
        buf = C.create_string_buffer(UNIQUE_ID_BYTES)
        _lib_check(lib, lib.lifeb200_multi_unique_id(buf))
        return buf.raw

    def _check(self, rc: int) -> None:
        if rc != 0:
            raise LifeB200Error(rc, self.lib.lifeb200_last_error().decode())

    def close(self) -> None:
        if self._h is not None:
            self.lib.lifeb200_multi_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def info(self) -> dict:
        w, f, k, v = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        self._check(self.lib.lifeb200_multi_info(self._h, C.byref(w), C.byref(f), C.byref(k), C.byref(v)))
        return {"world": w.value, "first_rank": f.value, "n_local": k.value, "nccl_version": v.value}

    def engine(self, local_index: int = 0) -> "Engine":
        h = _vp()
        self._check(self.lib.lifeb200_multi_ctx(self._h, local_index, C.byref(h)))
        return Engine(local_index, _borrowed=h)

    def sync(self) -> None:
        self._check(self.lib.lifeb200_multi_sync(self._h))

    def shard(self, first_id: int, n: int, rank: int) -> tuple[int, int]:
        b, e = C.c_uint64(), C.c_uint64()
        self._check(self.lib.lifeb200_multi_shard(self._h, first_id, n, rank, C.byref(b), C.byref(e)))
        return b.value, e.value

    def reserve(self, n: int, cap: int) -> None:
        self._check(self.lib.lifeb200_multi_reserve(self._h, n, cap))

    def set_gather(self, mode: int) -> None:
        self._check(self.lib.lifeb200_multi_set_gather(self._h, mode))

    def last_gather_ms(self) -> float:
        ms = C.c_float()
        self._check(self.lib.lifeb200_multi_last_gather_ms(self._h, C.byref(ms)))
        return ms.value

    def soup_survey(self, first_id: int, n: int, seed: int, dist: int, gens: int, cap: int, hits: np.ndarray | None = None):
        if hits is None:
            hits = np.zeros(max(cap, 1), dtype=HIT_DTYPE)
        nh, ps, hx = C.c_uint64(), C.c_uint64(), C.c_uint64()
        self._check(self.lib.lifeb200_multi_soup_survey(self._h, first_id, n, seed, dist, gens, _ptr(hits), cap, C.byref(nh),
                                                        C.byref(ps), C.byref(hx)))
        return hits[: nh.value], ps.value, hx.value

    def enumerate_collisions(self, catalysts: np.ndarray, projectiles: np.ndarray, id_begin: int, id_end: int, gens: int,
                             stable_interval: int, cap: int):
        catalysts = np.ascontiguousarray(catalysts, dtype=np.uint64)
        projectiles = np.ascontiguousarray(projectiles, dtype=np.uint64)
        hits = np.zeros(max(cap, 1), dtype=HIT_DTYPE)
        nh, ns = C.c_uint64(), C.c_uint64()
        self._check(self.lib.lifeb200_multi_enumerate_collisions(
            self._h, _ptr(catalysts), catalysts.size // WORDS, _ptr(projectiles), projectiles.size // WORDS, id_begin,
            id_end, gens, stable_interval, _ptr(hits), cap, C.byref(nh), C.byref(ns)))
        return hits[: nh.value].copy(), ns.value

    def step_batch_host(self, src: np.ndarray, gens: int, out: np.ndarray | None = None) -> np.ndarray:
        src = np.ascontiguousarray(src, dtype=np.uint64)
        if out is None:
            out = np.empty_like(src)
        self._check(self.lib.lifeb200_multi_step_batch_host(self._h, _ptr(src), _ptr(out), src.size // WORDS, gens))
        return out


# -- 8f rank 4: RLE in and out (host functions of the library; no context, no GPU) --------------------
def _lib_check(lib, rc: int) -> None:
    if rc != 0:
        raise LifeB200Error(rc, lib.lifeb200_last_error().decode())


def _rle_call(fn_name: str, planes: np.ndarray) -> str:
    lib = load_library()
    planes = np.ascontiguousarray(planes, dtype=np.uint64)
    cap = 1 << 12
    while True:
        buf = C.create_string_buffer(cap)
        n = C.c_size_t()
        rc = getattr(lib, fn_name)(_ptr(planes), buf, cap, C.byref(n))
        if rc == E_OVERFLOW:
            cap = n.value + 1
            continue
        _lib_check(lib, rc)
        return buf.value.decode()


def rle(state: np.ndarray) -> str:
    """LifeState::RLE()."""
    return _rle_call("lifeb200_rle", state)


def history_rle(record: np.ndarray) -> str:
    """LifeHistory::RLE() of a (4, 64) record: state, history, marked, original."""
    return _rle_call("lifeb200_history_rle", record)


def parse_rle(text: str) -> np.ndarray:
    """LifeState::Parse()."""
    lib = load_library()
    out = np.zeros(WORDS, dtype=np.uint64)
    _lib_check(lib, lib.lifeb200_parse_rle(text.encode(), _ptr(out)))
    return out


def history_parse_rle(text: str) -> np.ndarray:
    """LifeHistory::Parse() -> (4, 64) record."""
    lib = load_library()
    out = np.zeros((4, WORDS), dtype=np.uint64)
    _lib_check(lib, lib.lifeb200_history_parse_rle(text.encode(), _ptr(out)))
    return out
