"""lifeapi_b200 -- B200-native batched stepping engine behind LifeAPI's header API.

The product is the C-ABI shared library built from csrc/ (include/lifeapi_b200.h) plus the C++
drop-in header include/LifeAPI_b200.hpp.  This Python package is only the ctypes binding used by
the tests and bench.py, and the rank-sharding / hit-gather plumbing.  Nothing here computes Life
on the CPU, and nothing here imports oracle/.
"""
from .capi import (E_ARG, E_CUDA, E_NCCL, E_NODEVICE, E_OK, E_OVERFLOW, HIT_DTYPE, MATCH_BOUNDARY, MATCH_LIVE, MATCH_LIVE_AND_DEAD,
                   PATTERN_BYTES, WORDS, Engine, LifeB200Error, MultiEngine, declared_functions, history_parse_rle, history_rle,
                   load_library, parse_rle, rle)
from .shard import bind_host_to_device, gather_hits, merge_hits, shard_range

__all__ = ["Engine", "MultiEngine", "E_NCCL", "LifeB200Error", "HIT_DTYPE", "WORDS", "PATTERN_BYTES", "load_library", "declared_functions",
           "shard_range", "merge_hits", "gather_hits", "bind_host_to_device", "rle", "history_rle", "parse_rle", "history_parse_rle",
           "MATCH_LIVE_AND_DEAD", "MATCH_LIVE", "MATCH_BOUNDARY", "E_OK", "E_CUDA", "E_ARG", "E_OVERFLOW", "E_NODEVICE"]
