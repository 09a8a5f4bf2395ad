#!/usr/bin/env python
"""bench.py -- cell-updates/s of batched 64x64-torus Life stepping (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one pass of the hot path over one batch: BASELINE config 2, 1 Mi seeded random soups
stepped 1024 generations in registers (per GPU; weak scaling: every rank owns its own 1 Mi ids).
  value     Gcell/s, inputs resident in HBM, CUDA events on the launching stream, max over ranks
  e2e       same metric through lifeb200_step_batch_host with PINNED HOST buffers (H2D + D2H timed)
  roofline  integer (ALU) pipe: achieved lane-instr/s = cells/s x 0.375 (2 SHF + 10 LOP3
            per 32 cells per generation) vs the LOP3 rate measured in the same run; plus the
            HBM-bound G=1 launch vs MEASURED_PEAKS.json
  cpu_baseline  the reference's own LifeState::Step(n) (oracle/_ref) on all host cores, bounded sample
torch is used for torch.distributed (NCCL barrier / max / hit-counter gather) only.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SEED = 0x4C696665415049
N_SOUPS = 1 << 20
GENS = 1024
INSTR_PER_CELL = 12.0 / 32.0  # algorithmic 2 SHF + 10 LOP3 per 32 cells per generation
# What step_kernel_x<8> EXECUTES per 32 cells and generation is read from the committed SASS histogram of the shipped
# library (profiles/r02_sass_mix.json, made by tools/sass_mix.py from `cuobjdump -sass`); 7.75 is the fallback.
# DRAM traffic of one launch is read from the committed `ncu --set full` capture of this launch shape.
SASS_MIX_FILES = ["profiles/r02_sass_mix.json"]
NCU_RAW_FILES = ["profiles/r02v_step_kernel_x8_raw.csv", "profiles/r02_step_kernel_x8_raw.csv", "profiles/r01z_step_kernel_x8_raw.csv"]


def executed_instr_per_32_cells(kernel="step_kernel_x<8>"):
    for rel in SASS_MIX_FILES:
        try:
            k = json.load(open(os.path.join(ROOT, rel)))["kernels"][kernel]
            g = k["generation_loop"]
            return float(k["alu_instr_per_32_cells"]), {"file": rel, "loop_instructions": g["instructions"], "pipes": g["pipes"],
                                                         "ops": {o: c for o, c in list(g["ops"].items())[:6]}}
        except Exception:  # noqa: BLE001
            continue
    return 7.75, {"file": None, "note": "fallback constant (no committed SASS histogram found)"}


def ncu_dram_traffic(kernel_substr="step_kernel_x<8>", grid="(16384, 1, 1)"):
    """dram__bytes_read.sum + dram__bytes_write.sum of the committed capture of the same launch shape, or None."""
    import csv
    for rel in NCU_RAW_FILES:
        try:
            rows = list(csv.reader(open(os.path.join(ROOT, rel))))
            hdr = rows[0]
            ik, ig = hdr.index("Kernel Name"), hdr.index("Grid Size")
            ir, iw = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
            scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}  # rows[1] = units (ncu --print-units)
            for r in rows[2:]:
                if kernel_substr in r[ik] and r[ig] == grid:
                    return int(float(r[ir]) * scale[rows[1][ir]] + float(r[iw]) * scale[rows[1][iw]]), rel
        except Exception:  # noqa: BLE001
            continue
    return None, None
METRIC = "cell_updates_per_sec"
UNIT = "Gcell/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--soups", type=int, default=N_SOUPS, help="soups per GPU (default: BASELINE config 2)")
    ap.add_argument("--gens", type=int, default=GENS)
    ap.add_argument("--lpp", type=int, default=0, help="lanes per pattern override (tuning)")
    ap.add_argument("--layout", type=int, default=0, help="step-kernel register layout override 1/2/4/8 (tuning)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the config 3/4/5 side measurements")
    ap.add_argument("--no-survey", action="store_true", help="skip the config 5 survey + NCCL gather leg")
    ap.add_argument("--survey-soups", type=int, default=1 << 24, help="config 5 batch (whole job, every N)")
    return ap.parse_args()


def workload_config(args, world):
    return {"workload": f"config2: {args.soups} random 64x64 soups/GPU x {args.gens} generations in-register "
                        f"(splitmix64 seed 0x{SEED:X}, full-density)",
            "soups_per_gpu": args.soups, "generations": args.gens, "pattern_bytes": 512,
            "parallelism": f"batch-sharded x{world} (no data-path collective; counters gathered)",
            "l2_policy": f"inputs larger than L2 ({args.soups * 512 >> 20} MiB in + same out per step, out of place)"}


# ------------------------------------------------------------------------------------------------
# CPU arm: the reference's own Step(n) on the host cores (oracle/_ref), else the C port
# ------------------------------------------------------------------------------------------------
def load_cpu_checker():
    from oracle.loader import Port, Ref
    ref = Ref.load()
    return ref if ref is not None else Port()


def cpu_time_sample(chk, soups, gens, threads, keep=None):
    t0 = time.perf_counter()
    out = chk.step_batch(soups, gens, threads)
    dt = time.perf_counter() - t0
    if keep is not None:
        keep.append(out)
    return dt


def cpu_baseline(gens, first_id=0, n_max=1 << 18, target_s=10.0):
    """Times the reference on a bounded sample of the workload's soups AND returns what it computed, so that the
    same run can compare it with the GPU's rows (the bit-exact verdict of BASELINE.md's same-run rule)."""
    from oracle.loader import Port
    chk = load_cpu_checker()
    threads = chk.threads()
    port = Port()
    cal_n = 256 * max(1, threads)
    soups = port.soup(first_id, cal_n, SEED, 0)
    dt = cpu_time_sample(chk, soups, gens, threads)
    rate = cal_n / max(dt, 1e-6)
    n = int(min(max(rate * target_s, cal_n), n_max))
    soups = port.soup(first_id, n, SEED, 0)
    keep = []
    dt = cpu_time_sample(chk, soups, gens, threads, keep)
    cells = n * 4096.0 * gens
    return {"value": cells / dt / 1e9, "unit": UNIT, "cores": threads, "kind": chk.kind,
            "sample": f"{n} of the same seeded soups x {gens} generations, contiguous chunks over {threads} threads, "
                      f"{dt:.1f} s wall" + (f", isa {chk.isa()}, built -march=x86-64-v4/v3/v2 (oracle/Makefile)" if hasattr(chk, "isa") else "")}, keep[0]


def parity_against_cpu(eng, d_out, first_id, n, gens, cpu_rows=None, sample=4096):
    """Bit-exact verdict: the GPU's output rows against the reference's Step(n) of the same soups."""
    from oracle.loader import Port
    if cpu_rows is None:
        chk = load_cpu_checker()
        m = min(sample, n)
        cpu_rows = chk.step_batch(Port().soup(first_id, m, SEED, 0), gens, 0)
        kind = chk.kind
    else:
        kind = load_cpu_checker().kind
    m = min(len(cpu_rows), n)
    got = np.empty((m, 64), dtype=np.uint64)
    eng.d2h(got, d_out)
    eng.sync()
    return {"checked": int(m), "equal": bool((got == np.asarray(cpu_rows)[:m]).all()), "vs": kind,
            "what": f"every word of the first {m} output patterns of this rank against LifeState::Step({gens}) on the host"}


def run_reference_arm(args, rank, world):
    """--impl reference: the reference's CPU implementation of the path, all host threads, rank 0 only."""
    if rank != 0:
        return
    from oracle.loader import Port
    chk = load_cpu_checker()
    threads = chk.threads()
    port = Port()
    total_steps = args.steps + args.warmup
    per_step_s = max(1.0, min(15.0, 90.0 / max(1, total_steps)))
    cal_n = 128 * max(1, threads)
    soups = port.soup(0, cal_n, SEED, 0)
    dt = cpu_time_sample(chk, soups, args.gens, threads)
    n = int(min(max(cal_n / max(dt, 1e-6) * per_step_s, cal_n), args.soups))
    soups = port.soup(0, n, SEED, 0)
    for _ in range(args.warmup):
        cpu_time_sample(chk, soups, args.gens, threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        chk.step_batch(soups, args.gens, threads)
    dt = time.perf_counter() - t0
    cells = n * 4096.0 * args.gens * args.steps
    value = cells / dt / 1e9
    sample = (f"each step = {n} of the workload's seeded soups x {args.gens} generations on {threads} host threads "
              f"(bounded sample of the {args.soups}-soup batch)")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u64", "data": "synthetic", "config": workload_config(args, world),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": chk.kind, "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    _emit(line)


# ------------------------------------------------------------------------------------------------
# clocks sampler (NVML), runs during the timed region
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    REASONS = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
               0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz, self.ok = [], set(), None, False
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as e:  # noqa: BLE001
            self.err = repr(e)

    def _run(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:  # noqa: BLE001
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in self.REASONS.items():
                    if r & bit and name != "gpu_idle":
                        self.reasons.add(name)
            except Exception:  # noqa: BLE001
                pass
            self._stop.wait(0.02)

    def start(self):
        if self.ok:
            self._thread = threading.Thread(target=self._run, daemon=True)
            self._thread.start()

    def stop(self):
        self._stop.set()
        if self._thread:
            self._thread.join()
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "note": getattr(self, "err", "no samples")}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
_JSON_FD = None


def _claim_stdout():
    """Keep the process's stdout for the ONE JSON line: everything else any library writes to fd 1 (NCCL prints
    its version banner there at every NCCL_DEBUG level but NONE) goes to stderr from here on."""
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)


def _emit(line: dict):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_JSON_FD, data)


def run_single_process(args):
    """`bench.py --gpus N` WITHOUT torchrun: one process drives N GPUs through the C library alone (lifeb200_multi_*:
    ncclCommInitAll, one host thread per GPU inside the calls) -- no torch, no process group.  Same JSON line; the
    resident figure steps every GPU's own 1 Mi soups from one Python thread per GPU (ctypes releases the GIL), timed
    with CUDA events on each GPU's stream, max over GPUs."""
    import threading as th

    from lifeapi_b200 import HIT_DTYPE, MultiEngine
    n_gpus, n, gens = args.gpus, args.soups, args.gens
    group = MultiEngine(list(range(n_gpus)))
    engs = [group.engine(i) for i in range(n_gpus)]
    nbytes = n * 512
    bufs = []
    for i, e in enumerate(engs):
        d_in, d_out = e.malloc(nbytes), e.malloc(nbytes)
        e.soup_dev(d_in, i * n, n, SEED, 0)
        e.sync()
        bufs.append((d_in, d_out))
    p_int, mhz_eff = engs[0].measure_int_peak()

    def on_all(fn):
        res = [None] * n_gpus
        ts = [th.Thread(target=lambda i=i: res.__setitem__(i, fn(i))) for i in range(n_gpus)]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        return res

    def timed_steps(i, k):
        e = engs[i]
        e.timer_start()
        for _ in range(k):
            e.step_batch_dev(bufs[i][0], bufs[i][1], n, gens)
        return e.timer_stop()

    on_all(lambda i: timed_steps(i, args.warmup))
    sampler = ClockSampler(0)
    sampler.start()
    ms = max(on_all(lambda i: timed_steps(i, args.steps)))
    clocks = sampler.stop()
    cells_all = float(n_gpus) * n * 4096.0 * gens
    value = cells_all * args.steps / (ms * 1e-3) / 1e9
    folds = on_all(lambda i: engs[i].fold_dev(bufs[i][1], n, i * n))
    parity = [parity_against_cpu(engs[i], bufs[i][1], i * n, n, gens) for i in range(n_gpus)]
    # e2e: one pinned host batch of N x n soups through lifeb200_multi_step_batch_host
    e2e = None
    if not args.no_e2e:
        h_in = engs[0].host_alloc((n_gpus * n, 64))
        h_out = engs[0].host_alloc((n_gpus * n, 64))
        for i in range(n_gpus):
            engs[i].d2h(h_in[i * n:(i + 1) * n], bufs[i][0])
            engs[i].sync()
        for _ in range(min(args.warmup, 2)):
            group.step_batch_host(h_in, gens, out=h_out)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            group.step_batch_host(h_in, gens, out=h_out)
        dt = time.perf_counter() - t0
        ok = True
        for i in range(n_gpus):
            engs[i].h2d(bufs[i][1], h_out[i * n:(i + 1) * n])
            ok = ok and engs[i].fold_dev(bufs[i][1], n, i * n) == folds[i]
        e2e = {"value": cells_all * args.steps / dt / 1e9, "unit": UNIT, "h2d_bytes_per_step": nbytes * n_gpus,
               "d2h_bytes_per_step": nbytes * n_gpus, "ms_per_step": dt / args.steps * 1e3, "host_memory": "pinned (lifeb200_host_alloc)",
               "api": "lifeb200_multi_step_batch_host", "matches_resident_path": bool(ok)}
        engs[0].host_free(h_in)
        engs[0].host_free(h_out)
    survey = None
    if not args.no_survey:
        total5, gens5 = args.survey_soups, 4096
        hits = np.zeros(total5 // 8 + 4096, dtype=HIT_DTYPE)
        group.reserve(total5, len(hits))
        group.soup_survey(0, min(total5, n_gpus << 14), SEED, 0, 16, len(hits), hits=hits)
        t0 = time.perf_counter()
        merged, ps5, hx5 = group.soup_survey(0, total5, SEED, 0, gens5, len(hits), hits=hits)
        t_total = time.perf_counter() - t0
        lo = group.shard(0, total5, n_gpus - 1)[0]
        m = min(1 << 16, total5 - lo)
        own, _, _ = engs[0].soup_survey(lo, m, SEED, 0, gens5, m)
        sel = (merged["id"] >= lo) & (merged["id"] < lo + m)
        survey = {"workload": f"config5: ONE batch of {total5} on-device soups x {gens5} gens over {n_gpus} GPU(s) of this process",
                  "api": "lifeb200_multi_create + lifeb200_multi_soup_survey", "nccl_version": group.info()["nccl_version"],
                  "active_hits_total": int(len(merged)), "pop_sum": int(ps5), "hash_xor": f"{int(hx5):016x}", "total_s": t_total,
                  "gather_ms_device": group.last_gather_ms(), "value": total5 * 4096.0 * gens5 / t_total / 1e9, "unit": UNIT,
                  "cross_rank_parity": {"range": [int(lo), int(lo + m)], "owner_rank": n_gpus - 1, "records": int(sel.sum()),
                                        "equal": bool(own.tobytes() == merged[sel].tobytes())}}
    kernel_ms = ms / args.steps
    exec_per32, exec_src = executed_instr_per_32_cells("step_kernel_x<8>")
    per_gpu = n * 4096.0 * gens / (kernel_ms * 1e-3)
    pop_sum = sum(f[0] for f in folds) & 0xFFFFFFFFFFFFFFFF
    hash_xor = 0
    for f in folds:
        hash_xor ^= f[1]
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": n_gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": kernel_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": dict(workload_config(args, n_gpus), launch="single process, lifeb200_multi_create (no torch, no process group)"),
            "clocks": clocks, "e2e": e2e, "gpu_launches": args.steps * n_gpus,
            "roofline": {"bound": "int_pipe", "unit": "T lane-instr/s (LOP3+SHF)", "peak": p_int / 1e12,
                         "achieved": per_gpu * exec_per32 / 32.0 / 1e12, "frac": per_gpu * exec_per32 / 32.0 / p_int,
                         "executed_instr_per_32_cells": exec_per32, "executed_source": exec_src, "traffic": None, "kernel": "step_kernel_x<8>",
                         "kernel_ms": kernel_ms},
            "cpu_baseline": None,
            "parity": {"checked": sum(p["checked"] for p in parity), "equal": all(p["equal"] for p in parity), "vs": parity[0]["vs"],
                       "what": f"the first {parity[0]['checked']} output patterns of every GPU against the reference's Step({gens})"},
            "survey_gather": survey, "checksum": {"pop_sum": int(pop_sum), "hash_xor": f"{int(hash_xor):016x}"}}
    _emit(line)
    for i, e in enumerate(engs):
        e.free(bufs[i][0])
        e.free(bufs[i][1])
    group.close()


def main():
    args = parse_args()
    _claim_stdout()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return
    if args.gpus > 1 and "WORLD_SIZE" not in os.environ:
        run_single_process(args)  # one process, N GPUs, the C library's own multi-GPU group
        return

    import torch
    import torch.distributed as dist

    from lifeapi_b200 import HIT_DTYPE, Engine

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the b200 arm has no CPU fallback (use --impl reference)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    distributed = world > 1
    if distributed:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if distributed:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if not distributed:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # one process per GPU: keep each rank (and the pinned buffers it touches first) on its GPU's NUMA node
    host_affinity = "all cores (single process)"
    if distributed:
        from lifeapi_b200 import bind_host_to_device
        host_affinity = bind_host_to_device(local_rank)
    eng = Engine(local_rank)
    if args.lpp:
        eng.set_lanes_per_pattern(args.lpp)
    if args.layout:
        eng.set_step_layout(args.layout)
    info = eng.device_info()
    n, gens = args.soups, args.gens
    first_id = rank * n
    nbytes = n * 512
    d_in, d_out = eng.malloc(nbytes), eng.malloc(nbytes)
    eng.soup_dev(d_in, first_id, n, SEED, 0)
    eng.sync()

    # ---- integer-pipe peak, measured on this box in this run ----------------------------------
    p_int, mhz_eff = eng.measure_int_peak()

    def one_step():
        eng.step_batch_dev(d_in, d_out, n, gens)

    for _ in range(args.warmup):
        one_step()
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    l0 = eng.launch_count()
    barrier()
    eng.timer_start()
    for _ in range(args.steps):
        one_step()
    ms = eng.timer_stop()
    barrier()
    launches = eng.launch_count() - l0
    clocks = sampler.stop()
    ms = max_over_ranks(ms)
    cells_per_step_all = float(world) * n * 4096.0 * gens
    value = cells_per_step_all * args.steps / (ms * 1e-3) / 1e9
    kernel_ms = ms / args.steps

    # ---- counters / checksum of the headline batch ---------------------------------------------------------
    local_fold = eng.fold_dev(d_out, n, first_id)  # this rank's (pop_sum, hash_xor) of the resident result
    pop_sum, hash_xor = local_fold
    if distributed:
        t = torch.tensor([pop_sum & 0x7FFFFFFFFFFFFFFF, pop_sum >> 63, hash_xor & 0x7FFFFFFFFFFFFFFF, hash_xor >> 63],
                         dtype=torch.int64, device=dev)
        ts = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(ts, t)
        pop_sum = sum(int(x[0]) | (int(x[1]) << 63) for x in ts) & 0xFFFFFFFFFFFFFFFF
        hash_xor = 0
        for x in ts:
            hash_xor ^= int(x[2]) | (int(x[3]) << 63)

    # ---- BASELINE config 5: ONE batch of 16 Mi on-device soups x 4096 generations, sharded by id over the ranks
    # (strong scaling: the same batch at every N), hit lists and counters gathered over NCCL by the C library itself
    # (lifeb200_multi_soup_survey, csrc/multi.cu; torch only hands the 128-byte NCCL unique id to the ranks)
    survey_gather = None
    if not args.no_survey:
        from lifeapi_b200 import MultiEngine
        total5, gens5 = args.survey_soups, 4096
        uid = [MultiEngine.unique_id() if (distributed and rank == 0) else None]
        if distributed:
            dist.broadcast_object_list(uid, src=0)
        group = MultiEngine(device=local_rank, rank=rank, world=world, unique_id=uid[0])
        try:
            cap5 = total5
            h_hits = np.zeros(cap5 if rank == 0 else 1, dtype=HIT_DTYPE)
            group.reserve(total5, cap5)  # scratch and gather buffers at their final size: no allocation in the timed call
            group.soup_survey(0, min(total5, world << 14), SEED, 0, 16, cap5, hits=h_hits)  # untimed: kernel image
            barrier()
            t0 = time.perf_counter()
            merged, ps5, hx5 = group.soup_survey(0, total5, SEED, 0, gens5, cap5, hits=h_hits)
            t_total = max_over_ranks(time.perf_counter() - t0)   # survey + gather + download, slowest rank
            gather_ms = max_over_ranks(group.last_gather_ms())
            cross = None
            if rank == 0:
                # multi-rank parity: rank 0 re-surveys 64 Ki ids that belong to the LAST rank (and, at N = 1, to itself)
                # on its own GPU and compares them with the records the gather delivered for that range
                lo = group.shard(0, total5, world - 1)[0]
                m = min(1 << 16, total5 - lo)
                own, _, _ = eng.soup_survey(lo, m, SEED, 0, gens5, m)
                sel = (merged["id"] >= lo) & (merged["id"] < lo + m)
                cross = {"range": [int(lo), int(lo + m)], "owner_rank": world - 1, "records": int(sel.sum()),
                         "equal": bool(own.tobytes() == merged[sel].tobytes())}
            survey_gather = {"workload": f"config5: ONE batch of {total5} on-device soups x {gens5} gens sharded by id over {world} rank(s) "
                                         f"(strong scaling), hit lists gathered over NCCL inside the C library",
                             "api": "lifeb200_multi_create_rank + lifeb200_multi_soup_survey", "nccl_version": group.info()["nccl_version"],
                             "active_hits_total": int(len(merged)) if rank == 0 else None,
                             "hits_sorted_unique": bool((np.diff(merged["id"].astype(np.int64)) > 0).all()) if rank == 0 else None,
                             "pop_sum": int(ps5), "hash_xor": f"{int(hx5):016x}", "total_s": t_total, "gather_ms_device": gather_ms,
                             "gather_bytes": int(len(merged)) * 16 if rank == 0 else None,
                             "value": total5 * 4096.0 * gens5 / t_total / 1e9, "unit": UNIT,
                             "cross_rank_parity": cross}
        finally:
            group.close()

    # ---- HBM-bound companion: one generation per launch (load/store phases dominate) -----------
    for _ in range(3):
        eng.step_batch_dev(d_in, d_out, n, 1)
    eng.sync()
    eng.timer_start()
    reps1 = 10
    for _ in range(reps1):
        eng.step_batch_dev(d_in, d_out, n, 1)
    ms1 = eng.timer_stop() / reps1
    hbm_gbs = 2.0 * nbytes / (ms1 * 1e-3) / 1e9

    # ---- e2e: host buffers through the C ABI, copies inside the timed region -------------------
    e2e = None
    if not args.no_e2e:
        h_in = eng.host_alloc((n, 64))
        h_out = eng.host_alloc((n, 64))
        eng.d2h(h_in, d_in)
        eng.sync()
        for _ in range(min(args.warmup, 2)):
            eng.step_batch_host(h_in, gens, out=h_out)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            eng.step_batch_host(h_in, gens, out=h_out)
        torch.cuda.synchronize()
        dt_local = time.perf_counter() - t0
        dt = max_over_ranks(dt_local)
        by_rank = [dt_local]
        if distributed:
            t = torch.tensor([dt_local], dtype=torch.float64, device=dev)
            ts = [torch.zeros_like(t) for _ in range(world)]
            dist.all_gather(ts, t)
            by_rank = [float(x.item()) for x in ts]
        # result check: the host path's output equals the resident path's (same checksum)
        eng.h2d(d_out, h_out)
        chk = eng.fold_dev(d_out, n, first_id)
        e2e = {"value": cells_per_step_all * args.steps / dt / 1e9, "unit": UNIT, "h2d_bytes_per_step": nbytes * world,
               "d2h_bytes_per_step": nbytes * world, "ms_per_step": dt / args.steps * 1e3,
               "ms_per_step_by_rank": [round(x / args.steps * 1e3, 3) for x in by_rank],
               "host_memory": "pinned (lifeb200_host_alloc)", "host_affinity": host_affinity, "api": "lifeb200_step_batch_host",
               "matches_resident_path": None}
        ok = 1 if chk == local_fold else 0   # per rank: the host path's output folds to this rank's resident checksum
        if distributed:
            t = torch.tensor([ok], dtype=torch.int64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MIN)
            ok = int(t.item())
        e2e["matches_resident_path"] = bool(ok)
        eng.host_free(h_in)
        eng.host_free(h_out)

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:  # noqa: BLE001
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    hbm_src = "MEASURED_PEAKS.json" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md)"

    per_gpu_cells_s = n * 4096.0 * gens / (kernel_ms * 1e-3)
    layout8 = (args.layout in (0, 8)) and args.lpp == 0 and gens >= 16
    exec_per32, exec_src = executed_instr_per_32_cells("step_kernel_x<8>") if layout8 else (None, {"note": "non-default kernel"})
    traffic, traffic_file = ncu_dram_traffic() if (n, gens) == (N_SOUPS, GENS) and layout8 else (None, None)
    achieved_exec = per_gpu_cells_s * (exec_per32 / 32.0) if exec_per32 else None
    achieved_alg = per_gpu_cells_s * INSTR_PER_CELL
    roofline = {"bound": "int_pipe", "unit": "T lane-instr/s (LOP3+SHF)", "peak": p_int / 1e12,
                # achieved / frac: integer-pipe instructions the kernel EXECUTES (SASS histogram of the shipped .so) per
                # second against the LOP3 rate measured in this run -- a fraction of the pipe, <= 1 by construction
                "achieved": (achieved_exec or achieved_alg) / 1e12, "frac": (achieved_exec or achieved_alg) / p_int,
                "executed_instr_per_32_cells": exec_per32, "frac_executed": achieved_exec / p_int if achieved_exec else None,
                "executed_source": exec_src,
                # SURVEY 8(d)'s algorithmic figure (the literal two-full-adder form: 2 SHF + 10 LOP3 per 32 cells): what a
                # kernel issuing that form would need; above 1 because the shipped circuit is cheaper, not because work is skipped
                "algorithmic_equiv": {"instr_per_32_cells": 12.0, "achieved": achieved_alg / 1e12, "frac": achieved_alg / p_int,
                                      "what": "cells/s x 0.375 instr per cell-update (SURVEY 8d) / measured LOP3 peak"},
                "traffic": traffic,
                "traffic_source": f"dram__bytes_read.sum + dram__bytes_write.sum of this launch shape, read from {traffic_file} "
                                  "(ncu --set full)" if traffic_file else None,
                "algorithmic_bytes": 2 * nbytes,
                "kernel": "step_kernel_x<8>" if layout8 else "step_kernel", "kernel_ms": kernel_ms,
                "peak_source": f"LOP3 microbenchmark in this run (lifeb200_measure_int_peak): {p_int / 1e12:.2f} T/s = "
                               f"{info['sm_count']} SMs x 64 lanes x {mhz_eff:.0f} MHz effective",
                "hbm": {"bound": "hbm", "achieved": hbm_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": hbm_gbs / hbm_peak,
                        "what": "same kernel at 1 generation per launch (512 B in + 512 B out per pattern)",
                        "ms": ms1, "peak_source": hbm_src, "traffic": None}}

    # ---- CPU baseline (rank 0 at N = 1) and the bit-exact verdict of the same run (every rank) ----------------
    cpu, cpu_rows = None, None
    if rank == 0 and world == 1 and not args.no_cpu:
        cpu, cpu_rows = cpu_baseline(gens, first_id, n_max=min(1 << 18, n))
    eng.step_batch_dev(d_in, d_out, n, gens)  # d_out was reused by the e2e check: recompute the resident result
    parity = parity_against_cpu(eng, d_out, first_id, n, gens, cpu_rows)
    if distributed:
        t = torch.tensor([1 if parity["equal"] else 0, parity["checked"]], dtype=torch.int64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        parity["equal"] = bool(int(t[0].item()))
        parity["what"] += f"; every one of the {world} ranks checked its own first {parity['checked']} and all agree"
    extras = None
    if world == 1 and not args.no_extras:
        from bench_extras import run_extras
        extras = run_extras(eng, p_int)

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": kernel_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32",
                "data": "synthetic", "config": workload_config(args, world), "clocks": clocks, "e2e": e2e,
                "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu, "parity": parity, "other_configs": extras, "survey_gather": survey_gather,
                "checksum": {"pop_sum": int(pop_sum), "hash_xor": f"{int(hash_xor):016x}"},
                "lanes_per_pattern": args.lpp or "default", "step_layout": args.layout or "auto", "device": {"sms": info["sm_count"], "cc": list(info["cc"])}}
        _emit(line)

    eng.free(d_in)
    eng.free(d_out)
    eng.close()
    if distributed:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
