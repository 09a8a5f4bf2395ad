#!/usr/bin/env python
"""sass_mix.py -- per-kernel loop-body instruction histogram of the shipped library, from `cuobjdump -sass`.

    python tools/sass_mix.py [--lib lifeapi_b200/_lib/liblifeapi_b200.so] [--out profiles/r02_sass_mix]

Writes <out>.txt (readable table) and <out>.json (what bench.py reads for
`roofline.executed_instr_per_32_cells`).  For every kernel: code size, local-memory instructions
(LDL/STL = spills) and, for each loop (a backward branch: the instructions between the branch target
and the branch), the opcode histogram grouped by the pipe that issues it:
    alu   LOP3 SHF IADD3 PRMT ISETP SEL LEA IMNMX VIADD ...   (the integer pipe the step kernels are bound by)
    fma   IMAD (all forms incl. IMAD.MOV / IMAD.SHL)
    shfl  SHFL      lsu  LD* ST* ATOM* RED*      cf  BRA / BSSY / BSYNC / EXIT / WARPSYNC
    other everything else (S2R, VOTE, POPC, MOV, uniform-datapath U* ...)
The "generation loop" of a kernel is its largest innermost loop.  For the step kernels the figure the
roofline uses is   int-pipe instructions (alu) per 32-cell register per generation  =  alu / registers
stepped, with the registers per lane and iteration given in KERNEL_REGS below (64 for the layout-8
kernels: 8 columns x 8 row classes, 4 patterns bit-interleaved).
No GPU needed: cuobjdump reads the sm_100a cubin embedded in the .so.
"""
from __future__ import annotations

import argparse
import collections
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

ALU = {"LOP3", "SHF", "IADD3", "IADD", "PRMT", "ISETP", "SEL", "LEA", "IMNMX", "VIADD", "VIMNMX", "LOP", "BMSK", "SGXT",
       "IABS", "FLO", "BREV", "PLOP3", "R2P", "P2R", "ICMP", "FSEL", "FMNMX", "FSETP", "VABSDIFF", "VABSDIFF4", "IDP", "IADD32I",
       "LOP32I", "CSET", "CSETP", "FCHK", "PSETP"}
FMA = {"IMAD", "FFMA", "FMUL", "FADD", "IMUL", "IMUL32I", "IMAD32I"}
SHFL = {"SHFL"}
CF = {"BRA", "BSSY", "BSYNC", "EXIT", "WARPSYNC", "RET", "CALL", "BRX", "JMP", "BREAK", "NOP", "YIELD", "BPT", "KILL", "NANOSLEEP"}

# registers (32 cells each) one lane steps per iteration of the generation loop
KERNEL_REGS = {
    "step_kernel_x<8>": 64, "step_kernel_x<4>": 32, "survey_kernel_x<8>": 64, "enumerate_kernel_x": 64,
    "weld_step_kernel_x": 64, "step_match_kernel_x": 64,
    "step_kernel<8>": 16, "step_kernel<16>": 8, "step_kernel<4>": 32, "step_kernel<32>": 4,
    "unknown_step_kernel<16, false>": 8, "unknown_step_kernel<16, true>": 8,
    "unknown_step_kernel<32, false>": 4, "unknown_step_kernel<32, true>": 4,
    "unknown_step_kernel_x": 32,
}


def pipe_of(op: str) -> str:
    base = op.split(".")[0]
    if base in SHFL:
        return "shfl"
    if base in FMA:
        return "fma"
    if base in ALU:
        return "alu"
    if base in CF:
        return "cf"
    if base.startswith(("LD", "ST", "ATOM", "RED", "CCTL", "MEMBAR", "ERRBAR", "BAR")):
        return "lsu"
    return "other"


INSTR = re.compile(r"^\s*/\*([0-9a-f]{4,})\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)\b(.*?);")


def demangle(names: list[str]) -> list[str]:
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True, check=True).stdout.split("\n")
    res = []
    for n in out[: len(names)]:
        n = re.sub(r"^void\s+", "", n)
        n = n.replace("lifeb200::", "")
        n = re.sub(r"\(.*$", "", n)
        res.append(n)
    return res


def parse(lib: str):
    sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
    kernels, cur = [], None
    for line in sass.split("\n"):
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = {"mangled": m.group(1), "instr": []}
            kernels.append(cur)
            continue
        if cur is None:
            continue
        m = INSTR.match(line)
        if m:
            cur["instr"].append((int(m.group(1), 16), m.group(2), m.group(3)))
    names = demangle([k["mangled"] for k in kernels])
    for k, n in zip(kernels, names):
        k["name"] = n
    return kernels


def loops_of(instr):
    """Backward branches -> (start index, end index) of loop bodies, inclusive."""
    index = {a: i for i, (a, _, _) in enumerate(instr)}
    loops = []
    for i, (addr, op, rest) in enumerate(instr):
        if op.split(".")[0] != "BRA":
            continue
        m = re.search(r"0x([0-9a-f]+)", rest)
        if not m:
            continue
        tgt = int(m.group(1), 16)
        if tgt <= addr and tgt in index:
            loops.append((index[tgt], i))
    return loops


def summarise(kernel):
    instr = kernel["instr"]
    spills = sum(1 for _, op, _ in instr if op.split(".")[0] in ("LDL", "STL"))
    res = {"name": kernel["name"], "instructions": len(instr), "local_memory_instructions": spills, "loops": []}
    spans = loops_of(instr)
    for (s, e) in spans:
        body = instr[s:e + 1]
        ops = collections.Counter(op.split(".")[0] for _, op, _ in body)
        pipes = collections.Counter(pipe_of(op) for _, op, _ in body)
        res["loops"].append({"start": f"0x{instr[s][0]:04x}", "end": f"0x{instr[e][0]:04x}", "instructions": len(body),
                             "pipes": dict(pipes), "ops": dict(ops.most_common())})
    if res["loops"]:
        # the generation loop = the largest INNERMOST loop (persistent kernels wrap it in a pattern loop)
        inner = [l for l, (s0, e0) in zip(res["loops"], spans)
                 if not any((s1 >= s0 and e1 <= e0) and (s1, e1) != (s0, e0) for (s1, e1) in spans)]
        big = max(inner or res["loops"], key=lambda l: l["instructions"])
        res["generation_loop"] = big
        res["innermost_loops"] = [l for l in inner if l["instructions"] >= 40]
        regs = KERNEL_REGS.get(kernel["name"])
        if regs:
            res["registers_per_iteration"] = regs
            res["alu_instr_per_32_cells"] = big["pipes"].get("alu", 0) / regs
            res["all_instr_per_32_cells"] = big["instructions"] / regs
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--lib", default=os.path.join(ROOT, "lifeapi_b200", "_lib", "liblifeapi_b200.so"))
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "r02_sass_mix"))
    ap.add_argument("--only", default="", help="substring filter on kernel names")
    args = ap.parse_args()
    kernels = [summarise(k) for k in parse(args.lib) if args.only in k["name"]]
    kernels.sort(key=lambda k: k["name"])
    with open(args.out + ".json", "w") as f:
        json.dump({"lib": os.path.relpath(args.lib, ROOT), "tool": "tools/sass_mix.py (cuobjdump -sass)",
                   "kernels": {k["name"]: k for k in kernels}}, f, indent=1)
    lines = [f"# {os.path.relpath(args.lib, ROOT)}: loop-body instruction mix per kernel (tools/sass_mix.py, cuobjdump -sass)",
             "# kernel | total instr | LDL/STL | generation loop: instr = alu + fma + shfl + lsu + cf + other | alu per 32-cell register",
             ""]
    for k in kernels:
        g = k.get("generation_loop")
        if g:
            p = g["pipes"]
            per = f" | {k['alu_instr_per_32_cells']:.3f} alu / 32 cells ({k['registers_per_iteration']} registers per iteration)" \
                if "alu_instr_per_32_cells" in k else ""
            lines.append(f"{k['name']:44s} | {k['instructions']:6d} | {k['local_memory_instructions']:5d} | {g['instructions']:5d} = "
                         f"{p.get('alu', 0)} + {p.get('fma', 0)} + {p.get('shfl', 0)} + {p.get('lsu', 0)} + {p.get('cf', 0)} + {p.get('other', 0)}{per}")
            top = ", ".join(f"{o} {c}" for o, c in list(g["ops"].items())[:8])
            lines.append(f"{'':44s}   ops: {top}")
            for l in k.get("innermost_loops", []):
                if l is not g:
                    q = l["pipes"]
                    lines.append(f"{'':44s}   other loop {l['start']}: {l['instructions']} = {q.get('alu', 0)} + {q.get('fma', 0)} + "
                                 f"{q.get('shfl', 0)} + {q.get('lsu', 0)} + {q.get('cf', 0)} + {q.get('other', 0)}")
        else:
            lines.append(f"{k['name']:44s} | {k['instructions']:6d} | {k['local_memory_instructions']:5d} | no loop")
    with open(args.out + ".txt", "w") as f:
        f.write("\n".join(lines) + "\n")
    print("\n".join(lines[:3]))
    for k in kernels:
        if "alu_instr_per_32_cells" in k:
            print(f"{k['name']}: {k['alu_instr_per_32_cells']:.3f} alu instr / 32 cells, loop {k['generation_loop']['instructions']} instr")
    return 0


if __name__ == "__main__":
    sys.exit(main())
