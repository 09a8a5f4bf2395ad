"""CUDA emitter for LifeAPI's generated bitsliced circuits.

The reference generates its three-valued / stable-background cell logic offline: the scripts under
bitslicing/*.py enumerate truth tables, call the `espresso` minimiser and print sum-of-products C
fragments (bitslicing/common.py:285-335) that LifeStable.hpp #includes inside per-word loops.
This module is the CUDA back end of that flow:

  * `parse_hpp`       reads a committed fragment (the source of truth: espresso is not available
                      and two generators are stale w.r.t. their .hpp);
  * `from_espresso`   takes the same (lines, phase_line, innames, outnames) that
                      common.print_output / print_phase_correction consume, so the generators can
                      call it directly once espresso is present;
  * `synthesize`      maps the SOP onto a network of 3-input LUTs (LOP3): shared literal
                      triples/pairs are extracted greedily, product terms are folded three
                      factors per LUT, OR-accumulation absorbs a two-factor AND per LUT, and the
                      espresso output-phase inversions are folded into the last LUT;
  * `render_cuda`     prints `__device__ __forceinline__` functions over uint32_t lanes;
  * `verify`          proves network == SOP on ALL 2^k input combinations (k <= 17), so don't-care
                      behaviour is the .hpp's, not a re-minimisation's.

    python -m lifeapi_b200.bitslicing.emit_cuda [--ref /root/reference/bitslicing]

regenerates lifeapi_b200/bitslicing/circuits.json (parsed cubes + LUT netlists; the portable
artefact the tests check) and lifeapi_b200/csrc/generated/circuits.cuh.
"""
from __future__ import annotations

import argparse
import hashlib
import itertools
import json
import os
import re
from collections import Counter
from dataclasses import dataclass, field

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
JSON_PATH = os.path.join(HERE, "circuits.json")
CUH_PATH = os.path.join(ROOT, "lifeapi_b200", "csrc", "generated", "circuits.cuh")

# Input / output orders are the names the reference's loops declare (SURVEY.md 8a fact sheet).
CIRCUITS = {
    "unknown_step": dict(
        file="unknown_step.hpp",
        inputs="current_unknown current_on on2 on1 on0 unk3 unk2 unk1 unk0".split(),
        outputs="naive_next_unknown naive_next_on".split()),
    "stable_simple": dict(
        file="stable_simple.hpp",
        inputs="stateunk stateon on2 on1 on0 unk1 unk0".split(),
        outputs="set_off set_on signal_off signal_on abort".split()),
    "stable_count": dict(
        file="stable_count.hpp",
        inputs="known_off known_on on2 on1 on0 off3 off2 off1 off0".split(),
        outputs="l2 l3 d0 d1 d2 d4 d5 d6 abort".split()),
    "stable_signal": dict(
        file="stable_signal.hpp",
        inputs="stateunk stateon l2 l3 d0 d1 d2 d4 d5 d6 s2 s1 s0 m3 m2 m1 m0".split(),
        outputs="signalon signaloff centeron centeroff".split()),
    "stable_vulnerable": dict(
        file="stable_vulnerable.hpp",
        inputs="l2 l3 d0 d1 d2 d4 d5 d6 s2 s1 s0 unk3 unk2 unk1 unk0".split(),
        outputs="vulnerable_on vulnerable_off vulnerable_center_on vulnerable_center_off".split()),
    "unknown_step_refined": dict(
        file="unknown_step_refined.hpp",
        inputs="l2 l3 d0 d1 d2 d4 d5 d6 current_unknown current_on s2 s1 s0 on2 on1 on0".split(),
        outputs="next_on next_unknown next_unknown_stable".split()),
    "unknown_step_refined_m": dict(
        file="unknown_step_refined_m.hpp",
        inputs="l2 l3 d0 d1 d2 d4 d5 d6 current_on m3 m2 m1 m0".split(),
        outputs="next_on next_unknown next_unknown_stable".split()),
}


# ------------------------------------------------------------------------------------------------
# SOP representation
# ------------------------------------------------------------------------------------------------
@dataclass
class Circuit:
    """outputs[o] = (OR over terms t with o in t.outs of AND(t.literals)) ^ inverted[o]."""
    name: str
    inputs: list[str]
    outputs: list[str]
    terms: list[tuple[tuple[tuple[int, bool], ...], tuple[int, ...]]]  # ((input idx, positive)...), (output idx...)
    inverted: list[bool]
    source: str = ""
    sha256: str = ""

    def to_json(self) -> dict:
        return {"name": self.name, "inputs": self.inputs, "outputs": self.outputs, "inverted": self.inverted,
                "source": self.source, "sha256": self.sha256,
                "terms": [[[[i, int(p)] for i, p in lits], list(outs)] for lits, outs in self.terms]}

    @staticmethod
    def from_json(d: dict) -> "Circuit":
        terms = [(tuple((int(i), bool(p)) for i, p in lits), tuple(int(o) for o in outs)) for lits, outs in d["terms"]]
        return Circuit(d["name"], list(d["inputs"]), list(d["outputs"]), terms, [bool(x) for x in d["inverted"]],
                       d.get("source", ""), d.get("sha256", ""))


_LIT = re.compile(r"\(\s*~\s*(\w+)\s*\)|(\w+)")


def _parse_product(expr: str, index: dict[str, int]) -> tuple[tuple[int, bool], ...]:
    lits = []
    for part in expr.split("&"):
        m = _LIT.fullmatch(part.strip())
        if not m:
            raise ValueError(f"cannot parse literal {part!r}")
        name, pos = (m.group(1), False) if m.group(1) else (m.group(2), True)
        if name not in index:
            raise ValueError(f"unknown input {name!r}")
        lits.append((index[name], pos))
    if len({i for i, _ in lits}) != len(lits):
        raise ValueError(f"repeated variable in product {expr!r}")
    return tuple(sorted(lits))


def parse_hpp(name: str, text: str, inputs: list[str], outputs: list[str], source: str = "") -> Circuit:
    """Parse the text common.print_output / print_phase_correction print (common.py:285-309)."""
    iidx = {n: i for i, n in enumerate(inputs)}
    oidx = {n: i for i, n in enumerate(outputs)}
    terms, inverted = [], [False] * len(outputs)
    for raw in text.splitlines():
        line = raw.strip()
        if not line or line.startswith("//"):
            continue
        m = re.fullmatch(r"\{\s*uint64_t\s+temp\s*=\s*(.+?);\s*((?:\w+\s*\|=\s*temp;\s*)+)\}", line)
        if m:
            outs = tuple(oidx[o] for o in re.findall(r"(\w+)\s*\|=\s*temp;", m.group(2)))
            terms.append((_parse_product(m.group(1), iidx), outs))
            continue
        m = re.fullmatch(r"(\w+)\s*\|=\s*(.+);", line)
        if m:
            terms.append((_parse_product(m.group(2), iidx), (oidx[m.group(1)],)))
            continue
        m = re.fullmatch(r"(\w+)\s*=\s*~\s*(\w+)\s*;", line)
        if m and m.group(1) == m.group(2):
            inverted[oidx[m.group(1)]] = not inverted[oidx[m.group(1)]]
            continue
        raise ValueError(f"{name}: cannot parse line {raw!r}")
    return Circuit(name, list(inputs), list(outputs), terms, inverted, source,
                   hashlib.sha256(text.encode()).hexdigest())


def from_espresso(name: str, lines: list[str], phase_line: str, innames: list[str], outnames: list[str]) -> Circuit:
    """Same arguments as common.print_output + print_phase_correction (common.py:285-309,328-335)."""
    terms = []
    for term in lines:
        ins, outs = term.split(" ")
        lits = tuple((i, v == "1") for i, v in enumerate(ins) if v in "01")
        terms.append((lits, tuple(o for o, v in enumerate(outs) if v == "1")))
    inverted = [v == "0" for v in phase_line[8:8 + len(outnames)]]
    return Circuit(name, list(innames), list(outnames), terms, inverted, "espresso")


# ------------------------------------------------------------------------------------------------
# LUT3 network
# ------------------------------------------------------------------------------------------------
_A, _B, _C = 0xF0, 0xCC, 0xAA  # lop3 truth-table constants: lut = f(0xF0, 0xCC, 0xAA)


@dataclass
class Net:
    """Signals 0..n_in-1 are the inputs; signal n_in+k is nodes[k] = lop3<lut>(a, b, c)."""
    n_in: int
    nodes: list[tuple[int, int, int, int]] = field(default_factory=list)  # (lut, a, b, c)
    outputs: list[int] = field(default_factory=list)                     # signal per circuit output
    _cache: dict = field(default_factory=dict)

    def add(self, lut: int, a: int, b: int, c: int) -> int:
        key = (lut, a, b, c)
        if key not in self._cache:
            self.nodes.append(key)
            self._cache[key] = self.n_in + len(self.nodes) - 1
        return self._cache[key]

    def to_json(self) -> dict:
        return {"n_in": self.n_in, "nodes": [list(n) for n in self.nodes], "outputs": self.outputs}

    @staticmethod
    def from_json(d: dict) -> "Net":
        return Net(int(d["n_in"]), [tuple(int(x) for x in n) for n in d["nodes"]], [int(o) for o in d["outputs"]])


def _lut(fn, pols) -> int:
    """LUT of fn over three operands used with polarities pols (True = as is, False = inverted)."""
    vals = [v if p else (~v & 0xFF) for v, p in zip((_A, _B, _C), pols)]
    return fn(*vals) & 0xFF


Factor = tuple[int, bool]  # (signal, positive)


def _and(net: Net, fs: list[Factor]) -> Factor:
    """AND of 2 or 3 factors as one LUT (operands sorted so equal ANDs share a node)."""
    fs = sorted(fs)
    if len(fs) == 2:
        (a, pa), (b, pb) = fs
        return (net.add(_lut(lambda x, y, z: x & y, (pa, pb, True)), a, b, b), True)
    (a, pa), (b, pb), (c, pc) = fs
    return (net.add(_lut(lambda x, y, z: x & y & z, (pa, pb, pc)), a, b, c), True)


def synthesize(c: Circuit) -> Net:
    net = Net(len(c.inputs))
    terms = [[set(lits), outs] for lits, outs in c.terms]

    def substitute(group: tuple[Factor, ...], minlen: int) -> None:
        node = _and(net, list(group))
        g = set(group)
        for t in terms:
            if len(t[0]) >= minlen and g <= t[0]:
                t[0] = (t[0] - g) | {node}

    # 1. shared literal triples, then pairs that occur often enough to pay for their own LUT
    while True:
        cnt = Counter()
        for fs, _ in terms:
            if len(fs) >= 3:
                cnt.update(itertools.combinations(sorted(fs), 3))
        if not cnt:
            break
        best, n = max(cnt.items(), key=lambda kv: (kv[1], kv[0]))
        if n < 2:
            break
        substitute(best, 3)
    while True:
        cnt = Counter()
        for fs, _ in terms:
            if len(fs) >= 3:
                cnt.update(itertools.combinations(sorted(fs), 2))
        if not cnt:
            break
        best, n = max(cnt.items(), key=lambda kv: (kv[1], kv[0]))
        if n < 3:
            break
        substitute(best, 3)
    # 2. fold every product down to at most two factors, three factors per LUT
    for t in terms:
        fs = sorted(t[0])
        while len(fs) > 2:
            fs = sorted(fs[3:] + [_and(net, fs[:3])])
        if len(fs) == 2 and len(t[1]) >= 3:
            fs = [_and(net, fs)]
        t[0] = fs
    # 3. OR-accumulate per output; a two-factor product rides along inside the OR's LUT
    net.outputs = []
    for o in range(len(c.outputs)):
        mine = [t[0] for t in terms if o in t[1]]
        singles = sorted({f[0] for f in mine if len(f) == 1})
        doubles = sorted({tuple(f) for f in mine if len(f) == 2})
        acc: Factor | None = None
        fresh = False  # acc is a node created for this output only (its LUT may absorb the inversion)
        for (f1, f2) in doubles:
            if acc is None and singles:
                s = singles.pop(0)
                acc = (net.add(_lut(lambda x, y, z: (x & y) | z, (f1[1], f2[1], s[1])), f1[0], f2[0], s[0]), True)
            elif acc is None:
                acc = _and(net, [f1, f2])
            else:
                acc = (net.add(_lut(lambda x, y, z: (x & y) | z, (f1[1], f2[1], acc[1])), f1[0], f2[0], acc[0]), True)
            fresh = True
        while singles:
            take = singles[:2] if acc is not None else singles[:3]
            singles = singles[len(take):]
            ops = take + ([acc] if acc is not None else [])
            if len(ops) == 1:
                acc, fresh = ops[0], False
                continue
            while len(ops) < 3:
                ops.append(ops[-1])
            acc = (net.add(_lut(lambda x, y, z: x | y | z, tuple(p for _, p in ops)), *(s for s, _ in ops)), True)
            fresh = True
        if acc is None:
            raise ValueError(f"{c.name}: output {c.outputs[o]} has no product term")
        want_pos = not c.inverted[o]
        if acc[1] != want_pos or acc[0] < net.n_in:
            users = sum(1 for (_, a, b, cc) in net.nodes for s in (a, b, cc) if s == acc[0]) + net.outputs.count(acc[0])
            if fresh and acc[0] >= net.n_in and users == 0 and acc[0] == net.n_in + len(net.nodes) - 1:
                lut, a, b, cc = net.nodes.pop()
                del net._cache[(lut, a, b, cc)]
                acc = (net.add(~lut & 0xFF, a, b, cc), True)
            else:  # plain copy / inverter
                acc = (net.add(_lut(lambda x, y, z: x, (acc[1] == want_pos, True, True)), acc[0], acc[0], acc[0]), True)
        net.outputs.append(acc[0])
    return net


# ------------------------------------------------------------------------------------------------
# exhaustive evaluation (Python big ints as 2^k-row bit vectors)
# ------------------------------------------------------------------------------------------------
def _input_columns(k: int) -> tuple[list[int], int]:
    rows = 1 << k
    full = (1 << rows) - 1
    cols = []
    for i in range(k):
        period = 1 << (i + 1)
        block = ((1 << (1 << i)) - 1) << (1 << i)  # 2^i zeros then 2^i ones
        v = 0
        for j in range(rows // period):
            v |= block << (j * period)
        cols.append(v)
    return cols, full


def eval_sop(c: Circuit, cols: list[int], full: int) -> list[int]:
    outs = [0] * len(c.outputs)
    for lits, os_ in c.terms:
        t = full
        for i, pos in lits:
            t &= cols[i] if pos else (full ^ cols[i])
        for o in os_:
            outs[o] |= t
    return [(v ^ full) if inv else v for v, inv in zip(outs, c.inverted)]


def eval_net(net: Net, cols: list[int], full: int) -> list[int]:
    sig = list(cols)
    for lut, a, b, cc in net.nodes:
        x, y, z = sig[a], sig[b], sig[cc]
        r = 0
        for k in range(8):
            if lut >> k & 1:
                r |= (x if k & 4 else full ^ x) & (y if k & 2 else full ^ y) & (z if k & 1 else full ^ z)
        sig.append(r)
    return [sig[o] for o in net.outputs]


def verify(c: Circuit, net: Net) -> None:
    cols, full = _input_columns(len(c.inputs))
    want, got = eval_sop(c, cols, full), eval_net(net, cols, full)
    for o, (w, g) in enumerate(zip(want, got)):
        if w != g:
            raise AssertionError(f"{c.name}: LUT network differs from the SOP on output {c.outputs[o]}")


# ------------------------------------------------------------------------------------------------
# renderers
# ------------------------------------------------------------------------------------------------
def render_cuda(items: list[tuple[Circuit, Net]]) -> str:
    out = ["// GENERATED by lifeapi_b200/bitslicing/emit_cuda.py -- do not edit.",
           "// LOP3 networks equal, on every input combination, to the sum-of-products fragments the",
           "// reference's generators print (bitslicing/common.py:285-309); one function per fragment.",
           "// Outputs are plain values (the fragments OR into zero-initialised locals, e.g. LifeStable.hpp:442-455).",
           "#pragma once", "#include <cstdint>", '#include "../life_core.cuh"', "", "namespace lifeb200 {", "namespace circuits {", ""]
    for c, net in items:
        out.append(f"// {c.source}  sha256 {c.sha256[:16]}  {len(c.terms)} product terms -> {len(net.nodes)} LOP3")
        out.append(f"constexpr int k_{c.name}_lop3 = {len(net.nodes)};")
        args = [f"uint32_t {n}" for n in c.inputs] + [f"uint32_t& {n}" for n in c.outputs]
        out.append(f"__device__ __forceinline__ void {c.name}({', '.join(args)}) {{")
        names = list(c.inputs) + [f"t{k}" for k in range(len(net.nodes))]
        for k, (lut, a, b, cc) in enumerate(net.nodes):
            out.append(f"  const uint32_t t{k} = lop3<0x{lut:02X}>({names[a]}, {names[b]}, {names[cc]});")
        for n, s in zip(c.outputs, net.outputs):
            out.append(f"  {n} = {names[s]};")
        out += ["}", ""]
    out += ["}  // namespace circuits", "}  // namespace lifeb200", ""]
    return "\n".join(out)


def load_json(path: str = JSON_PATH) -> list[tuple[Circuit, Net]]:
    d = json.load(open(path))
    return [(Circuit.from_json(e["circuit"]), Net.from_json(e["net"])) for e in d["circuits"]]


def build_all(ref_dir: str | None) -> list[tuple[Circuit, Net]]:
    """Parse from the reference when it is present, else from the committed circuits.json."""
    if ref_dir and os.path.isdir(ref_dir):
        circuits = []
        for name, spec in CIRCUITS.items():
            text = open(os.path.join(ref_dir, spec["file"])).read()
            circuits.append(parse_hpp(name, text, spec["inputs"], spec["outputs"], f"bitslicing/{spec['file']}"))
    else:
        circuits = [c for c, _ in load_json()]
    items = []
    for c in circuits:
        net = synthesize(c)
        verify(c, net)
        items.append((c, net))
    return items


def write_all(items: list[tuple[Circuit, Net]]) -> None:
    os.makedirs(os.path.dirname(CUH_PATH), exist_ok=True)
    with open(JSON_PATH, "w") as f:
        json.dump({"circuits": [{"circuit": c.to_json(), "net": n.to_json()} for c, n in items]}, f, separators=(",", ":"))
        f.write("\n")
    with open(CUH_PATH, "w") as f:
        f.write(render_cuda(items))


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref", default="/root/reference/bitslicing", help="directory holding the committed .hpp fragments")
    args = ap.parse_args()
    items = build_all(args.ref)
    write_all(items)
    for c, n in items:
        lits = sum(len(l) for l, _ in c.terms)
        print(f"{c.name:24s} {len(c.inputs):2d} in {len(c.outputs)} out {len(c.terms):3d} terms {lits:4d} literals -> {len(n.nodes):3d} LOP3")


if __name__ == "__main__":
    main()
