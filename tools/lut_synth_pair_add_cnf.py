"""Exact synthesis of the pair-shared B3/S23 select with FREE constrained add gates (FMA pipe).

    python tools/lut_synth_pair_add_cnf.py <code 0|1|2> <pair kinds> <left kinds> <right kinds> [--timeout S]

Same problem as lut_synth_pair_cnf.py (--dc --raw form), with a second gate kind.  On sm_100 LOP3 /
SHF issue on the ALU pipe (64 lanes/clk/SM), IMAD on the FMA pipe, and the two pipes run side by
side (profiles/r01_ubench_pipes.txt: lop3 x8 + imad x4 keeps 1.94 of 1.99 LOP3/clk/SM).  An integer
add of two bit planes is a bitwise function only if no bit position ever carries or borrows:

    A gate:  out = x + y   needs x & y == 0 on every occurring input   (then out = x | y = x ^ y)
             out = x - y   needs y subset of x on every occurring input (then out = x & ~y = x ^ y)
             x or y may be the constant all-ones plane (out = ~y for free: -1 - y)
    L gate:  out = lop3<imm>(x, y, z), any three earlier signals

`kinds` is the gate order of a stage, e.g. LLAL: only L gates cost ALU-pipe issue slots.  The pair
stage reads the four count bits (A0 A1 B0 B1) and its own earlier gates; a tail reads the far
count (C0 C1), the cell, the four raw count bits, every pair gate and its own earlier gates; the
last gate of a tail is the next state.  Don't-cares: the cell is the middle bit of its own
column's triple (cell = 1 => T_own >= 1, T_own = 3 => cell = 1).  "-" is an empty stage.
"""
import itertools
import subprocess
import sys
import tempfile
import time

from lut_synth_pair_cnf import CODES, Cnf, solve


def build(code, kp, kl, kr, free=0, shape=False):
    enc = CODES[code]
    cnf = Cnf()
    pair_rows = list(itertools.product(range(4), range(4)))
    tail_rows = list(itertools.product(range(4), range(4), range(4), (0, 1)))
    gates = []

    def sel_inputs(cands, nrows, value_of, nsel, ordered):
        m = len(cands)
        sel = [[cnf.var() for _ in range(m)] for _ in range(nsel)]
        for k in range(nsel):
            cnf.exactly_one(sel[k])
        for k in range(nsel - 1):
            for i in range(m):
                for j in range(0, i + 1):
                    if ordered or j == i:
                        cnf.add(-sel[k][i], -sel[k + 1][j])
        xs = []
        for r in range(nrows):
            x = [cnf.var() for _ in range(nsel)]
            for k in range(nsel):
                for i, (_, sig) in enumerate(cands):
                    kind, v = value_of(sig, r)
                    if kind == "c":
                        cnf.add(-sel[k][i], x[k] if v else -x[k])
                    else:
                        cnf.add(-sel[k][i], -x[k], v)
                        cnf.add(-sel[k][i], x[k], -v)
            xs.append(x)
        return sel, xs

    def lut_gate(name, cands, nrows, value_of):
        cands = [c for c in cands if c[0] != "ONE"]
        sel, xs = sel_inputs(cands, nrows, value_of, 3, True)
        lut = [cnf.var() for _ in range(8)]
        out = [cnf.var() for _ in range(nrows)]
        for r in range(nrows):
            x = xs[r]
            for idx in range(8):
                bits = [(idx >> 2) & 1, (idx >> 1) & 1, idx & 1]
                pre = [(-x[k] if bits[k] else x[k]) for k in range(3)]
                cnf.add(*pre, -lut[idx], out[r])
                cnf.add(*pre, lut[idx], -out[r])
        gates.append((name, "L", [c[0] for c in cands], sel, lut))
        return out

    def add_gate(name, cands, nrows, value_of):
        sel, xs = sel_inputs(cands, nrows, value_of, 2, False)
        mode = cnf.var()  # 0: x + y (disjoint), 1: x - y (y subset of x)
        out = [cnf.var() for _ in range(nrows)]
        for r in range(nrows):
            x, y = xs[r]
            cnf.add(-x, -y, -out[r]); cnf.add(x, y, -out[r]); cnf.add(-x, y, out[r]); cnf.add(x, -y, out[r])
            cnf.add(mode, -x, -y)
            cnf.add(-mode, -y, x)
        gates.append((name, "A", [c[0] for c in cands], sel, mode))
        return out

    def pair_value(sig, r):
        kind, v = sig[0], sig[1]
        return ("c", v[r]) if kind == "const" else ("v", v[r])

    one = ("ONE", ("const", [1] * 128, "any"))
    code_sigs = [("A0", ("const", [enc[ca][0] for ca, cb in pair_rows], "pair")),
                 ("A1", ("const", [enc[ca][1] for ca, cb in pair_rows], "pair")),
                 ("B0", ("const", [enc[cb][0] for ca, cb in pair_rows], "pair")),
                 ("B1", ("const", [enc[cb][1] for ca, cb in pair_rows], "pair")), one]
    for g in range(free):  # --free K: K pair-level signals with an unconstrained truth table (lower-bound runs)
        fv = [cnf.var() for _ in pair_rows]
        code_sigs.append((f"f{g}", ("var", fv, "pair")))
        gates.append((f"f{g}", "F", [], [], fv))
    for g, kind in enumerate(kp):
        fn = lut_gate if kind == "L" else add_gate
        out = fn(f"p{g}", list(code_sigs), len(pair_rows), pair_value)
        code_sigs.append((f"p{g}", ("var", out, "pair")))

    def synth_tail(tag, rows, kinds):
        idx_pair = [pair_rows.index((ca, cb)) for ca, cb, co, cell in rows]

        def value(sig, r):
            kind, v, dom = sig
            if dom == "pair":
                r = idx_pair[r]
            return ("c", v[r]) if kind == "const" else ("v", v[r])

        cands = [("C0", ("const", [enc[co][0] for ca, cb, co, cell in rows], "tail")),
                 ("C1", ("const", [enc[co][1] for ca, cb, co, cell in rows], "tail")),
                 ("cell", ("const", [cell for ca, cb, co, cell in rows], "tail"))] + list(code_sigs)
        last = None
        for g, kind in enumerate(kinds):
            fn = lut_gate if kind == "L" else add_gate
            last = fn(f"{tag}{g}", list(cands), len(rows), value)
            cands.append((f"{tag}{g}", ("var", last, "tail")))
        for r, (ca, cb, co, cell) in enumerate(rows):
            n = ca + cb + co
            want = int(n == 3 or (cell and n == 4))
            cnf.add(last[r] if want else -last[r])

    def cares(own):
        return [(ca, cb, co, cell) for ca, cb, co, cell in tail_rows
                if not ((cell == 1 and (ca, cb)[own] == 0) or (cell == 0 and (ca, cb)[own] == 3))]

    if kl:
        synth_tail("l", cares(0), kl)
    if kr:
        synth_tail("r", cares(1), kr)

    for gi, (name, _, _, _, _) in enumerate(gates):  # no dead pair gates
        if not name.startswith("p"):
            continue
        users = []
        for (n2, _, cands2, sel2, _) in gates[gi + 1:]:
            if name in cands2:
                i = cands2.index(name)
                users += [s[i] for s in sel2]
        cnf.add(*users)
    if shape:  # --shape: the tail every solution so far has -- n0 = xor3(C0, A0, B0); out = L(cell, L(C1, z, L(C0, x, y)), n0)
        byname = {g[0]: g for g in gates}

        def force(name, ins, lutv=None):
            _, _, cands, sel, lut = byname[name]
            for k, nm in enumerate(ins):
                if nm is not None:
                    cnf.add(sel[k][cands.index(nm)])
            if lutv is not None:
                for i in range(8):
                    cnf.add(lut[i] if (lutv >> i) & 1 else -lut[i])

        for tag, kinds in (("l", kl), ("r", kr)):
            if "--gshape" in sys.argv and kinds.count("L") == 4:
                # the same topology with fewer pins: g1 = L(C0, x, y); g2 = L(C1, z, g1); g0 = L(C0, u, v) with a free
                # LUT; out = L(w, g2, g0), w = the cell or an add gate of the tail
                ga, gb, gc, gd = [f"{tag}{i}" for i, k in enumerate(kinds) if k == "L"]
                force(ga, ["C0", None, None])
                force(gb, ["C1", None, ga])
                force(gc, ["C0", None, None])
                force(gd, [None, gb, gc])
                continue
            if kinds == "LLLL" and not (tag == "r" and "--shape-left" in sys.argv):
                force(f"{tag}0", ["C0", None, None])
                force(f"{tag}1", ["C1", None, f"{tag}0"])
                force(f"{tag}2", ["C0", "A0", "B0"], 0x96)
                force(f"{tag}3", ["cell", f"{tag}1", f"{tag}2"])
    return cnf, gates


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    timeout = None
    if "--timeout" in sys.argv:
        timeout = float(sys.argv[sys.argv.index("--timeout") + 1])
        args.remove(sys.argv[sys.argv.index("--timeout") + 1])
    free = 0
    if "--free" in sys.argv:
        free = int(sys.argv[sys.argv.index("--free") + 1])
        args.remove(sys.argv[sys.argv.index("--free") + 1])
    code = int(args[0])
    kp, kl, kr = [("" if a == "-" else a) for a in args[1:4]]
    cnf, gates = build(code, kp, kl, kr, free, "--shape" in sys.argv or "--shape-left" in sys.argv or "--gshape" in sys.argv)
    t0 = time.time()
    status, model = solve(cnf, timeout)
    dt = time.time() - t0
    nl = sum(k.count("L") for k in (kp, kl, kr))
    print(f"free {free} " if free else "", end="")
    print("gshape " if "--gshape" in sys.argv else "shape " if "--shape" in sys.argv else "shape-left " if "--shape-left" in sys.argv else "", end="")
    print(f"code {code} pair {kp or '-'} left {kl} right {kr} ({nl} LOP3 per column pair): vars {cnf.n} clauses "
          f"{len(cnf.clauses)} -> {status} in {dt:.1f}s", flush=True)
    if model:
        for name, kind, cands, sel, extra in gates:
            ins = [cands[[model.get(v, False) for v in s].index(True)] for s in sel]
            if kind == "F":  # rows (Ta, Tb) = (0,0) (0,1) .. (3,3)
                print(f"  {name} = pair function, truth table over (Ta,Tb): " + "".join(str(int(model.get(v, False))) for v in extra))
            elif kind == "L":
                lutv = sum(1 << i for i in range(8) if model.get(extra[i], False))
                print(f"  {name} = lop3<0x{lutv:02X}>({', '.join(ins)})")
            else:
                print(f"  {name} = {ins[0]} {'-' if model.get(extra, False) else '+'} {ins[1]}")


if __name__ == "__main__":
    main()
