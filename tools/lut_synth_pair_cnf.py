"""Exact synthesis of the pair-shared B3/S23 select as a plain CNF (SSV encoding), solved by `z3 -dimacs`.

    python tools/lut_synth_pair_cnf.py <code 0|1|2> <GP> <GT> [--dc] [--timeout S]

Problem: per column the 2-bit count T of the vertical triple is already there
(2 LOP3).  GP gates turn (T_x, T_x+1) of a column pair into a code that BOTH columns use; GT gates
per column finish  next = [T_x + T_x+1 + T_far == 3] | cell & [.. == 4].  Cost per two half-words
= GP + 2*GT LOP3 (today: 4 + 2*4).

--raw tail gates may also read the pair's four raw count bits, not only the code gates.
--gtr N  (with --dc) the right column's tail has N gates instead of GT.
--dc  synthesise the left and the right column's tail separately (same gate count, their own
      LUT constants and wiring) and use the don't-cares each has: the cell is the middle bit of
      its own column's triple, so cell=1 implies T_own >= 1 and T_own = 3 implies cell = 1.

--weld  the LifeWeld::Step rule instead (LifeWeld.hpp:169-186): every tail row also carries the cell's frozen count
        f in 0..7 and  next = [T == 3] | cell & [T == 4]  with T = (sum + f) mod 8.  The tails read three FREE planes
        F0..F2 of f (launch constants: the kernel may park any three functions of the frozen planes; 8 truth-table
        bits per plane chosen by the solver, value at f = 0 fixed to 0).  Today: 2 x (4 + 6) LOP3 per column pair.

Gate g reads three earlier signals (selector variables, strictly increasing indices); code gates
may read the four count bits and earlier code gates, tail gates the far count, the cell, every
code gate and earlier gates of the same tail.  Prints the netlist of a model, or UNSAT.
"""
import itertools
import subprocess
import sys
import tempfile
import time

CODES = {0: {0: (0, 0), 1: (1, 0), 2: (0, 1), 3: (1, 1)},   # binary (xor3, maj)
         1: {0: (0, 0), 1: (1, 0), 2: (1, 1), 3: (0, 1)},   # gray (not-all-equal, maj)
         2: {0: (0, 0), 1: (1, 1), 2: (1, 0), 3: (0, 1)}}   # the third class


class Cnf:
    def __init__(self):
        self.n = 0
        self.clauses = []

    def var(self):
        self.n += 1
        return self.n

    def add(self, *lits):
        self.clauses.append(lits)

    def exactly_one(self, xs):
        self.add(*xs)
        for a, b in itertools.combinations(xs, 2):
            self.add(-a, -b)

    def write(self, f):
        f.write(f"p cnf {self.n} {len(self.clauses)}\n")
        for c in self.clauses:
            f.write(" ".join(map(str, c)) + " 0\n")


def build(code, GP, GT, dc, raw=False, GTR=None, weld=False):
    enc = CODES[code]
    cnf = Cnf()
    pair_rows = list(itertools.product(range(4), range(4)))            # (ca, cb)
    tail_rows = list(itertools.product(range(4), range(4), range(4), (0, 1)))  # (ca, cb, co, cell)
    fvals = range(8) if weld else (0,)
    free = {}  # (tag, plane) -> 8 variables

    # signal = ("const", [bits per row]) or ("var", [vars per row]); code signals live on pair rows,
    # tail signals on tail rows (code signals are looked up through the row's (ca, cb))
    def const_sig(bits):
        return ("const", bits)

    A0 = const_sig([enc[ca][0] for ca, cb in pair_rows])
    A1 = const_sig([enc[ca][1] for ca, cb in pair_rows])
    B0 = const_sig([enc[cb][0] for ca, cb in pair_rows])
    B1 = const_sig([enc[cb][1] for ca, cb in pair_rows])

    gates = []  # (name, candidates names, sel vars [3][m], lut vars[8])

    def make_gate(name, cands, nrows, value_of, is_output):
        """cands: list of (name, sig); value_of(sig, r) -> ('c', bit) or ('v', var)."""
        m = len(cands)
        sel = [[cnf.var() for _ in range(m)] for _ in range(3)]
        for k in range(3):
            cnf.exactly_one(sel[k])
        # strictly increasing: slot k+1 index > slot k index
        for k in range(2):
            for i in range(m):
                for j in range(0, i + 1):
                    cnf.add(-sel[k][i], -sel[k + 1][j])
        lut = [cnf.var() for _ in range(8)]
        if not is_output:
            cnf.add(-lut[0])  # polarity normal form: a downstream LUT absorbs the negation
        out = [cnf.var() for _ in range(nrows)]
        for r in range(nrows):
            x = [cnf.var() for _ in range(3)]
            for k in range(3):
                for i, (_, sig) in enumerate(cands):
                    kind, v = value_of(sig, r)
                    if kind == "c":
                        cnf.add(-sel[k][i], x[k] if v else -x[k])
                    else:
                        cnf.add(-sel[k][i], -x[k], v)
                        cnf.add(-sel[k][i], x[k], -v)
            for idx in range(8):
                bits = [(idx >> 2) & 1, (idx >> 1) & 1, idx & 1]  # lop3(a,b,c): a is bit 2
                pre = [(-x[k] if bits[k] else x[k]) for k in range(3)]
                cnf.add(*pre, -lut[idx], out[r])
                cnf.add(*pre, lut[idx], -out[r])
        gates.append((name, [c[0] for c in cands], sel, lut))
        return ("var", out)

    def pair_value(sig, r):
        kind, v = sig
        return ("c", v[r]) if kind == "const" else ("v", v[r])

    code_sigs = [("A0", A0), ("A1", A1), ("B0", B0), ("B1", B1)]
    for g in range(GP):
        s = make_gate(f"p{g}", list(code_sigs), len(pair_rows), pair_value, False)
        code_sigs.append((f"p{g}", s))
    code_gate_sigs = code_sigs[4:]

    def synth_tail(tag, rows, own, GT=GT):
        """own = 0: the cell belongs to column a (left), 1: to column b (right), None: no don't-cares."""
        rows = [(ca, cb, co, cell, f) for ca, cb, co, cell in rows for f in fvals]
        idx_pair = [pair_rows.index((ca, cb)) for ca, cb, co, cell, f in rows]

        def value(sig, r):
            kind, v, dom = sig
            if dom == "pair":
                rr = idx_pair[r]
                return ("c", v[rr]) if kind == "const" else ("v", v[rr])
            return ("c", v[r]) if kind == "const" else ("v", v[r])

        cands = [("C0", ("const", [enc[co][0] for ca, cb, co, cell, f in rows], "tail")),
                 ("C1", ("const", [enc[co][1] for ca, cb, co, cell, f in rows], "tail")),
                 ("cell", ("const", [cell for ca, cb, co, cell, f in rows], "tail"))]
        if weld:
            for pl in range(3):
                fv = [cnf.var() for _ in range(8)]
                cnf.add(-fv[0])
                free[(tag, pl)] = fv
                cands.append((f"F{pl}", ("var", [fv[f] for ca, cb, co, cell, f in rows], "tail")))
        cands += [(n, (s[0], s[1], "pair")) for n, s in (code_sigs if raw else code_gate_sigs)]
        last = None
        for g in range(GT):
            s = make_gate(f"{tag}{g}", list(cands), len(rows), value, g == GT - 1)
            cands.append((f"{tag}{g}", (s[0], s[1], "tail")))
            last = s
        for r, (ca, cb, co, cell, f) in enumerate(rows):
            n = (ca + cb + co + f) & 7
            want = int(n == 3 or (cell and n == 4))
            cnf.add(last[1][r] if want else -last[1][r])

    def cares(own):
        out = []
        for ca, cb, co, cell in tail_rows:
            t = (ca, cb)[own]
            if (cell == 1 and t == 0) or (cell == 0 and t == 3):
                continue
            out.append((ca, cb, co, cell))
        return out

    if dc:
        synth_tail("l", cares(0), 0)
        synth_tail("r", cares(1), 1, GTR if GTR is not None else GT)
    else:
        synth_tail("t", tail_rows, None)

    # every code gate is read by someone later (cheap symmetry breaking / no dead gates)
    names = [g[0] for g in gates]
    for gi, (name, _, _, _) in enumerate(gates):
        if not name.startswith("p"):
            continue
        users = []
        for (n2, cands2, sel2, _) in gates[gi + 1:]:
            if name in cands2:
                i = cands2.index(name)
                users += [sel2[k][i] for k in range(3)]
        cnf.add(*users)
    cnf.free = free
    return cnf, gates


def solve(cnf, timeout):
    with tempfile.NamedTemporaryFile("w", suffix=".cnf", delete=False) as f:
        cnf.write(f)
        path = f.name
    cmd = ["z3", "-dimacs", path]
    if timeout:
        cmd.insert(1, f"-T:{int(timeout)}")
    out = subprocess.run(cmd, capture_output=True, text=True).stdout
    status = "unknown"
    model = {}
    for line in out.splitlines():
        if line.startswith("s "):
            status = line[2:].strip()
        elif line.startswith("v "):
            for tok in line[2:].split():
                v = int(tok)
                if v:
                    model[abs(v)] = v > 0
        elif line.strip() in ("sat", "unsat", "timeout", "unknown"):
            status = line.strip()
        elif line and (line[0].isdigit() or line[0] == "-") and status in ("sat", "SATISFIABLE"):
            for tok in line.split():
                v = int(tok)
                if v:
                    model[abs(v)] = v > 0
    return status, model


def main():
    args = [a for i, a in enumerate(sys.argv[1:], 1) if not a.startswith("--") and sys.argv[i - 1] not in ("--timeout", "--gtr")]
    code, GP, GT = map(int, args[:3])
    dc = "--dc" in sys.argv
    timeout = None
    if "--timeout" in sys.argv:
        timeout = float(sys.argv[sys.argv.index("--timeout") + 1])
    raw = "--raw" in sys.argv
    gtr = int(sys.argv[sys.argv.index("--gtr") + 1]) if "--gtr" in sys.argv else None
    weld = "--weld" in sys.argv
    cnf, gates = build(code, GP, GT, dc, raw, gtr, weld)
    t0 = time.time()
    status, model = solve(cnf, timeout)
    dt = time.time() - t0
    print(f"{'weld ' if weld else ''}code {code} GP {GP} GT {GT} GTR {gtr} dc {dc} raw {raw}: vars {cnf.n} clauses {len(cnf.clauses)} -> {status} in {dt:.1f}s", flush=True)
    if model:
        for (tag, pl), fv in sorted(cnf.free.items()):
            print(f"  {tag}: F{pl}(f = 0..7) = {[int(model.get(v, False)) for v in fv]}")
        for name, cands, sel, lut in gates:
            ins = [cands[[model.get(v, False) for v in sel[k]].index(True)] for k in range(3)]
            lutv = sum(1 << i for i in range(8) if model.get(lut[i], False))
            print(f"  {name} = lop3<0x{lutv:02X}>({', '.join(ins)})")


if __name__ == "__main__":
    main()
