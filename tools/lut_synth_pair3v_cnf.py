"""Two-output variant of tools/lut_synth_pair_cnf.py for the three-valued step: the tail of a
column must deliver TWO signals of the 3x3 sum N (centre included) -- chosen with --outs from
  alive = [N == 3] | cell & [N == 4],  A = [N >= 3],  B = [N >= 5] | ~cell & [N >= 4]
-- as its last two gates (the first named is the LAST gate).  Everything else as in the parent tool:

Exact synthesis of the pair-shared B3/S23 select as a plain CNF (SSV encoding), solved by `z3 -dimacs`.

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


OUTS = ['alive', 'A']
JOINT = None
JOINT_ORDER = 0


def want_of(name, n, cell):
    if name == 'alive':
        return int(n == 3 or (cell and n == 4))
    if name == 'A':
        return int(n >= 3)
    if name == 'B':
        return int(n >= 5 or (not cell and n >= 4))
    raise ValueError(name)


def build(code, GP, GT, dc, raw=False, GTR=None):
    enc = CODES[code]
    cnf = Cnf()
    pair_rows = list(itertools.product(range(4), range(4)))            # (ca, cb)
    tail_rows = list(itertools.product(range(4), range(4), range(4), (0, 1)))  # (ca, cb, co, cell)

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
        idx_pair = [pair_rows.index((ca, cb)) for ca, cb, co, cell in rows]

        def value(sig, r):
            kind, v, dom = sig
            if dom == "pair":
                rr = idx_pair[r]
                return ("c", v[rr]) if kind == "const" else ("v", v[rr])
            return ("c", v[r]) if kind == "const" else ("v", v[r])

        cands = [("C0", ("const", [enc[co][0] for ca, cb, co, cell in rows], "tail")),
                 ("C1", ("const", [enc[co][1] for ca, cb, co, cell in rows], "tail")),
                 ("cell", ("const", [cell for ca, cb, co, cell in rows], "tail"))]
        cands += [(n, (s[0], s[1], "pair")) for n, s in (code_sigs if raw else code_gate_sigs)]
        sigs = []
        for g in range(GT):
            s = make_gate(f"{tag}{g}", list(cands), len(rows), value, g >= GT - len(OUTS))
            cands.append((f"{tag}{g}", (s[0], s[1], "tail")))
            sigs.append(s)
        for oi, oname in enumerate(OUTS):
            sig = sigs[GT - 1 - oi]
            for r, (ca, cb, co, cell) in enumerate(rows):
                want = want_of(oname, ca + cb + co, cell)
                cnf.add(sig[1][r] if want else -sig[1][r])

    def cares(own):
        out = []
        for ca, cb, co, cell in tail_rows:
            t = (ca, cb)[own]
            if (cell == 1 and t == 0) or (cell == 0 and t == 3):
                continue
            out.append((ca, cb, co, cell))
        return out

    def synth_joint(tag, own, GT):
        """--joint: the tail of the hi plane T = state | unknown also reads u (the unknown centre) and the two
        signals alive_lo, A_lo the lo plane's tail delivered, and ends in next_on (last gate) and next_unknown."""
        rows = []
        for ca, cb, co, cell in cares(own):
            n = ca + cb + co
            hc = 0 if n < 3 else (2 if (n >= 5 or (not cell and n >= 4)) else 1)
            for u in (0, 1):
                if u and not cell:
                    continue
                for lc in (0, 1, 2):
                    if not u and lc > hc:
                        continue
                    if u:
                        on, unk = 0, 1
                    else:
                        on = int(lc == 1 and hc == 1)
                        unk = int(not on and not (lc == hc))
                    rows.append((ca, cb, co, cell, u, lc, on, unk))
        idx_pair = [pair_rows.index((r[0], r[1])) for r in rows]

        def value(sig, r):
            kind, v, dom = sig
            if dom == "pair":
                rr = idx_pair[r]
                return ("c", v[rr]) if kind == "const" else ("v", v[rr])
            return ("c", v[r]) if kind == "const" else ("v", v[r])

        cands = [("C0", ("const", [enc[r[2]][0] for r in rows], "tail")),
                 ("C1", ("const", [enc[r[2]][1] for r in rows], "tail")),
                 ("cellT", ("const", [r[3] for r in rows], "tail")),
                 ("u", ("const", [r[4] for r in rows], "tail")),
                 ("alive_lo", ("const", [int(r[5] == 1) for r in rows], "tail")),
                 ("A_lo", ("const", [int(r[5] >= 1) for r in rows], "tail"))]
        cands += [(n, (s[0], s[1], "pair")) for n, s in (code_sigs if raw else code_gate_sigs)]
        sigs = []
        for g in range(GT):
            s = make_gate(f"{tag}{g}", list(cands), len(rows), value, g >= GT - 2)
            cands.append((f"{tag}{g}", (s[0], s[1], "tail")))
            sigs.append(s)
        order = (6, 7) if JOINT_ORDER == 0 else (7, 6)
        for oi, col in enumerate(order):
            sig = sigs[GT - 1 - oi]
            for r, row in enumerate(rows):
                cnf.add(sig[1][r] if row[col] else -sig[1][r])

    if JOINT is not None:
        synth_joint("l", 0, GT)
        synth_joint("r", 1, GTR if GTR is not None else GT)
    elif dc:
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
    global OUTS, JOINT, JOINT_ORDER
    if "--joint" in sys.argv:
        JOINT = True
        JOINT_ORDER = int(sys.argv[sys.argv.index("--joint") + 1])
    if "--outs" in sys.argv:
        OUTS = sys.argv[sys.argv.index("--outs") + 1].split(",")
    args = [a for i, a in enumerate(sys.argv[1:], 1) if not a.startswith("--") and sys.argv[i - 1] not in ("--timeout", "--gtr", "--outs", "--joint")]
    code, GP, GT = map(int, args[:3])
    dc = "--dc" in sys.argv
    timeout = None
    if "--timeout" in sys.argv:
        timeout = float(sys.argv[sys.argv.index("--timeout") + 1])
    raw = "--raw" in sys.argv
    gtr = int(sys.argv[sys.argv.index("--gtr") + 1]) if "--gtr" in sys.argv else None
    cnf, gates = build(code, GP, GT, dc, raw, gtr)
    t0 = time.time()
    status, model = solve(cnf, timeout)
    dt = time.time() - t0
    print(f"joint {JOINT} order {JOINT_ORDER} outs {OUTS} code {code} GP {GP} GT {GT} GTR {gtr} dc {dc} raw {raw}: vars {cnf.n} clauses {len(cnf.clauses)} -> {status} in {dt:.1f}s", flush=True)
    if model:
        for name, cands, sel, lut in gates:
            ins = [cands[[model.get(v, False) for v in sel[k]].index(True)] for k in range(3)]
            lutv = sum(1 << i for i in range(8) if model.get(lut[i], False))
            print(f"  {name} = lop3<0x{lutv:02X}>({', '.join(ins)})")


if __name__ == "__main__":
    main()
