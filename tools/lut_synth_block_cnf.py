"""Exact synthesis of a two-column block of the B3/S23 step over RAW cells (CNF, solved by `z3 -dimacs`).

    python tools/lut_synth_block_cnf.py <GS> <GL> <GR> [--timeout S]

A superset of tools/lut_synth_pair_cnf.py: the inputs are the six raw cells (l, a, r = rows y-1, y, y+1)
of columns x and x+1 and the binary 2-bit counts of the far columns x-1 (CL) and x+2 (CR).  The four
count bits of the block's own columns (xor3 / maj of their triples) are given for free -- they have
to exist anyway, as the far inputs of the neighbouring blocks (4 LOP3, not counted here).  GS shared
gates read raw cells, counts and earlier shared gates; GL (GR) gates of the left (right) tail also
read CL (CR) and earlier gates of the same tail; the last gate of each tail must equal the next
state of its column's centre cell.  Every raw combination is a care row, so the don't-cares of the
count formulation (cell = 1 implies count >= 1 ...) are built in.
Cost of a generation per two half-words = 4 + GS + GL + GR LOP3 (today 4 + 3 + 4 + 4 = 15).
"""
import itertools
import subprocess
import sys
import tempfile
import time


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


RAW = list(itertools.product((0, 1), repeat=6))  # la, aa, ra, lb, ab, rb


def build(GS, GL, GR):
    cnf = Cnf()
    gates = []

    def make_gate(name, cands, nrows, value_of, is_output):
        m = len(cands)
        sel = [[cnf.var() for _ in range(m)] for _ in range(3)]
        for k in range(3):
            cnf.exactly_one(sel[k])
        for k in range(2):
            for i in range(m):
                for j in range(0, i + 1):
                    cnf.add(-sel[k][i], -sel[k + 1][j])
        lut = [cnf.var() for _ in range(8)]
        if not is_output:
            cnf.add(-lut[0])
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
                bits = [(idx >> 2) & 1, (idx >> 1) & 1, idx & 1]
                pre = [(-x[k] if bits[k] else x[k]) for k in range(3)]
                cnf.add(*pre, -lut[idx], out[r])
                cnf.add(*pre, lut[idx], -out[r])
        gates.append((name, [c[0] for c in cands], sel, lut))
        return out

    # shared domain: 64 raw rows
    names = ["la", "aa", "ra", "lb", "ab", "rb"]
    shared = [(names[i], ("const", [row[i] for row in RAW], "s")) for i in range(6)]
    ta = [row[0] + row[1] + row[2] for row in RAW]
    tb = [row[3] + row[4] + row[5] for row in RAW]
    shared += [("A0", ("const", [t & 1 for t in ta], "s")), ("A1", ("const", [t >> 1 for t in ta], "s")),
               ("B0", ("const", [t & 1 for t in tb], "s")), ("B1", ("const", [t >> 1 for t in tb], "s"))]

    def sval(sig, r):
        kind, v, _ = sig
        return ("c", v[r]) if kind == "const" else ("v", v[r])

    for g in range(GS):
        out = make_gate(f"s{g}", list(shared), len(RAW), sval, False)
        shared.append((f"s{g}", ("var", out, "s")))

    def tail(tag, G, own_cell_index, own_t, other_t):
        rows = [(ri, c) for ri in range(len(RAW)) for c in range(4)]

        def value(sig, r):
            kind, v, dom = sig
            idx = rows[r][0] if dom == "s" else r
            return ("c", v[idx]) if kind == "const" else ("v", v[idx])

        cands = [("C0", ("const", [c & 1 for _, c in rows], "t")), ("C1", ("const", [c >> 1 for _, c in rows], "t"))]
        cands += list(shared)
        last = None
        for g in range(G):
            out = make_gate(f"{tag}{g}", list(cands), len(rows), value, g == G - 1)
            cands.append((f"{tag}{g}", ("var", out, "t")))
            last = out
        for r, (ri, c) in enumerate(rows):
            n = ta[ri] + tb[ri] + c
            cell = RAW[ri][own_cell_index]
            want = int(n == 3 or (cell and n == 4))
            cnf.add(last[r] if want else -last[r])

    tail("l", GL, 1, ta, tb)
    tail("r", GR, 4, tb, ta)
    for gi, (name, _, _, _) in enumerate(gates):
        if not name.startswith("s"):
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
        f.write(f"p cnf {cnf.n} {len(cnf.clauses)}\n")
        for c in cnf.clauses:
            f.write(" ".join(map(str, c)) + " 0\n")
        path = f.name
    cmd = ["z3", "-dimacs", path]
    if timeout:
        cmd.insert(1, f"-T:{int(timeout)}")
    out = subprocess.run(cmd, capture_output=True, text=True).stdout
    status, model = "unknown", {}
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
    return status, model


def main():
    GS, GL, GR = map(int, sys.argv[1:4])
    timeout = float(sys.argv[sys.argv.index("--timeout") + 1]) if "--timeout" in sys.argv else None
    cnf, gates = build(GS, GL, GR)
    t0 = time.time()
    status, model = solve(cnf, timeout)
    print(f"block GS {GS} GL {GL} GR {GR}: vars {cnf.n} clauses {len(cnf.clauses)} -> {status} in {time.time() - t0:.1f}s", flush=True)
    if model:
        for name, cands, sel, lut in gates:
            ins = [cands[[model.get(v, False) for v in sel[k]].index(True)] for k in range(3)]
            lutv = sum(1 << i for i in range(8) if model.get(lut[i], False))
            print(f"  {name} = lop3<0x{lutv:02X}>({', '.join(ins)})")


if __name__ == "__main__":
    main()
