"""Exact synthesis of a single-output 3-input-LUT network for a named specification (CNF, solved by `z3 -dimacs`).

    python tools/lut_synth_generic_cnf.py <spec> <gates> [--timeout S]

A specification is a list of named input columns over its CARE rows plus the wanted output column; the solver
looks for `gates` LOP3 gates, each reading three earlier signals (inputs or gates), the last gate being the
output.  Same SSV-style encoding as tools/lut_synth_pair_cnf.py (selection variables one-hot per operand,
ordered operands to break symmetry, normal gates: lut[0] = 0 except for the output gate).

Specs
  weld_tail   the part of LifeWeld::Step (LifeWeld.hpp:169-186) after the two full adders over the column counts:
              inputs s0, k0, s1, k1 (3x3 sum = s0 + 2 (k0 + s1) + 4 k1, centre included), cell, f0, f1, f2;
              output (T0 ^ T2) & (T1 ^ T2) & (cell | T0) with T = (sum + f) mod 8.  Rows: every (left, mid, right)
              column-count triple in 0..3 with cell <= [mid >= 1] and mid == 3 => cell, every f in 0..7.
              The ripple form is 9 gates; 8 were found (csrc/next.cu uses them).
  weld_tail_rK  weld_tail with the frozen count RECODED: the three frozen planes are constants of a launch, so the kernel
              may park ANY K functions of (f0, f1, f2) instead of the bits themselves (K = 3, 4); the solver chooses
              them (8 free truth-table bits per plane, plane value at f = 0 fixed to 0: a complement folds into the LUTs).
  weld_full   the same function from the raw column counts (10 inputs, 768 care rows); csrc/next.cu spends 12 gates.
"""
import itertools
import subprocess
import sys
import tempfile
import time


def spec_weld_tail():
    rows = []
    for l, m, r in itertools.product(range(4), repeat=3):
        for cell in (0, 1):
            if cell and m < 1:
                continue
            if m == 3 and not cell:
                continue
            l0, l1, m0, m1, r0, r1 = l & 1, l >> 1, m & 1, m >> 1, r & 1, r >> 1
            s0, k0 = l0 ^ m0 ^ r0, int(l0 + m0 + r0 >= 2)
            s1, k1 = l1 ^ m1 ^ r1, int(l1 + m1 + r1 >= 2)
            total = l + m + r
            assert total == s0 + 2 * (k0 + s1) + 4 * k1
            for f in range(8):
                t = (total + f) & 7
                out = int(t == 3 or (t == 4 and cell))
                rows.append(((s0, k0, s1, k1, cell, f & 1, (f >> 1) & 1, f >> 2), out))
    uniq = {}
    for ins, out in rows:
        assert uniq.setdefault(ins, out) == out, "inconsistent care rows"
    names = ["s0", "k0", "s1", "k1", "cell", "f0", "f1", "f2"]
    return names, sorted(uniq.items())


def spec_weld_tail_recoded(k):
    """weld_tail with K free planes F0..F(K-1) of the frozen count f: their row entries are ('f', f) and build() gives
    every plane one variable per value of f."""
    names, rows = spec_weld_tail()
    out = {}
    for ins, o in rows:
        f = ins[5] | (ins[6] << 1) | (ins[7] << 2)
        out[ins[:5] + tuple(("f", f) for _ in range(k))] = o
    return names[:5] + [f"F{p}" for p in range(k)], sorted(out.items(), key=lambda kv: str(kv[0]))


def spec_weld_full_recoded(k):
    """weld_full with K free planes of the frozen count (see spec_weld_tail_recoded)."""
    names, rows = spec_weld_full()
    out = {}
    for ins, o in rows:
        f = ins[7] | (ins[8] << 1) | (ins[9] << 2)
        out[ins[:7] + tuple(("f", f) for _ in range(k))] = o
    return names[:7] + [f"F{p}" for p in range(k)], sorted(out.items(), key=lambda kv: str(kv[0]))


def spec_weld_full():
    """The whole per-cell function of LifeWeld::Step from the three column counts (12 gates in csrc/next.cu: two full
    adders + the 8-gate tail): inputs l0 l1 m0 m1 r0 r1 cell f0 f1 f2."""
    rows = {}
    for l, m, r in itertools.product(range(4), repeat=3):
        for cell in (0, 1):
            if (cell and m < 1) or (m == 3 and not cell):
                continue
            for f in range(8):
                t = (l + m + r + f) & 7
                rows[(l & 1, l >> 1, m & 1, m >> 1, r & 1, r >> 1, cell, f & 1, (f >> 1) & 1, f >> 2)] = int(t == 3 or (t == 4 and cell))
    return ["l0", "l1", "m0", "m1", "r0", "r1", "cell", "f0", "f1", "f2"], sorted(rows.items())


def _lut(tt, a, b, c):
    return (tt >> ((a << 2) | (b << 1) | c)) & 1


def _three_valued(side):
    """The hi-plane tail of the three-valued step FUSED with the combine (csrc/life_three.cuh): inputs = the 4-signal
    pair code of the (state | unknown) plane, the far column's count there, that plane's centre cell, the lo plane's
    (alive, A) signals and the centre's unknown bit; outputs next_on, next_unknown.  Today: 5-gate tail + 3-gate combine.
    Rows: every 3 x 3 neighbourhood over {off, on, unknown} (disjoint planes), for the pair's left or right column."""
    rows = {}
    for cells in itertools.product(range(3), repeat=9):  # 0 off, 1 on, 2 unknown; column-major: far, own, other
        col = [cells[0:3], cells[3:6], cells[6:9]]
        far, own, other = col
        centre = own[1]
        cs = [sum(1 for v in c if v == 1) for c in (far, own, other)]   # lo plane (state)
        ct = [sum(1 for v in c if v != 0) for c in (far, own, other)]   # hi plane (state | unknown)
        cell_s, cell_t, u = int(centre == 1), int(centre != 0), int(centre == 2)
        n_lo, n_hi = sum(cs), sum(ct)
        alive = lambda n, c: int(n == 3 or (c and n == 4))  # noqa: E731
        a_lo, al_lo = int(n_lo >= 3), alive(n_lo, cell_s)
        a_hi, al_hi = int(n_hi >= 3), alive(n_hi, cell_t)
        next_on = int((not u) and al_lo and al_hi)
        next_unknown = int(u or (a_hi and not (a_lo and not al_lo) and not (al_lo and al_hi)))
        # pair code of the hi plane: (a, b) = (left column of the pair, right column of the pair)
        a, b = (ct[1], ct[2]) if side == "left" else (ct[2], ct[1])
        a0, a1, b0, b1 = a & 1, a >> 1, b & 1, b >> 1
        p0 = _lut(0x56, a0, b0, b1)
        p1 = _lut(0xBE, a1, b1, p0)
        p2 = _lut(0xE4, a1, b1, p0)
        p3 = _lut(0x92, a0, b0, p1)
        key = (p0, p1, p2, p3, ct[0] & 1, ct[0] >> 1, cell_t, al_lo, a_lo, u)
        val = (next_on, next_unknown)
        assert rows.setdefault(key, val) == val, "inconsistent rows: the inputs do not determine the outputs"
    names = ["p0", "p1", "p2", "p3", "c0", "c1", "cell_t", "alive_lo", "A_lo", "u"]
    return names, sorted(rows.items())


def spec_life_col():
    """One column of the plain B3/S23 step from the three binary column counts and the centre cell, no sharing with the
    neighbouring column: a (left), b (own, contains the cell), c (right).  The step kernel spends 11 gates per column
    PAIR; two unshared columns of 5 gates each would beat it."""
    rows = {}
    for a, b, c in itertools.product(range(4), repeat=3):
        for cell in (0, 1):
            if (cell and b < 1) or (b == 3 and not cell):
                continue
            n = a + b + c
            rows[(a & 1, a >> 1, b & 1, b >> 1, c & 1, c >> 1, cell)] = int(n == 3 or (cell and n == 4))
    return ["a0", "a1", "b0", "b1", "c0", "c1", "cell"], sorted(rows.items())


SPECS = {"weld_tail": spec_weld_tail, "weld_tail_r3": lambda: spec_weld_tail_recoded(3), "weld_tail_r4": lambda: spec_weld_tail_recoded(4),
         "weld_full_r3": lambda: spec_weld_full_recoded(3), "weld_full": spec_weld_full, "life_col": spec_life_col,
         "three_hi_left": lambda: _three_valued("left"), "three_hi_right": lambda: _three_valued("right")}


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


def build(names, rows, gates):
    cnf = Cnf()
    n_in, n_rows = len(names), len(rows)
    sig = []  # per signal: per-row constant or variable
    free = {}  # input index -> its 8 variables (one per value of the recoded quantity)
    for i in range(n_in):
        if isinstance(rows[0][0][i], tuple):
            free[i] = [cnf.var() for _ in range(8)]
            cnf.add(-free[i][0])
            sig.append([("v", free[i][rows[r][0][i][1]]) for r in range(n_rows)])
        else:
            sig.append([("c", rows[r][0][i]) for r in range(n_rows)])
    meta = []
    for g in range(gates):
        m = len(sig)
        sel = [[cnf.var() for _ in range(m)] for _ in range(3)]
        for k in range(3):
            cnf.exactly_one(sel[k])
        for k in range(2):  # operand k+1 strictly after operand k
            for i in range(m):
                for j in range(0, i + 1):
                    cnf.add(-sel[k][i], -sel[k + 1][j])
        if g > 0:  # every gate but the first must be used later or be the output: weak pruning, skip
            pass
        lut = [cnf.var() for _ in range(8)]
        n_out = len(rows[0][1]) if isinstance(rows[0][1], tuple) else 1
        if g < gates - n_out:
            cnf.add(-lut[0])
        out = [cnf.var() for _ in range(n_rows)]
        for r in range(n_rows):
            x = [cnf.var() for _ in range(3)]
            for k in range(3):
                for i in range(m):
                    kind, v = sig[i][r]
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
            want = rows[r][1] if isinstance(rows[r][1], tuple) else (rows[r][1],)
            k_out = len(want)
            if g >= gates - k_out:  # the last k gates are the outputs, in order
                w = want[g - (gates - k_out)]
                cnf.add(out[r] if w else -out[r])
        sig.append([("v", out[r]) for r in range(n_rows)])
        meta.append((sel, lut))
    # symmetry breaking: every gate that is not an output feeds some later gate
    n_out = len(rows[0][1]) if isinstance(rows[0][1], tuple) else 1
    for g in range(gates - n_out):
        users = [meta[h][0][k][n_in + g] for h in range(g + 1, gates) for k in range(3)]
        cnf.add(*users)
    cnf.free = free
    return cnf, meta


def solve(cnf, timeout):
    with tempfile.NamedTemporaryFile("w", suffix=".cnf", delete=False) as f:
        f.write(f"p cnf {cnf.n} {len(cnf.clauses)}\n")
        for c in cnf.clauses:
            f.write(" ".join(map(str, c)) + " 0\n")
        path = f.name
    t0 = time.time()
    try:
        out = subprocess.run(["z3", "-dimacs", f"-T:{timeout}", path], capture_output=True, text=True, timeout=timeout + 60).stdout
    except subprocess.TimeoutExpired:
        return "timeout", None, time.time() - t0
    dt = time.time() - t0
    if out.startswith("s SATISFIABLE") or out.startswith("sat"):
        model = set()
        for line in out.split("\n"):
            if line.startswith("v "):
                model.update(int(t) for t in line[2:].split() if int(t) > 0)
            elif line and line[0] in "-0123456789":
                model.update(int(t) for t in line.split() if int(t) > 0)
        return "sat", model, dt
    if "UNSAT" in out.upper():
        return "unsat", None, dt
    return "unknown/timeout", None, dt


def main():
    spec, gates = sys.argv[1], int(sys.argv[2])
    timeout = int(sys.argv[sys.argv.index("--timeout") + 1]) if "--timeout" in sys.argv else 3600
    names, rows = SPECS[spec]()
    cnf, meta = build(names, rows, gates)
    print(f"{spec}: {len(names)} inputs, {len(rows)} care rows, {gates} gates -> {cnf.n} vars, {len(cnf.clauses)} clauses", flush=True)
    status, model, dt = solve(cnf, timeout)
    print(f"{spec} gates={gates}: {status} in {dt:.0f} s", flush=True)
    if status == "sat":
        labels = list(names)
        for i, vs in sorted(cnf.free.items()):
            print(f"  {names[i]}(f = 0..7) = {[int(v in model) for v in vs]}")
        for g, (sel, lut) in enumerate(meta):
            ops = [labels[[i for i, v in enumerate(sel[k]) if v in model][0]] for k in range(3)]
            tt = sum(1 << idx for idx in range(8) if lut[idx] in model)
            # LOP3 immediate convention: a = 0xF0, b = 0xCC, c = 0xAA -> bit (a<<2 | b<<1 | c)
            print(f"  g{g} = lop3<0x{tt:02X}>({ops[0]}, {ops[1]}, {ops[2]})")
            labels.append(f"g{g}")


if __name__ == "__main__":
    main()
