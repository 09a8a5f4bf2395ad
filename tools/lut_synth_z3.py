"""Exact synthesis of LUT3 networks with z3.  usage: synth.py <problem> <k>"""
import sys, itertools, time
import z3

def synth(n, rows, outs, k, timeout_s=None, fixed=None):
    """rows: list of input tuples (care rows only); outs: list of (list of 0/1 per row) per output; k gates.
    Returns list of gates [(i,j,l,lut)] + output signal indices, or None."""
    s = z3.Solver()
    if timeout_s: s.set("timeout", int(timeout_s*1000))
    R = len(rows)
    val = [[z3.BoolVal(bool(rows[r][i])) for r in range(R)] for i in range(n)]
    gates = []
    for g in range(k):
        m = n + g
        triples = list(itertools.combinations(range(m), 3))
        S = [z3.Bool(f"s_{g}_{t[0]}_{t[1]}_{t[2]}") for t in triples]
        lut = [z3.Bool(f"l_{g}_{b}") for b in range(8)]
        v = [z3.Bool(f"v_{g}_{r}") for r in range(R)]
        s.add(z3.PbEq([(x,1) for x in S], 1))
        # normal form: lut[0] = 0 (output polarity absorbed downstream) -- only valid for non-output gates; skip for safety on last
        for ti,(i,j,l) in enumerate(triples):
            for r in range(R):
                # v[r] == lut[idx] where idx from val bits: use If-chains
                a,b,c = val[i][r], val[j][r], val[l][r]
                # idx bit2=a bit1=b bit0=c (LOP3 convention for lop3(a,b,c))
                sel = z3.If(a, z3.If(b, z3.If(c, lut[7], lut[6]), z3.If(c, lut[5], lut[4])),
                               z3.If(b, z3.If(c, lut[3], lut[2]), z3.If(c, lut[1], lut[0])))
                s.add(z3.Implies(S[ti], v[r] == sel))
        # every non-last gate must be used later: (weak symmetry breaking skipped)
        val.append(v)
        gates.append((triples, S, lut))
    # outputs: each output is some gate's signal (choose among gates) -- allow any gate
    osel = []
    for oi, o in enumerate(outs):
        if len(outs) == 1:
            for r in range(R):
                s.add(val[n+k-1][r] == bool(o[r]))
            osel.append(None)
        else:
            O = [z3.Bool(f"o_{oi}_{g}") for g in range(k)]
            s.add(z3.PbEq([(x,1) for x in O], 1))
            for g in range(k):
                for r in range(R):
                    s.add(z3.Implies(O[g], val[n+g][r] == bool(o[r])))
            osel.append(O)
    t0 = time.time()
    res = s.check()
    dt = time.time()-t0
    if res != z3.sat:
        return str(res), dt
    m = s.model()
    out = []
    for g,(triples,S,lut) in enumerate(gates):
        for ti,t in enumerate(triples):
            if z3.is_true(m[S[ti]]):
                lv = sum((1<<b) for b in range(8) if z3.is_true(m.eval(lut[b], model_completion=True)))
                out.append((t, lv))
    oidx = []
    for oi,O in enumerate(osel):
        if O is None: oidx.append(n+k-1)
        else: oidx.append(n + [g for g in range(k) if z3.is_true(m[O[g]])][0])
    return (out, oidx), dt

def life_rule_rows():
    # inputs: a,u0,c0,d0,u1,c1,d1
    rows, o = [], []
    for r in itertools.product((0,1), repeat=7):
        a,u0,c0,d0,u1,c1,d1 = r
        S = u0+c0+d0+2*(u1+c1+d1)
        rows.append(r); o.append(int(S==3 or (a and S==4)))
    return 7, rows, [o]

def pair_tail_rows():
    # inputs: a, c0, c1, p0, p1, p2 ; N = c + P, P in 0..6
    rows, o = [], []
    for r in itertools.product((0,1), repeat=6):
        a,c0,c1,p0,p1,p2 = r
        P = p0+2*p1+4*p2
        if P == 7: continue
        N = c0+2*c1+P
        rows.append(r); o.append(int(N==3 or (a and N==4)))
    return 6, rows, [o]

if __name__ == "__main__":
    prob, k = sys.argv[1], int(sys.argv[2])
    n, rows, outs = {"rule": life_rule_rows, "pairtail": pair_tail_rows}[prob]()
    res, dt = synth(n, rows, outs, k)
    print(prob, k, res, f"{dt:.1f}s", flush=True)
