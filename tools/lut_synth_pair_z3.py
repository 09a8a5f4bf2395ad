"""pair-shared synthesis: P gates over pair counts, tail gates over (P outs, outer count, cell)."""
import sys, itertools, time
import z3

CODES = {0: {0:(0,0),1:(1,0),2:(0,1),3:(1,1)},   # binary (xor, maj)
         1: {0:(0,0),1:(1,0),2:(1,1),3:(0,1)},   # gray (nae, maj)
         2: {0:(0,0),1:(1,1),2:(1,0),3:(0,1)}}   # third class

def run(code, GP, GT):
    enc = CODES[code]
    rows, outv = [], []
    for ca, cb, co, cell in itertools.product(range(4), range(4), range(4), (0,1)):
        N = ca+cb+co
        rows.append(enc[ca]+enc[cb]+enc[co]+(cell,))
        outv.append(int(N==3 or (cell and N==4)))
    n = 7; R = len(rows)
    s = z3.Solver()
    val = [[z3.BoolVal(bool(rows[r][i])) for r in range(R)] for i in range(n)]
    ginfo = []
    for g in range(GP+GT):
        if g < GP: allowed = [0,1,2,3] + [n+h for h in range(g)]
        else: allowed = [4,5,6] + [n+h for h in range(GP)] + [n+h for h in range(GP, g)]
        triples = list(itertools.combinations(allowed, 3))
        S = [z3.Bool(f"s_{g}_{ti}") for ti in range(len(triples))]
        lut = [z3.Bool(f"l_{g}_{b}") for b in range(8)]
        v = [z3.Bool(f"v_{g}_{r}") for r in range(R)]
        s.add(z3.PbEq([(x,1) for x in S], 1))
        for ti,(i,j,l) in enumerate(triples):
            for r in range(R):
                a,b,c = val[i][r], val[j][r], val[l][r]
                sel = z3.If(a, z3.If(b, z3.If(c, lut[7], lut[6]), z3.If(c, lut[5], lut[4])),
                               z3.If(b, z3.If(c, lut[3], lut[2]), z3.If(c, lut[1], lut[0])))
                s.add(z3.Implies(S[ti], v[r] == sel))
        val.append(v); ginfo.append((triples,S,lut))
    for r in range(R): s.add(val[n+GP+GT-1][r] == bool(outv[r]))
    t0=time.time(); res = s.check(); dt=time.time()-t0
    if res != z3.sat: return str(res), dt
    m = s.model(); out=[]
    for g,(triples,S,lut) in enumerate(ginfo):
        for ti,t in enumerate(triples):
            if z3.is_true(m[S[ti]]):
                out.append((t, sum((1<<b) for b in range(8) if z3.is_true(m.eval(lut[b], model_completion=True)))))
    return out, dt

if __name__ == "__main__":
    code, GP, GT = map(int, sys.argv[1:4])
    print("code",code,"GP",GP,"GT",GT, *run(code,GP,GT), flush=True)
