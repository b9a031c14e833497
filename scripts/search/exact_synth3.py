#!/usr/bin/env python
"""Variant of exact_synth.py: THREE signals per row (three symmetric functions of the row's 3-cell
sum, i.e. 3 horizontal LOP3 per row instead of 2).  Is there a vertical circuit with N gates over
the 9 row signals + the centre bit?  3 + N < 2 + 6 needs N <= 4.
usage: exact_synth3.py <gates> [timeout_s_per_case]"""
import itertools
import sys
import z3

NG = int(sys.argv[1]); tmo = int(sys.argv[2]) if len(sys.argv) > 2 else 600
# symmetric functions of the sum 0..3 as 4-bit masks, non-constant, up to complement (bit 0 clear)
FUNCS = [f for f in range(1, 15) if not f & 1]
def solve(trip):
    code = lambda v: tuple((f >> v) & 1 for f in trip)
    rows = []
    for A, B, Nn, m in itertools.product(range(4), range(4), range(4), range(2)):
        if m and B == 0:
            continue
        T = A + B + Nn
        rows.append((code(A) + code(B) + code(Nn) + (m,), int(T == 3 or (T == 4 and m))))
    # inconsistent (non-injective) encodings are simply unsatisfiable
    NI = 10
    s = z3.SolverFor("QF_FD"); s.set("timeout", tmo * 1000)
    sel = [[[z3.Bool(f"s_{i}_{k}_{j}") for j in range(NI + i)] for k in range(3)] for i in range(NG)]
    lut = [[z3.Bool(f"l_{i}_{b}") for b in range(8)] for i in range(NG)]
    val = [[z3.Bool(f"v_{i}_{r}") for r in range(len(rows))] for i in range(NG)]
    op = [[[z3.Bool(f"o_{i}_{k}_{r}") for r in range(len(rows))] for k in range(3)] for i in range(NG)]
    for i in range(NG):
        n = NI + i
        for k in range(3):
            s.add(z3.PbEq([(x, 1) for x in sel[i][k]], 1))
        for k in range(2):
            for j in range(n):
                for j2 in range(0, j + 1):
                    s.add(z3.Or(z3.Not(sel[i][k][j]), z3.Not(sel[i][k + 1][j2])))
        if i < NG - 1:
            s.add(z3.Not(lut[i][0]))
        for r, (bits, out) in enumerate(rows):
            for k in range(3):
                for j in range(n):
                    src = z3.BoolVal(bool(bits[j])) if j < NI else val[j - NI][r]
                    s.add(z3.Implies(sel[i][k][j], op[i][k][r] == src))
            for b in range(8):
                cond = [op[i][0][r] if b & 4 else z3.Not(op[i][0][r]), op[i][1][r] if b & 2 else z3.Not(op[i][1][r]),
                        op[i][2][r] if b & 1 else z3.Not(op[i][2][r])]
                s.add(z3.Implies(z3.And(*cond), val[i][r] == lut[i][b]))
    for i in range(NG - 1):
        s.add(z3.Or([sel[i2][k][NI + i] for i2 in range(i + 1, NG) for k in range(3)]))
    for r, (bits, out) in enumerate(rows):
        s.add(val[NG - 1][r] == z3.BoolVal(bool(out)))
    res = s.check()
    desc = ""
    if res == z3.sat:
        mdl = s.model()
        for i in range(NG):
            ins = [next(j for j in range(NI + i) if z3.is_true(mdl[sel[i][k][j]])) for k in range(3)]
            l = sum((1 << b) for b in range(8) if z3.is_true(mdl[lut[i][b]]))
            desc += f"\n  g{i} = LOP3<0x{l:02x}>({', '.join(('in%d' % j) if j < NI else ('g%d' % (j - NI)) for j in ins)})"
    return res, desc

part, nparts = (int(sys.argv[3]), int(sys.argv[4])) if len(sys.argv) > 4 else (0, 1)
for idx, trip in enumerate(itertools.combinations(FUNCS, 3)):
    if idx % nparts != part:
        continue
    res, desc = solve(trip)
    print(f"row functions {[bin(f) for f in trip]}, {NG} gates: {res}{desc}", flush=True)
