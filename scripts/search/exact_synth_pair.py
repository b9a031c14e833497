#!/usr/bin/env python
"""Exact synthesis (z3) of the vertical rule for TWO adjacent output rows at once.

Inputs: row sums A, B, C, D (two signals each, encoding as in exact_synth.py) and the centre bits
m1 (row B) and m2 (row C).  Outputs: next(B) from (A, B, C, m1) and next(C) from (B, C, D, m2).
Separately they cost 2 x 6 gates; a shared circuit with N < 12 gates would lower the cost per row.
usage: exact_synth_pair.py <row_encoding 0..2> <gates> [timeout_s]"""
import itertools
import sys
import z3

enc = int(sys.argv[1]); NG = int(sys.argv[2]); tmo = int(sys.argv[3]) if len(sys.argv) > 3 else 3600
CODES = [{0: (0, 0), 1: (1, 0), 2: (0, 1), 3: (1, 1)},
         {0: (0, 1), 1: (1, 0), 2: (0, 0), 3: (1, 1)},
         {0: (0, 1), 1: (0, 0), 2: (1, 0), 3: (1, 1)}][enc]
rows = []
for A, B, Cc, D, m1, m2 in itertools.product(range(4), range(4), range(4), range(4), range(2), range(2)):
    if (m1 and B == 0) or (m2 and Cc == 0):
        continue
    T1, T2 = A + B + Cc, B + Cc + D
    rows.append((CODES[A] + CODES[B] + CODES[Cc] + CODES[D] + (m1, m2),
                 int(T1 == 3 or (T1 == 4 and m1)), int(T2 == 3 or (T2 == 4 and m2))))
NI = 10
s = z3.SolverFor("QF_FD"); s.set("timeout", tmo * 1000)
sel = [[[z3.Bool(f"s_{i}_{k}_{j}") for j in range(NI + i)] for k in range(3)] for i in range(NG)]
lut = [[z3.Bool(f"l_{i}_{b}") for b in range(8)] for i in range(NG)]
val = [[z3.Bool(f"v_{i}_{r}") for r in range(len(rows))] for i in range(NG)]
op = [[[z3.Bool(f"o_{i}_{k}_{r}") for r in range(len(rows))] for k in range(3)] for i in range(NG)]
out1 = [z3.Bool(f"out1_{i}") for i in range(NG - 1)]      # which gate carries the first output
s.add(z3.PbEq([(x, 1) for x in out1], 1))
for i in range(NG):
    n = NI + i
    for k in range(3):
        s.add(z3.PbEq([(x, 1) for x in sel[i][k]], 1))
    for k in range(2):
        for j in range(n):
            for j2 in range(0, j + 1):
                s.add(z3.Or(z3.Not(sel[i][k][j]), z3.Not(sel[i][k + 1][j2])))
    if i < NG - 1:
        s.add(z3.Or(z3.Not(lut[i][0]), out1[i]))   # normal form unless the gate is an output
    for k, bit in enumerate((4, 2, 1)):
        s.add(z3.Or([z3.Xor(lut[i][b], lut[i][b | bit]) for b in range(8) if not b & bit]))
    for r, (bits, o1, o2) in enumerate(rows):
        for k in range(3):
            for j in range(n):
                src = z3.BoolVal(bool(bits[j])) if j < NI else val[j - NI][r]
                s.add(z3.Implies(sel[i][k][j], op[i][k][r] == src))
        for b in range(8):
            cond = [op[i][0][r] if b & 4 else z3.Not(op[i][0][r]), op[i][1][r] if b & 2 else z3.Not(op[i][1][r]),
                    op[i][2][r] if b & 1 else z3.Not(op[i][2][r])]
            s.add(z3.Implies(z3.And(*cond), val[i][r] == lut[i][b]))
for i in range(NG - 1):   # used by a later gate, or an output
    s.add(z3.Or([sel[i2][k][NI + i] for i2 in range(i + 1, NG) for k in range(3)] + [out1[i]]))
for r, (bits, o1, o2) in enumerate(rows):
    s.add(val[NG - 1][r] == z3.BoolVal(bool(o2)))
    for i in range(NG - 1):
        s.add(z3.Implies(out1[i], val[i][r] == z3.BoolVal(bool(o1))))
res = s.check()
print(f"pair, encoding {enc}, {NG} gates: {res}", flush=True)
if res == z3.sat:
    mdl = s.model()
    for i in range(NG):
        ins = [next(j for j in range(NI + i) if z3.is_true(mdl[sel[i][k][j]])) for k in range(3)]
        l = sum((1 << b) for b in range(8) if z3.is_true(mdl[lut[i][b]]))
        tag = " = out1" if i < NG - 1 and z3.is_true(mdl[out1[i]]) else (" = out2" if i == NG - 1 else "")
        print(f"  g{i} = LOP3<0x{l:02x}>({', '.join(('in%d' % j) if j < NI else ('g%d' % (j - NI)) for j in ins)}){tag}")
