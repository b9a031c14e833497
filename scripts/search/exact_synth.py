#!/usr/bin/env python
"""Exact synthesis of the vertical part of the B3/S23 rule as a LOP3 (3-input LUT) circuit with z3.

Inputs: the bit-sliced 3-cell sums of the rows above / at / below (each 0..3, two signals per row
in one of three symmetric encodings) and the centre bit m; m = 1 implies the middle sum >= 1.
Question answered: is there a circuit with N gates?   usage: exact_synth.py <row_encoding 0..2> <gates> [timeout_s]
"""
import sys
import itertools
import z3

enc = int(sys.argv[1]); NG = int(sys.argv[2]); tmo = int(sys.argv[3]) if len(sys.argv) > 3 else 3600
# value -> (signal0, signal1):  0: (parity, carry)  1: (parity, all-equal)  2: (carry, all-equal)
CODES = [{0: (0, 0), 1: (1, 0), 2: (0, 1), 3: (1, 1)},
         {0: (0, 1), 1: (1, 0), 2: (0, 0), 3: (1, 1)},
         {0: (0, 1), 1: (0, 0), 2: (1, 0), 3: (1, 1)}][enc]
rows = []   # (input bits tuple of 7, output)
for A, B, Nn, m in itertools.product(range(4), range(4), range(4), range(2)):
    if m and B == 0:
        continue
    T = A + B + Nn
    rows.append((CODES[A] + CODES[B] + CODES[Nn] + (m,), int(T == 3 or (T == 4 and m))))
NI = 7
s = z3.SolverFor("QF_FD")
s.set("timeout", tmo * 1000)
s.set("sat.random_seed", 7)
sel = [[[z3.Bool(f"s_{i}_{k}_{j}") for j in range(NI + i)] for k in range(3)] for i in range(NG)]
lut = [[z3.Bool(f"l_{i}_{b}") for b in range(8)] for i in range(NG)]
val = [[z3.Bool(f"v_{i}_{r}") for r in range(len(rows))] for i in range(NG)]
op = [[[z3.Bool(f"o_{i}_{k}_{r}") for r in range(len(rows))] for k in range(3)] for i in range(NG)]
for i in range(NG):
    n = NI + i
    for k in range(3):
        s.add(z3.PbEq([(x, 1) for x in sel[i][k]], 1))
    # operands strictly ordered (symmetry breaking; a LUT absorbs any permutation)
    for k in range(2):
        for j in range(n):
            for j2 in range(0, j + 1):
                s.add(z3.Or(z3.Not(sel[i][k][j]), z3.Not(sel[i][k + 1][j2])))
    if i < NG - 1:
        s.add(z3.Not(lut[i][0]))   # normal form: f(0,0,0) = 0 (a complement is absorbed downstream); not for the output gate
    for r, (bits, out) in enumerate(rows):
        for k in range(3):
            for j in range(n):
                src = z3.BoolVal(bool(bits[j])) if j < NI else val[j - NI][r]
                s.add(z3.Implies(sel[i][k][j], op[i][k][r] == src))
        for b in range(8):
            cond = [op[i][0][r] if b & 4 else z3.Not(op[i][0][r]), op[i][1][r] if b & 2 else z3.Not(op[i][1][r]),
                    op[i][2][r] if b & 1 else z3.Not(op[i][2][r])]
            s.add(z3.Implies(z3.And(*cond), val[i][r] == lut[i][b]))
# every gate depends on each of its three operands
for i in range(NG):
    for k, bit in enumerate((4, 2, 1)):
        s.add(z3.Or([z3.Xor(lut[i][b], lut[i][b | bit]) for b in range(8) if not b & bit]))
# every gate but the last is used by a later gate
for i in range(NG - 1):
    s.add(z3.Or([sel[i2][k][NI + i] for i2 in range(i + 1, NG) for k in range(3)]))
# output: last gate equals F or its complement is not allowed (we need F itself; drop the normal form there)
for r, (bits, out) in enumerate(rows):
    s.add(val[NG - 1][r] == z3.BoolVal(bool(out)))
res = s.check()
print(f"encoding {enc}, {NG} gates: {res}", flush=True)
if res == z3.sat:
    mdl = s.model()
    for i in range(NG):
        ins = [next(j for j in range(NI + i) if z3.is_true(mdl[sel[i][k][j]])) for k in range(3)]
        l = sum((1 << b) for b in range(8) if z3.is_true(mdl[lut[i][b]]))
        print(f"  g{i} = LOP3<0x{l:02x}>({', '.join(('in%d' % j) if j < NI else ('g%d' % (j - NI)) for j in ins)})")
