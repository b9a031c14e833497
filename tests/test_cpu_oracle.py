"""CPU-only tests: the bit-packed oracle against an independent numpy implementation and
against the UNMODIFIED serial reference program, host text I/O, generator, topology."""
import os

import numpy as np
import pytest

from gol_b200.ops import cpu
from gol_b200.utils import patterns, reference

SIZES = [(1, 1), (1, 5), (5, 1), (2, 2), (2, 3), (3, 7), (30, 30), (63, 9), (64, 64), (65, 33), (130, 70), (200, 3)]


@pytest.mark.parametrize("w,h", SIZES)
def test_step_matches_numpy_roll(w, h):
    rng = np.random.default_rng(w * 1000 + h)
    cells = (rng.random((h, w)) < 0.45).astype(np.uint8)
    if w >= 3 and h >= 3:  # np.roll double-counts differently from the reference only below 3
        expect = cells.copy()
        for _ in range(4):
            expect = patterns.conv_reference_step(expect)
        got = cpu.unpack(cpu.step(cpu.pack(cells), w, 4), w)
        assert np.array_equal(got, expect)
    # packing round trip and canonical tail for every size
    p = cpu.pack(cells)
    assert np.array_equal(cpu.unpack(p, w), cells)
    q = cpu.step(p, w, 3)
    if w % 64:
        assert int(q[:, -1].max()) < (1 << (w % 64))


def _run_reference(exe, cells, tmp_path):
    h, w = cells.shape
    inp = tmp_path / "in.txt"
    inp.write_bytes(cpu.format_text(cpu.pack(cells), w))
    return reference.run_program(exe, w, h, str(inp), str(tmp_path), "game_output.out", timeout=120)


FIXTURES = {
    "empty": (8, []), "cell": (8, [("cell", 3, 3)]), "block": (8, [("block", 2, 2)]),
    "tromino": (8, [("tromino", 2, 2)]), "blinker": (8, [("blinker", 2, 3)]),
    "beehive": (10, [("beehive", 3, 3)]), "diag3": (9, [("diag3", 3, 3)]),
    "glider": (12, [("glider", 1, 1)]), "pulsar": (17, [("pulsar", 2, 2)]),
    "rpent": (24, [("rpentomino", 10, 10)]), "seam": (10, [("block", 9, 9), ("blinker", 8, 0)]),
}


@pytest.mark.parametrize("name", sorted(FIXTURES))
def test_run_matches_unmodified_reference(name, ref_serial, tmp_path):
    n, stamps = FIXTURES[name]
    cells = patterns.grid_with(n, n, *stamps)
    gens, _, out, _ = _run_reference(ref_serial, cells, tmp_path)
    final, res = cpu.run(cpu.pack(cells), n)
    assert res.generations == gens
    assert cpu.format_text(final, n) == out


@pytest.mark.parametrize("n,seed", [(1, 1), (2, 2), (3, 3), (5, 4), (31, 5), (64, 6), (97, 7)])
def test_run_random_matches_unmodified_reference(n, seed, ref_serial, tmp_path):
    rng = np.random.default_rng(seed)
    cells = (rng.random((n, n)) < 0.4).astype(np.uint8)
    gens, _, out, _ = _run_reference(ref_serial, cells, tmp_path)
    final, res = cpu.run(cpu.pack(cells), n)
    assert res.generations == gens
    assert cpu.format_text(final, n) == out


def test_expected_generation_counts():
    # empty 0, single cell 1, block 2, tromino 2, blinker 1000
    def gens(n, *stamps, **kw):
        return cpu.run(cpu.pack(patterns.grid_with(n, n, *stamps)), n, **kw)[1]
    assert gens(8).generations == 0 and gens(8).exit_reason == "empty"
    assert gens(8, ("cell", 3, 3)).generations == 1
    assert gens(8, ("block", 2, 2)).generations == 2
    assert gens(8, ("tromino", 2, 2)).generations == 2
    assert gens(8, ("blinker", 2, 3)).generations == 1000
    assert gens(17, ("pulsar", 2, 2)).generations == 1000   # period 3 must not trigger the /3 check
    assert gens(8, ("block", 2, 2), similarity_frequency=0).generations == 1000
    assert gens(8, ("block", 2, 2), gen_limit=1).generations == 1
    r = gens(8, ("block", 2, 2), similarity_frequency=5)
    assert (r.generations, r.exit_reason) == (4, "similar")


def test_text_roundtrip_and_errors(tmp_path):
    g = cpu.random_grid(70, 13, 3)
    text = cpu.format_text(g, 70)
    assert len(text) == 13 * 71 and set(text) <= set(b"01\n")
    assert np.array_equal(cpu.parse_text(text, 70, 13), g)
    # line breaks may be anywhere (fgetc reader semantics)
    assert np.array_equal(cpu.parse_text(text.replace(b"\n", b""), 70, 13), g)
    with pytest.raises(Exception, match="too short"):
        cpu.parse_text(text[:-80], 70, 13)
    bad = bytearray(text)
    bad[5] = ord("x")
    with pytest.raises(Exception, match="invalid byte"):
        cpu.parse_text(bytes(bad), 70, 13)
    lenient = cpu.parse_text(bytes(bad), 70, 13, strict=False)
    assert cpu.unpack(lenient, 70)[0, 5] == 0
    path = tmp_path / "g.txt"
    cpu.write_file(str(path), g, 70)
    assert path.read_bytes() == text
    assert np.array_equal(cpu.read_file(str(path), 70, 13), g)


def test_random_generator(tmp_path):
    a, b = cpu.random_grid(300, 40, 11), cpu.random_grid(300, 40, 11)
    assert np.array_equal(a, b) and not np.array_equal(a, cpu.random_grid(300, 40, 12))
    frac = cpu.unpack(a, 300).mean()
    assert 0.45 < frac < 0.55
    path = tmp_path / "r.txt"
    cpu.random_text_file(str(path), 300, 40, 11)
    assert path.read_bytes() == cpu.format_text(a, 300)


def test_generated_file_is_accepted_by_reference(ref_serial, tmp_path):
    path = tmp_path / "r.txt"
    cpu.random_text_file(str(path), 48, 48, 5)
    gens, _, out, _ = reference.run_program(ref_serial, 48, 48, str(path), str(tmp_path), "game_output.out")
    final, res = cpu.run(cpu.random_grid(48, 48, 5), 48)
    assert gens == res.generations and out == cpu.format_text(final, 48)


def test_pbm_dump(tmp_path):
    g = cpu.pack(patterns.grid_with(10, 4, ("block", 0, 0)))
    p = tmp_path / "g.pbm"
    cpu.write_pbm(str(p), g, 10)
    data = p.read_bytes()
    assert data.startswith(b"P4\n10 4\n") and len(data) == len(b"P4\n10 4\n") + 4 * 2
    assert data[len(b"P4\n10 4\n")] == 0b11000000


def test_topology():
    from gol_b200.parallel import Topology, auto_grid, near_square_grid, OPPOSITE
    g = near_square_grid
    assert g(8) == (2, 4) and g(4) == (2, 2) and g(2) == (1, 2) and g(1) == (1, 1)
    assert auto_grid(8) == (8, 1) and auto_grid(2) == (2, 1) and auto_grid(1) == (1, 1) and auto_grid(64) == (8, 8)
    t = Topology(65536, 65536, 2, 4)
    assert t.neighbors(0) == [4, 4, 3, 1, 7, 5, 7, 5]
    for g in range(t.size):
        nb = t.neighbors(g)
        for d, n in enumerate(nb):
            assert t.neighbors(n)[OPPOSITE[d]] == g
    ext = [t.extent(g) for g in range(t.size)]
    assert sum(w * h for _, _, w, h in ext) == 65536 * 65536
    t2 = Topology(840, 840, 3, 3)
    assert [e[0] for e in (t2.extent(0), t2.extent(1), t2.extent(2))] == [0, 256, 576]
    assert t2.extent(2)[2] == 840 - 576


def test_kernel_rule_circuit_truth_table():
    """The LOP3 circuit of life_rule() in csrc/life_march.cuh (read from the source) equals B3/S23
    on all 4*4*4*2 combinations of the three row sums and the centre bit (reference rule: S:60-101)."""
    import itertools
    import os
    import re
    src = open(os.path.join(os.path.dirname(__file__), "..", "csrc", "life_march.cuh")).read()
    body = src[src.index("life_rule(uint32_t a0"):]
    body = body[:body.index("\n}\n")]
    stmts = re.findall(r"(\w+) = (xor3|maj3|lop3<0x([0-9A-Fa-f]+)>)\((\w+), (\w+), (\w+)\);", body)
    ret = re.search(r"return lop3<0x([0-9A-Fa-f]+)>\((\w+), (\w+), (\w+)\);", body)
    assert len(stmts) == 5 and ret, "life_rule is expected to be 6 LOP3"

    def lop3(lut, a, b, c):
        return (lut >> ((a << 2) | (b << 1) | c)) & 1

    for A, B, Nn, m in itertools.product(range(4), range(4), range(4), range(2)):
        if m and B == 0:
            continue   # the middle row sum includes the centre cell
        env = {"a0": A & 1, "a1": A >> 1, "b0": B & 1, "b1": B >> 1, "n0": Nn & 1, "n1": Nn >> 1, "m": m}
        for name, fn, lut, x, y, z in stmts:
            l = 0x96 if fn == "xor3" else 0xE8 if fn == "maj3" else int(lut, 16)
            env[name] = lop3(l, env[x], env[y], env[z])
        out = lop3(int(ret.group(1), 16), env[ret.group(2)], env[ret.group(3)], env[ret.group(4)])
        T = A + B + Nn
        assert out == int(T == 3 or (T == 4 and m)), (A, B, Nn, m)
