"""GPU tests of the engine: strip kernel (all temporal depths, ragged widths, multi-strip),
multi-domain decomposition emulated as one launch, exact early-exit semantics, file I/O."""
import os
import subprocess

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def make(w, h, **kw):
    import gol_b200
    return gol_b200.GameOfLife(w, h, device=0, **kw)


def check_steps(w, h, schedule, seed=1, **kw):
    from gol_b200.ops import cpu
    start = cpu.random_grid(w, h, seed)
    life = make(w, h, **kw)
    life.load_packed(start)
    want = start
    for n in schedule:
        life.step(n)
        want = cpu.step(want, w, n)
        got = life.packed_numpy()
        assert np.array_equal(got, want), f"{w}x{h} {kw} mismatch after +{n}"
    life.close()


@pytest.mark.parametrize("w,h", [(128, 16), (192, 40), (840, 840), (1000, 70), (129, 33), (191, 20), (200, 64),
                                 (2048, 64), (1984, 48), (3900, 50), (4096, 300), (7000, 37)])
def test_strip_kernel_matches_cpu(cuda_device, w, h):
    check_steps(w, h, [1, 2, 3, 5, 8, 16, 23])


@pytest.mark.parametrize("w,h,gens,grid", [(16384, 16384, 64, (1, 1)), (65536, 2048, 40, (1, 1)), (16384, 16384, 48, (2, 2))])
def test_benchmark_geometry_matches_oracle(cuda_device, w, h, gens, grid):
    # the shapes bench.py and BASELINE config 3 time: 512- / 2048-word rows (9 / 35 strips with the
    # right-aligned overlapping last strip), more chunks than resident warps, the AUTOTUNED chunk
    # height (only active from 256 rows up), programmatic dependent launch between the epochs
    from gol_b200.ops import cpu
    start = cpu.random_grid(w, h, 11)
    life = make(w, h, grid=grid, gen_limit=gens + 8)
    life.load_packed(start)
    assert life.uses_fast_path and life.chunk_rows >= 16
    life.step(gens)
    want = cpu.step(start, w, gens)
    assert np.array_equal(life.packed_numpy(), want)
    assert life.local_hash() == cpu.grid_hash(want, w)
    life.load_packed(start)                      # and the reference loop (flag epochs, gather, batches)
    res = life.run()
    wantg, wres = cpu.run(start, w, gen_limit=gens + 8)
    assert (res.generations, res.population) == (wres.generations, wres.population)
    assert life.local_hash() == cpu.grid_hash(wantg, w)
    life.close()


@pytest.mark.parametrize("k", [1, 2, 4, 8, 16])
def test_temporal_depths(cuda_device, k):
    check_steps(1100, 200, [k, 3 * k + 1, 37], k_max=k)


@pytest.mark.parametrize("grid", [(1, 2), (2, 1), (2, 2), (2, 4), (1, 8), (8, 1), (3, 3), (4, 2)])
def test_multi_domain_emulated(cuda_device, grid):
    # all R*C domains live in this process and are advanced by ONE launch per epoch; borders
    # travel through the same fused push that writes peer memory on a multi-GPU job
    check_steps(2240, 384, [1, 4, 7, 16, 30], grid=grid)
    check_steps(1500, 210, [5, 11], grid=grid, seed=3)


def test_glider_crosses_every_seam(cuda_device):
    from gol_b200.ops import cpu
    from gol_b200.utils import patterns
    w, h = 512, 128
    cells = patterns.grid_with(w, h, ("glider", w - 2, h - 2), ("glider", 254, 62), ("glider", 100, 10))
    start = cpu.pack(cells)
    for grid in [(1, 1), (2, 2), (2, 4)]:
        life = make(w, h, grid=grid)
        life.load_packed(start)
        life.step(4 * 70)
        assert np.array_equal(life.packed_numpy(), cpu.step(start, w, 4 * 70))
        assert life.population() == 15
        life.close()


def run_both(cells, **kw):
    from gol_b200.ops import cpu
    h, w = cells.shape
    start = cpu.pack(cells)
    cfg = {k: v for k, v in kw.items() if k in ("gen_limit", "similarity_frequency")}
    want, wres = cpu.run(start, w, **cfg)
    life = make(w, h, **kw)
    life.load_packed(start)
    res = life.run()
    got = life.packed_numpy()
    life.close()
    assert (res.generations, res.exit_reason) == (wres.generations, wres.exit_reason), (res, wres, kw)
    assert np.array_equal(got, want)
    assert res.population == wres.population
    return res


def test_run_semantics_fixtures(cuda_device):
    from gol_b200.utils import patterns
    w, h = 256, 64
    cases = {
        "empty": [], "cell": [("cell", 5, 5)], "block": [("block", 9, 9)], "tromino": [("tromino", 9, 9)],
        "blinker": [("blinker", 100, 30)], "beehive": [("beehive", 200, 40)], "diag3": [("diag3", 30, 30)],
        "pulsar": [("pulsar", 64, 20)], "rpent": [("rpentomino", 128, 32)],
        "seam": [("block", 255, 63), ("tromino", 126, 31)],
    }
    for name, stamps in cases.items():
        cells = patterns.grid_with(w, h, *stamps)
        for kw in ({}, {"grid": (2, 2)}, {"k_max": 8, "batch_gens": 24}, {"similarity_frequency": 0, "gen_limit": 50},
                   {"similarity_frequency": 5}, {"gen_limit": 7}):
            run_both(cells, **kw)


def test_late_exits_deep_inside_a_batch(cuda_device):
    # exits far from a batch boundary: the speculative driver overshoots, replays the batch one generation
    # at a time and must still report the serial loop's exact Generations value and grid
    from gol_b200.utils import patterns
    r = run_both(patterns.grid_with(1024, 1024, ("late590", 506, 506)))
    assert (r.generations, r.exit_reason) == (590, "similar")
    r = run_both(patterns.grid_with(1024, 512, ("diehard", 500, 250)), grid=(2, 1))
    assert (r.generations, r.exit_reason) == (130, "empty")


def test_run_stabilising_at_every_residue(cuda_device):
    # a tromino placed after d blinker-free generations of delay: block forms at generation 1;
    # shifting the similarity frequency walks the exit through every residue class
    from gol_b200.utils import patterns
    cells = patterns.grid_with(192, 32, ("tromino", 50, 10), ("diag3", 120, 12))
    for freq in (1, 2, 3, 4, 5, 7):
        for limit in (1, 2, 3, 4, 6, 1000):
            run_both(cells, similarity_frequency=freq, gen_limit=limit)


def test_run_random_sparse_grids(cuda_device):
    # sparse random grids die or stabilise at assorted generations
    rng = np.random.default_rng(5)
    for i in range(6):
        cells = (rng.random((48, 160)) < 0.04 + 0.01 * i).astype(np.uint8)
        run_both(cells, gen_limit=300)
        run_both(cells, gen_limit=300, grid=(1, 1), batch_gens=32, k_max=2)


def test_generic_path_small_grids(cuda_device):
    from gol_b200.utils import patterns
    rng = np.random.default_rng(11)
    for w, h in [(30, 30), (1, 1), (2, 2), (5, 3), (64, 8), (100, 10)]:
        cells = (rng.random((h, w)) < 0.35).astype(np.uint8)
        r = run_both(cells, gen_limit=200)
        assert r is not None
    life = make(30, 30)
    assert not life.uses_fast_path
    from gol_b200.ops import cpu
    g30 = cpu.random_grid(30, 30, 8)
    life.load_packed(g30)
    assert life.local_hash() == cpu.grid_hash(g30, 30) and life.population() == int(cpu.unpack(g30, 30).sum())
    blocks = [torch.empty(life.local_block_shape(0), dtype=torch.int64).pin_memory()]
    life.download_local(blocks)
    assert np.array_equal(blocks[0].numpy().view(np.uint64), g30)
    life.close()
    run_both(patterns.grid_with(30, 30, ("block", 3, 3)))
    run_both(patterns.grid_with(512, 64, ("block", 3, 3)), force_generic=True)


def test_tensor_views(cuda_device):
    from gol_b200.ops import cpu
    w, h = 840, 100
    start = cpu.random_grid(w, h, 4)
    life = make(w, h)
    life.load_packed(start)
    t = life.packed()
    assert t.dtype == torch.int64 and tuple(t.shape) == (h, 14) and t.is_cuda
    assert np.array_equal(t.cpu().numpy().view(np.uint64), start)
    assert np.array_equal(life.cells().cpu().numpy(), cpu.unpack(start, w))
    life.step(3)
    assert np.array_equal(life.packed().cpu().numpy().view(np.uint64), cpu.step(start, w, 3))
    life.close()


@pytest.mark.parametrize("grid", [(1, 1), (2, 2), (1, 3)])
def test_text_file_io(cuda_device, tmp_path, grid):
    from gol_b200.ops import cpu
    w, h = 840, 77
    inp, out = tmp_path / "in.txt", tmp_path / "out.txt"
    cpu.random_text_file(str(inp), w, h, 21)
    life = make(w, h, grid=grid)
    life.load_file(str(inp))
    assert np.array_equal(life.packed_numpy(), cpu.random_grid(w, h, 21))
    life.step(5)
    out.write_bytes(b"stale content that is longer than nothing" * 4000)
    life.save(str(out))
    assert out.read_bytes() == cpu.format_text(cpu.step(cpu.random_grid(w, h, 21), w, 5), w)
    short = tmp_path / "short.txt"
    short.write_bytes(inp.read_bytes()[: (w + 1) * 10])
    with pytest.raises(Exception, match="too short"):
        life.load_file(str(short))
    garbled = bytearray(inp.read_bytes())
    garbled[(w + 1) * 3 + 17] = ord("7")
    (tmp_path / "bad.txt").write_bytes(bytes(garbled))
    with pytest.raises(Exception, match="invalid byte"):
        life.load_file(str(tmp_path / "bad.txt"))
    life.close()


def test_cli_matches_unmodified_reference(cuda_device, ref_serial, tmp_path):
    from gol_b200.ops import cpu
    from gol_b200.utils import reference
    n = 256
    inp = tmp_path / "in.txt"
    cpu.random_text_file(str(inp), n, n, 33)
    gens, _, want, _ = reference.run_program(ref_serial, n, n, str(inp), str(tmp_path), "game_output.out")
    os.rename(tmp_path / "game_output.out", tmp_path / "ref.out")
    exe = os.path.join(ROOT, "bin", "gol")
    for variant, extra in [("game", []), ("cuda", []), ("collective", ["--ranks", "4"])]:
        p = subprocess.run([exe, str(n), str(n), str(inp), "--variant", variant] + extra, cwd=tmp_path,
                           capture_output=True, text=True, check=True)
        assert f"Generations:\t{gens}\n" in p.stdout, p.stdout
        assert "Execution time:\t" in p.stdout and p.stdout.rstrip().endswith("Finished")
        assert (tmp_path / f"{variant}_output.out").read_bytes() == want
        if variant == "collective":
            assert p.stdout.startswith("Reading file:\t") and "Writing file:\t" in p.stdout
    # fewer than three arguments: the C programs do nothing and print Finished (S:238-241)
    p = subprocess.run([exe, "10"], cwd=tmp_path, capture_output=True, text=True, check=True)
    assert p.stdout == "Finished\n"


def test_end_to_end_case(cuda_device):
    # scripts/sanitize_case.py walks every kernel of the library (deep-halo epoch pairs included) and
    # checks each against the CPU oracle; it is also the case meant for compute-sanitizer (closed on
    # this pool, so it is run plain here)
    import sys
    p = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "sanitize_case.py")], capture_output=True,
                       text=True, timeout=300)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    assert "sanitize case ok" in p.stdout


def test_resume_from_checkpoint_file(cuda_device, tmp_path):
    # output format == input format (README.md:61-63 of the reference), so a run can be resumed from
    # its output; start_generation restores the generation count and the similarity phase
    from gol_b200.ops import cpu
    from gol_b200.utils import patterns
    w, h = 192, 48
    cells = patterns.grid_with(w, h, ("rpentomino", 90, 20), ("glider", 10, 10))
    start = cpu.pack(cells)
    want, wres = cpu.run(start, w, gen_limit=400)
    for stop in (1, 4, 5, 97):
        a = make(w, h, gen_limit=stop, similarity_frequency=0)
        a.load_packed(start)
        assert a.run().generations == stop
        ck = tmp_path / f"ck{stop}.txt"
        a.save(str(ck))
        a.close()
        b = make(w, h, gen_limit=400, start_generation=stop)
        b.load_file(str(ck))
        res = b.run()
        assert (res.generations, res.exit_reason) == (wres.generations, wres.exit_reason), (stop, res, wres)
        assert np.array_equal(b.packed_numpy(), want)
        b.close()


def test_property_random_shapes_and_decompositions(cuda_device):
    # k temporally blocked generations == k single generations for random sizes / depths / grids
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=25, deadline=None)
    @given(st.integers(128, 2600), st.integers(16, 260), st.sampled_from([1, 2, 4, 8, 16]),
           st.sampled_from([(1, 1), (1, 2), (2, 1), (2, 2), (1, 3), (3, 1)]), st.integers(1, 40), st.integers(0, 2 ** 31))
    def prop(w, h, k, grid, n, seed):
        r, c = grid
        if w // c < 192 or h // r < 16:
            grid = (1, 1)
        check_steps(w, h, [n], seed=seed, k_max=k, grid=grid)

    prop()


def test_cli_one_process_per_gpu(cuda_device, tmp_path):
    # forked ranks + shared-memory rendezvous + CUDA IPC (needs two GPUs; the driver's box has one or eight)
    if torch.cuda.device_count() < 2:
        pytest.skip("needs at least two GPUs")
    from gol_b200.ops import cpu
    n = 1024
    inp = tmp_path / "in.txt"
    cpu.random_text_file(str(inp), n, n, 44)
    exe = os.path.join(ROOT, "bin", "gol")
    subprocess.run([exe, str(n), str(n), str(inp), "--variant", "cuda", "--gen-limit", "100"], cwd=tmp_path, check=True,
                   capture_output=True)
    p = subprocess.run([exe, str(n), str(n), str(inp), "--variant", "collective", "--gpus", "2", "--gen-limit", "100"],
                       cwd=tmp_path, check=True, capture_output=True, text=True)
    assert "Generations:\t100\n" in p.stdout and p.stdout.count("Finished") == 2
    assert (tmp_path / "collective_output.out").read_bytes() == (tmp_path / "cuda_output.out").read_bytes()
    want = cpu.format_text(cpu.step(cpu.random_grid(n, n, 44), n, 100), n)
    assert (tmp_path / "cuda_output.out").read_bytes() == want


@pytest.mark.parametrize("grid", [(1, 1), (1, 2), (2, 1), (2, 2), (2, 4), (8, 1), (3, 3)])
def test_deep_halo_pairs_emulated(cuda_device, monkeypatch, grid):
    # GOL_DEEP_HALO=1 switches the pair schedule on inside one process: the first epoch of a pair also
    # computes the ghost rows / ghost words next to every domain, the second one needs no exchange
    monkeypatch.setenv("GOL_DEEP_HALO", "1")
    life = make(2240, 384, grid=grid)
    assert life.uses_deep_halo
    life.close()
    check_steps(2240, 384, [16, 1, 32, 7, 48, 100], grid=grid)
    check_steps(1536, 210, [16, 33], grid=grid, seed=3, k_max=4)
    check_steps(1536, 100, [64], grid=(1, 1), seed=4)       # single-chunk domains (shorter than three edge chunks)
    from gol_b200.utils import patterns
    cells = patterns.grid_with(1024, 128, ("tromino", 510, 62), ("diag3", 700, 100))
    run_both(cells, gen_limit=200, grid=grid if 128 // grid[0] >= 16 else (1, 1))


def test_per_domain_stage_download_and_checksum(cuda_device):
    # a rank that keeps only its own blocks on the host: stage_local / download_local per local domain,
    # and the decomposition-independent checksum (sum over domains == host checksum of the global grid)
    from gol_b200.ops import cpu
    w, h = 2240, 384
    start = cpu.random_grid(w, h, 6)
    for grid in [(1, 1), (2, 2), (4, 1), (1, 3)]:
        life = make(w, h, grid=grid, gen_limit=24)
        ins, outs = [], []
        for i, g in enumerate(life.local_domains):
            x0, y0, dw, dh = life.domain_extent(g)
            blk = np.ascontiguousarray(start[y0:y0 + dh, x0 // 64:(x0 + dw + 63) // 64]).view(np.int64)
            ins.append(torch.from_numpy(blk).pin_memory())
            outs.append(torch.empty(life.local_block_shape(i), dtype=torch.int64).pin_memory())
        life.stage_local(ins).commit_staged()
        assert life.local_hash() == cpu.grid_hash(start, w)
        res = life.run()
        want, wres = cpu.run(start, w, gen_limit=24)
        assert res.generations == wres.generations and life.local_hash() == cpu.grid_hash(want, w)
        life.download_local(outs)
        for i, g in enumerate(life.local_domains):
            x0, y0, dw, dh = life.domain_extent(g)
            assert np.array_equal(outs[i].numpy().view(np.uint64), want[y0:y0 + dh, x0 // 64:(x0 + dw + 63) // 64]), (grid, g)
        life.close()


def test_staged_upload(cuda_device):
    from gol_b200.ops import cpu
    w, h = 1100, 200
    a, b = cpu.random_grid(w, h, 1), cpu.random_grid(w, h, 2)
    for grid in [(1, 1), (2, 2)]:
        life = make(w, h, grid=grid, gen_limit=20)
        ta = torch.from_numpy(a.view(np.int64)).pin_memory()
        tb = torch.from_numpy(b.view(np.int64)).pin_memory()
        life.stage_packed(ta).commit_staged()
        life.stage_packed(tb)                      # uploads while the first grid runs
        r1 = life.run()
        assert np.array_equal(life.packed_numpy(), cpu.run(a, w, gen_limit=20)[0]) and r1.generations == 20
        life.commit_staged(stage_next=ta)          # commit b and start re-uploading a in one call
        life.run()
        assert np.array_equal(life.packed_numpy(), cpu.run(b, w, gen_limit=20)[0])
        life.commit_staged()
        life.run()
        assert np.array_equal(life.packed_numpy(), cpu.run(a, w, gen_limit=20)[0])
        life.close()
