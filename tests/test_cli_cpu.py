"""The a.out-compatible CLI on a box without a GPU: `--cpu` (chosen automatically when no CUDA device
is usable) runs the whole program on the host, like the reference's serial program (src/game.c:224-245),
and must reproduce its stdout lines and output file byte for byte."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "bin", "gol")


def test_cli_cpu_matches_unmodified_reference(ref_serial, tmp_path, lib):
    from gol_b200.ops import cpu
    from gol_b200.utils import reference
    n = 96
    inp = tmp_path / "in.txt"
    cpu.random_text_file(str(inp), n, n, 12)
    gens, _, want, ref_stdout = reference.run_program(ref_serial, n, n, str(inp), str(tmp_path), "game_output.out")
    os.rename(tmp_path / "game_output.out", tmp_path / "ref.out")
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")          # no device even on a GPU box
    for extra in (["--cpu"], []):                            # explicit, and the automatic fallback
        p = subprocess.run([EXE, str(n), str(n), str(inp)] + extra, cwd=tmp_path, capture_output=True, text=True,
                           check=True, env=env)
        assert f"Generations:\t{gens}\n" in p.stdout and p.stdout.startswith("Finished.\n\n")
        assert p.stdout.rstrip().endswith("Finished")
        assert (tmp_path / "game_output.out").read_bytes() == want
        os.unlink(tmp_path / "game_output.out")
    # MPI-style report lines and output name; early exit on a still life
    from gol_b200.utils import patterns
    cells = patterns.grid_with(40, 40, ("block", 5, 5))
    (tmp_path / "block.txt").write_bytes(cpu.format_text(cpu.pack(cells), 40))
    p = subprocess.run([EXE, "40", "40", "block.txt", "--variant", "collective", "--cpu"], cwd=tmp_path,
                       capture_output=True, text=True, check=True, env=env)
    assert p.stdout.startswith("Reading file:\t") and "Generations:\t2\n" in p.stdout and "Writing file:\t" in p.stdout
    assert (tmp_path / "collective_output.out").read_bytes() == (tmp_path / "block.txt").read_bytes()


def test_cli_without_arguments_prints_finished(tmp_path):
    p = subprocess.run([EXE, "10"], cwd=tmp_path, capture_output=True, text=True, check=True)
    assert p.stdout == "Finished\n"
