"""GPU tests of the stand-alone device ops against plain PyTorch fp32 references."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def conv_oracle(cells: torch.Tensor) -> torch.Tensor:
    """Plain PyTorch fp32 reference: circular-padded 3x3 conv + B3/S23 rule."""
    x = cells.to(torch.float32)[None, None]
    h, w = cells.shape
    # explicit torus gather (works for every size >= 1, like the serial reference's wrap)
    ys = torch.arange(h, device=cells.device)
    xs = torch.arange(w, device=cells.device)
    n = torch.zeros_like(x)
    for dy in (-1, 0, 1):
        for dx in (-1, 0, 1):
            if dy == 0 and dx == 0:
                continue
            n += x[:, :, (ys + dy) % h][:, :, :, (xs + dx) % w]
    alive = (n == 3) | ((n == 2) & (x == 1))
    return alive[0, 0].to(torch.uint8)


SIZES = [(1, 1), (1, 4), (4, 1), (2, 2), (3, 5), (30, 30), (63, 7), (64, 64), (65, 9), (129, 33), (1000, 17)]


@pytest.mark.parametrize("w,h", SIZES)
def test_step_dense_matches_conv2d(cuda_device, w, h):
    from gol_b200.ops import dense
    g = torch.Generator(device="cpu").manual_seed(w * 131 + h)
    cells = (torch.rand((h, w), generator=g) < 0.4).to(torch.uint8).to(cuda_device)
    packed = dense.pack_cells(cells)
    assert torch.equal(dense.unpack_cells(packed, w), cells)
    cur = cells
    for _ in range(3):
        flags = torch.zeros(2, dtype=torch.int32, device=cuda_device)
        nxt = dense.step(packed, w, flags=flags)
        want = conv_oracle(cur)
        assert torch.equal(dense.unpack_cells(nxt, w), want)
        assert bool(flags[0].item()) == bool(want.any().item())
        assert bool(flags[1].item()) == (not torch.equal(want, cur))
        packed, cur = nxt, want


def test_strip_kernel_matches_conv2d(cuda_device):
    # the PRODUCT kernel (register-pipelined strips, temporal blocking, fused torus wrap) against the
    # plain PyTorch fp32 oracle, ragged width and several temporal depths
    import gol_b200
    g = torch.Generator(device="cpu").manual_seed(99)
    for w, h, k in [(840, 200, 8), (1000, 70, 4), (2112, 96, 8)]:
        cells = (torch.rand((h, w), generator=g) < 0.37).to(torch.uint8).to(cuda_device)
        life = gol_b200.GameOfLife(w, h, device=0, k_max=k)
        assert life.uses_fast_path
        life.load(cells)
        want = cells
        for n in (1, k, k + 3):
            life.step(n)
            for _ in range(n):
                want = conv_oracle(want)
            assert torch.equal(life.cells(), want), (w, h, k, n)
        life.close()


def test_flags_empty_similar(cuda_device):
    from gol_b200.ops import dense
    z = torch.zeros((5, 2), dtype=torch.int64, device=cuda_device)
    assert dense.empty(z, 100) and dense.similar(z, z.clone(), 100)
    a = z.clone()
    a[3, 1] = 1 << 20
    assert not dense.empty(a, 100) and not dense.similar(a, z, 100)
    assert dense.population(a, 100) == 1


def test_parse_format_random_match_cpu(cuda_device):
    from gol_b200.ops import cpu, dense
    for w, h in [(70, 13), (64, 3), (1, 1), (1025, 5)]:
        want = cpu.random_grid(w, h, 9)
        got = dense.random_grid(w, h, 9, device=cuda_device)
        assert np.array_equal(got.cpu().numpy().view(np.uint64), want)
        text = dense.format_text(got, w)
        assert bytes(text.cpu().numpy().tobytes()) == cpu.format_text(want, w)
        back = dense.parse_text(text, w, h)
        assert torch.equal(back, got)
        assert dense.population(got, w) == int(cpu.unpack(want, w).sum())
    bad = text.clone()
    bad[0, 3] = ord("x")
    with pytest.raises(Exception, match="invalid byte"):
        dense.parse_text(bad, 1025, 5)
    bad = text.clone()
    bad[2, 1025] = ord("0")
    with pytest.raises(Exception, match="newline"):
        dense.parse_text(bad, 1025, 5)


def test_text_kernels_every_alignment_and_ragged_tail(cuda_device):
    """The vectorised parse / format kernels against plain PyTorch: rows of W+1 bytes start at every
    byte alignment, widths that end inside a word / a 16-byte vector / a 8192-cell segment, buffers
    that are slices at odd offsets, and a last line without a newline (nothing past it is read)."""
    from gol_b200.ops import dense
    g = torch.Generator(device="cpu").manual_seed(77)
    for w, h in [(1, 3), (15, 4), (16, 5), (17, 33), (31, 2), (33, 17), (100, 9), (8191, 3), (8192, 2), (8193, 5), (20000, 3)]:
        cells = (torch.rand((h, w), generator=g) < 0.5).to(torch.uint8).to(cuda_device)
        packed = dense.pack_cells(cells)
        # plain PyTorch reference of the packing: bit i of word q = cell 64 q + i
        nq = (w + 63) // 64
        padded = torch.zeros((h, nq * 64), dtype=torch.int64, device=cuda_device)
        padded[:, :w] = cells
        want = (padded.view(h, nq, 64) << torch.arange(64, device=cuda_device)).sum(dim=2)   # wraps mod 2^64
        assert torch.equal(packed, want), (w, h)
        assert torch.equal(dense.unpack_cells(packed, w), cells)
        text = dense.format_text(packed, w)
        want_text = torch.full((h, w + 1), ord("\n"), dtype=torch.uint8, device=cuda_device)
        want_text[:, :w] = cells + ord("0")
        assert torch.equal(text, want_text), (w, h)
        for off in (0, 1, 3, 7, 13):                       # the buffer itself starts at an odd address
            big = torch.full((off + text.numel() + 64,), ord("x"), dtype=torch.uint8, device=cuda_device)
            big[off:off + text.numel()] = text.flatten()
            assert torch.equal(dense.parse_text(big[off:off + text.numel()], w, h), packed), (w, h, off)
            # no newline after the last line: the byte after the buffer is an 'x' that must not be looked at
            assert torch.equal(dense.parse_text(big[off:off + text.numel() - 1], w, h), packed), (w, h, off)
        assert torch.equal(dense.format_text(packed, w, newline=False), want_text[:, :w].contiguous())


def test_custom_op_and_module(cuda_device):
    from gol_b200.models import LifeStep
    from gol_b200.ops import dense
    g = torch.Generator(device="cpu").manual_seed(5)
    cells = (torch.rand((3, 40, 70), generator=g) < 0.4).to(torch.uint8).to(cuda_device)
    out = LifeStep(generations=2)(cells)
    for i in range(3):
        assert torch.equal(out[i], conv_oracle(conv_oracle(cells[i])))
    packed = dense.pack_cells(cells[0].contiguous())
    assert torch.equal(torch.ops.gol_b200.step_dense(packed, 70), dense.step(packed, 70))


def test_bench_contract(cuda_device):
    # the JSON line the driver parses (small grid so that the test stays short)
    import json
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    p = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--steps", "2", "--warmup", "3", "--size", "2048",
                        "--gens", "200"], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
                "vs_baseline", "dtype", "data", "config", "e2e", "gpu_launches", "clocks"):
        assert key in d, key
    assert d["n_gpus"] == 1 and d["steps"] == 2 and d["warmup"] == 3 and d["higher_is_better"] is True
    assert d["value"] > 0 and d["e2e"]["value"] > 0 and d["e2e"]["value"] < d["value"] * 1.001
    assert d["e2e"]["h2d_bytes_per_step"] == 2048 * 2048 // 8 and d["gpu_launches"] > 0
    assert d["details"]["generations_reported"] == 200 and d["details"]["fast_path"] is True
    assert d["e2e"]["d2h_bytes_per_step"] >= 2048 * 2048 // 8          # the final grid comes back every step
    assert d["verify"]["hash_equal"] and d["verify"]["population_equal"] and d["verify"]["downloaded_grid_population_equal"]
    assert d["cli_wall"]["generations"] == 200 and d["cli_wall"]["ms"] > 0
    assert d["details"]["chunk_rows"] > 0 and len(d["details"]["grid_hash"]) == 16
