"""The model: Conway's Game of Life (B3/S23) on a periodic grid, driven by the native engine.

``GameOfLife`` is the user-facing object (the counterpart of the reference's six programs):
it owns a native engine (bit-packed padded device buffers, strip kernel with fused halo push
and flag reduction, speculative epoch driver) and exposes the grid as PyTorch tensors.

    life = GameOfLife(16384, 16384)
    life.randomize(seed=1)                  # or life.load(cells) / life.load_file(path)
    result = life.run()                     # reference loop: Generations, exit reason, time
    cells = life.cells()                    # uint8 [H, W] CUDA tensor
    life.save("game_output.out")            # reference text format

Under ``torch.distributed`` (one process per GPU, NCCL) pass ``grid=(R, C)``: every rank owns
``R*C / world_size`` domains, the ranks exchange one CUDA IPC handle each through the process
group, and from then on halos and flags move by direct peer stores inside the generation
kernel - NCCL is bootstrap only.

Semantics are those of the serial reference (src/game.c:169-196); see SURVEY.md section 2.8.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from .. import _native as N
from ..ops.cpu import RunResult, words_per_row

# reference program -> output file written into the CWD (src/game.c:27, game_mpi.c:29, ...)
VARIANT_OUTPUT = {
    "game": "game_output.out", "mpi": "mpi_output.out", "async": "async_output.out",
    "collective": "collective_output.out", "openmp": "openmp_output.out", "cuda": "cuda_output.out",
}


@dataclass
class LifeConfig:
    """Run-time form of the reference's compile-time configuration (src/game.c:6-9)."""
    width: int = 30                    # the reference's default for non-positive sizes
    height: int = 30
    gen_limit: int = 1000              # GEN_LIMIT
    similarity_frequency: int = 3      # SIMILARITY_FREQUENCY; 0 = CHECK_SIMILARITY undefined
    grid: tuple[int, int] = (1, 1)     # R x C domain decomposition
    k_max: int = 0                     # temporal blocking depth (0 = auto)
    batch_gens: int = 0                # generations between flag checks (0 = auto)
    force_generic: bool = False
    start_generation: int = 0          # resume: the loaded grid is this generation (checkpoint/restart)


class _DeviceBuffer:
    """Minimal __cuda_array_interface__ carrier for a zero-copy tensor view."""

    def __init__(self, ptr: int, words32: int):
        self.__cuda_array_interface__ = {"shape": (words32,), "typestr": "<i4", "data": (ptr, False),
                                         "version": 2, "strides": None}


class GameOfLife:
    def __init__(self, width: int | None = None, height: int | None = None, *, config: LifeConfig | None = None,
                 device: int | None = None, process_group=None, distributed: bool | None = None, **overrides):
        """`distributed`: None = one rank of the default process group when torch.distributed is
        initialised and a decomposition is given (or `process_group` is passed); False = every
        domain lives in this process even inside a distributed job (no collective is issued)."""
        import torch
        N.require_cuda()
        cfg = config or LifeConfig()
        if width is not None:
            cfg.width = int(width)
        if height is not None:
            cfg.height = int(height)
        for k, v in overrides.items():
            if not hasattr(cfg, k):
                raise TypeError(f"unknown option {k}")
            setattr(cfg, k, v)
        self.config = cfg
        self._torch = torch
        self._group = process_group
        rank, world = 0, 1
        import torch.distributed as dist
        if distributed is None:
            distributed = process_group is not None or (dist.is_available() and dist.is_initialized() and cfg.grid != (1, 1))
        if distributed:
            rank, world = dist.get_rank(process_group), dist.get_world_size(process_group)
        self.rank, self.world = rank, world
        if device is None:
            device = torch.cuda.current_device()
        self.device = torch.device("cuda", int(device))
        nc = N.Config()
        lib = N.lib()
        lib.gol_config_default(C.byref(nc), cfg.width, cfg.height)
        nc.grid_rows, nc.grid_cols = cfg.grid
        nc.proc_rank, nc.proc_count = rank, world
        nc.device = self.device.index
        nc.gen_limit = cfg.gen_limit
        nc.similarity_frequency = cfg.similarity_frequency
        nc.k_max = cfg.k_max
        nc.batch_gens = cfg.batch_gens
        nc.flags = N.FLAG_FORCE_GENERIC if cfg.force_generic else 0
        nc.start_generation = cfg.start_generation
        handle = C.c_void_p()
        N.check(lib.gol_engine_create(C.byref(nc), C.byref(handle)), "gol_engine_create")
        self._h = handle
        self._lib = lib
        if world > 1:
            self._bootstrap()

    # ------------------------------------------------------------------ plumbing
    def _bootstrap(self):
        """All-gather one IPC blob per rank through the process group, then map the peers."""
        import torch.distributed as dist
        torch = self._torch
        size = int(self._lib.gol_engine_ipc_blob_size(self._h))
        blob = (C.c_char * size)()
        N.check(self._lib.gol_engine_ipc_export(self._h, blob), "ipc_export")
        backend = dist.get_backend(self._group)
        dev = self.device if backend == "nccl" else torch.device("cpu")
        mine = torch.frombuffer(bytearray(blob.raw), dtype=torch.uint8).to(dev)
        parts = [torch.empty_like(mine) for _ in range(self.world)]
        dist.all_gather(parts, mine, group=self._group)
        allb = b"".join(bytes(p.cpu().numpy().tobytes()) for p in parts)
        N.check(self._lib.gol_engine_ipc_import(self._h, allb), "ipc_import")
        dist.barrier(group=self._group)

    def _barrier(self):
        if self.world > 1:
            import torch.distributed as dist
            self._sync()
            dist.barrier(group=self._group)

    def _sync(self):
        N.check(self._lib.gol_engine_sync(self._h), "sync")

    def close(self):
        if getattr(self, "_h", None):
            self._lib.gol_engine_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ topology
    @property
    def uses_fast_path(self) -> bool:
        return bool(self._lib.gol_engine_uses_fast_path(self._h))

    @property
    def uses_deep_halo(self) -> bool:
        """Epochs run as pairs with one 16-deep halo exchange per pair (default between processes)."""
        return bool(self._lib.gol_engine_uses_deep_halo(self._h))

    @property
    def chunk_rows(self) -> int:
        """Height of the interior chunks at the engine's temporal depth (autotuned in prepare())."""
        return int(self._lib.gol_engine_chunk_rows(self._h))

    def trace_dump(self, path: str) -> None:
        """Per-launch device timeline of an engine created with GOL_TRACE=<n> in the environment."""
        N.check(self._lib.gol_engine_trace_dump(self._h, str(path).encode()), "trace_dump")

    @property
    def num_domains(self) -> int:
        return int(self._lib.gol_engine_num_domains(self._h))

    @property
    def local_domains(self) -> list[int]:
        n = self._lib.gol_engine_num_local_domains(self._h)
        return [int(self._lib.gol_engine_local_domain_id(self._h, i)) for i in range(n)]

    def domain_extent(self, gid: int) -> tuple[int, int, int, int]:
        x0, y0, w, h = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int64()
        N.check(self._lib.gol_engine_domain_extent(self._h, gid, C.byref(x0), C.byref(y0), C.byref(w), C.byref(h)))
        return x0.value, y0.value, w.value, h.value

    def domain_neighbors(self, gid: int) -> list[int]:
        nb = (C.c_int32 * 8)()
        N.check(self._lib.gol_engine_domain_neighbors(self._h, gid, C.byref(nb)))
        return list(nb)

    # ------------------------------------------------------------------ load / store
    def load_packed(self, packed) -> "GameOfLife":
        """Global packed grid: numpy uint64 / torch int64 CPU tensor [H, ceil(W/64)] (pinned memory
        makes the copy asynchronous-capable).  Every rank passes the same array."""
        torch = self._torch
        if isinstance(packed, torch.Tensor):
            if packed.is_cuda:
                if self.world > 1 or self.config.grid != (1, 1):
                    raise ValueError("device-resident load needs a single-domain engine")
                t = packed.contiguous()
                N.check(self._lib.gol_engine_load_packed_device(self._h, t.data_ptr(), 0, 0, self.config.width,
                                                                self.config.height), "load_packed_device")
                self._after_load()
                return self
            ptr = packed.contiguous().data_ptr()
            keep = packed
        else:
            keep = np.ascontiguousarray(packed, dtype=np.uint64)
            ptr = keep.ctypes.data
        if tuple(keep.shape) != (self.config.height, words_per_row(self.config.width)):
            raise ValueError(f"packed grid must have shape {(self.config.height, words_per_row(self.config.width))}")
        N.check(self._lib.gol_engine_load_packed(self._h, C.c_void_p(ptr)), "load_packed")
        self._after_load()
        return self

    def stage_packed(self, packed) -> "GameOfLife":
        """Start uploading the NEXT grid (pinned torch int64 CPU tensor or numpy uint64, global packed
        shape) on a second stream and return immediately; the engine may still be running the current
        grid.  Keep `packed` alive and unchanged until commit_staged()."""
        torch = self._torch
        if isinstance(packed, torch.Tensor):
            keep, ptr = packed, packed.contiguous().data_ptr()
        else:
            keep = np.ascontiguousarray(packed, dtype=np.uint64)
            ptr = keep.ctypes.data
        if tuple(keep.shape) != (self.config.height, words_per_row(self.config.width)):
            raise ValueError("packed grid has the wrong shape")
        self._staged_keepalive = keep
        N.check(self._lib.gol_engine_stage_packed(self._h, C.c_void_p(ptr)), "stage_packed")
        return self

    def local_block_shape(self, local_index: int = 0) -> tuple[int, int]:
        """(rows, 64-bit words per row) of a local domain's packed dense block."""
        _, _, w, h = self.domain_extent(self.local_domains[local_index])
        return h, words_per_row(w)

    def stage_local(self, blocks) -> "GameOfLife":
        """Like stage_packed, but from the rank's OWN blocks: one pinned int64 CPU tensor
        [h, ceil(w/64)] per local domain (local_block_shape), so a rank never holds the global grid."""
        if len(blocks) != len(self.local_domains):
            raise ValueError("one block per local domain")
        for i, b in enumerate(blocks):
            if tuple(b.shape) != self.local_block_shape(i) or not b.is_contiguous():
                raise ValueError(f"block {i} must be contiguous with shape {self.local_block_shape(i)}")
            N.check(self._lib.gol_engine_stage_domain(self._h, i, C.c_void_p(b.data_ptr())), "stage_domain")
        self._staged_keepalive = blocks
        return self

    def download_local(self, blocks) -> "GameOfLife":
        """Current grid of every local domain into the caller's blocks (pinned int64 CPU tensors,
        local_block_shape); returns when the data is in host memory."""
        if len(blocks) != len(self.local_domains):
            raise ValueError("one block per local domain")
        for i, b in enumerate(blocks):
            if tuple(b.shape) != self.local_block_shape(i) or not b.is_contiguous():
                raise ValueError(f"block {i} must be contiguous with shape {self.local_block_shape(i)}")
            N.check(self._lib.gol_engine_download_domain(self._h, i, C.c_void_p(b.data_ptr())), "download_domain")
        return self

    def local_hash(self) -> int:
        """Checksum of the local domains; the sum over all ranks (mod 2**64) identifies the global
        grid independently of the decomposition (gol_engine_hash)."""
        c = C.c_uint64()
        N.check(self._lib.gol_engine_hash(self._h, C.byref(c)), "hash")
        return int(c.value)

    def commit_staged(self, stage_next=None, stage_next_local=None) -> "GameOfLife":
        """Make the staged grid the current one (then run()/step() as after load_packed).
        `stage_next`: start uploading the grid after this one right away (same as calling
        stage_packed() next, but issued before the ranks synchronise, so the host time of the
        enqueue does not skew the ranks against each other at the start of the generation loop)."""
        N.check(self._lib.gol_engine_commit_staged(self._h), "commit_staged")
        if stage_next is not None:
            self.stage_packed(stage_next)
        if stage_next_local is not None:
            self.stage_local(stage_next_local)
        self._after_load()
        return self

    def load(self, cells) -> "GameOfLife":
        """uint8/bool [H, W] cells (numpy, or a torch tensor on CPU or GPU)."""
        torch = self._torch
        if isinstance(cells, torch.Tensor) and cells.is_cuda and self.world == 1 and self.config.grid == (1, 1):
            from ..ops import dense
            return self.load_packed(dense.pack_cells(cells.to(torch.uint8)))
        if isinstance(cells, torch.Tensor):
            cells = cells.cpu().numpy()
        from ..ops import cpu
        return self.load_packed(cpu.pack(cells))

    def load_file(self, path: str, strict: bool = True) -> "GameOfLife":
        """Parallel text read: each rank preads and GPU-parses only its own sub-blocks."""
        N.check(self._lib.gol_engine_load_text_file(self._h, str(path).encode(), int(strict)), "load_text_file")
        self._after_load()
        return self

    def randomize(self, seed: int = 1) -> "GameOfLife":
        N.check(self._lib.gol_engine_randomize(self._h, seed), "randomize")
        self._after_load()
        return self

    def _after_load(self):
        self._barrier()          # nobody pushes ghosts into a buffer that is still being loaded
        N.check(self._lib.gol_engine_prepare(self._h), "prepare")

    def save(self, path: str) -> None:
        """Parallel text write in the reference format (H lines of W chars + newline)."""
        import os
        if self.rank == 0 and os.path.exists(path):
            os.unlink(path)
        self._barrier()
        N.check(self._lib.gol_engine_store_text_file(self._h, str(path).encode()), "store_text_file")
        self._barrier()

    def packed_numpy(self) -> np.ndarray:
        """Global packed grid assembled on the host (local domains only under torch.distributed)."""
        out = np.zeros((self.config.height, words_per_row(self.config.width)), dtype=np.uint64)
        N.check(self._lib.gol_engine_download_packed(self._h, out.ctypes.data_as(C.c_void_p)), "download_packed")
        return out

    def cells_numpy(self) -> np.ndarray:
        from ..ops import cpu
        return cpu.unpack(self.packed_numpy(), self.config.width)

    # ------------------------------------------------------------------ tensors
    def padded_tensor(self, local_index: int = 0):
        """Zero-copy int32 view [rows_padded, pitch32] of a local domain's current device buffer
        (ghost rows/columns included; see gol_layout in include/gol_b200.h)."""
        torch = self._torch
        ptr, lay = C.c_void_p(), N.Layout()
        N.check(self._lib.gol_engine_domain_buffer(self._h, local_index, C.byref(ptr), C.byref(lay)), "domain_buffer")
        flat = torch.as_tensor(_DeviceBuffer(ptr.value, int(lay.buffer_words32)), device=self.device)
        return flat.view(int(lay.rows_padded), int(lay.pitch32)), lay

    def packed(self, local_index: int = 0):
        """Zero-copy int64 view [h, ceil(w/64)] of a local domain's interior (the grid as a tensor)."""
        t, lay = self.padded_tensor(local_index)
        g = int(lay.ghost_rows)
        off = N.GHOST_WORDS32 if g else 0
        inner = t[g:g + int(lay.height), off:off + int(lay.nw32)]
        return inner.view(self._torch.int64)

    def cells(self, local_index: int = 0):
        """uint8 [h, w] CUDA tensor of a local domain's cells."""
        from ..ops import dense
        _, lay = self.padded_tensor(local_index)
        return dense.unpack_cells(self.packed(local_index).contiguous(), int(lay.width))

    # ------------------------------------------------------------------ compute
    def step(self, generations: int = 1) -> float:
        """Advance exactly `generations` generations without exit checks; returns device ms."""
        N.check(self._lib.gol_engine_step(self._h, generations), "step")
        return float(self._lib.gol_engine_last_ms(self._h))

    def run(self) -> RunResult:
        """The reference generation loop with both early exits (exact Generations value)."""
        res = N.Result()
        N.check(self._lib.gol_engine_run(self._h, C.byref(res)), "run")
        return RunResult.from_native(res)

    def population(self) -> int:
        c = C.c_uint64()
        N.check(self._lib.gol_engine_population(self._h, C.byref(c)), "population")
        return int(c.value)

    @property
    def launch_count(self) -> int:
        return int(self._lib.gol_engine_launch_count(self._h))

    @property
    def last_ms(self) -> float:
        return float(self._lib.gol_engine_last_ms(self._h))
