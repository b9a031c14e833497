/*
 * gol_b200 — Blackwell-native Conway's Game of Life engine: public C ABI.
 *
 * One shared library (libgolb200.so, built in-tree for sm_100a) carries every
 * native component: bit-packed grid store, generation kernels, fused halo push
 * and flag reduction, text parse/format kernels, stripe file I/O, the epoch
 * driver with the reference's early-exit semantics, and a threaded CPU oracle.
 * The CLI (a.out), the Python/PyTorch surface and the tests all go through the
 * functions declared here.
 *
 * Conventions
 *  - every function returns 0 on success, non-zero on failure; the message for
 *    the last failure on the calling thread is gol_last_error().
 *  - cells are bits; bit i of 64-bit word q of a row is cell x = 64*q + i
 *    (LSB = lowest x).  Bits at x >= width are always zero ("canonical").
 *  - "packed dense" = H rows of ceil(W/64) uint64 words, no padding.
 *  - device pointers are plain CUDA device pointers; `stream` is a cudaStream_t
 *    passed as void* (NULL = the engine's / legacy default stream).
 *
 * Reference parity citations use S=src/game.c, M=src/game_mpi.c,
 * A=src/game_mpi_async.c, C=src/game_mpi_collective.c, O=src/game_openmp.c,
 * G=src/game_cuda.cu, T=include/timestamp.h of the reference repository.
 */
#ifndef GOL_B200_H_
#define GOL_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GOL_VERSION 100

/* Reference compile-time defaults (S:6-9, G:6-9); run-time overridable here. */
#define GOL_DEFAULT_GEN_LIMIT 1000
#define GOL_DEFAULT_SIMILARITY_FREQUENCY 3

/* Ghost depth of the padded device layout = largest temporal-blocking depth. */
#define GOL_GHOST_ROWS 16
/* Ghost column slot on each side of a row: 128 cells = 4 x 32-bit words. */
#define GOL_GHOST_WORDS32 4
/* Smallest local domain the strip kernel accepts; smaller grids use the generic kernel. */
#define GOL_MIN_FAST_WIDTH 128
#define GOL_MIN_FAST_HEIGHT GOL_GHOST_ROWS

const char *gol_last_error(void);
int gol_version(void);
/* 1 if a CUDA device is usable from this process, else 0 (never fails). */
int gol_cuda_available(void);

/* ------------------------------------------------------------------ layout */

/* Padded per-domain device layout (store/, replaces the reference's three grid
 * stores S:136-147, M:192-197, G:185-191).  Rows y in [-G, H+G) are stored at
 * padded row y+G; each row is
 *   [4 words left ghost][nw32 data words][4 words right ghost][pad]
 * in 32-bit words, nw32 = 2*ceil(W/64).  The left ghost slot holds the west
 * neighbour's last 128 cells; the right slot holds the east neighbour's first
 * cells pre-shifted by W%64 so that it lines up with this row's partial word. */
typedef struct gol_layout {
    int64_t width, height;
    int32_t ghost_rows;
    int32_t nq;         /* ceil(width/64): data words (64-bit) per row       */
    int32_t nqf;        /* floor(width/64): full 64-bit words                */
    int32_t rem;        /* width % 64                                        */
    int32_t nw32;       /* 2*nq                                              */
    int32_t pitch32;    /* row pitch in 32-bit words (multiple of 8)         */
    int64_t rows_padded;
    int64_t buffer_words32;
} gol_layout;

int gol_layout_init(gol_layout *L, int64_t width, int64_t height);

/* --------------------------------------------------------------- CPU oracle */

/* Threaded bit-packed CPU implementation of the serial reference semantics
 * (S:60-131, S:169-196).  `grid` is packed dense and is advanced in place. */
int gol_cpu_step(uint64_t *grid, int64_t width, int64_t height, int64_t generations, int threads);

typedef struct gol_result {
    int32_t generations;      /* value printed as "Generations:" (S:202)          */
    int32_t exit_reason;      /* GOL_EXIT_*                                        */
    int64_t generations_computed; /* generations actually evolved incl. speculation */
    double  loop_ms;          /* device (or CPU) time of the generation loop       */
    double  read_ms, write_ms;
    uint64_t population;      /* live cells in the final grid                      */
} gol_result;

#define GOL_EXIT_LIMIT 0
#define GOL_EXIT_EMPTY 1
#define GOL_EXIT_SIMILAR 2

/* Full serial loop (S:169-196) on the CPU oracle: exact Generations value. */
int gol_cpu_run(uint64_t *grid, int64_t width, int64_t height, int gen_limit,
                int similarity_frequency, int threads, gol_result *res);

/* ----------------------------------------------------------- host utilities */

/* Text <-> packed on the host (reader S:149-167 / writer S:25-40 semantics).
 * gol_text_parse: skips '\n' (and '\r'); '1' alive, '0' dead; with strict!=0 any
 * other byte is an error, otherwise it is a dead cell; a short input is an
 * error (the reference spins forever, defect D8). */
int gol_text_parse(const char *text, size_t len, int64_t width, int64_t height, int strict,
                   uint64_t *packed);
/* Writes height*(width+1) bytes: rows of '0'/'1' + '\n' (S:29-36, A:415-416). */
int gol_text_format(const uint64_t *packed, int64_t width, int64_t height, char *text);

int gol_file_read_packed(const char *path, int64_t width, int64_t height, int strict,
                         uint64_t *packed);
int gol_file_write_packed(const char *path, const uint64_t *packed, int64_t width, int64_t height);

/* Seeded replacement for generate.sh (generate.sh:3-13): i.i.d. P(1)=1/2 cells,
 * a pure function of (seed, x, y) so every decomposition produces the same grid. */
int gol_random_packed(uint64_t *packed, int64_t width, int64_t height, uint64_t seed);
/* Streams the same grid as text into `path` (height lines of width chars). */
int gol_random_text_file(const char *path, int64_t width, int64_t height, uint64_t seed);

/* Headless counterpart of the reference's dead VT100 show() (S:42-58): writes a
 * plain PBM (P4) image of the packed grid. */
int gol_write_pbm(const char *path, const uint64_t *packed, int64_t width, int64_t height);

/* ------------------------------------------------- stand-alone device ops */

/* One torus generation on a packed dense device grid (any width/height >= 1).
 * Replaces evolve + halo_rows + halo_cols (G:52-74, G:128-148).  If `flags`
 * (device, 2 x uint32) is non-NULL it is OR-ed with {alive(out), out != in}. */
int gol_op_step_dense(const uint64_t *d_in, uint64_t *d_out, int64_t width, int64_t height,
                      uint32_t *d_flags, void *stream);
/* empty (G:102-126) and compare (G:76-100) as one pass: d_flags[0] |= any alive
 * in a, d_flags[1] |= (a != b) when b != NULL. */
int gol_op_flags_dense(const uint64_t *d_a, const uint64_t *d_b, int64_t width, int64_t height,
                       uint32_t *d_flags, void *stream);
/* cells: uint8 [H][W] with 0/non-zero values <-> packed dense. */
int gol_op_pack_cells(const uint8_t *d_cells, uint64_t *d_packed, int64_t width, int64_t height,
                      void *stream);
int gol_op_unpack_cells(const uint64_t *d_packed, uint8_t *d_cells, int64_t width, int64_t height,
                        void *stream);
/* Text parse/format kernels on fixed-record text (width+1 bytes per line).
 * d_err (uint32) is OR-ed with 1 on a byte other than '0','1' and 2 on a
 * missing newline. */
int gol_op_parse_text(const char *d_text, uint64_t *d_packed, int64_t width, int64_t height,
                      int64_t text_pitch, uint32_t *d_err, void *stream);
/* Same with the length of the text buffer: the newline of a row is only checked where the buffer
 * has a byte at column `width` (a last line without a newline is accepted), and nothing outside
 * [d_text, d_text + text_bytes) is read. */
int gol_op_parse_text_n(const char *d_text, int64_t text_bytes, uint64_t *d_packed, int64_t width,
                        int64_t height, int64_t text_pitch, uint32_t *d_err, void *stream);
int gol_op_format_text(const uint64_t *d_packed, char *d_text, int64_t width, int64_t height,
                       int64_t text_pitch, int newline, void *stream);
int gol_op_random_dense(uint64_t *d_packed, int64_t width, int64_t height, int64_t x0, int64_t y0,
                        uint64_t seed, void *stream);
int gol_op_popcount_dense(const uint64_t *d_packed, int64_t width, int64_t height,
                          unsigned long long *d_count, void *stream);

/* ------------------------------------------------------------------ engine */

typedef struct gol_engine gol_engine;

#define GOL_FLAG_FORCE_GENERIC 1u   /* always use the 1-generation generic kernel   */
#define GOL_FLAG_NO_GRAPH      2u   /* do not capture epochs into CUDA graphs        */
#define GOL_FLAG_VERBOSE       4u
#define GOL_FLAG_NCCL_BASELINE 8u   /* (baseline only) halos are moved by the caller */

typedef struct gol_config {
    int64_t width, height;      /* global grid                                       */
    int32_t grid_rows, grid_cols; /* R x C domain grid (2D block decomposition,      */
                                  /* periodic; replaces MPI_Cart_create C:125-133)   */
    int32_t proc_rank, proc_count; /* this process / number of processes; domains    */
                                  /* are dealt to processes in contiguous blocks     */
    int32_t device;             /* CUDA device ordinal for this process              */
    int32_t gen_limit;          /* GEN_LIMIT (<=0: default 1000)                     */
    int32_t similarity_frequency; /* SIMILARITY_FREQUENCY (<0: default 3, 0: off)    */
    int32_t k_max;              /* temporal-blocking depth per launch (0 = auto)     */
    int32_t batch_gens;         /* generations between flag checks (0 = auto)        */
    uint32_t flags;
    int32_t start_generation;   /* resume support: the loaded grid is generation g (default 0); the   */
                                /* loop continues with evolve ordinal g+1, so the similarity phase     */
                                /* and the reported Generations carry on where a previous run stopped  */
    int32_t reserved;
} gol_config;

void gol_config_default(gol_config *cfg, int64_t width, int64_t height);

int gol_engine_create(const gol_config *cfg, gol_engine **out);
void gol_engine_destroy(gol_engine *e);
int gol_engine_set_stream(gol_engine *e, void *stream);

/* Topology queries (replace MPI_Cart_coords / MPI_Cart_rank, C:225-275). */
int gol_engine_num_domains(const gol_engine *e);          /* R*C                     */
int gol_engine_num_local_domains(const gol_engine *e);
int gol_engine_local_domain_id(const gol_engine *e, int local_index); /* global id   */
int gol_engine_domain_extent(const gol_engine *e, int domain_id, int64_t *x0, int64_t *y0,
                             int64_t *w, int64_t *h);
int gol_engine_domain_neighbors(const gol_engine *e, int domain_id, int32_t nb[8]);
int gol_engine_uses_fast_path(const gol_engine *e);

/* Multi-process bootstrap.  Each process exports one blob (CUDA IPC handles of
 * its buffers and flag tables), the caller all-gathers the blobs with any
 * transport (torch.distributed, the CLI's shared-memory rendezvous), and each
 * process imports the concatenation (proc_count * blob_size bytes, rank order). */
size_t gol_engine_ipc_blob_size(const gol_engine *e);
int gol_engine_ipc_export(gol_engine *e, void *blob);
int gol_engine_ipc_import(gol_engine *e, const void *all_blobs);

/* Grid load/store.  Every process passes the same arguments. */
int gol_engine_load_packed(gol_engine *e, const uint64_t *host_packed); /* global dense    */
/* Overlapped upload for a stream of grids: stage() starts the (asynchronous, if `host_packed` is
 * pinned) copy of the next global grid on a second stream and returns; commit() turns the staged
 * grid into the current one.  Between the two calls the engine may be running the previous grid. */
int gol_engine_stage_packed(gol_engine *e, const uint64_t *host_packed);
int gol_engine_commit_staged(gol_engine *e);
/* Per-domain variants for callers that keep only their own blocks on the host: local domain
 * `local_index` as packed dense rows [h][ceil(w/64)].  stage_domain is asynchronous like
 * stage_packed (stage every local domain, then commit); download_domain returns when the data is
 * in host memory. */
int gol_engine_stage_domain(gol_engine *e, int local_index, const uint64_t *host_dense);
int gol_engine_download_domain(gol_engine *e, int local_index, uint64_t *host_dense);
int gol_engine_load_packed_device(gol_engine *e, const uint64_t *d_packed, int64_t x0,
                                  int64_t y0, int64_t w, int64_t h);    /* one local block */
int gol_engine_randomize(gol_engine *e, uint64_t seed);
/* Parallel read: each process preads only the byte ranges of its domains and
 * parses them on the GPU (replaces M:201-254, A:174-201, C:174-198). */
int gol_engine_load_text_file(gol_engine *e, const char *path, int strict);
/* Parallel write: format kernel + disjoint pwrite stripes; the right-most
 * domain column owns the '\n' bytes (C:382-393, C:417-421).  An existing file
 * is replaced (A:432-439). */
int gol_engine_store_text_file(gol_engine *e, const char *path);
int gol_engine_download_packed(gol_engine *e, uint64_t *host_packed);   /* local domains   */
                                                                          /* into global grid */

/* Must be called (by every process) after a load and before step/run: fills all
 * ghost zones of the current buffer (first halo exchange). */
int gol_engine_prepare(gol_engine *e);

/* Advance exactly n generations, no early exit (similarity-off pure stencil). */
int gol_engine_step(gol_engine *e, int64_t n);
/* The reference generation loop (S:169-196): early exits on empty / similar,
 * exact Generations, final grid g_Generations left in the current buffer. */
int gol_engine_run(gol_engine *e, gol_result *res);
int gol_engine_sync(gol_engine *e);

/* Kernel launches issued by this engine since creation (for bench accounting). */
int64_t gol_engine_launch_count(const gol_engine *e);
/* Tuning facts for reports: height of the interior chunks the strip kernel uses at the engine's
 * temporal depth (after gol_engine_prepare: the autotuned value); whether epochs run as deep-halo
 * pairs (one halo exchange per two epochs). */
int gol_engine_chunk_rows(const gol_engine *e);
int gol_engine_uses_deep_halo(const gol_engine *e);
/* Per-launch device timeline (engines created with GOL_TRACE=<max launches> in the environment):
 * text file, one line per generation-kernel launch. */
int gol_engine_trace_dump(gol_engine *e, const char *path);
/* Local population / flags of the current buffer. */
/* gol_engine_hash: position-dependent 64-bit checksum of the local domains (sum over words of
 * hash(value, global word x, global row y), mod 2^64).  The sum over all processes identifies the
 * global grid independently of the decomposition; gol_hash_packed is the same number computed
 * on the host from a packed dense grid. */
int gol_engine_hash(gol_engine *e, uint64_t *hash);
int gol_hash_packed(const uint64_t *packed, int64_t width, int64_t height, uint64_t *hash);
int gol_engine_population(gol_engine *e, uint64_t *count);

/* Same result as gol_engine_population, for callers that may not allocate device memory: a
 * process whose GPU memory is owned wholesale by another allocator (a framework pool), where a
 * raw cudaMalloc fails or fights the pool, and latency-bound loops, where the device-wide
 * synchronisation inside cudaMalloc / cudaFree is not affordable.  (No counterpart in the
 * reference.)  It calls no cudaMalloc, cudaFree or allocating library routine: the dense
 * staging copy and the counter live in a static device array of GOL_STATIC_SCRATCH_BYTES + 256
 * bytes inside the library image, present on every GPU that loads the library until exit.
 *  - limit: each local domain must fit the staging area, ceil(w/64) * 8 * h <= 4 MiB
 *    (32 Mi cells, e.g. 4096 x 8192); larger domains are rejected with an error and nothing
 *    is launched.
 *  - the first call on a device looks the array's address up (cudaGetSymbolAddress) and keeps
 *    it; make that first call before any stream capture begins.
 *  - all calls share the one array and it keeps the last call's contents; the counter is
 *    cleared at the start of every call.  Calls that overlap on the GPU (from several threads
 *    or engines) race on it: serialise them.
 *  - runs on the calling thread's per-thread default stream after synchronising the engine. */
#define GOL_STATIC_SCRATCH_BYTES (4u << 20)
int gol_engine_population_static(gol_engine *e, uint64_t *count);

/* Zero-copy access to a local domain's current padded buffer (for the PyTorch
 * tensor view): device pointer to padded word (row 0, word 0) + its layout. */
int gol_engine_domain_buffer(gol_engine *e, int local_index, void **d_ptr, gol_layout *layout);

/* Device-timed duration (CUDA events) of the last gol_engine_step / run. */
double gol_engine_last_ms(const gol_engine *e);

#ifdef __cplusplus
}
#endif
#endif /* GOL_B200_H_ */
