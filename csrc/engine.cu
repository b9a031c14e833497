// Engine: domain decomposition, device arena, multi-process bootstrap, grid load/store and the
// epoch driver with the reference's exact early-exit semantics.
//
// Replaces: MPI_Cart_create / Cart_coords / Cart_rank topology (C:120-137, C:225-275), the
// generation loops (S:169-203, C:328-370, G:213-297), master scatter/gather (M:221-254,
// M:429-472), MPI-IO subarray read/write (A:174-201, C:174-198, C:379-445), timers
// (S:175, M:187.., T:8-20) and the compile-time configuration (S:6-9).
//
// Driver design.  Both early exits of the reference are fixed points (an empty grid stays
// empty; g_e == g_(e-1) implies every later generation is equal), so generations are run
// speculatively in batches of `batch_gens` without any host round trip: only the last epoch
// of a batch reduces the alive / changed bits (fused in the generation kernel), a one-CTA
// gather kernel all-reduces them across domains, and the host looks at one 8-byte word per
// batch.  When a batch ends dead or unchanged, the batch is replayed from its untouched
// input buffer one generation at a time with per-generation flags, and the serial loop is
// evaluated literally on those flags to obtain the exact "Generations:" value.
#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>

#include <algorithm>
#include <cerrno>
#include <cstdlib>
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>
#include <atomic>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "gol_internal.h"

namespace {

// NVTX range for the timeline of a profiler (phases: load / prepare / batch / replay / store).
struct NvtxRange {
    explicit NvtxRange(const char *name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
};

constexpr int G = GOL_GHOST_ROWS;
constexpr int kMaxBatch = 512;

struct LocalDomain {
    int gid = 0, r = 0, c = 0;
    int64_t x0 = 0, y0 = 0;
    gol_layout L{};
    int strips = 0;
};

// Device arena of one process: every process can compute every other process's arena layout
// from the configuration alone, so one IPC handle per process is all that has to be exchanged.
struct ArenaLayout {
    std::vector<int64_t> buf_off;     // [local domain][GOL_NBUF] byte offsets
    std::vector<int64_t> flags_off;   // [local domain]
    std::vector<int64_t> scratch_off; // [local domain]
    int64_t rec_off = 0;
    int64_t total = 0;
};

// Text I/O staging: one lane per worker thread, two slots per lane (see io_prepare).
struct IoSlot {
    char *h = nullptr, *d = nullptr;   // pinned host / device staging of one row block
    cudaEvent_t ev = nullptr;          // the device work that last used the slot
};
struct IoLane {
    cudaStream_t stream = nullptr;
    IoSlot slot[2];
};

struct IpcBlob {
    cudaIpcMemHandle_t handle;
    int32_t device;
    int32_t pid;
    int64_t arena_bytes;
};

}  // namespace

struct gol_engine {
    gol_config cfg{};
    int R = 1, C = 1, ndom = 1, P = 1, rank = 0, per_proc = 1;
    std::vector<int64_t> col_x, row_y;  // domain boundaries (C+1, R+1)
    bool fast = false;
    int num_sms = 148;

    // strip path
    std::vector<LocalDomain> local;
    ArenaLayout arena_layout;
    char *arena = nullptr;
    std::vector<char *> peer_arena;  // [P]; own entry = arena
    bool imported = false;
    GolDomainDev *d_doms = nullptr;
    unsigned long long **d_rec_tables = nullptr;
    unsigned long long *d_results = nullptr;  // [kMaxBatch + 1]
    unsigned long long *h_results = nullptr;  // pinned
    unsigned int *d_error = nullptr;
    uint64_t *stage = nullptr;      // dense staging buffer for linear host->device copies (all local domains)
    std::vector<size_t> stage_off;  // word offset of each local domain inside `stage`
    cudaStream_t copy_stream = nullptr;  // staged (overlapped) uploads run here
    cudaEvent_t ev_staged = nullptr, ev_stage_free = nullptr;
    bool staged = false, stage_busy = false;
    int max_strips = 1, max_strips_ext = 1;  // strips per row: plain / with the ghost words of a deep-halo epoch
    int64_t max_h = 1;
    int tuned_rows[17] = {0};       // measured best chunk height per temporal depth (0 = not tuned yet)
    bool autotune = true;           // GOL_AUTOTUNE=0 disables the one-off geometry timing
    bool deep_halo = false;         // epochs run as pairs with one 16-deep halo exchange per pair (see run_epochs)
    int pdl_mode = 2;               // GOL_PDL: 0 no programmatic dependent launch, 1 single-process engines only, 2 always
    long long wait_timeout = 0;     // device-side flag waits give up after this many clock cycles
    bool fault_wait = false;        // GOL_FAULT_INJECT=wait: epochs wait for a flag sequence nobody publishes (tests)
    // GOL_TRACE=<n>: per-launch timeline records (first CTA start, last CTA end, flag spin cycles)
    GolTrace *d_trace = nullptr;
    int trace_cap = 0, trace_n = 0;
    std::vector<int> trace_meta;    // per record: k | mode << 8 | flags << 16
    std::vector<IoLane> io;         // pipelined text I/O staging (allocated on first use)
    size_t io_slot_bytes = 0;
    uint64_t *scratch_dense = nullptr;  // engine-owned dense scratch (largest local domain): randomize, population
    unsigned long long *d_count = nullptr;

    // dense path (tiny grids, single domain)
    uint64_t *dense[3] = {nullptr, nullptr, nullptr};
    uint32_t *d_dflags = nullptr;  // [kMaxBatch + 1][2]
    uint32_t *h_dflags = nullptr;  // pinned

    int cur = 0, hist[2] = {-1, -1};  // current buffer; inputs of the last two epochs
    uint32_t seq = 0, rec_seq = 0;
    bool prepared = false;
    int64_t launches = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_slot[2] = {nullptr, nullptr}, ev_tune[2] = {nullptr, nullptr};
    cudaStream_t flag_stream = nullptr;   // flag gathers + their read-back run beside the generation stream
    cudaEvent_t ev_flag[2] = {nullptr, nullptr};
    double last_ms = 0.0;
    int kmax = 8, batch = 128;
};

namespace {

int64_t align_up(int64_t v, int64_t a) { return (v + a - 1) / a * a; }

void domain_extent(const gol_engine *e, int gid, int64_t *x0, int64_t *y0, int64_t *w, int64_t *h)
{
    const int r = gid / e->C, c = gid % e->C;
    *x0 = e->col_x[c];
    *y0 = e->row_y[r];
    *w = e->col_x[c + 1] - e->col_x[c];
    *h = e->row_y[r + 1] - e->row_y[r];
}

int neighbor_gid(const gol_engine *e, int gid, int dir)
{
    const int r = gid / e->C, c = gid % e->C;
    const int nr = (r + GOL_DIR_DY[dir] + e->R) % e->R, nc = (c + GOL_DIR_DX[dir] + e->C) % e->C;
    return nr * e->C + nc;
}

int owner_of(const gol_engine *e, int gid) { return gid / e->per_proc; }

// arena layout of process p
int build_arena_layout(const gol_engine *e, int p, ArenaLayout *A)
{
    A->buf_off.assign((size_t)e->per_proc * GOL_NBUF, 0);
    A->flags_off.assign((size_t)e->per_proc, 0);
    A->scratch_off.assign((size_t)e->per_proc, 0);
    int64_t off = 0;
    for (int i = 0; i < e->per_proc; ++i) {
        const int gid = p * e->per_proc + i;
        int64_t x0, y0, w, h;
        domain_extent(e, gid, &x0, &y0, &w, &h);
        gol_layout L;
        GOL_TRY(gol_layout_init(&L, w, h));
        for (int b = 0; b < GOL_NBUF; ++b) {
            A->buf_off[(size_t)i * GOL_NBUF + b] = off;
            off = align_up(off + L.buffer_words32 * 4, 256);
        }
    }
    for (int i = 0; i < e->per_proc; ++i) {
        A->flags_off[(size_t)i] = off;
        off += 256;
        A->scratch_off[(size_t)i] = off;
        off += 256;
    }
    A->rec_off = off;
    off = align_up(off + (int64_t)GOL_REC_SLOTS * e->ndom * 8, 256);
    A->total = off;
    return 0;
}

// (Re)build the device-side domain descriptors from the local + peer arenas.
int upload_domains(gol_engine *e)
{
    std::vector<GolDomainDev> doms((size_t)e->per_proc);
    std::vector<ArenaLayout> layouts((size_t)e->P);
    for (int p = 0; p < e->P; ++p) GOL_TRY(build_arena_layout(e, p, &layouts[(size_t)p]));
    std::vector<unsigned long long *> tables((size_t)e->P, nullptr);
    for (int p = 0; p < e->P; ++p)
        if (e->peer_arena[(size_t)p])
            tables[(size_t)p] = reinterpret_cast<unsigned long long *>(e->peer_arena[(size_t)p] + layouts[(size_t)p].rec_off);
    GOL_CUDA(cudaMemcpy(e->d_rec_tables, tables.data(), sizeof(void *) * (size_t)e->P, cudaMemcpyHostToDevice));

    for (int i = 0; i < e->per_proc; ++i) {
        const LocalDomain &ld = e->local[(size_t)i];
        GolDomainDev &D = doms[(size_t)i];
        memset(&D, 0, sizeof D);
        const ArenaLayout &A = layouts[(size_t)e->rank];
        for (int b = 0; b < GOL_NBUF; ++b)
            D.buf[b] = reinterpret_cast<uint32_t *>(e->arena + A.buf_off[(size_t)i * GOL_NBUF + b]);
        D.H = (int32_t)ld.L.height;
        D.nw32 = ld.L.nw32;
        D.nwf32 = 2 * ld.L.nqf;
        D.rem = ld.L.rem;
        D.pitch32 = ld.L.pitch32;
        D.global_id = ld.gid;
        D.epoch_flags = reinterpret_cast<uint32_t *>(e->arena + A.flags_off[(size_t)i]);
        D.scratch = reinterpret_cast<GolScratch *>(e->arena + A.scratch_off[(size_t)i]);
        D.rec_tables = e->d_rec_tables;
        D.proc_count = e->P;
        D.num_domains = e->ndom;
        D.wait_timeout = e->wait_timeout;
        for (int d = 0; d < GOL_NDIR; ++d) {
            const int ngid = neighbor_gid(e, ld.gid, d);
            if (ngid == ld.gid) D.self_mask |= 1u << d;
            const int q = owner_of(e, ngid), li = ngid - q * e->per_proc;
            char *base = e->peer_arena[(size_t)q];
            GolNeighbor &N = D.nb[d];
            int64_t x0, y0, w, h;
            domain_extent(e, ngid, &x0, &y0, &w, &h);
            gol_layout NL;
            GOL_TRY(gol_layout_init(&NL, w, h));
            N.H = (int32_t)h;
            N.nw32 = NL.nw32;
            N.rem = NL.rem;
            N.pitch32 = NL.pitch32;
            if (base) {
                const ArenaLayout &NA = layouts[(size_t)q];
                for (int b = 0; b < GOL_NBUF; ++b)
                    N.buf[b] = reinterpret_cast<uint32_t *>(base + NA.buf_off[(size_t)li * GOL_NBUF + b]);
                N.flag = reinterpret_cast<uint32_t *>(base + NA.flags_off[(size_t)li]) + GOL_DIR_OPP[d];
            }
        }
    }
    GOL_CUDA(cudaMemcpy(e->d_doms, doms.data(), sizeof(GolDomainDev) * doms.size(), cudaMemcpyHostToDevice));
    return 0;
}

bool multi(const gol_engine *e) { return e->P > 1; }
int uses_pdl(const gol_engine *e) { return e->pdl_mode >= 2 || (e->pdl_mode == 1 && !multi(e)) ? 1 : 0; }

// Launch geometry for temporal depth k.  Every domain is cut into a first and a last chunk of
// GOL_EDGE_ROWS rows (the rows that are exchanged with the north / south neighbours; short, so that
// they are done early in the launch) and interior chunks whose height trades apron redundancy
// (2k rows of pipeline fill per chunk) against the number of independent warps.
// `rows` > 0 forces the interior chunk height (autotune); ext_y / ext_x: deep-halo epoch.
void geometry(const gol_engine *e, int k, GolLaunch *lp, int ext_y = 0, int ext_x = 0, int rows = 0)
{
    static const int env_warps = getenv("GOL_WARPS_PER_SM") ? atoi(getenv("GOL_WARPS_PER_SM")) : 0;
    const int E = GOL_EDGE_ROWS;
    const int64_t S = ext_x ? e->max_strips_ext : e->max_strips;
    const int64_t span = e->max_h + 2 * ext_y;
    const bool single = span < 3 * E;
    const int64_t interior = single ? span : span - 2 * E;
    lp->edge_rows = E;
    lp->ext_y = ext_y;
    lp->ext_x = ext_x;
    static const int env_rows = getenv("GOL_CHUNK_ROWS") ? atoi(getenv("GOL_CHUNK_ROWS")) : 0;  // fixed height (profiling)
    if (!rows && env_rows > 0) rows = (int)std::min<int64_t>(env_rows, interior);
    if (!rows && !env_warps && k <= 16 && e->tuned_rows[k] > 0) rows = e->tuned_rows[k];
    if (!rows) {
        const int warps_per_sm = env_warps ? env_warps : 12;
        const int64_t target_items = (int64_t)e->num_sms * warps_per_sm / std::max(1, e->per_proc);
        const int64_t chunks = std::max<int64_t>(1, gol_div_up(target_items, S) - (single ? 0 : 2));
        rows = (int)std::max<int64_t>(gol_div_up(interior, chunks), std::min<int64_t>(interior, std::max(8, 6 * k)));
    }
    lp->chunk_rows = rows;
    const int64_t items = single ? S : (2 + gol_div_up(interior, rows)) * S;
    lp->ctas_per_domain = (int32_t)gol_div_up(items, GOL_STRIP_WARPS);
}

// largest temporal depth the strip kernel is instantiated for that does not exceed cap
int pick_depth(int64_t cap)
{
    if (cap >= 16) return 16;
    if (cap >= 8) return 8;
    return (int)std::max<int64_t>(1, cap);
}

int check_ready(gol_engine *e, const char *what)
{
    if (e->P > 1 && !e->imported) GOL_FAIL("%s: multi-process engine used before gol_engine_ipc_import", what);
    return 0;
}

GolTrace *next_trace(gol_engine *e, int k, int mode, bool flags)
{
    if (!e->d_trace || e->trace_n >= e->trace_cap) return nullptr;
    e->trace_meta.push_back(k | (mode << 8) | ((flags ? 1 : 0) << 16));
    return e->d_trace + e->trace_n++;
}

// Epoch kinds.  CLASSIC waits for the neighbours' pushes of the previous epoch and pushes its own
// borders.  A deep-halo pair exchanges halos once per two epochs: PAIR_FIRST waits, also computes
// the k ghost rows / 64 ghost cells next to the domain from the 16-row / 128-cell deep ghost zone,
// and pushes nothing; PAIR_SECOND needs no remote data and pushes a full-depth halo.
enum EpochMode { CLASSIC = 0, PAIR_FIRST = 1, PAIR_SECOND = 2 };

// one strip-kernel epoch: k generations from buffer `in` to buffer `out`
int launch_epoch(gol_engine *e, int k, int in, int out, bool flags, int rec_slot, uint32_t rec_seq, EpochMode mode)
{
    GolLaunch lp{};
    const int ext = mode == PAIR_FIRST ? k : 0;
    geometry(e, k, &lp, ext, ext ? 2 : 0);
    lp.in_idx = in;
    lp.out_idx = out;
    static const int dbg = getenv("GOL_DEBUG_COMM") ? atoi(getenv("GOL_DEBUG_COMM")) : 0;  // timing experiments only
    lp.wait_seq = (multi(e) && mode != PAIR_SECOND && !(dbg & 4)) ? e->seq : 0;
    if (e->fault_wait && mode != PAIR_SECOND) lp.wait_seq = e->seq + 1000;   // fault injection: the neighbours never get there
    lp.push = (mode != PAIR_FIRST ? (multi(e) ? 3 : 1) : 0) | ((dbg & 3) << 8);
    lp.seq = mode != PAIR_FIRST ? ++e->seq : e->seq;
    lp.push_rows = G;
    // waiting items go last (the neighbours' flags have a whole epoch to arrive), pushing items first
    lp.border_last = mode == PAIR_FIRST ? 1 : 0;
    lp.pdl = uses_pdl(e);
    lp.rec_slot = flags ? rec_slot : -1;
    lp.rec_seq = rec_seq;
    lp.trace = next_trace(e, k, mode, flags);
    GOL_TRY(gol_launch_strip(e->d_doms, e->per_proc, lp, k, flags ? 1 : 0, e->stream));
    ++e->launches;
    return 0;
}

// One-off geometry tuning for temporal depth k: the run time of an epoch is far from monotonic in
// the chunk height (apron rows per chunk against wave quantisation of the CTA scheduler), so a
// handful of candidate heights is timed on the real grid - pushes and flags switched off, output
// into a free buffer that the next real epoch overwrites - and the fastest is kept.
int autotune_geometry(gol_engine *e, int k, int in, int scratch)
{
    if (k > 16 || e->tuned_rows[k] || e->max_h < 256 || getenv("GOL_CHUNK_ROWS")) return 0;
    const int64_t interior = e->max_h - 2 * GOL_EDGE_ROWS;
    const int min_rows = std::max(8, 2 * k);
    std::vector<int> cands;
    auto add = [&](int64_t rows) {
        if (rows < min_rows || rows > interior) return;
        if (std::find(cands.begin(), cands.end(), (int)rows) == cands.end()) cands.push_back((int)rows);
    };
    for (int w = 10; w <= 26; ++w) {
        const int64_t target = (int64_t)e->num_sms * w / std::max(1, e->per_proc);
        add(gol_div_up(interior, std::max<int64_t>(1, gol_div_up(target, e->max_strips) - 2)));
    }
    // heights that fill the machine exactly: m full waves of resident CTAs, no partial last wave
    int occ = 0;
    if (gol_strip_ctas_per_sm(k, &occ) == 0 && occ > 0) {
        const int64_t slots = (int64_t)e->num_sms * occ * GOL_STRIP_WARPS / std::max(1, e->per_proc);  // resident work items
        for (int m = 1; m <= 3; ++m) {
            const int64_t chunks = m * slots / e->max_strips - 2;
            if (chunks < 1) continue;
            for (int extra = 0; extra <= 1; ++extra) add(gol_div_up(interior, chunks) + extra);
        }
    }
    float best = 1e30f;
    int best_rows = 0;
    auto time_rows = [&](int rows, float *ms_out) -> int {
        GolLaunch lp{};
        geometry(e, k, &lp, 0, 0, rows);
        lp.in_idx = in;
        lp.out_idx = scratch;
        lp.seq = e->seq;
        lp.wait_seq = 0;
        lp.push = 0;                 // no pushes, no flags
        lp.push_rows = G;
        lp.pdl = uses_pdl(e);        // same launch mode as the real epochs
        lp.rec_slot = -1;
        GOL_TRY(gol_launch_strip(e->d_doms, e->per_proc, lp, k, 0, e->stream));
        GOL_CUDA(cudaEventRecord(e->ev_tune[0], e->stream));
        for (int i = 0; i < 3; ++i) GOL_TRY(gol_launch_strip(e->d_doms, e->per_proc, lp, k, 0, e->stream));
        GOL_CUDA(cudaEventRecord(e->ev_tune[1], e->stream));
        GOL_CUDA(cudaEventSynchronize(e->ev_tune[1]));
        e->launches += 4;
        GOL_CUDA(cudaEventElapsedTime(ms_out, e->ev_tune[0], e->ev_tune[1]));
        if (e->cfg.flags & GOL_FLAG_VERBOSE)
            fprintf(stderr, "[gol autotune] k=%d chunk_rows=%d: %.2f us/epoch\n", k, rows, *ms_out * 1e3f / 3);
        return 0;
    };
    for (int rows : cands) {
        float ms = 0;
        GOL_TRY(time_rows(rows, &ms));
        if (ms < best) { best = ms; best_rows = rows; }
    }
    // local refinement around the coarse winner
    const int centre = best_rows;
    for (int d = -3; d <= 3; ++d) {
        const int rows = centre + d;
        if (d == 0 || rows < min_rows || rows > interior) continue;
        if (std::find(cands.begin(), cands.end(), rows) != cands.end()) continue;
        float ms = 0;
        GOL_TRY(time_rows(rows, &ms));
        if (ms < best) { best = ms; best_rows = rows; }
    }
    e->tuned_rows[k] = best_rows;
    return 0;
}

// Output buffer of the next epoch.  Never the epoch's input nor a checkpoint.  Between processes
// a rank may run ahead of its neighbours: its pushes land in the neighbour's buffer of the same
// index while that neighbour is still up to two epochs behind (deep-halo pairs: the pushing epoch
// follows an epoch that waited for pushes made two epochs earlier), so the inputs of the last
// two epochs are excluded as well - four rotating buffers.
int pick_out(const gol_engine *e, int in, int keep, int keep2, int *out)
{
    for (int relax = 0; relax <= (multi(e) ? 0 : 2); ++relax)
        for (int i = 0; i < GOL_NBUF; ++i) {
            if (i == in || i == keep || i == keep2) continue;
            if (relax < 2 && i == e->hist[0]) continue;
            if (relax < 1 && i == e->hist[1]) continue;
            *out = i;
            return 0;
        }
    GOL_FAIL("internal error: no free grid buffer for the next epoch");
}

// Run n generations starting from e->cur as epochs of at most kcap; buffers `keep`/`keep2` are never
// written.  With want_flags the last epoch publishes a record (rec slot/seq returned).
int run_epochs(gol_engine *e, int64_t n, int keep, bool want_flags, bool per_gen_flags, int kcap, int keep2 = -1)
{
    int in = e->cur;
    int64_t left = n;
    bool second = false;  // the next epoch completes a deep-halo pair
    int k = 1;
    while (left > 0) {
        EpochMode mode = CLASSIC;
        if (second) {
            mode = PAIR_SECOND;
            second = false;
        } else {
            k = per_gen_flags ? 1 : pick_depth(std::min<int64_t>(left, kcap));
            if (e->deep_halo && !per_gen_flags && 2 * k <= G && left >= 2 * k) {
                mode = PAIR_FIRST;
                second = true;
            }
        }
        int out = -1;
        GOL_TRY(pick_out(e, in, keep, keep2, &out));
        const bool last = left - k == 0;
        const bool f = per_gen_flags || (want_flags && last);
        uint32_t rs = 0;
        int slot = -1;
        if (f) {
            rs = ++e->rec_seq;
            slot = (int)(rs % GOL_REC_SLOTS);
        }
        GOL_TRY(launch_epoch(e, k, in, out, f, slot, rs, mode));
        if (f && per_gen_flags) {
            // in replay mode every generation is gathered (also keeps the record ring bounded)
            const ArenaLayout &A = e->arena_layout;
            GOL_TRY(gol_launch_gather(reinterpret_cast<unsigned long long *>(e->arena + A.rec_off), slot, e->ndom,
                                      rs, e->d_results + (n - left), e->d_error, e->wait_timeout * 4, e->stream));
            ++e->launches;
        }
        e->hist[1] = e->hist[0];
        e->hist[0] = in;
        in = out;
        left -= k;
    }
    e->cur = in;
    return 0;
}

// Flag all-reduce of the record published last + asynchronous read-back of the 8-byte result into
// pinned host slot `result_index`.  Both run on a second stream behind an event, so the generation
// stream goes straight on with the next batch (in-stream they cost a ~16 us bubble per batch: a
// one-CTA kernel, an 8-byte copy and two launch latencies between two epochs).
int gather_last(gol_engine *e, int result_index)
{
    const uint32_t rs = e->rec_seq;
    const ArenaLayout &A = e->arena_layout;
    cudaEvent_t ev = e->ev_flag[result_index & 1];
    GOL_CUDA(cudaEventRecord(ev, e->stream));
    GOL_CUDA(cudaStreamWaitEvent(e->flag_stream, ev, 0));
    GOL_TRY(gol_launch_gather(reinterpret_cast<unsigned long long *>(e->arena + A.rec_off),
                              (int)(rs % GOL_REC_SLOTS), e->ndom, rs, e->d_results + result_index, e->d_error,
                              e->wait_timeout * 4, e->flag_stream));
    ++e->launches;
    return 0;
}

// Sticky per-domain error words set by device-side waits that timed out.
int check_device_errors(gol_engine *e)
{
    for (int i = 0; i < e->per_proc; ++i) {
        GolScratch sc;
        GOL_CUDA(cudaMemcpy(&sc, e->arena + e->arena_layout.scratch_off[(size_t)i], sizeof sc, cudaMemcpyDeviceToHost));
        if (sc.error)
            GOL_FAIL("device-side wait timed out in domain %d (code %u): a neighbouring work item or rank stopped "
                     "making progress", e->local[(size_t)i].gid, sc.error);
    }
    return 0;
}

int fetch_results(gol_engine *e, int count)
{
    GOL_CUDA(cudaMemcpyAsync(e->h_results, e->d_results, sizeof(unsigned long long) * (size_t)count,
                             cudaMemcpyDeviceToHost, e->stream));
    unsigned int err = 0;
    GOL_CUDA(cudaMemcpyAsync(&e->h_results[kMaxBatch + 1], e->d_error, sizeof(unsigned int), cudaMemcpyDeviceToHost,
                             e->stream));
    GOL_CUDA(cudaStreamSynchronize(e->stream));
    err = (unsigned int)e->h_results[kMaxBatch + 1];
    if (err) GOL_FAIL("device-side wait timed out (code %u): a neighbouring rank stopped making progress", err);
    return check_device_errors(e);
}

// Literal evaluation of the serial loop (S:177-196) on per-generation flags.
// alive[i], diff[i] describe generation gen0+i (i = 0..n); generation gen0 is known alive.
void resolve_exit(int64_t gen0, int n, const std::vector<int> &alive, const std::vector<int> &diff, int limit,
                  int freq, int *generations, int *reason)
{
    for (int i = 1; i <= n; ++i) {
        const int64_t ord = gen0 + i;  // evolve ordinal
        if (!alive[(size_t)i - 1]) { *generations = (int)(ord - 1); *reason = GOL_EXIT_EMPTY; return; }
        if (freq > 0 && ord % freq == 0 && !diff[(size_t)i]) { *generations = (int)(ord - 1); *reason = GOL_EXIT_SIMILAR; return; }
    }
    const int64_t end = gen0 + n;
    if (!alive[(size_t)n]) { *generations = (int)end; *reason = GOL_EXIT_EMPTY; return; }
    if (freq > 0 && !diff[(size_t)n]) {
        const int64_t next = (end / freq + 1) * freq;  // next similarity check after `end`
        if (next <= limit) { *generations = (int)(next - 1); *reason = GOL_EXIT_SIMILAR; return; }
    }
    *generations = limit;
    *reason = GOL_EXIT_LIMIT;
}

// Population for gol_result without any allocation on the run path: counts straight from the
// padded buffers into a spare word of the result array.
int fill_population(gol_engine *e, gol_result *res)
{
    if (!e->fast) {
        uint64_t pop = 0;
        GOL_TRY(gol_engine_population(e, &pop));
        res->population = pop;
        return 0;
    }
    unsigned long long *d_count = e->d_results + kMaxBatch + 1;
    GOL_CUDA(cudaMemsetAsync(d_count, 0, sizeof(unsigned long long), e->stream));
    for (int i = 0; i < e->per_proc; ++i) {
        const char *src = e->arena + e->arena_layout.buf_off[(size_t)i * GOL_NBUF + e->cur];
        GOL_TRY(gol_launch_popcount_padded(reinterpret_cast<const uint32_t *>(src), e->local[(size_t)i].L, d_count, e->stream));
        ++e->launches;
    }
    GOL_CUDA(cudaMemcpyAsync(&e->h_results[kMaxBatch], d_count, sizeof(unsigned long long), cudaMemcpyDeviceToHost, e->stream));
    GOL_CUDA(cudaStreamSynchronize(e->stream));
    res->population = e->h_results[kMaxBatch];
    return 0;
}

// ------------------------------------------------------------ strip-path run
// Asynchronous read-back of one 8-byte flag word: device result slot -> pinned host slot, then an
// event the host can wait on while the GPU already runs the next batch.
int post_result(gol_engine *e, int slot)
{
    GOL_CUDA(cudaMemcpyAsync(&e->h_results[slot], e->d_results + slot, sizeof(unsigned long long), cudaMemcpyDeviceToHost,
                             e->flag_stream));
    GOL_CUDA(cudaEventRecord(e->ev_slot[slot & 1], e->flag_stream));
    return 0;
}

int run_fast(gol_engine *e, gol_result *res)
{
    const int limit = e->cfg.gen_limit, freq = e->cfg.similarity_frequency;
    GOL_CUDA(cudaEventRecord(e->ev0, e->stream));
    // empty() of generation 0: launched, not waited for - batch 0 runs behind it
    {
        GolLaunch lp{};
        geometry(e, 1, &lp);
        lp.in_idx = e->cur;
        lp.out_idx = e->cur;
        lp.seq = e->seq;  // no ghost traffic: sequence unchanged
        lp.push = 0;
        const uint32_t rs = ++e->rec_seq;
        lp.rec_slot = (int)(rs % GOL_REC_SLOTS);
        lp.rec_seq = rs;
        GOL_TRY(gol_launch_scan_alive(e->d_doms, e->per_proc, lp, e->stream));
        ++e->launches;
        GOL_TRY(gather_last(e, 1));
        GOL_TRY(post_result(e, 1));
    }
    int generations = limit, reason = GOL_EXIT_LIMIT;
    int64_t computed = 0;
    const int64_t gen_start = e->cfg.start_generation;

    // Batches are pipelined one deep: while batch b runs, the host looks at the flags of batch
    // b-1 (or of generation 0).  Exits are fixed points, so the batch that was launched "too
    // many" never changes the result; it only costs its run time when an exit is found.
    struct Pending { bool valid; int slot; int64_t gen0; int n; int ckpt; } pend = {true, 1, -1, 0, -1};
    int64_t gen = gen_start;
    int prev_ckpt = -1;
    bool done = false;
    while (!done) {
        Pending cur_batch = {false, 0, 0, 0, -1};
        if (gen < limit) {
            NvtxRange nv("gol batch");
            const int n = (int)std::min<int64_t>(e->batch, limit - gen);
            const int ckpt = e->cur;
            const int slot = pend.slot ^ 1;
            GOL_TRY(run_epochs(e, n, ckpt, true, false, e->kmax, prev_ckpt));
            GOL_TRY(gather_last(e, slot));
            GOL_TRY(post_result(e, slot));
            computed += n;
            cur_batch = {true, slot, gen, n, ckpt};
        }
        if (pend.valid) {
            GOL_CUDA(cudaEventSynchronize(e->ev_slot[pend.slot & 1]));
            const unsigned long long r = e->h_results[pend.slot];
            const bool alive_end = r & 1ull, diff_end = r & 2ull;
            if (pend.gen0 < 0) {  // generation 0
                if (!alive_end) {
                    generations = (int)gen_start;
                    reason = GOL_EXIT_EMPTY;
                    done = true;
                }
            } else if (!alive_end || (freq > 0 && !diff_end)) {
                // replay that batch one generation at a time from its untouched input
                NvtxRange nv("gol exit replay");
                GOL_CUDA(cudaStreamSynchronize(e->stream));
                GOL_CUDA(cudaStreamSynchronize(e->flag_stream));  // the speculative batch's gather: the replay reuses record slots
                e->cur = pend.ckpt;
                GOL_TRY(run_epochs(e, pend.n, pend.ckpt, true, true, 1));
                GOL_TRY(fetch_results(e, pend.n));
                computed += pend.n;
                std::vector<int> alive((size_t)pend.n + 1), diff((size_t)pend.n + 1);
                alive[0] = 1;
                diff[0] = 1;
                for (int i = 1; i <= pend.n; ++i) {
                    alive[(size_t)i] = (int)(e->h_results[i - 1] & 1ull);
                    diff[(size_t)i] = (int)((e->h_results[i - 1] >> 1) & 1ull);
                }
                resolve_exit(pend.gen0, pend.n, alive, diff, limit, freq, &generations, &reason);
                done = true;
            }
        }
        if (!cur_batch.valid) done = true;
        if (cur_batch.valid) {
            prev_ckpt = cur_batch.ckpt;
            gen += cur_batch.n;
        }
        pend = cur_batch;
    }
    GOL_CUDA(cudaEventRecord(e->ev1, e->stream));
    GOL_CUDA(cudaEventSynchronize(e->ev1));
    {
        unsigned int err = 0;
        GOL_CUDA(cudaMemcpy(&err, e->d_error, sizeof err, cudaMemcpyDeviceToHost));
        if (err) GOL_FAIL("device-side wait timed out (code %u): a neighbouring rank stopped making progress", err);
        GOL_TRY(check_device_errors(e));
    }
    float ms = 0;
    GOL_CUDA(cudaEventElapsedTime(&ms, e->ev0, e->ev1));
    e->last_ms = ms;
    if (res) {
        memset(res, 0, sizeof *res);
        res->generations = generations;
        res->exit_reason = reason;
        res->generations_computed = computed;
        res->loop_ms = ms;
    }
    return 0;
}

// ------------------------------------------------------------ dense-path run
int dense_steps(gol_engine *e, int n, bool flags)
{
    const int64_t W = e->cfg.width, H = e->cfg.height;
    for (int i = 0; i < n; ++i) {
        const int out = (e->cur + 1) % 3 == 2 ? 0 : (e->cur + 1) % 3;  // ping-pong 0 <-> 1 (2 = checkpoint, unused)
        GOL_TRY(gol_op_step_dense(e->dense[e->cur], e->dense[out], W, H, flags ? e->d_dflags + 2 * (i + 1) : nullptr,
                                  e->stream));
        ++e->launches;
        e->cur = out;
    }
    return 0;
}

int run_dense(gol_engine *e, gol_result *res)
{
    const int limit = e->cfg.gen_limit, freq = e->cfg.similarity_frequency;
    const int64_t W = e->cfg.width, H = e->cfg.height;
    GOL_CUDA(cudaEventRecord(e->ev0, e->stream));
    int generations = limit, reason = GOL_EXIT_LIMIT;
    int64_t computed = 0, gen = e->cfg.start_generation;
    bool known_alive = false;
    while (gen < limit) {
        const int n = (int)std::min<int64_t>(std::min(e->batch, kMaxBatch), limit - gen);
        GOL_CUDA(cudaMemsetAsync(e->d_dflags, 0, sizeof(uint32_t) * 2 * (size_t)(n + 1), e->stream));
        if (!known_alive) {
            GOL_TRY(gol_op_flags_dense(e->dense[e->cur], nullptr, W, H, e->d_dflags, e->stream));
            ++e->launches;
        }
        GOL_TRY(dense_steps(e, n, true));
        GOL_CUDA(cudaMemcpyAsync(e->h_dflags, e->d_dflags, sizeof(uint32_t) * 2 * (size_t)(n + 1),
                                 cudaMemcpyDeviceToHost, e->stream));
        GOL_CUDA(cudaStreamSynchronize(e->stream));
        computed += n;
        std::vector<int> alive((size_t)n + 1), diff((size_t)n + 1);
        for (int i = 0; i <= n; ++i) {
            alive[(size_t)i] = (int)e->h_dflags[2 * i];
            diff[(size_t)i] = (int)e->h_dflags[2 * i + 1];
        }
        if (known_alive) alive[0] = 1;
        diff[0] = 1;
        const bool alive_end = alive[(size_t)n], diff_end = diff[(size_t)n];
        if (!alive[0] || !alive_end || (freq > 0 && !diff_end)) {
            resolve_exit(gen, n, alive, diff, limit, freq, &generations, &reason);
            break;
        }
        known_alive = true;
        gen += n;
    }
    GOL_CUDA(cudaEventRecord(e->ev1, e->stream));
    GOL_CUDA(cudaEventSynchronize(e->ev1));
    float ms = 0;
    GOL_CUDA(cudaEventElapsedTime(&ms, e->ev0, e->ev1));
    e->last_ms = ms;
    if (res) {
        memset(res, 0, sizeof *res);
        res->generations = generations;
        res->exit_reason = reason;
        res->generations_computed = computed;
        res->loop_ms = ms;
    }
    return 0;
}

}  // namespace

// =============================================================== C ABI
extern "C" {

int gol_cuda_available(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n > 0 ? 1 : 0;
}

void gol_config_default(gol_config *cfg, int64_t width, int64_t height)
{
    memset(cfg, 0, sizeof *cfg);
    cfg->width = width;
    cfg->height = height;
    cfg->grid_rows = 1;
    cfg->grid_cols = 1;
    cfg->proc_rank = 0;
    cfg->proc_count = 1;
    cfg->device = 0;
    cfg->gen_limit = GOL_DEFAULT_GEN_LIMIT;
    cfg->similarity_frequency = GOL_DEFAULT_SIMILARITY_FREQUENCY;
}

void gol_engine_destroy(gol_engine *e)
{
    if (!e) return;
    cudaSetDevice(e->cfg.device);
    if (e->stream) cudaStreamSynchronize(e->stream);
    if (const char *tf = getenv("GOL_TRACE_FILE")) {
        const std::string path = std::string(tf) + "." + std::to_string(e->rank);
        gol_engine_trace_dump(e, path.c_str());
    }
    for (int p = 0; p < (int)e->peer_arena.size(); ++p)
        if (p != e->rank && e->peer_arena[(size_t)p]) cudaIpcCloseMemHandle(e->peer_arena[(size_t)p]);
    if (e->arena) cudaFree(e->arena);
    if (e->d_doms) cudaFree(e->d_doms);
    if (e->d_rec_tables) cudaFree(e->d_rec_tables);
    if (e->d_results) cudaFree(e->d_results);
    if (e->h_results) cudaFreeHost(e->h_results);
    if (e->d_error) cudaFree(e->d_error);
    if (e->stage) cudaFree(e->stage);
    for (IoLane &L : e->io) {
        for (IoSlot &S : L.slot) {
            if (S.h) cudaFreeHost(S.h);
            if (S.d) cudaFree(S.d);
            if (S.ev) cudaEventDestroy(S.ev);
        }
        if (L.stream) cudaStreamDestroy(L.stream);
    }
    if (e->d_trace) cudaFree(e->d_trace);
    if (e->scratch_dense) cudaFree(e->scratch_dense);
    if (e->d_count) cudaFree(e->d_count);
    if (e->copy_stream) cudaStreamDestroy(e->copy_stream);
    if (e->ev_staged) cudaEventDestroy(e->ev_staged);
    if (e->ev_stage_free) cudaEventDestroy(e->ev_stage_free);
    for (auto &d : e->dense)
        if (d) cudaFree(d);
    if (e->d_dflags) cudaFree(e->d_dflags);
    if (e->h_dflags) cudaFreeHost(e->h_dflags);
    if (e->ev0) cudaEventDestroy(e->ev0);
    if (e->ev1) cudaEventDestroy(e->ev1);
    for (auto &ev : e->ev_slot)
        if (ev) cudaEventDestroy(ev);
    for (auto &ev : e->ev_tune)
        if (ev) cudaEventDestroy(ev);
    for (auto &ev : e->ev_flag)
        if (ev) cudaEventDestroy(ev);
    if (e->flag_stream) cudaStreamDestroy(e->flag_stream);
    if (e->own_stream && e->stream) cudaStreamDestroy(e->stream);
    delete e;
}

static int engine_create_impl(gol_engine *e)
{
    const gol_config &cfg = e->cfg;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) {
        cudaGetLastError();
        GOL_FAIL("no usable CUDA device (the engine has no CPU fallback; gol_cpu_run is the CPU path)");
    }
    if (cfg.device < 0 || cfg.device >= ndev)
        GOL_FAIL("CUDA device %d requested but %d device(s) are visible (one process per GPU: the process count "
                 "must not exceed the GPU count)", cfg.device, ndev);
    GOL_CUDA(cudaSetDevice(cfg.device));
    // two attributes instead of cudaGetDeviceProperties (which queries everything and is slow at start-up)
    struct { int multiProcessorCount = 148, clockRate = 1965000; } prop;
    GOL_CUDA(cudaDeviceGetAttribute(&prop.multiProcessorCount, cudaDevAttrMultiProcessorCount, cfg.device));
    GOL_CUDA(cudaDeviceGetAttribute(&prop.clockRate, cudaDevAttrClockRate, cfg.device));   // kHz
    e->num_sms = prop.multiProcessorCount;
    GOL_CUDA(cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking));
    e->own_stream = true;
    GOL_CUDA(cudaEventCreate(&e->ev0));
    GOL_CUDA(cudaEventCreate(&e->ev1));
    GOL_CUDA(cudaEventCreate(&e->ev_tune[0]));
    GOL_CUDA(cudaEventCreate(&e->ev_tune[1]));
    GOL_CUDA(cudaStreamCreateWithFlags(&e->flag_stream, cudaStreamNonBlocking));
    GOL_CUDA(cudaEventCreateWithFlags(&e->ev_flag[0], cudaEventDisableTiming));
    GOL_CUDA(cudaEventCreateWithFlags(&e->ev_flag[1], cudaEventDisableTiming));
    GOL_CUDA(cudaEventCreateWithFlags(&e->ev_slot[0], cudaEventDisableTiming));
    GOL_CUDA(cudaEventCreateWithFlags(&e->ev_slot[1], cudaEventDisableTiming));

    GOL_CUDA(cudaMalloc(&e->d_count, 8));
    GOL_CUDA(cudaMalloc(&e->d_error, 2 * sizeof(unsigned int)));   // [0] sticky wait-timeout word, [1] text parse errors
    GOL_CUDA(cudaMemset(e->d_error, 0, 2 * sizeof(unsigned int)));
    if (!e->fast) {
        const size_t bytes = (size_t)(gol_div_up(cfg.width, 64) * cfg.height) * sizeof(uint64_t);
        for (auto &d : e->dense) {
            GOL_CUDA(cudaMalloc(&d, bytes));
            GOL_CUDA(cudaMemset(d, 0, bytes));
        }
        GOL_CUDA(cudaMalloc(&e->d_dflags, sizeof(uint32_t) * 2 * (kMaxBatch + 2)));
        GOL_CUDA(cudaMallocHost(&e->h_dflags, sizeof(uint32_t) * 2 * (kMaxBatch + 2)));
        return 0;
    }

    e->local.resize((size_t)e->per_proc);
    e->max_strips = 1;
    e->max_h = 1;
    for (int i = 0; i < e->per_proc; ++i) {
        LocalDomain &ld = e->local[(size_t)i];
        ld.gid = e->rank * e->per_proc + i;
        ld.r = ld.gid / e->C;
        ld.c = ld.gid % e->C;
        int64_t w, h;
        domain_extent(e, ld.gid, &ld.x0, &ld.y0, &w, &h);
        GOL_TRY(gol_layout_init(&ld.L, w, h));
        GOL_TRY(gol_strip_geometry(ld.L.nw32, 0, &ld.strips));
    }
    // launch geometry is uniform over all domains of the job (every process computes the same)
    bool aligned = true;
    for (int g = 0; g < e->ndom; ++g) {
        int64_t x0, y0, w, h;
        domain_extent(e, g, &x0, &y0, &w, &h);
        int s, sx;
        GOL_TRY(gol_strip_geometry(2 * (int)gol_div_up(w, 64), 0, &s));
        GOL_TRY(gol_strip_geometry(2 * (int)gol_div_up(w, 64), 2, &sx));
        e->max_strips = std::max(e->max_strips, s);
        e->max_strips_ext = std::max(e->max_strips_ext, sx);
        e->max_h = std::max(e->max_h, h);
        if (w % 64) aligned = false;
    }
    if (const char *v = getenv("GOL_AUTOTUNE")) e->autotune = atoi(v) != 0;
    // Deep-halo pairs: on by default between processes (GOL_DEEP_HALO=0 disables, =1 also enables them
    // inside one process).  A pair's first epoch computes 64 ghost cells per side as whole words, so
    // every domain width must be a multiple of 64, and two epochs must fit the 16-row ghost zone.
    {
        const char *v = getenv("GOL_DEEP_HALO");
        const int want = v ? atoi(v) : -1;
        e->deep_halo = aligned && 2 * e->kmax <= G && (want > 0 || (want < 0 && multi(e)));
    }
    if (const char *v = getenv("GOL_PDL")) e->pdl_mode = atoi(v);
    {
        // device-side flag waits: a rank that waits longer than this for a neighbour gives up with an error
        const char *v = getenv("GOL_WAIT_TIMEOUT_MS");
        const double ms = v ? atof(v) : 20000.0;
        e->wait_timeout = (long long)(std::max(1.0, ms) * (double)prop.clockRate);  // clockRate is in kHz = cycles per ms
    }
    if (const char *v = getenv("GOL_FAULT_INJECT")) e->fault_wait = strcmp(v, "wait") == 0;
    if (const char *v = getenv("GOL_TRACE")) {
        e->trace_cap = std::max(0, atoi(v));
        if (e->trace_cap) {
            GOL_CUDA(cudaMalloc(&e->d_trace, sizeof(GolTrace) * (size_t)e->trace_cap));
            std::vector<GolTrace> init((size_t)e->trace_cap, GolTrace{~0ull, 0ull, 0ull, 0ull});
            GOL_CUDA(cudaMemcpy(e->d_trace, init.data(), sizeof(GolTrace) * init.size(), cudaMemcpyHostToDevice));
        }
    }
    if (getenv("GOL_VERBOSE")) e->cfg.flags |= GOL_FLAG_VERBOSE;
    GOL_TRY(build_arena_layout(e, e->rank, &e->arena_layout));
    GOL_CUDA(cudaMalloc(&e->arena, (size_t)e->arena_layout.total));
    GOL_CUDA(cudaMemset(e->arena, 0, (size_t)e->arena_layout.total));
    e->peer_arena.assign((size_t)e->P, nullptr);
    e->peer_arena[(size_t)e->rank] = e->arena;
    GOL_CUDA(cudaMalloc(&e->d_doms, sizeof(GolDomainDev) * (size_t)e->per_proc));
    GOL_CUDA(cudaMalloc(&e->d_rec_tables, sizeof(void *) * (size_t)e->P));
    GOL_CUDA(cudaMalloc(&e->d_results, sizeof(unsigned long long) * (kMaxBatch + 2)));
    GOL_CUDA(cudaMemset(e->d_results, 0, sizeof(unsigned long long) * (kMaxBatch + 2)));
    GOL_CUDA(cudaMallocHost(&e->h_results, sizeof(unsigned long long) * (kMaxBatch + 2)));
    {
        size_t stage_bytes = 0;
        for (const LocalDomain &ld : e->local) {
            e->stage_off.push_back(stage_bytes / 8);
            stage_bytes += (size_t)ld.L.nq * (size_t)ld.L.height * 8;
        }
        GOL_CUDA(cudaMalloc(&e->stage, stage_bytes));
        size_t scratch_bytes = 8;
        for (const LocalDomain &ld : e->local) scratch_bytes = std::max(scratch_bytes, (size_t)ld.L.nq * (size_t)ld.L.height * 8);
        GOL_CUDA(cudaMalloc(&e->scratch_dense, scratch_bytes));
        GOL_CUDA(cudaStreamCreateWithFlags(&e->copy_stream, cudaStreamNonBlocking));
        GOL_CUDA(cudaEventCreateWithFlags(&e->ev_staged, cudaEventDisableTiming));
        GOL_CUDA(cudaEventCreateWithFlags(&e->ev_stage_free, cudaEventDisableTiming));
    }
    if (e->P == 1) {
        GOL_TRY(upload_domains(e));
        e->imported = true;
    }
    return 0;
}

int gol_engine_create(const gol_config *cfg_in, gol_engine **out)
{
    if (!cfg_in || !out) GOL_FAIL("gol_engine_create: null argument");
    gol_config cfg = *cfg_in;
    if (cfg.width < 1 || cfg.height < 1)
        GOL_FAIL("grid dimensions must be positive (got %lld x %lld)", (long long)cfg.width, (long long)cfg.height);
    if (cfg.grid_rows < 1) cfg.grid_rows = 1;
    if (cfg.grid_cols < 1) cfg.grid_cols = 1;
    if (cfg.proc_count < 1) cfg.proc_count = 1;
    if (cfg.gen_limit <= 0) cfg.gen_limit = GOL_DEFAULT_GEN_LIMIT;
    if (cfg.similarity_frequency < 0) cfg.similarity_frequency = GOL_DEFAULT_SIMILARITY_FREQUENCY;
    if (cfg.start_generation < 0) cfg.start_generation = 0;
    if (cfg.start_generation > cfg.gen_limit) cfg.start_generation = cfg.gen_limit;
    const int ndom = cfg.grid_rows * cfg.grid_cols;
    if (ndom % cfg.proc_count) GOL_FAIL("%d domains cannot be dealt evenly to %d processes", ndom, cfg.proc_count);
    if (cfg.proc_rank < 0 || cfg.proc_rank >= cfg.proc_count) GOL_FAIL("process rank %d out of range", cfg.proc_rank);

    gol_engine *e = new gol_engine();
    e->cfg = cfg;
    e->R = cfg.grid_rows;
    e->C = cfg.grid_cols;
    e->ndom = ndom;
    e->P = cfg.proc_count;
    e->rank = cfg.proc_rank;
    e->per_proc = ndom / cfg.proc_count;
    // columns split at 64-cell boundaries, the remainder goes to the last column; rows evenly
    const int64_t nq = gol_div_up(cfg.width, 64);
    e->col_x.resize((size_t)e->C + 1);
    for (int c = 0; c <= e->C; ++c) e->col_x[(size_t)c] = std::min<int64_t>(cfg.width, (nq * c / e->C) * 64);
    e->col_x[(size_t)e->C] = cfg.width;
    e->row_y.resize((size_t)e->R + 1);
    for (int r = 0; r <= e->R; ++r) e->row_y[(size_t)r] = cfg.height * r / e->R;

    bool fast = !(cfg.flags & GOL_FLAG_FORCE_GENERIC);
    for (int g = 0; g < ndom && fast; ++g) {
        int64_t x0, y0, w, h;
        domain_extent(e, g, &x0, &y0, &w, &h);
        if (w < GOL_MIN_FAST_WIDTH || h < GOL_MIN_FAST_HEIGHT) fast = false;
    }
    if (!fast && ndom > 1) {
        delete e;
        GOL_FAIL("grid %lld x %lld is too small for a %d x %d decomposition (every domain needs >= %d x %d cells)",
                 (long long)cfg.width, (long long)cfg.height, cfg.grid_rows, cfg.grid_cols, GOL_MIN_FAST_WIDTH,
                 GOL_MIN_FAST_HEIGHT);
    }
    e->fast = fast;
    e->kmax = cfg.k_max > 0 ? cfg.k_max : 8;
    if (e->kmax > GOL_GHOST_ROWS) e->kmax = GOL_GHOST_ROWS;
    e->kmax = pick_depth(e->kmax);
    e->batch = cfg.batch_gens > 0 ? cfg.batch_gens : 128;
    if (e->batch > kMaxBatch) e->batch = kMaxBatch;

    if (engine_create_impl(e)) {
        gol_engine_destroy(e);
        return 1;
    }
    *out = e;
    return 0;
}

int gol_engine_set_stream(gol_engine *e, void *stream)
{
    GOL_CUDA(cudaStreamSynchronize(e->stream));
    if (e->own_stream) cudaStreamDestroy(e->stream);
    e->own_stream = false;
    e->stream = (cudaStream_t)stream;
    return 0;
}

int gol_engine_num_domains(const gol_engine *e) { return e->ndom; }
int gol_engine_num_local_domains(const gol_engine *e) { return e->per_proc; }
int gol_engine_local_domain_id(const gol_engine *e, int i) { return e->rank * e->per_proc + i; }
int gol_engine_uses_fast_path(const gol_engine *e) { return e->fast ? 1 : 0; }

int gol_engine_domain_extent(const gol_engine *e, int gid, int64_t *x0, int64_t *y0, int64_t *w, int64_t *h)
{
    if (gid < 0 || gid >= e->ndom) GOL_FAIL("domain id %d out of range", gid);
    domain_extent(e, gid, x0, y0, w, h);
    return 0;
}

int gol_engine_domain_neighbors(const gol_engine *e, int gid, int32_t nb[8])
{
    if (gid < 0 || gid >= e->ndom) GOL_FAIL("domain id %d out of range", gid);
    for (int d = 0; d < GOL_NDIR; ++d) nb[d] = neighbor_gid(e, gid, d);
    return 0;
}

size_t gol_engine_ipc_blob_size(const gol_engine *) { return sizeof(IpcBlob); }

int gol_engine_ipc_export(gol_engine *e, void *blob)
{
    if (!e->fast) GOL_FAIL("gol_engine_ipc_export: engine is on the single-domain generic path");
    GOL_CUDA(cudaSetDevice(e->cfg.device));
    IpcBlob b;
    memset(&b, 0, sizeof b);
    GOL_CUDA(cudaIpcGetMemHandle(&b.handle, e->arena));
    b.device = e->cfg.device;
    b.pid = (int32_t)getpid();
    b.arena_bytes = e->arena_layout.total;
    memcpy(blob, &b, sizeof b);
    return 0;
}

int gol_engine_ipc_import(gol_engine *e, const void *all_blobs)
{
    if (!e->fast) GOL_FAIL("gol_engine_ipc_import: engine is on the single-domain generic path");
    GOL_CUDA(cudaSetDevice(e->cfg.device));
    const IpcBlob *blobs = static_cast<const IpcBlob *>(all_blobs);
    for (int p = 0; p < e->P; ++p) {
        if (p == e->rank) continue;
        IpcBlob b;
        memcpy(&b, blobs + p, sizeof b);
        ArenaLayout A;
        GOL_TRY(build_arena_layout(e, p, &A));
        if (A.total != b.arena_bytes)
            GOL_FAIL("process %d reports a %lld-byte arena, expected %lld: configurations differ", p,
                     (long long)b.arena_bytes, (long long)A.total);
        void *ptr = nullptr;
        GOL_CUDA(cudaIpcOpenMemHandle(&ptr, b.handle, cudaIpcMemLazyEnablePeerAccess));
        e->peer_arena[(size_t)p] = static_cast<char *>(ptr);
    }
    GOL_TRY(upload_domains(e));
    e->imported = true;
    return 0;
}

// -------------------------------------------------------------- load / store
static int clear_current(gol_engine *e)
{
    e->prepared = false;
    if (!e->fast) {
        e->cur = 0;
        return 0;
    }
    return 0;
}

int gol_engine_load_packed(gol_engine *e, const uint64_t *host)
{
    GOL_CUDA(cudaSetDevice(e->cfg.device));
    GOL_TRY(clear_current(e));
    const int64_t nqg = gol_div_up(e->cfg.width, 64);
    if (!e->fast) {
        GOL_CUDA(cudaMemcpyAsync(e->dense[e->cur], host, (size_t)(nqg * e->cfg.height) * 8, cudaMemcpyHostToDevice,
                                 e->stream));
        GOL_CUDA(cudaStreamSynchronize(e->stream));
        return 0;
    }
    for (int i = 0; i < e->per_proc; ++i) {
        const LocalDomain &ld = e->local[(size_t)i];
        char *dst = e->arena + e->arena_layout.buf_off[(size_t)i * GOL_NBUF + e->cur];
        GOL_CUDA(cudaMemsetAsync(dst, 0, (size_t)ld.L.buffer_words32 * 4, e->stream));
        if (ld.L.nq == nqg && e->stage) {
            // full-width domain: its rows are contiguous in the host grid - one linear DMA into the
            // staging buffer (a strided 2-D copy of 2 KiB rows runs far below PCIe speed), then a
            // kernel spreads the rows into the padded layout
            GOL_CUDA(cudaMemcpyAsync(e->stage + e->stage_off[(size_t)i], host + ld.y0 * nqg,
                                     (size_t)ld.L.nq * (size_t)ld.L.height * 8, cudaMemcpyHostToDevice, e->stream));
            GOL_TRY(gol_launch_dense_to_padded(e->stage + e->stage_off[(size_t)i], ld.L.nq, 0,
                                               reinterpret_cast<uint32_t *>(dst), ld.L, e->stream));
            ++e->launches;
        } else {
            GOL_CUDA(cudaMemcpy2DAsync(dst + ((int64_t)G * ld.L.pitch32 + GOL_GHOST_WORDS32) * 4, (size_t)ld.L.pitch32 * 4,
                                       host + ld.y0 * nqg + ld.x0 / 64, (size_t)nqg * 8, (size_t)ld.L.nq * 8,
                                       (size_t)ld.L.height, cudaMemcpyHostToDevice, e->stream));
        }
    }
    GOL_CUDA(cudaStreamSynchronize(e->stream));
    return 0;
}

// Overlapped upload for a stream of grids: stage() starts the host->device copy of the NEXT grid
// on a second stream into the staging buffer and returns at once (the host memory must be pinned
// for the copy to be asynchronous and must stay untouched until commit); commit() makes the staged
// grid the current one (a kernel spreads it into the padded layout).  While the engine runs grid i,
// grid i+1 is already crossing PCIe.
int gol_engine_stage_packed(gol_engine *e, const uint64_t *host)
{
    GOL_CUDA(cudaSetDevice(e->cfg.device));
    if (!e->fast) GOL_FAIL("gol_engine_stage_packed: only for engines on the strip path (use gol_engine_load_packed)");
    const int64_t nqg = gol_div_up(e->cfg.width, 64);
    if (e->stage_busy) GOL_CUDA(cudaStreamWaitEvent(e->copy_stream, e->ev_stage_free, 0));  // last commit still reads it
    for (int i = 0; i < e->per_proc; ++i) {
        const LocalDomain &ld = e->local[(size_t)i];
        uint64_t *dst = e->stage + e->stage_off[(size_t)i];
        if (ld.L.nq == nqg)
            GOL_CUDA(cudaMemcpyAsync(dst, host + ld.y0 * nqg, (size_t)ld.L.nq * (size_t)ld.L.height * 8,
                                     cudaMemcpyHostToDevice, e->copy_stream));
        else
            GOL_CUDA(cudaMemcpy2DAsync(dst, (size_t)ld.L.nq * 8, host + ld.y0 * nqg + ld.x0 / 64, (size_t)nqg * 8,
                                       (size_t)ld.L.nq * 8, (size_t)ld.L.height, cudaMemcpyHostToDevice, e->copy_stream));
    }
    GOL_CUDA(cudaEventRecord(e->ev_staged, e->copy_stream));
    e->staged = true;
    return 0;
}

// One local domain of the next grid from the caller's own dense block [h][ceil(w/64)] (pinned memory
// for an asynchronous copy): a rank keeps only its own blocks on the host.  Stage every local domain,
// then commit.
int gol_engine_stage_domain(gol_engine *e, int local_index, const uint64_t *host_dense)
{
    GOL_CUDA(cudaSetDevice(e->cfg.device));
    if (!e->fast) GOL_FAIL("gol_engine_stage_domain: only for engines on the strip path");
    if (local_index < 0 || local_index >= e->per_proc) GOL_FAIL("domain index %d out of range", local_index);
    if (e->stage_busy) GOL_CUDA(cudaStreamWaitEvent(e->copy_stream, e->ev_stage_free, 0));  // last commit still reads it
    const LocalDomain &ld = e->local[(size_t)local_index];
    GOL_CUDA(cudaMemcpyAsync(e->stage + e->stage_off[(size_t)local_index], host_dense,
                             (size_t)ld.L.nq * (size_t)ld.L.height * 8, cudaMemcpyHostToDevice, e->copy_stream));
    GOL_CUDA(cudaEventRecord(e->ev_staged, e->copy_stream));
    e->staged = true;
    return 0;
}

int gol_engine_commit_staged(gol_engine *e)
{
    GOL_CUDA(cudaSetDevice(e->cfg.device));
    if (!e->staged) GOL_FAIL("gol_engine_commit_staged: nothing staged");
    GOL_TRY(clear_current(e));
    GOL_CUDA(cudaStreamWaitEvent(e->stream, e->ev_staged, 0));
    for (int i = 0; i < e->per_proc; ++i) {
        const LocalDomain &ld = e->local[(size_t)i];
        char *dst = e->arena + e->arena_layout.buf_off[(size_t)i * GOL_NBUF + e->cur];
        GOL_CUDA(cudaMemsetAsync(dst, 0, (size_t)ld.L.buffer_words32 * 4, e->stream));
        GOL_TRY(gol_launch_dense_to_padded(e->stage + e->stage_off[(size_t)i], ld.L.nq, 0, reinterpret_cast<uint32_t *>(dst),
                                           ld.L, e->stream));
        ++e->launches;
    }
    GOL_CUDA(cudaEventRecord(e->ev_stage_free, e->stream));
    e->stage_busy = true;
    e->staged = false;
    return 0;
}

int gol_engine_load_packed_device(gol_engine *e, const uint64_t *d_packed, int64_t x0, int64_t y0, int64_t w,
                                  int64_t h)
{
    GOL_CUDA(cudaSetDevice(e->cfg.device));
    GOL_TRY(clear_current(e));
    const int64_t nq = gol_div_up(w, 64);
    if (!e->fast) {
        if (x0 || y0 || w != e->cfg.width || h != e->cfg.height) GOL_FAIL("load_packed_device: block must be the whole grid");
        GOL_CUDA(cudaMemcpyAsync(e->dense[e->cur], d_packed, (size_t)(nq * h) * 8, cudaMemcpyDeviceToDevice, e->stream));
        return 0;
    }
    for (int i = 0; i < e->per_proc; ++i) {
        const LocalDomain &ld = e->local[(size_t)i];
        if (ld.x0 != x0 || ld.y0 != y0) continue;
        if (ld.L.width != w || ld.L.height != h) GOL_FAIL("load_packed_device: block does not match domain %d", ld.gid);
        char *dst = e->arena + e->arena_layout.buf_off[(size_t)i * GOL_NBUF + e->cur];
        GOL_CUDA(cudaMemsetAsync(dst, 0, (size_t)ld.L.buffer_words32 * 4, e->stream));
        GOL_TRY(gol_launch_dense_to_padded(d_packed, nq, 0, reinterpret_cast<uint32_t *>(dst), ld.L, e->stream));
        ++e->launches;
        return 0;
    }
    GOL_FAIL("load_packed_device: no local domain starts at (%lld, %lld)", (long long)x0, (long long)y0);
}

int gol_engine_randomize(gol_engine *e, uint64_t seed)
{
    GOL_CUDA(cudaSetDevice(e->cfg.device));
    GOL_TRY(clear_current(e));
    if (!e->fast) {
        GOL_TRY(gol_op_random_dense(e->dense[e->cur], e->cfg.width, e->cfg.height, 0, 0, seed, e->stream));
        ++e->launches;
        return 0;
    }
    for (int i = 0; i < e->per_proc; ++i) {
        // the engine-owned dense scratch is reused by every local domain: stream order keeps them apart
        const LocalDomain &ld = e->local[(size_t)i];
        uint64_t *tmp = e->scratch_dense;
        GOL_TRY(gol_op_random_dense(tmp, ld.L.width, ld.L.height, ld.x0, ld.y0, seed, e->stream));
        char *dst = e->arena + e->arena_layout.buf_off[(size_t)i * GOL_NBUF + e->cur];
        GOL_CUDA(cudaMemsetAsync(dst, 0, (size_t)ld.L.buffer_words32 * 4, e->stream));
        GOL_TRY(gol_launch_dense_to_padded(tmp, ld.L.nq, 0, reinterpret_cast<uint32_t *>(dst), ld.L, e->stream));
        e->launches += 2;
    }
    GOL_CUDA(cudaStreamSynchronize(e->stream));
    return 0;
}

int gol_engine_download_packed(gol_engine *e, uint64_t *host)
{
    GOL_CUDA(cudaSetDevice(e->cfg.device));
    const int64_t nqg = gol_div_up(e->cfg.width, 64);
    if (!e->fast) {
        GOL_CUDA(cudaMemcpyAsync(host, e->dense[e->cur], (size_t)(nqg * e->cfg.height) * 8, cudaMemcpyDeviceToHost,
                                 e->stream));
        GOL_CUDA(cudaStreamSynchronize(e->stream));
        return 0;
    }
    for (int i = 0; i < e->per_proc; ++i) {
        const LocalDomain &ld = e->local[(size_t)i];
        const char *src = e->arena + e->arena_layout.buf_off[(size_t)i * GOL_NBUF + e->cur];
        GOL_CUDA(cudaMemcpy2DAsync(host + ld.y0 * nqg + ld.x0 / 64, (size_t)nqg * 8,
                                   src + ((int64_t)G * ld.L.pitch32 + GOL_GHOST_WORDS32) * 4, (size_t)ld.L.pitch32 * 4,
                                   (size_t)ld.L.nq * 8, (size_t)ld.L.height, cudaMemcpyDeviceToHost, e->stream));
    }
    GOL_CUDA(cudaStreamSynchronize(e->stream));
    return 0;
}

// One local domain of the current grid into the caller's dense block [h][ceil(w/64)]: a kernel
// gathers the padded rows into the engine's dense scratch, one linear copy brings them to the host.
int gol_engine_download_domain(gol_engine *e, int local_index, uint64_t *host_dense)
{
    GOL_CUDA(cudaSetDevice(e->cfg.device));
    if (!e->fast) {
        if (local_index != 0) GOL_FAIL("domain index %d out of range", local_index);
        GOL_CUDA(cudaMemcpyAsync(host_dense, e->dense[e->cur], (size_t)(gol_div_up(e->cfg.width, 64) * e->cfg.height) * 8,
                                 cudaMemcpyDeviceToHost, e->stream));
        GOL_CUDA(cudaStreamSynchronize(e->stream));
        return 0;
    }
    if (local_index < 0 || local_index >= e->per_proc) GOL_FAIL("domain index %d out of range", local_index);
    const LocalDomain &ld = e->local[(size_t)local_index];
    const char *src = e->arena + e->arena_layout.buf_off[(size_t)local_index * GOL_NBUF + e->cur];
    GOL_TRY(gol_launch_padded_to_dense(reinterpret_cast<const uint32_t *>(src), ld.L, e->scratch_dense, ld.L.nq, 0, 0, e->stream));
    ++e->launches;
    GOL_CUDA(cudaMemcpyAsync(host_dense, e->scratch_dense, (size_t)ld.L.nq * (size_t)ld.L.height * 8, cudaMemcpyDeviceToHost,
                             e->stream));
    GOL_CUDA(cudaStreamSynchronize(e->stream));
    return 0;
}

// Checksum of the local part of the current grid (see hash_padded_kernel): add the values of all
// processes (mod 2^64) to get a number that identifies the global grid whatever the decomposition.
int gol_engine_hash(gol_engine *e, uint64_t *hash)
{
    GOL_CUDA(cudaSetDevice(e->cfg.device));
    GOL_CUDA(cudaMemsetAsync(e->d_count, 0, 8, e->stream));
    if (!e->fast) {
        gol_layout L;
        GOL_TRY(gol_layout_init(&L, e->cfg.width, e->cfg.height));
        // dense buffer = a padded layout without ghosts: same kernel, shifted so that its offsets vanish
        L.pitch32 = L.nw32;
        GOL_TRY(gol_launch_hash_padded(reinterpret_cast<const uint32_t *>(e->dense[e->cur]) - ((int64_t)G * L.pitch32 + GOL_GHOST_WORDS32),
                                       L, 0, 0, e->d_count, e->stream));
        ++e->launches;
    } else {
        for (int i = 0; i < e->per_proc; ++i) {
            const LocalDomain &ld = e->local[(size_t)i];
            const char *src = e->arena + e->arena_layout.buf_off[(size_t)i * GOL_NBUF + e->cur];
            GOL_TRY(gol_launch_hash_padded(reinterpret_cast<const uint32_t *>(src), ld.L, ld.x0 / 64, ld.y0, e->d_count, e->stream));
            ++e->launches;
        }
    }
    unsigned long long c = 0;
    GOL_CUDA(cudaMemcpyAsync(&c, e->d_count, 8, cudaMemcpyDeviceToHost, e->stream));
    GOL_CUDA(cudaStreamSynchronize(e->stream));
    *hash = c;
    return 0;
}

// ------------------------------------------------------------ text file I/O
// Pipelined staging for the text reader and writer.  A few worker threads ("lanes") split the row
// blocks of a domain between them; each lane owns a CUDA stream and two pinned host / device slot
// pairs, so that on every lane the file system call of one block (pread / pwrite, a CPU-speed copy
// out of or into the page cache) overlaps the PCIe copy and the parse / format kernel of the next
// or previous block, and the lanes run their system calls in parallel.  The slots belong to the
// engine and are allocated on first use (replaces the non-blocking and collective MPI-IO of
// A:186-198, A:428-448, C:190-198, C:425-445: there the wait follows the request immediately).
}  // extern "C"

namespace {

int io_prepare(gol_engine *e)
{
    if (!e->io.empty()) return 0;
    int lanes = 4;
    if (const char *v = getenv("GOL_IO_THREADS")) lanes = std::max(1, std::min(16, atoi(v)));
    e->io_slot_bytes = 8u << 20;
    if (const char *v = getenv("GOL_IO_SLOT_MB")) e->io_slot_bytes = (size_t)std::max(1, atoi(v)) << 20;
    // a slot holds at least one full line of the file
    e->io_slot_bytes = std::max<size_t>(e->io_slot_bytes, (size_t)e->cfg.width + 64);
    e->io.resize((size_t)lanes);
    for (IoLane &L : e->io) {
        GOL_CUDA(cudaStreamCreateWithFlags(&L.stream, cudaStreamNonBlocking));
        for (IoSlot &S : L.slot) {
            GOL_CUDA(cudaMallocHost(&S.h, e->io_slot_bytes));
            GOL_CUDA(cudaMalloc(&S.d, e->io_slot_bytes));
            GOL_CUDA(cudaEventCreateWithFlags(&S.ev, cudaEventDisableTiming));
        }
    }
    return 0;
}

// read exactly `want` bytes at `pos`
bool pread_all(int fd, char *dst, int64_t want, int64_t pos, std::string *err, const char *path)
{
    int64_t off = 0;
    while (off < want) {
        const ssize_t got = pread(fd, dst + off, (size_t)(want - off), (off_t)(pos + off));
        if (got <= 0) {
            if (got < 0 && errno == EINTR) continue;
            *err = std::string("read: ") + path + ": " + (got < 0 ? strerror(errno) : "unexpected end of file");
            return false;
        }
        off += got;
    }
    return true;
}

struct IoJob {
    gol_engine *e;
    int fd;
    const char *path;
    int64_t file_size, W;            // global line length is W + 1
    int64_t x0, y0, w, h;            // the block of the grid this job moves
    uint64_t *d_dense;               // packed dense rows of the block on the device
    uint32_t *d_err;                 // parse errors (reader)
    std::mutex mu;
    std::string error;
    std::atomic<int> failed{0};
    void fail(const std::string &msg)
    {
        std::lock_guard<std::mutex> g(mu);
        if (!failed.exchange(1)) error = msg;
    }
};

void io_read_lane(IoJob *J, int lane_id, int lanes)
{
    gol_engine *e = J->e;
    IoLane &L = e->io[(size_t)lane_id];
    if (cudaSetDevice(e->cfg.device) != cudaSuccess) return J->fail("cudaSetDevice failed in an I/O thread");
    const bool last_col = J->x0 + J->w == J->W, whole_rows = J->x0 == 0 && last_col;
    const int64_t seg = J->w + (last_col ? 1 : 0);  // bytes per row of this block (newline owned by the last column)
    const int64_t rows_per = std::max<int64_t>(1, (int64_t)e->io_slot_bytes / seg);
    const int64_t chunks = gol_div_up(J->h, rows_per), nq = gol_div_up(J->w, 64);
    int use = 0;
    for (int64_t c = lane_id; c < chunks && !J->failed; c += lanes, use ^= 1) {
        IoSlot &S = L.slot[use];
        if (cudaEventSynchronize(S.ev) != cudaSuccess) return J->fail("I/O slot event failed");  // its last H2D + parse are done
        const int64_t r = c * rows_per, n = std::min(rows_per, J->h - r);
        std::string err;
        if (whole_rows) {  // one contiguous range of the file
            int64_t want = n * seg;
            const int64_t pos = (J->y0 + r) * (J->W + 1);
            if (pos + want > J->file_size) {  // no newline after the very last line
                want = J->file_size - pos;
                S.h[n * seg - 1] = '\n';
            }
            if (!pread_all(J->fd, S.h, want, pos, &err, J->path)) return J->fail(err);
        } else {
            for (int64_t k = 0; k < n; ++k) {
                int64_t want = seg;
                const int64_t pos = (J->y0 + r + k) * (J->W + 1) + J->x0;
                if (pos + want > J->file_size) { want = J->file_size - pos; S.h[k * seg + seg - 1] = '\n'; }
                if (!pread_all(J->fd, S.h + k * seg, want, pos, &err, J->path)) return J->fail(err);
            }
        }
        if (cudaMemcpyAsync(S.d, S.h, (size_t)(n * seg), cudaMemcpyHostToDevice, L.stream) != cudaSuccess)
            return J->fail("H2D copy of a text block failed");
        if (gol_op_parse_text_n(S.d, n * seg, J->d_dense + r * nq, J->w, n, seg, J->d_err, L.stream)) return J->fail(gol_last_error());
        cudaEventRecord(S.ev, L.stream);
    }
    if (cudaStreamSynchronize(L.stream) != cudaSuccess) J->fail("parse kernel failed");
}

void io_write_lane(IoJob *J, int lane_id, int lanes)
{
    gol_engine *e = J->e;
    IoLane &L = e->io[(size_t)lane_id];
    if (cudaSetDevice(e->cfg.device) != cudaSuccess) return J->fail("cudaSetDevice failed in an I/O thread");
    const bool last_col = J->x0 + J->w == J->W, whole_rows = J->x0 == 0 && last_col;
    const int64_t seg = J->w + (last_col ? 1 : 0);
    const int64_t rows_per = std::max<int64_t>(1, (int64_t)e->io_slot_bytes / seg);
    const int64_t chunks = gol_div_up(J->h, rows_per), nq = gol_div_up(J->w, 64);
    auto issue = [&](int64_t c, IoSlot &S) -> bool {  // format block c on the device and start bringing it home
        const int64_t r = c * rows_per, n = std::min(rows_per, J->h - r);
        if (gol_op_format_text(J->d_dense + r * nq, S.d, J->w, n, seg, last_col ? 1 : 0, L.stream)) { J->fail(gol_last_error()); return false; }
        if (cudaMemcpyAsync(S.h, S.d, (size_t)(n * seg), cudaMemcpyDeviceToHost, L.stream) != cudaSuccess) { J->fail("D2H copy of a text block failed"); return false; }
        cudaEventRecord(S.ev, L.stream);
        return true;
    };
    int use = 0;
    if (lane_id < chunks && !issue(lane_id, L.slot[0])) return;
    for (int64_t c = lane_id; c < chunks && !J->failed; c += lanes, use ^= 1) {
        if (c + lanes < chunks && !issue(c + lanes, L.slot[use ^ 1])) return;  // next block formats while this one is written
        IoSlot &S = L.slot[use];
        if (cudaEventSynchronize(S.ev) != cudaSuccess) return J->fail("format kernel / D2H copy failed");
        const int64_t r = c * rows_per, n = std::min(rows_per, J->h - r);
        if (whole_rows) {
            if (gol_write_all(J->fd, S.h, (size_t)(n * seg), (J->y0 + r) * (J->W + 1), J->path)) return J->fail(gol_last_error());
        } else {
            for (int64_t k = 0; k < n; ++k)
                if (gol_write_all(J->fd, S.h + k * seg, (size_t)seg, (J->y0 + r + k) * (J->W + 1) + J->x0, J->path))
                    return J->fail(gol_last_error());
        }
    }
}

int io_run(IoJob *J, bool reading)
{
    const int lanes = (int)J->e->io.size();
    std::vector<std::thread> pool;
    for (int t = 1; t < lanes; ++t) pool.emplace_back(reading ? io_read_lane : io_write_lane, J, t, lanes);
    (reading ? io_read_lane : io_write_lane)(J, 0, lanes);
    for (auto &th : pool) th.join();
    if (J->failed) GOL_FAIL("%s", J->error.c_str());
    return 0;
}

}  // namespace

extern "C" {

// Parallel text read: every process reads only its own byte ranges.  The file must be in
// the fixed-record layout (height lines of exactly width cells + '\n'), which is what
// generate.sh and every writer of the reference produce; other layouts go through
// gol_file_read_packed (host filter with the fgetc semantics) + gol_engine_load_packed.
int gol_engine_load_text_file(gol_engine *e, const char *path, int strict)
{
    NvtxRange nv("gol load_text_file");
    GOL_CUDA(cudaSetDevice(e->cfg.device));
    const int64_t W = e->cfg.width, H = e->cfg.height;
    int fd = open(path, O_RDONLY);
    if (fd < 0) GOL_FAIL("fopen: %s: %s", path, strerror(errno));
    struct stat st;
    if (fstat(fd, &st) != 0) {
        close(fd);
        GOL_FAIL("fstat: %s: %s", path, strerror(errno));
    }
    const bool fixed = st.st_size == H * (W + 1) || st.st_size == H * (W + 1) - 1;
    if (!fixed) {
        close(fd);
        if ((int64_t)st.st_size < W * H)
            GOL_FAIL("input is too short: %s has %lld bytes, a %lld x %lld grid needs at least %lld", path,
                     (long long)st.st_size, (long long)W, (long long)H, (long long)(W * H));
        if ((int64_t)st.st_size > (1ll << 32))
            GOL_FAIL("%s is not in fixed-record layout (%lld bytes, expected %lld) and too large for the host filter",
                     path, (long long)st.st_size, (long long)(H * (W + 1)));
        std::vector<uint64_t> packed((size_t)(gol_div_up(W, 64) * H));
        GOL_TRY(gol_file_read_packed(path, W, H, strict, packed.data()));
        return gol_engine_load_packed(e, packed.data());
    }
    GOL_TRY(clear_current(e));
    int rc = io_prepare(e);
    uint32_t *d_err = e->d_error + 1;
    if (!rc && cudaMemsetAsync(d_err, 0, 4, e->stream) != cudaSuccess) rc = 1;
    if (!rc && cudaStreamSynchronize(e->stream) != cudaSuccess) rc = 1;
    const int ndom_local = e->fast ? e->per_proc : 1;
    for (int i = 0; i < ndom_local && !rc; ++i) {
        IoJob J;
        J.e = e; J.fd = fd; J.path = path; J.file_size = st.st_size; J.W = W;
        J.x0 = 0; J.y0 = 0; J.w = W; J.h = H;
        J.d_err = d_err;
        if (e->fast) {
            const LocalDomain &ld = e->local[(size_t)i];
            J.x0 = ld.x0; J.y0 = ld.y0; J.w = ld.L.width; J.h = ld.L.height;
            J.d_dense = e->scratch_dense;
        } else {
            J.d_dense = e->dense[e->cur];
        }
        rc = io_run(&J, true);
        if (!rc && e->fast) {
            const LocalDomain &ld = e->local[(size_t)i];
            char *dst = e->arena + e->arena_layout.buf_off[(size_t)i * GOL_NBUF + e->cur];
            if (cudaMemsetAsync(dst, 0, (size_t)ld.L.buffer_words32 * 4, e->stream) != cudaSuccess) rc = 1;
            if (!rc) rc = gol_launch_dense_to_padded(e->scratch_dense, ld.L.nq, 0, reinterpret_cast<uint32_t *>(dst), ld.L, e->stream);
            ++e->launches;
            if (!rc && cudaStreamSynchronize(e->stream) != cudaSuccess) { gol_set_error("dense_to_padded failed"); rc = 1; }
        }
    }
    if (!rc) {
        uint32_t err = 0;
        cudaMemcpy(&err, d_err, 4, cudaMemcpyDeviceToHost);
        if ((err & 2) || ((err & 1) && strict)) {
            gol_set_error("%s: %s", path, (err & 2) ? "a line does not end in a newline at column width (garbled record layout)"
                                                   : "invalid byte: only '0', '1' and newlines are accepted");
            rc = 1;
        }
        cudaMemset(d_err, 0, 4);
    }
    close(fd);
    return rc;
}

int gol_engine_store_text_file(gol_engine *e, const char *path)
{
    NvtxRange nv("gol store_text_file");
    GOL_CUDA(cudaSetDevice(e->cfg.device));
    const int64_t W = e->cfg.width, H = e->cfg.height;
    int fd = open(path, O_WRONLY | O_CREAT, 0644);
    if (fd < 0) GOL_FAIL("fopen: %s: %s", path, strerror(errno));
    if (ftruncate(fd, (off_t)(H * (W + 1))) != 0) {
        close(fd);
        GOL_FAIL("ftruncate: %s: %s", path, strerror(errno));
    }
    int rc = io_prepare(e);
    const int ndom_local = e->fast ? e->per_proc : 1;
    for (int i = 0; i < ndom_local && !rc; ++i) {
        IoJob J;
        J.e = e; J.fd = fd; J.path = path; J.file_size = H * (W + 1); J.W = W;
        J.x0 = 0; J.y0 = 0; J.w = W; J.h = H;
        J.d_err = nullptr;
        if (e->fast) {
            const LocalDomain &ld = e->local[(size_t)i];
            J.x0 = ld.x0; J.y0 = ld.y0; J.w = ld.L.width; J.h = ld.L.height;
            const char *src = e->arena + e->arena_layout.buf_off[(size_t)i * GOL_NBUF + e->cur];
            rc = gol_launch_padded_to_dense(reinterpret_cast<const uint32_t *>(src), ld.L, e->scratch_dense, ld.L.nq, 0, 0, e->stream);
            ++e->launches;
            J.d_dense = e->scratch_dense;
        } else {
            J.d_dense = e->dense[e->cur];
        }
        if (!rc && cudaStreamSynchronize(e->stream) != cudaSuccess) { gol_set_error("padded_to_dense failed"); rc = 1; }
        if (!rc) rc = io_run(&J, false);
    }
    close(fd);
    return rc;
}

// ------------------------------------------------------------------- driver
int gol_engine_prepare(gol_engine *e)
{
    NvtxRange nv("gol prepare (first halo exchange)");
    GOL_CUDA(cudaSetDevice(e->cfg.device));
    GOL_TRY(check_ready(e, "gol_engine_prepare"));
    if (e->fast) {
        GolLaunch lp{};
        geometry(e, 1, &lp);
        lp.in_idx = e->cur;
        lp.out_idx = e->cur;
        lp.wait_seq = 0;
        lp.seq = ++e->seq;
        lp.push = multi(e) ? 3 : 1;
        lp.push_rows = G;
        lp.rec_slot = -1;
        GOL_TRY(gol_launch_halo(e->d_doms, e->per_proc, lp, e->stream));
        ++e->launches;
        // geometry tuning belongs to set-up, not to the timed generation loop
        if (e->autotune && e->kmax >= 2 && !e->tuned_rows[e->kmax])
            GOL_TRY(autotune_geometry(e, e->kmax, e->cur, (e->cur + 1) % GOL_NBUF));
    }
    e->prepared = true;
    return 0;
}

int gol_engine_step(gol_engine *e, int64_t n)
{
    GOL_CUDA(cudaSetDevice(e->cfg.device));
    GOL_TRY(check_ready(e, "gol_engine_step"));
    if (!e->prepared) GOL_TRY(gol_engine_prepare(e));
    GOL_CUDA(cudaEventRecord(e->ev0, e->stream));
    if (e->fast) {
        GOL_TRY(run_epochs(e, n, -1, false, false, e->kmax));
    } else {
        for (int64_t left = n; left > 0; left -= kMaxBatch) GOL_TRY(dense_steps(e, (int)std::min<int64_t>(left, kMaxBatch), false));
    }
    GOL_CUDA(cudaEventRecord(e->ev1, e->stream));
    GOL_CUDA(cudaEventSynchronize(e->ev1));
    float ms = 0;
    GOL_CUDA(cudaEventElapsedTime(&ms, e->ev0, e->ev1));
    e->last_ms = ms;
    if (e->fast) {
        unsigned int err = 0;
        GOL_CUDA(cudaMemcpy(&err, e->d_error, 4, cudaMemcpyDeviceToHost));
        if (err) GOL_FAIL("device-side wait timed out (code %u)", err);
        GOL_TRY(check_device_errors(e));
    }
    return 0;
}

int gol_engine_run(gol_engine *e, gol_result *res)
{
    NvtxRange nv("gol run");
    GOL_CUDA(cudaSetDevice(e->cfg.device));
    GOL_TRY(check_ready(e, "gol_engine_run"));
    if (!e->prepared) GOL_TRY(gol_engine_prepare(e));
    GOL_TRY(e->fast ? run_fast(e, res) : run_dense(e, res));
    if (res) GOL_TRY(fill_population(e, res));
    return 0;
}

int gol_engine_sync(gol_engine *e)
{
    GOL_CUDA(cudaSetDevice(e->cfg.device));
    GOL_CUDA(cudaStreamSynchronize(e->stream));
    return 0;
}

int64_t gol_engine_launch_count(const gol_engine *e) { return e->launches; }

int gol_engine_chunk_rows(const gol_engine *e)
{
    if (!e->fast) return 0;
    GolLaunch lp{};
    geometry(e, e->kmax, &lp);
    return lp.chunk_rows;
}

int gol_engine_uses_deep_halo(const gol_engine *e) { return e->deep_halo ? 1 : 0; }

// Writes the per-launch timeline (GOL_TRACE=<n> at creation) as text: one line per strip-kernel
// launch with its depth, epoch mode (0 classic, 1 pair-first, 2 pair-second), flag bit, duration
// (first CTA start to last CTA end), gap to the previous launch's end and flag spin cycles.
int gol_engine_trace_dump(gol_engine *e, const char *path)
{
    if (!e->d_trace || !e->trace_n) return 0;
    GOL_CUDA(cudaSetDevice(e->cfg.device));
    GOL_CUDA(cudaStreamSynchronize(e->stream));
    std::vector<GolTrace> t((size_t)e->trace_n);
    GOL_CUDA(cudaMemcpy(t.data(), e->d_trace, sizeof(GolTrace) * t.size(), cudaMemcpyDeviceToHost));
    FILE *f = fopen(path, "w");
    if (!f) GOL_FAIL("fopen: %s: %s", path, strerror(errno));
    fprintf(f, "# launch k mode flags start_ns dur_ns gap_ns spin_sum_cycles spin_max_cycles\n");
    for (size_t i = 0; i < t.size(); ++i) {
        const int m = e->trace_meta[i];
        const long long gap = i ? (long long)t[i].t0_min - (long long)t[i - 1].t1_max : 0;
        fprintf(f, "%zu %d %d %d %llu %lld %lld %llu %llu\n", i, m & 255, (m >> 8) & 255, (m >> 16) & 1,
                t[i].t0_min - t[0].t0_min, (long long)(t[i].t1_max - t[i].t0_min), gap, t[i].spin_sum, t[i].spin_max);
    }
    fclose(f);
    return 0;
}
double gol_engine_last_ms(const gol_engine *e) { return e->last_ms; }

int gol_engine_population(gol_engine *e, uint64_t *count)
{
    GOL_CUDA(cudaSetDevice(e->cfg.device));
    unsigned long long *d_count = e->d_count;   // engine-owned: no allocation on this path
    GOL_CUDA(cudaMemsetAsync(d_count, 0, 8, e->stream));
    if (!e->fast) {
        GOL_TRY(gol_op_popcount_dense(e->dense[e->cur], e->cfg.width, e->cfg.height, d_count, e->stream));
        ++e->launches;
    } else {
        for (int i = 0; i < e->per_proc; ++i) {
            const char *src = e->arena + e->arena_layout.buf_off[(size_t)i * GOL_NBUF + e->cur];
            GOL_TRY(gol_launch_popcount_padded(reinterpret_cast<const uint32_t *>(src), e->local[(size_t)i].L, d_count, e->stream));
            ++e->launches;
        }
    }
    unsigned long long c = 0;
    GOL_CUDA(cudaMemcpyAsync(&c, d_count, 8, cudaMemcpyDeviceToHost, e->stream));
    GOL_CUDA(cudaStreamSynchronize(e->stream));
    *count = c;
    return 0;
}

int gol_engine_domain_buffer(gol_engine *e, int i, void **d_ptr, gol_layout *layout)
{
    if (!e->fast) {
        if (i != 0) GOL_FAIL("domain index %d out of range", i);
        *d_ptr = e->dense[e->cur];
        GOL_TRY(gol_layout_init(layout, e->cfg.width, e->cfg.height));
        layout->ghost_rows = 0;       // dense buffer: no padding at all
        layout->pitch32 = layout->nw32;
        layout->rows_padded = layout->height;
        layout->buffer_words32 = layout->rows_padded * layout->pitch32;
        return 0;
    }
    if (i < 0 || i >= e->per_proc) GOL_FAIL("domain index %d out of range", i);
    *d_ptr = e->arena + e->arena_layout.buf_off[(size_t)i * GOL_NBUF + e->cur];
    *layout = e->local[(size_t)i].L;
    return 0;
}

}  // extern "C"
