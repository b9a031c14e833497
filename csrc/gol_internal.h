// Internal declarations shared by the native sources of libgolb200.
#pragma once

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>

#include "../include/gol_b200.h"

// ------------------------------------------------------------ error plumbing
void gol_set_error(const char *fmt, ...);

#define GOL_FAIL(...)               \
    do {                            \
        gol_set_error(__VA_ARGS__); \
        return 1;                   \
    } while (0)

// Checked CUDA call (the library never exits; cf. cudaSafeCall G:18-27).
#define GOL_CUDA(call)                                                                   \
    do {                                                                                 \
        cudaError_t err__ = (call);                                                      \
        if (err__ != cudaSuccess) {                                                      \
            gol_set_error("CUDA error %s at %s:%d: %s", cudaGetErrorName(err__), __FILE__, \
                          __LINE__, cudaGetErrorString(err__));                          \
            return 1;                                                                    \
        }                                                                                \
    } while (0)

#define GOL_TRY(expr)            \
    do {                         \
        if ((expr) != 0) return 1; \
    } while (0)

// ------------------------------------------------------------------- layout
static inline int64_t gol_div_up(int64_t a, int64_t b) { return (a + b - 1) / b; }

// Hash used by the seeded generator: one 64-bit word of cells per (seed, word x, row y).
// splitmix64 finaliser; a pure function so any decomposition produces the same grid.
#if defined(__CUDACC__)
__host__ __device__
#endif
static inline uint64_t gol_hash64(uint64_t seed, uint64_t qx, uint64_t y)
{
    uint64_t z = seed + 0x9E3779B97F4A7C15ull * (y * 0x100000001B3ull + qx + 1);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// ------------------------------------------------------- device descriptors
#define GOL_NBUF 5        // padded buffers per domain: 3 rotating (in / out / previous in) + the checkpoints of
                          // the batch being run and of the batch whose flags are still in flight
#define GOL_NDIR 8
// direction order: N, S, W, E, NW, NE, SW, SE  (dy,dx)
static const int GOL_DIR_DY[8] = {-1, 1, 0, 0, -1, -1, 1, 1};
static const int GOL_DIR_DX[8] = {0, 0, -1, 1, -1, 1, -1, 1};
// opposite direction index
static const int GOL_DIR_OPP[8] = {1, 0, 3, 2, 7, 6, 5, 4};

#define GOL_REC_SLOTS 64  // ring of flag records per process
#ifndef GOL_STRIP_WARPS
#define GOL_STRIP_WARPS 4  // warps (= work items) per CTA of the strip, halo and scan kernels
#endif

struct GolNeighbor {
    uint32_t *buf[GOL_NBUF];   // neighbour's padded buffers (local or peer-mapped)
    uint32_t *flag;            // slot in the neighbour's epoch-flag array that this domain writes
    int32_t H;                 // neighbour's local height
    int32_t nw32, rem;         // neighbour's data words / width % 64
    int32_t pitch32;
};

struct GolScratch {            // per-domain device scratch (zero between launches)
    unsigned int done;         // CTAs finished in the current launch
    unsigned int alive;        // OR of final-level cells
    unsigned int diff;         // OR of final-level (old ^ new)
    unsigned int error;        // sticky: flag-wait timeout etc.
    unsigned int border_done[2];  // border work items finished, per epoch parity (two epochs can be in flight)
};

struct GolDomainDev {
    uint32_t *buf[GOL_NBUF];
    int32_t H, nw32, nwf32, rem, pitch32;
    int32_t strips;            // strips of the strip kernel for this domain
    int32_t global_id;
    GolNeighbor nb[GOL_NDIR];
    uint32_t *epoch_flags;     // [8] written by the neighbours (slot d = neighbour in direction d)
    uint32_t *item_flags;      // [items] sequence of the last epoch each work item completed (persistent kernel)
    GolScratch *scratch;
    // flag records: rec_tables[p] = table of process p (local or peer-mapped),
    // entry [slot * num_domains + global_id]
    unsigned long long **rec_tables;   // device array [proc_count]
    int32_t proc_count, num_domains;
    uint32_t self_mask;        // bit d set: the neighbour in direction d is this domain itself (torus wrap)
    uint32_t mul2, mul4;       // the constants 2 and 4, kept opaque to ptxas so that shifts stay IMADs (FMA pipe)
};

struct GolLaunch {
    int32_t in_idx, out_idx;
    uint32_t seq;              // this launch's sequence number (ghost flags carry it)
    uint32_t wait_seq;         // sequence the input ghosts must have reached (0 = no wait)
    int32_t chunk_rows, chunks;
    int32_t ctas_per_domain;
    int32_t signal;            // 1: publish epoch flags to neighbours when done
    int32_t rec_slot;          // >=0: publish {alive,diff} record into this slot
    uint32_t rec_seq;
    int32_t pdl;               // 1: launch with programmatic stream serialisation (single-process runs)
};

// Persistent (dataflow) launch: `epochs` epochs of K generations in ONE kernel; work items wait
// for their 3x3 neighbourhood of the previous epoch through per-item flags instead of a
// kernel boundary.
struct GolPersist {
    int32_t epochs;
    uint32_t base_seq;         // epoch e carries sequence base_seq + 1 + e
    int32_t first_in;          // buffer read by epoch 0
    int32_t rot[3];            // epoch e writes rot[e % 3] (and reads rot[(e-1) % 3] for e > 0)
    int32_t chunk_rows;
    int32_t items_per_domain;  // uniform (padded) item count per domain
    int32_t ndoms;
    int32_t total_warps;
    uint32_t wait_seq0;        // ghost-flag sequence epoch 0's border items wait for (0 = none)
    unsigned int *work_counter; // device word, zero at launch: next (epoch, item) to hand out
};

// ------------------------------------------------------------ kernel entry
// (implemented in life_kernels.cu)
struct CUstream_st;
int gol_launch_strip(const GolDomainDev *d_doms, int ndoms, const GolLaunch &lp, int k, int want_flags,
                     CUstream_st *stream);
int gol_launch_halo(const GolDomainDev *d_doms, int ndoms, const GolLaunch &lp, CUstream_st *stream);
int gol_strip_geometry(int nw32, int *strips);
int gol_persistent_max_ctas(int k, int device, int *ctas);
int gol_strip_ctas_per_sm(int k, int *ctas_per_sm);
int gol_launch_pipe(const GolDomainDev *d_doms, int ndoms, const GolLaunch &lp, int lv, int d, int want_flags,
                    CUstream_st *stream);
int gol_pipe_ctas_per_sm(int lv, int d, int *ctas_per_sm);
int gol_launch_persistent(const GolDomainDev *d_doms, const GolPersist &pp, int k, int ctas, CUstream_st *stream);
int gol_launch_scan_alive(const GolDomainDev *d_doms, int ndoms, const GolLaunch &lp, CUstream_st *stream);
int gol_write_all(int fd, const char *buf, size_t len, int64_t offset, const char *path);
int gol_launch_gather(unsigned long long *rec_table, int slot, int num_domains, uint32_t seq,
                      unsigned long long *d_result, unsigned int *d_error, CUstream_st *stream);
int gol_launch_popcount_padded(const uint32_t *d_pad, const gol_layout &L, unsigned long long *d_count, CUstream_st *stream);
int gol_launch_dense_to_padded(const uint64_t *d_dense, int64_t dense_pitch_q, int64_t qx0, uint32_t *d_pad,
                               const gol_layout &L, CUstream_st *stream);
int gol_launch_padded_to_dense(const uint32_t *d_pad, const gol_layout &L, uint64_t *d_dense,
                               int64_t dense_pitch_q, int64_t qx0, int unused, CUstream_st *stream);
