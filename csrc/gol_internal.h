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
#define GOL_NBUF 6        // padded buffers per domain: 4 rotating (a neighbouring rank may still push into or read the
                          // outputs of the last three epochs, see run_epochs) + the checkpoints of the batch being
                          // run and of the batch whose flags are still in flight
#define GOL_NDIR 8
// direction order: N, S, W, E, NW, NE, SW, SE  (dy,dx)
static const int GOL_DIR_DY[8] = {-1, 1, 0, 0, -1, -1, 1, 1};
static const int GOL_DIR_DX[8] = {0, 0, -1, 1, -1, 1, -1, 1};
// opposite direction index
static const int GOL_DIR_OPP[8] = {1, 0, 3, 2, 7, 6, 5, 4};
#define GOL_DIRMASK_NS 0xF3u   // N, S and the four corners: data pushed by the top / bottom chunks
#define GOL_DIRMASK_WE 0x0Cu   // W, E: data pushed by the first / last strip of every chunk

#define GOL_REC_SLOTS 64  // ring of flag records per process
#ifndef GOL_STRIP_WARPS
#define GOL_STRIP_WARPS 4  // warps (= work items) per CTA of the strip, halo and scan kernels
#endif
#define GOL_EDGE_ROWS 16   // height of the first and last chunk of a domain (the rows that are pushed to N / S)

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
    unsigned int cnt_ns;       // top / bottom chunk items that have finished their pushes
    unsigned int cnt_we;       // first / last strip items that have finished their pushes
};

// Per-launch timeline record (GOL_TRACE=<launches>): first CTA start / last CTA end on the GPU's
// global timer (ns) and the clock cycles border warps spent spinning on neighbour flags.
struct GolTrace {
    unsigned long long t0_min, t1_max, spin_sum, spin_max;
};

struct GolDomainDev {
    uint32_t *buf[GOL_NBUF];
    int32_t H, nw32, nwf32, rem, pitch32;
    int32_t global_id;
    GolNeighbor nb[GOL_NDIR];
    uint32_t *epoch_flags;     // [8] written by the neighbours (slot d = neighbour in direction d)
    GolScratch *scratch;
    // flag records: rec_tables[p] = table of process p (local or peer-mapped),
    // entry [slot * num_domains + global_id]
    unsigned long long **rec_tables;   // device array [proc_count]
    int32_t proc_count, num_domains;
    uint32_t self_mask;        // bit d set: the neighbour in direction d is this domain itself (torus wrap)
    long long wait_timeout;    // clock cycles a border warp spins on a neighbour flag before it gives up
};

struct GolLaunch {
    int32_t in_idx, out_idx;
    uint32_t seq;              // sequence number published with this launch's pushes (ghost flags carry it)
    uint32_t wait_seq;         // sequence the input ghosts must have reached (0 = no wait)
    int32_t chunk_rows;        // height of the interior chunks
    int32_t edge_rows;         // height of the first and last chunk (domains shorter than 3x this are one chunk)
    int32_t ctas_per_domain;
    int32_t push;              // bit 0: push borders into the neighbours' ghost zones; bit 1: publish the epoch
                               // flags afterwards (several processes); bits 8..: timing experiments
                               // (1<<8 no column push, 2<<8 no row push)
    int32_t push_rows;         // ghost depth pushed to N / S (<= GOL_GHOST_ROWS)
    int32_t ext_y, ext_x;      // ghost rows / ghost words per side that are computed and stored as well
                               // (first epoch of a deep-halo pair: the next epoch then needs no exchange)
    int32_t border_last;       // 1: border items are numbered last instead of first
    int32_t rec_slot;          // >=0: publish {alive,diff} record into this slot
    uint32_t rec_seq;
    int32_t pdl;               // 1: launch with programmatic stream serialisation
    GolTrace *trace;           // optional timeline record of this launch
};

// ------------------------------------------------------------ kernel entry
// (implemented in life_kernels.cu)
struct CUstream_st;
int gol_launch_strip(const GolDomainDev *d_doms, int ndoms, const GolLaunch &lp, int k, int want_flags,
                     CUstream_st *stream);
int gol_launch_halo(const GolDomainDev *d_doms, int ndoms, const GolLaunch &lp, CUstream_st *stream);
int gol_strip_geometry(int nw32, int ext_x, int *strips);
int gol_strip_ctas_per_sm(int k, int *ctas_per_sm);
int gol_launch_scan_alive(const GolDomainDev *d_doms, int ndoms, const GolLaunch &lp, CUstream_st *stream);
int gol_write_all(int fd, const char *buf, size_t len, int64_t offset, const char *path);
int gol_launch_gather(unsigned long long *rec_table, int slot, int num_domains, uint32_t seq,
                      unsigned long long *d_result, unsigned int *d_error, long long timeout_cycles,
                      CUstream_st *stream);
int gol_launch_popcount_padded(const uint32_t *d_pad, const gol_layout &L, unsigned long long *d_count, CUstream_st *stream);
int gol_launch_hash_padded(const uint32_t *d_pad, const gol_layout &L, int64_t qx0, int64_t y0, unsigned long long *d_sum,
                           CUstream_st *stream);
int gol_launch_dense_to_padded(const uint64_t *d_dense, int64_t dense_pitch_q, int64_t qx0, uint32_t *d_pad,
                               const gol_layout &L, CUstream_st *stream);
int gol_launch_padded_to_dense(const uint32_t *d_pad, const gol_layout &L, uint64_t *d_dense,
                               int64_t dense_pitch_q, int64_t qx0, int unused, CUstream_st *stream);
