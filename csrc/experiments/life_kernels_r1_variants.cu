// Generation kernels for sm_100a.
//
//  life_strip_kernel<K, FLAGS>  — the product kernel.  One warp owns a vertical strip of the
//      padded bit grid (32 lanes x 64 cells, the two outer lanes are a redundant apron) and
//      streams down the rows.  K generations are pipelined in registers (temporal blocking):
//      level t consumes the row produced by level t-1 one step earlier, so a row is read from
//      memory once and written once per K generations.  Neighbour words move between lanes
//      with warp shuffles; the B3/S23 rule is an 8-LOP3 bit-sliced adder per 32 cells.
//      The same kernel pushes its border rows/columns straight into the neighbours' ghost
//      zones (its own for a single torus domain, peer-mapped memory over NVLink otherwise),
//      publishes release flags, and (FLAGS) reduces the alive / changed bits of the last level.
//      Replaces evolve (S:60-101, C:27-53, O:29-57, G:128-148), halo_rows/halo_cols (G:52-74),
//      the MPI halo exchange (C:280-344), compare/similarity (G:76-100, C:83-109) and
//      empty (G:102-126, C:55-81).
//  halo_push_kernel — the border push alone (first exchange after a load).
//  gather_flags_kernel — flag all-reduce: waits for every domain's record of a batch.
//  dense kernels — generic 1-generation torus step for any size (tiny grids, PyTorch op).
#include <cuda_runtime.h>

#include <algorithm>
#include <type_traits>

#include "gol_internal.h"

namespace {

constexpr int kWarps = GOL_STRIP_WARPS;  // warps per CTA
constexpr int kWL = 2;              // 32-bit words per lane (64 cells)
constexpr int kValidLanes = 30;     // lanes 1..30 own output, lanes 0 and 31 are apron
constexpr int kVW = kValidLanes * kWL;  // owned words per strip
constexpr int G = GOL_GHOST_ROWS;

// First owned word of a strip.  Strips are 60 words wide; the last one is right-aligned to the
// row end (it may overlap its left neighbour, which then computes identical words twice) so
// that it always owns the row's last 128 cells, which it pushes east.
__device__ __forceinline__ int strip_origin(int strip, int S, int nw)
{
    return (strip == S - 1 && S > 1) ? nw - kVW : strip * kVW;
}

template <int LUT>
__device__ __forceinline__ uint32_t lop3(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, %4;" : "=r"(d) : "r"(a), "r"(b), "r"(c), "n"(LUT));
    return d;
}
__device__ __forceinline__ uint32_t xor3(uint32_t a, uint32_t b, uint32_t c) { return lop3<0x96>(a, b, c); }
__device__ __forceinline__ uint32_t maj3(uint32_t a, uint32_t b, uint32_t c) { return lop3<0xE8>(a, b, c); }

// Next state of 32 cells from the bit-sliced 3-cell row sums (s0 + 2*s1, centre included) of
// the rows above (a), at (b) and below (n) and the centre word m.  T = 9-cell sum:
// alive' = (T == 3) | (m & T == 4).
// Six LOP3, found by exact synthesis (scripts/search/exact_synth.py, z3): the circuit is only
// required to be right where the centre bit is consistent with its row sum (m = 1 implies the
// middle sum b >= 1, which holds by construction because b includes the centre cell); it differs
// from the rule on two impossible inputs only ((a, b, n, m) = (0, 0, 0, 1) and (2, 0, 2, 1)).  Former forms, kept
// for comparison: GOL_RULE=7 (parity / all-equal encoding of the low column + 3 gates, found by
// scripts/search/lop3_search.c) and GOL_RULE=8 (two full adders + 4 gates).
#ifndef GOL_RULE
#define GOL_RULE 6
#endif
__device__ __forceinline__ uint32_t life_rule(uint32_t a0, uint32_t a1, uint32_t b0, uint32_t b1,
                                              uint32_t n0, uint32_t n1, uint32_t m)
{
#if GOL_RULE == 8
    const uint32_t x0 = xor3(a0, b0, n0);
    const uint32_t k0 = maj3(a0, b0, n0);
    const uint32_t x1 = xor3(a1, b1, n1);
    const uint32_t k1 = maj3(a1, b1, n1);
    const uint32_t g = lop3<0x14>(k0, x1, k1);   // (k0 ^ x1) & ~k1            : T in {2,3}
    const uint32_t h = lop3<0x42>(k0, x1, k1);   // ~(k0 ^ x1) & (k1 ^ (k0&x1)) : T in {4,5}
    const uint32_t u = lop3<0x08>(x0, h, m);     // ~x0 & h & m                 : T == 4, alive
    return lop3<0xEA>(x0, g, u);                 // (x0 & g) | u
#elif GOL_RULE == 7
    const uint32_t p0 = xor3(a0, b0, n0);        // low column: count odd
    const uint32_t e0 = lop3<0x81>(a0, b0, n0);  // low column: count 0 or 3
    const uint32_t p1 = xor3(a1, b1, n1);        // high column: count odd
    const uint32_t k1 = maj3(a1, b1, n1);        // high column: count >= 2
    const uint32_t g1 = lop3<0x56>(p0, e0, k1);
    const uint32_t g2 = lop3<0xA4>(p0, m, g1);
    return lop3<0x60>(g2, e0, p1);
#else
    const uint32_t p0 = xor3(a0, b0, n0);        // low column: count odd
    const uint32_t d0 = lop3<0x7E>(a0, b0, n0);  // low column: count 1 or 2
    const uint32_t o1 = lop3<0x16>(a1, b1, n1);  // high column: count exactly 1
    const uint32_t w = maj3(a1, n0, n1);         // (no symmetric meaning; part of the synthesised circuit)
    const uint32_t g = lop3<0x76>(w, o1, d0);
    return lop3<0x54>(m, p0, g);
#endif
}

// Experiment kept for the record (off): funnel shifts by one bit on the FMA pipe instead of the
// ALU pipe that the LOP3 adder saturates: (cur << 1) | (prev >> 31) = cur * 2 + hi32(prev * 2),
// (cur >> 1) | (next << 31) = hi32(cur * 2^31) + lo32(next * 2^31).  Measured on B200: slower
// (K=8: 11.03 vs 10.27 us/generation at 16384^2) - IMAD.HI costs more than the SHF it replaces.
#ifndef GOL_IMAD_SHIFTS
#define GOL_IMAD_SHIFTS 0
#endif
__device__ __forceinline__ uint32_t shl1_in(uint32_t prev, uint32_t cur)
{
#if GOL_IMAD_SHIFTS
    uint32_t t, d;
    asm("mul.hi.u32 %0, %1, 2;" : "=r"(t) : "r"(prev));
    asm("mad.lo.u32 %0, %1, 2, %2;" : "=r"(d) : "r"(cur), "r"(t));
    return d;
#else
    return __funnelshift_l(prev, cur, 1);
#endif
}
__device__ __forceinline__ uint32_t shr1_in(uint32_t cur, uint32_t next)
{
#if GOL_IMAD_SHIFTS
    uint32_t t, d;
    asm("mul.lo.u32 %0, %1, 0x80000000;" : "=r"(t) : "r"(next));
    asm("mad.hi.u32 %0, %1, 0x80000000, %2;" : "=r"(d) : "r"(cur), "r"(t));
    return d;
#else
    return __funnelshift_r(cur, next, 1);
#endif
}

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *p)
{
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t *p, uint32_t v)
{
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys64(const unsigned long long *p)
{
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

// ---------------------------------------------------------------- border push
// Copies the border rows / columns of rows [y0,y1) of buffer `o` owned by `strip` into the
// ghost zones of the 8 neighbours.  Executed by one warp; reads data this warp wrote earlier.
__device__ void push_borders(const GolDomainDev &D, int o, int strip, int y0, int y1, int lane, int dbg = 0)
{
    const int H = D.H, nw = D.nw32, pitch = D.pitch32;
    const uint32_t *src = D.buf[o];
    const int lo = strip_origin(strip, D.strips, nw), hi = min(nw, lo + kVW);
    const int npairs = (hi - lo) >> 1;  // uint2 elements per row owned by this strip (<= 30)

    // rows -> north / south neighbours' ghost rows.  All loads of a batch are issued before the
    // first store, so a batch costs one memory round trip instead of one per row.
    for (int pass = 0; pass < 2 && !(dbg & 2); ++pass) {
        const int ya = pass == 0 ? y0 : max(y0, H - G);
        const int yb = pass == 0 ? min(y1, G) : y1;
        if (ya >= yb) continue;
        const GolNeighbor &nb = D.nb[pass];  // 0 = N, 1 = S
        const int total = (yb - ya) * npairs;
        const int drow0 = pass == 0 ? (G + nb.H) : (G - H);
        for (int base = 0; base < total; base += 32 * 8) {
            uint2 v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int e = base + u * 32 + lane;
                if (e < total) {
                    const int row = e / npairs, col = e - row * npairs;
                    v[u] = *reinterpret_cast<const uint2 *>(src + (int64_t)(ya + row + G) * pitch + GOL_GHOST_WORDS32 + lo + 2 * col);
                }
            }
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int e = base + u * 32 + lane;
                if (e < total) {
                    const int row = e / npairs, col = e - row * npairs;
                    *reinterpret_cast<uint2 *>(nb.buf[o] + (int64_t)(drow0 + ya + row) * nb.pitch32 + GOL_GHOST_WORDS32 + lo + 2 * col) = v[u];
                }
            }
        }
    }

    const bool first = strip == 0, last = strip == D.strips - 1;
    if ((!first && !last) || (dbg & 1)) return;
    const int W_cells = D.nwf32 * 32 + D.rem;
    const int ob = W_cells - 128, wi0 = ob >> 5, bo = ob & 31;  // east: my last 128 cells start at word wi0, bit bo
    constexpr int RB = 4;  // rows per lane and batch
    for (int yb0 = y0; yb0 < y1; yb0 += 32 * RB) {
        uint32_t wv[RB][4], ev[RB][5];
#pragma unroll
        for (int u = 0; u < RB; ++u) {
            const int y = yb0 + u * 32 + lane;
            if (y < y1) {
                const uint32_t *s = src + (int64_t)(y + G) * pitch + GOL_GHOST_WORDS32;
                if (first) {
                    const uint4 q = *reinterpret_cast<const uint4 *>(s);
                    wv[u][0] = q.x; wv[u][1] = q.y; wv[u][2] = q.z; wv[u][3] = q.w;
                }
                if (last) {
#pragma unroll
                    for (int j = 0; j < 5; ++j) ev[u][j] = (wi0 + j < nw) ? s[wi0 + j] : 0u;
                }
            }
        }
#pragma unroll
        for (int u = 0; u < RB; ++u) {
            const int y = yb0 + u * 32 + lane;
            if (y >= y1) continue;
            if (first) {
                // my first cells -> west neighbours' right ghost slot, pre-shifted by their width % 64
                for (int v = 0; v < 3; ++v) {  // W, NW, SW
                    int dir, drow;
                    if (v == 0) { dir = 2; drow = G + y; }
                    else if (v == 1) { if (y >= G) continue; dir = 4; drow = G + D.nb[4].H + y; }
                    else { if (y < H - G) continue; dir = 6; drow = G + y - H; }
                    const GolNeighbor &nb = D.nb[dir];
                    const int sh = nb.rem;
                    uint32_t outw[4];
                    if (sh == 0) {
                        outw[0] = wv[u][0]; outw[1] = wv[u][1]; outw[2] = wv[u][2]; outw[3] = wv[u][3];
                    } else {
                        const int ws = sh >> 5, bs = sh & 31;
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            uint32_t val = 0;
#pragma unroll
                            for (int a = 0; a < 4; ++a) {
                                if (a == j - ws) val |= wv[u][a] << bs;
                                if (bs && a == j - ws - 1) val |= wv[u][a] >> (32 - bs);
                            }
                            outw[j] = val;
                        }
                    }
                    uint32_t *d = nb.buf[o] + (int64_t)drow * nb.pitch32 + GOL_GHOST_WORDS32 + nb.nw32;
                    // slot is 8-byte aligned (nw32 is even): two 64-bit stores
                    *reinterpret_cast<uint2 *>(d) = make_uint2(outw[0], outw[1]);
                    *reinterpret_cast<uint2 *>(d + 2) = make_uint2(outw[2], outw[3]);
                }
            }
            if (last) {
                // my last 128 cells -> east neighbours' left ghost slot
                uint32_t outw[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) outw[j] = bo ? ((ev[u][j] >> bo) | (ev[u][j + 1] << (32 - bo))) : ev[u][j];
                for (int v = 0; v < 3; ++v) {  // E, NE, SE
                    int dir, drow;
                    if (v == 0) { dir = 3; drow = G + y; }
                    else if (v == 1) { if (y >= G) continue; dir = 5; drow = G + D.nb[5].H + y; }
                    else { if (y < H - G) continue; dir = 7; drow = G + y - H; }
                    const GolNeighbor &nb = D.nb[dir];
                    uint32_t *d = nb.buf[o] + (int64_t)drow * nb.pitch32;
                    *reinterpret_cast<uint4 *>(d) = make_uint4(outw[0], outw[1], outw[2], outw[3]);
                }
            }
        }
    }
}

// Spin until the 8 neighbours have published sequence >= want into my epoch flags.
__device__ void wait_neighbors(const GolDomainDev &D, uint32_t want, int lane, uint32_t skip_mask = 0)
{
    if (lane < GOL_NDIR && !((skip_mask >> lane) & 1u)) {
        const uint32_t *f = D.epoch_flags + lane;
        long long t0 = clock64();
        while ((int32_t)(ld_acquire_sys(f) - want) < 0) {
            if (clock64() - t0 > (1ll << 32) || *(volatile unsigned int *)&D.scratch->error) {  // ~2 s: neighbour is gone
                atomicOr(&D.scratch->error, 1u);
                break;
            }
            __nanosleep(64);
        }
    }
    __syncwarp();
    __threadfence();  // acquire for the whole warp: drop stale L1 lines of the ghost zones
}

// Work items of a domain are (strip, chunk) pairs.  They are numbered so that every item that
// touches a ghost zone or pushes to a neighbour comes first: the top and bottom chunks, then
// the first/last strip of the middle chunks, then the interior.  Border pushes therefore leave
// at the very start of a launch and the neighbours' flags are published long before the
// interior is done - the halo exchange is hidden behind the interior items.
struct ItemMap {
    int S, chunks, top, first_bot, nborder;
};
__device__ __forceinline__ ItemMap item_map(const GolDomainDev &D, int chunk_rows)
{
    ItemMap m;
    m.S = D.strips;
    m.chunks = (D.H + chunk_rows - 1) / chunk_rows;
    m.top = min(m.chunks, (G + chunk_rows - 1) / chunk_rows);           // chunks with y0 < G
    m.first_bot = min(m.chunks, max(m.top, (D.H - G) / chunk_rows));    // first chunk with y1 > H - G
    const int mid = m.first_bot - m.top;
    m.nborder = (m.chunks - mid) * m.S + mid * min(m.S, 2);
    return m;
}
// returns false for idle items (beyond this domain's item count)
__device__ __forceinline__ bool decode_item(const ItemMap &m, int item, int &strip, int &chunk, bool &border)
{
    const int S = m.S, n1 = m.top * S, n2 = (m.chunks - m.first_bot) * S, mid = m.first_bot - m.top;
    const int e = min(S, 2), n3 = mid * e;
    border = true;
    if (item < n1) { chunk = item / S; strip = item - chunk * S; return true; }
    item -= n1;
    if (item < n2) { const int c = item / S; chunk = m.first_bot + c; strip = item - c * S; return true; }
    item -= n2;
    if (item < n3) { const int c = item / e; chunk = m.top + c; strip = (item - c * e) ? S - 1 : 0; return true; }
    item -= n3;
    border = false;
    if (S <= 2 || item >= mid * (S - 2)) return false;
    const int c = item / (S - 2);
    chunk = m.top + c;
    strip = 1 + item - c * (S - 2);
    return true;
}

// Called by every border warp after its pushes: the last one publishes the epoch flags.
__device__ __forceinline__ void border_item_done(const GolDomainDev &D, const GolLaunch &lp, int nborder, int lane)
{
    // Release pattern with one system-scope fence per epoch instead of one per border warp: every
    // border warp orders its pushes before its count at GPU scope; the warp that completes the
    // count issues the (cumulative) system-scope fence before it publishes the flags, which
    // carries all counted warps' pushes with it (PTX memory model: causality order is transitive
    // over the gpu-scope atomic and the sys-scope release).
    __threadfence();
    __syncwarp();
    if (lane == 0) {
        // In the persistent kernel border items of two consecutive epochs can be in flight (never
        // more: an item of epoch e+2 needs the neighbours' flags of e+1, which need ours of e).
        unsigned int *cnt = &D.scratch->border_done[lp.seq & 1u];
        const unsigned prev = atomicAdd(cnt, 1u);
        if (prev == (unsigned)nborder - 1) {
            *cnt = 0;
            __threadfence_system();
            for (int d = 0; d < GOL_NDIR; ++d) st_release_sys(D.nb[d].flag, lp.seq);
        }
    }
}

// CTA epilogue for launches that publish a flag record: the last CTA of the domain does it.
__device__ void finish_domain(const GolDomainDev &D, const GolLaunch &lp)
{
    __syncthreads();
    if (threadIdx.x != 0) return;
    __threadfence();
    const unsigned prev = atomicAdd(&D.scratch->done, 1u);
    if (prev != (unsigned)lp.ctas_per_domain - 1) return;
    __threadfence();
    const unsigned a = atomicExch(&D.scratch->alive, 0u);
    const unsigned d = atomicExch(&D.scratch->diff, 0u);
    const unsigned long long rec =
        ((unsigned long long)lp.rec_seq << 32) | (a ? 1ull : 0ull) | (d ? 2ull : 0ull);
    for (int p = 0; p < D.proc_count; ++p) {
        unsigned long long *t = D.rec_tables[p] + (int64_t)lp.rec_slot * D.num_domains + D.global_id;
        asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(t), "l"(rec) : "memory");
    }
    D.scratch->done = 0;
    __threadfence();
}

// ------------------------------------------------------------- strip kernel
__device__ __forceinline__ uint2 ldg_v2(const uint32_t *p)
{
    uint2 v;
    asm("ld.global.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p));
    return v;
}
__device__ __forceinline__ uint32_t ldg_u32(const uint32_t *p)
{
    uint32_t v;
    asm("ld.global.u32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ void stg_v2(uint32_t *p, uint32_t a, uint32_t b)
{
    asm volatile("st.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(a), "r"(b) : "memory");
}
__device__ __forceinline__ void stg_u32(uint32_t *p, uint32_t a)
{
    asm volatile("st.global.u32 [%0], %1;" ::"l"(p), "r"(a) : "memory");
}

// One-sided formulation (experiment, off by default): all three neighbour columns are taken from LEFT shifts,
//   C = row word, A = C*2 + carry (cell x-1 at bit x), B = C*4 + carry (cell x-2 at bit x),
// so the sums for cell x-1 appear at bit x and every level moves the picture one bit to the
// left; the K bits are shifted back once per epoch.  Multiplying by 2 and 4 and adding the
// carry is ONE full-rate IMAD each on the FMA pipe (the carries are IMAD.HI), which leaves the
// ALU pipe with exactly the 10 LOP3 of the adder instead of 10 LOP3 + 2 SHF.  The multipliers
// come from memory so that ptxas cannot turn them back into shifts.
// Measured on B200 (16384^2, K=8): 12.04 us/generation against 10.24 for the funnel-shift form:
// IMAD.HI issues at half rate (bin/gol-intpeak: 0.99 vs 1.98 warp-instr/clk/SM) and lengthens the
// per-level dependency chain, which costs more than the two SHF it saves.
#ifndef GOL_ONE_SIDED
#define GOL_ONE_SIDED 0
#endif
__device__ __forceinline__ uint32_t mul_hi(uint32_t a, uint32_t b)
{
    uint32_t d;
    asm("mul.hi.u32 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
    return d;
}
__device__ __forceinline__ uint32_t mad_lo(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t d;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// Shared-memory row FIFO between two consecutive pipeline stages (warps) of a CTA: kRingRows rows of
// 64 words; `head` counts rows produced, `tail` rows consumed.
constexpr int kRingRows = 8;
struct RingPort {
    uint32_t *data;
    volatile int *head, *tail;
};

struct MarchArgs {
    const uint32_t *in;   // padded input buffer (global)
    uint32_t *out;        // padded output buffer (global)
    int pitch, nw, nwf, rem;
    int vi;               // virtual word index of this lane's first word
    int y0, y1;           // output rows [y0, y1)
    uint32_t keep[kWL];   // per-word ownership / tail mask of this lane
    int mode;             // 1 plain vector load, 2 composed ragged tail, 0 nothing to load
    uint32_t two, four;   // run-time 2 and 4 (see GOL_ONE_SIDED)
    RingPort rin, rout;   // pipeline kernel only: row source / sink in shared memory
    int lane;
};

// One warp streams rows [y0-K, y1+K) of generation g through K register-resident levels and
// writes rows [y0, y1) of generation g+K.  FAST: every lane does a plain vector load and a
// (predicated) vector store - true for all strips except the ragged right edge of a row.
// Skewed pipeline (experiment, off by default): level t consumes the row level t-1 produced in the PREVIOUS step, so
// the K levels of one step are independent instruction streams (ILP = K x 2 words per thread)
// instead of one serial chain of K levels.  Costs K-1 extra drain steps per chunk and one
// pending row per level.  Measured on B200 (16384^2, K=8): 11.00 us/generation against 10.24 - the
// kernel is ALU-throughput bound, not latency bound, so the extra ILP buys nothing and the drain
// steps cost 7 %.
#ifndef GOL_SKEW
#define GOL_SKEW 0
#endif
#ifndef GOL_WARM_SPLIT
#define GOL_WARM_SPLIT 1   // prune the inactive levels of the pipeline fill (see march)
#endif

template <int K, bool FLAGS, bool FAST, bool SRC_RING = false, bool DST_RING = false>
__device__ __forceinline__ void march(const MarchArgs &A, uint32_t &flag_alive, uint32_t &flag_diff)
{
    constexpr int LAG = GOL_SKEW ? 2 * K - 1 : K;  // a step that consumes level-0 row r emits row r - LAG
    const int pitch = A.pitch;
    const bool store_all = (A.keep[0] & A.keep[1]) == 0xffffffffu;
    const bool store_any = (A.keep[0] | A.keep[1]) != 0u;
    const int rbeg = A.y0 - K, rload_end = A.y1 + K;  // level-0 rows consumed: [rbeg, rload_end)
    const int rend = A.y1 + LAG;                      // steps: r in [rbeg, rend)
    const uint32_t *pin = A.in + (int64_t)(rbeg + G) * pitch + GOL_GHOST_WORDS32;  // row being loaded
    uint32_t *pout = A.out + (int64_t)(A.y0 + G) * pitch + GOL_GHOST_WORDS32 + A.vi;

    int n_in = 0, seen_head = 0, n_out = 0, seen_tail = 0;  // ring bookkeeping (pipeline stages)
    auto load_row = [&](uint32_t(&w)[kWL]) {
        if (SRC_RING) {
            // rows before n_in have been copied out of the ring: hand their slots back
            // (the row fetched by the previous call may still be in flight; everything before it has
            // been moved on into the level-0 registers)
            __syncwarp();
            if (A.lane == 0 && n_in > 0) *A.rin.tail = n_in - 1;
            while (seen_head <= n_in) seen_head = *A.rin.head;
            __threadfence_block();
            const uint2 v = *reinterpret_cast<const uint2 *>(A.rin.data + (n_in % kRingRows) * 64 + A.lane * kWL);
            w[0] = v.x; w[1] = v.y;
            ++n_in;
            return;
        }
        if (FAST || A.mode == 1) {
            const uint2 v = ldg_v2(pin + A.vi);
            w[0] = v.x; w[1] = v.y;
        } else if (A.mode == 2) {
#pragma unroll
            for (int j = 0; j < kWL; ++j) {
                const int i = A.vi + j;
                uint32_t v = 0;
                if (i < A.nwf) v = ldg_u32(pin + i);
                else if (i < A.nwf + 4) v = (i < A.nw ? ldg_u32(pin + i) : 0u) | ldg_u32(pin + A.nw + (i - A.nwf));
                w[j] = v;
            }
        } else {
            w[0] = 0; w[1] = 0;
        }
        pin += pitch;
    };

    uint32_t h0[K][2][kWL], h1[K][2][kWL], ctr[K][kWL], pend[K][kWL];
#pragma unroll
    for (int t = 0; t < K; ++t)
#pragma unroll
        for (int j = 0; j < kWL; ++j) {
            h0[t][0][j] = h0[t][1][j] = h1[t][0][j] = h1[t][1][j] = 0;
            ctr[t][j] = 0;
            pend[t][j] = 0;
        }
    uint32_t acc_a[kWL] = {0, 0}, acc_d[kWL] = {0, 0};

    // two rows of prefetch in flight
    // q[p] holds the row for the next step of parity p: a row is requested two full steps before it
    // is touched again (no queue shifting - a move out of a register with a load in flight stalls)
    uint32_t q[2][kWL];
    load_row(q[0]);
    if (rbeg + 1 < rload_end) load_row(q[1]);
    else { q[1][0] = q[0][0]; q[1][1] = q[0][1]; }

    // One level: consumes row `x` of generation g+t, updates the level's 3-row window and returns
    // the row of generation g+t+1 that sits one row above x.
    auto level = [&](auto t_tag, auto par_tag, const uint32_t(&x)[kWL], uint32_t(&out)[kWL], bool flags_ok) {
        constexpr int t = decltype(t_tag)::value;
        constexpr int PAR = decltype(par_tag)::value;
#if GOL_ONE_SIDED
        uint32_t c4[kWL], c2[kWL];  // carries out of each word: its top two bits / top bit
#pragma unroll
        for (int j = 0; j < kWL; ++j) {
            c4[j] = mul_hi(x[j], A.four);
            c2[j] = mul_hi(x[j], A.two);
        }
        const uint32_t in4 = __shfl_up_sync(0xffffffffu, c4[kWL - 1], 1);
        const uint32_t in2 = __shfl_up_sync(0xffffffffu, c2[kWL - 1], 1);
#else
        const uint32_t left = __shfl_up_sync(0xffffffffu, x[kWL - 1], 1);
        const uint32_t right = __shfl_down_sync(0xffffffffu, x[0], 1);
#endif
#pragma unroll
        for (int j = 0; j < kWL; ++j) {
#if GOL_ONE_SIDED
            const uint32_t nL = mad_lo(x[j], A.four, j ? c4[j - 1] : in4);  // cell x-2 at bit x
            const uint32_t nC = mad_lo(x[j], A.two, j ? c2[j - 1] : in2);   // cell x-1 at bit x (centre)
            const uint32_t nR = x[j];
#elif GOL_TMIX
            // timing-only experiment (computes a different automaton): the instruction mix of a
            // strip-transposed bit layout, where x-neighbours are whole neighbouring words
            const uint32_t nL = j ? x[j - 1] : left * A.two;
            const uint32_t nC = x[j];
            const uint32_t nR = j < kWL - 1 ? x[j + 1] : (right >> (A.four & 1u));
#else
            const uint32_t nL = shl1_in(j ? x[j - 1] : left, x[j]);
            const uint32_t nC = x[j];
            const uint32_t nR = shr1_in(x[j], j < kWL - 1 ? x[j + 1] : right);
#endif
            const uint32_t s0 = xor3(nL, nC, nR);
            const uint32_t s1 = maj3(nL, nC, nR);
            // window: two rows up (slot PAR), one row up (slot PAR^1, centre ctr[t]) and this row
            const uint32_t v = life_rule(h0[t][PAR][j], h1[t][PAR][j], h0[t][PAR ^ 1][j], h1[t][PAR ^ 1][j], s0, s1,
                                         ctr[t][j]);
            h0[t][PAR][j] = s0;
            h1[t][PAR][j] = s1;
            if (FLAGS && t == K - 1 && flags_ok) {
                acc_a[j] |= v;
                acc_d[j] = lop3<0xF6>(acc_d[j], ctr[t][j], v);  // acc | (old ^ new)
            }
            ctr[t][j] = nC;
            out[j] = v;
        }
    };

    // WARM steps are the pipeline fill: level t only starts once its first valid input row
    // (y0-K+t) has come down the chain, i.e. at step r = rbeg + 2t.  The levels of a step form a
    // chain, so the inactive ones are a suffix and a warm step simply stops early; nothing is
    // emitted yet.  This removes K(K-1) of the 2K*K fill/drain level-steps of a chunk.
    auto step = [&](int r, auto par_tag, auto warm_tag) {
        constexpr bool WARM = decltype(warm_tag)::value;
        uint32_t x0[kWL], fin[kWL];
#pragma unroll
        for (int j = 0; j < kWL; ++j) { x0[j] = q[decltype(par_tag)::value][j]; fin[j] = 0; }
        if (r + 2 < rload_end) load_row(q[decltype(par_tag)::value]);
        const bool emit = !WARM && r - LAG >= A.y0;
#if GOL_SKEW
        // descending: level t+1 has already consumed pend[t+1] when level t overwrites it
        auto run = [&](auto self, auto t_tag) -> void {
            constexpr int t = decltype(t_tag)::value;
            uint32_t o[kWL];
            if constexpr (t == 0) level(t_tag, par_tag, x0, o, emit);
            else level(t_tag, par_tag, pend[t], o, emit);
#pragma unroll
            for (int j = 0; j < kWL; ++j) {
                if constexpr (t == K - 1) fin[j] = o[j];
                else pend[t + 1][j] = o[j];
            }
            if constexpr (t > 0) self(self, std::integral_constant<int, t - 1>{});
        };
        run(run, std::integral_constant<int, K - 1>{});
#else
        auto run = [&](auto self, auto t_tag, const uint32_t(&x)[kWL]) -> void {
            constexpr int t = decltype(t_tag)::value;
            uint32_t o[kWL];
            level(t_tag, par_tag, x, o, emit);
            if constexpr (t == K - 1) {
#pragma unroll
                for (int j = 0; j < kWL; ++j) fin[j] = o[j];
            } else {
                if (!WARM || r >= rbeg + 2 * (t + 1)) self(self, std::integral_constant<int, t + 1>{}, o);
            }
        };
        run(run, std::integral_constant<int, 0>{}, x0);
#endif
        if constexpr (!WARM) {
#if GOL_ONE_SIDED
            {   // the K levels moved the picture K bits to the left: shift it back
                const uint32_t up = __shfl_down_sync(0xffffffffu, fin[0], 1);
#pragma unroll
                for (int j = 0; j < kWL; ++j) fin[j] = __funnelshift_r(fin[j], j < kWL - 1 ? fin[j + 1] : up, K);
            }
#endif
            if (emit && DST_RING) {  // hand the row (all 64 words, aprons included) to the next stage
                while (n_out - seen_tail >= kRingRows) seen_tail = *A.rout.tail;
                *reinterpret_cast<uint2 *>(A.rout.data + (n_out % kRingRows) * 64 + A.lane * kWL) = make_uint2(fin[0], fin[1]);
                ++n_out;
                __threadfence_block();
                __syncwarp();
                if (A.lane == 0) *A.rout.head = n_out;
            } else if (emit) {  // fin = row (r - LAG) of generation g+K
                if (FAST) {
                    if (store_all) stg_v2(pout, fin[0], fin[1]);
                } else if (store_any) {
                    if (store_all) {
                        stg_v2(pout, fin[0], fin[1]);
                    } else {
#pragma unroll
                        for (int j = 0; j < kWL; ++j)
                            if (A.keep[j]) stg_u32(pout + j, fin[j] & A.keep[j]);
                    }
                }
                pout += pitch;
            }
        }
    };

    int r = rbeg;
#if !GOL_SKEW && GOL_WARM_SPLIT
    // fill: 2K-2 steps (always even, so the window parity stays aligned); chunks are >= 1 row, so
    // the loop below still has at least 3 steps left
    for (; r < rbeg + 2 * K - 2; r += 2) {
        step(r, std::integral_constant<int, 0>{}, std::true_type{});
        step(r + 1, std::integral_constant<int, 1>{}, std::true_type{});
    }
#endif
    for (; r + 1 < rend; r += 2) {
        step(r, std::integral_constant<int, 0>{}, std::false_type{});
        step(r + 1, std::integral_constant<int, 1>{}, std::false_type{});
    }
    if (r < rend) step(r, std::integral_constant<int, 0>{}, std::false_type{});

    if (FLAGS) {
#if GOL_ONE_SIDED
        const uint32_t ua = __shfl_down_sync(0xffffffffu, acc_a[0], 1), ud = __shfl_down_sync(0xffffffffu, acc_d[0], 1);
#pragma unroll
        for (int j = 0; j < kWL; ++j) {
            acc_a[j] = __funnelshift_r(acc_a[j], j < kWL - 1 ? acc_a[j + 1] : ua, K);
            acc_d[j] = __funnelshift_r(acc_d[j], j < kWL - 1 ? acc_d[j + 1] : ud, K);
        }
#endif
        flag_alive = (acc_a[0] & A.keep[0]) | (acc_a[1] & A.keep[1]);
        flag_diff = (acc_d[0] & A.keep[0]) | (acc_d[1] & A.keep[1]);
    }
}

#ifndef GOL_K8_MINBLOCKS
#define GOL_K8_MINBLOCKS 3   // keeps the K=8 kernels (flagged variant included) at 3 CTAs = 12 warps per SM
#endif

struct ItemCtx {
    int in_idx, out_idx;
    uint32_t seq, wait_seq;
    int signal, chunk_rows;
    uint32_t skip_mask;   // ghost flags not to wait for (directions covered by item-level waits)
};

// One work item: (optionally) wait for the neighbours' ghost pushes, stream the chunk through K
// levels, accumulate flags, push the borders and count the item towards the epoch flag.
template <int K, bool FLAGS>
__device__ __forceinline__ void run_item(const GolDomainDev &D, const ItemCtx &c, const ItemMap &im, int strip,
                                         int chunk, bool border, int lane)
{
    const int S = im.S, H = D.H;
    const int y0 = chunk * c.chunk_rows;
    const int y1 = min(H, y0 + c.chunk_rows);
    MarchArgs A;
    A.nw = D.nw32; A.nwf = D.nwf32; A.rem = D.rem; A.pitch = D.pitch32;
    A.in = reinterpret_cast<const uint32_t *>(__cvta_generic_to_global(D.buf[c.in_idx]));
    A.out = reinterpret_cast<uint32_t *>(__cvta_generic_to_global(D.buf[c.out_idx]));
    A.vi = strip_origin(strip, S, A.nw) - kWL + lane * kWL;
    A.y0 = y0; A.y1 = y1;
    A.lane = lane;
    A.two = D.mul2; A.four = D.mul4;
    if (c.wait_seq && border) wait_neighbors(D, c.wait_seq, lane, c.skip_mask);

    // per-lane constants
    const bool owned_lane = lane >= 1 && lane <= kValidLanes;
#pragma unroll
    for (int j = 0; j < kWL; ++j) {
        const int i = A.vi + j;
        uint32_t kmask = (owned_lane && i >= 0 && i < A.nw) ? 0xffffffffu : 0u;
        if (A.rem) {
            if (i == A.nwf) kmask &= A.rem >= 32 ? 0xffffffffu : ((1u << A.rem) - 1u);
            if (i == A.nwf + 1) kmask &= A.rem > 32 ? ((1u << (A.rem - 32)) - 1u) : 0u;
        }
        A.keep[j] = kmask;
    }
    // load mode: 1 = plain vector load, 2 = composed tail (partial word + shifted ghost), 0 = none
    if (A.rem == 0) A.mode = (A.vi >= -GOL_GHOST_WORDS32 && A.vi + kWL <= A.nw + GOL_GHOST_WORDS32) ? 1 : 0;
    else A.mode = (A.vi >= -GOL_GHOST_WORDS32 && A.vi + kWL <= A.nwf) ? 1 : (A.vi < A.nwf + 4 && A.vi + kWL > A.nwf ? 2 : 0);
    const bool all_or_none = (A.keep[0] & A.keep[1]) == 0xffffffffu || (A.keep[0] | A.keep[1]) == 0u;
    const bool fast = __all_sync(0xffffffffu, A.mode == 1 && all_or_none);

    uint32_t fa = 0, fd = 0;
    if (fast) march<K, FLAGS, true>(A, fa, fd);
    else march<K, FLAGS, false>(A, fa, fd);

    if (FLAGS) {
        const bool a = __any_sync(0xffffffffu, fa != 0);
        const bool d = __any_sync(0xffffffffu, fd != 0);
        if (lane == 0) {
            if (a) atomicOr(&D.scratch->alive, 1u);
            if (d) atomicOr(&D.scratch->diff, 1u);
        }
    }
    __syncwarp();
    if (border) {
        push_borders(D, c.out_idx, strip, y0, y1, lane, c.signal >> 8);
        if (c.signal & 1) {
            GolLaunch lp;
            lp.seq = c.seq;
            border_item_done(D, lp, im.nborder, lane);
        }
    }
}

template <int K, bool FLAGS>
__global__ void
#ifdef GOL_MAXNREG
__maxnreg__(GOL_MAXNREG)
#else
__launch_bounds__(kWarps * 32, K == 8 ? GOL_K8_MINBLOCKS : (K == 6 ? 4 : 1))
#endif
life_strip_kernel(const GolDomainDev *__restrict__ doms, const GolLaunch lp)
{
    const int dom_id = blockIdx.x / lp.ctas_per_domain;
    const int cta = blockIdx.x - dom_id * lp.ctas_per_domain;
    const GolDomainDev &D = doms[dom_id];
    const int lane = threadIdx.x & 31;
    const int item = cta * kWarps + (threadIdx.x >> 5);
    const ItemMap im = item_map(D, lp.chunk_rows);
    int strip = 0, chunk = 0;
    bool border = false;
    // Everything above reads only the launch parameters and the domain descriptor (constant after
    // upload).  Let the next launch become resident, then wait for the previous one to complete
    // before touching any grid buffer, flag or scratch word (no-ops without the PDL attribute).
    asm volatile("griddepcontrol.launch_dependents;");
    const bool have_item = decode_item(im, item, strip, chunk, border);
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (have_item) {
        ItemCtx c;
        c.in_idx = lp.in_idx; c.out_idx = lp.out_idx; c.seq = lp.seq; c.wait_seq = lp.wait_seq;
        c.signal = lp.signal; c.chunk_rows = lp.chunk_rows; c.skip_mask = 0;
        run_item<K, FLAGS>(D, c, im, strip, chunk, border, lane);
    }
    if (lp.rec_slot >= 0) finish_domain(D, lp);
}

// ---------------------------------------------------------- pipeline kernel
// One CTA per work item, D warps per CTA: warp s runs generations s*LV .. (s+1)*LV-1 of the item
// and hands each finished row to warp s+1 through a shared-memory row FIFO, so an item advances
// K = LV*D generations per pass while every warp only keeps LV levels in registers.  Compared
// with one warp doing all K levels this (a) lets K grow without the register file limiting
// occupancy and (b) makes a chunk D times taller for the same number of resident warps, which
// divides the relative cost of the pipeline fill ((K+1) rows per chunk) by D.
template <int LV, int D, bool FLAGS>
__global__ void __launch_bounds__(32 * D) life_pipe_kernel(const GolDomainDev *__restrict__ doms, const GolLaunch lp)
{
    __shared__ __align__(16) uint32_t ring[D > 1 ? D - 1 : 1][kRingRows * 64];
    __shared__ int heads[D], tails[D];
    if (threadIdx.x < D) { heads[threadIdx.x] = 0; tails[threadIdx.x] = 0; }
    __syncthreads();
    const int dom_id = blockIdx.x / lp.ctas_per_domain;
    const int item = blockIdx.x - dom_id * lp.ctas_per_domain;
    const GolDomainDev &Dm = doms[dom_id];
    const int lane = threadIdx.x & 31, stage = threadIdx.x >> 5;
    const ItemMap im = item_map(Dm, lp.chunk_rows);
    int strip = 0, chunk = 0;
    bool border = false;
    if (decode_item(im, item, strip, chunk, border)) {
        const int S = im.S, H = Dm.H;
        const int y0 = chunk * lp.chunk_rows, y1 = min(H, y0 + lp.chunk_rows);
        MarchArgs A;
        A.nw = Dm.nw32; A.nwf = Dm.nwf32; A.rem = Dm.rem; A.pitch = Dm.pitch32;
        A.in = reinterpret_cast<const uint32_t *>(__cvta_generic_to_global(Dm.buf[lp.in_idx]));
        A.out = reinterpret_cast<uint32_t *>(__cvta_generic_to_global(Dm.buf[lp.out_idx]));
        A.vi = strip_origin(strip, S, A.nw) - kWL + lane * kWL;
        A.lane = lane;
        A.two = Dm.mul2; A.four = Dm.mul4;
        // stage s produces generation (s+1)*LV for the rows the later stages still need
        A.y0 = y0 - (D - 1 - stage) * LV;
        A.y1 = y1 + (D - 1 - stage) * LV;
        if (stage > 0) { A.rin.data = ring[stage - 1]; A.rin.head = &heads[stage - 1]; A.rin.tail = &tails[stage - 1]; }
        if (stage < D - 1) { A.rout.data = ring[stage]; A.rout.head = &heads[stage]; A.rout.tail = &tails[stage]; }
        if (stage == 0 && lp.wait_seq && border) wait_neighbors(Dm, lp.wait_seq, lane);

        const bool owned_lane = lane >= 1 && lane <= kValidLanes;
#pragma unroll
        for (int j = 0; j < kWL; ++j) {
            const int i = A.vi + j;
            uint32_t kmask = (owned_lane && i >= 0 && i < A.nw) ? 0xffffffffu : 0u;
            if (A.rem) {
                if (i == A.nwf) kmask &= A.rem >= 32 ? 0xffffffffu : ((1u << A.rem) - 1u);
                if (i == A.nwf + 1) kmask &= A.rem > 32 ? ((1u << (A.rem - 32)) - 1u) : 0u;
            }
            A.keep[j] = kmask;
        }
        if (A.rem == 0) A.mode = (A.vi >= -GOL_GHOST_WORDS32 && A.vi + kWL <= A.nw + GOL_GHOST_WORDS32) ? 1 : 0;
        else A.mode = (A.vi >= -GOL_GHOST_WORDS32 && A.vi + kWL <= A.nwf) ? 1 : (A.vi < A.nwf + 4 && A.vi + kWL > A.nwf ? 2 : 0);
        const bool all_or_none = (A.keep[0] & A.keep[1]) == 0xffffffffu || (A.keep[0] | A.keep[1]) == 0u;
        const bool fast = __all_sync(0xffffffffu, A.mode == 1 && all_or_none);

        uint32_t fa = 0, fd = 0;
        if (D == 1) {
            if (fast) march<LV, FLAGS, true>(A, fa, fd); else march<LV, FLAGS, false>(A, fa, fd);
        } else if (stage == 0) {
            if (fast) march<LV, false, true, false, true>(A, fa, fd); else march<LV, false, false, false, true>(A, fa, fd);
        } else if (stage == D - 1) {
            if (fast) march<LV, FLAGS, true, true, false>(A, fa, fd); else march<LV, FLAGS, false, true, false>(A, fa, fd);
        } else {
            march<LV, false, true, true, true>(A, fa, fd);  // ring to ring: no global access, no special lanes
        }
        if (stage == D - 1) {
            if (FLAGS) {
                const bool a = __any_sync(0xffffffffu, fa != 0);
                const bool d = __any_sync(0xffffffffu, fd != 0);
                if (lane == 0) {
                    if (a) atomicOr(&Dm.scratch->alive, 1u);
                    if (d) atomicOr(&Dm.scratch->diff, 1u);
                }
            }
            __syncwarp();
            if (border) {
                push_borders(Dm, lp.out_idx, strip, y0, y1, lane, lp.signal >> 8);
                if (lp.signal & 1) border_item_done(Dm, lp, im.nborder, lane);
            }
        }
    }
    if (lp.rec_slot >= 0) finish_domain(Dm, lp);
}

// ------------------------------------------------------- persistent kernel
__device__ __forceinline__ int encode_item(const ItemMap &m, int strip, int chunk)
{
    const int S = m.S, n1 = m.top * S, n2 = (m.chunks - m.first_bot) * S, mid = m.first_bot - m.top;
    const int e = min(S, 2), n3 = mid * e;
    if (chunk < m.top) return chunk * S + strip;
    if (chunk >= m.first_bot) return n1 + (chunk - m.first_bot) * S + strip;
    const int c = chunk - m.top;
    if (strip == 0) return n1 + n2 + c * e;
    if (strip == S - 1) return n1 + n2 + c * e + (e - 1);
    return n1 + n2 + n3 + c * (S - 2) + (strip - 1);
}

__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t *p)
{
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// Wait until every item of this domain whose previous-epoch output this item reads has
// completed `want`: strips s-1..s+1 x the chunks covering rows [y0-K, y1+K).  Directions in
// which the domain is its own neighbour wrap around (single-domain torus); the others are
// covered by the ghost-zone flags of the neighbouring domains.
template <int K>
__device__ __forceinline__ void wait_items(const GolDomainDev &D, const ItemMap &im, int strip, int chunk,
                                           int chunk_rows, uint32_t want, int lane)
{
    const bool wrap_y = (D.self_mask & 3u) == 3u, wrap_x = (D.self_mask & 12u) == 12u;
    if (lane < 18) {
        const int H = D.H, y0 = chunk * chunk_rows, y1 = min(H, y0 + chunk_rows);
        const int ci = lane % 6;
        int c = -1;
        if (ci < 3) {  // chunks inside the domain that intersect [y0-K, y1+K)
            const int ca = max(0, y0 - K) / chunk_rows, cb = min(H - 1, y1 + K - 1) / chunk_rows;
            if (ca + ci <= cb) c = ca + ci;
        } else if (ci < 5) {  // rows above row 0 are the torus' last rows: up to two trailing chunks
            if (wrap_y && y0 - K < 0) {
                const int cw = (H - (K - y0)) / chunk_rows + (ci - 3);
                if (cw < im.chunks) c = cw;
            }
        } else {  // rows below the last row are the torus' first rows: chunk 0 (chunk_rows >= K)
            if (wrap_y && y1 + K > H) c = 0;
        }
        int s = strip + lane / 6 - 1;
        if (wrap_x) s = (s + im.S) % im.S;
        else if (s < 0 || s >= im.S) c = -1;
        if (c >= 0) {
            const uint32_t *f = D.item_flags + encode_item(im, s, c);
            long long t0 = clock64();
            while ((int32_t)(ld_acquire_gpu(f) - want) < 0) {
                if (clock64() - t0 > (1ll << 32) || *(volatile unsigned int *)&D.scratch->error) {
                    atomicOr(&D.scratch->error, 4u);
                    break;
                }
                __nanosleep(100);
            }
        }
    }
    __syncwarp();
}

// All epochs of a batch in one launch.  Warps draw (epoch, item) pairs from one atomic counter
// in epoch-major order; an item starts as soon as its own dependencies are met, so there is no
// grid-wide barrier, no launch gap and no per-epoch tail, and faster SMs simply take more items.
// Deadlock-free: the handed-out items are always a prefix of the global order, all of them are
// held by resident warps (cooperative launch), and an item only waits for smaller indices.
template <int K>
__global__ void __launch_bounds__(kWarps * 32, K >= 8 ? 3 : 1) life_persistent_kernel(const GolDomainDev *__restrict__ doms,
                                                                      const GolPersist P)
{
    const int lane = threadIdx.x & 31;
    const unsigned items_total = (unsigned)(P.ndoms * P.items_per_domain);
    const unsigned total = items_total * (unsigned)P.epochs;
    for (;;) {
        // dynamic hand-out in global (epoch-major, border-first) order: fast SMs take more items
        unsigned g = 0;
        if (lane == 0) g = atomicAdd(P.work_counter, 1u);
        g = __shfl_sync(0xffffffffu, g, 0);
        if (g >= total) break;
        const int e = (int)(g / items_total);
        const int i = (int)(g - (unsigned)e * items_total);
        const int dom = i / P.items_per_domain, item = i - dom * P.items_per_domain;
        const GolDomainDev &D = doms[dom];
        const ItemMap im = item_map(D, P.chunk_rows);
        int strip = 0, chunk = 0;
        bool border = false;
        if (!decode_item(im, item, strip, chunk, border)) continue;
        ItemCtx c;
        c.in_idx = e == 0 ? P.first_in : P.rot[(e - 1) % 3];
        c.out_idx = P.rot[e % 3];
        c.seq = P.base_seq + 1 + (uint32_t)e;
        c.wait_seq = e == 0 ? P.wait_seq0 : P.base_seq + (uint32_t)e;
        c.signal = 1;
        c.chunk_rows = P.chunk_rows;
        // inside the launch, self-directions are tracked item by item; ghost flags are only
        // awaited from directions where the neighbour is another domain
        c.skip_mask = e == 0 ? 0u : D.self_mask;
        if (e > 0) wait_items<K>(D, im, strip, chunk, P.chunk_rows, P.base_seq + (uint32_t)e, lane);
        run_item<K, false>(D, c, im, strip, chunk, border, lane);
        __threadfence();
        __syncwarp();
        if (lane == 0)
            asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(D.item_flags + item), "r"(c.seq) : "memory");
    }
}

// ------------------------------------------------------------- halo kernel
__global__ void __launch_bounds__(kWarps * 32) halo_push_kernel(const GolDomainDev *__restrict__ doms,
                                                                const GolLaunch lp)
{
    const int dom_id = blockIdx.x / lp.ctas_per_domain;
    const int cta = blockIdx.x - dom_id * lp.ctas_per_domain;
    const GolDomainDev &D = doms[dom_id];
    const int lane = threadIdx.x & 31;
    const int item = cta * kWarps + (threadIdx.x >> 5);
    const ItemMap im = item_map(D, lp.chunk_rows);
    int strip = 0, chunk = 0;
    bool border = false;
    if (!decode_item(im, item, strip, chunk, border) || !border) return;
    const int y0 = chunk * lp.chunk_rows;
    const int y1 = min(D.H, y0 + lp.chunk_rows);
    push_borders(D, lp.out_idx, strip, y0, y1, lane);
    if (lp.signal) border_item_done(D, lp, im.nborder, lane);
}

// ------------------------------------------------------------- alive scan
// empty() of the current generation (S:103-116, C:55-81) with the same record protocol.
__global__ void __launch_bounds__(kWarps * 32) scan_alive_kernel(const GolDomainDev *__restrict__ doms,
                                                                 const GolLaunch lp)
{
    const int dom_id = blockIdx.x / lp.ctas_per_domain;
    const int cta = blockIdx.x - dom_id * lp.ctas_per_domain;
    const GolDomainDev &D = doms[dom_id];
    const int lane = threadIdx.x & 31;
    const int item = cta * kWarps + (threadIdx.x >> 5);
    const ItemMap im = item_map(D, lp.chunk_rows);
    const int S = im.S;
    int strip = 0, chunk = 0;
    bool border = false;
    if (decode_item(im, item, strip, chunk, border)) {
        const int y0 = chunk * lp.chunk_rows;
        const int y1 = min(D.H, y0 + lp.chunk_rows);
        const int lo = strip_origin(strip, S, D.nw32), hi = min(D.nw32, lo + kVW);
        const uint32_t *src = D.buf[lp.in_idx];
        uint32_t acc = 0;
        for (int y = y0; y < y1; ++y) {
            const uint32_t *s = src + (int64_t)(y + G) * D.pitch32 + GOL_GHOST_WORDS32;
            for (int i = lo + lane; i < hi; i += 32) acc |= s[i];
        }
        if (__any_sync(0xffffffffu, acc != 0) && lane == 0) atomicOr(&D.scratch->alive, 1u);
    }
    finish_domain(D, lp);
}

// ------------------------------------------------- flag all-reduce (gather)
// One warp: waits until every domain's record for `seq` is in this process's table, ORs them.
__global__ void gather_flags_kernel(const unsigned long long *table, int num_domains, uint32_t seq,
                                    unsigned long long *result, unsigned int *error)
{
    unsigned long long acc = 0;
    for (int i = threadIdx.x; i < num_domains; i += blockDim.x) {
        const unsigned long long *p = table + i;
        long long t0 = clock64();
        unsigned long long v;
        while ((uint32_t)((v = ld_acquire_sys64(p)) >> 32) != seq) {
            if (clock64() - t0 > (1ll << 34)) {
                atomicOr(error, 2u);
                break;
            }
            __nanosleep(128);
        }
        acc |= v & 3ull;
    }
    for (int o = 16; o; o >>= 1) acc |= __shfl_xor_sync(0xffffffffu, acc, o);
    __shared__ unsigned long long sh[32];
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long r = 0;
        for (int i = 0; i < (int)(blockDim.x + 31) / 32; ++i) r |= sh[i];
        *result = ((unsigned long long)seq << 32) | r;
    }
}

// ------------------------------------------------ dense <-> padded copies
__global__ void dense_to_padded_kernel(const uint64_t *__restrict__ dense, int64_t dense_pitch_q, int64_t qx0,
                                       uint32_t *__restrict__ pad, int H, int nq, int rem, int pitch32)
{
    const int q = blockIdx.y * blockDim.x + threadIdx.x;
    const int y = blockIdx.x;
    if (q >= nq || y >= H) return;
    uint64_t v = dense[(int64_t)y * dense_pitch_q + qx0 + q];
    if (rem && q == nq - 1) v &= (1ull << rem) - 1ull;
    uint64_t *row = reinterpret_cast<uint64_t *>(pad + (int64_t)(y + G) * pitch32 + GOL_GHOST_WORDS32);
    row[q] = v;
}
__global__ void padded_to_dense_kernel(const uint32_t *__restrict__ pad, int H, int nq, int rem, int pitch32,
                                       uint64_t *__restrict__ dense, int64_t dense_pitch_q, int64_t qx0)
{
    const int q = blockIdx.y * blockDim.x + threadIdx.x;
    const int y = blockIdx.x;
    if (q >= nq || y >= H) return;
    const uint64_t *row = reinterpret_cast<const uint64_t *>(pad + (int64_t)(y + G) * pitch32 + GOL_GHOST_WORDS32);
    uint64_t v = row[q];
    if (rem && q == nq - 1) v &= (1ull << rem) - 1ull;
    dense[(int64_t)y * dense_pitch_q + qx0 + q] = v;
}

// live cells of a padded domain buffer (interior only), added to *count
__global__ void popcount_padded_kernel(const uint32_t *__restrict__ pad, int H, int nq, int pitch32,
                                       unsigned long long *count)
{
    unsigned long long c = 0;
    const int64_t total = (int64_t)H * nq;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t y = i / nq, q = i - y * nq;
        c += __popcll(reinterpret_cast<const unsigned long long *>(pad + (y + G) * pitch32 + GOL_GHOST_WORDS32)[q]);
    }
    for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(count, c);
}

template <int K, bool FLAGS>
cudaError_t launch_strip_t(const GolDomainDev *d, int grid, const GolLaunch &lp, cudaStream_t s)
{
    // Programmatic dependent launch: the next epoch's CTAs become resident (and run their
    // prologue) while this epoch drains; they block in griddepcontrol.wait until it has completed.
    static const int pdl = [] { const char *v = getenv("GOL_PDL"); return v ? atoi(v) : 1; }();  // 2: also across processes (experiment)
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(kWarps * 32);
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    // single-process runs only (lp.pdl): with halo flags between GPUs the early-resident CTAs cost more
    // than the launch gap they hide (2 GPUs, 16384^2 each: 9.06 us/generation without, 11.1 with)
    cfg.numAttrs = ((pdl && lp.pdl) || pdl == 2) ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, life_strip_kernel<K, FLAGS>, d, lp);
}

}  // namespace

int gol_strip_geometry(int nw32, int *strips)
{
    *strips = (int)gol_div_up(nw32, kVW);
    return 0;
}

int gol_launch_strip(const GolDomainDev *d_doms, int ndoms, const GolLaunch &lp, int k, int want_flags,
                     CUstream_st *stream)
{
    const int grid = ndoms * lp.ctas_per_domain;
    cudaError_t err;
#define GOL_CASE(KK)                                                                  \
    case KK:                                                                          \
        err = want_flags ? launch_strip_t<KK, true>(d_doms, grid, lp, stream)         \
                         : launch_strip_t<KK, false>(d_doms, grid, lp, stream);       \
        break;
    switch (k) {
        GOL_CASE(1)
        GOL_CASE(2)
        GOL_CASE(3)
        GOL_CASE(4)
        GOL_CASE(5)
        GOL_CASE(6)
        GOL_CASE(7)
        GOL_CASE(8)
        GOL_CASE(16)
    default:
        GOL_FAIL("unsupported temporal depth k=%d", k);
    }
#undef GOL_CASE
    if (err != cudaSuccess) GOL_FAIL("life_strip_kernel<k=%d> launch failed: %s", k, cudaGetErrorString(err));
    return 0;
}

template <int K>
static cudaError_t persistent_occupancy(int *ctas_per_sm)
{
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(ctas_per_sm, life_persistent_kernel<K>, kWarps * 32, 0);
}

template <int K>
static cudaError_t strip_occupancy(int *ctas_per_sm)
{
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(ctas_per_sm, life_strip_kernel<K, false>, kWarps * 32, 0);
}

// resident CTAs per SM of the (unflagged) strip kernel for depth k
int gol_strip_ctas_per_sm(int k, int *ctas_per_sm)
{
    cudaError_t err = cudaSuccess;
    switch (k) {
    case 1: err = strip_occupancy<1>(ctas_per_sm); break;
    case 2: err = strip_occupancy<2>(ctas_per_sm); break;
    case 3: err = strip_occupancy<3>(ctas_per_sm); break;
    case 4: err = strip_occupancy<4>(ctas_per_sm); break;
    case 5: err = strip_occupancy<5>(ctas_per_sm); break;
    case 6: err = strip_occupancy<6>(ctas_per_sm); break;
    case 7: err = strip_occupancy<7>(ctas_per_sm); break;
    case 8: err = strip_occupancy<8>(ctas_per_sm); break;
    case 16: err = strip_occupancy<16>(ctas_per_sm); break;
    default: GOL_FAIL("unsupported temporal depth k=%d", k);
    }
    if (err != cudaSuccess) GOL_FAIL("occupancy query failed: %s", cudaGetErrorString(err));
    return 0;
}

int gol_persistent_max_ctas(int k, int device, int *ctas)
{
    int per_sm = 0, sms = 0;
    cudaError_t err = cudaSuccess;
    switch (k) {
    case 1: err = persistent_occupancy<1>(&per_sm); break;
    case 2: err = persistent_occupancy<2>(&per_sm); break;
    case 4: err = persistent_occupancy<4>(&per_sm); break;
    case 8: err = persistent_occupancy<8>(&per_sm); break;
    default: GOL_FAIL("persistent kernel: unsupported temporal depth k=%d", k);
    }
    if (err == cudaSuccess) err = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    if (err != cudaSuccess) GOL_FAIL("occupancy query failed: %s", cudaGetErrorString(err));
    *ctas = per_sm * sms;
    return 0;
}

int gol_launch_persistent(const GolDomainDev *d_doms, const GolPersist &pp, int k, int ctas, CUstream_st *stream)
{
    void *args[] = {(void *)&d_doms, (void *)&pp};
    const void *fn = nullptr;
    switch (k) {
    case 1: fn = (const void *)life_persistent_kernel<1>; break;
    case 2: fn = (const void *)life_persistent_kernel<2>; break;
    case 4: fn = (const void *)life_persistent_kernel<4>; break;
    case 8: fn = (const void *)life_persistent_kernel<8>; break;
    default: GOL_FAIL("persistent kernel: unsupported temporal depth k=%d", k);
    }
    // cooperative launch: the runtime guarantees that all CTAs are resident at once, which the
    // item-to-item waits rely on
    cudaError_t err = cudaLaunchCooperativeKernel(fn, dim3((unsigned)ctas), dim3(kWarps * 32), args, 0, stream);
    if (err != cudaSuccess) GOL_FAIL("life_persistent_kernel<k=%d> launch failed: %s", k, cudaGetErrorString(err));
    return 0;
}

template <int LV, int D>
static cudaError_t launch_pipe_t(const GolDomainDev *d, int grid, const GolLaunch &lp, int want_flags, cudaStream_t s)
{
    if (want_flags) life_pipe_kernel<LV, D, true><<<grid, 32 * D, 0, s>>>(d, lp);
    else life_pipe_kernel<LV, D, false><<<grid, 32 * D, 0, s>>>(d, lp);
    return cudaGetLastError();
}
template <int LV, int D>
static cudaError_t pipe_occupancy(int *ctas_per_sm)
{
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(ctas_per_sm, life_pipe_kernel<LV, D, false>, 32 * D, 0);
}

#define GOL_PIPE_CONFIGS(X) X(4, 2) X(2, 4) X(4, 4) X(2, 8) X(8, 2) X(3, 4) X(4, 3) X(5, 2) X(6, 2) X(3, 3)

int gol_launch_pipe(const GolDomainDev *d_doms, int ndoms, const GolLaunch &lp, int lv, int d, int want_flags,
                    CUstream_st *stream)
{
    const int grid = ndoms * lp.ctas_per_domain;
    cudaError_t err = cudaErrorInvalidValue;
#define GOL_X(LV, DD) if (lv == LV && d == DD) err = launch_pipe_t<LV, DD>(d_doms, grid, lp, want_flags, stream);
    GOL_PIPE_CONFIGS(GOL_X)
#undef GOL_X
    if (err != cudaSuccess) GOL_FAIL("life_pipe_kernel<%d,%d> launch failed: %s", lv, d, cudaGetErrorString(err));
    return 0;
}

int gol_pipe_ctas_per_sm(int lv, int d, int *ctas_per_sm)
{
    cudaError_t err = cudaErrorInvalidValue;
#define GOL_X(LV, DD) if (lv == LV && d == DD) err = pipe_occupancy<LV, DD>(ctas_per_sm);
    GOL_PIPE_CONFIGS(GOL_X)
#undef GOL_X
    if (err != cudaSuccess) GOL_FAIL("pipeline kernel <%d,%d>: %s", lv, d, cudaGetErrorString(err));
    return 0;
}

int gol_launch_halo(const GolDomainDev *d_doms, int ndoms, const GolLaunch &lp, CUstream_st *stream)
{
    halo_push_kernel<<<ndoms * lp.ctas_per_domain, kWarps * 32, 0, stream>>>(d_doms, lp);
    cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) GOL_FAIL("halo_push_kernel launch failed: %s", cudaGetErrorString(err));
    return 0;
}

int gol_launch_scan_alive(const GolDomainDev *d_doms, int ndoms, const GolLaunch &lp, CUstream_st *stream)
{
    scan_alive_kernel<<<ndoms * lp.ctas_per_domain, kWarps * 32, 0, stream>>>(d_doms, lp);
    cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) GOL_FAIL("scan_alive_kernel launch failed: %s", cudaGetErrorString(err));
    return 0;
}

int gol_launch_gather(unsigned long long *rec_table, int slot, int num_domains, uint32_t seq,
                      unsigned long long *d_result, unsigned int *d_error, CUstream_st *stream)
{
    gather_flags_kernel<<<1, 64, 0, stream>>>(rec_table + (int64_t)slot * num_domains, num_domains, seq,
                                              d_result, d_error);
    cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) GOL_FAIL("gather_flags_kernel launch failed: %s", cudaGetErrorString(err));
    return 0;
}

int gol_launch_popcount_padded(const uint32_t *d_pad, const gol_layout &L, unsigned long long *d_count, CUstream_st *stream)
{
    const int64_t total = (int64_t)L.height * L.nq;
    const int blocks = (int)std::min<int64_t>(148 * 8, gol_div_up(total, 256));
    popcount_padded_kernel<<<blocks, 256, 0, stream>>>(d_pad, (int)L.height, L.nq, L.pitch32, d_count);
    cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) GOL_FAIL("popcount_padded launch failed: %s", cudaGetErrorString(err));
    return 0;
}

int gol_launch_dense_to_padded(const uint64_t *d_dense, int64_t dense_pitch_q, int64_t qx0, uint32_t *d_pad,
                               const gol_layout &L, CUstream_st *stream)
{
    dim3 grid((unsigned)L.height, (unsigned)gol_div_up(L.nq, 128));
    dense_to_padded_kernel<<<grid, 128, 0, stream>>>(d_dense, dense_pitch_q, qx0, d_pad, (int)L.height, L.nq,
                                                     L.rem, L.pitch32);
    cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) GOL_FAIL("dense_to_padded launch failed: %s", cudaGetErrorString(err));
    return 0;
}

int gol_launch_padded_to_dense(const uint32_t *d_pad, const gol_layout &L, uint64_t *d_dense,
                               int64_t dense_pitch_q, int64_t qx0, int, CUstream_st *stream)
{
    dim3 grid((unsigned)L.height, (unsigned)gol_div_up(L.nq, 128));
    padded_to_dense_kernel<<<grid, 128, 0, stream>>>(d_pad, (int)L.height, L.nq, L.rem, L.pitch32, d_dense,
                                                     dense_pitch_q, qx0);
    cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) GOL_FAIL("padded_to_dense launch failed: %s", cudaGetErrorString(err));
    return 0;
}
