// Generation kernels for sm_100a.
//
//  life_strip_kernel<K, FLAGS>  — the product kernel.  One warp owns a vertical strip of the
//      padded bit grid and streams down a chunk of rows with K generations pipelined in registers
//      (csrc/life_march.cuh: 6-LOP3 B3/S23 rule, bit-sliced adder).  The same kernel pushes its
//      border rows/columns straight into the neighbours' ghost zones (its own for a single torus
//      domain, peer-mapped memory over NVLink otherwise), publishes release flags per direction
//      group, and (FLAGS) reduces the alive / changed bits of the last level.
//      Replaces evolve (S:60-101, C:27-53, O:29-57, G:128-148), halo_rows/halo_cols (G:52-74),
//      the MPI halo exchange (C:280-344), compare/similarity (G:76-100, C:83-109) and
//      empty (G:102-126, C:55-81).
//  halo_push_kernel — the border push alone (first exchange after a load).
//  scan_alive_kernel — empty() of the current generation.
//  gather_flags_kernel — flag all-reduce: waits for every domain's record of a batch.
//  dense <-> padded copies, padded popcount.
//
// The kernel variants that were measured and lost (skewed levels, IMAD shifts, warp-pipelined
// shared-memory FIFOs, persistent dataflow kernel) live in csrc/experiments/.
#include <cuda_runtime.h>

#include <algorithm>

#include "gol_internal.h"
#include "life_march.cuh"
#include "life_items.h"

namespace {

using namespace golk;

constexpr int kWarps = GOL_STRIP_WARPS;  // warps per CTA

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
__device__ __forceinline__ unsigned long long global_timer()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// ---------------------------------------------------------------- border push
// Copies the border rows / columns of rows [y0,y1) of buffer `o` owned by `strip` into the
// ghost zones of the 8 neighbours: the top / bottom `depth` rows go north / south, the first /
// last 128 cells of every row go west / east (and to the corner neighbours for the top / bottom
// GOL_GHOST_ROWS rows).  Executed by one warp; reads data this warp wrote earlier.
__device__ void push_borders(const GolDomainDev &D, int o, int strip, int S, int y0, int y1, int depth, int lane,
                             int dbg = 0)
{
    const int H = D.H, nw = D.nw32, pitch = D.pitch32;
    const uint32_t *src = D.buf[o];
    const int lo = strip_origin(strip, S, nw, 0), hi = min(nw, lo + kVW);
    const int npairs = (hi - lo) >> 1;  // uint2 elements per row owned by this strip (<= 30)

    // rows -> north / south neighbours' ghost rows.  All loads of a batch are issued before the
    // first store, so a batch costs one memory round trip instead of one per row.
    for (int pass = 0; pass < 2 && !(dbg & 2); ++pass) {
        const int ya = pass == 0 ? y0 : max(y0, H - depth);
        const int yb = pass == 0 ? min(y1, depth) : y1;
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

    const bool first = strip == 0, last = strip == S - 1;
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

// Spin until the neighbours in `dir_mask` have published sequence >= want into my epoch flags.
// Returns false when a flag did not arrive within D.wait_timeout cycles (the sticky error word is
// set; the caller abandons its item, so no flags of this domain are published any more and the
// failure propagates to every rank instead of stale ghosts being consumed).
__device__ bool wait_neighbors(const GolDomainDev &D, uint32_t want, uint32_t dir_mask, int lane, GolTrace *trace)
{
    bool ok = true;
    long long spun = 0;
    if (lane < GOL_NDIR && ((dir_mask >> lane) & 1u)) {
        const uint32_t *f = D.epoch_flags + lane;
        if ((int32_t)(ld_acquire_sys(f) - want) < 0) {
            const long long t0 = clock64();
            while ((int32_t)(ld_acquire_sys(f) - want) < 0) {
                if (clock64() - t0 > D.wait_timeout || *(volatile unsigned int *)&D.scratch->error) {
                    atomicOr(&D.scratch->error, 1u);
                    ok = false;
                    break;
                }
                __nanosleep(64);
            }
            spun = clock64() - t0;
        }
    }
    ok = __all_sync(0xffffffffu, ok);
    if (trace) {
        for (int o = 4; o; o >>= 1) spun = max(spun, __shfl_xor_sync(0xffffffffu, spun, o));
        if (lane == 0 && spun > 0) {
            atomicAdd(&trace->spin_sum, (unsigned long long)spun);
            atomicMax(&trace->spin_max, (unsigned long long)spun);
        }
    }
    __threadfence();  // acquire for the whole warp: drop stale L1 lines of the ghost zones
    return ok;
}

// Called by every border warp after its pushes.  Two direction groups are counted separately:
// the edge-chunk items carry everything that goes N / S and to the corners, the first/last-strip
// items everything that goes W / E; the warp that completes a group publishes that group's flags.
// Release pattern with one system-scope fence per group instead of one per border warp: every
// border warp orders its pushes before its count at GPU scope; the completing warp issues the
// (cumulative) system-scope fence before it publishes the flags, which carries all counted
// warps' pushes with it (PTX memory model: causality order is transitive over the gpu-scope
// atomic and the sys-scope release).
__device__ __forceinline__ void border_item_done(const GolDomainDev &D, uint32_t seq, const ItemMap &m, const Item &it,
                                                 int lane)
{
    __threadfence();
    __syncwarp();
    if (lane != 0) return;
    uint32_t publish = 0;
    if (it.top || it.bot) {
        if (atomicAdd(&D.scratch->cnt_ns, 1u) == (unsigned)m.nA - 1) { D.scratch->cnt_ns = 0; publish |= GOL_DIRMASK_NS; }
    }
    if (it.first || it.last) {
        if (atomicAdd(&D.scratch->cnt_we, 1u) == (unsigned)((m.nedge + m.ci) * m.e) - 1) { D.scratch->cnt_we = 0; publish |= GOL_DIRMASK_WE; }
    }
    if (publish) {
        __threadfence_system();
        for (int d = 0; d < GOL_NDIR; ++d)
            if ((publish >> d) & 1u) st_release_sys(D.nb[d].flag, seq);
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
#ifndef GOL_K8_MINBLOCKS
#define GOL_K8_MINBLOCKS 3   // keeps the K=8 kernels (flagged variant included) at 3 CTAs = 12 warps per SM
#endif

// One work item: (optionally) wait for the neighbours' ghost pushes, stream the chunk through K
// levels, accumulate flags, push the borders and count the item towards the epoch flags.
template <int K, bool FLAGS>
__device__ __forceinline__ void run_item(const GolDomainDev &D, const GolLaunch &lp, const ItemMap &im, const Item &it,
                                         int lane)
{
    const int S = im.S, xe = lp.ext_x;
    const bool border = it.top || it.bot || it.first || it.last;
    MarchArgs A;
    A.nw = D.nw32; A.nwf = D.nwf32; A.rem = D.rem; A.pitch = D.pitch32;
    A.in = reinterpret_cast<const uint32_t *>(__cvta_generic_to_global(D.buf[lp.in_idx]));
    A.out = reinterpret_cast<uint32_t *>(__cvta_generic_to_global(D.buf[lp.out_idx]));
    A.vi = strip_origin(it.strip, S, A.nw, xe) - kWL + lane * kWL;
    A.y0 = it.y0; A.y1 = it.y1;
    A.lane = lane;
    if (lp.wait_seq && border && !wait_neighbors(D, lp.wait_seq, item_dir_mask(it), lane, lp.trace)) return;

    // per-lane constants
    const bool owned_lane = lane >= 1 && lane <= kValidLanes;
#pragma unroll
    for (int j = 0; j < kWL; ++j) {
        const int i = A.vi + j;
        uint32_t kmask = (owned_lane && i >= -xe && i < A.nw + xe) ? 0xffffffffu : 0u;
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
    GlobalRowSource src;
    if (fast) march<K, FLAGS, true>(A, src, fa, fd);
    else march<K, FLAGS, false>(A, src, fa, fd);

    if (FLAGS) {
        const bool a = __any_sync(0xffffffffu, fa != 0);
        const bool d = __any_sync(0xffffffffu, fd != 0);
        if (lane == 0) {
            if (a) atomicOr(&D.scratch->alive, 1u);
            if (d) atomicOr(&D.scratch->diff, 1u);
        }
    }
    __syncwarp();
    if (border && (lp.push & 1)) {
        push_borders(D, lp.out_idx, it.strip, S, it.y0, it.y1, lp.push_rows, lane, lp.push >> 8);
        if (lp.push & 2) border_item_done(D, lp.seq, im, it, lane);
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
    const ItemMap im = item_map(D.nw32, D.H, lp);
    Item it;
    // Everything above reads only the launch parameters and the domain descriptor (constant after
    // upload).  Let the next launch become resident, then wait for the previous one to complete
    // before touching any grid buffer, flag or scratch word (no-ops without the PDL attribute).
    asm volatile("griddepcontrol.launch_dependents;");
    const bool have_item = decode_item(im, item, lp.border_last != 0, it);
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (lp.trace && threadIdx.x == 0) atomicMin(&lp.trace->t0_min, global_timer());
    if (have_item) run_item<K, FLAGS>(D, lp, im, it, lane);
    if (lp.rec_slot >= 0) finish_domain(D, lp);
    if (lp.trace) {
        __syncthreads();
        if (threadIdx.x == 0) atomicMax(&lp.trace->t1_max, global_timer());
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
    const ItemMap im = item_map(D.nw32, D.H, lp);
    Item it;
    if (!decode_item(im, item, false, it) || !(it.top || it.bot || it.first || it.last)) return;
    push_borders(D, lp.out_idx, it.strip, im.S, it.y0, it.y1, lp.push_rows, lane);
    if (lp.push & 2) border_item_done(D, lp.seq, im, it, lane);
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
    const ItemMap im = item_map(D.nw32, D.H, lp);
    Item it;
    if (decode_item(im, item, false, it)) {
        const int lo = strip_origin(it.strip, im.S, D.nw32, 0), hi = min(D.nw32, lo + kVW);
        const uint32_t *src = D.buf[lp.in_idx];
        uint32_t acc = 0;
        for (int y = it.y0; y < it.y1; ++y) {
            const uint32_t *s = src + (int64_t)(y + G) * D.pitch32 + GOL_GHOST_WORDS32;
            for (int i = lo + lane; i < hi; i += 32) acc |= s[i];
            // the scan only has to find ONE live cell: a warp stops at the first rows that hold any
            // (a full scan of 16384^2 costs 78 us at the head of every run; a grid that is not empty
            // answers after four rows)
            if (((y - it.y0) & 3) == 3 && __any_sync(0xffffffffu, acc != 0)) break;
        }
        if (__any_sync(0xffffffffu, acc != 0) && lane == 0) atomicOr(&D.scratch->alive, 1u);
    }
    finish_domain(D, lp);
}

// ------------------------------------------------- flag all-reduce (gather)
// One CTA: waits until every domain's record for `seq` is in this process's table, ORs them.
__global__ void gather_flags_kernel(const unsigned long long *table, int num_domains, uint32_t seq,
                                    unsigned long long *result, unsigned int *error, long long timeout)
{
    unsigned long long acc = 0;
    for (int i = threadIdx.x; i < num_domains; i += blockDim.x) {
        const unsigned long long *p = table + i;
        long long t0 = clock64();
        unsigned long long v;
        while ((uint32_t)((v = ld_acquire_sys64(p)) >> 32) != seq) {
            if (clock64() - t0 > timeout) {
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

// Position-dependent 64-bit checksum of a padded domain buffer (interior only), added to *sum:
// the sum over all 64-bit words of hash(word value, global word x, global row y).  Addition
// commutes, so the checksum of a grid does not depend on how it is cut into domains.
__global__ void hash_padded_kernel(const uint32_t *__restrict__ pad, int H, int nq, int pitch32, int64_t qx0, int64_t y0,
                                   unsigned long long *sum)
{
    unsigned long long c = 0;
    const int64_t total = (int64_t)H * nq;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t y = i / nq, q = i - y * nq;
        const unsigned long long v = reinterpret_cast<const unsigned long long *>(pad + (y + G) * pitch32 + GOL_GHOST_WORDS32)[q];
        c += gol_hash64(v, (uint64_t)(qx0 + q), (uint64_t)(y0 + y));
    }
    for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0) atomicAdd(sum, c);
}

template <int K, bool FLAGS>
cudaError_t launch_strip_t(const GolDomainDev *d, int grid, const GolLaunch &lp, cudaStream_t s)
{
    // Programmatic dependent launch: the next epoch's CTAs become resident (and run their
    // prologue) while this epoch drains; they block in griddepcontrol.wait until it has completed.
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(kWarps * 32);
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = lp.pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, life_strip_kernel<K, FLAGS>, d, lp);
}

}  // namespace

int gol_strip_geometry(int nw32, int ext_x, int *strips)
{
    *strips = strip_count(nw32, ext_x);
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
                      unsigned long long *d_result, unsigned int *d_error, long long timeout_cycles,
                      CUstream_st *stream)
{
    gather_flags_kernel<<<1, 64, 0, stream>>>(rec_table + (int64_t)slot * num_domains, num_domains, seq,
                                              d_result, d_error, timeout_cycles);
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

int gol_launch_hash_padded(const uint32_t *d_pad, const gol_layout &L, int64_t qx0, int64_t y0, unsigned long long *d_sum,
                           CUstream_st *stream)
{
    const int64_t total = (int64_t)L.height * L.nq;
    const int blocks = (int)std::min<int64_t>(148 * 8, gol_div_up(total, 256));
    hash_padded_kernel<<<blocks, 256, 0, stream>>>(d_pad, (int)L.height, L.nq, L.pitch32, qx0, y0, d_sum);
    cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) GOL_FAIL("hash_padded launch failed: %s", cudaGetErrorString(err));
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
