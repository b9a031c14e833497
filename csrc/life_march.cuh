// Register-pipelined strip march: the arithmetic core of the generation kernel.
//
// One warp owns a vertical strip of the padded bit grid (32 lanes x 64 cells) and streams down
// the rows of a chunk.  K generations are pipelined in registers (temporal blocking): level t
// consumes the row level t-1 produced in the same step, so a row is read from memory once and
// written once per K generations.  Neighbour words move between lanes with warp shuffles; per 32
// cells and generation the work is 2 funnel shifts + 2 LOP3 (3-cell row sum, shared by three
// output rows) + 6 LOP3 (vertical add + B3/S23 rule).
//
// Replaces evolve (S:60-101, C:27-53, O:29-57, G:128-148) of the reference.
// Included by csrc/life_kernels.cu (product) and by the experiment TUs under csrc/experiments/.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <type_traits>

#include "gol_internal.h"

namespace golk {

constexpr int kWL = 2;              // 32-bit words per lane (64 cells)
constexpr int kValidLanes = 30;     // lanes 1..30 own output, lanes 0 and 31 are apron
constexpr int kVW = kValidLanes * kWL;  // owned words per strip
constexpr int G = GOL_GHOST_ROWS;

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
// from the rule on two impossible inputs only ((a, b, n, m) = (0, 0, 0, 1) and (2, 0, 2, 1)).
// The 7- and 8-LOP3 forms it replaced are in csrc/experiments/life_kernels_r1_variants.cu.
__device__ __forceinline__ uint32_t life_rule(uint32_t a0, uint32_t a1, uint32_t b0, uint32_t b1,
                                              uint32_t n0, uint32_t n1, uint32_t m)
{
    const uint32_t p0 = xor3(a0, b0, n0);        // low column: count odd
    const uint32_t d0 = lop3<0x7E>(a0, b0, n0);  // low column: count 1 or 2
    const uint32_t o1 = lop3<0x16>(a1, b1, n1);  // high column: count exactly 1
    const uint32_t w = maj3(a1, n0, n1);         // (no symmetric meaning; part of the synthesised circuit)
    const uint32_t g = lop3<0x76>(w, o1, d0);
    return lop3<0x54>(m, p0, g);
}

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

struct MarchArgs {
    const uint32_t *in;   // padded input buffer (global)
    uint32_t *out;        // padded output buffer (global)
    int pitch, nw, nwf, rem;
    int vi;               // virtual word index of this lane's first word
    int y0, y1;           // output rows [y0, y1)
    uint32_t keep[kWL];   // per-word ownership / tail mask of this lane
    int mode;             // 1 plain vector load, 2 composed ragged tail, 0 nothing to load
    int lane;
};

// Row source of the march: plain global loads with two rows of register prefetch in flight.
// (The TMA experiment supplies a shared-memory ring with the same interface.)
struct GlobalRowSource {
    const uint32_t *pin;
    int pitch;
    __device__ __forceinline__ void init(const MarchArgs &A, int rbeg)
    {
        pitch = A.pitch;
        pin = A.in + (int64_t)(rbeg + G) * A.pitch + GOL_GHOST_WORDS32;
    }
    template <bool FAST>
    __device__ __forceinline__ void load(const MarchArgs &A, uint32_t (&w)[kWL])
    {
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
    }
};

#ifndef GOL_WARM_SPLIT
#define GOL_WARM_SPLIT 1   // prune the inactive levels of the pipeline fill (see march)
#endif

// One warp streams rows [y0-K, y1+K) of generation g through K register-resident levels and
// writes rows [y0, y1) of generation g+K.  FAST: every lane does a plain vector load and a
// (predicated) vector store - true for all strips except the ragged right edge of a row.
template <int K, bool FLAGS, bool FAST, class Source>
__device__ __forceinline__ void march(const MarchArgs &A, Source &src, uint32_t &flag_alive, uint32_t &flag_diff)
{
    const int pitch = A.pitch;
    const bool store_all = (A.keep[0] & A.keep[1]) == 0xffffffffu;
    const bool store_any = (A.keep[0] | A.keep[1]) != 0u;
    const int rbeg = A.y0 - K, rload_end = A.y1 + K;  // level-0 rows consumed: [rbeg, rload_end)
    const int rend = A.y1 + K;                        // steps: r in [rbeg, rend)
    src.init(A, rbeg);
    uint32_t *pout = A.out + (int64_t)(A.y0 + G) * pitch + GOL_GHOST_WORDS32 + A.vi;

    uint32_t h0[K][2][kWL], h1[K][2][kWL], ctr[K][kWL];
#pragma unroll
    for (int t = 0; t < K; ++t)
#pragma unroll
        for (int j = 0; j < kWL; ++j) {
            h0[t][0][j] = h0[t][1][j] = h1[t][0][j] = h1[t][1][j] = 0;
            ctr[t][j] = 0;
        }
    uint32_t acc_a[kWL] = {0, 0}, acc_d[kWL] = {0, 0};

    // q[p] holds the row for the next step of parity p: a row is requested two full steps before it
    // is touched again (no queue shifting - a move out of a register with a load in flight stalls)
    uint32_t q[2][kWL];
    src.template load<FAST>(A, q[0]);
    if (rbeg + 1 < rload_end) src.template load<FAST>(A, q[1]);
    else { q[1][0] = q[0][0]; q[1][1] = q[0][1]; }

    // One level: consumes row `x` of generation g+t, updates the level's 3-row window and returns
    // the row of generation g+t+1 that sits one row above x.
    auto level = [&](auto t_tag, auto par_tag, const uint32_t(&x)[kWL], uint32_t(&out)[kWL], bool flags_ok) {
        constexpr int t = decltype(t_tag)::value;
        constexpr int PAR = decltype(par_tag)::value;
        const uint32_t left = __shfl_up_sync(0xffffffffu, x[kWL - 1], 1);
        const uint32_t right = __shfl_down_sync(0xffffffffu, x[0], 1);
#pragma unroll
        for (int j = 0; j < kWL; ++j) {
            const uint32_t nL = __funnelshift_l(j ? x[j - 1] : left, x[j], 1);
            const uint32_t nC = x[j];
            const uint32_t nR = __funnelshift_r(x[j], j < kWL - 1 ? x[j + 1] : right, 1);
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
    // emitted yet.  This removes K(K-1) of the 2K*K fill level-steps of a chunk.
    auto step = [&](int r, auto par_tag, auto warm_tag) {
        constexpr bool WARM = decltype(warm_tag)::value;
        uint32_t x0[kWL], fin[kWL];
#pragma unroll
        for (int j = 0; j < kWL; ++j) { x0[j] = q[decltype(par_tag)::value][j]; fin[j] = 0; }
        if (r + 2 < rload_end) src.template load<FAST>(A, q[decltype(par_tag)::value]);
        const bool emit = !WARM && r - K >= A.y0;
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
        if constexpr (!WARM) {
            if (emit) {  // fin = row (r - K) of generation g+K
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
#if GOL_WARM_SPLIT
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
        flag_alive = (acc_a[0] & A.keep[0]) | (acc_a[1] & A.keep[1]);
        flag_diff = (acc_d[0] & A.keep[0]) | (acc_d[1] & A.keep[1]);
    }
}

}  // namespace golk
