// Experiment: the generation kernel as a TMA-fed, cluster/DSMEM tile kernel (bin/gol-tilebench).
//
// The product kernel (csrc/life_kernels.cu) streams rows through a K-deep register pipeline and
// pays for temporal blocking with redundant rows at the top of every chunk (pipeline fill) and
// two apron lanes per strip.  This file measures the textbook alternative the design brief
// names - TMA tile loads, thread-block clusters that share tile halos through distributed shared
// memory, K generations per launch on a resident tile - on the same grid and against the same
// arithmetic core:
//
//   * a CTA of NW warps keeps a tile of 64 words x (NW * R) rows in REGISTERS (R rows x 2 words per
//     thread) for K generations;
//   * the tile arrives through the TMA unit into shared memory, completion on an mbarrier: as
//     cp.async.bulk.tensor.2d tile loads through a tensor map (TMA = 1) or as one 1-D cp.async.bulk
//     per tile row (TMA = 2); or with plain vector loads straight into registers (TMA = 0).  The
//     staged tile starts two words left of the strip so that every row starts on a 16-byte boundary;
//   * every generation the warps exchange their top / bottom rows through a small shared-memory
//     buffer; with a cluster of CY CTAs stacked in y the rows at CTA boundaries are read from the
//     neighbouring CTA's shared memory (DSMEM), so only the outer K rows of the whole cluster
//     tile are redundant instead of K rows per CTA;
//   * x neighbours move with warp shuffles exactly as in the product kernel (lanes 0 / 31 are apron).
//
// The tensor map covers this GPU's own buffer only and the load is not multicast.
// The binary checks every variant bit for bit against the product's march core and prints the
// time per epoch of K generations.  Results are recorded in profiles/README.md.
#include <cooperative_groups.h>
#include <cuda.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../life_march.cuh"

namespace cg = cooperative_groups;
using namespace golk;

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e__ = (call);                                                                  \
        if (e__ != cudaSuccess) {                                                                  \
            fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e__), __FILE__, __LINE__); \
            exit(1);                                                                               \
        }                                                                                          \
    } while (0)

struct Grid {
    int H, nw, pitch;      // rows, data words per row, row pitch (32-bit words)
    int rows_padded;
    int strips;
};

__host__ __device__ inline int strip_lo(int strip, int S, int nw) { return (strip == S - 1 && S > 1) ? nw - kVW : strip * kVW; }

// ------------------------------------------------------------ set-up kernels
__global__ void fill_random(uint32_t *buf, Grid g, uint64_t seed)
{
    const int y = blockIdx.x;
    for (int q = threadIdx.x; q < g.nw / 2; q += blockDim.x) {
        const uint64_t v = gol_hash64(seed, (uint64_t)q, (uint64_t)y);
        uint32_t *row = buf + (int64_t)(y + G) * g.pitch + GOL_GHOST_WORDS32;
        row[2 * q] = (uint32_t)v;
        row[2 * q + 1] = (uint32_t)(v >> 32);
    }
}
// torus ghost zones: 16 rows above / below, 4 words left / right (widths that are multiples of 64 cells)
__global__ void fill_ghosts(uint32_t *buf, Grid g)
{
    const int yp = blockIdx.x;  // padded row
    const int y = ((yp - G) % g.H + g.H) % g.H;
    uint32_t *dst = buf + (int64_t)yp * g.pitch + GOL_GHOST_WORDS32;
    const uint32_t *src = buf + (int64_t)(y + G) * g.pitch + GOL_GHOST_WORDS32;
    if (yp < G || yp >= g.H + G)
        for (int i = threadIdx.x; i < g.nw; i += blockDim.x) dst[i] = src[i];
    __syncthreads();
    if (threadIdx.x < GOL_GHOST_WORDS32) {
        dst[-GOL_GHOST_WORDS32 + (int)threadIdx.x] = src[g.nw - GOL_GHOST_WORDS32 + threadIdx.x];
        dst[g.nw + threadIdx.x] = src[threadIdx.x];
    }
}
__global__ void compare_interior(const uint32_t *a, const uint32_t *b, Grid g, unsigned long long *ndiff)
{
    const int y = blockIdx.x;
    unsigned long long c = 0;
    for (int i = threadIdx.x; i < g.nw; i += blockDim.x) {
        const int64_t o = (int64_t)(y + G) * g.pitch + GOL_GHOST_WORDS32 + i;
        c += a[o] != b[o];
    }
    if (c) atomicAdd(ndiff, c);
}

// ------------------------------------------------ reference: the product march
template <int K>
__global__ void __launch_bounds__(128, 3) march_ref_kernel(const uint32_t *in, uint32_t *out, Grid g, int chunk_rows)
{
    const int lane = threadIdx.x & 31;
    const int item = blockIdx.x * 4 + (threadIdx.x >> 5);
    const int chunks = (g.H + chunk_rows - 1) / chunk_rows;
    if (item >= chunks * g.strips) return;
    const int chunk = item / g.strips, strip = item - chunk * g.strips;
    MarchArgs A;
    A.nw = g.nw; A.nwf = g.nw; A.rem = 0; A.pitch = g.pitch;
    A.in = in; A.out = out;
    A.vi = strip_lo(strip, g.strips, g.nw) - kWL + lane * kWL;
    A.y0 = chunk * chunk_rows; A.y1 = min(g.H, A.y0 + chunk_rows);
    A.lane = lane;
    const bool owned = lane >= 1 && lane <= kValidLanes;
    for (int j = 0; j < kWL; ++j) A.keep[j] = (owned && A.vi + j >= 0 && A.vi + j < g.nw) ? 0xffffffffu : 0u;
    A.mode = 1;
    uint32_t fa, fd;
    GlobalRowSource src;
    march<K, false, true>(A, src, fa, fd);
}

// ---------------------------------------------------------------- tile kernel
__device__ __forceinline__ void hsum(const uint32_t (&x)[2], uint32_t (&s0)[2], uint32_t (&s1)[2])
{
    const uint32_t left = __shfl_up_sync(0xffffffffu, x[1], 1);
    const uint32_t right = __shfl_down_sync(0xffffffffu, x[0], 1);
    const uint32_t l0 = __funnelshift_l(left, x[0], 1), r0 = __funnelshift_r(x[0], x[1], 1);
    const uint32_t l1 = __funnelshift_l(x[0], x[1], 1), r1 = __funnelshift_r(x[1], right, 1);
    s0[0] = xor3(l0, x[0], r0); s1[0] = maj3(l0, x[0], r0);
    s0[1] = xor3(l1, x[1], r1); s1[1] = maj3(l1, x[1], r1);
}

// rows per tensor-map box: the largest divisor of the tile height that is at most 160
__host__ __device__ constexpr int box_rows(int th)
{
    int best = 1;
    for (int d = 1; d <= 160; ++d)
        if (th % d == 0) best = d;
    return best;
}

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int K, int R, int NW, int CY, int TMA, int MB>
__global__ void __launch_bounds__(NW * 32, MB)
life_tile_kernel(const __grid_constant__ CUtensorMap tmap, const uint32_t *__restrict__ in, uint32_t *__restrict__ out, Grid g)
{
    constexpr int TH = NW * R;              // rows per CTA
    constexpr int OUT = CY * TH - 2 * K;    // rows a cluster produces
    constexpr int BOXR = box_rows(TH);      // rows per tensor box (<= 256)
    constexpr int NBOX = TH / BOXR;
    static_assert(BOXR * NBOX == TH && BOXR <= 256, "tile height must split into boxes of <= 256 rows");
    constexpr int TW = 68;                  // staged words per row: 2 alignment words + the 64 words of the strip (+2)
    extern __shared__ __align__(128) unsigned char smem[];
    uint32_t *tile = reinterpret_cast<uint32_t *>(smem);                          // [TH][TW] (TMA only)
    constexpr int kTileBytes = TMA ? TH * TW * 4 : 0;
    uint2 *xch = reinterpret_cast<uint2 *>(smem + kTileBytes);                    // [2][NW][2][32]
    uint64_t *mbar = reinterpret_cast<uint64_t *>(smem + kTileBytes + 2 * NW * 2 * 32 * 8);

    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int strip = blockIdx.x;
    const int j = blockIdx.y / CY, cy = blockIdx.y - j * CY;   // cluster dims (1, CY, 1): cy = rank in the cluster
    const int ytop = j * OUT - K + cy * TH;                    // first grid row of this CTA's tile
    const int lo = strip_lo(strip, g.strips, g.nw);
    const int vi = lo - kWL + lane * kWL;                      // this lane's first word (virtual index)

    uint32_t c[R][2];
    if (TMA) {
        // rows of the tile that exist in the buffer (the last tiles hang over its end; those rows are never needed)
        const int nvalid = min(TH, g.rows_padded - (ytop + G));
        if (threadIdx.x == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(mbar)));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncthreads();
        if (TMA == 1) {
            if (threadIdx.x == 0) {
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(mbar)), "r"(TH * TW * 4) : "memory");
                // rows beyond the buffer come back as zeros (out-of-bounds fill); x starts on a 16-byte boundary
#pragma unroll
                for (int b = 0; b < NBOX; ++b)
                    asm volatile(
                        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                        ::"r"(smem_u32(tile + b * BOXR * TW)), "l"(&tmap), "r"(GOL_GHOST_WORDS32 + lo - 4), "r"(ytop + G + b * BOXR),
                        "r"(smem_u32(mbar))
                        : "memory");
            }
        } else if (w == 0) {
            if (lane == 0)
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(mbar)), "r"(nvalid * TW * 4) : "memory");
            __syncwarp();
            const uint32_t *src = in + (int64_t)(ytop + G) * g.pitch + GOL_GHOST_WORDS32 + lo - 4;
            for (int r = lane; r < nvalid; r += 32)
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(smem_u32(tile + r * TW)), "l"(src + (int64_t)r * g.pitch), "r"(TW * 4), "r"(smem_u32(mbar))
                             : "memory");
        }
        uint32_t done = 0;
        while (!done)
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }"
                         : "=r"(done) : "r"(smem_u32(mbar)) : "memory");
#pragma unroll
        for (int i = 0; i < R; ++i) {
            const uint2 v = *reinterpret_cast<const uint2 *>(tile + (w * R + i) * TW + 2 + lane * 2);
            c[i][0] = v.x; c[i][1] = v.y;
        }
    } else {
#pragma unroll
        for (int i = 0; i < R; ++i) {
            const int yp = ytop + w * R + i + G;
            uint2 v = make_uint2(0u, 0u);
            if (yp < g.rows_padded) v = ldg_v2(in + (int64_t)yp * g.pitch + GOL_GHOST_WORDS32 + vi);
            c[i][0] = v.x; c[i][1] = v.y;
        }
    }

    cg::cluster_group cluster = cg::this_cluster();
    const uint2 *xch_up = xch, *xch_dn = xch;
    if (CY > 1) {
        if (cy > 0) xch_up = cluster.map_shared_rank(xch, cy - 1);
        if (cy < CY - 1) xch_dn = cluster.map_shared_rank(xch, cy + 1);
    }

#pragma unroll 1
    for (int gen = 0; gen < K; ++gen) {
        const int par = gen & 1;
        // publish my top and bottom row, then pick up the rows next to mine
        xch[((par * NW + w) * 2 + 0) * 32 + lane] = make_uint2(c[0][0], c[0][1]);
        xch[((par * NW + w) * 2 + 1) * 32 + lane] = make_uint2(c[R - 1][0], c[R - 1][1]);
        if (CY > 1) cluster.sync();
        else __syncthreads();
        uint2 up = make_uint2(0u, 0u), dn = make_uint2(0u, 0u);
        if (w > 0) up = xch[((par * NW + w - 1) * 2 + 1) * 32 + lane];
        else if (CY > 1 && cy > 0) up = xch_up[((par * NW + NW - 1) * 2 + 1) * 32 + lane];
        if (w < NW - 1) dn = xch[((par * NW + w + 1) * 2 + 0) * 32 + lane];
        else if (CY > 1 && cy < CY - 1) dn = xch_dn[((par * NW + 0) * 2 + 0) * 32 + lane];

        uint32_t a0[2], a1[2], b0[2], b1[2], n0[2], n1[2];
        {
            const uint32_t u[2] = {up.x, up.y};
            hsum(u, a0, a1);
        }
        hsum(c[0], b0, b1);
#pragma unroll
        for (int i = 0; i < R; ++i) {
            if (i + 1 < R) {
                hsum(c[i + 1], n0, n1);
            } else {
                const uint32_t d[2] = {dn.x, dn.y};
                hsum(d, n0, n1);
            }
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                const uint32_t v = life_rule(a0[q], a1[q], b0[q], b1[q], n0[q], n1[q], c[i][q]);
                a0[q] = b0[q]; a1[q] = b1[q];
                b0[q] = n0[q]; b1[q] = n1[q];
                c[i][q] = v;
            }
        }
    }
    if (CY > 1) cluster.sync();   // nobody leaves while a neighbour may still read its exchange buffer

    const bool owned = lane >= 1 && lane <= kValidLanes && vi >= 0 && vi + kWL <= g.nw;
    const int ya = j * OUT, yb = min(g.H, ya + OUT);
    if (owned) {
#pragma unroll
        for (int i = 0; i < R; ++i) {
            const int y = ytop + w * R + i;
            if (y >= ya && y < yb) stg_v2(out + (int64_t)(y + G) * g.pitch + GOL_GHOST_WORDS32 + vi, c[i][0], c[i][1]);
        }
    }
}

// -------------------------------------------------------------------- host
typedef CUresult (*EncodeTiled)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static CUtensorMap make_map(const uint32_t *buf, const Grid &g, int box_rows)
{
    void *fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
    CUtensorMap m;
    const cuuint64_t dims[2] = {(cuuint64_t)g.pitch, (cuuint64_t)g.rows_padded};
    const cuuint64_t strides[1] = {(cuuint64_t)g.pitch * 4};
    const cuuint32_t box[2] = {68, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    CUresult r = ((EncodeTiled)fn)(&m, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, (void *)buf, dims, strides, box, estr,
                                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        fprintf(stderr, "cuTensorMapEncodeTiled failed: %d\n", (int)r);
        exit(1);
    }
    return m;
}

struct Result {
    const char *name;
    float us;
    unsigned long long diff;
    int ctas;
};

template <int K, int R, int NW, int CY, int TMA, int MB>
static Result run_tile(const char *name, const uint32_t *in, uint32_t *out, const uint32_t *ref, const Grid &g, int reps,
                       unsigned long long *d_diff)
{
    constexpr int TH = NW * R, OUT = CY * TH - 2 * K;
    constexpr int BOXR = box_rows(TH);
    const size_t smem = (TMA ? (size_t)TH * 68 * 4 : 0) + (size_t)2 * NW * 2 * 32 * 8 + 16;
    auto kern = life_tile_kernel<K, R, NW, CY, TMA, MB>;
    CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const CUtensorMap map = make_map(in, g, BOXR);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)g.strips, (unsigned)(((g.H + OUT - 1) / OUT) * CY));
    cfg.blockDim = dim3(NW * 32);
    cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 1; attr[0].val.clusterDim.y = CY; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    CK(cudaMemset(out, 0, (size_t)g.rows_padded * g.pitch * 4));
    CK(cudaLaunchKernelEx(&cfg, kern, map, in, out, g));
    CK(cudaDeviceSynchronize());
    CK(cudaMemset(d_diff, 0, 8));
    compare_interior<<<g.H, 256>>>(out, ref, g, d_diff);
    unsigned long long diff = 0;
    CK(cudaMemcpy(&diff, d_diff, 8, cudaMemcpyDeviceToHost));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for (int i = 0; i < 3; ++i) CK(cudaLaunchKernelEx(&cfg, kern, map, in, out, g));
    CK(cudaEventRecord(e0));
    for (int i = 0; i < reps; ++i) CK(cudaLaunchKernelEx(&cfg, kern, map, in, out, g));
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    return {name, ms * 1e3f / reps, diff, (int)(cfg.gridDim.x * cfg.gridDim.y)};
}

int main(int argc, char **argv)
{
    const int W = argc > 1 ? atoi(argv[1]) : 16384, H = argc > 2 ? atoi(argv[2]) : 16384, reps = argc > 3 ? atoi(argv[3]) : 20;
    if (W % 64 || W < 128 || H < 64) {
        fprintf(stderr, "usage: gol-tilebench [width (multiple of 64)] [height] [reps]\n");
        return 1;
    }
    constexpr int K = 8;
    Grid g;
    g.H = H; g.nw = W / 32;
    g.pitch = (2 * GOL_GHOST_WORDS32 + g.nw + 7) / 8 * 8;
    g.rows_padded = H + 2 * G;
    g.strips = (g.nw + kVW - 1) / kVW;
    const size_t bytes = (size_t)g.rows_padded * g.pitch * 4;
    uint32_t *in, *ref, *out;
    unsigned long long *d_diff;
    CK(cudaMalloc(&in, bytes)); CK(cudaMalloc(&ref, bytes)); CK(cudaMalloc(&out, bytes)); CK(cudaMalloc(&d_diff, 8));
    CK(cudaMemset(in, 0, bytes)); CK(cudaMemset(ref, 0, bytes));
    fill_random<<<H, 256>>>(in, g, 12345);
    fill_ghosts<<<g.rows_padded, 256>>>(in, g);
    CK(cudaDeviceSynchronize());

    // argv[4]: variant to run (-1 = header line only, 0 = the register pipeline, 1.. = tile variants).
    // One variant per process, so that a faulting variant cannot take the others with it.
    const int which = argc > 4 ? atoi(argv[4]) : 0;
    const int chunk_rows = argc > 5 ? atoi(argv[5]) : 85;
    std::vector<Result> rs;
    {   // reference output (and timing): the product's register-pipelined march at its tuned chunk height
        const int items = ((H + chunk_rows - 1) / chunk_rows) * g.strips;
        march_ref_kernel<K><<<(items + 3) / 4, 128>>>(in, ref, g, chunk_rows);
        CK(cudaDeviceSynchronize());
        if (which == 0) {
            cudaEvent_t e0, e1;
            CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
            for (int i = 0; i < 3; ++i) march_ref_kernel<K><<<(items + 3) / 4, 128>>>(in, out, g, chunk_rows);
            CK(cudaEventRecord(e0));
            for (int i = 0; i < reps; ++i) march_ref_kernel<K><<<(items + 3) / 4, 128>>>(in, out, g, chunk_rows);
            CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1));
            float ms = 0;
            CK(cudaEventElapsedTime(&ms, e0, e1));
            rs.push_back({"register pipeline (product march, no pushes)", ms * 1e3f / reps, 0ull, (items + 3) / 4});
        }
    }
    int idx = 0;
#define TILE(R, NW, CY, TMA, MB) \
    if (++idx == which)          \
    rs.push_back(run_tile<K, R, NW, CY, TMA, MB>("tile " #R " rows/thread x " #NW " warps, cluster " #CY ", tma " #TMA ", " #MB " CTAs/SM", in, out, ref, g, reps, d_diff))
    TILE(34, 8, 1, 0, 2);
    TILE(34, 8, 1, 1, 2);
    TILE(34, 8, 1, 2, 2);
    TILE(32, 8, 1, 1, 2);
    TILE(32, 8, 1, 2, 2);
    TILE(34, 8, 2, 0, 2);
    TILE(34, 8, 2, 2, 2);
    TILE(34, 8, 4, 0, 2);
    TILE(34, 8, 4, 2, 2);
    TILE(32, 8, 1, 0, 2);
    TILE(32, 8, 2, 0, 2);
    TILE(32, 8, 4, 0, 2);
    TILE(20, 8, 1, 0, 3);
    TILE(20, 8, 1, 2, 3);
    TILE(24, 8, 1, 0, 2);
    TILE(36, 8, 1, 0, 2);
    if (which < 0) {
        printf("# %d x %d, K = %d generations per launch, %d timed launches each, %d tile variants\n", W, H, K, reps, idx);
        printf("# %-58s %10s %14s %7s %s\n", "variant", "us/epoch", "us/generation", "CTAs", "words that differ from the reference");
    }
    for (const Result &r : rs) printf("  %-58s %10.2f %14.3f %7d %llu\n", r.name, r.us, r.us / K, r.ctas, r.diff);
    fflush(stdout);
    return 0;
}
