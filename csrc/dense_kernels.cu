// Stand-alone device ops on packed dense grids (gol_op_* of the C ABI):
//   generic 1-generation torus step for any size (tiny grids, PyTorch op; G:52-74,128-148),
//   empty / compare as one flag pass (G:76-126), cell <-> bit packing,
//   text parse / format kernels (replace the fgetc reader S:149-167 and fprintf writer S:25-40),
//   the seeded generator (generate.sh) and population count.
#include <cuda_runtime.h>

#include "gol_internal.h"
#include "life_rule.h"

namespace {

__global__ void step_dense_kernel(const uint64_t *__restrict__ in, uint64_t *__restrict__ out, int64_t W,
                                  int64_t H, int64_t nq, uint32_t *flags)
{
    const int64_t q = (int64_t)blockIdx.y * blockDim.x + threadIdx.x;
    const int64_t y = blockIdx.x;
    uint64_t alive = 0, diff = 0;
    if (q < nq) {
        const uint64_t v = gol_step_word(in, q, y, nq, W, H);
        out[y * nq + q] = v;
        alive = v;
        diff = v ^ in[y * nq + q];
    }
    if (flags) {
        const bool a = __any_sync(0xffffffffu, alive != 0), d = __any_sync(0xffffffffu, diff != 0);
        if ((threadIdx.x & 31) == 0) {
            if (a) atomicOr(flags + 0, 1u);
            if (d) atomicOr(flags + 1, 1u);
        }
    }
}

__global__ void flags_dense_kernel(const uint64_t *__restrict__ a, const uint64_t *__restrict__ b, int64_t n,
                                   uint32_t *flags)
{
    uint64_t alive = 0, diff = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const uint64_t v = a[i];
        alive |= v;
        if (b) diff |= v ^ b[i];
    }
    const bool fa = __any_sync(0xffffffffu, alive != 0), fd = __any_sync(0xffffffffu, diff != 0);
    if ((threadIdx.x & 31) == 0) {
        if (fa) atomicOr(flags + 0, 1u);
        if (fd) atomicOr(flags + 1, 1u);
    }
}

// One warp converts 32 consecutive 32-cell groups of a row: coalesced byte loads + ballot.
template <bool TEXT>
__global__ void pack_kernel(const uint8_t *__restrict__ src, int64_t src_pitch, uint32_t *__restrict__ dst,
                            int64_t W, int64_t H, int64_t nw32, uint32_t *err)
{
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.y * blockDim.x + threadIdx.x) >> 5;
    const int64_t y = blockIdx.x;
    const int64_t w0 = warp * 32;  // first 32-bit word handled by this warp
    if (w0 >= nw32) return;
    const uint8_t *row = src + y * src_pitch;
    uint32_t mine = 0, bad = 0;
#pragma unroll 4
    for (int k = 0; k < 32; ++k) {
        const int64_t x = (w0 + k) * 32 + lane;
        uint8_t c = 0;
        const bool in = x < W;
        if (in) c = row[x];
        bool bit;
        if (TEXT) {
            bit = c == '1';
            if (in && c != '1' && c != '0') bad = 1;
        } else {
            bit = c != 0;
        }
        const uint32_t word = __ballot_sync(0xffffffffu, bit);
        if (lane == k) mine = word;
    }
    if (w0 + lane < nw32) dst[y * nw32 + w0 + lane] = mine;
    if (TEXT) {
        if (warp == 0 && lane == 0 && src_pitch > W && row[W] != '\n') bad |= 2;
        if (bad && err) atomicOr(err, bad);
    }
}

template <bool TEXT>
__global__ void unpack_kernel(const uint32_t *__restrict__ src, uint8_t *__restrict__ dst, int64_t dst_pitch,
                              int64_t W, int64_t H, int64_t nw32, int newline)
{
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.y * blockDim.x + threadIdx.x) >> 5;
    const int64_t y = blockIdx.x;
    const int64_t w0 = warp * 32;
    if (w0 >= nw32) return;
    const uint32_t mine = (w0 + lane < nw32) ? src[y * nw32 + w0 + lane] : 0u;
    uint8_t *row = dst + y * dst_pitch;
#pragma unroll 4
    for (int k = 0; k < 32; ++k) {
        const uint32_t word = __shfl_sync(0xffffffffu, mine, k);
        const int64_t x = (w0 + k) * 32 + lane;
        if (x < W) row[x] = (uint8_t)((TEXT ? '0' : 0) + ((word >> lane) & 1u));
    }
    if (TEXT && newline && warp == 0 && lane == 0) row[W] = '\n';
}

__global__ void random_dense_kernel(uint64_t *out, int64_t W, int64_t H, int64_t nq, int64_t qx0, int64_t y0,
                                    uint64_t seed)
{
    const int64_t q = (int64_t)blockIdx.y * blockDim.x + threadIdx.x;
    const int64_t y = blockIdx.x;
    if (q >= nq) return;
    out[y * nq + q] = gol_hash64(seed, (uint64_t)(qx0 + q), (uint64_t)(y0 + y)) & gol_tail_mask(q, nq, W);
}

__global__ void popcount_kernel(const uint64_t *__restrict__ a, int64_t n, unsigned long long *count)
{
    unsigned long long c = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        c += __popcll(a[i]);
    for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(count, c);
}

int check_dims(int64_t W, int64_t H)
{
    if (W < 1 || H < 1) GOL_FAIL("grid dimensions must be positive (got %lld x %lld)", (long long)W, (long long)H);
    if (H > 0x7fffffffLL) GOL_FAIL("height %lld too large", (long long)H);
    return 0;
}
int check_launch(const char *what)
{
    cudaError_t err = cudaGetLastError();
    if (err != cudaSuccess) GOL_FAIL("%s launch failed: %s", what, cudaGetErrorString(err));
    return 0;
}

}  // namespace

extern "C" {

int gol_op_step_dense(const uint64_t *d_in, uint64_t *d_out, int64_t W, int64_t H, uint32_t *d_flags, void *stream)
{
    GOL_TRY(check_dims(W, H));
    if (d_in == d_out) GOL_FAIL("gol_op_step_dense: in-place step is not supported");
    const int64_t nq = gol_div_up(W, 64);
    const int threads = nq >= 128 ? 128 : 32 * (int)gol_div_up(nq, 32);
    dim3 grid((unsigned)H, (unsigned)gol_div_up(nq, threads));
    step_dense_kernel<<<grid, threads, 0, (cudaStream_t)stream>>>(d_in, d_out, W, H, nq, d_flags);
    return check_launch("step_dense_kernel");
}

int gol_op_flags_dense(const uint64_t *d_a, const uint64_t *d_b, int64_t W, int64_t H, uint32_t *d_flags,
                       void *stream)
{
    GOL_TRY(check_dims(W, H));
    const int64_t n = gol_div_up(W, 64) * H;
    const int blocks = (int)(n < 256 * 1184 ? gol_div_up(n, 256) : 1184);
    flags_dense_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(d_a, d_b, n, d_flags);
    return check_launch("flags_dense_kernel");
}

static dim3 warp_grid(int64_t H, int64_t nw32, int threads)
{
    const int64_t warps = gol_div_up(nw32, 32);
    return dim3((unsigned)H, (unsigned)gol_div_up(warps * 32, threads));
}

int gol_op_pack_cells(const uint8_t *d_cells, uint64_t *d_packed, int64_t W, int64_t H, void *stream)
{
    GOL_TRY(check_dims(W, H));
    const int64_t nw32 = 2 * gol_div_up(W, 64);
    pack_kernel<false><<<warp_grid(H, nw32, 128), 128, 0, (cudaStream_t)stream>>>(
        d_cells, W, reinterpret_cast<uint32_t *>(d_packed), W, H, nw32, nullptr);
    return check_launch("pack_kernel");
}

int gol_op_unpack_cells(const uint64_t *d_packed, uint8_t *d_cells, int64_t W, int64_t H, void *stream)
{
    GOL_TRY(check_dims(W, H));
    const int64_t nw32 = 2 * gol_div_up(W, 64);
    unpack_kernel<false><<<warp_grid(H, nw32, 128), 128, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const uint32_t *>(d_packed), d_cells, W, W, H, nw32, 0);
    return check_launch("unpack_kernel");
}

int gol_op_parse_text(const char *d_text, uint64_t *d_packed, int64_t W, int64_t H, int64_t text_pitch,
                      uint32_t *d_err, void *stream)
{
    GOL_TRY(check_dims(W, H));
    if (text_pitch < W) GOL_FAIL("gol_op_parse_text: text pitch %lld < width %lld", (long long)text_pitch, (long long)W);
    const int64_t nw32 = 2 * gol_div_up(W, 64);
    pack_kernel<true><<<warp_grid(H, nw32, 128), 128, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const uint8_t *>(d_text), text_pitch, reinterpret_cast<uint32_t *>(d_packed), W, H, nw32,
        d_err);
    return check_launch("parse_text_kernel");
}

int gol_op_format_text(const uint64_t *d_packed, char *d_text, int64_t W, int64_t H, int64_t text_pitch,
                       int newline, void *stream)
{
    GOL_TRY(check_dims(W, H));
    if (text_pitch < W + (newline ? 1 : 0)) GOL_FAIL("gol_op_format_text: text pitch too small");
    const int64_t nw32 = 2 * gol_div_up(W, 64);
    unpack_kernel<true><<<warp_grid(H, nw32, 128), 128, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const uint32_t *>(d_packed), reinterpret_cast<uint8_t *>(d_text), text_pitch, W, H, nw32,
        newline);
    return check_launch("format_text_kernel");
}

int gol_op_random_dense(uint64_t *d_packed, int64_t W, int64_t H, int64_t x0, int64_t y0, uint64_t seed,
                        void *stream)
{
    GOL_TRY(check_dims(W, H));
    if (x0 % 64) GOL_FAIL("gol_op_random_dense: x0 must be a multiple of 64");
    const int64_t nq = gol_div_up(W, 64);
    dim3 grid((unsigned)H, (unsigned)gol_div_up(nq, 128));
    random_dense_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(d_packed, W, H, nq, x0 / 64, y0, seed);
    return check_launch("random_dense_kernel");
}

int gol_op_popcount_dense(const uint64_t *d_packed, int64_t W, int64_t H, unsigned long long *d_count,
                          void *stream)
{
    GOL_TRY(check_dims(W, H));
    const int64_t n = gol_div_up(W, 64) * H;
    const int blocks = (int)(n < 256 * 1184 ? gol_div_up(n, 256) : 1184);
    popcount_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(d_packed, n, d_count);
    return check_launch("popcount_kernel");
}

}  // extern "C"
