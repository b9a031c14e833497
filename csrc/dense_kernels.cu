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

// ---------------------------------------------------------------- bytes <-> bits
// Text lines are W+1 bytes long, so a row starts at an arbitrary byte alignment.  Both kernels
// therefore move their bytes through shared memory: global traffic is 16-byte vectors on aligned
// addresses (fully coalesced), and the per-thread work is on aligned 32-bit words of the staged
// segment.  A CTA handles kSegWords consecutive 32-bit words (8192 cells) of one row.
constexpr int kSegThreads = 256;
constexpr int kSegWords = kSegThreads;          // one output / input word per thread
constexpr int kSegBytes = kSegWords * 32;       // cell bytes per segment

// 4 bytes holding 0/1 in their low bit -> 4 bits (byte 0 -> bit 0)
__device__ __forceinline__ uint32_t gather4(uint32_t v) { return ((v & 0x01010101u) * 0x01020408u) >> 24; }
// 4 bits -> 4 bytes holding 0/1
__device__ __forceinline__ uint32_t spread4(uint32_t nib) { return (nib * 0x00204081u) & 0x01010101u; }

// bytes -> bits.  TEXT: '1' alive, '0' dead, anything else sets error bit 1; a byte other than
// '\n' at column W (when the buffer has one there) sets error bit 2.  !TEXT: non-zero = alive.
// [buf, buf + len) is the whole source buffer: nothing outside it is touched.
template <bool TEXT>
__global__ void __launch_bounds__(kSegThreads) pack_kernel(const uint8_t *__restrict__ buf, int64_t len, int64_t src_pitch,
                                                           uint32_t *__restrict__ dst, int64_t W, int64_t nw32, uint32_t *err)
{
    __shared__ uint4 stage[kSegBytes / 16 + 2];
    const int64_t y = blockIdx.x, w0 = (int64_t)blockIdx.y * kSegWords;
    const int tid = threadIdx.x;
    const uint8_t *src = buf + y * src_pitch + w0 * 32;                       // first cell byte of the segment
    const int nbytes = (int)min((int64_t)kSegBytes, W - w0 * 32);              // cell bytes in this segment (>= 1)
    const int lead = (int)(reinterpret_cast<uintptr_t>(src) & 15);
    const uint8_t *a0 = src - lead;                                            // 16-byte aligned
    const int nvec = (lead + nbytes + 15) >> 4;
    const uint8_t *lo = buf, *hi = buf + len;
    for (int v = tid; v < nvec; v += kSegThreads) {
        const uint8_t *p = a0 + 16 * v;
        uint4 q;
        if (p >= lo && p + 16 <= hi) {
            q = *reinterpret_cast<const uint4 *>(p);
        } else {  // window straddles an end of the buffer: byte loads of what exists
            uint32_t w[4] = {0, 0, 0, 0};
            for (int i = 0; i < 16; ++i)
                if (p + i >= lo && p + i < hi) w[i >> 2] |= (uint32_t)p[i] << (8 * (i & 3));
            q = make_uint4(w[0], w[1], w[2], w[3]);
        }
        stage[v] = q;
    }
    __syncthreads();
    const int64_t wi = w0 + tid;
    uint32_t bad = 0;
    if (wi < nw32) {
        const uint32_t *sw = reinterpret_cast<const uint32_t *>(stage);
        const int o = lead + 32 * tid, k = o >> 2, sh = (o & 3) * 8;
        const int nb = max(0, min(32, nbytes - 32 * tid));                     // valid cell bytes of this word
        uint32_t bits = 0, prev = sw[k];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const uint32_t next = (4 * j < nb) ? sw[k + j + 1] : 0u;          // (never read past the staged bytes)
            uint32_t v = __funnelshift_r(prev, next, sh);
            prev = next;
            const int rest = nb - 4 * j;                                       // valid bytes in this group of 4
            const uint32_t mask = rest >= 4 ? 0xffffffffu : (rest <= 0 ? 0u : ((1u << (8 * rest)) - 1u));
            v &= mask;
            if (TEXT) {
                if (((v ^ (0x30303030u & mask)) & 0xfefefefeu) != 0) bad = 1;
                bits |= gather4(__vcmpeq4(v, 0x31313131u)) << (4 * j);   // exactly '1' is alive
            } else {
                bits |= gather4(__vcmpne4(v, 0u)) << (4 * j);
            }
        }
        dst[y * nw32 + wi] = bits;
    }
    if (TEXT) {
        if (blockIdx.y == 0 && tid == 0 && src_pitch > W) {
            const uint8_t *nl = buf + y * src_pitch + W;
            if (nl < hi && *nl != '\n') bad |= 2;
        }
        if (__any_sync(0xffffffffu, bad != 0)) {
            for (int o = 16; o; o >>= 1) bad |= __shfl_xor_sync(0xffffffffu, bad, o);
            if ((tid & 31) == 0 && err) atomicOr(err, bad);
        }
    }
}

// bits -> bytes ('0'/'1' text or 0/1 cells), optionally a '\n' at column W of every row.
template <bool TEXT>
__global__ void __launch_bounds__(kSegThreads) unpack_kernel(const uint32_t *__restrict__ src, uint8_t *__restrict__ dst,
                                                             int64_t dst_pitch, int64_t W, int64_t nw32, int newline)
{
    __shared__ uint4 stage[kSegBytes / 16 + 2];
    const int64_t y = blockIdx.x, w0 = (int64_t)blockIdx.y * kSegWords;
    const int tid = threadIdx.x;
    uint8_t *out = dst + y * dst_pitch + w0 * 32;                              // first byte this CTA writes
    const bool tail = w0 + kSegWords >= nw32;                                  // last segment of the row
    int nbytes = (int)min((int64_t)kSegBytes, W - w0 * 32);
    const int nl_at = (TEXT && newline && tail) ? nbytes : -1;                 // segment offset of the newline byte
    if (nl_at >= 0) ++nbytes;
    const int lead = (int)(reinterpret_cast<uintptr_t>(out) & 15);
    // Staged byte i (i >= lead) is segment byte i - lead.  Thread t fills the aligned staged words
    // 8t..8t+7, i.e. the cells 32t - lead .. 32t - lead + 31 of the segment: a 32-bit window of the
    // row's bit stream that straddles two source words.
    const uint32_t *row = src + y * nw32;
    uint32_t *sw = reinterpret_cast<uint32_t *>(stage);
    for (int t = tid; t <= kSegWords; t += kSegThreads) {
        const int64_t wi = w0 + t;
        const uint32_t cur = (t < kSegWords && wi < nw32) ? row[wi] : 0u;
        const uint32_t before = (t > 0 && wi - 1 < nw32) ? row[wi - 1] : 0u;
        const uint32_t win = __funnelshift_l(before, cur, lead);   // (cur << lead) | (before >> (32 - lead))
        const uint32_t fill = TEXT ? 0x30303030u : 0u;
        uint32_t o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = spread4((win >> (4 * j)) & 15u) | fill;
        if (8 * t + 8 <= (int)(sizeof(stage) / 4)) {
            *reinterpret_cast<uint4 *>(sw + 8 * t) = make_uint4(o[0], o[1], o[2], o[3]);
            *reinterpret_cast<uint4 *>(sw + 8 * t + 4) = make_uint4(o[4], o[5], o[6], o[7]);
        }
    }
    __syncthreads();
    if (nl_at >= 0 && tid == 0) reinterpret_cast<uint8_t *>(stage)[lead + nl_at] = '\n';
    __syncthreads();
    const int nvec = (lead + nbytes + 15) >> 4;
    uint8_t *a0 = out - lead;
    for (int v = tid; v < nvec; v += kSegThreads) {
        const int b0 = 16 * v, b1 = b0 + 16;                                   // staged byte range of this vector
        if (b0 >= lead && b1 <= lead + nbytes) {
            *reinterpret_cast<uint4 *>(a0 + b0) = stage[v];
        } else {  // partial vector at the head or tail of the segment: bytes
            const uint8_t *sb = reinterpret_cast<const uint8_t *>(stage);
            for (int i = max(b0, lead); i < min(b1, lead + nbytes); ++i) a0[i] = sb[i];
        }
    }
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

static dim3 seg_grid(int64_t H, int64_t nw32) { return dim3((unsigned)H, (unsigned)gol_div_up(nw32, kSegWords)); }

int gol_op_pack_cells(const uint8_t *d_cells, uint64_t *d_packed, int64_t W, int64_t H, void *stream)
{
    GOL_TRY(check_dims(W, H));
    const int64_t nw32 = 2 * gol_div_up(W, 64);
    pack_kernel<false><<<seg_grid(H, nw32), kSegThreads, 0, (cudaStream_t)stream>>>(
        d_cells, W * H, W, reinterpret_cast<uint32_t *>(d_packed), W, nw32, nullptr);
    return check_launch("pack_kernel");
}

int gol_op_unpack_cells(const uint64_t *d_packed, uint8_t *d_cells, int64_t W, int64_t H, void *stream)
{
    GOL_TRY(check_dims(W, H));
    const int64_t nw32 = 2 * gol_div_up(W, 64);
    unpack_kernel<false><<<seg_grid(H, nw32), kSegThreads, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const uint32_t *>(d_packed), d_cells, W, W, nw32, 0);
    return check_launch("unpack_kernel");
}

int gol_op_parse_text_n(const char *d_text, int64_t text_bytes, uint64_t *d_packed, int64_t W, int64_t H,
                        int64_t text_pitch, uint32_t *d_err, void *stream)
{
    GOL_TRY(check_dims(W, H));
    if (text_pitch < W) GOL_FAIL("gol_op_parse_text: text pitch %lld < width %lld", (long long)text_pitch, (long long)W);
    if (text_bytes < (H - 1) * text_pitch + W)
        GOL_FAIL("gol_op_parse_text: the text buffer holds %lld bytes, %lld rows of pitch %lld need at least %lld",
                 (long long)text_bytes, (long long)H, (long long)text_pitch, (long long)((H - 1) * text_pitch + W));
    const int64_t nw32 = 2 * gol_div_up(W, 64);
    pack_kernel<true><<<seg_grid(H, nw32), kSegThreads, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const uint8_t *>(d_text), text_bytes, text_pitch, reinterpret_cast<uint32_t *>(d_packed), W, nw32,
        d_err);
    return check_launch("parse_text_kernel");
}

int gol_op_parse_text(const char *d_text, uint64_t *d_packed, int64_t W, int64_t H, int64_t text_pitch,
                      uint32_t *d_err, void *stream)
{
    // without a length the buffer is taken to end with the last cell of the last row: a final line
    // without a newline is valid input and the byte after it is never touched
    return gol_op_parse_text_n(d_text, (H - 1) * text_pitch + W, d_packed, W, H, text_pitch, d_err, stream);
}

int gol_op_format_text(const uint64_t *d_packed, char *d_text, int64_t W, int64_t H, int64_t text_pitch,
                       int newline, void *stream)
{
    GOL_TRY(check_dims(W, H));
    if (text_pitch < W + (newline ? 1 : 0)) GOL_FAIL("gol_op_format_text: text pitch too small");
    const int64_t nw32 = 2 * gol_div_up(W, 64);
    unpack_kernel<true><<<seg_grid(H, nw32), kSegThreads, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const uint32_t *>(d_packed), reinterpret_cast<uint8_t *>(d_text), text_pitch, W, nw32,
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
