// Bit-sliced B3/S23 on 64-bit words, shared by the CPU oracle and the generic device kernel.
// Rule and torus wrap follow the serial reference exactly (S:60-101), including the degenerate
// widths/heights 1 and 2 where the wrap makes a cell its own neighbour several times.
#pragma once
#include <cstdint>

#if defined(__CUDACC__)
#define GOL_HD __host__ __device__ __forceinline__
#else
#define GOL_HD static inline
#endif

// word whose bit i is cell (64q + i - 1) of the row (torus in x)
GOL_HD uint64_t gol_row_left(const uint64_t *row, int64_t q, int64_t nq, int64_t W)
{
    const uint64_t cur = row[q];
    const uint64_t carry = q > 0 ? (row[q - 1] >> 63) : ((row[nq - 1] >> ((W - 1) & 63)) & 1ull);
    return (cur << 1) | carry;
}
// word whose bit i is cell (64q + i + 1) of the row (torus in x)
GOL_HD uint64_t gol_row_right(const uint64_t *row, int64_t q, int64_t nq, int64_t W)
{
    const uint64_t cur = row[q];
    if (q < nq - 1) return (cur >> 1) | (row[q + 1] << 63);
    return (cur >> 1) | ((row[0] & 1ull) << ((W - 1) & 63));
}
GOL_HD uint64_t gol_tail_mask(int64_t q, int64_t nq, int64_t W)
{
    const int rem = (int)(W & 63);
    return (rem && q == nq - 1) ? ((1ull << rem) - 1ull) : ~0ull;
}
// 3-cell horizontal sum (centre included) as two bit planes
GOL_HD void gol_hsum(const uint64_t *row, int64_t q, int64_t nq, int64_t W, uint64_t &s0, uint64_t &s1)
{
    const uint64_t l = gol_row_left(row, q, nq, W), c = row[q], r = gol_row_right(row, q, nq, W);
    s0 = l ^ c ^ r;
    s1 = (l & c) | (l & r) | (c & r);
}
// next state from the row sums above (a), at (b), below (n) and the centre word m
GOL_HD uint64_t gol_rule64(uint64_t a0, uint64_t a1, uint64_t b0, uint64_t b1, uint64_t n0, uint64_t n1,
                           uint64_t m)
{
    const uint64_t x0 = a0 ^ b0 ^ n0;
    const uint64_t k0 = (a0 & b0) | (a0 & n0) | (b0 & n0);
    const uint64_t x1 = a1 ^ b1 ^ n1;
    const uint64_t k1 = (a1 & b1) | (a1 & n1) | (b1 & n1);
    const uint64_t y1 = k0 ^ x1, k2 = k0 & x1;
    const uint64_t t3 = x0 & y1 & ~k1 & ~k2;         // 9-cell sum == 3
    const uint64_t t4 = ~x0 & ~y1 & (k1 ^ k2);       // 9-cell sum == 4
    return t3 | (t4 & m);
}
// one output word of generation g+1 from dense packed generation g
GOL_HD uint64_t gol_step_word(const uint64_t *grid, int64_t q, int64_t y, int64_t nq, int64_t W, int64_t H)
{
    const int64_t ym = y == 0 ? H - 1 : y - 1, yp = y == H - 1 ? 0 : y + 1;
    uint64_t a0, a1, b0, b1, n0, n1;
    gol_hsum(grid + ym * nq, q, nq, W, a0, a1);
    gol_hsum(grid + y * nq, q, nq, W, b0, b1);
    gol_hsum(grid + yp * nq, q, nq, W, n0, n1);
    return gol_rule64(a0, a1, b0, b1, n0, n1, grid[y * nq + q]) & gol_tail_mask(q, nq, W);
}
