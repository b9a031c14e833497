// Work-item geometry of the strip kernel: how a domain is cut into strips and chunks, in which
// order the (strip, chunk) items are numbered, and which ghost zones an item touches.
// Host + device: the kernels of csrc/life_kernels.cu use it on the GPU, and
// tests/native/items_check.cpp checks its invariants on the CPU for arbitrary domain shapes.
#pragma once
#include <cstdint>

#include "gol_internal.h"

#if defined(__CUDACC__)
#define GOL_ITEMS_HD __host__ __device__ __forceinline__
#else
#define GOL_ITEMS_HD static inline
#endif

namespace golk {

constexpr int kItemStripWords = 60;   // owned words per strip (= kVW of life_march.cuh)

GOL_ITEMS_HD int items_min(int a, int b) { return a < b ? a : b; }

// Strips are 60 owned words wide.  A launch may also compute `xe` ghost words on each side of
// the row (deep-halo pairs), i.e. the words [-xe, nw + xe).  The last strip is right-aligned to
// the end of that range (it may overlap its left neighbour, which then computes identical words
// twice) so that it always owns the row's last 128 cells, which it pushes east.
GOL_ITEMS_HD int strip_count(int nw, int xe) { return (nw + 2 * xe + kItemStripWords - 1) / kItemStripWords; }
GOL_ITEMS_HD int strip_origin(int strip, int S, int nw, int xe)
{
    return (strip == S - 1 && S > 1) ? nw + xe - kItemStripWords : strip * kItemStripWords - xe;
}

// Work items of a domain are (strip, chunk) pairs.  The computed row range [ylo, yhi) is cut into
// a short first and last chunk of `edge_rows` rows - exactly the rows that are pushed to the north
// and south neighbours, so those items finish (and their flags leave) long before the tall
// interior chunks of the same launch - and interior chunks of `chunk_rows` rows.  Domains shorter
// than three edge chunks are a single chunk.  Items are numbered so that everything that touches
// a ghost zone comes first: the two edge chunks, then the first/last strip of the interior
// chunks, then the interior (or in the reverse order: border_last).
struct ItemMap {
    int S, ylo, yhi, E, rows, ci, nedge;
    int nA, nB, e, n_items;
};
GOL_ITEMS_HD ItemMap item_map(int nw32, int H, const GolLaunch &lp)
{
    ItemMap m;
    m.S = strip_count(nw32, lp.ext_x);
    m.ylo = -lp.ext_y;
    m.yhi = H + lp.ext_y;
    m.rows = lp.chunk_rows;
    const int span = m.yhi - m.ylo;
    if (lp.edge_rows <= 0 || span < 3 * lp.edge_rows) {
        m.E = span; m.nedge = 1; m.ci = 0;
    } else {
        m.E = lp.edge_rows; m.nedge = 2;
        m.ci = (span - 2 * m.E + m.rows - 1) / m.rows;
    }
    m.e = items_min(m.S, 2);
    m.nA = m.nedge * m.S;
    m.nB = m.ci * m.e;
    m.n_items = m.nA + m.nB + m.ci * (m.S - m.e);
    return m;
}
struct Item {
    int strip, y0, y1;
    bool top, bot, first, last;   // chunk reads the north / south ghost rows; strip reads the west / east ghost slot
};
// returns false for idle items (beyond this domain's item count)
GOL_ITEMS_HD bool decode_item(const ItemMap &m, int item, bool border_last, Item &it)
{
    if (item >= m.n_items) return false;
    if (border_last) item = m.n_items - 1 - item;
    int chunk;  // 0 = first edge chunk, 1..ci = interior, ci+1 = last edge chunk
    if (item < m.nA) {
        const int which = item / m.S;
        it.strip = item - which * m.S;
        chunk = which == 0 ? 0 : m.ci + 1;
    } else if (item < m.nA + m.nB) {
        item -= m.nA;
        const int c = item / m.e;
        it.strip = (item - c * m.e) ? m.S - 1 : 0;
        chunk = 1 + c;
    } else {
        item -= m.nA + m.nB;
        const int c = item / (m.S - 2);
        it.strip = 1 + item - c * (m.S - 2);
        chunk = 1 + c;
    }
    if (chunk == 0) { it.y0 = m.ylo; it.y1 = m.ylo + m.E; }
    else if (chunk == m.ci + 1) { it.y0 = m.yhi - m.E; it.y1 = m.yhi; }
    else { it.y0 = m.ylo + m.E + (chunk - 1) * m.rows; it.y1 = items_min(m.yhi - m.E, it.y0 + m.rows); }
    it.top = chunk == 0;
    it.bot = m.nedge == 1 || chunk == m.ci + 1;
    it.first = it.strip == 0;
    it.last = it.strip == m.S - 1;
    return true;
}
// directions (bit d = GOL direction d: N, S, W, E, NW, NE, SW, SE) whose ghost zone the item reads
GOL_ITEMS_HD uint32_t item_dir_mask(const Item &it)
{
    return (it.top ? 1u : 0u) | (it.bot ? 2u : 0u) | (it.first ? 4u : 0u) | (it.last ? 8u : 0u) |
           (it.top && it.first ? 16u : 0u) | (it.top && it.last ? 32u : 0u) | (it.bot && it.first ? 64u : 0u) |
           (it.bot && it.last ? 128u : 0u);
}

}  // namespace golk
