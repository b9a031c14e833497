// Host-side check of the strip kernel's work-item geometry (csrc/life_items.h, the very functions the
// kernels run): for many domain shapes, chunk heights and both epoch kinds every cell of the computed
// region belongs to exactly one item, the items that may touch a ghost zone are exactly the ones flagged
// as such, the flag counters' targets equal the number of items that count, and the numbering puts the
// border items first (or last).  Built and run by tests/test_item_geometry.py; prints "items ok".
#include <cstdio>
#include <cstdlib>
#include <map>
#include <set>
#include <utility>
#include <vector>

#include "../../csrc/life_items.h"

using namespace golk;

static int failures = 0;
#define CHECK(cond, ...)                          \
    do {                                          \
        if (!(cond)) {                            \
            if (++failures < 20) {                \
                printf("FAIL %s: ", #cond);       \
                printf(__VA_ARGS__);              \
                printf("\n");                     \
            }                                     \
        }                                         \
    } while (0)

static void check_case(int nw, int H, int rows, int K, bool first_of_pair)
{
    GolLaunch lp{};
    lp.chunk_rows = rows;
    lp.edge_rows = GOL_EDGE_ROWS;
    lp.ext_y = first_of_pair ? K : 0;
    lp.ext_x = first_of_pair ? 2 : 0;
    const int xe = lp.ext_x, ye = lp.ext_y;
    const ItemMap m = item_map(nw, H, lp);
    // strips cover [-xe, nw + xe) with even origins inside the loadable range
    std::vector<int> covered(nw + 2 * xe, 0);
    for (int s = 0; s < m.S; ++s) {
        const int lo = strip_origin(s, m.S, nw, xe);
        CHECK(lo % 2 == 0, "odd strip origin %d (nw %d xe %d)", lo, nw, xe);
        CHECK(lo - 2 >= -GOL_GHOST_WORDS32 || m.S == 1, "strip %d starts at %d: apron lane outside the ghost slot", s, lo);
        for (int i = lo; i < lo + kItemStripWords; ++i)
            if (i >= -xe && i < nw + xe) covered[i + xe] = 1;
    }
    for (int i = 0; i < nw + 2 * xe; ++i) CHECK(covered[i], "word %d of [%d,%d) in no strip (nw %d)", i - xe, -xe, nw + xe, nw);
    CHECK(strip_origin(m.S - 1, m.S, nw, xe) + kItemStripWords >= nw + xe, "last strip does not reach the row end");

    for (int border_last = 0; border_last < 2; ++border_last) {
        std::map<int, std::vector<std::pair<int, int>>> rows_of_strip;
        std::set<std::pair<int, int>> seen;
        int n_edge = 0, n_we = 0, last_border_pos = -1, first_interior_pos = -1;
        for (int item = 0; item < m.n_items + 5; ++item) {
            Item it;
            const bool ok = decode_item(m, item, border_last != 0, it);
            CHECK(ok == (item < m.n_items), "item %d of %d: decode says %d", item, m.n_items, (int)ok);
            if (!ok) continue;
            CHECK(it.strip >= 0 && it.strip < m.S && it.y0 < it.y1, "item %d: strip %d rows [%d,%d)", item, it.strip, it.y0, it.y1);
            CHECK(seen.insert({it.strip, it.y0}).second, "item (%d,%d) handed out twice", it.strip, it.y0);
            rows_of_strip[it.strip].push_back({it.y0, it.y1});
            // an item that is not flagged must not read a ghost zone
            if (!it.top) CHECK(it.y0 - K >= 0, "item rows [%d,%d) reads north ghosts unflagged (K %d, H %d, rows %d)", it.y0, it.y1, K, H, rows);
            if (!it.bot) CHECK(it.y1 + K <= H, "item rows [%d,%d) reads south ghosts unflagged (K %d, H %d, rows %d)", it.y0, it.y1, K, H, rows);
            CHECK(it.first == (it.strip == 0) && it.last == (it.strip == m.S - 1), "first/last of strip %d", it.strip);
            // a pushing launch (no extension): the rows that go north / south live in flagged items only
            if (!first_of_pair) {
                if (!it.top) CHECK(it.y0 >= GOL_GHOST_ROWS, "rows [%d,%d) hold north push rows unflagged", it.y0, it.y1);
                if (!it.bot) CHECK(it.y1 <= H - GOL_GHOST_ROWS, "rows [%d,%d) hold south push rows unflagged", it.y0, it.y1);
            }
            const uint32_t mask = item_dir_mask(it);
            CHECK(((mask & 1u) != 0) == it.top && ((mask & 2u) != 0) == it.bot && ((mask & 4u) != 0) == it.first &&
                      ((mask & 8u) != 0) == it.last, "direction mask %x", mask);
            CHECK(((mask & 16u) != 0) == (it.top && it.first) && ((mask & 128u) != 0) == (it.bot && it.last), "corner mask %x", mask);
            if (it.top || it.bot) ++n_edge;
            if (it.first || it.last) ++n_we;
            const bool border = it.top || it.bot || it.first || it.last;
            if (border) last_border_pos = item;
            else if (first_interior_pos < 0) first_interior_pos = item;
        }
        CHECK(n_edge == m.nA, "edge-chunk items %d, counter target %d", n_edge, m.nA);
        CHECK(n_we == (m.nedge + m.ci) * m.e, "first/last-strip items %d, counter target %d", n_we, (m.nedge + m.ci) * m.e);
        if (first_interior_pos >= 0 && last_border_pos >= 0) {
            if (!border_last) CHECK(last_border_pos < first_interior_pos, "border item %d after interior item %d", last_border_pos, first_interior_pos);
            // reversed numbering: every interior item precedes every border item
            if (border_last) {
                int first_border = -1;
                for (int item = 0; item < m.n_items; ++item) {
                    Item it;
                    decode_item(m, item, true, it);
                    if (it.top || it.bot || it.first || it.last) { first_border = item; break; }
                }
                for (int item = first_border; item < m.n_items; ++item) {
                    Item it;
                    decode_item(m, item, true, it);
                    CHECK(it.top || it.bot || it.first || it.last, "interior item %d after the first border item %d", item, first_border);
                }
            }
        }
        // per strip: the chunks tile [ylo, yhi) exactly
        CHECK((int)rows_of_strip.size() == m.S, "%d strips have items, %d expected", (int)rows_of_strip.size(), m.S);
        for (auto &kv : rows_of_strip) {
            std::vector<int> hit(H + 2 * ye, 0);
            for (auto &r : kv.second)
                for (int y = r.first; y < r.second; ++y) {
                    CHECK(y >= -ye && y < H + ye, "row %d outside [%d,%d)", y, -ye, H + ye);
                    if (y >= -ye && y < H + ye) ++hit[y + ye];
                }
            for (int y = 0; y < H + 2 * ye; ++y) CHECK(hit[y] == 1, "strip %d row %d covered %d times (H %d rows %d)", kv.first, y - ye, hit[y], H, rows);
        }
    }
    // a shorter domain of the same job never needs more items (launch sizes come from the tallest domain)
    if (H > GOL_MIN_FAST_HEIGHT) CHECK(item_map(nw, H - 1, lp).n_items <= m.n_items, "H-1 has more items than H=%d", H);
}

int main()
{
    srand(12345);
    const int widths[] = {4, 6, 8, 58, 60, 62, 64, 118, 120, 122, 126, 180, 512, 514, 1024, 2048, 4096};
    const int heights[] = {16, 17, 31, 32, 40, 47, 48, 49, 63, 64, 65, 100, 192, 384, 1000, 2048, 8192, 16384};
    long cases = 0;
    for (int nw : widths)
        for (int H : heights)
            for (int K : {1, 2, 4, 8, 16})
                for (int t = 0; t < 6; ++t) {
                    const int min_rows = 8;
                    int rows = t == 0 ? min_rows : (t == 1 ? 85 : (t == 2 ? H : min_rows + rand() % 200));
                    check_case(nw, H, rows, K, false);
                    ++cases;
                    if (2 * K <= GOL_GHOST_ROWS && nw % 2 == 0) {   // deep-halo pair: first epoch computes K ghost rows / 2 ghost words
                        check_case(nw, H, rows, K, true);
                        ++cases;
                    }
                }
    for (int t = 0; t < 3000; ++t) {
        const int nw = 2 * (2 + rand() % 1100), H = 16 + rand() % 3000, rows = 8 + rand() % 300;
        const int K = 1 << (rand() % 4);
        check_case(nw, H, rows, K, false);
        check_case(nw, H, rows, K, true);
        cases += 2;
    }
    if (failures) {
        printf("%d failures in %ld cases\n", failures, cases);
        return 1;
    }
    printf("items ok (%ld cases)\n", cases);
    return 0;
}
