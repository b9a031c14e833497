// Error plumbing, layout arithmetic, the threaded CPU oracle and host-side text I/O.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cerrno>
#include <fcntl.h>
#include <sys/stat.h>
#include <thread>
#include <unistd.h>
#include <vector>

#include "gol_internal.h"
#include "life_rule.h"

static thread_local std::string g_last_error;

void gol_set_error(const char *fmt, ...)
{
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_last_error = buf;
}

extern "C" const char *gol_last_error(void) { return g_last_error.c_str(); }
extern "C" int gol_version(void) { return GOL_VERSION; }

extern "C" int gol_layout_init(gol_layout *L, int64_t width, int64_t height)
{
    if (!L) GOL_FAIL("gol_layout_init: null layout");
    if (width < 1 || height < 1) GOL_FAIL("grid dimensions must be positive (got %lld x %lld)", (long long)width, (long long)height);
    if (width > (1ll << 36) || height > (1ll << 30)) GOL_FAIL("grid %lld x %lld is too large", (long long)width, (long long)height);
    L->width = width;
    L->height = height;
    L->ghost_rows = GOL_GHOST_ROWS;
    L->nq = (int32_t)gol_div_up(width, 64);
    L->nqf = (int32_t)(width / 64);
    L->rem = (int32_t)(width % 64);
    L->nw32 = 2 * L->nq;
    L->pitch32 = (int32_t)(gol_div_up(2 * GOL_GHOST_WORDS32 + L->nw32, 8) * 8);
    L->rows_padded = height + 2 * GOL_GHOST_ROWS;
    L->buffer_words32 = L->rows_padded * (int64_t)L->pitch32;
    return 0;
}

// --------------------------------------------------------------- CPU oracle
namespace {

int resolve_threads(int threads, int64_t rows)
{
    if (threads <= 0) threads = (int)std::thread::hardware_concurrency();
    if (threads < 1) threads = 1;
    if (threads > rows) threads = (int)rows;
    return threads;
}

template <class F>
void parallel_rows(int64_t H, int threads, F &&fn)
{
    threads = resolve_threads(threads, H);
    if (threads == 1) {
        fn(0, H);
        return;
    }
    std::vector<std::thread> pool;
    for (int t = 0; t < threads; ++t) {
        const int64_t a = H * t / threads, b = H * (t + 1) / threads;
        pool.emplace_back([=, &fn] { fn(a, b); });
    }
    for (auto &th : pool) th.join();
}

// one generation src -> dst; returns {alive(dst), dst != src}
void cpu_generation(const uint64_t *src, uint64_t *dst, int64_t W, int64_t H, int threads, bool *alive, bool *diff)
{
    const int64_t nq = gol_div_up(W, 64);
    std::atomic<int> a{0}, d{0};
    parallel_rows(H, threads, [&](int64_t y0, int64_t y1) {
        uint64_t la = 0, ld = 0;
        for (int64_t y = y0; y < y1; ++y)
            for (int64_t q = 0; q < nq; ++q) {
                const uint64_t v = gol_step_word(src, q, y, nq, W, H);
                dst[y * nq + q] = v;
                la |= v;
                ld |= v ^ src[y * nq + q];
            }
        if (la) a.store(1);
        if (ld) d.store(1);
    });
    if (alive) *alive = a.load() != 0;
    if (diff) *diff = d.load() != 0;
}

bool cpu_any_alive(const uint64_t *g, int64_t n)
{
    for (int64_t i = 0; i < n; ++i)
        if (g[i]) return true;
    return false;
}

}  // namespace

extern "C" int gol_cpu_step(uint64_t *grid, int64_t W, int64_t H, int64_t generations, int threads)
{
    if (W < 1 || H < 1) GOL_FAIL("grid dimensions must be positive");
    const int64_t n = gol_div_up(W, 64) * H;
    std::vector<uint64_t> tmp((size_t)n);
    uint64_t *a = grid, *b = tmp.data();
    for (int64_t g = 0; g < generations; ++g) {
        cpu_generation(a, b, W, H, threads, nullptr, nullptr);
        std::swap(a, b);
    }
    if (a != grid) memcpy(grid, a, (size_t)n * sizeof(uint64_t));
    return 0;
}

// The serial reference loop, literally (S:169-196).
extern "C" int gol_cpu_run(uint64_t *grid, int64_t W, int64_t H, int gen_limit, int similarity_frequency,
                           int threads, gol_result *res)
{
    if (W < 1 || H < 1) GOL_FAIL("grid dimensions must be positive");
    if (gen_limit <= 0) gen_limit = GOL_DEFAULT_GEN_LIMIT;
    if (similarity_frequency < 0) similarity_frequency = GOL_DEFAULT_SIMILARITY_FREQUENCY;
    const int64_t n = gol_div_up(W, 64) * H;
    std::vector<uint64_t> tmp((size_t)n);
    uint64_t *univ = grid, *next = tmp.data();
    int generation = 1, counter = 0, reason = GOL_EXIT_LIMIT;
    const auto t0 = std::chrono::steady_clock::now();
    bool alive = cpu_any_alive(univ, n);
    int64_t computed = 0;
    while (true) {
        if (!alive) {
            reason = GOL_EXIT_EMPTY;
            break;
        }
        if (generation > gen_limit) break;
        bool next_alive, diff;
        cpu_generation(univ, next, W, H, threads, &next_alive, &diff);
        ++computed;
        if (similarity_frequency > 0 && ++counter == similarity_frequency) {
            if (!diff) {
                reason = GOL_EXIT_SIMILAR;
                break;
            }
            counter = 0;
        }
        std::swap(univ, next);
        alive = next_alive;
        ++generation;
    }
    const auto t1 = std::chrono::steady_clock::now();
    if (univ != grid) memcpy(grid, univ, (size_t)n * sizeof(uint64_t));
    if (res) {
        memset(res, 0, sizeof *res);
        res->generations = generation - 1;
        res->exit_reason = reason;
        res->generations_computed = computed;
        res->loop_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
        uint64_t pop = 0;
        for (int64_t i = 0; i < n; ++i) pop += (uint64_t)__builtin_popcountll(grid[i]);
        res->population = pop;
    }
    return 0;
}

// ------------------------------------------------------------ host text I/O
extern "C" int gol_text_parse(const char *text, size_t len, int64_t W, int64_t H, int strict, uint64_t *packed)
{
    if (W < 1 || H < 1) GOL_FAIL("grid dimensions must be positive");
    const int64_t nq = gol_div_up(W, 64);
    memset(packed, 0, (size_t)(nq * H) * sizeof(uint64_t));
    size_t pos = 0;
    for (int64_t y = 0; y < H; ++y) {
        uint64_t *row = packed + y * nq;
        for (int64_t x = 0; x < W;) {
            if (pos >= len)
                GOL_FAIL("input is too short: expected %lld x %lld cells, ran out at row %lld column %lld",
                         (long long)W, (long long)H, (long long)y, (long long)x);
            const char c = text[pos++];
            if (c == '\n' || c == '\r') continue;
            if (c == '1') row[x >> 6] |= 1ull << (x & 63);
            else if (c != '0' && strict)
                GOL_FAIL("invalid byte 0x%02x at offset %zu (row %lld column %lld): only '0', '1' and newlines are accepted",
                         (unsigned)(unsigned char)c, pos - 1, (long long)y, (long long)x);
            ++x;
        }
    }
    return 0;
}

extern "C" int gol_text_format(const uint64_t *packed, int64_t W, int64_t H, char *text)
{
    if (W < 1 || H < 1) GOL_FAIL("grid dimensions must be positive");
    const int64_t nq = gol_div_up(W, 64);
    for (int64_t y = 0; y < H; ++y) {
        const uint64_t *row = packed + y * nq;
        char *out = text + y * (W + 1);
        for (int64_t x = 0; x < W; ++x) out[x] = (char)('0' + ((row[x >> 6] >> (x & 63)) & 1ull));
        out[W] = '\n';
    }
    return 0;
}

namespace {
int read_whole_file(const char *path, std::vector<char> &data)
{
    int fd = open(path, O_RDONLY);
    if (fd < 0) GOL_FAIL("fopen: %s: %s", path, strerror(errno));
    struct stat st;
    if (fstat(fd, &st) != 0) {
        close(fd);
        GOL_FAIL("fstat: %s: %s", path, strerror(errno));
    }
    data.resize((size_t)st.st_size);
    size_t off = 0;
    while (off < data.size()) {
        ssize_t r = pread(fd, data.data() + off, data.size() - off, (off_t)off);
        if (r < 0) {
            if (errno == EINTR) continue;
            close(fd);
            GOL_FAIL("read: %s: %s", path, strerror(errno));
        }
        if (r == 0) break;
        off += (size_t)r;
    }
    data.resize(off);
    close(fd);
    return 0;
}
}  // namespace

int gol_write_all(int fd, const char *buf, size_t len, int64_t offset, const char *path)
{
    size_t off = 0;
    while (off < len) {
        ssize_t w = pwrite(fd, buf + off, len - off, (off_t)(offset + (int64_t)off));
        if (w < 0) {
            if (errno == EINTR) continue;
            GOL_FAIL("write: %s: %s", path, strerror(errno));
        }
        off += (size_t)w;
    }
    return 0;
}

extern "C" int gol_file_read_packed(const char *path, int64_t W, int64_t H, int strict, uint64_t *packed)
{
    std::vector<char> data;
    GOL_TRY(read_whole_file(path, data));
    return gol_text_parse(data.data(), data.size(), W, H, strict, packed);
}

extern "C" int gol_file_write_packed(const char *path, const uint64_t *packed, int64_t W, int64_t H)
{
    int fd = open(path, O_WRONLY | O_CREAT | O_TRUNC, 0644);
    if (fd < 0) GOL_FAIL("fopen: %s: %s", path, strerror(errno));
    const int64_t rows_per = std::max<int64_t>(1, (8 << 20) / (W + 1));
    std::vector<char> buf((size_t)(rows_per * (W + 1)));
    const int64_t nq = gol_div_up(W, 64);
    for (int64_t y = 0; y < H; y += rows_per) {
        const int64_t n = std::min(rows_per, H - y);
        gol_text_format(packed + y * nq, W, n, buf.data());
        if (gol_write_all(fd, buf.data(), (size_t)(n * (W + 1)), y * (W + 1), path)) {
            close(fd);
            return 1;
        }
    }
    close(fd);
    return 0;
}

extern "C" int gol_random_packed(uint64_t *packed, int64_t W, int64_t H, uint64_t seed)
{
    if (W < 1 || H < 1) GOL_FAIL("grid dimensions must be positive");
    const int64_t nq = gol_div_up(W, 64);
    parallel_rows(H, 0, [&](int64_t y0, int64_t y1) {
        for (int64_t y = y0; y < y1; ++y)
            for (int64_t q = 0; q < nq; ++q)
                packed[y * nq + q] = gol_hash64(seed, (uint64_t)q, (uint64_t)y) & gol_tail_mask(q, nq, W);
    });
    return 0;
}

extern "C" int gol_random_text_file(const char *path, int64_t W, int64_t H, uint64_t seed)
{
    if (W < 1 || H < 1) GOL_FAIL("grid dimensions must be positive");
    int fd = open(path, O_WRONLY | O_CREAT | O_TRUNC, 0644);
    if (fd < 0) GOL_FAIL("fopen: %s: %s", path, strerror(errno));
    const int64_t nq = gol_div_up(W, 64);
    const int64_t rows_per = std::max<int64_t>(1, (32 << 20) / (W + 1));
    std::vector<char> buf((size_t)(rows_per * (W + 1)));
    std::vector<uint64_t> words((size_t)(rows_per * nq));
    for (int64_t y = 0; y < H; y += rows_per) {
        const int64_t n = std::min(rows_per, H - y);
        parallel_rows(n, 0, [&](int64_t a, int64_t b) {
            for (int64_t r = a; r < b; ++r) {
                for (int64_t q = 0; q < nq; ++q)
                    words[r * nq + q] = gol_hash64(seed, (uint64_t)q, (uint64_t)(y + r)) & gol_tail_mask(q, nq, W);
                gol_text_format(words.data() + r * nq, W, 1, buf.data() + r * (W + 1));
            }
        });
        if (gol_write_all(fd, buf.data(), (size_t)(n * (W + 1)), y * (W + 1), path)) {
            close(fd);
            return 1;
        }
    }
    close(fd);
    return 0;
}

extern "C" int gol_hash_packed(const uint64_t *packed, int64_t W, int64_t H, uint64_t *hash)
{
    if (W < 1 || H < 1 || !packed || !hash) GOL_FAIL("gol_hash_packed: bad argument");
    const int64_t nq = gol_div_up(W, 64);
    std::atomic<uint64_t> total{0};
    parallel_rows(H, 0, [&](int64_t y0, int64_t y1) {
        uint64_t c = 0;
        for (int64_t y = y0; y < y1; ++y)
            for (int64_t q = 0; q < nq; ++q) c += gol_hash64(packed[y * nq + q], (uint64_t)q, (uint64_t)y);
        total += c;
    });
    *hash = total.load();
    return 0;
}

extern "C" int gol_write_pbm(const char *path, const uint64_t *packed, int64_t W, int64_t H)
{
    if (W < 1 || H < 1) GOL_FAIL("grid dimensions must be positive");
    FILE *f = fopen(path, "wb");
    if (!f) GOL_FAIL("fopen: %s: %s", path, strerror(errno));
    fprintf(f, "P4\n%lld %lld\n", (long long)W, (long long)H);
    const int64_t nq = gol_div_up(W, 64), nb = gol_div_up(W, 8);
    std::vector<unsigned char> row((size_t)nb);
    for (int64_t y = 0; y < H; ++y) {
        for (int64_t b = 0; b < nb; ++b) {
            unsigned v = (unsigned)((packed[y * nq + (b >> 3)] >> ((b & 7) * 8)) & 0xff), r = 0;
            for (int i = 0; i < 8; ++i) r |= ((v >> i) & 1u) << (7 - i);  // PBM is MSB-first
            row[(size_t)b] = (unsigned char)r;
        }
        fwrite(row.data(), 1, (size_t)nb, f);
    }
    fclose(f);
    return 0;
}
