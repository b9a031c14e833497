// Seeded replacement for the reference's generate.sh (generate.sh:3-13): prints <height> lines
// of <width> random '0'/'1' characters to stdout (or to a file with -o), i.i.d. P(1) = 1/2.
// The bash original takes 6 s for 840 x 840 and hours beyond 16384 x 16384; this streams
// gigabytes per second and is reproducible: the grid is a pure function of (seed, x, y).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../include/gol_b200.h"

int main(int argc, char **argv)
{
    long long width = 0, height = 0;
    unsigned long long seed = 1;
    std::string out;
    int npos = 0;
    for (int i = 1; i < argc; ++i) {
        if (!strcmp(argv[i], "-o") && i + 1 < argc) out = argv[++i];
        else if (!strcmp(argv[i], "--seed") && i + 1 < argc) seed = strtoull(argv[++i], nullptr, 10);
        else if (npos == 0) { width = atoll(argv[i]); ++npos; }
        else if (npos == 1) { height = atoll(argv[i]); ++npos; }
        else if (npos == 2) { seed = strtoull(argv[i], nullptr, 10); ++npos; }
    }
    if (width <= 0 || height <= 0) {
        fprintf(stderr, "usage: gol-generate <width> <height> [seed] [-o file]\n");
        return EXIT_FAILURE;
    }
    if (!out.empty()) {
        if (gol_random_text_file(out.c_str(), width, height, seed)) {
            fprintf(stderr, "%s\n", gol_last_error());
            return EXIT_FAILURE;
        }
        return 0;
    }
    const long long nq = (width + 63) / 64;
    const long long rows_per = std::max(1LL, (8LL << 20) / (width + 1));
    std::vector<uint64_t> packed((size_t)(nq * height));
    if (gol_random_packed(packed.data(), width, height, seed)) return EXIT_FAILURE;
    std::vector<char> text((size_t)(rows_per * (width + 1)));
    for (long long y = 0; y < height; y += rows_per) {
        const long long n = std::min(rows_per, height - y);
        gol_text_format(packed.data() + y * nq, width, n, text.data());
        fwrite(text.data(), 1, (size_t)(n * (width + 1)), stdout);
    }
    return 0;
}
