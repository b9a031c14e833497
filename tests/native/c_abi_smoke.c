/* Native (plain C) smoke test of the C ABI in include/gol_b200.h: CPU oracle, host text I/O and -
 * when a CUDA device is present - the engine, driven exactly as INTEGRATION.md shows.
 * Built and run by tests/test_native_abi.py.  Exit code 0 = pass. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "gol_b200.h"

#define CHECK(x)                                                             \
    do {                                                                     \
        if ((x) != 0) {                                                      \
            fprintf(stderr, "FAIL %s:%d: %s\n", __FILE__, __LINE__, gol_last_error()); \
            return 1;                                                        \
        }                                                                    \
    } while (0)

int main(int argc, char **argv)
{
    const int64_t w = 300, h = 64, nq = (w + 63) / 64;
    uint64_t *a = calloc((size_t)(nq * h), 8), *b = calloc((size_t)(nq * h), 8);
    char *text = malloc((size_t)(h * (w + 1)));
    if (gol_version() < 100) return 1;
    CHECK(gol_random_packed(a, w, h, 11));
    CHECK(gol_text_format(a, w, h, text));
    CHECK(gol_text_parse(text, (size_t)(h * (w + 1)), w, h, 1, b));
    if (memcmp(a, b, (size_t)(nq * h) * 8)) { fprintf(stderr, "text round trip differs\n"); return 1; }
    {   /* a short input is an error (the reference spins forever); it may clobber its output buffer */
        uint64_t *scratch = calloc((size_t)(nq * h), 8);
        if (gol_text_parse(text, 100, w, h, 1, scratch) == 0) { fprintf(stderr, "short input accepted\n"); return 1; }
        free(scratch);
    }
    gol_result cr;
    CHECK(gol_cpu_run(b, w, h, 50, 3, 2, &cr));           /* b := generation 50 */
    printf("cpu: generations=%d population=%llu\n", cr.generations, (unsigned long long)cr.population);

    if (argc > 1 && !strcmp(argv[1], "gpu")) {
        if (!gol_cuda_available()) { fprintf(stderr, "no CUDA device\n"); return 1; }
        gol_config cfg;
        gol_config_default(&cfg, w, h);
        cfg.gen_limit = 50;
        gol_engine *e = NULL;
        CHECK(gol_engine_create(&cfg, &e));
        CHECK(gol_engine_load_packed(e, a));
        CHECK(gol_engine_prepare(e));
        gol_result gr;
        CHECK(gol_engine_run(e, &gr));
        uint64_t *c = calloc((size_t)(nq * h), 8);
        CHECK(gol_engine_download_packed(e, c));
        if (gr.generations != cr.generations || gr.population != cr.population || memcmp(b, c, (size_t)(nq * h) * 8)) {
            fprintf(stderr, "engine result differs from the CPU oracle: generations %d/%d population %llu/%llu grid %s\n",
                    gr.generations, cr.generations, (unsigned long long)gr.population, (unsigned long long)cr.population,
                    memcmp(b, c, (size_t)(nq * h) * 8) ? "differs" : "equal");
            return 1;
        }
        uint64_t pop = 0;
        CHECK(gol_engine_population_static(e, &pop));
        if (pop != gr.population) { fprintf(stderr, "static population differs\n"); return 1; }
        printf("gpu: generations=%d population=%llu loop_ms=%.3f fast_path=%d\n", gr.generations,
               (unsigned long long)gr.population, gr.loop_ms, gol_engine_uses_fast_path(e));
        gol_engine_destroy(e);
        free(c);
    }
    free(a); free(b); free(text);
    printf("c abi ok\n");
    return 0;
}
