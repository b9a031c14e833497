// a.out-compatible command line driver.
//
//   ./a.out <width> <height> <input_file> [options]
//
// keeps the reference's public surface (S:224-245, G:150-157): positional width / height /
// input file, atoi parsing with the 30 x 30 default for non-positive values, the per-variant
// stdout lines (S:201-203, M:265/427/466/514, G:294-297) and the per-variant output file
// (./game_output.out, ./mpi_output.out, ./async_output.out, ./collective_output.out,
// ./openmp_output.out, ./cuda_output.out).  Everything after the three positionals is
// optional; the compile-time knobs of the reference (GEN_LIMIT, SIMILARITY_FREQUENCY,
// CHECK_SIMILARITY) become --gen-limit / --similarity-frequency / --no-similarity or the
// GOL_* environment variables.  With --gpus N the binary forks one process per GPU
// (the reference needs mpiexec for that, README.md:50-57); the processes rendezvous through
// an anonymous shared mapping and exchange CUDA IPC handles, no MPI or NCCL involved.
#include <cerrno>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <sys/mman.h>
#include <sys/wait.h>
#include <unistd.h>
#include <vector>

#include "../include/gol_b200.h"

#ifndef GOL_DEFAULT_VARIANT
#define GOL_DEFAULT_VARIANT "game"
#endif

namespace {

struct Options {
    long long width = 0, height = 0;
    std::string input, output, variant = GOL_DEFAULT_VARIANT, pbm;
    int gen_limit = GOL_DEFAULT_GEN_LIMIT;
    int sim_freq = GOL_DEFAULT_SIMILARITY_FREQUENCY;
    int gpus = 1, ranks = 0, grid_r = 0, grid_c = 0, k = 0, batch = 0, start_gen = 0;
    bool generic = false, lenient = false, no_write = false, stats = false, have_random = false, show = false, cpu = false;
    unsigned long long seed = 0;
};

struct Shared {  // rendezvous area shared by the forked ranks
    volatile int barrier_count;
    volatile int barrier_gen;
    volatile int failed;
    double loop_ms[64], read_ms[64], write_ms[64];
    unsigned long long population[64];
    char blobs[64][256];
};

bool g_is_parent = false;   // the process that forked the other ranks (rank 0)
double g_start_ms = 0;      // steady-clock time at the start of main (phase report of --stats)

// Has a forked rank died without reporting (signal, abort, non-zero exit)?  Only the parent can
// see that; WNOWAIT leaves the child for the final waitpid in main().
bool child_died()
{
    siginfo_t info;
    memset(&info, 0, sizeof info);
    if (waitid(P_ALL, 0, &info, WEXITED | WNOHANG | WNOWAIT) != 0 || info.si_pid == 0) return false;
    return info.si_code != CLD_EXITED || info.si_status != 0;
}

void barrier(Shared *sh, int n)
{
    if (n <= 1) return;
    static const double limit_s = getenv("GOL_BARRIER_TIMEOUT_S") ? atof(getenv("GOL_BARRIER_TIMEOUT_S")) : 1800.0;
    const int gen = sh->barrier_gen;
    if (__sync_add_and_fetch(&sh->barrier_count, 1) == n) {
        sh->barrier_count = 0;
        __sync_synchronize();
        __sync_add_and_fetch(&sh->barrier_gen, 1);
    } else {
        const auto t0 = std::chrono::steady_clock::now();
        long spins = 0;
        while (sh->barrier_gen == gen && !sh->failed) {
            usleep(50);
            if ((++spins & 1023) == 0) {
                // a rank killed by a signal never sets `failed`: the parent notices and tells everybody
                if (g_is_parent && child_died()) {
                    fprintf(stderr, "a rank terminated abnormally; aborting the job\n");
                    sh->failed = 1;
                }
                if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > limit_s) {
                    fprintf(stderr, "rank barrier timed out after %.0f s (GOL_BARRIER_TIMEOUT_S)\n", limit_s);
                    sh->failed = 1;
                }
            }
        }
    }
    if (sh->failed) exit(EXIT_FAILURE);  // another rank died: do not run on with half a job
}

[[noreturn]] void die(Shared *sh, const char *what)
{
    fprintf(stderr, "%s: %s\n", what, gol_last_error());
    if (sh) sh->failed = 1;
    exit(EXIT_FAILURE);
}

double now_ms()
{
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

bool is_mpi_variant(const std::string &v) { return v == "mpi" || v == "async" || v == "collective" || v == "openmp"; }

void usage()
{
    fprintf(stderr,
            "usage: a.out <width> <height> <input_file> [options]\n"
            "  --variant game|mpi|async|collective|openmp|cuda   report format + output file name\n"
            "  --output PATH            output file (default ./<variant>_output.out)\n"
            "  --gen-limit N            GEN_LIMIT (default 1000)\n"
            "  --similarity-frequency F SIMILARITY_FREQUENCY (default 3), --no-similarity disables\n"
            "  --gpus N                 one process per GPU (default 1)\n"
            "  --grid RxC               domain decomposition (default: N x 1 slabs up to 16 domains)\n"
            "  --ranks X                number of domains (like mpiexec -n X); may exceed --gpus\n"
            "  --k K --batch B          temporal depth per launch / generations per flag check\n"
            "  --start-generation G     resume: the input file is generation G of an earlier run\n"
            "  --generic                force the generic 1-generation kernel\n"
            "  --cpu                    run on the host (threaded bit-packed CPU implementation); chosen\n"
            "                           automatically when no CUDA device is usable\n"
            "  --lenient                bytes other than 0/1 are dead cells instead of an error\n"
            "  --random SEED            ignore the input file, start from a seeded random grid\n"
            "  --pbm PATH               also dump the final grid as a PBM image (headless show())\n"
            "  --show                   print the final grid with VT100 escape codes like the reference's show()\n"
            "  --no-write --stats\n");
}

// Default decomposition: horizontal slabs (n x 1) for up to 16 domains that leave every slab at least
// 48 rows - full-width rows, so only whole ghost rows travel between neighbours (measured faster than
// column splits, see profiles/README.md); otherwise near-square.
void auto_grid(int n, long long height, int *r, int *c)
{
    if (n <= 16 && height / n >= 48) {
        *r = n;
        *c = 1;
        return;
    }
    int best = 1;
    for (int i = 1; i * i <= n; ++i)
        if (n % i == 0) best = i;
    *r = best;
    *c = n / best;
}

// The whole program on the host: what the reference's serial program does (S:133-222) on a box
// without a GPU, with the library's threaded bit-packed implementation and the same report lines.
int run_cpu(const Options &opt)
{
    const bool mpi_like = is_mpi_variant(opt.variant);
    const long long nq = (opt.width + 63) / 64;
    std::vector<uint64_t> grid((size_t)(nq * opt.height));
    double t0 = now_ms();
    if (opt.have_random) {
        if (gol_random_packed(grid.data(), opt.width, opt.height, opt.seed)) die(nullptr, "randomize");
    } else if (gol_file_read_packed(opt.input.c_str(), opt.width, opt.height, opt.lenient ? 0 : 1, grid.data())) {
        die(nullptr, "input");
    }
    if (mpi_like) printf("Reading file:\t%.2lf msecs\n", now_ms() - t0);
    gol_result res;
    if (opt.start_gen) fprintf(stderr, "note: --start-generation is ignored on the CPU path\n");
    if (gol_cpu_run(grid.data(), opt.width, opt.height, opt.gen_limit, opt.sim_freq, 0, &res)) die(nullptr, "run");
    if (opt.variant == "game") printf("Finished.\n\n");
    if (mpi_like) printf("Generations:\t%d\nExecution time:\t%.2lf msecs\n", res.generations, res.loop_ms);
    else if (opt.variant == "game") printf("Generations:\t%d\nExecution time:\t%.2f msecs\n", res.generations, res.loop_ms);
    fflush(stdout);
    t0 = now_ms();
    if (!opt.no_write && gol_file_write_packed(opt.output.c_str(), grid.data(), opt.width, opt.height)) die(nullptr, "output");
    if (mpi_like) printf("Writing file:\t%.2lf msecs\n", now_ms() - t0);
    if (opt.variant == "cuda") printf("Generations:\t%d\nExecution time:\t%.2f msecs\n", res.generations, res.loop_ms);
    if (opt.stats)
        fprintf(stderr, "[gol] exit=%s generations=%d computed=%lld loop_ms=%.3f population=%llu path=cpu\n",
                res.exit_reason == GOL_EXIT_EMPTY ? "empty" : res.exit_reason == GOL_EXIT_SIMILAR ? "similar" : "limit",
                res.generations, (long long)res.generations_computed, res.loop_ms, (unsigned long long)res.population);
    if (!opt.pbm.empty() && gol_write_pbm(opt.pbm.c_str(), grid.data(), opt.width, opt.height)) die(nullptr, "pbm");
    return 0;
}

int run_rank(const Options &opt, int rank, int nproc, Shared *sh)
{
    const bool mpi_like = is_mpi_variant(opt.variant);
    const bool root = rank == 0;
    gol_config cfg;
    gol_config_default(&cfg, opt.width, opt.height);
    cfg.grid_rows = opt.grid_r;
    cfg.grid_cols = opt.grid_c;
    cfg.proc_rank = rank;
    cfg.proc_count = nproc;
    cfg.device = rank;  // one process per GPU
    cfg.gen_limit = opt.gen_limit;
    cfg.similarity_frequency = opt.sim_freq;
    cfg.k_max = opt.k;
    cfg.batch_gens = opt.batch;
    cfg.start_generation = opt.start_gen;
    cfg.flags = opt.generic ? GOL_FLAG_FORCE_GENERIC : 0;
    gol_engine *eng = nullptr;
    const double t_create0 = now_ms();
    if (gol_engine_create(&cfg, &eng)) {
        // a decomposition finer than the strip kernel's minimum domain: fall back to one domain
        // (only possible when all domains live in this process)
        if (nproc == 1 && cfg.grid_rows * cfg.grid_cols > 1) {
            fprintf(stderr, "note: %s; running as a single domain\n", gol_last_error());
            cfg.grid_rows = cfg.grid_cols = 1;
            if (gol_engine_create(&cfg, &eng)) die(sh, "engine");
        } else {
            die(sh, "engine");
        }
    }
    if (nproc > 1) {
        if (gol_engine_ipc_blob_size(eng) > sizeof sh->blobs[0]) {
            fprintf(stderr, "ipc blob too large\n");
            sh->failed = 1;
            exit(EXIT_FAILURE);
        }
        if (gol_engine_ipc_export(eng, sh->blobs[rank])) die(sh, "ipc export");
        barrier(sh, nproc);
        std::vector<char> all(gol_engine_ipc_blob_size(eng) * (size_t)nproc);
        for (int p = 0; p < nproc; ++p) memcpy(all.data() + gol_engine_ipc_blob_size(eng) * (size_t)p, sh->blobs[p], gol_engine_ipc_blob_size(eng));
        if (gol_engine_ipc_import(eng, all.data())) die(sh, "ipc import");
        barrier(sh, nproc);
    }

    const double create_ms = now_ms() - t_create0;
    double t0 = now_ms();
    if (opt.have_random) {
        if (gol_engine_randomize(eng, opt.seed)) die(sh, "randomize");
    } else if (gol_engine_load_text_file(eng, opt.input.c_str(), opt.lenient ? 0 : 1)) {
        die(sh, "input");
    }
    gol_engine_sync(eng);
    sh->read_ms[rank] = now_ms() - t0;
    barrier(sh, nproc);  // every rank has loaded before ghost zones are pushed
    if (mpi_like && root) {
        double ms = 0;
        for (int p = 0; p < nproc; ++p) ms = std::fmax(ms, sh->read_ms[p]);
        printf("Reading file:\t%.2lf msecs\n", ms);
    }

    const double t_prep0 = now_ms();
    if (gol_engine_prepare(eng)) die(sh, "prepare");
    gol_engine_sync(eng);
    const double prepare_ms = now_ms() - t_prep0;
    gol_result res;
    if (gol_engine_run(eng, &res)) die(sh, "run");
    sh->loop_ms[rank] = res.loop_ms;
    sh->population[rank] = res.population;
    barrier(sh, nproc);
    double loop_ms = 0;
    unsigned long long pop = 0;
    for (int p = 0; p < nproc; ++p) {
        loop_ms = std::fmax(loop_ms, sh->loop_ms[p]);
        pop += sh->population[p];
    }
    if (root) {
        if (opt.variant == "game") printf("Finished.\n\n");
        if (mpi_like) printf("Generations:\t%d\nExecution time:\t%.2lf msecs\n", res.generations, loop_ms);
        else if (opt.variant == "game") printf("Generations:\t%d\nExecution time:\t%.2f msecs\n", res.generations, loop_ms);
        fflush(stdout);
    }

    t0 = now_ms();
    if (!opt.no_write) {
        if (root) unlink(opt.output.c_str());  // replace an existing file (A:432-439)
        barrier(sh, nproc);
        if (gol_engine_store_text_file(eng, opt.output.c_str())) die(sh, "output");
    }
    sh->write_ms[rank] = now_ms() - t0;
    barrier(sh, nproc);
    if (root) {
        double ms = 0;
        for (int p = 0; p < nproc; ++p) ms = std::fmax(ms, sh->write_ms[p]);
        if (mpi_like) printf("Writing file:\t%.2lf msecs\n", ms);
        if (opt.variant == "cuda") printf("Generations:\t%d\nExecution time:\t%.2f msecs\n", res.generations, loop_ms);
        if (opt.stats) {
            const double cells = (double)opt.width * (double)opt.height;
            fprintf(stderr,
                    "[gol] exit=%s generations=%d computed=%lld loop_ms=%.3f cell_updates_per_s=%.4g population=%llu "
                    "fast_path=%d launches=%lld grid=%dx%d chunk_rows=%d deep_halo=%d | phases [ms]: since_start=%.0f create=%.0f "
                    "read=%.0f prepare=%.0f loop=%.1f write=%.0f\n",
                    res.exit_reason == GOL_EXIT_EMPTY ? "empty" : res.exit_reason == GOL_EXIT_SIMILAR ? "similar" : "limit",
                    res.generations, (long long)res.generations_computed, loop_ms,
                    cells * (double)res.generations_computed / (loop_ms * 1e-3), pop, gol_engine_uses_fast_path(eng),
                    (long long)gol_engine_launch_count(eng), cfg.grid_rows, cfg.grid_cols, gol_engine_chunk_rows(eng),
                    gol_engine_uses_deep_halo(eng), now_ms() - g_start_ms, create_ms, sh->read_ms[0], prepare_ms, loop_ms,
                    sh->write_ms[0]);
        }
        if (opt.show && nproc == 1) {
            // the reference's show() (S:42-58, dead code there): VT100 reverse-video blocks on stdout
            std::vector<uint64_t> packed((size_t)((opt.width + 63) / 64 * opt.height));
            if (gol_engine_download_packed(eng, packed.data())) die(sh, "show");
            const long long nq = (opt.width + 63) / 64;
            printf("\033[H");
            for (long long y = 0; y < opt.height; ++y) {
                for (long long x = 0; x < opt.width; ++x)
                    printf(((packed[(size_t)(y * nq + (x >> 6))] >> (x & 63)) & 1ull) ? "\033[07m  \033[m" : "  ");
                printf("\033[E");
            }
            fflush(stdout);
        }
        if (!opt.pbm.empty() && nproc == 1) {
            std::vector<uint64_t> packed((size_t)((opt.width + 63) / 64 * opt.height));
            if (gol_engine_download_packed(eng, packed.data()) || gol_write_pbm(opt.pbm.c_str(), packed.data(), opt.width, opt.height))
                die(sh, "pbm");
        }
    }
    barrier(sh, nproc);
    gol_engine_destroy(eng);
    return 0;
}

}  // namespace

int main(int argc, char **argv)
{
    g_start_ms = now_ms();
    Options opt;
    if (const char *v = getenv("GOL_VARIANT")) opt.variant = v;
    if (const char *v = getenv("GOL_GEN_LIMIT")) opt.gen_limit = atoi(v);
    if (const char *v = getenv("GOL_SIMILARITY_FREQUENCY")) opt.sim_freq = atoi(v);
    if (const char *v = getenv("GOL_GPUS")) opt.gpus = atoi(v);
    if (const char *v = getenv("GOL_RANKS")) opt.ranks = atoi(v);
    if (const char *v = getenv("GOL_GRID")) sscanf(v, "%dx%d", &opt.grid_r, &opt.grid_c);

    // the three positionals, exactly like the reference mains
    int npos = 0;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        auto next = [&](const char *name) -> const char * {
            if (i + 1 >= argc) {
                fprintf(stderr, "%s needs a value\n", name);
                exit(EXIT_FAILURE);
            }
            return argv[++i];
        };
        if (a.rfind("--", 0) != 0) {
            if (npos == 0) opt.width = atoi(argv[i]);
            else if (npos == 1) opt.height = atoi(argv[i]);
            else if (npos == 2) opt.input = argv[i];
            ++npos;
        } else if (a == "--variant") opt.variant = next("--variant");
        else if (a == "--output") opt.output = next("--output");
        else if (a == "--gen-limit") opt.gen_limit = atoi(next("--gen-limit"));
        else if (a == "--similarity-frequency") opt.sim_freq = atoi(next("--similarity-frequency"));
        else if (a == "--no-similarity") opt.sim_freq = 0;
        else if (a == "--gpus") opt.gpus = atoi(next("--gpus"));
        else if (a == "--ranks") opt.ranks = atoi(next("--ranks"));
        else if (a == "--grid") { if (sscanf(next("--grid"), "%dx%d", &opt.grid_r, &opt.grid_c) != 2) { usage(); return EXIT_FAILURE; } }
        else if (a == "--k") opt.k = atoi(next("--k"));
        else if (a == "--batch") opt.batch = atoi(next("--batch"));
        else if (a == "--start-generation") opt.start_gen = atoi(next("--start-generation"));
        else if (a == "--generic") opt.generic = true;
        else if (a == "--cpu") opt.cpu = true;
        else if (a == "--lenient") opt.lenient = true;
        else if (a == "--no-write") opt.no_write = true;
        else if (a == "--stats") opt.stats = true;
        else if (a == "--pbm") opt.pbm = next("--pbm");
        else if (a == "--show") opt.show = true;
        else if (a == "--random") { opt.have_random = true; opt.seed = strtoull(next("--random"), nullptr, 10); }
        else if (a == "--help") { usage(); return 0; }
        else { fprintf(stderr, "unknown option %s\n", a.c_str()); usage(); return EXIT_FAILURE; }
    }
    if (opt.width <= 0) opt.width = 30;
    if (opt.height <= 0) opt.height = 30;
    const bool mpi_like = is_mpi_variant(opt.variant);
    if (mpi_like) opt.height = opt.width;  // M:504, A:480, C:475, O:491
    if (opt.output.empty()) opt.output = "./" + opt.variant + "_output.out";
    if (opt.variant != "game" && opt.variant != "cuda" && !mpi_like) {
        fprintf(stderr, "unknown variant %s\n", opt.variant.c_str());
        return EXIT_FAILURE;
    }

    if ((npos >= 3 || opt.have_random) && !opt.cpu && !gol_cuda_available()) {
        fprintf(stderr, "note: no usable CUDA device; running on the host (--cpu)\n");
        opt.cpu = true;
    }
    if ((npos >= 3 || opt.have_random) && opt.cpu) {
        run_cpu(opt);
    } else if (npos >= 3 || opt.have_random) {
        if (opt.gpus < 1) opt.gpus = 1;
        if (opt.gpus > 64) opt.gpus = 64;
        int ndom = opt.grid_r > 0 && opt.grid_c > 0 ? opt.grid_r * opt.grid_c : (opt.ranks > 0 ? opt.ranks : opt.gpus);
        if (opt.grid_r <= 0 || opt.grid_c <= 0) auto_grid(ndom, opt.height, &opt.grid_r, &opt.grid_c);
        if (ndom % opt.gpus) {
            fprintf(stderr, "%d domains cannot be dealt evenly to %d processes\n", ndom, opt.gpus);
            return EXIT_FAILURE;
        }
        const int nproc = opt.gpus;
        Shared *sh = static_cast<Shared *>(mmap(nullptr, sizeof(Shared), PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0));
        if (sh == MAP_FAILED) {
            perror("mmap: ");
            return EXIT_FAILURE;
        }
        memset(sh, 0, sizeof *sh);
        if (nproc == 1) {
            run_rank(opt, 0, 1, sh);
        } else {
            // fork before any CUDA call: every rank creates its own context on its own GPU
            fflush(stdout);
            g_is_parent = true;
            std::vector<pid_t> kids;
            for (int r = 1; r < nproc; ++r) {
                pid_t pid = fork();
                if (pid < 0) {
                    perror("fork: ");
                    return EXIT_FAILURE;
                }
                if (pid == 0) {
                    g_is_parent = false;
                    run_rank(opt, r, nproc, sh);
                    if (mpi_like && opt.variant != "openmp") printf("Finished\n");
                    fflush(nullptr);
                    _exit(0);
                }
                kids.push_back(pid);
            }
            run_rank(opt, 0, nproc, sh);
            int bad = 0;
            for (pid_t k : kids) {
                int st = 0;
                waitpid(k, &st, 0);
                if (!WIFEXITED(st) || WEXITSTATUS(st) != 0) bad = 1;
            }
            if (bad) return EXIT_FAILURE;
        }
    }
    if (opt.variant != "openmp") printf("Finished\n");  // O:501-502 has this line commented out
    fflush(stdout);
    return 0;
}
