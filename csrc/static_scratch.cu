// gol_engine_population_static: the live-cell count of an engine's grid for callers that must not
// allocate device memory.
//
// gol_engine_population() keeps its counter in device memory that the engine obtains with
// cudaMalloc (the generic dense path also stages through cudaMalloc'ed memory).  Some hosts of this
// library cannot allocate: a process whose device memory has been handed over wholesale to another
// allocator (a framework memory pool that owns the whole GPU, so a raw cudaMalloc fails or fights
// the pool), or a loop with a latency budget, where the implicit device synchronisation of
// cudaMalloc / cudaFree stalls every stream of the process.  This variant runs the same two
// kernels on scratch space that is part of the library image instead.
//
// The scratch is one file-scope __device__ array of GOL_STATIC_SCRATCH_BYTES + 256 bytes.  Every
// GPU on which the library is loaded carries it from module load until the process exits.
#include <cuda_runtime.h>

#include <mutex>

#include "gol_internal.h"

namespace {

constexpr int kMaxDevices = 64;
constexpr size_t kCounterBytes = 256;  // first 256 bytes: the 64-bit counter (keeps the staging area aligned)

// In host code this name is only a handle for the runtime's symbol table, never a device address.
static __device__ __align__(256) unsigned char g_static_scratch[kCounterBytes + GOL_STATIC_SCRATCH_BYTES];

// Puts the caller's current device back when the entry point returns (on every path).
struct DeviceRestore {
    int prev = -1;
    DeviceRestore() { if (cudaGetDevice(&prev) != cudaSuccess) prev = -1; }
    ~DeviceRestore() { if (prev >= 0) cudaSetDevice(prev); }
};

std::mutex g_mutex;
unsigned char *g_device_ptr[kMaxDevices];  // resolved once per device

// Device address of the scratch array on `device` (which must be current).
int static_scratch_ptr(int device, unsigned char **out)
{
    if (device < 0 || device >= kMaxDevices) GOL_FAIL("device ordinal %d out of range", device);
    std::lock_guard<std::mutex> lock(g_mutex);
    if (!g_device_ptr[device]) {
        void *p = nullptr;
        cudaError_t err = cudaGetSymbolAddress(&p, g_static_scratch);
        if (err != cudaSuccess || !p)
            GOL_FAIL("cudaGetSymbolAddress(static scratch) failed on device %d: %s", device, cudaGetErrorString(err));
        g_device_ptr[device] = static_cast<unsigned char *>(p);
    }
    *out = g_device_ptr[device];
    return 0;
}

}  // namespace

extern "C" int gol_engine_population_static(gol_engine *e, uint64_t *count)
{
    if (!e || !count) GOL_FAIL("gol_engine_population_static: null argument");
    const int n = gol_engine_num_local_domains(e);
    // which device holds the grid: ask the runtime about the first domain's buffer
    void *d_first = nullptr;
    gol_layout L0;
    GOL_TRY(gol_engine_domain_buffer(e, 0, &d_first, &L0));
    cudaPointerAttributes attr;
    GOL_CUDA(cudaPointerGetAttributes(&attr, d_first));
    DeviceRestore restore;  // the grid's device is made current below; the caller's comes back on return
    GOL_CUDA(cudaSetDevice(attr.device));

    // refuse before touching anything if a domain does not fit the fixed staging area
    const int ndom = gol_engine_uses_fast_path(e) ? n : 1;
    for (int i = 0; i < ndom; ++i) {
        void *d = nullptr;
        gol_layout L;
        GOL_TRY(gol_engine_domain_buffer(e, i, &d, &L));
        const uint64_t need = (uint64_t)L.nq * (uint64_t)L.height * 8u;
        if (L.ghost_rows != 0 && need > GOL_STATIC_SCRATCH_BYTES)
            GOL_FAIL("gol_engine_population_static: local domain %d (%lld x %lld cells) needs %llu bytes of staging, "
                     "the static scratch holds %u bytes (%u cells); use gol_engine_population",
                     i, (long long)L.width, (long long)L.height, (unsigned long long)need,
                     (unsigned)GOL_STATIC_SCRATCH_BYTES, (unsigned)GOL_STATIC_SCRATCH_BYTES * 8u);
    }

    unsigned char *scratch = nullptr;
    GOL_TRY(static_scratch_ptr(attr.device, &scratch));  // first call on this device resolves the symbol
    unsigned long long *d_count = reinterpret_cast<unsigned long long *>(scratch);
    uint64_t *d_stage = reinterpret_cast<uint64_t *>(scratch + kCounterBytes);

    GOL_TRY(gol_engine_sync(e));  // the engine's own stream has produced the grid
    cudaStream_t stream = cudaStreamPerThread;
    // the array keeps whatever the previous call left: the counter has to start from zero,
    // the staging area is fully overwritten before it is read
    GOL_CUDA(cudaMemsetAsync(d_count, 0, sizeof(unsigned long long), stream));
    for (int i = 0; i < ndom; ++i) {
        void *d = nullptr;
        gol_layout L;
        GOL_TRY(gol_engine_domain_buffer(e, i, &d, &L));
        if (L.ghost_rows == 0) {  // tiny-grid engine: the buffer is already dense
            GOL_TRY(gol_op_popcount_dense(static_cast<const uint64_t *>(d), L.width, L.height, d_count, stream));
        } else {
            GOL_TRY(gol_launch_padded_to_dense(static_cast<const uint32_t *>(d), L, d_stage, L.nq, 0, 0, stream));
            GOL_TRY(gol_op_popcount_dense(d_stage, L.width, L.height, d_count, stream));
        }
    }
    unsigned long long c = 0;
    GOL_CUDA(cudaMemcpyAsync(&c, d_count, sizeof c, cudaMemcpyDeviceToHost, stream));
    GOL_CUDA(cudaStreamSynchronize(stream));
    *count = c;
    return 0;
}
