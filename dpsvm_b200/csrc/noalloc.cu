// noalloc.cu -- dpsvm_smo_iteration_noalloc: one complete SMO iteration on caller-owned arrays that
// performs no allocation of any kind.
//
// Who this is for.  Some hosts of this library are not allowed to, or cannot afford to, allocate
// device memory at the moment they want to advance the solver:
//   * a plugin living inside a framework whose caching allocator has already claimed the device --
//     a raw cudaMalloc from the plugin either fails or fragments what the framework thinks it owns;
//   * a service thread that must not stall its neighbours -- cudaMalloc / cudaFree synchronise the
//     whole device, so one allocation in a control loop freezes every other stream for its duration;
//   * code that runs under a memory budget fixed at start-up (all buffers sized and handed out once).
// The trainer context (dpsvm_create) needs a few small work buffers -- the staging area for the two
// selected rows and the iteration state the control kernel hands to the pass -- and gets them from
// cudaMalloc.  Here the same two kernels run with that work area placed in a fixed array that is
// part of the library image instead, so the caller passes only its inputs and outputs.
//
// The array costs kScratchBytes (33 024 bytes) of device memory on every GPU that loads the library,
// from load until process exit, whether or not this entry point is ever called.  That is why the
// limit is modest: rows of at most DPSVM_NOALLOC_MAX_PITCH floats.
#include <limits.h>

#include <mutex>

#include "dpsvm_internal.cuh"

namespace dpsvm {

constexpr int kStateBytes = 256;                                                  // DevState lives in the first 256 bytes
constexpr int kScratchBytes = kStateBytes + 2 * DPSVM_NOALLOC_MAX_PITCH * (int)sizeof(float);
static_assert(sizeof(DevState) <= kStateBytes, "DevState outgrew its slot");

// The work area.  In HOST code this name is only a handle for the runtime (cudaGetSymbolAddress);
// it is not a device address and must never be handed to a kernel as one.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

// The array keeps whatever the previous call left behind.  The control kernel refuses to run when it
// sees done != 0 or num_iter >= stop_at, so the state header is rewritten at the start of every call.
// (Device code may name the array directly.)
__global__ void noalloc_reset_kernel()
{
    DevState* st = reinterpret_cast<DevState*>(g_noalloc_scratch);
    st->num_iter = 0;
    st->done = 0;
    st->converged = 0;
    st->stop_at = INT_MAX;
    st->cache_hits = 0;
    st->cache_lookups = 0;
    st->pass.active = 0;
}

__global__ void noalloc_info_kernel(dpsvm_iteration_info* out)
{
    const DevState* st = reinterpret_cast<const DevState*>(g_noalloc_scratch);
    out->i_hi = st->i_hi;
    out->i_lo = st->i_lo;
    out->b_hi = st->b_hi;
    out->b_lo = st->b_lo;
    out->eta = st->eta;
    out->converged = st->converged;
}

constexpr int kMaxDevices = 64;
static std::mutex g_setup_mutex;
static unsigned char* g_scratch_addr[kMaxDevices] = {};

// Device address of the work area on the CURRENT device; looked up once per device.
static int scratch_address(unsigned char** out)
{
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail(DPSVM_ERR_CUDA, "dpsvm_smo_iteration_noalloc: cudaGetDevice -> %s", cudaGetErrorString(e));
    if (dev < 0 || dev >= kMaxDevices) return fail(DPSVM_ERR_LIMIT, "dpsvm_smo_iteration_noalloc: device ordinal %d beyond %d", dev, kMaxDevices);
    std::lock_guard<std::mutex> lock(g_setup_mutex);
    if (!g_scratch_addr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p)
            return fail(DPSVM_ERR_CUDA, "dpsvm_smo_iteration_noalloc: cudaGetSymbolAddress -> %s", cudaGetErrorString(e));
        g_scratch_addr[dev] = static_cast<unsigned char*>(p);
    }
    *out = g_scratch_addr[dev];
    return DPSVM_OK;
}

}  // namespace dpsvm

using namespace dpsvm;

extern "C" int dpsvm_smo_iteration_noalloc(const dpsvm_iteration_args* g, void* stream)
{
    if (!g || !g->d_x || !g->d_xsq || !g->d_y || !g->d_alpha || !g->d_f || !g->d_keys)
        return fail(DPSVM_ERR_INVALID, "dpsvm_smo_iteration_noalloc: null pointer");
    if (g->n <= 0 || g->d <= 0 || g->pitch < g->d || g->pitch % 4)
        return fail(DPSVM_ERR_INVALID, "dpsvm_smo_iteration_noalloc: bad shape (n > 0, pitch a multiple of 4 and >= d)");
    if ((reinterpret_cast<uintptr_t>(g->d_x) & 15u) != 0)
        return fail(DPSVM_ERR_INVALID, "dpsvm_smo_iteration_noalloc: d_x must be 16-byte aligned (rows are read as 128-bit words)");
    if (g->pitch > DPSVM_NOALLOC_MAX_PITCH)
        return fail(DPSVM_ERR_LIMIT,
                    "dpsvm_smo_iteration_noalloc: row pitch %d floats exceeds the built-in work area (DPSVM_NOALLOC_MAX_PITCH = %d); "
                    "use a trainer context (dpsvm_create) for wider rows",
                    g->pitch, DPSVM_NOALLOC_MAX_PITCH);
    unsigned char* base = nullptr;
    int rc = scratch_address(&base);
    if (rc != DPSVM_OK) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    DevState* st = reinterpret_cast<DevState*>(base);
    float* rows = reinterpret_cast<float*>(base + kStateBytes);

    noalloc_reset_kernel<<<1, 1, 0, s>>>();
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(DPSVM_ERR_CUDA, "dpsvm_smo_iteration_noalloc: reset launch -> %s", cudaGetErrorString(e));

    CtlDev c{};
    c.st = st;
    c.keys = reinterpret_cast<unsigned long long*>(g->d_keys);
    c.x = g->d_x;
    c.pitch = g->pitch;
    c.res0 = g->index_base;
    c.xsq = g->d_xsq;
    c.y = g->d_y;
    c.alpha = g->d_alpha;
    c.sel_rows = rows;
    c.c = g->c;
    c.gamma = g->gamma;
    c.epsilon = g->epsilon;
    c.max_iter = INT_MAX;
    c.cache.lines = 0;
    e = launch_control(c, s);
    if (e != cudaSuccess) return fail(DPSVM_ERR_CUDA, "dpsvm_smo_iteration_noalloc: control launch -> %s", cudaGetErrorString(e));

    PassDev p{};
    p.x4 = reinterpret_cast<const float4*>(g->d_x);
    p.n = g->n;
    p.pitch4 = g->pitch / 4;
    p.xsq = g->d_xsq;
    p.y = g->d_y;
    p.alpha = g->d_alpha;
    p.f = g->d_f;
    p.x_hi4 = reinterpret_cast<const float4*>(rows);
    p.x_lo4 = reinterpret_cast<const float4*>(rows + g->pitch);
    p.gamma = g->gamma;
    p.c = g->c;
    p.index_base = g->index_base;
    p.keys = reinterpret_cast<unsigned long long*>(g->d_keys);
    p.dev = &st->pass;
    p.early_load = 1;
    e = launch_pass(p, 0, s);
    if (e != cudaSuccess) return fail(DPSVM_ERR_CUDA, "dpsvm_smo_iteration_noalloc: pass launch -> %s", cudaGetErrorString(e));

    if (g->d_info) {
        noalloc_info_kernel<<<1, 1, 0, s>>>(g->d_info);
        e = cudaGetLastError();
        if (e != cudaSuccess) return fail(DPSVM_ERR_CUDA, "dpsvm_smo_iteration_noalloc: info launch -> %s", cudaGetErrorString(e));
    }
    return DPSVM_OK;
}
