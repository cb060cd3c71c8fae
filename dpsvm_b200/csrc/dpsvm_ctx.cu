// dpsvm_ctx.cu -- the trainer context behind include/dpsvm_cuda.h: owns one rank's device state,
// builds the CUDA graph of the device-resident SMO loop, and implements the extern "C" entry points.
#include <limits.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <new>
#include <vector>

#include "dpsvm_internal.cuh"

namespace dpsvm {

static thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

#define CU(call)                                                                                     \
    do {                                                                                             \
        cudaError_t e__ = (call);                                                                    \
        if (e__ != cudaSuccess)                                                                      \
            return fail(e__ == cudaErrorMemoryAllocation ? DPSVM_ERR_NOMEM : DPSVM_ERR_CUDA,         \
                        "%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__));       \
    } while (0)

static inline bool misaligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) != 0; }

static int env_int(const char* name, int dflt)
{
    const char* v = getenv(name);
    return (v && *v) ? atoi(v) : dflt;
}

// ---------------------------------------------------------------------------------------------
static void free_all(dpsvm_ctx* c)
{
    if (c->gexec) cudaGraphExecDestroy(c->gexec);
    if (c->graph) cudaGraphDestroy(c->graph);
    cudaFree(c->d_x); cudaFree(c->d_xsq); cudaFree(c->d_y); cudaFree(c->d_alpha); cudaFree(c->d_f);
    cudaFree(c->d_sel_rows); cudaFree(c->d_ext_rows); cudaFree(c->d_keys); cudaFree(c->d_keys_copy); cudaFree(c->d_st);
    cudaFree(c->cache.data); cudaFree(c->cache.line_of_key); cudaFree(c->cache.key_of_line); cudaFree(c->d_cache_stamp);
    cudaFree(c->cache.prev); cudaFree(c->cache.next); cudaFree(c->cache.meta);
    cudaFree(c->d_slots); cudaFree(c->d_barrier); cudaFree(c->d_lru); cudaFree(c->d_trace);
    sparse_release(c);
    predict_free(c);
    comm_free(c);
    if (c->h_st) cudaFreeHost(c->h_st);
    if (c->h_keys) cudaFreeHost(c->h_keys);
    if (c->h_small) cudaFreeHost(c->h_small);
    for (int i = 0; i < 2; ++i) if (c->ev_slot[i]) cudaEventDestroy(c->ev_slot[i]);
    if (c->ev_t0) cudaEventDestroy(c->ev_t0);
    if (c->ev_t1) cudaEventDestroy(c->ev_t1);
    if (c->stream) cudaStreamDestroy(c->stream);
}

PassDev make_pass_args(dpsvm_ctx* c)
{
    PassDev a{};
    const int off = c->p.shard_start - c->res0;   // owned rows inside the resident range
    a.x4 = reinterpret_cast<const float4*>(c->d_x + (size_t)off * c->pitch);
    a.n = c->p.shard_rows;
    a.pitch4 = c->pitch / 4;
    a.xsq = c->d_xsq + off;
    a.y = c->d_y + off;
    a.alpha = c->d_alpha + off;
    a.f = c->d_f;
    a.x_hi4 = reinterpret_cast<const float4*>(c->d_sel_rows);
    a.x_lo4 = reinterpret_cast<const float4*>(c->d_sel_rows + c->pitch);
    a.gamma = c->p.gamma;
    a.c = c->p.c;
    a.index_base = c->p.shard_start;
    a.keys = c->d_keys;
    a.dev = &c->d_st->pass;
    // with a large cache a both-rows-cached iteration (no X sweep) is likely: do not pre-issue X loads
    a.early_load = c->cache.lines < 64;
    a.pdl = c->pdl;
    a.push = c->d_push;
    a.sp = c->sp;
    return a;
}

CtlDev make_ctl_args(dpsvm_ctx* c)
{
    CtlDev a{};
    a.st = c->d_st;
    a.keys = c->d_keys;
    a.x = c->d_x;
    a.pitch = c->pitch;
    a.res0 = c->res0;
    a.xsq = c->d_xsq;
    a.y = c->d_y;
    a.alpha = c->d_alpha;
    a.sel_rows = c->d_sel_rows;
    a.c = c->p.c;
    a.gamma = c->p.gamma;
    a.epsilon = c->p.epsilon;
    a.max_iter = c->p.max_iter;
    a.cache = c->cache;
    a.pdl = c->pdl;
    a.timeout_ns = c->timeout_ns;
    return a;
}

// Variant word of the pass for iteration `parity` (bit 0 reverse, bit 1 L2 hints, >> 8 pinned MB).
// Default policy, from profiles/r01_l2_sweep.md: every SMO iteration re-reads the same X, so
//   * serpentine order -- odd iterations sweep back to front, starting on what L2 still holds;
//   * a 32 MB prefix of X is loaded evict_last, the rest evict_first, once X is too big for L2.
static int variant_for(const dpsvm_ctx* c, int parity)
{
    return c->variant ^ ((c->serpentine && (parity & 1)) ? 1 : 0);
}

static int launch_iteration(dpsvm_ctx* c, cudaStream_t s, int parity)
{
    if (c->world > 1) return comm_launch_iteration(c, s, variant_for(c, parity));
    CU(launch_control(make_ctl_args(c), s));
    CU(launch_pass(make_pass_args(c), variant_for(c, parity), s));
    return DPSVM_OK;
}

static int capture_graph(dpsvm_ctx* c)
{
    CU(cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal));
    int rc = DPSVM_OK;
    for (int i = 0; i < c->chunk && rc == DPSVM_OK; ++i) rc = launch_iteration(c, c->stream, i);
    cudaGraph_t g = nullptr;
    cudaError_t e = cudaStreamEndCapture(c->stream, &g);
    if (rc != DPSVM_OK) { if (g) cudaGraphDestroy(g); return rc; }
    if (e != cudaSuccess) return fail(DPSVM_ERR_CUDA, "graph capture failed: %s", cudaGetErrorString(e));
    e = cudaGraphInstantiate(&c->gexec, g, 0);
    if (e != cudaSuccess) {
        cudaGraphDestroy(g);
        c->gexec = nullptr;
        return fail(DPSVM_ERR_CUDA, "graph instantiate failed: %s", cudaGetErrorString(e));
    }
    c->graph = g;
    return DPSVM_OK;
}

static int build_graph(dpsvm_ctx* c)
{
    if (c->gexec) return DPSVM_OK;
    int rc = capture_graph(c);
    if (rc != DPSVM_OK && c->pdl) {
        // programmatic edges not accepted by this driver in a captured graph: plain edges instead
        cudaGetLastError();
        c->pdl = 0;
        rc = capture_graph(c);
    }
    return rc;
}

static int init_cache(dpsvm_ctx* c);

// The two loops keep the kernel-row cache's bookkeeping in different structures (recency lists +
// direct index for the split kernels, a replicated table for the resident loop).  Switching loops
// mid-run simply empties the cache: results never depend on cache contents.
static int ensure_cache_path(dpsvm_ctx* c, int path)
{
    if (c->cache.lines <= 0 || c->cache_path == path) { c->cache_path = path; return DPSVM_OK; }
    if (c->cache_path != 0) {
        int rc = init_cache(c);
        if (rc != DPSVM_OK) return rc;
    }
    c->cache_path = path;
    return DPSVM_OK;
}

constexpr int kResidentRefused = -1000;     // internal: launch_resident was refused before anything ran

static int launch_resident_chunk(dpsvm_ctx* c)
{
    int rc = ensure_cache_path(c, 2);
    if (rc != DPSVM_OK) return rc;
    CU(cudaMemsetAsync(c->d_barrier, 0, sizeof(unsigned int), c->stream));
    ResidentDev a{};
    const int off = c->p.shard_start - c->res0;
    a.x4 = reinterpret_cast<const float4*>(c->d_x + (size_t)off * c->pitch);
    a.n = c->p.shard_rows;
    a.pitch4 = c->pitch / 4;
    a.xsq = c->d_xsq + off;
    a.y = c->d_y + off;
    a.alpha = c->d_alpha + off;
    a.f = c->d_f;
    a.gamma = c->p.gamma; a.c = c->p.c; a.epsilon = c->p.epsilon;
    a.max_iter = c->p.max_iter;
    a.index_base = c->p.shard_start;
    a.empty_row = c->res0;
    a.st = c->d_st;
    a.keys = c->d_keys;
    a.slots = c->d_slots;
    a.barrier = c->d_barrier;
    a.chunk = c->chunk;
    a.cache_lines = c->cache.lines;
    a.cache_data = c->cache.data;
    a.line_stride = c->cache.line_stride;
    a.lru_key = c->d_lru;
    a.lru_stamp = c->d_lru + kResCacheLines;
    a.lru_meta = c->d_lru + 2 * kResCacheLines;
    a.cache_direct = (c->d_cache_stamp && c->cache.lines >= c->p.num_train_data) ? 1 : 0;
    a.cache_stamp = c->d_cache_stamp;
    a.sp = c->sp;
    a.world = 1;
    a.timeout_ns = c->timeout_ns;
    a.prefetch_mode = env_int("DPSVM_PREFETCH", 0);
    if (c->world > 1 && !comm_fill_resident(c, &a)) return fail(DPSVM_ERR_STATE, "resident loop needs the peer-memory exchange");
    const int trace_on = env_int("DPSVM_TRACE", 0);
    if (trace_on && !c->d_trace) { CU(cudaMalloc(&c->d_trace, sizeof(unsigned long long) * 16 * 1024)); CU(cudaMemset(c->d_trace, 0, sizeof(unsigned long long) * 16 * 1024)); }
    unsigned long long* d_trace = c->d_trace;
    a.trace = trace_on ? d_trace : nullptr;
    a.trace_iter = env_int("DPSVM_TRACE_ITER", 20);
    {
        const cudaError_t e = env_int("DPSVM_FAULT_RESIDENT_LAUNCH", 0) ? cudaErrorCooperativeLaunchTooLarge       // fault injection (tests)
                                                                       : launch_resident(a, variant_for(c, (int)(c->launched_iters & 1)), c->stream);
        if (e != cudaSuccess) {
            // a launch-time refusal (cooperative launch unavailable under MPS/MIG, shared-memory budget, grid that is not
            // co-resident): nothing has run; the caller may fall back to the graph loop, which is bit-identical
            cudaGetLastError();
            fail(DPSVM_ERR_CUDA, "resident loop launch -> %s", cudaGetErrorString(e));
            return kResidentRefused;
        }
    }
    c->launched_iters += c->chunk;
    if (trace_on && c->launched_iters == (long long)c->chunk * trace_on) {
        // debugging aid: phase times of one iteration, per CTA, in us.  Stamps (smo_resident.cu): 0 iteration start,
        // 1 pair rows landed, 2 |x_hi - x_lo|^2 partial sums done, 3 pass scalars ready, 4 sweep done, 5 grid barrier
        // passed, 6 slots merged, 7 candidates pushed (multi-GPU), 8 peers' candidates merged (multi-GPU)
        std::vector<unsigned long long> h(16 * 1024);
        CU(cudaStreamSynchronize(c->stream));
        CU(cudaMemcpy(h.data(), d_trace, sizeof(unsigned long long) * 16 * 1024, cudaMemcpyDeviceToHost));
        const int G = device_sm_count();
        static const char* names[8] = {"rows by TMA", "partial sums", "eta/alpha/LRU", "sweep", "barrier wait", "slot merge", "push", "remote wait+merge"};
        const int phases = c->world > 1 ? 8 : 6;
        double mx[8], mn[8], sum[8];
        for (int k = 0; k < 8; ++k) { mx[k] = 0; mn[k] = 1e30; sum[k] = 0; }
        unsigned long long t_first = ~0ull, t_last = 0;
        for (int b = 0; b < G; ++b) {
            for (int k = 0; k < phases; ++k) {
                if (k == 6 && b != 0) continue;       // only CTA 0 pushes
                const double us = (double)(long long)(h[b * 16 + k + 1] - h[b * 16 + k]) * 1e-3;
                mx[k] = us > mx[k] ? us : mx[k]; mn[k] = us < mn[k] ? us : mn[k]; sum[k] += us;
            }
            t_first = h[b * 16] < t_first ? h[b * 16] : t_first;
            t_last = h[b * 16 + 6] > t_last ? h[b * 16 + 6] : t_last;
        }
        if (const char* dump = getenv("DPSVM_TRACE_DUMP")) {
            if (FILE* fp = fopen(dump, "a")) {
                for (int b = 0; b < G; ++b) {
                    fprintf(fp, "%d %d", a.trace_iter, b);
                    for (int k = 0; k < phases; ++k) fprintf(fp, " %.3f", (double)(long long)(h[b * 16 + k + 1] - h[b * 16 + k]) * 1e-3);
                    fprintf(fp, "\n");
                }
                fclose(fp);
            }
        }
        fprintf(stderr, "[dpsvm trace] rank %d iteration %d over %d CTAs (min/avg/max us):", c->rank, a.trace_iter, G);
        for (int k = 0; k < phases; ++k)
            fprintf(stderr, " %s %.2f/%.2f/%.2f |", names[k], mn[k], k == 6 ? sum[k] : sum[k] / G, mx[k]);
        {
            double mn9 = 1e30, mx9 = 0, sum9 = 0, sp = 0, spmax = 0;
            for (int b = 0; b < G; ++b) {
                const double us = (double)(long long)(h[b * 16 + 9] - h[b * 16 + 4]) * 1e-3;
                mn9 = us < mn9 ? us : mn9; mx9 = us > mx9 ? us : mx9; sum9 += us;
                sp += (double)h[b * 16 + 10]; spmax = (double)h[b * 16 + 10] > spmax ? (double)h[b * 16 + 10] : spmax;
            }
            fprintf(stderr, " [slot stored + prefetch issued %.2f/%.2f/%.2f after sweep; thread 0 re-polled its slot %.1f times on average, %.0f at most] |", mn9, sum9 / G, mx9, sp / G, spmax);
        }
        fprintf(stderr, " first start -> last merge %.2f us\n", (double)(t_last - t_first) * 1e-3);
    }
    return DPSVM_OK;
}

static int launch_chunk(dpsvm_ctx* c)
{
    if (c->loop_resident && (c->world == 1 || c->exchange_p2p)) {
        const int rc = launch_resident_chunk(c);
        if (rc != kResidentRefused) return rc;
        // single GPU: carry on with the graph loop from the same state (the two loops are interchangeable).  Several
        // GPUs: every rank must run the same loop (they share the mailboxes' sequence numbers), so report the error.
        if (c->world > 1) return DPSVM_ERR_CUDA;
        c->loop_resident = 0;
    }
    {
        int rc = ensure_cache_path(c, 1);
        if (rc != DPSVM_OK) return rc;
    }
    if (c->p.flags & DPSVM_FLAG_NO_GRAPH) {
        for (int i = 0; i < c->chunk; ++i) {
            int rc = launch_iteration(c, c->stream, i);
            if (rc != DPSVM_OK) return rc;
        }
        return DPSVM_OK;
    }
    int rc = build_graph(c);
    if (rc != DPSVM_OK) return rc;
    CU(cudaGraphLaunch(c->gexec, c->stream));
    return DPSVM_OK;
}

static int set_stop_at(dpsvm_ctx* c, int stop_at)
{
    // pageable source: cudaMemcpyAsync stages the 4 bytes before returning, so a stack variable is
    // safe and the copy stays ordered with the graph launches on the stream.
    int v = stop_at;
    CU(cudaMemcpyAsync(&c->d_st->stop_at, &v, sizeof(int), cudaMemcpyHostToDevice, c->stream));
    c->stop_at = stop_at;
    return DPSVM_OK;
}

static void fill_status(const DevState& s, float ms, dpsvm_train_status* out)
{
    if (!out) return;
    out->num_iter = s.num_iter;
    out->converged = s.converged;
    out->b_hi = s.b_hi;
    out->b_lo = s.b_lo;
    out->b = (s.b_lo + s.b_hi) / 2;     // svmTrainMain.cpp:329
    out->cache_hits = s.cache_hits;
    out->cache_lookups = s.cache_lookups;
    out->last_chunk_ms = ms;
    out->i_hi = s.i_hi;
    out->i_lo = s.i_lo;
    out->eta = s.eta;
}

static int init_cache(dpsvm_ctx* c)
{
    CacheDev& k = c->cache;
    if (k.lines <= 0) return DPSVM_OK;
    CU(cudaMemsetAsync(k.line_of_key, 0xFF, sizeof(int) * (size_t)c->p.num_train_data, c->stream));
    if (c->d_cache_stamp) CU(cudaMemsetAsync(c->d_cache_stamp, 0, sizeof(int) * (size_t)c->p.num_train_data, c->stream));
    CU(cudaMemsetAsync(k.meta, 0, sizeof(int) * kCacheMetaInts, c->stream));
    CU(cudaMemsetAsync(c->d_lru, 0, sizeof(int) * (2 * kResCacheLines + 2), c->stream));
    const int minus1[2] = {-1, -1};   // pageable source: staged before the call returns
    CU(cudaMemcpyAsync(k.meta + kCacheHead, minus1, 2 * sizeof(int), cudaMemcpyHostToDevice, c->stream));
    return DPSVM_OK;
}

// alpha = 0, f = -y, first selection, state cleared.
static int reset_state(dpsvm_ctx* c)
{
    const int off = c->p.shard_start - c->res0;
    CU(cudaMemsetAsync(c->d_alpha, 0, sizeof(float) * (size_t)c->res_n, c->stream));
    CU(launch_init_state(c->d_y + off, c->p.shard_rows, nullptr, c->d_f, c->stream));
    CU(cudaMemsetAsync(c->d_st, 0, sizeof(DevState), c->stream));
    CU(cudaMemsetAsync(c->d_slots, 0, 2 * 1024 * sizeof(ResSlot), c->stream));     // iteration numbers restart: no stale slot may look valid
    int rc = set_stop_at(c, INT_MAX);
    if (rc != DPSVM_OK) return rc;
    rc = init_cache(c);
    if (rc != DPSVM_OK) return rc;
    c->cache_path = 0;
    CU(launch_select(c->d_alpha + off, c->d_y + off, c->d_f, c->p.shard_rows, c->p.c, c->p.shard_start,
                     c->d_keys, c->stream));
    CU(cudaMemcpyAsync(c->d_keys_copy, c->d_keys, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToDevice, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->target_iter = 0;
    if (c->world > 1) return comm_after_select(c);
    return DPSVM_OK;
}

}  // namespace dpsvm

using namespace dpsvm;

// =================================================================================================
// extern "C"
// =================================================================================================
extern "C" {

const char* dpsvm_last_error(void) { return g_err; }
int dpsvm_abi_version(void) { return DPSVM_ABI_VERSION; }

int dpsvm_row_norms(const float* d_x, int n, int d, int pitch, float* d_xsq, void* stream)
{
    if (!d_x || !d_xsq || n < 0 || d <= 0 || pitch < d || pitch % 4) return fail(DPSVM_ERR_INVALID, "dpsvm_row_norms: bad arguments");
    if (misaligned16(d_x)) return fail(DPSVM_ERR_INVALID, "dpsvm_row_norms: d_x must be 16-byte aligned (rows are read as 128-bit words)");
    CU(launch_row_norms(d_x, n, pitch, d_xsq, (cudaStream_t)stream));
    return DPSVM_OK;
}

int dpsvm_init_state(const int32_t* d_y, int n, float* d_alpha, float* d_f, void* stream)
{
    if (!d_y || !d_f || n < 0) return fail(DPSVM_ERR_INVALID, "dpsvm_init_state: bad arguments");
    CU(launch_init_state(d_y, n, d_alpha, d_f, (cudaStream_t)stream));
    return DPSVM_OK;
}

int dpsvm_select(const float* d_alpha, const int32_t* d_y, const float* d_f, int n, float c, int index_base,
                 uint64_t* d_keys, void* stream)
{
    if (!d_alpha || !d_y || !d_f || !d_keys || n < 0) return fail(DPSVM_ERR_INVALID, "dpsvm_select: bad arguments");
    CU(launch_select(d_alpha, d_y, d_f, n, c, index_base, reinterpret_cast<unsigned long long*>(d_keys), (cudaStream_t)stream));
    return DPSVM_OK;
}

int dpsvm_smo_pass(const dpsvm_pass_args* g, void* stream)
{
    if (!g || !g->d_x || !g->d_xsq || !g->d_y || !g->d_alpha || !g->d_f || !g->d_x_hi || !g->d_x_lo || !g->d_keys)
        return fail(DPSVM_ERR_INVALID, "dpsvm_smo_pass: null pointer");
    if (g->n < 0 || g->d <= 0 || g->pitch < g->d || g->pitch % 4)
        return fail(DPSVM_ERR_INVALID, "dpsvm_smo_pass: bad shape (pitch must be a multiple of 4 and >= d)");
    if (misaligned16(g->d_x) || misaligned16(g->d_x_hi) || misaligned16(g->d_x_lo))
        return fail(DPSVM_ERR_INVALID, "dpsvm_smo_pass: d_x, d_x_hi and d_x_lo must be 16-byte aligned (128-bit loads, TMA bulk copies)");
    PassDev a{};
    a.x4 = reinterpret_cast<const float4*>(g->d_x);
    a.n = g->n; a.pitch4 = g->pitch / 4;
    a.xsq = g->d_xsq; a.y = g->d_y; a.alpha = g->d_alpha; a.f = g->d_f;
    a.x_hi4 = reinterpret_cast<const float4*>(g->d_x_hi);
    a.x_lo4 = reinterpret_cast<const float4*>(g->d_x_lo);
    a.gamma = g->gamma; a.c = g->c; a.index_base = g->index_base;
    a.keys = reinterpret_cast<unsigned long long*>(g->d_keys);
    a.imm.active = 1; a.imm.both_hit = 0;
    a.imm.delta_hi = g->delta_hi; a.imm.delta_lo = g->delta_lo;
    a.imm.xsq_hi = g->xsq_hi; a.imm.xsq_lo = g->xsq_lo;
    a.imm.line_hi = g->d_line_hi; a.imm.line_lo = g->d_line_lo;
    a.dev = nullptr;
    a.early_load = 1;
    CU(launch_pass(a, g->variant, (cudaStream_t)stream));
    return DPSVM_OK;
}

int dpsvm_create(const dpsvm_params* p, dpsvm_ctx** out)
{
    if (!p || !out) return fail(DPSVM_ERR_INVALID, "dpsvm_create: null argument");
    if (p->num_attributes <= 0 || p->num_train_data <= 0) return fail(DPSVM_ERR_INVALID, "dpsvm_create: num_attributes and num_train_data must be positive");
    if (p->shard_rows <= 0 || p->shard_start < 0 || p->shard_start + (long long)p->shard_rows > p->num_train_data)
        return fail(DPSVM_ERR_INVALID, "dpsvm_create: shard [%d,+%d) outside [0,%d)", p->shard_start, p->shard_rows, p->num_train_data);
    if (p->cache_size < 0) return fail(DPSVM_ERR_INVALID, "dpsvm_create: cache_size < 0");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0)
        return fail(DPSVM_ERR_CUDA, "dpsvm_create: no CUDA device (this library has no CPU path)");
    dpsvm_ctx* c = new (std::nothrow) dpsvm_ctx();
    if (!c) return fail(DPSVM_ERR_NOMEM, "dpsvm_create: host allocation failed");
    c->p = *p;
    if (p->device >= 0) {
        cudaError_t e = cudaSetDevice(p->device);
        if (e != cudaSuccess) { delete c; return fail(DPSVM_ERR_CUDA, "cudaSetDevice(%d): %s", p->device, cudaGetErrorString(e)); }
    }
    cudaGetDevice(&c->device);
    c->pitch = dpsvm_pitch_for(p->num_attributes);
    const bool shard_only = (p->flags & DPSVM_FLAG_X_SHARD_ONLY) != 0;
    if ((p->flags & DPSVM_FLAG_X_MIRROR) && !shard_only) {
        delete c;
        return fail(DPSVM_ERR_INVALID, "dpsvm_create: DPSVM_FLAG_X_MIRROR goes with DPSVM_FLAG_X_SHARD_ONLY (the upload brings the owned rows, the peers bring the rest)");
    }
    // rows resident on this GPU: everything (reference layout, or the owned rows plus a mirror of the peers'), or the shard
    const bool keep_shard_only = shard_only && !(p->flags & DPSVM_FLAG_X_MIRROR);
    c->res0 = keep_shard_only ? p->shard_start : 0;
    c->res_n = keep_shard_only ? p->shard_rows : p->num_train_data;
    {
        const int ms = env_int("DPSVM_TIMEOUT_MS", 10000);
        c->timeout_ns = (unsigned long long)(ms > 0 ? ms : 10000) * 1000000ull;
    }
    c->chunk = env_int("DPSVM_CHUNK", 128);
    if (c->chunk < 1) c->chunk = 1;
    c->variant = env_int("DPSVM_VARIANT", dpsvm_pass_variant(p->shard_rows, c->pitch, 0));
    c->serpentine = env_int("DPSVM_SERPENTINE", 1);
    c->pdl = env_int("DPSVM_PDL", 1);
    // loop: the resident (persistent cooperative) kernel whenever it applies; DPSVM_FLAG_GRAPH_LOOP or
    // DPSVM_LOOP=graph keep the [control, pass] CUDA-graph loop (multi-GPU runs always use the graph loop)
    {
        const char* lp = getenv("DPSVM_LOOP");
        const bool want_graph = (p->flags & (DPSVM_FLAG_GRAPH_LOOP | DPSVM_FLAG_NO_GRAPH)) || (lp && strcmp(lp, "graph") == 0);
        c->loop_resident = (!want_graph && resident_supported(c->pitch / 4, p->cache_size >= 2 ? p->cache_size : 0, p->num_train_data)) ? 1 : 0;
    }
    if (c->chunk > 1 && (c->chunk & 1)) c->chunk += 1;   // even chunks keep the direction alternating across launches
    c->world = 1; c->rank = 0;
    // the reference with one line aliases the hi and lo buffers (latent bug, cache.cu:82-105): a
    // single line cannot hold both rows of an iteration, so 1 behaves as "cache off" here.
    c->cache.lines = p->cache_size >= 2 ? p->cache_size : 0;
    if (c->cache.lines > p->num_train_data) c->cache.lines = p->num_train_data;     // one line per row is all there is to cache
    c->cache.line_stride = ((size_t)p->shard_rows + 31) / 32 * 32;

#define CUC(call)                                                                                        \
    do {                                                                                                 \
        cudaError_t e__ = (call);                                                                        \
        if (e__ != cudaSuccess) {                                                                        \
            int code = fail(e__ == cudaErrorMemoryAllocation ? DPSVM_ERR_NOMEM : DPSVM_ERR_CUDA,         \
                            "dpsvm_create: %s -> %s", #call, cudaGetErrorString(e__));                   \
            free_all(c); delete c; return code;                                                          \
        }                                                                                                \
    } while (0)

    CUC(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    CUC(cudaEventCreate(&c->ev_t0));
    CUC(cudaEventCreate(&c->ev_t1));
    for (int i = 0; i < 2; ++i) CUC(cudaEventCreateWithFlags(&c->ev_slot[i], cudaEventDisableTiming));
    CUC(cudaMalloc(&c->d_x, sizeof(float) * (size_t)c->res_n * c->pitch));
    CUC(cudaMalloc(&c->d_xsq, sizeof(float) * (size_t)c->res_n));
    CUC(cudaMalloc(&c->d_y, sizeof(int) * (size_t)c->res_n));
    CUC(cudaMalloc(&c->d_alpha, sizeof(float) * (size_t)c->res_n));
    CUC(cudaMalloc(&c->d_f, sizeof(float) * (size_t)p->shard_rows));
    CUC(cudaMalloc(&c->d_sel_rows, sizeof(float) * 2 * (size_t)c->pitch));
    CUC(cudaMalloc(&c->d_ext_rows, sizeof(float) * 2 * (size_t)c->pitch));
    CUC(cudaMalloc(&c->d_keys, 2 * sizeof(unsigned long long)));
    CUC(cudaMalloc(&c->d_keys_copy, 2 * sizeof(unsigned long long)));
    CUC(cudaMalloc(&c->d_st, sizeof(DevState)));
    CUC(cudaMalloc(&c->d_slots, 2 * 1024 * sizeof(ResSlot)));
    CUC(cudaMalloc(&c->d_barrier, sizeof(unsigned int)));
    CUC(cudaMalloc(&c->d_lru, sizeof(int) * (2 * kResCacheLines + 2)));
    CUC(cudaMallocHost(&c->h_st, 2 * sizeof(DevState)));
    CUC(cudaMallocHost(&c->h_keys, 2 * sizeof(unsigned long long)));
    CUC(cudaMallocHost(&c->h_small, 16 * sizeof(int)));
    if (c->cache.lines > 0) {
        CUC(cudaMalloc(&c->cache.data, sizeof(float) * c->cache.line_stride * (size_t)c->cache.lines));
        CUC(cudaMalloc(&c->cache.line_of_key, sizeof(int) * (size_t)p->num_train_data));
        if (c->cache.lines >= p->num_train_data) CUC(cudaMalloc(&c->d_cache_stamp, sizeof(int) * (size_t)p->num_train_data));
        CUC(cudaMalloc(&c->cache.key_of_line, sizeof(int) * (size_t)c->cache.lines));
        CUC(cudaMalloc(&c->cache.prev, sizeof(int) * (size_t)c->cache.lines));
        CUC(cudaMalloc(&c->cache.next, sizeof(int) * (size_t)c->cache.lines));
        CUC(cudaMalloc(&c->cache.meta, sizeof(int) * kCacheMetaInts));
    }
#undef CUC
    *out = c;
    return DPSVM_OK;
}

int dpsvm_destroy(dpsvm_ctx* c)
{
    if (!c) return DPSVM_OK;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    free_all(c);
    delete c;
    return DPSVM_OK;
}

static int finish_upload(dpsvm_ctx* c)
{
    // a captured graph holds the previous upload's sparse-stream pointers: rebuild it lazily
    if (c->gexec) { cudaGraphExecDestroy(c->gexec); c->gexec = nullptr; }
    if (c->graph) { cudaGraphDestroy(c->graph); c->graph = nullptr; }
    int rc = sparse_setup(c);       // layout decision, sparse stream, |x|^2 in the matching order
    if (rc != DPSVM_OK) return rc;
    if (!getenv("DPSVM_VARIANT")) {
        // L2 policy follows what the sweep actually reads: hints and a pinned prefix once that outgrows L2
        const double mb = c->sp.val ? 1e-6 * (double)c->sp.step_bytes * (double)c->sp.total_steps : 4e-6 * (double)c->p.shard_rows * (double)c->pitch;
        // a stream about the size of L2 is kept there whole (measured at 123 MB: 25.9 us per iteration with 120 MB
        // pinned, 27.2 with 96, 28.3 with 64); a larger one keeps a 32 MB prefix (profiles/r01_l2_sweep.md)
        const int pin_mb = (c->sp.val && mb <= 130.0) ? 120 : 32;
        c->variant = mb > 64.0 ? (2 | (pin_mb << 8)) : 0;
    }
    c->uploaded = true;
    if (c->comm) {                  // DPSVM_FLAG_X_MIRROR: the peers' rows, labels and |x|^2 (collective; a no-op otherwise)
        rc = comm_mirror(c);
        if (rc != DPSVM_OK) return rc;
    }
    return reset_state(c);
}

// Rows an upload brings: all resident rows (reference layout) or the owned shard only (DPSVM_FLAG_X_SHARD_ONLY; with
// DPSVM_FLAG_X_MIRROR the other ranks' rows then arrive device to device).
static void upload_range(const dpsvm_ctx* c, size_t* first_row, size_t* rows)
{
    const bool shard_only = (c->p.flags & DPSVM_FLAG_X_SHARD_ONLY) != 0;
    *first_row = shard_only ? (size_t)(c->p.shard_start - c->res0) : 0;
    *rows = shard_only ? (size_t)c->p.shard_rows : (size_t)c->res_n;
}

int dpsvm_upload(dpsvm_ctx* c, const float* h_x, const int32_t* h_y)
{
    if (!c || !h_x || !h_y) return fail(DPSVM_ERR_INVALID, "dpsvm_upload: null argument");
    CU(cudaSetDevice(c->device));
    const int d = c->p.num_attributes;
    size_t row0, rows;
    upload_range(c, &row0, &rows);
    float* dst = c->d_x + row0 * (size_t)c->pitch;
    if (c->pitch != d) {
        CU(cudaMemsetAsync(dst, 0, sizeof(float) * rows * c->pitch, c->stream));
        CU(cudaMemcpy2DAsync(dst, sizeof(float) * (size_t)c->pitch, h_x, sizeof(float) * (size_t)d, sizeof(float) * (size_t)d,
                             rows, cudaMemcpyHostToDevice, c->stream));
    } else {
        CU(cudaMemcpyAsync(dst, h_x, sizeof(float) * rows * d, cudaMemcpyHostToDevice, c->stream));
    }
    CU(cudaMemcpyAsync(c->d_y + row0, h_y, sizeof(int) * rows, cudaMemcpyHostToDevice, c->stream));
    return finish_upload(c);
}

int dpsvm_upload_device(dpsvm_ctx* c, const float* d_x_shard, int pitch, const int32_t* d_y_shard)
{
    if (!c || !d_x_shard || !d_y_shard) return fail(DPSVM_ERR_INVALID, "dpsvm_upload_device: null argument");
    if (!(c->p.flags & DPSVM_FLAG_X_SHARD_ONLY) && c->p.shard_rows != c->p.num_train_data)
        return fail(DPSVM_ERR_STATE, "dpsvm_upload_device needs DPSVM_FLAG_X_SHARD_ONLY (or a single shard)");
    if (pitch != c->pitch) return fail(DPSVM_ERR_INVALID, "dpsvm_upload_device: pitch %d, expected %d", pitch, c->pitch);
    CU(cudaSetDevice(c->device));
    size_t row0, rows;
    upload_range(c, &row0, &rows);
    CU(cudaMemcpyAsync(c->d_x + row0 * (size_t)c->pitch, d_x_shard, sizeof(float) * rows * c->pitch, cudaMemcpyDeviceToDevice, c->stream));
    CU(cudaMemcpyAsync(c->d_y + row0, d_y_shard, sizeof(int) * rows, cudaMemcpyDeviceToDevice, c->stream));
    return finish_upload(c);
}

int dpsvm_reset(dpsvm_ctx* c)
{
    if (!c) return fail(DPSVM_ERR_INVALID, "dpsvm_reset: null context");
    if (!c->uploaded) return fail(DPSVM_ERR_STATE, "dpsvm_reset before dpsvm_upload");
    CU(cudaSetDevice(c->device));
    return reset_state(c);
}

int dpsvm_step1(dpsvm_ctx* c, float rv[4], uint64_t keys[2])
{
    if (!c || !rv) return fail(DPSVM_ERR_INVALID, "dpsvm_step1: null argument");
    if (!c->uploaded) return fail(DPSVM_ERR_STATE, "dpsvm_step1 before dpsvm_upload");
    // multi-GPU: the exchange consumes the live keys; the last published local selection is kept aside
    CU(cudaMemcpyAsync(c->h_keys, c->world > 1 ? c->d_keys_copy : c->d_keys, 2 * sizeof(unsigned long long),
                       cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    float f_hi, f_lo; int32_t i_hi, i_lo;
    unpack_key_up(c->h_keys[0], &f_hi, &i_hi);
    unpack_key_low(c->h_keys[1], &f_lo, &i_lo);
    // an empty local set reports what the reference's reduction init does: index -1, +-1e9 (svmTrain.cu:312-315)
    if (c->h_keys[0] == DPSVM_KEY_EMPTY) i_hi = -1;
    if (c->h_keys[1] == DPSVM_KEY_EMPTY) i_lo = -1;
    rv[0] = (float)i_hi; rv[1] = (float)i_lo; rv[2] = f_hi; rv[3] = f_lo;     // svmTrain.cu:478-481
    if (keys) { keys[0] = c->h_keys[0]; keys[1] = c->h_keys[1]; }
    return DPSVM_OK;
}

int dpsvm_step2(dpsvm_ctx* c, int i_hi, int i_lo, float alpha_hi_new, float alpha_lo_new,
                const float* h_x_hi, const float* h_x_lo)
{
    if (!c) return fail(DPSVM_ERR_INVALID, "dpsvm_step2: null context");
    if (!c->uploaded) return fail(DPSVM_ERR_STATE, "dpsvm_step2 before dpsvm_upload");
    const int N = c->p.num_train_data;
    if (i_hi < 0 || i_hi >= N || i_lo < 0 || i_lo >= N) return fail(DPSVM_ERR_INVALID, "dpsvm_step2: index out of range");
    const bool hi_res = i_hi >= c->res0 && i_hi < c->res0 + c->res_n;
    const bool lo_res = i_lo >= c->res0 && i_lo < c->res0 + c->res_n;
    if ((!hi_res && !h_x_hi) || (!lo_res && !h_x_lo))
        return fail(DPSVM_ERR_INVALID, "dpsvm_step2: row not resident on this rank and no host row supplied");
    if (!hi_res || !lo_res)
        return fail(DPSVM_ERR_STATE, "dpsvm_step2: non-resident rows need the fused multi-GPU loop (dpsvm_train)");
    CU(cudaSetDevice(c->device));
    {
        int rc = ensure_cache_path(c, 1);
        if (rc != DPSVM_OK) return rc;
    }
    CtlDev a = make_ctl_args(c);
    a.directed = 1; a.dir_i_hi = i_hi; a.dir_i_lo = i_lo; a.dir_a_hi = alpha_hi_new; a.dir_a_lo = alpha_lo_new;
    const int d = c->p.num_attributes;
    if (h_x_hi) {
        CU(cudaMemsetAsync(c->d_ext_rows, 0, sizeof(float) * (size_t)c->pitch, c->stream));
        CU(cudaMemcpyAsync(c->d_ext_rows, h_x_hi, sizeof(float) * (size_t)d, cudaMemcpyHostToDevice, c->stream));
        a.ext_row_hi = c->d_ext_rows;
    }
    if (h_x_lo) {
        CU(cudaMemsetAsync(c->d_ext_rows + c->pitch, 0, sizeof(float) * (size_t)c->pitch, c->stream));
        CU(cudaMemcpyAsync(c->d_ext_rows + c->pitch, h_x_lo, sizeof(float) * (size_t)d, cudaMemcpyHostToDevice, c->stream));
        a.ext_row_lo = c->d_ext_rows + c->pitch;
    }
    CU(launch_control(a, c->stream));
    CU(launch_pass(make_pass_args(c), variant_for(c, c->target_iter), c->stream));
    c->target_iter += 1;
    return DPSVM_OK;
}

int dpsvm_train_async(dpsvm_ctx* c, int iters)
{
    if (!c) return fail(DPSVM_ERR_INVALID, "dpsvm_train_async: null context");
    if (!c->uploaded) return fail(DPSVM_ERR_STATE, "dpsvm_train_async before dpsvm_upload");
    if (iters <= 0) return DPSVM_OK;
    CU(cudaSetDevice(c->device));
    long long target = (long long)c->target_iter + iters;
    if (target > INT_MAX) target = INT_MAX;
    c->target_iter = (int)target;
    int rc = set_stop_at(c, c->target_iter);
    if (rc != DPSVM_OK) return rc;
    const int launches = (iters + c->chunk - 1) / c->chunk;
    for (int i = 0; i < launches; ++i) {
        rc = launch_chunk(c);
        if (rc != DPSVM_OK) return rc;
    }
    return DPSVM_OK;
}

int dpsvm_sync(dpsvm_ctx* c, dpsvm_train_status* status)
{
    if (!c) return fail(DPSVM_ERR_INVALID, "dpsvm_sync: null context");
    CU(cudaSetDevice(c->device));
    CU(cudaMemcpyAsync(&c->h_st[0], c->d_st, sizeof(DevState), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->target_iter = c->h_st[0].num_iter;
    fill_status(c->h_st[0], 0.f, status);
    if (c->h_st[0].error) return fail(DPSVM_ERR_COMM, "multi-GPU exchange timed out waiting for a peer's candidates (iteration %d)", c->h_st[0].num_iter);
    return DPSVM_OK;
}

void* dpsvm_stream(dpsvm_ctx* c) { return c ? (void*)c->stream : nullptr; }

int dpsvm_loop_info(dpsvm_ctx* c, int* resident, int* chunk)
{
    if (!c) return fail(DPSVM_ERR_INVALID, "dpsvm_loop_info: null context");
    if (resident) *resident = (c->loop_resident && (c->world == 1 || c->exchange_p2p)) ? 1 : 0;
    if (chunk) *chunk = c->chunk;
    return DPSVM_OK;
}

int dpsvm_layout_info(dpsvm_ctx* c, int* lanes, long long* nnz, long long* sweep_bytes)
{
    if (!c) return fail(DPSVM_ERR_INVALID, "dpsvm_layout_info: null context");
    if (!c->uploaded) return fail(DPSVM_ERR_STATE, "dpsvm_layout_info before dpsvm_upload");
    if (lanes) *lanes = c->sp.val ? c->sp.lanes : 0;
    if (nnz) *nnz = c->sp_nnz;
    if (sweep_bytes) *sweep_bytes = c->sp.val ? (long long)c->sp.step_bytes * (long long)c->sp.total_steps : 4ll * c->p.shard_rows * (long long)c->pitch;
    return DPSVM_OK;
}

int dpsvm_train(dpsvm_ctx* c, int max_new_iter, dpsvm_train_status* status)
{
    if (!c) return fail(DPSVM_ERR_INVALID, "dpsvm_train: null context");
    if (!c->uploaded) return fail(DPSVM_ERR_STATE, "dpsvm_train before dpsvm_upload");
    CU(cudaSetDevice(c->device));
    dpsvm_train_status cur;
    int rc = dpsvm_sync(c, &cur);
    if (rc != DPSVM_OK) return rc;
    long long target = max_new_iter > 0 ? (long long)cur.num_iter + max_new_iter : (long long)INT_MAX;
    if (target > INT_MAX) target = INT_MAX;
    rc = set_stop_at(c, (int)target);
    if (rc != DPSVM_OK) return rc;
    CU(cudaStreamSynchronize(c->stream));

    // Chunks are launched one ahead of the status read-back so the GPU never waits for the host;
    // iterations past convergence / the pause point are device-side no-ops.
    float ms = 0.f;
    int slot = 0, pending = -1;
    bool finished = c->h_st[0].done || c->h_st[0].num_iter >= (int)target;
    while (!finished) {
        CU(cudaEventRecord(c->ev_t0, c->stream));
        rc = launch_chunk(c);
        if (rc != DPSVM_OK) return rc;
        CU(cudaEventRecord(c->ev_t1, c->stream));
        CU(cudaMemcpyAsync(&c->h_st[slot], c->d_st, sizeof(DevState), cudaMemcpyDeviceToHost, c->stream));
        CU(cudaEventRecord(c->ev_slot[slot], c->stream));
        if (pending >= 0) {
            CU(cudaEventSynchronize(c->ev_slot[pending]));
            const DevState& s = c->h_st[pending];
            if (s.done || s.num_iter >= (int)target) finished = true;
        }
        pending = slot;
        slot ^= 1;
    }
    CU(cudaStreamSynchronize(c->stream));
    if (pending >= 0) cudaEventElapsedTime(&ms, c->ev_t0, c->ev_t1);
    CU(cudaMemcpyAsync(&c->h_st[0], c->d_st, sizeof(DevState), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->target_iter = c->h_st[0].num_iter;
    fill_status(c->h_st[0], ms, status);
    if (c->h_st[0].error) return fail(DPSVM_ERR_COMM, "multi-GPU exchange timed out waiting for a peer's candidates (iteration %d)", c->h_st[0].num_iter);
    return DPSVM_OK;
}

int dpsvm_download(dpsvm_ctx* c, float* h_alpha, float* h_f, float* h_xsq)
{
    if (!c) return fail(DPSVM_ERR_INVALID, "dpsvm_download: null context");
    if (!c->uploaded) return fail(DPSVM_ERR_STATE, "dpsvm_download before dpsvm_upload");
    CU(cudaSetDevice(c->device));
    const int off = c->p.shard_start - c->res0;
    const size_t n = (size_t)c->p.shard_rows;
    if (h_alpha) CU(cudaMemcpyAsync(h_alpha, c->d_alpha + off, sizeof(float) * n, cudaMemcpyDeviceToHost, c->stream));
    if (h_f) CU(cudaMemcpyAsync(h_f, c->d_f, sizeof(float) * n, cudaMemcpyDeviceToHost, c->stream));
    if (h_xsq) CU(cudaMemcpyAsync(h_xsq, c->d_xsq + off, sizeof(float) * n, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return DPSVM_OK;
}

int dpsvm_device_state(dpsvm_ctx* c, const float** d_x, int* pitch, const float** d_xsq, const int32_t** d_y,
                       float** d_alpha, float** d_f)
{
    if (!c) return fail(DPSVM_ERR_INVALID, "dpsvm_device_state: null context");
    const int off = c->p.shard_start - c->res0;
    if (d_x) *d_x = c->d_x + (size_t)off * c->pitch;
    if (pitch) *pitch = c->pitch;
    if (d_xsq) *d_xsq = c->d_xsq + off;
    if (d_y) *d_y = c->d_y + off;
    if (d_alpha) *d_alpha = c->d_alpha + off;
    if (d_f) *d_f = c->d_f;
    return DPSVM_OK;
}

int dpsvm_cache_lookup(dpsvm_ctx* c, int key, float** d_line, int* hit)
{
    if (!c || !d_line || !hit) return fail(DPSVM_ERR_INVALID, "dpsvm_cache_lookup: null argument");
    if (c->cache.lines <= 0) return fail(DPSVM_ERR_STATE, "dpsvm_cache_lookup: cache disabled (cache_size < 2)");
    if (key < 0 || key >= c->p.num_train_data) return fail(DPSVM_ERR_INVALID, "dpsvm_cache_lookup: key out of range");
    CU(cudaSetDevice(c->device));
    int line = -1;
    int rc = ensure_cache_path(c, 1);
    if (rc != DPSVM_OK) return rc;
    rc = cache_lookup_host(c, key, &line, hit);
    if (rc != DPSVM_OK) return rc;
    *d_line = c->cache.data + (size_t)line * c->cache.line_stride;
    return DPSVM_OK;
}

}  // extern "C"
