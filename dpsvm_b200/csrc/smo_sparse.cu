// smo_sparse.cu -- building the sparse stream of the owned rows at upload time (layout and summation
// order: SparseDev in smo_kernels.cuh), and the decision whether a context sweeps it or the dense rows.
//
// The reference always multiplies dense rows (cublasSgemv over the whole shard, svmTrain.cu:222,247).
// Its own data sets are mostly zeros -- convert_mnist_to_odd_even.py / convert_adult.py write them out as
// dense CSV -- and the loop is bound by the bytes of X it re-reads every iteration, so the sweep reads
// 6 B per non-zero instead of 4 B per cell when that is the smaller stream.  Zero terms add nothing to a
// dot product; only the summation order of the non-zero terms differs from the dense kernels (both
// orders are fixed and documented; the parity tests run both).
#include <algorithm>
#include <stdlib.h>
#include <string.h>
#include <vector>

#include "dpsvm_internal.cuh"

namespace dpsvm {

#define CU(call)                                                                                     \
    do {                                                                                             \
        cudaError_t e__ = (call);                                                                    \
        if (e__ != cudaSuccess) return fail(e__ == cudaErrorMemoryAllocation ? DPSVM_ERR_NOMEM : DPSVM_ERR_CUDA, \
                                            "%s -> %s", #call, cudaGetErrorString(e__));            \
    } while (0)

// ---- non-zeros per row (warp per row) and their total -------------------------------------------
__global__ void __launch_bounds__(256) sparse_count_kernel(const float* __restrict__ x, int n, int pitch,
                                                           int* __restrict__ row_nnz, unsigned long long* __restrict__ total)
{
    const int lane = threadIdx.x & 31;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int n_warps = (gridDim.x * blockDim.x) >> 5;
    unsigned long long mine = 0;
    for (int row = warp; row < n; row += n_warps) {
        const float* rp = x + (size_t)row * pitch;
        int cnt = 0;
        for (int c0 = 0; c0 < pitch; c0 += 32) {
            const float v = (c0 + lane < pitch) ? __ldg(rp + c0 + lane) : 0.f;
            cnt += __popc(__ballot_sync(0xffffffffu, v != 0.f));
        }
        if (lane == 0) row_nnz[row] = cnt;
        mine += (unsigned long long)cnt;
    }
    if (lane == 0 && mine) atomicAdd(total, mine);
}

// ---- steps per unit: the longest row of the unit over G lanes, in whole chunks (at least one) ----------
__global__ void __launch_bounds__(256) sparse_unit_steps_kernel(const int* __restrict__ row_nnz, int n, int lanes,
                                                                unsigned int* __restrict__ unit_steps, int n_units)
{
    const int groups = 32 / lanes;
    for (int u = blockIdx.x * blockDim.x + threadIdx.x; u < n_units; u += gridDim.x * blockDim.x) {
        int m = 0;
        for (int r = 0; r < groups; ++r) {
            const int row = u * groups + r;
            if (row < n) m = max(m, row_nnz[row]);
        }
        const int steps = (m + lanes - 1) / lanes;
        unit_steps[u] = (unsigned int)(max(1, (steps + kSpChunk - 1) / kSpChunk) * kSpChunk);
    }
}

// ---- exclusive scan, one block (setup time; n_units is at most a few million) ----------------------
__global__ void __launch_bounds__(1024, 1) sparse_scan_kernel(const unsigned int* __restrict__ in, int n, unsigned int* __restrict__ out)
{
    __shared__ unsigned int s_warp[32];
    __shared__ unsigned int s_carry;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_carry = 0;
    __syncthreads();
    for (int base = 0; base < n; base += 1024) {
        const int i = base + tid;
        const unsigned int v = i < n ? in[i] : 0u;
        unsigned int incl = v;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const unsigned int t = __shfl_up_sync(0xffffffffu, incl, off);
            if (lane >= off) incl += t;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            unsigned int w = s_warp[lane];
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const unsigned int t = __shfl_up_sync(0xffffffffu, w, off);
                if (lane >= off) w += t;
            }
            s_warp[lane] = w;           // inclusive over warps
        }
        __syncthreads();
        const unsigned int before = s_carry + (warp ? s_warp[warp - 1] : 0u) + (incl - v);
        if (i < n) out[i] = before;
        __syncthreads();
        if (tid == 1023) s_carry = before + v;
        __syncthreads();
    }
    if (tid == 0) out[n] = s_carry;
}

// ---- fill: warp per row, ascending columns, entry e -> lane e % G of the row's group, step e / G -----
__global__ void __launch_bounds__(256) sparse_fill_kernel(const float* __restrict__ x, int n, int pitch, int lanes,
                                                          const unsigned int* __restrict__ unit_off,
                                                          float* __restrict__ val, unsigned short* __restrict__ col)
{
    const int lane = threadIdx.x & 31;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int n_warps = (gridDim.x * blockDim.x) >> 5;
    const int groups = 32 / lanes;
    for (int row = warp; row < n; row += n_warps) {
        const float* rp = x + (size_t)row * pitch;
        const int unit = row / groups, grp = row % groups;
        const unsigned int s0 = unit_off[unit];
        int base = 0;
        for (int c0 = 0; c0 < pitch; c0 += 32) {
            const int c = c0 + lane;
            const float v = (c < pitch) ? __ldg(rp + c) : 0.f;
            const bool nz = v != 0.f;
            const unsigned int mask = __ballot_sync(0xffffffffu, nz);
            if (nz) {
                const int e = base + __popc(mask & ((1u << lane) - 1u));
                const unsigned int step = s0 + (unsigned int)(e / lanes);
                const int slot = grp * lanes + (e % lanes);
                const size_t at = ((size_t)(step / kSpChunk) * 32 + slot) * kSpChunk + (step % kSpChunk);   // [chunk][lane][8]
                val[at] = v;
                col[at] = (unsigned short)(c * 8);
            }
            base += __popc(mask);
        }
    }
}

// ---- 8-bit column distances from the 16-bit offsets: a warp per unit, every lane walks its own entries in step order ----
// *too_far is raised when a distance does not fit in 8 bits (the context then keeps the 16-bit format).
__global__ void __launch_bounds__(256) sparse_col8_kernel(const float* __restrict__ val, const unsigned short* __restrict__ col,
                                                          const unsigned int* __restrict__ unit_off, int n_units,
                                                          unsigned char* __restrict__ col8, int* __restrict__ too_far)
{
    const int lane = threadIdx.x & 31;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int n_warps = (gridDim.x * blockDim.x) >> 5;
    bool bad = false;
    for (int u = warp; u < n_units; u += n_warps) {
        const unsigned int c0 = unit_off[u] / kSpChunk, c1 = unit_off[u + 1] / kSpChunk;
        int prev = 0;
        for (unsigned int q = c0; q < c1; ++q) {
            const size_t at = ((size_t)q * 32 + lane) * kSpChunk;
            const float4 va = __ldg(reinterpret_cast<const float4*>(val + at)), vb = __ldg(reinterpret_cast<const float4*>(val + at) + 1);
            const uint4 cw = __ldg(reinterpret_cast<const uint4*>(col + at));
            const float v[8] = {va.x, va.y, va.z, va.w, vb.x, vb.y, vb.z, vb.w};
            const unsigned int w[4] = {cw.x, cw.y, cw.z, cw.w};
            unsigned int out[2] = {0u, 0u};
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                int delta = 0;
                if (v[i] != 0.f) {                       // a padding entry (value 0) repeats the previous column
                    const int c = (int)((w[i >> 1] >> (16 * (i & 1))) & 0xFFFFu) >> 3;
                    delta = c - prev;
                    prev = c;
                    if (delta > 255) { bad = true; delta = 0; }
                }
                out[i >> 2] |= (unsigned int)delta << (8 * (i & 3));
            }
            *reinterpret_cast<uint2*>(col8 + at) = make_uint2(out[0], out[1]);
        }
    }
    if (bad) atomicOr(too_far, 1);
}

// ---- |x_i|^2 in the sparse sweep's order, from the stream itself ---------------------------------------
// A warp per unit, every lane walks its own entries exactly as the sweep does (sequential FMAs in step order, the
// value taking the place of both operands), then the same xor-butterfly over the G lanes of a row.
template <int G>
__global__ void __launch_bounds__(256) sparse_row_norms_kernel(const float* __restrict__ val, const unsigned int* __restrict__ unit_off,
                                                               int n_units, int n, float* __restrict__ out)
{
    constexpr int GROUPS = 32 / G;
    const int lane = threadIdx.x & 31;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int n_warps = (gridDim.x * blockDim.x) >> 5;
    for (int u = warp; u < n_units; u += n_warps) {
        const unsigned int c0 = unit_off[u] / kSpChunk, c1 = unit_off[u + 1] / kSpChunk;
        float acc = 0.f;
        for (unsigned int q = c0; q < c1; ++q) {
            const float4* vp = reinterpret_cast<const float4*>(val + ((size_t)q * 32 + lane) * kSpChunk);
            const float4 a = __ldg(vp), b = __ldg(vp + 1);
            acc = __fmaf_rn(a.x, a.x, acc); acc = __fmaf_rn(a.y, a.y, acc); acc = __fmaf_rn(a.z, a.z, acc); acc = __fmaf_rn(a.w, a.w, acc);
            acc = __fmaf_rn(b.x, b.x, acc); acc = __fmaf_rn(b.y, b.y, acc); acc = __fmaf_rn(b.z, b.z, acc); acc = __fmaf_rn(b.w, b.w, acc);
        }
#pragma unroll
        for (int off = G / 2; off >= 1; off >>= 1) acc = __fadd_rn(acc, __shfl_xor_sync(0xffffffffu, acc, off));
        const int row = u * GROUPS + lane / G;
        if (lane % G == 0 && row < n) out[row] = acc;
    }
}

static int warp_blocks(int n)
{
    int blocks = (n + 7) / 8;
    const int cap = device_sm_count() * 8;
    if (blocks > cap) blocks = cap;
    return blocks < 1 ? 1 : blocks;
}

cudaError_t launch_sparse_count(const float* x, int n, int pitch, int* row_nnz, unsigned long long* total, cudaStream_t s)
{
    cudaError_t e = cudaMemsetAsync(total, 0, sizeof(unsigned long long), s);
    if (e != cudaSuccess || n <= 0) return e;
    sparse_count_kernel<<<warp_blocks(n), 256, 0, s>>>(x, n, pitch, row_nnz, total);
    return cudaGetLastError();
}

cudaError_t launch_sparse_unit_steps(const int* row_nnz, int n, int lanes, unsigned int* unit_steps, int n_units, cudaStream_t s)
{
    if (n_units <= 0) return cudaSuccess;
    int blocks = (n_units + 255) / 256;
    const int cap = device_sm_count() * 8;
    if (blocks > cap) blocks = cap;
    sparse_unit_steps_kernel<<<blocks, 256, 0, s>>>(row_nnz, n, lanes, unit_steps, n_units);
    return cudaGetLastError();
}

cudaError_t launch_sparse_scan(const unsigned int* unit_steps, int n_units, unsigned int* unit_off, cudaStream_t s)
{
    sparse_scan_kernel<<<1, 1024, 0, s>>>(unit_steps, n_units, unit_off);
    return cudaGetLastError();
}

cudaError_t launch_sparse_fill(const float* x, int n, int pitch, int lanes, const unsigned int* unit_off,
                               float* val, unsigned short* col, cudaStream_t s)
{
    if (n <= 0) return cudaSuccess;
    sparse_fill_kernel<<<warp_blocks(n), 256, 0, s>>>(x, n, pitch, lanes, unit_off, val, col);
    return cudaGetLastError();
}

cudaError_t launch_sparse_col8(const float* val, const unsigned short* col, const unsigned int* unit_off, int n_units, unsigned char* col8,
                               int* too_far, cudaStream_t s)
{
    cudaError_t e = cudaMemsetAsync(too_far, 0, sizeof(int), s);
    if (e != cudaSuccess || n_units <= 0) return e;
    sparse_col8_kernel<<<warp_blocks(n_units), 256, 0, s>>>(val, col, unit_off, n_units, col8, too_far);
    return cudaGetLastError();
}

cudaError_t launch_sparse_row_norms(const float* val, const unsigned int* unit_off, int n_units, int n, int lanes, float* out, cudaStream_t s)
{
    if (n_units <= 0) return cudaSuccess;
    const int blocks = warp_blocks(n_units);
    switch (lanes) {
        case 4: sparse_row_norms_kernel<4><<<blocks, 256, 0, s>>>(val, unit_off, n_units, n, out); break;
        case 8: sparse_row_norms_kernel<8><<<blocks, 256, 0, s>>>(val, unit_off, n_units, n, out); break;
        case 16: sparse_row_norms_kernel<16><<<blocks, 256, 0, s>>>(val, unit_off, n_units, n, out); break;
        case 32: sparse_row_norms_kernel<32><<<blocks, 256, 0, s>>>(val, unit_off, n_units, n, out); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------
// Per-context setup: decide the layout, build the stream, compute the row norms in the matching order.
// ------------------------------------------------------------------------------------------------
// The stream's buffers live as long as the context and only ever grow: a re-upload (same shape, new data) must not
// pay for cudaFree / cudaMalloc of a hundred megabytes.
void sparse_release(dpsvm_ctx* c)
{
    cudaFree(c->d_sp_val); cudaFree(c->d_sp_col); cudaFree(c->d_sp_col8); cudaFree(c->d_sp_unit_off); cudaFree(c->d_sp_cta_first);
    c->d_sp_col8 = nullptr; c->sp_col8_cap = 0;
    cudaFree(c->d_sp_row_nnz); cudaFree(c->d_sp_unit_steps); cudaFree(c->d_sp_total); cudaFree(c->d_sp_perm);
    c->d_sp_perm = nullptr; c->sp_perm_cap = 0;
    c->d_sp_val = nullptr; c->d_sp_col = nullptr; c->d_sp_unit_off = nullptr; c->d_sp_cta_first = nullptr;
    c->d_sp_row_nnz = nullptr; c->d_sp_unit_steps = nullptr; c->d_sp_total = nullptr;
    c->sp_val_cap = c->sp_col_cap = c->sp_units_cap = 0;
    c->sp = SparseDev{};
    c->sp_nnz = -1;
}

template <class T>
static cudaError_t grow(T*& p, size_t& cap, size_t need)
{
    if (p && cap >= need) return cudaSuccess;
    cudaFree(p);
    p = nullptr;
    cap = 0;
    cudaError_t e = cudaMalloc(&p, need);
    if (e == cudaSuccess) cap = need;
    return e;
}

constexpr double kSparseAutoRatio = 0.7;
constexpr double kSparseMinDenseBytes = 32e6;

static int lanes_for(double mean_nnz)
{
    if (const char* v = getenv("DPSVM_SPARSE_LANES")) {
        const int g = atoi(v);
        if (g == 4 || g == 8 || g == 16 || g == 32) return g;
    }
    // measured at 60 000 x 784, 303 non-zeros per row: 4 and 8 lanes tie (27.5 / 27.2 us per iteration), 16 and 32
    // lose to the per-row epilogue (32.5 / 56.7 us); short rows want few lanes so that a step is not mostly padding
    return mean_nnz <= 32.0 ? 4 : 8;
}

// mode: 0 auto, 1 dense, 2 sparse (DPSVM_FLAG_DENSE / DPSVM_FLAG_SPARSE, or DPSVM_LAYOUT=dense|sparse|auto)
static int layout_mode(const dpsvm_ctx* c)
{
    int mode = (c->p.flags & DPSVM_FLAG_DENSE) ? 1 : ((c->p.flags & DPSVM_FLAG_SPARSE) ? 2 : 0);
    if (const char* v = getenv("DPSVM_LAYOUT")) {
        if (!strcmp(v, "dense")) mode = 1;
        else if (!strcmp(v, "sparse")) mode = 2;
        else if (!strcmp(v, "auto")) mode = 0;
    }
    return mode;
}

int sparse_setup(dpsvm_ctx* c)
{
    c->sp = SparseDev{};
    c->sp_nnz = -1;
    const int mode = layout_mode(c);
    const int n = c->p.shard_rows, pitch = c->pitch;
    const int off = c->p.shard_start - c->res0;
    const float* x_own = c->d_x + (size_t)off * pitch;
    cudaStream_t s = c->stream;
    // auto, small shard: dense rows that fit in the SMs' shared memory (or nearly) never leave the chip in the
    // resident loop and the iteration is bound by its dependency chain, not by bytes -- measured on Adult
    // 32 561 x 123 (16 MB): 0.274 s to converge dense, 0.291 s sparse.  Not worth counting non-zeros for.
    const bool small = 4.0 * (double)n * (double)pitch <= kSparseMinDenseBytes;
    // auto, drop-in layout with the whole of X on every rank (svmTrain.cu:344) but only a shard owned: |x|^2 of the
    // rows other ranks own would be summed in THIS rank's order, which need not be the owner's (the lane count
    // follows the shard's mean row length).  Keep the one order every rank shares.
    // (With DPSVM_FLAG_X_MIRROR the other ranks' rows arrive later, device to device, TOGETHER with the |x|^2 their
    // owners computed: only the owned rows are summed here, and no order is shared.)
    const bool mirror = (c->p.flags & DPSVM_FLAG_X_MIRROR) != 0;
    const bool foreign_rows = !mirror && c->res_n != n;
    auto dense_norms = [&]() -> cudaError_t {
        return mirror ? launch_row_norms(x_own, n, pitch, c->d_xsq + off, s) : launch_row_norms(c->d_x, c->res_n, pitch, c->d_xsq, s);
    };
    if (mode == 1 || pitch > kSparseMaxPitch || (mode == 0 && (small || foreign_rows))) {
        if (mode == 2) return fail(DPSVM_ERR_LIMIT, "sparse layout supports rows of at most %d floats (pitch %d)", kSparseMaxPitch, pitch);
        CU(dense_norms());
        return DPSVM_OK;
    }
    if (!c->d_sp_row_nnz) CU(cudaMalloc(&c->d_sp_row_nnz, sizeof(int) * (size_t)n));
    if (!c->d_sp_total) CU(cudaMalloc(&c->d_sp_total, sizeof(unsigned long long)));
    CU(launch_sparse_count(x_own, n, pitch, c->d_sp_row_nnz, c->d_sp_total, s));
    unsigned long long total = 0;
    CU(cudaMemcpyAsync(&total, c->d_sp_total, sizeof(total), cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    c->sp_nnz = (long long)total;
    const double dense_bytes = 4.0 * (double)n * (double)c->p.num_attributes;
    // auto: sparse when the stream is clearly the smaller one (6 B per non-zero plus padding against 4 B per cell).
    // Measured at 60 000 x 784 with 38.7 % non-zeros: stream 123 MB vs 188 MB of dense rows, 25.7 vs 32.7 us per
    // iteration; the sparse sweep sustains about three quarters of the dense sweep's byte rate, so the break-even
    // is near 0.75 and the threshold sits a little below it.
    if (mode == 0 && 6.0 * (double)total > kSparseAutoRatio * dense_bytes) {
        CU(dense_norms());
        return DPSVM_OK;
    }
    const int lanes = lanes_for((double)total / (double)n);
    const int groups = 32 / lanes;
    const int n_units = (n + groups - 1) / groups;
    {
        // unit_steps and unit_off are sized for the finest units (4 lanes per row would give the fewest; 32 the most)
        const size_t need = sizeof(unsigned int) * ((size_t)n + 1);
        size_t cap_a = c->sp_units_cap, cap_b = c->sp_units_cap;
        CU(grow(c->d_sp_unit_steps, cap_a, need));
        CU(grow(c->d_sp_unit_off, cap_b, need));
        c->sp_units_cap = need;
    }
    CU(launch_sparse_unit_steps(c->d_sp_row_nnz, n, lanes, c->d_sp_unit_steps, n_units, s));
    CU(launch_sparse_scan(c->d_sp_unit_steps, n_units, c->d_sp_unit_off, s));
    std::vector<unsigned int> h_off((size_t)n_units + 1);
    CU(cudaMemcpyAsync(h_off.data(), c->d_sp_unit_off, sizeof(unsigned int) * h_off.size(), cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    const unsigned int total_steps = h_off[n_units];
    const double stream_bytes = 192.0 * (double)total_steps;
    if (mode == 0 && (stream_bytes > kSparseAutoRatio * dense_bytes || total_steps >= kSpMaxStepsO)) {
        CU(dense_norms());
        return DPSVM_OK;
    }
    if (total_steps >= kSpMaxStepsO)
        return fail(DPSVM_ERR_LIMIT, "sparse layout: %u steps of 32 non-zeros in this shard, the sweep addresses fewer than %u", total_steps, kSpMaxStepsO);
    const size_t val_bytes = std::max<size_t>(256, (size_t)total_steps * 32 * sizeof(float));
    const size_t col_bytes = std::max<size_t>(256, (size_t)total_steps * 32 * sizeof(unsigned short));
    CU(grow(c->d_sp_val, c->sp_val_cap, val_bytes));
    CU(grow(c->d_sp_col, c->sp_col_cap, col_bytes));
    CU(cudaMemsetAsync(c->d_sp_val, 0, val_bytes, s));
    CU(cudaMemsetAsync(c->d_sp_col, 0, col_bytes, s));
    CU(launch_sparse_fill(x_own, n, pitch, lanes, c->d_sp_unit_off, c->d_sp_val, c->d_sp_col, s));
    // 8-bit column distances (a sixth fewer bytes per step) whenever every distance fits; DPSVM_SPARSE_COLS=16 keeps the offsets
    bool use_col8 = false;
    {
        const char* cols_env = getenv("DPSVM_SPARSE_COLS");
        if (!(cols_env && atoi(cols_env) == 16)) {
            const size_t col8_bytes = std::max<size_t>(256, (size_t)total_steps * 32);
            CU(grow(c->d_sp_col8, c->sp_col8_cap, col8_bytes + sizeof(int)));
            int* d_flag = reinterpret_cast<int*>(c->d_sp_col8 + ((c->sp_col8_cap - sizeof(int)) & ~(size_t)3));
            CU(launch_sparse_col8(c->d_sp_val, c->d_sp_col, c->d_sp_unit_off, n_units, c->d_sp_col8, d_flag, s));
            int too_far = 1;
            CU(cudaMemcpyAsync(&too_far, d_flag, sizeof(int), cudaMemcpyDeviceToHost, s));
            CU(cudaStreamSynchronize(s));
            use_col8 = too_far == 0;
        }
    }
    // resident loop: contiguous unit ranges per CTA, balanced by steps (rows differ in length)
    const int grid = resident_grid();
    std::vector<int> first((size_t)grid + 1);
    int max_units = 0;
    for (int i = 0; i <= grid; ++i) {
        const unsigned long long target = (unsigned long long)total_steps * (unsigned long long)i / (unsigned long long)grid;
        int u = (int)(std::lower_bound(h_off.begin(), h_off.end(), (unsigned int)target) - h_off.begin());
        if (total_steps == 0) u = (int)((long long)n_units * i / grid);
        if (u > n_units) u = n_units;
        first[i] = u;
    }
    first[0] = 0; first[grid] = n_units;
    for (int i = 0; i < grid; ++i) {
        if (first[i + 1] < first[i]) first[i + 1] = first[i];
        max_units = std::max(max_units, first[i + 1] - first[i]);
    }
    if (!c->d_sp_cta_first) CU(cudaMalloc(&c->d_sp_cta_first, sizeof(int) * 1025));
    if (grid > 1024) return fail(DPSVM_ERR_LIMIT, "resident grid of %d CTAs", grid);
    CU(cudaMemcpyAsync(c->d_sp_cta_first, first.data(), sizeof(int) * first.size(), cudaMemcpyHostToDevice, s));
    // ... and inside a CTA the units are dealt to its 16 warp slots longest first, each to the least loaded slot (units
    // come in whole chunks of 8 steps, typically 5 or 6 of them: seven short units weigh what six long ones do, where a
    // plain round-robin leaves some warps a whole unit behind).  A unit's weight is its steps plus a fixed part for the
    // lane sums and the per-row epilogue.
    constexpr int kSlots = 16;
    constexpr unsigned int kUnitFixedSteps = 8;
    if (max_units > 0xFFFF) return fail(DPSVM_ERR_LIMIT, "%d units in one CTA", max_units);
    std::vector<std::vector<unsigned short>> slot_units((size_t)grid * kSlots);
    int max_cnt = 0;
    {
        std::vector<int> order;
        std::vector<std::pair<unsigned long long, int>> load(kSlots);
        for (int b = 0; b < grid; ++b) {
            const int ub = first[b], m = first[b + 1] - first[b];
            order.resize((size_t)m);
            for (int i = 0; i < m; ++i) order[(size_t)i] = i;
            std::stable_sort(order.begin(), order.end(), [&](int p, int q) {
                return (h_off[(size_t)ub + p + 1] - h_off[(size_t)ub + p]) > (h_off[(size_t)ub + q + 1] - h_off[(size_t)ub + q]); });
            unsigned long long w[kSlots] = {};
            std::vector<unsigned short> tmp[kSlots];
            for (int i : order) {
                int best = 0;
                for (int t = 1; t < kSlots; ++t) if (w[t] < w[best]) best = t;
                tmp[best].push_back((unsigned short)i);
                w[best] += (unsigned long long)(h_off[(size_t)ub + i + 1] - h_off[(size_t)ub + i]) + kUnitFixedSteps;
            }
            for (int t = 0; t < kSlots; ++t) load[(size_t)t] = {w[t], t};
            std::stable_sort(load.begin(), load.end(), [](const std::pair<unsigned long long, int>& p, const std::pair<unsigned long long, int>& q) { return p.first > q.first; });
            for (int t = 0; t < kSlots; ++t) {
                std::vector<unsigned short>& v = tmp[load[(size_t)t].second];
                std::sort(v.begin(), v.end());            // stream order inside a slot (the sweep alternates its direction)
                max_cnt = std::max(max_cnt, (int)v.size());
                slot_units[(size_t)b * kSlots + t] = std::move(v);
            }
        }
    }
    const int perm_stride = kSlots * std::max(1, max_cnt);
    std::vector<unsigned short> h_perm((size_t)grid * perm_stride, 0), h_cnt((size_t)grid * kSlots, 0);
    for (int b = 0; b < grid; ++b)
        for (int t = 0; t < kSlots; ++t) {
            const std::vector<unsigned short>& v = slot_units[(size_t)b * kSlots + t];
            h_cnt[(size_t)b * kSlots + t] = (unsigned short)v.size();
            for (size_t j = 0; j < v.size(); ++j) h_perm[(size_t)b * perm_stride + j * kSlots + t] = v[j];
        }
    {
        size_t cap = c->sp_perm_cap;
        CU(grow(c->d_sp_perm, cap, sizeof(unsigned short) * (h_perm.size() + h_cnt.size())));
        c->sp_perm_cap = cap;
    }
    CU(cudaMemcpyAsync(c->d_sp_perm, h_perm.data(), sizeof(unsigned short) * h_perm.size(), cudaMemcpyHostToDevice, s));
    CU(cudaMemcpyAsync(c->d_sp_perm + h_perm.size(), h_cnt.data(), sizeof(unsigned short) * h_cnt.size(), cudaMemcpyHostToDevice, s));
    // the owned rows are all the rows this rank holds here (a forced sparse layout with foreign rows resident keeps
    // the dense-order norms for those: they only ever serve as the partner's |x|^2 in another row's kernel value)
    if (foreign_rows) CU(launch_row_norms(c->d_x, c->res_n, pitch, c->d_xsq, s));
    CU(launch_sparse_row_norms(c->d_sp_val, c->d_sp_unit_off, n_units, n, lanes, c->d_xsq + off, s));
    CU(cudaStreamSynchronize(s));       // `first` is pageable and goes out of scope
    SparseDev sp{};
    sp.val = c->d_sp_val; sp.col = c->d_sp_col; sp.unit_off = c->d_sp_unit_off; sp.cta_first = c->d_sp_cta_first;
    sp.col8 = use_col8 ? c->d_sp_col8 : nullptr;
    sp.step_bytes = use_col8 ? 160 : 192;
    sp.lanes = lanes; sp.n_units = n_units; sp.grid = grid; sp.max_cta_units = max_units;
    sp.total_steps = total_steps; sp.pin_steps = 0;
    sp.perm = c->d_sp_perm; sp.cnt = c->d_sp_perm + h_perm.size(); sp.perm_stride = perm_stride;
    c->sp = sp;
    return DPSVM_OK;
}

}  // namespace dpsvm
