// smo_kernels.cu -- hand-written sm_100a kernels of the SMO hot path and their launchers.
//
// Kernels (reference call sites they replace, paths in farshid83/dpsvm):
//   row_norms_kernel     svmTrain.cu:357-364   N x thrust::inner_product in a host loop
//   init_state_kernel    svmTrain.cu:349,379-380
//   select_kernel        svmTrain.cu:469-476   for_each(arbitrary_functor) + reduce(my_maxmin)
//   smo_pass_kernel      svmTrain.cu:222,247 (2 x cublasSgemv) + :276-278 (update_functor) +
//                        the next iteration's :469-476, fused into ONE sweep over the X shard
//   smo_control_kernel   svmTrainMain.cpp:246-302 (host merge, eta, alpha update, clip) +
//                        svmTrain.cu:142-156,487-494 / cache.cu:62-105 (cache lookup, alpha install)
//
// The pass is HBM-bound (2 dot products per row = GEMM with N=2, ~1 flop/byte): no tensor cores.
#include "smo_device.cuh"

#include <math.h>

namespace dpsvm {


// Stage the two selected rows in shared memory with TMA bulk copies and zero the tail up to `cap4`.
__device__ __forceinline__ void stage_selected_rows(float4* s_hi, float4* s_lo, const float4* g_hi, const float4* g_lo,
                                                    int pitch4, int cap4, uint64_t* bar)
{
    const int tid = threadIdx.x;
    if (tid == 0) mbar_init(bar, 1);
    __syncthreads();
    if (tid == 0) {
        const uint32_t bytes = (uint32_t)pitch4 * 16u;
        mbar_expect_tx(bar, 2u * bytes);
        tma_bulk_g2s(s_hi, g_hi, bytes, bar);
        tma_bulk_g2s(s_lo, g_lo, bytes, bar);
    }
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int i = pitch4 + tid; i < cap4; i += blockDim.x) { s_hi[i] = z; s_lo[i] = z; }
    mbar_wait(bar, 0);
    __syncthreads();
}

// Reduce the per-thread best keys over the block and publish with one atomicMin per key.
__device__ __forceinline__ void publish_keys(unsigned long long best_up, unsigned long long best_low,
                                             unsigned long long* keys, unsigned long long* s_red)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    best_up = warp_min_u64(best_up);
    best_low = warp_min_u64(best_low);
    if (lane == 0) { s_red[warp] = best_up; s_red[nwarps + warp] = best_low; }
    __syncthreads();
    if (warp == 0) {
        const unsigned long long u = warp_min_u64(lane < nwarps ? s_red[lane] : DPSVM_KEY_EMPTY);
        const unsigned long long l = warp_min_u64(lane < nwarps ? s_red[nwarps + lane] : DPSVM_KEY_EMPTY);
        if (lane == 0) {
            if (u != DPSVM_KEY_EMPTY) atomicMin(&keys[0], u);
            if (l != DPSVM_KEY_EMPTY) atomicMin(&keys[1], l);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Peer-memory candidate exchange (multi-GPU).  Runs in ONE block: builds this rank's package from
// the final local keys and stores it into slot (seq & 1) of every rank's mailbox -- stores to
// CUDA-IPC-mapped peer memory travel over NVLink -- then raises this rank's flag there.
// ------------------------------------------------------------------------------------------------

__device__ void pack_and_push(const PushDev& p)
{
    // final keys of this pass: every CTA fenced after its atomicMin and before taking its ticket
    const unsigned long long k_up = __ldcg(p.keys), k_low = __ldcg(p.keys + 1);
    const unsigned int seq = (unsigned int)__ldcg(&p.st->num_iter);
    push_package(p, k_up, k_low, 0.f, 0.f, false, seq);
    if (threadIdx.x == 0) {
        p.keys_copy[0] = k_up; p.keys_copy[1] = k_low;      // what train_step1 reports for this rank
        p.keys[0] = DPSVM_KEY_EMPTY; p.keys[1] = DPSVM_KEY_EMPTY;
    }
}

// After publish_keys: the last CTA of the pass to get here publishes the candidates.
__device__ __forceinline__ void push_if_last(const PushDev* push, int* s_flag)
{
    if (!push) return;
    if (threadIdx.x == 0) {
        __threadfence();                                   // my atomicMin before my ticket
        const unsigned int t = atomicAdd(push->counter, 1u);
        const int last = (t == gridDim.x - 1);
        if (last) { *push->counter = 0; __threadfence(); } // see everyone's atomicMin; re-arm for the next pass
        *s_flag = last;
    }
    __syncthreads();
    if (*s_flag) pack_and_push(*push);
}

__global__ void __launch_bounds__(kCtlThreads, 1) smo_push_kernel(const PushDev* push)
{
    pack_and_push(*push);
}

// The per-row tail shared by every pass variant: kernel values, f update (update_functor,
// svmTrain.cu:128-135), optional cache-line store, classification (arbitrary_functor, :53-91) and
// the running packed-key minima.
__device__ __forceinline__ void row_epilogue(int row, float dot_hi, float dot_lo, float xsq_j, float f_j,
                                             float alpha_j, int y_j, const PassScalars& sc, const PassDev& a,
                                             bool store_lines, unsigned long long& best_up, unsigned long long& best_low)
{
    if (store_lines) {
        if (sc.line_hi) sc.line_hi[row] = dot_hi;
        if (sc.line_lo) sc.line_lo[row] = dot_lo;
    }
    // expf(-1 * gamma * (xsq_j + x_sel_sq - (2*dot)))      svmTrain.cu:128,130
    const float arg_hi = __fsub_rn(__fadd_rn(xsq_j, sc.xsq_hi), __fmul_rn(2.0f, dot_hi));
    const float arg_lo = __fsub_rn(__fadd_rn(xsq_j, sc.xsq_lo), __fmul_rn(2.0f, dot_lo));
    const float rbf_hi = expf(__fmul_rn(-a.gamma, arg_hi));
    const float rbf_lo = expf(__fmul_rn(-a.gamma, arg_lo));
    // delta = (d_alpha_hi*y_hi*rbf_hi) + (d_alpha_lo*y_lo*rbf_lo);  f += delta       svmTrain.cu:133-135
    const float delta = __fadd_rn(__fmul_rn(sc.delta_hi, rbf_hi), __fmul_rn(sc.delta_lo, rbf_lo));
    const float f_new = __fadd_rn(f_j, delta);
    a.f[row] = f_new;
    bool up, low;
    iset_membership(alpha_j, y_j, a.c, up, low);
    const int gidx = a.index_base + row;
    if (up) best_up = umin64(best_up, pack_key_up(f_new, gidx));
    if (low) best_low = umin64(best_low, pack_key_low(f_new, gidx));
}

// Cache fast path: both dot-product lines are resident, X is not touched (28 B per row).
__device__ __forceinline__ void pass_both_hit(const PassScalars& sc, const PassDev& a,
                                              unsigned long long& best_up, unsigned long long& best_low)
{
    for (int row = blockIdx.x * blockDim.x + threadIdx.x; row < a.n; row += gridDim.x * blockDim.x) {
        const float dh = sc.line_hi[row], dl = sc.line_lo[row];
        row_epilogue(row, dh, dl, __ldg(a.xsq + row), a.f[row], __ldg(a.alpha + row), __ldg(a.y + row), sc, a,
                     false, best_up, best_low);
    }
}

// ------------------------------------------------------------------------------------------------
// Fused pass: direct 128-bit streaming loads of X, selected rows staged in smem by TMA.
//   LPR lanes cooperate on one row, ITERS float4 chunks per lane, R rows in flight per lane group.
//   HINT: loads carry an L2 eviction-priority policy -- units below a.pin_units are kept
//   (evict_last), the rest streamed (evict_first) -- so a prefix of X stays L2-resident from one
//   SMO iteration to the next (every iteration re-reads the same X).
//   a.reverse: sweep the units back to front (serpentine order across consecutive iterations: the
//   tail of the previous pass is what L2 still holds).
// ------------------------------------------------------------------------------------------------

template <int LPR, int ITERS, bool HINT>
__global__ void __launch_bounds__(kPassThreads, 2) smo_pass_kernel(const PassDev a)
{
    constexpr int R = kRowsInFlight;
    constexpr int GROUPS = 32 / LPR;
    constexpr int RPW = GROUPS * R;
    constexpr int CAP4 = LPR * ITERS;
    static_assert(LPR >= R, "epilogue owners are lanes sub < R");

    __shared__ __align__(16) float4 s_hi[CAP4];
    __shared__ __align__(16) float4 s_lo[CAP4];
    __shared__ __align__(8) uint64_t s_bar;
    __shared__ unsigned long long s_red[2 * (kPassThreads / 32)];
    __shared__ int s_pushflag;

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int sub = lane % LPR, grp = lane / LPR;
    const int n_units = (a.n + RPW - 1) / RPW;
    const int unit_stride = gridDim.x * (kPassThreads / 32);

    uint64_t pol_keep = 0, pol_stream = 0;
    if (HINT) {
        asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol_keep));
        asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol_stream));
    }

    // per-unit registers
    float4 xv[R][ITERS];
    float e_xsq = 0.f, e_f = 0.f, e_alpha = 0.f;
    int e_y = 0, erow = 0;
    bool eown = false;

    // constants of the unit (X, |x|^2, y): may be loaded before the dependency wait
    auto load_unit = [&](int unit) {
        const int u = a.reverse ? (n_units - 1 - unit) : unit;
        const int row0 = u * RPW + grp * R;
        const uint64_t pol = (u < a.pin_units) ? pol_keep : pol_stream;
        // the lane that will run the epilogue of row (row0 + sub) fetches its scalars up front
        erow = row0 + sub;
        eown = (sub < R) && (erow < a.n);
        if (eown) {
            e_xsq = __ldg(a.xsq + erow);
            e_y = __ldg(a.y + erow);
        }
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const float4* rowp = a.x4 + (size_t)(row0 + r) * a.pitch4;
            const bool row_ok = (row0 + r) < a.n;
#pragma unroll
            for (int it = 0; it < ITERS; ++it) {
                const int c4 = sub + it * LPR;
                xv[r][it] = (row_ok && c4 < a.pitch4) ? ldg_x4<HINT>(rowp + c4, pol) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
    };
    // state of the unit (f from the previous pass, alpha from the control step): only after the wait
    auto load_unit_state = [&]() {
        if (eown) {
            e_f = a.f[erow];
            e_alpha = a.alpha[erow];
        }
    };

    int unit = blockIdx.x * (kPassThreads / 32) + warp;
    // X does not depend on the pair: when a both-rows-cached iteration is unlikely (a.early_load),
    // put the first unit's loads in flight before waiting for anything else.
    bool loaded = false;
    if (a.early_load && unit < n_units) { load_unit(unit); loaded = true; }

    // Programmatic dependent launch: let the next kernel in the stream become resident early, then
    // wait until the previous one (the control step) has completed and its writes are visible.
    asm volatile("griddepcontrol.launch_dependents;");
    asm volatile("griddepcontrol.wait;" ::: "memory");

    const PassScalars sc = a.dev ? *a.dev : a.imm;
    if (!sc.active) return;

    unsigned long long best_up = DPSVM_KEY_EMPTY, best_low = DPSVM_KEY_EMPTY;
    if (sc.both_hit) {
        pass_both_hit(sc, a, best_up, best_low);
        publish_keys(best_up, best_low, a.keys, s_red);
        push_if_last(a.push, &s_pushflag);
        return;
    }

    stage_selected_rows(s_hi, s_lo, a.x_hi4, a.x_lo4, a.pitch4, CAP4, &s_bar);

    for (; unit < n_units; unit += unit_stride) {
        if (!loaded) load_unit(unit);
        loaded = false;
        load_unit_state();
        float acc_hi[R], acc_lo[R];
#pragma unroll
        for (int r = 0; r < R; ++r) { acc_hi[r] = 0.f; acc_lo[r] = 0.f; }
#pragma unroll
        for (int it = 0; it < ITERS; ++it) {
            const float4 h = s_hi[sub + it * LPR];
            const float4 l = s_lo[sub + it * LPR];
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const float4 v = xv[r][it];
                acc_hi[r] = __fmaf_rn(v.x, h.x, acc_hi[r]); acc_lo[r] = __fmaf_rn(v.x, l.x, acc_lo[r]);
                acc_hi[r] = __fmaf_rn(v.y, h.y, acc_hi[r]); acc_lo[r] = __fmaf_rn(v.y, l.y, acc_lo[r]);
                acc_hi[r] = __fmaf_rn(v.z, h.z, acc_hi[r]); acc_lo[r] = __fmaf_rn(v.z, l.z, acc_lo[r]);
                acc_hi[r] = __fmaf_rn(v.w, h.w, acc_hi[r]); acc_lo[r] = __fmaf_rn(v.w, l.w, acc_lo[r]);
            }
        }
#pragma unroll
        for (int off = LPR / 2; off >= 1; off >>= 1) {
#pragma unroll
            for (int r = 0; r < R; ++r) {
                acc_hi[r] = __fadd_rn(acc_hi[r], __shfl_xor_sync(0xffffffffu, acc_hi[r], off));
                acc_lo[r] = __fadd_rn(acc_lo[r], __shfl_xor_sync(0xffffffffu, acc_lo[r], off));
            }
        }
        if (eown) {
            float dh = acc_hi[0], dl = acc_lo[0];
#pragma unroll
            for (int r = 1; r < R; ++r) if (sub == r) { dh = acc_hi[r]; dl = acc_lo[r]; }
            row_epilogue(erow, dh, dl, e_xsq, e_f, e_alpha, e_y, sc, a, true, best_up, best_low);
        }
    }
    publish_keys(best_up, best_low, a.keys, s_red);
    push_if_last(a.push, &s_pushflag);
}

// Same algorithm for rows too long for the unrolled variants (pitch4 > 256): runtime chunk loop.
__global__ void __launch_bounds__(kPassThreads, 2) smo_pass_generic_kernel(const PassDev a)
{
    constexpr int R = kRowsInFlight;
    extern __shared__ __align__(16) unsigned char dyn_smem[];
    __shared__ __align__(8) uint64_t s_bar;
    __shared__ unsigned long long s_red[2 * (kPassThreads / 32)];
    __shared__ int s_pushflag;
    const int cap4 = ((a.pitch4 + 31) / 32) * 32;
    float4* s_hi = reinterpret_cast<float4*>(dyn_smem);
    float4* s_lo = s_hi + cap4;

    asm volatile("griddepcontrol.launch_dependents;");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const PassScalars sc = a.dev ? *a.dev : a.imm;
    if (!sc.active) return;
    unsigned long long best_up = DPSVM_KEY_EMPTY, best_low = DPSVM_KEY_EMPTY;
    if (sc.both_hit) {
        pass_both_hit(sc, a, best_up, best_low);
        publish_keys(best_up, best_low, a.keys, s_red);
        push_if_last(a.push, &s_pushflag);
        return;
    }
    stage_selected_rows(s_hi, s_lo, a.x_hi4, a.x_lo4, a.pitch4, cap4, &s_bar);

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n_units = (a.n + R - 1) / R;
    const int unit_stride = gridDim.x * (kPassThreads / 32);
    const int iters = cap4 / 32;
    for (int unit = blockIdx.x * (kPassThreads / 32) + warp; unit < n_units; unit += unit_stride) {
        const int row0 = unit * R;
        float acc_hi[R], acc_lo[R];
#pragma unroll
        for (int r = 0; r < R; ++r) { acc_hi[r] = 0.f; acc_lo[r] = 0.f; }
        for (int it = 0; it < iters; ++it) {
            const int c4 = lane + it * 32;
            const float4 h = s_hi[c4], l = s_lo[c4];
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const bool ok = (row0 + r) < a.n && c4 < a.pitch4;
                const float4 v = ok ? ldg_stream4(a.x4 + (size_t)(row0 + r) * a.pitch4 + c4) : make_float4(0.f, 0.f, 0.f, 0.f);
                acc_hi[r] = __fmaf_rn(v.x, h.x, acc_hi[r]); acc_lo[r] = __fmaf_rn(v.x, l.x, acc_lo[r]);
                acc_hi[r] = __fmaf_rn(v.y, h.y, acc_hi[r]); acc_lo[r] = __fmaf_rn(v.y, l.y, acc_lo[r]);
                acc_hi[r] = __fmaf_rn(v.z, h.z, acc_hi[r]); acc_lo[r] = __fmaf_rn(v.z, l.z, acc_lo[r]);
                acc_hi[r] = __fmaf_rn(v.w, h.w, acc_hi[r]); acc_lo[r] = __fmaf_rn(v.w, l.w, acc_lo[r]);
            }
        }
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) {
#pragma unroll
            for (int r = 0; r < R; ++r) {
                acc_hi[r] = __fadd_rn(acc_hi[r], __shfl_xor_sync(0xffffffffu, acc_hi[r], off));
                acc_lo[r] = __fadd_rn(acc_lo[r], __shfl_xor_sync(0xffffffffu, acc_lo[r], off));
            }
        }
        const int erow = row0 + lane;
        if (lane < R && erow < a.n) {
            float dh = acc_hi[0], dl = acc_lo[0];
#pragma unroll
            for (int r = 1; r < R; ++r) if (lane == r) { dh = acc_hi[r]; dl = acc_lo[r]; }
            row_epilogue(erow, dh, dl, __ldg(a.xsq + erow), a.f[erow], __ldg(a.alpha + erow), __ldg(a.y + erow), sc, a,
                         true, best_up, best_low);
        }
    }
    publish_keys(best_up, best_low, a.keys, s_red);
    push_if_last(a.push, &s_pushflag);
}

// ------------------------------------------------------------------------------------------------
// Fused pass over the sparse stream (layout and order: SparseDev, smo_kernels.cuh).  G lanes per row,
// 32/G rows per warp at a time; a warp walks the steps of its unit in register chunks, one chunk ahead
// in flight; the selected pair sits in shared memory interleaved as (x_hi[c], x_lo[c]) so that one
// 64-bit shared load serves both dot products of a non-zero.
// ------------------------------------------------------------------------------------------------
template <int G, bool C8>
__global__ void __launch_bounds__(kPassThreads, 2) smo_pass_sparse_kernel(const PassDev a)
{
    constexpr int GROUPS = 32 / G;
    __shared__ __align__(16) float2 s_hl[kSparseMaxPitch];
    __shared__ unsigned long long s_red[2 * (kPassThreads / 32)];
    __shared__ int s_pushflag;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int sub = lane % G, grp = lane / G;

    uint64_t pol_keep, pol_stream;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol_keep));
    if (a.sp.pin_steps) asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol_stream));
    else asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(pol_stream));

    SpSeqGrid seq;
    seq.k_first = blockIdx.x * (kPassThreads / 32) + warp;
    seq.k_stride = gridDim.x * (kPassThreads / 32);
    seq.k_end = a.sp.n_units;
    seq.unit_off = a.sp.unit_off;
    seq.rev = a.reverse != 0;
    seq.hint = a.sp.pin_steps != 0;
    seq.pin_steps = a.sp.pin_steps;

    // the stream does not depend on the pair: first chunks in flight before the dependency wait
    SpBuf ring[kSpRing];
    if (a.early_load) sp_prefetch<C8>(ring, a.sp, seq, lane, pol_keep, pol_stream);
    asm volatile("griddepcontrol.launch_dependents;");
    asm volatile("griddepcontrol.wait;" ::: "memory");

    const PassScalars sc = a.dev ? *a.dev : a.imm;
    if (!sc.active) return;
    unsigned long long best_up = DPSVM_KEY_EMPTY, best_low = DPSVM_KEY_EMPTY;
    if (sc.both_hit) {
        pass_both_hit(sc, a, best_up, best_low);
        publish_keys(best_up, best_low, a.keys, s_red);
        push_if_last(a.push, &s_pushflag);
        return;
    }
    for (int i = tid; i < a.pitch4; i += kPassThreads) {
        const float4 h = __ldcg(a.x_hi4 + i), l = __ldcg(a.x_lo4 + i);
        s_hl[4 * i + 0] = make_float2(h.x, l.x);
        s_hl[4 * i + 1] = make_float2(h.y, l.y);
        s_hl[4 * i + 2] = make_float2(h.z, l.z);
        s_hl[4 * i + 3] = make_float2(h.w, l.w);
    }
    __syncthreads();
    if (!a.early_load) sp_prefetch<C8>(ring, a.sp, seq, lane, pol_keep, pol_stream);

    float e_xsq = 0.f, e_f = 0.f, e_alpha = 0.f;
    int e_y = 0, erow = 0;
    bool eown = false;
    sp_sweep<G, C8>(ring, a.sp, seq, s_hl, lane, pol_keep, pol_stream,
        [&](int k) {
            erow = seq.local(k) * GROUPS + grp;
            eown = (sub == 0) && (erow < a.n);
            if (eown) {
                e_xsq = __ldg(a.xsq + erow); e_y = __ldg(a.y + erow);
                e_f = a.f[erow]; e_alpha = a.alpha[erow];
            }
        },
        [&](int, float dh, float dl) {
            if (eown) row_epilogue(erow, dh, dl, e_xsq, e_f, e_alpha, e_y, sc, a, true, best_up, best_low);
        });
    publish_keys(best_up, best_low, a.keys, s_red);
    push_if_last(a.push, &s_pushflag);
}

// ------------------------------------------------------------------------------------------------
// |x_i|^2 with the pass's accumulation order.
// ------------------------------------------------------------------------------------------------
template <int LPR, int ITERS>
__global__ void __launch_bounds__(256) row_norms_kernel(const float4* __restrict__ x4, int n, int pitch4, float* __restrict__ out)
{
    constexpr int GROUPS = 32 / LPR;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int sub = lane % LPR, grp = lane / LPR;
    const int n_units = (n + GROUPS - 1) / GROUPS;
    for (int unit = blockIdx.x * (blockDim.x >> 5) + warp; unit < n_units; unit += gridDim.x * (blockDim.x >> 5)) {
        const int row = unit * GROUPS + grp;
        float acc = 0.f;
#pragma unroll
        for (int it = 0; it < ITERS; ++it) {
            const int c4 = sub + it * LPR;
            const float4 v = (row < n && c4 < pitch4) ? ldg_stream4(x4 + (size_t)row * pitch4 + c4) : make_float4(0.f, 0.f, 0.f, 0.f);
            acc = __fmaf_rn(v.x, v.x, acc); acc = __fmaf_rn(v.y, v.y, acc);
            acc = __fmaf_rn(v.z, v.z, acc); acc = __fmaf_rn(v.w, v.w, acc);
        }
#pragma unroll
        for (int off = LPR / 2; off >= 1; off >>= 1) acc = __fadd_rn(acc, __shfl_xor_sync(0xffffffffu, acc, off));
        if (sub == 0 && row < n) out[row] = acc;
    }
}

__global__ void __launch_bounds__(256) row_norms_generic_kernel(const float4* __restrict__ x4, int n, int pitch4, float* __restrict__ out)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int iters = (pitch4 + 31) / 32;
    for (int row = blockIdx.x * (blockDim.x >> 5) + warp; row < n; row += gridDim.x * (blockDim.x >> 5)) {
        float acc = 0.f;
        for (int it = 0; it < iters; ++it) {
            const int c4 = lane + it * 32;
            const float4 v = c4 < pitch4 ? ldg_stream4(x4 + (size_t)row * pitch4 + c4) : make_float4(0.f, 0.f, 0.f, 0.f);
            acc = __fmaf_rn(v.x, v.x, acc); acc = __fmaf_rn(v.y, v.y, acc);
            acc = __fmaf_rn(v.z, v.z, acc); acc = __fmaf_rn(v.w, v.w, acc);
        }
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) acc = __fadd_rn(acc, __shfl_xor_sync(0xffffffffu, acc, off));
        if (lane == 0) out[row] = acc;
    }
}

// alpha = 0, f = -y   (svmTrain.cu:349, 379-380)
__global__ void init_state_kernel(const int* __restrict__ y, int n, float* __restrict__ alpha, float* __restrict__ f)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        if (alpha) alpha[i] = 0.f;
        f[i] = -(float)y[i];
    }
}

__global__ void __launch_bounds__(256) select_kernel(const float* __restrict__ alpha, const int* __restrict__ y,
                                                     const float* __restrict__ f, int n, float c, int index_base,
                                                     unsigned long long* keys)
{
    __shared__ unsigned long long s_red[2 * 8];
    unsigned long long best_up = DPSVM_KEY_EMPTY, best_low = DPSVM_KEY_EMPTY;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        bool up, low;
        iset_membership(alpha[i], y[i], c, up, low);
        const float fi = f[i];
        if (up) best_up = umin64(best_up, pack_key_up(fi, index_base + i));
        if (low) best_low = umin64(best_low, pack_key_low(fi, index_base + i));
    }
    publish_keys(best_up, best_low, keys, s_red);
}

// ------------------------------------------------------------------------------------------------
// Kernel-row cache bookkeeping (device-resident LRU; replaces myCache, cache.cu:49-105).
// Single-thread code: called by thread 0 of the control kernel.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void lru_unlink(const CacheDev& c, int line)
{
    const int p = c.prev[line], nx = c.next[line];
    if (p >= 0) c.next[p] = nx; else c.meta[kCacheHead] = nx;
    if (nx >= 0) c.prev[nx] = p; else c.meta[kCacheTail] = p;
}
__device__ __forceinline__ void lru_append(const CacheDev& c, int line)
{
    const int t = c.meta[kCacheTail];
    c.prev[line] = t; c.next[line] = -1;
    if (t >= 0) c.next[t] = line; else c.meta[kCacheHead] = line;
    c.meta[kCacheTail] = line;
}
// myCache::lookup (cache.cu:62-80) then, on a miss, myCache::get_new_cache_line (:82-105).
// `line` is line_of_key[key] as the caller already loaded it (-1 = not cached).
__device__ int cache_lookup_device(const CacheDev& c, int key, int line, int* hit)
{
    if (line >= 0) {
        if (c.meta[kCacheTail] != line) { lru_unlink(c, line); lru_append(c, line); }
        *hit = 1;
        return line;
    }
    *hit = 0;
    const int size = c.meta[kCacheSize];
    if (size < c.lines) {
        line = size;
        c.meta[kCacheSize] = size + 1;
    } else {
        line = c.meta[kCacheHead];            // least recently used
        c.line_of_key[c.key_of_line[line]] = -1;
        lru_unlink(c, line);
    }
    c.key_of_line[line] = key;
    c.line_of_key[key] = line;
    lru_append(c, line);
    return line;
}

// Host-callable lookup (SvmTrain::lookup_cache, svmTrain.cu:142-156): out[0] = line, out[1] = hit.
__global__ void cache_lookup_kernel(const CacheDev c, int key, int* out)
{
    int hit = 0;
    const int line = cache_lookup_device(c, key, c.line_of_key[key], &hit);
    out[0] = line; out[1] = hit;
}

// ------------------------------------------------------------------------------------------------
// Control step: one block.  Everything svmTrainMain.cpp:246-308 does on the host between
// train_step1 and train_step2, plus train_step2's alpha install and cache lookups -- on the device.
// directed != 0: the pair and the new alphas come from the host (SvmTrain::train_step2 drop-in).
//
// Built for latency (it sits on the critical path of every iteration):
//   * launched with programmatic dependent launch: it becomes resident while the previous pass is
//     still draining and lets the next pass do the same (griddepcontrol);
//   * every independent global load is issued at once: the two rows by the whole block, the eight
//     per-row scalars by eight lanes of warp 0, the recency lists of a small cache by the whole
//     block into shared memory -- so the serial part (eta, alpha update, LRU pointer chasing) runs
//     on registers and shared memory only.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kCtlThreads, 1) smo_control_kernel(const CtlDev a)
{
    extern __shared__ int s_cache[];                  // [3 * lines + kCacheMetaInts] when the cache is staged
    __shared__ float s_part[kCtlThreads / 32];
    __shared__ float s_scal[8];
    const int tid = threadIdx.x;
    DevState* st = a.st;

    asm volatile("griddepcontrol.launch_dependents;");
    asm volatile("griddepcontrol.wait;" ::: "memory");

    // uniform loads (one transaction per warp, no barrier needed to share them)
    const int done = st->done, num_iter = st->num_iter, stop_at = st->stop_at;
    const unsigned long long k_up = a.keys[0], k_low = a.keys[1];
    if (!a.directed && (done || num_iter >= stop_at)) {
        if (tid == 0) st->pass.active = 0;
        return;
    }
    int i_hi, i_lo;
    float b_hi, b_lo;
    const CandHeader* h_hi = nullptr;
    const CandHeader* h_lo = nullptr;
    bool ll_alpha = false;
    float ll_a_hi = 0.f, ll_a_lo = 0.f;
    if (a.directed) {
        i_hi = a.dir_i_hi; i_lo = a.dir_i_lo; b_hi = 0.f; b_lo = 0.f;
    } else if (a.world > 1 && a.ll_in) {
        // flagged-word exchange (every rank holds all of X): keys and alphas of every rank's candidates; the rows, y
        // and |x|^2 of the pair are read from local memory below, exactly as on one GPU
        __shared__ unsigned int s_ll[kMaxPeers * kLLWords];
        __shared__ int s_ok;
        if (!ll_wait(a.ll_in, a.world, (unsigned int)num_iter, s_ll, &s_ok, a.timeout_ns)) {
            if (tid == 0) { st->error = 1; st->done = 1; st->pass.active = 0; }
            return;
        }
        unsigned long long g_up, g_low;
        ll_merge(s_ll, a.world, g_up, ll_a_hi, g_low, ll_a_lo);
        ll_alpha = true;
        unpack_key_up(g_up, &b_hi, &i_hi);
        unpack_key_low(g_low, &b_lo, &i_lo);
        if (g_up == DPSVM_KEY_EMPTY) i_hi = a.res0;
        if (g_low == DPSVM_KEY_EMPTY) i_lo = a.res0;
    } else if (a.world > 1) {
        const unsigned char* cand = a.cand + (a.flags ? (size_t)(num_iter & 1) * a.world * a.cand_stride : 0);
        if (a.flags) {
            // peer-memory exchange: wait until every rank's package for this iteration has landed in
            // slot (num_iter & 1) of the local mailbox.  Bounded spin: a missing peer stops the loop
            // (DevState::error) instead of hanging the GPU.
            __shared__ int s_timeout;
            if (tid == 0) s_timeout = 0;
            __syncthreads();
            if (tid < a.world) {
                const unsigned int want = (unsigned int)num_iter + 1u;
                const unsigned int* fl = a.flags + (size_t)(num_iter & 1) * a.world + tid;
                unsigned long long t0 = 0;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
                while (ld_acquire_sys_u32(fl) != want) {
                    unsigned long long t1;
                    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
                    if (t1 - t0 > a.timeout_ns) { s_timeout = 1; break; }
                    __nanosleep(64);
                }
            }
            __syncthreads();
            if (s_timeout) {
                if (tid == 0) { st->error = 1; st->done = 1; st->pass.active = 0; }
                return;
            }
        }
        // Global merge of the gathered packages (svmTrainMain.cpp:255-274): the minimum packed key is
        // argmin f / argmax f with the lowest GLOBAL index among equal values, identical on every rank.
        unsigned long long g_up = DPSVM_KEY_EMPTY, g_low = DPSVM_KEY_EMPTY;
        int w_up = 0, w_low = 0;
        for (int r = 0; r < a.world; ++r) {
            const CandHeader* h = reinterpret_cast<const CandHeader*>(cand + (size_t)r * a.cand_stride);
            const unsigned long long u = __ldcg(&h->key_up), l = __ldcg(&h->key_low);
            if (u < g_up) { g_up = u; w_up = r; }
            if (l < g_low) { g_low = l; w_low = r; }
        }
        unpack_key_up(g_up, &b_hi, &i_hi);
        unpack_key_low(g_low, &b_lo, &i_lo);
        h_hi = reinterpret_cast<const CandHeader*>(cand + (size_t)w_up * a.cand_stride);
        h_lo = reinterpret_cast<const CandHeader*>(cand + (size_t)w_low * a.cand_stride);
    } else {
        unpack_key_up(k_up, &b_hi, &i_hi);
        unpack_key_low(k_low, &b_lo, &i_lo);
        // An empty set decodes to the reference's default, global row 0 (svmTrainMain.cpp:249-252).  When the rows
        // held here start at res0 > 0 that row is not resident: the first resident row stands in for it.
        if (k_up == DPSVM_KEY_EMPTY) i_hi = a.res0;
        if (k_low == DPSVM_KEY_EMPTY) i_lo = a.res0;
    }
    const int r_hi = i_hi - a.res0, r_lo = i_lo - a.res0;
    const float* src_hi = a.ext_row_hi ? a.ext_row_hi : a.x + (size_t)r_hi * a.pitch;
    const float* src_lo = a.ext_row_lo ? a.ext_row_lo : a.x + (size_t)r_lo * a.pitch;
    if (h_hi) {
        src_hi = reinterpret_cast<const float*>(h_hi + 1);                 // package: header, row_hi, row_lo
        src_lo = reinterpret_cast<const float*>(h_lo + 1) + a.pitch;
    }
    const bool cache_on = a.cache.lines > 0;
    const bool staged = cache_on && a.cache.lines <= kCacheSmemLines;

    // ---- everything that can be loaded now, in one round ----
    if (tid < 8) {
        float v;
        switch (tid) {
            case 0: v = h_hi ? (float)h_hi->y_hi : (float)a.y[r_hi]; break;
            case 1: v = h_lo ? (float)h_lo->y_lo : (float)a.y[r_lo]; break;
            case 2: v = ll_alpha ? ll_a_hi : (h_hi ? h_hi->a_hi : a.alpha[r_hi]); break;
            case 3: v = ll_alpha ? ll_a_lo : (h_lo ? h_lo->a_lo : a.alpha[r_lo]); break;
            case 4: v = h_hi ? h_hi->xsq_hi : a.xsq[r_hi]; break;
            case 5: v = h_lo ? h_lo->xsq_lo : a.xsq[r_lo]; break;
            case 6: v = cache_on ? __int_as_float(a.cache.line_of_key[i_hi]) : 0.f; break;
            default: v = cache_on ? __int_as_float(a.cache.line_of_key[i_lo]) : 0.f; break;
        }
        s_scal[tid] = v;
    }
    if (staged) {
        const int L = a.cache.lines;
        for (int i = tid; i < L; i += kCtlThreads) {
            s_cache[i] = a.cache.key_of_line[i];
            s_cache[L + i] = a.cache.prev[i];
            s_cache[2 * L + i] = a.cache.next[i];
        }
        if (tid < kCacheMetaInts) s_cache[3 * L + tid] = a.cache.meta[tid];
    }
    // copy the two rows to the staging buffer the pass reads, and |x_hi - x_lo|^2 on the way
    // (the host kernel of svmTrain.cu:696-714: saxpy(-1) then sdot of the difference).
    float part = 0.f;
    for (int i = tid; i < a.pitch; i += kCtlThreads) {
        const float h = src_hi[i], l = src_lo[i];
        a.sel_rows[i] = h;
        a.sel_rows[a.pitch + i] = l;
        const float df = __fsub_rn(h, l);
        part = __fmaf_rn(df, df, part);
    }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) part = __fadd_rn(part, __shfl_xor_sync(0xffffffffu, part, off));
    if ((tid & 31) == 0) s_part[tid >> 5] = part;
    __syncthreads();

    if (tid == 0) {
        float norm_sq = 0.f;
#pragma unroll
        for (int w = 0; w < kCtlThreads / 32; ++w) norm_sq = __fadd_rn(norm_sq, s_part[w]);

        const int y_hi = (int)s_scal[0], y_lo = (int)s_scal[1];
        const float a_hi_old = s_scal[2], a_lo_old = s_scal[3];
        float a_hi_new, a_lo_new, eta = 0.f;
        if (a.directed) {
            a_hi_new = a.dir_a_hi; a_lo_new = a.dir_a_lo;
        } else {
            // eta = K(hi,hi) + K(lo,lo) - 2 K(lo,hi)  (svmTrainMain.cpp:282); K(i,i) = exp(-gamma*0) = 1.
            // K = (float)exp(-1 * (float)gamma * norm_sq): float product, double exp (svmTrain.cu:709).
            const float k_lh = (float)exp((double)__fmul_rn(-a.gamma, norm_sq));
            eta = __fsub_rn(__fadd_rn(1.0f, 1.0f), __fmul_rn(2.0f, k_lh));
            // svmTrainMain.cpp:289-295; IEEE division (eta == 0 gives +-inf, clipped to 0 or C)
            const float s = (float)(y_lo * y_hi);
            const float a_lo_unc = __fadd_rn(a_lo_old, __fdiv_rn(__fmul_rn((float)y_lo, __fsub_rn(b_hi, b_lo)), eta));
            const float a_hi_unc = __fadd_rn(a_hi_old, __fmul_rn(s, __fsub_rn(a_lo_old, a_lo_unc)));
            a_lo_new = a_lo_unc < 0.0f ? 0.0f : (a_lo_unc > a.c ? a.c : a_lo_unc);
            a_hi_new = a_hi_unc < 0.0f ? 0.0f : (a_hi_unc > a.c ? a.c : a_hi_unc);
        }
        // alpha[max_idx] = alpha_lo_new; alpha[min_idx] = alpha_hi_new   (svmTrainMain.cpp:298-299)
        // (multi-GPU: every rank computes the same values, the owner of the row stores them)
        if (a.world <= 1 || (i_lo >= a.own0 && i_lo < a.own0 + a.own_n)) a.alpha[r_lo] = a_lo_new;
        if (a.world <= 1 || (i_hi >= a.own0 && i_hi < a.own0 + a.own_n)) a.alpha[r_hi] = a_hi_new;

        PassScalars ps;
        ps.active = 1;
        // when hi and lo are the same row the second store above wins; old values were read before both
        ps.delta_hi = __fmul_rn(__fsub_rn(a_hi_new, a_hi_old), (float)y_hi);
        ps.delta_lo = __fmul_rn(__fsub_rn(a_lo_new, a_lo_old), (float)y_lo);
        ps.xsq_hi = s_scal[4];
        ps.xsq_lo = s_scal[5];
        ps.line_hi = nullptr; ps.line_lo = nullptr; ps.both_hit = 0;
        if (cache_on) {
            CacheDev cd = a.cache;
            if (staged) {
                const int L = a.cache.lines;
                cd.key_of_line = s_cache; cd.prev = s_cache + L; cd.next = s_cache + 2 * L; cd.meta = s_cache + 3 * L;
            }
            // lookup order hi then lo (svmTrain.cu:203,238); the second lookup must see the first one's eviction
            int hit_hi, hit_lo;
            const int known_hi = __float_as_int(s_scal[6]);
            const int l_hi = cache_lookup_device(cd, i_hi, known_hi, &hit_hi);
            int known_lo = __float_as_int(s_scal[7]);
            if (i_lo == i_hi) known_lo = l_hi;
            else if (known_lo == l_hi) known_lo = -1;          // hi just took over lo's line
            const int l_lo = cache_lookup_device(cd, i_lo, known_lo, &hit_lo);
            st->cache_lookups += 2;
            st->cache_hits += hit_hi + hit_lo;
            float* p_hi = a.cache.data + (size_t)l_hi * a.cache.line_stride;
            float* p_lo = a.cache.data + (size_t)l_lo * a.cache.line_stride;
            if (hit_hi && hit_lo) { ps.both_hit = 1; ps.line_hi = p_hi; ps.line_lo = p_lo; }
            else { if (!hit_hi) ps.line_hi = p_hi; if (!hit_lo && l_lo != l_hi) ps.line_lo = p_lo; }
        }
        st->pass = ps;
        st->i_hi = i_hi; st->i_lo = i_lo; st->eta = eta;
        const int it = num_iter + 1;
        st->num_iter = it;
        if (!a.directed) {
            st->b_hi = b_hi; st->b_lo = b_lo;
            // while (b_lo > b_hi + 2*epsilon && num_iter < max_iter)   (svmTrainMain.cpp:310)
            const bool keep_going = b_lo > __fadd_rn(b_hi, __fmul_rn(2.0f, a.epsilon));
            if (!keep_going) { st->converged = 1; st->done = 1; }
            if (it >= a.max_iter) st->done = 1;
        }
        if (a.world <= 1) {               // (multi-GPU: the pack kernel consumed and reset the local keys)
            a.keys[0] = DPSVM_KEY_EMPTY;
            a.keys[1] = DPSVM_KEY_EMPTY;
        }
    }
    if (staged) {
        __syncthreads();
        const int L = a.cache.lines;
        for (int i = tid; i < L; i += kCtlThreads) {
            a.cache.key_of_line[i] = s_cache[i];
            a.cache.prev[i] = s_cache[L + i];
            a.cache.next[i] = s_cache[2 * L + i];
        }
        if (tid < kCacheMetaInts) a.cache.meta[tid] = s_cache[3 * L + tid];
    }
}

// ------------------------------------------------------------------------------------------------
// Multi-GPU: publish this rank's candidates (keys, scalars, both rows) as one package and reset the
// local keys.  One block.  An empty local set still ships a valid row (the first owned one) so the
// reference's "index 0, +-1e9" default (svmTrainMain.cpp:249-252) stays well defined.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kCtlThreads, 1) smo_pack_kernel(const PackDev a)
{
    const int tid = threadIdx.x;
    const unsigned long long k_up = a.keys[0], k_low = a.keys[1];
    float f_dummy;
    int i_hi = a.own0, i_lo = a.own0;
    if (k_up != DPSVM_KEY_EMPTY) unpack_key_up(k_up, &f_dummy, &i_hi);
    if (k_low != DPSVM_KEY_EMPTY) unpack_key_low(k_low, &f_dummy, &i_lo);
    const int r_hi = i_hi - a.res0, r_lo = i_lo - a.res0;
    CandHeader* h = reinterpret_cast<CandHeader*>(a.send);
    float* rows = reinterpret_cast<float*>(h + 1);
    const float* src_hi = a.x + (size_t)r_hi * a.pitch;
    const float* src_lo = a.x + (size_t)r_lo * a.pitch;
    for (int i = tid; i < a.pitch; i += kCtlThreads) {
        rows[i] = src_hi[i];
        rows[a.pitch + i] = src_lo[i];
    }
    if (tid == 0) {
        h->key_up = k_up; h->key_low = k_low;
        h->a_hi = a.alpha[r_hi]; h->a_lo = a.alpha[r_lo];
        h->y_hi = a.y[r_hi]; h->y_lo = a.y[r_lo];
        h->xsq_hi = a.xsq[r_hi]; h->xsq_lo = a.xsq[r_lo];
    }
    __syncthreads();
    if (tid == 0) {
        a.keys_copy[0] = k_up; a.keys_copy[1] = k_low;
        a.keys[0] = DPSVM_KEY_EMPTY; a.keys[1] = DPSVM_KEY_EMPTY;
    }
}

// ------------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------------
// Launch with (optionally) the programmatic-stream-serialization attribute: the kernel may become
// resident before its predecessor in the stream has finished; it orders itself with griddepcontrol.wait.
template <typename... KArgs, typename... Args>
static cudaError_t launch_ex(void (*kernel)(KArgs...), int blocks, int threads, size_t smem, cudaStream_t s, bool pdl, Args... args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)blocks);
    cfg.blockDim = dim3((unsigned)threads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, args...);
}

// SM count of the CURRENT device (cached per device; racing first calls store the same value).
int device_sm_count()
{
    static int s_sms[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (s_sms[dev] == 0) {
        int v = 0;
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) return 148;
        s_sms[dev] = v;
    }
    return s_sms[dev];
}

template <int LPR, int ITERS>
static cudaError_t launch_pass_t(PassDev a, int variant, cudaStream_t s)
{
    constexpr int RPW = (32 / LPR) * kRowsInFlight;
    const int n_units = (a.n + RPW - 1) / RPW;
    int blocks = (n_units + (kPassThreads / 32) - 1) / (kPassThreads / 32);
    const int cap = device_sm_count() * 2;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    // variant: bit 0 = reverse sweep, bit 1 = L2 eviction hints with (variant >> 8) MB of X pinned
    a.reverse = variant & 1;
    const bool hint = (variant & 2) != 0;
    if (hint) {
        const double pin_bytes = (double)((variant >> 8) & 0xFFFFF) * 1e6;
        const double unit_bytes = (double)RPW * a.pitch4 * 16.0;
        double pu = pin_bytes / unit_bytes;
        a.pin_units = pu > (double)n_units ? n_units : (int)pu;
        return launch_ex(smo_pass_kernel<LPR, ITERS, true>, blocks, kPassThreads, 0, s, a.pdl != 0, a);
    }
    a.pin_units = 0;
    return launch_ex(smo_pass_kernel<LPR, ITERS, false>, blocks, kPassThreads, 0, s, a.pdl != 0, a);
}

template <int LPR>
static cudaError_t launch_pass_lpr(int iters, const PassDev& a, int variant, cudaStream_t s)
{
    switch (iters) {
        case 1: return launch_pass_t<LPR, 1>(a, variant, s);
        case 2: return launch_pass_t<LPR, 2>(a, variant, s);
        case 3: return launch_pass_t<LPR, 3>(a, variant, s);
        case 4: return launch_pass_t<LPR, 4>(a, variant, s);
        case 5: return launch_pass_t<LPR, 5>(a, variant, s);
        case 6: return launch_pass_t<LPR, 6>(a, variant, s);
        case 7: return launch_pass_t<LPR, 7>(a, variant, s);
        case 8: return launch_pass_t<LPR, 8>(a, variant, s);
    }
    return cudaErrorInvalidValue;
}

cudaError_t launch_pass_sparse(const PassDev& a0, int variant, cudaStream_t s)
{
    PassDev a = a0;
    if (a.pitch4 * 4 > kSparseMaxPitch) return cudaErrorInvalidValue;
    int blocks = (a.sp.n_units + (kPassThreads / 32) - 1) / (kPassThreads / 32);
    const int cap = device_sm_count() * 2;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    a.reverse = variant & 1;
    // same L2 budget as the dense pass: a prefix of the stream is loaded evict_last once it outgrows L2
    a.sp.pin_steps = (variant & 2) ? (unsigned int)((double)((variant >> 8) & 0xFFFFF) * 1e6 / (double)a.sp.step_bytes) : 0u;
    if (a.sp.col8) {
        switch (a.sp.lanes) {
            case 4: return launch_ex(smo_pass_sparse_kernel<4, true>, blocks, kPassThreads, 0, s, a.pdl != 0, a);
            case 8: return launch_ex(smo_pass_sparse_kernel<8, true>, blocks, kPassThreads, 0, s, a.pdl != 0, a);
            case 16: return launch_ex(smo_pass_sparse_kernel<16, true>, blocks, kPassThreads, 0, s, a.pdl != 0, a);
            case 32: return launch_ex(smo_pass_sparse_kernel<32, true>, blocks, kPassThreads, 0, s, a.pdl != 0, a);
        }
        return cudaErrorInvalidValue;
    }
    switch (a.sp.lanes) {
        case 4: return launch_ex(smo_pass_sparse_kernel<4, false>, blocks, kPassThreads, 0, s, a.pdl != 0, a);
        case 8: return launch_ex(smo_pass_sparse_kernel<8, false>, blocks, kPassThreads, 0, s, a.pdl != 0, a);
        case 16: return launch_ex(smo_pass_sparse_kernel<16, false>, blocks, kPassThreads, 0, s, a.pdl != 0, a);
        case 32: return launch_ex(smo_pass_sparse_kernel<32, false>, blocks, kPassThreads, 0, s, a.pdl != 0, a);
    }
    return cudaErrorInvalidValue;
}

cudaError_t launch_pass(const PassDev& a, int variant, cudaStream_t s)
{
    if (a.n <= 0) return cudaSuccess;
    if (a.sp.val) return launch_pass_sparse(a, variant, s);
    if (a.pitch4 > 256) {
        const int cap4 = ((a.pitch4 + 31) / 32) * 32;
        const size_t smem = (size_t)cap4 * 2 * sizeof(float4);
        if (smem > 200 * 1024) return cudaErrorInvalidValue;
        static bool attr_set[64] = {false};      // function attributes belong to the device they were set on
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
        if (!attr_set[dev]) {
            cudaError_t e = cudaFuncSetAttribute(smo_pass_generic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
            if (e != cudaSuccess) return e;
            attr_set[dev] = true;
        }
        const int n_units = (a.n + kRowsInFlight - 1) / kRowsInFlight;
        int blocks = (n_units + 7) / 8;
        const int cap = device_sm_count() * 2;
        if (blocks > cap) blocks = cap;
        return launch_ex(smo_pass_generic_kernel, blocks, kPassThreads, smem, s, a.pdl != 0, a);
    }
    const DotGeom g = dot_geometry(a.pitch4);
    switch (g.lpr) {
        case 4: return launch_pass_lpr<4>(g.iters, a, variant, s);
        case 8: return launch_pass_lpr<8>(g.iters, a, variant, s);
        case 16: return launch_pass_lpr<16>(g.iters, a, variant, s);
        default: return launch_pass_lpr<32>(g.iters, a, variant, s);
    }
}

template <int LPR>
static cudaError_t launch_norms_lpr(int iters, const float4* x4, int n, int pitch4, float* out, cudaStream_t s)
{
    const int groups = 32 / LPR;
    const int n_units = (n + groups - 1) / groups;
    int blocks = (n_units + 7) / 8;
    const int cap = device_sm_count() * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    switch (iters) {
        case 1: row_norms_kernel<LPR, 1><<<blocks, 256, 0, s>>>(x4, n, pitch4, out); break;
        case 2: row_norms_kernel<LPR, 2><<<blocks, 256, 0, s>>>(x4, n, pitch4, out); break;
        case 3: row_norms_kernel<LPR, 3><<<blocks, 256, 0, s>>>(x4, n, pitch4, out); break;
        case 4: row_norms_kernel<LPR, 4><<<blocks, 256, 0, s>>>(x4, n, pitch4, out); break;
        case 5: row_norms_kernel<LPR, 5><<<blocks, 256, 0, s>>>(x4, n, pitch4, out); break;
        case 6: row_norms_kernel<LPR, 6><<<blocks, 256, 0, s>>>(x4, n, pitch4, out); break;
        case 7: row_norms_kernel<LPR, 7><<<blocks, 256, 0, s>>>(x4, n, pitch4, out); break;
        case 8: row_norms_kernel<LPR, 8><<<blocks, 256, 0, s>>>(x4, n, pitch4, out); break;
        default: return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

cudaError_t launch_row_norms(const float* x, int n, int pitch, float* out, cudaStream_t s)
{
    if (n <= 0) return cudaSuccess;
    const int pitch4 = pitch / 4;
    const float4* x4 = reinterpret_cast<const float4*>(x);
    if (pitch4 > 256) {
        int blocks = (n + 7) / 8;
        const int cap = device_sm_count() * 8;
        if (blocks > cap) blocks = cap;
        row_norms_generic_kernel<<<blocks, 256, 0, s>>>(x4, n, pitch4, out);
        return cudaGetLastError();
    }
    const DotGeom g = dot_geometry(pitch4);
    switch (g.lpr) {
        case 4: return launch_norms_lpr<4>(g.iters, x4, n, pitch4, out, s);
        case 8: return launch_norms_lpr<8>(g.iters, x4, n, pitch4, out, s);
        case 16: return launch_norms_lpr<16>(g.iters, x4, n, pitch4, out, s);
        default: return launch_norms_lpr<32>(g.iters, x4, n, pitch4, out, s);
    }
}

cudaError_t launch_init_state(const int* y, int n, float* alpha, float* f, cudaStream_t s)
{
    if (n <= 0) return cudaSuccess;
    int blocks = (n + 255) / 256;
    const int cap = device_sm_count() * 8;
    if (blocks > cap) blocks = cap;
    init_state_kernel<<<blocks, 256, 0, s>>>(y, n, alpha, f);
    return cudaGetLastError();
}

cudaError_t launch_select(const float* alpha, const int* y, const float* f, int n, float c, int index_base,
                          unsigned long long* keys, cudaStream_t s)
{
    cudaError_t e = cudaMemsetAsync(keys, 0xFF, 2 * sizeof(unsigned long long), s);
    if (e != cudaSuccess) return e;
    if (n <= 0) return cudaSuccess;
    int blocks = (n + 255) / 256;
    const int cap = device_sm_count() * 4;
    if (blocks > cap) blocks = cap;
    select_kernel<<<blocks, 256, 0, s>>>(alpha, y, f, n, c, index_base, keys);
    return cudaGetLastError();
}

cudaError_t launch_cache_lookup(const CacheDev& c, int key, int* d_out, cudaStream_t s)
{
    cache_lookup_kernel<<<1, 1, 0, s>>>(c, key, d_out);
    return cudaGetLastError();
}

cudaError_t launch_pack(const PackDev& a, cudaStream_t s)
{
    smo_pack_kernel<<<1, kCtlThreads, 0, s>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_push(const PushDev* d_push, cudaStream_t s)
{
    smo_push_kernel<<<1, kCtlThreads, 0, s>>>(d_push);
    return cudaGetLastError();
}

cudaError_t launch_control(const CtlDev& a, cudaStream_t s)
{
    const bool staged = a.cache.lines > 0 && a.cache.lines <= kCacheSmemLines;
    const size_t smem = staged ? sizeof(int) * (3 * (size_t)a.cache.lines + kCacheMetaInts) : 0;
    return launch_ex(smo_control_kernel, 1, kCtlThreads, smem, s, a.pdl != 0, a);
}

}  // namespace dpsvm
