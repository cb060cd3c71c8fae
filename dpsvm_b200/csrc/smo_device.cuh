// smo_device.cuh -- device-side PTX helpers shared by the kernel translation units (internal).
#pragma once

#include "smo_kernels.cuh"

namespace dpsvm {

// ------------------------------------------------------------------------------------------------
// small PTX helpers
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float4 ldg_stream4(const float4* p)
{
    // X is read exactly once per pass: keep it out of L1.
    float4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// TMA bulk copy global -> shared (1-D, no tensor map): bytes % 16 == 0, both sides 16 B aligned.
__device__ __forceinline__ void tma_bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(smem_dst)), "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ unsigned long long umin64(unsigned long long a, unsigned long long b) { return a < b ? a : b; }

__device__ __forceinline__ void st_release_sys_u32(unsigned int* p, unsigned int v)
{
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys_u32(const unsigned int* p)
{
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

template <bool HINT>
__device__ __forceinline__ float4 ldg_x4(const float4* p, uint64_t pol)
{
    float4 v;
    if (HINT) {
        asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.f32 {%0,%1,%2,%3}, [%4], %5;"
                     : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p), "l"(pol));
    } else {
        asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    }
    return v;
}

// ------------------------------------------------------------------------------------------------
// Sparse sweep building blocks (layout and order: SparseDev in smo_kernels.cuh).
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float ldg_f32_pol(const float* p, uint64_t pol)
{
    float v;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.f32 %0, [%1], %2;" : "=f"(v) : "l"(p), "l"(pol));
    return v;
}
__device__ __forceinline__ unsigned int ldg_u32_pol(const unsigned int* p, uint64_t pol)
{
    unsigned int v;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(v) : "l"(p), "l"(pol));
    return v;
}

struct SpChunk {
    float v[kSpChunk];
    unsigned int c2[kSpChunk / 2];
};

// Steps [step0, step0 + min(steps_left, kSpChunk)) of a unit for this lane; step0 is even.  What lies past
// the unit's end is value 0 / offset 0: an FMA with it leaves the accumulators as they are.
__device__ __forceinline__ void sp_load_chunk(SpChunk& ch, const SparseDev& sp, unsigned int step0, int steps_left,
                                              int lane, uint64_t pol)
{
    const float* vp = sp.val + (size_t)step0 * 32 + lane;
    const unsigned int* cp = sp.col2 + (size_t)(step0 >> 1) * 32 + lane;
#pragma unroll
    for (int j = 0; j < kSpChunk; j += 2) {
        const bool ok = j < steps_left;             // steps come in pairs (kSpStepAlign == 2)
        ch.v[j] = ok ? ldg_f32_pol(vp + j * 32, pol) : 0.f;
        ch.v[j + 1] = ok ? ldg_f32_pol(vp + (j + 1) * 32, pol) : 0.f;
        ch.c2[j / 2] = ok ? ldg_u32_pol(cp + (j / 2) * 32, pol) : 0u;
    }
}

// s_hl: the selected pair interleaved, s_hl[col] = (x_hi[col], x_lo[col]); c2 holds BYTE offsets into it.
__device__ __forceinline__ void sp_fma_chunk(const SpChunk& ch, const float2* s_hl, float& acc_hi, float& acc_lo)
{
    const unsigned char* base = reinterpret_cast<const unsigned char*>(s_hl);
#pragma unroll
    for (int j = 0; j < kSpChunk; j += 2) {
        const unsigned int w = ch.c2[j / 2];
        const float2 h0 = *reinterpret_cast<const float2*>(base + (w & 0xFFFFu));
        const float2 h1 = *reinterpret_cast<const float2*>(base + (w >> 16));
        acc_hi = __fmaf_rn(ch.v[j], h0.x, acc_hi);
        acc_lo = __fmaf_rn(ch.v[j], h0.y, acc_lo);
        acc_hi = __fmaf_rn(ch.v[j + 1], h1.x, acc_hi);
        acc_lo = __fmaf_rn(ch.v[j + 1], h1.y, acc_lo);
    }
}

template <int G>
__device__ __forceinline__ void sp_group_sum(float& acc_hi, float& acc_lo)
{
#pragma unroll
    for (int off = G / 2; off >= 1; off >>= 1) {
        acc_hi = __fadd_rn(acc_hi, __shfl_xor_sync(0xffffffffu, acc_hi, off));
        acc_lo = __fadd_rn(acc_lo, __shfl_xor_sync(0xffffffffu, acc_lo, off));
    }
}

// ------------------------------------------------------------------------------------------------
// Peer-memory candidate exchange (multi-GPU).  Runs in ONE block: builds this rank's package from
// its local selection and stores it into slot (seq & 1) of every rank's mailbox -- stores to
// CUDA-IPC-mapped peer memory travel over NVLink -- then raises this rank's flag there to seq + 1.
// have_alpha: the caller already holds the alpha of the two candidate rows (resident loop); else
// they are read from global memory.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void push_package(const PushDev& p, unsigned long long k_up, unsigned long long k_low,
                                             float a_up, float a_low, bool have_alpha, unsigned int seq)
{
    const int tid = threadIdx.x, nthr = blockDim.x;
    const unsigned int slot = seq & 1u;
    float f_dummy;
    int i_hi = p.own0, i_lo = p.own0;      // an empty set still ships a valid row (the first owned one), so that the
                                           // reference's "index 0, +-1e9" default (svmTrainMain.cpp:249-252) stays defined
    if (k_up != DPSVM_KEY_EMPTY) unpack_key_up(k_up, &f_dummy, &i_hi);
    if (k_low != DPSVM_KEY_EMPTY) unpack_key_low(k_low, &f_dummy, &i_lo);
    const int r_hi = i_hi - p.res0, r_lo = i_lo - p.res0;
    CandHeader h;
    h.key_up = k_up; h.key_low = k_low;
    h.a_hi = (have_alpha && k_up != DPSVM_KEY_EMPTY) ? a_up : __ldcg(p.alpha + r_hi);
    h.a_lo = (have_alpha && k_low != DPSVM_KEY_EMPTY) ? a_low : __ldcg(p.alpha + r_lo);
    h.y_hi = __ldg(p.y + r_hi); h.y_lo = __ldg(p.y + r_lo);
    h.xsq_hi = __ldg(p.xsq + r_hi); h.xsq_lo = __ldg(p.xsq + r_lo);
#pragma unroll
    for (int i = 0; i < 6; ++i) h.pad[i] = 0;
    const float4* src_hi = reinterpret_cast<const float4*>(p.x + (size_t)r_hi * p.pitch);
    const float4* src_lo = reinterpret_cast<const float4*>(p.x + (size_t)r_lo * p.pitch);
    const int pitch4 = p.pitch / 4;
    const size_t slot_off = ((size_t)slot * p.world + p.rank) * p.package;
    // each thread keeps its share of the two rows in registers and stores it to every peer
    for (int i = tid; i < 2 * pitch4; i += nthr) {
        const float4 v = (i < pitch4) ? __ldg(src_hi + i) : __ldg(src_lo + (i - pitch4));
        for (int q = 0; q < p.world; ++q) {
            float4* dst = reinterpret_cast<float4*>(p.peer_mail[q] + slot_off + sizeof(CandHeader));
            dst[i] = v;
        }
    }
    if (tid < p.world) {
        const uint4* hs = reinterpret_cast<const uint4*>(&h);
        uint4* dst = reinterpret_cast<uint4*>(p.peer_mail[tid] + slot_off);
#pragma unroll
        for (int i = 0; i < 4; ++i) dst[i] = hs[i];
    }
    // No per-thread system fence: the CTA barrier orders every thread's package stores before the flag
    // stores below, and st.release.sys is cumulative over what happens-before it (PTX memory model), so
    // a peer that acquires the flag sees the whole package.  (512 system fences cost 6-7 us here.)
    __syncthreads();
    if (tid < p.world) st_release_sys_u32(p.peer_flag[tid] + slot * p.world + p.rank, seq + 1u);
}

}  // namespace dpsvm
