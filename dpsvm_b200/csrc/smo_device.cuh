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
// Warp-wide minimum of 64-bit keys in two redux.sync: the high words, then the low words of the lanes that hold the
// minimal high word.  Every lane gets the result.
__device__ __forceinline__ unsigned long long warp_min_u64(unsigned long long k)
{
    const unsigned int hi = (unsigned int)(k >> 32), lo = (unsigned int)k;
    const unsigned int min_hi = __reduce_min_sync(0xffffffffu, hi);
    const unsigned int min_lo = __reduce_min_sync(0xffffffffu, hi == min_hi ? lo : 0xffffffffu);
    return ((unsigned long long)min_hi << 32) | min_lo;
}

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

__device__ __forceinline__ void st_relaxed_sys_u64(unsigned long long* p, unsigned long long v)
{
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_relaxed_sys_u64(const unsigned long long* p)
{
    unsigned long long v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_gpu_v2u64(unsigned long long* p, unsigned long long v0, unsigned long long v1)
{
    asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(v0), "l"(v1) : "memory");
}
__device__ __forceinline__ void ld_relaxed_gpu_v2u64(const unsigned long long* p, unsigned long long& v0, unsigned long long& v1)
{
    asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(v0), "=l"(v1) : "l"(p) : "memory");
}
__device__ __forceinline__ unsigned long long global_timer_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// ------------------------------------------------------------------------------------------------
// Flagged-word candidate exchange (LLDev, smo_kernels.cuh).  8-byte stores are single-copy atomic, and each word
// names the package it belongs to, so neither side needs a fence: the receiver spins on the words themselves.
// ------------------------------------------------------------------------------------------------
// data word `w` (0..5) of a rank's candidates
__device__ __forceinline__ unsigned int ll_word(int w, unsigned long long k_up, float a_up, unsigned long long k_low, float a_low)
{
    switch (w) {
        case 0: return (unsigned int)(k_up >> 32);
        case 1: return (unsigned int)k_up;
        case 2: return __float_as_uint(a_up);
        case 3: return (unsigned int)(k_low >> 32);
        case 4: return (unsigned int)k_low;
        default: return __float_as_uint(a_low);
    }
}
// store word `w` of package `seq` from rank `rank` into rank q's inbox
__device__ __forceinline__ void ll_send(const LLDev& ll, int q, int rank, unsigned int seq, int w, unsigned int data)
{
    unsigned long long* dst = ll.peer_in[q] + ((size_t)(seq & 1u) * kMaxPeers + rank) * kLLWords + w;
    st_relaxed_sys_u64(dst, ((unsigned long long)(seq + 1u) << 32) | data);
}
// Wait for package `seq` of every rank (threads tid < world * 8 take one word each and leave its data in
// s_ll[tid]); *s_ok is cleared on timeout.  Block-wide: contains two barriers.
__device__ __forceinline__ bool ll_wait(const unsigned long long* in, int world, unsigned int seq, unsigned int* s_ll, int* s_ok,
                                        unsigned long long timeout_ns)
{
    const int tid = threadIdx.x;
    if (tid == 0) *s_ok = 1;
    __syncthreads();
    if (tid < world * kLLWords && (tid & (kLLWords - 1)) < 6) {
        const unsigned long long* w = in + (size_t)(seq & 1u) * kMaxPeers * kLLWords + tid;
        const unsigned int want = seq + 1u;
        unsigned long long v = ld_relaxed_sys_u64(w);
        if ((unsigned int)(v >> 32) != want) {
            const unsigned long long t0 = global_timer_ns();
            unsigned int spins = 0;
            while ((unsigned int)((v = ld_relaxed_sys_u64(w)) >> 32) != want) {
                if ((++spins & 0xFFu) == 0 && global_timer_ns() - t0 > timeout_ns) { *s_ok = 0; break; }
            }
        }
        s_ll[tid] = (unsigned int)v;
    }
    __syncthreads();
    return *s_ok != 0;
}
// Global merge of the gathered candidates (svmTrainMain.cpp:255-274): minimum packed key = argmin f / argmax f with
// the lowest GLOBAL index among equal values; with every set empty the defaults are rank 0's (its first row is row 0).
__device__ __forceinline__ void ll_merge(const unsigned int* s_ll, int world, unsigned long long& g_up, float& a_up,
                                         unsigned long long& g_low, float& a_low)
{
    g_up = DPSVM_KEY_EMPTY; g_low = DPSVM_KEY_EMPTY;
    a_up = __uint_as_float(s_ll[2]); a_low = __uint_as_float(s_ll[5]);
    for (int r = 0; r < world; ++r) {
        const unsigned int* p = s_ll + r * kLLWords;
        const unsigned long long u = ((unsigned long long)p[0] << 32) | p[1], l = ((unsigned long long)p[3] << 32) | p[4];
        if (u < g_up) { g_up = u; a_up = __uint_as_float(p[2]); }
        if (l < g_low) { g_low = l; a_low = __uint_as_float(p[5]); }
    }
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
struct SpBuf {               // one chunk of kSpChunk steps for this lane
    unsigned int v[kSpChunk];   // fp32 bit patterns, exactly as the 256-bit load leaves them (no register moves)
    uint4 c;                 // 8 x 16-bit byte offsets into the interleaved (x_hi, x_lo) row
};
static_assert(kSpChunk == 8, "a chunk is one 256-bit value load and one 128-bit offset load per lane");

// 256-bit loads carry their L2 eviction priority in the opcode (LDG.E.NA.ELL2.256 / .EFL2.256): no policy
// register, no R2UR.  keep: 1 evict_last, 0 evict_first, -1 evict_normal.
template <bool C8>
__device__ __forceinline__ void sp_load(SpBuf& b, const SparseDev& sp, unsigned int chunk, int lane, int keep, uint64_t pol_keep, uint64_t pol_stream)
{
    // Two policies: keep (evict_last) for what should stay in L2 from one iteration to the next -- all of a stream that
    // fits, else its pinned prefix -- and evict_first for the rest of a stream that does not fit.  (keep < 0, "no hints",
    // is a stream small enough not to care: kept.)  The 256-bit value loads carry the policy in the opcode; narrower
    // loads cannot (ptxas), they take the policy register the caller selected.
    const float* vp = sp.val + ((size_t)chunk * 32 + lane) * kSpChunk;
    const void* cp = C8 ? static_cast<const void*>(reinterpret_cast<const uint2*>(sp.col8) + (size_t)chunk * 32 + lane)
                        : static_cast<const void*>(reinterpret_cast<const uint4*>(sp.col) + (size_t)chunk * 32 + lane);
    // (the policy register stays uniform inside each branch: no per-thread select, no move to a uniform register per load)
    if (keep != 0) {
        asm volatile("ld.global.nc.L1::no_allocate.L2::evict_last.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(b.v[0]), "=r"(b.v[1]), "=r"(b.v[2]), "=r"(b.v[3]), "=r"(b.v[4]), "=r"(b.v[5]), "=r"(b.v[6]), "=r"(b.v[7]) : "l"(vp));
        if (C8) asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.u32 {%0,%1}, [%2], %3;" : "=r"(b.c.x), "=r"(b.c.y) : "l"(cp), "l"(pol_keep));
        else asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
                          : "=r"(b.c.x), "=r"(b.c.y), "=r"(b.c.z), "=r"(b.c.w) : "l"(cp), "l"(pol_keep));
    } else {
        asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(b.v[0]), "=r"(b.v[1]), "=r"(b.v[2]), "=r"(b.v[3]), "=r"(b.v[4]), "=r"(b.v[5]), "=r"(b.v[6]), "=r"(b.v[7]) : "l"(vp));
        if (C8) asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.u32 {%0,%1}, [%2], %3;" : "=r"(b.c.x), "=r"(b.c.y) : "l"(cp), "l"(pol_stream));
        else asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
                          : "=r"(b.c.x), "=r"(b.c.y), "=r"(b.c.z), "=r"(b.c.w) : "l"(cp), "l"(pol_stream));
    }
}

// The resident loop's form of sp_load.  A chunk is named by `o`, THIS LANE's byte offset into the 8-bit column array,
// (chunk * 32 + lane) * 8: the lane's values sit at 4 * o and its 16-bit column offsets at 2 * o, so that the request side
// carries one running 32-bit number (+256 per chunk) instead of rebuilding two 64-bit addresses from a chunk index
// (kSpMaxStepsO bounds the stream so that it fits).
// KEEP_ALL: the whole stream fits in L2 -- one value load (evict_last in its opcode) and an unhinted column load, no
// policy registers; otherwise the two policies of sp_load.
template <bool C8, bool KEEP_ALL>
__device__ __forceinline__ void sp_load_at(SpBuf& b, const SparseDev& sp, unsigned int o, int keep, uint64_t pol_keep, uint64_t pol_stream)
{
    const unsigned char* vp = reinterpret_cast<const unsigned char*>(sp.val) + (size_t)o * 4u;
    const unsigned char* cp = C8 ? reinterpret_cast<const unsigned char*>(sp.col8) + o
                                 : reinterpret_cast<const unsigned char*>(sp.col) + (size_t)o * 2u;
    if (KEEP_ALL || keep != 0) {
        asm volatile("ld.global.nc.L1::no_allocate.L2::evict_last.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(b.v[0]), "=r"(b.v[1]), "=r"(b.v[2]), "=r"(b.v[3]), "=r"(b.v[4]), "=r"(b.v[5]), "=r"(b.v[6]), "=r"(b.v[7]) : "l"(vp));
        if (KEEP_ALL) {
            if (C8) asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(b.c.x), "=r"(b.c.y) : "l"(cp));
            else asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(b.c.x), "=r"(b.c.y), "=r"(b.c.z), "=r"(b.c.w) : "l"(cp));
        } else {
            if (C8) asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.u32 {%0,%1}, [%2], %3;" : "=r"(b.c.x), "=r"(b.c.y) : "l"(cp), "l"(pol_keep));
            else asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
                              : "=r"(b.c.x), "=r"(b.c.y), "=r"(b.c.z), "=r"(b.c.w) : "l"(cp), "l"(pol_keep));
        }
    } else {
        asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(b.v[0]), "=r"(b.v[1]), "=r"(b.v[2]), "=r"(b.v[3]), "=r"(b.v[4]), "=r"(b.v[5]), "=r"(b.v[6]), "=r"(b.v[7]) : "l"(vp));
        if (C8) asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.u32 {%0,%1}, [%2], %3;" : "=r"(b.c.x), "=r"(b.c.y) : "l"(cp), "l"(pol_stream));
        else asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
                          : "=r"(b.c.x), "=r"(b.c.y), "=r"(b.c.z), "=r"(b.c.w) : "l"(cp), "l"(pol_stream));
    }
}

// s_hl: the selected pair interleaved, s_hl[col] = (x_hi[col], x_lo[col]).
//   16-bit format: the stream carries BYTE offsets into it.
//   8-bit format (C8): it carries, per lane, the distance in columns from the lane's previous entry (from column 0 at
//   the start of a unit; padding repeats the last column); `pos` is the lane's running byte offset.
// REP = 4: the pair is replicated sixteen times, s_rep[col][16] float2 (128 bytes per column), and s_hl / pos point at THIS
// lane's copy, s_rep[0][lane & 15]: lane l of a half-warp only ever touches bank pair l, so every 64-bit gather is free of
// bank conflicts whatever the columns are (REP = 0: the plain interleaved row, sixteen random columns collide on 16 bank pairs).
template <bool C8, int REP = 0>
__device__ __forceinline__ void sp_fma(const SpBuf& b, const float2* s_hl, float& acc_hi, float& acc_lo, unsigned int& pos)
{
    const unsigned char* base = reinterpret_cast<const unsigned char*>(s_hl);
    if (C8) {
        // `pos` is a 32-bit shared-memory ADDRESS here (it starts at s_hl's own address at the first step of a unit): per
        // entry one byte permute, one multiply-add and the load -- left to itself the compiler builds a longer
        // shift / mask / add tree and adds the base separately
        const unsigned int w[2] = {b.c.x, b.c.y};
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            unsigned int dist;
            asm("prmt.b32 %0, %1, 0, %2;" : "=r"(dist) : "r"(w[j >> 2]), "r"(0x4440u | (unsigned int)(j & 3)));
            asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(pos) : "r"(dist), "n"(8 << REP));
            float2 h;
            asm("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(h.x), "=f"(h.y) : "r"(pos));
            const float v = __uint_as_float(b.v[j]);
            acc_hi = __fmaf_rn(v, h.x, acc_hi);
            acc_lo = __fmaf_rn(v, h.y, acc_lo);
        }
    } else {
        const unsigned int w[4] = {b.c.x, b.c.y, b.c.z, b.c.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 h0 = *reinterpret_cast<const float2*>(base + ((w[j] & 0xFFFFu) << REP));
            const float2 h1 = *reinterpret_cast<const float2*>(base + ((w[j] >> 16) << REP));
            const float v0 = __uint_as_float(b.v[2 * j]), v1 = __uint_as_float(b.v[2 * j + 1]);
            acc_hi = __fmaf_rn(v0, h0.x, acc_hi);
            acc_lo = __fmaf_rn(v0, h0.y, acc_lo);
            acc_hi = __fmaf_rn(v1, h1.x, acc_hi);
            acc_lo = __fmaf_rn(v1, h1.y, acc_lo);
        }
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

// (The resident loop's sparse sweep walks its units with one cursor and a tag per chunk in flight, smo_resident.cu; the
// cursor pair below serves the graph loop's pass kernel.)
// A warp's position in its chunk sequence: sequence unit k (k_first, k_first + k_stride, ... < k_end), chunk c of
// the unit's nc chunks.  SEQ supplies  void unit(int k, unsigned int& first_step, int& steps)  and the L2 policy
// int keep(unsigned int first_step).
struct SpCursor {
    int k, c, nc;
    unsigned int chunk0;
    int keep;
};
template <class SEQ>
__device__ __forceinline__ void sp_open(SpCursor& q, const SEQ& seq)
{
    q.c = 0;
    if (q.k < seq.k_end) {
        unsigned int beg;
        int steps;
        seq.unit(q.k, beg, steps);
        q.chunk0 = beg / kSpChunk;
        q.nc = steps / kSpChunk;
        q.keep = seq.keep(beg);
    }
}
template <class SEQ>
__device__ __forceinline__ void sp_next(SpCursor& q, const SEQ& seq)
{
    if (++q.c >= q.nc) { q.k += seq.k_stride; sp_open(q, seq); }
}

// Graph loop: units of the whole shard, first steps in global memory.
struct SpSeqGrid {
    int k_first, k_stride, k_end;       // k_end = units of the shard
    const unsigned int* unit_off;
    bool rev;
    int hint;
    unsigned int pin_steps;
    __device__ __forceinline__ int local(int k) const { return rev ? (k_end - 1 - k) : k; }
    __device__ __forceinline__ void unit(int k, unsigned int& beg, int& steps) const
    {
        const int u = local(k);
        beg = __ldg(unit_off + u);
        steps = (int)(__ldg(unit_off + u + 1) - beg);
    }
    __device__ __forceinline__ int keep(unsigned int beg) const { return hint ? (beg < pin_steps ? 1 : 0) : -1; }
};

// Put the first kSpRing chunks of a sequence in flight (ring slots 0 ..), e.g. for the NEXT iteration while the
// barrier and the merge of this one run.
template <bool C8, class SEQ>
__device__ __forceinline__ void sp_prefetch(SpBuf (&ring)[kSpRing], const SparseDev& sp, const SEQ& seq, int lane, uint64_t pol_keep, uint64_t pol_stream)
{
    SpCursor p;
    p.k = seq.k_first;
    sp_open(p, seq);
#pragma unroll
    for (int b = 0; b < kSpRing; ++b) {
        if (p.k < seq.k_end) {
            sp_load<C8>(ring[b], sp, p.chunk0 + (unsigned int)p.c, lane, p.keep, pol_keep, pol_stream);
            sp_next(p, seq);
        }
    }
}

// The sweep of one warp over its sequence.  The first kSpRing chunks are already in flight (sp_prefetch); every
// consumed slot is refilled at once with the chunk kSpRing ahead.  (Measured: rings of 2, 3, 4 and 6 chunks run
// at the same speed -- the sweep is bound by the shared-memory gathers, not by loads in flight.)  begin(k) runs when a unit starts (row scalars), end(k, dot_hi, dot_lo) after its last chunk, with the
// G-lane sums done.
template <int G, bool C8, class SEQ, class BEGIN, class END>
__device__ __forceinline__ void sp_sweep(SpBuf (&ring)[kSpRing], const SparseDev& sp, const SEQ& seq, const float2* s_hl, int lane,
                                         uint64_t pol_keep, uint64_t pol_stream, BEGIN begin, END end)
{
    SpCursor cns, prd;
    cns.k = seq.k_first;
    sp_open(cns, seq);
    prd = cns;
#pragma unroll
    for (int b = 0; b < kSpRing; ++b) if (prd.k < seq.k_end) sp_next(prd, seq);
    float acc_hi = 0.f, acc_lo = 0.f;
    const unsigned int pos0 = C8 ? smem_u32(s_hl) : 0u;
    unsigned int pos = pos0;                 // 8-bit format: the shared-memory address of this lane's current column in s_hl
    bool more = cns.k < seq.k_end;
    while (more) {
#pragma unroll
        for (int b = 0; b < kSpRing; ++b) {
            if (more) {
                if (cns.c == 0) { begin(cns.k); pos = pos0; }
                sp_fma<C8>(ring[b], s_hl, acc_hi, acc_lo, pos);
                if (cns.c + 1 >= cns.nc) {
                    sp_group_sum<G>(acc_hi, acc_lo);
                    end(cns.k, acc_hi, acc_lo);
                    acc_hi = 0.f; acc_lo = 0.f;
                }
                sp_next(cns, seq);
                more = cns.k < seq.k_end;
                if (prd.k < seq.k_end) {
                    sp_load<C8>(ring[b], sp, prd.chunk0 + (unsigned int)prd.c, lane, prd.keep, pol_keep, pol_stream);
                    sp_next(prd, seq);
                }
            }
        }
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
    if (p.ll.in) {
        // every rank holds all of X: only keys and alphas travel, as flagged words (one thread per word and peer)
        if (tid < p.world * kLLWords && (tid & (kLLWords - 1)) < 6) {
            int i_up = p.own0, i_low = p.own0;          // empty set: the alpha of the first owned row (rank 0: global row 0)
            if (k_up != DPSVM_KEY_EMPTY) unpack_key_up(k_up, &f_dummy, &i_up);
            if (k_low != DPSVM_KEY_EMPTY) unpack_key_low(k_low, &f_dummy, &i_low);
            const int w = tid & (kLLWords - 1);
            float au = a_up, al = a_low;
            if (w == 2 && !(have_alpha && k_up != DPSVM_KEY_EMPTY)) au = __ldcg(p.alpha + (i_up - p.res0));
            if (w == 5 && !(have_alpha && k_low != DPSVM_KEY_EMPTY)) al = __ldcg(p.alpha + (i_low - p.res0));
            ll_send(p.ll, tid / kLLWords, p.rank, seq, w, ll_word(w, k_up, au, k_low, al));
        }
        return;
    }
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
