// smo_resident.cu -- the "resident loop": up to `chunk` complete SMO iterations inside ONE persistent,
// cooperatively launched kernel (one CTA per SM), replacing the [control kernel, pass kernel] pair
// of the graph loop for the common case (rows up to 1 024 floats, kernel-row cache of at most 64 lines;
// one GPU, or several over the peer-memory exchange).
//
// Same arithmetic as the graph loop, bit for bit (same dot-product order, same eta reduction order,
// same epilogue), so the two loops are interchangeable and are tested against each other.
// Replaces, per iteration (paths in farshid83/dpsvm): train_step1 (svmTrain.cu:469-483), the host
// merge / eta / alpha update (svmTrainMain.cpp:246-302), train_step2 + update_f (svmTrain.cu:485-497,
// 190-302) and the cache lookups (cache.cu:62-105).
//
// Why it is faster than two kernels per iteration:
//   * No leader, no kernel boundary: after its sweep a CTA publishes its block-best candidates (keys
//     plus the alpha of those rows) in a per-CTA slot and arrives at a grid barrier; every CTA then
//     reads all slots and derives the next pair, eta and the alpha update REDUNDANTLY and
//     identically.  The dependency chain per iteration is barrier -> slots -> two rows by TMA ->
//     block reduction (2.8 us), instead of two kernel boundaries around a single-CTA control kernel
//     (5.6 us).
//   * On-chip slice: each CTA keeps some units of its row range -- rows AND their |x|^2, y, f, alpha --
//     in shared memory for the whole launch; they are never re-read from HBM/L2.  The whole shard when it
//     fits (Adult 32 561 x 123, or a 7 500-row shard of an 8-GPU run), otherwise 16 units per CTA: the
//     carve-out comes out of L1, and L1 capacity bounds the streaming loads an SM keeps in flight
//     (profiles/r01_resident_trace.md).
//   * The first streamed unit of the next iteration is loaded into registers before the barrier, so HBM
//     keeps streaming through the dependency chain.
//   * The kernel-row cache's LRU bookkeeping is a replicated state machine: every CTA holds the same
//     small table in shared memory and applies the same lookups, so no CTA waits for another's
//     decision.
//   * Multi-GPU: after the local barrier CTA 0 stores the rank's candidate package into every peer's
//     mailbox over NVLink; every CTA then waits on the arrival flags and merges the packages itself.
#include <algorithm>
#include <limits.h>
#include <mutex>
#include <math.h>
#include <stdlib.h>
#include <type_traits>

#include "smo_device.cuh"

namespace dpsvm {

__device__ __forceinline__ unsigned int ld_acquire_gpu_u32(const unsigned int* p)
{
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// All CTAs of the (co-resident) grid arrive; returns false on timeout.
__device__ __forceinline__ bool grid_barrier(unsigned int* counter, unsigned int target, int* s_flag, unsigned long long timeout_ns)
{
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        // the last CTA to arrive is the critical path: it learns from its own atomic that everyone is
        // here and skips the polling round trip
        const unsigned int before = atomicAdd(counter, 1u);
        int ok = 1;
        unsigned long long t0;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        unsigned int spins = 0;
        while (before + 1u < target && ld_acquire_gpu_u32(counter) < target) {
            if ((++spins & 0x3FFu) == 0) {
                unsigned long long t1;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
                if (t1 - t0 > timeout_ns) { ok = 0; break; }      // seconds: something is badly wrong
            }
        }
        *s_flag = ok;
    }
    __syncthreads();
    return *s_flag != 0;
}

struct Cand {
    unsigned long long key;
    float alpha;
};
__device__ __forceinline__ void cand_min(Cand& a, unsigned long long k, float al)
{
    if (k < a.key) { a.key = k; a.alpha = al; }
}
// Warp-wide minimum of the packed keys, with the alpha that belongs to the winner; every lane ends up with the
// result.  Two redux.sync (high word, then low word among the lanes that hold the minimal high word), a ballot and
// one shuffle instead of five butterfly levels of three shuffles each: this sits on the per-iteration dependency
// chain four times (block best and grid best, two stages each).
__device__ __forceinline__ void cand_warp_min(Cand& a)
{
    const unsigned int hi = (unsigned int)(a.key >> 32), lo = (unsigned int)a.key;
    const unsigned int min_hi = __reduce_min_sync(0xffffffffu, hi);
    const unsigned int min_lo = __reduce_min_sync(0xffffffffu, hi == min_hi ? lo : 0xffffffffu);
    const unsigned int who = __ballot_sync(0xffffffffu, hi == min_hi && lo == min_lo);     // never empty
    a.alpha = __shfl_sync(0xffffffffu, a.alpha, __ffs((int)who) - 1);
    a.key = ((unsigned long long)min_hi << 32) | min_lo;
}

// Exact LRU over a small table, replicated in every CTA (cache.cu:62-105 semantics: a lookup hit
// refreshes recency; a miss takes a free line, else evicts the least recently used one).
struct LruTable {
    int* key;      // [lines]
    int* stamp;    // [lines]
    int* meta;     // [0] = size, [1] = clock
};
// One lookup by a whole warp (lines <= 64: two table entries per lane): the match is a ballot, the victim of a full table
// the lane holding the smallest stamp (stamps are distinct).  Same decisions as the serial scan it replaces -- first free
// line, else least recently used -- in a few dozen cycles instead of a walk over the table by one thread, which sat
// on the critical path of every iteration (measured: the reference's default 10 lines cost 1.8 us per iteration).
__device__ __forceinline__ int lru_lookup_warp(const LruTable& t, int lines, int k, int* hit)
{
    const int lane = threadIdx.x & 31;
    const int size = t.meta[0];
    const bool m0 = lane < size && t.key[lane] == k, m1 = lane + 32 < size && t.key[lane + 32] == k;
    const unsigned int b0 = __ballot_sync(0xffffffffu, m0), b1 = __ballot_sync(0xffffffffu, m1);
    int line;
    *hit = (b0 | b1) != 0u;
    if (b0 | b1) {
        line = b0 ? __ffs((int)b0) - 1 : 32 + __ffs((int)b1) - 1;
    } else if (size < lines) {
        line = size;
    } else {
        const unsigned int s0 = lane < size ? (unsigned int)t.stamp[lane] : 0xffffffffu;
        const unsigned int s1 = lane + 32 < size ? (unsigned int)t.stamp[lane + 32] : 0xffffffffu;
        const unsigned int mn = __reduce_min_sync(0xffffffffu, min(s0, s1));
        const unsigned int v0 = __ballot_sync(0xffffffffu, s0 == mn), v1 = __ballot_sync(0xffffffffu, s1 == mn);
        line = v0 ? __ffs((int)v0) - 1 : 32 + __ffs((int)v1) - 1;
    }
    __syncwarp();
    if (lane == 0) {
        if (!*hit) { t.key[line] = k; if (size < lines) t.meta[0] = size + 1; }
        t.stamp[line] = ++t.meta[1];
    }
    __syncwarp();
    return line;
}

// SG == 0: the sweep reads the dense rows (LPR lanes per row, ITERS float4 chunks per lane).
// SG  > 0: the sweep reads the sparse stream with SG lanes per row (SparseDev, smo_kernels.cuh); LPR and
//          ITERS then only size the two selected-row buffers.  Everything outside the sweep is shared.
// C8: the sparse stream carries 8-bit column distances instead of 16-bit offsets (SparseDev::col8).
template <int LPR, int ITERS, int SG, bool C8 = false, bool KEEP_ALL = false>
__global__ void __launch_bounds__(kResThreads, 1) smo_resident_kernel(const ResidentDev a)
{
    constexpr bool SPARSE = SG > 0;
    constexpr int R = SPARSE ? 1 : kRowsInFlight;
    constexpr int GROUPS = SPARSE ? 32 / (SPARSE ? SG : 1) : 32 / LPR;
    constexpr int RPW = GROUPS * R;                    // rows per unit
    constexpr int CAP4 = LPR * ITERS;
    constexpr int NW = kResThreads / 32;
    constexpr int SGL = SPARSE ? SG : 32;              // lanes per row in the sparse sweep

    extern __shared__ __align__(128) unsigned char dyn_smem[];
    float4* s_hi = reinterpret_cast<float4*>(dyn_smem);
    float4* s_lo = s_hi + CAP4;
    float4* s_x = s_lo + CAP4;                         // dense: stationary rows [res_units * RPW][pitch4]
    // sparse: the pair interleaved, (x_hi[c], x_lo[c]), then the first step of each of this CTA's units
    // (sixteen copies of it, s_rep[column][16], so that the sweep's gathers are free of bank conflicts: sp_fma<., 4>)
    float4* s_rep = s_lo + CAP4;
    const float2* s_hl = reinterpret_cast<const float2*>(reinterpret_cast<const unsigned char*>(s_rep) + (threadIdx.x & 15) * 8);   // this lane's copy
    unsigned int* s_uoff = reinterpret_cast<unsigned int*>(s_rep + (size_t)a.pitch4 * 4 * 8);
    // ... the units each warp slot sweeps (SparseDev::perm, cnt) ...
    unsigned short* s_perm = reinterpret_cast<unsigned short*>(s_uoff + ((a.sp.max_cta_units + 1 + 3) & ~3));
    unsigned short* s_cnt = s_perm + ((a.sp.perm_stride + 7) & ~7);
    // (The stream itself stays in L2: a head of it held in shared memory for the whole launch, as the dense sweep holds
    // rows, cost the sweep more in per-chunk instructions than the L2 reads it saved -- measured at 15 000 to 60 000 rows.)
    // their per-row scalars live on chip as well (a warp round on a stationary unit then touches no
    // global memory at all): |x|^2, y, f, alpha -- [a.res_units * RPW] each, after the rows
    float* s_rxsq = reinterpret_cast<float*>(s_x + (size_t)a.res_units * RPW * a.pitch4);
    int* s_ry = reinterpret_cast<int*>(s_rxsq + a.res_units * RPW);
    float* s_rf = reinterpret_cast<float*>(s_ry + a.res_units * RPW);
    float* s_ralpha = s_rf + a.res_units * RPW;

    __shared__ __align__(8) uint64_t s_bar;
    __shared__ float s_part[8];
    __shared__ Cand s_cand[2][NW];
    __shared__ Cand s_cand2[2][NW];
    __shared__ int s_flag;
    __shared__ int s_lru_key[kResCacheLines], s_lru_stamp[kResCacheLines], s_lru_meta[2];
    __shared__ float s_delta[4];                       // delta_hi, delta_lo, alpha_hi_new, alpha_lo_new of this iteration
    __shared__ struct { int both_hit; int hits; float* line_hi; float* line_lo; float xsq_hi, xsq_lo; } s_early;   // known before eta
    __shared__ unsigned int s_ll[kMaxPeers * kLLWords];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int sub = SPARSE ? lane % SGL : lane % LPR, grp = SPARSE ? lane / SGL : lane / LPR;
    const int G = gridDim.x, cta = blockIdx.x;
    // units are dealt to the warps round-robin starting from the LAST warp: when the count is not a multiple of NW the
    // spare time falls to warp 0, whose thread 0 computes eta while the others already sweep
    const int wslot = NW - 1 - warp;
    DevState* st = a.st;

    // ---- launch-time state (identical in every CTA; nobody writes it before the first barrier) ----
    int num_iter = st->num_iter;
    {
        const int done = st->done, stop_at = st->stop_at;
        if (done || num_iter >= stop_at) return;
    }
    const int stop_at = st->stop_at;
    long long cache_hits = st->cache_hits, cache_lookups = st->cache_lookups;

    const int n_units = (a.n + RPW - 1) / RPW;
    // contiguous unit range of this CTA: equal unit counts (dense rows all cost the same), or the ranges
    // the host balanced by steps (sparse rows differ in length)
    const int u0 = SPARSE ? a.sp.cta_first[cta] : (int)((long long)n_units * cta / G);
    const int u1 = SPARSE ? a.sp.cta_first[cta + 1] : (int)((long long)n_units * (cta + 1) / G);
    const int res_units = SPARSE ? 0 : min(u1 - u0, a.res_units);   // units kept in shared memory
    const int us0 = u0 + res_units;                    // first streamed unit

    // the pair of the coming iteration (uniform over the CTA, identical in every CTA and every rank)
    int i_hi = 0, i_lo = 0;
    float b_hi = 0.f, b_lo = 0.f, a_hi_old = 0.f, a_lo_old = 0.f;
    // multi-GPU: scalars and rows of the pair come from the gathered candidate packages
    int m_y_hi = 0, m_y_lo = 0;
    float m_xsq_hi = 0.f, m_xsq_lo = 0.f;
    const float4* m_src_hi = nullptr;
    const float4* m_src_lo = nullptr;

    // Multi-GPU: wait until every rank's package number `seq` has landed in slot (seq & 1) of the local
    // mailbox, then merge them exactly as the control kernel does (svmTrainMain.cpp:255-274): the minimum
    // packed key is argmin f / argmax f with the lowest GLOBAL index among equal values.
    auto acquire_remote = [&](unsigned int seq) -> bool {
        if (tid == 0) s_flag = 1;
        __syncthreads();
        if (tid < a.world) {
            const unsigned int want = seq + 1u;
            const unsigned int* fl = a.flags + (size_t)(seq & 1u) * a.world + tid;
            unsigned long long t0;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
            unsigned int spins = 0;
            while (ld_acquire_sys_u32(fl) != want) {
                if ((++spins & 0xFFu) == 0) {
                    unsigned long long t1;
                    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
                    if (t1 - t0 > a.timeout_ns) { s_flag = 0; break; }
                }
            }
        }
        __syncthreads();
        if (!s_flag) return false;
        const unsigned char* cand = a.mail + (size_t)(seq & 1u) * a.world * a.package;
        unsigned long long g_up = DPSVM_KEY_EMPTY, g_low = DPSVM_KEY_EMPTY;
        int w_up = 0, w_low = 0;
        for (int r = 0; r < a.world; ++r) {
            const CandHeader* h = reinterpret_cast<const CandHeader*>(cand + (size_t)r * a.package);
            const unsigned long long u = __ldcg(&h->key_up), l = __ldcg(&h->key_low);
            if (u < g_up) { g_up = u; w_up = r; }
            if (l < g_low) { g_low = l; w_low = r; }
        }
        unpack_key_up(g_up, &b_hi, &i_hi);
        unpack_key_low(g_low, &b_lo, &i_lo);
        const CandHeader* h_hi = reinterpret_cast<const CandHeader*>(cand + (size_t)w_up * a.package);
        const CandHeader* h_lo = reinterpret_cast<const CandHeader*>(cand + (size_t)w_low * a.package);
        a_hi_old = __ldcg(&h_hi->a_hi); a_lo_old = __ldcg(&h_lo->a_lo);
        m_y_hi = __ldcg(&h_hi->y_hi); m_y_lo = __ldcg(&h_lo->y_lo);
        m_xsq_hi = __ldcg(&h_hi->xsq_hi); m_xsq_lo = __ldcg(&h_lo->xsq_lo);
        m_src_hi = reinterpret_cast<const float4*>(h_hi + 1);                  // package: header, row_hi, row_lo
        m_src_lo = reinterpret_cast<const float4*>(h_lo + 1) + a.pitch4;
        return true;
    };

    // Same, when every rank holds all of X (flagged-word exchange): only keys and alphas arrive; the rows, y and |x|^2 of
    // the pair are then read from local memory exactly as on one GPU.
    const bool ll_mode = a.world > 1 && a.ll.in != nullptr;
    auto acquire_ll = [&](unsigned int seq) -> bool {
        if (!ll_wait(a.ll.in, a.world, seq, s_ll, &s_flag, a.timeout_ns)) return false;
        unsigned long long g_up, g_low;
        ll_merge(s_ll, a.world, g_up, a_hi_old, g_low, a_lo_old);
        unpack_key_up(g_up, &b_hi, &i_hi);
        unpack_key_low(g_low, &b_lo, &i_lo);
        if (g_up == DPSVM_KEY_EMPTY) i_hi = a.empty_row;
        if (g_low == DPSVM_KEY_EMPTY) i_lo = a.empty_row;
        return true;
    };
    auto acquire = [&](unsigned int seq) -> bool { return ll_mode ? acquire_ll(seq) : acquire_remote(seq); };

    if (a.world <= 1) {
        // the selection left in the global keys by setup / the previous launch
        const unsigned long long k_up = a.keys[0], k_low = a.keys[1];
        unpack_key_up(k_up, &b_hi, &i_hi);
        unpack_key_low(k_low, &b_lo, &i_lo);
        // empty set: the reference's default is global row 0 (svmTrainMain.cpp:249-252); the first owned row stands
        // in for it when the rows resident here start later (same rule as smo_control_kernel)
        if (k_up == DPSVM_KEY_EMPTY) i_hi = a.empty_row;
        if (k_low == DPSVM_KEY_EMPTY) i_lo = a.empty_row;
        a_hi_old = a.alpha[i_hi - a.index_base];
        a_lo_old = a.alpha[i_lo - a.index_base];
    }
    if (SPARSE) {
        for (int i = tid; i <= u1 - u0; i += kResThreads) s_uoff[i] = a.sp.unit_off[u0 + i];
        for (int i = tid; i < a.sp.perm_stride; i += kResThreads) s_perm[i] = a.sp.perm[(size_t)cta * a.sp.perm_stride + i];
        if (tid < NW) s_cnt[tid] = a.sp.cnt[cta * NW + tid];
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int i = a.pitch4 + tid; i < CAP4; i += kResThreads) { s_hi[i] = z; s_lo[i] = z; }
    } else {
        // stationary slice -> shared memory; zero the tails of the two selected-row buffers once
        const int total4 = res_units * RPW * a.pitch4;
        const float4* src = a.x4 + (size_t)u0 * RPW * a.pitch4;
        const long long valid4 = (long long)(a.n - (long long)u0 * RPW) * a.pitch4;     // rows past n do not exist
        for (int i = tid; i < total4; i += kResThreads)
            s_x[i] = (i < valid4) ? ldg_stream4(src + i) : make_float4(0.f, 0.f, 0.f, 0.f);
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int i = a.pitch4 + tid; i < CAP4; i += kResThreads) { s_hi[i] = z; s_lo[i] = z; }
        const int row_b = u0 * RPW;
        for (int i = tid; i < res_units * RPW; i += kResThreads) {
            const int row = row_b + i;
            const bool ok = row < a.n;
            s_rxsq[i] = ok ? a.xsq[row] : 0.f;
            s_ry[i] = ok ? a.y[row] : 0;
            s_rf[i] = ok ? a.f[row] : 0.f;
            s_ralpha[i] = ok ? a.alpha[row] : 0.f;
        }
    }
    if (a.cache_lines > 0 && !a.cache_direct) {
        for (int i = tid; i < a.cache_lines; i += kResThreads) { s_lru_key[i] = a.lru_key[i]; s_lru_stamp[i] = a.lru_stamp[i]; }
        if (tid < 2) s_lru_meta[tid] = a.lru_meta[tid];
    }
    if (tid == 0) mbar_init(&s_bar, 1);
    unsigned int bar_target = (unsigned int)G;
    // everyone has read alpha of the initial pair before any owner overwrites it
    if (!grid_barrier(a.barrier, bar_target, &s_flag, a.timeout_ns)) { if (cta == 0 && tid == 0) { st->error = 1; st->done = 1; } return; }
    bar_target += (unsigned int)G;
    if (a.world > 1 && !acquire((unsigned int)num_iter)) { if (cta == 0 && tid == 0) { st->error = 1; st->done = 1; } return; }

    uint64_t pol_keep, pol_stream;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol_keep));
    if (a.stream_evict_first) asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol_stream));
    else asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(pol_stream));

    // per-unit registers; they survive from the end of one iteration to the start of the next (prefetch)
    const int my_units = u1 - u0;
    const int n_stream = my_units - res_units;
    float4 xv[R][ITERS];
    SpBuf sp_ring[kSpRing];                            // sparse: the chunks in flight ...
    unsigned int sp_tag[kSpRing];                      // ... each with a tag: the unit's first row | kTagLast on a unit's last chunk, else 0
    int sp_total = 0;                                  // chunks in this warp slot's sequence (the same in every iteration)
    if (SPARSE) {
        const int cnt = (int)s_cnt[wslot];
        for (int j = lane; j < cnt; j += 32) {
            const int lu = (int)s_perm[j * NW + wslot];
            sp_total += (int)(s_uoff[lu + 1] - s_uoff[lu]) / kSpChunk;
        }
        sp_total = __reduce_add_sync(0xffffffffu, sp_total);
    }
    // the ONE cursor of the sparse sweep: where the next chunk to REQUEST is (unit pj of this warp slot's list, sp_left of
    // its chunks to go, the next one at this lane's offset sp_po, see sp_load_at); the consumer needs no cursor of its own,
    // a chunk's tag says where its unit ends (and which rows it was)
    int sp_pj = 0, sp_left = 0, sp_pkeep = 0, sp_cnt = 0;
    unsigned int sp_po = 0, sp_utag = 0;
    bool sp_prev = false;
    float e_xsq = 0.f, e_f = 0.f, e_alpha = 0.f;
    int e_y = 0, erow = 0;
    bool eown = false, pre = false, unit_in_smem = false;
    int unit_first_row = 0;                            // dense sweep: first row of the unit whose X is in xv
    // constants of the k-th unit of this warp's sequence: streamed units first (forward or reverse),
    // the shared-memory slice last (it fills the tail while the slower SMs finish streaming)
    auto load_const = [&](int k, bool rv) {
        const bool in_smem = k >= n_stream;
        unit_in_smem = in_smem;
        const int u = in_smem ? (u0 + (k - n_stream)) : (rv ? (u1 - 1 - k) : (us0 + k));
        const int row0 = u * RPW + grp * R;
        unit_first_row = u * RPW;
        if constexpr (RPW > 4) {             // geometries whose epilogue runs per unit (see the dense sweep)
            erow = row0 + sub;
            eown = (sub < R) && (erow < a.n);
            if (eown) {
                if (in_smem) { e_xsq = s_rxsq[erow - u0 * RPW]; e_y = s_ry[erow - u0 * RPW]; }
                else { e_xsq = __ldg(a.xsq + erow); e_y = __ldg(a.y + erow); }
            }
        }
        if (in_smem) {
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const float4* rowp = s_x + (size_t)((u - u0) * RPW + grp * R + r) * a.pitch4;
#pragma unroll
                for (int itc = 0; itc < ITERS; ++itc) {
                    const int c4 = sub + itc * LPR;
                    xv[r][itc] = (c4 < a.pitch4) ? rowp[c4] : make_float4(0.f, 0.f, 0.f, 0.f);
                }
            }
        } else {
            const uint64_t pol = (u - us0 < a.pin_units) ? pol_keep : pol_stream;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const float4* rowp = a.x4 + (size_t)(row0 + r) * a.pitch4;
                const bool row_ok = (row0 + r) < a.n;
#pragma unroll
                for (int itc = 0; itc < ITERS; ++itc) {
                    const int c4 = sub + itc * LPR;
                    xv[r][itc] = (row_ok && c4 < a.pitch4) ? ldg_x4<true>(rowp + c4, pol) : make_float4(0.f, 0.f, 0.f, 0.f);
                }
            }
        }
    };

    auto stamp = [&](int it, int which) {
        if (a.trace && it == a.trace_iter && tid == 0) {
            unsigned long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            a.trace[(size_t)cta * 16 + which] = t;
        }
    };

    for (int it = 0;; ++it) {
        stamp(it, 0);
        // ================= pair -> rows, eta, alpha update (every CTA, identically) =================
        const int r_hi = i_hi - a.index_base, r_lo = i_lo - a.index_base;     // may lie outside [0, n) on a multi-GPU run
        int y_hi, y_lo;
        float xsq_hi, xsq_lo;
        // kernel-row cache, decided while the rows are in flight (it needs the pair only): a table of up to 64 lines is a
        // replicated state machine run by thread 0; a cache with a line for EVERY row never evicts, its line is the
        // key itself, and "filled in an earlier iteration" is read from a stamp by every thread for itself
        int hit_hi = 0, hit_lo = 0;
        if (a.cache_direct) {
            // stamp = 1 + the iteration that filled the line; this iteration's own stamp (written below by CTA 0, possibly
            // before another CTA reads it) does not count as a hit
            const int s_h = __ldcg(a.cache_stamp + i_hi), s_l = __ldcg(a.cache_stamp + i_lo);
            hit_hi = (s_h != 0 && s_h <= num_iter);
            hit_lo = (i_lo == i_hi) || (s_l != 0 && s_l <= num_iter);      // hi then lo (svmTrain.cu:203,238): the second lookup finds the first
        }
        if (a.world > 1 && !ll_mode) {
            // rows travelled in the packages (written by peers over NVLink): plain L2 loads into smem
            for (int i = tid; i < a.pitch4; i += kResThreads) { s_hi[i] = __ldcg(m_src_hi + i); s_lo[i] = __ldcg(m_src_lo + i); }
            y_hi = m_y_hi; y_lo = m_y_lo; xsq_hi = m_xsq_hi; xsq_lo = m_xsq_lo;
        } else {
            if (tid == 0) {
                const uint32_t bytes = (uint32_t)a.pitch4 * 16u;
                mbar_expect_tx(&s_bar, 2u * bytes);
                tma_bulk_g2s(s_hi, a.x4 + (size_t)r_hi * a.pitch4, bytes, &s_bar);
                tma_bulk_g2s(s_lo, a.x4 + (size_t)r_lo * a.pitch4, bytes, &s_bar);
            }
            y_hi = __ldg(a.y + r_hi); y_lo = __ldg(a.y + r_lo);
            xsq_hi = __ldg(a.xsq + r_hi); xsq_lo = __ldg(a.xsq + r_lo);
        }
        if (a.cache_direct) {
            if (tid == 0) {
                float* p_hi = a.cache_data + (size_t)i_hi * a.line_stride;
                float* p_lo = a.cache_data + (size_t)i_lo * a.line_stride;
                float* l_hi_p = nullptr;
                float* l_lo_p = nullptr;
                int both = 0;
                if (hit_hi && hit_lo) { both = 1; l_hi_p = p_hi; l_lo_p = p_lo; }
                else { if (!hit_hi) l_hi_p = p_hi; if (!hit_lo) l_lo_p = p_lo; }
                s_early.both_hit = both; s_early.line_hi = l_hi_p; s_early.line_lo = l_lo_p; s_early.hits = hit_hi + hit_lo;
                if (cta == 0 && !both) {
                    if (!hit_hi) a.cache_stamp[i_hi] = num_iter + 1;
                    if (!hit_lo) a.cache_stamp[i_lo] = num_iter + 1;
                    __threadfence();        // ordered before this thread's slot store, which is what the other CTAs wait for
                }
            }
        } else if (a.cache_lines > 0) {
            if (warp == 1) {                // a whole warp per lookup, off thread 0's path (it is issuing the row copies)
                LruTable t{s_lru_key, s_lru_stamp, s_lru_meta};
                const int l_hi = lru_lookup_warp(t, a.cache_lines, i_hi, &hit_hi);      // hi then lo (svmTrain.cu:203,238)
                const int l_lo = lru_lookup_warp(t, a.cache_lines, i_lo, &hit_lo);
                if (lane == 0) {
                    float* p_hi = a.cache_data + (size_t)l_hi * a.line_stride;
                    float* p_lo = a.cache_data + (size_t)l_lo * a.line_stride;
                    float* l_hi_p = nullptr;
                    float* l_lo_p = nullptr;
                    int both = 0;
                    if (hit_hi && hit_lo) { both = 1; l_hi_p = p_hi; l_lo_p = p_lo; }
                    else { if (!hit_hi) l_hi_p = p_hi; if (!hit_lo && l_lo != l_hi) l_lo_p = p_lo; }
                    s_early.both_hit = both; s_early.line_hi = l_hi_p; s_early.line_lo = l_lo_p; s_early.hits = hit_hi + hit_lo;
                }
            }
        } else if (tid == 0) {
            s_early.both_hit = 0; s_early.line_hi = nullptr; s_early.line_lo = nullptr; s_early.hits = 0;
        }
        if (tid == 0) { s_early.xsq_hi = xsq_hi; s_early.xsq_lo = xsq_lo; }
        if (a.world > 1 && !ll_mode) __syncthreads();
        else mbar_wait(&s_bar, (uint32_t)(it & 1));
        stamp(it, 1);
        // |x_hi - x_lo|^2 exactly as the control kernel sums it: 256 strided partials, warp butterfly,
        // then the eight warp sums in order (svmTrain.cu:696-714 is the host original)
        if (tid < kCtlThreads) {
            const float* fh = reinterpret_cast<const float*>(s_hi);
            const float* fl = reinterpret_cast<const float*>(s_lo);
            float part = 0.f;
            for (int i = tid; i < a.pitch4 * 4; i += kCtlThreads) {
                const float vh = fh[i], vl = fl[i];
                const float df = __fsub_rn(vh, vl);
                part = __fmaf_rn(df, df, part);
            }
#pragma unroll
            for (int off = 16; off >= 1; off >>= 1) part = __fadd_rn(part, __shfl_xor_sync(0xffffffffu, part, off));
            if (lane == 0) s_part[warp] = part;
        }
        if (SPARSE && !(a.cache_direct && hit_hi && hit_lo)) {
            // the pair, interleaved and replicated: 8 x 16 bytes per column, by all threads (not when both kernel rows are
            // cached: that iteration has no sweep; with a line per row every thread knows the two hits for itself)
            const float* fh = reinterpret_cast<const float*>(s_hi);
            const float* fl = reinterpret_cast<const float*>(s_lo);
            for (int i = tid; i < a.pitch4 * 4 * 8; i += kResThreads) {
                const int col = i >> 3;
                const float vh = fh[col], vl = fl[col];
                s_rep[i] = make_float4(vh, vl, vh, vl);
            }
        }
        __syncthreads();
        stamp(it, 2);
        const int both_hit = s_early.both_hit;
        // From here every warp may sweep: the dot products need the pair only.  What the per-row epilogue needs on top --
        // the two alpha deltas, i.e. eta -- is computed by thread 0 meanwhile and handed over through named barrier 1:
        // warp 0 arrives once the values are in shared memory, every other warp joins right before its first epilogue.
        bool have_deltas = false;
        if (warp == 0) {
            if (tid == 0) {
                if (a.cache_lines > 0) { cache_lookups += 2; cache_hits += s_early.hits; }
                float norm_sq = 0.f;
#pragma unroll
                for (int w = 0; w < kCtlThreads / 32; ++w) norm_sq = __fadd_rn(norm_sq, s_part[w]);
                const float k_lh = (float)exp((double)__fmul_rn(-a.gamma, norm_sq));                 // svmTrain.cu:709
                const float eta = __fsub_rn(__fadd_rn(1.0f, 1.0f), __fmul_rn(2.0f, k_lh));          // svmTrainMain.cpp:282
                const float s = (float)(y_lo * y_hi);                                                // :289-295
                const float a_lo_unc = __fadd_rn(a_lo_old, __fdiv_rn(__fmul_rn((float)y_lo, __fsub_rn(b_hi, b_lo)), eta));
                const float a_hi_unc = __fadd_rn(a_hi_old, __fmul_rn(s, __fsub_rn(a_lo_old, a_lo_unc)));
                const float lo_new = a_lo_unc < 0.0f ? 0.0f : (a_lo_unc > a.c ? a.c : a_lo_unc);
                const float hi_new = a_hi_unc < 0.0f ? 0.0f : (a_hi_unc > a.c ? a.c : a_hi_unc);
                // the CTA that owns a row stores its alpha (only that CTA ever reads it again)
                const int row_lo0 = u0 * RPW, row_hi0 = min(u1 * RPW, a.n);
                const int row_res = row_lo0 + res_units * RPW;       // rows below this are the on-chip slice
                if (r_lo >= row_lo0 && r_lo < row_hi0) { a.alpha[r_lo] = lo_new; if (r_lo < row_res) s_ralpha[r_lo - row_lo0] = lo_new; }   // svmTrainMain.cpp:298-299
                if (r_hi >= row_lo0 && r_hi < row_hi0) { a.alpha[r_hi] = hi_new; if (r_hi < row_res) s_ralpha[r_hi - row_lo0] = hi_new; }
                s_delta[0] = __fmul_rn(__fsub_rn(hi_new, a_hi_old), (float)y_hi);
                s_delta[1] = __fmul_rn(__fsub_rn(lo_new, a_lo_old), (float)y_lo);
                s_delta[2] = hi_new; s_delta[3] = lo_new;
                if (cta == 0) {
                    st->i_hi = i_hi; st->i_lo = i_lo; st->eta = eta; st->b_hi = b_hi; st->b_lo = b_lo;
                    st->num_iter = num_iter + 1;
                    st->cache_hits = cache_hits; st->cache_lookups = cache_lookups;
                }
            }
            __syncwarp();
            // (no fence: when a named barrier completes, what its arriving threads wrote before is visible to the threads it releases)
            asm volatile("bar.arrive 1, %0;" ::"n"(kResThreads) : "memory");
            have_deltas = true;
        }
        auto need_deltas = [&]() {          // warp-uniform; every warp runs it exactly once per iteration
            if (!have_deltas) {
                asm volatile("bar.sync 1, %0;" ::"n"(kResThreads) : "memory");
                have_deltas = true;
            }
        };
        stamp(it, 3);

        // ================= sweep over this CTA's rows =================
        Cand best_up{DPSVM_KEY_EMPTY, 0.f}, best_low{DPSVM_KEY_EMPTY, 0.f};
        const bool rev = ((it + a.parity0) & 1) != 0;

        const int row_res_end = (u0 + res_units) * RPW;      // rows below this (and >= u0*RPW) are on chip
        auto epilogue = [&](int row, float dh, float dl, float xs, float fj, float al, int yj, bool store) {
            if (store) {
                float* const l_hi_p = s_early.line_hi;
                float* const l_lo_p = s_early.line_lo;
                if (l_hi_p) l_hi_p[row] = dh;
                if (l_lo_p) l_lo_p[row] = dl;
            }
            // the row's alpha may have been read before thread 0 installed this iteration's two new values
            if (row == r_lo) al = s_delta[3];                                                    // svmTrainMain.cpp:298-299,
            if (row == r_hi) al = s_delta[2];                                                    // in that order
            const float arg_hi = __fsub_rn(__fadd_rn(xs, s_early.xsq_hi), __fmul_rn(2.0f, dh));  // svmTrain.cu:128,130
            const float arg_lo = __fsub_rn(__fadd_rn(xs, s_early.xsq_lo), __fmul_rn(2.0f, dl));
            const float rbf_hi = expf(__fmul_rn(-a.gamma, arg_hi));
            const float rbf_lo = expf(__fmul_rn(-a.gamma, arg_lo));
            const float delta = __fadd_rn(__fmul_rn(s_delta[0], rbf_hi), __fmul_rn(s_delta[1], rbf_lo));   // :133-135
            const float f_new = __fadd_rn(fj, delta);
            a.f[row] = f_new;
            if (row < row_res_end) s_rf[row - u0 * RPW] = f_new;
            bool up, low;
            iset_membership(al, yj, a.c, up, low);
            const int gidx = a.index_base + row;
            if (up) cand_min(best_up, pack_key_up(f_new, gidx), al);
            if (low) cand_min(best_low, pack_key_low(f_new, gidx), al);
        };

        // do-while (svmTrainMain.cpp:310): this iteration's update is applied in any case; is it the last?
        const bool keep_going = b_lo > __fadd_rn(b_hi, __fmul_rn(2.0f, a.epsilon));
        const bool last_iter = !keep_going || (num_iter + 1) >= a.max_iter || (num_iter + 1) >= stop_at || (it + 1) >= a.chunk;

        // ---- sparse sweep: request side (SparseDev::perm: the units of this warp slot, first steps in shared memory) ----
        constexpr unsigned int kTagLast = 1u << 31, kTagRow = kTagLast - 1u;     // tag of a unit's last chunk: its first row | kTagLast; other chunks: 0
        auto sp_open = [&]() {                       // cursor -> unit sp_pj of the list, in the cursor's direction
            const int lu = (int)s_perm[(sp_prev ? (sp_cnt - 1 - sp_pj) : sp_pj) * NW + wslot];
            const unsigned int beg = s_uoff[lu];
            sp_left = (int)(s_uoff[lu + 1] - beg) / kSpChunk;
            sp_po = beg * 32u + (unsigned int)lane * 8u;                 // chunk beg / 8, 256 bytes of column distances per chunk
            if (!KEEP_ALL) sp_pkeep = a.stream_evict_first ? ((beg - s_uoff[0]) < a.sp.pin_steps ? 1 : 0) : -1;
            sp_utag = (unsigned int)((u0 + lu) * GROUPS) | kTagLast;
        };
        auto sp_start = [&](bool direction) {        // a new sequence: nothing requested yet
            sp_prev = direction;
            sp_cnt = (int)s_cnt[wslot];
            sp_pj = 0;
            if (sp_cnt > 0) sp_open();
        };
        auto sp_request = [&](int b, bool sure = false) {    // next chunk of the sequence -> ring slot b (no-op once the list is exhausted;
            if (sure || sp_pj < sp_cnt) {                    //  sure: the caller knows that it is not)
                const bool last = --sp_left == 0;
                sp_tag[b] = last ? sp_utag : 0u;
                sp_load_at<C8, KEEP_ALL>(sp_ring[b], a.sp, sp_po, sp_pkeep, pol_keep, pol_stream);
                sp_po += 256u;
                if (last) { if (++sp_pj < sp_cnt) sp_open(); }
            }
        };
        auto sp_prime = [&](bool direction) {        // first kSpRing chunks of a sequence in flight
            sp_start(direction);
#pragma unroll
            for (int b = 0; b < kSpRing; ++b) sp_request(b);
        };
        if (both_hit) {
            // both kernel rows cached: no X, 28 B per row
            need_deltas();
            pre = false;
            const int row_b = u0 * RPW, row_e = min(u1 * RPW, a.n);
            for (int row = row_b + tid; row < row_e; row += kResThreads)
                epilogue(row, s_early.line_hi[row], s_early.line_lo[row], __ldg(a.xsq + row), a.f[row], a.alpha[row], __ldg(a.y + row), false);
        } else if constexpr (SPARSE) {
            if (!pre) sp_prime(rev);
            pre = false;
            // Per-row epilogues are batched: a finished unit leaves one (dot_hi, dot_lo) pair per row in the G lanes of that
            // row's group, i.e. GROUPS = 32 / G results in a warp.  Running the epilogue there would keep GROUPS of 32
            // lanes busy for ~200 instructions per unit (measured: a fifth of the sweep's instruction stream).  Instead
            // lane L parks the result of group L % GROUPS of the (L / GROUPS)-th unit since the last flush -- and requests
            // that row's scalars -- and after G units every lane holds one row: ONE epilogue pass with all lanes active.
            // Same values, same stores; the candidate minima are order-independent (distinct keys).
            int n_parked = 0;
            int p_row = -1;
            float p_dh = 0.f, p_dl = 0.f;
            auto flush = [&]() {
                need_deltas();
                if (p_row >= 0) epilogue(p_row, p_dh, p_dl, e_xsq, e_f, e_alpha, e_y, true);
                p_row = -1;
                n_parked = 0;
            };
            {
                unsigned int pos0 = C8 ? smem_u32(s_hl) : 0u;
                asm volatile("" : "+r"(pos0));   // held in a register: rebuilt from the CTA's shared window it costs a special-register read per chunk
                unsigned int pos = pos0;         // 8-bit format: shared-memory address of this lane's current column in s_hl
                float acc_hi = 0.f, acc_lo = 0.f;
                auto consume = [&](auto slot, auto req) {
                    constexpr int b = decltype(slot)::value;
                    constexpr int REQ = decltype(req)::value;        // 2: a chunk kSpRing ahead exists; 1: it may; 0: it does not
                    const unsigned int tag = sp_tag[b];
                    sp_fma<C8, 4>(sp_ring[b], s_hl, acc_hi, acc_lo, pos);
                    if constexpr (REQ > 0) sp_request(b, REQ == 2);  // the consumed slot takes the chunk kSpRing ahead at once
                    if (tag & kTagLast) {
                        const int unit_row0 = (int)(tag & kTagRow);
                        sp_group_sum<SGL>(acc_hi, acc_lo);
                        // after the xor-butterfly every lane of a group holds the group's two sums: read group L % GROUPS
                        const int src = (lane % GROUPS) * SGL;
                        const float vh = __shfl_sync(0xffffffffu, acc_hi, src), vl = __shfl_sync(0xffffffffu, acc_lo, src);
                        if (lane / GROUPS == n_parked) {
                            const int row = unit_row0 + lane % GROUPS;
                            if (row < a.n) {
                                p_row = row; p_dh = vh; p_dl = vl;
                                e_xsq = __ldg(a.xsq + row); e_y = __ldg(a.y + row);
                                e_f = a.f[row]; e_alpha = a.alpha[row];
                            }
                        }
                        if (++n_parked == SGL) flush();
                        acc_hi = 0.f; acc_lo = 0.f;
                        pos = pos0;              // the next chunk starts a unit: column distances count from column 0 again
                    }
                };
                // Whole rounds of the ring without per-chunk end tests -- while at least two rounds are left every request
                // is for a chunk that exists -- then one round whose requests may run off the list, then the up to three
                // chunks left, which request nothing.
                static_assert(kSpRing == 4, "the sweep is unrolled over the four ring slots");
                const std::integral_constant<int, 0> b0, none; const std::integral_constant<int, 1> b1, maybe;
                const std::integral_constant<int, 2> b2, sure; const std::integral_constant<int, 3> b3;
                int left = sp_total;
                for (; left >= 2 * kSpRing; left -= kSpRing) { consume(b0, sure); consume(b1, sure); consume(b2, sure); consume(b3, sure); }
                if (left >= kSpRing) { consume(b0, maybe); consume(b1, maybe); consume(b2, maybe); consume(b3, maybe); left -= kSpRing; }
                if (left > 0) {
                    consume(b0, none);
                    if (left > 1) { consume(b1, none); if (left > 2) consume(b2, none); }
                }
            }
            flush();
        } else {
            if constexpr (RPW <= 4) {
            // Per-row epilogues are batched as in the sparse sweep: a finished unit leaves RPW = (32 / LPR) * R results in a
            // warp (two for rows of 784 floats); lane L parks result L % RPW of the (L / RPW)-th unit since the last flush and
            // requests that row's scalars; after 32 / RPW units every lane holds one row and the epilogue runs once, all lanes busy.
            constexpr int BATCH = 32 / RPW;
            int n_parked = 0, p_row = -1;
            float p_dh = 0.f, p_dl = 0.f;
            auto flush = [&]() {
                need_deltas();
                if (p_row >= 0) epilogue(p_row, p_dh, p_dl, e_xsq, e_f, e_alpha, e_y, true);
                p_row = -1;
                n_parked = 0;
            };
            for (int k = wslot; k < my_units; k += NW) {
                if (!pre) load_const(k, rev);
                pre = false;
                const int this_first_row = unit_first_row;
                const bool this_in_smem = unit_in_smem;
                float acc_hi[R], acc_lo[R];
#pragma unroll
                for (int r = 0; r < R; ++r) { acc_hi[r] = 0.f; acc_lo[r] = 0.f; }
#pragma unroll
                for (int itc = 0; itc < ITERS; ++itc) {
                    const float4 h = s_hi[sub + itc * LPR];
                    const float4 l = s_lo[sub + itc * LPR];
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        const float4 v = xv[r][itc];
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
                // every lane of a group now holds the group's R pairs of sums: result q of the unit is row (q / R)'s group, its r = q % R
                {
                    const int q = lane % RPW, src = (q / R) * LPR;
                    float vh = 0.f, vl = 0.f;
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        const float th = __shfl_sync(0xffffffffu, acc_hi[r], src), tl = __shfl_sync(0xffffffffu, acc_lo[r], src);
                        if (q % R == r) { vh = th; vl = tl; }
                    }
                    if (lane / RPW == n_parked) {
                        const int row = this_first_row + q;
                        if (row < a.n) {
                            p_row = row; p_dh = vh; p_dl = vl;
                            if (this_in_smem) {      // state written by the previous iteration / this prologue
                                const int li = row - u0 * RPW;
                                e_xsq = s_rxsq[li]; e_y = s_ry[li]; e_f = s_rf[li]; e_alpha = s_ralpha[li];
                            } else {
                                e_xsq = __ldg(a.xsq + row); e_y = __ldg(a.y + row);
                                e_f = a.f[row]; e_alpha = a.alpha[row];
                            }
                        }
                    }
                    if (++n_parked == BATCH) flush();
                }
            }
            flush();
            } else {
                // narrow rows (8 or 16 rows per unit, as many lanes busy in the epilogue): epilogue per unit, its row scalars
                // requested together with the unit's X (measured at 4 000 000 x 128, HBM-bound: batching costs 5 % there)
            for (int k = wslot; k < my_units; k += NW) {
                if (!pre) load_const(k, rev);
                pre = false;
                if (eown) {                      // state written by the previous iteration / this prologue
                    if (unit_in_smem) { e_f = s_rf[erow - u0 * RPW]; e_alpha = s_ralpha[erow - u0 * RPW]; }
                    else { e_f = a.f[erow]; e_alpha = a.alpha[erow]; }
                }
                float acc_hi[R], acc_lo[R];
#pragma unroll
                for (int r = 0; r < R; ++r) { acc_hi[r] = 0.f; acc_lo[r] = 0.f; }
#pragma unroll
                for (int itc = 0; itc < ITERS; ++itc) {
                    const float4 h = s_hi[sub + itc * LPR];
                    const float4 l = s_lo[sub + itc * LPR];
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        const float4 v = xv[r][itc];
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
                need_deltas();
                if (eown) {
                    float dh = acc_hi[0], dl = acc_lo[0];
#pragma unroll
                    for (int r = 1; r < R; ++r) if (sub == r) { dh = acc_hi[r]; dl = acc_lo[r]; }
                    epilogue(erow, dh, dl, e_xsq, e_f, e_alpha, e_y, true);
                }
            }
            }
        }
        need_deltas();          // a warp without rows this iteration still takes part in the hand-over
        // ================= block best -> slot; all slots -> next pair =================
        cand_warp_min(best_up);
        cand_warp_min(best_low);
        if (lane == 0) { s_cand[0][warp] = best_up; s_cand[1][warp] = best_low; }
        if (tid == 0) s_flag = 1;
        __syncthreads();
        stamp(it, 4);
        num_iter += 1;
        const unsigned int seq = (unsigned int)num_iter;            // >= 1: a cleared slot never looks valid
        const unsigned long long tag = (unsigned long long)seq << 32;
        ResSlot* const slot_set = a.slots + (size_t)(seq & 1u) * G;
        if (warp == 0) {
            Cand u = lane < NW ? s_cand[0][lane] : Cand{DPSVM_KEY_EMPTY, 0.f};
            Cand l = lane < NW ? s_cand[1][lane] : Cand{DPSVM_KEY_EMPTY, 0.f};
            cand_warp_min(u);
            cand_warp_min(l);
            // three 16-byte stores of two flagged words each; nothing to order, nothing to count
            if (lane < 3) {
                unsigned int d0, d1;
                if (lane == 0) { d0 = (unsigned int)(u.key >> 32); d1 = (unsigned int)u.key; }
                else if (lane == 1) { d0 = (unsigned int)(l.key >> 32); d1 = (unsigned int)l.key; }
                else { d0 = __float_as_uint(u.alpha); d1 = __float_as_uint(l.alpha); }
                st_relaxed_gpu_v2u64(&slot_set[cta].w[2 * lane], tag | d0, tag | d1);
            }
        }
        // The next sweep's first loads do not depend on the pair: they go in flight now -- behind this CTA's slot, so
        // that the one message every other SM is waiting for is not queued after them -- and land during the merge.
        auto prefetch_next = [&]() {
            if constexpr (SPARSE) {
                if (!last_iter) {
                    sp_prime(!rev);
                    pre = true;
                }
            } else {
                if (!last_iter && wslot < my_units) { load_const(wslot, !rev); pre = true; }
            }
        };
        if (a.prefetch_mode == 0) prefetch_next();
        stamp(it, 9);

        // every CTA reads every slot until all of them carry this iteration's number, and merges them: minimum packed key =
        // lowest global index among equal f
        Cand nxt_up, nxt_low;
        {
            Cand u{DPSVM_KEY_EMPTY, 0.f}, l{DPSVM_KEY_EMPTY, 0.f};
            for (int c = tid; c < G; c += kResThreads) {
                const unsigned long long* w = slot_set[c].w;
                unsigned long long v0, v1, v2, v3, v4, v5;
                unsigned long long t0 = 0;
                unsigned int spins = 0;
                for (;;) {
                    ld_relaxed_gpu_v2u64(w, v0, v1);
                    ld_relaxed_gpu_v2u64(w + 2, v2, v3);
                    ld_relaxed_gpu_v2u64(w + 4, v4, v5);
                    const unsigned long long bad = ((v0 ^ tag) | (v1 ^ tag) | (v2 ^ tag) | (v3 ^ tag) | (v4 ^ tag) | (v5 ^ tag)) >> 32;
                    if (bad == 0) break;
                    if ((++spins & 0xFFu) == 0) {
                        if (t0 == 0) t0 = global_timer_ns();
                        else if (global_timer_ns() - t0 > a.timeout_ns) { s_flag = 0; break; }
                    }
                }
                if (a.trace && it == a.trace_iter && tid == 0) a.trace[(size_t)cta * 16 + 10] = spins;
                cand_min(u, (v0 << 32) | (unsigned int)v1, __uint_as_float((unsigned int)v4));
                cand_min(l, (v2 << 32) | (unsigned int)v3, __uint_as_float((unsigned int)v5));
            }
            if (tid == 0) stamp(it, 5);
            if (a.prefetch_mode == 1) prefetch_next();
            cand_warp_min(u);
            cand_warp_min(l);
            if (lane == 0) { s_cand2[0][warp] = u; s_cand2[1][warp] = l; }
            __syncthreads();
            if (!s_flag) { if (cta == 0 && tid == 0) { st->error = 1; st->done = 1; } return; }
            // every warp finishes the reduction for itself: no second barrier, no broadcast through shared memory
            nxt_up = lane < NW ? s_cand2[0][lane] : Cand{DPSVM_KEY_EMPTY, 0.f};
            nxt_low = lane < NW ? s_cand2[1][lane] : Cand{DPSVM_KEY_EMPTY, 0.f};
            cand_warp_min(nxt_up);
            cand_warp_min(nxt_low);
        }
        // multi-GPU: CTA 0 publishes this rank's candidates to every rank (also on the last iteration of the
        // launch: the next launch starts by waiting for exactly these packages)
        stamp(it, 6);
        if (a.world > 1 && cta == 0) {
            if (ll_mode) {
                // one 8-byte flagged word per thread and peer, straight from registers: nothing to fence, nothing to read
                if (tid < a.world * kLLWords && (tid & (kLLWords - 1)) < 6) {
                    // an empty set ships the alpha of the first owned row (on rank 0 that is global row 0, the default)
                    const float au = nxt_up.key == DPSVM_KEY_EMPTY ? __ldcg(a.alpha) : nxt_up.alpha;
                    const float al = nxt_low.key == DPSVM_KEY_EMPTY ? __ldcg(a.alpha) : nxt_low.alpha;
                    const int w = tid & (kLLWords - 1);
                    ll_send(a.ll, tid / kLLWords, a.rank, (unsigned int)num_iter, w, ll_word(w, nxt_up.key, au, nxt_low.key, al));
                }
            } else {
                push_package(*a.push, nxt_up.key, nxt_low.key, nxt_up.alpha, nxt_low.alpha, true, (unsigned int)num_iter);
            }
        }
        stamp(it, 7);
        if (last_iter) {
            if (cta == 0 && tid == 0) {
                a.keys[0] = nxt_up.key;
                a.keys[1] = nxt_low.key;
                if (a.world > 1) { a.push->keys_copy[0] = nxt_up.key; a.push->keys_copy[1] = nxt_low.key; }
                if (!keep_going) { st->converged = 1; st->done = 1; }
                if (num_iter >= a.max_iter) st->done = 1;
            }
            break;
        }
        if (a.world > 1) {
            if (!acquire((unsigned int)num_iter)) { if (cta == 0 && tid == 0) { st->error = 1; st->done = 1; } return; }
            stamp(it, 8);
        } else {
            unpack_key_up(nxt_up.key, &b_hi, &i_hi);
            unpack_key_low(nxt_low.key, &b_lo, &i_lo);
            a_hi_old = nxt_up.alpha;
            a_lo_old = nxt_low.alpha;
            // an empty set has no candidate row: default row and ITS alpha, exactly as at launch and in the graph loop
            // (the owner's store of this iteration is ordered before the grid barrier above)
            if (nxt_up.key == DPSVM_KEY_EMPTY) { i_hi = a.empty_row; a_hi_old = __ldcg(a.alpha + (a.empty_row - a.index_base)); }
            if (nxt_low.key == DPSVM_KEY_EMPTY) { i_lo = a.empty_row; a_lo_old = __ldcg(a.alpha + (a.empty_row - a.index_base)); }
        }
        // (s_cand is next written after a whole sweep, with several block barriers in between)
    }
    if (cta == 0 && a.cache_lines > 0 && !a.cache_direct) {
        __syncthreads();
        for (int i = tid; i < a.cache_lines; i += kResThreads) { a.lru_key[i] = s_lru_key[i]; a.lru_stamp[i] = s_lru_stamp[i]; }
        if (tid < 2) a.lru_meta[tid] = s_lru_meta[tid];
    }
}

// ------------------------------------------------------------------------------------------------
// launcher
// ------------------------------------------------------------------------------------------------
// Per kernel instantiation and per device: the dynamic shared-memory budget (function attributes belong to the
// device they were set on) and the grid.  Filled once under a lock; host threads may share the library.
struct ResCfg {
    int smem_max = -1;
    int grid = 0;
};
constexpr int kMaxDevices = 64;

static cudaError_t resident_config(const void* kernel, int min_budget, ResCfg (&cache)[kMaxDevices], std::mutex& mu, ResCfg* out)
{
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(mu);
    ResCfg& c = cache[dev];
    if (c.smem_max < 0) {
        int optin = 0, sms = 0;
        e = cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
        if (e != cudaSuccess) return e;
        e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) return e;
        cudaFuncAttributes fa;
        e = cudaFuncGetAttributes(&fa, kernel);
        if (e != cudaSuccess) return e;
        const int budget = optin - (int)fa.sharedSizeBytes - 1024;
        if (budget < min_budget) return cudaErrorInvalidValue;
        e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, budget);
        if (e != cudaSuccess) return e;
        int per_sm = 0;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kResThreads, budget);
        if (e != cudaSuccess) return e;
        if (per_sm < 1) return cudaErrorLaunchOutOfResources;
        if (sms < 1 || sms > kResThreads) return cudaErrorInvalidValue;      // thread c of a CTA reads CTA c's slot
        c.grid = sms;
        c.smem_max = budget;
    }
    *out = c;
    return cudaSuccess;
}

constexpr int kResSliceUnits = 16;     // on-chip units per CTA when the shard does not fit on chip entirely

template <int LPR, int ITERS>
static cudaError_t launch_resident_t(ResidentDev a, int variant, cudaStream_t s)
{
    constexpr int RPW = (32 / LPR) * kRowsInFlight;
    constexpr int CAP4 = LPR * ITERS;
    auto kernel = smo_resident_kernel<LPR, ITERS, 0>;
    static ResCfg s_cache[kMaxDevices];
    static std::mutex s_mu;
    ResCfg cfg;
    {
        const cudaError_t e = resident_config((const void*)kernel, 2 * CAP4 * 16, s_cache, s_mu, &cfg);
        if (e != cudaSuccess) return e;
    }
    const int s_smem_max = cfg.smem_max, s_grid = cfg.grid;
    const size_t unit_bytes = (size_t)RPW * a.pitch4 * 16;
    const size_t unit_smem = unit_bytes + (size_t)RPW * 16;          // rows + |x|^2, y, f, alpha
    const size_t avail = (size_t)s_smem_max - 2 * (size_t)CAP4 * 16;
    a.res_units = (int)(avail / unit_smem);
    {
        // Measured (profiles/r01_resident_trace.md): the shared-memory carve-out comes out of L1, and L1
        // capacity bounds how many streaming loads an SM keeps in flight -- with the full carve-out the
        // sweep of a streamed X slows from 24 to 33 us.  So: keep the whole shard on chip when it fits
        // (nothing streams then), otherwise only a small slice.
        const int n_units = (a.n + RPW - 1) / RPW;
        const int per_cta = (n_units + s_grid - 1) / s_grid;
        if (per_cta > a.res_units && a.res_units > kResSliceUnits) a.res_units = kResSliceUnits;
    }
    {
        static const char* cap_env = getenv("DPSVM_RES_UNITS");     // tuning aid: cap the shared-memory slice
        if (cap_env && *cap_env) { const int cap = atoi(cap_env); if (cap >= 0 && cap < a.res_units) a.res_units = cap; }
    }
    // L2 policy of the streamed part: same budget as the graph loop's pinned prefix, split over the CTAs
    const bool hint = (variant & 2) != 0;
    a.stream_evict_first = hint ? 1 : 0;
    a.pin_units = hint ? (int)(((double)((variant >> 8) & 0xFFFFF) * 1e6 / s_grid) / (double)unit_bytes) : 0;
    a.parity0 = variant & 1;
    size_t smem = 2 * (size_t)CAP4 * 16 + (size_t)a.res_units * unit_smem;
    {
        static const char* pad_env = getenv("DPSVM_RES_PAD");       // experiment: allocate the full carve-out regardless
        if (pad_env && atoi(pad_env)) smem = (size_t)s_smem_max;
    }
    void* args[] = {&a};
    return cudaLaunchCooperativeKernel((void*)kernel, dim3((unsigned)s_grid), dim3(kResThreads), args, smem, s);
}

template <int LPR>
static cudaError_t launch_resident_lpr(int iters, const ResidentDev& a, int variant, cudaStream_t s)
{
    switch (iters) {
        case 1: return launch_resident_t<LPR, 1>(a, variant, s);
        case 2: return launch_resident_t<LPR, 2>(a, variant, s);
        case 3: return launch_resident_t<LPR, 3>(a, variant, s);
        case 4: return launch_resident_t<LPR, 4>(a, variant, s);
        case 5: return launch_resident_t<LPR, 5>(a, variant, s);
        case 6: return launch_resident_t<LPR, 6>(a, variant, s);
        case 7: return launch_resident_t<LPR, 7>(a, variant, s);
        case 8: return launch_resident_t<LPR, 8>(a, variant, s);
    }
    return cudaErrorInvalidValue;
}

// Sparse sweep: the two selected-row buffers are sized for the widest supported row (LPR 32 x ITERS 8).
template <int SG, bool C8, bool KEEP_ALL>
static cudaError_t launch_resident_sparse_k(ResidentDev a, int variant, cudaStream_t s)
{
    constexpr int CAP4 = 256;
    static_assert(CAP4 * 4 == kSparseMaxPitch, "selected-row buffers cover the widest sparse row");
    auto kernel = smo_resident_kernel<32, 8, SG, C8, KEEP_ALL>;
    static ResCfg s_cache[kMaxDevices];
    static std::mutex s_mu;
    ResCfg cfg;
    {
        const cudaError_t e = resident_config((const void*)kernel, 64 * 1024, s_cache, s_mu, &cfg);
        if (e != cudaSuccess) return e;
    }
    const int s_smem_max = cfg.smem_max, s_grid = cfg.grid;
    if (a.sp.grid != s_grid) return cudaErrorInvalidValue;      // cta_first was balanced for another grid
    static_assert(kResThreads / 32 == 16, "SparseDev::perm deals the units of a CTA to 16 warp slots");
    const size_t fixed = 2 * (size_t)CAP4 * 16 + (size_t)a.pitch4 * 4 * 128 + (((size_t)a.sp.max_cta_units + 1 + 3) & ~(size_t)3) * sizeof(unsigned int) +
                         (((size_t)a.sp.perm_stride + 7) & ~(size_t)7) * sizeof(unsigned short) + 16 * sizeof(unsigned short);
    if (fixed > (size_t)s_smem_max) return cudaErrorInvalidValue;
    const size_t smem = fixed;
    a.res_units = 0;
    a.pin_units = 0;
    const bool hint = (variant & 2) != 0;
    a.stream_evict_first = hint ? 1 : 0;
    a.sp.pin_steps = hint ? (unsigned int)(((double)((variant >> 8) & 0xFFFFF) * 1e6 / s_grid) / (double)a.sp.step_bytes) : 0u;
    a.parity0 = variant & 1;
    void* args[] = {&a};
    return cudaLaunchCooperativeKernel((void*)kernel, dim3((unsigned)s_grid), dim3(kResThreads), args, smem, s);
}

// A stream that fits in L2 is kept there as a whole (variant bit 1 clear): the sweep then needs neither the second
// eviction policy nor the per-unit choice between the two -- KEEP_ALL, built for the lane counts the layout picks
// by itself (4 and 8; 16 and 32 exist for DPSVM_SPARSE_LANES only).
template <int SG, bool C8>
static cudaError_t launch_resident_sparse_t(const ResidentDev& a, int variant, cudaStream_t s)
{
    if constexpr (SG <= 8) {
        if (!(variant & 2)) return launch_resident_sparse_k<SG, C8, true>(a, variant, s);
    }
    return launch_resident_sparse_k<SG, C8, false>(a, variant, s);
}

cudaError_t launch_resident_sparse(const ResidentDev& a, int variant, cudaStream_t s)
{
    if (a.sp.col8) {
        switch (a.sp.lanes) {
            case 4: return launch_resident_sparse_t<4, true>(a, variant, s);
            case 8: return launch_resident_sparse_t<8, true>(a, variant, s);
            case 16: return launch_resident_sparse_t<16, true>(a, variant, s);
            case 32: return launch_resident_sparse_t<32, true>(a, variant, s);
        }
        return cudaErrorInvalidValue;
    }
    switch (a.sp.lanes) {
        case 4: return launch_resident_sparse_t<4, false>(a, variant, s);
        case 8: return launch_resident_sparse_t<8, false>(a, variant, s);
        case 16: return launch_resident_sparse_t<16, false>(a, variant, s);
        case 32: return launch_resident_sparse_t<32, false>(a, variant, s);
    }
    return cudaErrorInvalidValue;
}

int resident_grid() { return device_sm_count(); }

// Kernel-row caches the resident loop can run: none, a small one (replicated LRU table), or one with a line for
// every row of the set (nothing is ever evicted: direct index).  In between, eviction needs one writer: graph loop.
bool resident_supported(int pitch4, int cache_lines, int n_total)
{
    return pitch4 <= 256 && (cache_lines <= kResCacheLines || cache_lines >= n_total);
}

cudaError_t launch_resident(const ResidentDev& a, int variant, cudaStream_t s)
{
    if (a.n <= 0) return cudaSuccess;
    if (a.pitch4 > 256 || (a.cache_lines > kResCacheLines && !a.cache_direct)) return cudaErrorInvalidValue;
    if (a.sp.val) return launch_resident_sparse(a, variant, s);
    const DotGeom g = dot_geometry(a.pitch4);
    switch (g.lpr) {
        case 4: return launch_resident_lpr<4>(g.iters, a, variant, s);
        case 8: return launch_resident_lpr<8>(g.iters, a, variant, s);
        case 16: return launch_resident_lpr<16>(g.iters, a, variant, s);
        default: return launch_resident_lpr<32>(g.iters, a, variant, s);
    }
}

}  // namespace dpsvm
