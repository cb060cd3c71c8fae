// smo_common.cuh -- shared host/device definitions for the SMO hot path (sm_100a).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

#include "../../include/dpsvm_cuda.h"

namespace dpsvm {

// ------------------------------------------------------------------------------------------------
// Packed selection keys.  Replaces i_helper (reference svmTrain.h:24-37) + my_maxmin
// (svmTrain.cu:400-467): one unsigned 64-bit MIN gives argmin/argmax with lowest-index ties.
// ------------------------------------------------------------------------------------------------
__host__ __device__ __forceinline__ uint32_t float_bits(float f)
{
#ifdef __CUDA_ARCH__
    return __float_as_uint(f);
#else
    uint32_t u; memcpy(&u, &f, 4); return u;
#endif
}
__host__ __device__ __forceinline__ float bits_float(uint32_t u)
{
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    float f; memcpy(&f, &u, 4); return f;
#endif
}
// monotone float -> uint32 (a < b  <=>  ord(a) < ord(b)); -0 folded onto +0 first.
__host__ __device__ __forceinline__ uint32_t float_ord(float f)
{
    if (f == 0.0f) f = 0.0f;
    uint32_t u = float_bits(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ __forceinline__ float ord_float(uint32_t o)
{
    return bits_float((o & 0x80000000u) ? (o & 0x7FFFFFFFu) : ~o);
}
__host__ __device__ __forceinline__ uint64_t pack_key_up(float f, int32_t idx)
{
    return ((uint64_t)float_ord(f) << 32) | (uint32_t)idx;
}
__host__ __device__ __forceinline__ uint64_t pack_key_low(float f, int32_t idx)
{
    return ((uint64_t)float_ord(-f) << 32) | (uint32_t)idx;
}
// Empty key -> the reference's defaults (svmTrainMain.cpp:249-252): index 0, +-1e9.
__host__ __device__ __forceinline__ void unpack_key_up(uint64_t k, float* f, int32_t* idx)
{
    if (k == DPSVM_KEY_EMPTY) { *f = DPSVM_F_SENTINEL; *idx = 0; return; }
    *f = ord_float((uint32_t)(k >> 32)); *idx = (int32_t)(uint32_t)k;
}
__host__ __device__ __forceinline__ void unpack_key_low(uint64_t k, float* f, int32_t* idx)
{
    if (k == DPSVM_KEY_EMPTY) { *f = -DPSVM_F_SENTINEL; *idx = 0; return; }
    float nf = ord_float((uint32_t)(k >> 32));
    *f = (nf == 0.0f) ? 0.0f : -nf; *idx = (int32_t)(uint32_t)k;
}

// I-set membership with the reference's exact tests (svmTrain.cu:56-91 == seq.cpp:477-491).
__host__ __device__ __forceinline__ void iset_membership(float alpha, int y, float c, bool& up, bool& low)
{
    if (alpha == 0.0f)   { up = (y == 1);  low = !up; }
    else if (alpha == c) { up = (y == -1); low = !up; }
    else                 { up = true; low = true; }
}

// ------------------------------------------------------------------------------------------------
// Dot-product lane geometry.  A row of pitch4 = pitch/4 float4 chunks is split over LPR lanes
// (LPR in {4,8,16,32}); lane `sub` takes chunks sub, sub+LPR, ... (ITERS of them), accumulates
// with sequential FMAs (x,y,z,w order), then an xor-butterfly over the LPR lanes sums the partials.
// dpsvm_row_norms and every pass variant use this same order, which makes
// <x_i,x_i> == |x_i|^2 bit-exact (so K(i,i) == 1 exactly, as on the reference's host path).
// ------------------------------------------------------------------------------------------------
struct DotGeom { int lpr; int iters; };

__host__ __device__ inline DotGeom dot_geometry(int pitch4)
{
    DotGeom g;
    if (pitch4 <= 32) g.lpr = 4;
    else if (pitch4 <= 64) g.lpr = 8;
    else if (pitch4 <= 128) g.lpr = 16;
    else g.lpr = 32;
    g.iters = (pitch4 + g.lpr - 1) / g.lpr;
    return g;
}

// Scalars of one fused pass; written by the control kernel (device-resident loop) or passed by value.
struct PassScalars {
    int active;            // 0: the pass is a no-op (converged / paused)
    int both_hit;          // 1: both dot-product lines come from the cache, X is not read
    float delta_hi;        // (alpha_hi_new - alpha_hi_old) * y_hi      (svmTrain.cu:133)
    float delta_lo;
    float xsq_hi, xsq_lo;
    float* line_hi;        // cache line to write (miss) or read (both_hit); may be null
    float* line_lo;
};

// Device-resident trainer state (one per context).
struct DevState {
    int num_iter;
    int done;              // terminal: converged or max_iter reached
    int converged;
    int stop_at;           // pause when num_iter reaches this (host-controlled chunking)
    float b_hi, b_lo;
    int i_hi, i_lo;
    float eta;
    int error;             // 1: the peer exchange timed out (multi-GPU); the loop is stopped
    long long cache_hits;
    long long cache_lookups;
    PassScalars pass;
};

}  // namespace dpsvm
