// smo_kernels.cuh -- kernel argument blocks and launcher declarations (internal).
#pragma once

#include "smo_common.cuh"

namespace dpsvm {

constexpr int kPassThreads = 256;
constexpr int kRowsInFlight = 2;   // rows per lane group whose loads are issued before any FMA
constexpr int kCtlThreads = 256;

// Arguments of the fused pass kernel (by value; small enough for the parameter space).
struct PassDev {
    const float4* x4;       // [n x pitch4] owned rows
    int n, pitch4;
    const float* xsq;       // [n]
    const int* y;           // [n]
    const float* alpha;     // [n] (already holds this iteration's new values)
    float* f;               // [n]
    const float4* x_hi4;    // [pitch4]
    const float4* x_lo4;
    float gamma, c;
    int index_base;
    unsigned long long* keys;   // [2]
    PassScalars imm;            // used when dev == nullptr
    const PassScalars* dev;     // device-resident scalars written by the control kernel
};

// Kernel-row cache (device-resident LRU).  line_of_key is a direct index over the GLOBAL row
// number (a perfect hash; 4 N bytes), prev/next a doubly linked recency list over the lines.
enum { kCacheSize = 0, kCacheHead = 1, kCacheTail = 2, kCacheMetaInts = 4 };
struct CacheDev {
    int lines;              // capacity; 0 = cache disabled
    float* data;            // [lines x line_stride]
    size_t line_stride;     // floats per line (>= owned rows)
    int* line_of_key;       // [N global] -> line or -1
    int* key_of_line;       // [lines]
    int* prev;              // [lines] towards LRU
    int* next;              // [lines] towards MRU
    int* meta;              // [kCacheMetaInts]: size, head (LRU), tail (MRU)
};

struct CtlDev {
    DevState* st;
    unsigned long long* keys;   // [2] local selection result of the previous pass
    const float* x;             // resident rows, global row res0 first
    int pitch;
    int res0;                   // global index of resident row 0 (x, xsq, y, alpha all use it)
    const float* xsq;
    const int* y;
    float* alpha;
    float* sel_rows;            // [2 x pitch] staging for the pass
    float c, gamma, epsilon;
    int max_iter;
    CacheDev cache;
    int directed;               // host-directed update (SvmTrain::train_step2)
    int dir_i_hi, dir_i_lo;
    float dir_a_hi, dir_a_lo;
    const float* ext_row_hi;    // optional externally supplied rows (device pointers)
    const float* ext_row_lo;
};

int device_sm_count();
cudaError_t launch_pass(const PassDev& a, int variant, cudaStream_t s);
cudaError_t launch_row_norms(const float* x, int n, int pitch, float* out, cudaStream_t s);
cudaError_t launch_init_state(const int* y, int n, float* alpha, float* f, cudaStream_t s);
cudaError_t launch_select(const float* alpha, const int* y, const float* f, int n, float c, int index_base,
                          unsigned long long* keys, cudaStream_t s);
cudaError_t launch_control(const CtlDev& a, cudaStream_t s);
cudaError_t launch_cache_lookup(const CacheDev& c, int key, int* d_out, cudaStream_t s);

}  // namespace dpsvm
