// predict.cu -- the test/predict entry behind dpsvm_predict.
//
// Replaces SvmTrain::test_setup / aggregate_sv / get_train_accuracy (reference svmTrain.cu:569-665):
//   * support-vector compaction: thrust::remove_if(is_not_sv) keeps alpha > 0 (:538-546, :597-602)
//   * per sample: cublasSgemv over the SV matrix + transform_reduce(test_functor) (:637-662), i.e.
//       dual_i = sum_j y_j alpha_j expf(-gamma (|x_j|^2 + |t_i|^2 - 2 <x_j, t_i>))  [- b]
// here as ONE kernel over tiles of test rows (fp32 throughout; the sum order is per-warp strided
// then a fixed tree, the reference's transform_reduce order is unspecified).
// Also hosts the host-callable kernel-row cache lookup.
#include <vector>

#include "dpsvm_internal.cuh"

namespace dpsvm {

constexpr int kPredTile = 8;       // test rows per block
constexpr int kPredThreads = 256;

struct PredictScratch {
    float* d_t = nullptr;      size_t t_cap = 0;        // test rows [n_test x pitch]
    int* d_sv = nullptr;       size_t sv_cap = 0;       // SV row indices (resident-relative)
    float* d_coef = nullptr;                            // y_j * alpha_j
    float* d_dual = nullptr;   size_t dual_cap = 0;
    int* d_lookup = nullptr;                            // cache lookup result
};

__device__ __forceinline__ float4 ldg4(const float4* p) { return __ldg(p); }

// grid.x = ceil(n_test / kPredTile).  smem: kPredTile rows of pitch floats.
__global__ void __launch_bounds__(kPredThreads) predict_kernel(const float4* __restrict__ x4, const float* __restrict__ xsq,
                                                               int pitch4, const int* __restrict__ sv, const float* __restrict__ coef,
                                                               int nsv, const float4* __restrict__ t4, int n_test, float gamma,
                                                               float b_sub, float* __restrict__ dual)
{
    extern __shared__ __align__(16) unsigned char dyn[];
    float4* s_t = reinterpret_cast<float4*>(dyn);                  // [kPredTile][pitch4]
    __shared__ float s_tsq[kPredTile];
    __shared__ float s_part[kPredThreads / 32][kPredTile];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int t0 = blockIdx.x * kPredTile;
    for (int i = tid; i < kPredTile * pitch4; i += kPredThreads) {
        const int r = i / pitch4, c = i - r * pitch4;
        s_t[i] = (t0 + r < n_test) ? ldg4(t4 + (size_t)(t0 + r) * pitch4 + c) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    __syncthreads();
    // |t|^2 per tile row (warp r handles row r), same lane-strided order as the dot products below
    if (warp < kPredTile) {
        float acc = 0.f;
        for (int c = lane; c < pitch4; c += 32) {
            const float4 v = s_t[warp * pitch4 + c];
            acc = __fmaf_rn(v.x, v.x, acc); acc = __fmaf_rn(v.y, v.y, acc);
            acc = __fmaf_rn(v.z, v.z, acc); acc = __fmaf_rn(v.w, v.w, acc);
        }
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
        if (lane == 0) s_tsq[warp] = acc;
    }
    __syncthreads();

    float my_dual = 0.f;   // lane t < kPredTile accumulates the tile row t
    for (int j = warp; j < nsv; j += kPredThreads / 32) {
        const int row = sv[j];
        const float4* xr = x4 + (size_t)row * pitch4;
        float acc[kPredTile];
#pragma unroll
        for (int t = 0; t < kPredTile; ++t) acc[t] = 0.f;
        for (int c = lane; c < pitch4; c += 32) {
            const float4 v = ldg4(xr + c);
#pragma unroll
            for (int t = 0; t < kPredTile; ++t) {
                const float4 w = s_t[t * pitch4 + c];
                acc[t] = __fmaf_rn(v.x, w.x, acc[t]); acc[t] = __fmaf_rn(v.y, w.y, acc[t]);
                acc[t] = __fmaf_rn(v.z, w.z, acc[t]); acc[t] = __fmaf_rn(v.w, w.w, acc[t]);
            }
        }
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) {
#pragma unroll
            for (int t = 0; t < kPredTile; ++t) acc[t] += __shfl_xor_sync(0xffffffffu, acc[t], off);
        }
        if (lane < kPredTile) {
            float dp = acc[0];
#pragma unroll
            for (int t = 1; t < kPredTile; ++t) if (lane == t) dp = acc[t];
            // y * alpha * expf(-1 * gamma * (xsq_j + i_sq - (2*dp)))          svmTrain.cu:564
            const float arg = __fsub_rn(__fadd_rn(xsq[row], s_tsq[lane]), __fmul_rn(2.0f, dp));
            my_dual = __fmaf_rn(coef[j], expf(__fmul_rn(-gamma, arg)), my_dual);
        }
    }
    if (lane < kPredTile) s_part[warp][lane] = my_dual;
    __syncthreads();
    if (tid < kPredTile && t0 + tid < n_test) {
        float d = 0.f;
#pragma unroll
        for (int w = 0; w < kPredThreads / 32; ++w) d += s_part[w][tid];
        dual[t0 + tid] = d - b_sub;      // dual -= b   svmTrain.cu:652
    }
}

void predict_free(dpsvm_ctx* c)
{
    PredictScratch* s = static_cast<PredictScratch*>(c->predict);
    if (!s) return;
    cudaFree(s->d_t); cudaFree(s->d_sv); cudaFree(s->d_coef); cudaFree(s->d_dual); cudaFree(s->d_lookup);
    delete s;
    c->predict = nullptr;
}

static PredictScratch* scratch(dpsvm_ctx* c)
{
    if (!c->predict) c->predict = new PredictScratch();
    return static_cast<PredictScratch*>(c->predict);
}

int cache_lookup_host(dpsvm_ctx* c, int key, int* line, int* hit)
{
    PredictScratch* s = scratch(c);
    if (!s->d_lookup && cudaMalloc(&s->d_lookup, 2 * sizeof(int)) != cudaSuccess)
        return fail(DPSVM_ERR_NOMEM, "cache_lookup: cudaMalloc failed");
    if (launch_cache_lookup(c->cache, key, s->d_lookup, c->stream) != cudaSuccess)
        return fail(DPSVM_ERR_CUDA, "cache_lookup: launch failed");
    int out[2] = {-1, 0};
    if (cudaMemcpyAsync(out, s->d_lookup, sizeof(out), cudaMemcpyDeviceToHost, c->stream) != cudaSuccess ||
        cudaStreamSynchronize(c->stream) != cudaSuccess)
        return fail(DPSVM_ERR_CUDA, "cache_lookup: %s", cudaGetErrorString(cudaGetLastError()));
    *line = out[0]; *hit = out[1];
    return DPSVM_OK;
}

}  // namespace dpsvm

using namespace dpsvm;

#define CU(call)                                                                                     \
    do {                                                                                             \
        cudaError_t e__ = (call);                                                                    \
        if (e__ != cudaSuccess)                                                                      \
            return fail(e__ == cudaErrorMemoryAllocation ? DPSVM_ERR_NOMEM : DPSVM_ERR_CUDA,         \
                        "%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__));       \
    } while (0)

extern "C" int dpsvm_predict(dpsvm_ctx* c, const float* h_alpha, const int32_t* h_y, float b, int subtract_b,
                             const float* h_t, int n_test, const int32_t* h_labels, float* h_dual, int32_t* h_pred,
                             int* num_sv, float* accuracy)
{
    if (!c || !h_alpha || !h_y) return fail(DPSVM_ERR_INVALID, "dpsvm_predict: null argument");
    if (!c->uploaded) return fail(DPSVM_ERR_STATE, "dpsvm_predict before dpsvm_upload");
    if (n_test < 0) return fail(DPSVM_ERR_INVALID, "dpsvm_predict: n_test < 0");
    CU(cudaSetDevice(c->device));
    const int N = c->p.num_train_data, d = c->p.num_attributes, pitch = c->pitch;
    // aggregate_sv: keep alpha > 0 (is_not_sv removes alpha <= 0), original order
    std::vector<int> sv; std::vector<float> coef;
    for (int i = 0; i < N; ++i) {
        if (!(h_alpha[i] <= 0)) {
            if (i < c->res0 || i >= c->res0 + c->res_n)
                return fail(DPSVM_ERR_STATE, "dpsvm_predict: support vector %d is not resident on this rank", i);
            sv.push_back(i - c->res0);
            coef.push_back((float)h_y[i] * h_alpha[i]);
        }
    }
    const int nsv = (int)sv.size();
    if (num_sv) *num_sv = nsv;
    // h_t == NULL: evaluate on the resident training rows themselves (get_train_accuracy)
    const bool on_train = (h_t == nullptr);
    if (on_train) n_test = c->res_n;
    if (n_test == 0) { if (accuracy) *accuracy = 0.f; return DPSVM_OK; }

    PredictScratch* s = scratch(c);
    const size_t t_floats = (size_t)n_test * pitch;
    if (!on_train && s->t_cap < t_floats) {
        cudaFree(s->d_t); s->d_t = nullptr; s->t_cap = 0;
        CU(cudaMalloc(&s->d_t, sizeof(float) * t_floats));
        s->t_cap = t_floats;
    }
    if (s->sv_cap < (size_t)nsv + 1) {
        cudaFree(s->d_sv); cudaFree(s->d_coef); s->d_sv = nullptr; s->d_coef = nullptr; s->sv_cap = 0;
        CU(cudaMalloc(&s->d_sv, sizeof(int) * ((size_t)nsv + 1)));
        CU(cudaMalloc(&s->d_coef, sizeof(float) * ((size_t)nsv + 1)));
        s->sv_cap = (size_t)nsv + 1;
    }
    if (s->dual_cap < (size_t)n_test) {
        cudaFree(s->d_dual); s->d_dual = nullptr; s->dual_cap = 0;
        CU(cudaMalloc(&s->d_dual, sizeof(float) * (size_t)n_test));
        s->dual_cap = (size_t)n_test;
    }
    const float* d_t = c->d_x;
    if (!on_train) {
        if (pitch != d) CU(cudaMemsetAsync(s->d_t, 0, sizeof(float) * t_floats, c->stream));
        CU(cudaMemcpy2DAsync(s->d_t, sizeof(float) * (size_t)pitch, h_t, sizeof(float) * (size_t)d, sizeof(float) * (size_t)d,
                             (size_t)n_test, cudaMemcpyHostToDevice, c->stream));
        d_t = s->d_t;
    }
    if (nsv > 0) {
        CU(cudaMemcpyAsync(s->d_sv, sv.data(), sizeof(int) * (size_t)nsv, cudaMemcpyHostToDevice, c->stream));
        CU(cudaMemcpyAsync(s->d_coef, coef.data(), sizeof(float) * (size_t)nsv, cudaMemcpyHostToDevice, c->stream));
    }
    const size_t smem = sizeof(float) * (size_t)kPredTile * pitch;
    if (smem > 200 * 1024) return fail(DPSVM_ERR_LIMIT, "dpsvm_predict: num_attributes %d too large for the predict tile", d);
    CU(cudaFuncSetAttribute(predict_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(200 * 1024)));
    const int blocks = (n_test + kPredTile - 1) / kPredTile;
    predict_kernel<<<blocks, kPredThreads, smem, c->stream>>>(
        reinterpret_cast<const float4*>(c->d_x), c->d_xsq, pitch / 4, s->d_sv, s->d_coef, nsv,
        reinterpret_cast<const float4*>(d_t), n_test, c->p.gamma, subtract_b ? b : 0.0f, s->d_dual);
    CU(cudaGetLastError());
    std::vector<float> dual((size_t)n_test);
    CU(cudaMemcpyAsync(dual.data(), s->d_dual, sizeof(float) * (size_t)n_test, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    int correct = 0;
    for (int i = 0; i < n_test; ++i) {
        const int result = dual[i] < 0.0f ? -1 : 1;           // svmTrain.cu:654-657
        if (h_dual) h_dual[i] = dual[i];
        if (h_pred) h_pred[i] = result;
        const int32_t* lab = h_labels ? h_labels : (on_train ? h_y + c->res0 : nullptr);
        if (lab && result == lab[i]) ++correct;
    }
    if (accuracy) *accuracy = (float)correct / (float)n_test;  // svmTrain.cu:664
    return DPSVM_OK;
}
