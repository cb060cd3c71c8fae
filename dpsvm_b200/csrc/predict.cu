// predict.cu -- the test/predict entry behind dpsvm_predict.
//
// Replaces SvmTrain::test_setup / aggregate_sv / get_train_accuracy (reference svmTrain.cu:569-665):
//   * support-vector compaction: thrust::remove_if(is_not_sv) keeps alpha > 0 (:538-546, :597-602)
//   * per sample: cublasSgemv over the SV matrix + transform_reduce(test_functor) (:637-662), i.e.
//       dual_i = sum_j y_j alpha_j expf(-gamma (|x_j|^2 + |t_i|^2 - 2 <x_j, t_i>))  [- b]
// here as ONE kernel over tiles of test rows (fp32 throughout; the sum order is per-warp strided
// then a fixed tree, the reference's transform_reduce order is unspecified).
// Also hosts the host-callable kernel-row cache lookup.
#include <algorithm>
#include <vector>

#include "dpsvm_internal.cuh"

namespace dpsvm {

// The evaluation is a dense contraction -- n_test x n_sv dot products of length d, then expf, a coefficient and a row
// sum -- so it is tiled like one: a CTA owns 128 test rows and walks the support vectors in tiles of 128, 16 columns
// of both operands at a time through shared memory (transposed on the way in, so that a thread reads its 8 + 8 operands
// as four float4), 8 x 8 accumulators per thread, next chunk's global loads in flight during the FMAs.  Every SV row is
// read once per 128 test rows instead of once per 8.  fp32 FFMA throughout: the labels must equal seq_test's.
// The support vectors of a run are split over gridDim.y CTAs per row tile when there are few test rows; each writes
// its partial sums and a second kernel adds them in a fixed order (no atomics: the result is deterministic).
constexpr int kPM = 128, kPN = 128, kPK = 16, kPThreads = 256, kPPad = 4;

struct PredictScratch {
    float* d_t = nullptr;      size_t t_cap = 0;        // test rows [n_test x pitch]
    float* d_tsq = nullptr;    size_t tsq_cap = 0;      // their squared norms
    int* d_sv = nullptr;       size_t sv_cap = 0;       // SV row indices (resident-relative)
    float* d_coef = nullptr;                            // y_j * alpha_j
    float* d_dual = nullptr;   size_t dual_cap = 0;
    float* d_part = nullptr;   size_t part_cap = 0;     // [splits][n_test] partial sums
    int* d_lookup = nullptr;                            // cache lookup result
};

__global__ void __launch_bounds__(kPThreads, 2)
predict_tile_kernel(const float* __restrict__ x, const float* __restrict__ xsq, int pitch, const int* __restrict__ sv,
                    const float* __restrict__ coef, int nsv, const float* __restrict__ t, const float* __restrict__ tsq, int n_test,
                    float gamma, int tiles_per_split, float* __restrict__ partial)
{
    __shared__ __align__(16) float sA[2][kPK][kPM + kPPad];
    __shared__ __align__(16) float sB[2][kPK][kPN + kPPad];
    __shared__ int sRow[kPN];
    __shared__ float sCoef[kPN], sXsq[kPN];
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int m0 = blockIdx.x * kPM;
    const int n_tiles = (nsv + kPN - 1) / kPN;
    const int tile_begin = blockIdx.y * tiles_per_split, tile_end = min(n_tiles, tile_begin + tiles_per_split);
    const int nk = (pitch + kPK - 1) / kPK;
    // global -> shared: thread (lr, kq) moves one float4 (4 consecutive columns) of rows lr and lr + 64 of each operand
    const int lr = tid >> 2, kq = tid & 3;
    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
    __shared__ float sTsq[kPM];
    if (tid < kPM) sTsq[tid] = (m0 + tid < n_test) ? tsq[m0 + tid] : 0.f;
    // rows beyond the end read row 0 and are never stored; their columns beyond pitch read zero
    const float* a_row0 = t + (size_t)((m0 + lr < n_test) ? m0 + lr : 0) * pitch;
    const float* a_row1 = t + (size_t)((m0 + lr + 64 < n_test) ? m0 + lr + 64 : 0) * pitch;
    float dual_acc[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) dual_acc[r] = 0.f;

    for (int tile = tile_begin; tile < tile_end; ++tile) {
        const int n0 = tile * kPN;
        __syncthreads();                      // previous tile's epilogue has read sCoef / sXsq, its last chunk sA / sB
        if (tid < kPN) {
            const int j = n0 + tid;
            const bool ok = j < nsv;
            const int row = ok ? sv[j] : 0;
            sRow[tid] = ok ? row : -1;
            sCoef[tid] = ok ? coef[j] : 0.f;
            sXsq[tid] = xsq[row];             // (a column past the end stands in as row 0 with coefficient 0: a finite kernel value)
        }
        __syncthreads();
        // (columns of a tile beyond the last support vector read row 0 with coefficient 0)
        const float* b_row0 = x + (size_t)max(sRow[lr], 0) * pitch;
        const float* b_row1 = x + (size_t)max(sRow[lr + 64], 0) * pitch;
        float acc[8][8];
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int q = 0; q < 8; ++q) acc[r][q] = 0.f;

        float4 ga0, ga1, gb0, gb1;
        auto gload = [&](int kc) {
            const int k = kc * kPK + kq * 4;
            const bool kin = k < pitch;
            ga0 = kin ? __ldg(reinterpret_cast<const float4*>(a_row0 + k)) : zero4;
            ga1 = kin ? __ldg(reinterpret_cast<const float4*>(a_row1 + k)) : zero4;
            gb0 = kin ? __ldg(reinterpret_cast<const float4*>(b_row0 + k)) : zero4;
            gb1 = kin ? __ldg(reinterpret_cast<const float4*>(b_row1 + k)) : zero4;
        };
        auto sstore = [&](int buf) {
            float(*A)[kPM + kPPad] = sA[buf];
            float(*B)[kPN + kPPad] = sB[buf];
            const int k = kq * 4;
            A[k][lr] = ga0.x; A[k + 1][lr] = ga0.y; A[k + 2][lr] = ga0.z; A[k + 3][lr] = ga0.w;
            A[k][lr + 64] = ga1.x; A[k + 1][lr + 64] = ga1.y; A[k + 2][lr + 64] = ga1.z; A[k + 3][lr + 64] = ga1.w;
            B[k][lr] = gb0.x; B[k + 1][lr] = gb0.y; B[k + 2][lr] = gb0.z; B[k + 3][lr] = gb0.w;
            B[k][lr + 64] = gb1.x; B[k + 1][lr + 64] = gb1.y; B[k + 2][lr + 64] = gb1.z; B[k + 3][lr + 64] = gb1.w;
        };
        gload(0);
        sstore(0);
        __syncthreads();
        for (int kc = 0; kc < nk; ++kc) {
            const int buf = kc & 1;
            if (kc + 1 < nk) gload(kc + 1);
#pragma unroll
            for (int k = 0; k < kPK; ++k) {
                const float4 a_lo = *reinterpret_cast<const float4*>(&sA[buf][k][ty * 4]);
                const float4 a_hi = *reinterpret_cast<const float4*>(&sA[buf][k][64 + ty * 4]);
                const float4 b_lo = *reinterpret_cast<const float4*>(&sB[buf][k][tx * 4]);
                const float4 b_hi = *reinterpret_cast<const float4*>(&sB[buf][k][64 + tx * 4]);
                const float av[8] = {a_lo.x, a_lo.y, a_lo.z, a_lo.w, a_hi.x, a_hi.y, a_hi.z, a_hi.w};
                const float bv[8] = {b_lo.x, b_lo.y, b_lo.z, b_lo.w, b_hi.x, b_hi.y, b_hi.z, b_hi.w};
#pragma unroll
                for (int r = 0; r < 8; ++r)
#pragma unroll
                    for (int q = 0; q < 8; ++q) acc[r][q] = __fmaf_rn(av[r], bv[q], acc[r][q]);
            }
            if (kc + 1 < nk) {
                sstore(buf ^ 1);              // the other buffer was last read one chunk ago, before the barrier below
                __syncthreads();
            }
        }
        // y * alpha * expf(-1 * gamma * (xsq_j + i_sq - (2*dp)))          svmTrain.cu:564
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const int col = q < 4 ? tx * 4 + q : 64 + tx * 4 + (q - 4);
            const float cj = sCoef[col], xj = sXsq[col];
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                const float ti = sTsq[r < 4 ? ty * 4 + r : 64 + ty * 4 + (r - 4)];
                const float arg = __fsub_rn(__fadd_rn(xj, ti), __fmul_rn(2.0f, acc[r][q]));
                dual_acc[r] = __fmaf_rn(cj, expf(__fmul_rn(-gamma, arg)), dual_acc[r]);
            }
        }
    }
    // the 16 threads of a half-warp share their 8 rows: xor-butterfly, then one of them stores
#pragma unroll
    for (int off = 8; off >= 1; off >>= 1)
#pragma unroll
        for (int r = 0; r < 8; ++r) dual_acc[r] += __shfl_xor_sync(0xffffffffu, dual_acc[r], off);
    if (tx == 0) {
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            const int m = m0 + (r < 4 ? ty * 4 + r : 64 + ty * 4 + (r - 4));
            if (m < n_test) partial[(size_t)blockIdx.y * n_test + m] = dual_acc[r];
        }
    }
}

// dual_i = sum of the partial sums in split order, minus b (svmTrain.cu:652)
__global__ void predict_finish_kernel(const float* __restrict__ partial, int splits, int n_test, float b_sub, float* __restrict__ dual)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_test) return;
    float d = 0.f;
    for (int s = 0; s < splits; ++s) d += partial[(size_t)s * n_test + i];
    dual[i] = d - b_sub;
}

void predict_free(dpsvm_ctx* c)
{
    PredictScratch* s = static_cast<PredictScratch*>(c->predict);
    if (!s) return;
    cudaFree(s->d_t); cudaFree(s->d_tsq); cudaFree(s->d_sv); cudaFree(s->d_coef); cudaFree(s->d_dual); cudaFree(s->d_part); cudaFree(s->d_lookup);
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
    // |t_i|^2: the training rows' own norms, or one pass over the uploaded test rows
    const float* d_tsq = c->d_xsq;
    if (!on_train) {
        if (s->tsq_cap < (size_t)n_test) {
            cudaFree(s->d_tsq); s->d_tsq = nullptr; s->tsq_cap = 0;
            CU(cudaMalloc(&s->d_tsq, sizeof(float) * (size_t)n_test));
            s->tsq_cap = (size_t)n_test;
        }
        CU(launch_row_norms(d_t, n_test, pitch, s->d_tsq, c->stream));
        d_tsq = s->d_tsq;
    }
    const int row_tiles = (n_test + kPM - 1) / kPM, sv_tiles = (nsv + kPN - 1) / kPN;
    // enough CTAs for two waves: split the support vectors when there are few row tiles
    int splits = 1;
    if (sv_tiles > 0) {
        const int want = (2 * 2 * device_sm_count() + row_tiles - 1) / row_tiles;
        splits = std::max(1, std::min(sv_tiles, want));
    }
    const int tiles_per_split = sv_tiles > 0 ? (sv_tiles + splits - 1) / splits : 1;
    splits = sv_tiles > 0 ? (sv_tiles + tiles_per_split - 1) / tiles_per_split : 1;
    if (s->part_cap < (size_t)splits * n_test) {
        cudaFree(s->d_part); s->d_part = nullptr; s->part_cap = 0;
        CU(cudaMalloc(&s->d_part, sizeof(float) * (size_t)splits * n_test));
        s->part_cap = (size_t)splits * n_test;
    }
    if (sv_tiles == 0) {
        CU(cudaMemsetAsync(s->d_part, 0, sizeof(float) * (size_t)n_test, c->stream));
    } else {
        predict_tile_kernel<<<dim3((unsigned)row_tiles, (unsigned)splits), kPThreads, 0, c->stream>>>(
            c->d_x, c->d_xsq, pitch, s->d_sv, s->d_coef, nsv, d_t, d_tsq, n_test, c->p.gamma, tiles_per_split, s->d_part);
        CU(cudaGetLastError());
    }
    predict_finish_kernel<<<(n_test + 255) / 256, 256, 0, c->stream>>>(s->d_part, splits, n_test, subtract_b ? b : 0.0f, s->d_dual);
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
