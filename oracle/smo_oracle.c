/*
 * oracle/smo_oracle.c -- TEST INFRASTRUCTURE ("the oracle"), not product code.
 *
 * A plain-C CPU restatement of the reference's SMO training path.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it;
 * nothing under dpsvm_b200/ or include/ links, imports or calls it.
 *
 * Parity status: PINNED against the reference itself.  tests/test_oracle_ref.py runs this
 * file and oracle/_ref/libseqref.so (the reference's own seq.cpp compiled from
 * /root/reference, see ref_harness.cpp / Makefile) on the same inputs and requires
 * bit-identical (I_hi, I_low) traces, iteration counts, alpha[], f[] and b in kernel_form 0 /
 * select_rule 0; tests/golden/ holds outputs of that reference build for the GPU box, where
 * /root/reference does not exist.  What stays unpinned is the third-party BLAS boundary
 * (netlib CBLAS, not vendored in the reference) -- see cblas_shim/cblas.h.
 *
 * Each function cites the reference lines it follows (paths under /root/reference).
 *
 * kernel_form:  0 = seq.cpp's:  K = (float)exp(-(double)gamma * nrm2(x1-x2)^2)   (seq.cpp:398-415)
 *               1 = the GPU path's: K = expf(-gamma*(|xj|^2 + |xi|^2 - 2<xj,xi>))  (svmTrain.cu:128-130)
 *                   with eta from the host kernel (float)exp(-gamma*sdot(diff,diff)) (svmTrain.cu:696-714)
 * select_rule:  0 = seq.cpp's scan order I0, I1, I2 / I0, I3, I4, strict compares (seq.cpp:495-553)
 *               1 = lowest global index among equal f (north_star's rule; what the CUDA path does)
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "cblas_shim/cblas.h"

#ifdef _OPENMP
#include <omp.h>
#endif

/* ---- I-set membership: seq.cpp:469-493 == svmTrain.cu:41-95 (arbitrary_functor) ---------- */
static inline void iset_membership(float alpha, int y, float c, int* up, int* low)
{
    if (alpha == 0)      { *up = (y == 1);  *low = !*up; }   /* I1 (up) / I4 (low) */
    else if (alpha == c) { *up = (y == -1); *low = !*up; }   /* I2 (up) / I3 (low) */
    else                 { *up = 1; *low = 1; }              /* I0 */
}

/* Which of seq.cpp's five sets a row falls in (seq.cpp:476-491). */
static inline int iset_id(float alpha, int y, float c)
{
    if (alpha == 0) return y == 1 ? 1 : 4;
    if (alpha == c) return y == -1 ? 2 : 3;
    return 0;
}

/*
 * Working-set selection over one snapshot.
 * rule 0: seq.cpp:495-553 -- scan set I0 then I1 then I2 with strict '<' (resp. I0, I3, I4
 *         with strict '>'), each set in ascending index order; defaults index 0, +-1e9.
 * rule 1: single ascending-index scan with strict compares == lowest index among equals.
 *         Same defaults (svmTrainMain.cpp:249-252).
 * idx_base is added to reported indices (global index of row 0 of this shard).
 */
void oracle_select(const float* alpha, const int* y, const float* f, int n, float c, int rule,
                   int idx_base, int* i_hi, int* i_low, float* b_hi, float* b_low)
{
    int ih = 0, il = 0;
    float mn = 1000000000, mx = -1000000000;
    if (rule == 0) {
        static const int up_sets[3] = {0, 1, 2};
        static const int low_sets[3] = {0, 3, 4};
        for (int s = 0; s < 3; s++)
            for (int i = 0; i < n; i++)
                if (iset_id(alpha[i], y[i], c) == up_sets[s] && f[i] < mn) { mn = f[i]; ih = i + idx_base; }
        for (int s = 0; s < 3; s++)
            for (int i = 0; i < n; i++)
                if (iset_id(alpha[i], y[i], c) == low_sets[s] && f[i] > mx) { mx = f[i]; il = i + idx_base; }
    } else {
        for (int i = 0; i < n; i++) {
            int up, low;
            iset_membership(alpha[i], y[i], c, &up, &low);
            if (up && f[i] < mn)  { mn = f[i]; ih = i + idx_base; }
            if (low && f[i] > mx) { mx = f[i]; il = i + idx_base; }
        }
    }
    *i_hi = ih; *i_low = il; *b_hi = mn; *b_low = mx;
}

/* seq.cpp:398-415 -- copy, saxpy(-1), snrm2, exp in double.  scratch holds d floats. */
static float rbf_seq(const float* x1, const float* x2, int d, float gamma, float* scratch)
{
    memcpy(scratch, x1, sizeof(float) * (size_t)d);
    cblas_saxpy(d, -1, x2, 1, scratch, 1);
    float norm = cblas_snrm2(d, scratch, 1);
    return (float)exp(-1 * (double)gamma * norm * norm);
}

/* svmTrain.cu:696-714 -- host kernel of the GPU driver: i2 - i1, sdot, exp of a float product. */
static float rbf_gpu_host(const float* x_i1, const float* x_i2, int d, float gamma, float* scratch)
{
    memcpy(scratch, x_i2, sizeof(float) * (size_t)d);
    cblas_saxpy(d, -1, x_i1, 1, scratch, 1);
    float norm_sq = cblas_sdot(d, scratch, 1, scratch, 1);
    return (float)exp(-1 * (float)gamma * norm_sq);
}

float oracle_rbf(const float* x1, const float* x2, int d, float gamma, int kernel_form)
{
    float* s = (float*)malloc(sizeof(float) * (size_t)(d > 0 ? d : 1));
    float r = kernel_form == 0 ? rbf_seq(x1, x2, d, gamma, s) : rbf_gpu_host(x1, x2, d, gamma, s);
    free(s);
    return r;
}

/* svmTrain.cu:361-364 -- |x_i|^2 as a sequential fp32 inner product starting from 0.0f. */
void oracle_row_norms(const float* x, int n, int d, float* out)
{
    for (int i = 0; i < n; i++) out[i] = cblas_sdot(d, x + (size_t)i * d, 1, x + (size_t)i * d, 1);
}

/* seq.cpp:388-396 == svmTrain.cu:674-682 */
static inline float clip_value(float num, float low, float high)
{
    if (num < low) return low;
    else if (num > high) return high;
    return num;
}

/*
 * The alpha pair update shared by both drivers: seq.cpp:228-250, svmTrainMain.cpp:279-299.
 * Note alpha_hi_new is formed from the UNCLIPPED alpha_low_new; eta is not guarded.
 */
void oracle_alpha_update(float eta, int y_hi, int y_low, float b_hi, float b_low,
                         float alpha_hi_old, float alpha_low_old, float c,
                         float* alpha_hi_new, float* alpha_low_new)
{
    float s = y_low * y_hi;
    float aln = alpha_low_old + (y_low * (b_hi - b_low) / eta);
    float ahn = alpha_hi_old + (s * (alpha_low_old - aln));
    *alpha_low_new = clip_value(aln, 0.0, c);
    *alpha_hi_new = clip_value(ahn, 0.0, c);
}

/*
 * Full training loop: seq.cpp:181-260 (kernel_form 0) / svmTrainMain.cpp:235-310 with
 * svmTrain.cu:98-137 for the f update (kernel_form 1).
 * warm_start != 0: alpha[] and f[] hold the starting state; else alpha=0, f=-y (seq.cpp:181-184,463-467).
 * nthreads > 1 parallelises ONLY the per-row f update (rows are independent, results identical).
 */
int oracle_smo_train(const float* x, const int* y, int n, int d,
                     float c, float gamma, float epsilon, int max_iter,
                     int kernel_form, int select_rule, int warm_start, int nthreads,
                     float* alpha, float* f,
                     int* trace_idx, float* trace_b, int trace_cap,
                     int* num_iter_out, float* b_hi_out, float* b_low_out,
                     int* rule_disagreements)
{
    if (!warm_start) {
        for (int i = 0; i < n; i++) { alpha[i] = 0; f[i] = -1 * y[i]; }
    }
    if (nthreads < 1) nthreads = 1;
    float* scratch = (float*)malloc(sizeof(float) * (size_t)d * (size_t)nthreads);
    float* xsq = NULL;
    if (kernel_form == 1) { xsq = (float*)malloc(sizeof(float) * (size_t)n); oracle_row_norms(x, n, d, xsq); }

    float b_low = 0, b_hi = 0;
    int num_iter = 0, disagreements = 0;
    do {
        int I_hi, I_low;
        oracle_select(alpha, y, f, n, c, select_rule, 0, &I_hi, &I_low, &b_hi, &b_low);
        if (rule_disagreements) {
            int oh, ol; float t1, t2;
            oracle_select(alpha, y, f, n, c, !select_rule, 0, &oh, &ol, &t1, &t2);
            if (oh != I_hi || ol != I_low) disagreements++;
        }
        if (num_iter < trace_cap) {
            if (trace_idx) { trace_idx[2 * num_iter] = I_hi; trace_idx[2 * num_iter + 1] = I_low; }
            if (trace_b)   { trace_b[2 * num_iter] = b_hi;   trace_b[2 * num_iter + 1] = b_low; }
        }
        const float* x_hi = x + (size_t)I_hi * d;
        const float* x_low = x + (size_t)I_low * d;
        int y_low = y[I_low], y_hi = y[I_hi];

        float eta;
        if (kernel_form == 0)
            eta = rbf_seq(x_hi, x_hi, d, gamma, scratch) + rbf_seq(x_low, x_low, d, gamma, scratch)
                  - (2 * rbf_seq(x_low, x_hi, d, gamma, scratch));                 /* seq.cpp:228 */
        else
            eta = rbf_gpu_host(x_hi, x_hi, d, gamma, scratch) + rbf_gpu_host(x_low, x_low, d, gamma, scratch)
                  - (2 * rbf_gpu_host(x_low, x_hi, d, gamma, scratch));            /* svmTrainMain.cpp:282 */

        float alpha_low_old = alpha[I_low], alpha_hi_old = alpha[I_hi];
        float alpha_low_new, alpha_hi_new;
        oracle_alpha_update(eta, y_hi, y_low, b_hi, b_low, alpha_hi_old, alpha_low_old, c,
                            &alpha_hi_new, &alpha_low_new);
        alpha[I_low] = alpha_low_new;                                              /* seq.cpp:249-250 */
        alpha[I_hi] = alpha_hi_new;

        if (kernel_form == 0) {
            /* seq.cpp:378-386 */
#ifdef _OPENMP
#pragma omp parallel for num_threads(nthreads) schedule(static)
#endif
            for (int i = 0; i < n; i++) {
#ifdef _OPENMP
                float* s = scratch + (size_t)omp_get_thread_num() * d;
#else
                float* s = scratch;
#endif
                const float* xi = x + (size_t)i * d;
                float delta = (((alpha_hi_new - alpha_hi_old) * y_hi * rbf_seq(x_hi, xi, d, gamma, s))
                             + ((alpha_low_new - alpha_low_old) * y_low * rbf_seq(x_low, xi, d, gamma, s)));
                f[i] += delta;
            }
        } else {
            /* svmTrain.cu:222,247 (two gemv) + :98-137 (update_functor) */
            float x_hi_sq = xsq[I_hi], x_lo_sq = xsq[I_low];
#ifdef _OPENMP
#pragma omp parallel for num_threads(nthreads) schedule(static)
#endif
            for (int i = 0; i < n; i++) {
                const float* xi = x + (size_t)i * d;
                float dot_hi = cblas_sdot(d, xi, 1, x_hi, 1);
                float dot_lo = cblas_sdot(d, xi, 1, x_low, 1);
                float rbf_hi = expf(-1 * gamma * (xsq[i] + x_hi_sq - (2 * dot_hi)));
                float rbf_lo = expf(-1 * gamma * (xsq[i] + x_lo_sq - (2 * dot_lo)));
                float delta = (((alpha_hi_new - alpha_hi_old) * y_hi * rbf_hi)
                             + ((alpha_low_new - alpha_low_old) * y_low * rbf_lo));
                f[i] += delta;
            }
        }
        num_iter++;
    } while ((b_low > (b_hi + (2 * epsilon))) && num_iter < max_iter);   /* seq.cpp:260 */

    if (num_iter_out) *num_iter_out = num_iter;
    if (b_hi_out) *b_hi_out = b_hi;
    if (b_low_out) *b_low_out = b_low;
    if (rule_disagreements) *rule_disagreements = disagreements;
    free(scratch);
    free(xsq);
    return 0;
}

/*
 * Decision values over a test set.
 * convention 0: seq.cpp:323-350 / seq_test.cpp:187-210 -- sum over alpha!=0, seq kernel, NO b.
 * convention 1: svmTrain.cu:633-665 -- sum over alpha>0 (is_not_sv removes alpha<=0, :538-546),
 *               expansion-form expf kernel (:548-566), then dual -= b.
 * label = dual<0 ? -1 : +1 in both.
 */
int oracle_predict(const float* x, const int* y, const float* alpha, int n, int d, float gamma, float b,
                   const float* t, int n_test, int convention, float* dual_out, int* labels_out)
{
    float* scratch = (float*)malloc(sizeof(float) * (size_t)(d > 0 ? d : 1));
    float* xsq = NULL;
    if (convention == 1) { xsq = (float*)malloc(sizeof(float) * (size_t)n); oracle_row_norms(x, n, d, xsq); }
    for (int i = 0; i < n_test; i++) {
        const float* ti = t + (size_t)i * d;
        float dual = 0;
        if (convention == 0) {
            for (int j = 0; j < n; j++)
                if (alpha[j] != 0) dual += y[j] * alpha[j] * rbf_seq(x + (size_t)j * d, ti, d, gamma, scratch);
        } else {
            float i_sq = cblas_sdot(d, ti, 1, ti, 1);
            for (int j = 0; j < n; j++)
                if (!(alpha[j] <= 0)) {
                    float dp = cblas_sdot(d, x + (size_t)j * d, 1, ti, 1);
                    dual += (y[j] * alpha[j] * expf(-1 * gamma * (xsq[j] + i_sq - (2 * dp))));
                }
            dual -= b;
        }
        if (dual_out) dual_out[i] = dual;
        if (labels_out) labels_out[i] = dual < 0 ? -1 : 1;
    }
    free(scratch);
    free(xsq);
    return 0;
}

/* svmTrainMain.cpp:367-384 -- ceil(N/P) rows per rank, remainder (possibly <= 0) to the last. */
void oracle_shards(int n_total, int world, int* shard_size, int* shard_disp)
{
    int per = (int)ceil((double)n_total / (double)world);
    for (int i = 0; i < world - 1; i++) shard_size[i] = per;
    shard_size[world - 1] = n_total - ((world - 1) * per);
    shard_disp[0] = 0;
    for (int i = 1; i < world; i++) shard_disp[i] = shard_disp[i - 1] + shard_size[i - 1];
}

/*
 * svmTrainMain.cpp:249-277 -- merge per-rank {I_hi, I_lo, f_hi, f_lo} (4 floats per rank, indices
 * carried as floats) with strict compares in ascending rank order; defaults index 0, +-1e9.
 */
void oracle_merge_candidates(const float* recv, int world, int* i_hi, int* i_lo, float* b_hi, float* b_lo)
{
    float max = -1000000000, min = 1000000000;
    int max_idx = 0, min_idx = 0;
    for (int i = 0; i < world; i++) {
        int ih = (int)recv[4 * i], il = (int)recv[4 * i + 1];
        float fh = recv[4 * i + 2], fl = recv[4 * i + 3];
        if (fl > max) { max = fl; max_idx = il; }
        if (fh < min) { min = fh; min_idx = ih; }
    }
    *i_hi = min_idx; *i_lo = max_idx; *b_hi = min; *b_lo = max;
}

/*
 * cache.cu:49-105 (myCache) replayed on a key sequence: LRU over `capacity` lines, lookup refreshes
 * recency (cache.cu:62-80), miss takes a free line or evicts the least recently used (:82-105).
 * hit_out[k] = 1 on hit; line_out[k] = line index used.  Returns number of hits.
 * O(capacity) per access -- fine for test sizes.
 */
int oracle_lru_replay(const int* keys, int num, int capacity, unsigned char* hit_out, int* line_out)
{
    int* line_key = (int*)malloc(sizeof(int) * (size_t)(capacity > 0 ? capacity : 1));
    long long* stamp = (long long*)malloc(sizeof(long long) * (size_t)(capacity > 0 ? capacity : 1));
    int size = 0, hits = 0;
    long long clock = 0;
    for (int k = 0; k < num; k++) {
        int found = -1;
        for (int l = 0; l < size; l++) if (line_key[l] == keys[k]) { found = l; break; }
        if (found >= 0) {
            stamp[found] = ++clock; hits++;
            if (hit_out) hit_out[k] = 1;
            if (line_out) line_out[k] = found;
            continue;
        }
        int line;
        if (size == capacity) {
            line = 0;
            for (int l = 1; l < size; l++) if (stamp[l] < stamp[line]) line = l;
        } else {
            line = size++;
        }
        line_key[line] = keys[k]; stamp[line] = ++clock;
        if (hit_out) hit_out[k] = 0;
        if (line_out) line_out[k] = line;
    }
    free(line_key); free(stamp);
    return hits;
}

/* ------------------------------------------------------------------------------------------------
 * The two dot-product ORDERS of the CUDA path, restated on the CPU so that the kernels can be checked bit
 * for bit against something that is not themselves (tests/test_gpu_parity.py reads the dot products back
 * from the kernel-row cache).  These are specifications of this repository, not of the reference, whose
 * cublasSgemv (svmTrain.cu:222,247) fixes no order; fmaf() is the correctly rounded fused multiply-add.
 * ---------------------------------------------------------------------------------------------- */

/* Dense rows (DotGeom, dpsvm_b200/csrc/smo_common.cuh): the row, zero padded to pitch4 = ceil(d/4) float4
 * chunks, is split over LPR lanes (4/8/16/32 by pitch4); lane s takes chunks s, s+LPR, ... and adds their
 * x,y,z,w products with sequential FMAs; an xor-butterfly (offsets LPR/2 .. 1) sums the lanes. */
float oracle_dot_dense_order(const float* x, const float* h, int d)
{
    const int pitch4 = (d + 3) / 4;
    const int lpr = pitch4 <= 32 ? 4 : pitch4 <= 64 ? 8 : pitch4 <= 128 ? 16 : 32;
    const int iters = (pitch4 + lpr - 1) / lpr;
    float acc[32];
    for (int s = 0; s < lpr; ++s) {
        float a = 0.0f;
        for (int it = 0; it < iters; ++it) {
            const int c4 = s + it * lpr;
            for (int k = 0; k < 4; ++k) {
                const int c = 4 * c4 + k;
                const float xv = (c4 < pitch4 && c < d) ? x[c] : 0.0f;
                const float hv = (c4 < pitch4 && c < d) ? h[c] : 0.0f;
                a = fmaf(xv, hv, a);
            }
        }
        acc[s] = a;
    }
    for (int off = lpr / 2; off >= 1; off >>= 1) {
        float nxt[32];
        for (int s = 0; s < lpr; ++s) nxt[s] = acc[s] + acc[s ^ off];
        memcpy(acc, nxt, sizeof(float) * (size_t)lpr);
    }
    return acc[0];
}

/* Sparse stream (SparseDev, dpsvm_b200/csrc/smo_kernels.cuh): the non-zeros of the row in ascending column
 * order are dealt round-robin to `lanes` lanes (entry e -> lane e % lanes); each lane adds its products with
 * sequential FMAs in that order; the same xor-butterfly sums the lanes.  -0.0 counts as zero. */
float oracle_dot_sparse_order(const float* x, const float* h, int d, int lanes)
{
    float acc[32] = {0};
    int e = 0;
    for (int c = 0; c < d; ++c) {
        if (x[c] != 0.0f) {
            acc[e % lanes] = fmaf(x[c], h[c], acc[e % lanes]);
            ++e;
        }
    }
    for (int off = lanes / 2; off >= 1; off >>= 1) {
        float nxt[32];
        for (int s = 0; s < lanes; ++s) nxt[s] = acc[s] + acc[s ^ off];
        memcpy(acc, nxt, sizeof(float) * (size_t)lanes);
    }
    return acc[0];
}

int oracle_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
