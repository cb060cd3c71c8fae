// oracle/ref_harness.cpp -- TEST INFRASTRUCTURE, not product code.
//
// Compiles the reference's own sequential SMO (/root/reference/seq.cpp) *from where it
// lies* (the Makefile passes -I/root/reference; no reference source is copied into this
// repo) and exposes its functions through a small C interface so tests can drive them
// with in-memory data and read back full-precision results.  The build output goes to
// oracle/_ref/libseqref.so (git-ignored, travels to the GPU box).
//
// What is the reference's and what is this file's:
//   * set_I_arrays / get_I_up / get_I_low / rbf_kernel / update_f / clip_value /
//     initialize_f_array / get_train_accuracy are the reference's, unmodified
//     (seq.cpp:469-553, 398-415, 378-386, 388-396, 463-467, 323-350).
//   * the do-while below restates seq.cpp:195-260 line for line (main() itself cannot be
//     called in-process: it reads a CSV, prints every iteration and only writes a
//     6-significant-digit model file), adding only the trace capture.
//   * BLAS calls resolve to oracle/cblas_shim/cblas.h (see its header for provenance).
//
// The unmodified reference main() is ALSO built, as oracle/_ref/seq (see Makefile), and
// tests/test_oracle_ref.py checks that this loop and that binary agree on iteration
// count, b and the model file.

#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <getopt.h>
#include <vector>
#include <iostream>
#include <fstream>
#include <sstream>
#include <string>
#include <cblas.h>
#include "CycleTimer.h"

namespace seqref {
#define main seq_reference_main
#include "seq.cpp"
#undef main
}  // namespace seqref

extern "C" {

// Runs the reference SMO on row-major x[n*d], y[n] (+1/-1).
// trace_idx[2*k+0]=I_hi, [2*k+1]=I_low; trace_b[2*k+0]=b_hi, [2*k+1]=b_low for iteration k
// (first trace_cap iterations; pass NULL / 0 to skip).
// rule_disagreements counts iterations where a pure lowest-index tie-break would have picked
// a different (I_hi, I_low) than seq.cpp's "I0 then I1/I3 then I2/I4" scan order.
// Returns 0.
int seqref_train(const float* x_flat, const int* y_in, int n, int d,
                 float c, float gamma, float epsilon, int max_iter,
                 float* alpha_out, float* f_out,
                 int* trace_idx, float* trace_b, int trace_cap,
                 int* num_iter_out, float* b_hi_out, float* b_low_out,
                 int* rule_disagreements)
{
    using namespace seqref;
    state.num_attributes = d;
    state.num_train_data = n;
    state.c = c;
    state.gamma = gamma;
    state.epsilon = epsilon;
    state.max_iter = max_iter;

    float** x = new float*[n];
    for (int i = 0; i < n; i++) {
        x[i] = new float[d];
        memcpy(x[i], x_flat + (size_t)i * d, sizeof(float) * d);
    }
    int* y = new int[n];
    memcpy(y, y_in, sizeof(int) * n);

    float* alpha = new float[n]();
    float* f = new float[n];
    initialize_f_array(f, y);                       // seq.cpp:184

    std::vector<int> I[5];
    float b_low = 0, b_hi = 0;
    int num_iter = 0;
    int disagreements = 0;

    do {                                            // seq.cpp:195
        set_I_arrays(alpha, y, I);                  // :198
        int I_hi = get_I_up(f, I);                  // :201
        b_hi = f[I_hi];
        int I_low = get_I_low(f, I);                // :204
        b_low = f[I_low];

        if (rule_disagreements) {
            // lowest-index rule over the same snapshot (north_star's selection rule)
            int li_hi = 0, li_lo = 0; float mn = 1000000000, mx = -1000000000;
            for (int i = 0; i < n; i++) {
                bool up, low;
                if (alpha[i] == 0)            { up = (y[i] == 1);  low = !up; }
                else if (alpha[i] == state.c) { up = (y[i] == -1); low = !up; }
                else                          { up = low = true; }
                if (up  && f[i] < mn) { mn = f[i]; li_hi = i; }
                if (low && f[i] > mx) { mx = f[i]; li_lo = i; }
            }
            if (li_hi != I_hi || li_lo != I_low) disagreements++;
        }
        if (num_iter < trace_cap) {
            if (trace_idx) { trace_idx[2 * num_iter] = I_hi; trace_idx[2 * num_iter + 1] = I_low; }
            if (trace_b)   { trace_b[2 * num_iter] = b_hi;   trace_b[2 * num_iter + 1] = b_low; }
        }

        int y_low = y[I_low];                       // :222
        int y_hi = y[I_hi];
        float eta = rbf_kernel(x[I_hi], x[I_hi]) + rbf_kernel(x[I_low], x[I_low])
                    - (2 * rbf_kernel(x[I_low], x[I_hi]));                       // :228
        float alpha_low_old = alpha[I_low];
        float alpha_hi_old = alpha[I_hi];
        float s = y_low * y_hi;                                                    // :237
        float alpha_low_new = alpha_low_old + (y_low * (b_hi - b_low) / eta);      // :238
        float alpha_hi_new = alpha_hi_old + (s * (alpha_low_old - alpha_low_new)); // :239
        alpha_low_new = clip_value(alpha_low_new, 0.0, state.c);                   // :242
        alpha_hi_new = clip_value(alpha_hi_new, 0.0, state.c);
        alpha[I_low] = alpha_low_new;                                              // :249
        alpha[I_hi] = alpha_hi_new;
        update_f(f, x, I_low, I_hi, y_low, y_hi, alpha_low_old, alpha_hi_old,
                 alpha_low_new, alpha_hi_new);                                     // :253
        num_iter++;
    } while ((b_low > (b_hi + (2 * state.epsilon))) && num_iter < state.max_iter); // :260

    if (alpha_out) memcpy(alpha_out, alpha, sizeof(float) * n);
    if (f_out) memcpy(f_out, f, sizeof(float) * n);
    if (num_iter_out) *num_iter_out = num_iter;
    if (b_hi_out) *b_hi_out = b_hi;
    if (b_low_out) *b_low_out = b_low;
    if (rule_disagreements) *rule_disagreements = disagreements;

    for (int i = 0; i < n; i++) delete[] x[i];
    delete[] x; delete[] y; delete[] alpha; delete[] f;
    return 0;
}

// The reference's RBF kernel on two rows (seq.cpp:398-415).
float seqref_rbf(const float* x1, const float* x2, int d, float gamma)
{
    seqref::state.num_attributes = d;
    seqref::state.gamma = gamma;
    return seqref::rbf_kernel(const_cast<float*>(x1), const_cast<float*>(x2));
}

// Decision values the way seq.cpp:323-350 / seq_test.cpp:187-210 form them (NO b term):
// dual_i = sum_{j: alpha_j != 0} y_j alpha_j K(x_j, t_i), accumulated in fp32 in index order.
// labels_out[i] = dual<0 ? -1 : +1.  Uses the reference's rbf_kernel.
int seqref_predict(const float* x_flat, const int* y, const float* alpha, int n, int d, float gamma,
                   const float* t_flat, int n_test, float* dual_out, int* labels_out)
{
    using namespace seqref;
    state.num_attributes = d;
    state.gamma = gamma;
    for (int i = 0; i < n_test; i++) {
        float dual = 0;
        for (int j = 0; j < n; j++) {
            if (alpha[j] != 0) {
                dual += y[j] * alpha[j] * rbf_kernel(const_cast<float*>(x_flat + (size_t)j * d),
                                                      const_cast<float*>(t_flat + (size_t)i * d));
            }
        }
        if (dual_out) dual_out[i] = dual;
        if (labels_out) labels_out[i] = dual < 0 ? -1 : 1;
    }
    return 0;
}

}  // extern "C"
