// svmTrain.cpp -- host side of class SvmTrain (include/svmTrain.h) on top of the C ABI
// (include/dpsvm_cuda.h).  No CUDA code here: every device operation is a call into libdpsvm.so.
#include "svmTrain.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include <iostream>

#include "dpsvm_cuda.h"
#include "svmTrainMain.hpp"

using namespace std;

struct SvmTrain::Impl {
    dpsvm_ctx* ctx = nullptr;
    std::vector<float> x;        // host copy of the full X (reference svmTrain.cu:338) for rbf_kernel()
    std::vector<int> y;          // host copy of the labels (:339)
    std::vector<float> alpha;    // global-length host alpha
    long long hits = 0, lookups = 0;
    int new_size = 0;            // number of support vectors (reference svmTrain.h:101)
    bool test_ready = false;
};

// Reference behaviour on failure: print and exit (svmTrain.cu:166-179).
static void die_on(int rc, const char* what)
{
    if (rc == DPSVM_OK) return;
    cout << what << " failed: " << dpsvm_last_error() << "\n";
    exit(EXIT_FAILURE);
}

SvmTrain::SvmTrain(int n_data, int d)
{
    impl = new Impl();
    num_train_data = n_data;
    start = d;
    end = d + n_data;
    rv = new float[4];           // reference allocates this in setup (svmTrain.cu:388)
    rv[0] = rv[1] = -1; rv[2] = DPSVM_F_SENTINEL; rv[3] = -DPSVM_F_SENTINEL;   // init, svmTrain.cu:312-315
    b = 0; b_hi = 0; b_lo = 0; converged = false;
}

SvmTrain::~SvmTrain()
{
    if (impl) {
        if (impl->ctx) dpsvm_destroy(impl->ctx);
        delete impl;
    }
    delete[] rv;
}

void SvmTrain::setup(std::vector<float>& raw_x, std::vector<int>& raw_y)
{
    dpsvm_params p;
    p.num_attributes = state.num_attributes;
    p.num_train_data = state.num_train_data;
    p.c = state.c;
    p.gamma = state.gamma;
    p.epsilon = state.epsilon;
    p.max_iter = state.max_iter;
    p.cache_size = state.cache_size;
    p.shard_rows = num_train_data;
    p.shard_start = start;
    p.device = -1;
    const char* dev = getenv("DPSVM_DEVICE");
    if (dev && *dev) p.device = atoi(dev);
    p.flags = DPSVM_FLAG_NONE;
    if ((long long)raw_x.size() < (long long)state.num_train_data * state.num_attributes ||
        (long long)raw_y.size() < (long long)state.num_train_data) {
        cout << "SvmTrain::setup: raw_x / raw_y smaller than num_train_data x num_attributes\n";
        exit(EXIT_FAILURE);
    }
    die_on(dpsvm_create(&p, &impl->ctx), "CUDA initialization");
    impl->x = raw_x;             // the caller keeps raw_x / raw_y; setup copies (svmTrain.cu:338-339)
    impl->y = raw_y;
    impl->alpha.assign((size_t)state.num_train_data, 0.0f);
    die_on(dpsvm_upload(impl->ctx, raw_x.data(), raw_y.data()), "SvmTrain::setup");
}

void SvmTrain::init_cuda_handles() {}
void SvmTrain::destroy_cuda_handles() {}
void SvmTrain::destroy_t_cuda_handles() {}

void SvmTrain::train_step1()
{
    die_on(dpsvm_step1(impl->ctx, rv, nullptr), "SvmTrain::train_step1");
}

void SvmTrain::train_step2(int I_hi, int I_lo, float alpha_hi_new, float alpha_lo_new)
{
    die_on(dpsvm_step2(impl->ctx, I_hi, I_lo, alpha_hi_new, alpha_lo_new, nullptr, nullptr), "SvmTrain::train_step2");
    if (I_lo >= 0 && I_lo < (int)impl->alpha.size()) impl->alpha[I_lo] = alpha_lo_new;
    if (I_hi >= 0 && I_hi < (int)impl->alpha.size()) impl->alpha[I_hi] = alpha_hi_new;
}

// The reference's update_f takes the old alphas explicitly; the device keeps them, so only the
// new values are forwarded (y_lo / y_hi and the old alphas are implied by the device state).
int SvmTrain::update_f(int I_lo, int I_hi, int, int, float, float, float alpha_lo_new, float alpha_hi_new)
{
    train_step2(I_hi, I_lo, alpha_hi_new, alpha_lo_new);
    return 0;
}

float* SvmTrain::lookup_cache(int I_idx, bool& cache_hit)
{
    float* line = nullptr;
    int hit = 0;
    die_on(dpsvm_cache_lookup(impl->ctx, I_idx, &line, &hit), "SvmTrain::lookup_cache");
    cache_hit = hit != 0;
    return line;
}

float SvmTrain::clip_value(float num, float low, float high)
{
    if (num < low) {
        return low;
    } else if (num > high) {
        return high;
    }
    return num;
}

void SvmTrain::get_x(float* x, float* x_copy, int idx, int num_attributes)
{
    const float* src = x + (size_t)idx * num_attributes;
    for (int i = 0; i < num_attributes; i++) x_copy[i] = src[i];
}

// Host RBF kernel on the host copy of X: difference, squared norm in fp32 in index order, exp of
// a float argument evaluated in double (reference svmTrain.cu:696-714).
float SvmTrain::rbf_kernel(int i1, int i2)
{
    const int d = state.num_attributes;
    const float* a = &impl->x[(size_t)i1 * d];
    const float* c = &impl->x[(size_t)i2 * d];
    float norm_sq = 0.0f;
    for (int k = 0; k < d; k++) {
        const float diff = c[k] + (-1.0f) * a[k];
        norm_sq = norm_sq + diff * diff;
    }
    return (float)exp(-1 * (float)state.gamma * norm_sq);
}

int SvmTrain::train(int max_new_iter)
{
    dpsvm_train_status st;
    die_on(dpsvm_train(impl->ctx, max_new_iter, &st), "SvmTrain::train");
    b_hi = st.b_hi; b_lo = st.b_lo; converged = st.converged != 0;
    impl->hits = st.cache_hits; impl->lookups = st.cache_lookups;
    die_on(dpsvm_download(impl->ctx, impl->alpha.data() + start, nullptr, nullptr), "SvmTrain::train (download)");
    return st.num_iter;
}

const std::vector<float>& SvmTrain::alpha() const { return impl->alpha; }
void SvmTrain::set_alpha(const float* a) { impl->alpha.assign(a, a + state.num_train_data); }
long long SvmTrain::cache_hits() const { return impl->hits; }
long long SvmTrain::cache_lookups() const { return impl->lookups; }
int SvmTrain::num_sv() const { return impl->new_size; }

void SvmTrain::test_setup()
{
    // g_alpha_c = g_alpha (svmTrain.cu:571): the device alpha of the owned rows is authoritative
    die_on(dpsvm_download(impl->ctx, impl->alpha.data() + start, nullptr, nullptr), "SvmTrain::test_setup");
    aggregate_sv();
    impl->test_ready = true;
}

void SvmTrain::aggregate_sv()
{
    int n = 0;
    for (size_t i = 0; i < impl->alpha.size(); i++) if (!(impl->alpha[i] <= 0)) n++;   // is_not_sv, svmTrain.cu:538-546
    impl->new_size = n;
    cout << "Number of SVs: " << n << "\n";                                           // svmTrain.cu:604
}

float SvmTrain::get_train_accuracy()
{
    if (!impl->test_ready) test_setup();
    int nsv = 0;
    float acc = 0.f;
    die_on(dpsvm_predict(impl->ctx, impl->alpha.data(), impl->y.data(), b, 1, nullptr, 0, nullptr, nullptr, nullptr,
                         &nsv, &acc), "SvmTrain::get_train_accuracy");
    return acc;
}
