// host_utils.cu -- host-only entry points of include/dpsvm_cuda.h (no GPU touched).
#include <math.h>

#include "dpsvm_internal.cuh"

using namespace dpsvm;

extern "C" {

// svmTrainMain.cpp:367-384
int dpsvm_shard_layout(int n_total, int world, int* shard_size, int* shard_disp)
{
    if (n_total <= 0 || world <= 0 || !shard_size || !shard_disp) return fail(DPSVM_ERR_INVALID, "dpsvm_shard_layout: bad arguments");
    const int per = (int)ceil((double)n_total / (double)world);
    for (int i = 0; i < world - 1; ++i) shard_size[i] = per;
    shard_size[world - 1] = n_total - ((world - 1) * per);
    shard_disp[0] = 0;
    for (int i = 1; i < world; ++i) shard_disp[i] = shard_disp[i - 1] + shard_size[i - 1];
    // the reference happily hands a rank zero or negative rows (e.g. N=9, P=8); refuse instead
    if (shard_size[world - 1] <= 0)
        return fail(DPSVM_ERR_INVALID, "dpsvm_shard_layout: %d rows over %d ranks leaves the last rank %d rows", n_total, world,
                    shard_size[world - 1]);
    return DPSVM_OK;
}

// svmTrainMain.cpp:249-277
int dpsvm_merge_candidates(const float* recv, int world, int* i_hi, int* i_lo, float* b_hi, float* b_lo)
{
    if (!recv || world <= 0 || !i_hi || !i_lo || !b_hi || !b_lo) return fail(DPSVM_ERR_INVALID, "dpsvm_merge_candidates: bad arguments");
    float max = -DPSVM_F_SENTINEL, min = DPSVM_F_SENTINEL;
    int max_idx = 0, min_idx = 0;
    for (int r = 0; r < world; ++r) {
        const int ih = (int)recv[4 * r], il = (int)recv[4 * r + 1];
        const float fh = recv[4 * r + 2], fl = recv[4 * r + 3];
        if (fl > max) { max = fl; max_idx = il; }
        if (fh < min) { min = fh; min_idx = ih; }
    }
    *i_hi = min_idx; *i_lo = max_idx; *b_hi = min; *b_lo = max;
    return DPSVM_OK;
}

// svmTrainMain.cpp:289-295 with clip_value, svmTrain.cu:674-682
int dpsvm_alpha_update(float eta, int y_hi, int y_lo, float b_hi, float b_lo, float alpha_hi_old, float alpha_lo_old,
                       float c, float* alpha_hi_new, float* alpha_lo_new)
{
    if (!alpha_hi_new || !alpha_lo_new) return fail(DPSVM_ERR_INVALID, "dpsvm_alpha_update: null output");
    const float s = (float)(y_lo * y_hi);
    const float lo = alpha_lo_old + ((float)y_lo * (b_hi - b_lo) / eta);
    const float hi = alpha_hi_old + (s * (alpha_lo_old - lo));
    *alpha_lo_new = lo < 0.0f ? 0.0f : (lo > c ? c : lo);
    *alpha_hi_new = hi < 0.0f ? 0.0f : (hi > c ? c : hi);
    return DPSVM_OK;
}

int dpsvm_pitch_for(int d) { return d <= 0 ? 0 : (d + 3) / 4 * 4; }

// Default L2 policy of the fused pass for a shard of n rows (see dpsvm_ctx.cu: variant_for).
int dpsvm_pass_variant(int n, int pitch, int iteration_parity)
{
    const double x_mb = 4e-6 * (double)n * (double)pitch;
    int v = 0;
    if (x_mb > 64.0) v = 2 | (32 << 8);          // L2 hints, 32 MB pinned prefix
    if (iteration_parity & 1) v ^= 1;            // serpentine: odd iterations sweep in reverse
    return v;
}

uint64_t dpsvm_pack_key_up(float f, int32_t index) { return pack_key_up(f, index); }
uint64_t dpsvm_pack_key_low(float f, int32_t index) { return pack_key_low(f, index); }
void dpsvm_unpack_key_up(uint64_t key, float* f, int32_t* index) { unpack_key_up(key, f, index); }
void dpsvm_unpack_key_low(uint64_t key, float* f, int32_t* index) { unpack_key_low(key, f, index); }

}  // extern "C"
