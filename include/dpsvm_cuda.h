/*
 * dpsvm_cuda.h -- the C ABI of libdpsvm.so: DPSVM's SMO training hot path on B200 (sm_100a).
 *
 * This is the drop-in boundary.  Everything above it (include/svmTrain.h, the CLI, Python glue) is
 * host code that only calls these functions; everything below it is hand-written CUDA.  Signatures
 * use plain pointers and sizes only.  Every entry point cites the reference interface it replaces
 * (paths relative to the reference checkout, farshid83/dpsvm).
 *
 * Conventions
 *   - return value: 0 (DPSVM_OK) or a negative dpsvm_status; dpsvm_last_error() gives the text of
 *     the last failure on the calling thread.  Nothing here calls exit() (the reference prints and
 *     exits, svmTrain.cu:166-179; include/svmTrain.h keeps that behaviour on top of these codes).
 *   - "d_" pointers are device pointers on the current CUDA device; "h_" / unprefixed are host.
 *   - stream arguments are cudaStream_t passed as void* (NULL = legacy default stream).
 *   - X is row-major fp32 with a row pitch (in floats) that is a multiple of 4, padding zero-filled
 *     (dpsvm_pitch_for()); labels are int32 +1/-1; indices are int32 GLOBAL row numbers.
 *   - device pointers to X rows (d_x, d_x_hi, d_x_lo) must be 16-byte aligned: rows are read as 128-bit words and
 *     staged by TMA bulk copies.  A misaligned pointer is rejected with DPSVM_ERR_INVALID.
 *   - an EMPTY selection key (an I-set with no member) stands for the reference's default "index 0, +-1e9"
 *     (svmTrainMain.cpp:249-252).  Where the rows a call holds start at a global index > 0 (index_base, or a
 *     trainer created with DPSVM_FLAG_X_SHARD_ONLY and shard_start > 0 on a single GPU), the first row held
 *     stands in for row 0.
 *   - there is no CPU fallback: without a usable CUDA device every compute entry point returns
 *     DPSVM_ERR_CUDA.
 */
#ifndef DPSVM_CUDA_H
#define DPSVM_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DPSVM_ABI_VERSION 1

typedef enum {
    DPSVM_OK = 0,
    DPSVM_ERR_INVALID = -1,   /* bad argument */
    DPSVM_ERR_CUDA = -2,      /* CUDA runtime/driver error (no device, launch failure, ...) */
    DPSVM_ERR_NOMEM = -3,     /* device or host allocation failed */
    DPSVM_ERR_STATE = -4,     /* call out of order (e.g. train before upload) */
    DPSVM_ERR_COMM = -5,      /* NCCL / multi-GPU exchange failure */
    DPSVM_ERR_LIMIT = -6      /* size beyond a documented maximum */
} dpsvm_status;

/* Empty-set sentinels of the reference: svmTrain.cu:61,67,314-315; svmTrainMain.cpp:249-252. */
#define DPSVM_F_SENTINEL 1000000000.0f
#define DPSVM_KEY_EMPTY 0xFFFFFFFFFFFFFFFFull

const char* dpsvm_last_error(void);
int dpsvm_abi_version(void);

/* ------------------------------------------------------------------------------------------------
 * Host-only helpers (no GPU touched)
 * ---------------------------------------------------------------------------------------------- */

/* svmTrainMain.cpp:367-384 initialize_shard_sizes / initialize_shard_disp: ceil(N/P) rows per rank,
 * the remainder to the last rank.  Unlike the reference a non-positive last shard is an error. */
int dpsvm_shard_layout(int n_total, int world, int* shard_size, int* shard_disp);

/* svmTrainMain.cpp:249-277: merge per-rank {I_hi, I_lo, f_hi, f_lo} quadruples (the rv[4] floats of
 * svmTrain.h:108, gathered by the Allgather at svmTrainMain.cpp:244) into the global pair. Strict
 * compares in rank order; defaults index 0 and +-1e9. */
int dpsvm_merge_candidates(const float* recv, int world, int* i_hi, int* i_lo, float* b_hi, float* b_lo);

/* svmTrainMain.cpp:282-295 + svmTrain.cu:674-682: the alpha pair update for a given eta
 * (alpha_hi from the UNCLIPPED alpha_lo, then both clipped to [0,c]). */
int dpsvm_alpha_update(float eta, int y_hi, int y_lo, float b_hi, float b_lo, float alpha_hi_old,
                       float alpha_lo_old, float c, float* alpha_hi_new, float* alpha_lo_new);

/* Row pitch (floats) used for a given attribute count: d rounded up to a multiple of 4. */
int dpsvm_pitch_for(int d);

/* The `variant` word the trainer uses for the fused pass of iteration number `iteration_parity` on a
 * shard of n rows: bit 0 = sweep back to front (serpentine order across iterations), bit 1 = L2
 * eviction-priority hints with (variant >> 8) MB of X kept resident.  Results do not depend on it. */
int dpsvm_pass_variant(int n, int pitch, int iteration_parity);

/* Packed selection keys: (orderable(f) << 32) | index for I_up, (orderable(-f) << 32) | index for
 * I_low, so that an unsigned 64-bit MIN yields argmin f / argmax f with the lowest index among
 * equal values (north_star's tie rule; replaces i_helper, svmTrain.h:24-37, and my_maxmin,
 * svmTrain.cu:400-467).  -0.0f is folded onto +0.0f so key order equals float '<' order. */
uint64_t dpsvm_pack_key_up(float f, int32_t index);
uint64_t dpsvm_pack_key_low(float f, int32_t index);
/* Inverse; DPSVM_KEY_EMPTY decodes to index 0 and +1e9 / -1e9 (svmTrainMain.cpp:249-252). */
void dpsvm_unpack_key_up(uint64_t key, float* f, int32_t* index);
void dpsvm_unpack_key_low(uint64_t key, float* f, int32_t* index);

/* ------------------------------------------------------------------------------------------------
 * Stateless kernel launchers on caller-owned device memory
 * ---------------------------------------------------------------------------------------------- */

/* svmTrain.cu:357-364 (host loop of N thrust::inner_product) as one kernel:
 * d_xsq[i] = <x_i, x_i>, accumulated in exactly the order dpsvm_smo_pass accumulates <x_i, x_sel>,
 * so that |x_i|^2 + |x_i|^2 - 2<x_i,x_i> == 0 bit-exactly. */
int dpsvm_row_norms(const float* d_x, int n, int d, int pitch, float* d_xsq, void* stream);

/* svmTrain.cu:349,379-380: alpha = 0, f = -y. */
int dpsvm_init_state(const int32_t* d_y, int n, float* d_alpha, float* d_f, void* stream);

/* train_step1's device work, svmTrain.cu:469-476 (arbitrary_functor :41-95 + reduce my_maxmin
 * :400-467) as one kernel: classify every row into I_up / I_low from (alpha, y, c) with exact
 * ==0 / ==c tests and reduce to two packed keys.  d_keys[0] = key_up (argmin f over I_up),
 * d_keys[1] = key_low (argmax f over I_low); both are (re)initialised to DPSVM_KEY_EMPTY by this
 * call.  index_base = global index of row 0 (SvmTrain::start, svmTrain.cu:307). */
int dpsvm_select(const float* d_alpha, const int32_t* d_y, const float* d_f, int n, float c,
                 int index_base, uint64_t* d_keys, void* stream);

/* Arguments of one fused pass.  Replaces train_step2's device work: the two cublasSgemv
 * (svmTrain.cu:222,247), update_functor (:98-137, for_each :276-278) and the following
 * train_step1 (:469-476), in ONE sweep over the X shard:
 *   dot_hi[j] = <x_j, x_hi>, dot_lo[j] = <x_j, x_lo>
 *   f[j] += delta_hi * expf(-gamma*(xsq[j] + xsq_hi - 2 dot_hi[j]))
 *         + delta_lo * expf(-gamma*(xsq[j] + xsq_lo - 2 dot_lo[j]))
 *   then classify row j with the CURRENT alpha and reduce packed keys into d_keys (which the
 *   caller must have set to DPSVM_KEY_EMPTY, e.g. via dpsvm_select or cudaMemset 0xFF).
 * delta_hi = (alpha_hi_new - alpha_hi_old) * y_hi, delta_lo likewise (svmTrain.cu:133).
 * d_alpha must already hold the new alpha values.  d_line_hi / d_line_lo, when non-NULL, receive
 * the n dot products (the kernel-row cache line contents, cache.cu / svmTrain.cu:203-249). */
typedef struct {
    const float* d_x;        /* [n x pitch] rows of this shard */
    int n, d, pitch;
    const float* d_xsq;      /* [n] */
    const int32_t* d_y;      /* [n] */
    const float* d_alpha;    /* [n] */
    float* d_f;              /* [n] in/out */
    const float* d_x_hi;     /* [pitch] selected row I_hi (zero padded) */
    const float* d_x_lo;     /* [pitch] selected row I_lo */
    float xsq_hi, xsq_lo;
    float delta_hi, delta_lo;
    float gamma, c;
    int index_base;
    uint64_t* d_keys;        /* [2] in/out */
    float* d_line_hi;        /* optional [n] */
    float* d_line_lo;        /* optional [n] */
    int variant;             /* 0 = plain forward sweep; see dpsvm_pass_variant */
} dpsvm_pass_args;

int dpsvm_smo_pass(const dpsvm_pass_args* args, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Trainer context: owns the device state of one rank (one GPU)
 * ---------------------------------------------------------------------------------------------- */

typedef struct dpsvm_ctx dpsvm_ctx;

/* Field for field the training parameters of state_model (svmTrainMain.hpp:4-19) plus the shard
 * geometry SvmTrain's constructor takes (svmTrain.cu:305-316) and an explicit device ordinal (the
 * reference never calls cudaSetDevice). */
typedef struct {
    int num_attributes;   /* state.num_attributes (-a) */
    int num_train_data;   /* state.num_train_data (-x): GLOBAL row count N */
    float c;              /* state.c (-c) */
    float gamma;          /* state.gamma (-g) */
    float epsilon;        /* state.epsilon (-e) */
    int max_iter;         /* state.max_iter (-n) */
    int cache_size;       /* state.cache_size (-s): kernel-row cache lines; 0 disables the cache */
    int shard_rows;       /* SvmTrain(n_data, .) : rows owned by this rank */
    int shard_start;      /* SvmTrain(., d)      : global index of the first owned row */
    int device;           /* CUDA device ordinal; -1 = current device */
    int flags;            /* DPSVM_FLAG_* */
} dpsvm_params;

#define DPSVM_FLAG_NONE 0
#define DPSVM_FLAG_NO_GRAPH 1        /* launch kernels one by one instead of through a CUDA graph */
#define DPSVM_FLAG_X_SHARD_ONLY 2    /* dpsvm_upload receives / keeps only the owned rows of X */
#define DPSVM_FLAG_GRAPH_LOOP 4      /* run the loop as a CUDA graph of [control, pass] kernel pairs instead of the
                                        resident (persistent, cooperative) kernel; results are identical */
#define DPSVM_FLAG_DENSE 8           /* the sweep always reads the dense rows (what the reference's cublasSgemv does,
                                        svmTrain.cu:222,247) */
#define DPSVM_FLAG_SPARSE 16         /* the sweep always reads the sparse stream of the owned rows (rows of at most
                                        1024 floats, fewer than 2^27 steps of 32 non-zero slots per shard: DPSVM_ERR_LIMIT
                                        at upload otherwise).  Neither flag: decided per upload from the share of non-zeros.
                                        The two layouts add the same non-zero products in a different, fixed order
                                        (see dpsvm_layout_info); results agree to fp32 rounding of the dot products. */

#define DPSVM_FLAG_X_MIRROR 32       /* with DPSVM_FLAG_X_SHARD_ONLY on a multi-GPU run: the upload still brings only the owned
                                        rows, but every rank then copies the OTHER ranks' rows (with their labels and |x|^2) device
                                        to device over NVLink, so that each GPU holds all of X -- what the reference does through the
                                        host (svmTrain.cu:344).  The sweep still reads only the owned shard; the two rows of the
                                        selected pair then come from local memory and the per-iteration exchange shrinks to 48 bytes
                                        per rank (see dpsvm_comm_transport).  Costs 4*N*pitch bytes per GPU; every rank or none. */

int dpsvm_create(const dpsvm_params* params, dpsvm_ctx** out);
int dpsvm_destroy(dpsvm_ctx* ctx);

/* SvmTrain::setup (svmTrain.cu:319-395): copy X and y to the device, alpha = 0, f = -y, |x|^2,
 * cache lines, first selection.  h_x is the FULL N x d row-major matrix and h_y the N labels, as
 * the reference passes them -- unless DPSVM_FLAG_X_SHARD_ONLY, where both hold only the owned
 * rows.  Buffers are borrowed for the duration of the call. */
int dpsvm_upload(dpsvm_ctx* ctx, const float* h_x, const int32_t* h_y);

/* Same, from device memory already holding the owned rows ([shard_rows x pitch], pitch from
 * dpsvm_pitch_for, zero padded) and labels; copied device-to-device. */
int dpsvm_upload_device(dpsvm_ctx* ctx, const float* d_x_shard, int pitch, const int32_t* d_y_shard);

/* alpha = 0, f = -y, iteration counter and cache cleared (a fresh SvmTrain::setup without the copy). */
int dpsvm_reset(dpsvm_ctx* ctx);

/* SvmTrain::train_step1 (svmTrain.cu:469-483): rv = {I_hi, I_lo, b_hi, b_lo} with the indices
 * carried as floats exactly as svmTrain.cu:478-481 does (rv, svmTrain.h:108).  keys, when non-NULL,
 * also receives the two packed 64-bit keys (exact indices beyond 2^24). */
int dpsvm_step1(dpsvm_ctx* ctx, float rv[4], uint64_t keys[2]);

/* SvmTrain::train_step2 (svmTrain.cu:485-497) -> update_f (:190-302): install the two new alpha
 * values and run the fused pass (which also leaves the next selection ready for dpsvm_step1).
 * For rows outside the X range resident on this rank pass the row contents in h_x_hi / h_x_lo
 * (d floats each); NULL means "take it from device X". */
int dpsvm_step2(dpsvm_ctx* ctx, int i_hi, int i_lo, float alpha_hi_new, float alpha_lo_new,
                const float* h_x_hi, const float* h_x_lo);

/* The whole loop of svmTrainMain.cpp:235-310 kept on the device: selection merge, eta
 * (svmTrain.cu:696-714), alpha update and clip (:674-682), cache lookup (:142-156), fused pass,
 * convergence test -- as CUDA graph launches of `chunk` iterations with no host round trip inside.
 * Runs until converged (b_lo <= b_hi + 2 epsilon), or the TOTAL iteration count reaches
 * params.max_iter, or `max_new_iter` more iterations were done (<=0: no extra bound).  Do-while
 * semantics: the iteration that detects convergence still applies its update. */
typedef struct {
    int num_iter;          /* total SMO iterations so far */
    int converged;         /* 1 once b_lo <= b_hi + 2*epsilon was observed */
    float b_hi, b_lo;      /* of the last executed iteration (svmTrainMain.cpp:276-277) */
    float b;               /* (b_lo + b_hi) / 2, svmTrainMain.cpp:329 */
    long long cache_hits;  /* kernel-row cache hits / lookups (2 lookups per iteration) */
    long long cache_lookups;
    float last_chunk_ms;   /* device time of the last graph launch (CUDA events) */
    int i_hi, i_lo;        /* pair of the last executed iteration (global indices) */
    float eta;             /* its eta (svmTrainMain.cpp:282) */
} dpsvm_train_status;

int dpsvm_train(dpsvm_ctx* ctx, int max_new_iter, dpsvm_train_status* status);

/* Run exactly `iters` iterations asynchronously on the context's stream (no host sync, no early
 * exit on the host side; iterations after convergence are no-ops on the device).  For timing. */
int dpsvm_train_async(dpsvm_ctx* ctx, int iters);
/* How the loop runs on this context: *resident = 1 for the persistent cooperative kernel (one launch =
 * up to *chunk iterations), 0 for the CUDA graph of [control, pass] pairs (one graph launch = *chunk
 * iterations).  Either output may be NULL. */
int dpsvm_loop_info(dpsvm_ctx* ctx, int* resident, int* chunk);
int dpsvm_sync(dpsvm_ctx* ctx, dpsvm_train_status* status);
void* dpsvm_stream(dpsvm_ctx* ctx);
/* What the per-iteration sweep reads on this context (decided by the last upload).  No reference counterpart:
 * the reference multiplies dense rows whatever the data (svmTrain.cu:222,247), although the data sets it was
 * measured on are 80-90 % zeros (convert_mnist_to_odd_even.py, convert_adult.py).
 *   *lanes        0: dense rows, dot products in the order of dpsvm_row_norms;
 *                 4/8/16/32: the sparse stream -- the non-zeros of a row in ascending column order are dealt
 *                 round-robin to that many lanes, each lane adds its products with sequential FMAs, an
 *                 xor-butterfly sums the lanes; |x|^2 is summed in the same order.
 *   *nnz          non-zeros of the owned rows (-1 when they were not counted: dense layout forced or rows too wide)
 *   *sweep_bytes  bytes of X (or of its sparse stream) one sweep reads, padding included.
 * Outputs may be NULL. */
int dpsvm_layout_info(dpsvm_ctx* ctx, int* lanes, long long* nnz, long long* sweep_bytes);

/* Copy state back: global alpha of the owned rows (the caller-side alpha[] of
 * svmTrainMain.cpp:223,298-299), f, and |x|^2. Each may be NULL. */
int dpsvm_download(dpsvm_ctx* ctx, float* h_alpha_shard, float* h_f_shard, float* h_xsq_shard);
/* Device pointers of the owned state, for zero-copy inspection by tests (do not free). */
int dpsvm_device_state(dpsvm_ctx* ctx, const float** d_x, int* pitch, const float** d_xsq,
                       const int32_t** d_y, float** d_alpha, float** d_f);

/* SvmTrain::lookup_cache (svmTrain.cu:142-156) over myCache (cache.cu:62-105): returns the device
 * line holding <x_j, x_key> for the owned rows, and whether it was already present.  The LRU
 * bookkeeping lives on the device. */
int dpsvm_cache_lookup(dpsvm_ctx* ctx, int key, float** d_line, int* hit);

/* SvmTrain::test_setup / aggregate_sv / get_train_accuracy (svmTrain.cu:569-665): support-vector
 * compaction (alpha > 0) and sign(sum_j y_j alpha_j K(x_j, t_i) - b) over a host test set.
 * subtract_b = 1 is the GPU reference's convention (svmTrain.cu:652); 0 is seq.cpp's /
 * seq_test.cpp's (seq.cpp:337, seq_test.cpp:197).  h_labels (optional) is compared for accuracy.
 * Outputs may be NULL. */
int dpsvm_predict(dpsvm_ctx* ctx, const float* h_alpha_full, const int32_t* h_y_full, float b,
                  int subtract_b, const float* h_t, int n_test, const int32_t* h_labels,
                  float* h_dual, int32_t* h_pred, int* num_sv, float* accuracy);

/* ------------------------------------------------------------------------------------------------
 * One SMO iteration that allocates nothing
 * ---------------------------------------------------------------------------------------------- */

/* This entry point has no counterpart in the reference.
 *
 * It exists for callers that cannot call cudaMalloc when they need to advance the solver: a plugin
 * inside a framework whose own allocator already owns the device, a thread that must not trigger
 * the device-wide synchronisation cudaMalloc / cudaFree imply, or a program running under a memory
 * budget that was fixed at start-up.  It runs exactly the two kernels one iteration of dpsvm_train
 * runs (control step, then the fused pass) on arrays the CALLER owns, and keeps the small work area
 * those kernels need (the two selected rows and the iteration state) in a fixed array inside the
 * library image instead of in cudaMalloc'ed memory.  The call itself performs no cudaMalloc,
 * cudaFree or any other allocating routine.
 *
 * Limits and rules
 *   - pitch <= DPSVM_NOALLOC_MAX_PITCH floats; wider rows are rejected with DPSVM_ERR_LIMIT.  The
 *     work area is 33 024 bytes per GPU, held by every GPU that loads the library until exit.
 *   - no kernel-row cache (every iteration recomputes both kernel rows).
 *   - d_keys must hold the current selection on entry (dpsvm_select, or the previous call's
 *     output) and receives the next one; d_alpha and d_f are updated in place; all rows referenced
 *     by the keys must lie in [index_base, index_base + n).
 *   - the first call on each device also looks up the work area's device address; make that first
 *     call before starting any stream capture on that device.
 *   - every call on a device shares the one work area, which keeps the previous call's contents
 *     (the call clears what it needs cleared).  Two calls that overlap on the GPU -- different
 *     streams, or different host threads -- race on it; serialise them on one stream.
 *   - asynchronous on `stream`; outputs are valid once the stream has drained.  */
#define DPSVM_NOALLOC_MAX_PITCH 4096

typedef struct {
    int32_t i_hi, i_lo;      /* the pair this iteration worked on (global indices) */
    float b_hi, b_lo;        /* svmTrainMain.cpp:276-277 */
    float eta;               /* svmTrainMain.cpp:282 */
    int32_t converged;       /* 1 if b_lo <= b_hi + 2*epsilon was observed (the update is still applied) */
} dpsvm_iteration_info;

typedef struct {
    const float* d_x;        /* [n x pitch] rows, zero padded */
    int n, d, pitch;
    const float* d_xsq;      /* [n] from dpsvm_row_norms */
    const int32_t* d_y;      /* [n] */
    float* d_alpha;          /* [n] in/out */
    float* d_f;              /* [n] in/out */
    uint64_t* d_keys;        /* [2] in/out */
    float c, gamma, epsilon;
    int index_base;          /* global index of row 0 */
    dpsvm_iteration_info* d_info;   /* optional DEVICE pointer, filled on the stream */
} dpsvm_iteration_args;

int dpsvm_smo_iteration_noalloc(const dpsvm_iteration_args* args, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Multi-GPU (one process per GPU; replaces the OpenMPI Allgather of svmTrainMain.cpp:244)
 * ---------------------------------------------------------------------------------------------- */

#define DPSVM_COMM_ID_BYTES 128
/* rank 0 creates the id (ncclGetUniqueId) and distributes the bytes by any out-of-band means
 * (torch.distributed broadcast, a file, MPI ...); every rank then joins. */
int dpsvm_comm_unique_id(unsigned char id[DPSVM_COMM_ID_BYTES]);
int dpsvm_comm_init(dpsvm_ctx* ctx, const unsigned char id[DPSVM_COMM_ID_BYTES], int rank, int world);
/* Which transport the per-iteration exchange uses: 0 = none (single GPU), 1 = NCCL all-gather
 * captured in the graph, 2 = peer memory with the candidate rows in the package (the last CTA of the pass, or CTA 0
 * of the resident loop, stores keys, scalars and both candidate rows into every rank's CUDA-IPC-mapped mailbox over
 * NVLink and raises a flag), 3 = peer memory with X mirrored (DPSVM_FLAG_X_MIRROR): only the packed keys and the two
 * alphas travel, as six 8-byte words per rank that each carry their own sequence number -- no fence, no flag, one
 * one-way trip.  Peer memory is the default and is chosen only if every rank could map every peer; DPSVM_EXCHANGE=nccl
 * on any rank forces 1 on all.
 * Every device-side wait for another rank (or another CTA) is bounded by DPSVM_TIMEOUT_MS (default 10 000): ranks
 * must enter dpsvm_train / dpsvm_train_async within that window of each other (put a host-side barrier before the
 * call if their setup times differ); on expiry the loop stops and the call returns DPSVM_ERR_COMM. */
int dpsvm_comm_transport(dpsvm_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* DPSVM_CUDA_H */
