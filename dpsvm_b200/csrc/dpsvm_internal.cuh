// dpsvm_internal.cuh -- the context structure shared by the library's translation units (internal).
#pragma once

#include "smo_kernels.cuh"

struct dpsvm_ctx {
    dpsvm_params p{};
    int device = 0;
    int pitch = 0;               // floats per row, multiple of 4
    int res0 = 0, res_n = 0;     // global row range resident on this GPU (X, xsq, y, alpha)
    int chunk = 128;             // SMO iterations per CUDA-graph launch
    int variant = 0;             // pass kernel variant word (DPSVM_VARIANT): bit0 reverse, bit1 L2 hints, >>8 pinned MB
    int serpentine = 0;          // alternate the sweep direction between consecutive iterations
    int pdl = 1;                 // programmatic dependent launch between control and pass kernels
    bool uploaded = false;
    int target_iter = 0;         // host view of how many iterations have been requested so far
    int stop_at = 0;

    cudaStream_t stream = nullptr;
    cudaEvent_t ev_t0 = nullptr, ev_t1 = nullptr;
    cudaEvent_t ev_slot[2] = {nullptr, nullptr};

    float* d_x = nullptr;        // [res_n x pitch]
    float* d_xsq = nullptr;      // [res_n]
    int* d_y = nullptr;          // [res_n]
    float* d_alpha = nullptr;    // [res_n]
    float* d_f = nullptr;        // [shard_rows]
    float* d_sel_rows = nullptr; // [2 x pitch] rows of the current pair, read by the pass
    float* d_ext_rows = nullptr; // [2 x pitch] rows supplied by the host (step2)
    unsigned long long* d_keys = nullptr;   // [2]
    unsigned long long* d_keys_copy = nullptr;   // [2] multi-GPU: last published local selection (train_step1)
    dpsvm::DevState* d_st = nullptr;
    dpsvm::DevState* h_st = nullptr;        // pinned, 2 slots
    unsigned long long* h_keys = nullptr;   // pinned
    int* h_small = nullptr;                 // pinned scratch

    dpsvm::CacheDev cache{};
    int* d_cache_stamp = nullptr;   // resident loop, cache with a line per row: [N] 0 = empty, else 1 + filling iteration

    // sparse stream of the owned rows (smo_sparse.cu); sp.val == null: the sweep reads the dense rows
    dpsvm::SparseDev sp{};
    float* d_sp_val = nullptr;
    unsigned short* d_sp_col = nullptr;
    unsigned char* d_sp_col8 = nullptr;      // 8-bit column distances (+ a flag word at the end)
    size_t sp_col8_cap = 0;
    unsigned int* d_sp_unit_off = nullptr;
    int* d_sp_cta_first = nullptr;
    unsigned short* d_sp_perm = nullptr;     // SparseDev::perm followed by SparseDev::cnt
    size_t sp_perm_cap = 0;
    int* d_sp_row_nnz = nullptr;             // build scratch, kept for the next upload
    unsigned int* d_sp_unit_steps = nullptr;
    unsigned long long* d_sp_total = nullptr;
    size_t sp_val_cap = 0, sp_col_cap = 0, sp_units_cap = 0;   // bytes allocated (buffers only grow)
    long long sp_nnz = -1;       // non-zeros of the owned rows (-1: not counted, dense layout forced)

    // resident loop (smo_resident.cu)
    int loop_resident = 0;       // 1: the persistent cooperative kernel runs the iterations
    dpsvm::ResSlot* d_slots = nullptr;
    unsigned int* d_barrier = nullptr;
    int* d_lru = nullptr;        // [kResCacheLines keys][kResCacheLines stamps][2 meta]
    int cache_path = 0;          // which loop last owned the kernel-row cache: 0 none, 1 graph/split, 2 resident
    long long launched_iters = 0;
    unsigned long long* d_trace = nullptr;   // DPSVM_TRACE: per-CTA phase stamps of one iteration (debugging aid)

    cudaGraph_t graph = nullptr;
    cudaGraphExec_t gexec = nullptr;

    // multi-GPU
    int rank = 0, world = 1;
    void* comm = nullptr;        // opaque (comm.cu)
    const dpsvm::PushDev* d_push = nullptr;   // peer-memory exchange descriptor (device), null otherwise
    int exchange_p2p = 0;
    unsigned long long timeout_ns = 10000000000ull;   // bound of every device-side wait for other CTAs / ranks (DPSVM_TIMEOUT_MS)

    // predict scratch (predict.cu)
    void* predict = nullptr;
};

namespace dpsvm {

int fail(int code, const char* fmt, ...);
PassDev make_pass_args(dpsvm_ctx* c);
CtlDev make_ctl_args(dpsvm_ctx* c);

// cache_host.cu
int cache_lookup_host(dpsvm_ctx* c, int key, int* line, int* hit);
// smo_sparse.cu: layout decision, sparse stream, row norms in the matching order (called by every upload)
int sparse_setup(dpsvm_ctx* c);
void sparse_release(dpsvm_ctx* c);
// predict.cu
void predict_free(dpsvm_ctx* c);
// comm.cu
void comm_free(dpsvm_ctx* c);
int comm_launch_iteration(dpsvm_ctx* c, cudaStream_t s, int variant);
int comm_after_select(dpsvm_ctx* c);
int comm_mirror(dpsvm_ctx* c);        // DPSVM_FLAG_X_MIRROR: copy the other ranks' rows, labels and |x|^2 in (collective)
bool comm_fill_resident(dpsvm_ctx* c, ResidentDev* a);   // false: this exchange cannot run inside the resident loop

}  // namespace dpsvm
