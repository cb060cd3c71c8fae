// comm.cu -- multi-GPU candidate exchange (one process per GPU).
//
// Replaces the per-iteration OpenMPI Allgather of the reference (svmTrainMain.cpp:244: 4 floats per
// rank over TCP, indices carried as floats) and the host merge that follows it (:255-274).
// Per iteration, captured in the same CUDA graph as the kernels:
//     pack kernel     -> this rank's package: packed keys + scalars + the two candidate rows
//     ncclAllGather   -> every rank holds all P packages (NVLink / NVSwitch)
//     control kernel  -> identical merge + eta + alpha update on every rank (as the reference does
//                        redundantly on every host, svmTrainMain.cpp:279-299); owners store alpha
//     fused pass      -> over the rank's own shard
// The rows ride in the package, so no second, data-dependent broadcast of the selected feature
// vectors is needed (its root would only be known on the device).
//
// NCCL is bound at run time with dlopen("libnccl.so.2"): the library has no link-time dependency on
// it, single-GPU users never load it, and inside a PyTorch process the already-loaded NCCL is reused.
#include <dlfcn.h>
#include <string.h>

#include "dpsvm_internal.cuh"

namespace dpsvm {

namespace {

typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;
enum { ncclChar = 0 };

struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
};

NcclApi g_nccl;

int load_nccl()
{
    if (g_nccl.lib) return DPSVM_OK;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    void* h = nullptr;
    for (const char* n : names) {
        h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (h) break;
    }
    if (!h) return fail(DPSVM_ERR_COMM, "cannot load NCCL (libnccl.so.2): %s", dlerror());
    NcclApi a;
    a.lib = h;
    a.GetUniqueId = reinterpret_cast<decltype(a.GetUniqueId)>(dlsym(h, "ncclGetUniqueId"));
    a.CommInitRank = reinterpret_cast<decltype(a.CommInitRank)>(dlsym(h, "ncclCommInitRank"));
    a.CommDestroy = reinterpret_cast<decltype(a.CommDestroy)>(dlsym(h, "ncclCommDestroy"));
    a.AllGather = reinterpret_cast<decltype(a.AllGather)>(dlsym(h, "ncclAllGather"));
    a.GetErrorString = reinterpret_cast<decltype(a.GetErrorString)>(dlsym(h, "ncclGetErrorString"));
    if (!a.GetUniqueId || !a.CommInitRank || !a.CommDestroy || !a.AllGather || !a.GetErrorString)
        return fail(DPSVM_ERR_COMM, "NCCL library lacks a required symbol");
    g_nccl = a;
    return DPSVM_OK;
}

struct CommState {
    ncclComm_t comm = nullptr;
    unsigned char* d_send = nullptr;     // one package
    unsigned char* d_recv = nullptr;     // world packages
    size_t package = 0;
};

}  // namespace

void comm_free(dpsvm_ctx* c)
{
    CommState* s = static_cast<CommState*>(c->comm);
    if (!s) return;
    if (s->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(s->comm);
    cudaFree(s->d_send);
    cudaFree(s->d_recv);
    delete s;
    c->comm = nullptr;
}

// One iteration of the sharded loop: pack -> all-gather -> control -> pass.
int comm_launch_iteration(dpsvm_ctx* c, cudaStream_t stream, int variant)
{
    CommState* s = static_cast<CommState*>(c->comm);
    if (!s || !s->comm) return fail(DPSVM_ERR_COMM, "multi-GPU exchange not initialised (dpsvm_comm_init)");
    PackDev p{};
    p.keys = c->d_keys;
    p.x = c->d_x;
    p.pitch = c->pitch;
    p.res0 = c->res0;
    p.own0 = c->p.shard_start;
    p.xsq = c->d_xsq;
    p.y = c->d_y;
    p.alpha = c->d_alpha;
    p.send = s->d_send;
    cudaError_t e = launch_pack(p, stream);
    if (e != cudaSuccess) return fail(DPSVM_ERR_CUDA, "pack launch: %s", cudaGetErrorString(e));
    ncclResult_t r = g_nccl.AllGather(s->d_send, s->d_recv, s->package, ncclChar, s->comm, stream);
    if (r != 0) return fail(DPSVM_ERR_COMM, "ncclAllGather: %s", g_nccl.GetErrorString(r));
    CtlDev ctl = make_ctl_args(c);
    ctl.world = c->world;
    ctl.cand = s->d_recv;
    ctl.cand_stride = s->package;
    ctl.own0 = c->p.shard_start;
    ctl.own_n = c->p.shard_rows;
    ctl.pdl = 0;                       // its predecessor is an NCCL kernel: plain stream order
    e = launch_control(ctl, stream);
    if (e != cudaSuccess) return fail(DPSVM_ERR_CUDA, "control launch: %s", cudaGetErrorString(e));
    e = launch_pass(make_pass_args(c), variant, stream);
    if (e != cudaSuccess) return fail(DPSVM_ERR_CUDA, "pass launch: %s", cudaGetErrorString(e));
    return DPSVM_OK;
}

}  // namespace dpsvm

using namespace dpsvm;

extern "C" int dpsvm_comm_unique_id(unsigned char id[DPSVM_COMM_ID_BYTES])
{
    if (!id) return fail(DPSVM_ERR_INVALID, "dpsvm_comm_unique_id: null argument");
    int rc = load_nccl();
    if (rc != DPSVM_OK) return rc;
    ncclUniqueId u;
    ncclResult_t r = g_nccl.GetUniqueId(&u);
    if (r != 0) return fail(DPSVM_ERR_COMM, "ncclGetUniqueId: %s", g_nccl.GetErrorString(r));
    static_assert(sizeof(u) == DPSVM_COMM_ID_BYTES, "id size");
    memcpy(id, u.internal, DPSVM_COMM_ID_BYTES);
    return DPSVM_OK;
}

extern "C" int dpsvm_comm_init(dpsvm_ctx* c, const unsigned char id[DPSVM_COMM_ID_BYTES], int rank, int world)
{
    if (!c || !id) return fail(DPSVM_ERR_INVALID, "dpsvm_comm_init: null argument");
    if (world < 1 || rank < 0 || rank >= world) return fail(DPSVM_ERR_INVALID, "dpsvm_comm_init: rank %d of %d", rank, world);
    if (c->comm) return fail(DPSVM_ERR_STATE, "dpsvm_comm_init: already initialised");
    if (c->gexec) return fail(DPSVM_ERR_STATE, "dpsvm_comm_init must precede the first training call");
    if (world == 1) { c->world = 1; c->rank = 0; return DPSVM_OK; }
    if (!(c->p.flags & DPSVM_FLAG_X_SHARD_ONLY))
        return fail(DPSVM_ERR_STATE, "dpsvm_comm_init needs a context created with DPSVM_FLAG_X_SHARD_ONLY "
                                     "(each rank keeps only its rows; candidates travel in the exchange)");
    int rc = load_nccl();
    if (rc != DPSVM_OK) return rc;
    cudaError_t e = cudaSetDevice(c->device);
    if (e != cudaSuccess) return fail(DPSVM_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));
    CommState* s = new CommState();
    s->package = cand_package_bytes(c->pitch);
    if (cudaMalloc(&s->d_send, s->package) != cudaSuccess || cudaMalloc(&s->d_recv, s->package * (size_t)world) != cudaSuccess) {
        cudaFree(s->d_send); cudaFree(s->d_recv); delete s;
        return fail(DPSVM_ERR_NOMEM, "dpsvm_comm_init: cudaMalloc failed");
    }
    ncclUniqueId u;
    memcpy(u.internal, id, DPSVM_COMM_ID_BYTES);
    ncclResult_t r = g_nccl.CommInitRank(&s->comm, world, u, rank);
    if (r != 0) {
        cudaFree(s->d_send); cudaFree(s->d_recv); delete s;
        return fail(DPSVM_ERR_COMM, "ncclCommInitRank: %s", g_nccl.GetErrorString(r));
    }
    c->comm = s;
    c->world = world;
    c->rank = rank;
    return DPSVM_OK;
}
