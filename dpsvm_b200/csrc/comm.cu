// comm.cu -- multi-GPU candidate exchange (one process per GPU).
//
// Replaces the per-iteration OpenMPI Allgather of the reference (svmTrainMain.cpp:244: 4 floats per
// rank over TCP, indices carried as floats) and the host merge that follows it (:255-274).
// Every rank publishes, once per iteration, a "candidate package": its packed local selection keys,
// the scalars of its two candidate rows and the rows themselves -- so no second, data-dependent
// broadcast of the selected feature vectors is needed (its root would only be known on the device).
// Every rank then runs the identical merge + eta + alpha update (as the reference does redundantly
// on every host, svmTrainMain.cpp:279-299); the owner of a row stores its new alpha.
//
// Two transports:
//   p2p  (default)  fused with the pass: the last CTA of the pass to finish stores the package
//                   straight into every rank's mailbox (CUDA-IPC-mapped peer memory, NVLink /
//                   NVSwitch) and raises a flag; the next control kernel waits on the flags.
//                   Per iteration: control, pass -- two kernels, no collective call.
//   nccl            pack kernel -> ncclAllGather -> control -> pass, all captured in the graph.
//                   (DPSVM_EXCHANGE=nccl, or automatically when peer mapping is unavailable.)
// NCCL is always used for the bootstrap (IPC handle exchange, barriers).  It is bound at run time
// with dlopen("libnccl.so.2"): no link-time dependency, single-GPU users never load it, and inside a
// PyTorch process the already-loaded NCCL is reused.
#include <dlfcn.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

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
    ncclResult_t (*Broadcast)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
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
    a.Broadcast = reinterpret_cast<decltype(a.Broadcast)>(dlsym(h, "ncclBroadcast"));
    a.GroupStart = reinterpret_cast<decltype(a.GroupStart)>(dlsym(h, "ncclGroupStart"));
    a.GroupEnd = reinterpret_cast<decltype(a.GroupEnd)>(dlsym(h, "ncclGroupEnd"));
    a.GetErrorString = reinterpret_cast<decltype(a.GetErrorString)>(dlsym(h, "ncclGetErrorString"));
    if (!a.GetUniqueId || !a.CommInitRank || !a.CommDestroy || !a.AllGather || !a.Broadcast || !a.GroupStart || !a.GroupEnd || !a.GetErrorString)
        return fail(DPSVM_ERR_COMM, "NCCL library lacks a required symbol");
    g_nccl = a;
    return DPSVM_OK;
}

constexpr size_t kFlagBytes = 1024;      // flags [2][world] u32 at the head of the mailbox allocation
// then the flagged-word inbox (kLLBytes, see LLDev), then the mailbox proper [2][world][package]
constexpr size_t kMailOffset = kFlagBytes + kLLBytes;

struct CommState {
    ncclComm_t comm = nullptr;
    size_t package = 0;
    // nccl transport
    unsigned char* d_send = nullptr;     // one package
    unsigned char* d_recv = nullptr;     // world packages
    // p2p transport
    bool p2p = false;
    unsigned char* d_box = nullptr;      // [kFlagBytes flags][kLLBytes flagged words][2][world][package]
    bool ll = false;                     // X mirrored on every rank: candidates travel as flagged words, no rows
    bool mirrored = false;               // the peers' rows of the current upload have been copied in
    std::vector<int> shard_start, shard_rows;   // every rank's owned range (all-gathered at init)
    unsigned int* d_counter = nullptr;
    PushDev* d_push = nullptr;
    PushDev h_push{};                    // host copy of *d_push
    void* opened[kMaxPeers] = {};
    unsigned char* d_tmp = nullptr;      // bootstrap all-gather buffer
};

int nccl_gather_bytes(dpsvm_ctx* c, CommState* s, const void* h_in, void* h_out, size_t bytes)
{
    // bootstrap helper: all-gather `bytes` host bytes per rank through device memory; synchronous
    const size_t total = bytes * (size_t)c->world;
    unsigned char* d = nullptr;
    if (cudaMalloc(&d, total + bytes) != cudaSuccess) return fail(DPSVM_ERR_NOMEM, "comm bootstrap: cudaMalloc");
    cudaError_t e = cudaMemcpyAsync(d + total, h_in, bytes, cudaMemcpyHostToDevice, c->stream);
    ncclResult_t r = 0;
    if (e == cudaSuccess) r = g_nccl.AllGather(d + total, d, bytes, ncclChar, s->comm, c->stream);
    if (e == cudaSuccess && r == 0) e = cudaMemcpyAsync(h_out, d, total, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess && r == 0) e = cudaStreamSynchronize(c->stream);
    cudaFree(d);
    if (r != 0) return fail(DPSVM_ERR_COMM, "comm bootstrap all-gather: %s", g_nccl.GetErrorString(r));
    if (e != cudaSuccess) return fail(DPSVM_ERR_CUDA, "comm bootstrap: %s", cudaGetErrorString(e));
    return DPSVM_OK;
}

// Map every peer's mailbox.  Returns 1 when this rank succeeded, 0 otherwise (never an error: the
// caller falls back to the NCCL transport when any rank reports 0).
int try_setup_p2p(dpsvm_ctx* c, CommState* s)
{
    const int world = c->world;
    if (world > kMaxPeers) return 0;
    const size_t box_bytes = kMailOffset + 2 * (size_t)world * s->package;
    // every rank performs the same collective calls whatever happens locally (a local failure only
    // turns this rank's vote into 0)
    int ok = 1;
    if (cudaMalloc(&s->d_box, box_bytes) != cudaSuccess || cudaMalloc(&s->d_counter, sizeof(unsigned int)) != cudaSuccess ||
        cudaMalloc(&s->d_push, sizeof(PushDev)) != cudaSuccess) {
        cudaGetLastError();
        ok = 0;
    }
    cudaIpcMemHandle_t mine;
    memset(&mine, 0, sizeof(mine));
    if (ok) {
        cudaMemset(s->d_box, 0, box_bytes);
        cudaMemset(s->d_counter, 0, sizeof(unsigned int));
        if (cudaIpcGetMemHandle(&mine, s->d_box) != cudaSuccess) { cudaGetLastError(); ok = 0; }
    }
    std::vector<cudaIpcMemHandle_t> all((size_t)world);
    if (nccl_gather_bytes(c, s, &mine, all.data(), sizeof(mine)) != DPSVM_OK) return 0;
    PushDev h{};
    h.counter = s->d_counter;
    h.keys = c->d_keys;
    h.keys_copy = c->d_keys_copy;
    h.x = c->d_x;
    h.pitch = c->pitch;
    h.res0 = c->res0;
    h.own0 = c->p.shard_start;
    h.xsq = c->d_xsq;
    h.y = c->d_y;
    h.alpha = c->d_alpha;
    h.st = c->d_st;
    h.world = world;
    h.rank = c->rank;
    h.package = s->package;
    h.timeout_ns = c->timeout_ns;
    const bool mirror = (c->p.flags & DPSVM_FLAG_X_MIRROR) != 0;
    for (int q = 0; q < world && ok; ++q) {
        unsigned char* base = nullptr;
        if (q == c->rank) {
            base = s->d_box;
        } else {
            void* p = nullptr;
            if (cudaIpcOpenMemHandle(&p, all[(size_t)q], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                cudaGetLastError();
                ok = 0;
                break;
            }
            s->opened[q] = p;
            base = static_cast<unsigned char*>(p);
        }
        h.peer_flag[q] = reinterpret_cast<unsigned int*>(base);
        h.peer_mail[q] = base + kMailOffset;
        if (mirror) h.ll.peer_in[q] = reinterpret_cast<unsigned long long*>(base + kFlagBytes);
    }
    if (mirror && ok) h.ll.in = reinterpret_cast<unsigned long long*>(s->d_box + kFlagBytes);
    s->h_push = h;
    if (ok && cudaMemcpy(s->d_push, &h, sizeof(h), cudaMemcpyHostToDevice) != cudaSuccess) { cudaGetLastError(); ok = 0; }
    return ok;
}

}  // namespace

void comm_free(dpsvm_ctx* c)
{
    CommState* s = static_cast<CommState*>(c->comm);
    if (!s) return;
    for (int q = 0; q < kMaxPeers; ++q) if (s->opened[q]) cudaIpcCloseMemHandle(s->opened[q]);
    if (s->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(s->comm);
    cudaFree(s->d_send); cudaFree(s->d_recv); cudaFree(s->d_box); cudaFree(s->d_counter); cudaFree(s->d_push);
    delete s;
    c->comm = nullptr;
    c->d_push = nullptr;
}

// DPSVM_FLAG_X_MIRROR: after an upload every rank copies the rows the OTHER ranks own -- with the labels and the |x|^2
// their owners computed (each owner sums in the order of its own sweep) -- device to device over NVLink, one
// broadcast per owner.  The sweep still reads only the owned shard; the copies exist so that the two rows of the
// selected pair, whoever owns them, are one local TMA copy away, as on a single GPU.
int comm_mirror(dpsvm_ctx* c)
{
    CommState* s = static_cast<CommState*>(c->comm);
    if (!s || !s->comm || !(c->p.flags & DPSVM_FLAG_X_MIRROR)) return DPSVM_OK;
    ncclResult_t r = g_nccl.GroupStart();
    for (int q = 0; q < c->world && r == 0; ++q) {
        const size_t row0 = (size_t)(s->shard_start[(size_t)q] - c->res0), rows = (size_t)s->shard_rows[(size_t)q];
        float* x = c->d_x + row0 * (size_t)c->pitch;
        r = g_nccl.Broadcast(x, x, rows * (size_t)c->pitch * sizeof(float), ncclChar, q, s->comm, c->stream);
        if (r == 0) r = g_nccl.Broadcast(c->d_y + row0, c->d_y + row0, rows * sizeof(int), ncclChar, q, s->comm, c->stream);
        if (r == 0) r = g_nccl.Broadcast(c->d_xsq + row0, c->d_xsq + row0, rows * sizeof(float), ncclChar, q, s->comm, c->stream);
    }
    const ncclResult_t r2 = g_nccl.GroupEnd();
    if (r != 0 || r2 != 0) return fail(DPSVM_ERR_COMM, "mirroring X: %s", g_nccl.GetErrorString(r != 0 ? r : r2));
    const cudaError_t e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) return fail(DPSVM_ERR_CUDA, "mirroring X: %s", cudaGetErrorString(e));
    s->mirrored = true;
    return DPSVM_OK;
}

// After the local (re)selection at setup/reset: make the first package available to every rank.
int comm_after_select(dpsvm_ctx* c)
{
    CommState* s = static_cast<CommState*>(c->comm);
    if (!s || !s->p2p) return DPSVM_OK;
    // clear my flags, then a barrier (an all-gather completes only after every rank has contributed,
    // i.e. after every rank's clear is ordered before its contribution), then the initial push
    cudaError_t e = cudaMemsetAsync(s->d_box, 0, kMailOffset, c->stream);      // arrival flags and flagged words
    if (e == cudaSuccess) e = cudaMemsetAsync(s->d_counter, 0, sizeof(unsigned int), c->stream);
    if (e != cudaSuccess) return fail(DPSVM_ERR_CUDA, "comm reset: %s", cudaGetErrorString(e));
    ncclResult_t r = g_nccl.AllGather(s->d_send, s->d_recv, 8, ncclChar, s->comm, c->stream);
    if (r != 0) return fail(DPSVM_ERR_COMM, "comm barrier: %s", g_nccl.GetErrorString(r));
    e = launch_push(s->d_push, c->stream);
    if (e != cudaSuccess) return fail(DPSVM_ERR_CUDA, "initial push: %s", cudaGetErrorString(e));
    e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) return fail(DPSVM_ERR_CUDA, "initial push: %s", cudaGetErrorString(e));
    return DPSVM_OK;
}

bool comm_fill_resident(dpsvm_ctx* c, ResidentDev* a)
{
    CommState* s = static_cast<CommState*>(c->comm);
    if (!s || !s->p2p) return false;
    a->world = c->world;
    a->rank = c->rank;
    a->push = s->d_push;
    if (s->ll) a->ll = s->h_push.ll;
    a->mail = s->d_box + kMailOffset;
    a->flags = reinterpret_cast<const unsigned int*>(s->d_box);
    a->package = s->package;
    return true;
}

// One iteration of the sharded loop.
int comm_launch_iteration(dpsvm_ctx* c, cudaStream_t stream, int variant)
{
    CommState* s = static_cast<CommState*>(c->comm);
    if (!s || !s->comm) return fail(DPSVM_ERR_COMM, "multi-GPU exchange not initialised (dpsvm_comm_init)");
    CtlDev ctl = make_ctl_args(c);
    ctl.world = c->world;
    ctl.cand_stride = s->package;
    ctl.own0 = c->p.shard_start;
    ctl.own_n = c->p.shard_rows;
    cudaError_t e;
    ctl.timeout_ns = c->timeout_ns;
    if (s->p2p) {
        ctl.cand = s->d_box + kMailOffset;
        ctl.flags = reinterpret_cast<const unsigned int*>(s->d_box);
        if (s->ll) ctl.ll_in = s->h_push.ll.in;
    } else {
        PackDev p{};
        p.keys = c->d_keys;
        p.keys_copy = c->d_keys_copy;
        p.x = c->d_x;
        p.pitch = c->pitch;
        p.res0 = c->res0;
        p.own0 = c->p.shard_start;
        p.xsq = c->d_xsq;
        p.y = c->d_y;
        p.alpha = c->d_alpha;
        p.send = s->d_send;
        e = launch_pack(p, stream);
        if (e != cudaSuccess) return fail(DPSVM_ERR_CUDA, "pack launch: %s", cudaGetErrorString(e));
        ncclResult_t r = g_nccl.AllGather(s->d_send, s->d_recv, s->package, ncclChar, s->comm, stream);
        if (r != 0) return fail(DPSVM_ERR_COMM, "ncclAllGather: %s", g_nccl.GetErrorString(r));
        ctl.cand = s->d_recv;
        ctl.flags = nullptr;
        ctl.pdl = 0;                   // its predecessor is an NCCL kernel: plain stream order
    }
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
                                     "(each rank uploads only its rows; candidates travel in the exchange)");
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

    // What every rank must agree on is gathered FIRST, by a collective every rank joins whatever its environment
    // says: the owned row ranges, whether this rank asks for the NCCL transport (DPSVM_EXCHANGE=nccl), its flags.
    const char* ex = getenv("DPSVM_EXCHANGE");
    int mine4[4] = {c->p.shard_start, c->p.shard_rows, (ex && strcmp(ex, "nccl") == 0) ? 1 : 0, c->p.flags & DPSVM_FLAG_X_MIRROR};
    std::vector<int> all4((size_t)world * 4, 0);
    rc = nccl_gather_bytes(c, s, mine4, all4.data(), sizeof(mine4));
    if (rc != DPSVM_OK) return rc;
    bool any_nccl = false, mirror_all = true, mirror_any = false;
    s->shard_start.resize((size_t)world);
    s->shard_rows.resize((size_t)world);
    for (int q = 0; q < world; ++q) {
        s->shard_start[(size_t)q] = all4[(size_t)q * 4];
        s->shard_rows[(size_t)q] = all4[(size_t)q * 4 + 1];
        any_nccl = any_nccl || all4[(size_t)q * 4 + 2] != 0;
        mirror_all = mirror_all && all4[(size_t)q * 4 + 3] != 0;
        mirror_any = mirror_any || all4[(size_t)q * 4 + 3] != 0;
    }
    if (mirror_any != mirror_all) return fail(DPSVM_ERR_INVALID, "dpsvm_comm_init: DPSVM_FLAG_X_MIRROR must be set on every rank or on none");

    // transport: peer memory unless any rank asked otherwise; every rank maps its peers (or skips it) alike, and the
    // outcome is voted on
    int mine = any_nccl ? 0 : try_setup_p2p(c, s);
    std::vector<int> votes((size_t)world, 0);
    rc = nccl_gather_bytes(c, s, &mine, votes.data(), sizeof(int));
    if (rc != DPSVM_OK) return rc;
    bool all_ok = true;
    for (int v : votes) all_ok = all_ok && v != 0;
    s->p2p = all_ok;
    s->ll = all_ok && mirror_all;
    c->d_push = all_ok ? s->d_push : nullptr;
    c->exchange_p2p = all_ok ? 1 : 0;
    if (c->uploaded) {
        rc = comm_mirror(c);
        if (rc != DPSVM_OK) return rc;
        return comm_after_select(c);
    }
    return DPSVM_OK;
}

extern "C" int dpsvm_comm_transport(dpsvm_ctx* c)
{
    if (!c || c->world <= 1 || !c->comm) return 0;
    if (!c->exchange_p2p) return 1;
    return static_cast<CommState*>(c->comm)->ll ? 3 : 2;
}
