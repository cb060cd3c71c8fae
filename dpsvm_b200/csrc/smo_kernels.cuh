// smo_kernels.cuh -- kernel argument blocks and launcher declarations (internal).
#pragma once

#include "smo_common.cuh"

namespace dpsvm {

constexpr int kPassThreads = 256;
constexpr int kRowsInFlight = 2;   // rows per lane group whose loads are issued before any FMA
constexpr int kCtlThreads = 256;
constexpr int kCacheSmemLines = 3584;   // caches up to this many lines keep their recency lists in smem during the control step

struct PushDev;

// ------------------------------------------------------------------------------------------------
// Sparse stream of the owned rows (smo_sparse.cu).  X itself stays in HBM as dense rows -- the selected
// pair, the candidate packages and predict read those -- but the per-iteration sweep, which is what
// the loop is bound by, reads this compressed copy instead when the data is sparse enough (MNIST is
// 19 % non-zero, Adult 11 %): value (4 B) + column (2 B) per non-zero instead of 4 B per cell.
//
// Layout ("sliced ELL", slices of 32/G rows, G lanes per row):  a UNIT is 32/G consecutive rows; row
// r of the unit belongs to lane group r.  The non-zeros of a row, in ascending column order, are dealt
// round-robin to the G lanes of its group: entry e goes to lane e % G, step e / G.  A unit has
// T = unit_off[u+1] - unit_off[u] steps: the longest row of the unit, rounded up to a whole number of
// CHUNKS of kSpChunk = 8 steps (at least one); shorter rows are padded with value 0 / column 0, which
// leave an FMA accumulator unchanged.  Chunk q of the stream is, for each of the 32 lanes, 8 consecutive
// values  val[(q * 32 + lane) * 8 + i]  -- one 256-bit load per lane, 1 KB per warp -- and 8 16-bit byte
// offsets  col[(q * 32 + lane) * 8 + i] = 8 * column  into the interleaved (x_hi, x_lo) float2 row in
// shared memory -- one 128-bit load per lane.
//
// Dot-product order of the sparse layout (the counterpart of DotGeom): per lane sequential FMAs over
// its entries in step order, then an xor-butterfly over the G lanes.  The row norms of a sparse
// context are summed from the stream in the same order (sparse_row_norms_kernel), so <x_i,x_i> == |x_i|^2
// stays bit-exact.
// ------------------------------------------------------------------------------------------------
constexpr int kSpChunk = 8;             // steps per chunk = alignment of a unit's step count
constexpr int kSpRing = 4;              // register chunks per warp: one being consumed, three in flight (what an HBM-sized stream needs)
constexpr unsigned int kSpMaxStepsO = 1u << 27;   // steps of a stream the resident loop's 32-bit lane offsets (sp_load_at) can address: 32 bytes of column distances per step
struct SparseDev {
    const float* val;           // null: dense sweep
    const unsigned short* col;  // 16-bit format: byte offsets 8 * column
    const unsigned char* col8;  // 8-bit format (null: not in use): per lane, the distance in columns from the lane's previous
                                // entry (from column 0 at the first step of a unit; padding entries carry 0).  Same chunk
                                // geometry, col8[(q * 32 + lane) * 8 + i]: one 64-bit load per lane and chunk, 160 instead of
                                // 192 bytes per step.  Built when no distance exceeds 255 (always so for d <= 256).
    int step_bytes;             // 192 or 160
    const unsigned int* unit_off;   // [n_units + 1] first step of each unit
    const int* cta_first;       // [grid + 1] first unit of each CTA of the resident loop (balanced by steps)
    // resident loop: which units of its CTA each warp sweeps.  Units differ in length (whole chunks of 8 steps), so the
    // host deals them longest-first to the least loaded of the CTA's 16 warp slots; slot s sweeps
    // perm[cta][j * 16 + s], j = 0 .. cnt[cta][s] - 1 (local unit numbers, ascending), slot 0 carrying the most.
    const unsigned short* perm; // [grid][perm_stride]
    const unsigned short* cnt;  // [grid][16]
    int perm_stride;            // 16 * the largest cnt
    int lanes;                  // G: 4, 8, 16 or 32
    int n_units;
    int grid;                   // CTAs cta_first was computed for
    int max_cta_units;          // largest cta_first[i+1] - cta_first[i]
    unsigned int total_steps;
    unsigned int pin_steps;     // filled by the launcher: steps per CTA loaded L2::evict_last
};

// Arguments of the fused pass kernel (by value; small enough for the parameter space).
struct PassDev {
    const float4* x4;       // [n x pitch4] owned rows
    int n, pitch4;
    const float* xsq;       // [n]
    const int* y;           // [n]
    const float* alpha;     // [n] (already holds this iteration's new values)
    float* f;               // [n]
    const float4* x_hi4;    // [pitch4]
    const float4* x_lo4;
    float gamma, c;
    int index_base;
    unsigned long long* keys;   // [2]
    int reverse;                // sweep units back to front (set by the launcher from `variant`)
    int pin_units;              // units below this index are loaded L2::evict_last (HINT kernels)
    int early_load;             // issue the first unit's X loads before reading the pass scalars
    int pdl;                    // launch with programmatic dependent launch (host-side flag)
    PassScalars imm;            // used when dev == nullptr
    const PassScalars* dev;     // device-resident scalars written by the control kernel
    const PushDev* push;        // multi-GPU peer-memory exchange: last CTA publishes the candidates (else null)
    SparseDev sp;               // sp.val != null: sweep the sparse stream instead of x4
};

// Kernel-row cache (device-resident LRU).  line_of_key is a direct index over the GLOBAL row
// number (a perfect hash; 4 N bytes), prev/next a doubly linked recency list over the lines.
enum { kCacheSize = 0, kCacheHead = 1, kCacheTail = 2, kCacheMetaInts = 4 };
struct CacheDev {
    int lines;              // capacity; 0 = cache disabled
    float* data;            // [lines x line_stride]
    size_t line_stride;     // floats per line (>= owned rows)
    int* line_of_key;       // [N global] -> line or -1
    int* key_of_line;       // [lines]
    int* prev;              // [lines] towards LRU
    int* next;              // [lines] towards MRU
    int* meta;              // [kCacheMetaInts]: size, head (LRU), tail (MRU)
};

// Candidate package one rank publishes per iteration in the multi-GPU loop: its local selection
// (packed keys), the scalars of the two candidate rows and the rows themselves.  Replaces the
// 4-float rv[] of the reference's Allgather (svmTrainMain.cpp:244) -- the rows ride along so that no
// second, data-dependent broadcast is needed.
struct CandHeader {
    unsigned long long key_up, key_low;
    float a_hi, a_lo;        // alpha of the local candidates
    int y_hi, y_lo;
    float xsq_hi, xsq_lo;
    int pad[6];
};
static_assert(sizeof(CandHeader) == 64, "CandHeader is 64 bytes");
inline size_t cand_package_bytes(int pitch) { return sizeof(CandHeader) + 2 * sizeof(float) * (size_t)pitch; }

struct PackDev {
    unsigned long long* keys;   // [2] local selection; reset to empty by the pack kernel
    unsigned long long* keys_copy;   // [2] receives what was packed (train_step1 reads it)
    const float* x;             // resident rows (global row res0 first)
    int pitch, res0, own0;      // own0 = first owned global row (fallback candidate when a set is empty)
    const float* xsq;
    const int* y;
    const float* alpha;
    unsigned char* send;        // one package
};

// Peer-memory exchange (multi-GPU, one process per GPU): every rank owns a mailbox
//   mail[2 slots][world][package]  and  flag[2 slots][world]
// mapped into every peer with CUDA IPC.  The last CTA of a pass to finish writes this rank's package
// straight into slot (seq & 1) of EVERY rank's mailbox over NVLink (plain stores to mapped peer
// memory), fences, and raises its flag there to seq + 1; the next control kernel waits for all
// `world` flags of that slot.  seq = DevState::num_iter, which every rank advances identically.
constexpr int kMaxPeers = 16;

// Flagged-word exchange ("LL": every 8-byte word carries 4 bytes of data and the sequence number it belongs to, so
// a word is valid the moment it is seen -- no fence, no separate flag, ONE one-way trip over NVLink).  Used when
// every rank holds all of X (DPSVM_FLAG_X_MIRROR): only the candidates travel, 6 words per rank and iteration:
//   word 0/1/2 = ord(f_up), index_up, alpha_up      word 3/4/5 = ord(-f_low), index_low, alpha_low
// Inbox of a rank: in[2 parities][kMaxPeers sources][8 words]; package `seq` goes to parity seq & 1 with flag seq + 1.
constexpr int kLLWords = 8;             // words per (parity, source) slot: 6 used, 64-byte slots
constexpr size_t kLLBytes = 2 * (size_t)kMaxPeers * kLLWords * sizeof(unsigned long long);
struct LLDev {
    unsigned long long* in;                     // this rank's inbox (null: the flagged-word exchange is not in use)
    unsigned long long* peer_in[kMaxPeers];     // rank q's inbox as mapped here (q == rank: local)
};

struct PushDev {
    unsigned int* counter;          // CTAs of the current pass that have finished
    unsigned long long* keys;       // [2] local selection; reset to empty after packing
    unsigned long long* keys_copy;  // [2] the last published local selection (what train_step1 reports)
    const float* x;                 // resident rows (global row res0 first)
    int pitch, res0, own0;
    const float* xsq;
    const int* y;
    const float* alpha;
    const DevState* st;             // for the sequence number
    int world, rank;
    size_t package;                 // bytes per package
    unsigned char* peer_mail[kMaxPeers];   // base of rank q's mailbox as mapped here (q == rank: local)
    unsigned int* peer_flag[kMaxPeers];
    LLDev ll;                       // ll.in != null: candidates travel as flagged words, rows come from the local mirror of X
    unsigned long long timeout_ns;
};

struct CtlDev {
    DevState* st;
    unsigned long long* keys;   // [2] local selection result of the previous pass
    const float* x;             // resident rows, global row res0 first
    int pitch;
    int res0;                   // global index of resident row 0 (x, xsq, y, alpha all use it)
    const float* xsq;
    const int* y;
    float* alpha;
    float* sel_rows;            // [2 x pitch] staging for the pass
    float c, gamma, epsilon;
    int max_iter;
    CacheDev cache;
    int directed;               // host-directed update (SvmTrain::train_step2)
    int dir_i_hi, dir_i_lo;
    float dir_a_hi, dir_a_lo;
    int pdl;                    // launch with programmatic dependent launch (host-side flag)
    // multi-GPU: world > 1 -> the selection comes from `world` gathered candidate packages
    int world;
    const unsigned char* cand;  // [world x cand_stride]
    size_t cand_stride;
    int own0, own_n;            // owned global row range: only the owner installs a new alpha
    const unsigned int* flags;  // peer-memory exchange: [2 slots][world] arrival flags (null: cand is already complete)
    const float* ext_row_hi;    // optional externally supplied rows (device pointers)
    const float* ext_row_lo;
    const unsigned long long* ll_in;   // flagged-word exchange: this rank's inbox (rows, y, |x|^2 then come from the local mirror)
    unsigned long long timeout_ns;     // bound of the wait for the peers' candidates
};

// ---- resident loop (smo_resident.cu): persistent cooperative kernel, one CTA per SM ----
constexpr int kResThreads = 512;
constexpr int kResCacheLines = 64;      // largest kernel-row cache whose LRU table is replicated per CTA

// What a CTA publishes after its sweep: its block-best packed keys and the alpha of those two rows, as six flagged
// 8-byte words (data in the low half, 1 + the global iteration number in the high half; see LLDev).  A word is valid
// the moment a reader sees the right number in it, so the writer needs no fence and the readers no barrier: the
// slot array IS the grid barrier.  w[0],w[1] = key_up (high, low word)  w[2],w[3] = key_low  w[4],w[5] = alpha_up, alpha_low.
struct ResSlot {
    unsigned long long w[8];
};
static_assert(sizeof(ResSlot) == 64, "ResSlot is 64 bytes");

struct ResidentDev {
    const float4* x4;           // [n x pitch4] owned rows
    int n, pitch4;
    const float* xsq;           // [n]
    const int* y;               // [n]
    float* alpha;               // [n]
    float* f;                   // [n]
    float gamma, c, epsilon;
    int max_iter, index_base;
    int empty_row;              // global row an empty I-set decodes to: the first RESIDENT row (row 0 when X is held from row 0)
    DevState* st;
    unsigned long long* keys;   // [2] selection in (launch) / out (exit)
    ResSlot* slots;             // [2][grid]
    unsigned int* barrier;      // arrival counter, zero at launch
    int chunk;                  // at most this many iterations per launch
    // kernel-row cache (cache_lines <= kResCacheLines): replicated LRU table, lines in global memory
    int cache_lines;
    float* cache_data;
    size_t line_stride;
    int* lru_key;               // [cache_lines]
    int* lru_stamp;             // [cache_lines]
    int* lru_meta;              // [2] size, clock
    int cache_direct;           // 1: one line per global row (cache_lines >= N): line = key, nothing is ever evicted
    int* cache_stamp;           // [N global] 0 = line not filled, else 1 + the iteration that filled it
    // filled by the launcher
    int res_units;              // units of X each CTA keeps in shared memory
    int pin_units;              // streamed units per CTA loaded L2::evict_last
    int stream_evict_first;
    int parity0;                // sweep direction of the launch's first iteration
    int prefetch_mode;          // next sweep's first loads: 0 issued right after the slot store, 1 after the slots have been read, 2 not at all
    // multi-GPU (world > 1): the pair comes from the peer-memory mailbox (see PushDev); CTA 0 publishes
    int world, rank;
    const PushDev* push;
    const unsigned char* mail;  // local mailbox [2][world][package]
    const unsigned int* flags;  // local arrival flags [2][world]
    size_t package;
    LLDev ll;                   // ll.in != null: flagged-word exchange (X mirrored on every rank)
    unsigned long long timeout_ns;   // bound of every spin (grid barrier, peers' candidates)
    unsigned long long* trace;  // optional [grid][4] globaltimer stamps of iteration trace_iter (DPSVM_TRACE)
    int trace_iter;
    SparseDev sp;               // sp.val != null: sweep the sparse stream instead of x4
};

bool resident_supported(int pitch4, int cache_lines, int n_total);
int resident_grid();            // CTAs of the resident loop on the current device (one per SM)
cudaError_t launch_resident_sparse(const ResidentDev& a, int variant, cudaStream_t s);
cudaError_t launch_resident(const ResidentDev& a, int variant, cudaStream_t s);

int device_sm_count();
cudaError_t launch_pass(const PassDev& a, int variant, cudaStream_t s);
cudaError_t launch_row_norms(const float* x, int n, int pitch, float* out, cudaStream_t s);
cudaError_t launch_init_state(const int* y, int n, float* alpha, float* f, cudaStream_t s);
cudaError_t launch_select(const float* alpha, const int* y, const float* f, int n, float c, int index_base,
                          unsigned long long* keys, cudaStream_t s);
cudaError_t launch_control(const CtlDev& a, cudaStream_t s);
cudaError_t launch_pack(const PackDev& a, cudaStream_t s);
cudaError_t launch_push(const PushDev* d_push, cudaStream_t s);   // publish the current local selection (initial package)
cudaError_t launch_cache_lookup(const CacheDev& c, int key, int* d_out, cudaStream_t s);
cudaError_t launch_pass_sparse(const PassDev& a, int variant, cudaStream_t s);

// smo_sparse.cu -- building the sparse stream from the dense rows (setup time)
constexpr int kSparseMaxPitch = 1024;   // floats per row the sparse sweep supports (one 8 KB float2 row in smem)
cudaError_t launch_sparse_count(const float* x, int n, int pitch, int* row_nnz, unsigned long long* total, cudaStream_t s);
cudaError_t launch_sparse_unit_steps(const int* row_nnz, int n, int lanes, unsigned int* unit_steps, int n_units, cudaStream_t s);
cudaError_t launch_sparse_scan(const unsigned int* unit_steps, int n_units, unsigned int* unit_off, cudaStream_t s);
cudaError_t launch_sparse_fill(const float* x, int n, int pitch, int lanes, const unsigned int* unit_off,
                               float* val, unsigned short* col, cudaStream_t s);
cudaError_t launch_sparse_col8(const float* val, const unsigned short* col, const unsigned int* unit_off, int n_units, unsigned char* col8,
                               int* too_far, cudaStream_t s);
cudaError_t launch_sparse_row_norms(const float* val, const unsigned int* unit_off, int n_units, int n, int lanes, float* out, cudaStream_t s);

}  // namespace dpsvm
