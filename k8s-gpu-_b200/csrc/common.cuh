// common.cuh -- context, device buffers and launch bookkeeping shared by the .cu files of
// libk8sgpu.so.  sm_100a only.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <thread>
#include <time.h>

#include "../../include/k8sgpu.h"

namespace k8sgpu {

constexpr int kMaxTimedRounds = 512;   // scoring-kernel event pairs kept per run
#ifndef K8S_NODE_TILE
#define K8S_NODE_TILE 256
#endif
#ifndef K8S_STAGES
#define K8S_STAGES 3
#endif
constexpr uint32_t kNodeTile = K8S_NODE_TILE;  // nodes per shared-memory tile (32 B each)
constexpr uint32_t kSortTile = 2048;   // pairs per radix-sort block
constexpr int kSortThreads = 256;
constexpr int kScanThreads = 256;
constexpr uint32_t kScanTile = 2048;   // elements per scan block

// A grow-only device allocation.
struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
    template <class T> T *as() const { return reinterpret_cast<T *>(p); }
};

struct SortScratch {
    DevBuf keys_alt, vals_alt, hist, scan_tmp;
};

struct PlaceState {
    uint32_t n = 0, m = 0, n_pad = 0;
    bool uploaded = false;
    // uploaded tables
    DevBuf node_cols;  // 8 x n u32 (column major: free_gpu, free_mcpu, free_mib, cap_gpu, sku, zone, taint, flags)
    DevBuf pod_cols;   // 8 x m u32
    // run state
    DevBuf free_cols;  // 3 x n u32: current free_gpu, free_mcpu, free_mib
    DevBuf nodeq;      // n_pad x 32 B packed node records (rebuilt every round)
    DevBuf podq;       // m x 32 B packed pod records (built at begin)
    DevBuf assign;     // m i32
    DevBuf active[2];  // m u32 each
    DevBuf order[2];   // m u32: the active pods in scoring order (class, rot bucket): the scoring kernel's slot -> pod map
    DevBuf inv[2];     // m u32 each: [0] posof[pod] = position in the active list, [1] slotof[pod] = slot in order[cur]
    DevBuf slots;      // n x 8 u32: slot admission, first proposers of the round per node
    bool want_order = false;   // this run sorts its scoring order by class (the in-library loops do)
    bool order_valid = false;  // order[cur] / inv[cur] describe active[cur]
    DevBuf best_key;   // m u64, indexed by active position
    DevBuf best;       // m i32, indexed by active position
    DevBuf sort_keys, sort_vals;  // m u32
    DevBuf seg;        // segmented-scan block aggregates
    DevBuf flags;      // m u32 (compaction flags / scan output)
    DevBuf counters;   // small u32 block: [0]=error flag, [1]=next active count, [2]=placed
    DevBuf res_min;    // 2 x 128 u32: per-round resource summary of the node table (resmin_note), by round parity
    DevBuf like_seen;  // 65536 u32: which (class, hash) values occur among the pods (scoring order, per run)
    SortScratch sort;
    DevBuf scan_tmp;
    int cur = 0;          // which active[] buffer is current
    uint32_t n_active = 0;
    uint32_t rounds = 0;
    uint64_t evals = 0;
    uint32_t policy_flags = 0, max_rounds = 0;
    uint32_t n_gangs = 0, gang_iters = 0;  // gang placement: table size, round loops run so far
    int score_res = 2;  // eval_pair variant the pod records are laid out for
    uint32_t best_cleared_len = 0;  // best_key[0 .. this) is known to be zero (left so by the admission kernel)
    bool nodeq_stale = true;  // packed node records do not reflect free_cols yet
    DevBuf peer_local;        // exchange buffer other ranks store their proposals into (IPC-exported)
    int32_t *peer_ptr[16] = {};  // every rank's exchange buffer as seen from this GPU (own included)
    int peer_world = 0, peer_rank = -1;
    uint32_t peer_m = 0;         // proposals per half of the exchange buffer
    DevBuf gang_id, gang_min, gang_count, gang_state;  // gang placement: [m], [g], [g], [2g] (rejected, newly)
    uint32_t last_score_slots = 0;        // local slots of the last scoring launch
    uint32_t coop_seq = 0, coop_grid = 0;  // admission launch in flight: mailbox sequence number, CTAs
    int timed = 0;        // scoring launches timed so far in this run
    uint64_t timed_evals = 0;
};

struct ReconcileState {
    uint32_t p = 0, v = 0, n_owner_ids = 0;
    bool uploaded = false, ran = false, has_ordinal = false, want_members = false, want_ordinals = false;
    // One block per direction, so that a caller whose host columns lie back to back gets one copy.
    // in_slab : desired[P] | pool_owner[P] | vm_owner[V] | vm_flags[V] | vm_ordinal[V]   (each padded to 16 B)
    // out_slab: the k8sgpu_reconcile_layout sections, same order and padding as k8sgpu_reconcile_out_size
    DevBuf in_slab, out_slab;
    uint64_t in_off[6] = {}, out_off[12] = {};  // element offsets of the sections (out_off[11] = total)
    uint64_t create_cap = 0;                     // elements reserved for create_ordinal (sum of desired)
    DevBuf lut;       // owner id -> pool row, built at upload
    DevBuf pool_tmp;  // kvict[P] | span_words[P] | word_off[P+1] | need[P] | cand_cnt[P] | cand_off[P+1] | bits[P/32+1]
    DevBuf tot;       // small path: per-CTA member counts, pool major [P][G]
    DevBuf bitmap;    // used-ordinal bits of the pools that scale up
    DevBuf sort_keys, sort_vals;  // general path: (pool, row) pairs
    DevBuf cta_cnt;   // pruning pass: candidates per CTA of the current segment
    DevBuf counters;  // [0]=error flags, [1]=arrival ticket, [2]=pools still short of victims, [3]=candidates, [4..7]=totals
    SortScratch sort;
    DevBuf scan_tmp;
    uint32_t n_victims = 0, n_members = 0, n_creates = 0;
    const uint32_t *members_src = nullptr;  // rows grouped by pool after the last run (member_vm)
    int small_occ = 0;                      // occupancy of the one-launch kernel at small_occ_smem bytes (0 = not queried)
    size_t small_occ_smem = 0;
    bool small_attr_set = false;
    bool prune_attr_set = false;
    uint32_t last_path = 0, last_grid = 0, last_segments = 0, last_candidates = 0;  // what the last run used
};

// Developer / test knobs.  Seeded once from K8SGPU_* environment variables at k8sgpu_create and
// changed only through k8sgpu_set_tuning: nothing reads the environment after that.  -1 = automatic.
struct Tuning {
    int res = -1;        // eval_pair variant 0..4
    int ppt = -1;        // pods per thread 1/2/4
    int splits = -1;     // node splits of the scoring grid
    int coop = -1;       // 1: one-launch admission, 0: kernel chain
    int coop_grid = -1;  // CTAs of the cooperative admission launch
    int rec_path = -1;   // 0: one-launch small-pool kernel, 1: pruned ranking, 2: full stable sort
    int rec_grid = -1;   // CTAs of the cooperative reconcile launches
    int rec_segments = -1;  // row segments of the pruned ranking pass
    int coop_prof = 0, rec_prof = 0;  // print phase times (stderr)
    int shard_upload = -1;  // multi-GPU: 0 = every rank copies the whole pod table from the host
    int shard_admit = -1;   // multi-GPU: 1 = admission sharded by node range, 0 = replicated
    int slot_admit = -1;    // 0: always the sort-based admission
    int order = -1;         // 0: do not sort the scoring order by pod class
    int force_cls = -1;     // 0..11: eval_pair4 class (CLS | TT << 2) for every warp -- timing experiments only, results are wrong; 16: toleration clause always evaluated (results stay right)
    int exchange = -1;      // multi-GPU: 0 = NCCL all-gather, 1 = peer stores, -1 = per round by list length
};

}  // namespace k8sgpu

struct k8sgpu_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev_begin = nullptr, ev_end = nullptr;
    cudaEvent_t ev_score[2 * k8sgpu::kMaxTimedRounds] = {};
    uint32_t *h_pinned = nullptr;  // 64 u32 of pinned, device-mapped host memory for small read-backs
    uint32_t *d_mailbox = nullptr; // device view of h_pinned[32..]: {seq, next active count, error flags}
    uint32_t mail_seq = 0;         // last sequence number handed to a kernel
    int sm_count = 148;
    int coop_blocks_per_sm = 0;  // occupancy of admit_coop_kernel (0 = not queried yet, <0 = unusable)
    char err[512] = {0};
    k8sgpu_stats stats = {};
    k8sgpu::PlaceState place;
    k8sgpu::ReconcileState rec;
    k8sgpu::DevBuf dbg[6];
    k8sgpu::SortScratch dbg_sort;
    k8sgpu::Tuning tune;
    // Placement group: the contexts (one per GPU) that shard one placement between them.  A context
    // made by k8sgpu_create with ndev > 1 leads ndev - 1 more contexts of this process (local[0] is
    // the leader itself); a context joined to other processes by k8sgpu_comm_init_rank is one rank
    // of `world`.  comm is this device's NCCL communicator (ncclComm_t), null while single.
    int world = 1, rank = 0, nlocal = 1;
    k8sgpu_ctx *local[16] = {};
    void *comm = nullptr;
    k8sgpu::DevBuf xchg;      // proposals of all ranks (all-gather target / peer-store target, two halves)
    k8sgpu::DevBuf xsend;     // pod-table staging of a sharded upload (8 columns x padded rows)
    k8sgpu::DevBuf xsmall;    // barrier word, IPC handles
    uint32_t xchg_m = 0;      // proposals per half of xchg
    int32_t *xpeer[16] = {};  // every rank's xchg as seen from this GPU (peer-store exchange)
    bool xpeer_ready = false, xpeer_ipc = false, xpeer_failed = false;
};

namespace k8sgpu {

inline int fail(k8sgpu_ctx *ctx, int code, const char *fmt, ...) {
    if (ctx) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(ctx->err, sizeof ctx->err, fmt, ap);
        va_end(ap);
    }
    return code;
}

#define K8S_CUDA(ctx, expr)                                                                       \
    do {                                                                                          \
        cudaError_t e__ = (expr);                                                                 \
        if (e__ != cudaSuccess)                                                                   \
            return ::k8sgpu::fail((ctx), K8SGPU_E_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #expr, \
                                  cudaGetErrorString(e__));                                       \
    } while (0)

// Every kernel launch goes through this so that stats.kernel_launches is a count, not a guess.
#define K8S_LAUNCH(ctx, kernel, grid, block, smem, stream, ...)                                    \
    do {                                                                                          \
        kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__);                               \
        (ctx)->stats.kernel_launches++;                                                           \
        cudaError_t e__ = cudaGetLastError();                                                     \
        if (e__ != cudaSuccess)                                                                   \
            return ::k8sgpu::fail((ctx), K8SGPU_E_CUDA, "%s:%d launch %s: %s", __FILE__, __LINE__, \
                                  #kernel, cudaGetErrorString(e__));                              \
    } while (0)

#define K8S_ENTER(ctx)                                                                          \
    if (!(ctx)) return K8SGPU_E_ARG;                                                            \
    (ctx)->err[0] = 0;                                                                          \
    {                                                                                           \
        cudaError_t e__ = cudaSetDevice((ctx)->device);                                         \
        if (e__ != cudaSuccess)                                                                 \
            return ::k8sgpu::fail((ctx), K8SGPU_E_CUDA, "cudaSetDevice(%d): %s", (ctx)->device, \
                                  cudaGetErrorString(e__));                                     \
    }

inline void reset_call_stats(k8sgpu_ctx *ctx) {
    ctx->stats.kernel_launches = 0;
    ctx->stats.h2d_bytes = 0;
    ctx->stats.d2h_bytes = 0;
}

// Stream arguments of the C-ABI: a cudaStream_t as is (0 = legacy default stream), or
// K8SGPU_STREAM_OWN for the context's internal stream.
inline cudaStream_t pick_stream(const k8sgpu_ctx *ctx, void *stream) {
    return stream == K8SGPU_STREAM_OWN ? ctx->stream : (cudaStream_t)stream;
}

// Kernels post small results to mapped, pinned host memory ({seq, values...}: the values first,
// a system-wide fence, then the sequence number) so that the host need not wait for the whole
// kernel, or pay a copy plus a stream synchronise, to learn them.  This is the host side: a short
// busy wait (the post usually lands within microseconds), then yields, then sleeps, so that a
// controller sharing its cores with other work is not starved; the stream is queried now and then
// so that a failed or finished kernel cannot leave the caller waiting for ever.
inline int wait_mailbox(k8sgpu_ctx *ctx, volatile uint32_t *mb, uint32_t seq, cudaStream_t s, const char *what) {
    for (uint32_t spins = 0; mb[0] != seq; ++spins) {
        if (spins < 20000u) {
#if defined(__x86_64__) || defined(__i386__)
            __builtin_ia32_pause();
#endif
            if ((spins & 0x3FFu) != 0x3FFu) continue;
        } else if (spins < 24000u) {
            std::this_thread::yield();
        } else {
            timespec ts = {0, 50000};  // 50 us
            nanosleep(&ts, nullptr);
        }
        cudaError_t q = cudaStreamQuery(s);
        if (q == cudaSuccess) {
            if (mb[0] == seq) break;
            return fail(ctx, K8SGPU_E_CUDA, "%s: kernel finished without posting its result", what);
        }
        if (q != cudaErrorNotReady) return fail(ctx, K8SGPU_E_CUDA, "%s: kernel failed: %s", what, cudaGetErrorString(q));
    }
    std::atomic_thread_fence(std::memory_order_acquire);  // the values were written before the sequence number
    return K8SGPU_OK;
}

inline uint32_t ceil_div(uint32_t a, uint32_t b) { return (a + b - 1) / b; }
inline int key_bits(uint32_t max_key) {  // bits needed to hold values 0..max_key
    int b = 1;
    while (b < 32 && (max_key >> b)) ++b;
    return b;
}

// place.cu (host side of the round loop; group.cu drives them across GPUs)
int check_tables(k8sgpu_ctx *ctx, const k8sgpu_node_table *nodes, const k8sgpu_pod_table *pods);
struct PodShare {  // one rank's share of a pod-table upload (see place_upload)
    uint32_t lo, hi, stride;
    uint32_t *staging;
};
int place_upload(k8sgpu_ctx *ctx, const k8sgpu_node_table *nodes, const k8sgpu_pod_table *pods,
                 const PodShare *share = nullptr);
int place_download(k8sgpu_ctx *ctx, k8sgpu_place_out *out);
int place_run(k8sgpu_ctx *ctx, const k8sgpu_policy *policy);
int round_begin(k8sgpu_ctx *ctx, const k8sgpu_policy *policy, cudaStream_t s);
int round_score(k8sgpu_ctx *ctx, uint32_t lo, uint32_t hi, int32_t *d_best, cudaStream_t s, bool prep = true,
                bool finalize = true, uint32_t xw = 1, uint32_t xrank = 0);
int round_finalize_peers(k8sgpu_ctx *ctx, uint32_t lo, uint32_t hi, int32_t *const *peer, int world, cudaStream_t s,
                         uint32_t xw, uint32_t xrank);
uint32_t group_slot_count(uint32_t na, uint32_t xw);
int round_admit(k8sgpu_ctx *ctx, const int32_t *d_best_all, cudaStream_t s, uint32_t *n_active_out);
int round_admit_coop(k8sgpu_ctx *ctx, cudaStream_t s, const int32_t *proposals);
int round_admit_coop_launch(k8sgpu_ctx *ctx, cudaStream_t s, const int32_t *proposals, uint32_t xw = 1);
int round_admit_coop_finish(k8sgpu_ctx *ctx, cudaStream_t s);
int round_end(k8sgpu_ctx *ctx, cudaStream_t s);
bool coop_admission(const k8sgpu_ctx *ctx);
// group.cu
int group_place_run(k8sgpu_ctx *ctx, const k8sgpu_policy *policy, const k8sgpu_gang_table *gangs /* may be null */);
int group_upload(k8sgpu_ctx *ctx, const k8sgpu_node_table *nodes, const k8sgpu_pod_table *pods);
int group_sync(k8sgpu_ctx *ctx);
int group_init_local(k8sgpu_ctx *ctx);
void group_destroy(k8sgpu_ctx *ctx);

// primitives.cu
int exclusive_scan_u32(k8sgpu_ctx *ctx, cudaStream_t s, const uint32_t *d_in, uint32_t *d_out, uint32_t n,
                       DevBuf &tmp, uint32_t *d_total /* may be null */);
// Sorts n pairs by the low `bits` bits of the key, stable.  Result ends up in (keys, vals).
int sort_pairs(k8sgpu_ctx *ctx, cudaStream_t s, uint32_t *keys, uint32_t *vals, uint32_t n, int bits,
               SortScratch &scr);

}  // namespace k8sgpu
