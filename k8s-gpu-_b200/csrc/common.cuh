// common.cuh -- context, device buffers and launch bookkeeping shared by the .cu files of
// libk8sgpu.so.  sm_100a only.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdarg>
#include <cstdio>
#include <cstring>

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
    DevBuf best_key;   // m u64, indexed by active position
    DevBuf best;       // m i32, indexed by active position
    DevBuf sort_keys, sort_vals;  // m u32
    DevBuf seg;        // segmented-scan block aggregates
    DevBuf flags;      // m u32 (compaction flags / scan output)
    DevBuf counters;   // small u32 block: [0]=error flag, [1]=next active count, [2]=placed
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
    int timed = 0;        // scoring launches timed so far in this run
    uint64_t timed_evals = 0;
};

struct ReconcileState {
    uint32_t p = 0, v = 0, n_owner_ids = 0;
    bool uploaded = false, ran = false, has_ordinal = false, want_members = false, want_ordinals = false;
    DevBuf desired, pool_owner, vm_owner, vm_flags, vm_ordinal;
    DevBuf lut, pool_of, current, delta, create_count, kvict, victim_off, member_off, create_off, span_off;
    DevBuf sort_keys, sort_vals, victim_vm, create_ordinal, bitmap;
    DevBuf counters;  // [0]=error flag
    SortScratch sort;
    DevBuf scan_tmp;
    uint32_t n_victims = 0, n_members = 0, n_creates = 0;
    const uint32_t *sorted_vals = nullptr;  // rows grouped by pool (member_vm) after the last run
    int coop_blocks_per_sm = 0;             // occupancy of reconcile_coop_kernel (0 = not queried)
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

inline uint32_t ceil_div(uint32_t a, uint32_t b) { return (a + b - 1) / b; }
inline int key_bits(uint32_t max_key) {  // bits needed to hold values 0..max_key
    int b = 1;
    while (b < 32 && (max_key >> b)) ++b;
    return b;
}

// primitives.cu
int exclusive_scan_u32(k8sgpu_ctx *ctx, cudaStream_t s, const uint32_t *d_in, uint32_t *d_out, uint32_t n,
                       DevBuf &tmp, uint32_t *d_total /* may be null */);
// Sorts n pairs by the low `bits` bits of the key, stable.  Result ends up in (keys, vals).
int sort_pairs(k8sgpu_ctx *ctx, cudaStream_t s, uint32_t *keys, uint32_t *vals, uint32_t n, int bits,
               SortScratch &scr);

}  // namespace k8sgpu
