// reconcile.cu -- steps 3-5 of AzureVmPoolReconciler.Reconcile (README.md:187-230) for every
// pool of a snapshot at once.
//
//   tag filter (listManagedVMs, README.md:188-189,238)     -> match_kernel (128-bit SoA loads)
//   existingVMs of each pool, in list order                -> stable sort of rows by pool
//   currentCount := len(existingVMs)  (README.md:195)      -> per-pool counters filled by the tag filter
//                                                             (chain of kernels: segment bounds of the sort)
//   scaleUp / scaleDown counts (README.md:199-201,210-212) -> pool_kernel
//   vmsToDelete := existingVMs[:scaleDownCount] (:214)     -> victim_gather_kernel
//   names of new VMs (README.md:204)         -> ordinal bitmap kernels
//   Status.ReadyReplicas = currentCount (README.md:225)    -> ready == current
// createVM / deleteVM themselves are ARM calls (network) and stay on the host.
#include <cstdlib>

#include "prims.cuh"

namespace k8sgpu {

constexpr uint32_t kErrNegDesired = 1u, kErrDupOwner = 2u;

__global__ void pool_lut_kernel(const int32_t *__restrict__ desired, const uint32_t *__restrict__ owner, uint32_t P,
                                uint32_t n_owner_ids, int32_t *__restrict__ lut, uint32_t *__restrict__ err) {
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    if (desired[p] < 0) atomicOr(err, kErrNegDesired);  // +kubebuilder:validation:Minimum=0, README.md:94
    uint32_t o = owner[p];
    if (o < n_owner_ids) {  // (K8SGPU_NONE_U32 is never < n_owner_ids)
        int32_t old = atomicCAS(&lut[o], -1, (int32_t)p);
        if (old != -1) atomicOr(err, kErrDupOwner);  // an owner id maps to at most one pool
    }
}

__device__ __forceinline__ int32_t match_one(uint32_t owner, uint32_t flags, uint32_t n_owner_ids,
                                             const int32_t *__restrict__ lut) {
    // README.md:238: managed-by must match AND owner must be "<namespace>-<name>" of a pool
    return ((flags & 1u) && owner < n_owner_ids) ? __ldg(&lut[owner]) : -1;
}

// 4 VM rows per thread: one 128-bit load per input column, one 128-bit store per output column.
__global__ void __launch_bounds__(256) match_kernel(const uint32_t *__restrict__ owner,
                                                    const uint32_t *__restrict__ flags, uint32_t V,
                                                    uint32_t n_owner_ids, const int32_t *__restrict__ lut, uint32_t P,
                                                    int32_t *__restrict__ pool_of, uint32_t *__restrict__ keys,
                                                    uint32_t *__restrict__ vals) {
    const uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;  // quad index
    const uint32_t v0 = q * 4u;
    if (v0 >= V) return;
    if (v0 + 4u <= V) {
        const uint4 o = __ldg(reinterpret_cast<const uint4 *>(owner) + q);
        const uint4 f = __ldg(reinterpret_cast<const uint4 *>(flags) + q);
        int4 p;
        p.x = match_one(o.x, f.x, n_owner_ids, lut);
        p.y = match_one(o.y, f.y, n_owner_ids, lut);
        p.z = match_one(o.z, f.z, n_owner_ids, lut);
        p.w = match_one(o.w, f.w, n_owner_ids, lut);
        reinterpret_cast<int4 *>(pool_of)[q] = p;
        uint4 k;
        k.x = p.x < 0 ? P : (uint32_t)p.x;
        k.y = p.y < 0 ? P : (uint32_t)p.y;
        k.z = p.z < 0 ? P : (uint32_t)p.z;
        k.w = p.w < 0 ? P : (uint32_t)p.w;
        reinterpret_cast<uint4 *>(keys)[q] = k;
        reinterpret_cast<uint4 *>(vals)[q] = make_uint4(v0, v0 + 1, v0 + 2, v0 + 3);
    } else {
        for (uint32_t v = v0; v < V; ++v) {
            int32_t p = match_one(owner[v], flags[v], n_owner_ids, lut);
            pool_of[v] = p;
            keys[v] = p < 0 ? P : (uint32_t)p;
            vals[v] = v;
        }
    }
}

// member_off[p] = first sorted position whose key >= p, for p in [0, P].
__global__ void segment_bounds_kernel(const uint32_t *__restrict__ keys, uint32_t V, uint32_t P,
                                      int32_t *__restrict__ member_off) {
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p > P) return;
    uint32_t lo = 0, hi = V;
    while (lo < hi) {
        uint32_t mid = lo + ((hi - lo) >> 1);
        if (__ldg(&keys[mid]) < p) lo = mid + 1;
        else hi = mid;
    }
    member_off[p] = (int32_t)lo;
}

__global__ void pool_kernel(const int32_t *__restrict__ desired, const int32_t *__restrict__ member_off, uint32_t P,
                            int32_t *__restrict__ current, int32_t *__restrict__ delta,
                            int32_t *__restrict__ create_count, uint32_t *__restrict__ kvict,
                            uint32_t *__restrict__ span_words) {
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    int32_t cur = member_off[p + 1] - member_off[p];  // README.md:195
    int32_t des = desired[p];                         // README.md:196
    int32_t d = des - cur;
    current[p] = cur;
    delta[p] = d;
    int32_t cc = d > 0 ? d : 0;                       // scaleUpCount, README.md:201
    create_count[p] = cc;
    kvict[p] = d < 0 ? (uint32_t)(-d) : 0u;           // scaleDownCount, README.md:212
    // the cc lowest free ordinals all lie in [0, cur + cc)
    span_words[p] = cc ? ((uint32_t)cur + (uint32_t)cc + 31u) / 32u : 0u;
}

// vmsToDelete := existingVMs[:scaleDownCount]  (README.md:214)
__global__ void victim_gather_kernel(const uint32_t *__restrict__ keys, const uint32_t *__restrict__ vals,
                                     uint32_t n_members_upper, uint32_t P, const int32_t *__restrict__ member_off,
                                     const uint32_t *__restrict__ kvict, const int32_t *__restrict__ victim_off,
                                     uint32_t *__restrict__ victim_vm) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_members_upper) return;
    uint32_t p = keys[i];
    if (p >= P) return;  // unmanaged rows sort behind every pool
    uint32_t r = i - (uint32_t)member_off[p];
    if (r < kvict[p]) victim_vm[(uint32_t)victim_off[p] + r] = vals[i];
}

__global__ void ordinal_mark_kernel(const uint32_t *__restrict__ keys, const uint32_t *__restrict__ vals, uint32_t V,
                                    uint32_t P, const uint32_t *__restrict__ ordinal,
                                    const int32_t *__restrict__ current, const int32_t *__restrict__ create_count,
                                    const uint32_t *__restrict__ word_off, uint32_t *__restrict__ bitmap) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= V) return;
    uint32_t p = keys[i];
    if (p >= P) return;
    uint32_t cc = (uint32_t)create_count[p];
    if (!cc) return;
    uint32_t o = ordinal[vals[i]];
    if (o < (uint32_t)current[p] + cc) atomicOr(&bitmap[word_off[p] + (o >> 5)], 1u << (o & 31u));
}

// One warp per pool: emit the create_count lowest clear bits of the pool's ordinal bitmap.
__device__ __forceinline__ void ordinal_emit_pool(uint32_t p, const int32_t *__restrict__ current,
                                                  const int32_t *__restrict__ create_count,
                                                  const uint32_t *__restrict__ word_off,
                                                  const uint32_t *__restrict__ bitmap,
                                                  const int32_t *__restrict__ create_off,
                                                  uint32_t *__restrict__ create_ordinal) {
    const int lane = threadIdx.x & 31;
    const uint32_t cc = (uint32_t)create_count[p];
    if (!cc) return;
    const uint32_t span = (uint32_t)current[p] + cc, words = (span + 31u) / 32u;
    const uint32_t *bm = bitmap + word_off[p];
    uint32_t *out = create_ordinal + (uint32_t)create_off[p];
    uint32_t emitted = 0;
    for (uint32_t w0 = 0; w0 < words && emitted < cc; w0 += 32) {
        uint32_t w = w0 + lane;
        uint32_t free_bits = 0;
        if (w < words) {
            free_bits = ~bm[w];
            uint32_t rem = span - w * 32u;
            if (rem < 32u) free_bits &= (1u << rem) - 1u;
        }
        uint32_t c = __popc(free_bits), inc = c;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            uint32_t y = __shfl_up_sync(0xFFFFFFFFu, inc, d);
            if (lane >= d) inc += y;
        }
        uint32_t pos = emitted + inc - c;
        while (free_bits && pos < cc) {
            uint32_t b = __ffs(free_bits) - 1u;
            free_bits &= free_bits - 1u;
            out[pos++] = w * 32u + b;
        }
        emitted += __shfl_sync(0xFFFFFFFFu, inc, 31);
    }
}

__global__ void ordinal_emit_kernel(uint32_t P, const int32_t *__restrict__ current,
                                    const int32_t *__restrict__ create_count, const uint32_t *__restrict__ word_off,
                                    const uint32_t *__restrict__ bitmap, const int32_t *__restrict__ create_off,
                                    uint32_t *__restrict__ create_ordinal) {
    const uint32_t p = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (p < P) ordinal_emit_pool(p, current, create_count, word_off, bitmap, create_off, create_ordinal);
}

// ------------------------------------------------------------------ one launch for the whole pass
//
// Config 2 (1k pools x 100k VMs) is 1.2 MB of work: as ~16 dependent kernels it is pure launch
// latency.  reconcile_coop_kernel runs the same phases inside ONE cooperative launch with grid
// barriers between them (same tile functions as the chain, so the results are identical).

struct RecArgs {
    const int32_t *desired;
    const uint32_t *pool_owner, *vm_owner, *vm_flags, *vm_ordinal;  // vm_ordinal may be null
    int32_t *lut;  // pre-set to -1 by the host
    int32_t *pool_of, *current, *delta, *create_count, *member_off, *victim_off, *create_off;
    uint32_t *keys, *vals, *keys_alt, *vals_alt, *hist, *hist_scan, *scan_sums;
    uint32_t *kvict, *span_words, *word_off, *victim_vm, *bitmap /* pre-zeroed */, *create_ordinal, *counters;
    uint32_t P, V, n_owner_ids;
    int bits, want_ordinals;
    unsigned long long *prof;  // optional (K8SGPU_REC_PROF): %globaltimer at each phase boundary, else null
};

__device__ __forceinline__ void rec_mark(const RecArgs &a, int slot) {
    if (a.prof && blockIdx.x == 0 && threadIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        a.prof[slot] = t;
    }
}

// README.md:195-212 for pool p, from its member count.
__device__ __forceinline__ void pool_math(const RecArgs &a, uint32_t p) {
    int32_t cur = a.current[p], d = a.desired[p] - cur;
    a.delta[p] = d;
    int32_t cc = d > 0 ? d : 0;
    a.create_count[p] = cc;
    a.kvict[p] = d < 0 ? (uint32_t)(-d) : 0u;
    a.span_words[p] = cc ? ((uint32_t)cur + (uint32_t)cc + 31u) / 32u : 0u;
}

// Exclusive scan of n <= 16 tiles by the calling CTA alone; out[n] = total.
__device__ __forceinline__ void cta_exclusive_scan(const uint32_t *in, uint32_t *out, uint32_t n) {
    uint32_t carry = 0;
    for (uint32_t t = 0; t * kScanTile < n; ++t) carry = block_scan_tile(in, out, t * kScanTile, n, carry, true);
    if (threadIdx.x == 0) out[n] = carry;
}

// Phases (each line ends in a grid barrier):
//   1. owner id -> pool row table; member counters zeroed
//   2. tag filter, sort keys, digit-0 histogram, per-pool member count (warp-aggregated atomics:
//      counts only, order-free)
//   3. per 8-bit digit: histogram (after digit 0), its scan, the stable scatter.  While CTA 0
//      scans the first histogram, the LAST CTA does everything that needs only the counts --
//      currentCount, the diff, and the offsets of every output list -- so none of that costs a
//      barrier of its own (when P is too large for one CTA, the whole grid does it instead).
//   4. victims = first -delta rows of each pool's run of the sorted list; used-ordinal bits
//   5. names of the VMs to create (lowest free ordinals)
__global__ void __launch_bounds__(256) reconcile_coop_kernel(RecArgs a) {
    cg::grid_group grid = cg::this_grid();
    const uint32_t tid = threadIdx.x, cta = blockIdx.x, G = gridDim.x, gtid = cta * 256 + tid, gsz = G * 256;
    const uint32_t P = a.P, V = a.V;
    const uint32_t nb = (V + kSortTile - 1) / kSortTile;
    const bool one_cta_pools = (P + kScanTile - 1) / kScanTile <= 16;
    constexpr uint32_t kSmallP = 4096;  // pools whose member counters fit in shared memory
    __shared__ uint32_t pcount[kSmallP];
    const bool small_p = P <= kSmallP;
    if (small_p)
        for (uint32_t p = tid; p < P; p += 256) pcount[p] = 0;
    rec_mark(a, 0);

    for (uint32_t p = gtid; p < P; p += gsz) {  // owner id -> pool row
        if (a.desired[p] < 0) atomicOr(a.counters, kErrNegDesired);
        uint32_t o = a.pool_owner[p];
        if (o < a.n_owner_ids && atomicCAS(&a.lut[o], -1, (int32_t)p) != -1) atomicOr(a.counters, kErrDupOwner);
        a.current[p] = 0;
    }
    grid.sync();
    rec_mark(a, 1);
    for (uint32_t vt = cta; vt < nb; vt += G) {  // tag filter + sort keys + digit-0 histogram + member counts
        // all loads of the tile first (the stores below would otherwise fence each row's loads
        // behind the previous row's stores: 8 dependent round trips instead of 2)
        uint32_t ow[kSortTile / 256], fl[kSortTile / 256];
        int32_t pp[kSortTile / 256];
#pragma unroll
        for (uint32_t r = 0; r < kSortTile / 256; ++r) {
            uint32_t v = vt * kSortTile + r * 256 + tid;
            ow[r] = v < V ? a.vm_owner[v] : K8SGPU_NONE_U32;
            fl[r] = v < V ? a.vm_flags[v] : 0u;
        }
#pragma unroll
        for (uint32_t r = 0; r < kSortTile / 256; ++r)  // README.md:238; the table was written in phase 1: L2 load
            pp[r] = ((fl[r] & 1u) && ow[r] < a.n_owner_ids) ? __ldcg(&a.lut[ow[r]]) : -1;
#pragma unroll
        for (uint32_t r = 0; r < kSortTile / 256; ++r) {
            uint32_t v = vt * kSortTile + r * 256 + tid;
            uint32_t key = 0xFFFFFFFFu;  // rows past the end: no pool, nothing stored
            if (v < V) {
                int32_t p = pp[r];
                a.pool_of[v] = p;
                key = p < 0 ? P : (uint32_t)p;
                a.keys[v] = key;
                a.vals[v] = v;
            }
            // currentCount := len(existingVMs), README.md:195 (counts only: order-free)
            if (small_p) {
                if (key < P) atomicAdd(&pcount[key], 1u);
            } else {  // one global atomic per distinct pool per warp
                const uint32_t peers = __match_any_sync(0xFFFFFFFFu, key);
                if (key < P && (tid & 31u) == (uint32_t)__ffs(peers) - 1u)
                    atomicAdd(&a.current[key], (int32_t)__popc(peers));
            }
        }
        if (small_p) {  // flush this tile's per-pool counts
            __syncthreads();
            for (uint32_t p = tid; p < P; p += 256) {
                const uint32_t c = pcount[p];
                if (c) {
                    atomicAdd(&a.current[p], (int32_t)c);
                    pcount[p] = 0;
                }
            }
        }
        __syncthreads();
        if (vt == 0) rec_mark(a, 6);
        sort_hist_tile(a.keys, V, 0, 0xFFu, a.hist, nb, vt);
        if (vt == 0) rec_mark(a, 7);
    }
    grid.sync();
    rec_mark(a, 2);
    if (one_cta_pools) {
        if (cta == G - 1) {  // overlaps the first histogram scan (CTA 0) unless the grid is one CTA
            for (uint32_t p = tid; p < P; p += 256) pool_math(a, p);
            __syncthreads();
            cta_exclusive_scan((const uint32_t *)a.current, (uint32_t *)a.member_off, P);
            cta_exclusive_scan(a.kvict, (uint32_t *)a.victim_off, P);
            cta_exclusive_scan((const uint32_t *)a.create_count, (uint32_t *)a.create_off, P);
            if (a.want_ordinals) {
                cta_exclusive_scan(a.span_words, a.word_off, P);
                if (tid == 0) a.counters[4] = a.word_off[P];
            }
        }
    } else {
        for (uint32_t p = gtid; p < P; p += gsz) pool_math(a, p);
        grid.sync();
        coop_exclusive_scan(grid, (const uint32_t *)a.current, (uint32_t *)a.member_off, P, a.scan_sums,
                            (uint32_t *)a.member_off + P);
        coop_exclusive_scan(grid, a.kvict, (uint32_t *)a.victim_off, P, a.scan_sums, (uint32_t *)a.victim_off + P);
        coop_exclusive_scan(grid, (const uint32_t *)a.create_count, (uint32_t *)a.create_off, P, a.scan_sums,
                            (uint32_t *)a.create_off + P);
        if (a.want_ordinals) coop_exclusive_scan(grid, a.span_words, a.word_off, P, a.scan_sums, a.counters + 4);
    }
    uint32_t *sk = a.keys, *sv = a.vals, *dk = a.keys_alt, *dv = a.vals_alt;
    if (V > 0) {
        for (int shift = 0; shift < a.bits; shift += 8) {
            const int w = a.bits - shift < 8 ? a.bits - shift : 8;
            const uint32_t mask = (1u << w) - 1u;
            if (shift > 0) {
                for (uint32_t vt = cta; vt < nb; vt += G) sort_hist_tile(sk, V, shift, mask, a.hist, nb, vt);
                grid.sync();
            }
            if (nb <= kLocalScanMaxTiles) {  // every CTA scans the small histogram itself: no scan phase, no barrier
                for (uint32_t vt = cta; vt < nb; vt += G) sort_scatter_tile(sk, sv, dk, dv, V, shift, mask, a.hist, nb, vt, true);
            } else {
                coop_exclusive_scan(grid, a.hist, a.hist_scan, nb * 256u, a.scan_sums, nullptr);
                for (uint32_t vt = cta; vt < nb; vt += G) sort_scatter_tile(sk, sv, dk, dv, V, shift, mask, a.hist_scan, nb, vt);
            }
            grid.sync();
            uint32_t *t;
            t = sk; sk = dk; dk = t;
            t = sv; sv = dv; dv = t;
        }
    } else {
        grid.sync();  // the offsets written by the last CTA become visible to everyone
    }
    // (the host knows where the sorted pairs ended up: in keys/vals after an even number of passes)
    rec_mark(a, 3);
    for (uint32_t i = gtid; i < V; i += gsz) {  // victims (README.md:214) and used-ordinal bits
        uint32_t p = sk[i];
        if (p >= P) continue;
        uint32_t r = i - (uint32_t)a.member_off[p];
        if (r < a.kvict[p]) a.victim_vm[(uint32_t)a.victim_off[p] + r] = sv[i];
        if (a.want_ordinals && a.vm_ordinal) {
            uint32_t cc = (uint32_t)a.create_count[p];
            if (cc) {
                uint32_t o = a.vm_ordinal[sv[i]];
                if (o < (uint32_t)a.current[p] + cc) atomicOr(&a.bitmap[a.word_off[p] + (o >> 5)], 1u << (o & 31u));
            }
        }
    }
    if (!a.want_ordinals) return;
    grid.sync();
    rec_mark(a, 4);
    for (uint32_t p = gtid >> 5; p < P; p += gsz >> 5)  // one warp per pool
        ordinal_emit_pool(p, a.current, a.create_count, a.word_off, a.bitmap, a.create_off, a.create_ordinal);
    rec_mark(a, 5);
}

static int rec_upload(k8sgpu_ctx *ctx, const k8sgpu_pool_table *pools, const k8sgpu_vm_table *vms,
                      uint32_t n_owner_ids) {
    ReconcileState &st = ctx->rec;
    cudaStream_t s = ctx->stream;
    st.uploaded = st.ran = false;
    if (!pools || !vms) return fail(ctx, K8SGPU_E_ARG, "reconcile: NULL table");
    if (pools->p >= 0x7FFFFFFFu || vms->v >= 0x7FFFFFF0u)
        return fail(ctx, K8SGPU_E_ARG, "reconcile: table too large (p=%u v=%u)", pools->p, vms->v);
    if (pools->p && (!pools->desired || !pools->owner_id)) return fail(ctx, K8SGPU_E_ARG, "reconcile: NULL pool column");
    if (vms->v && (!vms->owner_id || !vms->flags)) return fail(ctx, K8SGPU_E_ARG, "reconcile: NULL VM column");
    const size_t P = pools->p, V = vms->v;
    st.p = pools->p;
    st.v = vms->v;
    st.n_owner_ids = n_owner_ids;
    st.has_ordinal = vms->ordinal != nullptr;
    K8S_CUDA(ctx, st.desired.reserve(P * 4 + 16));
    K8S_CUDA(ctx, st.pool_owner.reserve(P * 4 + 16));
    K8S_CUDA(ctx, st.vm_owner.reserve(V * 4 + 16));
    K8S_CUDA(ctx, st.vm_flags.reserve(V * 4 + 16));
    if (P) {
        K8S_CUDA(ctx, cudaMemcpyAsync(st.desired.p, pools->desired, P * 4, cudaMemcpyHostToDevice, s));
        K8S_CUDA(ctx, cudaMemcpyAsync(st.pool_owner.p, pools->owner_id, P * 4, cudaMemcpyHostToDevice, s));
    }
    if (V) {
        K8S_CUDA(ctx, cudaMemcpyAsync(st.vm_owner.p, vms->owner_id, V * 4, cudaMemcpyHostToDevice, s));
        K8S_CUDA(ctx, cudaMemcpyAsync(st.vm_flags.p, vms->flags, V * 4, cudaMemcpyHostToDevice, s));
    }
    ctx->stats.h2d_bytes += (2 * P + 2 * V) * 4;
    if (st.has_ordinal) {
        K8S_CUDA(ctx, st.vm_ordinal.reserve(V * 4 + 16));
        if (V) K8S_CUDA(ctx, cudaMemcpyAsync(st.vm_ordinal.p, vms->ordinal, V * 4, cudaMemcpyHostToDevice, s));
        ctx->stats.h2d_bytes += V * 4;
    }
    // upper bound of the ordinal bitmap: sum over pools of ceil((cur + cc)/32) <= (V + sum desired)/32 + P
    uint64_t sum_desired = 0;
    for (size_t p = 0; p < P; ++p) sum_desired += pools->desired[p] > 0 ? (uint64_t)pools->desired[p] : 0u;
    if (sum_desired > 0x7FFFFFFFull) return fail(ctx, K8SGPU_E_ARG, "reconcile: sum of desired replicas exceeds 2^31");
    const size_t bitmap_words = (V + sum_desired) / 32 + P + 1;
    K8S_CUDA(ctx, st.lut.reserve(((size_t)n_owner_ids + 1) * 4));
    K8S_CUDA(ctx, st.pool_of.reserve(V * 4 + 16));
    K8S_CUDA(ctx, st.current.reserve(P * 4 + 16));
    K8S_CUDA(ctx, st.delta.reserve(P * 4 + 16));
    K8S_CUDA(ctx, st.create_count.reserve(P * 4 + 16));
    K8S_CUDA(ctx, st.kvict.reserve(P * 4 + 16));
    K8S_CUDA(ctx, st.victim_off.reserve((P + 1) * 4 + 16));
    K8S_CUDA(ctx, st.member_off.reserve((P + 2) * 4 + 16));
    K8S_CUDA(ctx, st.create_off.reserve((P + 1) * 4 + 16));
    K8S_CUDA(ctx, st.span_off.reserve((2 * P + 2) * 4 + 16));
    K8S_CUDA(ctx, st.sort_keys.reserve(V * 4 + 16));
    K8S_CUDA(ctx, st.sort_vals.reserve(V * 4 + 16));
    K8S_CUDA(ctx, st.victim_vm.reserve(V * 4 + 16));
    K8S_CUDA(ctx, st.create_ordinal.reserve((size_t)sum_desired * 4 + 16));
    K8S_CUDA(ctx, st.bitmap.reserve(bitmap_words * 4));
    K8S_CUDA(ctx, st.counters.reserve(64 * 4));
    st.uploaded = true;
    return K8SGPU_OK;
}

static int rec_finish(k8sgpu_ctx *ctx, cudaStream_t s);

static int rec_run_coop(k8sgpu_ctx *ctx, cudaStream_t s) {
    ReconcileState &st = ctx->rec;
    const uint32_t P = st.p, V = st.v;
    uint32_t *cnts = st.counters.as<uint32_t>();
    if (st.coop_blocks_per_sm == 0) {
        int per_sm = 0;
        K8S_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, reconcile_coop_kernel, 256, 0));
        st.coop_blocks_per_sm = per_sm > 0 ? per_sm : -1;
    }
    if (st.coop_blocks_per_sm < 0) return fail(ctx, K8SGPU_E_CUDA, "cooperative reconcile kernel cannot be resident");
    const uint32_t nb = ceil_div(V ? V : 1u, kSortTile);
    K8S_CUDA(ctx, st.sort.keys_alt.reserve((size_t)V * 4 + 16));
    K8S_CUDA(ctx, st.sort.vals_alt.reserve((size_t)V * 4 + 16));
    K8S_CUDA(ctx, st.sort.hist.reserve(((size_t)nb * 256 * 2 + 256) * 4));
    K8S_CUDA(ctx, st.sort.scan_tmp.reserve(((size_t)nb * 256 / kScanTile + (size_t)P / kScanTile + 128) * 4));
    K8S_CUDA(ctx, cudaEventRecord(ctx->ev_begin, s));
    K8S_CUDA(ctx, cudaMemsetAsync(cnts, 0, 64 * 4, s));
    K8S_CUDA(ctx, cudaMemsetAsync(st.lut.p, 0xFF, ((size_t)st.n_owner_ids + 1) * 4, s));
    if (st.want_ordinals) K8S_CUDA(ctx, cudaMemsetAsync(st.bitmap.p, 0, st.bitmap.cap, s));
    RecArgs a;
    a.desired = st.desired.as<int32_t>();
    a.pool_owner = st.pool_owner.as<uint32_t>();
    a.vm_owner = st.vm_owner.as<uint32_t>();
    a.vm_flags = st.vm_flags.as<uint32_t>();
    a.vm_ordinal = st.has_ordinal ? st.vm_ordinal.as<uint32_t>() : nullptr;
    a.lut = st.lut.as<int32_t>();
    a.pool_of = st.pool_of.as<int32_t>();
    a.current = st.current.as<int32_t>();
    a.delta = st.delta.as<int32_t>();
    a.create_count = st.create_count.as<int32_t>();
    a.member_off = st.member_off.as<int32_t>();
    a.victim_off = st.victim_off.as<int32_t>();
    a.create_off = st.create_off.as<int32_t>();
    a.keys = st.sort_keys.as<uint32_t>();
    a.vals = st.sort_vals.as<uint32_t>();
    a.keys_alt = st.sort.keys_alt.as<uint32_t>();
    a.vals_alt = st.sort.vals_alt.as<uint32_t>();
    a.hist = st.sort.hist.as<uint32_t>();
    a.hist_scan = a.hist + (size_t)nb * 256;
    a.scan_sums = st.sort.scan_tmp.as<uint32_t>();
    a.kvict = st.kvict.as<uint32_t>();
    a.span_words = st.span_off.as<uint32_t>();
    a.word_off = a.span_words + P + 1;
    a.victim_vm = st.victim_vm.as<uint32_t>();
    a.bitmap = st.bitmap.as<uint32_t>();
    a.create_ordinal = st.create_ordinal.as<uint32_t>();
    a.counters = cnts;
    a.P = P;
    a.V = V;
    a.n_owner_ids = st.n_owner_ids;
    a.bits = key_bits(P);
    a.want_ordinals = st.want_ordinals ? 1 : 0;
    const int passes = V ? (a.bits + 7) / 8 : 0;
    st.sorted_vals = (passes & 1) ? a.vals_alt : a.vals;
    // CTAs: enough to spread the tiles (+1 for the pool offsets), few enough that the grid barriers
    // stay cheap -- at most one per SM while the pass is latency bound (measured at config 2:
    // 98-148 CTAs 57-58 us, 196 CTAs 62 us), two per SM beyond that.
    uint32_t want = ceil_div(V ? V : 1u, 1024u) + 1u, pw = ceil_div(P + 1, 256u);
    if (pw > want) want = pw;
    const uint32_t per_sm = (V <= (256u << 10) || st.coop_blocks_per_sm < 2) ? 1u : 2u;
    uint32_t cap = (uint32_t)ctx->sm_count * per_sm;
    uint32_t grid = want > cap ? cap : want;
    if (const char *e = getenv("K8SGPU_REC_GRID")) {  // developer knob: CTAs in the cooperative grid
        uint32_t g = (uint32_t)atoi(e);
        if (g >= 1 && g <= (uint32_t)ctx->sm_count * (uint32_t)st.coop_blocks_per_sm) grid = g;
    }
    static const bool want_prof = getenv("K8SGPU_REC_PROF") != nullptr;
    a.prof = want_prof ? reinterpret_cast<unsigned long long *>(cnts + 16) : nullptr;  // counters[16..28)
    void *args[] = {&a};
    K8S_CUDA(ctx, cudaLaunchCooperativeKernel((void *)reconcile_coop_kernel, dim3(grid), dim3(256), args, 0, s));
    ctx->stats.kernel_launches++;
    int rc = rec_finish(ctx, s);
    if (want_prof && rc == K8SGPU_OK) {  // developer aid: phase boundaries (CTA 0's clock), microseconds
        unsigned long long t[8];
        K8S_CUDA(ctx, cudaMemcpy(t, cnts + 16, sizeof t, cudaMemcpyDeviceToHost));
        fprintf(stderr, "[reconcile coop CTA 0, tile 0] match %.1f hist %.1f barrier %.1f us\n", (t[6] - t[1]) / 1e3,
                (t[7] - t[6]) / 1e3, (t[2] - t[7]) / 1e3);
        fprintf(stderr, "[reconcile coop P=%u V=%u grid=%u] lut %.1f match+hist %.1f pools+sort %.1f gather %.1f "
                        "ordinals %.1f kernel %.1f us, events %.1f us\n",
                P, V, grid, (t[1] - t[0]) / 1e3, (t[2] - t[1]) / 1e3, (t[3] - t[2]) / 1e3, (t[4] - t[3]) / 1e3,
                (t[5] - t[4]) / 1e3, (t[5] - t[0]) / 1e3, ctx->stats.device_ms * 1e3);
    }
    return rc;
}

static int rec_run(k8sgpu_ctx *ctx, int want_members, int want_ordinals) {
    ReconcileState &st = ctx->rec;
    cudaStream_t s = ctx->stream;
    if (!st.uploaded) return fail(ctx, K8SGPU_E_ARG, "reconcile: no tables uploaded");
    const uint32_t P = st.p, V = st.v;
    st.want_members = want_members != 0;
    st.want_ordinals = want_ordinals != 0;
    uint32_t *cnts = st.counters.as<uint32_t>();
    st.sorted_vals = st.sort_vals.as<uint32_t>();
    // One cooperative launch for the whole pass while it is latency bound (config 2 is); the chain
    // of kernels below once the tables are big enough to be bandwidth bound.  K8SGPU_REC_COOP=0/1
    // forces either.
    bool coop = V <= (4u << 20);
    if (const char *e = getenv("K8SGPU_REC_COOP")) coop = atoi(e) != 0;
    if (coop) return rec_run_coop(ctx, s);
    K8S_CUDA(ctx, cudaEventRecord(ctx->ev_begin, s));
    K8S_CUDA(ctx, cudaMemsetAsync(cnts, 0, 64 * 4, s));
    K8S_CUDA(ctx, cudaMemsetAsync(st.lut.p, 0xFF, ((size_t)st.n_owner_ids + 1) * 4, s));
    if (P)
        K8S_LAUNCH(ctx, pool_lut_kernel, ceil_div(P, 256), 256, 0, s, st.desired.as<int32_t>(),
                   st.pool_owner.as<uint32_t>(), P, st.n_owner_ids, st.lut.as<int32_t>(), cnts);
    uint32_t *keys = st.sort_keys.as<uint32_t>(), *vals = st.sort_vals.as<uint32_t>();
    if (V) {
        K8S_LAUNCH(ctx, match_kernel, ceil_div(ceil_div(V, 4), 256), 256, 0, s, st.vm_owner.as<uint32_t>(),
                   st.vm_flags.as<uint32_t>(), V, st.n_owner_ids, st.lut.as<int32_t>(), P, st.pool_of.as<int32_t>(),
                   keys, vals);
        int rc = sort_pairs(ctx, s, keys, vals, V, key_bits(P), st.sort);
        if (rc) return rc;
    }
    int32_t *member_off = st.member_off.as<int32_t>();
    K8S_LAUNCH(ctx, segment_bounds_kernel, ceil_div(P + 1, 256), 256, 0, s, keys, V, P, member_off);
    uint32_t *kvict = st.kvict.as<uint32_t>();
    uint32_t *span_words = st.span_off.as<uint32_t>(), *word_off = span_words + P + 1;
    int32_t *victim_off = st.victim_off.as<int32_t>(), *create_off = st.create_off.as<int32_t>();
    if (P) {
        K8S_LAUNCH(ctx, pool_kernel, ceil_div(P, 256), 256, 0, s, st.desired.as<int32_t>(), member_off, P,
                   st.current.as<int32_t>(), st.delta.as<int32_t>(), st.create_count.as<int32_t>(), kvict, span_words);
        int rc = exclusive_scan_u32(ctx, s, kvict, (uint32_t *)victim_off, P, st.scan_tmp, (uint32_t *)victim_off + P);
        if (rc) return rc;
        rc = exclusive_scan_u32(ctx, s, (const uint32_t *)st.create_count.p, (uint32_t *)create_off, P, st.scan_tmp,
                                (uint32_t *)create_off + P);
        if (rc) return rc;
    } else {
        K8S_CUDA(ctx, cudaMemsetAsync(victim_off, 0, 4, s));
        K8S_CUDA(ctx, cudaMemsetAsync(create_off, 0, 4, s));
    }
    if (V && P)
        K8S_LAUNCH(ctx, victim_gather_kernel, ceil_div(V, 256), 256, 0, s, keys, vals, V, P, member_off, kvict,
                   victim_off, st.victim_vm.as<uint32_t>());
    if (st.want_ordinals && P) {
        int rc = exclusive_scan_u32(ctx, s, span_words, word_off, P, st.scan_tmp, cnts + 4);
        if (rc) return rc;
        K8S_CUDA(ctx, cudaMemsetAsync(st.bitmap.p, 0, st.bitmap.cap, s));
        if (st.has_ordinal && V)
            K8S_LAUNCH(ctx, ordinal_mark_kernel, ceil_div(V, 256), 256, 0, s, keys, vals, V, P,
                       st.vm_ordinal.as<uint32_t>(), st.current.as<int32_t>(), st.create_count.as<int32_t>(),
                       word_off, st.bitmap.as<uint32_t>());
        K8S_LAUNCH(ctx, ordinal_emit_kernel, ceil_div(P, 8), 256, 0, s, P, st.current.as<int32_t>(),
                   st.create_count.as<int32_t>(), word_off, st.bitmap.as<uint32_t>(), create_off,
                   st.create_ordinal.as<uint32_t>());
    }
    return rec_finish(ctx, s);
}

static int rec_finish(k8sgpu_ctx *ctx, cudaStream_t s) {
    ReconcileState &st = ctx->rec;
    const uint32_t P = st.p;
    uint32_t *cnts = st.counters.as<uint32_t>();
    int32_t *victim_off = st.victim_off.as<int32_t>(), *create_off = st.create_off.as<int32_t>();
    int32_t *member_off = st.member_off.as<int32_t>();
    K8S_CUDA(ctx, cudaEventRecord(ctx->ev_end, s));
    // totals + error flag
    K8S_CUDA(ctx, cudaMemcpyAsync(ctx->h_pinned, cnts, 4, cudaMemcpyDeviceToHost, s));
    K8S_CUDA(ctx, cudaMemcpyAsync(ctx->h_pinned + 1, victim_off + P, 4, cudaMemcpyDeviceToHost, s));
    K8S_CUDA(ctx, cudaMemcpyAsync(ctx->h_pinned + 2, member_off + P, 4, cudaMemcpyDeviceToHost, s));
    K8S_CUDA(ctx, cudaMemcpyAsync(ctx->h_pinned + 3, create_off + P, 4, cudaMemcpyDeviceToHost, s));
    K8S_CUDA(ctx, cudaStreamSynchronize(s));
    if (ctx->h_pinned[0] & kErrNegDesired) return fail(ctx, K8SGPU_E_ARG, "reconcile: negative desired replicas");
    if (ctx->h_pinned[0] & kErrDupOwner) return fail(ctx, K8SGPU_E_ARG, "reconcile: two pools share an owner id");
    st.n_victims = ctx->h_pinned[1];
    st.n_members = ctx->h_pinned[2];
    st.n_creates = ctx->h_pinned[3];
    float ms = 0.f;
    K8S_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev_begin, ctx->ev_end));
    ctx->stats.device_ms = ms;
    st.ran = true;
    return K8SGPU_OK;
}

static int rec_download(k8sgpu_ctx *ctx, k8sgpu_reconcile_out *out) {
    ReconcileState &st = ctx->rec;
    cudaStream_t s = ctx->stream;
    if (!st.ran) return fail(ctx, K8SGPU_E_ARG, "reconcile: nothing to download");
    if (!out) return fail(ctx, K8SGPU_E_ARG, "reconcile: NULL out");
    const size_t P = st.p, V = st.v;
    if ((V && !out->pool_of) || (P && (!out->current || !out->delta || !out->ready || !out->create_count)) ||
        !out->victim_off || (st.n_victims && !out->victim_vm))
        return fail(ctx, K8SGPU_E_ARG, "reconcile: NULL output column");
    if (out->create_off && out->create_ordinal && st.want_ordinals && st.n_creates > out->create_capacity)
        return fail(ctx, K8SGPU_E_CAPACITY, "reconcile: create_ordinal needs %u elements, caller gave %llu",
                    st.n_creates, (unsigned long long)out->create_capacity);
    uint64_t bytes = 0;
#define D2H(dst, src, n)                                                                         \
    if ((n) > 0) {                                                                               \
        K8S_CUDA(ctx, cudaMemcpyAsync((dst), (src), (size_t)(n)*4, cudaMemcpyDeviceToHost, s)); \
        bytes += (uint64_t)(n)*4;                                                                \
    }
    D2H(out->pool_of, st.pool_of.p, V);
    D2H(out->current, st.current.p, P);
    D2H(out->ready, st.current.p, P);  // README.md:225: ReadyReplicas = currentCount
    D2H(out->delta, st.delta.p, P);
    D2H(out->create_count, st.create_count.p, P);
    D2H(out->victim_off, st.victim_off.p, P + 1);
    D2H(out->victim_vm, st.victim_vm.p, st.n_victims);
    if (out->member_off && out->member_vm && st.want_members) {
        D2H(out->member_off, st.member_off.p, P + 1);
        D2H(out->member_vm, st.sorted_vals, st.n_members);
    }
    if (out->create_off && out->create_ordinal && st.want_ordinals) {
        D2H(out->create_off, st.create_off.p, P + 1);
        D2H(out->create_ordinal, st.create_ordinal.p, st.n_creates);
    }
#undef D2H
    ctx->stats.d2h_bytes += bytes;
    K8S_CUDA(ctx, cudaStreamSynchronize(s));
    return K8SGPU_OK;
}

}  // namespace k8sgpu

using namespace k8sgpu;

extern "C" int k8sgpu_reconcile_upload(k8sgpu_ctx *ctx, const k8sgpu_pool_table *pools, const k8sgpu_vm_table *vms,
                                       uint32_t n_owner_ids) {
    K8S_ENTER(ctx);
    reset_call_stats(ctx);
    int rc = rec_upload(ctx, pools, vms, n_owner_ids);
    if (rc) return rc;
    K8S_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return K8SGPU_OK;
}

extern "C" int k8sgpu_reconcile_run(k8sgpu_ctx *ctx, int want_members, int want_ordinals) {
    K8S_ENTER(ctx);
    reset_call_stats(ctx);
    return rec_run(ctx, want_members, want_ordinals);
}

extern "C" int k8sgpu_reconcile_download(k8sgpu_ctx *ctx, k8sgpu_reconcile_out *out) {
    K8S_ENTER(ctx);
    ctx->stats.d2h_bytes = 0;
    return rec_download(ctx, out);
}

extern "C" int k8sgpu_reconcile_batch(k8sgpu_ctx *ctx, const k8sgpu_pool_table *pools, const k8sgpu_vm_table *vms,
                                      uint32_t n_owner_ids, k8sgpu_reconcile_out *out) {
    K8S_ENTER(ctx);
    reset_call_stats(ctx);
    if (!out) return fail(ctx, K8SGPU_E_ARG, "reconcile: NULL out");
    int rc = rec_upload(ctx, pools, vms, n_owner_ids);
    if (rc) return rc;
    rc = rec_run(ctx, out->member_off && out->member_vm, out->create_off && out->create_ordinal);
    if (rc) return rc;
    return rec_download(ctx, out);
}
