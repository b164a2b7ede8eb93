// reconcile.cu -- steps 3-5 of AzureVmPoolReconciler.Reconcile (README.md:187-230) for every
// pool of a snapshot at once.
//
//   tag filter (listManagedVMs, README.md:188-189,238)     -> owner id -> pool row table, one gather per VM row
//   currentCount := len(existingVMs)  (README.md:195)      -> per-pool member counts (counts only: order-free)
//   scaleUp / scaleDown counts (README.md:199-201,210-212) -> pool_math
//   vmsToDelete := existingVMs[:scaleDownCount] (:214)     -> RANK of each row among the earlier rows of its
//                                                             pool, in list order; rank < scaleDownCount = victim
//   names of new VMs (README.md:204)         -> ordinal bitmap kernels
//   Status.ReadyReplicas = currentCount (README.md:225)    -> ready == current
// createVM / deleteVM themselves are ARM calls (network) and stay on the host.
//
// The rank is all that list order is needed for, so nothing is sorted unless the caller asks for
// the full member lists of a large table.  Three paths, all bit-identical (tests run each):
//   0  one cooperative launch (pools <= kSmallMaxPools): every warp owns a contiguous run of rows
//      and a private per-pool counter array in shared memory; per-CTA totals are prefix-summed over
//      CTAs per pool, which turns the private counters into running global ranks.  Two grid
//      barriers (three with create names).  Config 2 (1k pools x 100k VMs) runs here.
//   1  many pools, victims only: global per-pool counts (RED), then a pruning pass over row
//      segments that keeps only rows of pools still short of victims (a deterministic prefix of
//      each such pool's rows), and the stable sort of just those candidates.
//   2  many pools with member lists: the full stable sort by pool (rows grouped = Status.VMs).
#include "prims.cuh"

namespace k8sgpu {

constexpr uint32_t kErrNegDesired = 1u, kErrDupOwner = 2u;
constexpr uint32_t kSmallWarps = 8;          // warps per CTA of the one-launch kernel
constexpr uint32_t kSmallMaxPools = 6144;    // 8 warps x 6144 pools x 4 B = 192 KB of shared memory
constexpr uint32_t kRecMailbox = 8;          // reconcile's words of the mapped mailbox start here

enum { kOutPoolOf = 0, kOutCurrent, kOutReady, kOutDelta, kOutCreateCount, kOutVictimOff, kOutMemberOff,
       kOutCreateOff, kOutVictimVm, kOutMemberVm, kOutCreateOrdinal, kOutTotal };
enum { kInVmOwner = 0, kInVmFlags, kInVmOrdinal, kInDesired, kInPoolOwner, kInTotal };

// Section offsets (4-byte elements) of the output block: the V-sized tag-filter result first (so
// that it is 16-byte aligned), then everything whose length is fixed by P, then the three lists
// whose used length depends on the data.  Tightly packed: a caller that lays its host buffers out
// the same way (k8sgpu_reconcile_out_size) gets one device->host copy for the fixed part.
static void rec_out_layout(uint64_t P, uint64_t V, uint64_t C, uint64_t off[12]) {
    uint64_t o = 0;
    const uint64_t len[11] = {V, P, P, P, P, P + 1, P + 1, P + 1, V, V, C};
    for (int i = 0; i < 11; ++i) {
        off[i] = o;
        o += len[i];
    }
    off[kOutTotal] = o;
}

static void rec_in_layout(uint64_t P, uint64_t V, uint64_t off[6]) {
    uint64_t o = 0;
    const uint64_t len[5] = {V, V, V, P, P};
    for (int i = 0; i < 5; ++i) {
        off[i] = o;
        o += len[i];
    }
    off[kInTotal] = o;
}

// Copies queued back to back on one stream; entries whose source AND destination are adjacent
// are merged into one transfer.
struct CopyList {
    struct E {
        char *dst;
        const char *src;
        size_t bytes;
    } e[16];
    int n = 0;
    uint64_t total = 0;
    void add(void *dst, const void *src, size_t bytes) {
        if (!bytes) return;
        total += bytes;
        if (n && e[n - 1].dst + e[n - 1].bytes == (char *)dst && e[n - 1].src + e[n - 1].bytes == (const char *)src) {
            e[n - 1].bytes += bytes;
            return;
        }
        e[n++] = {(char *)dst, (const char *)src, bytes};
    }
    cudaError_t issue(cudaMemcpyKind kind, cudaStream_t s) const {
        for (int i = 0; i < n; ++i) {
            cudaError_t err = cudaMemcpyAsync(e[i].dst, e[i].src, e[i].bytes, kind, s);
            if (err != cudaSuccess) return err;
        }
        return cudaSuccess;
    }
};

// ------------------------------------------------------------------ upload time: owner id -> pool row

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

// README.md:195-212 for one pool, from its member count.  scaleDownCount can never exceed the
// member count for a valid (non-negative) desired; the clamp keeps every later offset inside its
// buffer even if an invalid table slipped through.
struct PoolVerdict {
    int32_t delta, create_count;
    uint32_t kvict, span_words;
};
__device__ __forceinline__ PoolVerdict pool_math(int32_t cur, int32_t des) {
    PoolVerdict r;
    const long long d = (long long)des - (long long)cur;
    r.delta = des - cur;                                        // README.md:201 / :212
    r.create_count = d > 0 ? (int32_t)d : 0;                    // scaleUpCount
    const long long k = d < 0 ? -d : 0;                         // scaleDownCount
    r.kvict = (uint32_t)(k > cur ? cur : k);
    // the create_count lowest free ordinals all lie in [0, cur + create_count)
    r.span_words = r.create_count ? ((uint32_t)cur + (uint32_t)r.create_count + 31u) / 32u : 0u;
    return r;
}

// One warp per pool: emit the create_count lowest clear bits of the pool's ordinal bitmap.
// (No __restrict__ / read-only loads here: the one-launch kernel calls this on arrays other CTAs
// wrote earlier in the same launch, so every input is read from L2.)
__device__ __forceinline__ void ordinal_emit_pool(uint32_t p, const int32_t *current, const int32_t *create_count,
                                                  const uint32_t *word_off, const uint32_t *bitmap,
                                                  const int32_t *create_off, uint32_t *create_ordinal) {
    const int lane = threadIdx.x & 31;
    const uint32_t cc = (uint32_t)__ldcg(&create_count[p]);
    if (!cc) return;
    const uint32_t span = (uint32_t)__ldcg(&current[p]) + cc, words = (span + 31u) / 32u;
    const uint32_t *bm = bitmap + __ldcg(&word_off[p]);
    uint32_t *out = create_ordinal + (uint32_t)__ldcg(&create_off[p]);
    uint32_t emitted = 0;
    for (uint32_t w0 = 0; w0 < words && emitted < cc; w0 += 32) {
        uint32_t w = w0 + lane;
        uint32_t free_bits = 0;
        if (w < words) {
            free_bits = ~__ldcg(&bm[w]);  // set by atomics of this launch: read it where they landed (L2)
            uint32_t rem = span - w * 32u;
            if (rem < 32u) free_bits &= (1u << rem) - 1u;
        }
        uint32_t c = __popc(free_bits), inc = warp_incl_scan(c, lane);
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

// Exclusive scan of n elements by the calling CTA alone (256 threads); out[n] = total.
__device__ __forceinline__ uint32_t cta_exclusive_scan(const uint32_t *in, uint32_t *out, uint32_t n) {
    uint32_t carry = 0;
    for (uint32_t t = 0; t * kScanTile < n; ++t) carry = block_scan_tile(in, out, t * kScanTile, n, carry, true);
    if (threadIdx.x == 0) out[n] = carry;
    return carry;
}

// ------------------------------------------------------------------ path 0: one launch, few pools

struct RecSmallArgs {
    const uint32_t *vm_owner, *vm_flags, *vm_ordinal;  // vm_ordinal may be null
    const int32_t *desired, *lut;
    int32_t *pool_of, *current, *ready, *delta, *create_count, *victim_off, *member_off, *create_off;
    uint32_t *victim_vm, *member_vm, *create_ordinal;
    uint32_t *tot;  // [P][G] members of pool p among CTA c's rows; prefix-summed over c in place
    uint32_t *kvict, *span_words, *word_off, *bitmap, *counters, *mailbox;
    uint32_t P, V, n_owner_ids, rows_per_warp /* multiple of 32 */, bitmap_words, seq;
    int want_members, want_ordinals;
    unsigned long long *prof;  // optional: %globaltimer at each phase boundary (CTA 0), else null
};

__device__ __forceinline__ void rec_mark(unsigned long long *prof, int slot) {
    if (prof && blockIdx.x == 0 && threadIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        prof[slot] = t;
    }
}

// Phases (each line ends in a grid barrier):
//   A. tag filter of this warp's rows -> pool_of; members per (warp, pool) in shared memory;
//      members per (CTA, pool) to global memory; ordinal bitmap cleared
//   B. one warp per pool: prefix sum of the per-CTA counts (= rank base of each CTA's rows),
//      currentCount, the diff, scale-up / scale-down counts.  The CTA that finishes last also
//      turns the per-pool counts into the offsets of every output list and tells the host the
//      list lengths -- no barrier of its own.
//   C. the per-(warp, pool) counts become running ranks; second walk over the same rows in list
//      order: rank < scaleDownCount -> victim slot (README.md:214); member slot; used-ordinal bit
//   D. names of the VMs to create (lowest free ordinals), one warp per pool
__global__ void __launch_bounds__(kSmallWarps * 32) reconcile_small_kernel(RecSmallArgs a) {
    cg::grid_group grid = cg::this_grid();
    extern __shared__ uint32_t cnt[];  // [kSmallWarps][P]
    __shared__ uint32_t is_last;
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5, cta = blockIdx.x, G = gridDim.x;
    const uint32_t P = a.P, V = a.V;
    const uint32_t lt = (1u << lane) - 1u;
    rec_mark(a.prof, 0);

    for (uint32_t i = tid; i < kSmallWarps * P; i += blockDim.x) cnt[i] = 0;
    if (a.want_ordinals)
        for (uint32_t i = cta * blockDim.x + tid; i < a.bitmap_words; i += G * blockDim.x) a.bitmap[i] = 0;
    if (cta == 0 && tid == 0) a.counters[1] = 0;  // arrival ticket of phase B
    __syncthreads();

    uint32_t *mine = cnt + warp * P;
    const uint64_t gw = (uint64_t)cta * kSmallWarps + warp;
    const uint64_t r0_64 = gw * a.rows_per_warp;
    const uint32_t r0 = r0_64 < V ? (uint32_t)r0_64 : V;
    const uint32_t r1 = (uint64_t)r0 + a.rows_per_warp < V ? r0 + a.rows_per_warp : V;
    constexpr int U = 4;  // steps of 32 rows whose loads are issued together
    for (uint32_t v0 = r0; v0 < r1; v0 += 32 * U) {
        uint32_t ow[U], fl[U];
        int32_t pp[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint32_t v = v0 + u * 32 + lane;
            ow[u] = v < r1 ? __ldg(&a.vm_owner[v]) : K8SGPU_NONE_U32;
            fl[u] = v < r1 ? __ldg(&a.vm_flags[v]) : 0u;
        }
#pragma unroll
        for (int u = 0; u < U; ++u) pp[u] = match_one(ow[u], fl[u], a.n_owner_ids, a.lut);
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint32_t v = v0 + u * 32 + lane;
            if (v < r1) a.pool_of[v] = pp[u];
            const uint32_t key = (v < r1 && pp[u] >= 0) ? (uint32_t)pp[u] : K8SGPU_NONE_U32;
            const uint32_t peers = __match_any_sync(0xFFFFFFFFu, key);
            // currentCount := len(existingVMs), README.md:195.  The lowest lane of each pool's
            // group adds the group's size: distinct pools, distinct words, no atomics.
            if (key != K8SGPU_NONE_U32 && (peers & lt) == 0u) mine[key] += (uint32_t)__popc(peers);
            __syncwarp();
        }
    }
    __syncthreads();
    for (uint32_t p = tid; p < P; p += blockDim.x) {
        uint32_t t = 0;
#pragma unroll
        for (uint32_t w = 0; w < kSmallWarps; ++w) t += cnt[w * P + p];
        a.tot[(size_t)p * G + cta] = t;
    }
    grid.sync();
    rec_mark(a.prof, 1);

    for (uint32_t p = cta * kSmallWarps + warp; p < P; p += G * kSmallWarps) {
        uint32_t *row = a.tot + (size_t)p * G;
        uint32_t run = 0;
        for (uint32_t c0 = 0; c0 < G; c0 += 32) {
            const uint32_t c = c0 + lane;
            const uint32_t x = c < G ? __ldcg(&row[c]) : 0u;
            const uint32_t inc = warp_incl_scan(x, lane);
            if (c < G) row[c] = run + inc - x;
            run += __shfl_sync(0xFFFFFFFFu, inc, 31);
        }
        if (lane == 0) {
            const int32_t cur = (int32_t)run;
            const PoolVerdict r = pool_math(cur, a.desired[p]);
            a.current[p] = cur;
            a.ready[p] = cur;  // README.md:225: ReadyReplicas = currentCount, the pre-action count
            a.delta[p] = r.delta;
            a.create_count[p] = r.create_count;
            a.kvict[p] = r.kvict;
            a.span_words[p] = r.span_words;
        }
    }
    __syncthreads();
    if (tid == 0) {
        __threadfence();
        is_last = atomicAdd(&a.counters[1], 1u) == G - 1 ? 1u : 0u;
    }
    __syncthreads();
    if (is_last) {  // every pool's counts are in global memory, and visible: offsets of the output lists
        __threadfence();
        const uint32_t nm = cta_exclusive_scan((const uint32_t *)a.current, (uint32_t *)a.member_off, P);
        const uint32_t nv = cta_exclusive_scan(a.kvict, (uint32_t *)a.victim_off, P);
        const uint32_t nc = cta_exclusive_scan((const uint32_t *)a.create_count, (uint32_t *)a.create_off, P);
        if (a.want_ordinals) cta_exclusive_scan(a.span_words, a.word_off, P);
        if (tid == 0) {  // the host can size its copies while phases C and D run
            volatile uint32_t *mb = a.mailbox;
            mb[1] = a.counters[0];
            mb[2] = nv;
            mb[3] = nm;
            mb[4] = nc;
            __threadfence_system();
            mb[0] = a.seq;
        }
    }
    grid.sync();
    rec_mark(a.prof, 2);

    for (uint32_t p = tid; p < P; p += blockDim.x) {  // counts -> running ranks of each warp's next row
        uint32_t b = __ldcg(&a.tot[(size_t)p * G + cta]);
#pragma unroll
        for (uint32_t w = 0; w < kSmallWarps; ++w) {
            const uint32_t t = cnt[w * P + p];
            cnt[w * P + p] = b;
            b += t;
        }
    }
    __syncthreads();
    const bool ords = a.want_ordinals && a.vm_ordinal;
    for (uint32_t v0 = r0; v0 < r1; v0 += 32 * U) {
        int32_t pp[U];
        uint32_t kv[U], voff[U], moff[U], cc[U], cur[U], woff[U], ord[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint32_t v = v0 + u * 32 + lane;
            pp[u] = v < r1 ? a.pool_of[v] : -1;  // this thread's own store of phase A
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const bool m = pp[u] >= 0;
            const uint32_t p = m ? (uint32_t)pp[u] : 0u, v = v0 + u * 32 + lane;
            kv[u] = m ? __ldcg(&a.kvict[p]) : 0u;
            voff[u] = m ? (uint32_t)__ldcg(&a.victim_off[p]) : 0u;
            moff[u] = (m && a.want_members) ? (uint32_t)__ldcg(&a.member_off[p]) : 0u;
            cc[u] = (m && ords) ? (uint32_t)__ldcg(&a.create_count[p]) : 0u;
            cur[u] = (m && ords) ? (uint32_t)__ldcg(&a.current[p]) : 0u;
            woff[u] = (m && ords) ? __ldcg(&a.word_off[p]) : 0u;
            ord[u] = (m && ords) ? __ldg(&a.vm_ordinal[v]) : K8SGPU_NONE_U32;
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint32_t v = v0 + u * 32 + lane;
            const uint32_t key = pp[u] >= 0 ? (uint32_t)pp[u] : K8SGPU_NONE_U32;
            const uint32_t peers = __match_any_sync(0xFFFFFFFFu, key);
            uint32_t rank = 0;
            if (key != K8SGPU_NONE_U32) rank = mine[key] + (uint32_t)__popc(peers & lt);  // rows of this pool before v
            __syncwarp();
            if (key != K8SGPU_NONE_U32) {
                if ((peers & lt) == 0u) mine[key] += (uint32_t)__popc(peers);
                if (rank < kv[u]) a.victim_vm[voff[u] + rank] = v;  // existingVMs[:scaleDownCount], README.md:214
                if (a.want_members) a.member_vm[moff[u] + rank] = v;
                if (cc[u] && ord[u] < cur[u] + cc[u]) atomicOr(&a.bitmap[woff[u] + (ord[u] >> 5)], 1u << (ord[u] & 31u));
            }
            __syncwarp();
        }
    }
    if (!a.want_ordinals) {
        rec_mark(a.prof, 3);
        return;
    }
    grid.sync();
    rec_mark(a.prof, 3);
    for (uint32_t p = cta * kSmallWarps + warp; p < P; p += G * kSmallWarps)  // one warp per pool
        ordinal_emit_pool(p, a.current, a.create_count, a.word_off, a.bitmap, a.create_off, a.create_ordinal);
    rec_mark(a.prof, 4);
}

// ------------------------------------------------------------------ paths 1 and 2: many pools

// Tag filter, 4 VM rows per thread (128-bit loads and stores when the columns are 16-byte
// aligned), and currentCount by one RED per managed row (counts only: order-free).  With ~10^6
// pools the pass is bound by the L2 atomic rate (measured ~1.4 x 10^11 per second on B200), not by
// HBM.  (Folding the owner-id gather and the count into one 64-bit atomic that returns the pool row
// was tried: an atomic with a return value costs what the gather plus the RED cost -- no gain.)
__global__ void __launch_bounds__(256) match_count_kernel(const uint32_t *__restrict__ owner,
                                                          const uint32_t *__restrict__ flags, uint32_t V,
                                                          uint32_t n_owner_ids, const int32_t *__restrict__ lut,
                                                          int32_t *__restrict__ pool_of, int32_t *__restrict__ current,
                                                          int vec) {
    const uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;  // quad index
    const uint32_t v0 = q * 4u;
    if (v0 >= V) return;
    if (vec && v0 + 4u <= V) {
        const uint4 o = __ldg(reinterpret_cast<const uint4 *>(owner) + q);
        const uint4 f = __ldg(reinterpret_cast<const uint4 *>(flags) + q);
        int4 p;
        p.x = match_one(o.x, f.x, n_owner_ids, lut);
        p.y = match_one(o.y, f.y, n_owner_ids, lut);
        p.z = match_one(o.z, f.z, n_owner_ids, lut);
        p.w = match_one(o.w, f.w, n_owner_ids, lut);
        reinterpret_cast<int4 *>(pool_of)[q] = p;
        if (p.x >= 0) atomicAdd(&current[p.x], 1);
        if (p.y >= 0) atomicAdd(&current[p.y], 1);
        if (p.z >= 0) atomicAdd(&current[p.z], 1);
        if (p.w >= 0) atomicAdd(&current[p.w], 1);
    } else {
        for (uint32_t v = v0; v < V && v < v0 + 4u; ++v) {
            int32_t p = match_one(owner[v], flags[v], n_owner_ids, lut);
            pool_of[v] = p;
            if (p >= 0) atomicAdd(&current[p], 1);
        }
    }
}

// counters[2] += pools that scale down (the pruning pass stops when it reaches 0)
__global__ void pool_kernel(const int32_t *__restrict__ desired, uint32_t P, const int32_t *__restrict__ current,
                            int32_t *__restrict__ ready, int32_t *__restrict__ delta,
                            int32_t *__restrict__ create_count, uint32_t *__restrict__ kvict,
                            uint32_t *__restrict__ span_words, int32_t *__restrict__ need,
                            uint32_t *__restrict__ bits, uint32_t *__restrict__ counters) {
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t down = 0;
    if (p < P) {
        const int32_t cur = current[p];  // README.md:195
        const PoolVerdict r = pool_math(cur, desired[p]);
        ready[p] = cur;  // README.md:225
        delta[p] = r.delta;
        create_count[p] = r.create_count;
        kvict[p] = r.kvict;
        span_words[p] = r.span_words;
        need[p] = (int32_t)r.kvict;
        down = r.kvict ? 1u : 0u;
    }
    // thread p of a warp holds pool p: the ballot IS word p / 32 of the "still short" bitmap
    const uint32_t word = __ballot_sync(0xFFFFFFFFu, down);
    if ((threadIdx.x & 31u) == 0) {
        if (p < P) bits[p >> 5] = word;
        if (word) atomicAdd(&counters[2], (uint32_t)__popc(word));
    }
}

__global__ void keys_from_pool_kernel(const int32_t *__restrict__ pool_of, uint32_t V, uint32_t P,
                                      uint32_t *__restrict__ keys, uint32_t *__restrict__ vals) {
    uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= V) return;
    const int32_t p = pool_of[v];
    keys[v] = p < 0 ? P : (uint32_t)p;  // unmanaged rows sort behind every pool
    vals[v] = v;
}

// vmsToDelete := existingVMs[:scaleDownCount]  (README.md:214).  (keys, vals) = rows grouped by
// pool in list order; first[p] = where pool p's run starts.
__global__ void victim_gather_kernel(const uint32_t *__restrict__ keys, const uint32_t *__restrict__ vals, uint32_t n,
                                     uint32_t P, const uint32_t *__restrict__ first,
                                     const uint32_t *__restrict__ kvict, const int32_t *__restrict__ victim_off,
                                     uint32_t *__restrict__ victim_vm) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t p = keys[i];
    if (p >= P) return;
    uint32_t r = i - first[p];
    if (r < kvict[p]) victim_vm[(uint32_t)victim_off[p] + r] = vals[i];
}

__global__ void ordinal_mark_kernel(const int32_t *__restrict__ pool_of, uint32_t V,
                                    const uint32_t *__restrict__ ordinal, const int32_t *__restrict__ current,
                                    const int32_t *__restrict__ create_count, const uint32_t *__restrict__ word_off,
                                    uint32_t *__restrict__ bitmap) {
    uint32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= V) return;
    const int32_t p = pool_of[v];
    if (p < 0) return;
    const uint32_t cc = (uint32_t)create_count[p];
    if (!cc) return;
    const uint32_t o = ordinal[v];
    if (o < (uint32_t)current[p] + cc) atomicOr(&bitmap[word_off[p] + (o >> 5)], 1u << (o & 31u));
}

// Pruning pass of path 1.  A row can only be a victim while its pool has seen fewer than
// scaleDownCount rows, so the table is walked in S row segments, in order, and a row of segment s
// is kept as a CANDIDATE iff its pool was still short of victims when segment s began -- one bit
// per pool, a value that depends only on the counts of earlier segments.  The candidates of a pool
// are therefore a prefix, in list order, of that pool's rows, at least scaleDownCount long: their
// stable sort by pool yields the victims.  The pass stops at the first segment boundary where no
// pool is short any more -- on config-2-like data after about half the table, with ~8 % of the
// rows kept.
//   start of a segment: every CTA copies the "still short" bitmap (P bits) into shared memory, so
//            the test of a row costs no L2 request (SMEM_BITS; pools beyond ~1.6 M: read from L2)
//   phase 1: candidate flags (bit mask in shared memory), need[p] -= 1 per candidate (RED)
//   phase 2: candidates written in list order (CTA bases from the per-CTA counts); a pool whose
//            need reached 0 has its bit cleared, and whoever cleared it takes it off counters[2]
struct PruneArgs {
    const int32_t *pool_of;
    int32_t *need;     // scaleDownCount minus the pool's candidates so far (may end below 0)
    uint32_t *bits;    // bit p: pool p is still short of victims
    uint32_t *cta_cnt, *cand_keys, *cand_vals, *counters;
    uint32_t V, segments, rows_per_cta /* multiple of 1024, per segment */, bit_words;
};
constexpr uint32_t kPruneThreads = 1024, kPruneWarps = kPruneThreads / 32;
constexpr uint32_t kPruneMaxRowsPerCta = 65536;  // 2048 mask words
constexpr uint32_t kPruneMaxSmemBitWords = 50 * 1024;  // 200 KB of bitmap next to 16 KB of masks: 1.6 M pools

template <bool SMEM_BITS>
__global__ void __launch_bounds__(kPruneThreads, 1) prune_kernel(PruneArgs a) {
    cg::grid_group grid = cg::this_grid();
    extern __shared__ uint32_t sbits[];                  // copy of a.bits (SMEM_BITS)
    __shared__ uint32_t mask[kPruneMaxRowsPerCta / 32];  // candidate bits: word w = rows lo + 32 w ..
    __shared__ uint32_t wpre[kPruneMaxRowsPerCta / 32];  // candidates of this CTA before that word
    __shared__ uint32_t red[kPruneWarps], red2[kPruneWarps], wtot[kPruneWarps];
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5, cta = blockIdx.x, G = gridDim.x;
    const uint32_t lt = (1u << lane) - 1u;
    const uint32_t tiles = a.rows_per_cta / kPruneThreads, words = tiles * kPruneWarps;
    const uint64_t seg_rows = (uint64_t)a.rows_per_cta * G;
    constexpr uint32_t U = 4;  // tiles whose loads are issued together
    uint32_t total = 0;  // candidates of all earlier segments (every CTA computes the same number)
    for (uint32_t s = 0; s < a.segments; ++s) {
        const uint64_t lo64 = (uint64_t)s * seg_rows + (uint64_t)cta * a.rows_per_cta;
        const uint32_t lo = lo64 < a.V ? (uint32_t)lo64 : a.V;
        const uint32_t hi = (uint64_t)lo + a.rows_per_cta < a.V ? lo + a.rows_per_cta : a.V;
        if (SMEM_BITS) {
            for (uint32_t i = tid; i < a.bit_words; i += kPruneThreads) sbits[i] = __ldcg(&a.bits[i]);
            __syncthreads();
        }
        uint32_t mine = 0;
        for (uint32_t t0 = 0; t0 < tiles; t0 += U) {
            int32_t p[U];
            bool cand[U];
#pragma unroll
            for (uint32_t u = 0; u < U; ++u) {
                const uint32_t v = lo + (t0 + u) * kPruneThreads + tid;
                p[u] = (t0 + u < tiles && v < hi) ? __ldg(&a.pool_of[v]) : -1;
            }
#pragma unroll
            for (uint32_t u = 0; u < U; ++u) {
                const uint32_t w = p[u] >= 0 ? ((uint32_t)p[u] >> 5) : 0u;
                const uint32_t word = p[u] >= 0 ? (SMEM_BITS ? sbits[w] : __ldcg(&a.bits[w])) : 0u;
                cand[u] = (word >> ((uint32_t)p[u] & 31u)) & 1u;
            }
#pragma unroll
            for (uint32_t u = 0; u < U; ++u) {
                if (cand[u]) atomicSub(&a.need[p[u]], 1);
                const uint32_t m = __ballot_sync(0xFFFFFFFFu, cand[u]);
                if (lane == 0 && t0 + u < tiles) {
                    mask[(t0 + u) * kPruneWarps + warp] = m;
                    mine += (uint32_t)__popc(m);
                }
            }
        }
        if (lane == 0) red[warp] = mine;
        __syncthreads();
        if (tid == 0) {
            uint32_t c = 0;
            for (uint32_t w = 0; w < kPruneWarps; ++w) c += red[w];
            a.cta_cnt[cta] = c;
        }
        grid.sync();
        // my base = candidates so far + those of the CTAs before me in this segment
        uint32_t before = 0, all = 0;
        for (uint32_t c = tid; c < G; c += kPruneThreads) {
            const uint32_t x = __ldcg(&a.cta_cnt[c]);
            all += x;
            before += c < cta ? x : 0u;
        }
        before = __reduce_add_sync(0xFFFFFFFFu, before);
        all = __reduce_add_sync(0xFFFFFFFFu, all);
        // exclusive prefix of the per-word candidate counts: thread i owns words 2i and 2i + 1
        const uint32_t w0 = tid * 2u;
        const uint32_t c0 = w0 < words ? (uint32_t)__popc(mask[w0]) : 0u;
        const uint32_t c1 = w0 + 1u < words ? (uint32_t)__popc(mask[w0 + 1u]) : 0u;
        const uint32_t winc = warp_incl_scan(c0 + c1, (int)lane);
        if (lane == 0) {
            red[warp] = before;
            red2[warp] = all;
        }
        if (lane == 31) wtot[warp] = winc;
        __syncthreads();
        before = all = 0;
        uint32_t woff = 0;
        for (uint32_t k = 0; k < kPruneWarps; ++k) {
            before += red[k];
            all += red2[k];
            woff += k < warp ? wtot[k] : 0u;
        }
        if (w0 < words) wpre[w0] = woff + winc - c0 - c1;
        if (w0 + 1u < words) wpre[w0 + 1u] = woff + winc - c1;
        __syncthreads();
        const uint32_t base = total + before;
        for (uint32_t t0 = 0; t0 < tiles; t0 += U) {
            bool is[U];
            uint32_t v[U], pos[U], p[U];
            int32_t nd[U];
#pragma unroll
            for (uint32_t u = 0; u < U; ++u) {
                const uint32_t w = (t0 + u) * kPruneWarps + warp;
                const uint32_t m = t0 + u < tiles ? mask[w] : 0u;
                is[u] = (m >> lane) & 1u;
                v[u] = lo + (t0 + u) * kPruneThreads + tid;
                pos[u] = is[u] ? base + wpre[w] + (uint32_t)__popc(m & lt) : 0u;
                p[u] = is[u] ? (uint32_t)__ldg(&a.pool_of[v[u]]) : 0u;
            }
#pragma unroll
            for (uint32_t u = 0; u < U; ++u) nd[u] = is[u] ? __ldcg(&a.need[p[u]]) : 1;
#pragma unroll
            for (uint32_t u = 0; u < U; ++u) {
                if (!is[u]) continue;
                a.cand_keys[pos[u]] = p[u];
                a.cand_vals[pos[u]] = v[u];
                if (nd[u] <= 0) {  // the pool has its victims: no later row of it is a candidate
                    const uint32_t bit = 1u << (p[u] & 31u);
                    if (atomicAnd(&a.bits[p[u] >> 5], ~bit) & bit) atomicSub(&a.counters[2], 1u);
                }
            }
        }
        total += all;
        grid.sync();
        if (__ldcg(&a.counters[2]) == 0u) break;  // no pool is short of victims: the rest of the table cannot matter
    }
    if (cta == 0 && tid == 0) a.counters[3] = total;
}

// candidates kept per pool = scaleDownCount - need (need may have gone below zero)
__global__ void cand_count_kernel(const uint32_t *__restrict__ kvict, const int32_t *__restrict__ need, uint32_t P,
                                  uint32_t *__restrict__ count) {
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p < P) count[p] = (uint32_t)((int32_t)kvict[p] - need[p]);
}

// ------------------------------------------------------------------ host side

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
    // Spec.Replicas carries +kubebuilder:validation:Minimum=0 (README.md:94).  Checked here, before
    // anything is launched: every list offset below is derived from it.
    uint64_t sum_desired = 0;
    for (size_t p = 0; p < P; ++p) {
        if (pools->desired[p] < 0)
            return fail(ctx, K8SGPU_E_ARG, "reconcile: negative desired replicas (pool row %zu: %d)", p, pools->desired[p]);
        sum_desired += (uint64_t)pools->desired[p];
    }
    if (sum_desired > 0x7FFFFFFFull) return fail(ctx, K8SGPU_E_ARG, "reconcile: sum of desired replicas exceeds 2^31");
    st.p = pools->p;
    st.v = vms->v;
    st.n_owner_ids = n_owner_ids;
    st.has_ordinal = vms->ordinal != nullptr;
    st.create_cap = sum_desired;
    rec_in_layout(P, V, st.in_off);
    rec_out_layout(P, V, sum_desired, st.out_off);
    K8S_CUDA(ctx, st.in_slab.reserve(st.in_off[kInTotal] * 4 + 64));
    K8S_CUDA(ctx, st.out_slab.reserve(st.out_off[kOutTotal] * 4 + 64));
    uint32_t *in = st.in_slab.as<uint32_t>();
    CopyList cl;
    cl.add(in + st.in_off[kInVmOwner], vms->owner_id, V * 4);
    cl.add(in + st.in_off[kInVmFlags], vms->flags, V * 4);
    if (st.has_ordinal) cl.add(in + st.in_off[kInVmOrdinal], vms->ordinal, V * 4);
    cl.add(in + st.in_off[kInDesired], pools->desired, P * 4);
    cl.add(in + st.in_off[kInPoolOwner], pools->owner_id, P * 4);
    K8S_CUDA(ctx, cl.issue(cudaMemcpyHostToDevice, s));
    ctx->stats.h2d_bytes += cl.total;
    // upper bound of the ordinal bitmap: sum over pools of ceil((cur + cc)/32) <= (V + sum desired)/32 + P
    const size_t bitmap_words = (V + sum_desired) / 32 + P + 1;
    K8S_CUDA(ctx, st.lut.reserve(((size_t)n_owner_ids + 1) * 4));
    K8S_CUDA(ctx, st.pool_tmp.reserve((6 * P + P / 32 + 16) * 4));
    K8S_CUDA(ctx, st.bitmap.reserve(bitmap_words * 4));
    K8S_CUDA(ctx, st.counters.reserve(64 * 4));
    // owner id -> pool row: part of the uploaded state (every run of these tables uses it)
    K8S_CUDA(ctx, cudaMemsetAsync(st.counters.p, 0, 64 * 4, s));
    K8S_CUDA(ctx, cudaMemsetAsync(st.lut.p, 0xFF, ((size_t)n_owner_ids + 1) * 4, s));
    if (P)
        K8S_LAUNCH(ctx, pool_lut_kernel, ceil_div((uint32_t)P, 256), 256, 0, s,
                   reinterpret_cast<const int32_t *>(in + st.in_off[kInDesired]), in + st.in_off[kInPoolOwner], (uint32_t)P,
                   n_owner_ids, st.lut.as<int32_t>(), st.counters.as<uint32_t>());
    st.uploaded = true;
    return K8SGPU_OK;
}

struct RecPtrs {
    const uint32_t *vm_owner, *vm_flags, *vm_ordinal, *pool_owner;
    const int32_t *desired;
    int32_t *pool_of, *current, *ready, *delta, *create_count, *victim_off, *member_off, *create_off;
    uint32_t *victim_vm, *member_vm, *create_ordinal;
    uint32_t *kvict, *span_words, *word_off, *cand_cnt, *cand_off, *bits;
    int32_t *need;
};
static RecPtrs rec_ptrs(const ReconcileState &st) {
    RecPtrs r;
    const uint32_t *in = st.in_slab.as<uint32_t>();
    uint32_t *out = st.out_slab.as<uint32_t>();
    r.vm_owner = in + st.in_off[kInVmOwner];
    r.vm_flags = in + st.in_off[kInVmFlags];
    r.vm_ordinal = st.has_ordinal ? in + st.in_off[kInVmOrdinal] : nullptr;
    r.desired = reinterpret_cast<const int32_t *>(in + st.in_off[kInDesired]);
    r.pool_owner = in + st.in_off[kInPoolOwner];
    r.pool_of = reinterpret_cast<int32_t *>(out + st.out_off[kOutPoolOf]);
    r.current = reinterpret_cast<int32_t *>(out + st.out_off[kOutCurrent]);
    r.ready = reinterpret_cast<int32_t *>(out + st.out_off[kOutReady]);
    r.delta = reinterpret_cast<int32_t *>(out + st.out_off[kOutDelta]);
    r.create_count = reinterpret_cast<int32_t *>(out + st.out_off[kOutCreateCount]);
    r.victim_off = reinterpret_cast<int32_t *>(out + st.out_off[kOutVictimOff]);
    r.member_off = reinterpret_cast<int32_t *>(out + st.out_off[kOutMemberOff]);
    r.create_off = reinterpret_cast<int32_t *>(out + st.out_off[kOutCreateOff]);
    r.victim_vm = out + st.out_off[kOutVictimVm];
    r.member_vm = out + st.out_off[kOutMemberVm];
    r.create_ordinal = out + st.out_off[kOutCreateOrdinal];
    const size_t P = st.p;
    uint32_t *t = st.pool_tmp.as<uint32_t>();
    r.kvict = t;
    r.span_words = t + P;
    r.word_off = t + 2 * P;          // P + 1
    r.need = reinterpret_cast<int32_t *>(t + 3 * P + 1);
    r.cand_cnt = t + 4 * P + 1;
    r.cand_off = t + 5 * P + 1;      // P + 1
    r.bits = t + ((6 * P + 2 + 3) & ~(size_t)3);  // P / 32 + 1 words, 16-byte aligned
    return r;
}

static int rec_check_flags(k8sgpu_ctx *ctx, uint32_t err) {
    if (err & kErrNegDesired) return fail(ctx, K8SGPU_E_ARG, "reconcile: negative desired replicas");
    if (err & kErrDupOwner) return fail(ctx, K8SGPU_E_ARG, "reconcile: two pools share an owner id");
    return K8SGPU_OK;
}

// Path 0: the whole pass as one cooperative launch.
static int rec_run_small(k8sgpu_ctx *ctx, cudaStream_t s) {
    ReconcileState &st = ctx->rec;
    const uint32_t P = st.p, V = st.v;
    const RecPtrs q = rec_ptrs(st);
    const size_t smem = (size_t)kSmallWarps * P * 4;
    if (!st.small_attr_set) {  // more than 48 KB of dynamic shared memory needs the opt-in
        K8S_CUDA(ctx, cudaFuncSetAttribute(reconcile_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)(kSmallWarps * kSmallMaxPools * 4)));
        st.small_attr_set = true;
    }
    if (st.small_occ == 0 || st.small_occ_smem != smem) {
        int occ = 0;
        K8S_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, reconcile_small_kernel, kSmallWarps * 32, smem));
        st.small_occ = occ > 0 ? occ : -1;
        st.small_occ_smem = smem;
    }
    const int per_sm = st.small_occ;
    if (per_sm < 1) return fail(ctx, K8SGPU_E_CUDA, "cooperative reconcile kernel cannot be resident (P=%u)", P);
    // CTAs: about 1024 rows each while the pass is latency bound (one per SM at config 2: measured
    // best, the grid barriers stay cheap), up to four per SM once there is bandwidth to fill.
    uint32_t cap = (uint32_t)ctx->sm_count * (uint32_t)(per_sm < 4 ? per_sm : 4);
    uint32_t grid = ceil_div(V ? V : 1u, 1024u);
    if (grid > cap) grid = cap;
    if (ctx->tune.rec_grid >= 1 && (uint32_t)ctx->tune.rec_grid <= (uint32_t)ctx->sm_count * (uint32_t)per_sm)
        grid = (uint32_t)ctx->tune.rec_grid;
    K8S_CUDA(ctx, st.tot.reserve(((size_t)P * grid + 4) * 4));
    RecSmallArgs a;
    a.vm_owner = q.vm_owner; a.vm_flags = q.vm_flags; a.vm_ordinal = q.vm_ordinal;
    a.desired = q.desired; a.lut = st.lut.as<int32_t>();
    a.pool_of = q.pool_of; a.current = q.current; a.ready = q.ready; a.delta = q.delta;
    a.create_count = q.create_count; a.victim_off = q.victim_off; a.member_off = q.member_off;
    a.create_off = q.create_off; a.victim_vm = q.victim_vm; a.member_vm = q.member_vm;
    a.create_ordinal = q.create_ordinal;
    a.tot = st.tot.as<uint32_t>();
    a.kvict = q.kvict; a.span_words = q.span_words; a.word_off = q.word_off;
    a.bitmap = st.bitmap.as<uint32_t>();
    a.counters = st.counters.as<uint32_t>();
    a.mailbox = ctx->d_mailbox + kRecMailbox;
    a.P = P; a.V = V; a.n_owner_ids = st.n_owner_ids;
    a.rows_per_warp = ceil_div(ceil_div(V ? V : 1u, grid * kSmallWarps), 32u) * 32u;
    a.bitmap_words = (uint32_t)(((size_t)V + st.create_cap) / 32 + P + 1);
    a.seq = ++ctx->mail_seq;
    a.want_members = st.want_members ? 1 : 0;
    a.want_ordinals = st.want_ordinals ? 1 : 0;
    a.prof = ctx->tune.rec_prof ? reinterpret_cast<unsigned long long *>(a.counters + 16) : nullptr;
    st.last_path = 0;
    st.last_grid = grid;
    st.members_src = q.member_vm;
    K8S_CUDA(ctx, cudaEventRecord(ctx->ev_begin, s));
    void *args[] = {&a};
    K8S_CUDA(ctx, cudaLaunchCooperativeKernel((void *)reconcile_small_kernel, dim3(grid), dim3(kSmallWarps * 32), args,
                                              smem, s));
    ctx->stats.kernel_launches++;
    K8S_CUDA(ctx, cudaEventRecord(ctx->ev_end, s));
    // the list lengths arrive in mapped host memory half way through the kernel
    volatile uint32_t *mb = ctx->h_pinned + 32 + kRecMailbox;
    int rc = wait_mailbox(ctx, mb, a.seq, s, "reconcile");
    if (rc) return rc;
    rc = rec_check_flags(ctx, mb[1]);
    if (rc) return rc;
    st.n_victims = mb[2];
    st.n_members = mb[3];
    st.n_creates = mb[4];
    if (ctx->tune.rec_prof) {  // developer aid: phase boundaries (CTA 0's clock), microseconds
        unsigned long long t[5];
        K8S_CUDA(ctx, cudaMemcpyAsync(t, a.counters + 16, sizeof t, cudaMemcpyDeviceToHost, s));
        K8S_CUDA(ctx, cudaStreamSynchronize(s));
        fprintf(stderr, "[reconcile one-launch P=%u V=%u grid=%u] filter+count %.1f pools %.1f rank %.1f names %.1f "
                        "kernel %.1f us\n",
                P, V, grid, (t[1] - t[0]) / 1e3, (t[2] - t[1]) / 1e3, (t[3] - t[2]) / 1e3,
                st.want_ordinals ? (t[4] - t[3]) / 1e3 : 0.0, (t[st.want_ordinals ? 4 : 3] - t[0]) / 1e3);
    }
    return K8SGPU_OK;
}

// Paths 1 and 2: chains of kernels for tables with many pools.
static int rec_run_large(k8sgpu_ctx *ctx, cudaStream_t s, bool full_sort) {
    ReconcileState &st = ctx->rec;
    const uint32_t P = st.p, V = st.v;
    const RecPtrs q = rec_ptrs(st);
    uint32_t *cnts = st.counters.as<uint32_t>();
    K8S_CUDA(ctx, st.sort_keys.reserve((size_t)V * 4 + 16));
    K8S_CUDA(ctx, st.sort_vals.reserve((size_t)V * 4 + 16));
    uint32_t *keys = st.sort_keys.as<uint32_t>(), *vals = st.sort_vals.as<uint32_t>();
    K8S_CUDA(ctx, cudaEventRecord(ctx->ev_begin, s));
    K8S_CUDA(ctx, cudaMemsetAsync(cnts + 1, 0, 7 * 4, s));  // [0] keeps the upload-time error flags
    if (P) K8S_CUDA(ctx, cudaMemsetAsync(q.current, 0, (size_t)P * 4, s));
    if (V) {
        const int vec = ((((uintptr_t)q.vm_owner | (uintptr_t)q.vm_flags | (uintptr_t)q.pool_of) & 15u) == 0) ? 1 : 0;
        K8S_LAUNCH(ctx, match_count_kernel, ceil_div(ceil_div(V, 4), 256), 256, 0, s, q.vm_owner, q.vm_flags, V,
                   st.n_owner_ids, st.lut.as<int32_t>(), q.pool_of, q.current, vec);
    }
    if (P) {
        K8S_LAUNCH(ctx, pool_kernel, ceil_div(P, 256), 256, 0, s, q.desired, P, q.current, q.ready, q.delta,
                   q.create_count, q.kvict, q.span_words, q.need, q.bits, cnts);
        int rc = exclusive_scan_u32(ctx, s, q.kvict, (uint32_t *)q.victim_off, P, st.scan_tmp, (uint32_t *)q.victim_off + P);
        if (rc) return rc;
        rc = exclusive_scan_u32(ctx, s, (const uint32_t *)q.create_count, (uint32_t *)q.create_off, P, st.scan_tmp,
                                (uint32_t *)q.create_off + P);
        if (rc) return rc;
        rc = exclusive_scan_u32(ctx, s, (const uint32_t *)q.current, (uint32_t *)q.member_off, P, st.scan_tmp,
                                (uint32_t *)q.member_off + P);
        if (rc) return rc;
    } else {
        K8S_CUDA(ctx, cudaMemsetAsync(q.victim_off, 0, 4, s));
        K8S_CUDA(ctx, cudaMemsetAsync(q.create_off, 0, 4, s));
        K8S_CUDA(ctx, cudaMemsetAsync(q.member_off, 0, 4, s));
    }
    st.last_segments = st.last_candidates = 0;
    if (V && P && full_sort) {
        // rows grouped by pool, in list order: Status.VMs of every pool, and the victims at the
        // head of each run
        st.last_path = 2;
        K8S_LAUNCH(ctx, keys_from_pool_kernel, ceil_div(V, 256), 256, 0, s, q.pool_of, V, P, keys, vals);
        int rc = sort_pairs(ctx, s, keys, vals, V, key_bits(P), st.sort);
        if (rc) return rc;
        K8S_LAUNCH(ctx, victim_gather_kernel, ceil_div(V, 256), 256, 0, s, keys, vals, V, P,
                   (const uint32_t *)q.member_off, q.kvict, q.victim_off, q.victim_vm);
        st.members_src = vals;
    } else if (V && P) {
        st.last_path = 1;
        const uint32_t bit_words = P / 32u + 1u;
        const bool smem_bits = bit_words <= kPruneMaxSmemBitWords;
        const size_t smem = smem_bits ? (size_t)bit_words * 4 : 0;
        void *kern = smem_bits ? (void *)prune_kernel<true> : (void *)prune_kernel<false>;
        if (smem_bits && !st.prune_attr_set) {  // more than 48 KB of dynamic shared memory needs the opt-in
            K8S_CUDA(ctx, cudaFuncSetAttribute(prune_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                               (int)(kPruneMaxSmemBitWords * 4)));
            st.prune_attr_set = true;
        }
        // one CTA of 1024 threads per SM (the bitmap copy may fill its shared memory)
        uint32_t grid = (uint32_t)ctx->sm_count;
        if (grid > ceil_div(V, kPruneThreads)) grid = ceil_div(V, kPruneThreads);
        if (ctx->tune.rec_grid >= 1 && (uint32_t)ctx->tune.rec_grid <= (uint32_t)ctx->sm_count) grid = (uint32_t)ctx->tune.rec_grid;
        // Segments: few enough that the two grid barriers per segment stay small beside the rows
        // walked, many enough that the pass can stop early and that a pool overshoots its victim
        // count by little; a CTA's share of a segment must fit the shared-memory mask.
        uint32_t segs = V / (grid * 32768u);
        segs = segs < 4u ? 4u : (segs > 32u ? 32u : segs);
        if (ctx->tune.rec_segments >= 1) segs = (uint32_t)ctx->tune.rec_segments;
        uint32_t rows_per_cta = ceil_div(ceil_div(V, segs), grid);
        rows_per_cta = ceil_div(rows_per_cta, kPruneThreads) * kPruneThreads;
        if (rows_per_cta > kPruneMaxRowsPerCta) rows_per_cta = kPruneMaxRowsPerCta;
        segs = ceil_div(V, rows_per_cta * grid);  // V < 2^31, grid >= 1: fits
        K8S_CUDA(ctx, st.cta_cnt.reserve((size_t)grid * 4 + 16));
        PruneArgs a;
        a.pool_of = q.pool_of;
        a.need = q.need;
        a.bits = q.bits;
        a.cta_cnt = st.cta_cnt.as<uint32_t>();
        a.cand_keys = keys;
        a.cand_vals = vals;
        a.counters = cnts;
        a.V = V;
        a.segments = segs;
        a.rows_per_cta = rows_per_cta;
        a.bit_words = bit_words;
        void *args[] = {&a};
        K8S_CUDA(ctx, cudaLaunchCooperativeKernel(kern, dim3(grid), dim3(kPruneThreads), args, smem, s));
        ctx->stats.kernel_launches++;
        K8S_CUDA(ctx, cudaMemcpyAsync(ctx->h_pinned + 8, cnts + 3, 4, cudaMemcpyDeviceToHost, s));
        K8S_CUDA(ctx, cudaStreamSynchronize(s));  // the sort's launch sizes depend on the candidate count
        const uint32_t n_cand = ctx->h_pinned[8];
        st.last_grid = grid;
        st.last_segments = segs;
        st.last_candidates = n_cand;
        int rc = sort_pairs(ctx, s, keys, vals, n_cand, key_bits(P), st.sort);
        if (rc) return rc;
        K8S_LAUNCH(ctx, cand_count_kernel, ceil_div(P, 256), 256, 0, s, q.kvict, q.need, P, q.cand_cnt);
        rc = exclusive_scan_u32(ctx, s, q.cand_cnt, q.cand_off, P, st.scan_tmp, q.cand_off + P);
        if (rc) return rc;
        if (n_cand)
            K8S_LAUNCH(ctx, victim_gather_kernel, ceil_div(n_cand, 256), 256, 0, s, keys, vals, n_cand, P, q.cand_off,
                       q.kvict, q.victim_off, q.victim_vm);
        st.members_src = nullptr;
    } else {
        st.last_path = full_sort ? 2 : 1;
        st.members_src = vals;  // (no managed rows: nothing is read from it)
    }
    if (st.want_ordinals && P) {
        int rc = exclusive_scan_u32(ctx, s, q.span_words, q.word_off, P, st.scan_tmp, nullptr);
        if (rc) return rc;
        K8S_CUDA(ctx, cudaMemsetAsync(st.bitmap.p, 0, (((size_t)V + st.create_cap) / 32 + P + 1) * 4, s));
        if (st.has_ordinal && V)
            K8S_LAUNCH(ctx, ordinal_mark_kernel, ceil_div(V, 256), 256, 0, s, q.pool_of, V, q.vm_ordinal, q.current,
                       q.create_count, q.word_off, st.bitmap.as<uint32_t>());
        K8S_LAUNCH(ctx, ordinal_emit_kernel, ceil_div(P, 8), 256, 0, s, P, q.current, q.create_count, q.word_off,
                   st.bitmap.as<uint32_t>(), q.create_off, q.create_ordinal);
    }
    K8S_CUDA(ctx, cudaEventRecord(ctx->ev_end, s));
    // totals + error flag: one small copy each (this path is for large tables; latency is not the issue)
    K8S_CUDA(ctx, cudaMemcpyAsync(ctx->h_pinned, cnts, 4, cudaMemcpyDeviceToHost, s));
    K8S_CUDA(ctx, cudaMemcpyAsync(ctx->h_pinned + 1, q.victim_off + P, 4, cudaMemcpyDeviceToHost, s));
    K8S_CUDA(ctx, cudaMemcpyAsync(ctx->h_pinned + 2, q.member_off + P, 4, cudaMemcpyDeviceToHost, s));
    K8S_CUDA(ctx, cudaMemcpyAsync(ctx->h_pinned + 3, q.create_off + P, 4, cudaMemcpyDeviceToHost, s));
    K8S_CUDA(ctx, cudaStreamSynchronize(s));
    int rc = rec_check_flags(ctx, ctx->h_pinned[0]);
    if (rc) return rc;
    st.n_victims = ctx->h_pinned[1];
    st.n_members = ctx->h_pinned[2];
    st.n_creates = ctx->h_pinned[3];
    return K8SGPU_OK;
}

static int rec_run(k8sgpu_ctx *ctx, int want_members, int want_ordinals) {
    ReconcileState &st = ctx->rec;
    cudaStream_t s = ctx->stream;
    if (!st.uploaded) return fail(ctx, K8SGPU_E_ARG, "reconcile: no tables uploaded");
    st.want_members = want_members != 0;
    st.want_ordinals = want_ordinals != 0;
    st.ran = false;
    // Few pools: one launch.  Many pools: the pruned ranking, or the full stable sort when the
    // caller wants every pool's member list (Status.VMs).  k8sgpu_set_tuning("rec_path") forces one.
    int path = st.p <= kSmallMaxPools ? 0 : (st.want_members ? 2 : 1);
    if (ctx->tune.rec_path >= 0 && ctx->tune.rec_path <= 2) path = ctx->tune.rec_path;
    if (path == 0 && st.p > kSmallMaxPools) path = st.want_members ? 2 : 1;
    if (path == 1 && st.want_members) path = 2;
    int rc = path == 0 ? rec_run_small(ctx, s) : rec_run_large(ctx, s, path == 2);
    if (rc) return rc;
    st.ran = true;
    return K8SGPU_OK;
}

// Fills stats.device_ms (needs the run's kernels to have finished).
static int rec_time(k8sgpu_ctx *ctx) {
    float ms = 0.f;
    K8S_CUDA(ctx, cudaEventSynchronize(ctx->ev_end));
    K8S_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev_begin, ctx->ev_end));
    ctx->stats.device_ms = ms;
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
    const bool members = out->member_off && out->member_vm && st.want_members;
    const bool ordinals = out->create_off && out->create_ordinal && st.want_ordinals;
    if (ordinals && st.n_creates > out->create_capacity)
        return fail(ctx, K8SGPU_E_CAPACITY, "reconcile: create_ordinal needs %u elements, caller gave %llu",
                    st.n_creates, (unsigned long long)out->create_capacity);
    if (members && !st.members_src) return fail(ctx, K8SGPU_E_ARG, "reconcile: member lists were not produced by the last run");
    const RecPtrs q = rec_ptrs(st);
    CopyList cl;  // in the order of the device block: adjacent host buffers merge into one copy
    cl.add(out->pool_of, q.pool_of, V * 4);
    cl.add(out->current, q.current, P * 4);
    cl.add(out->ready, q.ready, P * 4);  // README.md:225: ReadyReplicas = currentCount
    cl.add(out->delta, q.delta, P * 4);
    cl.add(out->create_count, q.create_count, P * 4);
    cl.add(out->victim_off, q.victim_off, (P + 1) * 4);
    if (members) cl.add(out->member_off, q.member_off, (P + 1) * 4);
    if (ordinals) cl.add(out->create_off, q.create_off, (P + 1) * 4);
    cl.add(out->victim_vm, q.victim_vm, (size_t)st.n_victims * 4);
    if (members) cl.add(out->member_vm, st.members_src, (size_t)st.n_members * 4);
    if (ordinals) cl.add(out->create_ordinal, q.create_ordinal, (size_t)st.n_creates * 4);
    K8S_CUDA(ctx, cl.issue(cudaMemcpyDeviceToHost, s));
    ctx->stats.d2h_bytes += cl.total;
    K8S_CUDA(ctx, cudaStreamSynchronize(s));
    return rec_time(ctx);
}

}  // namespace k8sgpu

using namespace k8sgpu;

extern "C" int k8sgpu_reconcile_out_size(uint32_t p, uint32_t v, uint64_t create_capacity, k8sgpu_reconcile_layout *out) {
    if (!out) return K8SGPU_E_ARG;
    uint64_t off[12];
    rec_out_layout(p, v, create_capacity, off);
    out->pool_of = off[kOutPoolOf];
    out->current = off[kOutCurrent];
    out->ready = off[kOutReady];
    out->delta = off[kOutDelta];
    out->create_count = off[kOutCreateCount];
    out->victim_off = off[kOutVictimOff];
    out->member_off = off[kOutMemberOff];
    out->create_off = off[kOutCreateOff];
    out->victim_vm = off[kOutVictimVm];
    out->member_vm = off[kOutMemberVm];
    out->create_ordinal = off[kOutCreateOrdinal];
    out->total = off[kOutTotal];
    return K8SGPU_OK;
}

extern "C" int k8sgpu_reconcile_upload(k8sgpu_ctx *ctx, const k8sgpu_pool_table *pools, const k8sgpu_vm_table *vms,
                                       uint32_t n_owner_ids) {
    K8S_ENTER(ctx);
    reset_call_stats(ctx);
    int rc = rec_upload(ctx, pools, vms, n_owner_ids);
    if (rc) return rc;
    K8S_CUDA(ctx, cudaMemcpyAsync(ctx->h_pinned, ctx->rec.counters.p, 4, cudaMemcpyDeviceToHost, ctx->stream));
    K8S_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    rc = rec_check_flags(ctx, ctx->h_pinned[0]);
    if (rc) ctx->rec.uploaded = false;
    return rc;
}

extern "C" int k8sgpu_reconcile_run(k8sgpu_ctx *ctx, int want_members, int want_ordinals) {
    K8S_ENTER(ctx);
    reset_call_stats(ctx);
    int rc = rec_run(ctx, want_members, want_ordinals);
    if (rc) return rc;
    return rec_time(ctx);
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
