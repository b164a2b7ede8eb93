// place.cu -- round-based pod->node placement on one B200 (DESIGN.md policy v0).
//
// The reference has no scheduler source: pods are handed to Volcano (schedulerName: volcano,
// GPU调度平台搭建.md:649).  Per round:
//   node_prep   : current free columns -> packed 32-B node records (one per node, padded to tiles)
//   score       : every active pod x every node: filter (a11) + key (a12) + running argmax.
//                 Node records are staged through shared memory in 8 KB tiles by TMA bulk copies
//                 (cp.async.bulk + mbarrier, double buffered); each thread keeps its pods in
//                 registers and reads the tile by warp-broadcast LDS.128.
//   finalize    : packed best key -> node row
//   admit       : stable sort of proposals by node, segmented prefix sums in pod order,
//                 S <= free test (a13), free -= admitted, stable compaction of the active list.
// In the in-library loop (k8sgpu_place / _run) finalize, admit and the next round's node_prep are
// ONE cooperative launch (admit_coop_kernel); the round-at-a-time API used across GPUs feeds the
// same kernel the exchanged proposals (K8SGPU_COOP=0: separate kernels).  k8sgpu_place_gang wraps the same loop in reject / release / retry
// iterations.  Integer only; no atomics decide any result except commutative max / add.
#include "prims.cuh"

namespace k8sgpu {

struct __align__(16) NodeQ {  // what the inner loop reads; rebuilt from the free columns each round
    // first half: read only by warps with a pod that restricts SKUs or zones, or whose toleration
    // clause has to be evaluated (eval_pair4 CLS bit 0, TT = 2)
    uint32_t sku_bit;     // 1 << sku_id
    uint32_t zone_bit;    // 1 << (31 - zone_id): bit-reversed one-hot (see eval_pair RES = 3)
    uint32_t taint;
    uint32_t free_gpu_s;  // free_gpu, or 0xFFFFFFFF (= -1 signed) when not schedulable (RES < 4 only)
    // second half: everything the other warps need -- one 128-bit shared-memory load per node
    uint32_t free_mcpu_s; // free_mcpu, or 0xFFFFFFFF (= -1 signed) when the node is not schedulable
    uint32_t free_mib;
    uint32_t khi;         // 1<<31 | (0x7FFF - free_gpu) << 16 | (255 - cap_gpu) << 8 | (63 - n % 64) << 1
                          // (RES = 4 also takes the free-GPU test from here: see eval_pair4)
    uint32_t zrot;        // bits 0..4: (zone_id - 7) mod 32, rotates the pod's one-hot preferred zone onto bit 7
                          // (the funnel shift reads only these five bits)
                          // bits 5..12: min(free_gpu, 127), or -1 when the node is not schedulable (eval R, TT = 0)
                          // bits 13..31: free_gpu, or -1 when not schedulable OR tainted (TT = 1)
};
static_assert(sizeof(NodeQ) == 32, "node record is 32 B: the algorithmic bytes per eval");

struct __align__(16) PodQ {
    uint32_t req_gpu, req_mcpu, req_mib;
    uint32_t nsku, nzone, ntol;  // complements of sku_mask / zone_mask (bit-reversed) / tol_mask
    uint32_t pzb;                // 1 << pref_zone, or 0
    uint32_t rot;                // rot(p), DESIGN.md a12
};
static_assert(sizeof(PodQ) == 32, "pod record is 32 B");

constexpr uint32_t kErrNode = 1u, kErrPod = 2u, kErrProposal = 4u;
constexpr uint32_t kSlotsPerNode = 8;  // slot admission (admit_coop_kernel): first proposers kept per node
// counters[5] != 0 rules the slot admission out for this run: 1 = some pod asks for 0 GPUs,
// 2 = some node has more than kSlotsPerNode free GPUs
constexpr int kSlotVeto = 5;
constexpr int kTaintUnion = 6;  // counters[6]: OR of every node's taint mask (per run, before pod_prep)
constexpr int kLikeRuns = 7;    // counters[7]: how many distinct (class, hash) values the pods' sort keys hold
constexpr uint32_t kLikeBits = 12, kLikeMinRun = 1024;
constexpr int kResWide = 8;     // counters[8] != 0: some schedulable node has >= kResGpus free GPUs (no resource summary)
constexpr uint32_t kResGpus = 64;  // the summary below covers free_gpu 0 .. 63

__device__ __forceinline__ uint64_t mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// ------------------------------------------------------------------ table preparation

__global__ void taint_union_kernel(const uint32_t *__restrict__ taint, uint32_t n, uint32_t *__restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t v = __reduce_or_sync(0xFFFFFFFFu, i < n ? taint[i] : 0u);
    if ((threadIdx.x & 31u) == 0u && v) atomicOr(out, v);
}
// How a pod's toleration test can be scored (eval_pair4 TT), given the union of the node taints:
// 0 = it tolerates every taint that occurs (the clause is always true), 1 = it tolerates none of
// them (the clause is "the node has no taint"), 2 = anything else (the clause is evaluated).
__device__ __forceinline__ uint32_t taint_class(uint32_t ntol, uint32_t tunion) {
    const uint32_t x = ntol & tunion;
    return x == 0u ? 0u : (x == tunion ? 1u : 2u);
}

// Also, when the run keeps a scoring order (see "scoring order" below): the pod's sort key
// (eval_pair4 class, CLS | TT << 2; optionally a hash of the pod's masks and preferred zone; where in
// the node table rot(p) falls -- 24 bits), its own index as the
// sort payload, and posof[p] = p (every pod starts active at its own position).
__global__ void __launch_bounds__(256) pod_prep_kernel(const uint32_t *__restrict__ cols, uint32_t m, uint32_t n_nodes,
                                                       uint32_t policy, int mul_masks, PodQ *__restrict__ podq,
                                                       int32_t *__restrict__ assign, uint32_t *__restrict__ active,
                                                       uint32_t *__restrict__ err, uint32_t *__restrict__ sort_key,
                                                       uint32_t *__restrict__ sort_val, uint32_t *__restrict__ posof,
                                                       uint32_t hash_bits, uint32_t *__restrict__ seen) {
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= m) return;
    PodQ q;
    q.req_gpu = cols[0 * (size_t)m + p];
    q.req_mcpu = cols[1 * (size_t)m + p];
    q.req_mib = cols[2 * (size_t)m + p];
    q.nsku = ~cols[3 * (size_t)m + p];
    q.nzone = __brev(~cols[4 * (size_t)m + p]);  // bit-reversed, like NodeQ::zone_bit
    if (mul_masks) {  // eval_pair RES = 4: each mask is mirrored against its node-side one-hot word
        q.nsku = __brev(q.nsku);
        q.nzone = __brev(q.nzone);
    }
    q.ntol = ~cols[5 * (size_t)m + p];
    uint32_t pz = cols[6 * (size_t)m + p];
    q.pzb = pz < 32u ? (1u << pz) : 0u;
    q.rot = ((policy & K8SGPU_POLICY_ROT_ZERO) || n_nodes == 0)
                ? 0u
                : (uint32_t)(mix64((uint64_t)p + 0x9E3779B97F4A7C15ull) % (uint64_t)n_nodes);
    if ((q.req_gpu | q.req_mcpu | q.req_mib) & 0x80000000u) atomicOr(err, kErrPod);
    if (q.req_gpu == 0u) atomicOr(err + kSlotVeto, 1u);
    if (pz >= 32u && pz != K8SGPU_NONE_U32) atomicOr(err, kErrPod);
    podq[p] = q;
    assign[p] = K8SGPU_ASSIGN_PENDING;
    active[p] = p;
    if (sort_key) {
        const uint32_t tt = taint_class(q.ntol, err[kTaintUnion]);
        const uint32_t cls = ((q.nsku | q.nzone) ? 1u : 0u) | (q.pzb ? 2u : 0u) | (tt << 2);
        // pods that agree on the words eval_pair4_u2 shares sort together (12 hash bits inside the class)
        uint32_t h = q.nsku * 0x9E3779B1u ^ q.nzone * 0x85EBCA77u ^ q.pzb * 0xC2B2AE3Du ^ (tt == 2u ? q.ntol : 0u) * 0x27D4EB2Fu;
        h ^= h >> 15;
        h *= 0x2C1B3C6Du;
        h ^= h >> 12;
        // 20 bits under the class: hash_bits of the hash, the rest says where in the node table rot(p) falls
        const uint32_t rb = 20u - hash_bits;
        const uint32_t where = n_nodes ? (uint32_t)(((unsigned long long)q.rot << rb) / n_nodes) : 0u;
        const uint32_t hh = hash_bits ? h >> (32u - hash_bits) : 0u;
        // (dearest classes first: the grid's last CTAs, which no other work can hide, are then the cheap ones)
        sort_key[p] = ((15u - cls) << 20) | (hh << rb) | where;
        if (seen) {  // count the distinct (class, hash) values: one atomic per value, then plain reads
            uint32_t *sn = seen + ((cls << kLikeBits) | hh);
            if (*(volatile uint32_t *)sn == 0u && atomicExch(sn, 1u) == 0u) atomicAdd(err + kLikeRuns, 1u);
        }
        sort_val[p] = p;
        posof[p] = p;
    }
}

// The hash in the sort key pays while the runs of like pods stay long (see round_begin): with
// fewer than kLikeMinRun pods per distinct value on average it is cleared again, which leaves
// (class, the top 8 bits of where rot(p) falls).
__global__ void like_hash_fix_kernel(uint32_t *__restrict__ sort_key, uint32_t m, const uint32_t *__restrict__ counters) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p < m && (unsigned long long)counters[kLikeRuns] * kLikeMinRun > m) sort_key[p] &= ~(((1u << kLikeBits) - 1u) << (20u - kLikeBits));
}

__device__ __forceinline__ NodeQ make_nodeq(const uint32_t *__restrict__ cols, const uint32_t *__restrict__ freec,
                                            uint32_t n, uint32_t i, uint32_t *__restrict__ err, int validate) {
    NodeQ q;
    if (i < n) {
        uint32_t fg = freec[i], fc = freec[(size_t)n + i], fm = freec[2 * (size_t)n + i];
        uint32_t cap = cols[3 * (size_t)n + i], sku = cols[4 * (size_t)n + i], zone = cols[5 * (size_t)n + i];
        uint32_t taint = cols[6 * (size_t)n + i], flags = cols[7 * (size_t)n + i];
        if (validate && (fg > 0x7FFFu || cap > 255u || sku >= 32u || zone >= 32u || ((fc | fm) & 0x80000000u)))
            atomicOr(err, kErrNode);
        if (validate && fg > kSlotsPerNode) atomicOr(err + kSlotVeto, 2u);
        q.free_gpu_s = (flags & 1u) ? fg : 0xFFFFFFFFu;
        q.free_mcpu_s = (flags & 1u) ? fc : 0xFFFFFFFFu;
        q.free_mib = fm;
        q.sku_bit = 1u << (sku & 31u);
        q.zone_bit = 0x80000000u >> (zone & 31u);
        q.taint = taint;
        q.khi = 0x80000000u | ((0x7FFFu - (fg & 0x7FFFu)) << 16) | ((255u - (cap & 255u)) << 8) |
                ((63u - (i & 63u)) << 1);
        q.zrot = ((zone - 7u) & 31u) | ((((flags & 1u) ? min(fg, 127u) : 0xFFu) & 0xFFu) << 5) |
                 ((((flags & 1u) && taint == 0u) ? fg : 0xFFFFFFFFu) << 13);
    } else {  // padding up to a whole tile: never feasible
        q.free_gpu_s = q.free_mcpu_s = 0xFFFFFFFFu;
        q.free_mib = 0;
        q.sku_bit = q.zone_bit = q.taint = 0;
        q.khi = 0;
        q.zrot = 0xFFFFFFE0u;
    }
    return q;
}

// Resource summary of a round's node table: tab[g] = the least free_mcpu, tab[kResGpus + g] the
// least free_mib, among the schedulable nodes with exactly g free GPUs.  From it the scoring kernel
// decides, per pod, whether "enough GPUs" already implies "enough CPU and memory" at EVERY node
// (req_mcpu <= min over g' >= req_gpu of tab[g'], likewise memory): then the two other resource
// differences of that pod's evals are never negative where the first is not, and are left out.
// Called by whole warps (one node per lane); one atomic pair per distinct g per warp.
__device__ __forceinline__ void resmin_note(uint32_t *__restrict__ tab, uint32_t *__restrict__ counters, const NodeQ &q) {
    const bool sched = q.free_mcpu_s != 0xFFFFFFFFu;
    uint32_t g = sched ? q.free_gpu_s : 0xFFFFFFFFu;
    if (sched && g >= kResGpus) {
        atomicOr(counters + kResWide, 1u);
        g = 0xFFFFFFFFu;
    }
    const uint32_t peers = __match_any_sync(0xFFFFFFFFu, g);
    const uint32_t mc = __reduce_min_sync(peers, q.free_mcpu_s), mm = __reduce_min_sync(peers, q.free_mib);
    if (g != 0xFFFFFFFFu && (threadIdx.x & 31u) == (uint32_t)__ffs((int)peers) - 1u) {
        atomicMin(tab + g, mc);
        atomicMin(tab + kResGpus + g, mm);
    }
}

__global__ void __launch_bounds__(256) node_prep_kernel(const uint32_t *__restrict__ cols, const uint32_t *__restrict__ freec, uint32_t n,
                                 uint32_t n_pad, NodeQ *__restrict__ nodeq, uint32_t *__restrict__ err,
                                 int validate, uint32_t *__restrict__ resmin) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;  // (n_pad is a multiple of 256: whole warps)
    if (i >= n_pad) return;
    const NodeQ q = make_nodeq(cols, freec, n, i, err, validate);
    nodeq[i] = q;
    resmin_note(resmin, err, q);
}

// ------------------------------------------------------------------ TMA bulk copy + mbarrier

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// ------------------------------------------------------------------ scoring

template <int PPT> __device__ __forceinline__ uint4 ldrec(const uint4 *p);
// One 128-bit shared-memory load, kept whole: the class-specialised loops use only some words of a
// node record, and left to itself the compiler splits the load into narrower ones (3 LDS per node
// instead of 2 -- issue slots the loop does not have).
__device__ __forceinline__ uint4 lds128(const uint4 *p) {
    uint4 v;
    asm volatile("ld.volatile.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(smem_u32(p)));
    return v;
}
// Half a node record for the inner loops.  With four pods per thread one record feeds four evals,
// the loads are a smaller share of the issue slots, and the freedom to schedule them matters more
// (measured: config 4 is 6 % slower with the pinned loads), so that variant keeps plain loads.
template <int PPT> __device__ __forceinline__ uint4 ldrec(const uint4 *p) { return PPT >= 4 ? *p : lds128(p); }

// One pod x one node: filter (a11), key (a12), running max.  Written as PTX so that the
// feasibility clauses stay one predicate chain with no separate predicate-combine instruction:
// the loop is bound by the integer ALU pipe (ncu: >90 % busy), so every ALU instruction removed
// here is throughput.
//   a = {sku_bit, zone_bit, taint, free_gpu or -1}   b = {free_mcpu or -1, free_mib, khi, zrot}
//   (the two halves of a NodeQ; the RES < 4 bodies below were written for an older word order
//   and get the words renamed on the way in)
//   key = khi | (rotr(pzb, zrot) & 0x80): bit 7 of the rotated one-hot preferred-zone word is
//   set exactly when the node's zone is the preferred one (zrot = (zone_id - 7) mod 32).
//   khi carries 63 - (n mod 64) in bits 1..6, so inside a 64-node group the lowest n wins ties
//   and no node index has to be tracked per eval.
// RES = 0: three chained integer compares (ALU pipe).
// RES = 1: three differences on the FMA pipe (mad by a runtime -1), OR-ed; the sign bit of the
//          OR says "some resource is short".  One ALU instruction fewer per eval.
// RES = 2: as 1, and the running best is updated by a predicated mad (key * runtime 1 + 0), which
//          ptxas keeps on the FMA pipe instead of turning it into an ALU select: 8 ALU ops/eval.
// RES = 4: as 2, and the SKU and zone allow-masks are tested on the FMA pipe too: the pod's mask,
//          mirrored against the node's one-hot word, times that word is the mask shifted so that
//          the node's own bit lands on bit 31 -- a sign bit that joins the resource differences'
//          OR.  7 ALU + 6 FMA-pipe ops per eval.  (PodQ::nsku / nzone are stored mirrored for it.)
// RES = 3: as 2, and the zone-preference bit comes from two multiplies instead of a funnel shift:
//          pzb * zone_bit = 2^(31 + pref - zone) has bit 31 set exactly when pref == zone, and
//          mul.hi by 2^8 brings that bit to position 7.  7 ALU + 6 FMA-pipe ops per eval.
// CLS (RES = 4 only) drops clauses that cannot matter for ANY pod of the calling warp -- the host
// sorts each round's scoring order by pod class so that most warps are uniform:
//   bit 0: some pod restricts SKUs or zones (else the two mask multiplies and their OR go: -3)
//   bit 1: some pod prefers a zone         (else the funnel shift and the key LOP3 go:      -2)
// A pod without the feature has an all-zero mask / one-hot word, for which the dropped instructions
// compute the identity, so every CLS gives the same bits as CLS = 3 for the pods it is used on.
// The RES = 4 body, assembled from its clauses.  Operands (the same list for every variant; a
// variant that leaves a clause out is handed a register it uses anyway in that operand's place):
//   %0 best | %1 sku_bit %2 zone_bit %3 taint | %4 free_mcpu %5 free_mib %6 khi %7 zrot
//   %8 rq (the GPU request, see below) %9 rc %10 rm | %11 nsku %12 nzone %13 ntol %14 pzb
//   %15 runtime -1, %16 runtime 1
// The free-GPU test needs no word of its own.  TT != 1: khi holds 0x7FFF - free_gpu in bits 16..30
// under a set bit 31, so khi + (0x80000000 + (req << 16)) has bit 31 set exactly when
// 0x7FFF - free + req carries out of the field, i.e. when free < req (rq = minus that constant,
// req clamped to 0x8000 -- no node has more than 0x7FFF free).  TT = 1: zrot's bits 13..31 hold
// free_gpu, or -1 on a tainted node; rq = req << 13.  An unschedulable node has free_mcpu = -1.
#define K8S_EP4(FREEW, MASKMUL, TAND, TSETP, MASKOR, SIGN, PREF, KEY)                                              \
    asm("{\n\t"                                                                                                    \
        ".reg .pred pf;\n\t"                                                                                       \
        ".reg .b32 m, z, key, d1, d2, d3, d4, d5;\n\t"                                                             \
        "mad.lo.s32 d1, %8, %15, " FREEW ";\n\t"                                                                   \
        "mad.lo.s32 d2, %9, %15, %4;\n\t"                                                                          \
        "mad.lo.s32 d3, %10, %15, %5;\n\t" MASKMUL TAND "lop3.b32 d1, d1, d2, d3, 0xFE;\n\t" TSETP MASKOR SIGN PREF \
        "setp.gt.and.u32 pf, " KEY ", %0, pf;\n\t"                                                                 \
        "@pf mad.lo.u32 %0, " KEY ", %16, 0;\n\t"                                                                  \
        "}"                                                                                                        \
        : "+r"(best)                                                                                               \
        : "r"(o_sku), "r"(o_zone), "r"(o_taint), "r"(b.x), "r"(b.y), "r"(b.z), "r"(b.w), "r"(rq), "r"(rc), "r"(rm), \
          "r"(o_nsku), "r"(o_nzone), "r"(o_ntol), "r"(o_pzb), "r"(negone), "r"(one))
#define K8S_EP4_MUL "mul.lo.u32 d4, %11, %1;\n\tmul.lo.u32 d5, %12, %2;\n\t"
#define K8S_EP4_MOR "lop3.b32 d1, d1, d4, d5, 0xFE;\n\t"
#define K8S_EP4_TAND "and.b32 m, %3, %13;\n\t"
#define K8S_EP4_TSETP "setp.eq.u32 pf, m, 0;\n\t"
#define K8S_EP4_SIGN_AND "setp.ge.and.s32 pf, d1, 0, pf;\n\t"
#define K8S_EP4_SIGN "setp.ge.s32 pf, d1, 0;\n\t"
#define K8S_EP4_PREF "shf.r.wrap.b32 z, %14, %14, %7;\n\tlop3.b32 key, z, 0x80, %6, 0xEA;\n\t"
// CLS: which clauses some pod of the warp needs (bit 0: SKU / zone masks, bit 1: zone preference).
// TT: the taint clause -- 2: tested (and + setp); 0: left out (every pod of the warp tolerates every
// taint any node carries); 1: left out, free GPUs read from the field that is -1 on a tainted node
// (every pod of the warp tolerates none of the taints that occur).
// Instructions per eval: 14 - 3 [no masks] - 2 [no preference] - 2 [TT < 2], and with CLS bit 0
// clear and TT < 2 the first half of the node record is not even loaded.
template <int CLS, int TT> __host__ __device__ constexpr bool ep4_needs_first_half() { return (CLS & 1) || TT == 2; }
__device__ __forceinline__ uint32_t ep4_request(uint32_t req_gpu, bool tainted_field) {
    const uint32_t r = min(req_gpu, 0x8000u);
    return tainted_field ? r << 13 : 0u - (0x80000000u + (r << 16));
}
template <int CLS, int TT>
__device__ __forceinline__ void eval_pair4(const uint4 &a, const uint4 &b, uint32_t rq, uint32_t rc, uint32_t rm,
                                           uint32_t nsku, uint32_t nzone, uint32_t ntol, uint32_t pzb,
                                           uint32_t &best, uint32_t negone, uint32_t one) {
    const uint32_t o_sku = (CLS & 1) ? a.x : b.x, o_zone = (CLS & 1) ? a.y : b.x, o_taint = TT == 2 ? a.z : b.x;
    const uint32_t o_nsku = (CLS & 1) ? nsku : rc, o_nzone = (CLS & 1) ? nzone : rc, o_ntol = TT == 2 ? ntol : rc;
    const uint32_t o_pzb = (CLS & 2) ? pzb : rc;
#define K8S_EP4_T(MASKMUL, MASKOR, PREF, KEY)                                                                  \
    if (TT == 2) K8S_EP4("%6", MASKMUL, K8S_EP4_TAND, K8S_EP4_TSETP, MASKOR, K8S_EP4_SIGN_AND, PREF, KEY);     \
    else if (TT == 1) K8S_EP4("%7", MASKMUL, "", "", MASKOR, K8S_EP4_SIGN, PREF, KEY);                         \
    else K8S_EP4("%6", MASKMUL, "", "", MASKOR, K8S_EP4_SIGN, PREF, KEY)
    if (CLS == 0) {
        K8S_EP4_T("", "", "", "%6");
    } else if (CLS == 1) {
        K8S_EP4_T(K8S_EP4_MUL, K8S_EP4_MOR, "", "%6");
    } else if (CLS == 2) {
        K8S_EP4_T("", "", K8S_EP4_PREF, "key");
    } else {
        K8S_EP4_T(K8S_EP4_MUL, K8S_EP4_MOR, K8S_EP4_PREF, "key");
    }
#undef K8S_EP4_T
}

// Two pods of ONE thread against a node, when the two agree on their masks, preferred zone and
// (TT = 2) tolerations -- the scoring order sorts like pods together, so almost every thread
// qualifies.  What depends on the node and on those shared words alone is computed once: the
// mask / toleration test becomes one predicate (plain AND tests on un-mirrored masks: up to three
// terms, one LOP3 each), the key one funnel shift + LOP3; per pod remain the three resource
// differences, their OR, the sign test, the compare and the predicated update (7 instructions).
// One asm body for both pods, because the shared predicate cannot leave it.
//   %0 %1 best | %2 sku_bit %3 zone_bit %4 taint | %5 free_mcpu %6 free_mib %7 khi %8 zrot
//   %9 %10 rq | %11 %12 rc | %13 %14 rm | %15 sku mask %16 zone mask (complements, plain and
//   bit-reversed like the node words) %17 ntol %18 pzb | %19 runtime -1 | %20 %21 runtime 1
#define K8S_EP4U(FREEW, SHARED, SIGN0, SIGN1, PREF, KEY)                                                       \
    asm("{\n\t"                                                                                                \
        ".reg .pred pf0, pf1, pm;\n\t"                                                                         \
        ".reg .b32 m, z, key, x1, x2, x3, y1, y2, y3;\n\t"                                                     \
        "mad.lo.s32 x1, %9, %19, " FREEW ";\n\t"                                                               \
        "mad.lo.s32 y1, %10, %19, " FREEW ";\n\t"                                                              \
        "mad.lo.s32 x2, %11, %19, %5;\n\t"                                                                     \
        "mad.lo.s32 y2, %12, %19, %5;\n\t"                                                                     \
        "mad.lo.s32 x3, %13, %19, %6;\n\t"                                                                     \
        "mad.lo.s32 y3, %14, %19, %6;\n\t" SHARED "lop3.b32 x1, x1, x2, x3, 0xFE;\n\t"                         \
        "lop3.b32 y1, y1, y2, y3, 0xFE;\n\t" SIGN0 SIGN1 PREF "setp.gt.and.u32 pf0, " KEY ", %0, pf0;\n\t"      \
        "setp.gt.and.u32 pf1, " KEY ", %1, pf1;\n\t"                                                           \
        "@pf0 mad.lo.u32 %0, " KEY ", %20, 0;\n\t"                                                             \
        "@pf1 mad.lo.u32 %1, " KEY ", %21, 0;\n\t"                                                             \
        "}"                                                                                                    \
        : "+r"(best0), "+r"(best1)                                                                             \
        : "r"(o_sku), "r"(o_zone), "r"(o_taint), "r"(b.x), "r"(b.y), "r"(b.z), "r"(b.w), "r"(rq0), "r"(rq1),  \
          "r"(rc0), "r"(rc1), "r"(rm0), "r"(rm1), "r"(o_msku), "r"(o_mzone), "r"(o_ntol), "r"(o_pzb),         \
          "r"(negone), "r"(one0), "r"(one1))
#define K8S_EP4U_MASK "and.b32 m, %16, %3;\n\tlop3.b32 m, %15, %2, m, 0xEA;\n\t"
#define K8S_EP4U_TOL_MORE "lop3.b32 m, %4, %17, m, 0xEA;\n\t"
#define K8S_EP4U_TOL_ONLY "and.b32 m, %4, %17;\n\t"
#define K8S_EP4U_PM "setp.eq.u32 pm, m, 0;\n\t"
#define K8S_EP4U_PREF "shf.r.wrap.b32 z, %18, %18, %8;\n\tlop3.b32 key, z, 0x80, %7, 0xEA;\n\t"
template <int CLS, int TT> __host__ __device__ constexpr bool ep4u_exists() { return CLS != 0 || TT == 2; }
template <int CLS, int TT>
__device__ __forceinline__ void eval_pair4_u2(const uint4 &a, const uint4 &b, uint32_t rq0, uint32_t rq1, uint32_t rc0,
                                              uint32_t rc1, uint32_t rm0, uint32_t rm1, uint32_t msku, uint32_t mzone,
                                              uint32_t ntol, uint32_t pzb, uint32_t &best0, uint32_t &best1,
                                              uint32_t negone, uint32_t one0, uint32_t one1) {
    const uint32_t o_sku = (CLS & 1) ? a.x : b.x, o_zone = (CLS & 1) ? a.y : b.x, o_taint = TT == 2 ? a.z : b.x;
    const uint32_t o_msku = (CLS & 1) ? msku : rc0, o_mzone = (CLS & 1) ? mzone : rc0, o_ntol = TT == 2 ? ntol : rc0;
    const uint32_t o_pzb = (CLS & 2) ? pzb : rc0;
#define K8S_SG0 "setp.ge.s32 pf0, x1, 0;\n\t"
#define K8S_SG1 "setp.ge.s32 pf1, y1, 0;\n\t"
#define K8S_SG0P "setp.ge.and.s32 pf0, x1, 0, pm;\n\t"
#define K8S_SG1P "setp.ge.and.s32 pf1, y1, 0, pm;\n\t"
#define K8S_EP4U_T(PREF, KEY)                                                                                  \
    if ((CLS & 1) && TT == 2) K8S_EP4U("%7", K8S_EP4U_MASK K8S_EP4U_TOL_MORE K8S_EP4U_PM, K8S_SG0P, K8S_SG1P, PREF, KEY); \
    else if ((CLS & 1) && TT == 1) K8S_EP4U("%8", K8S_EP4U_MASK K8S_EP4U_PM, K8S_SG0P, K8S_SG1P, PREF, KEY);   \
    else if (CLS & 1) K8S_EP4U("%7", K8S_EP4U_MASK K8S_EP4U_PM, K8S_SG0P, K8S_SG1P, PREF, KEY);                \
    else if (TT == 2) K8S_EP4U("%7", K8S_EP4U_TOL_ONLY K8S_EP4U_PM, K8S_SG0P, K8S_SG1P, PREF, KEY);            \
    else if (TT == 1) K8S_EP4U("%8", "", K8S_SG0, K8S_SG1, PREF, KEY);                                         \
    else K8S_EP4U("%7", "", K8S_SG0, K8S_SG1, PREF, KEY)
    if (CLS & 2) {
        K8S_EP4U_T(K8S_EP4U_PREF, "key");
    } else {
        K8S_EP4U_T("", "%7");
    }
#undef K8S_EP4U_T
#undef K8S_SG0
#undef K8S_SG1
#undef K8S_SG0P
#undef K8S_SG1P
}

// R: evals of pods for which, this round, "enough GPUs" implies "enough CPU and memory" at every
// node (resmin_note).  Only the GPU difference is left: against the field of NodeQ::zrot that is
// -1 on an unschedulable (TT = 0: bits 5..12, brought to the top by the multiply) or unschedulable
// or tainted node (TT = 1: bits 13..31).  rq: ep4_request_r.
//   %0 best | %1 khi %2 zrot | %3 rq | %4 runtime -1 | %5 runtime 1 | %6 runtime 2^19
template <int TT>
__device__ __forceinline__ void eval_pair4_r(const uint4 &b, uint32_t rq, uint32_t &best, uint32_t negone,
                                             uint32_t one, uint32_t k19) {
    if (TT == 1) {
        asm("{\n\t"
            ".reg .pred pf;\n\t"
            ".reg .b32 d1;\n\t"
            "mad.lo.s32 d1, %3, %4, %2;\n\t"
            "setp.ge.s32 pf, d1, 0;\n\t"
            "setp.gt.and.u32 pf, %1, %0, pf;\n\t"
            "@pf mad.lo.u32 %0, %1, %5, 0;\n\t"
            "}"
            : "+r"(best)
            : "r"(b.z), "r"(b.w), "r"(rq), "r"(negone), "r"(one), "r"(k19));
    } else {
        asm("{\n\t"
            ".reg .pred pf;\n\t"
            ".reg .b32 d1;\n\t"
            "mad.lo.s32 d1, %2, %6, %3;\n\t"
            "setp.ge.s32 pf, d1, 0;\n\t"
            "setp.gt.and.u32 pf, %1, %0, pf;\n\t"
            "@pf mad.lo.u32 %0, %1, %5, 0;\n\t"
            "}"
            : "+r"(best)
            : "r"(b.z), "r"(b.w), "r"(rq), "r"(negone), "r"(one), "r"(k19));
    }
}
// the GPU request in the form eval_pair4_r takes (free_gpu < kResGpus at every node when it runs)
__device__ __forceinline__ uint32_t ep4_request_r(uint32_t req_gpu, bool tainted_field) {
    return tainted_field ? min(req_gpu, 0x8000u) << 13 : 0u - (min(req_gpu, kResGpus) << 24);
}
// ... and for a thread's two like pods (eval_pair4_u2): masks and preference once per node.
//   %0 %1 best | %2 sku_bit %3 zone_bit | %4 khi %5 zrot | %6 %7 rq | %8 sku mask %9 zone mask
//   %10 pzb | %11 runtime -1 | %12 %13 runtime 1 | %14 runtime 2^19
#define K8S_EP4UR(D0, D1, SHARED, SIGN0, SIGN1, PREF, KEY)                                                     \
    asm("{\n\t"                                                                                                \
        ".reg .pred pf0, pf1, pm;\n\t"                                                                         \
        ".reg .b32 m, z, key, x1, y1;\n\t" D0 D1 SHARED SIGN0 SIGN1 PREF                                       \
        "setp.gt.and.u32 pf0, " KEY ", %0, pf0;\n\t"                                                           \
        "setp.gt.and.u32 pf1, " KEY ", %1, pf1;\n\t"                                                           \
        "@pf0 mad.lo.u32 %0, " KEY ", %12, 0;\n\t"                                                             \
        "@pf1 mad.lo.u32 %1, " KEY ", %13, 0;\n\t"                                                             \
        "}"                                                                                                    \
        : "+r"(best0), "+r"(best1)                                                                             \
        : "r"(o_sku), "r"(o_zone), "r"(b.z), "r"(b.w), "r"(rq0), "r"(rq1), "r"(o_msku), "r"(o_mzone),         \
          "r"(o_pzb), "r"(negone), "r"(one0), "r"(one1), "r"(k19))
template <int CLS, int TT>
__device__ __forceinline__ void eval_pair4_ur2(const uint4 &a, const uint4 &b, uint32_t rq0, uint32_t rq1,
                                               uint32_t msku, uint32_t mzone, uint32_t pzb, uint32_t &best0,
                                               uint32_t &best1, uint32_t negone, uint32_t one0, uint32_t one1,
                                               uint32_t k19) {
    const uint32_t o_sku = (CLS & 1) ? a.x : b.z, o_zone = (CLS & 1) ? a.y : b.z;
    const uint32_t o_msku = (CLS & 1) ? msku : rq0, o_mzone = (CLS & 1) ? mzone : rq0, o_pzb = (CLS & 2) ? pzb : rq0;
#define K8S_R_D0_T1 "mad.lo.s32 x1, %6, %11, %5;\n\t"
#define K8S_R_D1_T1 "mad.lo.s32 y1, %7, %11, %5;\n\t"
#define K8S_R_D0_T0 "mad.lo.s32 x1, %5, %14, %6;\n\t"
#define K8S_R_D1_T0 "mad.lo.s32 y1, %5, %14, %7;\n\t"
#define K8S_R_MASK "and.b32 m, %9, %3;\n\tlop3.b32 m, %8, %2, m, 0xEA;\n\tsetp.eq.u32 pm, m, 0;\n\t"
#define K8S_R_PREF "shf.r.wrap.b32 z, %10, %10, %5;\n\tlop3.b32 key, z, 0x80, %4, 0xEA;\n\t"
#define K8S_R_SG0 "setp.ge.s32 pf0, x1, 0;\n\t"
#define K8S_R_SG1 "setp.ge.s32 pf1, y1, 0;\n\t"
#define K8S_R_SG0P "setp.ge.and.s32 pf0, x1, 0, pm;\n\t"
#define K8S_R_SG1P "setp.ge.and.s32 pf1, y1, 0, pm;\n\t"
#define K8S_EP4UR_T(SHARED, S0, S1, PREF, KEY)                                                                 \
    if (TT == 1) K8S_EP4UR(K8S_R_D0_T1, K8S_R_D1_T1, SHARED, S0, S1, PREF, KEY);                               \
    else K8S_EP4UR(K8S_R_D0_T0, K8S_R_D1_T0, SHARED, S0, S1, PREF, KEY)
    if (CLS == 1) {
        K8S_EP4UR_T(K8S_R_MASK, K8S_R_SG0P, K8S_R_SG1P, "", "%4");
    } else if (CLS == 2) {
        K8S_EP4UR_T("", K8S_R_SG0, K8S_R_SG1, K8S_R_PREF, "key");
    } else {
        K8S_EP4UR_T(K8S_R_MASK, K8S_R_SG0P, K8S_R_SG1P, K8S_R_PREF, "key");
    }
#undef K8S_EP4UR_T
}

// (RES = 4: rg is ep4_request(req_gpu, false), as the kernel keeps it)
template <int RES, int CLS = 3>
__device__ __forceinline__ void eval_pair(const uint4 &na, const uint4 &nb, uint32_t rg, uint32_t rc, uint32_t rm,
                                          uint32_t nsku, uint32_t nzone, uint32_t ntol, uint32_t pzb,
                                          uint32_t &best, uint32_t negone) {
    const uint4 a = make_uint4(na.w, nb.x, nb.y, na.x), b = make_uint4(na.y, na.z, nb.z, nb.w);
    if (RES == 4) {
        eval_pair4<CLS, 2>(na, nb, rg, rc, rm, nsku, nzone, ntol, pzb, best, negone, 0u - negone);
    } else if (RES == 0) {
        asm("{\n\t"
            ".reg .pred pf;\n\t"
            ".reg .b32 m, z, key;\n\t"
            "and.b32 m, %1, %8;\n\t"
            "lop3.b32 m, %5, %9, m, 0xEA;\n\t"
            "lop3.b32 m, %6, %10, m, 0xEA;\n\t"
            "setp.eq.u32 pf, m, 0;\n\t"
            "setp.ge.and.u32 pf, %3, %13, pf;\n\t"
            "setp.ge.and.u32 pf, %4, %14, pf;\n\t"
            "setp.ge.and.s32 pf, %2, %12, pf;\n\t"
            "shf.r.wrap.b32 z, %11, %11, %7;\n\t"
            "lop3.b32 key, z, 0x80, %15, 0xEA;\n\t"
            "setp.gt.and.u32 pf, key, %0, pf;\n\t"
            "selp.u32 %0, key, %0, pf;\n\t"
            "}"
            : "+r"(best)
            : "r"(a.w), "r"(a.x), "r"(a.y), "r"(a.z), "r"(b.x), "r"(b.y), "r"(b.w), "r"(nsku), "r"(nzone),
              "r"(ntol), "r"(pzb), "r"(rg), "r"(rc), "r"(rm), "r"(b.z));
    } else if (RES == 1) {
        asm("{\n\t"
            ".reg .pred pf;\n\t"
            ".reg .b32 m, z, key, d1, d2, d3;\n\t"
            "and.b32 m, %1, %8;\n\t"
            "lop3.b32 m, %5, %9, m, 0xEA;\n\t"
            "lop3.b32 m, %6, %10, m, 0xEA;\n\t"
            "mad.lo.s32 d1, %12, %16, %2;\n\t"
            "mad.lo.s32 d2, %13, %16, %3;\n\t"
            "mad.lo.s32 d3, %14, %16, %4;\n\t"
            "setp.eq.u32 pf, m, 0;\n\t"
            "lop3.b32 d1, d1, d2, d3, 0xFE;\n\t"
            "setp.ge.and.s32 pf, d1, 0, pf;\n\t"
            "shf.r.wrap.b32 z, %11, %11, %7;\n\t"
            "lop3.b32 key, z, 0x80, %15, 0xEA;\n\t"
            "setp.gt.and.u32 pf, key, %0, pf;\n\t"
            "selp.u32 %0, key, %0, pf;\n\t"
            "}"
            : "+r"(best)
            : "r"(a.w), "r"(a.x), "r"(a.y), "r"(a.z), "r"(b.x), "r"(b.y), "r"(b.w), "r"(nsku), "r"(nzone),
              "r"(ntol), "r"(pzb), "r"(rg), "r"(rc), "r"(rm), "r"(b.z), "r"(negone));
    } else if (RES == 2) {
        asm("{\n\t"
            ".reg .pred pf;\n\t"
            ".reg .b32 m, z, key, d1, d2, d3;\n\t"
            "and.b32 m, %1, %8;\n\t"
            "lop3.b32 m, %5, %9, m, 0xEA;\n\t"
            "lop3.b32 m, %6, %10, m, 0xEA;\n\t"
            "mad.lo.s32 d1, %12, %16, %2;\n\t"
            "mad.lo.s32 d2, %13, %16, %3;\n\t"
            "mad.lo.s32 d3, %14, %16, %4;\n\t"
            "setp.eq.u32 pf, m, 0;\n\t"
            "lop3.b32 d1, d1, d2, d3, 0xFE;\n\t"
            "setp.ge.and.s32 pf, d1, 0, pf;\n\t"
            "shf.r.wrap.b32 z, %11, %11, %7;\n\t"
            "lop3.b32 key, z, 0x80, %15, 0xEA;\n\t"
            "setp.gt.and.u32 pf, key, %0, pf;\n\t"
            "@pf mad.lo.u32 %0, key, %17, 0;\n\t"
            "}"
            : "+r"(best)
            : "r"(a.w), "r"(a.x), "r"(a.y), "r"(a.z), "r"(b.x), "r"(b.y), "r"(b.w), "r"(nsku), "r"(nzone),
              "r"(ntol), "r"(pzb), "r"(rg), "r"(rc), "r"(rm), "r"(b.z), "r"(negone), "r"(0u - negone));
    } else {
        asm("{\n\t"
            ".reg .pred pf;\n\t"
            ".reg .b32 m, z, key, d1, d2, d3;\n\t"
            "and.b32 m, %1, %8;\n\t"
            "lop3.b32 m, %5, %9, m, 0xEA;\n\t"
            "lop3.b32 m, %6, %10, m, 0xEA;\n\t"
            "mad.lo.s32 d1, %12, %16, %2;\n\t"
            "mad.lo.s32 d2, %13, %16, %3;\n\t"
            "mad.lo.s32 d3, %14, %16, %4;\n\t"
            "setp.eq.u32 pf, m, 0;\n\t"
            "lop3.b32 d1, d1, d2, d3, 0xFE;\n\t"
            "setp.ge.and.s32 pf, d1, 0, pf;\n\t"
            "mul.lo.u32 z, %11, %5;\n\t"
            "mul.hi.u32 z, z, %18;\n\t"
            "lop3.b32 key, z, 0x80, %15, 0xEA;\n\t"
            "setp.gt.and.u32 pf, key, %0, pf;\n\t"
            "@pf mad.lo.u32 %0, key, %17, 0;\n\t"
            "}"
            : "+r"(best)
            : "r"(a.w), "r"(a.x), "r"(a.y), "r"(a.z), "r"(b.x), "r"(b.y), "r"(b.w), "r"(nsku), "r"(nzone),
              "r"(ntol), "r"(pzb), "r"(rg), "r"(rc), "r"(rm), "r"(b.z), "r"(negone), "r"(0u - negone), "r"((0u - negone) << 8));
    }
}

// Which scoring slots a launch covers.  One GPU, or the round-at-a-time API: the contiguous range
// [lo, hi).  A placement group of xw ranks: blocks of kXBlock slots dealt round-robin, rank r
// taking blocks r, r + xw, ... -- the slots are sorted by pod class (dearest classes first), so
// contiguous ranges would give the ranks unequal work.  A rank's slots are numbered 0 .. xcnt by
// their local index jj; in the exchanged proposal array rank r's slot jj sits at r * xcnt + jj.
constexpr uint32_t kXBlock = 2048;
struct SlotMap {
    uint32_t lo, hi, xw, xrank, xcnt;
};
__device__ __forceinline__ bool slot_of(const SlotMap &m, uint32_t jj, uint32_t *slot) {
    if (m.xw <= 1u) {
        *slot = m.lo + jj;
        return *slot < m.hi;
    }
    *slot = ((jj / kXBlock) * m.xw + m.xrank) * kXBlock + (jj % kXBlock);
    return jj < m.xcnt && *slot < m.hi;
}
// where slot s of the round sits in the exchanged proposal array
__device__ __forceinline__ uint32_t exchange_index(uint32_t s, uint32_t xw, uint32_t xcnt) {
    if (xw <= 1u) return s;
    const uint32_t b = s / kXBlock;
    return (b % xw) * xcnt + (b / xw) * kXBlock + (s % kXBlock);
}

constexpr int kDefaultRes = 4;
constexpr uint32_t kStages = K8S_STAGES;  // node tiles in flight per CTA

constexpr uint32_t kGroup = 64;        // nodes per index group (index bits 1..6 of the key)
constexpr uint32_t kIdxMask = 0x7Eu;   // those bits

// grid.x: pod tiles of 256*PPT active positions; grid.y: node splits (runs of whole tiles).
// best_key[pos] = (key32 << 32) | (0xFFFFFFFF - ((n - rot) mod N)), 0 when no node is feasible.
// key32 orders nodes exactly like the spec's upper key bits for a fixed pod (the pod's req_gpu
// only shifts the 15-bit field by a constant).  The rotated-index tie-break costs nothing per
// eval: nodes are scanned in ascending n with a strict '>', so among equal keys the lowest n
// wins; after each 64-node group the index bits of the running best are saturated, so a later
// group needs a strictly larger key; and at the one point where the scan crosses n == rot(p)
// the running best is lowered to just below its own key, which lets a later node with an EQUAL
// key take over -- exactly "nodes at or after rot come first".  Only a 16-node block that
// contains rot takes the slow path.
// UNR: nodes per trip of the class loops (divides 16).  16 amortises the loop bookkeeping best; 8
// halves the 22 loop bodies, which matters when CTAs are short-lived and the SM's instruction
// cache holds many classes at once (measured: config 3, one tile per CTA, 4.49 vs 4.61 ms per
// placement; config 4, 79 tiles per CTA, 273.9 vs 270.7).
template <int PPT, int RES, int MB, int UNR = 16>
__global__ void __launch_bounds__(256, MB) score_kernel(const NodeQ *__restrict__ nodeq, uint32_t n_tiles,
                                                    uint32_t tiles_per_split, const PodQ *__restrict__ podq,
                                                    const uint32_t *__restrict__ active, SlotMap sm,
                                                    unsigned long long *__restrict__ best_key, uint32_t n_nodes,
                                                    int use_atomic, uint32_t negone, int cls_force,
                                                    const uint32_t *__restrict__ spod,
                                                    const uint32_t *__restrict__ posof,
                                                    const uint32_t *__restrict__ ctr,
                                                    const uint32_t *__restrict__ resmin, uint32_t big_rows) {
    // kStages-deep ring of node tiles.  full[b]: the TMA copy into buffer b has landed (one
    // arrival + its byte count); empty[b]: all 8 warps are done reading buffer b (one arrival per
    // warp).  No __syncthreads in the tile loop: a warp may run a tile ahead of its slowest sibling.
    __shared__ __align__(128) uint4 tile[kStages][kNodeTile * 2];
    __shared__ __align__(8) uint64_t full[kStages], empty[kStages];
    // ressuf[0][g]: the least free_mcpu among the schedulable nodes with >= g free GPUs (resmin_note),
    // ressuf[1][g]: the least free_mib; [kResGpus]: no such node
    __shared__ uint32_t ressuf[2][kResGpus + 1];
    const int tid = threadIdx.x;
    // grid.y: big_rows splits of tiles_per_split node tiles each, then one split per remaining tile
    // (short CTAs that the hardware dispatches last: they fill the launch's final wave finely)
    const uint32_t t0 = blockIdx.y < big_rows ? blockIdx.y * tiles_per_split
                                              : big_rows * tiles_per_split + (blockIdx.y - big_rows);
    const uint32_t t1 = min(t0 + (blockIdx.y < big_rows ? tiles_per_split : 1u), n_tiles);
    constexpr uint32_t kTileBytes = kNodeTile * sizeof(NodeQ);

    if (tid == 0) {
#pragma unroll
        for (uint32_t b = 0; b < kStages; ++b) {
            mbar_init(&full[b], 1);
            mbar_init(&empty[b], 256 / 32);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (RES == 4 && PPT == 2 && tid >= 32 && tid < 64) {  // warp 1: suffix minima of the round's resource summary
        const uint32_t lane = tid - 32;
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            uint32_t lo = __ldg(resmin + k * kResGpus + lane), hi = __ldg(resmin + k * kResGpus + 32 + lane);
#pragma unroll
            for (uint32_t off = 1; off < 32; off <<= 1) {
                const uint32_t l2 = __shfl_down_sync(0xFFFFFFFFu, lo, off), h2 = __shfl_down_sync(0xFFFFFFFFu, hi, off);
                if (lane + off < 32) lo = min(lo, l2), hi = min(hi, h2);
            }
            lo = min(lo, __shfl_sync(0xFFFFFFFFu, hi, 0));
            ressuf[k][lane] = lo;
            ressuf[k][32 + lane] = hi;
            if (lane == 0) ressuf[k][kResGpus] = 0xFFFFFFFFu;
        }
    }
    __syncthreads();
    if (tid == 0) {
        for (uint32_t b = 0; b + 1 < kStages && t0 + b < t1; ++b) {
            mbar_expect_tx(&full[b], kTileBytes);
            bulk_g2s(&tile[b][0], nodeq + (size_t)(t0 + b) * kNodeTile, kTileBytes, &full[b]);
        }
    }

    uint32_t pos[PPT], rg[PPT], rc[PPT], rm[PPT], nsku[PPT], nzone[PPT], ntol[PPT], pzb[PPT], rot[PPT];
    uint32_t best[PPT], bestn[PPT];
#pragma unroll
    for (int j = 0; j < PPT; ++j) {
        // thread -> pod: the active position itself, or the slot of the round's scoring order
        uint32_t slot;
        const bool live = slot_of(sm, (blockIdx.x * 256u + tid) * PPT + j, &slot);  // a thread takes neighbours: like pods, close rot(p)
        uint32_t pod = 0;
        pos[j] = 0xFFFFFFFFu;
        if (live) {
            if (spod) {
                pod = __ldg(&spod[slot]);
                pos[j] = __ldg(&posof[pod]);
            } else {
                pos[j] = slot;
                pod = active[slot];
            }
        }
        best[j] = 0;
        bestn[j] = 0;
        if (live) {
            const uint4 *q = reinterpret_cast<const uint4 *>(podq + pod);
            uint4 a = __ldg(q), b = __ldg(q + 1);
            rg[j] = a.x; rc[j] = a.y; rm[j] = a.z; nsku[j] = a.w;
            nzone[j] = b.x; ntol[j] = b.y; pzb[j] = b.z; rot[j] = b.w;
        } else {  // idle lane: no node has 2^31 - 1 free GPUs, so nothing is ever feasible (and its class is 0)
            rg[j] = 0x7FFFFFFFu; rc[j] = rm[j] = 0; nsku[j] = nzone[j] = 0; ntol[j] = 0xFFFFFFFFu; pzb[j] = 0; rot[j] = 0;
        }
    }
    // Class of this warp's pods (see eval_pair4): which clauses can be left out of the inner loop.
    // wv = CLS | TT << 2.  Idle lanes fit every variant (their request exceeds any free count).
    uint32_t wv = 3u | (2u << 2);
    // R (eval_pair4_r): does "enough GPUs" imply "enough CPU and memory" at every node, for each of my pods?
    bool implied = RES == 4 && PPT == 2 && __ldg(ctr + kResWide) == 0u;
    uint32_t rgr0[PPT], rgr1[PPT];  // ... and the request in the forms those evals take
#pragma unroll
    for (int j = 0; j < PPT; ++j) {
        const uint32_t g = min(rg[j], kResGpus);
        if (RES == 4 && PPT == 2) implied = implied && rc[j] <= ressuf[0][g] && rm[j] <= ressuf[1][g];
        rgr0[j] = ep4_request_r(rg[j], false);
        rgr1[j] = ep4_request_r(rg[j], true);
    }
    uint32_t rg5[PPT];  // RES = 4: the GPU request in the two forms eval_pair4 takes (rg[] itself holds the other)
    // The multiplier of the predicated "best = key" mad, one copy per pod the compiler cannot tell
    // apart: with a shared one, two pods whose key is the node's own word (no preference clause)
    // share ONE multiply and each gets an ALU select instead -- half an instruction more per eval.
    uint32_t one[PPT];
#pragma unroll
    for (int j = 0; j < PPT; ++j) {
        rg5[j] = ep4_request(rg[j], true);
        if (RES == 4) rg[j] = ep4_request(rg[j], false);
        // (all of these are 1 for negone = 0xFFFFFFFF, which is what the host passes)
        one[j] = j == 0 ? 0u - negone : j == 1 ? negone * negone : j == 2 ? negone >> 31 : negone & 1u;
    }
    if (RES == 4) {
        const uint32_t tunion = __ldg(ctr + kTaintUnion);
        uint32_t mine = 0;
#pragma unroll
        for (int j = 0; j < PPT; ++j) {
            mine |= ((nsku[j] | nzone[j]) ? 1u : 0u) | (pzb[j] ? 2u : 0u);
            if (pos[j] != 0xFFFFFFFFu) mine |= 4u << taint_class(ntol[j], tunion);
        }
        const uint32_t w = __reduce_or_sync(0xFFFFFFFFu, mine);
        const uint32_t tset = w >> 2;  // bit t: some pod of the warp has taint class t
        const uint32_t wtt = ((tset & 4u) || (tset & 3u) == 3u) ? 2u : ((tset & 2u) ? 1u : 0u);
        wv = (w & 3u) | (wtt << 2);
        // bit 4: in every thread of the warp the two pods share masks, preferred zone and (if the
        // clause is evaluated) tolerations -- eval_pair4_u2.  An idle second pod takes the first's.
        if (PPT == 2) {
            if (pos[1] == 0xFFFFFFFFu) nsku[1] = nsku[0], nzone[1] = nzone[0], pzb[1] = pzb[0], ntol[1] = ntol[0];
            const bool alike = nsku[0] == nsku[1] && nzone[0] == nzone[1] && pzb[0] == pzb[1] &&
                               (wtt != 2u || ntol[0] == ntol[1]);
            if (__all_sync(0xFFFFFFFFu, alike) && ((wv & 3u) != 0u || wtt == 2u) && cls_force != 17) wv |= 16u;
        }
        // bit 5: and for every pod of the warp the CPU and memory clauses are implied (R)
        if (PPT == 2 && __all_sync(0xFFFFFFFFu, implied) && wtt < 2u && ((wv & 3u) == 0u || (wv & 16u)) &&
            cls_force != 18 && cls_force != 16)
            wv |= 32u;
        // timing experiments only (results are wrong): 0..11 forces CLS | TT << 2.  Results stay
        // right with 16 (no taint specialisation, nothing shared, no R), 17 (nothing shared), 18 (no R)
        if (cls_force >= 0 && cls_force < 12) wv = (uint32_t)cls_force;
        if (cls_force == 16) wv = (wv & 3u) | (2u << 2);
    }

    const uint32_t msku = __brev(nsku[0]), mzone = __brev(nzone[0]);  // eval_pair4_u2: un-mirrored, for AND tests
    const uint32_t k19 = one[0] << 19;                                   // eval_pair4_r: a runtime 2^19

    for (uint32_t t = t0; t < t1; ++t) {
        const uint32_t it = t - t0, buf = it % kStages;
        for (uint32_t spins = 0; !mbar_try_wait(&full[buf], (it / kStages) & 1u);)
            if (++spins > (1u << 26)) __trap();  // a lost TMA completion must fail loudly, not hang the GPU
        const uint32_t nbase = t * kNodeTile;
        uint32_t rrel[PPT];  // rot - nbase: the scan crosses rot inside this tile iff rrel < 256
#pragma unroll
        for (int j = 0; j < PPT; ++j) rrel[j] = rot[j] - nbase;
        const uint4 *tl = tile[buf];
#pragma unroll 1
        for (uint32_t g = 0; g < kNodeTile; g += kGroup) {
            uint32_t prev[PPT];
#pragma unroll
            for (int j = 0; j < PPT; ++j) prev[j] = best[j];
            // Which of the group's four 16-node blocks hold rot(p) of SOME pod of this warp?  (Warp-
            // uniform, so the two code paths never both run for one warp.)  Almost always none: the
            // whole group goes through the class's loop, 64 evals per pod with no checks at all.
            // Otherwise the runs of blocks without a rot(p) still do -- the same loop, so no second
            // copy of it competes for the instruction cache -- and each block with one is walked node
            // by node.  The pods of a warp are neighbours in rot(p) (the scoring order sorts them so),
            // so a warp meets one or two such groups per pass over the node table.
            uint32_t xb = 0;
            bool gx = false;
#pragma unroll
            for (int j = 0; j < PPT; ++j) gx |= (rrel[j] - g) < kGroup;
            if (__any_sync(0xFFFFFFFFu, gx)) {
#pragma unroll
                for (int j = 0; j < PPT; ++j) xb |= (rrel[j] - g) < kGroup ? 1u << ((rrel[j] - g) >> 4) : 0u;
                xb = __reduce_or_sync(0xFFFFFFFFu, xb);
            }
            for (uint32_t k0 = g; k0 < g + kGroup;) {
                const uint32_t bi = (k0 - g) >> 4;
                if (xb & (1u << bi)) {  // rare: rot(p) of some pod of this warp falls inside this block
                    // (a compact loop with branches: written with selects and unrolled it measured
                    // 5 % slower on config 3)
#pragma unroll 1
                    for (uint32_t i = k0; i < k0 + 16u; ++i) {
                        const uint4 a = tl[2 * i], b = tl[2 * i + 1];
#pragma unroll
                        for (int j = 0; j < PPT; ++j) {
                            if (rrel[j] == i) {  // entering n >= rot: bank the winner so far, lower it
                                if (best[j] != prev[j])
                                    bestn[j] = nbase + g + (kGroup - 1u) - ((best[j] & kIdxMask) >> 1);
                                if (best[j] != 0u) best[j] = (best[j] & ~(kIdxMask | 1u)) - 1u;
                                prev[j] = best[j];
                            }
                            eval_pair<RES>(a, b, rg[j], rc[j], rm[j], nsku[j], nzone[j], ntol[j], pzb[j], best[j],
                                           negone);
                        }
                    }
                    k0 += 16u;
                    continue;
                }
                // the run of blocks up to the next one that holds a rot(p)
                const uint32_t later = xb >> bi;  // bit 0 is clear
                const uint32_t k1 = later ? k0 + ((uint32_t)__ffs((int)later) - 1u) * 16u : g + kGroup;
#define K8S_GROUP(CLS, TT)                                                                                       \
    _Pragma("unroll 1") for (uint32_t kb = k0; kb < k1; kb += UNR) {                                              \
        _Pragma("unroll") for (uint32_t i = 0; i < UNR; ++i) {                                                   \
            const uint4 b = ldrec<PPT>(&tl[2 * (kb + i) + 1]);                                                   \
            const uint4 a = ep4_needs_first_half<CLS, TT>() ? ldrec<PPT>(&tl[2 * (kb + i)]) : b;                 \
            _Pragma("unroll") for (int j = 0; j < PPT; ++j)                                                      \
                eval_pair4<CLS, TT>(a, b, TT == 1 ? rg5[j] : rg[j], rc[j], rm[j], nsku[j], nzone[j], ntol[j],    \
                                    pzb[j], best[j], negone, one[j]);                                            \
        }                                                                                                        \
    }
#define K8S_GROUP_RES(CLS)                                                                                       \
    _Pragma("unroll 1") for (uint32_t kb = k0; kb < k1; kb += UNR) {                                              \
        _Pragma("unroll") for (uint32_t i = 0; i < UNR; ++i) {                                                   \
            const uint4 a = ldrec<PPT>(&tl[2 * (kb + i)]), b = ldrec<PPT>(&tl[2 * (kb + i) + 1]);                \
            _Pragma("unroll") for (int j = 0; j < PPT; ++j)                                                      \
                eval_pair<RES, CLS>(a, b, rg[j], rc[j], rm[j], nsku[j], nzone[j], ntol[j], pzb[j], best[j], negone); \
        }                                                                                                        \
    }
#define K8S_GROUP_U(CLS, TT)                                                                                     \
    _Pragma("unroll 1") for (uint32_t kb = k0; kb < k1; kb += UNR) {                                              \
        _Pragma("unroll") for (uint32_t i = 0; i < UNR; ++i) {                                                   \
            const uint4 b = ldrec<PPT>(&tl[2 * (kb + i) + 1]);                                                   \
            const uint4 a = ep4_needs_first_half<CLS, TT>() ? ldrec<PPT>(&tl[2 * (kb + i)]) : b;                 \
            eval_pair4_u2<CLS, TT>(a, b, TT == 1 ? rg5[0] : rg[0], TT == 1 ? rg5[PPT - 1] : rg[PPT - 1], rc[0],  \
                                   rc[PPT - 1], rm[0], rm[PPT - 1], msku, mzone, ntol[0], pzb[0], best[0],        \
                                   best[PPT - 1], negone, one[0], one[PPT - 1]);                                 \
        }                                                                                                        \
    }
#define K8S_GROUP_R(TT) /* (short bodies: always 16 nodes per trip) */                                           \
    _Pragma("unroll 1") for (uint32_t kb = k0; kb < k1; kb += 16) {                                              \
        _Pragma("unroll") for (uint32_t i = 0; i < 16; ++i) {                                                    \
            const uint4 b = ldrec<PPT>(&tl[2 * (kb + i) + 1]);                                                   \
            _Pragma("unroll") for (int j = 0; j < PPT; ++j)                                                      \
                eval_pair4_r<TT>(b, TT == 1 ? rgr1[j] : rgr0[j], best[j], negone, one[j], k19);                  \
        }                                                                                                        \
    }
#define K8S_GROUP_UR(CLS, TT)                                                                                    \
    _Pragma("unroll 1") for (uint32_t kb = k0; kb < k1; kb += 16) {                                              \
        _Pragma("unroll") for (uint32_t i = 0; i < 16; ++i) {                                                    \
            const uint4 b = ldrec<PPT>(&tl[2 * (kb + i) + 1]);                                                   \
            const uint4 a = (CLS & 1) ? ldrec<PPT>(&tl[2 * (kb + i)]) : b;                                       \
            eval_pair4_ur2<CLS, TT>(a, b, TT == 1 ? rgr1[0] : rgr0[0], TT == 1 ? rgr1[PPT - 1] : rgr0[PPT - 1],  \
                                    msku, mzone, pzb[0], best[0], best[PPT - 1], negone, one[0], one[PPT - 1],   \
                                    k19);                                                                        \
        }                                                                                                        \
    }
                if (RES != 4) {
                    K8S_GROUP_RES(3)
                } else if (PPT == 2 && (wv & 32u)) {
                    switch (wv & 7u) {
                    case 0: K8S_GROUP_R(0) break;
                    case 1: K8S_GROUP_UR(1, 0) break;
                    case 2: K8S_GROUP_UR(2, 0) break;
                    case 3: K8S_GROUP_UR(3, 0) break;
                    case 4: K8S_GROUP_R(1) break;
                    case 5: K8S_GROUP_UR(1, 1) break;
                    case 6: K8S_GROUP_UR(2, 1) break;
                    default: K8S_GROUP_UR(3, 1) break;
                    }
                } else if (PPT == 2 && (wv & 16u)) {
                    switch (wv & 15u) {
                    case 1: K8S_GROUP_U(1, 0) break;
                    case 2: K8S_GROUP_U(2, 0) break;
                    case 3: K8S_GROUP_U(3, 0) break;
                    case 5: K8S_GROUP_U(1, 1) break;
                    case 6: K8S_GROUP_U(2, 1) break;
                    case 7: K8S_GROUP_U(3, 1) break;
                    case 8: K8S_GROUP_U(0, 2) break;
                    case 9: K8S_GROUP_U(1, 2) break;
                    case 10: K8S_GROUP_U(2, 2) break;
                    default: K8S_GROUP_U(3, 2) break;
                    }
                } else {
                    switch (wv & 15u) {
                    case 0: K8S_GROUP(0, 0) break;
                    case 1: K8S_GROUP(1, 0) break;
                    case 2: K8S_GROUP(2, 0) break;
                    case 3: K8S_GROUP(3, 0) break;
                    case 4: K8S_GROUP(0, 1) break;
                    case 5: K8S_GROUP(1, 1) break;
                    case 6: K8S_GROUP(2, 1) break;
                    case 7: K8S_GROUP(3, 1) break;
                    case 8: K8S_GROUP(0, 2) break;
                    case 9: K8S_GROUP(1, 2) break;
                    case 10: K8S_GROUP(2, 2) break;
                    default: K8S_GROUP(3, 2) break;
                    }
                }
#undef K8S_GROUP_RES
#undef K8S_GROUP_U
#undef K8S_GROUP_R
#undef K8S_GROUP_UR
#undef K8S_GROUP
                k0 = k1;
            }
#pragma unroll
            for (int j = 0; j < PPT; ++j)
                if (best[j] != prev[j]) {  // a node of this group took the lead: decode it, saturate the index bits
                    bestn[j] = nbase + g + (kGroup - 1u) - ((best[j] & kIdxMask) >> 1);
                    best[j] |= kIdxMask;
                }
        }
        __syncwarp();
        if ((tid & 31) == 0) mbar_arrive(&empty[buf]);  // this warp is done with tile[buf]
        const uint32_t nxt = it + kStages - 1u;          // refill the buffer that was read one tile ago
        if (t0 + nxt < t1) {
            const uint32_t b2 = nxt % kStages;
            // Every thread waits (block-uniform control flow keeps the loop counters and tile
            // addresses in uniform registers); the buffer was read a whole tile ago, so this
            // almost never spins -- the warps may drift apart by one tile.
            if (nxt >= kStages)
                for (uint32_t spins = 0; !mbar_try_wait(&empty[b2], (nxt / kStages - 1u) & 1u);)
                    if (++spins > (1u << 26)) __trap();
            if (tid == 0) {
                mbar_expect_tx(&full[b2], kTileBytes);
                bulk_g2s(&tile[b2][0], nodeq + (size_t)(t0 + nxt) * kNodeTile, kTileBytes, &full[b2]);
            }
        }
    }

#pragma unroll
    for (int j = 0; j < PPT; ++j) {
        if (pos[j] == 0xFFFFFFFFu) continue;
        unsigned long long key64 = 0ull;
        if (best[j] != 0u) {
            // strip the index bits and undo a crossing decrement: the spec's upper key bits
            uint32_t key32 = (best[j] + 1u) & ~(kIdxMask | 1u);
            uint32_t ridx = bestn[j] >= rot[j] ? bestn[j] - rot[j] : bestn[j] + (n_nodes - rot[j]);
            key64 = ((unsigned long long)key32 << 32) | (unsigned long long)(0xFFFFFFFFu - ridx);
        }
        if (use_atomic) {
            if (key64) atomicMax(&best_key[pos[j]], key64);
        } else {
            best_key[pos[j]] = key64;
        }
    }
}

// Proposals travel by SLOT (the index the scoring kernel's threads are numbered by): slot s is the
// pod spod[s] of the round's scoring order, at active position posof[pod]; without an order, slot
// and active position are the same thing.
struct SlotView {
    const uint32_t *active, *spod, *posof;
};
__device__ __forceinline__ int32_t decode_proposal(const unsigned long long *__restrict__ best_key,
                                                   const PodQ *__restrict__ podq, const SlotView &v, uint32_t slot,
                                                   uint32_t n_nodes) {
    uint32_t pod, pos;
    if (v.spod) {
        pod = __ldg(&v.spod[slot]);
        pos = __ldg(&v.posof[pod]);
    } else {
        pos = slot;
        pod = v.active[slot];
    }
    const unsigned long long key = best_key[pos];
    if (key == 0ull) return -1;
    const uint32_t n = (0xFFFFFFFFu - (uint32_t)key) + podq[pod].rot;  // < 2N < 2^32
    return (int32_t)(n >= n_nodes ? n - n_nodes : n);
}

__global__ void score_finalize_kernel(const unsigned long long *__restrict__ best_key, const PodQ *__restrict__ podq,
                                      SlotView view, SlotMap sm, uint32_t cnt, uint32_t n_nodes,
                                      int32_t *__restrict__ out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x, slot;
    if (i >= cnt) return;
    out[i] = slot_of(sm, i, &slot) ? decode_proposal(best_key, podq, view, slot, n_nodes) : -1;
}

// Decode the packed best key of this launch's slots and store each proposal straight into EVERY
// rank's exchange buffer (peer stores over NVLink; the own buffer is just another pointer): the
// all-gather of the round is done by the kernel that produces the values.
struct PeerPtrs {
    int32_t *p[16];
    int world;
};
__global__ void score_finalize_p2p_kernel(const unsigned long long *__restrict__ best_key,
                                          const PodQ *__restrict__ podq, SlotView view, SlotMap sm, uint32_t cnt,
                                          uint32_t n_nodes, uint32_t base, PeerPtrs peers) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x, slot;
    if (i >= cnt || !slot_of(sm, i, &slot)) return;
    const int32_t node = decode_proposal(best_key, podq, view, slot, n_nodes);
#pragma unroll 1
    for (int r = 0; r < peers.world; ++r) peers.p[r][base + i] = node;
}

// ------------------------------------------------------------------ admission (a13)

__global__ void admit_keys_kernel(const int32_t *__restrict__ best, const uint32_t *__restrict__ active,
                                  uint32_t na, uint32_t n_nodes, int32_t *__restrict__ assign,
                                  uint32_t *__restrict__ keys, uint32_t *__restrict__ vals,
                                  uint32_t *__restrict__ err) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= na) return;
    int32_t b = best[i];
    if (b < -1 || b >= (int32_t)n_nodes) {  // not something score_kernel can produce: a stale or
        atomicOr(err, kErrProposal);        // unsynchronised proposal buffer was handed in
        b = -1;
    }
    if (b < 0) assign[active[i]] = K8SGPU_ASSIGN_NONE;  // no feasible node: final
    keys[i] = b < 0 ? n_nodes : (uint32_t)b;           // non-proposers sort behind every node
    vals[i] = active[i];  // payload = pod id: the stable sort keeps pod order inside each node
}

struct Seg {  // running (saturating) demand of one node's proposers, in pod order
    uint32_t g, c, m, head;
};
__device__ __forceinline__ uint32_t sat_add(uint32_t a, uint32_t b) {
    uint32_t s = a + b;
    return s < a ? 0xFFFFFFFFu : s;
}
// segmented-scan operator: right operand wins outright if it starts a segment
__device__ __forceinline__ Seg seg_combine(const Seg &l, const Seg &r) {
    Seg o;
    o.head = l.head | r.head;
    o.g = r.head ? r.g : sat_add(l.g, r.g);
    o.c = r.head ? r.c : sat_add(l.c, r.c);
    o.m = r.head ? r.m : sat_add(l.m, r.m);
    return o;
}
__device__ __forceinline__ Seg seg_shfl_up(const Seg &x, int d) {
    Seg y;
    y.g = __shfl_up_sync(0xFFFFFFFFu, x.g, d);
    y.c = __shfl_up_sync(0xFFFFFFFFu, x.c, d);
    y.m = __shfl_up_sync(0xFFFFFFFFu, x.m, d);
    y.head = __shfl_up_sync(0xFFFFFFFFu, x.head, d);
    return y;
}
__device__ __forceinline__ Seg seg_warp_scan(Seg x, int lane) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        Seg y = seg_shfl_up(x, d);
        if (lane >= d) x = seg_combine(y, x);
    }
    return x;
}

constexpr int kSegThreads = 256;

// Inclusive segmented scan of one 256-element block of the sorted proposal list.
// Returns this thread's inclusive value (without any carry from earlier blocks).
__device__ __forceinline__ Seg seg_block_scan(Seg x, Seg *block_total) {
    __shared__ Seg wtot[8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    Seg inc = seg_warp_scan(x, lane);
    if (lane == 31) wtot[warp] = inc;
    __syncthreads();
    Seg pre;
    pre.g = pre.c = pre.m = pre.head = 0;
    for (int w = 0; w < warp; ++w) pre = seg_combine(pre, wtot[w]);
    Seg out = seg_combine(pre, inc);
    if (block_total && threadIdx.x == kSegThreads - 1) *block_total = out;
    return out;
}

// Exclusive carry of block vb (everything in blocks [0, vb)), computed by the calling CTA alone
// from the per-block totals: thread t folds its contiguous share, one block scan orders the
// shares.  All threads return the same value.  Replaces the one-CTA carry phase and its grid
// barrier while the list is short (kLocalCarryMaxBlocks).
constexpr uint32_t kLocalCarryMaxBlocks = 1024;
__device__ __forceinline__ Seg seg_carry_local(const Seg *__restrict__ block_tot, uint32_t vb) {
    __shared__ Seg total;
    const uint32_t per = (vb + kSegThreads - 1) / kSegThreads;
    Seg run;
    run.g = run.c = run.m = run.head = 0;
    for (uint32_t j = 0; j < per; ++j) {
        const uint32_t b = threadIdx.x * per + j;
        if (b < vb) run = seg_combine(run, block_tot[b]);
    }
    Seg inc = seg_block_scan(run, nullptr);
    if (threadIdx.x == kSegThreads - 1) total = inc;
    __syncthreads();  // also: seg_block_scan's shared scratch is free again
    return total;
}

__device__ __forceinline__ Seg load_seg(const uint32_t *__restrict__ keys, const uint32_t *__restrict__ vals,
                                        const uint32_t *__restrict__ active, const PodQ *__restrict__ podq,
                                        uint32_t i, uint32_t na, uint32_t *key_out) {
    Seg x;
    x.g = x.c = x.m = 0;
    x.head = 1;  // out-of-range slots start their own (ignored) segment
    *key_out = 0xFFFFFFFFu;
    if (i < na) {
        uint32_t k = keys[i];
        const PodQ *q = podq + vals[i];
        x.g = q->req_gpu;
        x.c = q->req_mcpu;
        x.m = q->req_mib;
        x.head = (i == 0 || keys[i - 1] != k) ? 1u : 0u;
        *key_out = k;
    }
    return x;
}

__global__ void __launch_bounds__(kSegThreads) admit_seg_reduce_kernel(
    const uint32_t *__restrict__ keys, const uint32_t *__restrict__ vals, const uint32_t *__restrict__ active,
    const PodQ *__restrict__ podq, uint32_t na, Seg *__restrict__ block_tot) {
    uint32_t i = blockIdx.x * kSegThreads + threadIdx.x, key;
    Seg x = load_seg(keys, vals, active, podq, i, na, &key);
    seg_block_scan(x, &block_tot[blockIdx.x]);
}

// One CTA turns block totals into exclusive carries (carry[b] = everything before block b),
// 256 blocks per step.
__device__ __forceinline__ void seg_carry_block(Seg *__restrict__ block_tot, uint32_t nb) {
    __shared__ Seg buf[kSegThreads];
    Seg run;
    run.g = run.c = run.m = run.head = 0;
    for (uint32_t base = 0; base < nb; base += kSegThreads) {
        const uint32_t b = base + threadIdx.x;
        Seg x;
        x.g = x.c = x.m = x.head = 0;
        if (b < nb) x = block_tot[b];
        Seg with = seg_combine(run, seg_block_scan(x, nullptr));  // inclusive through block b
        buf[threadIdx.x] = with;
        __syncthreads();
        if (b < nb) block_tot[b] = threadIdx.x ? buf[threadIdx.x - 1] : run;  // exclusive
        run = buf[kSegThreads - 1];
        __syncthreads();
    }
}
__global__ void __launch_bounds__(kSegThreads) admit_seg_carry_kernel(Seg *__restrict__ block_tot, uint32_t nb) {
    seg_carry_block(block_tot, nb);
}

// S <= free test on the inclusive prefix sums; admitted pods get their node.
__global__ void __launch_bounds__(kSegThreads) admit_apply_kernel(
    const uint32_t *__restrict__ keys, const uint32_t *__restrict__ vals, const uint32_t *__restrict__ active,
    const PodQ *__restrict__ podq, uint32_t na, uint32_t n_nodes, const Seg *__restrict__ carry,
    const uint32_t *__restrict__ freec, int32_t *__restrict__ assign, uint32_t *__restrict__ admitted) {
    uint32_t i = blockIdx.x * kSegThreads + threadIdx.x, key;
    Seg x = load_seg(keys, vals, active, podq, i, na, &key);
    Seg inc = seg_block_scan(x, nullptr);
    Seg s = seg_combine(carry[blockIdx.x], inc);
    if (i >= na) return;
    uint32_t ok = 0;
    if (key < n_nodes) {
        ok = (s.g <= freec[key]) & (s.c <= freec[(size_t)n_nodes + key]) & (s.m <= freec[2 * (size_t)n_nodes + key]);
        if (ok) assign[vals[i]] = (int32_t)key;
    }
    admitted[i] = ok;
}

// free -= admitted requests.  Integer atomics commute, so the result is order independent;
// it runs after admit_apply so no admission test can see a half-updated column.
__global__ void admit_free_update_kernel(const uint32_t *__restrict__ keys, const uint32_t *__restrict__ vals,
                                         const uint32_t *__restrict__ active, const PodQ *__restrict__ podq,
                                         const uint32_t *__restrict__ admitted, uint32_t na, uint32_t n_nodes,
                                         uint32_t *__restrict__ freec) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= na || !admitted[i]) return;
    uint32_t key = keys[i];
    const PodQ *q = podq + vals[i];
    if (q->req_gpu) atomicSub(&freec[key], q->req_gpu);
    if (q->req_mcpu) atomicSub(&freec[(size_t)n_nodes + key], q->req_mcpu);
    if (q->req_mib) atomicSub(&freec[2 * (size_t)n_nodes + key], q->req_mib);
}

__global__ void pending_flags_kernel(const uint32_t *__restrict__ active, const int32_t *__restrict__ assign,
                                     uint32_t na, uint32_t *__restrict__ flags) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < na) flags[i] = assign[active[i]] == K8SGPU_ASSIGN_PENDING ? 1u : 0u;
}

__global__ void compact_kernel(const uint32_t *__restrict__ active, const uint32_t *__restrict__ flags,
                               const uint32_t *__restrict__ offs, uint32_t na, uint32_t *__restrict__ next) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < na && flags[i]) next[offs[i]] = active[i];
}

__global__ void count_placed_kernel(const int32_t *__restrict__ assign, uint32_t m, uint32_t *__restrict__ out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t ok = (i < m && assign[i] >= 0) ? 1u : 0u;
    uint32_t cnt = __popc(__ballot_sync(0xFFFFFFFFu, ok));
    if ((threadIdx.x & 31) == 0 && cnt) atomicAdd(out, cnt);
}

// ------------------------------------------------------------------ scoring order
//
// eval_pair leaves out the clauses no pod of a warp needs (its CLS parameter), and the scan takes
// its slow path only in the 64-node groups that hold some pod's rot(p).  Both pay only if the pods
// of a warp are alike, so the scoring kernel does not walk the active list in pod order: it walks
// `spod`, the active pods sorted ONCE per run by (class, which 256th of the node table rot(p)
// falls into) -- one 10-bit stable sort -- and compacted every round next to the active list.
// Warps are then class-uniform, and the 64 pods of a warp cross into "n >= rot(p)" inside one or
// two groups instead of in four groups out of ten (config 3: 11 % of the scoring time).
// Admission order is untouched -- the active list stays in pod order; only which THREAD scores
// which pod changes.  posof[pod] = the pod's position in the active list, slotof[pod] = its slot
// in spod (proposals are exchanged between GPUs by slot).
//
// One tile (2048 entries) of a stable compaction: dst[q] = src[i] for the flagged entries, in
// order, and index_of[src[i]] = q.  chunk_tot[c] = flagged entries of chunk c (256 entries).
// Returns the number of flagged entries up to and including this tile (all threads).
__device__ __forceinline__ uint32_t compact_tile(uint32_t vt, uint32_t na, const uint32_t *flags, const uint32_t *src,
                                                 const uint32_t *chunk_tot, uint32_t *dst, uint32_t *index_of) {
    __shared__ uint32_t part[8], wtot[8];
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const uint32_t first = vt * (kScanTile / 256u);
    uint32_t before = 0;
    for (uint32_t c = tid; c < first; c += 256u) before += __ldcg(&chunk_tot[c]);
    const uint32_t i0 = vt * kScanTile + tid * 8u;
    uint32_t f[8], v[8], mine = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const bool in = i0 + k < na;
        f[k] = in ? flags[i0 + k] : 0u;
        v[k] = in ? src[i0 + k] : 0u;
        mine += f[k];
    }
    before = __reduce_add_sync(0xFFFFFFFFu, before);
    const uint32_t inc = warp_incl_scan(mine, (int)lane);
    __syncthreads();  // shared scratch of the previous tile / phase is no longer read
    if (lane == 0) part[warp] = before;
    if (lane == 31) wtot[warp] = inc;
    __syncthreads();
    uint32_t woff = 0, tile = 0;
    before = 0;
#pragma unroll
    for (uint32_t w = 0; w < 8; ++w) {
        before += part[w];
        woff += w < warp ? wtot[w] : 0u;
        tile += wtot[w];
    }
    uint32_t q = before + woff + inc - mine;
#pragma unroll
    for (int k = 0; k < 8; ++k)
        if (f[k]) {
            dst[q] = v[k];
            index_of[v[k]] = q;
            ++q;
        }
    return before + tile;
}

// slotof[spod[j]] = j after the one-time sort of a run.
__global__ void slot_init_kernel(const uint32_t *__restrict__ spod, uint32_t m, uint32_t *__restrict__ slotof) {
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < m) slotof[spod[j]] = j;
}

// ------------------------------------------------------------------ admission, one launch
//
// The admission step above is ~20 small dependent kernels (each a few microseconds of launch and
// drain) per round.  admit_coop_kernel runs the same phases inside ONE cooperative launch, with
// grid-wide barriers between them, and also builds the next round's packed node records, so a
// single-GPU round is two launches: score_kernel and this.  Same tile functions, same results.

struct AdmitArgs {
    const unsigned long long *best_key;  // by active position (score_kernel output)
    unsigned long long *best_key_clear;  // the same array: cleared at the end for the next round's atomicMax
    const int32_t *proposals;            // or: node row per scoring slot (exchanged between GPUs); else null
    const uint32_t *spod;                // this round's scoring order (pods; null: none, slot = active position)
    uint32_t *spod_next;                 // ... the next round's, compacted here
    uint32_t *posof, *slotof;            // per pod: position in the active list / slot in spod (updated here)
    uint32_t *gflags;                    // na u32 scratch: pending flag per slot
    uint32_t xw, xcnt;                   // placement group: ranks, local slots per rank (see SlotMap); else 1, 0
    uint32_t *slots;                     // kSlotsPerNode per node: first proposers of the round (0xFFFFFFFF = empty)
    const PodQ *podq;
    const uint32_t *active;
    uint32_t *next_active;
    int32_t *assign;
    uint32_t *keys, *vals, *keys_alt, *vals_alt;  // na each
    uint32_t *hist, *hist_scan, *scan_sums;       // 256 * ceil(na/2048) each; sums: one per 2048 of those
    Seg *seg;                                     // ceil(na/256)
    uint32_t *flags, *offs;                       // na each
    uint32_t *freec;                              // 3 x n_nodes, updated in place
    const uint32_t *node_cols;
    NodeQ *nodeq;
    uint32_t *counters;                           // [1] <- next active count
    uint32_t na, n_nodes, n_pad;
    uint32_t *resmin_done, *resmin_next;  // resource summaries (resmin_note): this round's (cleared here), the next round's (built here)
    int bits;
    unsigned long long *prof;  // optional: %globaltimer at each phase boundary (CTA 0), else null
    uint32_t *mailbox;         // mapped host memory: {seq, next active count, error flags}
    uint32_t seq;
};

__device__ __forceinline__ void prof_mark(const AdmitArgs &a, int slot) {
    if (a.prof && blockIdx.x == 0 && threadIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        a.prof[slot] = t;
    }
}

__global__ void __launch_bounds__(256) admit_coop_kernel(AdmitArgs a) {
    cg::grid_group grid = cg::this_grid();
    const uint32_t tid = threadIdx.x, cta = blockIdx.x, G = gridDim.x;
    prof_mark(a, 0);
    // the summary this round's scoring read is done with: empty it for the round after the next
    if (cta == 0 && tid < 2 * kResGpus) a.resmin_done[tid] = 0xFFFFFFFFu;
    const uint32_t na = a.na, N = a.n_nodes;
    const uint32_t nb = (na + kSortTile - 1) / kSortTile;

    // Which admission.  A node with f free GPUs can admit at most f pods that each ask for at
    // least one GPU, and the admitted pods are a PREFIX of its proposers in pod order (the sums S
    // only grow).  So when every pod asks for >= 1 GPU and no node has more than kSlotsPerNode
    // free (checked once per run on the device: counters[5]), the round needs, per node, only its
    // first free_gpu proposers: they are collected in a small sorted slot array per node by a
    // cascade of atomicMin (keep the smaller, carry the larger down), and one thread per node
    // walks its slots.  No sort, no segmented scan, 3 grid barriers instead of 7.  The cascade is
    // exact whatever the interleaving: every operation conserves the multiset {slot, carried
    // value} and slots only decrease, so a value among the k smallest can never be carried past
    // slot k - 1, and what reaches slot s + 1 is larger than slot s ends up.
    const bool slot_path = a.slots != nullptr && __ldcg(&a.counters[kSlotVeto]) == 0u;
    uint32_t *sk = a.keys, *sv = a.vals;
    if (slot_path) {
        for (uint32_t i0 = cta * 256; i0 < na; i0 += G * 256) {  // (whole warps: the votes below need every lane)
            const uint32_t i = i0 + tid;
            int32_t b = -1;
            if (i < na) {
                const uint32_t pod = a.active[i];
                if (a.proposals) {
                    b = a.proposals[exchange_index(a.spod ? a.slotof[pod] : i, a.xw, a.xcnt)];
                    if (b < -1 || b >= (int32_t)N) {
                        atomicOr(&a.counters[0], kErrProposal);
                        b = -1;
                    }
                } else {
                    const unsigned long long k64 = a.best_key[i];
                    if (k64 != 0ull) {
                        const uint32_t n = (0xFFFFFFFFu - (uint32_t)k64) + a.podq[pod].rot;
                        b = (int32_t)(n >= N ? n - N : n);
                    }
                }
                if (b < 0) a.assign[pod] = K8SGPU_ASSIGN_NONE;
            }
            // Lanes are consecutive positions: of the lanes of this warp that propose to the same
            // node only the first k can be among that node's first k proposers.  On a contested
            // node this takes most of the traffic off its eight slot words.
            const uint32_t peers = __match_any_sync(0xFFFFFFFFu, b);
            if (b < 0) continue;
            uint32_t *sl = a.slots + (size_t)b * kSlotsPerNode;
            const uint32_t k = a.freec[b];  // 1 .. kSlotsPerNode: the node was feasible for this pod
            if (k - 1u >= kSlotsPerNode) {  // a proposal the scoring kernel cannot have made for these tables
                atomicOr(&a.counters[0], kErrProposal);
                continue;
            }
            if ((uint32_t)__popc(peers & ((1u << (tid & 31u)) - 1u)) >= k) continue;
            // the k-th smallest so far is already before me: I cannot be among the first k
            if (__ldcg(&sl[k - 1]) < i) continue;
            uint32_t v = i;
            for (uint32_t j = 0; j < k; ++j) {
                const uint32_t old = atomicMin(&sl[j], v);
                if (old > v) {
                    if (old == 0xFFFFFFFFu) break;  // took an empty slot: nothing to carry
                    v = old;
                }
            }
        }
        grid.sync();
        prof_mark(a, 1);
        for (uint32_t n = cta * 256 + tid; n < N; n += G * 256) {
            uint4 *sl = reinterpret_cast<uint4 *>(a.slots + (size_t)n * kSlotsPerNode);
            const uint4 s0 = __ldcg(sl), s1 = __ldcg(sl + 1);
            if (s0.x == 0xFFFFFFFFu) continue;  // nobody proposed to this node
            const uint32_t pos[kSlotsPerNode] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
            uint32_t pod[kSlotsPerNode];
            uint4 rq[kSlotsPerNode];
#pragma unroll
            for (uint32_t j = 0; j < kSlotsPerNode; ++j) pod[j] = pos[j] != 0xFFFFFFFFu ? a.active[pos[j]] : 0u;
#pragma unroll
            for (uint32_t j = 0; j < kSlotsPerNode; ++j)
                rq[j] = pos[j] != 0xFFFFFFFFu ? *reinterpret_cast<const uint4 *>(a.podq + pod[j]) : make_uint4(0, 0, 0, 0);
            const uint32_t fg = a.freec[n], fc = a.freec[(size_t)N + n], fm = a.freec[2 * (size_t)N + n];
            uint32_t sg = 0, sc = 0, sm = 0, ag = 0, ac = 0, am = 0;
            bool open = true;
#pragma unroll
            for (uint32_t j = 0; j < kSlotsPerNode; ++j) {
                if (pos[j] == 0xFFFFFFFFu) open = false;
                if (!open) continue;
                sg = sat_add(sg, rq[j].x);  // a13: inclusive sums over the proposers, in pod order
                sc = sat_add(sc, rq[j].y);
                sm = sat_add(sm, rq[j].z);
                if (sg <= fg && sc <= fc && sm <= fm) {
                    a.assign[pod[j]] = (int32_t)n;
                    ag += rq[j].x;
                    ac += rq[j].y;
                    am += rq[j].z;
                } else {
                    open = false;  // the sums only grow: nobody after this one fits either
                }
            }
            a.freec[n] = fg - ag;  // this thread owns the node: plain stores
            a.freec[(size_t)N + n] = fc - ac;
            a.freec[2 * (size_t)N + n] = fm - am;
            sl[0] = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu);  // empty again for the next round
            sl[1] = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu);
        }
        grid.sync();
        prof_mark(a, 2);
        prof_mark(a, 3);
        prof_mark(a, 4);
    } else {
    // phase 0: packed best key -> proposal; sort key = node (N for "no feasible node"); digit-0 histogram
    for (uint32_t vt = cta; vt < nb; vt += G) {
        // all loads of the tile first: the stores below would otherwise order each row's loads
        // behind the previous row's stores (8 dependent round trips instead of 2)
        unsigned long long k64[kSortTile / 256];
        uint32_t pod[kSortTile / 256], rot[kSortTile / 256];
#pragma unroll
        for (uint32_t r = 0; r < kSortTile / 256; ++r) {
            uint32_t i = vt * kSortTile + r * 256 + tid;
            pod[r] = i < na ? a.active[i] : 0u;
            if (a.proposals) {  // already decoded: re-encode as "distance 0 from rot = node" below
                // proposals arrive by slot, and in a group by (rank, local slot)
                int32_t b = i < na ? a.proposals[exchange_index(a.spod ? a.slotof[pod[r]] : i, a.xw, a.xcnt)] : -1;
                if (b < -1 || b >= (int32_t)N) {  // not something score_kernel can produce: a stale or
                    atomicOr(&a.counters[0], kErrProposal);  // unsynchronised buffer was handed in
                    b = -1;
                }
                k64[r] = b < 0 ? 0ull : (0x100000000ull | (unsigned long long)(0xFFFFFFFFu - (uint32_t)b));
            } else {
                k64[r] = i < na ? a.best_key[i] : 0ull;
            }
        }
#pragma unroll
        for (uint32_t r = 0; r < kSortTile / 256; ++r)
            rot[r] = (k64[r] != 0ull && !a.proposals) ? a.podq[pod[r]].rot : 0u;
#pragma unroll
        for (uint32_t r = 0; r < kSortTile / 256; ++r) {
            uint32_t i = vt * kSortTile + r * 256 + tid;
            if (i < na) {
                uint32_t key = N;
                if (k64[r] != 0ull) {
                    uint32_t n = (0xFFFFFFFFu - (uint32_t)k64[r]) + rot[r];
                    key = n >= N ? n - N : n;
                } else {
                    a.assign[pod[r]] = K8SGPU_ASSIGN_NONE;
                }
                a.keys[i] = key;
                a.vals[i] = pod[r];  // payload = pod id
            }
        }
        __syncthreads();
        sort_hist_tile(a.keys, na, 0, 0xFFu, a.hist, nb, vt);
    }
    grid.sync();
    prof_mark(a, 1);

    // stable LSD radix sort by node, 8 bits a pass
    uint32_t *dk = a.keys_alt, *dv = a.vals_alt;
    for (int shift = 0; shift < a.bits; shift += 8) {
        const int w = a.bits - shift < 8 ? a.bits - shift : 8;
        const uint32_t mask = (1u << w) - 1u;
        if (shift > 0) {
            for (uint32_t vt = cta; vt < nb; vt += G) sort_hist_tile(sk, na, shift, mask, a.hist, nb, vt);
            grid.sync();
        }
        if (nb <= kLocalScanMaxTiles) {  // every CTA scans the small histogram itself: no scan phase, no barrier
            for (uint32_t vt = cta; vt < nb; vt += G)
                sort_scatter_tile(sk, sv, dk, dv, na, shift, mask, a.hist, nb, vt, true);
        } else {
            coop_exclusive_scan(grid, a.hist, a.hist_scan, nb * 256u, a.scan_sums, nullptr);
            for (uint32_t vt = cta; vt < nb; vt += G)
                sort_scatter_tile(sk, sv, dk, dv, na, shift, mask, a.hist_scan, nb, vt);
        }
        grid.sync();
        uint32_t *t;
        t = sk; sk = dk; dk = t;
        t = sv; sv = dv; dv = t;
    }
    prof_mark(a, 2);

    // segmented prefix sums over the proposers of each node, in pod order
    const uint32_t nb2 = (na + kSegThreads - 1) / kSegThreads;
    for (uint32_t vb = cta; vb < nb2; vb += G) {
        uint32_t key;
        Seg x = load_seg(sk, sv, a.active, a.podq, vb * kSegThreads + tid, na, &key);
        seg_block_scan(x, &a.seg[vb]);
        __syncthreads();
    }
    grid.sync();
    prof_mark(a, 3);
    const bool local_carry = nb2 <= kLocalCarryMaxBlocks;
    if (!local_carry) {
        if (cta == 0) seg_carry_block(a.seg, nb2);
        grid.sync();
    }
    prof_mark(a, 4);
    for (uint32_t vb = cta; vb < nb2; vb += G) {  // S <= free (a13); admitted flags go to offs[] for now
        const uint32_t i = vb * kSegThreads + tid;
        // this thread's loads first, so that they are in flight while the carry is folded
        uint32_t key;
        Seg x = load_seg(sk, sv, a.active, a.podq, i, na, &key);
        const bool real = i < na && key < N;
        const uint32_t pod = i < na ? sv[i] : 0u;
        const uint32_t fg = real ? a.freec[key] : 0u, fc = real ? a.freec[(size_t)N + key] : 0u,
                       fm = real ? a.freec[2 * (size_t)N + key] : 0u;
        const Seg carry = local_carry ? seg_carry_local(a.seg, vb) : a.seg[vb];
        Seg s = seg_combine(carry, seg_block_scan(x, nullptr));
        if (i < na) {
            uint32_t ok = 0;
            if (real) {
                ok = (s.g <= fg) & (s.c <= fc) & (s.m <= fm);
                if (ok) a.assign[pod] = (int32_t)key;
            }
            a.offs[i] = ok;
        }
        __syncthreads();
    }
    grid.sync();
    }  // (sort-based admission)
    prof_mark(a, 5);
    // free -= admitted (commutative integer atomics), pending flags in ORIGINAL active order, and
    // the number of still-pending pods per 256-element chunk (in hist_scan: the sort is done with it)
    const uint32_t nt = (na + kScanTile - 1) / kScanTile, nchunks = (na + 255u) / 256u;
    const bool local_scan = nt <= kLocalScanMaxTiles;
    uint32_t *chunk_tot = a.hist_scan;
    __shared__ uint32_t red[8];
    for (uint32_t c = cta; c < nchunks; c += G) {
        const uint32_t i = c * 256 + tid;
        uint32_t f = 0;
        if (i < na) {
            // independent loads first (the two chains then overlap instead of running back to back)
            const uint32_t pod_here = a.active[i];
            const uint32_t admitted = slot_path ? 0u : a.offs[i], key = slot_path ? 0u : sk[i],
                           pod_sorted = slot_path ? 0u : sv[i];
            const int32_t state = a.assign[pod_here];
            if (admitted) {  // (sort-based admission; the slot walk has updated its nodes already)
                const PodQ *q = a.podq + pod_sorted;
                const uint32_t rg = q->req_gpu, rc = q->req_mcpu, rm = q->req_mib;
                if (rg) atomicSub(&a.freec[key], rg);
                if (rc) atomicSub(&a.freec[(size_t)N + key], rc);
                if (rm) atomicSub(&a.freec[2 * (size_t)N + key], rm);
            }
            f = state == K8SGPU_ASSIGN_PENDING ? 1u : 0u;
            a.flags[i] = f;
        }
        const uint32_t cnt = (uint32_t)__syncthreads_count((int)f);
        if (tid == 0) chunk_tot[c] = cnt;
    }
    uint32_t *chunk_tot2 = chunk_tot + nchunks;  // the same count per chunk of the scoring order
    if (a.spod) {
        for (uint32_t c = cta; c < nchunks; c += G) {
            const uint32_t j = c * 256 + tid;
            uint32_t g = 0;
            if (j < na) {
                g = a.assign[a.spod[j]] == K8SGPU_ASSIGN_PENDING ? 1u : 0u;
                a.gflags[j] = g;
            }
            const uint32_t cnt = (uint32_t)__syncthreads_count((int)g);
            if (tid == 0) chunk_tot2[c] = cnt;
        }
    }
    grid.sync();
    prof_mark(a, 6);
    if (a.spod) {
        // Next active list (pod order) and next scoring order (its own order), each a stable
        // compaction by the same pending test; the CTA that owns the last tile of the active list
        // knows the next round's length and tells the host before it goes on.
        for (uint32_t vt = cta; vt < nt; vt += G) {
            const uint32_t upto = compact_tile(vt, na, a.flags, a.active, chunk_tot, a.next_active, a.posof);
            if (vt == nt - 1 && tid == 0) {
                a.counters[1] = upto;
                volatile uint32_t *mb = a.mailbox;
                mb[1] = upto;
                mb[2] = a.counters[0];
                __threadfence_system();
                mb[0] = a.seq;
            }
        }
        for (uint32_t vt = cta; vt < nt; vt += G)
            compact_tile(vt, na, a.gflags, a.spod, chunk_tot2, a.spod_next, a.slotof);
        prof_mark(a, 7);
    } else if (local_scan) {
        // Stable compaction without a scan phase: each CTA adds up the pending counts of the tiles
        // before its own, scans its tile in registers and writes the survivors.  The CTA that owns
        // the last tile knows the next round's length and tells the host before it scatters.
        for (uint32_t vt = cta; vt < nt; vt += G) {
            uint32_t before = 0;
            if (tid < 32) {
                for (uint32_t t = tid; t < vt * (kScanTile / 256u); t += 32) before += chunk_tot[t];
                before = __reduce_add_sync(0xFFFFFFFFu, before);
                if (tid == 0) red[0] = before;
            }
            const uint32_t i0 = vt * kScanTile + tid * 8u;
            uint32_t f[8], act[8], mine = 0;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const bool in = i0 + k < na;
                f[k] = in ? a.flags[i0 + k] : 0u;
                act[k] = in ? a.active[i0 + k] : 0u;
                mine += f[k];
            }
            const uint32_t inc = warp_incl_scan(mine, tid & 31);
            __syncthreads();  // red[0] written; red[] of the previous phase / tile no longer read
            before = red[0];
            __syncthreads();
            if ((tid & 31u) == 31u) red[tid >> 5] = inc;
            __syncthreads();
            uint32_t woff = 0, all = 0;
#pragma unroll
            for (int w = 0; w < 8; ++w) {
                const uint32_t t = red[w];
                if (w < (int)(tid >> 5)) woff += t;
                all += t;
            }
            if (vt == nt - 1 && tid == 0) {  // tell the host now: it can queue the next round while we finish
                a.counters[1] = before + all;
                volatile uint32_t *mb = a.mailbox;
                mb[1] = before + all;
                mb[2] = a.counters[0];
                __threadfence_system();
                mb[0] = a.seq;
            }
            uint32_t pos = before + woff + inc - mine;
#pragma unroll
            for (int k = 0; k < 8; ++k)
                if (f[k]) a.next_active[pos++] = act[k];
            __syncthreads();
        }
        if (nt == 0 && cta == 0 && tid == 0) {  // (not reached: the host never launches with na == 0)
            a.counters[1] = 0;
            volatile uint32_t *mb = a.mailbox;
            mb[1] = 0;
            mb[2] = a.counters[0];
            __threadfence_system();
            mb[0] = a.seq;
        }
        prof_mark(a, 7);
    } else {
        coop_exclusive_scan(grid, a.flags, a.offs, na, a.scan_sums, a.counters + 1);
        prof_mark(a, 7);
        if (cta == 0 && tid == 0) {  // tell the host now: it can queue the next round while we finish
            volatile uint32_t *mb = a.mailbox;
            mb[1] = a.counters[1];
            mb[2] = a.counters[0];
            __threadfence_system();
            mb[0] = a.seq;
        }
        // stable compaction of the active list
        for (uint32_t i = cta * 256 + tid; i < na; i += G * 256)
            if (a.flags[i]) a.next_active[a.offs[i]] = a.active[i];
    }
    // the next round's packed node records, and its (shorter) key array cleared: one stream
    // operation fewer between this kernel and the next scoring launch
    for (uint32_t i = cta * 256 + tid; i < a.n_pad; i += G * 256) {  // (whole warps: n_pad is a multiple of 256)
        const NodeQ q = make_nodeq(a.node_cols, a.freec, N, i, nullptr, 0);
        a.nodeq[i] = q;
        resmin_note(a.resmin_next, a.counters, q);
    }
    for (uint32_t i = cta * 256 + tid; i < na; i += G * 256) a.best_key_clear[i] = 0ull;
    prof_mark(a, 8);
}

// ------------------------------------------------------------------ host side

int check_tables(k8sgpu_ctx *ctx, const k8sgpu_node_table *nodes, const k8sgpu_pod_table *pods) {
    if (!nodes || !pods) return fail(ctx, K8SGPU_E_ARG, "place: NULL table");
    if (nodes->n >= 0x80000000u - 4096u || pods->m >= 0x80000000u)
        return fail(ctx, K8SGPU_E_ARG, "place: table too large (n=%u m=%u)", nodes->n, pods->m);
    const uint32_t *nc[8] = {nodes->free_gpu, nodes->free_mcpu, nodes->free_mib, nodes->cap_gpu,
                             nodes->sku_id,   nodes->zone_id,   nodes->taint_mask, nodes->flags};
    const uint32_t *pc[8] = {pods->req_gpu,  pods->req_mcpu, pods->req_mib,   pods->sku_mask,
                             pods->zone_mask, pods->tol_mask, pods->pref_zone, pods->priority};
    for (int c = 0; c < 8; ++c) {
        if (nodes->n && !nc[c]) return fail(ctx, K8SGPU_E_ARG, "place: node column %d is NULL", c);
        if (pods->m && !pc[c] && c != 7)  // priority (column 7) is reserved and may be NULL
            return fail(ctx, K8SGPU_E_ARG, "place: pod column %d is NULL", c);
    }
    return K8SGPU_OK;
}

// share != null (placement groups): only pod rows [lo, hi) are copied from the host, into a staging
// block with `stride` rows per column; the group's ranks then all-gather the staging block over
// NVLink and copy it into place (group.cu), so each rank moves 1/world of the pod table over PCIe.
int place_upload(k8sgpu_ctx *ctx, const k8sgpu_node_table *nodes, const k8sgpu_pod_table *pods, const PodShare *share) {
    int rc = check_tables(ctx, nodes, pods);
    if (rc) return rc;
    PlaceState &st = ctx->place;
    cudaStream_t s = ctx->stream;
    st.uploaded = false;
    st.n = nodes->n;
    st.m = pods->m;
    st.n_pad = ceil_div(st.n ? st.n : 1u, kNodeTile) * kNodeTile;
    const size_t n = st.n, m = st.m;
    K8S_CUDA(ctx, st.node_cols.reserve(8 * n * 4 + 16));
    K8S_CUDA(ctx, st.pod_cols.reserve(8 * m * 4 + 16));
    const uint32_t *nc[8] = {nodes->free_gpu, nodes->free_mcpu, nodes->free_mib, nodes->cap_gpu,
                             nodes->sku_id,   nodes->zone_id,   nodes->taint_mask, nodes->flags};
    const uint32_t *pc[8] = {pods->req_gpu,  pods->req_mcpu, pods->req_mib,   pods->sku_mask,
                             pods->zone_mask, pods->tol_mask, pods->pref_zone, pods->priority};
    // A caller that keeps the eight columns of a table back to back in one block (the layout
    // the device uses) gets one copy per table instead of eight: fewer, larger transfers.
    const int npc = pc[7] ? 8 : 7;  // the reserved priority column is not read by any kernel: NULL = not sent
    bool nslab = n > 0, pslab = m > 0;
    for (int c = 0; c + 1 < 8; ++c) nslab = nslab && nc[c + 1] == nc[c] + n;
    for (int c = 0; c + 1 < npc; ++c) pslab = pslab && pc[c + 1] == pc[c] + m;
    if (nslab) K8S_CUDA(ctx, cudaMemcpyAsync(st.node_cols.p, nc[0], 8 * n * 4, cudaMemcpyHostToDevice, s));
    if (share) pslab = false;
    if (pslab) K8S_CUDA(ctx, cudaMemcpyAsync(st.pod_cols.p, pc[0], (size_t)npc * m * 4, cudaMemcpyHostToDevice, s));
    for (int c = 0; c < 8; ++c) {
        if (n && !nslab)
            K8S_CUDA(ctx, cudaMemcpyAsync(st.node_cols.as<uint32_t>() + c * n, nc[c], n * 4, cudaMemcpyHostToDevice, s));
        if (m && !pslab && c < npc && !share)
            K8S_CUDA(ctx, cudaMemcpyAsync(st.pod_cols.as<uint32_t>() + c * m, pc[c], m * 4, cudaMemcpyHostToDevice, s));
        if (share && c < npc && share->hi > share->lo)
            K8S_CUDA(ctx, cudaMemcpyAsync(share->staging + (size_t)c * share->stride + share->lo, pc[c] + share->lo,
                                          (size_t)(share->hi - share->lo) * 4, cudaMemcpyHostToDevice, s));
    }
    ctx->stats.h2d_bytes += (8 * n + (size_t)npc * (share ? share->hi - share->lo : m)) * 4;
    // run-state buffers
    K8S_CUDA(ctx, st.free_cols.reserve(3 * n * 4 + 16));
    K8S_CUDA(ctx, st.nodeq.reserve((size_t)st.n_pad * sizeof(NodeQ)));
    K8S_CUDA(ctx, st.podq.reserve(m * sizeof(PodQ) + 32));
    K8S_CUDA(ctx, st.assign.reserve(m * 4 + 16));
    K8S_CUDA(ctx, st.active[0].reserve(m * 4 + 16));
    K8S_CUDA(ctx, st.active[1].reserve(m * 4 + 16));
    for (int b = 0; b < 2; ++b) {
        K8S_CUDA(ctx, st.order[b].reserve(m * 4 + 16));
        K8S_CUDA(ctx, st.inv[b].reserve(m * 4 + 16));
    }
    K8S_CUDA(ctx, st.slots.reserve(((size_t)n + 1) * kSlotsPerNode * 4));
    K8S_CUDA(ctx, st.best_key.reserve(m * 8 + 16));
    K8S_CUDA(ctx, st.best.reserve(m * 4 + 16));
    K8S_CUDA(ctx, st.sort_keys.reserve(m * 4 + 16));
    K8S_CUDA(ctx, st.sort_vals.reserve(m * 4 + 16));
    K8S_CUDA(ctx, st.flags.reserve(2 * m * 4 + 16));
    K8S_CUDA(ctx, st.seg.reserve(((size_t)ceil_div(st.m ? st.m : 1u, kSegThreads) + 1) * sizeof(Seg)));
    K8S_CUDA(ctx, st.counters.reserve(64 * 4));
    K8S_CUDA(ctx, st.like_seen.reserve((size_t)(16u << kLikeBits) * 4));
    K8S_CUDA(ctx, st.res_min.reserve(2 * 2 * kResGpus * 4));
    {  // scratch of the single-launch admission path, sized for all m pods up front
        const size_t nbmax = ceil_div(st.m ? st.m : 1u, kSortTile);
        K8S_CUDA(ctx, st.sort.keys_alt.reserve(m * 4 + 16));
        K8S_CUDA(ctx, st.sort.vals_alt.reserve(m * 4 + 16));
        K8S_CUDA(ctx, st.sort.hist.reserve((nbmax * 256 * 2 + 256) * 4));
        K8S_CUDA(ctx, st.sort.scan_tmp.reserve((nbmax * 256 / kScanTile + m / kScanTile + 128) * 4));
    }
    st.uploaded = true;
    return K8SGPU_OK;
}

int round_begin(k8sgpu_ctx *ctx, const k8sgpu_policy *policy, cudaStream_t s) {
    PlaceState &st = ctx->place;
    if (!st.uploaded) return fail(ctx, K8SGPU_E_ARG, "place: no tables uploaded");
    st.policy_flags = policy ? policy->flags : 0u;
    st.max_rounds = policy ? policy->max_rounds : 0u;
    st.cur = 0;
    st.n_active = st.m;
    st.rounds = 0;
    st.evals = 0;
    st.timed = 0;
    st.timed_evals = 0;
    st.nodeq_stale = true;
    st.best_cleared_len = 0;
    st.score_res = kDefaultRes;  // which eval_pair variant (see its comment); all are bit-identical
    if (ctx->tune.res >= 0) st.score_res = ctx->tune.res;
    K8S_CUDA(ctx, cudaMemsetAsync(st.counters.p, 0, 64 * 4, s));
    K8S_CUDA(ctx, cudaMemsetAsync(st.res_min.p, 0xFF, 2 * 2 * kResGpus * 4, s));
    K8S_CUDA(ctx, cudaMemsetAsync(st.slots.p, 0xFF, ((size_t)st.n + 1) * kSlotsPerNode * 4, s));  // every node's slots empty
    if (st.n)
        K8S_CUDA(ctx, cudaMemcpyAsync(st.free_cols.p, st.node_cols.p, 3 * (size_t)st.n * 4, cudaMemcpyDeviceToDevice, s));
    // Scoring order (see "scoring order" above): only with the class-specialised eval_pair (RES = 4)
    // and the one-launch admission, which compacts it every round next to the active list.
    const bool ordered = st.want_order && st.score_res == 4 && coop_admission(ctx) && ctx->tune.order != 0 && st.m > 0;
    st.order_valid = false;
    if (st.n)  // which taints occur at all: decides how each pod's toleration clause is scored
        K8S_LAUNCH(ctx, taint_union_kernel, ceil_div(st.n, 256), 256, 0, s, st.node_cols.as<uint32_t>() + 6 * (size_t)st.n,
                   st.n, st.counters.as<uint32_t>() + kTaintUnion);
    // Sorting like pods together (a hash of their masks and preferred zone, under the class) lets a
    // thread share clauses between its two pods (eval_pair4_u2: config 4 272 against 309 ms,
    // config 3 4.56 against 4.69), but splits each class into runs whose pods' rot(p) lie further
    // apart, and every 64-node group that holds some rot(p) of a warp costs it about twice a plain
    // one: a warp meets 64 / (pods per run) + 64 / N of its groups that way.  So the hash is kept
    // only while the runs average kLikeMinRun pods (counted by pod_prep, applied by
    // like_hash_fix_kernel -- no host round trip).  order = 2 + b forces b hash bits.
    const bool adaptive = ctx->tune.order < 2;
    const uint32_t hash_bits = adaptive ? kLikeBits : (uint32_t)(ctx->tune.order - 2 > 12 ? 12 : ctx->tune.order - 2);
    if (st.m && ordered && adaptive) K8S_CUDA(ctx, cudaMemsetAsync(st.like_seen.p, 0, (size_t)(16u << kLikeBits) * 4, s));
    if (st.m) {
        uint32_t *spod = st.order[0].as<uint32_t>(), *posof = st.inv[0].as<uint32_t>(), *slotof = st.inv[1].as<uint32_t>();
        uint32_t *keys = st.sort_keys.as<uint32_t>();
        K8S_LAUNCH(ctx, pod_prep_kernel, ceil_div(st.m, 256), 256, 0, s, st.pod_cols.as<uint32_t>(), st.m, st.n,
                   st.policy_flags, st.score_res == 4 ? 1 : 0, st.podq.as<PodQ>(), st.assign.as<int32_t>(),
                   st.active[0].as<uint32_t>(), st.counters.as<uint32_t>(), ordered ? keys : nullptr, spod, posof,
                   hash_bits, ordered && adaptive ? st.like_seen.as<uint32_t>() : nullptr);
        if (ordered) {
            if (adaptive) K8S_LAUNCH(ctx, like_hash_fix_kernel, ceil_div(st.m, 256), 256, 0, s, keys, st.m, st.counters.as<uint32_t>());
            int rc = sort_pairs(ctx, s, keys, spod, st.m, 24, st.sort);  // by (class, like pods, where rot(p) falls), pod order inside
            if (rc) return rc;
            K8S_LAUNCH(ctx, slot_init_kernel, ceil_div(st.m, 256), 256, 0, s, spod, st.m, slotof);
            st.order_valid = true;
        }
    }
    return K8SGPU_OK;
}

// Scores active positions [lo,hi) into d_best[0 .. hi-lo).
// prep: rebuild the packed node records first; finalize: decode the packed keys into d_best.
// (The single-launch admission path needs neither after round 0.)
// Local slots of one rank of xw when the round's na slots are dealt in blocks (see SlotMap).
uint32_t group_slot_count(uint32_t na, uint32_t xw) { return ceil_div(ceil_div(na, kXBlock), xw) * kXBlock; }

int round_score(k8sgpu_ctx *ctx, uint32_t lo, uint32_t hi, int32_t *d_best, cudaStream_t s, bool prep, bool finalize,
                uint32_t xw, uint32_t xrank) {
    PlaceState &st = ctx->place;
    if (hi > st.n_active || lo > hi) return fail(ctx, K8SGPU_E_ARG, "score: bad slice [%u,%u) of %u", lo, hi, st.n_active);
    if (xw > 1 && (lo != 0 || hi != st.n_active || xrank >= xw))
        return fail(ctx, K8SGPU_E_ARG, "score: a group round covers the whole active list");
    // node records from the current free columns (validated on the first round)
    uint32_t *resmin = st.res_min.as<uint32_t>() + (st.rounds & 1u) * 2 * kResGpus;  // this round's summary
    if (prep) {
        if (st.rounds) K8S_CUDA(ctx, cudaMemsetAsync(resmin, 0xFF, 2 * kResGpus * 4, s));
        K8S_LAUNCH(ctx, node_prep_kernel, ceil_div(st.n_pad, 256), 256, 0, s, st.node_cols.as<uint32_t>(),
                   st.free_cols.as<uint32_t>(), st.n, st.n_pad, st.nodeq.as<NodeQ>(), st.counters.as<uint32_t>(),
                   st.rounds == 0 ? 1 : 0, resmin);
        st.nodeq_stale = false;
    }
    SlotMap sm;
    sm.lo = lo;
    sm.hi = hi;
    sm.xw = xw;
    sm.xrank = xrank;
    sm.xcnt = xw > 1 ? group_slot_count(hi, xw) : hi - lo;
    const uint32_t cnt = sm.xcnt;  // local slots of this launch (a group rank's last block may be short or empty)
    uint32_t live = hi - lo;        // ... and how many of them hold a pod
    if (xw > 1) {
        live = 0;
        for (uint32_t b = xrank; (uint64_t)b * kXBlock < hi; b += xw) live += hi - b * kXBlock < kXBlock ? hi - b * kXBlock : kXBlock;
    }
    st.last_score_slots = cnt;
    if (cnt == 0) return K8SGPU_OK;
    const uint32_t n_tiles = st.n_pad / kNodeTile;
    // Grid shape.  The kernel is bound by the ALU pipe and the issue slots, so what matters is an
    // even spread over the 148 SMs, a short tail, and enough node tiles per CTA to pay for a CTA's
    // start-up (a tenth of a one-tile CTA's instructions since the loops got short): aim for ~128
    // CTAs per SM but at least two tiles per CTA.  Measured (profiles/r02_order_knobs.txt):
    // config 3, 40 tiles: 20 splits 3.55 ms, 40 3.62, 10 3.73; config 4, 391 tiles: 10 splits
    // 181.2 ms, 5 182.3, 3 185.1.
    const uint32_t sms = (uint32_t)ctx->sm_count;
    // Pods per thread: two share every shared-memory read of a node record and, when they are
    // alike, the clauses that depend on the node and on what they have in common (eval_pair4_u2,
    // eval_pair4_ur2; only the two-pod kernel has those loops).  Four per thread is a knob: 80
    // registers, three CTAs per SM (config 4: 355 vs 272 ms per placement when last measured).
    int ppt = st.score_res == 4 ? 2 : 1;
    if (ctx->tune.ppt >= 1) ppt = ctx->tune.ppt >= 4 ? 4 : ctx->tune.ppt >= 2 ? 2 : 1;
    if (ppt == 4 && st.score_res != 4) ppt = 2;
    const uint32_t xb = ceil_div(cnt, 256u * ppt);
    uint32_t splits = ceil_div(128u * sms, xb);
    uint32_t cap = n_tiles / 2u;
    if ((uint64_t)xb * cap < 4ull * sms) cap = n_tiles;  // (few pods: the machine is not full otherwise)
    if (splits > cap) splits = cap;
    if (ctx->tune.splits >= 1) splits = (uint32_t)ctx->tune.splits;
    if (splits < 1) splits = 1;
    if (splits > n_tiles) splits = n_tiles;
    uint32_t tps = ceil_div(n_tiles, splits);
    splits = ceil_div(n_tiles, tps);
    if (splits > 65535u) {
        splits = 65535u;
        tps = ceil_div(n_tiles, splits);
        splits = ceil_div(n_tiles, tps);
    }
    // The last wave.  With CTAs of a few tiles the launch is a handful of waves over the resident
    // slots (4 per SM), and the SMs that run out of work first idle for up to a whole CTA (ncu,
    // config 3, 20 splits of 2 tiles: SM sub-partitions active 86 % of the elapsed cycles).  So the
    // last tiles -- about a wave and a half of CTAs -- are given one CTA each.
    uint32_t big_rows = splits;
    if (tps >= 2 && tps <= 4 && ctx->tune.splits < 1) {
        uint32_t small = ceil_div(6u * sms, xb);  // tiles handled one per CTA
        if (small >= n_tiles) {
            small = n_tiles;
        } else {
            small += (n_tiles - small) % tps;  // the rest in whole splits
        }
        big_rows = (n_tiles - small) / tps;
        if (big_rows + small <= 65535u) splits = big_rows + small;
        else big_rows = splits;
    }
    const int use_atomic = splits > 1;
    unsigned long long *bk = st.best_key.as<unsigned long long>();
    if (use_atomic && !(st.best_cleared_len >= ((st.order_valid || xw > 1) ? st.n_active : hi))) {  // (the single-launch admission leaves it cleared)
        if (st.order_valid || xw > 1)  // the launch's slots are scattered positions: clear the whole list
            K8S_CUDA(ctx, cudaMemsetAsync(bk, 0, (size_t)st.n_active * 8, s));
        else
            K8S_CUDA(ctx, cudaMemsetAsync(bk + lo, 0, (size_t)cnt * 8, s));
    }
    st.best_cleared_len = 0;
    SlotView view;  // scoring order of this round, if the run keeps one
    view.active = st.active[st.cur].as<uint32_t>();
    view.spod = st.order_valid ? st.order[st.cur].as<uint32_t>() : nullptr;
    view.posof = st.inv[0].as<uint32_t>();
    const bool timed = st.timed < kMaxTimedRounds;  // (the two event records cost ~4 us per round)
    if (timed) K8S_CUDA(ctx, cudaEventRecord(ctx->ev_score[2 * st.timed], s));
    dim3 grid(xb, splits);
    const int res = st.score_res;  // fixed at round_begin: the pod records are laid out for it
#define K8S_SCORE(P, R, B) K8S_SCORE_U(P, R, B, 16)
#define K8S_SCORE_U(P, R, B, U)                                                                                    \
    K8S_LAUNCH(ctx, (score_kernel<P, R, B, U>), grid, 256, 0, s, st.nodeq.as<NodeQ>(), n_tiles, tps, st.podq.as<PodQ>(), \
               st.active[st.cur].as<uint32_t>(), sm, bk, st.n, use_atomic, 0xFFFFFFFFu, ctx->tune.force_cls, view.spod,  \
               view.posof, st.counters.as<uint32_t>(), resmin, big_rows)
    // third argument: resident CTAs per SM the register budget is set for (measured best per variant)
    if (ppt == 4) K8S_SCORE(4, 4, 3);
    else if (ppt == 2 && res >= 4 && tps <= 4) K8S_SCORE_U(2, 4, 4, 8);  // short-lived CTAs: see UNR
    else if (ppt == 2 && res >= 4) K8S_SCORE(2, 4, 4);
    else if (ppt == 2 && res == 3) K8S_SCORE(2, 3, 5);
    else if (ppt == 2 && res == 2) K8S_SCORE(2, 2, 5);
    else if (ppt == 2 && res == 1) K8S_SCORE(2, 1, 5);
    else if (ppt == 2) K8S_SCORE(2, 0, 5);
    else if (res >= 4) K8S_SCORE(1, 4, 6);
    else if (res == 3) K8S_SCORE(1, 3, 6);
    else if (res == 2) K8S_SCORE(1, 2, 6);
    else if (res == 1) K8S_SCORE(1, 1, 6);
    else K8S_SCORE(1, 0, 6);
#undef K8S_SCORE
#undef K8S_SCORE_U
    if (timed) {
        K8S_CUDA(ctx, cudaEventRecord(ctx->ev_score[2 * st.timed + 1], s));
        st.timed++;
        st.timed_evals += (uint64_t)live * st.n;
    }
    if (finalize)
        K8S_LAUNCH(ctx, score_finalize_kernel, ceil_div(cnt, 256), 256, 0, s, bk, st.podq.as<PodQ>(), view, sm, cnt, st.n,
                   d_best);
    return K8SGPU_OK;
}

// Admission over all n_active proposals; leaves the next active list in active[cur^1] and its
// length (device) in counters[1].  Synchronises the stream to read the count back.
int round_admit(k8sgpu_ctx *ctx, const int32_t *d_best_all, cudaStream_t s, uint32_t *n_active_out) {
    PlaceState &st = ctx->place;
    const uint32_t na = st.n_active;
    uint32_t *cnts = st.counters.as<uint32_t>();
    if (na) {
        uint32_t *keys = st.sort_keys.as<uint32_t>(), *vals = st.sort_vals.as<uint32_t>();
        uint32_t *active = st.active[st.cur].as<uint32_t>(), *next = st.active[st.cur ^ 1].as<uint32_t>();
        uint32_t *flags = st.flags.as<uint32_t>(), *offs = flags + st.m;
        const uint32_t g = ceil_div(na, 256);
        K8S_LAUNCH(ctx, admit_keys_kernel, g, 256, 0, s, d_best_all, active, na, st.n, st.assign.as<int32_t>(), keys, vals,
                   cnts);
        int rc = sort_pairs(ctx, s, keys, vals, na, key_bits(st.n), st.sort);
        if (rc) return rc;
        const uint32_t nb = ceil_div(na, kSegThreads);
        Seg *seg = st.seg.as<Seg>();
        K8S_LAUNCH(ctx, admit_seg_reduce_kernel, nb, kSegThreads, 0, s, keys, vals, active, st.podq.as<PodQ>(), na, seg);
        K8S_LAUNCH(ctx, admit_seg_carry_kernel, 1, kSegThreads, 0, s, seg, nb);
        K8S_LAUNCH(ctx, admit_apply_kernel, nb, kSegThreads, 0, s, keys, vals, active, st.podq.as<PodQ>(), na, st.n,
                   seg, st.free_cols.as<uint32_t>(), st.assign.as<int32_t>(), flags);
        K8S_LAUNCH(ctx, admit_free_update_kernel, g, 256, 0, s, keys, vals, active, st.podq.as<PodQ>(), flags, na,
                   st.n, st.free_cols.as<uint32_t>());
        K8S_LAUNCH(ctx, pending_flags_kernel, g, 256, 0, s, active, st.assign.as<int32_t>(), na, flags);
        rc = exclusive_scan_u32(ctx, s, flags, offs, na, st.scan_tmp, cnts + 1);
        if (rc) return rc;
        K8S_LAUNCH(ctx, compact_kernel, g, 256, 0, s, active, flags, offs, na, next);
    } else {
        K8S_CUDA(ctx, cudaMemsetAsync(cnts + 1, 0, 4, s));
    }
    K8S_CUDA(ctx, cudaMemcpyAsync(ctx->h_pinned, cnts, 8, cudaMemcpyDeviceToHost, s));
    K8S_CUDA(ctx, cudaStreamSynchronize(s));
    if (ctx->h_pinned[0] & kErrProposal)
        return fail(ctx, K8SGPU_E_ARG, "place: proposal buffer holds values outside [-1, n): was it written on "
                                       "another stream than the one passed here?");
    if (ctx->h_pinned[0] != 0u)
        return fail(ctx, K8SGPU_E_ARG, "place: %s table outside the input domain (see k8sgpu.h)",
                    (ctx->h_pinned[0] & kErrNode) ? "node" : "pod");
    if (na && ctx->h_pinned[1] >= na)  // every round with active pods settles at least one of them
        return fail(ctx, K8SGPU_E_ARG, "place: round %u made no progress (%u pods still active): proposals are "
                                       "inconsistent with the tables", st.rounds, na);
    st.evals += (uint64_t)na * st.n;
    st.rounds++;
    st.cur ^= 1;
    st.n_active = ctx->h_pinned[1];
    st.nodeq_stale = true;  // the free columns changed under the packed node records
    st.order_valid = false; // (this path does not rebuild the class-sorted scoring order)
    if (n_active_out) *n_active_out = st.n_active;
    return K8SGPU_OK;
}

// Single-launch admission (admit_coop_kernel) for the in-library single-GPU loop: consumes
// best_key[0..n_active) as score_kernel left it, and leaves the next round's node records ready.
// Launch half: returns as soon as the kernel is queued, so that a caller driving several GPUs can
// queue all of them before it waits for any (round_admit_coop_finish).
int round_admit_coop_launch(k8sgpu_ctx *ctx, cudaStream_t s, const int32_t *proposals, uint32_t xw) {
    PlaceState &st = ctx->place;
    const uint32_t na = st.n_active;
    uint32_t *cnts = st.counters.as<uint32_t>();
    if (ctx->coop_blocks_per_sm == 0) {
        int per_sm = 0;
        K8S_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, admit_coop_kernel, 256, 0));
        ctx->coop_blocks_per_sm = per_sm > 0 ? per_sm : -1;
    }
    if (ctx->coop_blocks_per_sm < 0) return fail(ctx, K8SGPU_E_CUDA, "cooperative admission kernel cannot be resident");
    const uint32_t nb = ceil_div(na, kSortTile);
    AdmitArgs a;
    a.best_key = st.best_key.as<unsigned long long>();
    a.best_key_clear = st.best_key.as<unsigned long long>();
    a.proposals = proposals;
    a.spod = st.order_valid ? st.order[st.cur].as<uint32_t>() : nullptr;
    a.spod_next = st.order[st.cur ^ 1].as<uint32_t>();
    a.posof = st.inv[0].as<uint32_t>();
    a.slotof = st.inv[1].as<uint32_t>();
    a.gflags = st.best.as<uint32_t>();  // (m words the one-launch admission does not use otherwise)
    a.xw = xw;
    a.xcnt = xw > 1 ? group_slot_count(st.n_active, xw) : 0;
    a.slots = ctx->tune.slot_admit != 0 ? st.slots.as<uint32_t>() : nullptr;
    a.podq = st.podq.as<PodQ>();
    a.active = st.active[st.cur].as<uint32_t>();
    a.next_active = st.active[st.cur ^ 1].as<uint32_t>();
    a.assign = st.assign.as<int32_t>();
    a.keys = st.sort_keys.as<uint32_t>();
    a.vals = st.sort_vals.as<uint32_t>();
    a.keys_alt = st.sort.keys_alt.as<uint32_t>();
    a.vals_alt = st.sort.vals_alt.as<uint32_t>();
    a.hist = st.sort.hist.as<uint32_t>();
    a.hist_scan = a.hist + (size_t)nb * 256;
    a.scan_sums = st.sort.scan_tmp.as<uint32_t>();
    a.seg = st.seg.as<Seg>();
    a.flags = st.flags.as<uint32_t>();
    a.offs = a.flags + st.m;
    a.freec = st.free_cols.as<uint32_t>();
    a.node_cols = st.node_cols.as<uint32_t>();
    a.nodeq = st.nodeq.as<NodeQ>();
    a.counters = cnts;
    a.na = na;
    a.n_nodes = st.n;
    a.n_pad = st.n_pad;
    a.resmin_done = st.res_min.as<uint32_t>() + (st.rounds & 1u) * 2 * kResGpus;
    a.resmin_next = st.res_min.as<uint32_t>() + ((st.rounds + 1u) & 1u) * 2 * kResGpus;
    a.bits = key_bits(st.n);
    a.prof = nullptr;
    const bool want_prof = ctx->tune.coop_prof != 0;
    if (want_prof) a.prof = reinterpret_cast<unsigned long long *>(cnts + 16);  // counters[16..34)
    // enough CTAs to spread the tiles, few enough that the grid barriers stay cheap
    uint32_t want = ceil_div(na, 512u), prep = ceil_div(st.n_pad, 2048u);
    if (prep > want) want = prep;
    uint32_t cap = (uint32_t)ctx->sm_count * (uint32_t)(ctx->coop_blocks_per_sm < 2 ? ctx->coop_blocks_per_sm : 2);
    uint32_t grid = want < 1 ? 1 : (want > cap ? cap : want);
    if (ctx->tune.coop_grid >= 1 && (uint32_t)ctx->tune.coop_grid <= cap) grid = (uint32_t)ctx->tune.coop_grid;
    a.mailbox = ctx->d_mailbox;
    a.seq = ++ctx->mail_seq;
    void *args[] = {&a};
    K8S_CUDA(ctx, cudaLaunchCooperativeKernel((void *)admit_coop_kernel, dim3(grid), dim3(256), args, 0, s));
    ctx->stats.kernel_launches++;
    st.coop_seq = a.seq;
    st.coop_grid = grid;
    return K8SGPU_OK;
}

// Wait half: the kernel posts {seq, count, flags} to mapped host memory as soon as the count is
// known; the host waits for that instead of a copy + stream synchronise.
int round_admit_coop_finish(k8sgpu_ctx *ctx, cudaStream_t s) {
    PlaceState &st = ctx->place;
    const uint32_t na = st.n_active, grid = st.coop_grid;
    uint32_t *cnts = st.counters.as<uint32_t>();
    const bool want_prof = ctx->tune.coop_prof != 0;
    volatile uint32_t *mb = ctx->h_pinned + 32;
    int wrc = wait_mailbox(ctx, mb, st.coop_seq, s, "place: admission");
    if (wrc) return wrc;
    const uint32_t next_count = mb[1], err_flags = mb[2];
    if (err_flags & kErrProposal)
        return fail(ctx, K8SGPU_E_ARG, "place: proposal buffer holds values outside [-1, n): was it written on "
                                       "another stream than the one passed here?");
    if (err_flags != 0u)
        return fail(ctx, K8SGPU_E_ARG, "place: %s table outside the input domain (see k8sgpu.h)",
                    (err_flags & kErrNode) ? "node" : "pod");
    if (next_count >= na)
        return fail(ctx, K8SGPU_E_CUDA, "place: round %u made no progress (%u pods still active)", st.rounds, na);
    ctx->h_pinned[1] = next_count;
    if (want_prof) {  // developer aid: phase boundaries of this round's admission, microseconds
        unsigned long long t[9];
        K8S_CUDA(ctx, cudaMemcpy(t, cnts + 16, sizeof t, cudaMemcpyDeviceToHost));
        fprintf(stderr, "[coop round %u na=%u grid=%u] keys+hist %.1f sort %.1f segreduce %.1f carry %.1f apply %.1f "
                        "free+flags %.1f scan %.1f compact+prep %.1f total %.1f us\n",
                st.rounds, na, grid, (t[1] - t[0]) / 1e3, (t[2] - t[1]) / 1e3, (t[3] - t[2]) / 1e3, (t[4] - t[3]) / 1e3,
                (t[5] - t[4]) / 1e3, (t[6] - t[5]) / 1e3, (t[7] - t[6]) / 1e3, (t[8] - t[7]) / 1e3, (t[8] - t[0]) / 1e3);
    }
    st.evals += (uint64_t)na * st.n;
    st.rounds++;
    st.cur ^= 1;
    st.n_active = ctx->h_pinned[1];
    st.nodeq_stale = false;  // the kernel rebuilt the packed node records from the new free columns
    st.best_cleared_len = na;  // ... and cleared best_key[0 .. na)
    return K8SGPU_OK;
}


int round_admit_coop(k8sgpu_ctx *ctx, cudaStream_t s, const int32_t *proposals) {
    int rc = round_admit_coop_launch(ctx, s, proposals, 1);
    if (rc) return rc;
    return round_admit_coop_finish(ctx, s);
}

// Decode the packed best keys of active positions [lo, hi) and store each proposal at its
// position in every buffer of `peer` (this GPU's own and the other ranks', mapped peer memory).
int round_finalize_peers(k8sgpu_ctx *ctx, uint32_t lo, uint32_t hi, int32_t *const *peer, int world, cudaStream_t s,
                         uint32_t xw, uint32_t xrank) {
    PlaceState &st = ctx->place;
    SlotMap sm;
    sm.lo = lo;
    sm.hi = hi;
    sm.xw = xw;
    sm.xrank = xrank;
    sm.xcnt = xw > 1 ? group_slot_count(hi, xw) : hi - lo;
    if (sm.xcnt == 0) return K8SGPU_OK;
    PeerPtrs pp;
    for (int r = 0; r < 16; ++r) pp.p[r] = r < world ? peer[r] : nullptr;
    pp.world = world;
    // where this launch's first slot sits in every exchange buffer
    const uint32_t base = xw > 1 ? xrank * sm.xcnt : lo;
    SlotView view;
    view.active = st.active[st.cur].as<uint32_t>();
    view.spod = st.order_valid ? st.order[st.cur].as<uint32_t>() : nullptr;
    view.posof = st.inv[0].as<uint32_t>();
    K8S_LAUNCH(ctx, score_finalize_p2p_kernel, ceil_div(sm.xcnt, 256), 256, 0, s, st.best_key.as<unsigned long long>(),
               st.podq.as<PodQ>(), view, sm, sm.xcnt, st.n, base, pp);
    return K8SGPU_OK;
}

int round_end(k8sgpu_ctx *ctx, cudaStream_t s) {
    PlaceState &st = ctx->place;
    uint32_t *cnts = st.counters.as<uint32_t>();
    K8S_CUDA(ctx, cudaMemsetAsync(cnts + 2, 0, 4, s));
    if (st.m)
        K8S_LAUNCH(ctx, count_placed_kernel, ceil_div(st.m, 256), 256, 0, s, st.assign.as<int32_t>(), st.m, cnts + 2);
    K8S_CUDA(ctx, cudaMemcpyAsync(ctx->h_pinned, cnts, 16, cudaMemcpyDeviceToHost, s));
    K8S_CUDA(ctx, cudaStreamSynchronize(s));
    if (ctx->h_pinned[0] != 0u)
        return fail(ctx, K8SGPU_E_ARG, "place: %s table outside the input domain (see k8sgpu.h)",
                    (ctx->h_pinned[0] & kErrNode) ? "node" : "pod");
    ctx->stats.rounds = st.rounds;
    ctx->stats.evals = st.evals;
    ctx->stats.placed = ctx->h_pinned[2];
    float ms = 0.f, tot = 0.f;
    for (int i = 0; i < st.timed; ++i) {
        if (cudaEventElapsedTime(&ms, ctx->ev_score[2 * i], ctx->ev_score[2 * i + 1]) == cudaSuccess) tot += ms;
    }
    ctx->stats.score_ms = tot;
    ctx->stats.score_evals = st.timed_evals;
    ctx->stats.score_launches = (uint32_t)st.timed;
    return K8SGPU_OK;
}

bool coop_admission(const k8sgpu_ctx *ctx) { return ctx->tune.coop < 0 ? true : ctx->tune.coop != 0; }

// Rounds of score + admit until no pod is active (or max_rounds is reached).
static int run_rounds(k8sgpu_ctx *ctx, cudaStream_t s) {
    PlaceState &st = ctx->place;
    // Admission as ONE cooperative launch per round (two launches per round in all): measured
    // ~3 % faster per config-3 step than the chain of ~20 small kernels.  K8SGPU_COOP=0 selects
    // the chain (the round-at-a-time API always uses it); both give identical results.
    const bool coop = coop_admission(ctx);
    int rc = K8SGPU_OK;
    while (st.n_active > 0 && (st.max_rounds == 0 || st.rounds < st.max_rounds)) {
        if (coop) {
            rc = round_score(ctx, 0, st.n_active, nullptr, s, /*prep=*/st.nodeq_stale, /*finalize=*/false);
            if (rc) return rc;
            rc = round_admit_coop(ctx, s, nullptr);  // leaves the next round's node records ready
        } else {
            rc = round_score(ctx, 0, st.n_active, st.best.as<int32_t>(), s);
            if (rc) return rc;
            rc = round_admit(ctx, st.best.as<int32_t>(), s, nullptr);
        }
        if (rc) return rc;
    }
    return K8SGPU_OK;
}

int place_run(k8sgpu_ctx *ctx, const k8sgpu_policy *policy) {
    cudaStream_t s = ctx->stream;
    ctx->place.want_order = true;
    K8S_CUDA(ctx, cudaEventRecord(ctx->ev_begin, s));
    int rc = round_begin(ctx, policy, s);
    if (rc) return rc;
    rc = run_rounds(ctx, s);
    if (rc) return rc;
    K8S_CUDA(ctx, cudaEventRecord(ctx->ev_end, s));
    rc = round_end(ctx, s);
    if (rc) return rc;
    float ms = 0.f;
    K8S_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev_begin, ctx->ev_end));
    ctx->stats.device_ms = ms;
    return K8SGPU_OK;
}

// ------------------------------------------------------------------ gang placement (f4)

__global__ void gang_count_kernel(const int32_t *__restrict__ assign, const uint32_t *__restrict__ gang_id, uint32_t m,
                                  uint32_t n_gangs, uint32_t *__restrict__ count, uint32_t *__restrict__ err) {
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= m) return;
    uint32_t g = gang_id[p];
    if (g == K8SGPU_NONE_U32) return;
    if (g >= n_gangs) {
        atomicOr(err, kErrPod);
        return;
    }
    if (assign[p] >= 0) atomicAdd(&count[g], 1u);  // counts only: order-free
}

// state[g] = rejected so far, state[n_gangs + g] = rejected in this iteration; counters[3] += new rejections
__global__ void gang_decide_kernel(const uint32_t *__restrict__ count, const uint32_t *__restrict__ min_available,
                                   uint32_t n_gangs, uint32_t *__restrict__ state, uint32_t *__restrict__ counters) {
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n_gangs) return;
    uint32_t newly = (!state[g] && count[g] < min_available[g]) ? 1u : 0u;
    state[n_gangs + g] = newly;
    if (newly) {
        state[g] = 1u;
        atomicAdd(&counters[3], 1u);
    }
}

// Pods of a gang rejected in this iteration give back what they reserved and are marked; pods
// still unplaced whose gang survives are flagged for the retry list.
__global__ void gang_release_kernel(int32_t *__restrict__ assign, const uint32_t *__restrict__ gang_id,
                                    const PodQ *__restrict__ podq, uint32_t m, uint32_t n_gangs, uint32_t n_nodes,
                                    const uint32_t *__restrict__ state, uint32_t *__restrict__ freec,
                                    uint32_t *__restrict__ retry_flags) {
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= m) return;
    const uint32_t g = gang_id[p];
    int32_t a = assign[p];
    if (g != K8SGPU_NONE_U32 && g < n_gangs && state[n_gangs + g]) {
        if (a >= 0) {  // commutative integer atomics: the result does not depend on order
            const PodQ *q = podq + p;
            if (q->req_gpu) atomicAdd(&freec[a], q->req_gpu);
            if (q->req_mcpu) atomicAdd(&freec[(size_t)n_nodes + a], q->req_mcpu);
            if (q->req_mib) atomicAdd(&freec[2 * (size_t)n_nodes + a], q->req_mib);
        }
        assign[p] = a = K8SGPU_ASSIGN_GANG;
    }
    const bool survives = g == K8SGPU_NONE_U32 || (g < n_gangs && !state[g]);
    const uint32_t retry = (a == K8SGPU_ASSIGN_NONE && survives) ? 1u : 0u;
    retry_flags[p] = retry;
    if (retry) assign[p] = K8SGPU_ASSIGN_PENDING;
}

__global__ void gang_retry_compact_kernel(const uint32_t *__restrict__ flags, const uint32_t *__restrict__ offs,
                                          uint32_t m, uint32_t *__restrict__ active) {
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p < m && flags[p]) active[offs[p]] = p;
}

__global__ void gang_unretry_kernel(int32_t *__restrict__ assign, const uint32_t *__restrict__ flags, uint32_t m) {
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p < m && flags[p]) assign[p] = K8SGPU_ASSIGN_NONE;
}

__global__ void gang_ok_kernel(const uint32_t *__restrict__ state, uint32_t n_gangs, int32_t *__restrict__ ok) {
    uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g < n_gangs) ok[g] = state[g] ? 0 : 1;
}

// Gang table to the device, verdict state cleared.
int gang_setup(k8sgpu_ctx *ctx, const k8sgpu_gang_table *gangs, cudaStream_t s) {
    PlaceState &st = ctx->place;
    if (!st.uploaded) return fail(ctx, K8SGPU_E_ARG, "place_gang: no tables uploaded");
    const uint32_t G = gangs->n_gangs, M = st.m;
    if (M && !gangs->gang_id) return fail(ctx, K8SGPU_E_ARG, "place_gang: gang_id is NULL");
    if (G && !gangs->min_available) return fail(ctx, K8SGPU_E_ARG, "place_gang: min_available is NULL");
    K8S_CUDA(ctx, st.gang_id.reserve((size_t)M * 4 + 16));
    K8S_CUDA(ctx, st.gang_min.reserve((size_t)G * 4 + 16));
    K8S_CUDA(ctx, st.gang_count.reserve((size_t)G * 4 + 16));
    K8S_CUDA(ctx, st.gang_state.reserve((size_t)G * 8 + 16));
    if (M) K8S_CUDA(ctx, cudaMemcpyAsync(st.gang_id.p, gangs->gang_id, (size_t)M * 4, cudaMemcpyHostToDevice, s));
    if (G) K8S_CUDA(ctx, cudaMemcpyAsync(st.gang_min.p, gangs->min_available, (size_t)G * 4, cudaMemcpyHostToDevice, s));
    ctx->stats.h2d_bytes += ((uint64_t)M + G) * 4;
    K8S_CUDA(ctx, cudaMemsetAsync(st.gang_state.p, 0, (size_t)G * 8 + 16, s));
    st.n_gangs = G;
    st.gang_iters = 1;
    return K8SGPU_OK;
}

// After a round loop that left no pod active: count placed members, reject gangs below
// min_available for good, release their reservations, and build the retry list (unplaced pods of
// surviving gangs, ascending pod index).  *n_retry == 0 means the placement is final.
int gang_verdicts(k8sgpu_ctx *ctx, cudaStream_t s, uint32_t *n_retry_out) {
    PlaceState &st = ctx->place;
    const uint32_t G = st.n_gangs, M = st.m;
    uint32_t *cnts = st.counters.as<uint32_t>();
    *n_retry_out = 0;
    if (M == 0) return K8SGPU_OK;
    K8S_CUDA(ctx, cudaMemsetAsync(st.gang_count.p, 0, (size_t)G * 4 + 16, s));
    K8S_CUDA(ctx, cudaMemsetAsync(cnts + 3, 0, 4, s));
    K8S_LAUNCH(ctx, gang_count_kernel, ceil_div(M, 256), 256, 0, s, st.assign.as<int32_t>(),
               st.gang_id.as<uint32_t>(), M, G, st.gang_count.as<uint32_t>(), cnts);
    if (G)
        K8S_LAUNCH(ctx, gang_decide_kernel, ceil_div(G, 256), 256, 0, s, st.gang_count.as<uint32_t>(),
                   st.gang_min.as<uint32_t>(), G, st.gang_state.as<uint32_t>(), cnts);
    uint32_t *flags = st.flags.as<uint32_t>(), *offs = flags + st.m;
    K8S_LAUNCH(ctx, gang_release_kernel, ceil_div(M, 256), 256, 0, s, st.assign.as<int32_t>(),
               st.gang_id.as<uint32_t>(), st.podq.as<PodQ>(), M, G, st.n, st.gang_state.as<uint32_t>(),
               st.free_cols.as<uint32_t>(), flags);
    int rc = exclusive_scan_u32(ctx, s, flags, offs, M, st.scan_tmp, cnts + 1);
    if (rc) return rc;
    K8S_LAUNCH(ctx, gang_retry_compact_kernel, ceil_div(M, 256), 256, 0, s, flags, offs, M,
               st.active[st.cur].as<uint32_t>());
    K8S_CUDA(ctx, cudaMemcpyAsync(ctx->h_pinned, cnts, 16, cudaMemcpyDeviceToHost, s));
    K8S_CUDA(ctx, cudaStreamSynchronize(s));
    if (ctx->h_pinned[0] != 0u)
        return fail(ctx, K8SGPU_E_ARG, "place_gang: a gang id is >= n_gangs, or a table is outside the input domain");
    const uint32_t n_new = ctx->h_pinned[3], n_retry = ctx->h_pinned[1];
    if (n_new == 0) {  // nothing was rejected: the flagged pods keep their verdict (-1)
        if (n_retry)   // they were marked PENDING for a retry that will not happen: restore -1
            K8S_LAUNCH(ctx, gang_unretry_kernel, ceil_div(M, 256), 256, 0, s, st.assign.as<int32_t>(), flags, M);
        return K8SGPU_OK;
    }
    st.n_active = n_retry;
    st.nodeq_stale = true;  // free columns changed under the packed records
    st.order_valid = false; // the retry list is a new active list: scored in plain order
    if (n_retry) st.gang_iters++;
    *n_retry_out = n_retry;
    return K8SGPU_OK;
}

static int place_gang_run(k8sgpu_ctx *ctx, const k8sgpu_gang_table *gangs, const k8sgpu_policy *policy) {
    PlaceState &st = ctx->place;
    cudaStream_t s = ctx->stream;
    int rc = gang_setup(ctx, gangs, s);
    if (rc) return rc;
    K8S_CUDA(ctx, cudaEventRecord(ctx->ev_begin, s));
    st.want_order = true;
    rc = round_begin(ctx, policy, s);
    if (rc) return rc;
    for (;;) {
        rc = run_rounds(ctx, s);  // step 1: policy v0 rounds over the current candidates
        if (rc) return rc;
        if (st.n_active > 0) break;  // max_rounds reached: no verdicts on half a placement
        uint32_t n_retry = 0;
        rc = gang_verdicts(ctx, s, &n_retry);  // steps 2 and 3
        if (rc) return rc;
        if (n_retry == 0) break;
    }
    K8S_CUDA(ctx, cudaEventRecord(ctx->ev_end, s));
    rc = round_end(ctx, s);
    if (rc) return rc;
    ctx->stats.gang_iterations = st.gang_iters;
    float ms = 0.f;
    K8S_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev_begin, ctx->ev_end));
    ctx->stats.device_ms = ms;
    return K8SGPU_OK;
}

// gang_ok[g] = 1 unless gang g was rejected; host buffer of n_gangs entries.
static int gang_ok_download(k8sgpu_ctx *ctx, cudaStream_t s, int32_t *gang_ok) {
    PlaceState &st = ctx->place;
    if (!gang_ok || !st.n_gangs) return K8SGPU_OK;
    int32_t *d_ok = reinterpret_cast<int32_t *>(st.gang_count.p);  // reuse: counts are no longer needed
    K8S_LAUNCH(ctx, gang_ok_kernel, ceil_div(st.n_gangs, 256), 256, 0, s, st.gang_state.as<uint32_t>(), st.n_gangs, d_ok);
    K8S_CUDA(ctx, cudaMemcpyAsync(gang_ok, d_ok, (size_t)st.n_gangs * 4, cudaMemcpyDeviceToHost, s));
    K8S_CUDA(ctx, cudaStreamSynchronize(s));
    ctx->stats.d2h_bytes += (uint64_t)st.n_gangs * 4;
    return K8SGPU_OK;
}

int place_download(k8sgpu_ctx *ctx, k8sgpu_place_out *out) {
    PlaceState &st = ctx->place;
    cudaStream_t s = ctx->stream;
    if (!out) return fail(ctx, K8SGPU_E_ARG, "place: NULL out");
    if (st.m && !out->assign) return fail(ctx, K8SGPU_E_ARG, "place: out->assign is NULL");
    if (st.m) {
        K8S_CUDA(ctx, cudaMemcpyAsync(out->assign, st.assign.p, (size_t)st.m * 4, cudaMemcpyDeviceToHost, s));
        ctx->stats.d2h_bytes += (uint64_t)st.m * 4;
    }
    uint32_t *dst[3] = {out->free_gpu, out->free_mcpu, out->free_mib};
    if (st.n && dst[0] && dst[1] == dst[0] + st.n && dst[2] == dst[1] + st.n) {  // one block: one copy
        K8S_CUDA(ctx, cudaMemcpyAsync(dst[0], st.free_cols.p, 3 * (size_t)st.n * 4, cudaMemcpyDeviceToHost, s));
        ctx->stats.d2h_bytes += 3 * (uint64_t)st.n * 4;
        dst[0] = dst[1] = dst[2] = nullptr;
    }
    for (int c = 0; c < 3; ++c)
        if (dst[c] && st.n) {
            K8S_CUDA(ctx, cudaMemcpyAsync(dst[c], st.free_cols.as<uint32_t>() + (size_t)c * st.n, (size_t)st.n * 4,
                                          cudaMemcpyDeviceToHost, s));
            ctx->stats.d2h_bytes += (uint64_t)st.n * 4;
        }
    K8S_CUDA(ctx, cudaStreamSynchronize(s));
    return K8SGPU_OK;
}

}  // namespace k8sgpu

using namespace k8sgpu;

extern "C" int k8sgpu_place_upload(k8sgpu_ctx *ctx, const k8sgpu_node_table *nodes, const k8sgpu_pod_table *pods) {
    K8S_ENTER(ctx);
    reset_call_stats(ctx);
    int rc = group_upload(ctx, nodes, pods);  // every GPU of the group gets a replica (one GPU: just this one)
    if (rc) return rc;
    return group_sync(ctx);
}

extern "C" int k8sgpu_place_run(k8sgpu_ctx *ctx, const k8sgpu_policy *policy) {
    K8S_ENTER(ctx);
    reset_call_stats(ctx);
    return ctx->world > 1 ? group_place_run(ctx, policy, nullptr) : place_run(ctx, policy);
}

extern "C" int k8sgpu_place_download(k8sgpu_ctx *ctx, k8sgpu_place_out *out) {
    K8S_ENTER(ctx);
    ctx->stats.d2h_bytes = 0;
    return place_download(ctx, out);
}

extern "C" int k8sgpu_place(k8sgpu_ctx *ctx, const k8sgpu_node_table *nodes, const k8sgpu_pod_table *pods,
                            const k8sgpu_policy *policy, k8sgpu_place_out *out) {
    K8S_ENTER(ctx);
    reset_call_stats(ctx);
    int rc = group_upload(ctx, nodes, pods);
    if (rc) return rc;
    rc = ctx->world > 1 ? group_place_run(ctx, policy, nullptr) : place_run(ctx, policy);
    if (rc) return rc;
    return place_download(ctx, out);  // the replicas are identical: the leader's copy is the result
}

extern "C" int k8sgpu_place_gang(k8sgpu_ctx *ctx, const k8sgpu_node_table *nodes, const k8sgpu_pod_table *pods,
                                 const k8sgpu_gang_table *gangs, const k8sgpu_policy *policy, k8sgpu_place_out *out,
                                 int32_t *gang_ok) {
    K8S_ENTER(ctx);
    reset_call_stats(ctx);
    if (!gangs) return fail(ctx, K8SGPU_E_ARG, "place_gang: NULL gang table");
    int rc = group_upload(ctx, nodes, pods);
    if (rc) return rc;
    rc = ctx->world > 1 ? group_place_run(ctx, policy, gangs) : place_gang_run(ctx, gangs, policy);
    if (rc) return rc;
    rc = gang_ok_download(ctx, ctx->stream, gang_ok);
    if (rc) return rc;
    return place_download(ctx, out);
}

// Gang placement one step at a time (one process per GPU): begin, then rounds through the
// k8sgpu_place_round_* entries until no pod is active, then verdicts; repeat while verdicts
// returns a retry list; end.  Every rank runs the verdict kernels on its own (identical) replica.
extern "C" int k8sgpu_place_gang_begin(k8sgpu_ctx *ctx, const k8sgpu_gang_table *gangs, const k8sgpu_policy *policy,
                                       void *stream, uint32_t *n_active) {
    K8S_ENTER(ctx);
    reset_call_stats(ctx);
    if (!gangs) return fail(ctx, K8SGPU_E_ARG, "place_gang: NULL gang table");
    cudaStream_t s = pick_stream(ctx, stream);
    int rc = gang_setup(ctx, gangs, s);
    if (rc) return rc;
    ctx->place.want_order = false;  // the caller exchanges proposals by active position
    rc = round_begin(ctx, policy, s);
    if (rc) return rc;
    if (n_active) *n_active = ctx->place.n_active;
    return K8SGPU_OK;
}

extern "C" int k8sgpu_place_gang_verdicts(k8sgpu_ctx *ctx, void *stream, uint32_t *n_active) {
    K8S_ENTER(ctx);
    if (!n_active) return fail(ctx, K8SGPU_E_ARG, "place_gang_verdicts: NULL n_active");
    if (ctx->place.n_active != 0) return fail(ctx, K8SGPU_E_ARG, "place_gang_verdicts: %u pods are still active", ctx->place.n_active);
    return gang_verdicts(ctx, pick_stream(ctx, stream), n_active);
}

extern "C" int k8sgpu_place_gang_end(k8sgpu_ctx *ctx, void *stream, int32_t *gang_ok) {
    K8S_ENTER(ctx);
    cudaStream_t s = pick_stream(ctx, stream);
    int rc = round_end(ctx, s);
    if (rc) return rc;
    ctx->stats.gang_iterations = ctx->place.gang_iters;
    return gang_ok_download(ctx, s, gang_ok);
}

extern "C" int k8sgpu_score_once(k8sgpu_ctx *ctx, const k8sgpu_policy *policy, int32_t *best_node) {
    K8S_ENTER(ctx);
    reset_call_stats(ctx);
    PlaceState &st = ctx->place;
    cudaStream_t s = ctx->stream;
    if (!best_node && st.m) return fail(ctx, K8SGPU_E_ARG, "score_once: NULL output");
    K8S_CUDA(ctx, cudaEventRecord(ctx->ev_begin, s));
    st.want_order = false;  // best_node[] is indexed by pod
    int rc = round_begin(ctx, policy, s);
    if (rc) return rc;
    rc = round_score(ctx, 0, st.n_active, st.best.as<int32_t>(), s);
    if (rc) return rc;
    K8S_CUDA(ctx, cudaEventRecord(ctx->ev_end, s));
    if (st.m) K8S_CUDA(ctx, cudaMemcpyAsync(best_node, st.best.p, (size_t)st.m * 4, cudaMemcpyDeviceToHost, s));
    ctx->stats.d2h_bytes = (uint64_t)st.m * 4;
    st.evals = (uint64_t)st.m * st.n;
    st.rounds = 1;
    rc = round_end(ctx, s);
    if (rc) return rc;
    float ms = 0.f;
    K8S_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev_begin, ctx->ev_end));
    ctx->stats.device_ms = ms;
    return K8SGPU_OK;
}

extern "C" int k8sgpu_place_round_begin(k8sgpu_ctx *ctx, const k8sgpu_policy *policy, void *stream,
                                        uint32_t *n_active) {
    K8S_ENTER(ctx);
    reset_call_stats(ctx);
    cudaStream_t s = ::k8sgpu::pick_stream(ctx, stream);
    ctx->place.want_order = false;  // the caller exchanges proposals by active position
    int rc = round_begin(ctx, policy, s);
    if (rc) return rc;
    if (n_active) *n_active = ctx->place.n_active;
    return K8SGPU_OK;
}

extern "C" int k8sgpu_place_round_score(k8sgpu_ctx *ctx, uint32_t lo, uint32_t hi, int32_t *d_best, void *stream) {
    K8S_ENTER(ctx);
    if (!d_best && hi > lo) return fail(ctx, K8SGPU_E_ARG, "round_score: NULL device buffer");
    return round_score(ctx, lo, hi, d_best, ::k8sgpu::pick_stream(ctx, stream), /*prep=*/ctx->place.nodeq_stale);
}

extern "C" int k8sgpu_place_round_admit(k8sgpu_ctx *ctx, const int32_t *d_best_all, void *stream,
                                        uint32_t *n_active) {
    K8S_ENTER(ctx);
    if (!d_best_all && ctx->place.n_active) return fail(ctx, K8SGPU_E_ARG, "round_admit: NULL device buffer");
    cudaStream_t s = ::k8sgpu::pick_stream(ctx, stream);
    if (!coop_admission(ctx) || ctx->place.n_active == 0) return round_admit(ctx, d_best_all, s, n_active);
    int rc = round_admit_coop(ctx, s, d_best_all);  // one launch; also rebuilds the packed node records
    if (rc == K8SGPU_OK && n_active) *n_active = ctx->place.n_active;
    return rc;
}

extern "C" int k8sgpu_place_round_end(k8sgpu_ctx *ctx, void *stream) {
    K8S_ENTER(ctx);
    return round_end(ctx, ::k8sgpu::pick_stream(ctx, stream));
}

// ------------------------------------------------------------------ peer exchange (one process per GPU)

// Unmaps every peer's exchange buffer.  Memory exported over CUDA IPC must not be freed while
// another process still maps it, so before a group re-exports (k8sgpu_peer_export frees the old
// buffer) every rank calls this and the ranks meet at a barrier of their own.
extern "C" int k8sgpu_peer_close(k8sgpu_ctx *ctx) {
    K8S_ENTER(ctx);
    PlaceState &st = ctx->place;
    for (int r = 0; r < 16; ++r) {
        if (r != st.peer_rank && st.peer_ptr[r]) cudaIpcCloseMemHandle(st.peer_ptr[r]);
        st.peer_ptr[r] = nullptr;
    }
    st.peer_world = 0;
    st.peer_rank = -1;
    return K8SGPU_OK;
}

extern "C" int k8sgpu_peer_export(k8sgpu_ctx *ctx, uint32_t m, void *handle64, void **d_local) {
    K8S_ENTER(ctx);
    if (!handle64) return fail(ctx, K8SGPU_E_ARG, "peer_export: NULL handle buffer");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is 64 bytes");
    PlaceState &st = ctx->place;
    k8sgpu_peer_close(ctx);  // (our own mappings; see k8sgpu_peer_close for the other direction)
    st.peer_local.release();  // a fresh allocation: its handle must match what peers will open
    K8S_CUDA(ctx, st.peer_local.reserve((size_t)(m ? m : 1u) * 2 * 4));  // two halves, used by round parity
    st.peer_m = m ? m : 1u;
    cudaIpcMemHandle_t h;
    K8S_CUDA(ctx, cudaIpcGetMemHandle(&h, st.peer_local.p));
    memcpy(handle64, &h, 64);
    if (d_local) *d_local = st.peer_local.p;
    return K8SGPU_OK;
}

extern "C" int k8sgpu_peer_import(k8sgpu_ctx *ctx, int rank, int world, const void *handles) {
    K8S_ENTER(ctx);
    PlaceState &st = ctx->place;
    if (world < 1 || world > 16 || rank < 0 || rank >= world || !handles)
        return fail(ctx, K8SGPU_E_ARG, "peer_import: bad rank/world (%d/%d; at most 16 ranks)", rank, world);
    if (!st.peer_local.p) return fail(ctx, K8SGPU_E_ARG, "peer_import: call k8sgpu_peer_export first");
    for (int r = 0; r < 16; ++r) st.peer_ptr[r] = nullptr;
    for (int r = 0; r < world; ++r) {
        if (r == rank) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, static_cast<const char *>(handles) + 64 * (size_t)r, 64);
        void *p = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {  // leave nothing half-mapped behind
            for (int q = 0; q < r; ++q) {
                if (st.peer_ptr[q]) cudaIpcCloseMemHandle(st.peer_ptr[q]);
                st.peer_ptr[q] = nullptr;
            }
            return fail(ctx, K8SGPU_E_CUDA, "peer_import: mapping rank %d's buffer: %s", r, cudaGetErrorString(e));
        }
        st.peer_ptr[r] = static_cast<int32_t *>(p);
    }
    st.peer_ptr[rank] = st.peer_local.as<int32_t>();
    st.peer_world = world;
    st.peer_rank = rank;
    return K8SGPU_OK;
}

extern "C" int k8sgpu_place_round_score_p2p(k8sgpu_ctx *ctx, uint32_t lo, uint32_t hi, void *stream) {
    K8S_ENTER(ctx);
    PlaceState &st = ctx->place;
    if (st.peer_world < 1) return fail(ctx, K8SGPU_E_ARG, "round_score_p2p: no peers imported");
    if (st.m > st.peer_m) return fail(ctx, K8SGPU_E_CAPACITY, "round_score_p2p: exchange buffer holds %u proposals, tables have %u pods", st.peer_m, st.m);
    cudaStream_t s = pick_stream(ctx, stream);
    int rc = round_score(ctx, lo, hi, nullptr, s, /*prep=*/st.nodeq_stale, /*finalize=*/false);
    if (rc) return rc;
    if (hi > lo) {
        // Rounds alternate between the two halves of every exchange buffer: a fast rank's stores of
        // round r+1 must not land in the half a slower rank is still reading for round r.
        const size_t half = (size_t)(st.rounds & 1u) * st.peer_m;
        int32_t *pp[16];
        for (int r = 0; r < 16; ++r) pp[r] = r < st.peer_world ? st.peer_ptr[r] + half : nullptr;
        rc = round_finalize_peers(ctx, lo, hi, pp, st.peer_world, s, 1, 0);
        if (rc) return rc;
    }
    return K8SGPU_OK;
}
