// group.cu -- one placement sharded over several B200s, behind the same C-ABI calls.
//
// Pods are sharded, node tables are replicated: every GPU of the group holds both
// tables and the same active list; per round each GPU scores only its contiguous slice of the
// active list, the slices' proposals are exchanged, and admission then runs on every GPU over the
// full proposal list (a13 is a pure prefix rule over global pod order, so the replicas stay
// bit-identical and the result cannot depend on the number of GPUs).
//
// A group is either
//   * the ndev devices one process passed to k8sgpu_create (the Go controller's case: one call,
//     eight GPUs; NCCL communicators from ncclCommInitAll, peer access enabled directly), or
//   * one process per GPU (torchrun), each with a one-device context joined to the others by
//     k8sgpu_comm_init_rank from a shared ncclUniqueId.
// Both run the same loop below; "local" ranks are the contexts this process drives.
//
// Exchange of a round's proposals (4 B per active pod), knob "exchange":
//   0  ncclAllGather, in place in the exchange buffer
//   1  peer stores: the kernel that decodes the scores writes each proposal into every rank's
//      exchange buffer over NVLink peer memory, followed by a 1-word ncclAllReduce as the barrier
//  -1  (default) peer stores when every rank can map every other rank's buffer, else ncclAllGather
// NCCL is loaded with dlopen when a group is first formed: single-GPU users never need it, and a
// process that already carries an NCCL (PyTorch's) shares that copy.
#include <dlfcn.h>
#include <nccl.h>

#include <mutex>
#include <new>

#include "common.cuh"

namespace k8sgpu {

struct NcclApi {
    ncclResult_t (*GetVersion)(int *);
    ncclResult_t (*GetUniqueId)(ncclUniqueId *);
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int);
    ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *);
    ncclResult_t (*CommDestroy)(ncclComm_t);
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t);
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t);
    ncclResult_t (*GroupStart)();
    ncclResult_t (*GroupEnd)();
    const char *(*GetErrorString)(ncclResult_t);
    void *handle;
    char err[256];
};

static NcclApi g_nccl = {};
static std::mutex g_nccl_mu;

static const NcclApi *nccl_api(k8sgpu_ctx *ctx) {
    std::lock_guard<std::mutex> lock(g_nccl_mu);
    if (g_nccl.handle) return &g_nccl;
    void *h = nullptr;
    for (const char *name : {"libnccl.so.2", "libnccl.so"}) {
        h = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
        if (h) break;
    }
    if (!h) {
        fail(ctx, K8SGPU_E_NCCL, "NCCL is not loadable (libnccl.so.2): %s", dlerror());
        return nullptr;
    }
    NcclApi a = {};
    bool ok = true;
#define K8S_SYM(field, sym) ok = ok && (*(void **)(&a.field) = dlsym(h, sym)) != nullptr
    K8S_SYM(GetVersion, "ncclGetVersion");
    K8S_SYM(GetUniqueId, "ncclGetUniqueId");
    K8S_SYM(CommInitRank, "ncclCommInitRank");
    K8S_SYM(CommInitAll, "ncclCommInitAll");
    K8S_SYM(CommDestroy, "ncclCommDestroy");
    K8S_SYM(AllGather, "ncclAllGather");
    K8S_SYM(AllReduce, "ncclAllReduce");
    K8S_SYM(GroupStart, "ncclGroupStart");
    K8S_SYM(GroupEnd, "ncclGroupEnd");
    K8S_SYM(GetErrorString, "ncclGetErrorString");
#undef K8S_SYM
    if (!ok) {
        fail(ctx, K8SGPU_E_NCCL, "libnccl lacks a required symbol");
        dlclose(h);
        return nullptr;
    }
    a.handle = h;
    g_nccl = a;
    return &g_nccl;
}

#define K8S_NCCL(ctx, api, expr)                                                                      \
    do {                                                                                              \
        ncclResult_t r__ = (expr);                                                                    \
        if (r__ != ncclSuccess)                                                                       \
            return ::k8sgpu::fail((ctx), K8SGPU_E_NCCL, "%s:%d %s: %s", __FILE__, __LINE__, #expr,    \
                                  (api)->GetErrorString(r__));                                        \
    } while (0)

#define K8S_DEV(ctx, c)                                                                               \
    do {                                                                                              \
        cudaError_t e__ = cudaSetDevice((c)->device);                                                 \
        if (e__ != cudaSuccess)                                                                       \
            return ::k8sgpu::fail((ctx), K8SGPU_E_CUDA, "cudaSetDevice(%d): %s", (c)->device,         \
                                  cudaGetErrorString(e__));                                           \
    } while (0)

// A failure on a follower context is reported on the leader (the one the caller holds).
static int lift(k8sgpu_ctx *lead, k8sgpu_ctx *c, int rc) {
    if (rc && c != lead) memcpy(lead->err, c->err, sizeof lead->err);
    return rc;
}

void group_destroy(k8sgpu_ctx *ctx) {
    const NcclApi *api = g_nccl.handle ? &g_nccl : nullptr;
    if (ctx->xpeer_ipc)
        for (int r = 0; r < ctx->world; ++r)
            if (r != ctx->rank && ctx->xpeer[r]) cudaIpcCloseMemHandle(ctx->xpeer[r]);
    if (ctx->comm && api) api->CommDestroy((ncclComm_t)ctx->comm);
    ctx->comm = nullptr;
    ctx->xchg.release();
    ctx->xsend.release();
    ctx->xsmall.release();
}

// Communicators for the devices of one process (k8sgpu_create with ndev > 1) and direct peer
// access between them.  ctx->local[] is already filled.
int group_init_local(k8sgpu_ctx *ctx) {
    const NcclApi *api = nccl_api(ctx);
    if (!api) return K8SGPU_E_NCCL;
    const int n = ctx->nlocal;
    int devs[16];
    ncclComm_t comms[16];
    for (int i = 0; i < n; ++i) devs[i] = ctx->local[i]->device;
    K8S_NCCL(ctx, api, api->CommInitAll(comms, n, devs));
    bool peers = true;
    for (int i = 0; i < n; ++i) {
        k8sgpu_ctx *c = ctx->local[i];
        c->comm = comms[i];
        c->world = n;
        c->rank = i;
        K8S_DEV(ctx, c);
        for (int j = 0; j < n; ++j) {
            if (j == i) continue;
            int can = 0;
            cudaDeviceCanAccessPeer(&can, devs[i], devs[j]);
            cudaError_t e = can ? cudaDeviceEnablePeerAccess(devs[j], 0) : cudaErrorPeerAccessUnsupported;
            if (e == cudaErrorPeerAccessAlreadyEnabled) {
                cudaGetLastError();
                e = cudaSuccess;
            }
            if (e != cudaSuccess) {
                cudaGetLastError();
                peers = false;
            }
        }
    }
    for (int i = 0; i < n; ++i) ctx->local[i]->xpeer_ready = peers;  // else: the NCCL exchange only
    return K8SGPU_OK;
}

static int group_barrier(k8sgpu_ctx *ctx, const NcclApi *api);

// Exchange buffers for m pods on every local rank; peer pointers for the peer-store exchange.
static int group_reserve(k8sgpu_ctx *ctx, uint32_t m) {
    const NcclApi *api = nccl_api(ctx);
    if (!api) return K8SGPU_E_NCCL;
    const uint32_t world = (uint32_t)ctx->world;
    // >= the exchanged array of any round: world runs of whole 2048-slot blocks (see SlotMap)
    const uint32_t half = group_slot_count(m ? m : 1u, world) * world + world;
    bool grew = false;
    if (ctx->xpeer_ipc && ctx->xchg_m < half) {
        // The exchange buffers are about to be replaced (every rank takes this branch in the same
        // call: same tables).  Memory exported over CUDA IPC must not be freed while another
        // process still maps it: close our mappings, meet the other ranks, only then free.
        for (int r = 0; r < ctx->world; ++r)
            if (r != ctx->rank && ctx->xpeer[r]) cudaIpcCloseMemHandle(ctx->xpeer[r]);
        ctx->xpeer_ipc = false;
        ctx->xpeer_ready = false;
        int rc = group_barrier(ctx, api);
        if (rc) return rc;
        K8S_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    for (int i = 0; i < ctx->nlocal; ++i) {
        k8sgpu_ctx *c = ctx->local[i];
        K8S_DEV(ctx, c);
        if (c->xchg_m < half || !c->xchg.p) {
            c->xchg.release();
            K8S_CUDA(ctx, c->xchg.reserve((size_t)half * 2 * 4));
            c->xchg_m = half;
            grew = true;
        }
        K8S_CUDA(ctx, c->xsmall.reserve(64 * 16 + 256));
        K8S_CUDA(ctx, cudaMemsetAsync(c->xsmall.p, 0, 64, c->stream));
    }
    if (ctx->nlocal > 1) {  // one process: the other devices' buffers are directly addressable
        for (int i = 0; i < ctx->nlocal; ++i)
            for (int j = 0; j < ctx->nlocal; ++j) ctx->local[i]->xpeer[j] = ctx->local[j]->xchg.as<int32_t>();
    } else if (ctx->world > 1 && ctx->tune.exchange != 0 && !ctx->xpeer_failed && (grew || !ctx->xpeer_ready)) {
        // one process per GPU: swap CUDA IPC handles of the exchange buffers through an all-gather
        cudaIpcMemHandle_t mine, all[16];
        char *d = ctx->xsmall.as<char>() + 256;
        K8S_CUDA(ctx, cudaIpcGetMemHandle(&mine, ctx->xchg.p));
        K8S_CUDA(ctx, cudaMemcpyAsync(d + 64 * ctx->rank, &mine, 64, cudaMemcpyHostToDevice, ctx->stream));
        K8S_NCCL(ctx, api, api->AllGather(d + 64 * ctx->rank, d, 64, ncclUint8, (ncclComm_t)ctx->comm, ctx->stream));
        K8S_CUDA(ctx, cudaMemcpyAsync(all, d, 64 * (size_t)ctx->world, cudaMemcpyDeviceToHost, ctx->stream));
        K8S_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        int opened_ok = 1;
        for (int r = 0; r < ctx->world; ++r) ctx->xpeer[r] = nullptr;
        for (int r = 0; r < ctx->world && opened_ok; ++r) {
            if (r == ctx->rank) {
                ctx->xpeer[r] = ctx->xchg.as<int32_t>();
                continue;
            }
            void *p = nullptr;
            if (cudaIpcOpenMemHandle(&p, all[r], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                cudaGetLastError();
                opened_ok = 0;
            } else {
                ctx->xpeer[r] = static_cast<int32_t *>(p);
            }
        }
        // every rank must take the same exchange: the mapping counts only if it worked everywhere
        int32_t *flag = ctx->xsmall.as<int32_t>() + 8;
        K8S_CUDA(ctx, cudaMemcpyAsync(flag, &opened_ok, 4, cudaMemcpyHostToDevice, ctx->stream));
        K8S_NCCL(ctx, api, api->AllReduce(flag, flag, 1, ncclInt32, ncclMin, (ncclComm_t)ctx->comm, ctx->stream));
        int all_ok = 0;
        K8S_CUDA(ctx, cudaMemcpyAsync(&all_ok, flag, 4, cudaMemcpyDeviceToHost, ctx->stream));
        K8S_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        if (!all_ok) {  // no peer mapping between some pair of these GPUs: NCCL carries every round
            for (int r = 0; r < ctx->world; ++r)
                if (r != ctx->rank && ctx->xpeer[r]) cudaIpcCloseMemHandle(ctx->xpeer[r]);
            ctx->xpeer_failed = true;
            ctx->xpeer_ready = false;
            return K8SGPU_OK;
        }
        ctx->xpeer_ipc = true;
        ctx->xpeer_ready = true;
    }
    return K8SGPU_OK;
}

// All local ranks meet all other ranks: nothing queued after this on any rank's stream starts
// before everything queued before it on every rank's stream has finished.
static int group_barrier(k8sgpu_ctx *ctx, const NcclApi *api) {
    K8S_NCCL(ctx, api, api->GroupStart());
    for (int i = 0; i < ctx->nlocal; ++i) {
        k8sgpu_ctx *c = ctx->local[i];
        ncclResult_t r = api->AllReduce(c->xsmall.p, c->xsmall.p, 1, ncclInt32, ncclSum, (ncclComm_t)c->comm, c->stream);
        if (r != ncclSuccess) {
            api->GroupEnd();
            return fail(ctx, K8SGPU_E_NCCL, "ncclAllReduce (barrier): %s", api->GetErrorString(r));
        }
    }
    K8S_NCCL(ctx, api, api->GroupEnd());
    return K8SGPU_OK;
}

// Rounds of score / exchange / admit on every local rank until no pod is active.
static int group_rounds(k8sgpu_ctx *ctx, const NcclApi *api) {
    PlaceState &lead = ctx->place;
    const uint32_t world = (uint32_t)ctx->world;
    const bool coop = coop_admission(ctx);
    while (lead.n_active > 0 && (lead.max_rounds == 0 || lead.rounds < lead.max_rounds)) {
        const uint32_t na = lead.n_active;
        // How the round's scoring slots are shared out.  With the one-launch admission: blocks of
        // 2048 slots dealt round-robin (the slots are sorted by pod class, dearest classes first, so
        // contiguous shares would be unequal work); each rank's proposals are then one contiguous
        // run of `slice` entries in the exchanged array.  With the kernel chain (a test path, no
        // class order): contiguous shares of the active list.
        const uint32_t xw = coop ? world : 1u;
        const uint32_t slice = coop ? group_slot_count(na, world) : ceil_div(na, world);
        // Which exchange this round (the same on every rank).  Forced by the "exchange" knob, else
        // peer stores wherever every rank could map every other rank's buffer -- the kernel that
        // decodes the scores does the exchange, and the only collective left is a 1-word barrier:
        // measured 75 us instead of 220 us per short round on 2 GPUs, and equal to ncclAllGather
        // at 10^6 proposals (profiles/r02_group_ranks_x2_check.log) -- and ncclAllGather otherwise.
        const bool p2p = ctx->xpeer_ready && ctx->tune.exchange != 0;
        // Rounds alternate between the two halves of the exchange buffer, whichever exchange they
        // use: a fast rank's peer stores of round r+1 must not land where a slower rank is still
        // reading round r.
        const size_t half = (size_t)(lead.rounds & 1u) * ctx->xchg_m;
        for (int i = 0; i < ctx->nlocal; ++i) {
            k8sgpu_ctx *c = ctx->local[i];
            K8S_DEV(ctx, c);
            const uint32_t rank = (uint32_t)c->rank;
            uint32_t lo = 0, hi = na;
            if (!coop) {
                lo = rank * slice < na ? rank * slice : na;
                hi = lo + slice < na ? lo + slice : na;
            }
            int rc;
            if (p2p) {
                rc = round_score(c, lo, hi, nullptr, c->stream, c->place.nodeq_stale, /*finalize=*/false, xw, rank);
                if (!rc) {
                    int32_t *pp[16];
                    for (uint32_t r = 0; r < 16; ++r) pp[r] = r < world ? c->xpeer[r] + half : nullptr;
                    rc = round_finalize_peers(c, lo, hi, pp, (int)world, c->stream, xw, rank);
                }
            } else {  // this rank's run lands where the in-place all-gather expects it
                rc = round_score(c, lo, hi, c->xchg.as<int32_t>() + half + (size_t)rank * slice, c->stream,
                                 c->place.nodeq_stale, true, xw, rank);
            }
            if (rc) return lift(ctx, c, rc);
        }
        if (p2p) {
            int rc = group_barrier(ctx, api);
            if (rc) return rc;
        } else {
            K8S_NCCL(ctx, api, api->GroupStart());
            for (int i = 0; i < ctx->nlocal; ++i) {
                k8sgpu_ctx *c = ctx->local[i];
                int32_t *buf = c->xchg.as<int32_t>() + half;
                ncclResult_t r = api->AllGather(buf + (size_t)c->rank * slice, buf, slice, ncclInt32, (ncclComm_t)c->comm, c->stream);
                if (r != ncclSuccess) {
                    api->GroupEnd();
                    return fail(ctx, K8SGPU_E_NCCL, "ncclAllGather: %s", api->GetErrorString(r));
                }
            }
            K8S_NCCL(ctx, api, api->GroupEnd());
        }
        // admission on every replica: queue all of them, then wait for each
        for (int i = 0; i < ctx->nlocal; ++i) {
            k8sgpu_ctx *c = ctx->local[i];
            K8S_DEV(ctx, c);
            const int32_t *prop = c->xchg.as<int32_t>() + half;
            int rc = coop ? round_admit_coop_launch(c, c->stream, prop, xw) : round_admit(c, prop, c->stream, nullptr);
            if (rc) return lift(ctx, c, rc);
        }
        for (int i = 0; i < ctx->nlocal; ++i) {
            k8sgpu_ctx *c = ctx->local[i];
            K8S_DEV(ctx, c);
            if (coop) {
                int rc = round_admit_coop_finish(c, c->stream);
                if (rc) return lift(ctx, c, rc);
            }
            if (c->place.n_active != lead.n_active)
                return fail(ctx, K8SGPU_E_CUDA, "replicas diverged in round %u: device %d has %u active pods, device %d has %u",
                            lead.rounds, c->device, c->place.n_active, ctx->device, lead.n_active);
        }
    }
    return K8SGPU_OK;
}

int gang_setup(k8sgpu_ctx *ctx, const k8sgpu_gang_table *gangs, cudaStream_t s);
int gang_verdicts(k8sgpu_ctx *ctx, cudaStream_t s, uint32_t *n_retry_out);

// One whole placement (optionally with gangs) from the uploaded tables, on every rank of the group.
int group_place_run(k8sgpu_ctx *ctx, const k8sgpu_policy *policy, const k8sgpu_gang_table *gangs) {
    const NcclApi *api = nccl_api(ctx);
    if (!api) return K8SGPU_E_NCCL;
    if (!ctx->comm) return fail(ctx, K8SGPU_E_NCCL, "placement group has no communicator");
    int rc = group_reserve(ctx, ctx->place.m);
    if (rc) return rc;
    K8S_DEV(ctx, ctx);
    K8S_CUDA(ctx, cudaEventRecord(ctx->ev_begin, ctx->stream));
    if (ctx->tune.exchange != 0 && ctx->xpeer_ready) {
        // no rank may store into exchange buffers a slower rank is still reading from the previous run
        rc = group_barrier(ctx, api);
        if (rc) return rc;
    }
    for (int i = 0; i < ctx->nlocal; ++i) {
        k8sgpu_ctx *c = ctx->local[i];
        K8S_DEV(ctx, c);
        if (c != ctx) reset_call_stats(c);
        if (gangs) rc = gang_setup(c, gangs, c->stream);
        c->place.want_order = true;  // slots of the class-sorted scoring order are what the ranks share out
        if (!rc) rc = round_begin(c, policy, c->stream);
        if (rc) return lift(ctx, c, rc);
    }
    for (;;) {
        rc = group_rounds(ctx, api);
        if (rc) return rc;
        if (!gangs || ctx->place.n_active > 0) break;  // (max_rounds reached: no verdicts on half a placement)
        uint32_t n_retry = 0;
        for (int i = 0; i < ctx->nlocal; ++i) {  // deterministic on every replica: no collective needed
            k8sgpu_ctx *c = ctx->local[i];
            K8S_DEV(ctx, c);
            uint32_t nr = 0;
            rc = gang_verdicts(c, c->stream, &nr);
            if (rc) return lift(ctx, c, rc);
            if (i == 0) n_retry = nr;
            else if (nr != n_retry) return fail(ctx, K8SGPU_E_CUDA, "replicas diverged in the gang verdicts");
        }
        if (n_retry == 0) break;
    }
    K8S_DEV(ctx, ctx);
    K8S_CUDA(ctx, cudaEventRecord(ctx->ev_end, ctx->stream));
    for (int i = ctx->nlocal - 1; i >= 0; --i) {  // the leader last: its device stays current
        k8sgpu_ctx *c = ctx->local[i];
        K8S_DEV(ctx, c);
        rc = round_end(c, c->stream);
        if (rc) return lift(ctx, c, rc);
        if (c != ctx) ctx->stats.kernel_launches += c->stats.kernel_launches;
    }
    if (gangs) ctx->stats.gang_iterations = ctx->place.gang_iters;
    float ms = 0.f;
    K8S_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev_begin, ctx->ev_end));
    ctx->stats.device_ms = ms;
    return K8SGPU_OK;
}

// Tables to every GPU of the group.  The node table (32 B x N) goes to each GPU from the host; of
// the pod table each rank copies only its 1/world of the rows over PCIe, and the ranks complete
// each other's copies with one ncclAllGather per column over NVLink.
int group_upload(k8sgpu_ctx *ctx, const k8sgpu_node_table *nodes, const k8sgpu_pod_table *pods) {
    if (ctx->world == 1 || !pods || pods->m < 4096u * (uint32_t)ctx->world || ctx->tune.shard_upload == 0) {
        for (int i = 0; i < ctx->nlocal; ++i) {  // queued on every device's own stream: the copies overlap
            k8sgpu_ctx *c = ctx->local[i];
            K8S_DEV(ctx, c);
            if (c != ctx) reset_call_stats(c);
            int rc = place_upload(c, nodes, pods);
            if (rc) return lift(ctx, c, rc);
            if (c != ctx) ctx->stats.h2d_bytes += c->stats.h2d_bytes;
        }
        K8S_DEV(ctx, ctx);
        return K8SGPU_OK;
    }
    const NcclApi *api = nccl_api(ctx);
    if (!api) return K8SGPU_E_NCCL;
    const uint32_t world = (uint32_t)ctx->world, m = pods->m, slice = ceil_div(m, world), stride = slice * world;
    const int npc = pods->priority ? 8 : 7;
    for (int i = 0; i < ctx->nlocal; ++i) {
        k8sgpu_ctx *c = ctx->local[i];
        K8S_DEV(ctx, c);
        if (c != ctx) reset_call_stats(c);
        K8S_CUDA(ctx, c->xsend.reserve((size_t)8 * stride * 4 + 64));
        PodShare sh;
        sh.lo = (uint32_t)c->rank * slice < m ? (uint32_t)c->rank * slice : m;
        sh.hi = sh.lo + slice < m ? sh.lo + slice : m;
        sh.stride = stride;
        sh.staging = c->xsend.as<uint32_t>();
        int rc = place_upload(c, nodes, pods, &sh);
        if (rc) return lift(ctx, c, rc);
        if (c != ctx) ctx->stats.h2d_bytes += c->stats.h2d_bytes;
    }
    K8S_NCCL(ctx, api, api->GroupStart());
    for (int i = 0; i < ctx->nlocal; ++i) {
        k8sgpu_ctx *c = ctx->local[i];
        for (int col = 0; col < npc; ++col) {
            uint32_t *base = c->xsend.as<uint32_t>() + (size_t)col * stride;
            ncclResult_t r = api->AllGather(base + (size_t)c->rank * slice, base, slice, ncclUint32, (ncclComm_t)c->comm, c->stream);
            if (r != ncclSuccess) {
                api->GroupEnd();
                return fail(ctx, K8SGPU_E_NCCL, "ncclAllGather (pod table): %s", api->GetErrorString(r));
            }
        }
    }
    K8S_NCCL(ctx, api, api->GroupEnd());
    for (int i = 0; i < ctx->nlocal; ++i) {
        k8sgpu_ctx *c = ctx->local[i];
        K8S_DEV(ctx, c);
        for (int col = 0; col < npc; ++col)
            K8S_CUDA(ctx, cudaMemcpyAsync(c->place.pod_cols.as<uint32_t>() + (size_t)col * m,
                                          c->xsend.as<uint32_t>() + (size_t)col * stride, (size_t)m * 4,
                                          cudaMemcpyDeviceToDevice, c->stream));
    }
    K8S_DEV(ctx, ctx);
    return K8SGPU_OK;
}

int group_sync(k8sgpu_ctx *ctx) {
    for (int i = ctx->nlocal - 1; i >= 0; --i) {
        k8sgpu_ctx *c = ctx->local[i];
        K8S_DEV(ctx, c);
        K8S_CUDA(ctx, cudaStreamSynchronize(c->stream));
    }
    return K8SGPU_OK;
}

}  // namespace k8sgpu

using namespace k8sgpu;

extern "C" int k8sgpu_comm_unique_id(void *id128) {
    if (!id128) return K8SGPU_E_ARG;
    const NcclApi *api = nccl_api(nullptr);
    if (!api) return K8SGPU_E_NCCL;
    static_assert(sizeof(ncclUniqueId) == K8SGPU_UNIQUE_ID_BYTES, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    if (api->GetUniqueId(&id) != ncclSuccess) return K8SGPU_E_NCCL;
    memcpy(id128, &id, sizeof id);
    return K8SGPU_OK;
}

extern "C" int k8sgpu_comm_init_rank(k8sgpu_ctx *ctx, int rank, int world, const void *id128) {
    K8S_ENTER(ctx);
    if (!id128 || world < 1 || world > 16 || rank < 0 || rank >= world)
        return fail(ctx, K8SGPU_E_ARG, "comm_init_rank: bad rank/world (%d/%d; at most 16 ranks)", rank, world);
    if (ctx->nlocal != 1 || ctx->comm)
        return fail(ctx, K8SGPU_E_ARG, "comm_init_rank: this context already belongs to a placement group");
    const NcclApi *api = nccl_api(ctx);
    if (!api) return K8SGPU_E_NCCL;
    ncclUniqueId id;
    memcpy(&id, id128, sizeof id);
    ncclComm_t comm = nullptr;
    K8S_NCCL(ctx, api, api->CommInitRank(&comm, world, id, rank));
    ctx->comm = comm;
    ctx->world = world;
    ctx->rank = rank;
    return K8SGPU_OK;
}

extern "C" int k8sgpu_group_info(const k8sgpu_ctx *ctx, int *world, int *rank, int *nlocal, int *nccl_version) {
    if (!ctx) return K8SGPU_E_ARG;
    if (world) *world = ctx->world;
    if (rank) *rank = ctx->rank;
    if (nlocal) *nlocal = ctx->nlocal;
    if (nccl_version) {
        *nccl_version = 0;
        if (g_nccl.handle) g_nccl.GetVersion(nccl_version);
    }
    return K8SGPU_OK;
}
