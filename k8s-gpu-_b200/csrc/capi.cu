// capi.cu -- context lifetime, errors, statistics and the primitive test hooks of libk8sgpu.so.
#include <cstdlib>
#include <mutex>
#include <new>

#include "common.cuh"

using namespace k8sgpu;

static char g_create_err[512] = "";

extern "C" int k8sgpu_abi_version(void) { return K8SGPU_ABI_VERSION; }

// Developer / test knobs, by name.  The K8SGPU_<NAME> environment variables seed them once, when
// the context is created; after that only k8sgpu_set_tuning changes them.
static int *tuning_slot(Tuning &t, const char *key) {
    struct {
        const char *name;
        int *slot;
    } table[] = {{"res", &t.res}, {"ppt", &t.ppt}, {"splits", &t.splits}, {"coop", &t.coop},
                 {"coop_grid", &t.coop_grid}, {"rec_path", &t.rec_path}, {"rec_grid", &t.rec_grid},
                 {"rec_segments", &t.rec_segments}, {"coop_prof", &t.coop_prof}, {"rec_prof", &t.rec_prof},
                 {"shard_admit", &t.shard_admit}, {"exchange", &t.exchange}, {"force_cls", &t.force_cls}, {"order", &t.order}, {"shard_upload", &t.shard_upload}, {"slot_admit", &t.slot_admit}};
    for (auto &e : table)
        if (strcmp(e.name, key) == 0) return e.slot;
    return nullptr;
}

static void tuning_from_env(Tuning &t) {
    static const char *names[] = {"res", "ppt", "splits", "coop", "coop_grid", "rec_path", "rec_grid",
                                  "rec_segments", "coop_prof", "rec_prof", "shard_admit", "exchange", "force_cls", "order", "shard_upload", "slot_admit"};
    for (const char *n : names) {
        char env[64] = "K8SGPU_";
        size_t k = strlen(env);
        for (const char *c = n; *c && k + 1 < sizeof env; ++c) env[k++] = (char)(*c >= 'a' && *c <= 'z' ? *c - 32 : *c);
        env[k] = 0;
        if (const char *v = getenv(env)) *tuning_slot(t, n) = atoi(v);
    }
}

extern "C" int k8sgpu_set_tuning(k8sgpu_ctx *ctx, const char *key, int value) {
    if (!ctx || !key) return K8SGPU_E_ARG;
    int *slot = tuning_slot(ctx->tune, key);
    if (!slot) return fail(ctx, K8SGPU_E_ARG, "set_tuning: unknown knob '%s'", key);
    *slot = value;
    for (int i = 1; i < ctx->nlocal; ++i) *tuning_slot(ctx->local[i]->tune, key) = value;
    return K8SGPU_OK;
}

// With ctx == NULL: the message of the most recent failed k8sgpu_create in this process (a
// process-wide buffer, written under a lock; concurrent creators should use k8sgpu_create_ex).
extern "C" const char *k8sgpu_last_error(const k8sgpu_ctx *ctx) { return ctx ? ctx->err : g_create_err; }

static std::mutex g_create_mu;  // guards g_create_err (messages of failed k8sgpu_create calls)

static void create_err(char *errbuf, size_t errcap, const char *fmt, ...) {
    char msg[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(msg, sizeof msg, fmt, ap);
    va_end(ap);
    if (errbuf && errcap) snprintf(errbuf, errcap, "%s", msg);
    std::lock_guard<std::mutex> lock(g_create_mu);
    snprintf(g_create_err, sizeof g_create_err, "%s", msg);
}

static int create_one(int device, k8sgpu_ctx **out, char *errbuf, size_t errcap) {
    cudaError_t e;
    k8sgpu_ctx *ctx = new (std::nothrow) k8sgpu_ctx();
    if (!ctx) return K8SGPU_E_CUDA;
    ctx->device = device;
    ctx->local[0] = ctx;
#define CREATE_TRY(expr)                                                                   \
    if ((e = (expr)) != cudaSuccess) {                                                     \
        create_err(errbuf, errcap, "k8sgpu_create: %s: %s", #expr, cudaGetErrorString(e)); \
        k8sgpu_destroy(ctx);                                                               \
        return K8SGPU_E_CUDA;                                                              \
    }
    CREATE_TRY(cudaSetDevice(ctx->device));
    cudaDeviceProp prop;
    CREATE_TRY(cudaGetDeviceProperties(&prop, ctx->device));
    if (prop.major < 10) {
        create_err(errbuf, errcap, "k8sgpu_create: device %d is sm_%d%d; this library is sm_100a only", ctx->device,
                   prop.major, prop.minor);
        k8sgpu_destroy(ctx);
        return K8SGPU_E_CUDA;
    }
    ctx->sm_count = prop.multiProcessorCount;
    CREATE_TRY(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    CREATE_TRY(cudaEventCreate(&ctx->ev_begin));
    CREATE_TRY(cudaEventCreate(&ctx->ev_end));
    for (int i = 0; i < 2 * kMaxTimedRounds; ++i) CREATE_TRY(cudaEventCreate(&ctx->ev_score[i]));
    CREATE_TRY(cudaHostAlloc((void **)&ctx->h_pinned, 64 * sizeof(uint32_t), cudaHostAllocMapped));
    memset(ctx->h_pinned, 0, 64 * sizeof(uint32_t));
    CREATE_TRY(cudaHostGetDevicePointer((void **)&ctx->d_mailbox, ctx->h_pinned + 32, 0));
#undef CREATE_TRY
    tuning_from_env(ctx->tune);  // the only place the environment is read
    *out = ctx;
    return K8SGPU_OK;
}

extern "C" int k8sgpu_create_ex(const int *devices, int ndev, k8sgpu_ctx **out, char *errbuf, size_t errcap) {
    if (!out) return K8SGPU_E_ARG;
    *out = nullptr;
    if (errbuf && errcap) errbuf[0] = 0;
    if (ndev < 1 || ndev > 16 || !devices) {
        create_err(errbuf, errcap, "k8sgpu_create: ndev must be 1..16 (got %d)", ndev);
        return K8SGPU_E_ARG;
    }
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count <= 0) {  // no CPU fallback: this is an error, not a mode
        create_err(errbuf, errcap, "k8sgpu_create: no CUDA device (%s)",
                   e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
        return K8SGPU_E_CUDA;
    }
    for (int i = 0; i < ndev; ++i) {
        if (devices[i] < 0 || devices[i] >= count) {
            create_err(errbuf, errcap, "k8sgpu_create: device %d out of range [0,%d)", devices[i], count);
            return K8SGPU_E_ARG;
        }
        for (int j = 0; j < i; ++j)
            if (devices[j] == devices[i]) {
                create_err(errbuf, errcap, "k8sgpu_create: device %d listed twice", devices[i]);
                return K8SGPU_E_ARG;
            }
    }
    k8sgpu_ctx *lead = nullptr;
    int rc = create_one(devices[0], &lead, errbuf, errcap);
    if (rc) return rc;
    for (int i = 1; i < ndev; ++i) {
        k8sgpu_ctx *c = nullptr;
        rc = create_one(devices[i], &c, errbuf, errcap);
        if (rc) {
            k8sgpu_destroy(lead);
            return rc;
        }
        lead->local[lead->nlocal++] = c;
    }
    if (ndev > 1) {  // one process drives all the GPUs: communicators and peer access now
        rc = group_init_local(lead);
        if (rc) {
            create_err(errbuf, errcap, "%s", lead->err);
            k8sgpu_destroy(lead);
            return rc;
        }
        cudaSetDevice(lead->device);
    }
    *out = lead;
    return K8SGPU_OK;
}

extern "C" int k8sgpu_create(const int *devices, int ndev, k8sgpu_ctx **out) {
    return k8sgpu_create_ex(devices, ndev, out, nullptr, 0);
}

static void release_sort(SortScratch &s) {
    s.keys_alt.release();
    s.vals_alt.release();
    s.hist.release();
    s.scan_tmp.release();
}

extern "C" int k8sgpu_destroy(k8sgpu_ctx *ctx) {
    if (!ctx) return K8SGPU_OK;
    for (int i = ctx->nlocal - 1; i >= 1; --i) {  // the other GPUs of an in-process group first
        k8sgpu_ctx *c = ctx->local[i];
        ctx->local[i] = nullptr;
        if (c) k8sgpu_destroy(c);
    }
    ctx->nlocal = 1;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    group_destroy(ctx);
    PlaceState &p = ctx->place;
    DevBuf *pb[] = {&p.node_cols, &p.pod_cols, &p.free_cols, &p.nodeq,     &p.podq,      &p.assign, &p.active[0],
                    &p.active[1], &p.order[0], &p.order[1], &p.inv[0], &p.inv[1], &p.slots, &p.best_key, &p.best,      &p.sort_keys, &p.sort_vals, &p.seg,    &p.flags,
                    &p.counters,  &p.scan_tmp,  &p.gang_id,  &p.gang_min,  &p.gang_count, &p.gang_state};
    for (int r = 0; r < p.peer_world; ++r)
        if (r != p.peer_rank && p.peer_ptr[r]) cudaIpcCloseMemHandle(p.peer_ptr[r]);
    p.peer_local.release();
    for (DevBuf *b : pb) b->release();
    release_sort(p.sort);
    ReconcileState &r = ctx->rec;
    DevBuf *rb[] = {&r.in_slab, &r.out_slab,  &r.lut,       &r.pool_tmp, &r.tot,     &r.bitmap,
                    &r.sort_keys, &r.sort_vals, &r.cta_cnt, &r.counters, &r.scan_tmp};
    for (DevBuf *b : rb) b->release();
    release_sort(r.sort);
    for (DevBuf &b : ctx->dbg) b.release();
    release_sort(ctx->dbg_sort);
    if (ctx->h_pinned) cudaFreeHost(ctx->h_pinned);
    for (int i = 0; i < 2 * kMaxTimedRounds; ++i)
        if (ctx->ev_score[i]) cudaEventDestroy(ctx->ev_score[i]);
    if (ctx->ev_begin) cudaEventDestroy(ctx->ev_begin);
    if (ctx->ev_end) cudaEventDestroy(ctx->ev_end);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
    return K8SGPU_OK;
}

// Page-locked host memory for the caller's tables and results: a cgo or ctypes caller gets the
// asynchronous, full-rate PCIe copies of pinned buffers without binding the CUDA runtime itself.
extern "C" int k8sgpu_host_alloc(size_t bytes, void **out) {
    if (!out) return K8SGPU_E_ARG;
    *out = nullptr;
    cudaError_t e = cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocDefault);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return K8SGPU_E_CUDA;
    }
    return K8SGPU_OK;
}

extern "C" int k8sgpu_host_free(void *p) {
    if (!p) return K8SGPU_OK;
    return cudaFreeHost(p) == cudaSuccess ? K8SGPU_OK : K8SGPU_E_CUDA;
}

extern "C" int k8sgpu_get_stats(const k8sgpu_ctx *ctx, k8sgpu_stats *out) {
    if (!ctx || !out) return K8SGPU_E_ARG;
    *out = ctx->stats;
    return K8SGPU_OK;
}

extern "C" int k8sgpu_debug_sort_pairs(k8sgpu_ctx *ctx, const uint32_t *keys, const uint32_t *vals, uint32_t n,
                                       int bits, uint32_t *out_keys, uint32_t *out_vals) {
    K8S_ENTER(ctx);
    reset_call_stats(ctx);
    if (n && (!keys || !vals || !out_keys || !out_vals)) return fail(ctx, K8SGPU_E_ARG, "sort_pairs: NULL buffer");
    if (bits < 0 || bits > 32) return fail(ctx, K8SGPU_E_ARG, "sort_pairs: bits=%d", bits);
    cudaStream_t s = ctx->stream;
    K8S_CUDA(ctx, ctx->dbg[0].reserve((size_t)n * 4 + 16));
    K8S_CUDA(ctx, ctx->dbg[1].reserve((size_t)n * 4 + 16));
    if (n) {
        K8S_CUDA(ctx, cudaMemcpyAsync(ctx->dbg[0].p, keys, (size_t)n * 4, cudaMemcpyHostToDevice, s));
        K8S_CUDA(ctx, cudaMemcpyAsync(ctx->dbg[1].p, vals, (size_t)n * 4, cudaMemcpyHostToDevice, s));
    }
    K8S_CUDA(ctx, cudaEventRecord(ctx->ev_begin, s));
    int rc = sort_pairs(ctx, s, ctx->dbg[0].as<uint32_t>(), ctx->dbg[1].as<uint32_t>(), n, bits, ctx->dbg_sort);
    if (rc) return rc;
    K8S_CUDA(ctx, cudaEventRecord(ctx->ev_end, s));
    if (n) {
        K8S_CUDA(ctx, cudaMemcpyAsync(out_keys, ctx->dbg[0].p, (size_t)n * 4, cudaMemcpyDeviceToHost, s));
        K8S_CUDA(ctx, cudaMemcpyAsync(out_vals, ctx->dbg[1].p, (size_t)n * 4, cudaMemcpyDeviceToHost, s));
    }
    K8S_CUDA(ctx, cudaStreamSynchronize(s));
    float ms = 0.f;
    K8S_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev_begin, ctx->ev_end));
    ctx->stats.device_ms = ms;
    return K8SGPU_OK;
}

extern "C" int k8sgpu_debug_exclusive_scan(k8sgpu_ctx *ctx, const uint32_t *in, uint32_t n, uint32_t *out,
                                           uint32_t *total) {
    K8S_ENTER(ctx);
    reset_call_stats(ctx);
    if (n && (!in || !out)) return fail(ctx, K8SGPU_E_ARG, "exclusive_scan: NULL buffer");
    cudaStream_t s = ctx->stream;
    K8S_CUDA(ctx, ctx->dbg[2].reserve((size_t)n * 4 + 16));
    K8S_CUDA(ctx, ctx->dbg[3].reserve((size_t)n * 4 + 16));
    K8S_CUDA(ctx, ctx->dbg[4].reserve(64));
    if (n) K8S_CUDA(ctx, cudaMemcpyAsync(ctx->dbg[2].p, in, (size_t)n * 4, cudaMemcpyHostToDevice, s));
    int rc = exclusive_scan_u32(ctx, s, ctx->dbg[2].as<uint32_t>(), ctx->dbg[3].as<uint32_t>(), n, ctx->dbg[5],
                                ctx->dbg[4].as<uint32_t>());
    if (rc) return rc;
    if (n) K8S_CUDA(ctx, cudaMemcpyAsync(out, ctx->dbg[3].p, (size_t)n * 4, cudaMemcpyDeviceToHost, s));
    K8S_CUDA(ctx, cudaMemcpyAsync(ctx->h_pinned, ctx->dbg[4].p, 4, cudaMemcpyDeviceToHost, s));
    K8S_CUDA(ctx, cudaStreamSynchronize(s));
    if (total) *total = ctx->h_pinned[0];
    return K8SGPU_OK;
}
