/*
 * k8sgpu.h -- C-ABI of libk8sgpu.so: the batch form of the AzureVmPool operator's decision
 * logic and GPU pod->node placement, computed on one B200 (sm_100a) per process.
 *
 * This is the boundary a cgo shim binds (INTEGRATION.md shows it).  Plain C types, caller-
 * owned host buffers, no callbacks, no thread-local state; every entry point selects its CUDA
 * device itself, so it may be called from any OS thread (goroutines migrate between threads).
 * A context is NOT re-entrant: one call at a time per context.  There is NO CPU fallback: if
 * no CUDA device can be opened every call fails with K8SGPU_E_CUDA.
 *
 * Citations are file:line under the reference tree (Andy-ckm/K8S-GPU-).  The reference has no
 * FFI of its own (it is Go text inside Markdown); each entry point names the reference
 * statements it replaces.
 */
#ifndef K8SGPU_H
#define K8SGPU_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define K8SGPU_ABI_VERSION 2

/* error codes: every export returns 0 or one of these; text via k8sgpu_last_error */
#define K8SGPU_OK 0
#define K8SGPU_E_ARG (-1)      /* NULL pointer, table outside the input domain, duplicate owner id */
#define K8SGPU_E_CUDA (-2)     /* CUDA runtime error or no device */
#define K8SGPU_E_NCCL (-3)     /* NCCL missing, or a communicator / collective call failed */
#define K8SGPU_E_CAPACITY (-4) /* a caller-sized output buffer is too small */

#define K8SGPU_NONE_U32 0xFFFFFFFFu
#define K8SGPU_ASSIGN_NONE (-1)    /* pod has no feasible node                       */
#define K8SGPU_ASSIGN_PENDING (-2) /* pod still proposing when max_rounds was reached */
#define K8SGPU_POLICY_ROT_ZERO 1u  /* rot(p) = 0: plain lowest-index tie-break        */
/* `stream` arguments are cudaStream_t values.  0 keeps its CUDA meaning (the legacy default
 * stream); this sentinel selects the context's own internal non-blocking stream instead. */
#define K8SGPU_STREAM_OWN ((void *)(uintptr_t)-1)

typedef struct k8sgpu_ctx k8sgpu_ctx;

/* ------------------------------------------------------------------ tables (SoA, host) */

/* One row per AzureVmPool.  desired = Spec.Replicas (README.md:93-95,196).  owner_id = the
 * interned id of the pool's owner tag value "<namespace>-<name>" (README.md:238). */
typedef struct {
    uint32_t p;
    const int32_t *desired;
    const uint32_t *owner_id;
} k8sgpu_pool_table;

/* One row per VM observed in the resource group, in list order (README.md:189,214).
 * owner_id: interned value of the VM's `owner` tag, K8SGPU_NONE_U32 if absent.
 * flags bit0: tag managed-by == "azurevmpool-operator" (README.md:238).
 * ordinal: k when the VM is named "<pool-name>-<k>", else K8SGPU_NONE_U32; may be NULL. */
typedef struct {
    uint32_t v;
    const uint32_t *owner_id;
    const uint32_t *flags;
    const uint32_t *ordinal;
} k8sgpu_vm_table;

/* Caller-sized outputs.  Capacities: pool_of V; current/delta/ready/create_count P;
 * *_off P+1; victim_vm and member_vm V; create_ordinal create_capacity.
 * member_*, create_off/create_ordinal may be NULL (then they are not produced). */
typedef struct {
    int32_t *pool_of;         /* listManagedVMs tag filter, README.md:188-189,238              */
    int32_t *current;         /* currentCount := len(existingVMs), README.md:195               */
    int32_t *delta;           /* desired - current, README.md:199-201,210-212                  */
    int32_t *ready;           /* Status.ReadyReplicas = currentCount (pre-action), :225        */
    int32_t *create_count;    /* scaleUpCount, README.md:201                                   */
    int32_t *victim_off;      /* CSR over pools                                                */
    uint32_t *victim_vm;      /* existingVMs[:scaleDownCount] as VM rows, README.md:214        */
    int32_t *member_off;      /* CSR of existingVMs per pool (Status.VMs, README.md:124-125)   */
    uint32_t *member_vm;
    int32_t *create_off;      /* CSR of the ordinals createVM would pick, README.md:204-205    */
    uint32_t *create_ordinal;
    uint64_t create_capacity; /* elements available in create_ordinal                          */
} k8sgpu_reconcile_out;

/* Where each output array starts (in 4-byte elements) inside ONE caller block of `total` elements
 * that holds them all, capacities as listed above.  A caller that allocates its outputs this way
 * (k8sgpu_reconcile_out_size) gets them back in as few device->host copies as possible: the library
 * keeps the same layout in HBM and merges adjacent transfers -- the tag-filter result and everything
 * sized by P travel as one copy.  The three lists (victim_vm, member_vm, create_ordinal) come last
 * because only their used prefixes are copied.  Any other placement of the buffers works too. */
typedef struct {
    uint64_t pool_of, current, ready, delta, create_count, victim_off, member_off, create_off;
    uint64_t victim_vm, member_vm, create_ordinal, total;
} k8sgpu_reconcile_layout;

/* Node and pod records: 8 u32 columns each (DESIGN.md "tables").  The reference holds no
 * scheduler source; the pod shape comes from GPU调度平台搭建.md:665-667 (nvidia.com/gpu: 1)
 * and :534 (gpu-1x-16c-32g-1gpu), the node attributes from README.md:99-100 (location, vmSize).
 * Input domain (else K8SGPU_E_ARG): free_gpu <= 0x7FFF, cap_gpu <= 255, sku_id < 32,
 * zone_id < 32, every free_* and req_* < 2^31, pref_zone < 32 or K8SGPU_NONE_U32,
 * n < 2^31 - 4096, m < 2^31. */
typedef struct {
    uint32_t n;
    const uint32_t *free_gpu, *free_mcpu, *free_mib, *cap_gpu, *sku_id, *zone_id, *taint_mask, *flags;
} k8sgpu_node_table;

/* priority is RESERVED: policy v0 admits in plain pod-index order and never reads it.  It may be
 * NULL; a caller that sets priorities must order its pod table accordingly (row order IS the
 * admission order). */
typedef struct {
    uint32_t m;
    const uint32_t *req_gpu, *req_mcpu, *req_mib, *sku_mask, *zone_mask, *tol_mask, *pref_zone, *priority;
} k8sgpu_pod_table;

typedef struct {
    uint32_t flags;      /* K8SGPU_POLICY_* */
    uint32_t max_rounds; /* 0 = run until no pod is active */
} k8sgpu_policy;

typedef struct {
    int32_t *assign;     /* [m] node row, K8SGPU_ASSIGN_NONE or K8SGPU_ASSIGN_PENDING */
    uint32_t *free_gpu;  /* [n] free columns after placement; each may be NULL */
    uint32_t *free_mcpu;
    uint32_t *free_mib;
} k8sgpu_place_out;

typedef struct {
    uint32_t rounds;          /* placement rounds of the last place call                         */
    uint32_t placed;          /* pods with assign >= 0                                           */
    uint64_t evals;           /* sum over rounds of active pods x nodes                          */
    uint64_t kernel_launches; /* kernels this library launched in the last call                  */
    uint64_t h2d_bytes;       /* bytes copied host->device / device->host in the last call       */
    uint64_t d2h_bytes;
    float device_ms;          /* CUDA-event time of the device portion (no H2D/D2H) of last call */
    float score_ms;           /* CUDA-event time spent in the scoring kernel (sum over rounds)   */
    uint64_t score_evals;     /* evals covered by score_ms                                       */
    uint32_t score_launches;  /* scoring-kernel launches covered by score_ms                     */
    uint32_t gang_iterations; /* place/retry iterations of the last k8sgpu_place_gang call       */
} k8sgpu_stats;

/* ------------------------------------------------------------------ context */

/* Opens the listed CUDA devices (1 <= ndev <= 16, all distinct, all sm_100) as ONE context.
 * Replaces nothing in the reference; it is what SetupWithManager would call once.
 *   ndev == 1: one GPU.  Several such contexts, one per process, can still share a placement
 *              (k8sgpu_comm_init_rank below).
 *   ndev  > 1: a placement group inside this process -- what a single controller process with
 *              eight B200s uses.  The library owns one NCCL communicator per device (NCCL is
 *              loaded at this point; K8SGPU_E_NCCL if it cannot be) and enables peer access.
 *              k8sgpu_place / _place_gang / _place_upload / _place_run then shard the pods of
 *              every round across the devices (tables replicated, proposals exchanged by
 *              ncclAllGather) and return byte-identical results to ndev == 1.  Batch reconcile
 *              and the other entry points run on devices[0] ("replicas only").
 * k8sgpu_create_ex is the same with the failure text copied into a caller buffer (errbuf may be
 * NULL), for callers that create contexts from several threads at once. */
int k8sgpu_create(const int *devices, int ndev, k8sgpu_ctx **out);
int k8sgpu_create_ex(const int *devices, int ndev, k8sgpu_ctx **out, char *errbuf, size_t errcap);
int k8sgpu_destroy(k8sgpu_ctx *ctx);
/* Message of the last failing call on ctx.  ctx == NULL: the most recent k8sgpu_create failure
 * of this process (one process-wide buffer; use k8sgpu_create_ex when creations can overlap). */
const char *k8sgpu_last_error(const k8sgpu_ctx *ctx);
int k8sgpu_get_stats(const k8sgpu_ctx *ctx, k8sgpu_stats *out);
int k8sgpu_abi_version(void);

/* One process per GPU (torchrun, MPI): every process creates a one-device context, rank 0 makes
 * an id with k8sgpu_comm_unique_id (an ncclUniqueId, K8SGPU_UNIQUE_ID_BYTES bytes), hands it to the
 * others by whatever channel the job has, and all call k8sgpu_comm_init_rank -- a collective call:
 * it returns once all `world` ranks have joined.  From then on k8sgpu_place and friends on these
 * contexts are collective too: every rank passes the same tables and gets the same result.
 * K8SGPU_E_NCCL when NCCL cannot be loaded or reports an error.  No reference counterpart. */
#define K8SGPU_UNIQUE_ID_BYTES 128
int k8sgpu_comm_unique_id(void *id128);
int k8sgpu_comm_init_rank(k8sgpu_ctx *ctx, int rank, int world, const void *id128);
/* world / rank / devices driven by this process / NCCL version in use (0 while NCCL is not loaded);
 * any pointer may be NULL. */
int k8sgpu_group_info(const k8sgpu_ctx *ctx, int *world, int *rank, int *nlocal, int *nccl_version);

/* Page-locked ("pinned") host memory for tables and results.  Not required -- any host memory works
 * -- but copies from and to pinned buffers are asynchronous and run at the full PCIe rate; a cgo
 * caller gets that without binding the CUDA runtime itself (snapshot staging).  Free
 * with k8sgpu_host_free.  K8SGPU_E_CUDA when no device / no memory.  No reference counterpart. */
int k8sgpu_host_alloc(size_t bytes, void **out);
int k8sgpu_host_free(void *p);

/* Developer and test knobs by name ("res", "ppt", "splits", "order", "force_cls" (16 / 17 / 18
 * only), "slot_admit", "coop", "coop_grid", "rec_path", "rec_grid", "rec_segments", "coop_prof",
 * "rec_prof", "exchange", "shard_upload"; -1 = automatic; README.md lists what each selects).
 * Every setting gives bit-identical results: they select kernel variants and launch shapes
 * ("force_cls" below 12 is the exception: timing experiments only, results are wrong).  The
 * K8SGPU_<NAME> environment variables seed them once, inside k8sgpu_create; nothing reads the
 * environment afterwards.  No reference counterpart. */
int k8sgpu_set_tuning(k8sgpu_ctx *ctx, const char *key, int value);

/* ------------------------------------------------------------------ batch reconcile */

/* Steps 3 and 4 of Reconcile for every pool at once (README.md:187-222): tag-filtered list,
 * currentCount, the desired-vs-actual diff, the scale-down victim slice and the scale-up
 * count, plus the status fields of step 5 (README.md:225).  n_owner_ids bounds the interned
 * owner ids (ids >= n_owner_ids are foreign).  Host buffers in, host buffers out. */
int k8sgpu_reconcile_batch(k8sgpu_ctx *ctx, const k8sgpu_pool_table *pools, const k8sgpu_vm_table *vms,
                           uint32_t n_owner_ids, k8sgpu_reconcile_out *out);

/* Capacities and the single-block layout of k8sgpu_reconcile_out for p pools, v VM rows and
 * room for create_capacity new-VM ordinals (the sum of the scale-up counts; the sum of all desired
 * replicas is always enough).  See k8sgpu_reconcile_layout.  Needs no context and no GPU. */
int k8sgpu_reconcile_out_size(uint32_t p, uint32_t v, uint64_t create_capacity, k8sgpu_reconcile_layout *out);

/* The same split in three: tables stay resident in HBM between calls (bench `value`). */
int k8sgpu_reconcile_upload(k8sgpu_ctx *ctx, const k8sgpu_pool_table *pools, const k8sgpu_vm_table *vms,
                            uint32_t n_owner_ids);
int k8sgpu_reconcile_run(k8sgpu_ctx *ctx, int want_members, int want_ordinals);
int k8sgpu_reconcile_download(k8sgpu_ctx *ctx, k8sgpu_reconcile_out *out);

/* ------------------------------------------------------------------ placement */

/* Filter / score / bin-pack of every pending pod onto the node table, round based
 * (DESIGN.md policy v0).  Stands where the reference hands pods to a third-party scheduler
 * (schedulerName: volcano, GPU调度平台搭建.md:649; helm installs :275-298). */
int k8sgpu_place(k8sgpu_ctx *ctx, const k8sgpu_node_table *nodes, const k8sgpu_pod_table *pods,
                 const k8sgpu_policy *policy, k8sgpu_place_out *out);

/* Resident form: upload once, run any number of times (each run restarts from the uploaded
 * tables), download results. */
int k8sgpu_place_upload(k8sgpu_ctx *ctx, const k8sgpu_node_table *nodes, const k8sgpu_pod_table *pods);
int k8sgpu_place_run(k8sgpu_ctx *ctx, const k8sgpu_policy *policy);
int k8sgpu_place_download(k8sgpu_ctx *ctx, k8sgpu_place_out *out);

/* Gang placement: all-or-nothing groups, the reference's `minAvailable` (GPU调度平台搭建.md:648,
 * value 1 in its only example) generalised.  Layered on the round loop of k8sgpu_place:
 *   1. every candidate pod is placed by the normal rounds (tentatively);
 *   2. a gang with fewer placed members than min_available[g] is rejected for good: its pods
 *      get K8SGPU_ASSIGN_GANG and what they had reserved is released;
 *   3. pods still unplaced (-1) whose gang is not rejected are retried against the released
 *      capacity; back to 2 until an iteration rejects no new gang.
 * gang_id[p] == K8SGPU_NONE_U32: the pod is on its own.  gang_ok[g] (may be NULL) = 1 if kept.
 * Queues (`queue: default`, :650) are not modelled.  k8sgpu_place_gang is the one-call form on
 * one GPU; the stepwise form below lets one process per GPU shard the rounds. */
#define K8SGPU_ASSIGN_GANG (-3)
typedef struct {
    uint32_t n_gangs;
    const uint32_t *gang_id;        /* [m]        */
    const uint32_t *min_available;  /* [n_gangs]  */
} k8sgpu_gang_table;
int k8sgpu_place_gang(k8sgpu_ctx *ctx, const k8sgpu_node_table *nodes, const k8sgpu_pod_table *pods,
                      const k8sgpu_gang_table *gangs, const k8sgpu_policy *policy, k8sgpu_place_out *out,
                      int32_t *gang_ok);

/* One scoring pass of all pods against the uploaded node table (no admission):
 * best_node[m] = argmax node per pod, -1 if none.  Host buffer out. */
int k8sgpu_score_once(k8sgpu_ctx *ctx, const k8sgpu_policy *policy, int32_t *best_node);

/* Round-at-a-time form for several cooperating processes (one per GPU): every process
 * uploads the same tables, scores its own slice [lo,hi) of the active list into a DEVICE
 * buffer, the caller exchanges the slices (all-gather), and every process then runs the
 * admission step on the full proposal list, which keeps all replicas identical.
 * `stream` is a cudaStream_t, with CUDA's usual meaning of 0 (the legacy default stream, which is
 * what torch.cuda.current_stream().cuda_stream is unless a side stream was entered);
 * K8SGPU_STREAM_OWN selects the context's internal stream.
 *   begin : reset state from the uploaded tables; *n_active = m
 *   score : d_best[i - lo] = best node of active[i], lo <= i < hi        (asynchronous)
 *   admit : d_best_all[i] for 0 <= i < n_active; returns the next active count (synchronises)
 *   end   : fills rounds / placed / evals of k8sgpu_get_stats */
int k8sgpu_place_round_begin(k8sgpu_ctx *ctx, const k8sgpu_policy *policy, void *stream, uint32_t *n_active);
int k8sgpu_place_round_score(k8sgpu_ctx *ctx, uint32_t lo, uint32_t hi, int32_t *d_best, void *stream);
int k8sgpu_place_round_admit(k8sgpu_ctx *ctx, const int32_t *d_best_all, void *stream, uint32_t *n_active);
int k8sgpu_place_round_end(k8sgpu_ctx *ctx, void *stream);

/* Gang placement one step at a time, for the same cooperating processes.  After
 * k8sgpu_place_upload:
 *   gang_begin    : gang table in, state reset; *n_active = m
 *   (rounds)      : k8sgpu_place_round_score / _score_p2p + _admit until the active count is 0
 *   gang_verdicts : steps 2 and 3 of k8sgpu_place_gang on this replica; *n_active = length of the
 *                   retry list -- 0 means final, otherwise run the rounds again and call it again
 *   gang_end      : as k8sgpu_place_round_end; gang_ok[n_gangs] (host, may be NULL) = 1 if kept.
 * The verdict kernels are deterministic functions of the replicated state, so every rank reaches
 * the same verdicts without a collective. */
int k8sgpu_place_gang_begin(k8sgpu_ctx *ctx, const k8sgpu_gang_table *gangs, const k8sgpu_policy *policy,
                            void *stream, uint32_t *n_active);
int k8sgpu_place_gang_verdicts(k8sgpu_ctx *ctx, void *stream, uint32_t *n_active);
int k8sgpu_place_gang_end(k8sgpu_ctx *ctx, void *stream, int32_t *gang_ok);

/* Peer exchange for the round-at-a-time form: instead of an NCCL all-gather of the proposal
 * slices, the kernel that decodes the scores stores every proposal directly into EVERY rank's
 * exchange buffer over NVLink peer memory (CUDA IPC; one process per GPU on one node).
 *   export : allocates this rank's exchange buffer (two halves of m proposals, used alternately
 *            by even and odd rounds), returns its 64-byte IPC handle and its local device pointer
 *            (round r's proposals are at d_local + (r & 1) * m for k8sgpu_place_round_admit);
 *   import : handles = world x 64 bytes, rank r's handle at offset 64 r (gathered by the host);
 *   score_p2p : like k8sgpu_place_round_score, but the results land at positions [lo,hi) of all
 *            ranks' buffers.  The caller then needs only a barrier across ranks (e.g. a
 *            1-element all-reduce on the same stream) before k8sgpu_place_round_admit.
 * No reference counterpart. */
int k8sgpu_peer_export(k8sgpu_ctx *ctx, uint32_t m, void *handle64, void **d_local);
/* Unmaps the peers' buffers.  Before a group exports NEW buffers (a larger pod table), every rank
 * calls this and the ranks meet at a barrier of their own: exported memory must not be freed
 * while another process still maps it. */
int k8sgpu_peer_close(k8sgpu_ctx *ctx);
int k8sgpu_peer_import(k8sgpu_ctx *ctx, int rank, int world, const void *handles);
int k8sgpu_place_round_score_p2p(k8sgpu_ctx *ctx, uint32_t lo, uint32_t hi, void *stream);

/* ------------------------------------------------------------------ primitives (tests) */

/* Stable LSD radix sort of (key,val) pairs on `bits` low key bits, and an exclusive u32
 * prefix sum: the two device primitives victim order and admission order are built on.
 * Host buffers in and out.  No reference counterpart. */
int k8sgpu_debug_sort_pairs(k8sgpu_ctx *ctx, const uint32_t *keys, const uint32_t *vals, uint32_t n, int bits,
                            uint32_t *out_keys, uint32_t *out_vals);
int k8sgpu_debug_exclusive_scan(k8sgpu_ctx *ctx, const uint32_t *in, uint32_t n, uint32_t *out, uint32_t *total);

/* Exclusive u32 prefix sum that allocates nothing.  This entry point has no counterpart in the
 * reference.  It is for callers that may not allocate device memory once they are running: a
 * controller that shares its GPU with the jobs it schedules and lives inside a memory slice
 * fixed at start-up (a late cudaMalloc can fail there), and callers with GPU work in flight
 * that a device-wide stall would hurt (cudaMalloc and cudaFree synchronise the device).
 *
 * d_in, d_out (n elements each) and d_total (one element, may be NULL) are DEVICE pointers the
 * caller owns; d_out may equal d_in.  `stream` is a cudaStream_t (0 = legacy default stream, K8SGPU_STREAM_OWN = the context's); the
 * call only enqueues work and returns.  The block sums the scan needs live in a fixed 32 KiB
 * array inside the library image on every GPU that loads it, which limits n to
 * K8SGPU_SCAN_NOALLOC_MAX_N; a larger n returns K8SGPU_E_CAPACITY and launches nothing.
 *
 * The first call on each device also looks up the device address of that array, so make that
 * first call before starting any stream capture.  All calls on a device share the one array
 * and it keeps whatever the previous call wrote (the scan overwrites every slot before reading
 * it, so no clearing is needed); two calls whose kernels may overlap on the GPU, e.g. on
 * different streams, race on it and must be ordered by the caller. */
#define K8SGPU_SCAN_NOALLOC_MAX_N (1u << 24)
int k8sgpu_exclusive_scan_noalloc(k8sgpu_ctx *ctx, const uint32_t *d_in, uint32_t n, uint32_t *d_out,
                                  uint32_t *d_total, void *stream);

#ifdef __cplusplus
}
#endif
#endif
