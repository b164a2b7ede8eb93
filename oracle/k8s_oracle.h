/*
 * k8s_oracle.h -- CPU oracle for the k8sgpu hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library.  Nothing under k8s-gpu-_b200/ links, imports or calls it.
 *
 * What it restates (all citations relative to /root/reference):
 *   - the single-pool Reconcile state machine, README.md:170-235, statement by statement
 *     (orc_reconcile_one in reconcile_sm.c);
 *   - its batch form: tag filter (README.md:188,238), currentCount (:195), diff (:199-212),
 *     scale-down victims = first k VMs in list order (:214), ReadyReplicas = pre-action
 *     count (:225)                                                 (orc_reconcile_batch);
 *   - the round-based pod->node placement policy v0.  The reference has NO source for this
 *     (placement is delegated to Volcano/HAMi, GPU调度平台搭建.md:275-298); the policy is
 *     build-defined in SURVEY.md section 8 rows a9-a13 and written out in DESIGN.md.
 *
 * PARITY UNPINNED: the reference is Markdown only -- no runnable code, no tests, no golden
 * vectors and no Go toolchain exists in this image.  The oracle is
 * pinned only to the known-answer tests K1-K14 that tests/ derives from the cited lines.
 *
 * Plain C11, integer only, u64 accumulators, no UB on overflow.
 */
#ifndef K8S_ORACLE_H
#define K8S_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORC_NONE_U32 0xFFFFFFFFu
#define ORC_ASSIGN_NONE (-1)    /* no feasible node */
#define ORC_ASSIGN_PENDING (-2) /* still proposing when max_rounds was reached */
#define ORC_POLICY_ROT_ZERO 1u  /* rot(p) == 0: plain lowest-index tie-break */

/* ---- PRNG: splitmix64.  i-th output (i from 0) of stream `seed` = mix(seed + (i+1)*GAMMA) */
uint64_t orc_mix64(uint64_t z);
uint64_t orc_splitmix64_at(uint64_t seed, uint64_t i);
/* rot(p) = splitmix64_at(p, 0) mod N   */
uint32_t orc_rot(uint64_t pod_index, uint32_t n_nodes, uint32_t policy_flags);

/* ---- tables: structure of arrays, u32 columns */
typedef struct {
    uint32_t n;
    const uint32_t *free_gpu, *free_mcpu, *free_mib, *cap_gpu, *sku_id, *zone_id, *taint_mask, *flags;
} orc_node_table;

typedef struct {
    uint32_t m;
    const uint32_t *req_gpu, *req_mcpu, *req_mib, *sku_mask, *zone_mask, *tol_mask, *pref_zone, *priority;
} orc_pod_table;

typedef struct {
    uint32_t rounds;
    uint32_t placed;
    uint64_t evals; /* sum over rounds of active_pods * n_nodes */
} orc_place_stats;

/* Returns 0 when the tables are inside the input domain DESIGN.md states, else -1. */
int orc_validate_tables(const orc_node_table *nodes, const orc_pod_table *pods);

/* a11 + a12: the packed u64 key of pod p on node n; 0 = infeasible. */
uint64_t orc_score_key(const orc_node_table *nodes, const uint32_t *free_gpu, const uint32_t *free_mcpu,
                       const uint32_t *free_mib, const orc_pod_table *pods, uint32_t p, uint32_t n,
                       uint32_t rot);

/* One scoring pass: best[i] = argmax_n key(active[i], n), -1 if every key is 0.
 * free_* are the CURRENT free columns (nodes->free_* are ignored here).
 * nthreads > 1 shards the active list over pthreads; output is identical. */
void orc_score_round(const orc_node_table *nodes, const uint32_t *free_gpu, const uint32_t *free_mcpu,
                     const uint32_t *free_mib, const orc_pod_table *pods, const uint32_t *active,
                     uint32_t n_active, uint32_t policy_flags, int nthreads, int32_t *best);

/* a13: admission for one round.  active[] is ascending in global pod index; best[] is
 * aligned with it.  Updates assign[] (indexed by global pod), the free columns, and writes
 * the next active list (ascending).  Returns the next active count. */
uint32_t orc_admit_round(uint32_t n_nodes, uint32_t *free_gpu, uint32_t *free_mcpu, uint32_t *free_mib,
                         const orc_pod_table *pods, const uint32_t *active, uint32_t n_active,
                         const int32_t *best, int32_t *assign, uint32_t *next_active);

/* Whole placement: rounds of score + admit until no pod is active (or max_rounds != 0 hit).
 * out_free_* receive the final free columns.  Returns 0, or -1 on invalid tables. */
int orc_place(const orc_node_table *nodes, const orc_pod_table *pods, uint32_t policy_flags,
              uint32_t max_rounds, int nthreads, int32_t *assign, uint32_t *out_free_gpu,
              uint32_t *out_free_mcpu, uint32_t *out_free_mib, orc_place_stats *stats);

/* Gang placement (the reference only shows minAvailable: 1,
 * GPU调度平台搭建.md:648).  Build-defined, layered on orc_place's rounds:
 *   1. place every candidate pod (policy v0 rounds) -- tentative;
 *   2. a gang with fewer placed members than min_available[g] is REJECTED for good: its pods get
 *      ORC_ASSIGN_GANG and what they had reserved is released;
 *   3. pods left unplaced (-1) whose gang is not rejected are retried against the released
 *      capacity; repeat from 2 until an iteration rejects no new gang.
 * gang_id[p] == ORC_NONE_U32: pod belongs to no gang.  gang_ok[g] = 1 if gang g was kept.
 * Returns 0, -1 on invalid tables or a gang id >= n_gangs. */
#define ORC_ASSIGN_GANG (-3)
int orc_place_gang(const orc_node_table *nodes, const orc_pod_table *pods, const uint32_t *gang_id,
                   const uint32_t *min_available, uint32_t n_gangs, uint32_t policy_flags, uint32_t max_rounds,
                   int nthreads, int32_t *assign, uint32_t *out_free_gpu, uint32_t *out_free_mcpu,
                   uint32_t *out_free_mib, int32_t *gang_ok, orc_place_stats *stats, uint32_t *iterations);

/* ---- batch reconcile */
typedef struct {
    uint32_t p;
    const int32_t *desired;        /* Spec.Replicas, >= 0 (README.md:93-95) */
    const uint32_t *owner_id;      /* interned id of "<namespace>-<name>" (README.md:238) */
} orc_pool_table;

typedef struct {
    uint32_t v;
    const uint32_t *owner_id; /* interned `owner` tag value; ORC_NONE_U32 = tag absent */
    const uint32_t *flags;    /* bit0: tag managed-by == "azurevmpool-operator" */
    const uint32_t *ordinal;  /* k if the VM is named "<pool-name>-<k>", else ORC_NONE_U32; may be NULL */
} orc_vm_table;

typedef struct {
    int32_t *pool_of;       /* [V]  pool row of each VM, -1 = unmanaged/foreign      */
    int32_t *current;       /* [P]  README.md:195                                     */
    int32_t *delta;         /* [P]  desired - current, README.md:201,212              */
    int32_t *ready;         /* [P]  README.md:225 (pre-action count)                  */
    int32_t *create_count;  /* [P]  max(delta,0), README.md:201                       */
    int32_t *victim_off;    /* [P+1] CSR offsets                                      */
    uint32_t *victim_vm;    /* [sum k] VM rows to delete, README.md:214 order         */
    int32_t *member_off;    /* [P+1] CSR of every listed VM per pool (Status.VMs), may be NULL */
    uint32_t *member_vm;    /* [sum current] rows ascending within pool, may be NULL  */
    int32_t *create_off;    /* [P+1] CSR of new ordinals (a6), may be NULL            */
    uint32_t *create_ordinal; /* [sum create_count], may be NULL                       */
} orc_reconcile_out;

/* n_owner_ids: owner ids >= this are foreign.  nthreads shards VM rows (output identical).
 * Returns 0; -1 if two pools share an owner id or a desired count is negative. */
int orc_reconcile_batch(const orc_pool_table *pools, const orc_vm_table *vms, uint32_t n_owner_ids,
                        int nthreads, orc_reconcile_out *out);

/* ---- single-pool Reconcile state machine against a fake in-memory ARM client
 *      (reconcile_sm.c; README.md:170-235) */
typedef struct orc_arm orc_arm;     /* fake Azure resource group */
typedef struct orc_cluster orc_cluster; /* fake API server: AzureVmPool objects + status */

orc_arm *orc_arm_new(void);
void orc_arm_free(orc_arm *);
/* Adds a VM with tags; managed_by / owner may be NULL for "tag absent".  Returns row index. */
int orc_arm_add_vm(orc_arm *, const char *name, const char *managed_by, const char *owner);
int orc_arm_vm_count(const orc_arm *);
const char *orc_arm_vm_name(const orc_arm *, int i);
/* Failure injection: the k-th (1-based) upcoming call of that kind fails; 0 = never. */
void orc_arm_fail_list_at(orc_arm *, int k);
void orc_arm_fail_create_at(orc_arm *, int k);
void orc_arm_fail_delete_at(orc_arm *, int k);
void orc_arm_fail_client(orc_arm *, int on);
int orc_arm_calls_list(const orc_arm *);
int orc_arm_calls_create(const orc_arm *);
int orc_arm_calls_delete(const orc_arm *);

orc_cluster *orc_cluster_new(void);
void orc_cluster_free(orc_cluster *);
int orc_cluster_put_pool(orc_cluster *, const char *ns, const char *name, int32_t replicas);
int orc_cluster_set_replicas(orc_cluster *, const char *ns, const char *name, int32_t replicas);
int orc_cluster_delete_pool(orc_cluster *, const char *ns, const char *name);
void orc_cluster_fail_status_update(orc_cluster *, int on);
int32_t orc_cluster_ready_replicas(const orc_cluster *, const char *ns, const char *name);
int orc_cluster_status_vm_count(const orc_cluster *, const char *ns, const char *name);
const char *orc_cluster_status_vm(const orc_cluster *, const char *ns, const char *name, int i);
int orc_cluster_status_writes(const orc_cluster *);

enum {
    ORC_OK = 0,
    ORC_ERR_CLIENT = 1, /* README.md:184 */
    ORC_ERR_LIST = 2,   /* README.md:192 */
    ORC_ERR_CREATE = 3, /* README.md:207 */
    ORC_ERR_DELETE = 4, /* README.md:219 */
    ORC_ERR_STATUS = 5  /* README.md:229 */
};

typedef struct {
    int64_t requeue_after_s; /* 0 = ctrl.Result{} */
    int err;                 /* ORC_OK or ORC_ERR_* */
} orc_result;

orc_result orc_reconcile_one(orc_cluster *, orc_arm *, const char *ns, const char *name);
/* BASELINE.json config 1 (replicas 3 -> 5 -> 2, K1..K5) `iterations` times; returns Reconcile calls made, -1 on a wrong outcome */
long orc_config1_loop(int iterations);

#ifdef __cplusplus
}
#endif
#endif
