/*
 * k8s_oracle.c -- CPU oracle: batch reconcile + round-based placement.
 * TEST INFRASTRUCTURE ONLY (see k8s_oracle.h).  PARITY UNPINNED (reference has no code/tests).
 *
 * Everything here is the literal, slow, obviously-right form of DESIGN.md's policy v0.
 */
#include "k8s_oracle.h"

#include <pthread.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ PRNG */
#define ORC_GAMMA 0x9E3779B97F4A7C15ull

uint64_t orc_mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

uint64_t orc_splitmix64_at(uint64_t seed, uint64_t i) { return orc_mix64(seed + (i + 1) * ORC_GAMMA); }

uint32_t orc_rot(uint64_t pod_index, uint32_t n_nodes, uint32_t policy_flags) {
    if ((policy_flags & ORC_POLICY_ROT_ZERO) || n_nodes == 0) return 0;
    return (uint32_t)(orc_splitmix64_at(pod_index, 0) % (uint64_t)n_nodes);
}

/* ------------------------------------------------------------------ placement */
int orc_validate_tables(const orc_node_table *nodes, const orc_pod_table *pods) {
    if (nodes->n >= 0x80000000u || pods->m >= 0x80000000u) return -1;
    for (uint32_t n = 0; n < nodes->n; ++n) {
        if (nodes->free_gpu[n] > 0x7FFFu || nodes->cap_gpu[n] > 255u) return -1;
        if (nodes->free_mcpu[n] >= 0x80000000u || nodes->free_mib[n] >= 0x80000000u) return -1;
        if (nodes->sku_id[n] >= 32u || nodes->zone_id[n] >= 32u) return -1;
    }
    for (uint32_t p = 0; p < pods->m; ++p) {
        if (pods->req_gpu[p] >= 0x80000000u || pods->req_mcpu[p] >= 0x80000000u ||
            pods->req_mib[p] >= 0x80000000u)
            return -1;
        if (pods->pref_zone[p] >= 32u && pods->pref_zone[p] != ORC_NONE_U32) return -1;
    }
    return 0;
}

/* SURVEY.md 8 a11 (filter) and a12 (key).  No reference source exists for either. */
uint64_t orc_score_key(const orc_node_table *nodes, const uint32_t *free_gpu, const uint32_t *free_mcpu,
                       const uint32_t *free_mib, const orc_pod_table *pods, uint32_t p, uint32_t n,
                       uint32_t rot) {
    /* a11: every clause must hold */
    if (!(nodes->flags[n] & 1u)) return 0;
    if (free_gpu[n] < pods->req_gpu[p]) return 0;
    if (free_mcpu[n] < pods->req_mcpu[p]) return 0;
    if (free_mib[n] < pods->req_mib[p]) return 0;
    if (!((pods->sku_mask[p] >> nodes->sku_id[n]) & 1u)) return 0;
    if (!((pods->zone_mask[p] >> nodes->zone_id[n]) & 1u)) return 0;
    if ((nodes->taint_mask[n] & ~pods->tol_mask[p]) != 0u) return 0;

    /* a12: lexicographic u64, max wins */
    uint32_t left = free_gpu[n] - pods->req_gpu[p];
    if (left > 0x7FFFu) left = 0x7FFFu;
    uint64_t key = 1ull << 63;
    key |= (uint64_t)(0x7FFFu - left) << 48;              /* best fit: least GPUs left over      */
    key |= (uint64_t)(255u - nodes->cap_gpu[n]) << 40;    /* prefer the smallest SKU that fits   */
    key |= (uint64_t)(nodes->zone_id[n] == pods->pref_zone[p] ? 1u : 0u) << 39;
    uint32_t N = nodes->n;
    uint32_t ridx = (uint32_t)(((uint64_t)n + N - rot) % N); /* (n - rot) mod N */
    key |= (uint64_t)(0xFFFFFFFFu - ridx);
    return key;
}

typedef struct {
    const orc_node_table *nodes;
    const uint32_t *fg, *fc, *fm;
    const orc_pod_table *pods;
    const uint32_t *active;
    uint32_t n_active, flags, chunk;
    int32_t *best;
} score_job;

/* best[i] for active positions [lo, hi): every node, in ascending order, strict '>' */
static void score_range(const score_job *j, uint32_t lo, uint32_t hi) {
    const uint32_t N = j->nodes->n;
    for (uint32_t i = lo; i < hi; ++i) {
        uint32_t p = j->active[i];
        uint32_t rot = orc_rot(p, N, j->flags);
        uint64_t bk = 0;
        int32_t bn = -1;
        for (uint32_t n = 0; n < N; ++n) {
            uint64_t k = orc_score_key(j->nodes, j->fg, j->fc, j->fm, j->pods, p, n, rot);
            if (k > bk) {
                bk = k;
                bn = (int32_t)n;
            }
        }
        j->best[i] = bn;
    }
}

/* A persistent pool of worker threads (created on first use, kept for the life of the process):
 * a round hands out chunks of the active list through one shared counter, so a scoring pass costs
 * two condition-variable hand-shakes instead of nthreads pthread_create/join pairs, and threads
 * that draw cheap pods (early filter exits) simply take more chunks.  Which thread scores a pod
 * cannot change best[i]: every entry is a function of the tables alone. */
static struct {
    pthread_mutex_t mu;
    pthread_cond_t work, done;
    pthread_t *tid;
    int nworkers;      /* threads created so far                                  */
    int want;          /* workers [0, want) take part in the current job          */
    int pending;       /* workers that have not finished the current job yet      */
    uint64_t gen;      /* job generation: bumped once per posted job              */
    const score_job *job;
    uint32_t next;     /* next unclaimed active position (atomic)                 */
} g_pool = {PTHREAD_MUTEX_INITIALIZER, PTHREAD_COND_INITIALIZER, PTHREAD_COND_INITIALIZER, NULL, 0, 0, 0, 0, NULL, 0};

static void pool_drain(const score_job *j) {
    for (;;) {
        uint32_t lo = __atomic_fetch_add(&g_pool.next, j->chunk, __ATOMIC_RELAXED);
        if (lo >= j->n_active) return;
        uint32_t hi = lo + j->chunk;
        score_range(j, lo, hi < j->n_active ? hi : j->n_active);
    }
}

static void *pool_worker(void *arg) {
    const int me = (int)(intptr_t)arg;
    uint64_t seen = 0;
    pthread_mutex_lock(&g_pool.mu);
    for (;;) {
        while (g_pool.gen == seen) pthread_cond_wait(&g_pool.work, &g_pool.mu);
        seen = g_pool.gen;
        if (me >= g_pool.want) continue;  /* not part of this job */
        const score_job *j = g_pool.job;
        pthread_mutex_unlock(&g_pool.mu);
        pool_drain(j);
        pthread_mutex_lock(&g_pool.mu);
        if (--g_pool.pending == 0) pthread_cond_signal(&g_pool.done);
    }
    return NULL;
}

void orc_score_round(const orc_node_table *nodes, const uint32_t *free_gpu, const uint32_t *free_mcpu,
                     const uint32_t *free_mib, const orc_pod_table *pods, const uint32_t *active,
                     uint32_t n_active, uint32_t policy_flags, int nthreads, int32_t *best) {
    score_job j;
    j.nodes = nodes; j.fg = free_gpu; j.fc = free_mcpu; j.fm = free_mib;
    j.pods = pods; j.active = active; j.n_active = n_active; j.flags = policy_flags; j.best = best;
    if (nthreads < 1) nthreads = 1;
    if ((uint32_t)nthreads > n_active) nthreads = n_active ? (int)n_active : 1;
    if (nthreads == 1) {
        score_range(&j, 0, n_active);
        return;
    }
    /* ~16 chunks per thread, at most 256 pods each: fine enough to balance, coarse enough that the
     * shared counter is touched once per ~10^6 evals at config-3 sizes */
    uint32_t chunk = n_active / ((uint32_t)nthreads * 16u);
    j.chunk = chunk < 1 ? 1 : (chunk > 256 ? 256 : chunk);
    pthread_mutex_lock(&g_pool.mu);
    const int helpers = nthreads - 1;  /* the calling thread works too */
    if (g_pool.nworkers < helpers) {
        g_pool.tid = (pthread_t *)realloc(g_pool.tid, (size_t)helpers * sizeof(pthread_t));
        while (g_pool.nworkers < helpers) {
            if (pthread_create(&g_pool.tid[g_pool.nworkers], NULL, pool_worker, (void *)(intptr_t)g_pool.nworkers) != 0)
                break;  /* fewer threads than asked for: the job still completes */
            pthread_detach(g_pool.tid[g_pool.nworkers]);
            g_pool.nworkers++;
        }
    }
    g_pool.want = helpers < g_pool.nworkers ? helpers : g_pool.nworkers;
    g_pool.pending = g_pool.want;
    g_pool.job = &j;
    __atomic_store_n(&g_pool.next, 0u, __ATOMIC_RELAXED);
    g_pool.gen++;
    pthread_cond_broadcast(&g_pool.work);
    pthread_mutex_unlock(&g_pool.mu);
    pool_drain(&j);
    pthread_mutex_lock(&g_pool.mu);
    while (g_pool.pending > 0) pthread_cond_wait(&g_pool.done, &g_pool.mu);
    pthread_mutex_unlock(&g_pool.mu);
}

/* prefix sums over ALL proposers of a node in global pod order. */
uint32_t orc_admit_round(uint32_t n_nodes, uint32_t *free_gpu, uint32_t *free_mcpu, uint32_t *free_mib,
                         const orc_pod_table *pods, const uint32_t *active, uint32_t n_active,
                         const int32_t *best, int32_t *assign, uint32_t *next_active) {
    uint64_t *S = (uint64_t *)calloc((size_t)n_nodes * 3 + 1, sizeof(uint64_t));  /* running demand */
    uint64_t *A = (uint64_t *)calloc((size_t)n_nodes * 3 + 1, sizeof(uint64_t));  /* admitted       */
    uint32_t next = 0;
    for (uint32_t i = 0; i < n_active; ++i) {
        uint32_t p = active[i];
        int32_t n = best[i];
        if (n < 0) {
            assign[p] = ORC_ASSIGN_NONE;
            continue;
        }
        uint64_t *s = &S[(size_t)n * 3];
        s[0] += pods->req_gpu[p];
        s[1] += pods->req_mcpu[p];
        s[2] += pods->req_mib[p];
        if (s[0] <= free_gpu[n] && s[1] <= free_mcpu[n] && s[2] <= free_mib[n]) {
            assign[p] = n;
            uint64_t *a = &A[(size_t)n * 3];
            a[0] += pods->req_gpu[p];
            a[1] += pods->req_mcpu[p];
            a[2] += pods->req_mib[p];
        } else {
            next_active[next++] = p; /* retries next round against the updated tables */
        }
    }
    for (uint32_t n = 0; n < n_nodes; ++n) {
        free_gpu[n] -= (uint32_t)A[(size_t)n * 3 + 0];
        free_mcpu[n] -= (uint32_t)A[(size_t)n * 3 + 1];
        free_mib[n] -= (uint32_t)A[(size_t)n * 3 + 2];
    }
    free(S);
    free(A);
    return next;
}

int orc_place(const orc_node_table *nodes, const orc_pod_table *pods, uint32_t policy_flags,
              uint32_t max_rounds, int nthreads, int32_t *assign, uint32_t *fg, uint32_t *fc,
              uint32_t *fm, orc_place_stats *stats) {
    if (orc_validate_tables(nodes, pods) != 0) return -1;
    const uint32_t N = nodes->n, M = pods->m;
    memcpy(fg, nodes->free_gpu, (size_t)N * 4);
    memcpy(fc, nodes->free_mcpu, (size_t)N * 4);
    memcpy(fm, nodes->free_mib, (size_t)N * 4);
    uint32_t *act = (uint32_t *)malloc(((size_t)M + 1) * 4);
    uint32_t *nxt = (uint32_t *)malloc(((size_t)M + 1) * 4);
    int32_t *best = (int32_t *)malloc(((size_t)M + 1) * 4);
    uint32_t na = M, rounds = 0;
    uint64_t evals = 0;
    for (uint32_t p = 0; p < M; ++p) {
        act[p] = p;
        assign[p] = ORC_ASSIGN_PENDING;
    }
    while (na > 0 && (max_rounds == 0 || rounds < max_rounds)) {
        orc_score_round(nodes, fg, fc, fm, pods, act, na, policy_flags, nthreads, best);
        evals += (uint64_t)na * N;
        uint32_t nn = orc_admit_round(N, fg, fc, fm, pods, act, na, best, assign, nxt);
        uint32_t *t = act; act = nxt; nxt = t;
        na = nn;
        ++rounds;
    }
    uint32_t placed = 0;
    for (uint32_t p = 0; p < M; ++p) placed += assign[p] >= 0;
    if (stats) {
        stats->rounds = rounds;
        stats->placed = placed;
        stats->evals = evals;
    }
    free(act); free(nxt); free(best);
    return 0;
}

/* all-or-nothing gangs on top of the round loop (see k8s_oracle.h). */
int orc_place_gang(const orc_node_table *nodes, const orc_pod_table *pods, const uint32_t *gang_id,
                   const uint32_t *min_available, uint32_t n_gangs, uint32_t policy_flags, uint32_t max_rounds,
                   int nthreads, int32_t *assign, uint32_t *fg, uint32_t *fc, uint32_t *fm, int32_t *gang_ok,
                   orc_place_stats *stats, uint32_t *iterations) {
    if (orc_validate_tables(nodes, pods) != 0) return -1;
    const uint32_t N = nodes->n, M = pods->m;
    for (uint32_t p = 0; p < M; ++p)
        if (gang_id[p] != ORC_NONE_U32 && gang_id[p] >= n_gangs) return -1;
    memcpy(fg, nodes->free_gpu, (size_t)N * 4);
    memcpy(fc, nodes->free_mcpu, (size_t)N * 4);
    memcpy(fm, nodes->free_mib, (size_t)N * 4);
    uint32_t *act = (uint32_t *)malloc(((size_t)M + 1) * 4);
    uint32_t *nxt = (uint32_t *)malloc(((size_t)M + 1) * 4);
    int32_t *best = (int32_t *)malloc(((size_t)M + 1) * 4);
    uint32_t *count = (uint32_t *)calloc((size_t)n_gangs + 1, 4);
    uint8_t *rejected = (uint8_t *)calloc((size_t)n_gangs + 1, 1);
    uint8_t *newly = (uint8_t *)calloc((size_t)n_gangs + 1, 1);
    uint32_t na = M, rounds = 0, iters = 0;
    uint64_t evals = 0;
    for (uint32_t p = 0; p < M; ++p) {
        act[p] = p;
        assign[p] = ORC_ASSIGN_PENDING;
    }
    for (;;) {
        while (na > 0 && (max_rounds == 0 || rounds < max_rounds)) {  /* step 1: policy v0 rounds */
            orc_score_round(nodes, fg, fc, fm, pods, act, na, policy_flags, nthreads, best);
            evals += (uint64_t)na * N;
            uint32_t nn = orc_admit_round(N, fg, fc, fm, pods, act, na, best, assign, nxt);
            uint32_t *t = act; act = nxt; nxt = t;
            na = nn;
            ++rounds;
        }
        ++iters;
        if (na > 0) break; /* max_rounds reached: leave the rest pending, no verdicts on half a placement */
        for (uint32_t g = 0; g < n_gangs; ++g) count[g] = 0;
        for (uint32_t p = 0; p < M; ++p)
            if (gang_id[p] != ORC_NONE_U32 && assign[p] >= 0) count[gang_id[p]]++;
        uint32_t n_new = 0;
        for (uint32_t g = 0; g < n_gangs; ++g) { /* step 2 */
            newly[g] = (!rejected[g] && count[g] < min_available[g]) ? 1 : 0;
            if (newly[g]) { rejected[g] = 1; ++n_new; }
        }
        if (n_new == 0) break;
        for (uint32_t p = 0; p < M; ++p) {
            uint32_t g = gang_id[p];
            if (g == ORC_NONE_U32 || !newly[g]) continue;
            if (assign[p] >= 0) { /* release the reservation */
                fg[assign[p]] += pods->req_gpu[p];
                fc[assign[p]] += pods->req_mcpu[p];
                fm[assign[p]] += pods->req_mib[p];
            }
            assign[p] = ORC_ASSIGN_GANG;
        }
        na = 0; /* step 3: retry what is still unplaced and not rejected */
        for (uint32_t p = 0; p < M; ++p) {
            uint32_t g = gang_id[p];
            if (assign[p] == ORC_ASSIGN_NONE && (g == ORC_NONE_U32 || !rejected[g])) {
                act[na++] = p;
                assign[p] = ORC_ASSIGN_PENDING;
            }
        }
        if (na == 0) break;
    }
    uint32_t placed = 0;
    for (uint32_t p = 0; p < M; ++p) placed += assign[p] >= 0;
    for (uint32_t g = 0; g < n_gangs; ++g) gang_ok[g] = rejected[g] ? 0 : 1;
    if (stats) { stats->rounds = rounds; stats->placed = placed; stats->evals = evals; }
    if (iterations) *iterations = iters;
    free(act); free(nxt); free(best); free(count); free(rejected); free(newly);
    return 0;
}

/* ------------------------------------------------------------------ batch reconcile */
typedef struct {
    const orc_vm_table *vms;
    const int32_t *lut;
    uint32_t n_owner_ids, lo, hi;
    int32_t *pool_of;
} match_job;

/* README.md:238 -- managed-by must match AND owner must be this pool's "<namespace>-<name>" */
static void *match_worker(void *arg) {
    match_job *j = (match_job *)arg;
    for (uint32_t v = j->lo; v < j->hi; ++v) {
        int32_t pool = -1;
        uint32_t o = j->vms->owner_id[v];
        if ((j->vms->flags[v] & 1u) && o != ORC_NONE_U32 && o < j->n_owner_ids) pool = j->lut[o];
        j->pool_of[v] = pool;
    }
    return NULL;
}

int orc_reconcile_batch(const orc_pool_table *pools, const orc_vm_table *vms, uint32_t n_owner_ids,
                        int nthreads, orc_reconcile_out *out) {
    const uint32_t P = pools->p, V = vms->v;
    int32_t *lut = (int32_t *)malloc(((size_t)n_owner_ids + 1) * 4);
    for (uint32_t i = 0; i < n_owner_ids; ++i) lut[i] = -1;
    for (uint32_t p = 0; p < P; ++p) {
        if (pools->desired[p] < 0) { free(lut); return -1; }
        uint32_t o = pools->owner_id[p];
        if (o == ORC_NONE_U32 || o >= n_owner_ids) continue; /* pool can never match a VM */
        if (lut[o] != -1) { free(lut); return -1; }          /* owner id maps to <= 1 pool */
        lut[o] = (int32_t)p;
    }
    /* a2: tag filter (the only part worth threading) */
    if (nthreads < 1) nthreads = 1;
    if ((uint32_t)nthreads > V) nthreads = V ? (int)V : 1;
    match_job *jobs = (match_job *)calloc((size_t)nthreads, sizeof(match_job));
    pthread_t *tid = (pthread_t *)calloc((size_t)nthreads, sizeof(pthread_t));
    for (int t = 0; t < nthreads; ++t) {
        jobs[t].vms = vms; jobs[t].lut = lut; jobs[t].n_owner_ids = n_owner_ids;
        jobs[t].pool_of = out->pool_of;
        jobs[t].lo = (uint32_t)((uint64_t)V * (uint64_t)t / (uint64_t)nthreads);
        jobs[t].hi = (uint32_t)((uint64_t)V * (uint64_t)(t + 1) / (uint64_t)nthreads);
    }
    if (nthreads == 1) match_worker(&jobs[0]);
    else {
        for (int t = 0; t < nthreads; ++t) pthread_create(&tid[t], NULL, match_worker, &jobs[t]);
        for (int t = 0; t < nthreads; ++t) pthread_join(tid[t], NULL);
    }
    free(jobs); free(tid); free(lut);

    /* a3: currentCount := len(existingVMs)  (README.md:195) */
    for (uint32_t p = 0; p < P; ++p) out->current[p] = 0;
    for (uint32_t v = 0; v < V; ++v)
        if (out->pool_of[v] >= 0) out->current[out->pool_of[v]]++;

    /* a4, a6, a7 */
    int64_t vtotal = 0, mtotal = 0, ctotal = 0;
    for (uint32_t p = 0; p < P; ++p) {
        int32_t cur = out->current[p], des = pools->desired[p];
        out->delta[p] = des - cur;                       /* README.md:201 / :212 */
        out->ready[p] = cur;                             /* README.md:225, pre-action */
        out->create_count[p] = des > cur ? des - cur : 0;
        out->victim_off[p] = (int32_t)vtotal;
        vtotal += cur > des ? cur - des : 0;
        if (out->member_off) { out->member_off[p] = (int32_t)mtotal; mtotal += cur; }
        if (out->create_off) { out->create_off[p] = (int32_t)ctotal; ctotal += out->create_count[p]; }
    }
    out->victim_off[P] = (int32_t)vtotal;
    if (out->member_off) out->member_off[P] = (int32_t)mtotal;
    if (out->create_off) out->create_off[P] = (int32_t)ctotal;

    /* a5: vmsToDelete := existingVMs[:scaleDownCount]  (README.md:214); list order = row order */
    int32_t *moff = (int32_t *)malloc(((size_t)P + 1) * 4); /* existingVMs of every pool, in list order */
    uint32_t *mvm = (uint32_t *)malloc(((size_t)V + 1) * 4);
    int32_t *seen = (int32_t *)calloc((size_t)P + 1, 4);
    moff[0] = 0;
    for (uint32_t p = 0; p < P; ++p) moff[p + 1] = moff[p] + out->current[p];
    for (uint32_t v = 0; v < V; ++v) {
        int32_t p = out->pool_of[v];
        if (p >= 0) mvm[moff[p] + seen[p]++] = v;
    }
    for (uint32_t p = 0; p < P; ++p) {
        int32_t k = out->victim_off[p + 1] - out->victim_off[p];
        for (int32_t r = 0; r < k; ++r) out->victim_vm[out->victim_off[p] + r] = mvm[moff[p] + r];
        if (out->member_vm)
            for (int32_t r = 0; r < out->current[p]; ++r) out->member_vm[out->member_off[p] + r] = mvm[moff[p] + r];
    }
    free(seen);

    /* a6: new VM names "<pool-name>-<ordinal>": the lowest ordinals not in use, ascending */
    if (out->create_off && out->create_ordinal) {
        for (uint32_t p = 0; p < P; ++p) {
            int32_t c = out->create_count[p];
            if (c == 0) continue;
            uint32_t span = (uint32_t)out->current[p] + (uint32_t)c; /* answers lie in [0, span) */
            uint8_t *used = (uint8_t *)calloc((size_t)span + 1, 1);
            if (vms->ordinal)
                for (int32_t r = 0; r < out->current[p]; ++r) {
                    uint32_t o = vms->ordinal[mvm[moff[p] + r]];
                    if (o < span) used[o] = 1;
                }
            int32_t w = 0;
            for (uint32_t o = 0; o < span && w < c; ++o)
                if (!used[o]) out->create_ordinal[out->create_off[p] + w++] = o;
            free(used);
        }
    }
    free(moff);
    free(mvm);
    return 0;
}
