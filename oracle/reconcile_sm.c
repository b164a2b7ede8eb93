/*
 * reconcile_sm.c -- the single-pool Reconcile of README.md:170-235 restated statement by
 * statement in C, against a fake in-memory ARM client and a fake API server.
 * TEST INFRASTRUCTURE ONLY (see k8s_oracle.h).  PARITY UNPINNED (reference has no code/tests).
 *
 * The four helpers the reference names but never defines (README.md:181,189,205,217) follow
 * the prose contract at README.md:237-240: tags managed-by / owner, strict tag filter on
 * list, unique name per create, idempotent create/delete.  List order is insertion order.
 */
#include "k8s_oracle.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define NAME_MAX_ 128
#define MANAGED_BY_VALUE "azurevmpool-operator" /* README.md:238 */

typedef struct {
    char name[NAME_MAX_];
    char managed_by[NAME_MAX_];
    char owner[NAME_MAX_];
    int has_managed_by, has_owner;
} arm_vm;

struct orc_arm {
    arm_vm *vms;
    int n, cap;
    int fail_list_at, fail_create_at, fail_delete_at, fail_client;
    int calls_list, calls_create, calls_delete;
};

typedef struct {
    char ns[NAME_MAX_], name[NAME_MAX_];
    int32_t replicas;
    int32_t ready_replicas;
    char (*status_vms)[NAME_MAX_];
    int n_status_vms;
    int live;
} pool_obj;

struct orc_cluster {
    pool_obj *pools;
    int n, cap;
    int fail_status;
    int status_writes;
};

/* ---------------------------------------------------------------- fake ARM */
orc_arm *orc_arm_new(void) { return (orc_arm *)calloc(1, sizeof(orc_arm)); }
void orc_arm_free(orc_arm *a) {
    if (!a) return;
    free(a->vms);
    free(a);
}
static void copy_str(char *dst, const char *src) {
    size_t n = strlen(src);
    if (n > NAME_MAX_ - 1) n = NAME_MAX_ - 1;
    memcpy(dst, src, n);
    dst[n] = 0;
}
int orc_arm_add_vm(orc_arm *a, const char *name, const char *managed_by, const char *owner) {
    if (a->n == a->cap) {
        a->cap = a->cap ? a->cap * 2 : 16;
        a->vms = (arm_vm *)realloc(a->vms, (size_t)a->cap * sizeof(arm_vm));
    }
    arm_vm *v = &a->vms[a->n];
    memset(v, 0, sizeof(*v));
    copy_str(v->name, name);
    if (managed_by) { copy_str(v->managed_by, managed_by); v->has_managed_by = 1; }
    if (owner) { copy_str(v->owner, owner); v->has_owner = 1; }
    return a->n++;
}
int orc_arm_vm_count(const orc_arm *a) { return a->n; }
const char *orc_arm_vm_name(const orc_arm *a, int i) { return (i >= 0 && i < a->n) ? a->vms[i].name : NULL; }
void orc_arm_fail_list_at(orc_arm *a, int k) { a->fail_list_at = k; }
void orc_arm_fail_create_at(orc_arm *a, int k) { a->fail_create_at = k; }
void orc_arm_fail_delete_at(orc_arm *a, int k) { a->fail_delete_at = k; }
void orc_arm_fail_client(orc_arm *a, int on) { a->fail_client = on; }
int orc_arm_calls_list(const orc_arm *a) { return a->calls_list; }
int orc_arm_calls_create(const orc_arm *a) { return a->calls_create; }
int orc_arm_calls_delete(const orc_arm *a) { return a->calls_delete; }

static int countdown(int *at) { /* 1 = this call fails */
    if (*at <= 0) return 0;
    if (--*at == 0) return 1;
    return 0;
}

/* ---------------------------------------------------------------- fake API server */
orc_cluster *orc_cluster_new(void) { return (orc_cluster *)calloc(1, sizeof(orc_cluster)); }
void orc_cluster_free(orc_cluster *c) {
    if (!c) return;
    for (int i = 0; i < c->n; ++i) free(c->pools[i].status_vms);
    free(c->pools);
    free(c);
}
static pool_obj *find_pool(const orc_cluster *c, const char *ns, const char *name) {
    for (int i = 0; i < c->n; ++i)
        if (c->pools[i].live && !strcmp(c->pools[i].ns, ns) && !strcmp(c->pools[i].name, name))
            return &c->pools[i];
    return NULL;
}
int orc_cluster_put_pool(orc_cluster *c, const char *ns, const char *name, int32_t replicas) {
    if (replicas < 0) return -1; /* +kubebuilder:validation:Minimum=0, README.md:94 */
    if (find_pool(c, ns, name)) return orc_cluster_set_replicas(c, ns, name, replicas);
    if (c->n == c->cap) {
        c->cap = c->cap ? c->cap * 2 : 8;
        c->pools = (pool_obj *)realloc(c->pools, (size_t)c->cap * sizeof(pool_obj));
    }
    pool_obj *p = &c->pools[c->n++];
    memset(p, 0, sizeof(*p));
    copy_str(p->ns, ns);
    copy_str(p->name, name);
    p->replicas = replicas;
    p->live = 1;
    return 0;
}
int orc_cluster_set_replicas(orc_cluster *c, const char *ns, const char *name, int32_t replicas) {
    pool_obj *p = find_pool(c, ns, name);
    if (!p || replicas < 0) return -1;
    p->replicas = replicas;
    return 0;
}
int orc_cluster_delete_pool(orc_cluster *c, const char *ns, const char *name) {
    pool_obj *p = find_pool(c, ns, name);
    if (!p) return -1;
    p->live = 0;
    return 0;
}
void orc_cluster_fail_status_update(orc_cluster *c, int on) { c->fail_status = on; }
int32_t orc_cluster_ready_replicas(const orc_cluster *c, const char *ns, const char *name) {
    pool_obj *p = find_pool(c, ns, name);
    return p ? p->ready_replicas : -1;
}
int orc_cluster_status_vm_count(const orc_cluster *c, const char *ns, const char *name) {
    pool_obj *p = find_pool(c, ns, name);
    return p ? p->n_status_vms : -1;
}
const char *orc_cluster_status_vm(const orc_cluster *c, const char *ns, const char *name, int i) {
    pool_obj *p = find_pool(c, ns, name);
    return (p && i >= 0 && i < p->n_status_vms) ? p->status_vms[i] : NULL;
}
int orc_cluster_status_writes(const orc_cluster *c) { return c->status_writes; }

/* ---------------------------------------------------------------- the four helpers */
static void owner_tag(const pool_obj *p, char *out) { /* "<namespace>-<name>", README.md:238 */
    snprintf(out, NAME_MAX_, "%s-%s", p->ns, p->name);
}

/* listManagedVMs, README.md:189 + :238: strict filter on BOTH tags; returns indices. */
static int list_managed(orc_arm *a, const pool_obj *p, int **out_idx, int *out_n) {
    a->calls_list++;
    if (countdown(&a->fail_list_at)) return -1;
    char owner[NAME_MAX_];
    owner_tag(p, owner);
    int *idx = (int *)malloc(((size_t)a->n + 1) * sizeof(int));
    int k = 0;
    for (int i = 0; i < a->n; ++i) {
        const arm_vm *v = &a->vms[i];
        if (v->has_managed_by && !strcmp(v->managed_by, MANAGED_BY_VALUE) && v->has_owner &&
            !strcmp(v->owner, owner))
            idx[k++] = i;
    }
    *out_idx = idx;
    *out_n = k;
    return 0;
}

/* createVM, README.md:205: unique name "<pool-name>-<ordinal>", lowest ordinal not in use by
 * this pool's VMs; both tags set; create-if-absent (idempotent, :240). */
static int create_vm(orc_arm *a, const pool_obj *p) {
    a->calls_create++;
    if (countdown(&a->fail_create_at)) return -1;
    char owner[NAME_MAX_], name[NAME_MAX_ + 32];
    owner_tag(p, owner);
    for (uint32_t ord = 0;; ++ord) {
        snprintf(name, sizeof name, "%s-%u", p->name, ord);
        int taken = 0;
        for (int i = 0; i < a->n && !taken; ++i) {
            const arm_vm *v = &a->vms[i];
            if (v->has_managed_by && !strcmp(v->managed_by, MANAGED_BY_VALUE) && v->has_owner &&
                !strcmp(v->owner, owner) && !strcmp(v->name, name))
                taken = 1;
        }
        if (!taken) break;
    }
    orc_arm_add_vm(a, name, MANAGED_BY_VALUE, owner);
    return 0;
}

/* deleteVM, README.md:217: by name; VM + NIC + OS disk go together (:239); delete of a
 * missing VM is a no-op success (:240). */
static int delete_vm(orc_arm *a, const pool_obj *p, const char *name) {
    (void)p;
    a->calls_delete++;
    if (countdown(&a->fail_delete_at)) return -1;
    for (int i = 0; i < a->n; ++i)
        if (!strcmp(a->vms[i].name, name)) {
            memmove(&a->vms[i], &a->vms[i + 1], (size_t)(a->n - i - 1) * sizeof(arm_vm));
            a->n--;
            break;
        }
    return 0;
}

/* ---------------------------------------------------------------- Reconcile, README.md:170-235 */
orc_result orc_reconcile_one(orc_cluster *c, orc_arm *a, const char *ns, const char *name) {
    orc_result r = {0, ORC_OK};

    /* 1. r.Get -- NotFound is swallowed: ({}, nil), README.md:175-177 */
    pool_obj *pool = find_pool(c, ns, name);
    if (!pool) return r;

    /* 2. getAzureVMClient, README.md:181-185 */
    if (a->fail_client) {
        r.requeue_after_s = 30;
        r.err = ORC_ERR_CLIENT;
        return r;
    }

    /* 3. listManagedVMs, README.md:189-193 */
    int *existing = NULL, n_existing = 0;
    if (list_managed(a, pool, &existing, &n_existing) != 0) {
        r.requeue_after_s = 20;
        r.err = ORC_ERR_LIST;
        return r;
    }
    /* names are captured now: the slice the reference ranges over is fixed at list time */
    char(*names)[NAME_MAX_] = (char(*)[NAME_MAX_])malloc(((size_t)n_existing + 1) * NAME_MAX_);
    for (int i = 0; i < n_existing; ++i) copy_str(names[i], a->vms[existing[i]].name);
    free(existing);

    int32_t currentCount = (int32_t)n_existing; /* README.md:195 */
    int32_t desiredCount = pool->replicas;      /* README.md:196 */

    /* 4. scale up or down, README.md:199-222 */
    if (currentCount < desiredCount) {
        int32_t scaleUpCount = desiredCount - currentCount; /* :201 */
        for (int32_t i = 0; i < scaleUpCount; ++i) {
            if (create_vm(a, pool) != 0) { /* :205-208: return mid-loop, earlier creates stand */
                free(names);
                r.requeue_after_s = 40;
                r.err = ORC_ERR_CREATE;
                return r;
            }
        }
    } else if (currentCount > desiredCount) {
        int32_t scaleDownCount = currentCount - desiredCount; /* :212 */
        /* vmsToDelete := existingVMs[:scaleDownCount], :214 */
        for (int32_t i = 0; i < scaleDownCount; ++i) {
            if (delete_vm(a, pool, names[i]) != 0) { /* :217-220 */
                free(names);
                r.requeue_after_s = 40;
                r.err = ORC_ERR_DELETE;
                return r;
            }
        }
    }

    /* 5. status: ReadyReplicas = currentCount, the PRE-action count, README.md:225.
     *    Status.VMs = the listed names (README.md:226 is elided). */
    if (c->fail_status) { /* r.Status().Update failed, README.md:227-230: ({}, err) */
        free(names);
        r.err = ORC_ERR_STATUS;
        return r;
    }
    pool->ready_replicas = currentCount;
    free(pool->status_vms);
    pool->status_vms = names;
    pool->n_status_vms = n_existing;
    c->status_writes++;

    r.requeue_after_s = 60; /* README.md:234 */
    return r;
}

/* BASELINE.json config 1 as a timing loop (bench.py cpu leg): a fresh fake cluster per iteration,
 * 3 VMs, replicas 3 -> 5 -> 2 with a settle Reconcile after each change (K1..K5: six Reconciles).
 * Returns the number of Reconcile calls made, or -1 if a step did not end as K1..K5 say. */
long orc_config1_loop(int iterations) {
    long calls = 0;
    for (int it = 0; it < iterations; ++it) {
        orc_arm *a = orc_arm_new();
        orc_cluster *c = orc_cluster_new();
        if (!a || !c) return -1;
        orc_cluster_put_pool(c, "default", "pool", 3);
        orc_arm_add_vm(a, "v0", "azurevmpool-operator", "default-pool");
        orc_arm_add_vm(a, "v1", "azurevmpool-operator", "default-pool");
        orc_arm_add_vm(a, "v2", "azurevmpool-operator", "default-pool");
        static const int32_t seq[3] = {3, 5, 2};
        static const int want_vms[3] = {3, 5, 2};
        for (int s = 0; s < 3; ++s) {
            orc_cluster_set_replicas(c, "default", "pool", seq[s]);
            for (int k = 0; k < 2; ++k) { /* the change, then the settle pass */
                orc_result r = orc_reconcile_one(c, a, "default", "pool");
                ++calls;
                if (r.err != ORC_OK || r.requeue_after_s != 60) return -1;
            }
            if (orc_arm_vm_count(a) != want_vms[s] || orc_cluster_ready_replicas(c, "default", "pool") != want_vms[s])
                return -1;
        }
        orc_arm_free(a);
        orc_cluster_free(c);
    }
    return calls;
}
