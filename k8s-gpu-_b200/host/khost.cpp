// khost.cpp -- C entry points over the C++ host mirror so that tests (Python ctypes) can drive
// AzureVmPoolReconciler and BatchReconciler the way the reference's envtest-style tests would:
// put CRs, seed the fake resource group, reconcile, read back ARM state and status.
#include <cstring>
#include <memory>

#include "reconciler.hpp"

using namespace k8sgpu_host;

struct khost_world {
    FakeApiServer api;
    FakeArmClient arm;
    AzureVmPoolReconciler rec{&api, &arm};
    k8sgpu_ctx *gpu = nullptr;
    std::unique_ptr<BatchReconciler> batch;
    std::string scratch;
};

extern "C" {

khost_world *khost_new() { return new khost_world(); }
void khost_free(khost_world *w) {
    if (!w) return;
    w->batch.reset();
    if (w->gpu) k8sgpu_destroy(w->gpu);
    delete w;
}

int khost_put_pool(khost_world *w, const char *ns, const char *name, int32_t replicas, const char *rg) {
    if (replicas < 0) return -1;  // +kubebuilder:validation:Minimum=0 (README.md:94)
    computev1alpha1::AzureVmPool p;
    p.Metadata.Namespace = ns;
    p.Metadata.Name = name;
    p.Spec.Replicas = replicas;
    p.Spec.ResourceGroupName = rg;
    p.Spec.Location = "eastus";                 // sample CR, README.md:275-276
    p.Spec.VMSize = "Standard_NC4as_T4_v3";
    p.Spec.AzureCredentialSecret = "azure-credentials";
    w->api.Put(p);
    return 0;
}
int khost_delete_pool(khost_world *w, const char *ns, const char *name) { return w->api.Delete(ns, name) ? 0 : -1; }

int khost_add_vm(khost_world *w, const char *rg, const char *name, const char *managed_by, const char *owner) {
    VM vm;
    vm.Name = name;
    if (managed_by) vm.Tags[kManagedByKey] = managed_by;
    if (owner) vm.Tags[kOwnerKey] = owner;
    w->arm.Seed(rg, vm);
    return 0;
}
// kind: 0 client, 1 list, 2 create, 3 delete (k-th upcoming call fails), 4 status update (on/off)
void khost_fail(khost_world *w, int kind, int k) {
    if (kind == 4) w->api.FailStatus(k != 0);
    else w->arm.FailAt(static_cast<FailKind>(kind), k);
}
int khost_calls(khost_world *w, int kind) { return w->arm.Calls(static_cast<FailKind>(kind)); }
int khost_status_writes(khost_world *w) { return w->api.StatusWrites(); }

int khost_vm_count(khost_world *w, const char *rg) { return static_cast<int>(w->arm.Group(rg).size()); }
const char *khost_vm_name(khost_world *w, const char *rg, int i) {
    const auto &g = w->arm.Group(rg);
    return (i >= 0 && static_cast<size_t>(i) < g.size()) ? g[static_cast<size_t>(i)].Name.c_str() : nullptr;
}
int khost_status_ready(khost_world *w, const char *ns, const char *name) {
    auto *p = w->api.Get(ns, name);
    return p ? p->Status.ReadyReplicas : -1;
}
int khost_status_vm_count(khost_world *w, const char *ns, const char *name) {
    auto *p = w->api.Get(ns, name);
    return p ? static_cast<int>(p->Status.VMs.size()) : -1;
}
const char *khost_status_vm(khost_world *w, const char *ns, const char *name, int i) {
    auto *p = w->api.Get(ns, name);
    return (p && i >= 0 && static_cast<size_t>(i) < p->Status.VMs.size()) ? p->Status.VMs[static_cast<size_t>(i)].c_str() : nullptr;
}
const char *khost_status_condition(khost_world *w, const char *ns, const char *name) {
    auto *p = w->api.Get(ns, name);
    if (!p || p->Status.Conditions.empty()) return nullptr;
    w->scratch = p->Status.Conditions[0].Type + "=" + p->Status.Conditions[0].Status + ":" + p->Status.Conditions[0].Reason;
    return w->scratch.c_str();
}

// every condition of the pool's status, "Type=Status:Reason" joined by ';'
const char *khost_status_conditions(khost_world *w, const char *ns, const char *name) {
    auto *p = w->api.Get(ns, name);
    if (!p) return nullptr;
    w->scratch.clear();
    for (const auto &c : p->Status.Conditions) {
        if (!w->scratch.empty()) w->scratch += ";";
        w->scratch += c.Type + "=" + c.Status + ":" + c.Reason;
    }
    return w->scratch.c_str();
}

// the reference's entry point for one pool (host only)
int khost_reconcile(khost_world *w, const char *ns, const char *name, int64_t *requeue_s, int *err) {
    auto [res, e] = w->rec.Reconcile(ctrl::Request{ns, name});
    *requeue_s = res.RequeueAfter.count();
    *err = static_cast<int>(e.kind);
    return 0;
}

// every pool named in (ns[i], name[i]) in one pass through the GPU; returns a k8sgpu error code
int khost_reconcile_batch(khost_world *w, int device, int n, const char *const *ns, const char *const *name,
                          int64_t *requeue_s, int *err) {
    if (!w->gpu) {
        int rc = k8sgpu_create(&device, 1, &w->gpu);
        if (rc != K8SGPU_OK) return rc;  // no GPU: an error, never a CPU fallback
        w->batch = std::make_unique<BatchReconciler>(&w->api, &w->arm, w->gpu);
    }
    std::vector<ctrl::Request> reqs;
    for (int i = 0; i < n; ++i) reqs.push_back(ctrl::Request{ns[i], name[i]});
    auto out = w->batch->ReconcileAll(reqs);
    for (int i = 0; i < n; ++i) {
        requeue_s[i] = out[static_cast<size_t>(i)].result.RequeueAfter.count();
        err[i] = static_cast<int>(out[static_cast<size_t>(i)].err.kind);
    }
    return K8SGPU_OK;
}
uint64_t khost_last_gpu_launches(khost_world *w) { return w->batch ? w->batch->LastStats().kernel_launches : 0; }

}  // extern "C"

// ------------------------------------------------------------------ snapshot ingestion (f2)
#include "snapshot.hpp"

struct khost_snap {
    SnapshotBuilder b;
    ReconcileTables rec;
    NodeTable nodes;
    PodTable pods;
    std::string err;
};

extern "C" {

khost_snap *khost_snap_new() { return new khost_snap(); }
void khost_snap_free(khost_snap *s) { delete s; }
const char *khost_snap_error(khost_snap *s) { return s->err.c_str(); }

static int snap_result(khost_snap *s, bool ok, const ParseError &e) {
    s->err = ok ? "" : (e.msg + " (line " + std::to_string(e.line) + ")");
    return ok ? 0 : -1;
}
int khost_snap_add_pools_yaml(khost_snap *s, const char *text) {
    ParseError e;
    return snap_result(s, s->b.AddPoolsYaml(text, &e), e);
}
int khost_snap_add_vms_json(khost_snap *s, const char *text) {
    ParseError e;
    return snap_result(s, s->b.AddVmsJson(text, &e), e);
}
int khost_snap_add_jobs_yaml(khost_snap *s, const char *text) {
    ParseError e;
    return snap_result(s, s->b.AddJobsYaml(text, &e), e);
}
void khost_snap_register_sku(khost_snap *s, const char *vm_size, uint32_t gpu, uint32_t mcpu, uint32_t mib) {
    s->b.RegisterSku(vm_size, gpu, mcpu, mib);
}
void khost_snap_build(khost_snap *s) {
    s->rec = s->b.BuildReconcileTables();
    s->nodes = s->b.BuildNodeTable();  // interns the zones the pod table then refers to
    s->pods = s->b.BuildPodTable();
}
// table: 0 pools, 1 vms, 2 nodes, 3 pods
uint32_t khost_snap_rows(khost_snap *s, int table) {
    switch (table) {
        case 0: return static_cast<uint32_t>(s->rec.desired.size());
        case 1: return static_cast<uint32_t>(s->rec.vm_owner.size());
        case 2: return static_cast<uint32_t>(s->nodes.flags.size());
        default: return static_cast<uint32_t>(s->pods.req_gpu.size());
    }
}
uint32_t khost_snap_n_owner_ids(khost_snap *s) { return s->rec.n_owner_ids; }
// columns in the order of include/k8sgpu.h; pools: desired (as i32 bits), owner_id
const uint32_t *khost_snap_col(khost_snap *s, int table, int col) {
    const std::vector<uint32_t> *v = nullptr;
    if (table == 0) {
        if (col == 0) return reinterpret_cast<const uint32_t *>(s->rec.desired.data());
        v = &s->rec.pool_owner;
    } else if (table == 1) {
        v = col == 0 ? &s->rec.vm_owner : col == 1 ? &s->rec.vm_flags : &s->rec.vm_ordinal;
    } else if (table == 2) {
        const std::vector<uint32_t> *c[8] = {&s->nodes.free_gpu, &s->nodes.free_mcpu, &s->nodes.free_mib, &s->nodes.cap_gpu,
                                             &s->nodes.sku_id,   &s->nodes.zone_id,   &s->nodes.taint_mask, &s->nodes.flags};
        v = c[col & 7];
    } else {
        const std::vector<uint32_t> *c[8] = {&s->pods.req_gpu,  &s->pods.req_mcpu, &s->pods.req_mib,   &s->pods.sku_mask,
                                             &s->pods.zone_mask, &s->pods.tol_mask, &s->pods.pref_zone, &s->pods.priority};
        v = c[col & 7];
    }
    return v->data();
}
const char *khost_snap_name(khost_snap *s, int table, uint32_t row) {
    const std::vector<std::string> *v = table == 0 ? &s->rec.pool_key : table == 1 ? &s->rec.vm_name
                                        : table == 2 ? &s->nodes.name : &s->pods.name;
    return row < v->size() ? (*v)[row].c_str() : nullptr;
}
// i-th pool left out of the reconcile tables because its owner tag is already claimed:
// "<dropped ns/name>|<ns/name of the pool that keeps the tag>", nullptr past the end
const char *khost_snap_conflict(khost_snap *s, uint32_t i) {
    if (i >= s->rec.conflicts.size()) return nullptr;
    s->err = s->rec.conflicts[i].first + "|" + s->rec.conflicts[i].second;  // (err doubles as the string scratch)
    return s->err.c_str();
}
const char *khost_snap_pool_field(khost_snap *s, uint32_t i, const char *field) {
    const auto &pools = s->b.Pools();
    if (i >= pools.size()) return nullptr;
    const auto &p = pools[i];
    std::string f = field;
    const std::string *v = f == "name" ? &p.Metadata.Name : f == "namespace" ? &p.Metadata.Namespace
                           : f == "resourceGroupName" ? &p.Spec.ResourceGroupName : f == "location" ? &p.Spec.Location
                           : f == "vmSize" ? &p.Spec.VMSize : f == "vnetName" ? &p.Spec.VnetName
                           : f == "subnetName" ? &p.Spec.SubnetName : f == "azureCredentialSecret" ? &p.Spec.AzureCredentialSecret
                           : f == "imageReference.publisher" ? &p.Spec.ImageReference.Publisher
                           : f == "imageReference.offer" ? &p.Spec.ImageReference.Offer
                           : f == "imageReference.sku" ? &p.Spec.ImageReference.SKU
                           : f == "imageReference.version" ? &p.Spec.ImageReference.Version : nullptr;
    return v ? v->c_str() : nullptr;
}

}  // extern "C"
