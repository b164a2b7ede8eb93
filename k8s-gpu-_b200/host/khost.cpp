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
