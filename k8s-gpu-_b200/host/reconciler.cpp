// reconciler.cpp -- see reconciler.hpp.  Citations: /root/reference/README.md.
#include "reconciler.hpp"

#include <map>

namespace k8sgpu_host {

using computev1alpha1::AzureVmPool;
using std::chrono::seconds;

bool AzureVmPoolReconciler::IsManagedBy(const VM &vm, const std::string &owner) {
    auto mb = vm.Tags.find(kManagedByKey);
    auto ow = vm.Tags.find(kOwnerKey);
    return mb != vm.Tags.end() && mb->second == kManagedByValue && ow != vm.Tags.end() && ow->second == owner;
}

std::optional<uint32_t> AzureVmPoolReconciler::OrdinalOf(const std::string &vm_name, const std::string &pool_name) {
    if (vm_name.size() <= pool_name.size() + 1 || vm_name.compare(0, pool_name.size(), pool_name) != 0 ||
        vm_name[pool_name.size()] != '-')
        return std::nullopt;
    uint64_t v = 0;
    size_t digits = 0;
    for (size_t i = pool_name.size() + 1; i < vm_name.size(); ++i, ++digits) {
        char c = vm_name[i];
        if (c < '0' || c > '9') return std::nullopt;
        v = v * 10 + static_cast<uint64_t>(c - '0');
        if (v >= 0xFFFFFFFFull) return std::nullopt;
    }
    if (digits > 1 && vm_name[pool_name.size() + 1] == '0') return std::nullopt;  // "-007" is not what createVM writes
    return static_cast<uint32_t>(v);
}

// README.md:181 -- reads the credential Secret; the fake only knows "works" or "fails".
std::pair<FakeArmClient *, Error> AzureVmPoolReconciler::getAzureVMClient(const AzureVmPool &) {
    if (arm_->ClientFails()) return {nullptr, Error{ErrKind::Client, "cannot build the Azure VM client"}};
    return {arm_, Error{}};
}

// README.md:189 + :238 -- strict filter on BOTH tags.
std::pair<std::vector<VM>, Error> AzureVmPoolReconciler::listManagedVMs(FakeArmClient *vmClient, const AzureVmPool &pool) {
    std::vector<VM> all, mine;
    if (auto e = vmClient->List(pool.Spec.ResourceGroupName, &all)) return {{}, Error{ErrKind::List, e->msg}};
    const std::string owner = OwnerTag(pool);
    for (const VM &vm : all)
        if (IsManagedBy(vm, owner)) mine.push_back(vm);
    return {mine, Error{}};
}

Error AzureVmPoolReconciler::createVMNamed(const AzureVmPool &pool, const std::string &name) {
    VM vm;
    vm.Name = name;
    vm.Tags[kManagedByKey] = kManagedByValue;  // README.md:238: both tags on everything we create
    vm.Tags[kOwnerKey] = OwnerTag(pool);
    if (auto e = arm_->Create(pool.Spec.ResourceGroupName, vm)) return Error{ErrKind::Create, e->msg};
    return Error{};
}

// README.md:205 -- unique name per create: "<pool-name>-<k>", k the lowest ordinal not used by
// this pool's VMs.
Error AzureVmPoolReconciler::createVM(const AzureVmPool &pool) {
    const std::string owner = OwnerTag(pool);
    std::map<uint32_t, bool> used;
    for (const VM &vm : arm_->Group(pool.Spec.ResourceGroupName))
        if (IsManagedBy(vm, owner))
            if (auto k = OrdinalOf(vm.Name, pool.Metadata.Name)) used[*k] = true;
    uint32_t k = 0;
    while (used.count(k)) ++k;
    return createVMNamed(pool, pool.Metadata.Name + "-" + std::to_string(k));
}

// README.md:217 -- VM + NIC + OS disk (:239), idempotent (:240).
Error AzureVmPoolReconciler::deleteVM(const AzureVmPool &pool, const std::string &name) {
    if (auto e = arm_->Delete(pool.Spec.ResourceGroupName, name)) return Error{ErrKind::Delete, e->msg};
    return Error{};
}

const char *ErrKindName(ErrKind k) {
    switch (k) {
        case ErrKind::Client: return "ClientFailed";
        case ErrKind::List: return "ListFailed";
        case ErrKind::Create: return "CreateFailed";
        case ErrKind::Delete: return "DeleteFailed";
        case ErrKind::Status: return "StatusUpdateFailed";
        case ErrKind::Gpu: return "GpuFailed";
        case ErrKind::Conflict: return "OwnerTagConflict";
        default: return "None";
    }
}

// Step 5's status (README.md:225-226) with the four condition types README.md:310 proposes:
// Ready, Provisioning, Deleting, Failed -- in that order, all four always present.
static void fill_status(AzureVmPool *pool, int32_t currentCount, const std::vector<std::string> &names, ErrKind prev) {
    pool->Status.ReadyReplicas = currentCount;  // README.md:225: the PRE-action count
    pool->Status.VMs = names;                   // README.md:124-125, :226 (elided in the reference)
    const int32_t desired = pool->Spec.Replicas;
    auto cond = [](const char *type, bool on, const char *reason) {
        computev1alpha1::Condition c;
        c.Type = type;
        c.Status = on ? "True" : "False";
        c.Reason = reason;
        return c;
    };
    pool->Status.Conditions = {
        cond("Ready", currentCount == desired,
             currentCount == desired ? "Reconciled" : (currentCount < desired ? "Provisioning" : "Deleting")),
        cond("Provisioning", currentCount < desired, currentCount < desired ? "ScalingUp" : "NotScalingUp"),
        cond("Deleting", currentCount > desired, currentCount > desired ? "ScalingDown" : "NotScalingDown"),
        // the pass before this one ended in an error (and so wrote no status, README.md:207,219)
        cond("Failed", prev != ErrKind::None, prev != ErrKind::None ? ErrKindName(prev) : "NoError"),
    };
}

ErrKind AzureVmPoolReconciler::lastFailure(const ctrl::Request &req) const {
    auto it = last_failure_.find(std::make_pair(req.Namespace, req.Name));
    return it == last_failure_.end() ? ErrKind::None : it->second;
}

void AzureVmPoolReconciler::remember(const ctrl::Request &req, const Error &e, bool pool_exists) {
    auto key = std::make_pair(req.Namespace, req.Name);
    if (e && pool_exists) last_failure_[key] = e.kind;
    else last_failure_.erase(key);
}

std::pair<ctrl::Result, Error> AzureVmPoolReconciler::Reconcile(const ctrl::Request &req) {
    auto out = reconcileOnce(req, lastFailure(req));
    remember(req, out.second, api_->Get(req.Namespace, req.Name) != nullptr);
    return out;
}

std::pair<ctrl::Result, Error> AzureVmPoolReconciler::reconcileOnce(const ctrl::Request &req, ErrKind prev) {
    // 1. r.Get; NotFound is swallowed (README.md:174-177)
    AzureVmPool *found = api_->Get(req.Namespace, req.Name);
    if (!found) return {ctrl::Result{}, Error{}};
    AzureVmPool azureVmPool = *found;

    // 2. Azure client (README.md:181-185)
    auto [vmClient, cerr] = getAzureVMClient(azureVmPool);
    if (cerr) return {ctrl::Result{seconds(30)}, cerr};

    // 3. VMs this pool manages (README.md:189-193)
    auto [existingVMs, lerr] = listManagedVMs(vmClient, azureVmPool);
    if (lerr) return {ctrl::Result{seconds(20)}, lerr};

    const int32_t currentCount = static_cast<int32_t>(existingVMs.size());  // :195
    const int32_t desiredCount = azureVmPool.Spec.Replicas;                 // :196

    // 4. scale up or down (README.md:199-222)
    if (currentCount < desiredCount) {
        const int32_t scaleUpCount = desiredCount - currentCount;  // :201
        for (int32_t i = 0; i < scaleUpCount; ++i)
            if (Error e = createVM(azureVmPool)) return {ctrl::Result{seconds(40)}, e};  // :205-208
    } else if (currentCount > desiredCount) {
        const int32_t scaleDownCount = currentCount - desiredCount;  // :212
        for (int32_t i = 0; i < scaleDownCount; ++i)                 // vmsToDelete := existingVMs[:scaleDownCount], :214
            if (Error e = deleteVM(azureVmPool, existingVMs[static_cast<size_t>(i)].Name))
                return {ctrl::Result{seconds(40)}, e};  // :217-220
    }

    // 5. status (README.md:225-230)
    std::vector<std::string> names;
    for (const VM &vm : existingVMs) names.push_back(vm.Name);
    fill_status(&azureVmPool, currentCount, names, prev);
    if (!api_->StatusUpdate(azureVmPool)) return {ctrl::Result{}, Error{ErrKind::Status, "status update failed"}};

    return {ctrl::Result{seconds(60)}, Error{}};  // :234
}

// ------------------------------------------------------------------ batch

std::vector<BatchOutcome> BatchReconciler::ReconcileAll(const std::vector<ctrl::Request> &reqs) {
    std::vector<BatchOutcome> out(reqs.size());
    struct Live {
        size_t req;         // first request index of this pool
        AzureVmPool pool;   // working copy
        bool rg_failed = false;
    };
    std::vector<Live> live;
    std::map<std::pair<std::string, std::string>, size_t> seen;  // (ns,name) -> index in live, or npos
    std::map<std::pair<std::string, std::string>, size_t> first_req;  // (ns,name) -> first request naming it
    std::vector<long> alias(reqs.size(), -1);                    // request -> live index
    std::vector<long> same_as(reqs.size(), -1);                  // request -> earlier request with the same outcome
    constexpr size_t npos = static_cast<size_t>(-1);

    // steps 1 and 2 per request
    for (size_t r = 0; r < reqs.size(); ++r) {
        auto key = std::make_pair(reqs[r].Namespace, reqs[r].Name);
        auto it = seen.find(key);
        if (it != seen.end()) {  // same pool again: one pass, same outcome (whatever it was)
            if (it->second != npos) alias[r] = static_cast<long>(it->second);
            else same_as[r] = static_cast<long>(first_req[key]);
            continue;
        }
        first_req[key] = r;
        AzureVmPool *found = api_->Get(reqs[r].Namespace, reqs[r].Name);
        if (!found) {  // README.md:176
            seen[key] = npos;
            remember(reqs[r], Error{}, false);
            continue;
        }
        auto [vmClient, cerr] = getAzureVMClient(*found);
        (void)vmClient;
        if (cerr) {  // README.md:184
            out[r] = BatchOutcome{ctrl::Result{seconds(30)}, cerr};
            seen[key] = npos;
            remember(reqs[r], cerr, true);
            continue;
        }
        seen[key] = live.size();
        alias[r] = static_cast<long>(live.size());
        live.push_back(Live{r, *found});
    }

    // step 3, once per resource group instead of once per pool
    std::map<std::string, std::vector<VM>> listed;
    std::map<std::string, bool> rg_ok;
    for (Live &l : live) {
        const std::string &rg = l.pool.Spec.ResourceGroupName;
        if (!rg_ok.count(rg)) {
            std::vector<VM> vms;
            rg_ok[rg] = !arm_->List(rg, &vms).has_value();
            listed[rg] = std::move(vms);
        }
        l.rg_failed = !rg_ok[rg];
    }

    // tables for the GPU: pool p's owner id is p; a VM gets the id of the pool of ITS resource
    // group whose owner tag it carries, else "foreign"
    std::vector<size_t> pool_of_live;  // live index of table row p
    std::map<std::pair<std::string, std::string>, uint32_t> owner_id;  // (rg, owner tag) -> pool row
    std::vector<int32_t> desired;
    std::vector<uint32_t> pool_owner;
    std::vector<long> conflicts_with(live.size(), -1);
    for (size_t i = 0; i < live.size(); ++i) {
        if (live[i].rg_failed) continue;
        auto key = std::make_pair(live[i].pool.Spec.ResourceGroupName, OwnerTag(live[i].pool));
        auto dup = owner_id.find(key);
        if (dup != owner_id.end()) {
            // "<namespace>-<name>" (README.md:238) is not injective: team-a/pool and team/a-pool share
            // the tag "team-a-pool", so both would claim the same VMs.  The first keeps the tag; this
            // pool gets an explicit error instead of a table row (its VMs cannot be told apart).
            conflicts_with[i] = static_cast<long>(pool_of_live[dup->second]);
            continue;
        }
        owner_id[key] = static_cast<uint32_t>(pool_of_live.size());
        pool_owner.push_back(static_cast<uint32_t>(pool_of_live.size()));
        desired.push_back(live[i].pool.Spec.Replicas);
        pool_of_live.push_back(i);
    }
    const uint32_t P = static_cast<uint32_t>(pool_of_live.size());
    std::vector<uint32_t> vm_owner, vm_flags, vm_ordinal;
    std::vector<const VM *> vm_row;
    for (auto &kv : listed) {
        if (!rg_ok[kv.first]) continue;
        for (const VM &vm : kv.second) {
            uint32_t oid = K8SGPU_NONE_U32, ord = K8SGPU_NONE_U32;
            auto ow = vm.Tags.find(kOwnerKey);
            if (ow != vm.Tags.end()) {
                auto it = owner_id.find(std::make_pair(kv.first, ow->second));
                if (it != owner_id.end()) {
                    oid = it->second;
                    if (auto k = OrdinalOf(vm.Name, live[pool_of_live[oid]].pool.Metadata.Name)) ord = *k;
                } else {
                    oid = P;  // somebody else's owner tag
                }
            }
            auto mb = vm.Tags.find(kManagedByKey);
            vm_owner.push_back(oid);
            vm_flags.push_back(mb != vm.Tags.end() && mb->second == kManagedByValue ? 1u : 0u);
            vm_ordinal.push_back(ord);
            vm_row.push_back(&vm);
        }
    }
    const uint32_t V = static_cast<uint32_t>(vm_row.size());

    // steps 3-5 arithmetic for every pool: ONE call into the CUDA library
    uint64_t create_cap = 1;
    for (int32_t d : desired) create_cap += d > 0 ? static_cast<uint64_t>(d) : 0;
    std::vector<int32_t> pool_of(V + 1), current(P + 1), delta(P + 1), ready(P + 1), create_count(P + 1),
        victim_off(P + 2), member_off(P + 2), create_off(P + 2);
    std::vector<uint32_t> victim_vm(V + 1), member_vm(V + 1), create_ordinal(create_cap);
    k8sgpu_pool_table pools{P, desired.data(), pool_owner.data()};
    k8sgpu_vm_table vms{V, vm_owner.data(), vm_flags.data(), vm_ordinal.data()};
    k8sgpu_reconcile_out ro{pool_of.data(),    current.data(),    delta.data(),      ready.data(),
                            create_count.data(), victim_off.data(), victim_vm.data(),  member_off.data(),
                            member_vm.data(),  create_off.data(), create_ordinal.data(), create_cap};
    int rc = K8SGPU_OK;
    if (P > 0) {
        rc = k8sgpu_reconcile_batch(gpu_, &pools, &vms, P, &ro);
        k8sgpu_get_stats(gpu_, &stats_);
    }

    // per pool: the ARM calls and the status write, README.md:203-230
    std::vector<BatchOutcome> live_out(live.size());
    std::vector<long> row_of_live(live.size(), -1);
    for (uint32_t p = 0; p < P; ++p) row_of_live[pool_of_live[p]] = p;
    for (size_t i = 0; i < live.size(); ++i) {
        Live &l = live[i];
        if (l.rg_failed) {  // README.md:192
            live_out[i] = BatchOutcome{ctrl::Result{seconds(20)}, Error{ErrKind::List, "ARM list failed"}};
            continue;
        }
        if (conflicts_with[i] >= 0) {
            const AzureVmPool &o = live[static_cast<size_t>(conflicts_with[i])].pool;
            live_out[i] = BatchOutcome{ctrl::Result{seconds(30)},
                                       Error{ErrKind::Conflict, "owner tag '" + OwnerTag(l.pool) + "' is already claimed by pool " +
                                                                    o.Metadata.Namespace + "/" + o.Metadata.Name}};
            continue;
        }
        if (rc != K8SGPU_OK) {  // no CPU path: report, requeue like a client failure
            live_out[i] = BatchOutcome{ctrl::Result{seconds(30)}, Error{ErrKind::Gpu, k8sgpu_last_error(gpu_)}};
            continue;
        }
        const size_t p = static_cast<size_t>(row_of_live[i]);
        Error err;
        for (int32_t j = create_off[p]; j < create_off[p + 1] && !err; ++j)  // :203-209
            err = createVMNamed(l.pool, l.pool.Metadata.Name + "-" + std::to_string(create_ordinal[static_cast<size_t>(j)]));
        for (int32_t j = victim_off[p]; j < victim_off[p + 1] && !err; ++j)  // :214-221
            err = deleteVM(l.pool, vm_row[victim_vm[static_cast<size_t>(j)]]->Name);
        if (err) {
            live_out[i] = BatchOutcome{ctrl::Result{seconds(40)}, err};  // :207, :219: status NOT written
            continue;
        }
        std::vector<std::string> names;
        for (int32_t j = member_off[p]; j < member_off[p + 1]; ++j) names.push_back(vm_row[member_vm[static_cast<size_t>(j)]]->Name);
        fill_status(&l.pool, ready[p], names, lastFailure(reqs[l.req]));  // :225
        if (!api_->StatusUpdate(l.pool)) {
            live_out[i] = BatchOutcome{ctrl::Result{}, Error{ErrKind::Status, "status update failed"}};  // :229
            continue;
        }
        live_out[i] = BatchOutcome{ctrl::Result{seconds(60)}, Error{}};  // :234
    }
    for (size_t i = 0; i < live.size(); ++i) remember(reqs[live[i].req], live_out[i].err, true);
    for (size_t r = 0; r < reqs.size(); ++r) {
        if (alias[r] >= 0) out[r] = live_out[static_cast<size_t>(alias[r])];
        else if (same_as[r] >= 0) out[r] = out[static_cast<size_t>(same_as[r])];  // e.g. the client failure of the first one
    }
    return out;
}

}  // namespace k8sgpu_host
