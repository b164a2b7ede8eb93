// Fake in-memory Azure Resource Manager and API server, standing in for the Azure SDK for Go
// and for controller-runtime's client.Client (there is no network here).  Semantics follow the
// reference's prose contract for its four helpers (README.md:237-240): tags
// managed-by / owner on every created VM, list order = insertion order, create and delete are
// idempotent, a delete removes the VM together with its NIC and OS disk (modelled as one row).
// Failures can be injected per call kind, which the reference's error paths need
// (README.md:184,192,207,219,229).
#pragma once
#include <map>
#include <optional>
#include <string>
#include <vector>

#include "azurevmpool_types.hpp"

namespace k8sgpu_host {

inline constexpr const char *kManagedByKey = "managed-by";
inline constexpr const char *kManagedByValue = "azurevmpool-operator";  // README.md:238
inline constexpr const char *kOwnerKey = "owner";                       // value "<namespace>-<name>"

struct VM {
    std::string Name;  // the reference dereferences *vm.Name (README.md:217)
    std::map<std::string, std::string> Tags;
};

struct ArmError {
    std::string msg;
};

enum class FailKind { Client = 0, List = 1, Create = 2, Delete = 3, Status = 4 };

// One Azure subscription: resource group name -> VMs in creation order.
class FakeArmClient {
  public:
    // k-th upcoming call of that kind fails (1-based); 0 = never.  Client/Status are on/off.
    void FailAt(FailKind kind, int k) { fail_at_[static_cast<int>(kind)] = k; }
    int Calls(FailKind kind) const { return calls_[static_cast<int>(kind)]; }

    std::optional<ArmError> List(const std::string &rg, std::vector<VM> *out) {
        if (Tick(FailKind::List)) return ArmError{"ARM list failed (injected)"};
        *out = groups_[rg];
        return std::nullopt;
    }
    // create-if-absent by name (README.md:240)
    std::optional<ArmError> Create(const std::string &rg, const VM &vm) {
        if (Tick(FailKind::Create)) return ArmError{"ARM create failed (injected)"};
        auto &g = groups_[rg];
        for (const VM &v : g)
            if (v.Name == vm.Name) return std::nullopt;
        g.push_back(vm);
        return std::nullopt;
    }
    // delete VM + NIC + OS disk; deleting a missing VM succeeds (README.md:239-240)
    std::optional<ArmError> Delete(const std::string &rg, const std::string &name) {
        if (Tick(FailKind::Delete)) return ArmError{"ARM delete failed (injected)"};
        auto &g = groups_[rg];
        for (size_t i = 0; i < g.size(); ++i)
            if (g[i].Name == name) {
                g.erase(g.begin() + static_cast<long>(i));
                break;
            }
        return std::nullopt;
    }
    bool ClientFails() const { return fail_at_[static_cast<int>(FailKind::Client)] != 0; }
    const std::vector<VM> &Group(const std::string &rg) { return groups_[rg]; }
    void Seed(const std::string &rg, VM vm) { groups_[rg].push_back(std::move(vm)); }

  private:
    bool Tick(FailKind kind) {
        int i = static_cast<int>(kind);
        calls_[i]++;
        if (fail_at_[i] <= 0) return false;
        return --fail_at_[i] == 0;
    }
    std::map<std::string, std::vector<VM>> groups_;
    int fail_at_[5] = {0, 0, 0, 0, 0};
    int calls_[5] = {0, 0, 0, 0, 0};
};

// What the reconciler uses of client.Client: Get (README.md:175) and Status().Update (:227).
class FakeApiServer {
  public:
    void Put(const computev1alpha1::AzureVmPool &p) {
        for (auto &q : pools_)
            if (q.Metadata.Namespace == p.Metadata.Namespace && q.Metadata.Name == p.Metadata.Name) {
                q.Spec = p.Spec;
                return;
            }
        pools_.push_back(p);
    }
    bool Delete(const std::string &ns, const std::string &name) {
        for (size_t i = 0; i < pools_.size(); ++i)
            if (pools_[i].Metadata.Namespace == ns && pools_[i].Metadata.Name == name) {
                pools_.erase(pools_.begin() + static_cast<long>(i));
                return true;
            }
        return false;
    }
    // nullptr == NotFound
    computev1alpha1::AzureVmPool *Get(const std::string &ns, const std::string &name) {
        for (auto &q : pools_)
            if (q.Metadata.Namespace == ns && q.Metadata.Name == name) return &q;
        return nullptr;
    }
    bool StatusUpdate(const computev1alpha1::AzureVmPool &p) {
        if (fail_status_) return false;
        if (auto *q = Get(p.Metadata.Namespace, p.Metadata.Name)) {
            q->Status = p.Status;
            status_writes_++;
            return true;
        }
        return false;
    }
    void FailStatus(bool on) { fail_status_ = on; }
    int StatusWrites() const { return status_writes_; }
    const std::vector<computev1alpha1::AzureVmPool> &Pools() const { return pools_; }

  private:
    std::vector<computev1alpha1::AzureVmPool> pools_;
    bool fail_status_ = false;
    int status_writes_ = 0;
};

}  // namespace k8sgpu_host
