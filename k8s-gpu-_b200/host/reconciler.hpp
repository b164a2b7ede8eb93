// AzureVmPoolReconciler -- C++ mirror of the reference's controller (README.md:170-235).
// The reference is Go on controller-runtime; no Go toolchain exists on this image, so the host
// side above the C-ABI is C++ with the same names, argument meaning and error behaviour:
//   Reconcile(ctx, req) -> (ctrl.Result, error)
//   getAzureVMClient / listManagedVMs / createVM / deleteVM           (README.md:181,189,205,217)
//   requeue 30 s / 20 s / 40 s / 40 s on client / list / create / delete error, 1 min on success,
//   ({} , nil) when the CR is gone, ({}, err) when the status update fails.
//
// BatchReconciler is the drop-in for many pools at once: one ARM list per resource group, one
// k8sgpu_reconcile_batch call for steps 3-5's arithmetic (the CUDA path), then the ARM calls and
// status writes of README.md:203-230 per pool, in the same order and with the same early returns.
#pragma once
#include <chrono>
#include <map>
#include <optional>
#include <string>
#include <utility>
#include <vector>

#include "../../include/k8sgpu.h"
#include "arm_client.hpp"

namespace k8sgpu_host {

namespace ctrl {
struct Result {
    std::chrono::seconds RequeueAfter{0};  // 0 == ctrl.Result{}
};
struct Request {
    std::string Namespace, Name;  // req.NamespacedName
};
}  // namespace ctrl

enum class ErrKind { None = 0, Client = 1, List = 2, Create = 3, Delete = 4, Status = 5, Gpu = 6, Conflict = 7 };
const char *ErrKindName(ErrKind k);  // "ClientFailed", "ListFailed", ... (condition reasons)
struct Error {
    ErrKind kind = ErrKind::None;
    std::string msg;
    explicit operator bool() const { return kind != ErrKind::None; }
};

class AzureVmPoolReconciler {
  public:
    AzureVmPoolReconciler(FakeApiServer *api, FakeArmClient *arm) : api_(api), arm_(arm) {}

    // README.md:170-235, statement by statement.
    std::pair<ctrl::Result, Error> Reconcile(const ctrl::Request &req);

    // the four helpers the reference names (README.md:181,189,205,217)
    std::pair<FakeArmClient *, Error> getAzureVMClient(const computev1alpha1::AzureVmPool &pool);
    std::pair<std::vector<VM>, Error> listManagedVMs(FakeArmClient *vmClient, const computev1alpha1::AzureVmPool &pool);
    Error createVM(const computev1alpha1::AzureVmPool &pool);
    Error createVMNamed(const computev1alpha1::AzureVmPool &pool, const std::string &name);
    Error deleteVM(const computev1alpha1::AzureVmPool &pool, const std::string &name);

    static std::string OwnerTag(const computev1alpha1::AzureVmPool &pool) {  // README.md:238
        return pool.Metadata.Namespace + "-" + pool.Metadata.Name;
    }
    static bool IsManagedBy(const VM &vm, const std::string &owner);
    // k when vm is named "<pool-name>-<k>" (decimal, no sign, no leading junk), else nullopt
    static std::optional<uint32_t> OrdinalOf(const std::string &vm_name, const std::string &pool_name);

  protected:
    // Steps 1-5 for one pool; prev = how this pool's previous pass ended (for the Failed condition).
    std::pair<ctrl::Result, Error> reconcileOnce(const ctrl::Request &req, ErrKind prev);
    // The four condition types README.md:310 names.  A pass that fails returns before the status
    // write (README.md:207,219), so Failed=True can only be written by the NEXT pass that gets as
    // far as step 5: the reconciler remembers how each pool's last pass ended.
    ErrKind lastFailure(const ctrl::Request &req) const;
    void remember(const ctrl::Request &req, const Error &e, bool pool_exists);

    FakeApiServer *api_;
    FakeArmClient *arm_;
    std::map<std::pair<std::string, std::string>, ErrKind> last_failure_;
};

// Per-pool outcome of one batch pass, in request order.
struct BatchOutcome {
    ctrl::Result result;
    Error err;
};

class BatchReconciler : public AzureVmPoolReconciler {
  public:
    BatchReconciler(FakeApiServer *api, FakeArmClient *arm, k8sgpu_ctx *gpu) : AzureVmPoolReconciler(api, arm), gpu_(gpu) {}

    // Reconciles every request in one pass.  The decision arithmetic of all pools runs in ONE
    // k8sgpu_reconcile_batch call; there is no CPU path for it (a failing GPU call is reported
    // as ErrKind::Gpu for every pool of the pass).
    std::vector<BatchOutcome> ReconcileAll(const std::vector<ctrl::Request> &reqs);

    const k8sgpu_stats &LastStats() const { return stats_; }

  private:
    k8sgpu_ctx *gpu_;
    k8sgpu_stats stats_{};
};

}  // namespace k8sgpu_host
