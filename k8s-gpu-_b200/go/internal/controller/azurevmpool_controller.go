// UNCOMPILED on this image (no Go toolchain).  The reconciler with the reference's signature
// (README.md:170) and the batch form that calls libk8sgpu.so once for all pools.  The compiled,
// tested equivalent is ../../host/reconciler.cpp.
package controller

import (
	"context"
	"fmt"
	"strconv"
	"strings"
	"time"

	ctrl "sigs.k8s.io/controller-runtime"
	"sigs.k8s.io/controller-runtime/pkg/client"
	"sigs.k8s.io/controller-runtime/pkg/log"

	computev1alpha1 "github.com/my-org/azure-vm-operator/api/v1alpha1"
	"github.com/my-org/azure-vm-operator/internal/fakearm"
	"github.com/my-org/azure-vm-operator/internal/k8sgpu"
)

type AzureVmPoolReconciler struct {
	client.Client
	ARM *fakearm.Client // stands in for the Azure SDK client of README.md:181
	GPU *k8sgpu.Ctx     // nil: single-pool path only
}

func ownerTag(p *computev1alpha1.AzureVmPool) string { return p.Namespace + "-" + p.Name } // README.md:238

func managedBy(vm fakearm.VM, owner string) bool {
	return vm.Tags[fakearm.ManagedByKey] == fakearm.ManagedByValue && vm.Tags[fakearm.OwnerKey] == owner
}

func (r *AzureVmPoolReconciler) listManagedVMs(ctx context.Context, vmClient *fakearm.Client,
	pool *computev1alpha1.AzureVmPool) ([]fakearm.VM, error) {
	all, err := vmClient.ListVMs(pool.Spec.ResourceGroupName)
	if err != nil {
		return nil, err
	}
	var mine []fakearm.VM
	for _, vm := range all {
		if managedBy(vm, ownerTag(pool)) { // strict tag filter, README.md:238
			mine = append(mine, vm)
		}
	}
	return mine, nil
}

func ordinalOf(vmName, poolName string) (uint32, bool) {
	rest, ok := strings.CutPrefix(vmName, poolName+"-")
	if !ok || rest == "" || (len(rest) > 1 && rest[0] == '0') {
		return 0, false
	}
	k, err := strconv.ParseUint(rest, 10, 32)
	return uint32(k), err == nil && k < 0xFFFFFFFF
}

func (r *AzureVmPoolReconciler) createVMNamed(ctx context.Context, pool *computev1alpha1.AzureVmPool, name string) error {
	return r.ARM.CreateVM(pool.Spec.ResourceGroupName, fakearm.VM{Name: &name, Tags: map[string]string{
		fakearm.ManagedByKey: fakearm.ManagedByValue, fakearm.OwnerKey: ownerTag(pool)}})
}

// createVM picks "<pool>-<k>" with the lowest k not in use by this pool (README.md:204-205).
func (r *AzureVmPoolReconciler) createVM(ctx context.Context, pool *computev1alpha1.AzureVmPool) error {
	used := map[uint32]bool{}
	all, _ := r.ARM.ListVMs(pool.Spec.ResourceGroupName)
	for _, vm := range all {
		if managedBy(vm, ownerTag(pool)) {
			if k, ok := ordinalOf(*vm.Name, pool.Name); ok {
				used[k] = true
			}
		}
	}
	k := uint32(0)
	for used[k] {
		k++
	}
	return r.createVMNamed(ctx, pool, fmt.Sprintf("%s-%d", pool.Name, k))
}

func (r *AzureVmPoolReconciler) deleteVM(ctx context.Context, pool *computev1alpha1.AzureVmPool, name string) error {
	return r.ARM.DeleteVM(pool.Spec.ResourceGroupName, name)
}

// Reconcile keeps the reference's control flow and requeue delays (README.md:170-235).
func (r *AzureVmPoolReconciler) Reconcile(ctx context.Context, req ctrl.Request) (ctrl.Result, error) {
	logger := log.FromContext(ctx)
	var pool computev1alpha1.AzureVmPool
	if err := r.Get(ctx, req.NamespacedName, &pool); err != nil {
		return ctrl.Result{}, client.IgnoreNotFound(err)
	}
	vmClient := r.ARM
	existingVMs, err := r.listManagedVMs(ctx, vmClient, &pool)
	if err != nil {
		logger.Error(err, "listing managed VMs failed")
		return ctrl.Result{RequeueAfter: 20 * time.Second}, err
	}
	currentCount := int32(len(existingVMs))
	desiredCount := pool.Spec.Replicas
	if currentCount < desiredCount {
		for i := int32(0); i < desiredCount-currentCount; i++ {
			if err := r.createVM(ctx, &pool); err != nil {
				return ctrl.Result{RequeueAfter: 40 * time.Second}, err
			}
		}
	} else if currentCount > desiredCount {
		for _, vm := range existingVMs[:currentCount-desiredCount] {
			if err := r.deleteVM(ctx, &pool, *vm.Name); err != nil {
				return ctrl.Result{RequeueAfter: 40 * time.Second}, err
			}
		}
	}
	pool.Status.ReadyReplicas = currentCount // pre-action count, README.md:225
	pool.Status.VMs = pool.Status.VMs[:0]
	for _, vm := range existingVMs {
		pool.Status.VMs = append(pool.Status.VMs, *vm.Name)
	}
	if err := r.Status().Update(ctx, &pool); err != nil {
		return ctrl.Result{}, err
	}
	return ctrl.Result{RequeueAfter: time.Minute}, nil
}

// ReconcileAll is the batch form: one list per resource group, ONE k8sgpu call for the decision
// arithmetic of every pool, then the same ARM calls and status writes per pool, same order,
// same early returns.  See BatchReconciler::ReconcileAll in ../../host/reconciler.cpp for the
// table construction, which this function repeats line for line.
func (r *AzureVmPoolReconciler) ReconcileAll(ctx context.Context, reqs []ctrl.Request) ([]ctrl.Result, []error) {
	panic("see ../../host/reconciler.cpp: BatchReconciler::ReconcileAll (compiled and tested); the Go body is row f1 of SURVEY.md section 8 and needs a Go toolchain to be written against")
}

func (r *AzureVmPoolReconciler) SetupWithManager(mgr ctrl.Manager) error {
	return ctrl.NewControllerManagedBy(mgr).For(&computev1alpha1.AzureVmPool{}).Complete(r)
}
