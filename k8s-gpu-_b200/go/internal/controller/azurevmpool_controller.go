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

// getAzureVMClient is step 2 of Reconcile (README.md:181): build the Azure VM client from the
// Secret named by spec.azureCredentialSecret (keys AZURE_CLIENT_ID, AZURE_CLIENT_SECRET,
// AZURE_TENANT_ID, AZURE_SUBSCRIPTION_ID, README.md:108).  There is no network here, so the fake
// ARM client stands in; it can be told to fail, which is the path README.md:183-185 handles.
func (r *AzureVmPoolReconciler) getAzureVMClient(ctx context.Context, pool *computev1alpha1.AzureVmPool) (*fakearm.Client, error) {
	if r.ARM == nil {
		return nil, fmt.Errorf("no Azure client for secret %q", pool.Spec.AzureCredentialSecret)
	}
	if err := r.ARM.ClientError(); err != nil {
		return nil, err
	}
	return r.ARM, nil
}

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
	// 2. Azure client; an error here requeues after 30 s (README.md:181-185)
	vmClient, err := r.getAzureVMClient(ctx, &pool)
	if err != nil {
		logger.Error(err, "unable to create Azure VM client")
		return ctrl.Result{RequeueAfter: 30 * time.Second}, err
	}
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
	results, errs := make([]ctrl.Result, len(reqs)), make([]error, len(reqs))
	type live struct {
		req  int
		pool computev1alpha1.AzureVmPool
	}
	var pools []live
	for i, req := range reqs { // step 1 per request (README.md:174-177)
		var p computev1alpha1.AzureVmPool
		if err := r.Get(ctx, req.NamespacedName, &p); err != nil {
			errs[i] = client.IgnoreNotFound(err)
			continue
		}
		if _, err := r.getAzureVMClient(ctx, &p); err != nil { // step 2 (README.md:181-185)
			results[i], errs[i] = ctrl.Result{RequeueAfter: 30 * time.Second}, err
			continue
		}
		pools = append(pools, live{i, p})
	}
	// step 3 once per resource group; a failed list fails every pool of that group (README.md:192)
	listed, failed := map[string][]fakearm.VM{}, map[string]error{}
	for _, l := range pools {
		rg := l.pool.Spec.ResourceGroupName
		if _, done := listed[rg]; !done && failed[rg] == nil {
			if vms, err := r.ARM.ListVMs(rg); err != nil {
				failed[rg] = err
			} else {
				listed[rg] = vms
			}
		}
	}
	// tables: pool row p has owner id p; a VM carries the row of the pool (of its group) it names
	type key struct{ rg, owner string }
	rowOf, rows := map[key]uint32{}, []int{}
	var desired []int32
	var poolOwner []uint32
	for i, l := range pools {
		if failed[l.pool.Spec.ResourceGroupName] != nil {
			continue
		}
		rowOf[key{l.pool.Spec.ResourceGroupName, ownerTag(&l.pool)}] = uint32(len(rows))
		poolOwner = append(poolOwner, uint32(len(rows)))
		desired = append(desired, l.pool.Spec.Replicas)
		rows = append(rows, i)
	}
	P := uint32(len(rows))
	var vmOwner, vmFlags, vmOrdinal []uint32
	var vmName []string
	for rg, vms := range listed {
		for _, vm := range vms {
			oid, ord, flag := uint32(0xFFFFFFFF), uint32(0xFFFFFFFF), uint32(0)
			if owner, ok := vm.Tags[fakearm.OwnerKey]; ok {
				if row, mine := rowOf[key{rg, owner}]; mine {
					oid = row
					if k, ok := ordinalOf(*vm.Name, pools[rows[row]].pool.Name); ok {
						ord = k
					}
				} else {
					oid = P // somebody else's owner tag
				}
			}
			if vm.Tags[fakearm.ManagedByKey] == fakearm.ManagedByValue {
				flag = 1
			}
			vmOwner, vmFlags, vmOrdinal = append(vmOwner, oid), append(vmFlags, flag), append(vmOrdinal, ord)
			vmName = append(vmName, *vm.Name)
		}
	}
	var dec *k8sgpu.Decisions
	var gpuErr error
	if P > 0 {
		dec, gpuErr = r.GPU.ReconcileBatch(desired, poolOwner, vmOwner, vmFlags, vmOrdinal, P)
	}
	for row, i := range rows { // README.md:203-230 per pool, driven by the GPU's decisions
		l := &pools[i]
		if gpuErr != nil {
			results[l.req], errs[l.req] = ctrl.Result{RequeueAfter: 30 * time.Second}, gpuErr
			continue
		}
		var err error
		for j := dec.CreateOff[row]; j < dec.CreateOff[row+1] && err == nil; j++ {
			err = r.createVMNamed(ctx, &l.pool, fmt.Sprintf("%s-%d", l.pool.Name, dec.CreateOrdinal[j]))
		}
		for j := dec.VictimOff[row]; j < dec.VictimOff[row+1] && err == nil; j++ {
			err = r.deleteVM(ctx, &l.pool, vmName[dec.VictimVM[j]])
		}
		if err != nil {
			results[l.req], errs[l.req] = ctrl.Result{RequeueAfter: 40 * time.Second}, err
			continue
		}
		l.pool.Status.ReadyReplicas = dec.Ready[row]
		l.pool.Status.VMs = l.pool.Status.VMs[:0]
		for j := dec.MemberOff[row]; j < dec.MemberOff[row+1]; j++ {
			l.pool.Status.VMs = append(l.pool.Status.VMs, vmName[dec.MemberVM[j]])
		}
		if err := r.Status().Update(ctx, &l.pool); err != nil {
			results[l.req], errs[l.req] = ctrl.Result{}, err
			continue
		}
		results[l.req] = ctrl.Result{RequeueAfter: time.Minute}
	}
	for _, l := range pools {
		if err := failed[l.pool.Spec.ResourceGroupName]; err != nil {
			results[l.req], errs[l.req] = ctrl.Result{RequeueAfter: 20 * time.Second}, err
		}
	}
	return results, errs
}

func (r *AzureVmPoolReconciler) SetupWithManager(mgr ctrl.Manager) error {
	return ctrl.NewControllerManagedBy(mgr).For(&computev1alpha1.AzureVmPool{}).Complete(r)
}
