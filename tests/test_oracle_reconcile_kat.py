"""Known-answer tests K1-K11 for the single-pool Reconcile state machine.

The reference has no tests; every vector below is derived from the
cited lines of /root/reference/README.md and pins the oracle (oracle/reconcile_sm.c).
BASELINE.json config 1: one AzureVmPool, replicas 3 -> 5 -> 2 against a fake ARM client.
"""
import pytest

NS, NAME = "default", "gpu-pool-prod"  # sample CR name, README.md:271
OWNER = f"{NS}-{NAME}"                 # owner tag, README.md:238
ERR = dict(OK=0, CLIENT=1, LIST=2, CREATE=3, DELETE=4, STATUS=5)


def fresh(orc, replicas=3, n_vms=3):
    cl, arm = orc.FakeCluster(), orc.FakeArm()
    cl.put_pool(NS, NAME, replicas)
    for i in range(n_vms):
        arm.add_vm(f"{NAME}-{i}", owner=OWNER)
    return cl, arm


def test_k1_to_k5_config1_replicas_3_5_2(orc):
    cl, arm = fresh(orc)
    # K1: desired 3, three VMs listed -> no ARM mutation, ReadyReplicas 3, requeue 1 min (:234)
    assert orc.reconcile_one(cl, arm, NS, NAME) == (60, ERR["OK"])
    assert arm.calls == {"list": 1, "create": 0, "delete": 0}
    assert cl.ready_replicas(NS, NAME) == 3
    assert cl.status_vms(NS, NAME) == [f"{NAME}-0", f"{NAME}-1", f"{NAME}-2"]
    # K2: desired 5 -> 2 creates (:201-209); status carries the PRE-action count (:225)
    cl.set_replicas(NS, NAME, 5)
    assert orc.reconcile_one(cl, arm, NS, NAME) == (60, ERR["OK"])
    assert arm.calls["create"] == 2
    assert arm.names() == [f"{NAME}-{i}" for i in range(5)]
    assert cl.ready_replicas(NS, NAME) == 3
    # K3: settle -> no-op, ReadyReplicas 5
    assert orc.reconcile_one(cl, arm, NS, NAME) == (60, ERR["OK"])
    assert arm.calls["create"] == 2 and arm.calls["delete"] == 0
    assert cl.ready_replicas(NS, NAME) == 5
    # K4: desired 2 -> deletes existingVMs[:3] in list order (:214)
    cl.set_replicas(NS, NAME, 2)
    assert orc.reconcile_one(cl, arm, NS, NAME) == (60, ERR["OK"])
    assert arm.calls["delete"] == 3
    assert arm.names() == [f"{NAME}-3", f"{NAME}-4"]
    assert cl.ready_replicas(NS, NAME) == 5
    # K5: settle -> no-op, ReadyReplicas 2, survivors are the two newest
    assert orc.reconcile_one(cl, arm, NS, NAME) == (60, ERR["OK"])
    assert arm.calls["delete"] == 3
    assert cl.ready_replicas(NS, NAME) == 2
    assert cl.status_vms(NS, NAME) == [f"{NAME}-3", f"{NAME}-4"]
    assert cl.status_writes == 5


def test_k6_missing_cr_is_swallowed(orc):
    cl, arm = fresh(orc)
    cl.delete_pool(NS, NAME)
    # client.IgnoreNotFound -> ({}, nil), no ARM traffic (:175-177)
    assert orc.reconcile_one(cl, arm, NS, NAME) == (0, ERR["OK"])
    assert arm.calls == {"list": 0, "create": 0, "delete": 0}
    assert cl.status_writes == 0


def test_client_error_requeues_30s(orc):
    cl, arm = fresh(orc)
    arm.fail_client()
    assert orc.reconcile_one(cl, arm, NS, NAME) == (30, ERR["CLIENT"])  # :184
    assert arm.calls["list"] == 0 and cl.status_writes == 0


def test_k7_list_error_requeues_20s_without_status(orc):
    cl, arm = fresh(orc)
    arm.fail_list_at(1)
    assert orc.reconcile_one(cl, arm, NS, NAME) == (20, ERR["LIST"])  # :192
    assert cl.status_writes == 0


def test_k8_second_of_three_creates_fails(orc):
    cl, arm = fresh(orc, replicas=6)
    arm.fail_create_at(2)
    assert orc.reconcile_one(cl, arm, NS, NAME) == (40, ERR["CREATE"])  # :207
    assert arm.names() == [f"{NAME}-{i}" for i in range(4)]  # the first create stands
    assert cl.status_writes == 0 and cl.ready_replicas(NS, NAME) == 0  # status untouched
    # level-triggered: the next pass lists 4 and creates the missing 2
    assert orc.reconcile_one(cl, arm, NS, NAME) == (60, ERR["OK"])
    assert len(arm.names()) == 6 and cl.ready_replicas(NS, NAME) == 4


def test_k9_second_of_three_deletes_fails(orc):
    cl, arm = fresh(orc, replicas=0)
    arm.fail_delete_at(2)
    assert orc.reconcile_one(cl, arm, NS, NAME) == (40, ERR["DELETE"])  # :219
    assert arm.names() == [f"{NAME}-1", f"{NAME}-2"]  # first victim gone, loop returned
    assert cl.status_writes == 0


def test_k10_foreign_vms_are_never_touched(orc):
    cl, arm = fresh(orc, replicas=1, n_vms=0)
    arm.add_vm("other-0", owner="default-other")                     # wrong owner
    arm.add_vm(f"{NAME}-0", owner=OWNER)
    arm.add_vm("bare-vm", managed_by=None, owner=OWNER)              # managed-by absent
    arm.add_vm("impostor", managed_by="someone-else", owner=OWNER)   # managed-by wrong
    arm.add_vm(f"{NAME}-1", owner=OWNER)
    arm.add_vm("untagged", managed_by=None, owner=None)
    assert orc.reconcile_one(cl, arm, NS, NAME) == (60, ERR["OK"])
    assert cl.ready_replicas(NS, NAME) == 2  # only the two strictly tagged VMs count (:238)
    assert arm.names() == ["other-0", "bare-vm", "impostor", f"{NAME}-1", "untagged"]


def test_k11_desired_zero_deletes_all_in_list_order(orc):
    cl, arm = fresh(orc, replicas=0, n_vms=4)
    assert orc.reconcile_one(cl, arm, NS, NAME) == (60, ERR["OK"])
    assert arm.names() == [] and arm.calls["delete"] == 4
    assert cl.ready_replicas(NS, NAME) == 4


def test_status_update_error_returns_empty_result(orc):
    cl, arm = fresh(orc)
    cl.fail_status_update()
    assert orc.reconcile_one(cl, arm, NS, NAME) == (0, ERR["STATUS"])  # :229


def test_create_fills_lowest_free_ordinals(orc):
    cl, arm = fresh(orc, replicas=5, n_vms=0)
    for o in (1, 3):
        arm.add_vm(f"{NAME}-{o}", owner=OWNER)
    assert orc.reconcile_one(cl, arm, NS, NAME) == (60, ERR["OK"])
    assert arm.names() == [f"{NAME}-1", f"{NAME}-3", f"{NAME}-0", f"{NAME}-2", f"{NAME}-4"]


def test_negative_replicas_rejected_by_crd_validation(orc):
    cl = orc.FakeCluster()
    assert cl.put_pool(NS, NAME, -1) == -1  # +kubebuilder:validation:Minimum=0, README.md:94
