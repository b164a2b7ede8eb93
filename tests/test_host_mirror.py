"""The C++ host mirror of the reference's reconciler (k8s-gpu-_b200/host/).

CPU part: AzureVmPoolReconciler::Reconcile, the statement-by-statement mirror of README.md:170-235,
replays K1-K11 and must agree with the CPU oracle's state machine on every observable: return
value, ARM contents, status, call counts.
GPU part (marked gpu): BatchReconciler (one k8sgpu_reconcile_batch call for all pools) must leave
the same world behind as running the single-pool Reconcile once per pool.
"""
import numpy as np
import pytest

NS = "default"


@pytest.fixture(scope="module")
def hm():
    import __graft_entry__ as g

    g.build()
    import k8sgpu_b200  # noqa: F401
    from k8sgpu_b200 import hostmirror

    return hostmirror


def both_worlds(hm, orc, pools, vms):
    """pools: [(name, replicas)], vms: [(vm_name, managed_by|None, owner|None)] -> (mirror, oracle)"""
    w = hm.World()
    cl, arm = orc.FakeCluster(), orc.FakeArm()
    for name, rep in pools:
        w.put_pool(NS, name, rep)
        cl.put_pool(NS, name, rep)
    for name, mb, owner in vms:
        w.add_vm(name, managed_by=mb, owner=owner)
        arm.add_vm(name, managed_by=mb, owner=owner)
    return w, (cl, arm)


def step_both(hm, orc, w, o, name):
    cl, arm = o
    got = w.reconcile(NS, name)
    want = orc.reconcile_one(cl, arm, NS, name)
    assert got == want, (got, want)
    assert w.vm_names() == arm.names()
    assert w.ready_replicas(NS, name) == cl.ready_replicas(NS, name)
    assert w.status_vms(NS, name) == cl.status_vms(NS, name)
    assert w.status_writes == cl.status_writes
    assert {k: w.calls(k) for k in ("list", "create", "delete")} == arm.calls
    return got


OP = "azurevmpool-operator"


def test_config1_replicas_3_5_2(hm, orc):
    name = "gpu-pool-prod"
    owner = f"{NS}-{name}"
    w, o = both_worlds(hm, orc, [(name, 3)], [(f"{name}-{i}", OP, owner) for i in range(3)])
    assert step_both(hm, orc, w, o, name) == (60, 0)                      # K1
    assert w.condition(NS, name) == "Ready=True:Reconciled"
    w.put_pool(NS, name, 5); o[0].set_replicas(NS, name, 5)
    assert step_both(hm, orc, w, o, name) == (60, 0)                      # K2
    assert w.vm_names() == [f"{name}-{i}" for i in range(5)] and w.ready_replicas(NS, name) == 3
    assert w.condition(NS, name) == "Ready=False:Provisioning"
    assert step_both(hm, orc, w, o, name) == (60, 0)                      # K3
    w.put_pool(NS, name, 2); o[0].set_replicas(NS, name, 2)
    assert step_both(hm, orc, w, o, name) == (60, 0)                      # K4
    assert w.vm_names() == [f"{name}-3", f"{name}-4"] and w.ready_replicas(NS, name) == 5
    assert step_both(hm, orc, w, o, name) == (60, 0)                      # K5
    assert w.ready_replicas(NS, name) == 2


def test_error_paths_match_reference_requeue_delays(hm, orc):
    name, owner = "p", f"{NS}-p"
    seed = [(f"p-{i}", OP, owner) for i in range(3)]
    # K6 missing CR
    w, o = both_worlds(hm, orc, [], seed)
    assert step_both(hm, orc, w, o, name) == (0, 0)
    # client error: 30 s (README.md:184)
    w, o = both_worlds(hm, orc, [(name, 3)], seed)
    w.fail("client"); o[1].fail_client()
    assert step_both(hm, orc, w, o, name) == (30, 1)
    # K7 list error: 20 s, no status write (README.md:192)
    w, o = both_worlds(hm, orc, [(name, 3)], seed)
    w.fail("list", 1); o[1].fail_list_at(1)
    assert step_both(hm, orc, w, o, name) == (20, 2)
    # K8 2nd of 3 creates fails: 40 s, first create stands, status untouched (README.md:207)
    w, o = both_worlds(hm, orc, [(name, 6)], seed)
    w.fail("create", 2); o[1].fail_create_at(2)
    assert step_both(hm, orc, w, o, name) == (40, 3)
    assert step_both(hm, orc, w, o, name) == (60, 0)
    # K9 2nd of 3 deletes fails (README.md:219)
    w, o = both_worlds(hm, orc, [(name, 0)], seed)
    w.fail("delete", 2); o[1].fail_delete_at(2)
    assert step_both(hm, orc, w, o, name) == (40, 4)
    assert w.vm_names() == ["p-1", "p-2"]
    # status update error: ({}, err) (README.md:229)
    w, o = both_worlds(hm, orc, [(name, 3)], seed)
    w.fail("status", 1); o[0].fail_status_update()
    assert step_both(hm, orc, w, o, name) == (0, 5)


def test_four_condition_types(hm, orc):
    """README.md:310 names Provisioning / Ready / Deleting / Failed.  All four are written on every
    status update; Failed=True appears on the first pass AFTER a failed one, because the failing
    pass returns before the status write (README.md:207,219) -- K8 and K9 are unchanged by this."""
    name, owner = "p", f"{NS}-p"
    w, o = both_worlds(hm, orc, [(name, 3)], [(f"p-{i}", OP, owner) for i in range(3)])
    assert step_both(hm, orc, w, o, name) == (60, 0)
    assert w.conditions(NS, name) == {"Ready": ("True", "Reconciled"), "Provisioning": ("False", "NotScalingUp"),
                                      "Deleting": ("False", "NotScalingDown"), "Failed": ("False", "NoError")}
    w.put_pool(NS, name, 6); o[0].set_replicas(NS, name, 6)
    w.fail("create", 2); o[1].fail_create_at(2)
    assert step_both(hm, orc, w, o, name) == (40, 3)                      # K8: status NOT written
    assert w.conditions(NS, name)["Failed"] == ("False", "NoError")       # ... so it still says what it said
    assert step_both(hm, orc, w, o, name) == (60, 0)                      # next pass gets to step 5
    c = w.conditions(NS, name)
    assert c["Failed"] == ("True", "CreateFailed") and c["Provisioning"] == ("True", "ScalingUp")
    assert c["Ready"] == ("False", "Provisioning")                         # ReadyReplicas is the pre-action count (4)
    assert step_both(hm, orc, w, o, name) == (60, 0)
    c = w.conditions(NS, name)
    assert c["Failed"] == ("False", "NoError") and c["Ready"] == ("True", "Reconciled")
    w.put_pool(NS, name, 1); o[0].set_replicas(NS, name, 1)
    w.fail("delete", 3); o[1].fail_delete_at(3)
    assert step_both(hm, orc, w, o, name) == (40, 4)                      # K9
    assert step_both(hm, orc, w, o, name) == (60, 0)
    c = w.conditions(NS, name)
    assert c["Failed"] == ("True", "DeleteFailed") and c["Deleting"] == ("True", "ScalingDown")


def test_k10_k11_foreign_vms_and_scale_to_zero(hm, orc):
    name, owner = "p", f"{NS}-p"
    vms = [("other-0", OP, "default-other"), ("p-0", OP, owner), ("bare", None, owner), ("impostor", "someone", owner),
           ("p-1", OP, owner), ("untagged", None, None)]
    w, o = both_worlds(hm, orc, [(name, 1)], vms)
    assert step_both(hm, orc, w, o, name) == (60, 0)
    assert w.vm_names() == ["other-0", "bare", "impostor", "p-1", "untagged"] and w.ready_replicas(NS, name) == 2
    w, o = both_worlds(hm, orc, [(name, 0)], [(f"p-{i}", OP, owner) for i in range(4)])
    assert step_both(hm, orc, w, o, name) == (60, 0)
    assert w.vm_names() == []


def random_world(rng, hm, n_pools, n_vms, two_groups):
    """Same random world twice (for the sequential and the batch reconciler)."""
    pools = [(f"pool{p}", int(rng.integers(0, 9)), ("rgA" if (p % 2 == 0 or not two_groups) else "rgB"))
             for p in range(n_pools)]
    vms, used = [], set()
    for i in range(n_vms):
        kind = rng.random()
        p = int(rng.integers(0, n_pools))
        rg = pools[p][2]
        if kind < 0.75:
            k = int(rng.integers(0, 12))
            nm = f"pool{p}-{k}"
            if (rg, nm) in used:
                nm = f"pool{p}-extra{i}"
            vms.append((nm, OP, f"{NS}-pool{p}", rg))
        elif kind < 0.85:
            nm = f"foreign-{i}"
            vms.append((nm, OP, "default-somebody-else", rg))
        elif kind < 0.95:
            nm = f"notours-{i}"
            vms.append((nm, "other-operator", f"{NS}-pool{p}", rg))
        else:
            nm = f"untagged-{i}"
            vms.append((nm, None, None, rg))
        used.add((rg, nm))
    worlds = []
    for _ in range(2):
        w = hm.World()
        for name, rep, rg in pools:
            w.put_pool(NS, name, rep, rg)
        for nm, mb, owner, rg in vms:
            w.add_vm(nm, managed_by=mb, owner=owner, rg=rg)
        worlds.append(w)
    return pools, worlds


def snapshot(w, pools):
    return ({rg: w.vm_names(rg) for rg in ("rgA", "rgB")},
            {name: (w.ready_replicas(NS, name), w.status_vms(NS, name), w.conditions(NS, name)) for name, _, _ in pools},
            w.status_writes)


@pytest.mark.gpu
@pytest.mark.parametrize("seed", range(6))
def test_batch_reconciler_equals_sequential(hm, seed):
    rng = np.random.default_rng(seed)
    pools, (seq, bat) = random_world(rng, hm, int(rng.integers(1, 12)), int(rng.integers(0, 90)), seed % 2 == 1)
    fail = [None, ("create", 2), ("delete", 3), None, ("status", 1), None][seed]
    if fail:
        seq.fail(*fail)
        bat.fail(*fail)
    keys = [(NS, name) for name, _, _ in pools] + [(NS, "no-such-pool")]
    want = [seq.reconcile(ns, nm) for ns, nm in keys]
    got = bat.reconcile_batch(keys)
    assert got == want
    assert snapshot(bat, pools) == snapshot(seq, pools)
    assert bat.last_gpu_launches > 0  # the decisions came from the CUDA library
    # a second pass settles both worlds identically (level-triggered reconcile)
    if fail and fail[0] == "status":
        seq.fail("status", 0)
        bat.fail("status", 0)
    want = [seq.reconcile(ns, nm) for ns, nm in keys]
    assert bat.reconcile_batch(keys) == want
    assert snapshot(bat, pools) == snapshot(seq, pools)


@pytest.mark.gpu
def test_batch_owner_tag_collision_and_repeated_requests(hm):
    """"<namespace>-<name>" is not injective: team-a/pool and team/a-pool both own the tag
    "team-a-pool".  The batch pass gives the second pool an explicit Conflict outcome (no table row,
    no out-of-range row index) and still reconciles the first.  A pool named twice in one pass gets
    the same outcome both times -- also when that outcome is the 30 s client failure."""
    w = hm.World()
    w.put_pool("team-a", "pool", 2)
    w.put_pool("team", "a-pool", 5)
    w.put_pool(NS, "solo", 1)
    for i in range(3):
        w.add_vm(f"pool-{i}", managed_by=OP, owner="team-a-pool")
    out = w.reconcile_batch([("team-a", "pool"), ("team", "a-pool"), (NS, "solo"), ("team", "a-pool")])
    assert out[0] == (60, 0) and out[2] == (60, 0)
    assert out[1] == (30, hm.ERR["Conflict"]) and out[3] == out[1]
    assert w.vm_names() == ["pool-1", "pool-2", "solo-0"]  # first victim in list order; solo scaled up
    w2 = hm.World()
    w2.put_pool(NS, "p", 1)
    w2.fail("client")
    assert w2.reconcile_batch([(NS, "p"), (NS, "p"), (NS, "gone")]) == [(30, 1), (30, 1), (0, 0)]


def test_batch_reconciler_needs_a_gpu(hm):
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    w = hm.World()
    w.put_pool(NS, "p", 1)
    with pytest.raises(RuntimeError):
        w.reconcile_batch([(NS, "p")])
