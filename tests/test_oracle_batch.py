"""Batch reconcile oracle == the single-pool state machine run once per pool (SURVEY.md 8(c)
property tests), plus hand-derived vectors for the CSR outputs."""
import numpy as np
from hypothesis import given, settings, strategies as st

NONE = 0xFFFFFFFF


def run_single_pool_machine(orc, n_pools, desired, vm_rows):
    """vm_rows: list of (owner_pool or None/'foreign', managed_ok, ordinal or None).
    Builds one fake resource group shared by all pools and reconciles every pool once."""
    cl, arm = orc.FakeCluster(), orc.FakeArm()
    for p in range(n_pools):
        cl.put_pool("ns", f"pool{p}", int(desired[p]))
    names = []
    for row, (owner, ok, ordinal) in enumerate(vm_rows):
        if owner == "foreign":
            owner_tag, base = "ns-somebody-else", "x"
        elif owner is None:
            owner_tag, base = None, "x"
        else:
            owner_tag, base = f"ns-pool{owner}", f"pool{owner}"
        # only a VM this operator manages carries an operator-style name (unique per resource group)
        managed = ok and isinstance(owner, int)
        name = f"{base}-{ordinal}" if (ordinal is not None and managed) else f"vm-row{row}"
        if name in names:  # VM names are unique inside a resource group
            name = f"vm-row{row}"
        names.append(name)
        arm.add_vm(name, managed_by="azurevmpool-operator" if ok else "not-us", owner=owner_tag)
    before = arm.names()
    res = []
    for p in range(n_pools):
        pre = arm.names()
        assert orc.reconcile_one(cl, arm, "ns", f"pool{p}") == (60, 0)
        post = arm.names()
        deleted = [n for n in pre if n not in post]
        created = [n for n in post if n not in pre]
        res.append(dict(ready=cl.ready_replicas("ns", f"pool{p}"), members=cl.status_vms("ns", f"pool{p}"),
                        deleted=deleted, created=created))
    return before, res


def to_tables(n_pools, desired, vm_rows, names):
    owner = np.array([NONE if o is None else (n_pools + 7 if o == "foreign" else o) for o, _, _ in vm_rows],
                     dtype=np.uint32)
    flags = np.array([1 if ok else 0 for _, ok, _ in vm_rows], dtype=np.uint32)
    ordinal = []
    for row, (o, ok, k) in enumerate(vm_rows):
        # the ordinal column only means something when the VM is really named <pool>-<k>
        real = isinstance(o, int) and names[row] == f"pool{o}-{k}"
        ordinal.append(k if real else NONE)
    return dict(desired=np.array(desired, np.int32), pool_owner_id=np.arange(n_pools, dtype=np.uint32),
                vm_owner_id=owner, vm_flags=flags, vm_ordinal=np.array(ordinal, np.uint32),
                n_owner_ids=n_pools + 16)


vm_row = st.tuples(st.one_of(st.none(), st.just("foreign"), st.integers(0, 3)), st.booleans(),
                   st.one_of(st.none(), st.integers(0, 9)))


@settings(max_examples=150, deadline=None)
@given(desired=st.lists(st.integers(0, 8), min_size=4, max_size=4), rows=st.lists(vm_row, max_size=30))
def test_batch_equals_single_pool_machine(orc, desired, rows):
    names, single = run_single_pool_machine(orc, 4, desired, rows)
    t = to_tables(4, desired, rows, names)
    out = orc.reconcile_batch(t["desired"], t["pool_owner_id"], t["vm_owner_id"], t["vm_flags"],
                              t["n_owner_ids"], t["vm_ordinal"])
    for p in range(4):
        s = single[p]
        assert out["ready"][p] == s["ready"] == out["current"][p]
        assert out["delta"][p] == desired[p] - s["ready"]
        mem = out["member_vm"][out["member_off"][p]:out["member_off"][p + 1]]
        assert [names[v] for v in mem] == s["members"]                      # Status.VMs
        vic = out["victim_vm"][out["victim_off"][p]:out["victim_off"][p + 1]]
        assert [names[v] for v in vic] == s["deleted"]                      # README.md:214 order
        cre = out["create_ordinal"][out["create_off"][p]:out["create_off"][p + 1]]
        assert [f"pool{p}-{k}" for k in cre] == s["created"]               # a6 names
        assert out["create_count"][p] == len(s["created"])
    managed = out["pool_of"] >= 0
    assert managed.sum() == out["current"].sum()


def test_batch_hand_vector(orc):
    # rows:        0  1  2  3     4  5  6  7  8
    owner = [1, 0, 1, NONE, 0, 1, 9, 1, 0]
    flags = [1, 1, 1, 1,    1, 0, 1, 1, 1]
    out = orc.reconcile_batch([1, 1, 4], [0, 1, 2], owner, flags, 16, [0, 0, 1, 0, 2, 5, 0, 3, 7])
    assert out["pool_of"].tolist() == [1, 0, 1, -1, 0, -1, -1, 1, 0]
    assert out["current"].tolist() == [3, 3, 0]
    assert out["delta"].tolist() == [-2, -2, 4]
    assert out["ready"].tolist() == [3, 3, 0]
    assert out["create_count"].tolist() == [0, 0, 4]
    assert out["victim_off"].tolist() == [0, 2, 4, 4]
    assert out["victim_vm"].tolist() == [1, 4, 0, 2]
    assert out["member_vm"].tolist() == [1, 4, 8, 0, 2, 7]
    assert out["create_ordinal"].tolist() == [0, 1, 2, 3]


def test_batch_rejects_shared_owner_and_negative_desired(orc):
    import pytest
    with pytest.raises(ValueError):
        orc.reconcile_batch([1, 1], [3, 3], [3], [1], 8)
    with pytest.raises(ValueError):
        orc.reconcile_batch([-1], [0], [0], [1], 8)


def test_batch_thread_invariance_and_sums(orc, synth):
    t = synth.gen_reconcile(200, 20000)
    a = orc.reconcile_batch(t["desired"], t["pool_owner_id"], t["vm_owner_id"], t["vm_flags"],
                            t["n_owner_ids"], t["vm_ordinal"], nthreads=1)
    b = orc.reconcile_batch(t["desired"], t["pool_owner_id"], t["vm_owner_id"], t["vm_flags"],
                            t["n_owner_ids"], t["vm_ordinal"], nthreads=5)
    for k in a:
        assert np.array_equal(a[k], b[k]), k
    assert len(a["victim_vm"]) == np.maximum(-a["delta"], 0).sum()
    # victims are each pool's lowest row ids
    for p in range(200):
        rows = np.nonzero(a["pool_of"] == p)[0]
        k = max(-a["delta"][p], 0)
        assert a["victim_vm"][a["victim_off"][p]:a["victim_off"][p + 1]].tolist() == rows[:k].tolist()
