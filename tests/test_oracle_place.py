"""Placement policy v0 on the oracle: key layout, K12-K14, invariants.
The reference has no source for placement (it installs Volcano/HAMi by helm,
GPU调度平台搭建.md:275-298), so these vectors pin the build-defined policy."""
import numpy as np
import pytest

NONE = 0xFFFFFFFF


def one_node(**kw):
    d = dict(free_gpu=4, free_mcpu=48000, free_mib=900000, cap_gpu=8, sku_id=1, zone_id=2, taint_mask=0, flags=1)
    d.update(kw)
    return d


def one_pod(**kw):
    d = dict(req_gpu=1, req_mcpu=4000, req_mib=28672, sku_mask=NONE, zone_mask=NONE, tol_mask=0, pref_zone=NONE,
             priority=0)
    d.update(kw)
    return d


def tables(nodes, pods):
    n = {k: np.array([x[k] for x in nodes], np.uint32) for k in nodes[0]}
    p = {k: np.array([x[k] for x in pods], np.uint32) for k in pods[0]}
    return n, p


def test_key_bit_layout(orc):
    nodes, pods = tables([one_node(), one_node(free_gpu=1, cap_gpu=1, zone_id=0)], [one_pod(pref_zone=2)])
    N = 2
    rot = orc.rot(0, N)
    for n, (free, cap, zone) in enumerate([(4, 8, 2), (1, 1, 0)]):
        want = (1 << 63) | ((0x7FFF - (free - 1)) << 48) | ((255 - cap) << 40) | (int(zone == 2) << 39) \
            | (0xFFFFFFFF - ((n - rot) % N))
        assert orc.score_key(nodes, pods, 0, n) == want


@pytest.mark.parametrize("node_kw,pod_kw", [
    (dict(flags=0), {}), (dict(free_gpu=0), {}), (dict(free_mcpu=3999), {}), (dict(free_mib=28671), {}),
    ({}, dict(sku_mask=1)), ({}, dict(zone_mask=3)), (dict(taint_mask=2), dict(tol_mask=1)),
])
def test_filter_clauses_each_veto(orc, node_kw, pod_kw):
    nodes, pods = tables([one_node(**node_kw)], [one_pod(**pod_kw)])
    assert orc.score_key(nodes, pods, 0, 0) == 0
    nodes, pods = tables([one_node()], [one_pod()])
    assert orc.score_key(nodes, pods, 0, 0) >> 63 == 1


def test_k12_identical_ties_rot_zero(orc, synth):
    # N identical nodes with 8 GPUs, M identical 1-GPU pods, lowest-index tie-break:
    # pod i lands on node floor(i / 8); one round per node.
    nodes, pods = synth.gen_identical(5, 40)
    r = orc.place(nodes, pods, flags=orc.POLICY_ROT_ZERO)
    assert r["assign"].tolist() == [i // 8 for i in range(40)]
    assert r["rounds"] == 5 and r["placed"] == 40
    assert r["free_gpu"].tolist() == [0] * 5
    # one pod too many stays unplaced (-1) after a final all-infeasible round
    nodes, pods = synth.gen_identical(5, 41)
    r = orc.place(nodes, pods, flags=orc.POLICY_ROT_ZERO)
    assert r["assign"].tolist() == [i // 8 for i in range(40)] + [-1]
    assert r["rounds"] == 6


def test_k12_rotation_spreads_and_is_thread_invariant(orc, synth):
    nodes, pods = synth.gen_identical(64, 512)
    a = orc.place(nodes, pods, nthreads=1)
    b = orc.place(nodes, pods, nthreads=7)
    assert np.array_equal(a["assign"], b["assign"]) and a["rounds"] == b["rounds"]
    assert a["placed"] == 512 and a["rounds"] < 64  # rotation keeps the round count low
    assert np.bincount(a["assign"], minlength=64).tolist() == [8] * 64


def test_k13_taint_without_toleration(orc):
    nodes, pods = tables([one_node(taint_mask=1)], [one_pod(), one_pod(tol_mask=1)])
    r = orc.place(nodes, pods)
    assert r["assign"].tolist() == [-1, 0]


def test_k14_pod_larger_than_any_node(orc):
    nodes, pods = tables([one_node(), one_node()], [one_pod(req_gpu=9)])
    r = orc.place(nodes, pods)
    assert r["assign"].tolist() == [-1] and r["rounds"] == 1


def test_best_fit_then_small_sku_then_zone(orc):
    nodes, pods = tables(
        [one_node(free_gpu=4, cap_gpu=8), one_node(free_gpu=2, cap_gpu=8, zone_id=0),
         one_node(free_gpu=2, cap_gpu=4, zone_id=0), one_node(free_gpu=2, cap_gpu=4, zone_id=1)],
        [one_pod(req_gpu=2, pref_zone=1), one_pod(req_gpu=4)])
    r = orc.place(nodes, pods, flags=orc.POLICY_ROT_ZERO)
    assert r["assign"].tolist() == [3, 0]  # leftover 0 beats 2; cap 4 beats 8; preferred zone wins


def test_admission_prefix_rule_counts_rejected_proposers(orc):
    # three pods propose the only node (3 GPUs free): 2 + 2 + 1.  Prefix sums 2, 4, 5:
    # pod 0 admitted, pod 1 rejected, pod 2 ALSO rejected this round (S counts everyone),
    # and is admitted in round 2.
    nodes, pods = tables([one_node(free_gpu=3)], [one_pod(req_gpu=2), one_pod(req_gpu=2), one_pod(req_gpu=1)])
    r = orc.place(nodes, pods, max_rounds=1)
    assert r["assign"].tolist() == [0, -2, -2]
    r = orc.place(nodes, pods)
    assert r["assign"].tolist() == [0, -1, 0] and r["rounds"] == 2


def test_never_overcommits_and_threads_agree(orc, synth):
    nodes, pods = synth.gen_nodes(700, config_id=5), synth.gen_pods(4000, config_id=5)
    a = orc.place(nodes, pods, nthreads=1)
    b = orc.place(nodes, pods, nthreads=6)
    assert np.array_equal(a["assign"], b["assign"])
    ok = a["assign"] >= 0
    for col, req in (("free_gpu", "req_gpu"), ("free_mcpu", "req_mcpu"), ("free_mib", "req_mib")):
        used = np.bincount(a["assign"][ok], weights=pods[req][ok].astype(np.float64), minlength=700)
        assert np.array_equal(nodes[col] - used.astype(np.uint32), a[col])
        assert (used <= nodes[col]).all()
    # placed pods satisfy every filter clause on their node
    n = a["assign"][ok]
    assert (nodes["flags"][n] & 1).all()
    assert ((pods["sku_mask"][ok] >> nodes["sku_id"][n]) & 1).all()
    assert ((pods["zone_mask"][ok] >> nodes["zone_id"][n]) & 1).all()
    assert ((nodes["taint_mask"][n] & ~pods["tol_mask"][ok]) == 0).all()


def test_domain_validation(orc):
    nodes, pods = tables([one_node(cap_gpu=256)], [one_pod()])
    assert orc.validate(nodes, pods) == -1
    nodes, pods = tables([one_node(sku_id=32)], [one_pod()])
    assert orc.validate(nodes, pods) == -1
    nodes, pods = tables([one_node()], [one_pod(pref_zone=40)])
    assert orc.validate(nodes, pods) == -1
    nodes, pods = tables([one_node()], [one_pod()])
    assert orc.validate(nodes, pods) == 0


def test_empty_tables(orc):
    nodes, pods = tables([one_node()], [one_pod()])
    e_nodes = {k: v[:0] for k, v in nodes.items()}
    e_pods = {k: v[:0] for k, v in pods.items()}
    r = orc.place(e_nodes, pods)
    assert r["assign"].tolist() == [-1] and r["rounds"] == 1
    r = orc.place(nodes, e_pods)
    assert r["assign"].tolist() == [] and r["rounds"] == 0
