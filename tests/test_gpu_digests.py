"""BASELINE.json's configs at FULL size against the CPU oracle, through committed digests.

tests/golden/digests.json holds one sha256 per case (assign + final free columns, or every reconcile
output array) computed by the oracle alone (tests/golden/make_digests.py; config 4 takes ~12 minutes
of 8 host cores, so it is not re-run here).  These tests run the CUDA path through the C-ABI on the
same regenerated inputs and compare digests and round / placed / eval counts."""
import importlib.util
import json
import os

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
spec = importlib.util.spec_from_file_location("make_digests", os.path.join(HERE, "golden", "make_digests.py"))
md = importlib.util.module_from_spec(spec)
spec.loader.exec_module(md)
DOC = json.load(open(os.path.join(HERE, "golden", "digests.json")))


@pytest.fixture(scope="module")
def ctx():
    import k8sgpu_b200  # noqa: F401
    from k8sgpu_b200 import capi

    c = capi.Context(0)
    yield c
    c.close()


def test_every_case_has_a_digest():
    assert set(md.PLACE_CASES) == set(DOC["place"]) and set(md.RECONCILE_CASES) == set(DOC["reconcile"])
    for name, c in md.PLACE_CASES.items():
        assert DOC["place"][name]["inputs"] == c, name


@pytest.mark.parametrize("name", ["c3", "c5_ties_rot0_2k_16k"])
def test_oracle_reproduces_committed_digest(orc, name):
    """The committed digests are what oracle/ computes today (the two cases that take seconds)."""
    from k8sgpu_b200.digest import place_digest

    c = md.PLACE_CASES[name]
    nodes, pods = md.place_inputs(c)
    r = orc.place(nodes, pods, flags=c["flags"], nthreads=os.cpu_count() or 1)
    want = DOC["place"][name]
    assert place_digest(r) == want["sha256"]
    assert (r["rounds"], r["placed"], r["evals"]) == (want["rounds"], want["placed"], want["evals"])


def test_oracle_reproduces_reconcile_digest(orc):
    from k8sgpu_b200.digest import reconcile_digest

    assert reconcile_digest(orc.reconcile_batch(*md.reconcile_inputs(md.RECONCILE_CASES["c2"]))) == \
        DOC["reconcile"]["c2"]["sha256"]


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(md.PLACE_CASES))
def test_gpu_placement_digest(ctx, name):
    from k8sgpu_b200.digest import place_digest

    c = md.PLACE_CASES[name]
    nodes, pods = md.place_inputs(c)
    got = ctx.place(nodes, pods, flags=c["flags"])
    want = DOC["place"][name]
    assert (got["rounds"], got["placed"], got["evals"]) == (want["rounds"], want["placed"], want["evals"])
    assert place_digest(got) == want["sha256"]


@pytest.mark.gpu
@pytest.mark.parametrize("path", [None, 1, 2])
@pytest.mark.parametrize("name", list(md.RECONCILE_CASES))
def test_gpu_reconcile_digest(ctx, name, path):
    from k8sgpu_b200.digest import reconcile_digest

    args = md.reconcile_inputs(md.RECONCILE_CASES[name])
    ctx.set_tuning(rec_path=path)
    try:
        got = ctx.reconcile_batch(*args, members=path != 1)
    finally:
        ctx.reset_tuning()
    want = DOC["reconcile"][name]
    if path == 1:  # victims only: compare everything but the member lists with a full run's arrays
        full = ctx.reconcile_batch(*args)
        import numpy as np

        for k in ("pool_of", "current", "delta", "ready", "create_count", "victim_off", "victim_vm", "create_off",
                  "create_ordinal"):
            assert np.array_equal(got[k], full[k]), k
        got = full
    assert len(got["victim_vm"]) == want["victims"] and len(got["create_ordinal"]) == want["creates"]
    assert reconcile_digest(got) == want["sha256"]
