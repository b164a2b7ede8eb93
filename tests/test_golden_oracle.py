"""The oracle must reproduce the committed golden fixtures (tests/golden/make_golden.py)."""
import importlib.util
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
spec = importlib.util.spec_from_file_location("make_golden", os.path.join(HERE, "golden", "make_golden.py"))
mg = importlib.util.module_from_spec(spec)
spec.loader.exec_module(mg)


def load(name):
    return np.load(os.path.join(HERE, "golden", name))


@pytest.mark.parametrize("name", list(mg.PLACE_CASES) + list(mg.TIES_CASES))
def test_oracle_place_matches_golden(orc, name):
    nodes, pods, flags = mg.place_inputs(name)
    g = load(f"place_{name}.npz")
    r = orc.place(nodes, pods, flags=flags, nthreads=4)
    for k in ("assign", "free_gpu", "free_mcpu", "free_mib"):
        assert np.array_equal(r[k], g[k]), k
    assert (r["rounds"], r["placed"], r["evals"]) == (int(g["rounds"]), int(g["placed"]), int(g["evals"]))


@pytest.mark.parametrize("name", list(mg.RECONCILE_CASES))
def test_oracle_reconcile_matches_golden(orc, synth, name):
    c = mg.RECONCILE_CASES[name]
    t = synth.gen_reconcile(c["p"], c["v"], c["config_id"])
    o = orc.reconcile_batch(t["desired"], t["pool_owner_id"], t["vm_owner_id"], t["vm_flags"], t["n_owner_ids"],
                            t["vm_ordinal"], nthreads=3)
    g = load(f"reconcile_{name}.npz")
    for k in g.files:
        assert np.array_equal(o[k], g[k]), k
