"""A fixed, short slice of tests/fuzz_gpu.py inside the collected GPU suite: 200 seeded random
cases (placement with gangs, every kernel variant and cooperative grid size; batch reconcile),
CUDA path through the C-ABI against the CPU oracle, byte for byte."""
import importlib.util
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def fuzz():
    spec = importlib.util.spec_from_file_location("fuzz_gpu", os.path.join(HERE, "fuzz_gpu.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.mark.parametrize("seed0", [70_000, 71_000, 72_000, 73_000])
def test_fifty_random_cases(fuzz, seed0):
    bad = []
    with fuzz.capi.Context(0) as ctx:
        for seed in range(seed0, seed0 + 50):
            rng = np.random.default_rng(seed)
            ok, desc = (fuzz.case_place if rng.random() < 0.7 else fuzz.case_reconcile)(rng, ctx)
            if not ok:
                bad.append((seed, desc))
    assert not bad, bad
