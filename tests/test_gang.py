"""Gang placement: all-or-nothing groups with minAvailable layered on the
placement rounds.  The reference shows only `minAvailable: 1` (GPU调度平台搭建.md:648), which
degenerates to per-pod placement; the policy for larger values is build-defined (k8sgpu.h).
CPU tests pin the oracle with hand-derived cases; gpu tests compare the CUDA path with it."""
import numpy as np
import pytest

NONE = 0xFFFFFFFF
NCOLS = ("free_gpu", "free_mcpu", "free_mib", "cap_gpu", "sku_id", "zone_id", "taint_mask", "flags")
PCOLS = ("req_gpu", "req_mcpu", "req_mib", "sku_mask", "zone_mask", "tol_mask", "pref_zone", "priority")


def nodes_of(free_gpus):
    n = len(free_gpus)
    g = np.array(free_gpus, np.uint32)
    return dict(free_gpu=g, free_mcpu=g * 12000, free_mib=g * 243200, cap_gpu=np.full(n, 8, np.uint32),
                sku_id=np.ones(n, np.uint32), zone_id=np.zeros(n, np.uint32), taint_mask=np.zeros(n, np.uint32),
                flags=np.ones(n, np.uint32))


def pods_of(req_gpus):
    m = len(req_gpus)
    g = np.array(req_gpus, np.uint32)
    return dict(req_gpu=g, req_mcpu=g * 4000, req_mib=g * 28672, sku_mask=np.full(m, NONE, np.uint32),
                zone_mask=np.full(m, NONE, np.uint32), tol_mask=np.zeros(m, np.uint32),
                pref_zone=np.full(m, NONE, np.uint32), priority=np.zeros(m, np.uint32))


def test_min_available_one_is_plain_placement(orc, synth):
    # the reference's only value: every pod its own decision
    nodes, pods = synth.gen_nodes(300, 5), synth.gen_pods(1500, 5)
    gang = (np.arange(1500) // 3).astype(np.uint32)
    plain = orc.place(nodes, pods)
    g = orc.place_gang(nodes, pods, gang, np.ones(500, np.uint32))
    placed = plain["assign"] >= 0
    assert np.array_equal(g["assign"][placed], plain["assign"][placed])
    # a gang with no member placed fails even minAvailable 1: its pods read -3 instead of -1
    assert set(np.unique(g["assign"][~placed])) <= {-1, -3}
    for k in ("free_gpu", "free_mcpu", "free_mib"):
        assert np.array_equal(g[k], plain[k])
    assert g["iterations"] <= 2


def test_rejected_gang_releases_capacity_for_others(orc):
    # one node, 8 GPUs.  gang 0 = pods 0-2 (4 GPUs each, needs all 3 = 12 GPUs: impossible),
    # pods 3 and 4 (4 GPUs each) are on their own.  Round-based placement first gives the node to
    # pods 0 and 1 (lowest indices); gang 0 then fails its minimum, releases, and 3 and 4 get in.
    nodes, pods = nodes_of([8]), pods_of([4, 4, 4, 4, 4])
    gang = np.array([0, 0, 0, NONE, NONE], np.uint32)
    r = orc.place_gang(nodes, pods, gang, np.array([3], np.uint32), flags=1)
    assert r["assign"].tolist() == [-3, -3, -3, 0, 0]
    assert r["gang_ok"].tolist() == [0] and r["free_gpu"].tolist() == [0] and r["iterations"] == 2
    # with minAvailable 2 the gang is satisfied by its first two members and keeps the node
    r = orc.place_gang(nodes, pods, gang, np.array([2], np.uint32), flags=1)
    assert r["assign"].tolist() == [0, 0, -1, -1, -1] and r["gang_ok"].tolist() == [1] and r["iterations"] == 1


def test_cascade_and_zero_minimum(orc):
    # two nodes of 4 GPUs.  gang 0: pods 0,1 (4 GPUs each, min 2) fits; gang 1: pods 2,3 (4 GPUs each, min 2)
    # cannot (no capacity left) -> rejected, nothing to release; gang 2 (min 0) with one 8-GPU pod never fits
    # anywhere but min 0 keeps it "ok".
    nodes, pods = nodes_of([4, 4]), pods_of([4, 4, 4, 4, 8])
    gang = np.array([0, 0, 1, 1, 2], np.uint32)
    r = orc.place_gang(nodes, pods, gang, np.array([2, 2, 0], np.uint32), flags=1)
    assert r["assign"].tolist() == [0, 1, -3, -3, -1]
    assert r["gang_ok"].tolist() == [1, 0, 1]


def test_bad_gang_id_rejected(orc):
    with pytest.raises(ValueError):
        orc.place_gang(nodes_of([8]), pods_of([1]), np.array([5], np.uint32), np.array([1], np.uint32))


def random_case(rng, n, m, n_gangs):
    nodes = nodes_of(rng.integers(0, 9, n))
    pods = pods_of(rng.choice([1, 1, 1, 2, 4, 8], m))
    gang = rng.integers(0, n_gangs, m).astype(np.uint32)
    gang[rng.random(m) < 0.3] = NONE
    size = np.bincount(gang[gang != NONE], minlength=n_gangs)
    mins = np.minimum(size, rng.integers(0, 6, n_gangs)).astype(np.uint32)
    mins[rng.random(n_gangs) < 0.1] += 50  # some impossible ones
    return nodes, pods, gang, mins


def test_oracle_invariants(orc):
    rng = np.random.default_rng(11)
    for _ in range(20):
        nodes, pods, gang, mins = random_case(rng, 40, 300, 25)
        a = orc.place_gang(nodes, pods, gang, mins, nthreads=1)
        b = orc.place_gang(nodes, pods, gang, mins, nthreads=4)
        assert np.array_equal(a["assign"], b["assign"])
        ok = a["assign"] >= 0
        used = np.bincount(a["assign"][ok], weights=pods["req_gpu"][ok].astype(np.float64), minlength=40)
        assert np.array_equal(nodes["free_gpu"] - used.astype(np.uint32), a["free_gpu"])   # releases are exact
        for g in range(25):
            members = gang == g
            placed = int((a["assign"][members] >= 0).sum())
            if a["gang_ok"][g]:
                assert placed >= mins[g] and not (a["assign"][members] == -3).any()
            else:
                assert (a["assign"][members] == -3).all()


@pytest.mark.gpu
def test_gpu_gang_matches_oracle(orc):
    import k8sgpu_b200  # noqa: F401
    from k8sgpu_b200 import capi

    rng = np.random.default_rng(5)
    with capi.Context(0) as ctx:
        cases = [random_case(rng, 40, 300, 25) for _ in range(8)] + [random_case(rng, 700, 9000, 400)]
        cases.append((nodes_of([8]), pods_of([4, 4, 4, 4, 4]), np.array([0, 0, 0, NONE, NONE], np.uint32),
                      np.array([3], np.uint32)))
        for nodes, pods, gang, mins in cases:
            want = orc.place_gang(nodes, pods, gang, mins, nthreads=8)
            got = ctx.place_gang(nodes, pods, gang, mins)
            for k in ("assign", "free_gpu", "free_mcpu", "free_mib", "gang_ok"):
                assert np.array_equal(got[k], want[k]), k
            assert (got["rounds"], got["placed"], got["evals"], got["gang_iterations"]) == \
                (want["rounds"], want["placed"], want["evals"], want["iterations"])
        with pytest.raises(capi.K8sGpuError) as e:
            ctx.place_gang(nodes_of([8]), pods_of([1]), np.array([5], np.uint32), np.array([1], np.uint32))
        assert e.value.code == capi.E_ARG


@pytest.mark.gpu
def test_gpu_gang_chain_path(orc, monkeypatch):
    import k8sgpu_b200  # noqa: F401
    from k8sgpu_b200 import capi

    monkeypatch.setenv("K8SGPU_COOP", "0")
    rng = np.random.default_rng(6)
    nodes, pods, gang, mins = random_case(rng, 120, 2000, 90)
    with capi.Context(0) as ctx:
        got, want = ctx.place_gang(nodes, pods, gang, mins), orc.place_gang(nodes, pods, gang, mins)
        assert np.array_equal(got["assign"], want["assign"]) and np.array_equal(got["gang_ok"], want["gang_ok"])


@pytest.mark.gpu
def test_gpu_gang_stepwise_equals_one_call(orc):
    """k8sgpu_place_gang_begin / rounds / _verdicts / _end (what one process per GPU runs, here with
    a world of one and the peer-store scoring entry as well) equals the one-call k8sgpu_place_gang."""
    import k8sgpu_b200  # noqa: F401
    from k8sgpu_b200 import capi, placement

    rng = np.random.default_rng(9)
    with capi.Context(0) as ctx:
        for n, m, g in ((40, 300, 25), (300, 5000, 200)):
            nodes, pods, gang, mins = random_case(rng, n, m, g)
            want = orc.place_gang(nodes, pods, gang, mins, nthreads=8)
            pl = placement.ShardedPlacer(ctx)
            pl.upload(nodes, pods)
            st = pl.run_gang(gang, mins)
            got = pl.download()
            for k in ("assign", "free_gpu", "free_mcpu", "free_mib"):
                assert np.array_equal(got[k], want[k]), k
            assert np.array_equal(st["gang_ok"], want["gang_ok"])
            assert (st["rounds"], st["placed"], st["evals"], st["gang_iterations"]) == \
                (want["rounds"], want["placed"], want["evals"], want["iterations"])
        with pytest.raises(capi.K8sGpuError):  # verdicts are refused while pods are still active
            ctx.gang_begin(gang, mins)
            ctx.gang_verdicts()
