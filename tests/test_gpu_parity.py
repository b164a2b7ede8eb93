"""Parity of the CUDA path (through the C-ABI, libk8sgpu.so) against the CPU oracle.
Bit-exact: everything on this path is integer / index work.  Run on the B200 box:
    python -m pytest tests -m gpu -x -q
"""
import importlib.util
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
spec = importlib.util.spec_from_file_location("make_golden", os.path.join(HERE, "golden", "make_golden.py"))
mg = importlib.util.module_from_spec(spec)
spec.loader.exec_module(mg)

NONE = 0xFFFFFFFF
NCPU = os.cpu_count() or 1


@pytest.fixture(scope="module")
def ctx():
    import k8sgpu_b200  # noqa: F401
    from k8sgpu_b200 import capi

    c = capi.Context(0)
    yield c
    c.close()


@pytest.fixture(autouse=True)
def _default_tuning(request):
    """Kernel-variant knobs set by a test (ctx.set_tuning) never leak into the next one."""
    yield
    if "ctx" in request.fixturenames:
        request.getfixturevalue("ctx").reset_tuning()


@pytest.fixture(scope="module")
def capi():
    import k8sgpu_b200  # noqa: F401
    from k8sgpu_b200 import capi as c

    return c


def same_placement(got, want):
    for k in ("assign", "free_gpu", "free_mcpu", "free_mib"):
        assert np.array_equal(got[k], want[k]), k
    assert (got["rounds"], got["placed"], got["evals"]) == (want["rounds"], want["placed"], want["evals"])


# ------------------------------------------------------------------ primitives
@pytest.mark.parametrize("n", [0, 1, 31, 2047, 2048, 2049, 16384, 16385, 100_003, 1_200_000])
def test_exclusive_scan(ctx, n):
    rng = np.random.default_rng(n)
    a = rng.integers(0, 1000, n, dtype=np.uint32)
    out, tot = ctx.exclusive_scan(a)
    want = np.concatenate([[0], np.cumsum(a, dtype=np.uint64)[:-1]]).astype(np.uint32) if n else a
    assert np.array_equal(out, want) and tot == int(a.sum())


@pytest.mark.parametrize("n,bits", [(0, 8), (1, 8), (33, 3), (2048, 8), (2049, 9), (5000, 1), (100_003, 17),
                                    (1_000_000, 20), (300_000, 32)])
def test_sort_pairs_is_stable(ctx, n, bits):
    rng = np.random.default_rng(n + bits)
    keys = rng.integers(0, 1 << min(bits, 31), n, dtype=np.uint32)
    if n > 10:
        keys[: n // 3] = keys[0]  # one very long bucket
    vals = np.arange(n, dtype=np.uint32)
    k, v = ctx.sort_pairs(keys, vals, bits)
    order = np.argsort(keys, kind="stable")
    assert np.array_equal(k, keys[order]) and np.array_equal(v, vals[order])


# ------------------------------------------------------------------ scoring kernel alone
@pytest.mark.parametrize("ppt,splits,res", [(None, None, None), ("1", "1", "4"), ("2", "1", "4"), ("1", "3", "2"),
                                            ("2", "7", "2"), ("4", "2", "4"), ("4", "1", "4"), ("1", "2", "0"),
                                            ("2", "3", "1"), ("1", "5", "3")])
@pytest.mark.parametrize("flags", [0, 1])
def test_score_once_matches_oracle(ctx, orc, synth, ppt, splits, res, flags):
    """Every kernel variant (pods per thread, node splits, eval_pair RES 0-4) is bit-identical."""
    if ppt:
        ctx.set_tuning(ppt=int(ppt))
        ctx.set_tuning(splits=int(splits))
        ctx.set_tuning(res=int(res))
    nodes, pods = synth.gen_nodes(1500, config_id=5), synth.gen_pods(3333, config_id=5)
    ctx.place_upload(nodes, pods)
    got = ctx.score_once(flags)
    want = orc.score_round(nodes, (nodes["free_gpu"], nodes["free_mcpu"], nodes["free_mib"]), pods,
                           np.arange(3333, dtype=np.uint32), flags=flags, nthreads=NCPU)
    assert np.array_equal(got, want)


def test_score_once_identical_ties(ctx, orc, synth):
    # every node ties on the upper key bits: the result is decided by the rotated index alone
    nodes, pods = synth.gen_identical(777, 2000)
    ctx.place_upload(nodes, pods)
    free = (nodes["free_gpu"], nodes["free_mcpu"], nodes["free_mib"])
    act = np.arange(2000, dtype=np.uint32)
    assert np.array_equal(ctx.score_once(0), orc.score_round(nodes, free, pods, act, flags=0))
    assert np.array_equal(ctx.score_once(1), np.zeros(2000, np.int32))


@pytest.mark.parametrize("n", [1, 7, 8, 9, 255, 256, 257, 511, 513, 1031])
@pytest.mark.parametrize("splits", [None, "1", "5"])
def test_rotation_tiebreak_at_chunk_and_tile_edges(ctx, orc, synth, n, splits):
    # identical nodes: only rot(p) decides, and rot(p) lands on every chunk/tile boundary case
    if splits:
        ctx.set_tuning(splits=int(splits))
    nodes, pods = synth.gen_identical(n, 700)
    ctx.place_upload(nodes, pods)
    free = (nodes["free_gpu"], nodes["free_mcpu"], nodes["free_mib"])
    want = orc.score_round(nodes, free, pods, np.arange(700, dtype=np.uint32), flags=0)
    assert np.array_equal(ctx.score_once(0), want)
    # and with only every third node feasible (equal keys separated by infeasible nodes)
    nodes["flags"] = (np.arange(n) % 3 == 0).astype(np.uint32)
    ctx.place_upload(nodes, pods)
    want = orc.score_round(nodes, free, pods, np.arange(700, dtype=np.uint32), flags=0)
    assert np.array_equal(ctx.score_once(0), want)


# ------------------------------------------------------------------ whole placement
@pytest.mark.parametrize("name", list(mg.PLACE_CASES) + list(mg.TIES_CASES))
def test_place_matches_golden_and_oracle(ctx, orc, name):
    nodes, pods, flags = mg.place_inputs(name)
    g = np.load(os.path.join(HERE, "golden", f"place_{name}.npz"))
    got = ctx.place(nodes, pods, flags=flags)
    for k in ("assign", "free_gpu", "free_mcpu", "free_mib"):
        assert np.array_equal(got[k], g[k]), k
    assert (got["rounds"], got["placed"], got["evals"]) == (int(g["rounds"]), int(g["placed"]), int(g["evals"]))
    same_placement(got, orc.place(nodes, pods, flags=flags, nthreads=NCPU))


def test_place_tables_in_one_block_and_caller_outputs(ctx, synth):
    """Columns kept back to back in one block travel as one copy per table; results land in the
    caller's arrays (also one block for the three free columns).  Same answer as separate columns."""
    nodes, pods = synth.gen_nodes(777, config_id=5), synth.gen_pods(4321, config_id=5)
    want = ctx.place(nodes, pods)

    def block(cols, order):
        n = len(cols[order[0]])
        blk = np.empty(len(order) * n, np.uint32)
        out = {}
        for c, k in enumerate(order):
            out[k] = blk[c * n:(c + 1) * n]
            out[k][:] = cols[k]
        return out

    bn = block(nodes, ("free_gpu", "free_mcpu", "free_mib", "cap_gpu", "sku_id", "zone_id", "taint_mask", "flags"))
    bp = block(pods, ("req_gpu", "req_mcpu", "req_mib", "sku_mask", "zone_mask", "tol_mask", "pref_zone", "priority"))
    out = block({k: np.zeros(777, np.uint32) for k in ("free_gpu", "free_mcpu", "free_mib")},
                ("free_gpu", "free_mcpu", "free_mib"))
    out["assign"] = np.full(4321, 99, np.int32)
    got = ctx.place(bn, bp, out=out)
    for k in ("assign", "free_gpu", "free_mcpu", "free_mib"):
        assert got[k] is out[k] and np.array_equal(got[k], want[k]), k
    with pytest.raises(ValueError):
        ctx.place(bn, bp, out={**out, "assign": np.zeros(5, np.int32)})


@pytest.mark.parametrize("h100,taint,m", [(0, 0, 10), (50, 5, 100), (100, 50, 1000), (30, 5, 10_000),
                                          (50, 0, 40_000)])
def test_place_heterogeneous_sweep(ctx, orc, synth, h100, taint, m):
    # BASELINE.json config 5 at sizes the oracle finishes in seconds
    nodes = synth.gen_nodes(2000, config_id=5, h100_pct=h100, taint_pct=taint)
    pods = synth.gen_pods(m, config_id=5)
    same_placement(ctx.place(nodes, pods), orc.place(nodes, pods, nthreads=NCPU))


@pytest.mark.parametrize("n,m", [(200, 140_000), (150, 300_000)])
def test_place_long_active_list(ctx, orc, synth, n, m):
    """More proposals than the single-launch admission handles with its per-CTA local scans
    (> 64 sort tiles; > 1024 segment blocks): the grid-wide scan / carry phases, against the oracle."""
    nodes, pods = synth.gen_nodes(n, config_id=5), synth.gen_pods(m, config_id=5)
    same_placement(ctx.place(nodes, pods), orc.place(nodes, pods, nthreads=NCPU))


@pytest.mark.parametrize("grid", ["1", "3", "7"])
def test_place_few_ctas_many_tiles(ctx, orc, synth, grid):
    """30 sort tiles / 240 segment blocks handled by 1, 3 or 7 CTAs: every per-CTA loop of the
    single-launch admission runs many iterations (shared scratch is reused across them)."""
    ctx.set_tuning(coop_grid=int(grid))
    nodes, pods = synth.gen_nodes(500, config_id=5), synth.gen_pods(60_000, config_id=5)
    same_placement(ctx.place(nodes, pods), orc.place(nodes, pods, nthreads=NCPU))


@pytest.mark.parametrize("grid", [None, "1", "3", "40"])
@pytest.mark.parametrize("name", ["c3_small", "ties_rot_zero", "c5_all_h100"])
def test_place_single_launch_admission(ctx, name, grid):
    """K8SGPU_COOP=1: the cooperative one-launch admission path gives the golden result too --
    also with so few CTAs that each one loops over several tiles in every phase."""
    ctx.set_tuning(coop=int("1"))
    if grid:
        ctx.set_tuning(coop_grid=int(grid))
    nodes, pods, flags = mg.place_inputs(name)
    g = np.load(os.path.join(HERE, "golden", f"place_{name}.npz"))
    got = ctx.place(nodes, pods, flags=flags)
    for k in ("assign", "free_gpu", "free_mcpu", "free_mib"):
        assert np.array_equal(got[k], g[k]), k
    assert (got["rounds"], got["placed"], got["evals"]) == (int(g["rounds"]), int(g["placed"]), int(g["evals"]))
    # two launches per round, plus the once-per-run table preparation (pod and node records, the
    # taint union, 24-bit sort of the scoring order: 3 passes x histogram / scan / scatter, slot table, count)
    assert got["kernel_launches"] <= 2 * got["rounds"] + 20


@pytest.mark.parametrize("order", [0, 1])
@pytest.mark.parametrize("n,m,coop_grid", [(500, 60_000, 3), (1500, 9_000, None), (200, 140_000, None), (37, 300, 1)])
def test_place_class_sorted_scoring_order(ctx, orc, synth, order, n, m, coop_grid):
    """The scoring kernel walks each round's active list through an order sorted by pod class, so
    that whole warps can drop the clauses none of their pods need.  With the order on and off the
    result is the oracle's: short and long lists, few CTAs (many tiles per CTA in the partition)."""
    nodes, pods = synth.gen_nodes(n, config_id=5), synth.gen_pods(m, config_id=5)
    ctx.set_tuning(order=order, coop_grid=coop_grid)
    same_placement(ctx.place(nodes, pods), orc.place(nodes, pods, nthreads=NCPU))


@pytest.mark.parametrize("force", [None, 16, 17])
@pytest.mark.parametrize("case", ["all_or_nothing", "partial", "no_taints", "high_bits", "everything_tainted"])
def test_toleration_classes(ctx, orc, synth, case, force):
    """The scoring kernel leaves the toleration clause out for warps whose pods tolerate every
    taint that occurs (or none of them: those read free GPUs from a field that is -1 on tainted
    nodes) and takes the free-GPU test from the key word.  Every mix against the oracle, next to
    the kernel with that specialisation ruled out (force_cls = 16): partial tolerations (clause
    evaluated), no taint at all, taint bit 31, requests of 0 and beyond any node, unschedulable
    nodes with free capacity."""
    rng = np.random.default_rng(hash(case) % 1000)
    n, m = 1100, 9000
    nodes, pods = synth.gen_nodes(n, config_id=5), synth.gen_pods(m, config_id=5)
    taints = {"all_or_nothing": [0, 0, 0, 1, 2, 3], "partial": [0, 0, 1, 2, 4, 7], "no_taints": [0],
              "high_bits": [0, 0, 0x80000000, 0x80000001, 0x40000000], "everything_tainted": [1, 2, 3]}[case]
    nodes["taint_mask"] = rng.choice(np.array(taints, np.uint32), n).astype(np.uint32)
    union = int(np.bitwise_or.reduce(nodes["taint_mask"]))
    tols = {"all_or_nothing": [0, 0, union, 0xFFFFFFFF], "partial": [0, 1, 2, 3, 5, 7, 0xFFFFFFFF],
            "no_taints": [0, 1, 0xFFFFFFFF], "high_bits": [0, 0x80000000, 0xC0000001, 1],
            "everything_tainted": [0, 1, 2, 3]}[case]
    pods["tol_mask"] = rng.choice(np.array(tols, np.uint32), m).astype(np.uint32)
    pods["req_gpu"] = pods["req_gpu"].copy()
    pods["req_gpu"][::97] = 0
    pods["req_gpu"][5::101] = 0x8000
    pods["req_gpu"][7::103] = 0x7FFFFFFF
    nodes["flags"] = (rng.random(n) > 0.1).astype(np.uint32)  # unschedulable nodes keep their free capacity
    ctx.set_tuning(force_cls=force)
    ctx.place_upload(nodes, pods)
    free = (nodes["free_gpu"], nodes["free_mcpu"], nodes["free_mib"])
    want = orc.score_round(nodes, free, pods, np.arange(m, dtype=np.uint32), flags=0, nthreads=NCPU)
    assert np.array_equal(ctx.score_once(0), want)
    same_placement(ctx.place(nodes, pods), orc.place(nodes, pods, nthreads=NCPU))


@pytest.mark.parametrize("force", [None, 18])
@pytest.mark.parametrize("case", ["implied", "cpu_scarce", "mem_scarce_some", "mixed_requests", "wide_node", "zero_gpu"])
def test_resource_implication(ctx, orc, synth, case, force):
    """Each round the library summarises the node table (least free CPU / memory among the nodes
    with >= g free GPUs) and leaves the CPU and memory differences out of the evals of pods for
    which enough GPUs implies enough of both at every node.  Tables where that holds for all pods,
    for none, for some, where it stops holding as nodes fill up, with a node beyond the summary's
    64 GPUs, and with pods asking for 0 GPUs -- against the oracle, and against the kernel with
    the shortcut ruled out (force_cls = 18)."""
    rng = np.random.default_rng(len(case))
    n, m = 900, 12000
    nodes, pods = synth.gen_nodes(n, config_id=5), synth.gen_pods(m, config_id=5)
    nodes = {k: v.copy() for k, v in nodes.items()}
    pods = {k: v.copy() for k, v in pods.items()}
    if case == "cpu_scarce":
        nodes["free_mcpu"] = (nodes["free_mcpu"] // 3).astype(np.uint32)
    elif case == "mem_scarce_some":
        tight = rng.random(n) < 0.2
        nodes["free_mib"] = np.where(tight, nodes["free_mib"] // 2, nodes["free_mib"]).astype(np.uint32)
    elif case == "mixed_requests":
        pods["req_mcpu"] = (pods["req_gpu"] * rng.choice([1000, 4000, 13000], m)).astype(np.uint32)
        pods["req_mib"] = (pods["req_gpu"] * rng.choice([1024, 28672, 300000], m)).astype(np.uint32)
    elif case == "wide_node":
        nodes["free_gpu"][3], nodes["cap_gpu"][3] = 200, 255
        nodes["free_mcpu"][3], nodes["free_mib"][3] = 100, 100
    elif case == "zero_gpu":
        pods["req_gpu"][::5] = 0
        pods["req_mcpu"][::5] = rng.choice([0, 500, 50_000], len(pods["req_gpu"][::5])).astype(np.uint32)
    ctx.set_tuning(force_cls=force)
    ctx.place_upload(nodes, pods)
    free = (nodes["free_gpu"], nodes["free_mcpu"], nodes["free_mib"])
    want = orc.score_round(nodes, free, pods, np.arange(m, dtype=np.uint32), flags=0, nthreads=NCPU)
    assert np.array_equal(ctx.score_once(0), want)
    same_placement(ctx.place(nodes, pods), orc.place(nodes, pods, nthreads=NCPU))


@pytest.mark.parametrize("slot_admit", [0, 1])
@pytest.mark.parametrize("case", ["hetero", "ties_rot0", "zero_gpu_pod", "big_node", "one_node", "coop_grid_1"])
def test_place_slot_admission(ctx, orc, synth, slot_admit, case):
    """Admission by per-node slot arrays (the first free_gpu proposers of each node, no sort) against
    the oracle, next to the sort-based admission: mixed tables, everyone proposing to one node, and
    the two inputs that rule the slot path out for a run (a pod asking for 0 GPUs; a node with more
    than 8 free GPUs), where the kernel must take the sort on its own."""
    flags = 0
    if case == "ties_rot0":
        nodes, pods = synth.gen_identical(40, 700)
        flags = 1
    elif case == "one_node":
        nodes, pods = synth.gen_identical(1, 3000)
    else:
        nodes, pods = synth.gen_nodes(900, config_id=5), synth.gen_pods(30_000, config_id=5)
    if case == "zero_gpu_pod":
        pods["req_gpu"] = pods["req_gpu"].copy()
        pods["req_gpu"][::7] = 0
    if case == "big_node":
        for k, per in (("free_gpu", 1), ("free_mcpu", 12000), ("free_mib", 243200)):
            nodes[k] = nodes[k].copy()
            nodes[k][5] = 12 * per
        nodes["cap_gpu"] = nodes["cap_gpu"].copy()
        nodes["cap_gpu"][5] = 16
    ctx.set_tuning(slot_admit=slot_admit, coop_grid=1 if case == "coop_grid_1" else None)
    same_placement(ctx.place(nodes, pods, flags=flags), orc.place(nodes, pods, flags=flags, nthreads=NCPU))


def test_context_creation_and_knob_errors(ctx, capi):
    """k8sgpu_create_ex reports its failures in the caller's buffer; an unknown knob is an error,
    not a silent no-op; a one-GPU context is a group of one."""
    with pytest.raises(capi.K8sGpuError) as e:
        capi.Context([0, 0])
    assert e.value.code == capi.E_ARG and "twice" in str(e.value)
    with pytest.raises(capi.K8sGpuError) as e:
        capi.Context(4096)
    assert e.value.code == capi.E_ARG and "out of range" in str(e.value)
    with pytest.raises(capi.K8sGpuError) as e:
        ctx.set_tuning(no_such_knob=1)
    assert e.value.code == capi.E_ARG and "no_such_knob" in str(e.value)
    info = ctx.group_info()
    assert (info["world"], info["rank"], info["nlocal"]) == (1, 0, 1)
    with pytest.raises(capi.K8sGpuError):  # a context cannot join a group with a bad rank
        ctx.comm_init_rank(3, 2, b"\0" * capi.UNIQUE_ID_BYTES)


def test_pinned_host_blocks_from_the_library(ctx, capi, orc, synth):
    """Tables and results in page-locked memory handed out by k8sgpu_host_alloc (what a cgo caller
    would use), columns back to back: same bytes as from ordinary numpy arrays."""
    nodes, pods = synth.gen_nodes(500, config_id=5), synth.gen_pods(6000, config_id=5)
    nb, pb, ob = capi.PinnedBlock(8 * 500), capi.PinnedBlock(8 * 6000), capi.PinnedBlock(3 * 500 + 6000)
    hn = {k: nb.array[c * 500:(c + 1) * 500] for c, k in enumerate(capi.NODE_COLS)}
    hp = {k: pb.array[c * 6000:(c + 1) * 6000] for c, k in enumerate(capi.POD_COLS)}
    for k in hn:
        hn[k][:] = nodes[k]
    for k in hp:
        hp[k][:] = pods[k]
    out = {k: ob.array[c * 500:(c + 1) * 500] for c, k in enumerate(("free_gpu", "free_mcpu", "free_mib"))}
    out["assign"] = ob.array[1500:].view(np.int32)
    got = ctx.place(hn, hp, out=out)
    same_placement(got, orc.place(nodes, pods, nthreads=NCPU))
    for b in (nb, pb, ob):
        b.close()


def test_place_without_priority_column(ctx, orc, synth):
    """priority is reserved (policy v0 admits in pod-index order): it may be left out."""
    nodes, pods = synth.gen_nodes(300, config_id=5), synth.gen_pods(4000, config_id=5)
    want = orc.place(nodes, pods, nthreads=NCPU)
    slim = {k: v for k, v in pods.items() if k != "priority"}
    slim["priority"] = None
    same_placement(ctx.place(nodes, slim), want)


def test_k12_ties_rot_zero(ctx, synth):
    nodes, pods = synth.gen_identical(5, 41)
    r = ctx.place(nodes, pods, flags=1)
    assert r["assign"].tolist() == [i // 8 for i in range(40)] + [-1]
    assert r["rounds"] == 6 and r["free_gpu"].tolist() == [0] * 5


def test_k13_k14_and_prefix_rule(ctx, orc):
    def tab(rows, cols):
        return {c: np.array([r[i] for r in rows], np.uint32) for i, c in enumerate(cols)}

    ncols = ("free_gpu", "free_mcpu", "free_mib", "cap_gpu", "sku_id", "zone_id", "taint_mask", "flags")
    pcols = ("req_gpu", "req_mcpu", "req_mib", "sku_mask", "zone_mask", "tol_mask", "pref_zone", "priority")
    # K13: taint without toleration -> -1; with toleration -> placed
    nodes = tab([(4, 48000, 900000, 8, 1, 2, 1, 1)], ncols)
    pods = tab([(1, 4000, 28672, NONE, NONE, 0, NONE, 0), (1, 4000, 28672, NONE, NONE, 1, NONE, 0)], pcols)
    assert ctx.place(nodes, pods)["assign"].tolist() == [-1, 0]
    # K14: pod larger than any node -> -1 in one round
    pods = tab([(9, 4000, 28672, NONE, NONE, 1, NONE, 0)], pcols)
    r = ctx.place(nodes, pods)
    assert r["assign"].tolist() == [-1] and r["rounds"] == 1
    # a13: rejected proposers still count in the prefix sums
    nodes = tab([(3, 48000, 900000, 8, 1, 2, 0, 1)], ncols)
    pods = tab([(2, 1, 1, NONE, NONE, 0, NONE, 0), (2, 1, 1, NONE, NONE, 0, NONE, 0), (1, 1, 1, NONE, NONE, 0, NONE, 0)],
               pcols)
    assert ctx.place(nodes, pods, max_rounds=1)["assign"].tolist() == [0, -2, -2]
    r = ctx.place(nodes, pods)
    assert r["assign"].tolist() == [0, -1, 0] and r["rounds"] == 2
    same_placement(r, orc.place(nodes, pods))


def test_empty_tables(ctx, synth):
    nodes, pods = synth.gen_nodes(10), synth.gen_pods(7)
    e_nodes = {k: v[:0] for k, v in nodes.items()}
    e_pods = {k: v[:0] for k, v in pods.items()}
    r = ctx.place(e_nodes, pods)
    assert r["assign"].tolist() == [-1] * 7 and r["rounds"] == 1
    r = ctx.place(nodes, e_pods)
    assert r["assign"].tolist() == [] and r["rounds"] == 0 and np.array_equal(r["free_gpu"], nodes["free_gpu"])


def test_input_domain_is_enforced(ctx, capi, synth):
    nodes, pods = synth.gen_nodes(100), synth.gen_pods(100)
    bad = dict(nodes)
    bad["cap_gpu"] = nodes["cap_gpu"].copy()
    bad["cap_gpu"][7] = 256
    with pytest.raises(capi.K8sGpuError) as e:
        ctx.place(bad, pods)
    assert e.value.code == capi.E_ARG
    badp = dict(pods)
    badp["pref_zone"] = pods["pref_zone"].copy()
    badp["pref_zone"][3] = 40
    with pytest.raises(capi.K8sGpuError) as e:
        ctx.place(nodes, badp)
    assert e.value.code == capi.E_ARG
    ctx.place(nodes, pods)  # the context stays usable


def test_resident_run_is_repeatable(ctx, orc, synth):
    nodes, pods = synth.gen_nodes(900), synth.gen_pods(5000)
    ctx.place_upload(nodes, pods)
    want = orc.place(nodes, pods, nthreads=NCPU)
    for _ in range(2):
        st = ctx.place_run()
        got = ctx.place_download()
        got.update(st)
        same_placement(got, want)
    assert st["kernel_launches"] > 0 and st["score_launches"] == st["rounds"]


def test_round_api_with_two_emulated_shards(ctx, orc, synth):
    """The round-at-a-time entry points with the active list scored in two slices (what two
    cooperating processes do) give the single-call result."""
    import torch

    nodes, pods = synth.gen_nodes(800), synth.gen_pods(4000)
    ctx.place_upload(nodes, pods)
    stream = torch.cuda.current_stream().cuda_stream
    na = ctx.round_begin(stream=stream)
    buf = torch.empty(4000, dtype=torch.int32, device="cuda:0")
    while na:
        half = (na + 1) // 2
        ctx.round_score(0, half, buf.data_ptr(), stream)
        ctx.round_score(half, na, buf.data_ptr() + 4 * half, stream)
        na = ctx.round_admit(buf.data_ptr(), stream)
    st = ctx.round_end(stream)
    got = ctx.place_download()
    got.update(st)
    same_placement(got, orc.place(nodes, pods, nthreads=NCPU))


def test_round_api_peer_store_world_of_one(ctx, orc, synth):
    """The peer-store scoring entry with a world of one rank (its own exchange buffer is the
    only destination): two slices per round, admission from the half the round's parity selects.
    The multi-process form is checked by tests/run_sharded_gpu.py on 2+ GPUs."""
    import torch

    nodes, pods = synth.gen_nodes(700, 5), synth.gen_pods(3000, 5)
    ctx.place_upload(nodes, pods)
    handle, local = ctx.peer_export(3000)
    assert len(handle) == 64 and local
    ctx.peer_import(0, 1, handle)
    stream = torch.cuda.current_stream().cuda_stream
    na, r = ctx.round_begin(stream=stream), 0
    while na:
        half = (na + 1) // 2
        ctx.round_score_p2p(0, half, stream)
        ctx.round_score_p2p(half, na, stream)
        na = ctx.round_admit(local + 4 * 3000 * (r & 1), stream)
        r += 1
    st = ctx.round_end(stream)
    got = ctx.place_download()
    got.update(st)
    same_placement(got, orc.place(nodes, pods, nthreads=NCPU))
    # a table larger than the exported buffer is refused, not overrun
    ctx.place_upload(nodes, synth.gen_pods(3001, 5))
    ctx.round_begin(stream=stream)
    with pytest.raises(Exception, match="exchange buffer"):
        ctx.round_score_p2p(0, 10, stream)
    ctx.round_end(stream)


def test_config3_full_size(ctx, orc, synth):
    """BASELINE.json config 3 at full size: 10k nodes x 100k pods, bit-exact vs the oracle
    (the oracle needs roughly 100 core-seconds; it is threaded over pods)."""
    nodes, pods = synth.gen_nodes(10_000), synth.gen_pods(100_000)
    got = ctx.place(nodes, pods)
    same_placement(got, orc.place(nodes, pods, nthreads=NCPU))
    # size-independent properties: no over-commit, free = initial - used
    ok = got["assign"] >= 0
    for col, req in (("free_gpu", "req_gpu"), ("free_mcpu", "req_mcpu"), ("free_mib", "req_mib")):
        used = np.bincount(got["assign"][ok], weights=pods[req][ok].astype(np.float64), minlength=10_000)
        assert (used <= nodes[col]).all()
        assert np.array_equal(nodes[col] - used.astype(np.uint32), got[col])


def test_config4_shape_properties(ctx, synth):
    """BASELINE.json config 4 node count (100k) with 200k pods: too big for the oracle in test
    time, so check the invariants the policy guarantees and idempotence of a second pass."""
    nodes, pods = synth.gen_nodes(100_000, config_id=4), synth.gen_pods(200_000, config_id=4)
    got = ctx.place(nodes, pods)
    a = got["assign"]
    ok = a >= 0
    assert ((a >= 0) | (a == -1)).all()
    for col, req in (("free_gpu", "req_gpu"), ("free_mcpu", "req_mcpu"), ("free_mib", "req_mib")):
        used = np.bincount(a[ok], weights=pods[req][ok].astype(np.float64), minlength=100_000)
        assert np.array_equal(nodes[col] - used.astype(np.uint32), got[col])
    n = a[ok]
    assert (nodes["flags"][n] & 1).all()
    assert ((pods["sku_mask"][ok] >> nodes["sku_id"][n]) & 1).all()
    assert ((pods["zone_mask"][ok] >> nodes["zone_id"][n]) & 1).all()
    assert ((nodes["taint_mask"][n] & ~pods["tol_mask"][ok]) == 0).all()
    # pods left unplaced stay unplaced against the final tables (nothing feasible remains)
    left = {k: v[~ok] for k, v in pods.items()}
    after = dict(nodes, free_gpu=got["free_gpu"], free_mcpu=got["free_mcpu"], free_mib=got["free_mib"])
    again = ctx.place(after, left)
    assert (again["assign"] == -1).all() and again["rounds"] == 1


# ------------------------------------------------------------------ batch reconcile
def same_reconcile(got, want):
    for k in want:
        assert np.array_equal(got[k], want[k]), k


@pytest.mark.parametrize("grid", [None, "1", "3"])
@pytest.mark.parametrize("name", list(mg.RECONCILE_CASES))
def test_reconcile_matches_golden(ctx, synth, name, grid):
    if grid:  # few CTAs in the one-launch pass: every CTA loops over several tiles
        ctx.set_tuning(rec_grid=int(grid))
    c = mg.RECONCILE_CASES[name]
    t = synth.gen_reconcile(c["p"], c["v"], c["config_id"])
    got = ctx.reconcile_batch(t["desired"], t["pool_owner_id"], t["vm_owner_id"], t["vm_flags"], t["n_owner_ids"],
                              t["vm_ordinal"])
    g = np.load(os.path.join(HERE, "golden", f"reconcile_{name}.npz"))
    for k in g.files:
        assert np.array_equal(got[k], g[k]), k


@pytest.mark.parametrize("p,v", [(1000, 100_000), (1, 1000), (7, 3), (5000, 1_000_003), (6144, 70_000),
                                 (6145, 70_000), (40_000, 400_000)])
def test_reconcile_matches_oracle(ctx, orc, synth, p, v):
    # (1000, 100000) is BASELINE.json config 2 at full size; 6144 / 6145 pools straddle the limit of
    # the one-launch path (automatic path selection on both sides of it)
    t = synth.gen_reconcile(p, v)
    args = (t["desired"], t["pool_owner_id"], t["vm_owner_id"], t["vm_flags"], t["n_owner_ids"], t["vm_ordinal"])
    same_reconcile(ctx.reconcile_batch(*args), orc.reconcile_batch(*args, nthreads=NCPU))


@pytest.mark.parametrize("path,members", [(0, True), (0, False), (1, False), (2, True), (2, False)])
@pytest.mark.parametrize("p,v,segments,grid", [(1000, 100_000, None, None), (300, 77_777, 3, 5), (3000, 250_001, 9, 17),
                                               (17, 4097, 1, 1), (2, 100_000, 40, 64), (5000, 300, None, None)])
def test_reconcile_every_path(ctx, orc, synth, path, members, p, v, segments, grid):
    """The three ways of ranking rows inside their pool -- one cooperative launch, pruned candidate
    sort, full stable sort -- give the oracle's bytes, with and without member lists, with few and
    many CTAs / row segments (every per-CTA loop runs several tiles; early exit of the pruning pass)."""
    t = synth.gen_reconcile(p, v)
    args = (t["desired"], t["pool_owner_id"], t["vm_owner_id"], t["vm_flags"], t["n_owner_ids"], t["vm_ordinal"])
    ctx.set_tuning(rec_path=path, rec_segments=segments, rec_grid=grid)
    got, want = ctx.reconcile_batch(*args, members=members), orc.reconcile_batch(*args, nthreads=NCPU)
    for k in want:
        if members or not k.startswith("member_"):
            assert np.array_equal(got[k], want[k]), k


@pytest.mark.parametrize("path", [0, 1, 2])
def test_reconcile_scale_to_zero_and_one_giant_pool(ctx, orc, path):
    """K11 in batch form: desired 0 everywhere (every managed row is a victim, in list order), and
    one pool that owns nearly every row (the longest possible rank chain)."""
    rng = np.random.default_rng(11)
    V, P = 120_000, 40
    vowner = rng.integers(0, P + 5, V).astype(np.uint32)
    vflags = (rng.random(V) < 0.95).astype(np.uint32)
    ctx.set_tuning(rec_path=path, rec_segments=5)
    for desired, owner in ((np.zeros(P, np.int32), vowner), (np.full(P, 3, np.int32), np.where(rng.random(V) < 0.97, 7, vowner).astype(np.uint32))):
        args = (desired, np.arange(P, dtype=np.uint32), owner, vflags, P + 5, None)
        got, want = ctx.reconcile_batch(*args, members=path != 1), orc.reconcile_batch(*args, nthreads=NCPU)
        for k in want:
            if path != 1 or not k.startswith("member_"):
                assert np.array_equal(got[k], want[k]), k


def test_reconcile_single_block_outputs(ctx, capi, orc, synth):
    """Outputs laid out in one caller block as k8sgpu_reconcile_out_size says (the layout the
    library keeps in HBM, so the copies merge): same bytes as separately allocated arrays."""
    t = synth.gen_reconcile(1000, 100_000)
    args = (t["desired"], t["pool_owner_id"], t["vm_owner_id"], t["vm_flags"], t["n_owner_ids"], t["vm_ordinal"])
    P, V, cap = 1000, 100_000, int(t["desired"].sum()) + 1
    lay = capi.reconcile_out_size(P, V, cap)
    assert lay["pool_of"] == 0 and lay["current"] == V and lay["total"] == 3 * V + 7 * P + 3 + cap
    block = np.full(lay["total"], 0xDEADBEEF, np.uint32)
    sizes = {"pool_of": V, "current": P, "ready": P, "delta": P, "create_count": P, "victim_off": P + 1,
             "member_off": P + 1, "create_off": P + 1, "victim_vm": V, "member_vm": V, "create_ordinal": cap}
    order = iter(["pool_of", "current", "delta", "ready", "create_count", "victim_off", "victim_vm", "member_off",
                  "member_vm", "create_off", "create_ordinal"])  # the order Context._rec_out asks for buffers in

    def alloc(n, dt):
        k = next(order)
        return block[lay[k]: lay[k] + sizes[k]].view(dt)

    got = ctx.reconcile_batch(*args, alloc=alloc, create_capacity=cap - 1)
    same_reconcile(got, orc.reconcile_batch(*args, nthreads=NCPU))


def test_reconcile_hand_vector_and_config1(ctx):
    owner = [1, 0, 1, NONE, 0, 1, 9, 1, 0]
    flags = [1, 1, 1, 1, 1, 0, 1, 1, 1]
    out = ctx.reconcile_batch([1, 1, 4], [0, 1, 2], owner, flags, 16, [0, 0, 1, 0, 2, 5, 0, 3, 7])
    assert out["pool_of"].tolist() == [1, 0, 1, -1, 0, -1, -1, 1, 0]
    assert out["current"].tolist() == [3, 3, 0] and out["delta"].tolist() == [-2, -2, 4]
    assert out["victim_off"].tolist() == [0, 2, 4, 4] and out["victim_vm"].tolist() == [1, 4, 0, 2]
    assert out["member_vm"].tolist() == [1, 4, 8, 0, 2, 7]
    assert out["create_ordinal"].tolist() == [0, 1, 2, 3]
    # BASELINE.json config 1 as a batch of one pool: replicas 3 -> 5 -> 2 (K1, K2, K4)
    one = lambda desired, ords: ctx.reconcile_batch([desired], [0], [0] * len(ords), [1] * len(ords), 1, ords)
    k1 = one(3, [0, 1, 2])
    assert (k1["delta"].tolist(), k1["ready"].tolist(), k1["victim_vm"].tolist()) == ([0], [3], [])
    k2 = one(5, [0, 1, 2])
    assert (k2["create_count"].tolist(), k2["create_ordinal"].tolist(), k2["ready"].tolist()) == ([2], [3, 4], [3])
    k4 = one(2, [0, 1, 2, 3, 4])
    assert (k4["delta"].tolist(), k4["victim_vm"].tolist(), k4["ready"].tolist()) == ([-3], [0, 1, 2], [5])


def test_reconcile_errors_and_empty(ctx, capi):
    with pytest.raises(capi.K8sGpuError) as e:
        ctx.reconcile_batch([1, 1], [3, 3], [3], [1], 8)
    assert e.value.code == capi.E_ARG
    with pytest.raises(capi.K8sGpuError):
        ctx.reconcile_batch([-1], [0], [0], [1], 8)
    with pytest.raises(capi.K8sGpuError) as e:  # far outside any buffer if it were ever used as a count
        ctx.reconcile_batch([-1_000_000, 0, -(2 ** 31)], [0, 1, 2], [1] * 10, [1] * 10, 8, list(range(10)))
    assert e.value.code == capi.E_ARG and "negative" in str(e.value)
    ok = ctx.reconcile_batch([0, 12], [0, 1], [1] * 10, [1] * 10, 8, list(range(10)))  # the context still works
    assert ok["delta"].tolist() == [0, 2] and ok["create_ordinal"].tolist() == [10, 11]
    out = ctx.reconcile_batch([2, 0], [0, 1], [], [], 4)
    assert out["delta"].tolist() == [2, 0] and out["victim_off"].tolist() == [0, 0, 0]
    assert out["create_ordinal"].tolist() == [0, 1]
    out = ctx.reconcile_batch([], [], [0, 1], [1, 1], 4)
    assert out["pool_of"].tolist() == [-1, -1]
    with pytest.raises(capi.K8sGpuError) as e:  # caller-sized buffer too small
        ctx.reconcile_batch([5], [0], [], [], 1, create_capacity=1)
    assert e.value.code == capi.E_CAPACITY


def test_reconcile_random_small(ctx, orc):
    rng = np.random.default_rng(7)
    for _ in range(40):
        P, V = int(rng.integers(1, 9)), int(rng.integers(0, 80))
        desired = rng.integers(0, 12, P)
        powner = rng.permutation(12)[:P].astype(np.uint32)
        vowner = rng.integers(0, 14, V).astype(np.uint32)
        vowner[rng.random(V) < 0.1] = NONE
        vflags = (rng.random(V) < 0.85).astype(np.uint32)
        ords = rng.integers(0, 16, V).astype(np.uint32)
        ords[rng.random(V) < 0.1] = NONE
        args = (desired, powner, vowner, vflags, 12, ords)
        same_reconcile(ctx.reconcile_batch(*args), orc.reconcile_batch(*args))
