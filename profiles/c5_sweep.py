"""BASELINE.json config 5: heterogeneous sweep at N = 10 000 nodes -- M in {10 .. 10^6} pods, SKU mix
0/50/100 % H100-like, taints on 0/5/50 % of the nodes, and the adversarial case (all nodes identical,
all pods identical) with and without the rotated tie-break.  Prints one JSON line per point:
rounds, evals, device ms, evals/s, and `oracle_equal`: for M <= 10^4 the result is compared bit for bit
with the CPU oracle run here; for the larger cases with the oracle's committed digest
(tests/golden/digests.json, computed by tests/golden/make_digests.py).   python profiles/c5_sweep.py [max_exp]
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)
import k8sgpu_b200  # noqa: E402,F401
from k8sgpu_b200 import capi, synth  # noqa: E402
import oracle as orc  # noqa: E402  (checker only)

from k8sgpu_b200.digest import place_digest  # noqa: E402

DIGESTS = json.load(open(os.path.join(ROOT, "tests", "golden", "digests.json")))["place"]
N = 10_000
max_exp = int(sys.argv[1]) if len(sys.argv) > 1 else 6
NCPU = os.cpu_count() or 1


def run(ctx, tag, nodes, pods, flags=0, check=False, digest=None):
    ctx.place_upload(nodes, pods)
    ctx.place_run(flags)
    st = ctx.place_run(flags)
    out = {"case": tag, "nodes": len(nodes["flags"]), "pods": len(pods["req_gpu"]), "rounds": st["rounds"],
           "placed": st["placed"], "evals": st["evals"], "device_ms": st["device_ms"], "score_ms": st["score_ms"],
           "evals_per_s": st["evals"] / (st["device_ms"] * 1e-3) if st["device_ms"] else None,
           "launches": st["kernel_launches"]}
    if check:
        got = ctx.place_download()
        want = orc.place(nodes, pods, flags=flags, nthreads=NCPU)
        out["oracle_equal"] = bool(all(np.array_equal(got[k], want[k]) for k in ("assign", "free_gpu", "free_mcpu", "free_mib"))
                                   and st["rounds"] == want["rounds"])
    if digest:
        d, got = DIGESTS[digest], ctx.place_download()
        out["oracle_equal"] = bool(place_digest(got) == d["sha256"] and (st["rounds"], st["placed"], st["evals"]) ==
                                   (d["rounds"], d["placed"], d["evals"]))
        out["oracle_digest"] = digest
    print(json.dumps(out), flush=True)


with capi.Context(0) as ctx:
    for e in range(1, max_exp + 1):
        M = 10 ** e
        for h100 in (0, 50, 100):
            for taint in (0, 5, 50):
                if M >= 10 ** 5 and (h100, taint) not in ((0, 0), (50, 5), (100, 50)):
                    continue  # keep the big sizes to the diagonal of the grid
                nodes = synth.gen_nodes(N, config_id=5, h100_pct=h100, taint_pct=taint)
                pods = synth.gen_pods(M, config_id=5)
                dg = f"c5_1e{e}_h{h100}_t{taint}" if M >= 10 ** 5 else None
                run(ctx, f"hetero h100={h100}% taint={taint}%", nodes, pods, check=M <= 10_000, digest=dg)
    # adversarial ties: 2 000 identical 8-GPU nodes, 16 000 identical 1-GPU pods (exactly fills them)
    nodes, pods = synth.gen_identical(2_000, 16_000)
    run(ctx, "ties rot=0 (lowest index first: one node per round)", nodes, pods, flags=1, check=True)
    run(ctx, "ties rotated", nodes, pods, flags=0, check=True)
    nodes, pods = synth.gen_identical(N, 100_000)
    run(ctx, "ties rotated, 10k nodes x 100k pods", nodes, pods, flags=0, digest="c5_ties_rotated_10k_100k")
