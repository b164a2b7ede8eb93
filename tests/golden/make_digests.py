"""Full-size result digests from the CPU oracle -> tests/golden/digests.json.

BASELINE.json's configs 3, 4 and 5 at the sizes it quotes are too large to keep as fixtures
(config 4: 1 M assignments + 3 x 100 k free columns), so what is committed is one sha256 per case
(k8sgpu_b200/digest.py) plus rounds / placed / evals, all computed here by the CPU ORACLE ONLY --
no GPU is involved in producing them.  tests/test_gpu_digests.py recomputes the digest of the
CUDA path's result on the GPU box and compares; bench.py prints the digest of what it timed.

The reference holds no vectors of its own (Markdown only): these digests
freeze the oracle's answer, they do not pin the oracle to the reference.

    python tests/golden/make_digests.py [case ...]     (no arguments: every case; ~15-25 min on 8 cores)

Cases already present in digests.json are kept unless named on the command line.
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle")]
import k8sgpu_b200  # noqa: E402,F401
from k8sgpu_b200 import synth  # noqa: E402
from k8sgpu_b200.digest import place_digest, reconcile_digest  # noqa: E402
import oracle as orc  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "digests.json")

# name -> how to regenerate the inputs (k8sgpu_b200.synth) and the policy flags
PLACE_CASES = {
    "c3": dict(kind="hetero", n=10_000, m=100_000, config_id=3, h100_pct=30, taint_pct=5, flags=0),
    "c4": dict(kind="hetero", n=100_000, m=1_000_000, config_id=4, h100_pct=30, taint_pct=5, flags=0),
    "c5_1e5_h0_t0": dict(kind="hetero", n=10_000, m=100_000, config_id=5, h100_pct=0, taint_pct=0, flags=0),
    "c5_1e5_h50_t5": dict(kind="hetero", n=10_000, m=100_000, config_id=5, h100_pct=50, taint_pct=5, flags=0),
    "c5_1e5_h100_t50": dict(kind="hetero", n=10_000, m=100_000, config_id=5, h100_pct=100, taint_pct=50, flags=0),
    "c5_1e6_h0_t0": dict(kind="hetero", n=10_000, m=1_000_000, config_id=5, h100_pct=0, taint_pct=0, flags=0),
    "c5_1e6_h50_t5": dict(kind="hetero", n=10_000, m=1_000_000, config_id=5, h100_pct=50, taint_pct=5, flags=0),
    "c5_1e6_h100_t50": dict(kind="hetero", n=10_000, m=1_000_000, config_id=5, h100_pct=100, taint_pct=50, flags=0),
    "c5_ties_rotated_10k_100k": dict(kind="ties", n=10_000, m=100_000, flags=0),
    "c5_ties_rot0_2k_16k": dict(kind="ties", n=2_000, m=16_000, flags=1),
}
RECONCILE_CASES = {
    "c2": dict(p=1_000, v=100_000, config_id=2),
    "c2_x100": dict(p=100_000, v=10_000_000, config_id=2),
}


def place_inputs(c):
    if c["kind"] == "ties":
        nodes, pods = synth.gen_identical(c["n"], c["m"])
        return nodes, pods
    return (synth.gen_nodes(c["n"], c["config_id"], h100_pct=c["h100_pct"], taint_pct=c["taint_pct"]),
            synth.gen_pods(c["m"], c["config_id"]))


def reconcile_inputs(c):
    t = synth.gen_reconcile(c["p"], c["v"], c["config_id"])
    return (t["desired"], t["pool_owner_id"], t["vm_owner_id"], t["vm_flags"], t["n_owner_ids"], t["vm_ordinal"])


def main():
    want = sys.argv[1:]
    try:
        with open(OUT) as f:
            doc = json.load(f)
    except Exception:
        doc = {"about": "sha256 digests of the CPU oracle's results (tests/golden/make_digests.py)",
               "place": {}, "reconcile": {}}
    nt = os.cpu_count() or 1
    for name, c in PLACE_CASES.items():
        if (want and name not in want) or (not want and name in doc["place"]):
            continue
        nodes, pods = place_inputs(c)
        t0 = time.perf_counter()
        r = orc.place(nodes, pods, flags=c["flags"], nthreads=nt)
        dt = time.perf_counter() - t0
        doc["place"][name] = {"inputs": c, "sha256": place_digest(r), "rounds": int(r["rounds"]),
                              "placed": int(r["placed"]), "evals": int(r["evals"]),
                              "oracle_seconds": round(dt, 1), "oracle_threads": nt}
        print(name, doc["place"][name], flush=True)
        with open(OUT, "w") as f:
            json.dump(doc, f, indent=1, sort_keys=True)
    for name, c in RECONCILE_CASES.items():
        if (want and name not in want) or (not want and name in doc["reconcile"]):
            continue
        o = orc.reconcile_batch(*reconcile_inputs(c))
        doc["reconcile"][name] = {"inputs": c, "sha256": reconcile_digest(o), "victims": int(len(o["victim_vm"])),
                                  "creates": int(len(o["create_ordinal"])), "members": int(len(o["member_vm"]))}
        print(name, doc["reconcile"][name], flush=True)
        with open(OUT, "w") as f:
            json.dump(doc, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
