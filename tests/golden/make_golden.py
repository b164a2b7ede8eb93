"""Regenerates tests/golden/*.npz from the CPU oracle (oracle/liboracle.so).

The reference holds no golden vectors and cannot be run or imported
(Markdown only), so these fixtures freeze policy v0 as the oracle computes it today: a change
to either the oracle or the CUDA path that alters any output shows up as a diff here.
Run from the repo root:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle")]
import k8sgpu_b200  # noqa: E402,F401
from k8sgpu_b200 import synth  # noqa: E402
import oracle as orc  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))

# name -> (generator description, kwargs).  Inputs are regenerated from these parameters.
PLACE_CASES = {
    "c3_small": dict(n=600, m=3000, config_id=3, h100_pct=30, taint_pct=5, flags=0),
    "c5_all_h100": dict(n=300, m=2500, config_id=5, h100_pct=100, taint_pct=0, flags=0),
    "c5_all_t4_taint50": dict(n=500, m=1500, config_id=5, h100_pct=0, taint_pct=50, flags=0),
    "c5_rot_zero": dict(n=257, m=1111, config_id=5, h100_pct=50, taint_pct=5, flags=1),
}
TIES_CASES = {
    "ties_rot_zero": dict(n=33, m=300, flags=1),
    "ties_rotated": dict(n=300, m=2500, flags=0),
}
RECONCILE_CASES = {"c2_small": dict(p=50, v=5000, config_id=2), "c2_tiny_pools": dict(p=3, v=700, config_id=2)}


def place_inputs(name):
    if name in PLACE_CASES:
        c = PLACE_CASES[name]
        return (synth.gen_nodes(c["n"], c["config_id"], h100_pct=c["h100_pct"], taint_pct=c["taint_pct"]),
                synth.gen_pods(c["m"], c["config_id"]), c["flags"])
    c = TIES_CASES[name]
    nodes, pods = synth.gen_identical(c["n"], c["m"])
    return nodes, pods, c["flags"]


def main():
    for name in list(PLACE_CASES) + list(TIES_CASES):
        nodes, pods, flags = place_inputs(name)
        r = orc.place(nodes, pods, flags=flags)
        best0 = orc.score_round(nodes, (nodes["free_gpu"], nodes["free_mcpu"], nodes["free_mib"]), pods,
                                np.arange(len(pods["req_gpu"]), dtype=np.uint32), flags=flags)
        np.savez_compressed(os.path.join(HERE, f"place_{name}.npz"), assign=r["assign"], free_gpu=r["free_gpu"],
                            free_mcpu=r["free_mcpu"], free_mib=r["free_mib"], best_round0=best0,
                            rounds=np.int64(r["rounds"]), placed=np.int64(r["placed"]), evals=np.int64(r["evals"]))
    for name, c in RECONCILE_CASES.items():
        t = synth.gen_reconcile(c["p"], c["v"], c["config_id"])
        o = orc.reconcile_batch(t["desired"], t["pool_owner_id"], t["vm_owner_id"], t["vm_flags"], t["n_owner_ids"],
                                t["vm_ordinal"])
        np.savez_compressed(os.path.join(HERE, f"reconcile_{name}.npz"), **o)
    print("wrote", sorted(f for f in os.listdir(HERE) if f.endswith(".npz")))


if __name__ == "__main__":
    main()
