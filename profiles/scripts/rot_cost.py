"""How much of the scoring time is the rot(p) crossing path?  Scoring rate of whole placements with the
rotated tie-break (default) and with rot = 0 for every pod (K8SGPU_POLICY_ROT_ZERO: the crossing
happens in the first 8-node chunk only), config 3 and config 4."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import k8sgpu_b200  # noqa
from k8sgpu_b200 import capi, synth
for name, n, m, cid in (("c3", 10_000, 100_000, 3), ("c4", 100_000, 1_000_000, 4)):
    nodes, pods = synth.gen_nodes(n, cid), synth.gen_pods(m, cid)
    with capi.Context(0) as c:
        c.place_upload(nodes, pods)
        for flags in (0, 1):
            c.place_run(flags, 3)
            st = c.place_run(flags, 3)  # the first three rounds only: the same pods are active in both
            print(f"{name} flags={flags}: {st['score_evals'] / st['score_ms'] / 1e9:.4f} T evals/s in the scoring launches "
                  f"({st['score_ms']:.3f} ms, {st['score_evals']:.3e} evals, rounds {st['rounds']})", flush=True)
