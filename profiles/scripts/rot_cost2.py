"""Config 3, first round only (the same pods are active either way): scoring rate with and without the rotated tie-break under several
scoring orders / kernel variants (what the rot(p) crossing path costs in each)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import k8sgpu_b200  # noqa
from k8sgpu_b200 import capi, synth
nodes, pods = synth.gen_nodes(10_000, 3), synth.gen_pods(100_000, 3)
with capi.Context(0) as c:
    c.place_upload(nodes, pods)
    for kn in (dict(), dict(order=14), dict(order=14, force_cls=17), dict(force_cls=16), dict(order=6)):
        for flags in (0, 1):
            c.reset_tuning()
            c.set_tuning(**kn)
            c.place_run(flags, 1)
            st = c.place_run(flags, 1)
            print(f"{kn} flags={flags}: {st['score_evals'] / st['score_ms'] / 1e9:.4f} T evals/s ({st['score_ms']:.3f} ms)", flush=True)
