"""Workload for the ncu passes: config 3 (10k nodes x 100k pods), tables resident in HBM,
one warm placement then one more.  Same code path as bench.py's timed step.
    python profiles/profile_step.py [workload]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import k8sgpu_b200  # noqa: E402,F401
from k8sgpu_b200 import capi, synth  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "c3"
n, m, cid = {"c3": (10_000, 100_000, 3), "c4s": (100_000, 100_000, 4), "tiny": (1000, 5000, 3)}[wl]
nodes, pods = synth.gen_nodes(n, cid), synth.gen_pods(m, cid)
with capi.Context(0) as ctx:
    ctx.place_upload(nodes, pods)
    for i in range(2):
        st = ctx.place_run()
        print(i, {k: st[k] for k in ("rounds", "placed", "evals", "kernel_launches", "device_ms", "score_ms")})
