"""Config 3 on one GPU with the admission kernel's phase clock on (coop_prof): microseconds per phase, per round (stderr)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import k8sgpu_b200  # noqa
from k8sgpu_b200 import capi, synth
nodes, pods = synth.gen_nodes(10_000, 3), synth.gen_pods(100_000, 3)
with capi.Context(0) as c:
    c.place_upload(nodes, pods)
    c.place_run()
    c.set_tuning(coop_prof=1)
    st = c.place_run()
    print(st["device_ms"], st["score_ms"], st["rounds"])
