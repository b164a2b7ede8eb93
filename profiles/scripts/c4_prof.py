"""Config 4 on one GPU with the admission kernel's phase times (stderr), order on and off."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import k8sgpu_b200  # noqa
from k8sgpu_b200 import capi, synth
nodes, pods = synth.gen_nodes(100_000, 4), synth.gen_pods(1_000_000, 4)
with capi.Context(0) as c:
    c.place_upload(nodes, pods)
    for order in (1, 0):
        c.set_tuning(order=order, coop_prof=0)
        c.place_run()
        st = c.place_run()
        print(f"order={order}: {st['device_ms']:.2f} ms, score {st['score_ms']:.2f} ms, rounds {st['rounds']}", flush=True)
        c.set_tuning(coop_prof=1)
        c.place_run()
