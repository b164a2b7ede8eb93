"""Config 4 on one GPU: pods per thread 2 vs 4 (whole placement, device ms)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import k8sgpu_b200  # noqa
from k8sgpu_b200 import capi, synth
nodes, pods = synth.gen_nodes(100_000, 4), synth.gen_pods(1_000_000, 4)
with capi.Context(0) as c:
    c.place_upload(nodes, pods)
    for ppt, splits in ((None, None), (2, None), (4, None), (2, 10), (2, 20), (2, 40)):
        c.set_tuning(ppt=ppt, splits=splits)
        c.place_run()
        st = c.place_run()
        print(f"ppt={ppt} splits={splits}: {st['device_ms']:.2f} ms, score {st['score_ms']:.2f} ms", flush=True)
