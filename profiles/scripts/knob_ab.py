"""Config 3 (or 4 with argv[1] = 4) on one GPU under a list of knob settings: whole placement, device ms."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import k8sgpu_b200  # noqa
from k8sgpu_b200 import capi, synth
cfg = int(sys.argv[1]) if len(sys.argv) > 1 else 3
n, m = (10_000, 100_000) if cfg == 3 else (100_000, 1_000_000)
nodes, pods = synth.gen_nodes(n, cfg), synth.gen_pods(m, cfg)
SETS = [dict(), dict(splits=20), dict(splits=40), dict()]
if cfg == 4:
    SETS = [dict(), dict(splits=10)]
with capi.Context(0) as c:
    c.place_upload(nodes, pods)
    for kn in SETS:
        c.reset_tuning()
        c.set_tuning(**kn)
        c.place_run()
        ms = []
        for _ in range(5):
            st = c.place_run()
            ms.append((st["device_ms"], st["score_ms"]))
        ms.sort()
        print(f"{kn}: {ms[2][0]:.3f} ms, score {ms[2][1]:.3f} ms", flush=True)
