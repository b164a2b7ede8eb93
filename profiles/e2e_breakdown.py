"""Where the end-to-end time of the drop-in call goes (config 3): upload / run / download timed
separately through the resident-table entry points, against the one-call k8sgpu_place."""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import k8sgpu_b200  # noqa: E402,F401
from k8sgpu_b200 import capi, synth  # noqa: E402

nodes, pods = synth.gen_nodes(10_000, 3), synth.gen_pods(100_000, 3)
pin = lambda d: {k: torch.from_numpy(v).pin_memory().numpy() for k, v in d.items()}
hn, hp = pin(nodes), pin(pods)
R = 20
with capi.Context(0) as ctx:
    for name, n_, p_ in (("pinned", hn, hp), ("pageable", nodes, pods)):
        ctx.place(n_, p_)
        t = [0.0] * 4
        for _ in range(R):
            t0 = time.perf_counter(); ctx.place_upload(n_, p_)
            t1 = time.perf_counter(); st = ctx.place_run()
            t2 = time.perf_counter(); ctx.place_download()
            t3 = time.perf_counter(); ctx.place(n_, p_)
            t4 = time.perf_counter()
            for i, d in enumerate((t1 - t0, t2 - t1, t3 - t2, t4 - t3)):
                t[i] += d
        print(f"{name}: upload {1e3*t[0]/R:.3f} ms  run {1e3*t[1]/R:.3f} ms (device {st['device_ms']:.3f})  "
              f"download {1e3*t[2]/R:.3f} ms  | one-call place {1e3*t[3]/R:.3f} ms")
