"""Times ONE scoring pass (round 1 of config 3 or 4, every pod active) through the round API with
CUDA events, for kernel-variant sweeps (env K8SGPU_RES / K8SGPU_PPT / K8SGPU_SPLITS).
    python profiles/score_only.py [c3|c4s] [reps]
Prints evals/s of the scoring launch alone.  Not a bench line."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import k8sgpu_b200  # noqa: E402,F401
from k8sgpu_b200 import capi, synth  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "c3"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
n, m, cid = {"c3": (10_000, 100_000, 3), "c4s": (100_000, 200_000, 4)}[wl]
nodes, pods = synth.gen_nodes(n, cid), synth.gen_pods(m, cid)
with capi.Context(0) as ctx:
    ctx.place_upload(nodes, pods)
    s = torch.cuda.current_stream().cuda_stream
    buf = torch.empty(m, dtype=torch.int32, device="cuda:0")
    ctx.round_begin(stream=s)
    for _ in range(3):
        ctx.round_score(0, m, buf.data_ptr(), s)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        ctx.round_score(0, m, buf.data_ptr(), s)
    e1.record()
    e1.synchronize()
    ms = e0.elapsed_time(e1) / reps
    ctx.round_end(s)
    tag = " ".join(f"{k}={os.environ[k]}" for k in sorted(os.environ) if k.startswith("K8SGPU_"))
    print(f"{wl} {tag}: {ms:.4f} ms per scoring pass (incl. node prep + decode), {n * m / ms / 1e9:.1f} G evals/s, "
          f"checksum {int(buf.sum().item())}")
