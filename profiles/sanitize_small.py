"""Small end-to-end pass over every kernel for compute-sanitizer (one tool per gpurun call):
    compute-sanitizer --tool memcheck python profiles/sanitize_small.py
Sizes are tiny on purpose (the tools slow kernels down 10-100x)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import k8sgpu_b200  # noqa: E402,F401
from k8sgpu_b200 import capi, synth  # noqa: E402

NONE = 0xFFFFFFFF
with capi.Context(0) as ctx:
    for n, m in ((300, 700), (257, 2049), (1, 5)):
        nodes, pods = synth.gen_nodes(n, 5), synth.gen_pods(m, 5)
        for coop in ("1", "0"):
            os.environ["K8SGPU_COOP"] = coop
            r = ctx.place(nodes, pods)
            print("place", n, m, "coop", coop, r["rounds"], r["placed"])
        gang = (np.arange(m) // 4).astype(np.uint32)
        r = ctx.place_gang(nodes, pods, gang, np.full(m // 4 + 1, 3, np.uint32))
        print("gang", r["gang_iterations"], r["placed"])
    ctx.place_upload(synth.gen_nodes(513, 5), synth.gen_pods(900, 5))
    for res in ("0", "1", "2", "3"):
        os.environ["K8SGPU_RES"] = res
        os.environ["K8SGPU_SPLITS"] = "3"
        print("score_once res", res, int(ctx.score_once().sum()))
    for p, v in ((50, 5000), (3, 10), (300, 2049)):
        t = synth.gen_reconcile(p, v)
        o = ctx.reconcile_batch(t["desired"], t["pool_owner_id"], t["vm_owner_id"], t["vm_flags"], t["n_owner_ids"],
                                t["vm_ordinal"])
        print("reconcile", p, v, len(o["victim_vm"]), len(o["create_ordinal"]))
    k, v = ctx.sort_pairs(np.arange(5000, dtype=np.uint32)[::-1] % 300, np.arange(5000, dtype=np.uint32), 9)
    print("sort ok", bool((np.diff(k.astype(np.int64)) >= 0).all()))
    out, tot = ctx.exclusive_scan(np.ones(20000, np.uint32))
    print("scan", tot)
print("done")
