"""Batch reconcile at V rows, P = V/100 pools, tables resident: victims only, then with member lists.
    python profiles/scripts/rec_large_run.py V [runs]     (run it under ncu for the per-kernel launch list)"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import k8sgpu_b200  # noqa: E402,F401
from k8sgpu_b200 import capi, synth  # noqa: E402

V = int(float(sys.argv[1])) if len(sys.argv) > 1 else 10 ** 7
runs = int(sys.argv[2]) if len(sys.argv) > 2 else 3
t = synth.gen_reconcile(V // 100, V)
with capi.Context(0) as c:
    c.reconcile_upload(t["desired"], t["pool_owner_id"], t["vm_owner_id"], t["vm_flags"], t["n_owner_ids"], t["vm_ordinal"])
    for mode in ((False, False), (True, True)):
        for i in range(runs):
            st = c.reconcile_run(members=mode[0], ordinals=mode[1])
            print(f"V={V} members={mode[0]} ordinals={mode[1]} run {i}: {st['device_ms']:.3f} ms, {st['kernel_launches']} launches", flush=True)
