"""Scaled sweep of the batch reconcile: V in {1e5 .. 1e8} VM rows, P = V/100
pools, tables resident in HBM.  Config 2 itself (1k x 100k) moves 1.2 MB and is launch-latency
bound, so the HBM-roofline fraction is read off the large sizes.
    python profiles/reconcile_sweep.py [max_exp]      (default 8 -> up to 1e8 rows)
Prints one JSON line per size: device ms (CUDA events), algorithmic GB/s = (12 V + 16 P) / t,
fraction of the measured HBM peak.
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import k8sgpu_b200  # noqa: E402,F401
from k8sgpu_b200 import capi, synth  # noqa: E402

try:
    PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    PEAK = 6650.0
max_exp = int(sys.argv[1]) if len(sys.argv) > 1 else 8
with capi.Context(0) as ctx:
    for e in range(5, max_exp + 1):
        V = 10 ** e
        P = V // 100
        t0 = time.time()
        t = synth.gen_reconcile(P, V)
        gen_s = time.time() - t0
        ctx.reconcile_upload(t["desired"], t["pool_owner_id"], t["vm_owner_id"], t["vm_flags"], t["n_owner_ids"],
                             t["vm_ordinal"])
        for mode, (mem, ords) in {"victims_only": (False, False), "full": (True, True)}.items():
            ms = []
            for i in range(6):
                st = ctx.reconcile_run(members=mem, ordinals=ords)
                if i:
                    ms.append(st["device_ms"])
            best = min(ms)
            alg = 12 * V + 16 * P
            print(json.dumps({"V": V, "P": P, "mode": mode, "device_ms_best": best, "device_ms_median": float(np.median(ms)),
                              "launches": st["kernel_launches"], "decisions_per_s": P / (best * 1e-3),
                              "vm_rows_per_s": V / (best * 1e-3), "algorithmic_bytes": alg,
                              "achieved_gbs": alg / (best * 1e-3) / 1e9, "frac_of_hbm_peak": alg / (best * 1e-3) / 1e9 / PEAK,
                              "hbm_peak_gbs": PEAK, "gen_seconds": round(gen_s, 1)}), flush=True)
