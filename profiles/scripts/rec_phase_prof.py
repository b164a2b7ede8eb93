"""Phase times of the one-launch reconcile kernel at config 2 (stderr lines from the library) and
its device time for several cooperative grid sizes.   python profiles/scripts/rec_phase_prof.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import k8sgpu_b200  # noqa: E402,F401
from k8sgpu_b200 import capi, synth  # noqa: E402

t = synth.gen_reconcile(1000, 100000)
with capi.Context(0) as c:
    c.reconcile_upload(t["desired"], t["pool_owner_id"], t["vm_owner_id"], t["vm_flags"], t["n_owner_ids"], t["vm_ordinal"])
    c.set_tuning(rec_prof=1)
    for i in range(3):
        print("auto grid", c.reconcile_run()["device_ms"], flush=True)
    c.set_tuning(rec_prof=0)
    for members, ordinals in ((True, True), (False, False)):
        for g in (None, 37, 74, 98, 148, 296):
            c.set_tuning(rec_grid=g)
            ms = sorted(c.reconcile_run(members=members, ordinals=ordinals)["device_ms"] for _ in range(9))
            print(f"members={members} ordinals={ordinals} grid={g}: best {ms[0]*1e3:.1f} us, median {ms[4]*1e3:.1f} us", flush=True)
