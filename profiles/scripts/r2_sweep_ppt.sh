#!/bin/bash
# scoring-grid sweep at config 3 with the class-sorted order (whole placement, device ms)
for v in "" "K8SGPU_PPT=4" "K8SGPU_PPT=1" "K8SGPU_SPLITS=20" "K8SGPU_SPLITS=80" "K8SGPU_PPT=4 K8SGPU_SPLITS=40" "K8SGPU_PPT=4 K8SGPU_SPLITS=80"; do
env $v python - <<PY
import sys, os; sys.path.insert(0,".")
import k8sgpu_b200
from k8sgpu_b200 import capi, synth
nodes,pods=synth.gen_nodes(10000,3),synth.gen_pods(100000,3)
with capi.Context(0) as c:
    c.place_upload(nodes,pods)
    ms=sorted((c.place_run()["device_ms"], c.stats()["score_ms"]) for _ in range(8))
    print("$v", "best %.3f ms, median %.3f ms, score %.3f" % (ms[0][0], ms[4][0], ms[4][1]), flush=True)
PY
done
