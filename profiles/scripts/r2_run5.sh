#!/bin/bash
# scoring-kernel class experiment: time of one scoring pass when every warp runs eval_pair class 0..3
for c in -1 0 1 2 3; do K8SGPU_FORCE_CLS=$c python profiles/score_only.py c3 20 2>&1 | tail -1; done
for c in -1 0 3; do K8SGPU_FORCE_CLS=$c python profiles/score_only.py c4s 5 2>&1 | tail -1; done
python -m pytest tests/test_gpu_parity.py -x -q 2>&1 | tail -3
