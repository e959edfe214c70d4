#!/bin/bash
# learner evidence at batch 4096 (tensor-core path): launch list of whole updates + full ncu of the fused kernels and a split-k product
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
O=gpurun_out
python tools/prof_learner.py 4096 1 > $O/r2l_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $O/r2l_learner.csv python tools/prof_learner.py 4096 1 > $O/r2l_ncu.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"mlp_.wd_tc|gemm_tc|dw_img_tc" -s 30 -c 14 -o $O/r2l_fused python tools/prof_learner.py 4096 1 > $O/r2l_ncu2.log 2>&1
tail -2 $O/r2l_ncu2.log
python tools/exp_learner.py 2>&1 | tail -8
