#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
O=gpurun_out
ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 40 -c 6 -o $O/r2l_tc python tools/prof_learner.py 4096 1 > $O/r2l_ncu.log 2>&1
tail -2 $O/r2l_ncu.log
