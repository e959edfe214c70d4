#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
O=gpurun_out
ncu --set full --clock-control none --import-source on -k regex:env_step_x2_kernel -s 30 -c 1 -o $O/r2r_rollx2 python tools/prof_run.py roll iters=12 warm_iters=20 > $O/r2r_ncu.log 2>&1
tail -2 $O/r2r_ncu.log
SIT_NO_ENV_X2=1 ncu --set full --clock-control none --import-source on -k regex:env_step_kernel -s 30 -c 1 -o $O/r2r_rollsc python tools/prof_run.py roll iters=12 warm_iters=20 > $O/r2r_ncu2.log 2>&1
tail -2 $O/r2r_ncu2.log
