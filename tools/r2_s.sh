#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
O=gpurun_out
SIT_ENV_X2=1 ncu --set full --clock-control none --import-source on -k regex:env_step_x2_kernel -s 11 -c 1 -o $O/r2s_x2late python tools/prof_run.py env32 10 > $O/r2s_ncu.log 2>&1
tail -1 $O/r2s_ncu.log | cut -c1-200
SIT_ENV_X2=0 ncu --set full --clock-control none --import-source on -k regex:env_step_kernel -s 11 -c 1 -o $O/r2s_sclate python tools/prof_run.py env32 10 > $O/r2s_ncu2.log 2>&1
tail -1 $O/r2s_ncu2.log | cut -c1-200
