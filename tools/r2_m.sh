#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
O=gpurun_out
python tools/prof_run.py env32 > $O/r2m_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:env_step_kernel -s 4 -c 1 -o $O/r2m_env32 python tools/prof_run.py env32 > $O/r2m_ncu.log 2>&1
tail -2 $O/r2m_ncu.log; tail -1 $O/r2m_plain.log | head -c 600
