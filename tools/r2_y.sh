#!/bin/bash
# env32 full ncu capture (K2 at the bench env leg) + launch list of the rollout leg
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
O=gpurun_out
bash tools/r2_m.sh
python tools/prof_run.py roll > $O/r2y_roll_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file $O/r2y_roll_launches.csv python tools/prof_run.py roll > $O/r2y_roll_ncu.log 2>&1
tail -1 $O/r2y_roll_plain.log | head -c 400
