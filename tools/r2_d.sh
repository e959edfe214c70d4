#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
O=gpurun_out
python bench.py --steps 3 --warmup 3 --no-extras > $O/r2d_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:fleet_step_x2 -s 3 -c 1 -o $O/r2d_x2 python bench.py --steps 3 --warmup 3 --no-extras > $O/r2d_ncu.log 2>&1
tail -2 $O/r2d_ncu.log
