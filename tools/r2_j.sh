#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
O=gpurun_out
python tools/prof_learner.py 4096 1 > $O/r2j_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none --csv --log-file $O/r2j_learner_launches.csv python tools/prof_learner.py 4096 1 > $O/r2j_ncu.log 2>&1
tail -3 $O/r2j_ncu.log
python - <<'PY'
import csv
rows = [r for r in csv.reader(open("gpurun_out/r2j_learner_launches.csv")) if len(r) > 10 and r[0].isdigit()]
# last update only: find kernel names
names = [(r[4][:70], r[7], r[8], float(r[-1])) for r in rows]
# print the last 45 launches
tot = 0
for n in names[-42:]:
    print(f"{n[3]/1000:9.1f} us  grid {n[2]:>16s} block {n[1]:>12s}  {n[0]}")
    tot += n[3]
print("sum us", tot / 1000)
PY
