#!/bin/bash
# round 2 re-entry: full GPU suite, bench (both arms), launch list + full ncu of the packed K1 kernel
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
O=gpurun_out
python -m pytest tests -q -m gpu -s > $O/r2g_pytest.log 2>&1
echo "pytest rc=$?" >> $O/r2g_pytest.log
tail -8 $O/r2g_pytest.log
python bench.py > $O/r2g_bench.json 2> $O/r2g_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 20 --warmup 3 > $O/r2g_ref.json 2> $O/r2g_ref.err; echo "ref rc=$?"
python bench.py --steps 3 --warmup 3 --no-extras > $O/r2g_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2g_launches.csv python bench.py --steps 3 --warmup 3 --no-extras > $O/r2g_ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:fleet_step_x2 -s 3 -c 1 -o $O/r2g_x2 python bench.py --steps 3 --warmup 3 --no-extras > $O/r2g_ncu.log 2>&1
tail -2 $O/r2g_ncu.log
cat $O/r2g_bench.json | head -c 6000
