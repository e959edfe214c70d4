#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
O=gpurun_out
python -m pytest tests/test_gpu_bench_parity.py tests/test_gpu_fleet.py tests/test_gpu_edges.py -q -m gpu -s > $O/r2b_pytest.log 2>&1
echo "pytest rc=$?" >> $O/r2b_pytest.log
python bench.py --steps 3 --warmup 3 --no-extras > $O/r2b_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:fleet_step_x2 -s 3 -c 1 -o $O/r2b_x2 python bench.py --steps 3 --warmup 3 --no-extras > $O/r2b_ncu.log 2>&1
tail -5 $O/r2b_pytest.log; tail -3 $O/r2b_ncu.log
