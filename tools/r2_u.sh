#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
O=gpurun_out
python -m pytest tests/test_gpu_bench_parity.py tests/test_gpu_fleet.py tests/test_gpu_edges.py tests/test_gpu_results.py -q -m gpu -s > $O/r2u_pytest.log 2>&1
echo "pytest rc=$?" >> $O/r2u_pytest.log
grep "flip rate\|bench parity\|passed\|failed\|rc=" $O/r2u_pytest.log
SIT_ENV_X2=1 python -m pytest tests/test_gpu_env.py -q -m gpu -k packed 2>&1 | tail -2
python bench.py --steps 20 --warmup 3 --no-extras 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['roofline_hbm_fp32']['frac'] if 'roofline_hbm_fp32' in d else None)"
