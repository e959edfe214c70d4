#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_env.py tests/test_gpu_rollout.py tests/test_gpu_edges.py tests/test_gpu_results.py tests/test_gpu_variants.py tests/test_gpu_train_loop.py tests/test_gpu_sac.py -q -m gpu > $O/r2p_pytest.log 2>&1
echo "pytest rc=$?" >> $O/r2p_pytest.log
grep -v "^$" $O/r2p_pytest.log | tail -40
for v in x2 scalar; do
  if [ $v = scalar ]; then export SIT_NO_ENV_X2=1; else unset SIT_NO_ENV_X2; fi
  python tools/prof_run.py env32 2>&1 | tail -1 | python -c "import sys,ast; d=ast.literal_eval(sys.stdin.read()); print('$v env', d['value'], d['episodes_done'], d['status_counts'])"
  python tools/prof_run.py roll 2>&1 | tail -1 | python -c "import sys,re; t=sys.stdin.read(); print('$v roll', re.findall(r\"'value': ([0-9.e+]+)\", t)[0], re.findall(r\"'ms_rollout_only': ([0-9.e+]+)\", t), re.findall(r\"'samples': ([0-9]+)\", t), re.findall(r\"'resamples': ([0-9]+)\", t))"
done
