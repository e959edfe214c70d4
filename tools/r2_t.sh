#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
O=gpurun_out
SIT_ENV_X2=1 timeout 900 python -m pytest tests/test_gpu_env.py tests/test_gpu_rollout.py tests/test_gpu_train_loop.py tests/test_gpu_edges.py -q -m gpu > $O/r2t_pytest.log 2>&1
echo "pytest rc=$?" >> $O/r2t_pytest.log
grep -v "^$" $O/r2t_pytest.log | tail -15
for v in scalar m2 m3 m4; do
  if [ $v = scalar ]; then export SIT_ENV_X2=0 SIT_LIBSHIPSIM=$PWD/sac_maritime_ast_b200/exp/lib_ex2m4.so; else export SIT_ENV_X2=1 SIT_LIBSHIPSIM=$PWD/sac_maritime_ast_b200/exp/lib_ex2$v.so; fi
  python tools/prof_run.py env32 3 2>&1 | tail -1 | python -c "import sys,ast; d=ast.literal_eval(sys.stdin.read()); print('$v env-young', d['value'])"
  python tools/prof_run.py env32 10 2>&1 | tail -1 | python -c "import sys,ast; d=ast.literal_eval(sys.stdin.read()); print('$v env-bench', d['value'], d['episodes_done'])"
  python tools/prof_run.py roll 2>&1 | tail -1 | python -c "import sys,re; t=sys.stdin.read(); print('$v roll', re.findall(r\"'value': ([0-9.e+]+)\", t)[0], re.findall(r\"'ms_rollout_only': ([0-9.e+]+)\", t), re.findall(r\"'samples': ([0-9]+)\", t))"
done
