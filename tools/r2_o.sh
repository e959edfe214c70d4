#!/bin/bash
# K2 register-cap experiments + host copy floor at N=1
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
O=gpurun_out
python tools/h2d_probe.py > $O/r2o_h2d_n1.log 2>&1; tail -1 $O/r2o_h2d_n1.log
for v in base minb5 minb6; do
  if [ $v = base ]; then L=$PWD/sac_maritime_ast_b200/libshipsim.so; else L=$PWD/sac_maritime_ast_b200/exp/lib_$v.so; fi
  SIT_LIBSHIPSIM=$L python tools/prof_run.py env32 2>&1 | tail -1 | python -c "import sys,ast; d=ast.literal_eval(sys.stdin.read()); print('$v env', d['value'])"
  SIT_LIBSHIPSIM=$L python tools/prof_run.py roll 2>&1 | tail -1 | python -c "import sys,re; t=sys.stdin.read(); print('$v roll', re.findall(r\"'value': ([0-9.e+]+)\", t)[0], re.findall(r\"'ms_rollout_only': ([0-9.e+]+)\", t))"
done
