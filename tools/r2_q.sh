#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
for v in m2 m3 m4; do
 for b in 128 64; do
  export SIT_LIBSHIPSIM=$PWD/sac_maritime_ast_b200/exp/lib_ex2$v.so SIT_ENVX2_BLOCK=$b
  python tools/prof_run.py env32 2>&1 | tail -1 | python -c "import sys,ast; d=ast.literal_eval(sys.stdin.read()); print('$v b$b env', d['value'])"
  python tools/prof_run.py roll 2>&1 | tail -1 | python -c "import sys,re; t=sys.stdin.read(); print('$v b$b roll', re.findall(r\"'value': ([0-9.e+]+)\", t)[0], re.findall(r\"'ms_rollout_only': ([0-9.e+]+)\", t))"
 done
done
