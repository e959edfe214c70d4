#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
python -m pytest tests/test_gpu_env.py tests/test_gpu_rollout.py tests/test_gpu_results.py tests/test_gpu_edges.py -q -m gpu 2>&1 | tail -3
python tools/prof_run.py env32 10 2>&1 | tail -1 | python -c "import sys,ast; d=ast.literal_eval(sys.stdin.read()); print('env-bench', d['value'], d['episodes_done'])"
python tools/prof_run.py env32 3 2>&1 | tail -1 | python -c "import sys,ast; d=ast.literal_eval(sys.stdin.read()); print('env-young', d['value'])"
python tools/prof_run.py roll 2>&1 | tail -1 | python -c "import sys,re; t=sys.stdin.read(); print('roll', re.findall(r\"'value': ([0-9.e+]+)\", t)[0], re.findall(r\"'samples': ([0-9]+)\", t))"
