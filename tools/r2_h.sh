#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
O=gpurun_out
python -m pytest tests/test_gpu_rollout.py tests/test_gpu_sac.py tests/test_gpu_train_loop.py tests/test_gpu_env.py -q -m gpu -x > $O/r2h_pytest.log 2>&1
echo "pytest rc=$?" >> $O/r2h_pytest.log
tail -25 $O/r2h_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/r2h_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 $O/r2h_smoke.log
python bench.py --steps 5 --warmup 3 > $O/r2h_bench.json 2> $O/r2h_bench.err; echo "bench rc=$?"; tail -5 $O/r2h_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2h_bench.json'))
print('rollout', json.dumps(d['rollout']))
print('learner', json.dumps(d['learner']))
PY
