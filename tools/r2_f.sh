#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
O=gpurun_out
python -m pytest tests -q -m gpu -x > $O/r2f_pytest.log 2>&1
echo "pytest rc=$?" >> $O/r2f_pytest.log
tail -5 $O/r2f_pytest.log
python bench.py > $O/r2f_bench.json 2> $O/r2f_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 20 --warmup 3 > $O/r2f_ref.json 2> $O/r2f_ref.err; echo "ref rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2f_bench.json'))
for k in ('value','ms_per_step','e2e','roofline','strong','cpu_baseline','roofline_hbm_fp32','roofline_hbm'):
    print(k, d.get(k))
print('ast_env', d['ast_env']['value'], 'rollout', d['rollout']['value'], 'learner', d['learner']['batch_64']['updates_per_s'], d['learner']['batch_4096']['updates_per_s'])
r=json.load(open('gpurun_out/r2f_ref.json')); print('ref', r['value'], r['cpu_baseline']['kind'], r['cpu_baseline']['cores'], r['cpu_baseline'].get('port',{}).get('value'))
PY
