#!/bin/bash
# N-GPU bench line + training loop only (the host copy probe and the reference arm are unchanged since r2_n8.sh)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
N=${1:-8}
O=gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 900 $T bench.py --gpus $N > $O/r2n${N}_bench.json 2> $O/r2n${N}_bench.err; echo "bench rc=$?"; tail -c 300 $O/r2n${N}_bench.err
timeout 400 $T test_beds/main_ast_b200.py --n_envs 65536 --num_steps 1024 --start_steps 512 --k_steps 80 --collect_iterations 8 --max_updates_per_segment 16 > $O/r2n${N}_train.log 2>&1; tail -3 $O/r2n${N}_train.log
python - <<PY
import json
d=json.load(open('gpurun_out/r2n${N}_bench.json'))
for k in ('value','ms_per_step','e2e','strong'): print(k, d.get(k))
print('rollout', {k: d['rollout'][k] for k in ('value','value_without_gather','ms_exposed_exchange','gather_complete','gathered_rows','exchange_block_rows','rows_left_behind')})
print('long', d['long_horizon']['value'])
PY
