#!/bin/bash
# round 2: N-GPU evidence (bench both arms, host copy floor, the training loop with the side-stream exchange)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
N=${1:-8}
O=gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
timeout 900 $T bench.py --gpus $N > $O/r2n${N}_bench.json 2> $O/r2n${N}_bench.err; echo "bench rc=$?"; tail -c 400 $O/r2n${N}_bench.err
timeout 300 $T tools/h2d_probe.py > $O/r2n${N}_h2d.log 2>&1; tail -1 $O/r2n${N}_h2d.log
timeout 400 $T test_beds/main_ast_b200.py --n_envs 65536 --num_steps 1024 --start_steps 512 --k_steps 80 --collect_iterations 8 --max_updates_per_segment 16 > $O/r2n${N}_train.log 2>&1; tail -3 $O/r2n${N}_train.log
timeout 400 $T bench.py --impl reference --gpus $N --steps 20 --warmup 3 > $O/r2n${N}_ref.json 2> $O/r2n${N}_ref.err; echo "ref rc=$?"
python - <<PY
import json
d=json.load(open('gpurun_out/r2n${N}_bench.json'))
for k in ('value','ms_per_step','e2e','strong'): print(k, d.get(k))
print('rollout', {k: d['rollout'][k] for k in ('value','value_without_gather','ms_exposed_exchange','gather_complete','gathered_rows','exchange_block_rows')})
print('cpu', d.get('cpu_baseline',{}).get('value'))
PY
