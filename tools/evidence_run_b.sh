# round evidence, part B: full captures of the env-step and rollout kernels
mkdir -p gpurun_out
T=${1:-s42}
F="--set full --clock-control none --import-source on"
timeout 300 ncu $F -k regex:env_step_kernel -s 3 -c 2 -o gpurun_out/${T}_env32_full -f python tools/prof_run.py env32 > gpurun_out/${T}_ncu_e32.log 2>&1
timeout 400 ncu $F -k regex:"rollout_|env_step" -s 112 -c 11 -o gpurun_out/${T}_roll_full -f python tools/prof_run.py roll iters=10 warm_iters=10 > gpurun_out/${T}_ncu_roll_full.log 2>&1
ls -la gpurun_out | tail -20
