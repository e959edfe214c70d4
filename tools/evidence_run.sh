# round evidence, part A: tests, smoke, both bench arms, ncu launch lists, full captures of the K1 kernels
# (gpurun copies back at most 64 MiB per call, hence two parts)
mkdir -p gpurun_out
T=${1:-s41}
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; tail -3 gpurun_out/${T}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; tail -1 gpurun_out/${T}_smoke.log
timeout 400 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err
timeout 600 python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; tail -c 300 gpurun_out/${T}_bench.err
M="--metrics gpu__time_duration.sum --clock-control none --csv"
timeout 300 ncu $M -c 800 --log-file gpurun_out/${T}_launches_bench.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_ncu_bench.log 2>&1
timeout 300 ncu $M -c 800 --log-file gpurun_out/${T}_launches_roll.csv python tools/prof_run.py roll iters=10 warm_iters=10 > gpurun_out/${T}_ncu_roll.log 2>&1
F="--set full --clock-control none --import-source on"
timeout 300 ncu $F -k regex:fleet_step_fast -s 3 -c 2 -o gpurun_out/${T}_fleet32_full -f python bench.py --steps 3 --warmup 3 --no-extras > gpurun_out/${T}_ncu_f32.log 2>&1
timeout 300 ncu $F -k regex:fleet_step_fast -s 3 -c 2 -o gpurun_out/${T}_traj64_full -f python tools/prof_run.py traj64 > gpurun_out/${T}_ncu_t64.log 2>&1
timeout 300 ncu $F -k regex:fleet_step_fast -s 3 -c 2 -o gpurun_out/${T}_traj32_full -f python tools/prof_run.py traj32 > gpurun_out/${T}_ncu_t32.log 2>&1
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > gpurun_out/${T}_smi.txt
ls -la gpurun_out | tail -20
