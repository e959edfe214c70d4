"""Small workload touching every kernel once, for compute-sanitizer (memcheck / racecheck / synccheck):
    compute-sanitizer --tool memcheck python tools/sanitize_run.py
(compute-sanitizer is closed on the shared B200 pool; tests/test_gpu_edges.py checks the output bounds with guard bands instead.)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import torch
import scenarios as sc
import gpu_helpers as G
import sac_maritime_ast_b200 as sit
from sac_maritime_ast_b200 import rollout as ro

for prec in ("fp32", "fp64"):
    fl = G.make_fleet(sc.SHIP_SCENARIOS["short_segments"], n_ships=300, precision=prec)
    fl.step(40); fl.step(40, traj=True); fl.step(10, traj=True, log27=True, cmd=True)
    fl.reset(); fl.reset(hard=True)
    env = G.make_env(sc.ENV_SCENARIOS["default"], n_envs=70, precision=prec)
    env.reset(); env.init_step()
    env.step_k(50, (8300.0, 8600.0), True, init=True, auto_reset="hard", reward_parts=True)
    env.step_k(50, (8300.0, 8600.0), True, init=None, auto_reset=True, record_all=False, want_sums=True)
    n, e = torch.rand(500, device="cuda", dtype=env.real) * 1e4, torch.rand(500, device="cuda", dtype=env.real) * 1e4
    env.map_query(n, e, want_dist=True)
big = G.make_fleet(sc.SHIP_SCENARIOS["default"], n_ships=140_001, precision="fp32")      # chunked host pipeline, ragged tail
d = torch.zeros((4, 140_001), dtype=torch.float32).pin_memory(); p = torch.empty((8, 140_001), dtype=torch.float32).pin_memory()
big.step_host(8, d, p)
for tc in (False, True):
    pol = ro.GaussianPolicy(hidden_size=64, device="cuda:0", seed=5, tensor_cores=tc); pol.w0.mul_(1e-4)
    for groups in ("1", "3"):
        os.environ["SIT_ROLLOUT_GROUPS"] = groups
        env = G.make_env(sc.ENV_SCENARIOS["default"], n_envs=900, precision="fp32")
        smp = ro.IntermediateWaypointSampler(env, pol, capacity=8192, seed=1, num_steps_episode=100, auto_reset="hard")
        smp.begin(); smp.collect(8, 16); smp.collect(3, 16, graph=False)
torch.cuda.synchronize()
print("sanitize_run ok")
