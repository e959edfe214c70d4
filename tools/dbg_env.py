import sys, os
sys.path[:0] = ["/root/repo", "/root/repo/tests"]
import torch, numpy as np, bench
import sac_maritime_ast_b200 as sit
dev = torch.device("cuda", 0)
env, g = bench.make_bench_env(sit, dev, 8, "fp32", 1)
a = bench.draw_waypoints(env, g, dev)
print("act", a.cpu().numpy())
sac = torch.ones(8, dtype=torch.uint8, device=dev)
for t in range(14):
    out = env.step_k(1, a, sac if t == 0 else torch.zeros_like(sac), init=None, auto_reset=True, record_all=True)
    o = out["obs"][0].cpu().numpy()
    print(t, "obsN", o[6, :4], "obsE", o[7, :4], "st", out["status"][0, :4].cpu().numpy(), "nwp", env.ship[1].n_wp[:4].cpu().numpy(), "k", env.ship[1].k[:4].cpu().numpy())
