"""How much of the rollout's env-step time is the partial last wave (65 536 envs = 1.73 waves of 592 CTA slots)?"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import torch
import bench
import sac_maritime_ast_b200 as sit
dev = torch.device("cuda", 0)
for B, K in ((262144, 64), (262144, 32), (65536, 64), (65536, 32), (37888, 32), (75776, 32), (56832, 32)):
    r = bench.env_leg(sit, dev, B=B, K=K, steps=20, warmup=5)
    print(f"env  B={B:7d} K={K:3d}  {r['value']:.4e} env-steps/s  {r['ms_per_launch']*1e3:8.1f} us/launch", flush=True)
for B in (65536, 75776, 37888, 56832):
    r = bench.rollout_leg(sit, dev, B=B)
    print(f"roll B={B:7d}  {r['value']:.4e} samples/s  {r['env_steps_per_s']:.4e} env-steps/s  {r['ms_per_iteration']*1e3:7.1f} us/iter "
          f"samples {r['samples']}", flush=True)
