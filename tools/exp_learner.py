"""Learner updates/s at several batch sizes (tensor-core path): python tools/exp_learner.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sac_maritime_ast_b200 import sac
rows = torch.zeros(262144, 24); rows[:, :23] = torch.randn(262144, 23); rows[:, 10].clamp_(-1, 1); rows[:, 22] = 1
mem = sac.ReplayMemory(262144, device="cuda:0"); mem.extend(rows.cuda())
for b, tcv in ((64, True), (64, 2), (256, True), (256, 2), (1024, True), (2048, True), (4096, True), (8192, True)):
    a = sac.SAC(10, None, None, device="cuda:0", hidden_size=256, batch_size=b, seed=1, tensor_cores=tcv, obs_scale=None)
    a.update(mem, 50); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); a.update(mem, 400); e1.record(); torch.cuda.synchronize()
    print(f"batch {b} tc={tcv} kernels {a.kernels_per_update}: {400 / (e0.elapsed_time(e1) * 1e-3):.0f} updates/s", flush=True)
