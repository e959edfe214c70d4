"""Launch list of one large-batch learner update (run under ncu --metrics gpu__time_duration.sum)."""
import sys
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from sac_maritime_ast_b200 import sac
b = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
tc = bool(int(sys.argv[2])) if len(sys.argv) > 2 else True
rows = torch.zeros(65536, 24); rows[:, :23] = torch.randn(65536, 23); rows[:, 10].clamp_(-1, 1); rows[:, 22] = 1
mem = sac.ReplayMemory(65536, device="cuda:0"); mem.extend(rows.cuda())
a = sac.SAC(10, None, None, device="cuda:0", hidden_size=256, batch_size=b, seed=1, tensor_cores=tc, obs_scale=None, use_graph=False)
a.update(mem, 3)
torch.cuda.synchronize()
