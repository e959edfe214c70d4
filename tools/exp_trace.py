"""Timeline of a chained collect (SIT_ROLLOUT_TRACE=1, plain launches): python tools/exp_trace.py [K] [groups]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
K = int(sys.argv[1]) if len(sys.argv) > 1 else 48
os.environ["SIT_ROLLOUT_GROUPS"] = sys.argv[2] if len(sys.argv) > 2 else "3"
import ctypes as C
import torch
import bench
import sac_maritime_ast_b200 as sit
from sac_maritime_ast_b200 import _lib
dev = torch.device("cuda", 0)
r = bench.rollout_leg(sit, dev, K=K)
s = r["sampler"]
print({k: r[k] for k in ("value", "ms_per_iteration")}, flush=True)
os.environ["SIT_ROLLOUT_TRACE"] = "1"
st = torch.cuda.Stream()
torch.cuda.synchronize()
for _ in range(2):
    _lib.check(s.env.L.sit_rollout_collect(s.env.h, s.env.state_ptr(), 8, K, C.byref(s._pol), C.byref(s.ro), C.byref(s.io), 0, st.cuda_stream))
torch.cuda.synchronize()
