"""Small single-GPU driver for ncu captures: a few launches of each hot kernel at bench sizes.
    python tools/prof_run.py fleet32|traj64|env32|env32rec|env64"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import numpy as np
import torch
import bench
import sac_maritime_ast_b200 as sit

which = sys.argv[1]
dev = torch.device("cuda", 0)
if which == "traj64":
    print(bench.traj_roofline(sit, dev))
elif which == "traj32":
    print(bench.traj_roofline(sit, dev, "fp32"))
elif which.startswith("roll"):
    kw = {}
    for a in sys.argv[2:]:
        k, v = a.split("=")
        kw[k] = v if k in ("precision", "mode") else int(v)
    r = bench.rollout_leg(sit, dev, **kw); smp = r.pop("sampler"); r.pop("memory", None); print(r)
    if which == "roll_learn":
        print(bench.learner_leg(sit, dev, smp))
elif which.startswith("env"):
    prec = "fp64" if "64" in which else "fp32"
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    print(bench.env_leg(sit, dev, B=1 << 18 if prec == "fp32" else 1 << 16, K=64, steps=steps, warmup=3, precision=prec,
                        record_all=which.endswith("rec")))
else:
    raise SystemExit("fleet32 is `python bench.py --steps 3 --warmup 3 --no-extras`")
