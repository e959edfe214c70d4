"""Times the REAL reference (pure Python, /root/reference through oracle/ref_harness.py) in the build container: bare ship-step
loop and full MultiShipRLEnv.step.  It cannot travel to the GPU box, so this number is context for the CPU arm of bench.py
(which times the C port there), not a bench line.    python tools/time_reference_here.py"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import numpy as np
import scenarios as sc
from oracle import ref_harness as rh

assert rh.available(), "reference tree not found"
s = sc.SHIP_SCENARIOS["default"]
ship = rh.build_ship(s["route"], s["psi0"], s["u_sp"])
for _ in range(300):
    rh.bare_ship_step(ship)
t0 = time.perf_counter(); n = 4000
for _ in range(n):
    rh.bare_ship_step(ship)
dt = time.perf_counter() - t0
print(f"reference bare ship-step: {n / dt:,.0f} ship-steps/s on one core")
e = sc.ENV_SCENARIOS["default"]
env = rh.build_env(e["test_route"], e["obs_route"], e["test_psi0"], e["obs_psi0"], e["test_u_sp"], e["obs_u_sp"],
                   polygons=sc.DEFAULT_MAP, theta=e.get("theta", 2))
env.reset(); env.init_step()
a_n, a_e, sac = sc.action_at(e["actions"], 0)
env.step((a_n, a_e), sac, True)
t0 = time.perf_counter(); n = 600
for t in range(1, n + 1):
    env.step((a_n, a_e), False, False)
dt = time.perf_counter() - t0
print(f"reference MultiShipRLEnv.step (geometry restated in numpy, shapely absent): {n / dt:,.0f} env-steps/s on one core")
print(f"cpu count here: {os.cpu_count()}")
