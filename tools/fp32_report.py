"""Per-channel fp32-vs-reference error report over the golden ship scenarios (run on the GPU box)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import numpy as np
import helpers as H, scenarios as sc, gpu_helpers as G

names = ["N", "E", "psi", "u", "v", "r", "omega"]
for name, s in sc.SHIP_SCENARIOS.items():
    g = H.golden("ship_" + name)
    out = G.run_fleet_scenario(s, n_ships=2, precision="fp32", log27=False)
    traj = out["traj"].astype(np.float64)
    rows = g["rows"]
    err = H.rel_err(traj[rows][:, :7, 0], g["state"][:, :7], 1.0)
    kdiff = int((traj[:, 7, 0].astype(np.int8) != g["k"]).sum())
    worst = err.argmax(axis=0)
    print(f"{name:22s} kdiff={kdiff} " + " ".join(f"{n}:{err[:, i].max():.1e}@{rows[worst[i]]}" for i, n in enumerate(names)))
s = sc.SHIP_SCENARIOS["mode_switching"]; g = H.golden("ship_mode_switching")
o32 = G.run_fleet_scenario(s, n_ships=2, precision="fp32", log27=False)["traj"].astype(np.float64)
o64 = G.run_fleet_scenario(s, n_ships=2, precision="fp64", log27=False)["traj"]
for t in (100, 200, 296, 300, 304, 320, 360, 392, 500, 800, 1199):
    print(t, "f64", o64[t, :7, 0], "\n    d32", o32[t, :7, 0] - o64[t, :7, 0])
