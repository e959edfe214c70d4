"""The benchmark's own workload (bench.make_routes / bench.make_dist: BASELINE config 3) mirrored on the C oracle and on a
ShipFleet, so that the parity tests run the exact configuration the headline number is measured on."""
from __future__ import annotations

import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import bench  # noqa: E402  (numpy-only at import time)
import helpers as H  # noqa: E402
from oracle import cport  # noqa: E402

ALL_MODES = ("PTO", "PTI", "MEC")


def make_inputs(n, seed, max_wp=8, n_launches=4, multi_mode=False, u_sp_range=None):
    """Routes, set points and one disturbance block per launch -- float32-representable, as bench.py draws them."""
    rng = np.random.default_rng(seed)
    n_wp, wp_n, wp_e, psi0, u_sp = bench.make_routes(n, max_wp, rng)
    if u_sp_range is not None:
        u_sp = rng.uniform(u_sp_range[0], u_sp_range[1], n).astype(np.float32)
    dists = [bench.make_dist(n, rng) for _ in range(n_launches)]
    mode = rng.integers(0, 3, n).astype(np.int32) if multi_mode else np.zeros(n, np.int32)
    return dict(n=n, max_wp=max_wp, n_wp=n_wp, wp_n=wp_n, wp_e=wp_e, psi0=psi0, u_sp=u_sp, dists=dists, mode=mode,
                multi_mode=multi_mode)


def oracle_ships(inp):
    """(OrcParams, OrcShip array) holding the same fleet in float64."""
    n, max_wp = inp["n"], inp["max_wp"]
    cfg = H.default_cfg()
    p = cport.make_params(cfg, mode_names=ALL_MODES if inp["multi_mode"] else ("PTI",))
    ships = (cport.OrcShip * n)()
    proto = cport.OrcShip()
    cport.init_ship(proto, cfg, [(0, 0), (1, 1)], 0.0, 8.0)
    arr = np.ctypeslib.as_array(ships)
    arr[:] = np.ctypeslib.as_array((cport.OrcShip * 1)(proto))[0]
    arr["north"], arr["east"] = inp["wp_n"][0], inp["wp_e"][0]
    arr["psi"], arr["u_sp"] = inp["psi0"], inp["u_sp"]
    arr["wp_n"][:, :max_wp] = inp["wp_n"].T; arr["wp_e"][:, :max_wp] = inp["wp_e"].T
    arr["n_wp"] = inp["n_wp"]
    arr["mode"] = inp["mode"]
    return p, ships, arr


def oracle_run(p, ships, arr, dist, K, avoid=False, traj=False, threads=0):
    """One launch worth of steps: disturbances held for K steps.  Returns [K, 8, n] when traj."""
    n = len(arr)
    d = dist.astype(np.float64)
    arr["vc_n"], arr["vc_e"], arr["wind_speed"], arr["wind_dir"] = d[0], d[1], d[2], d[3]
    out = np.zeros((K, 8, n)) if traj else None
    cport.lib().orc_run_ships(C.byref(p), ships, n, K, int(avoid), None, out.ctypes.data if traj else None, threads)
    return out


def oracle_pose(arr):
    return np.stack([arr["north"], arr["east"], arr["psi"], arr["u"], arr["v"], arr["r"], arr["omega"],
                     arr["k"].astype(np.float64)])


def make_fleet(inp, precision, device="cuda:0"):
    import torch
    import sac_maritime_ast_b200 as sit
    kw = dict(mode_names=ALL_MODES, mode_index=0) if inp["multi_mode"] else {}
    sm, tc, ap = sit.default_ship([(0, 0), (1000, 1000)], 0.0, **kw)
    fleet = sit.ShipFleet(sm, tc, ap, 8.0, n_ships=inp["n"], device=device, precision=precision, max_wp=inp["max_wp"])
    real = fleet.real
    fleet.set_routes(torch.from_numpy(inp["wp_n"]).to(device, real), torch.from_numpy(inp["wp_e"]).to(device, real),
                     torch.from_numpy(inp["n_wp"]).to(device))
    fleet.f["north"].copy_(torch.from_numpy(inp["wp_n"][0].astype(np.float64)))
    fleet.f["east"].copy_(torch.from_numpy(inp["wp_e"][0].astype(np.float64)))
    fleet.f["psi"].copy_(torch.from_numpy(inp["psi0"]).to(real)); fleet.f["u_sp"].copy_(torch.from_numpy(inp["u_sp"]).to(real))
    fleet.f["mode"].copy_(torch.from_numpy(inp["mode"]))
    fleet.capture_init()
    return fleet
