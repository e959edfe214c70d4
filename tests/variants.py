"""Scenario constants of the model-variant fixtures (shared by tests/golden/make_golden_variants.py and the tests)."""
import ctypes as C

import numpy as np

import helpers as H
from oracle import cport

SIMPLIFIED = dict(route=[(0.0, 0.0), (2500.0, 1500.0), (3000.0, 5000.0), (6000.0, 7000.0)], psi0=0.4, u_sp=7.0, nsteps=2400,
                  modes=("MEC", "PTO"), mode=0, mode_schedule=[(1200, 1)], thrust_time_constant=30.0, kp=0.25, ki=0.012,
                  initial_thrust_force=150e3)


def oracle_simplified():
    """C oracle run of the SIMPLIFIED scenario: dict(state [T, 7], cmd [T, 2], k [T], fuel [T, 3], ei_speed [T])."""
    s = SIMPLIFIED
    L = cport.lib()
    cfg = H.default_cfg()
    p = cport.make_params(cfg, mode_names=s["modes"])
    p.mach_kind, p.thrust_time_constant, p.kp_simplified, p.ki_simplified = 1, s["thrust_time_constant"], s["kp"], s["ki"]
    ship = cport.OrcShip()
    cport.init_ship(ship, cfg, s["route"], s["psi0"], s["u_sp"], mode=s["mode"])
    ship.omega, ship.ei_shaft = s["initial_thrust_force"], 0.0
    n = s["nsteps"]
    sched = dict(s["mode_schedule"])
    state = np.zeros((n, 7)); cmd = np.zeros((n, 2)); k = np.zeros(n, np.int8); fuel = np.zeros((n, 3)); ei = np.zeros(n)
    logs = np.zeros((n, cport.N_LOG))
    rud, thr = C.c_double(), C.c_double()
    lg = (C.c_double * cport.N_LOG)()
    for t in range(n):
        if t in sched:
            ship.mode = sched[t]
        L.orc_ship_step(C.byref(p), C.byref(ship), 0, 1, lg, C.byref(rud), C.byref(thr))
        state[t] = (ship.north, ship.east, ship.psi, ship.u, ship.v, ship.r, ship.omega)
        cmd[t] = (rud.value, thr.value); k[t] = ship.k
        fuel[t] = (ship.fuel_me, ship.fuel_el, ship.fuel); ei[t] = ship.ei_speed
        logs[t] = lg[:]
    return dict(state=state, cmd=cmd, k=k, fuel=fuel, ei_speed=ei, log27=logs)
