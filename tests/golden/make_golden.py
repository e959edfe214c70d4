"""Generate the golden fixtures by running the REAL reference (imported from /root/reference
behind the shims of oracle/ref_harness.py).  Build-container only; the fixtures it writes are
committed so that every test runs without the reference tree.

    python tests/golden/make_golden.py

Writes tests/golden/ship_<name>.npz and tests/golden/env_<name>.npz for the scenarios in
tests/scenarios.py.  Ship fixtures keep the waypoint index for EVERY step and the float state for
the rows listed in `rows` (every step up to 64, then every 4th, plus the last).  Env fixtures keep
every step.  The polygon sub-path goes through oracle/geometry.py (shapely is not installable
here): that part is PARITY UNPINNED.
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import ref_harness as rh  # noqa: E402
from oracle.cport import status_bits_from_string  # noqa: E402
import scenarios as sc  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))

LOG_KEYS = ['time [s]', 'north position [m]', 'east position [m]', 'yaw angle [deg]', 'rudder angle [deg]',
            'forward speed [m/s]', 'sideways speed [m/s]', 'yaw rate [deg/sec]', 'propeller shaft speed [rpm]',
            'commanded load fraction me [-]', 'commanded load fraction hsg [-]', 'power me [kw]',
            'available power me [kw]', 'power electrical [kw]', 'available power electrical [kw]', 'power [kw]',
            'propulsion power [kw]', 'fuel rate me [kg/s]', 'fuel rate hsg [kg/s]', 'fuel rate [kg/s]',
            'fuel consumption me [kg]', 'fuel consumption hsg [kg]', 'fuel consumption [kg]',
            'motor torque [Nm]', 'thrust force [kN]', 'cross track error [m]', 'heading error [deg]']


def ship_rows(nsteps):
    rows = list(range(min(64, nsteps))) + list(range(64, nsteps, 4)) + [nsteps - 1]
    return np.array(sorted(set(rows)), dtype=np.int32)


def run_ship(name, s):
    n = s["nsteps"]
    ship = rh.build_ship(s["route"], s["psi0"], s["u_sp"], mode_names=s.get("modes", ("PTI",)),
                         mode_index=s.get("mode", 0), environment=s.get("environment"), start=s.get("start"))
    dist = sc.disturbance_sequence(s["disturbance"]["seed"], n, s["disturbance"]["hold"]) if "disturbance" in s else None
    sched = dict(s.get("mode_schedule", []))
    rows = ship_rows(n)
    rowset = set(rows.tolist())
    states, cmds, ks = [], [], np.zeros(n, np.int8)
    for t in range(n):
        if t in sched:
            ship.ship_model.ship_machinery_model.mode_selector(sched[t])
        if dist is not None:
            ship.ship_model.vel_c = np.array([dist[t, 0], dist[t, 1], 0.0])
            ship.ship_model.wind_speed = dist[t, 2]
            ship.ship_model.wind_dir = dist[t, 3]
        rud, thr = rh.bare_ship_step(ship, avoid=s.get("avoid", False))
        ks[t] = ship.auto_pilot.next_wpt
        if t in rowset:
            states.append(rh.ship_state_vector(ship))
            cmds.append((rud, thr))
    res = ship.ship_model.simulation_results
    log_rows = np.array([0, 1, n // 2, n - 1])
    log27 = np.array([[float(res[k][r]) for k in LOG_KEYS] for r in log_rows])
    fuel3 = np.array([ship.ship_model.ship_machinery_model.fuel_cons_me,
                      ship.ship_model.ship_machinery_model.fuel_cons_electrical,
                      ship.ship_model.ship_machinery_model.fuel_cons])
    np.savez_compressed(os.path.join(OUT, f"ship_{name}.npz"), rows=rows, state=np.array(states),
                        cmd=np.array(cmds), k=ks, log_rows=log_rows, log27=log27, fuel3=fuel3,
                        time=ship.ship_model.int.time)
    print(f"ship_{name}: {n} steps, final k={ks[-1]}, N={states[-1][0]:.3f} E={states[-1][1]:.3f}")


def run_env_episode(env, actions, nsteps):
    obs = np.zeros((nsteps, 10)); rew = np.zeros(nsteps); done = np.zeros(nsteps, np.uint8)
    status = np.zeros(nsteps, np.int32); dist = np.zeros((nsteps, 2)); stop = np.zeros((nsteps, 2), np.uint8)
    strings = []
    for t in range(nsteps):
        a_n, a_e, sac = sc.action_at(actions, t)
        ns, r, d, st = env.step((a_n, a_e), sac, t == 0)
        obs[t], rew[t], done[t] = np.array(ns, dtype=np.float64), r, d
        status[t] = status_bits_from_string(st)
        dist[t] = (env.eps_distance_travelled, env.sampling_distance_travelled)
        stop[t] = (env.test.stop_flag, env.obs.stop_flag)
        strings.append(st)
    return dict(obs=obs, reward=rew, done=done, status=status, dist=dist, stop=stop), strings


def run_env(name, s):
    environment = sc.NO_CURRENT_ENV if s.get("no_current") else None
    env = rh.build_env(s["test_route"], s["obs_route"], s["test_psi0"], s["obs_psi0"], s["test_u_sp"],
                       s["obs_u_sp"], polygons=sc.DEFAULT_MAP, theta=s.get("theta", 2), environment=environment,
                       test_start=s.get("test_start"), obs_start=s.get("obs_start"),
                       mode_names=s.get("modes", ("PTI",)), mode_index=s.get("mode", 0))
    state0 = env.reset()
    env.init_step()
    ep, strings = run_env_episode(env, s["actions"], s["nsteps"])
    out = {f"ep1_{k}": v for k, v in ep.items()}
    out["reset_state"] = np.asarray(state0, dtype=np.float64)
    first_done = int(np.argmax(ep["done"])) if ep["done"].any() else -1
    msg = f"env_{name}: first done at step {first_done + 1 if first_done >= 0 else None}"
    if first_done >= 0:
        msg += f" [{strings[first_done].strip()}] reward {ep['reward'][first_done]:.4f}"
    if "second_episode" in s:
        state1 = env.reset()
        env.init_step()
        ep2, _ = run_env_episode(env, s["second_episode"]["actions"], s["second_episode"]["nsteps"])
        out.update({f"ep2_{k}": v for k, v in ep2.items()})
        out["reset_state2"] = np.asarray(state1, dtype=np.float64)
    np.savez_compressed(os.path.join(OUT, f"env_{name}.npz"), **out)
    seen = 0
    for st in ep["status"]:
        seen |= int(st)
    print(msg, f"; status bits seen 0x{seen:03x}")


if __name__ == "__main__":
    assert rh.available(), "reference tree not found"
    for name, s in sc.SHIP_SCENARIOS.items():
        run_ship(name, s)
    for name, s in sc.ENV_SCENARIOS.items():
        run_env(name, s)
