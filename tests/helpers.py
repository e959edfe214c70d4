"""Shared test helpers: default parameter table, scenario -> oracle objects, error metrics."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

import scenarios as sc
from oracle import cport

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

# per-channel floors for relative error: v and r are ~1e-4..1e-8 early on
STATE_FLOORS = np.array([1.0, 1.0, 1.0, 1.0, 1e-3, 1e-3, 1.0, 1.0, 1.0, 1.0, 1.0, 1.0, 1.0, 1e-3, 1.0])


def default_cfg():
    """Default parameter table (values of test_beds/main_ast.py:106-237) as a plain dict."""
    return dict(
        ship=dict(dead_weight_tonnage=3850000, coefficient_of_deadweight_to_displacement=0.7,
                  bunkers=200000, ballast=200000, length_of_ship=80, width_of_ship=16,
                  added_mass_coefficient_in_surge=0.4, added_mass_coefficient_in_sway=0.4,
                  added_mass_coefficient_in_yaw=0.4,
                  mass_over_linear_friction_coefficient_in_surge=130,
                  mass_over_linear_friction_coefficient_in_sway=18,
                  mass_over_linear_friction_coefficient_in_yaw=90,
                  nonlinear_friction_coefficient__in_surge=2400,
                  nonlinear_friction_coefficient__in_sway=4000,
                  nonlinear_friction_coefficient__in_yaw=400),
        environment=dict(current_velocity_component_from_north=-2.0,
                         current_velocity_component_from_east=-2.0,
                         wind_speed=2.0, wind_direction=-np.pi / 4),
        modes=dict(PTO=(2160e3, 0.0, "GEN"), PTI=(0.0, 2 * 510e3, "MOTOR"), MEC=(2160e3, 510e3, "OFF")),
        machinery=dict(hotel_load=200000, rated_speed_main_engine_rpm=1000,
                       linear_friction_main_engine=68, linear_friction_hybrid_shaft_generator=57,
                       gear_ratio_between_main_engine_and_propeller=0.6,
                       gear_ratio_between_hybrid_shaft_generator_and_propeller=0.6,
                       propeller_inertia=6000, propeller_speed_to_torque_coefficient=7.5,
                       propeller_diameter=3.1, propeller_speed_to_thrust_force_coefficient=1.7,
                       rudder_angle_to_sway_force_coefficient=50e3,
                       rudder_angle_to_yaw_force_coefficient=500e3, max_rudder_angle_degrees=30),
        sfc_me=(128.9, -168.9, 246.8), sfc_dg=(108.7, -289.9, 324.9),
        throttle=dict(kp_ship_speed=7, ki_ship_speed=0.13, kp_shaft_speed=0.05, ki_shaft_speed=0.005),
        initial_shaft_speed_integral_error=114,
        heading=dict(kp=1, kd=90, ki=0.01),
        los=dict(radius_of_acceptance=300, lookahead_distance=1000, integral_gain=0.002,
                 integrator_windup_limit=4000),
        initial_propeller_shaft_speed_rad_per_s=400 * np.pi / 30,
        dt=0.5, simulation_time=3600,
    )


def golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


def rel_err(a, b, floors):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return np.abs(a - b) / np.maximum(np.abs(b), floors)


def oracle_ship(s):
    cfg = default_cfg()
    p = cport.make_params(cfg, mode_names=s.get("modes", ("PTI",)))
    ship = cport.OrcShip()
    cport.init_ship(ship, cfg, s["route"], s["psi0"], s["u_sp"], mode=s.get("mode", 0),
                    environment=s.get("environment"), start=s.get("start"))
    return p, ship


def oracle_env(s):
    cfg = default_cfg()
    p = cport.make_params(cfg, mode_names=s.get("modes", ("PTI",)))
    m = cport.make_map(sc.DEFAULT_MAP)
    e = cport.OrcEnv()
    cport.init_env(e, cfg, s["test_route"], s["obs_route"], s["test_psi0"], s["obs_psi0"], s["test_u_sp"],
                   s["obs_u_sp"], theta=s.get("theta", 2), mode=s.get("mode", 0),
                   environment=sc.NO_CURRENT_ENV if s.get("no_current") else None,
                   test_start=s.get("test_start"), obs_start=s.get("obs_start"))
    return p, m, e


def run_oracle_ship(s):
    """Returns dict(state[T,15], cmd[T,2], k[T]) for every step (oracle, C float64)."""
    L = cport.lib()
    p, ship = oracle_ship(s)
    n = s["nsteps"]
    dist = sc.disturbance_sequence(s["disturbance"]["seed"], n, s["disturbance"]["hold"]) if "disturbance" in s else None
    sched = dict(s.get("mode_schedule", []))
    state = np.zeros((n, 15)); cmd = np.zeros((n, 2)); k = np.zeros(n, np.int8)
    logs = np.zeros((n, cport.N_LOG))
    rud, thr = C.c_double(), C.c_double()
    lg = (C.c_double * cport.N_LOG)()
    for t in range(n):
        if t in sched:
            ship.mode = sched[t]
        if dist is not None:
            ship.vc_n, ship.vc_e, ship.wind_speed, ship.wind_dir = dist[t]
        L.orc_ship_step(C.byref(p), C.byref(ship), int(s.get("avoid", False)), 1, lg, C.byref(rud), C.byref(thr))
        state[t] = cport.ship_state_vector(ship)
        cmd[t] = (rud.value, thr.value)
        k[t] = ship.k
        logs[t] = lg[:]
    return dict(state=state, cmd=cmd, k=k, log27=logs, fuel3=np.array([ship.fuel_me, ship.fuel_el, ship.fuel]),
                time=ship.time)


def run_oracle_env_episode(p, m, e, actions, nsteps):
    L = cport.lib()
    obs = np.zeros((nsteps, 10)); rew = np.zeros(nsteps); done = np.zeros(nsteps, np.uint8)
    status = np.zeros(nsteps, np.int32); dist = np.zeros((nsteps, 2)); stop = np.zeros((nsteps, 2), np.uint8)
    o10 = (C.c_double * 10)(); r = C.c_double(); d = C.c_int(); st = C.c_int()
    for t in range(nsteps):
        a_n, a_e, sac = sc.action_at(actions, t)
        L.orc_env_step(C.byref(p), C.byref(p), C.byref(m), C.byref(e), a_n, a_e, int(sac), int(t == 0),
                       o10, C.byref(r), C.byref(d), C.byref(st))
        obs[t], rew[t], done[t], status[t] = o10[:], r.value, d.value, st.value
        dist[t] = (e.eps_distance_travelled, e.sampling_distance_travelled)
        stop[t] = (e.stop_test, e.stop_obs)
    return dict(obs=obs, reward=rew, done=done, status=status, dist=dist, stop=stop)
