"""ctypes binding of the plain-C float64 oracle (oracle/ship_oracle.c).  TEST INFRASTRUCTURE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this.  The library is built by ``oracle/build.py`` (gcc -O2 -fopenmp, no fast-math, no FMA
contraction) into ``oracle/_build/libshiporacle.so``.
"""
from __future__ import annotations

import ctypes as C
import math
import os

import numpy as np

MAX_WP = 16
MAX_MODES = 4
N_LOG = 27
SG = {"MOTOR": 0, "GEN": 1, "OFF": 2}

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_build", "libshiporacle.so")

D = C.c_double


class OrcParams(C.Structure):
    _fields_ = [
        ("dt", D),
        ("dead_weight_tonnage", D), ("coefficient_of_deadweight_to_displacement", D), ("bunkers", D), ("ballast", D),
        ("length_of_ship", D), ("width_of_ship", D),
        ("added_mass_coeff_surge", D), ("added_mass_coeff_sway", D), ("added_mass_coeff_yaw", D),
        ("t_surge", D), ("t_sway", D), ("t_yaw", D), ("ku", D), ("kv", D), ("kr", D),
        ("n_modes", C.c_int),
        ("mode_me_cap", D * MAX_MODES), ("mode_el_cap", D * MAX_MODES), ("mode_sg_state", C.c_int * MAX_MODES),
        ("hotel_load", D), ("d_me", D), ("d_hsg", D), ("r_me", D), ("r_hsg", D), ("jp", D), ("kp_torque", D),
        ("dp", D), ("kt", D),
        ("c_rudder_v", D), ("c_rudder_r", D), ("max_rudder_angle_deg", D),
        ("sfc_me", D * 3), ("sfc_dg", D * 3),
        ("kp_speed", D), ("ki_speed", D), ("kp_shaft", D), ("ki_shaft", D),
        ("kp_heading", D), ("kd_heading", D), ("ki_heading", D),
        ("ra", D), ("lookahead", D), ("ki_los", D), ("windup_limit", D),
        ("mass", D), ("i_z", D), ("x_du", D), ("y_dv", D), ("n_dr", D), ("proj_area_f", D), ("proj_area_l", D),
        ("rho_a", D), ("cx", D), ("cy", D), ("cn", D), ("rudder_max", D),
        ("avail_total", D * MAX_MODES), ("avail_me", D * MAX_MODES), ("avail_el", D * MAX_MODES),
        ("mach_kind", C.c_int), ("_pad_mach", C.c_int),
        ("thrust_time_constant", D), ("kp_simplified", D), ("ki_simplified", D),
    ]


class OrcShip(C.Structure):
    _fields_ = [
        ("north", D), ("east", D), ("psi", D), ("u", D), ("v", D), ("r", D), ("omega", D),
        ("ei_speed", D), ("ei_shaft", D), ("ei_heading", D), ("e_prev", D),
        ("e_ct", D), ("e_ct_int", D), ("heading_ref", D), ("heading_mea", D),
        ("fuel_me", D), ("fuel_el", D), ("fuel", D), ("time", D), ("u_sp", D),
        ("vc_n", D), ("vc_e", D), ("wind_speed", D), ("wind_dir", D),
        ("wp_n", D * MAX_WP), ("wp_e", D * MAX_WP),
        ("k", C.c_int), ("n_wp", C.c_int), ("mode", C.c_int), ("n_wp0", C.c_int),
        ("init_north", D), ("init_east", D), ("init_psi", D), ("init_u", D), ("init_v", D), ("init_r", D),
        ("init_vc_n", D), ("init_vc_e", D), ("init_wind_speed", D), ("init_wind_dir", D), ("init_u_sp", D),
        ("wp0_n", D * MAX_WP), ("wp0_e", D * MAX_WP),
    ]


class OrcMap(C.Structure):
    _fields_ = [
        ("n_edges", C.c_int), ("n_poly", C.c_int),
        ("min_north", D), ("max_north", D), ("min_east", D), ("max_east", D),
        ("x1", D * 256), ("y1", D * 256), ("x2", D * 256), ("y2", D * 256),
        ("poly", C.c_int * 256),
    ]


class OrcEnv(C.Structure):
    _fields_ = [
        ("test", OrcShip), ("obs", OrcShip),
        ("stop_test", C.c_int), ("stop_obs", C.c_int), ("n_stored_obs", C.c_int),
        ("stored_n", D * 2), ("stored_e", D * 2),
        ("last_rpm", D * 2), ("last_ect", D * 2), ("last_power_me", D * 2), ("last_avail_me", D),
        ("eps_distance_travelled", D), ("sampling_distance_travelled", D),
        ("ab_segment_length", D), ("ab_alpha", D), ("theta", D),
        ("sampling_frequency", D), ("e_tolerance", D),
    ]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            from oracle import build as _b
            _b.build_oracle()
        L = C.CDLL(LIB_PATH)
        for name, st in (("orc_sizeof_params", OrcParams), ("orc_sizeof_ship", OrcShip),
                         ("orc_sizeof_env", OrcEnv), ("orc_sizeof_map", OrcMap)):
            got = getattr(L, name)()
            if got != C.sizeof(st):
                raise RuntimeError(f"oracle ABI mismatch for {st.__name__}: C {got} vs ctypes {C.sizeof(st)}")
        P = C.POINTER
        L.orc_derive.argtypes = [P(OrcParams)]
        L.orc_ship_step.argtypes = [P(OrcParams), P(OrcShip), C.c_int, C.c_int, P(D), P(D), P(D)]
        L.orc_run_ships.argtypes = [P(OrcParams), P(OrcShip), C.c_long, C.c_int, C.c_int, C.c_void_p,
                                    C.c_void_p, C.c_int]
        L.orc_map_contains.argtypes = [P(OrcMap), D, D]
        L.orc_map_distance.argtypes = [P(OrcMap), D, D]
        L.orc_map_distance.restype = D
        L.orc_ship_capture_init.argtypes = [P(OrcShip)]
        L.orc_ship_reset.argtypes = [P(OrcShip)]
        L.orc_env_reset_params.argtypes = [P(OrcEnv)]
        L.orc_env_reset.argtypes = [P(OrcEnv)]
        L.orc_env_init_step.argtypes = [P(OrcParams), P(OrcParams), P(OrcEnv)]
        L.orc_env_step.argtypes = [P(OrcParams), P(OrcParams), P(OrcMap), P(OrcEnv), D, D, C.c_int, C.c_int,
                                   P(D), P(D), P(C.c_int), P(C.c_int)]
        L.orc_run_envs.argtypes = [P(OrcParams), P(OrcParams), P(OrcMap), P(OrcEnv), C.c_long, C.c_int,
                                   C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                   C.c_void_p, C.c_int]
        for f in ("orc_derive", "orc_ship_step", "orc_run_ships", "orc_ship_capture_init", "orc_ship_reset",
                  "orc_env_reset_params", "orc_env_reset", "orc_env_init_step", "orc_env_step", "orc_run_envs"):
            getattr(L, f).restype = None
        _lib = L
    return _lib


# ----------------------------------------------------------------------------------------------
def make_params(cfg: dict, mode_names=("PTI",), dt=None) -> OrcParams:
    """cfg = the plain dict of ref_harness.default_config_dict() (values of main_ast.py:106-237)."""
    p = OrcParams()
    s, m = cfg["ship"], cfg["machinery"]
    p.dt = cfg["dt"] if dt is None else dt
    p.dead_weight_tonnage = s["dead_weight_tonnage"]
    p.coefficient_of_deadweight_to_displacement = s["coefficient_of_deadweight_to_displacement"]
    p.bunkers, p.ballast = s["bunkers"], s["ballast"]
    p.length_of_ship, p.width_of_ship = s["length_of_ship"], s["width_of_ship"]
    p.added_mass_coeff_surge = s["added_mass_coefficient_in_surge"]
    p.added_mass_coeff_sway = s["added_mass_coefficient_in_sway"]
    p.added_mass_coeff_yaw = s["added_mass_coefficient_in_yaw"]
    p.t_surge = s["mass_over_linear_friction_coefficient_in_surge"]
    p.t_sway = s["mass_over_linear_friction_coefficient_in_sway"]
    p.t_yaw = s["mass_over_linear_friction_coefficient_in_yaw"]
    p.ku = s["nonlinear_friction_coefficient__in_surge"]
    p.kv = s["nonlinear_friction_coefficient__in_sway"]
    p.kr = s["nonlinear_friction_coefficient__in_yaw"]
    p.n_modes = len(mode_names)
    for i, n in enumerate(mode_names):
        me, el, sg = cfg["modes"][n]
        p.mode_me_cap[i], p.mode_el_cap[i], p.mode_sg_state[i] = me, el, SG[sg]
    p.hotel_load = m["hotel_load"]
    p.d_me, p.d_hsg = m["linear_friction_main_engine"], m["linear_friction_hybrid_shaft_generator"]
    p.r_me = m["gear_ratio_between_main_engine_and_propeller"]
    p.r_hsg = m["gear_ratio_between_hybrid_shaft_generator_and_propeller"]
    p.jp = m["propeller_inertia"]
    p.kp_torque = m["propeller_speed_to_torque_coefficient"]
    p.dp = m["propeller_diameter"]
    p.kt = m["propeller_speed_to_thrust_force_coefficient"]
    p.c_rudder_v = m["rudder_angle_to_sway_force_coefficient"]
    p.c_rudder_r = m["rudder_angle_to_yaw_force_coefficient"]
    p.max_rudder_angle_deg = m["max_rudder_angle_degrees"]
    for i in range(3):
        p.sfc_me[i], p.sfc_dg[i] = cfg["sfc_me"][i], cfg["sfc_dg"][i]
    t, h, l = cfg["throttle"], cfg["heading"], cfg["los"]
    p.kp_speed, p.ki_speed = t["kp_ship_speed"], t["ki_ship_speed"]
    p.kp_shaft, p.ki_shaft = t["kp_shaft_speed"], t["ki_shaft_speed"]
    p.kp_heading, p.kd_heading, p.ki_heading = h["kp"], h["kd"], h["ki"]
    p.ra, p.lookahead = l["radius_of_acceptance"], l["lookahead_distance"]
    p.ki_los, p.windup_limit = l["integral_gain"], l["integrator_windup_limit"]
    lib().orc_derive(C.byref(p))
    return p


def init_ship(s: OrcShip, cfg: dict, route, psi0, u_sp, mode=0, environment=None, u0=0.0, v0=0.0, r0=0.0,
              start=None):
    """Fill one OrcShip the way the reference constructors do (ship starts on route[0])."""
    env = environment or cfg["environment"]
    route = np.asarray(route, dtype=np.float64)
    s.north, s.east = (float(route[0][0]), float(route[0][1])) if start is None else (float(start[0]), float(start[1]))
    s.psi = float(psi0)
    s.u, s.v, s.r = u0, v0, r0
    s.omega = cfg["initial_propeller_shaft_speed_rad_per_s"]
    s.ei_speed, s.ei_shaft = 0.0, float(cfg["initial_shaft_speed_integral_error"])
    s.ei_heading = s.e_prev = 0.0
    s.e_ct = s.e_ct_int = s.heading_ref = s.heading_mea = 0.0
    s.fuel_me = s.fuel_el = s.fuel = 0.0
    s.time = 0.0
    s.u_sp = float(u_sp)
    s.vc_n = env["current_velocity_component_from_north"]
    s.vc_e = env["current_velocity_component_from_east"]
    s.wind_speed, s.wind_dir = env["wind_speed"], env["wind_direction"]
    assert 2 <= len(route) <= MAX_WP
    for i, (n, e) in enumerate(route):
        s.wp_n[i], s.wp_e[i] = float(n), float(e)
    s.k, s.n_wp, s.mode = 1, len(route), mode
    lib().orc_ship_capture_init(C.byref(s))
    return s


def make_map(polygons) -> OrcMap:
    """polygons: list of [(east, north), ...] exactly as passed to PolygonObstacle (obstacle.py:98-109)."""
    from oracle.geometry import flatten_edges
    m = OrcMap()
    edges = flatten_edges(polygons)
    assert len(edges) <= 256
    m.n_edges, m.n_poly = len(edges), len(polygons)
    for i, (x1, y1, x2, y2, pid) in enumerate(edges):
        m.x1[i], m.y1[i], m.x2[i], m.y2[i], m.poly[i] = x1, y1, x2, y2, pid
    pts = [p for poly in polygons for p in poly]
    m.min_east, m.max_east = float(min(p[0] for p in pts)), float(max(p[0] for p in pts))
    m.min_north, m.max_north = float(min(p[1] for p in pts)), float(max(p[1] for p in pts))
    return m


def init_env(e: OrcEnv, cfg, test_route, obs_route, test_psi0, obs_psi0, test_u_sp, obs_u_sp,
             sampling_frequency=7, theta=2, mode=0, environment=None, test_start=None, obs_start=None):
    init_ship(e.test, cfg, test_route, test_psi0, test_u_sp, mode=mode, environment=environment, start=test_start)
    init_ship(e.obs, cfg, obs_route, obs_psi0, obs_u_sp, mode=mode, environment=environment, start=obs_start)
    e.stop_test = e.stop_obs = e.n_stored_obs = 0
    e.eps_distance_travelled = e.sampling_distance_travelled = 0.0
    e.sampling_frequency, e.theta = float(sampling_frequency), float(theta)
    lib().orc_env_reset_params(C.byref(e))
    return e


SHIP_STATE_FIELDS = ("north", "east", "psi", "u", "v", "r", "omega", "k", "e_ct", "e_ct_int",
                     "ei_speed", "ei_shaft", "ei_heading", "e_prev", "fuel")


def ship_state_vector(s: OrcShip) -> np.ndarray:
    """Same ordering as ref_harness.ship_state_vector."""
    return np.array([getattr(s, f) for f in SHIP_STATE_FIELDS], dtype=np.float64)


STATUS_STRINGS = (
    "|Test ship reaches endpoint|", "|Test ship hits map horizon|", "|Test ship collides with the terrain|",
    "|Test ship mechanical failure|", "|Test ship navigation failure|", "|Test ship blackout failure|",
    "|Obstacle ship reaches endpoint|", "|Obstacle ship hits map horizon|",
    "|Obstacle ship collides with the terrain|", "|Obstacle ship IW sampled in terminal state|",
    "|Obstacle ship navigation failure|", "|Ship collision|")


def status_bits_from_string(status: str) -> int:
    """Map a reference status string (MSRL_env_ex.py:760-899) to the oracle's bitmask."""
    bits = 0
    for i, s in enumerate(STATUS_STRINGS):
        if s in status:
            bits |= 1 << i
    return bits
