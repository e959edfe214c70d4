"""Run the REAL reference simulator/env from /root/reference behind import shims.

TEST / BASELINE INFRASTRUCTURE.  ``/root/reference`` does not exist on the GPU box; there the
files staged by ``oracle/stage_ref.py`` under ``oracle/_ref/`` (git-ignored, shipped with the
snapshot) are used instead, and only by ``bench.py``'s CPU-baseline legs (``oracle/ref_timing.py``),
never by ``-m gpu`` tests or ``smoke()``.  In the build container it is used by
``tests/golden/make_golden.py`` to freeze reference outputs as fixtures and by CPU tests
(skipped when the tree is absent) that pin the C / Python restatements.

Shims: empty ``matplotlib{,.pyplot,.patches}`` modules (imported at
ship_model.py:8, LOS_guidance.py:8, obstacle.py:2,5 but only used for plotting); a
``gymnasium`` stub with ``Env``, ``spaces.Box``, ``utils.seeding`` (MSRL_env_ex.py:5-7);
``shapely.geometry`` -> oracle.geometry (obstacle.py:1); package alias
``simulator`` -> ``simulators.ship_in_transit`` (MSRL_env_ex.py:13-15).
Default parameters follow test_beds/main_ast.py:106-237.
"""
from __future__ import annotations

import os
import sys
import types
from types import SimpleNamespace

import numpy as np

def _reference_root() -> str:
    """/root/reference in the build container; on the GPU box the copy staged by oracle/stage_ref.py (oracle/_ref/)."""
    env = os.environ.get("SIT_REFERENCE_ROOT")
    if env:
        return env
    if os.path.isdir("/root/reference/simulators/ship_in_transit"):
        return "/root/reference"
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


REFERENCE_ROOT = _reference_root()

DEFAULT_MAP = [
    [(0, 10000), (5500, 10000), (5300, 9000), (4800, 8500), (4200, 7300), (4000, 5700), (4300, 4900),
     (4900, 4400), (4400, 4000), (3200, 4100), (2000, 4500), (1000, 4000), (900, 3500), (500, 2600), (0, 2350)],
    [(10000, 0), (4000, 0), (4250, 250), (5000, 400), (6000, 900), (8000, 1100), (8500, 1500), (9000, 2250),
     (9500, 3500), (10000, 4000)],
    [(5500, 5500), (5700, 7000), (6200, 8100), (7500, 8000), (7800, 7000), (7600, 5500), (6900, 4700), (6000, 5000)],
    [(2000, 2000), (2500, 2300), (4000, 2500), (5000, 3000), (4200, 2100), (3400, 1900)],
]


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "simulators", "ship_in_transit"))


_installed = False


def install_shims():
    global _installed
    if _installed:
        return
    if not available():
        raise RuntimeError(f"reference tree not found at {REFERENCE_ROOT}")
    for name in ("matplotlib", "matplotlib.pyplot", "matplotlib.patches"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
    sys.modules["matplotlib"].patches = sys.modules["matplotlib.patches"]

    if "gymnasium" not in sys.modules:
        gym = types.ModuleType("gymnasium")
        spaces = types.ModuleType("gymnasium.spaces")
        utils = types.ModuleType("gymnasium.utils")
        seeding = types.ModuleType("gymnasium.utils.seeding")

        class Env:  # metadata only on this path
            def __init__(self):
                pass

        class Box:
            def __init__(self, low, high, dtype=np.float32):
                self.low, self.high = np.asarray(low), np.asarray(high)
                self.shape, self.dtype = self.low.shape, self.low.dtype

        def np_random(seed=None):
            return np.random.default_rng(seed), seed

        gym.Env, spaces.Box, seeding.np_random = Env, Box, np_random
        gym.spaces, gym.utils, utils.seeding = spaces, utils, seeding
        sys.modules.update({"gymnasium": gym, "gymnasium.spaces": spaces,
                            "gymnasium.utils": utils, "gymnasium.utils.seeding": seeding})

    if "shapely" not in sys.modules:
        from oracle import geometry as _g
        shp = types.ModuleType("shapely")
        geo = types.ModuleType("shapely.geometry")
        geo.Polygon, geo.Point = _g.Polygon, _g.Point
        shp.geometry = geo
        sys.modules.update({"shapely": shp, "shapely.geometry": geo})

    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import importlib
    sit = importlib.import_module("simulators.ship_in_transit")
    sys.modules["simulator"] = sit
    for sub in ("ship_model", "controllers", "obstacle", "LOS_guidance", "ship_engine", "utils"):
        sys.modules[f"simulator.{sub}"] = importlib.import_module(f"simulators.ship_in_transit.{sub}")
    _installed = True


def ref_modules():
    install_shims()
    import importlib
    m = SimpleNamespace()
    for sub in ("ship_model", "controllers", "obstacle", "LOS_guidance", "ship_engine", "utils"):
        setattr(m, sub, importlib.import_module(f"simulators.ship_in_transit.{sub}"))
    m.env = importlib.import_module("RLEnv.MSRL_env_ex")
    return m


# ----------------------------------------------------------------------------------------------
# Builders for the default parameter set (test_beds/main_ast.py:106-237).  These only CALL the
# reference's constructors; every number is a configuration value, not reference code.
# ----------------------------------------------------------------------------------------------
def default_config_dict():
    """Plain-dict form of the default parameter set; shared with the product-side tests so both
    sides are configured from one table."""
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


def build_ship(route, psi0, u_sp, mode_names=("PTI",), mode_index=0, dt=0.5, environment=None,
               u0=0.0, v0=0.0, r0=0.0, cfg=None, start=None):
    """One reference ship (ShipModelAST + throttle controller + sampled-route autopilot).

    route: sequence of (north, east); the ship starts on route[0] unless ``start`` is given.
    """
    m = ref_modules()
    cfg = cfg or default_config_dict()
    sm, se, ct = m.ship_model, m.ship_engine, m.controllers
    ship_config = sm.ShipConfiguration(**cfg["ship"])
    env_config = sm.EnvironmentConfiguration(**(environment or cfg["environment"]))
    modes = se.MachineryModes([
        se.MachineryMode(se.MachineryModeParams(main_engine_capacity=cfg["modes"][n][0],
                                                electrical_capacity=cfg["modes"][n][1],
                                                shaft_generator_state=cfg["modes"][n][2]))
        for n in mode_names])
    machinery_config = se.MachinerySystemConfiguration(
        machinery_modes=modes, machinery_operating_mode=mode_index,
        specific_fuel_consumption_coefficients_me=se.FuelConsumptionCoefficients(*cfg["sfc_me"]),
        specific_fuel_consumption_coefficients_dg=se.FuelConsumptionCoefficients(*cfg["sfc_dg"]),
        **cfg["machinery"])
    route = np.asarray(route, dtype=np.float64)
    sim_config = sm.SimulationConfiguration(
        initial_north_position_m=float(route[0][0] if start is None else start[0]),
        initial_east_position_m=float(route[0][1] if start is None else start[1]),
        initial_yaw_angle_rad=psi0, initial_forward_speed_m_per_s=u0,
        initial_sideways_speed_m_per_s=v0, initial_yaw_rate_rad_per_s=r0,
        integration_step=dt, simulation_time=cfg["simulation_time"])
    ship_model = sm.ShipModelAST(ship_config=ship_config, machinery_config=machinery_config,
                                 environment_config=env_config, simulation_config=sim_config,
                                 initial_propeller_shaft_speed_rad_per_s=cfg["initial_propeller_shaft_speed_rad_per_s"])
    throttle = ct.EngineThrottleFromSpeedSetPoint(
        gains=ct.ThrottleControllerGains(**cfg["throttle"]),
        max_shaft_speed=ship_model.ship_machinery_model.shaft_speed_max, time_step=dt,
        initial_shaft_speed_integral_error=cfg["initial_shaft_speed_integral_error"])
    auto_pilot = ct.HeadingBySampledRouteController(
        route, heading_controller_gains=ct.HeadingControllerGains(**cfg["heading"]),
        los_parameters=ct.LosParameters(**cfg["los"]), time_step=dt,
        max_rudder_angle=cfg["machinery"]["max_rudder_angle_degrees"] * np.pi / 180, num_of_samplings=2)
    return SimpleNamespace(ship_model=ship_model, throttle_controller=throttle, auto_pilot=auto_pilot,
                           desired_forward_speed=u_sp)


def bare_ship_step(s, avoid=False):
    """One ship-step in the order MSRL_env_ex.py:346-398 uses (guidance -> throttle ->
    store -> update -> integrate -> next_time).  Returns (rudder, throttle) commanded."""
    sm = s.ship_model
    rud = s.auto_pilot.rudder_angle_from_sampled_route(sm.north, sm.east, sm.yaw_angle)
    thr = s.throttle_controller.throttle(s.desired_forward_speed, sm.forward_speed, sm.forward_speed)
    if avoid:  # MSRL_env_ex.py:240-249 (branch always taken for the test ship)
        thr = np.clip(0.5 * thr, 0.0, 1.1)
        mx = s.auto_pilot.heading_controller.max_rudder_angle
        rud = np.clip(rud + np.deg2rad(3), -mx, mx)
    sm.store_simulation_data(thr, rud, s.auto_pilot.get_cross_track_error(), s.auto_pilot.get_heading_error())
    sm.update_differentials(engine_throttle=thr, rudder_angle=rud)
    sm.integrate_differentials()
    sm.int.next_time()
    return float(rud), float(thr)


def ship_state_vector(s):
    """(N, E, psi, u, v, r, omega, k, e_ct, e_ct_int, ei_speed, ei_shaft, ei_heading, e_prev, fuel)."""
    sm, ap, tc = s.ship_model, s.auto_pilot, s.throttle_controller
    pid = ap.heading_controller.ship_heading_controller
    return np.array([sm.north, sm.east, sm.yaw_angle, sm.forward_speed, sm.sideways_speed, sm.yaw_rate,
                     sm.ship_machinery_model.omega, ap.next_wpt, ap.navigate.e_ct, ap.navigate.e_ct_int,
                     tc.ship_speed_controller.error_i, tc.shaft_speed_controller.error_i,
                     pid.error_i, pid.prev_error, sm.ship_machinery_model.fuel_cons], dtype=np.float64)


def build_env(test_route, obs_route, test_psi0, obs_psi0, test_u_sp, obs_u_sp, polygons=None,
              sampling_frequency=7, theta=2, test_start=None, obs_start=None, **ship_kw):
    """Reference MultiShipRLEnv (RLEnv/MSRL_env_ex.py) on the default map."""
    m = ref_modules()
    t = build_ship(test_route, test_psi0, test_u_sp, start=test_start, **ship_kw)
    o = build_ship(obs_route, obs_psi0, obs_u_sp, start=obs_start, **ship_kw)
    A = m.env.ShipAssets
    assets = [A(t.ship_model, t.throttle_controller, t.auto_pilot, test_u_sp, [], [], "test_ship", False),
              A(o.ship_model, o.throttle_controller, o.auto_pilot, obs_u_sp, [], [], "obs_ship", False)]
    the_map = m.obstacle.PolygonObstacle(polygons or DEFAULT_MAP)
    args = SimpleNamespace(sampling_frequency=sampling_frequency, theta=theta)
    return m.env.MultiShipRLEnv(assets=assets, map=the_map, ship_draw=False,
                                time_since_last_ship_drawing=30, args=args)
