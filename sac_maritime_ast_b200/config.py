"""Host-side mirror of the reference's configuration surface.

Same class names, field names and constructor signatures as
``simulators/ship_in_transit/{ship_model,ship_engine,controllers,LOS_guidance,obstacle}.py`` so a
caller of the reference can build the same objects; here they only *hold configuration* (the
simulation itself runs in the CUDA kernels behind ``libshipsim.so``) and flatten into the C ABI's
``SitShipParams`` / ``SitMap``.  Citations are relative to /root/reference/.
"""
from __future__ import annotations

from typing import List, NamedTuple, Sequence, Union

import numpy as np

from . import _lib


# ---- ship_model.py:20-53 ---------------------------------------------------------------------------
class ShipConfiguration(NamedTuple):
    dead_weight_tonnage: float
    coefficient_of_deadweight_to_displacement: float
    bunkers: float
    ballast: float
    length_of_ship: float
    width_of_ship: float
    added_mass_coefficient_in_surge: float
    added_mass_coefficient_in_sway: float
    added_mass_coefficient_in_yaw: float
    mass_over_linear_friction_coefficient_in_surge: float
    mass_over_linear_friction_coefficient_in_sway: float
    mass_over_linear_friction_coefficient_in_yaw: float
    nonlinear_friction_coefficient__in_surge: float
    nonlinear_friction_coefficient__in_sway: float
    nonlinear_friction_coefficient__in_yaw: float


class EnvironmentConfiguration(NamedTuple):
    current_velocity_component_from_north: float
    current_velocity_component_from_east: float
    wind_speed: float
    wind_direction: float


class SimulationConfiguration(NamedTuple):
    initial_north_position_m: float
    initial_east_position_m: float
    initial_yaw_angle_rad: float
    initial_forward_speed_m_per_s: float
    initial_sideways_speed_m_per_s: float
    initial_yaw_rate_rad_per_s: float
    integration_step: float
    simulation_time: float


# ---- ship_engine.py:17-170 -------------------------------------------------------------------------
class MachineryModeParams(NamedTuple):
    main_engine_capacity: float
    electrical_capacity: float
    shaft_generator_state: str


class MachineryMode:
    def __init__(self, params: MachineryModeParams):
        if params.shaft_generator_state not in _lib.SG:
            raise ValueError("shaft_generator_state must be 'MOTOR', 'GEN' or 'OFF'")
        self.main_engine_capacity = params.main_engine_capacity
        self.electrical_capacity = params.electrical_capacity
        self.shaft_generator_state = params.shaft_generator_state


class MachineryModes:
    def __init__(self, list_of_modes: List[MachineryMode]):
        self.list_of_modes = list_of_modes


class FuelConsumptionCoefficients(NamedTuple):
    a: float
    b: float
    c: float


class SpecificFuelConsumptionWartila6L26:
    def __init__(self):
        self.a, self.b, self.c = 128.9, -168.9, 246.8

    def fuel_consumption_coefficients(self):
        return FuelConsumptionCoefficients(a=self.a, b=self.b, c=self.c)


class SpecificFuelConsumptionBaudouin6M26Dot3:
    def __init__(self):
        self.a, self.b, self.c = 108.7, -289.9, 324.9

    def fuel_consumption_coefficients(self):
        return FuelConsumptionCoefficients(a=self.a, b=self.b, c=self.c)


class MachinerySystemConfiguration(NamedTuple):
    hotel_load: float
    machinery_modes: MachineryModes
    machinery_operating_mode: int
    rated_speed_main_engine_rpm: float
    linear_friction_main_engine: float
    linear_friction_hybrid_shaft_generator: float
    gear_ratio_between_main_engine_and_propeller: float
    gear_ratio_between_hybrid_shaft_generator_and_propeller: float
    propeller_inertia: float
    propeller_speed_to_torque_coefficient: float
    propeller_diameter: float
    propeller_speed_to_thrust_force_coefficient: float
    rudder_angle_to_sway_force_coefficient: float
    rudder_angle_to_yaw_force_coefficient: float
    max_rudder_angle_degrees: float
    specific_fuel_consumption_coefficients_me: FuelConsumptionCoefficients
    specific_fuel_consumption_coefficients_dg: FuelConsumptionCoefficients


class SimplifiedPropulsionMachinerySystemConfiguration(NamedTuple):
    """ship_engine.py:148-157."""
    hotel_load: float
    machinery_modes: MachineryModes
    machinery_operating_mode: int
    specific_fuel_consumption_coefficients_me: FuelConsumptionCoefficients
    specific_fuel_consumption_coefficients_dg: FuelConsumptionCoefficients
    thrust_force_dynamic_time_constant: float
    rudder_angle_to_sway_force_coefficient: float
    rudder_angle_to_yaw_force_coefficient: float
    max_rudder_angle_degrees: float


# ---- controllers.py:16-38, LOS_guidance.py:15-19 ---------------------------------------------------
class ThrottleControllerGains(NamedTuple):
    kp_ship_speed: float
    ki_ship_speed: float
    kp_shaft_speed: float
    ki_shaft_speed: float


class HeadingControllerGains(NamedTuple):
    kp: float
    kd: float
    ki: float


class LosParameters(NamedTuple):
    radius_of_acceptance: float
    lookahead_distance: float
    integral_gain: float
    integrator_windup_limit: float


# ---- configuration holders with the reference's constructor signatures -----------------------------
class ShipModelAST:
    """ship_model.py:545-574 -- holds the configuration of one ship; the state lives on the device."""

    def __init__(self, ship_config: ShipConfiguration, simulation_config: SimulationConfiguration,
                 environment_config: EnvironmentConfiguration, machinery_config: MachinerySystemConfiguration,
                 initial_propeller_shaft_speed_rad_per_s):
        self.ship_config = ship_config
        self.simulation_config = simulation_config
        self.environment_config = environment_config
        self.machinery_config = machinery_config
        self.initial_propeller_shaft_speed_rad_per_s = float(initial_propeller_shaft_speed_rad_per_s)
        # the attributes the env reads at construction (MSRL_env_ex.py:86-90)
        self.north = float(simulation_config.initial_north_position_m)
        self.east = float(simulation_config.initial_east_position_m)
        self.yaw_angle = float(simulation_config.initial_yaw_angle_rad)
        self.forward_speed = float(simulation_config.initial_forward_speed_m_per_s)
        self.sideways_speed = float(simulation_config.initial_sideways_speed_m_per_s)
        self.yaw_rate = float(simulation_config.initial_yaw_rate_rad_per_s)
        self.l_ship = ship_config.length_of_ship
        # ship_engine.py:325: shaft_speed_max = 1.1 * rated speed [rad/s] * gear ratio
        self.shaft_speed_max = (1.1 * machinery_config.rated_speed_main_engine_rpm * np.pi / 30 *
                                machinery_config.gear_ratio_between_main_engine_and_propeller)
        self.ship_machinery_model = self  # reference callers read ship_model.ship_machinery_model.shaft_speed_max


ShipModel = ShipModelAST  # ship_model.py:406-542: identical on the step path


class SimplifiedMachineryModel:
    """ship_engine.py:398-433 -- configuration holder: thrust force as a first-order lag of the commanded power,
    ``d_thrust = (-k_thrust thrust + load_perc (P_me + P_el)) / T`` with ``k_thrust = 2160 / 790``."""

    def __init__(self, machinery_config: SimplifiedPropulsionMachinerySystemConfiguration, time_step: float,
                 initial_thrust_force: float):
        self.machinery_config, self.time_step = machinery_config, float(time_step)
        self.thrust = float(initial_thrust_force)
        self.k_thrust = 2160 / 790
        self.thrust_time_constant = machinery_config.thrust_force_dynamic_time_constant


class ShipModelSimplifiedPropulsion:
    """A hull (``BaseShipModel`` kinematics / ``three_dof_kinetics(thrust_force=..., rudder_angle=...)``,
    ship_model.py:233-250,436-466) propelled by a ``SimplifiedMachineryModel``.  The reference ships the machinery and the
    throttle controller classes (ship_engine.py:398-433, controllers.py:154-172) but no ship class that joins them; this
    holder is that join, in the order of ``ShipModel.update_differentials`` (ship_model.py:484-490): kinematics,
    ``update_thrust_force(throttle)``, kinetics with the pre-update thrust, Euler step of hull and thrust.
    Runs as SIT_MACH_SIMPLIFIED in a ``ShipFleet`` (the state field ``omega`` then holds the thrust force [N])."""

    def __init__(self, ship_config: ShipConfiguration, simulation_config: SimulationConfiguration,
                 environment_config: EnvironmentConfiguration,
                 machinery_config: SimplifiedPropulsionMachinerySystemConfiguration, initial_thrust_force: float):
        self.ship_config, self.simulation_config = ship_config, simulation_config
        self.environment_config, self.machinery_config = environment_config, machinery_config
        self.ship_machinery_model = SimplifiedMachineryModel(machinery_config, simulation_config.integration_step,
                                                             initial_thrust_force)
        # ShipFleet.init_from_config writes this into the machinery state slot
        self.initial_propeller_shaft_speed_rad_per_s = float(initial_thrust_force)
        self.north = float(simulation_config.initial_north_position_m)
        self.east = float(simulation_config.initial_east_position_m)
        self.yaw_angle = float(simulation_config.initial_yaw_angle_rad)
        self.l_ship = ship_config.length_of_ship


class EngineThrottleFromSpeedSetPoint:
    """controllers.py:108-136."""

    def __init__(self, gains: ThrottleControllerGains, max_shaft_speed: float, time_step: float,
                 initial_shaft_speed_integral_error: float):
        self.gains = gains
        self.max_shaft_speed = max_shaft_speed
        self.time_step = time_step
        self.initial_shaft_speed_integral_error = float(initial_shaft_speed_integral_error)


class ThrottleFromSpeedSetPointSimplifiedPropulsion:
    """controllers.py:154-172: ``sat(PI(speed set point, speed), 0, 1.1)``."""

    def __init__(self, kp: float, ki: float, time_step: float):
        self.kp, self.ki, self.time_step = float(kp), float(ki), float(time_step)
        self.initial_shaft_speed_integral_error = 0.0      # no shaft loop: the slot stays 0


class HeadingBySampledRouteController:
    """controllers.py:253-296.  ``route_name``: an array / list of (north, east) rows, or a text file
    with one ``north east`` pair per line (LOS_guidance.py:65-86)."""

    def __init__(self, route_name, heading_controller_gains: HeadingControllerGains,
                 los_parameters: LosParameters, time_step: float, max_rudder_angle: float,
                 num_of_samplings: int = 2):
        self.heading_controller_gains = heading_controller_gains
        self.los_parameters = los_parameters
        self.time_step = time_step
        self.max_rudder_angle = max_rudder_angle
        self.num_of_samplings = num_of_samplings
        if isinstance(route_name, (str, bytes)):
            route = np.loadtxt(route_name)
        else:
            route = np.asarray(route_name, dtype=np.float64)
        if route.ndim != 2 or route.shape[1] != 2 or route.shape[0] < 2:
            raise ValueError("route must be [n_wp >= 2, 2] (north, east)")
        self.route = route


HeadingByRouteController = HeadingBySampledRouteController  # controllers.py:192-250, fixed-route variant


class PolygonObstacle:
    """obstacle.py:92-148: list of polygons, each a list of (east, north) vertices."""

    def __init__(self, list_of_vertices_list: Sequence[Sequence[Sequence[float]]]):
        self.list_of_vertices_list = [[(float(e), float(n)) for e, n in poly] for poly in list_of_vertices_list]
        self.num_obstacles = len(self.list_of_vertices_list)
        pts = [p for poly in self.list_of_vertices_list for p in poly]
        # obstacle.py:111-124 map_boundaries
        self.min_east, self.max_east = min(p[0] for p in pts), max(p[0] for p in pts)
        self.min_north, self.max_north = min(p[1] for p in pts), max(p[1] for p in pts)

    def to_sit_map(self) -> "_lib.SitMap":
        m = _lib.SitMap()
        i = 0
        for pid, poly in enumerate(self.list_of_vertices_list):
            n = len(poly)
            for j in range(n):
                if i >= _lib.MAX_EDGES:
                    raise ValueError(f"map has more than {_lib.MAX_EDGES} ring segments")
                (x1, y1), (x2, y2) = poly[j], poly[(j + 1) % n]
                m.x1[i], m.y1[i], m.x2[i], m.y2[i], m.poly[i] = x1, y1, x2, y2, pid
                i += 1
        m.n_edges, m.n_poly = i, self.num_obstacles
        return m


def ship_params(ship_model: ShipModelAST, throttle_controller: EngineThrottleFromSpeedSetPoint,
                auto_pilot: HeadingBySampledRouteController) -> "_lib.SitShipParams":
    """Flatten one ship's configuration objects into the C ABI's SitShipParams."""
    sc, mc = ship_model.ship_config, ship_model.machinery_config
    p = _lib.SitShipParams()
    simplified = isinstance(mc, SimplifiedPropulsionMachinerySystemConfiguration)
    if simplified != isinstance(throttle_controller, ThrottleFromSpeedSetPointSimplifiedPropulsion):
        raise TypeError("SimplifiedMachineryModel goes with ThrottleFromSpeedSetPointSimplifiedPropulsion (no shaft speed to "
                        "feed back), ShipMachineryModel with EngineThrottleFromSpeedSetPoint")
    p.machinery_kind = _lib.MACH_SIMPLIFIED if simplified else _lib.MACH_FULL
    p.dt = float(ship_model.simulation_config.integration_step)
    for name in ShipConfiguration._fields:
        setattr(p, name, float(getattr(sc, name)))
    modes = mc.machinery_modes.list_of_modes
    if not 1 <= len(modes) <= _lib.MAX_MODES:
        raise ValueError(f"1..{_lib.MAX_MODES} machinery modes supported")
    p.n_modes = len(modes)
    for i, m in enumerate(modes):
        p.mode_main_engine_capacity[i] = float(m.main_engine_capacity)
        p.mode_electrical_capacity[i] = float(m.electrical_capacity)
        p.mode_shaft_generator_state[i] = _lib.SG[m.shaft_generator_state]
    names = ("hotel_load", "rudder_angle_to_sway_force_coefficient", "rudder_angle_to_yaw_force_coefficient",
             "max_rudder_angle_degrees")
    if simplified:
        p.thrust_force_dynamic_time_constant = float(mc.thrust_force_dynamic_time_constant)
        p.kp_simplified, p.ki_simplified = throttle_controller.kp, throttle_controller.ki
    else:
        names += ("linear_friction_main_engine", "linear_friction_hybrid_shaft_generator",
                  "gear_ratio_between_main_engine_and_propeller",
                  "gear_ratio_between_hybrid_shaft_generator_and_propeller", "propeller_inertia",
                  "propeller_speed_to_torque_coefficient", "propeller_diameter",
                  "propeller_speed_to_thrust_force_coefficient")
    for name in names:
        setattr(p, name, float(getattr(mc, name)))
    for i in range(3):
        p.sfc_me[i] = float(mc.specific_fuel_consumption_coefficients_me[i])
        p.sfc_dg[i] = float(mc.specific_fuel_consumption_coefficients_dg[i])
    if not simplified:
        g = throttle_controller.gains
        p.kp_ship_speed, p.ki_ship_speed = float(g.kp_ship_speed), float(g.ki_ship_speed)
        p.kp_shaft_speed, p.ki_shaft_speed = float(g.kp_shaft_speed), float(g.ki_shaft_speed)
    h, l = auto_pilot.heading_controller_gains, auto_pilot.los_parameters
    p.kp_heading, p.kd_heading, p.ki_heading = float(h.kp), float(h.kd), float(h.ki)
    p.radius_of_acceptance, p.lookahead_distance = float(l.radius_of_acceptance), float(l.lookahead_distance)
    p.los_integral_gain, p.los_integrator_windup_limit = float(l.integral_gain), float(l.integrator_windup_limit)
    return p


class StaticObstacle:
    """obstacle.py:8-90: circular obstacles, rows (north, east, radius) of an array or a text file.  The three queries keep
    the reference's names and run batched on the device (``sit_circle_query``): scalars or equal-shaped tensors / arrays of
    positions in, a tensor per query out.  ``if_ship_inside_obstacles`` / ``if_route_inside_obstacles``: d^2 <= r^2 for any
    circle (boundary counts as inside, :69-85); ``obstacles_distance``: min over circles of sqrt(|d^2|) - r (:55-67)."""

    def __init__(self, obstacle_data, print_init_msg=False, device="cuda", precision="fp64"):
        data = np.loadtxt(obstacle_data) if isinstance(obstacle_data, (str, bytes)) else np.asarray(obstacle_data, dtype=np.float64)
        if data.ndim == 1 and data.shape[0] == 3:
            data = data.reshape(1, 3)
        if data.ndim != 2 or data.shape[1] != 3 or not 1 <= data.shape[0] <= _lib.MAX_CIRCLES:
            raise ValueError(f"obstacle_data must be [1..{_lib.MAX_CIRCLES}, 3] rows of (north, east, radius)")
        self.data = np.ascontiguousarray(data)
        self.num_obstacles = data.shape[0]
        self.n_obs, self.e_obs, self.r_obs = data[:, 0].tolist(), data[:, 1].tolist(), data[:, 2].tolist()
        self.device, self.precision = device, precision

    def _query(self, n_pos, e_pos, want_dist):
        import ctypes as C
        import torch
        real = torch.float64 if self.precision == "fp64" else torch.float32
        dev = torch.device(self.device)
        if dev.type != "cuda":
            raise _lib.ShipSimError("StaticObstacle queries run on a CUDA device only (no CPU fallback)")
        n = torch.as_tensor(n_pos, dtype=real, device=dev).contiguous()
        e = torch.as_tensor(e_pos, dtype=real, device=dev).contiguous()
        if n.shape != e.shape:
            raise ValueError("north and east must have the same shape")
        flags = torch.empty(n.shape, dtype=torch.uint8, device=dev)
        dist = torch.empty(n.shape, dtype=real, device=dev) if want_dist else None
        _lib.check(_lib.lib().sit_circle_query(_lib.F64 if self.precision == "fp64" else _lib.F32, self.num_obstacles,
                                               self.data.ctypes.data_as(C.c_void_p), n.numel(), n.data_ptr(), e.data_ptr(),
                                               flags.data_ptr(), dist.data_ptr() if want_dist else None,
                                               torch.cuda.current_stream(dev).cuda_stream))
        return flags, dist

    def obstacles_distance(self, n_ship, e_ship):
        return self._query(n_ship, e_ship, True)[1]

    def if_ship_inside_obstacles(self, n_ship, e_ship):
        return self._query(n_ship, e_ship, False)[0].bool()

    def if_route_inside_obstacles(self, n_route, e_route):
        return self._query(n_route, e_route, False)[0].bool()


class ShipDraw:
    """utils.py:56-101: the six outline points (closed pentagon) of an 80 m x 20 m ship seen from above.
    ``ship_snap_shot`` data (ship_model.py:326-338) for a batch of poses: ``outline(north, east, psi)`` ->
    (x_ned [n, 6], y_ned [n, 6]), the rows the reference appends to ``ship_drawings[0]`` / ``[1]``; device kernel
    ``sit_ship_outline``."""

    def __init__(self):
        self.l, self.b = 80.0, 20.0

    def local_coords(self):
        l, b = self.l, self.b
        x = np.array([-l / 2, l / 4, l / 2, l / 4, -l / 2, -l / 2])
        y = np.array([-b / 2, -b / 2, 0.0, b / 2, b / 2, -b / 2])
        return x, y

    def outline(self, north, east, psi):
        import torch
        n = torch.as_tensor(north)
        if n.device.type != "cuda":
            raise _lib.ShipSimError("ShipDraw.outline runs on a CUDA device only (no CPU fallback)")
        real = n.dtype
        if real not in (torch.float32, torch.float64):
            raise TypeError("poses must be float32 or float64 tensors")
        n = n.contiguous()
        e = torch.as_tensor(east, dtype=real, device=n.device).contiguous()
        p = torch.as_tensor(psi, dtype=real, device=n.device).contiguous()
        x = torch.empty((n.numel(), 6), dtype=real, device=n.device)
        y = torch.empty_like(x)
        _lib.check(_lib.lib().sit_ship_outline(_lib.F64 if real == torch.float64 else _lib.F32, n.numel(), n.data_ptr(),
                                               e.data_ptr(), p.data_ptr(), x.data_ptr(), y.data_ptr()))
        return x, y


# ---- the default parameter set, test_beds/main_ast.py:106-237 --------------------------------------
DEFAULT_MAP = [
    [(0, 10000), (5500, 10000), (5300, 9000), (4800, 8500), (4200, 7300), (4000, 5700), (4300, 4900),
     (4900, 4400), (4400, 4000), (3200, 4100), (2000, 4500), (1000, 4000), (900, 3500), (500, 2600), (0, 2350)],
    [(10000, 0), (4000, 0), (4250, 250), (5000, 400), (6000, 900), (8000, 1100), (8500, 1500), (9000, 2250),
     (9500, 3500), (10000, 4000)],
    [(5500, 5500), (5700, 7000), (6200, 8100), (7500, 8000), (7800, 7000), (7600, 5500), (6900, 4700), (6000, 5000)],
    [(2000, 2000), (2500, 2300), (4000, 2500), (5000, 3000), (4200, 2100), (3400, 1900)],
]

_MODE_TABLE = {"PTO": (2160e3, 0.0, "GEN"), "PTI": (0.0, 2 * 510e3, "MOTOR"), "MEC": (2160e3, 510e3, "OFF")}


def default_ship(route, psi0: float, mode_names: Sequence[str] = ("PTI",), mode_index: int = 0, dt: float = 0.5,
                 environment: Union[EnvironmentConfiguration, None] = None, start=None,
                 u0: float = 0.0, v0: float = 0.0, r0: float = 0.0):
    """(ship_model, throttle_controller, auto_pilot) built with the values of main_ast.py:113-237."""
    route = np.asarray(route, dtype=np.float64)
    ship_config = ShipConfiguration(
        coefficient_of_deadweight_to_displacement=0.7, bunkers=200000, ballast=200000, length_of_ship=80,
        width_of_ship=16, added_mass_coefficient_in_surge=0.4, added_mass_coefficient_in_sway=0.4,
        added_mass_coefficient_in_yaw=0.4, dead_weight_tonnage=3850000,
        mass_over_linear_friction_coefficient_in_surge=130, mass_over_linear_friction_coefficient_in_sway=18,
        mass_over_linear_friction_coefficient_in_yaw=90, nonlinear_friction_coefficient__in_surge=2400,
        nonlinear_friction_coefficient__in_sway=4000, nonlinear_friction_coefficient__in_yaw=400)
    env_config = environment or EnvironmentConfiguration(
        current_velocity_component_from_north=-2.0, current_velocity_component_from_east=-2.0,
        wind_speed=2.0, wind_direction=-np.pi / 4)
    modes = MachineryModes([MachineryMode(MachineryModeParams(*_MODE_TABLE[n])) for n in mode_names])
    machinery_config = MachinerySystemConfiguration(
        machinery_modes=modes, machinery_operating_mode=mode_index, linear_friction_main_engine=68,
        linear_friction_hybrid_shaft_generator=57, gear_ratio_between_main_engine_and_propeller=0.6,
        gear_ratio_between_hybrid_shaft_generator_and_propeller=0.6, propeller_inertia=6000,
        propeller_diameter=3.1, propeller_speed_to_torque_coefficient=7.5,
        propeller_speed_to_thrust_force_coefficient=1.7, hotel_load=200000, rated_speed_main_engine_rpm=1000,
        rudder_angle_to_sway_force_coefficient=50e3, rudder_angle_to_yaw_force_coefficient=500e3,
        max_rudder_angle_degrees=30,
        specific_fuel_consumption_coefficients_me=SpecificFuelConsumptionWartila6L26().fuel_consumption_coefficients(),
        specific_fuel_consumption_coefficients_dg=SpecificFuelConsumptionBaudouin6M26Dot3().fuel_consumption_coefficients())
    n0, e0 = (route[0] if start is None else start)
    simulation_setup = SimulationConfiguration(
        initial_north_position_m=float(n0), initial_east_position_m=float(e0), initial_yaw_angle_rad=float(psi0),
        initial_forward_speed_m_per_s=u0, initial_sideways_speed_m_per_s=v0, initial_yaw_rate_rad_per_s=r0,
        integration_step=dt, simulation_time=3600)
    ship_model = ShipModelAST(ship_config=ship_config, machinery_config=machinery_config,
                              environment_config=env_config, simulation_config=simulation_setup,
                              initial_propeller_shaft_speed_rad_per_s=400 * np.pi / 30)
    throttle_controller = EngineThrottleFromSpeedSetPoint(
        gains=ThrottleControllerGains(kp_ship_speed=7, ki_ship_speed=0.13, kp_shaft_speed=0.05, ki_shaft_speed=0.005),
        max_shaft_speed=ship_model.shaft_speed_max, time_step=dt, initial_shaft_speed_integral_error=114)
    auto_pilot = HeadingBySampledRouteController(
        route, heading_controller_gains=HeadingControllerGains(kp=1, kd=90, ki=0.01),
        los_parameters=LosParameters(radius_of_acceptance=300, lookahead_distance=1000, integral_gain=0.002,
                                     integrator_windup_limit=4000),
        time_step=dt, max_rudder_angle=30 * np.pi / 180, num_of_samplings=2)
    return ship_model, throttle_controller, auto_pilot


def simplified_ship(route, psi0: float, thrust_force_dynamic_time_constant: float, kp: float, ki: float,
                    initial_thrust_force: float, mode_names: Sequence[str] = ("MEC",), mode_index: int = 0, **kw):
    """(ship_model, throttle_controller, auto_pilot) of the default hull propelled by a SimplifiedMachineryModel
    (ship_engine.py:398-433) under ThrottleFromSpeedSetPointSimplifiedPropulsion (controllers.py:154-172)."""
    full, _, auto_pilot = default_ship(route, psi0, mode_names=mode_names, mode_index=mode_index, **kw)
    mc = full.machinery_config
    machinery_config = SimplifiedPropulsionMachinerySystemConfiguration(
        hotel_load=mc.hotel_load, machinery_modes=mc.machinery_modes, machinery_operating_mode=mode_index,
        specific_fuel_consumption_coefficients_me=mc.specific_fuel_consumption_coefficients_me,
        specific_fuel_consumption_coefficients_dg=mc.specific_fuel_consumption_coefficients_dg,
        thrust_force_dynamic_time_constant=thrust_force_dynamic_time_constant,
        rudder_angle_to_sway_force_coefficient=mc.rudder_angle_to_sway_force_coefficient,
        rudder_angle_to_yaw_force_coefficient=mc.rudder_angle_to_yaw_force_coefficient,
        max_rudder_angle_degrees=mc.max_rudder_angle_degrees)
    ship_model = ShipModelSimplifiedPropulsion(full.ship_config, full.simulation_config, full.environment_config,
                                               machinery_config, initial_thrust_force)
    dt = full.simulation_config.integration_step
    return ship_model, ThrottleFromSpeedSetPointSimplifiedPropulsion(kp=kp, ki=ki, time_step=dt), auto_pilot
