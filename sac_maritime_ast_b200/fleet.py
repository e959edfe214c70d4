"""ShipFleet: n independent bare ships advanced K steps per launch (kernel K1).

Batched stand-in for the reference's per-ship loop
``rudder_angle_from_sampled_route -> throttle -> store_simulation_data -> update_differentials ->
integrate_differentials -> next_time`` (MSRL_env_ex.py:346-398 over
simulators/ship_in_transit/*).  State fields are torch views of the device state buffer
(``fleet.f["north"]`` ...), so callers can randomise or read any of them in place.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import numpy as np
import torch

from . import _lib, config
from .core import SimCore, _precision_code, nvtx_range, ptr


class ShipFleet(SimCore):
    def __init__(self, ship_model: config.ShipModelAST, throttle_controller: config.EngineThrottleFromSpeedSetPoint,
                 auto_pilot: config.HeadingBySampledRouteController, desired_forward_speed: float,
                 n_ships: int = 1, device=None, precision="fp64", max_wp: Optional[int] = None):
        cfg = _lib.SitConfig()
        cfg.abi_version, cfg.kind, cfg.precision = _lib.ABI_VERSION, _lib.KIND_FLEET, _precision_code(precision)
        route = auto_pilot.route
        cfg.max_wp = int(max_wp if max_wp is not None else max(2, min(_lib.MAX_WP, len(route))))
        cfg.ship[0] = config.ship_params(ship_model, throttle_controller, auto_pilot)
        cfg.ship[1] = cfg.ship[0]
        cfg.sampling_frequency, cfg.theta = 1.0, 1.0
        super().__init__(cfg, n_ships, device)
        self.init_from_config(ship_model, throttle_controller, auto_pilot, desired_forward_speed)

    # ---- initial state: what the reference constructors set (ship_model.py:132-191, ship_engine.py:327,
    #      controllers.py:49,77-78,275, LOS_guidance.py:59-86) -------------------------------------------
    def init_from_config(self, ship_model, throttle_controller, auto_pilot, desired_forward_speed, sel=slice(None)):
        f, sc, ec = self.f, ship_model.simulation_config, ship_model.environment_config
        f["north"][sel] = sc.initial_north_position_m
        f["east"][sel] = sc.initial_east_position_m
        f["psi"][sel] = sc.initial_yaw_angle_rad
        f["u"][sel] = sc.initial_forward_speed_m_per_s
        f["v"][sel] = sc.initial_sideways_speed_m_per_s
        f["r"][sel] = sc.initial_yaw_rate_rad_per_s
        f["omega"][sel] = ship_model.initial_propeller_shaft_speed_rad_per_s
        for name in ("ei_speed", "ei_heading", "e_prev", "e_ct", "e_ct_int", "fuel_me", "fuel_el", "fuel"):
            f[name][sel] = 0
        f["ei_shaft"][sel] = throttle_controller.initial_shaft_speed_integral_error
        f["u_sp"][sel] = desired_forward_speed
        f["vc_n"][sel] = ec.current_velocity_component_from_north
        f["vc_e"][sel] = ec.current_velocity_component_from_east
        f["wind_speed"][sel] = ec.wind_speed
        f["wind_dir"][sel] = ec.wind_direction
        f["k"][sel] = 1
        f["mode"][sel] = ship_model.machinery_config.machinery_operating_mode
        f["steps"][sel] = 0
        self.set_route(auto_pilot.route, sel)
        self.capture_init()

    def set_route(self, route, sel=slice(None)):
        """route: [n_wp, 2] (north, east) shared by the selected ships."""
        route = np.asarray(route, dtype=np.float64)
        n = len(route)
        if n > self.cfg.max_wp:
            raise ValueError(f"route has {n} waypoints, capacity max_wp={self.cfg.max_wp}")
        r = torch.as_tensor(route, dtype=self.real, device=self.device)
        self.f["wp_n"][:, sel] = 0
        self.f["wp_e"][:, sel] = 0
        self.f["wp_n"][:n, sel] = r[:, 0:1]
        self.f["wp_e"][:n, sel] = r[:, 1:2]
        self.f["n_wp"][sel] = n

    def set_routes(self, wp_n: torch.Tensor, wp_e: torch.Tensor, n_wp: torch.Tensor):
        """Per-ship routes: wp_n / wp_e [max_wp, n_ships], n_wp [n_ships]."""
        self.f["wp_n"].copy_(wp_n)
        self.f["wp_e"].copy_(wp_e)
        self.f["n_wp"].copy_(n_wp)

    # ---- stepping -------------------------------------------------------------------------------------
    @nvtx_range("sit.fleet.step")
    def step(self, K: int = 1, dist: Optional[torch.Tensor] = None, avoid: bool = False, store: bool = True,
             traj: bool = False, log27: bool = False, cmd: bool = False, out: Optional[dict] = None):
        """Advance every ship K steps in one launch.

        dist: None, or [4, n_ships] (held) / [K, 4, n_ships] (per step) of (current north, current east,
        wind speed, wind dir).  Returns a dict with the requested outputs: ``traj`` [K, 8, n],
        ``log27`` [K, 27, n], ``cmd`` [K, 2, n] (reals of the fleet's precision).
        """
        io = _lib.SitFleetIO()
        io.avoid, io.store = int(avoid), int(store)
        if dist is not None:
            dist = dist.to(device=self.device, dtype=self.real).contiguous()
            if dist.dim() == 2:
                dist = dist.unsqueeze(0)
            assert dist.shape[1:] == (4, self.n_ships) and dist.shape[0] in (1, K)
            io.dist, io.dist_steps = dist.data_ptr(), dist.shape[0]
        res = out if out is not None else {}
        for name, want, ch in (("traj", traj, _lib.N_TRAJ), ("log27", log27, _lib.N_LOG), ("cmd", cmd, 2)):
            if want:
                t = res.get(name)
                if t is None or t.shape != (K, ch, self.n_ships):
                    t = torch.empty((K, ch, self.n_ships), dtype=self.real, device=self.device)
                    res[name] = t
                setattr(io, name, t.data_ptr())
        io.traj_steps = K if traj else 0
        _lib.check(self.L.sit_fleet_step_k(self.h, self.state_ptr(), int(K), C.byref(io), self.stream()))
        self.launches += 1
        return res

    CHANNELS = ("north", "east", "psi", "u", "v", "r", "omega", "k")

    @nvtx_range("sit.fleet.step_host")
    def step_host(self, K: int, dist_host: Optional[torch.Tensor], pose_host: Optional[torch.Tensor],
                  avoid: bool = False, channels=None, k8_host: Optional[torch.Tensor] = None):
        """One call through ``sit_fleet_step_host[_sel]``: HOST (pinned) disturbances [4, n_ships] in, HOST post-step
        (N, E, psi, u, v, r, omega, k) [8, n_ships] of the last step out; copies and synchronisation inside.
        ``channels``: names (or a bit mask) of the channels the caller reads -- pose_host is then
        [len(channels), n_ships] in channel order; ``k8_host``: int8 [n_ships] waypoint index."""
        for t in (dist_host, pose_host):
            assert t is None or (t.device.type == "cpu" and t.dtype == self.real and t.is_contiguous())
        if channels is None and k8_host is None:
            _lib.check(self.L.sit_fleet_step_host(self.h, self.state_ptr(), int(K), ptr(dist_host), int(avoid),
                                                  ptr(pose_host), self.stream()))
        else:
            mask = 0xff if channels is None else (channels if isinstance(channels, int) else
                                                  sum(1 << self.CHANNELS.index(c) for c in channels))
            assert pose_host is None or pose_host.shape == (bin(mask & 0xff).count("1"), self.n_ships)
            assert k8_host is None or (k8_host.dtype == torch.int8 and k8_host.numel() == self.n_ships and k8_host.is_contiguous())
            _lib.check(self.L.sit_fleet_step_host_sel(self.h, self.state_ptr(), int(K), ptr(dist_host), int(avoid),
                                                      ptr(pose_host), int(mask), ptr(k8_host), self.stream()))
        self.launches += 1
        return pose_host
