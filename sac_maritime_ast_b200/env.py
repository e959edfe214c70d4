"""Batched drop-in for ``RLEnv/MSRL_env_ex.py`` ``MultiShipRLEnv`` (kernel K2).

Same constructor arguments, method names and argument order as the reference
(``MultiShipRLEnv(assets, map, ship_draw, time_since_last_ship_drawing, args)``, ``reset()``,
``init_step()``, ``step(converted_action, SAC_update, init)``, ``seed()``), batched over ``n_envs``
independent copies held on one GPU.  With ``n_envs == 1`` the 4-tuple returned by ``step`` can be
``.item()``-ed into the reference's scalars; ``decode_status`` rebuilds the reference's status
strings from the bitmask.
"""
from __future__ import annotations

import copy
import ctypes as C
from dataclasses import dataclass, field
from types import SimpleNamespace
from typing import List, Optional, Union

import numpy as np
import torch

from . import _lib, config
from .core import SimCore, _precision_code, nvtx_range, ptr


@dataclass
class ShipAssets:
    """MSRL_env_ex.py:23-33."""
    ship_model: config.ShipModelAST
    throttle_controller: config.EngineThrottleFromSpeedSetPoint
    auto_pilot: config.HeadingBySampledRouteController
    desired_forward_speed: float
    integrator_term: List[float]
    time_list: List[float]
    type_tag: str
    stop_flag: bool
    init_copy: "ShipAssets" = field(default=None, repr=False, compare=False)


class Box:
    """The two attributes of gymnasium.spaces.Box the reference's callers read (MSRL_env_ex.py:67-83)."""

    def __init__(self, low, high, dtype=np.float32):
        self.low, self.high = np.asarray(low, dtype=dtype), np.asarray(high, dtype=dtype)
        self.shape, self.dtype = self.low.shape, np.dtype(dtype)


class _ShipView:
    """Per-role strided views (test = even ship slots, obstacle = odd) of the ship state arrays."""

    def __init__(self, core: SimCore, role: int):
        for name, t in core.f.items():
            if t.shape[-1] == core.n_ships:
                object.__setattr__(self, name, t[..., role::2])


class MultiShipRLEnv(SimCore):
    def __init__(self, assets: List[ShipAssets], map: config.PolygonObstacle, ship_draw: bool = False,
                 time_since_last_ship_drawing: float = 30.0, args=None, n_envs: int = 1, device=None,
                 precision="fp64", max_wp: Optional[int] = None):
        self.test, self.obs = assets
        self.assets = [self.test, self.obs]
        for asset in self.assets:  # MSRL_env_ex.py:55-56
            asset.init_copy = copy.deepcopy(asset)
        self.ship_draw, self.time_since_last_ship_drawing = ship_draw, time_since_last_ship_drawing
        self.map = map
        self.args = args if args is not None else SimpleNamespace(sampling_frequency=7, theta=2)
        pi = np.pi
        self.observation_space = Box(low=[0, 0, -pi, -3000, 0, 0, 0, 0, -pi, 0],
                                     high=[10000, 20000, pi, 3000, 1000, 2000, 10000, 20000, pi, 1000])
        self.action_space = Box(low=[-pi / 6], high=[pi / 6])

        cfg = _lib.SitConfig()
        cfg.abi_version, cfg.kind, cfg.precision = _lib.ABI_VERSION, _lib.KIND_ENV, _precision_code(precision)
        longest = max(len(a.auto_pilot.route) for a in self.assets)
        cfg.max_wp = int(max_wp if max_wp is not None else
                         min(_lib.MAX_WP, max(longest, 2 + int(self.args.sampling_frequency))))
        for r, a in enumerate(self.assets):
            cfg.ship[r] = config.ship_params(a.ship_model, a.throttle_controller, a.auto_pilot)
        cfg.map = map.to_sit_map()
        cfg.sampling_frequency, cfg.theta = float(self.args.sampling_frequency), float(self.args.theta)
        super().__init__(cfg, n_envs, device)
        self.ship = [_ShipView(self, 0), _ShipView(self, 1)]
        self._init_state()
        self.e_tolerance = 1000
        self.stats = torch.zeros(_lib.N_STATS, dtype=torch.int64, device=self.device)
        self._last_action = None
        self.np_random = None

    # ---- construction-time state (the reference's constructors) ------------------------------------------
    def _init_state(self):
        for role, a in enumerate(self.assets):
            v, sm = self.ship[role], a.ship_model
            sc, ec = sm.simulation_config, sm.environment_config
            v.north[:] = sc.initial_north_position_m; v.east[:] = sc.initial_east_position_m
            v.psi[:] = sc.initial_yaw_angle_rad; v.u[:] = sc.initial_forward_speed_m_per_s
            v.v[:] = sc.initial_sideways_speed_m_per_s; v.r[:] = sc.initial_yaw_rate_rad_per_s
            v.omega[:] = sm.initial_propeller_shaft_speed_rad_per_s
            for name in ("ei_speed", "ei_heading", "e_prev", "e_ct", "e_ct_int", "fuel_me", "fuel_el", "fuel"):
                getattr(v, name)[:] = 0
            v.ei_shaft[:] = a.throttle_controller.initial_shaft_speed_integral_error
            v.u_sp[:] = a.desired_forward_speed
            v.vc_n[:] = ec.current_velocity_component_from_north
            v.vc_e[:] = ec.current_velocity_component_from_east
            v.wind_speed[:] = ec.wind_speed; v.wind_dir[:] = ec.wind_direction
            v.k[:] = 1; v.mode[:] = sm.machinery_config.machinery_operating_mode; v.steps[:] = 0
            route = torch.as_tensor(np.asarray(a.auto_pilot.route), dtype=self.real, device=self.device)
            n = route.shape[0]
            if n > self.cfg.max_wp:
                raise ValueError("route longer than max_wp")
            v.wp_n[:] = 0; v.wp_e[:] = 0
            v.wp_n[:n] = route[:, 0:1]; v.wp_e[:n] = route[:, 1:2]
            v.n_wp[:] = n
        for name in ("prev_n", "prev_e", "eps_dist", "samp_dist", "act_n", "act_e", "ep_return", "flags", "ep_steps"):
            self.f[name].zero_()
        # MSRL_env_ex.py:59-60: the snapshot timer starts at the constructor argument and is never reset by reset()
        self.f["draw_t"].fill_(float(self.time_since_last_ship_drawing))
        # ship_drawings[0] / [1] of each ship (ship_model.py:120,181,337-338) for the batch: one entry per snapshot step,
        # (step index, env ids [m], x_ned [m, 6], y_ned [m, 6]); cleared by reset() like ShipModelAST.reset (:390)
        self.ship_drawings = {"test": [], "obs": []}
        self._draw_step = 0
        self.capture_init()   # also derives AB_segment_length (MSRL_env_ex.py:122-127)

    # ---- reference attributes read by callers ----------------------------------------------------------------
    @property
    def eps_distance_travelled(self):
        return self.f["eps_dist"]

    @property
    def sampling_distance_travelled(self):
        return self.f["samp_dist"]

    @property
    def AB_segment_length(self):
        return self.f["ab_seg"]

    @property
    def stop_flags(self):
        """[n_envs, 2] bool: (test.stop_flag, obs.stop_flag)."""
        fl = self.f["flags"]
        return torch.stack([(fl & 1) != 0, (fl & 2) != 0], dim=1)

    def initial_state(self):
        """MSRL_env_ex.py:86-90: float32 [n_envs, 10] = [N_t, E_t, psi_t, 0, 0, 0, N_o, E_o, psi_o, 0]."""
        s = torch.zeros((self.n_envs, _lib.N_OBS), dtype=torch.float32, device=self.device)
        t, o = self.ship
        s[:, 0], s[:, 1], s[:, 2] = self.f["init_north"][0::2], self.f["init_east"][0::2], t.init_psi
        s[:, 6], s[:, 7], s[:, 8] = self.f["init_north"][1::2], self.f["init_east"][1::2], o.init_psi
        return s

    # ---- the reference interface -------------------------------------------------------------------------------
    def seed(self, seed=None):
        """MSRL_env_ex.py:442-444 (nothing on the step path draws random numbers)."""
        self.np_random = np.random.default_rng(seed)
        return seed

    def reset(self, mask: Optional[torch.Tensor] = None):
        """MSRL_env_ex.py:145-186.  Returns the float32 initial state [n_envs, 10]."""
        SimCore.reset(self, mask=mask, hard=False)
        for a in self.assets:
            a.stop_flag = False
        if mask is None:
            self.ship_drawings = {"test": [], "obs": []}
        return self.initial_state()

    def init_step(self, mask: Optional[torch.Tensor] = None):
        """MSRL_env_ex.py:188-215."""
        if mask is not None:
            mask = mask.to(device=self.device, dtype=torch.uint8).contiguous()
        _lib.check(self.L.sit_env_init_step(self.h, self.state_ptr(), ptr(mask), self.stream()))
        self.launches += 1

    def _as_action(self, converted_action, steps=1):
        if converted_action is None:
            raise TypeError("converted_action must be a (north, east) pair (the reference unpacks it at "
                            "MSRL_env_ex.py:919 and fails the same way)")
        a = torch.as_tensor(converted_action, dtype=self.real, device=self.device)
        if a.dim() == 1:                       # one (north, east) pair for every env
            a = a.view(1, 2).expand(self.n_envs, 2)
        if a.dim() == 2:                       # [n_envs, 2] -> [1][2][n_envs]
            assert a.shape == (self.n_envs, 2)
            return a.t().contiguous().view(1, 2, self.n_envs)
        assert a.shape == (steps, self.n_envs, 2)
        return a.permute(0, 2, 1).contiguous()

    def _as_mask(self, m, steps=1):
        if isinstance(m, (bool, int, np.bool_)):
            return torch.full((1, self.n_envs), int(bool(m)), dtype=torch.uint8, device=self.device)
        m = torch.as_tensor(m, device=self.device).to(torch.uint8)
        if m.dim() == 1:
            m = m.view(1, self.n_envs)
        return m.contiguous()

    def step(self, converted_action, SAC_update, init=False):
        """MSRL_env_ex.py:402-440 for every env.

        converted_action: (north, east) or [n_envs, 2]; SAC_update: bool or [n_envs]; init: bool.
        Returns (next_state [n_envs, 10], reward [n_envs], done [n_envs] bool, status [n_envs] int32).
        """
        out = self.step_k(1, converted_action, SAC_update, init=init, record_all=False)
        return out["obs"][0].t(), out["reward"][0], out["done"][0].bool(), out["status"][0]

    @nvtx_range("sit.env.step_k")
    def step_k(self, K, converted_action, SAC_update=False, init=None, auto_reset=False, record_all=True,
               want_sums=False, out: Optional[dict] = None, reward_parts=False):
        """K env steps in ONE launch.  converted_action [n_envs, 2] (held for the launch; SAC_update applies to
        its first step) or [K, n_envs, 2] with SAC_update [K, n_envs].  init: True / False for the first
        step, None = per-env first-step-since-reset flag.  auto_reset: False, True (reference reset semantics,
        quirk Q3) or "hard" (integrators, shaft speed and fuel restored too).  Outputs are SoA: obs [rec, 10, n_envs],
        reward / done / status [rec, n_envs] with rec = K (record_all) or 1 (last step); reward_parts adds
        ``reward_parts`` [rec, 8, n_envs], the per-step increments of the reference's ``reward_results`` lists."""
        io = _lib.SitEnvIO()
        act = self._as_action(converted_action, K)
        sac = self._as_mask(SAC_update, K)
        assert sac.shape[0] == act.shape[0] or sac.shape[0] == 1
        if sac.shape[0] != act.shape[0]:
            sac = sac.expand(act.shape[0], self.n_envs).contiguous()
        io.action, io.sac_update, io.act_steps = act.data_ptr(), sac.data_ptr(), act.shape[0]
        io.init = -1 if init is None else int(bool(init))
        io.auto_reset = 2 if auto_reset == "hard" else int(bool(auto_reset))
        rec = K if record_all else 1
        io.rec_steps = rec
        res = out if out is not None else {}
        B = self.n_envs
        spec = (("obs", (rec, _lib.N_OBS, B), self.real), ("reward", (rec, B), self.real),
                ("done", (rec, B), torch.uint8), ("status", (rec, B), torch.int32))
        if reward_parts:
            spec += (("reward_parts", (rec, _lib.N_REWARD_PARTS, B), self.real),)
        if self.ship_draw:
            if not (record_all or K == 1):
                raise ValueError("ship_draw needs every step recorded (record_all=True): a snapshot is tied to its step's pose")
            spec += (("snap", (rec, B), torch.uint8),)
        for name, shape, dt in spec:
            t = res.get(name)
            if t is None or tuple(t.shape) != shape:
                t = torch.empty(shape, dtype=dt, device=self.device)
                res[name] = t
            setattr(io, name, t.data_ptr())
        if want_sums:
            for name, dt in (("reward_sum", self.real), ("done_any", torch.uint8)):
                t = res.get(name)
                if t is None:
                    t = torch.empty(B, dtype=dt, device=self.device)
                    res[name] = t
                setattr(io, name, t.data_ptr())
        io.stats = self.stats.data_ptr()
        _lib.check(self.L.sit_env_step_k(self.h, self.state_ptr(), int(K), C.byref(io), self.stream()))
        self.launches += 1
        self._keep = (act, sac)   # keep inputs alive until the stream has consumed them
        if self.ship_draw:
            self._collect_snapshots(res, rec)
        return res

    def _collect_snapshots(self, res, rec):
        """ship_snap_shot of both ships for the steps the kernel flagged (MSRL_env_ex.py:416-421): outline rows from that
        step's observation pose (post-step north, east, yaw), appended to ``ship_drawings`` (data only, no plotting)."""
        snap = res["snap"]
        for t in range(rec):
            ids = torch.nonzero(snap[t], as_tuple=False).flatten()
            if ids.numel():
                ob = res["obs"][t]
                draw = config.ShipDraw()
                for name, c0 in (("test", 0), ("obs", 6)):
                    x, y = draw.outline(ob[c0, ids], ob[c0 + 1, ids], ob[c0 + 2, ids])
                    self.ship_drawings[name].append((self._draw_step + t, ids, x, y))
        self._draw_step += rec

    @nvtx_range("sit.env.step_host")
    def step_host(self, K, action_host: torch.Tensor, sac_update_host: torch.Tensor, init, out_host: dict):
        """One call through ``sit_env_step_host``: HOST (pinned) action [2, n_envs] / sac_update [n_envs] in,
        HOST obs [10, n_envs] / reward / done / status out; copies and synchronisation inside."""
        assert action_host.device.type == "cpu" and action_host.dtype == self.real
        _lib.check(self.L.sit_env_step_host(
            self.h, self.state_ptr(), int(K), ptr(action_host), ptr(sac_update_host), int(bool(init)),
            ptr(out_host["obs"]), ptr(out_host["reward"]), ptr(out_host["done"]), ptr(out_host["status"]),
            self.stream()))
        self.launches += 1
        return out_host

    def step_host_noalloc(self, K, action_host: torch.Tensor, sac_update_host: torch.Tensor, init, out_host: dict):
        """``step_host`` through ``sit_env_step_host_noalloc``: same arguments and results, the device staging in the
        library's fixed 2 MiB array (no cudaMalloc in the call; at most ``_lib.NOALLOC_MAX_ENVS`` envs; calls on one
        device share the array, so they must not overlap on the GPU).  No counterpart in the reference."""
        assert action_host.device.type == "cpu" and action_host.dtype == self.real
        _lib.check(self.L.sit_env_step_host_noalloc(
            self.h, self.state_ptr(), int(K), ptr(action_host), ptr(sac_update_host), int(bool(init)),
            ptr(out_host["obs"]), ptr(out_host["reward"]), ptr(out_host["done"]), ptr(out_host["status"]),
            self.stream()))
        self.launches += 1
        return out_host

    def map_query(self, north: torch.Tensor, east: torch.Tensor, want_dist: bool = False):
        """Waypoint validity for n points: flags uint8 (bit 0 = inside an obstacle, MSRL_env_ex.py:544-552; bit 1 =
        outside the map, :517-542) and optionally the distance to the nearest obstacle ring (obstacle.py:138-141)."""
        north = north.to(device=self.device, dtype=self.real).contiguous()
        east = east.to(device=self.device, dtype=self.real).contiguous()
        assert north.shape == east.shape
        flags = torch.empty(north.shape, dtype=torch.uint8, device=self.device)
        dist = torch.empty(north.shape, dtype=self.real, device=self.device) if want_dist else None
        _lib.check(self.L.sit_map_query(self.h, north.numel(), ptr(north), ptr(east), ptr(flags), ptr(dist), self.stream()))
        self.launches += 1
        return (flags, dist) if want_dist else flags

    # ---- status ---------------------------------------------------------------------------------------------------
    @staticmethod
    def decode_status(mask: int) -> str:
        """Rebuild the reference's status string (MSRL_env_ex.py:760-879,899) from the bitmask."""
        s = _lib.status_strings()
        mask = int(mask)
        test = " " + "".join(s[b] for b in range(0, 6) if mask >> b & 1)
        if not mask & 0x3F:                       # :806-807 appended when the test-ship check set no done
            test += "|Test ship not in terminal state|"
        obs = " " + "".join(s[b] for b in range(6, 11) if mask >> b & 1)
        if not mask & 0x780:                      # :878-879; reaching the end point alone sets no done (:830-834)
            obs += "|Obstacle ship not in terminal state|"
        col = " " + (s[11] if mask >> 11 & 1 else "")
        return test + obs + col                   # :975
