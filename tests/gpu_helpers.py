"""Scenario -> product objects (sac_maritime_ast_b200) for the GPU parity tests."""
from __future__ import annotations

from types import SimpleNamespace

import numpy as np
import torch

import scenarios as sc
import sac_maritime_ast_b200 as sit


def _env_cfg(d):
    if d is None:
        return None
    return sit.EnvironmentConfiguration(**d)


def make_fleet(s, n_ships=1, precision="fp64", device="cuda:0", max_wp=None):
    sm, tc, ap = sit.default_ship(s["route"], s["psi0"], mode_names=s.get("modes", ("PTI",)),
                                  mode_index=s.get("mode", 0), environment=_env_cfg(s.get("environment")),
                                  start=s.get("start"))
    return sit.ShipFleet(sm, tc, ap, s["u_sp"], n_ships=n_ships, device=device, precision=precision, max_wp=max_wp)


def run_fleet_scenario(s, n_ships=1, precision="fp64", chunk=97, log27=True):
    """Run one ship scenario on the GPU in launches of <= `chunk` steps (cut at mode-schedule points).
    Returns numpy dict: traj [T, 8, n], cmd [T, 2, n], log27 [T, 27, n]."""
    fleet = make_fleet(s, n_ships, precision)
    n = s["nsteps"]
    dist = None
    if "disturbance" in s:
        d = sc.disturbance_sequence(s["disturbance"]["seed"], n, s["disturbance"]["hold"])   # [T, 4]
        dist = torch.as_tensor(d, dtype=fleet.real, device=fleet.device).unsqueeze(-1).expand(n, 4, n_ships).contiguous()
    sched = dict(s.get("mode_schedule", []))
    cuts = sorted(set([0, n] + list(sched) + list(range(0, n, chunk))))
    trajs, cmds, logs = [], [], []
    for a, b in zip(cuts[:-1], cuts[1:]):
        if a in sched:
            fleet.f["mode"][:] = sched[a]
        out = fleet.step(b - a, dist=None if dist is None else dist[a:b], avoid=s.get("avoid", False),
                         traj=True, cmd=True, log27=log27)
        trajs.append(out["traj"].cpu().numpy()); cmds.append(out["cmd"].cpu().numpy())
        if log27:
            logs.append(out["log27"].cpu().numpy())
    res = dict(traj=np.concatenate(trajs), cmd=np.concatenate(cmds), fleet=fleet)
    if log27:
        res["log27"] = np.concatenate(logs)
    return res


def make_env(s, n_envs=1, precision="fp64", device="cuda:0"):
    environment = _env_cfg(sc.NO_CURRENT_ENV) if s.get("no_current") else None
    kw = dict(mode_names=s.get("modes", ("PTI",)), mode_index=s.get("mode", 0), environment=environment)
    t = sit.default_ship(s["test_route"], s["test_psi0"], start=s.get("test_start"), **kw)
    o = sit.default_ship(s["obs_route"], s["obs_psi0"], start=s.get("obs_start"), **kw)
    assets = [sit.ShipAssets(*t, s["test_u_sp"], [], [], "test_ship", False),
              sit.ShipAssets(*o, s["obs_u_sp"], [], [], "obs_ship", False)]
    args = SimpleNamespace(sampling_frequency=7, theta=s.get("theta", 2))
    return sit.MultiShipRLEnv(assets=assets, map=sit.PolygonObstacle(sc.DEFAULT_MAP), ship_draw=False,
                              time_since_last_ship_drawing=30, args=args, n_envs=n_envs, device=device,
                              precision=precision)


def run_env_episode(env, actions, nsteps, per_step=False):
    """Drive the env like tests/golden/make_golden.py drives the reference.  per_step=True calls
    env.step() once per step (the reference's call pattern); otherwise one launch per action segment.
    Returns numpy dict like the golden files (env 0 only unless n_envs > 1: arrays keep the env axis)."""
    B = env.n_envs
    obs, rew, done, status, dist, stop = [], [], [], [], [], []
    if per_step:
        for t in range(nsteps):
            a_n, a_e, sac = sc.action_at(actions, t)
            ns, r, d, st = env.step((a_n, a_e), sac, t == 0)
            obs.append(ns.t().cpu().numpy()[None]); rew.append(r.cpu().numpy()[None]); done.append(d.cpu().numpy()[None])
            status.append(st.cpu().numpy()[None])
            dist.append(np.stack([env.eps_distance_travelled.cpu().numpy(), env.sampling_distance_travelled.cpu().numpy()])[None])
            stop.append(env.stop_flags.cpu().numpy().T[None])
    else:
        K = nsteps
        act = torch.zeros((K, B, 2), dtype=env.real, device=env.device)
        sac = torch.zeros((K, B), dtype=torch.uint8, device=env.device)
        for t in range(K):
            a_n, a_e, su = sc.action_at(actions, t)
            act[t, :, 0], act[t, :, 1], sac[t] = a_n, a_e, int(su)
        out = env.step_k(K, act, sac, init=True, record_all=True)
        obs.append(out["obs"].cpu().numpy()); rew.append(out["reward"].cpu().numpy())
        done.append(out["done"].cpu().numpy()); status.append(out["status"].cpu().numpy())
    res = dict(obs=np.concatenate(obs), reward=np.concatenate(rew), done=np.concatenate(done).astype(np.uint8),
               status=np.concatenate(status))
    if per_step:
        res["dist"] = np.concatenate(dist); res["stop"] = np.concatenate(stop).astype(np.uint8)
    return res
