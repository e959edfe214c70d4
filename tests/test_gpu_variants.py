"""GPU parity of the model variants and snapshot data through the C ABI, against the fixtures frozen from
the live reference classes (tests/golden/make_golden_variants.py) and the C oracle:
  * SimplifiedMachineryModel + ThrottleFromSpeedSetPointSimplifiedPropulsion as SIT_MACH_SIMPLIFIED in a ShipFleet
    (fp64 <= 1e-9 per step, waypoint index exact; fp32 <= 1e-4 on pose and speed at the end);
  * StaticObstacle circles (sit_circle_query): fp64 distances within 1e-9 m of the reference, inside flags exact;
  * ship_snap_shot: snapshot steps exact, outline rows <= 1e-9 (fp64) against the rows the real env appended."""
import numpy as np
import pytest
import torch

import gpu_helpers as G
import helpers as H
import scenarios as sc
import variants as V
import sac_maritime_ast_b200 as sit

pytestmark = pytest.mark.gpu
FLOORS7 = np.array([1.0, 1.0, 1.0, 1.0, 1e-3, 1e-3, 1.0])[None, :, None]


def _simplified_fleet(n, precision):
    s = V.SIMPLIFIED
    sm, tc, ap = sit.simplified_ship(s["route"], s["psi0"], s["thrust_time_constant"], s["kp"], s["ki"], s["initial_thrust_force"],
                                     mode_names=s["modes"], mode_index=s["mode"])
    return sit.ShipFleet(sm, tc, ap, s["u_sp"], n_ships=n, device="cuda:0", precision=precision)


def _run_simplified(precision, log27=False):
    s = V.SIMPLIFIED
    fleet = _simplified_fleet(33, precision)
    sched = dict(s["mode_schedule"])
    cuts = sorted(set([0, s["nsteps"]] + list(sched) + list(range(0, s["nsteps"], 500))))
    trajs, cmds, logs = [], [], []
    for a, b in zip(cuts[:-1], cuts[1:]):
        if a in sched:
            fleet.f["mode"][:] = sched[a]
        out = fleet.step(b - a, traj=True, cmd=True, log27=log27)
        trajs.append(out["traj"].double().cpu().numpy()); cmds.append(out["cmd"].double().cpu().numpy())
        if log27:
            logs.append(out["log27"].double().cpu().numpy())
    return fleet, np.concatenate(trajs), np.concatenate(cmds), (np.concatenate(logs) if log27 else None)


def test_simplified_machinery_fp64_matches_reference_and_oracle():
    g, o = H.golden("variant_simplified"), V.oracle_simplified()
    fleet, traj, cmd, lg = _run_simplified("fp64", log27=True)
    for ship in (0, 17, 32):
        assert np.array_equal(traj[:, 7, ship].astype(np.int8), g["k"])
        for ref in (g["state"], o["state"]):
            assert H.rel_err(traj[:, :7, ship], ref, FLOORS7[0, :, 0]).max() <= 1e-9
        assert H.rel_err(cmd[:, :, ship], g["cmd"], 1e-2).max() <= 1e-8
        assert H.rel_err(lg[:, :, ship], o["log27"], np.maximum(np.abs(o["log27"]).max(0) * 1e-6, 1e-9)).max() <= 1e-8
    fuel = np.stack([fleet.f[k].cpu().numpy() for k in ("fuel_me", "fuel_el", "fuel")])[:, 5]
    assert H.rel_err(fuel, g["fuel"][-1], 1e-6).max() <= 1e-9
    assert H.rel_err(fleet.f["ei_speed"][5].item(), g["ei_speed"][-1], 1.0) <= 1e-9
    # the machinery state slot holds the thrust force [N]
    assert abs(fleet.f["omega"][0].item() - g["state"][-1, 6]) <= 1e-9 * g["state"][-1, 6]


def test_simplified_machinery_fp32_close_to_reference():
    g = H.golden("variant_simplified")
    _, traj, _, _ = _run_simplified("fp32")
    err = H.rel_err(traj[-1, :4, 3], g["state"][-1, :4], np.array([100.0, 100.0, 1.0, 1.0]))
    assert err.max() <= 1e-4, err
    assert np.array_equal(traj[:, 7, 3].astype(np.int8), g["k"])
    assert abs(traj[-1, 6, 3] / g["state"][-1, 6] - 1.0) <= 1e-4


def test_simplified_machinery_is_fleet_only_and_needs_its_controller():
    s = sc.ENV_SCENARIOS["default"]
    v = V.SIMPLIFIED
    simp = sit.simplified_ship(s["test_route"], s["test_psi0"], v["thrust_time_constant"], v["kp"], v["ki"], 1e5)
    full = sit.default_ship(s["obs_route"], s["obs_psi0"])
    assets = [sit.ShipAssets(*simp, 8.0, [], [], "test_ship", False), sit.ShipAssets(*full, 8.0, [], [], "obs_ship", False)]
    with pytest.raises(sit.ShipSimError, match="fleet-only"):
        sit.MultiShipRLEnv(assets=assets, map=sit.PolygonObstacle(sc.DEFAULT_MAP), n_envs=2, device="cuda:0")
    with pytest.raises(TypeError):
        sit.ShipFleet(simp[0], full[1], simp[2], 8.0, n_ships=2, device="cuda:0")


@pytest.mark.parametrize("precision", ["fp64", "fp32"])
def test_static_obstacle_circles_match_reference(precision):
    g = H.golden("variant_circles")
    obs = sit.StaticObstacle(g["circles"], device="cuda:0", precision=precision)
    n, e = torch.from_numpy(g["points"][:, 0]), torch.from_numpy(g["points"][:, 1])
    dist = obs.obstacles_distance(n, e).double().cpu().numpy()
    inside = obs.if_ship_inside_obstacles(n, e).cpu().numpy()
    route = obs.if_route_inside_obstacles(n, e).cpu().numpy()
    assert np.array_equal(inside, route)
    if precision == "fp64":
        # obstacles_distance squares Python floats with `** 2` (libm pow: within an ulp of x * x, obstacle.py:62), the inside
        # tests square numpy arrays (x * x exactly, :73,81): distances to 1e-9 m, flags exact -- also ON the circles
        assert np.abs(dist - g["dist"]).max() <= 1e-9
        assert np.array_equal(inside, g["inside"])
    else:
        assert np.abs(dist - g["dist"]).max() <= 2e-3                      # metres; float32 positions up to 1e4 m
        clear = np.abs(g["dist"]) > 1e-2                                   # points within a centimetre of a circle may flip
        assert np.array_equal(inside[clear], g["inside"][clear]) and clear.sum() > 3000
    # scalar call, like the reference's callers
    assert bool(obs.if_ship_inside_obstacles(2000.0, 3000.0)) and float(obs.obstacles_distance(2000.0, 3000.0)) == -500.0
    with pytest.raises(ValueError):
        sit.StaticObstacle(np.zeros((2, 2)))


@pytest.mark.parametrize("t0,tag", [(30.0, "t30"), (0.0, "t0")])
def test_ship_snap_shot_schedule_and_outlines_match_reference(t0, tag):
    g = H.golden("variant_snapshots")
    s = sc.ENV_SCENARIOS["default"]
    kw = dict(mode_names=s.get("modes", ("PTI",)), mode_index=s.get("mode", 0))
    t = sit.default_ship(s["test_route"], s["test_psi0"], start=s.get("test_start"), **kw)
    o = sit.default_ship(s["obs_route"], s["obs_psi0"], start=s.get("obs_start"), **kw)
    assets = [sit.ShipAssets(*t, s["test_u_sp"], [], [], "test_ship", False), sit.ShipAssets(*o, s["obs_u_sp"], [], [], "obs_ship", False)]
    from types import SimpleNamespace
    env = sit.MultiShipRLEnv(assets=assets, map=sit.PolygonObstacle(sc.DEFAULT_MAP), ship_draw=True, time_since_last_ship_drawing=t0,
                             args=SimpleNamespace(sampling_frequency=7, theta=2), n_envs=5, device="cuda:0", precision="fp64")
    env.reset(); env.init_step()
    # 150 single steps (the reference's call pattern), then one 250-step launch
    for k in range(150):
        a_n, a_e, sac = sc.action_at(s["actions"], k)
        env.step((a_n, a_e), sac, k == 0)
    K = 250
    act = torch.zeros((K, 5, 2), dtype=env.real, device=env.device); su = torch.zeros((K, 5), dtype=torch.uint8, device=env.device)
    for k in range(K):
        a_n, a_e, sac = sc.action_at(s["actions"], 150 + k)
        act[k, :, 0], act[k, :, 1], su[k] = a_n, a_e, int(sac)
    env.step_k(K, act, su, init=False, record_all=True)
    for name in ("test", "obs"):
        dr = env.ship_drawings[name]
        assert [d[0] for d in dr] == g[f"{tag}_steps"].tolist(), "snapshot steps"
        for j, (_, ids, x, y) in enumerate(dr):
            assert ids.tolist() == [0, 1, 2, 3, 4]
            for e in (0, 4):
                assert H.rel_err(x[e].cpu().numpy(), g[f"{tag}_{name}_x"][j], 1.0).max() <= 1e-9
                assert H.rel_err(y[e].cpu().numpy(), g[f"{tag}_{name}_y"][j], 1.0).max() <= 1e-9
    # reset() clears the drawings (ShipModelAST.reset, ship_model.py:390) but not the timer (MSRL_env_ex.py:145-186)
    t_before = env.f["draw_t"].clone()
    env.reset()
    assert env.ship_drawings == {"test": [], "obs": []} and torch.equal(env.f["draw_t"], t_before)
