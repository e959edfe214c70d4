"""Scenario table shared by the golden-fixture generator (runs the REAL reference), the oracle
tests and the GPU parity tests.  Pure data -- no simulator code.

Every scenario uses the default parameter set of the reference (test_beds/main_ast.py:106-237)
unless it overrides something here.  Routes are (north, east) lists.
"""
from __future__ import annotations

import numpy as np

DEFAULT_MAP = [
    [(0, 10000), (5500, 10000), (5300, 9000), (4800, 8500), (4200, 7300), (4000, 5700), (4300, 4900),
     (4900, 4400), (4400, 4000), (3200, 4100), (2000, 4500), (1000, 4000), (900, 3500), (500, 2600), (0, 2350)],
    [(10000, 0), (4000, 0), (4250, 250), (5000, 400), (6000, 900), (8000, 1100), (8500, 1500), (9000, 2250),
     (9500, 3500), (10000, 4000)],
    [(5500, 5500), (5700, 7000), (6200, 8100), (7500, 8000), (7800, 7000), (7600, 5500), (6900, 4700), (6000, 5000)],
    [(2000, 2000), (2500, 2300), (4000, 2500), (5000, 3000), (4200, 2100), (3400, 1900)],
]

ALL_MODES = ("PTO", "PTI", "MEC")


def disturbance_sequence(seed: int, nsteps: int, hold: int) -> np.ndarray:
    """[nsteps, 4] = (current north, current east, wind speed, wind dir), re-drawn every `hold` steps."""
    rng = np.random.default_rng(seed)
    nblk = (nsteps + hold - 1) // hold
    blk = np.stack([rng.uniform(-2, 2, nblk), rng.uniform(-2, 2, nblk),
                    rng.uniform(0, 15, nblk), rng.uniform(-np.pi, np.pi, nblk)], axis=1)
    return np.repeat(blk, hold, axis=0)[:nsteps].copy()


# ---- single-ship scenarios (bare ship-step loop, the K1 kernel) --------------------------------
SHIP_SCENARIOS = {
    # SURVEY.md Appendix D / BASELINE config 1
    "default": dict(route=[(0, 0), (3000, 3000), (6000, 9000)], psi0=np.pi / 4, u_sp=8.0, nsteps=2000),
    # MEC mode, no current, strong wind, sharp turns
    "mec_wind_turns": dict(route=[(0, 0), (1500, 500), (1200, 2500), (3000, 3000)], psi0=0.3, u_sp=14.0,
                           modes=ALL_MODES, mode=2, nsteps=1500,
                           environment=dict(current_velocity_component_from_north=0.0,
                                            current_velocity_component_from_east=0.0,
                                            wind_speed=10.0, wind_direction=1.0)),
    # PTO mode, collision-avoidance branch (MSRL_env_ex.py:240-249), per-step disturbance sequence
    "pto_avoid_disturbed": dict(route=[(1000, 1000), (2500, 4000), (6000, 5000)], psi0=1.2, u_sp=5.0,
                                modes=ALL_MODES, mode=0, avoid=True, nsteps=1500,
                                disturbance=dict(seed=7, hold=50)),
    # machinery-mode switching at run time (ship_engine.py:233-237)
    "mode_switching": dict(route=[(0, 0), (4000, 1000), (8000, 5000)], psi0=0.2, u_sp=7.0, modes=ALL_MODES,
                           mode=1, nsteps=1200, mode_schedule=[(300, 2), (600, 0), (900, 1)]),
    # far off track: LOS saturation (quirk Q8), unwrapped heading error (Q9)
    "los_saturation": dict(route=[(0, 0), (0, 6000)], start=(2500, 500), psi0=-2.5, u_sp=6.0, nsteps=1500),
    # many short segments -> several waypoint switches, end-of-route hold
    "short_segments": dict(route=[(0, 0), (600, 200), (1100, 700), (1300, 1400), (1900, 1800), (2600, 1900)],
                           psi0=0.0, u_sp=7.0, nsteps=2000,
                           environment=dict(current_velocity_component_from_north=0.5,
                                            current_velocity_component_from_east=-0.5,
                                            wind_speed=5.0, wind_direction=2.0)),
}

# ---- two-ship env scenarios (MultiShipRLEnv.step, the K2 kernel) -------------------------------
# actions: list of (step index, north, east); SAC_update is True at those steps and the most recent
# action is passed as `converted_action` at every step (MSRL_env_ex.py:402-440).
ENV_SCENARIOS = {
    # BASELINE config 1: ends at step 959 with "Obstacle ship collides with the terrain"; keeps
    # stepping past done, then reset() + init_step() and a second, shorter episode (quirk Q3)
    "default": dict(test_route=[(500, 500), (9500, 9500)], obs_route=[(9500, 9000), (1000, 1000)],
                    test_psi0=np.pi / 4, obs_psi0=-3 * np.pi / 4, test_u_sp=8.0, obs_u_sp=8.0,
                    actions=[(0, 8300.0, 8600.0)], nsteps=1000,
                    second_episode=dict(actions=[(0, 8600.0, 8200.0)], nsteps=300)),
    # several decision events, obstacle ship threads the channel
    "multi_waypoint": dict(test_route=[(500, 500), (9500, 9500)], obs_route=[(9500, 9000), (1000, 1000)],
                           test_psi0=np.pi / 4, obs_psi0=-3 * np.pi / 4, test_u_sp=8.0, obs_u_sp=8.0,
                           actions=[(0, 8800.0, 8700.0), (150, 8200.0, 8500.0), (300, 7800.0, 8300.0)],
                           nsteps=700),
    # intermediate waypoint sampled inside an island -> terminal at the first step
    "iw_in_terrain": dict(test_route=[(500, 500), (9500, 9500)], obs_route=[(9500, 9000), (1000, 1000)],
                          test_psi0=np.pi / 4, obs_psi0=-3 * np.pi / 4, test_u_sp=8.0, obs_u_sp=8.0,
                          actions=[(0, 7000.0, 6500.0)], nsteps=40),
    # intermediate waypoint outside the map
    "iw_outside_map": dict(test_route=[(500, 500), (9500, 9500)], obs_route=[(9500, 9000), (1000, 1000)],
                           test_psi0=np.pi / 4, obs_psi0=-3 * np.pi / 4, test_u_sp=8.0, obs_u_sp=8.0,
                           actions=[(0, 10500.0, 8600.0)], nsteps=20),
    # head-on ships in the open channel (north 3000-4000) -> ship collision (+2000), both stop flags
    "ship_collision": dict(test_route=[(3500, 1500), (3500, 8000)], obs_route=[(3500, 4000), (3500, 1000)],
                           test_psi0=np.pi / 2, obs_psi0=-np.pi / 2, test_u_sp=8.0, obs_u_sp=8.0,
                           actions=[(0, 3500.0, 2500.0)], nsteps=600, no_current=True, theta=10),
    # short obstacle route: obstacle ship reaches its end point -> frozen path (quirk Q5), then the
    # test ship reaches its own end point
    "endpoints": dict(test_route=[(3000, 1500), (3000, 2800)], obs_route=[(3600, 6000), (3600, 4800)],
                      test_psi0=np.pi / 2, obs_psi0=-np.pi / 2, test_u_sp=8.0, obs_u_sp=8.0,
                      actions=[(0, 3600.0, 5400.0)], nsteps=900, no_current=True, theta=50),
    # test ship passes its end point and sails off the map; obstacle ship exceeds the sampling
    # distance (navigation failure, MSRL_env_ex.py:566-576)
    "horizon_and_sampling": dict(test_route=[(4500, 8800), (4500, 9990)], obs_route=[(3500, 2000), (3500, 8000)],
                                 test_psi0=np.pi / 2, obs_psi0=np.pi / 2, test_u_sp=8.0, obs_u_sp=8.0,
                                 actions=[(0, 3600.0, 3000.0)], nsteps=900, no_current=True, theta=0.5),
    # test ship steers into island 4 -> "Test ship collides with the terrain" (+1000)
    "test_terrain": dict(test_route=[(3500, 2500), (2000, 2600)], obs_route=[(9500, 9000), (1000, 1000)],
                         test_psi0=np.pi, obs_psi0=-3 * np.pi / 4, test_u_sp=8.0, obs_u_sp=8.0,
                         actions=[(0, 8300.0, 8600.0)], nsteps=700),
    # test ship starts 1300 m off its track -> navigation failure at once
    "test_nav_failure": dict(test_route=[(5000, 8400), (9500, 8400)], test_start=(5000, 9700),
                             obs_route=[(9500, 9000), (1000, 1000)],
                             test_psi0=0.0, obs_psi0=-3 * np.pi / 4, test_u_sp=8.0, obs_u_sp=8.0,
                             actions=[(0, 8300.0, 8600.0)], nsteps=60),
    # PTO mode: commanded load 1.1 * (2160 - 200) + 200 kW exceeds the 2160 kW main engine -> blackout
    "blackout_pto": dict(test_route=[(500, 500), (9500, 9500)], obs_route=[(9500, 9000), (1000, 1000)],
                         test_psi0=np.pi / 4, obs_psi0=-3 * np.pi / 4, test_u_sp=8.0, obs_u_sp=8.0,
                         modes=ALL_MODES, mode=0, actions=[(0, 8300.0, 8600.0)], nsteps=200),
}

NO_CURRENT_ENV = dict(current_velocity_component_from_north=0.0, current_velocity_component_from_east=0.0,
                      wind_speed=0.0, wind_direction=0.0)


def action_at(actions, t):
    """(act_n, act_e, sac_update) for step t given the event list."""
    cur = None
    sac = False
    for (s, n, e) in actions:
        if s <= t:
            cur = (n, e)
        if s == t:
            sac = True
    assert cur is not None, "first action must be at step 0 (quirk Q6)"
    return cur[0], cur[1], sac
