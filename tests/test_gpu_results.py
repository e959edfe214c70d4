"""N4: the 27-channel ``simulation_results`` log as a columnar device buffer and its exporters,
checked against the log the live reference wrote for the same scenario (tests/golden/ship_*.npz, frozen from
ship_model.py:645-684 by make_golden.py)."""
import numpy as np
import pytest

import helpers as H
import scenarios as sc

pytestmark = pytest.mark.gpu


def test_simulation_results_match_reference_log(tmp_path):
    import gpu_helpers as G
    import sac_maritime_ast_b200 as sit
    import torch
    name = "default"
    s, g = sc.SHIP_SCENARIOS[name], H.golden("ship_" + name)
    fleet = G.make_fleet(s, n_ships=5, precision="fp64")
    rec = sit.SimulationResults(fleet, ships=torch.tensor([0, 4]))
    n, done = s["nsteps"], 0
    for K in (1, 96, 400):
        rec.step(K)
        done += K
    rec.step(n - done)
    assert len(rec) == n
    col = rec.columnar()
    assert col.shape == (n, 27, 2) and col.is_cuda
    d = rec.to_dict(1)
    assert tuple(d) == sit.LOG_KEYS and all(len(v) == n for v in d.values())
    got = np.stack([np.asarray(d[k]) for k in sit.LOG_KEYS], axis=1)          # [T, 27]
    err = H.rel_err(got[g["log_rows"]], g["log27"], 1e-6)
    assert err.max() <= 1e-9, f"log rel err {err.max():.3e}"
    assert d['time [s]'][:3] == [0.0, 0.5, 1.0]
    df = rec.to_dataframe(0)
    assert list(df.columns) == list(sit.LOG_KEYS) and len(df) == n
    assert np.array_equal(df['north position [m]'].to_numpy(), got[:, 1])     # replicated ships agree bit-for-bit
    path = tmp_path / "log.npz"
    rec.save_npz(str(path))
    z = np.load(path)
    assert z["log"].shape == (n, 27, 2) and list(z["ships"]) == [0, 4] and list(z["keys"]) == list(sit.LOG_KEYS)
    rec.reset()
    assert len(rec) == 0 and rec.columnar().shape == (0, 27, 2)


@pytest.mark.parametrize("name", ["default", "multi_waypoint", "test_terrain", "ship_collision"])
def test_reward_results_match_reference(name):
    """``MultiShipRLEnv.reward_results`` (cumulative non-terminal reward terms, MSRL_env_ex.py:133-141,924-964) rebuilt from
    the env-step kernel's per-step increments against the lists the live reference produced
    (tests/golden/env_reward_results.npz, make_golden_rewards.py), both episodes where the scenario has two."""
    import gpu_helpers as G
    import sac_maritime_ast_b200 as sit
    import torch
    s = sc.ENV_SCENARIOS[name]
    g = H.golden("env_reward_results")
    env = G.make_env(s, n_envs=3, precision="fp64")
    rec = sit.RewardResults(env, envs=torch.tensor([2]))
    episodes = [(s["actions"], s["nsteps"])]
    if "second_episode" in s:
        episodes.append((s["second_episode"]["actions"], s["second_episode"]["nsteps"]))
    for ep, (actions, nsteps) in enumerate(episodes, 1):
        rec.reset()
        env.init_step()
        t = 0
        while t < nsteps:                      # one launch per held action segment, cut at 256 steps
            a_n, a_e, sac = sc.action_at(actions, t)
            K = 1
            while t + K < nsteps and K < 256 and sc.action_at(actions, t + K) == (a_n, a_e, False):
                K += 1
            rec.step_k(K, (a_n, a_e), sac, init=(t == 0))
            t += K
        assert len(rec) == nsteps
        d = rec.to_dict(0)
        want = g[f"{name}_ep{ep}"]
        got = np.stack([np.asarray(d[a][b]) for a, b in sit.REWARD_KEYS], axis=1)
        err = H.rel_err(got, want, 1e-3)
        assert err.max() <= 1e-9, f"{name} episode {ep}: reward_results rel err {err.max():.3e} at {np.unravel_index(err.argmax(), err.shape)}"
