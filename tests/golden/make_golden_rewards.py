"""Freeze the reference's `reward_results` lists (MSRL_env_ex.py:133-141, appended in reward_function :924-964) for a few
env scenarios into tests/golden/env_reward_results.npz.  Run in the build container, where /root/reference exists:
    python tests/golden/make_golden_rewards.py
Same driving pattern as make_golden.py (reset, init_step, nsteps x step; a second episode where the scenario has one:
reset() -> reward_function_params() :174 re-creates the lists, so every episode starts its running sums at zero)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]

import scenarios as sc  # noqa: E402
from oracle import ref_harness as rh  # noqa: E402

NAMES = ("default", "multi_waypoint", "test_terrain", "ship_collision")
KEYS = (("test_ship", "reward_e_ct"), ("test_ship", "reward_near_col"), ("test_ship", "total_non_terminal"),
        ("obs_ship", "reward_base"), ("obs_ship", "reward_e_ct"), ("obs_ship", "reward_near_col"),
        ("obs_ship", "total_non_terminal"), ("shared", "total_non_terminal"))


def run(name, s):
    environment = sc.NO_CURRENT_ENV if s.get("no_current") else None
    env = rh.build_env(s["test_route"], s["obs_route"], s["test_psi0"], s["obs_psi0"], s["test_u_sp"],
                       s["obs_u_sp"], polygons=sc.DEFAULT_MAP, theta=s.get("theta", 2), environment=environment,
                       test_start=s.get("test_start"), obs_start=s.get("obs_start"),
                       mode_names=s.get("modes", ("PTI",)), mode_index=s.get("mode", 0))
    episodes = [(s["actions"], s["nsteps"])]
    if "second_episode" in s:
        episodes.append((s["second_episode"]["actions"], s["second_episode"]["nsteps"]))
    res = {}
    for ep, (actions, nsteps) in enumerate(episodes, 1):
        env.reset()
        env.init_step()
        for t in range(nsteps):
            a_n, a_e, sac = sc.action_at(actions, t)
            env.step((a_n, a_e), sac, t == 0)
        cum = np.stack([np.asarray(env.reward_results[a][b], dtype=np.float64) for a, b in KEYS], axis=1)
        assert cum.shape[0] == nsteps
        print(f"{name} episode {ep}: {cum.shape[0]} steps; final sums {cum[-1]}")
        res[f"{name}_ep{ep}"] = cum
    return res


if __name__ == "__main__":
    assert rh.available(), "reference tree not found"
    out = {}
    for name in NAMES:
        out.update(run(name, sc.ENV_SCENARIOS[name]))
    np.savez_compressed(os.path.join(HERE, "env_reward_results.npz"), **out)
