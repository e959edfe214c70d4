"""The batched training loop (test_beds/main_ast_b200.py: the reference's main_ast.py:310-412 loop on the device) runs end to
end: uniform warm-up, switch to the learned policy, replay growth, finite losses, and the learner really changes the policy the
rollout kernels read (they share the workspace)."""
import importlib.util
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_training_loop_smoke(tmp_path):
    spec = importlib.util.spec_from_file_location("main_ast_b200", os.path.join(ROOT, "test_beds", "main_ast_b200.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    log = tmp_path / "hist.json"
    hist, agent, memory = mod.main(["--n_envs", "300", "--num_steps", "384", "--start_steps", "128", "--k_steps", "16",
                                    "--collect_iterations", "4", "--hidden_size", "64", "--batch_size", "32",
                                    "--num_steps_episode", "100", "--truncate_episodes", "1",
                                    "--max_updates_per_segment", "6", "--log_json", str(log)])
    assert len(hist) == 6 and [h["policy"] for h in hist] == ["uniform"] * 2 + ["learned"] * 4
    assert hist[0]["memory"] >= 300 and hist[-1]["memory"] > hist[0]["memory"], "one decision per env at start, more later"
    assert all(h["env_steps"] > 0 for h in hist) and sum(h["episodes"] for h in hist) >= 300
    assert hist[-1]["updates"] == 36 and agent.updates == 36
    for h in hist:
        for k in ("critic_1_loss", "critic_2_loss", "policy_loss", "alpha"):
            assert np.isfinite(h[k]), (k, h)
    rows = memory.buffer[:len(memory)]
    assert torch.isfinite(rows[:, :23]).all() and (rows[:, 10].abs() <= 1).all()
    assert log.exists()
    # the learner really trains: the observations reach the networks scaled by 1 / observation_space.high, so tanh(mu) is not
    # saturated, the policy gradient is not identically zero and the parameters the rollout reads have moved
    import ctypes as C
    from sac_maritime_ast_b200.sac import SAC
    fresh = SAC(10, None, None, device=agent.device, hidden_size=64, batch_size=32, seed=25450)
    torch.cuda.synchronize()
    assert not torch.equal(fresh.f["policy"], agent.f["policy"]), "policy parameters did not move"
    assert float(agent.f["adam_v"][:agent.f["policy"].numel()].abs().max()) > 0.0, "every policy gradient was exactly zero"
    learned = [h for h in hist if h["policy"] == "learned" and h["share_of_unsaturated_actions"] is not None]
    assert learned and min(h["share_of_unsaturated_actions"] for h in learned) > 0.5, [h["share_of_unsaturated_actions"] for h in hist]
    assert all(h["rows_left_behind"] == 0 for h in hist)
