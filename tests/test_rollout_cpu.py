"""CPU tests of the rollout checker (oracle/rollout_ref.py) and of the host-side sharding helpers."""
import numpy as np

import helpers as H
import scenarios as sc


def test_philox_known_answers():
    """Random123 kat_vectors for philox4x32_10."""
    from oracle.rollout_ref import philox4x32_10
    kat = [((0, 0, 0, 0), (0, 0), "6627e8d5 e169c58d bc57ac4c 9b00dbd8"),
           ((0xffffffff,) * 4, (0xffffffff,) * 2, "408f276d 41c83b0e a20bc7c6 6d5451fd"),
           ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
            "d16cfe09 94fdcceb 5001e420 24126ea1")]
    for c, k, want in kat:
        assert " ".join("%08x" % x for x in philox4x32_10(c, k)) == want


def test_rollout_ref_loop_contract():
    """main_ast.py:310-412: first decision at the first step on reset()'s float32 state, a transition per decision
    step only, at most sampling_frequency decisions per episode, mask rule of :387."""
    from oracle import rollout_ref as rr
    p, m, e = H.oracle_env(sc.ENV_SCENARIOS["default"])
    ref = rr.RolloutRef(p, m, e, gid=3, mode="uniform", num_steps_episode=250, truncate=True)
    eps_decisions, n_in_ep = [], 0
    for t in range(1500):
        before = len(ref.transitions)
        ep_before = ref.ep_steps
        ref.step()
        if len(ref.transitions) > before:
            n_in_ep += 1
        if ref.ep_steps < ep_before + 1:     # episode restarted
            eps_decisions.append(n_in_ep); n_in_ep = 0
    assert len(ref.transitions) == len(ref.decisions) >= 6
    assert all(1 <= n <= 7 for n in eps_decisions)
    s0 = ref.transitions[0][0]
    assert s0[3:6] == [0.0, 0.0, 0.0] and s0[9] == 0.0 and s0[0] == float(np.float32(500.0))
    for (s, a, r, s2, mask) in ref.transitions:
        assert -1.0 <= a <= 1.0 and mask in (0.0, 1.0)
    for (_, a, wn, we, tries) in ref.decisions:
        assert tries >= 1


def test_shard_range_partitions_every_env_once():
    from sac_maritime_ast_b200.rollout import shard_range
    for n, w in [(65536, 8), (10, 3), (7, 8), (1 << 20, 4)]:
        blocks = [shard_range(n, r, w) for r in range(w)]
        assert blocks[0][0] == 0 and blocks[-1][1] == n
        assert all(blocks[i][1] == blocks[i + 1][0] for i in range(w - 1))
        sizes = [hi - lo for lo, hi in blocks]
        assert max(sizes) - min(sizes) <= 1
