"""GPU parity of the device-resident rollout (sit_rollout_begin / _policy / _step_k) against the CPU restatement
oracle/rollout_ref.py, which wraps the pinned C oracle env step in the loop contract of main_ast.py:310-412.
Decisions (action bits, masks, number of resamples) exact; float fields to float32 storage precision."""
import numpy as np
import pytest
import torch

import helpers as H
import scenarios as sc

pytestmark = pytest.mark.gpu


def _by_env(rows):
    rows = rows.cpu().numpy()
    ids = rows[:, 23].copy().view(np.int32)
    return {int(g): rows[ids == g] for g in np.unique(ids)}


def _make(n_envs, precision, policy, offset=0, cap=600, scenario="default", **kw):
    import gpu_helpers as G
    from sac_maritime_ast_b200 import rollout as ro
    env = G.make_env(sc.ENV_SCENARIOS[scenario], n_envs=n_envs, precision=precision)
    # tests that shorten the episode (cap != 600) were written for truncated episodes: opt in explicitly
    kw.setdefault("truncate_episodes", cap != 600)
    return ro.IntermediateWaypointSampler(env, policy, capacity=max(4096, n_envs), seed=25450, env_id_offset=offset,
                                          num_steps_episode=cap, **kw)


def _oracle_rows(gid, nsteps, mode="uniform", weights=None, cap=600, scenario="default", truncate=None):
    from oracle import rollout_ref as rr
    p, m, e = H.oracle_env(sc.ENV_SCENARIOS[scenario])
    ref = rr.RolloutRef(p, m, e, gid, seed=25450, mode=mode, weights=weights, num_steps_episode=cap,
                        truncate=(cap != 600) if truncate is None else truncate)
    for _ in range(nsteps):
        ref.step()
    return ref


def _check_rows(got, ref_tr, tol):
    n = len(got)
    assert n >= 2 and n <= len(ref_tr) + 0
    for j in range(n):
        s, a, r, s2, mask = ref_tr[j]
        row = got[j]
        assert row[10] == np.float32(a), (j, row[10], a)                      # action: exact float32
        assert row[22] == mask, (j, row[22], mask)
        assert H.rel_err(row[0:10], s, 1.0).max() <= tol, (j, "state")
        assert H.rel_err(row[12:22], s2, 1.0).max() <= tol, (j, "next_state")
        assert H.rel_err(row[11], r, 1.0).max() <= tol, (j, "reward", row[11], r)


@pytest.mark.parametrize("cap,truncate", [(600, False), (150, True), (150, False), (40, False)])
def test_uniform_policy_rollout_matches_oracle_fp64(cap, truncate):
    """fp64 env, UniformPolicy: every env's transition sequence equals the CPU loop's (envs pause at decision events,
    so each env is compared on its own step count).  truncate False is the reference's loop (main_ast.py:322,387): the row
    closing at episode step num_steps_episode gets mask 1 and the episode runs on to a terminal state; True also resets
    the episode there (opt-in extension)."""
    from sac_maritime_ast_b200 import rollout as ro
    B = 6
    smp = _make(B, "fp64", ro.UniformPolicy(), offset=100, cap=cap, truncate_episodes=truncate)
    smp.begin()
    smp.collect(iterations=60, K=32)
    torch.cuda.synchronize()
    rows = smp.transitions()
    assert smp.n_transitions() == rows.shape[0] == int(smp.counters[1].item())
    per_env = _by_env(rows)
    assert sorted(per_env) == list(range(100, 100 + B))
    for gid, got in per_env.items():
        ref = _oracle_rows(gid, 60 * 32, cap=cap, truncate=truncate)
        # replay rows of one env are in order of their slots within a launch sequence; sort by nothing: slots grow in time
        _check_rows(got, ref.transitions, 2e-7)
    assert len({len(v) for v in per_env.values()}) >= 1
    assert int(smp.env.stats[13].item()) > 0


@pytest.mark.parametrize("tc,hidden", [(False, 256), (True, 256), (True, 96), (True, 32)])
def test_gaussian_policy_actions_match_torch_reference(tc, hidden):
    """First decision of every env: kernel MLP + tanh-Gaussian sample against a plain torch fp32 forward of the same
    weights and the same Philox normals; waypoint conversion against numpy.  tc: the tcgen05 hidden layer (bf16 operands,
    fp32 accumulate) against the same torch forward with both operands of that product rounded to bf16 (tolerance 2e-4:
    an fp32 rounding difference in layer 0 can flip a bf16 rounding)."""
    from oracle import rollout_ref as rr
    from sac_maritime_ast_b200 import rollout as ro
    B = 300   # more than one CTA tile, not a multiple of 16 or 128
    pol = ro.GaussianPolicy(hidden_size=hidden, device="cuda:0", seed=3, tensor_cores=tc)
    # scale the first layer so that activations stay O(1) on raw observations (~1e4 m) and log-sigma is unclipped
    pol.w0.mul_(1e-4); pol.w2.mul_(0.3)
    smp = _make(B, "fp64", pol, offset=7)
    # different observations per env: shift the initial positions held for the policy
    smp.begin()
    co = smp.env.f["cur_obs"]
    co.add_(torch.linspace(0, 500, B, device=co.device, dtype=co.dtype))
    obs = co.t().contiguous().float()
    smp.policy_step()
    torch.cuda.synchronize()
    rows = _by_env(smp.replay[:smp.n_opened()])          # opened rows: state + action are written, the step closes them
    mu, logsig = pol.mu_logsig(obs)
    mu, sig = mu.cpu().numpy(), torch.exp(logsig).cpu().numpy()
    e_ab = smp.env.f["ab_seg"].cpu().numpy(), smp.env.f["ab_alpha"].cpu().numpy()
    for b in range(B):
        row = rows[7 + b][0]
        tries = 0
        while True:
            u1, u2 = rr.draw(25450, 7 + b, 0, tries)
            eps = np.float32(np.sqrt(-2.0 * np.log(u1)) * np.cos(2.0 * np.pi * u2))
            a = np.tanh(np.float32(mu[b] + sig[b] * eps))
            ang = e_ab[1][b] + float(a) * np.pi / 6
            wn, we = 9500.0 + e_ab[0][b] * np.cos(ang), 9000.0 + e_ab[0][b] * np.sin(ang)
            from oracle import geometry as geo
            bad = not (0 <= wn <= 10000 and 0 <= we <= 10000) or any(geo.Polygon(p).contains(geo.Point(we, wn)) for p in sc.DEFAULT_MAP)
            tries += 1
            if not bad:
                break
        assert abs(row[10] - a) <= (2e-4 if tc else 2e-5), (b, row[10], a)
        assert np.allclose(row[0:10], obs[b].cpu().numpy(), rtol=1e-6)
    act_n, act_e = smp.env.f["act_n"].cpu().numpy(), smp.env.f["act_e"].cpu().numpy()
    assert np.isfinite(act_n).all() and np.isfinite(act_e).all()
    assert (smp.env.f["flags"] & 256).ne(0).all()           # every env holds an action now


def test_shard_invariance_and_fp32_consistency():
    """The same global envs split over two handles (env_id_offset) give bit-identical transitions; fp32 run is finite
    and its counters are consistent."""
    from sac_maritime_ast_b200 import rollout as ro
    whole = _make(8, "fp64", ro.UniformPolicy(), offset=0, cap=200)
    parts = [_make(4, "fp64", ro.UniformPolicy(), offset=0, cap=200), _make(4, "fp64", ro.UniformPolicy(), offset=4, cap=200)]
    for s in [whole] + parts:
        s.begin(); s.collect(iterations=20, K=32)
    torch.cuda.synchronize()
    w = _by_env(whole.transitions())
    for part in parts:
        for gid, got in _by_env(part.transitions()).items():
            assert np.array_equal(got.view(np.int32), w[gid].view(np.int32)), gid
    pol = ro.GaussianPolicy(hidden_size=128, device="cuda:0", seed=1)
    pol.w0.mul_(1e-4)
    s32 = _make(512, "fp32", pol, cap=300, auto_reset="hard")
    s32.begin(); s32.collect(iterations=40, K=16)
    torch.cuda.synchronize()
    rows = s32.transitions().cpu().numpy()
    assert np.isfinite(rows[:, :23]).all()
    assert set(np.unique(rows[:, 22])) <= {0.0, 1.0}
    assert (np.abs(rows[:, 10]) <= 1.0).all()
    assert 0 < int(s32.env.stats[13].item()) <= 512 * 40 * 16


def test_graph_collect_equals_eager_collect():
    """sit_rollout_collect (CUDA graph of one iteration, replayed) == the same launches issued one by one: replay rows,
    counters and env state bit-identical; changing K re-captures the graph."""
    from sac_maritime_ast_b200 import rollout as ro
    pol = ro.GaussianPolicy(hidden_size=64, device="cuda:0", seed=5)
    pol.w0.mul_(1e-4)
    a, b = _make(96, "fp32", pol, cap=200, auto_reset="hard"), _make(96, "fp32", pol, cap=200, auto_reset="hard")
    for s, graph in ((a, True), (b, False)):
        s.begin()
        s.collect(iterations=25, K=16, graph=graph)
        s.collect(iterations=10, K=8, graph=graph)
    torch.cuda.synchronize()
    assert a.n_transitions() == b.n_transitions() > 96
    assert torch.equal(a.counters, b.counters) and torch.equal(a.env.stats, b.env.stats)
    assert torch.equal(a.replay.view(torch.int32), b.replay.view(torch.int32))
    for k in ("north", "east", "psi", "u", "omega", "flags", "ep_steps"):
        assert torch.equal(a.env.f[k], b.env.f[k]), k


def test_env_group_pipelining_does_not_change_any_env(monkeypatch):
    """sit_rollout_collect cuts an iteration into env groups whose policy chains overlap the other groups' env steps
    (internal streams, one CUDA graph).  Size-independent property: every env's transition sequence, the counters and the
    env state are the same for 1, 2 and 3 groups (ragged last group), graph or eager."""
    from sac_maritime_ast_b200 import rollout as ro
    pol = ro.GaussianPolicy(hidden_size=64, device="cuda:0", seed=5, tensor_cores=True)
    pol.w0.mul_(1e-4)
    runs = []
    for groups, graph in ((1, True), (2, True), (3, True), (3, False)):
        monkeypatch.setenv("SIT_ROLLOUT_GROUPS", str(groups))
        s = _make(800, "fp32", pol, cap=200, auto_reset="hard")
        s.begin()
        if graph:
            s.collect(iterations=26, K=16)
        else:   # eager launches through the same entry point
            from sac_maritime_ast_b200 import _lib
            import ctypes as C
            st = torch.cuda.Stream()
            st.wait_stream(torch.cuda.current_stream())
            _lib.check(s.env.L.sit_rollout_collect(s.env.h, s.env.state_ptr(), 26, 16, C.byref(s._pol), C.byref(s.ro),
                                                   C.byref(s.io), 0, st.cuda_stream))
            torch.cuda.current_stream().wait_stream(st)
        torch.cuda.synchronize()
        runs.append(s)
    a = runs[0]
    assert 2 * 800 <= a.n_transitions() <= a.capacity
    wa = _by_env(a.transitions())
    for b in runs[1:]:
        assert torch.equal(a.counters, b.counters) and torch.equal(a.env.stats, b.env.stats)
        for k in ("north", "east", "psi", "u", "omega", "flags", "ep_steps", "act_n", "n_dec"):
            assert torch.equal(a.env.f[k], b.env.f[k]), k
        wb = _by_env(b.transitions())
        assert sorted(wa) == sorted(wb)
        for gid in wa:
            assert np.array_equal(wa[gid].view(np.int32), wb[gid].view(np.int32)), gid


def test_full_size_rollout_properties_and_shard_invariance():
    """BASELINE config 4 size (65 536 envs, fp32, Gaussian policy on the tensor cores): size-independent properties --
    counters agree with the ring, every row is complete and finite, and the two half shards (env_id_offset 0 / 32 768)
    reproduce the whole batch's transitions bit for bit, env by env."""
    from sac_maritime_ast_b200 import rollout as ro
    B, iters, K = 65536, 6, 32
    pol = ro.GaussianPolicy(hidden_size=256, device="cuda:0", seed=11, tensor_cores=True)
    pol.w0.mul_(1e-4)

    def run(n, offset):
        import gpu_helpers as G
        env = G.make_env(sc.ENV_SCENARIOS["default"], n_envs=n, precision="fp32")
        gen = torch.Generator(device="cuda").manual_seed(5)
        usp = 9.5 + 2.5 * torch.rand(B, generator=gen, device="cuda")           # per GLOBAL env, so shards see the same values
        env.ship[1].u_sp.copy_(usp[offset:offset + n])
        env.capture_init()
        s = ro.IntermediateWaypointSampler(env, pol, capacity=4 * n, seed=25450, env_id_offset=offset, num_steps_episode=150,
                                           auto_reset="hard", truncate_episodes=True)
        s.begin(); s.collect(iters, K)
        torch.cuda.synchronize()
        return s
    whole = run(B, 0)
    c = whole.counters.cpu().numpy()
    n = whole.n_transitions()
    assert B < n <= whole.capacity and c[0] == c[1] == n, "every decision opened exactly one replay row"
    rows = whole.transitions()
    assert torch.isfinite(rows[:, :23]).all() and (rows[:, 10].abs() <= 1).all()
    ids = rows[:, 23].contiguous().view(torch.int32)
    assert int(ids.min()) == 0 and int(ids.max()) == B - 1
    assert torch.bincount(ids, minlength=B).min() >= 1, "every env took its first decision"
    assert 0 < int(whole.env.stats[13].item()) <= B * K * iters
    assert set(torch.unique(rows[:, 22]).tolist()) <= {0.0, 1.0}
    # shard invariance at full size: sort rows by (env id, slot order) and compare bitwise
    def keyed(s, lo, hi):
        r = s.transitions()
        i = r[:, 23].contiguous().view(torch.int32).to(torch.int64)
        order = torch.argsort(i * (1 << 22) + torch.arange(r.shape[0], device=r.device), stable=True)
        r, i = r[order], i[order]
        m = (i >= lo) & (i < hi)
        return r[m].view(torch.int32)
    for lo in (0, B // 2):
        part = run(B // 2, lo)
        assert torch.equal(keyed(part, lo, lo + B // 2), keyed(whole, lo, lo + B // 2)), f"shard at {lo} differs from the whole batch"


@pytest.mark.parametrize("groups", [2, 3])
def test_chained_collect_graph_equals_plain_launches_and_fills_the_ring_in_one_order(monkeypatch, groups):
    """Several env groups: sit_rollout_collect runs the whole call as independent per-group chains (policy -> K steps -> policy
    ...), joined at the end of the call only, and replays it from ONE graph from the second call with the same arguments
    on.  The groups reserve their replay rows in a fixed order (events between the count kernels), so the ring is identical
    row for row between the graph replay, plain launches and a second run -- whatever the overlap on the GPU was."""
    from sac_maritime_ast_b200 import _lib, rollout as ro
    monkeypatch.setenv("SIT_ROLLOUT_GROUPS", str(groups))
    pol = ro.GaussianPolicy(hidden_size=64, device="cuda:0", seed=3, obs_scale="space", tensor_cores=True)
    B = 1500                                                  # ragged: groups of 768 + 732, or 512 + 512 + 476
    rings = []
    for graph in (True, False, True):
        smp = _make(B, "fp32", pol, cap=200, auto_reset="hard")
        smp.begin()
        for j in range(4):                                    # plain launches, capture, replay, replay
            smp.collect(12, 16, graph=graph)
            if j == 0:                                        # before the 4096-row ring wraps
                torch.cuda.synchronize()
                assert B <= smp.n_opened() <= smp.capacity
                ids = smp.replay[:B, 23].contiguous().view(torch.int32)
                assert torch.equal(ids, torch.arange(B, dtype=torch.int32, device=ids.device)), \
                    "first decisions: group after group, env order"
        smp.collect(5, 16, graph=graph)                       # other arguments: a new graph on their second use
        smp.collect(5, 16, graph=graph)
        smp.collect(70, 4, graph=graph)                       # a long call: 64 iterations per graph, the rest as plain launches
        smp.collect(70, 4, graph=graph)
        torch.cuda.synchronize()
        c = smp.counters.cpu().numpy()
        assert c[_lib.RC_CLOSED] == c[_lib.RC_TRANSITIONS] == c[_lib.RC_DECISIONS] > 2 * B
        rings.append((smp.replay.clone(), c.copy(), smp.env.stats.clone(), {k: smp.env.f[k].clone() for k in ("north", "psi", "flags", "ep_steps")}))
    for other in rings[1:]:
        assert np.array_equal(rings[0][1], other[1]) and torch.equal(rings[0][2], other[2])
        assert torch.equal(rings[0][0].view(torch.int32), other[0].view(torch.int32)), "ring contents differ"
        for k, v in rings[0][3].items():
            assert torch.equal(v, other[3][k]), k


def test_replay_ring_order_is_deterministic_and_closed_rows_are_counted():
    """Two runs with one seed fill the replay ring IDENTICALLY (row for row, not only as a set): the compaction orders
    the pending envs by env index whatever the CTA scheduling.  SIT_RC_CLOSED trails SIT_RC_TRANSITIONS between a policy
    call and the step launch that closes the rows it opened, and catches up after it."""
    from sac_maritime_ast_b200 import _lib, rollout as ro
    pol = ro.GaussianPolicy(hidden_size=64, device="cuda:0", seed=3, obs_scale="space")
    rings = []
    for _ in range(2):
        smp = _make(5000, "fp32", pol, cap=200, auto_reset="hard")
        smp.begin()
        smp.collect(20, 32, graph=True)
        torch.cuda.synchronize()
        c = smp.counters.cpu().numpy()
        assert c[_lib.RC_CLOSED] == c[_lib.RC_TRANSITIONS] == c[_lib.RC_DECISIONS] > 5000
        rings.append((smp.replay.clone(), c.copy()))
        ids = smp.replay[:5000, 23].contiguous().view(torch.int32)
        assert torch.equal(ids, torch.arange(5000, dtype=torch.int32, device=ids.device)), "first decisions are in env order"
        # a policy call opens rows that only the next step launch completes (step on until some env reaches a decision)
        opened = 0
        for _ in range(200):
            c = smp.counters.cpu().numpy()
            smp.policy_step()
            torch.cuda.synchronize()
            c2 = smp.counters.cpu().numpy()
            opened = int(c2[_lib.RC_TRANSITIONS] - c[_lib.RC_TRANSITIONS])
            if opened:
                break
            smp.env_steps(8)
        assert opened > 0 and c2[_lib.RC_CLOSED] == c[_lib.RC_CLOSED] and smp.n_transitions() == int(c[_lib.RC_CLOSED])
        smp.env_steps(1)
        torch.cuda.synchronize()
        c3 = smp.counters.cpu().numpy()
        assert c3[_lib.RC_CLOSED] == c3[_lib.RC_TRANSITIONS]
    assert np.array_equal(rings[0][1], rings[1][1])
    assert torch.equal(rings[0][0].view(torch.int32), rings[1][0].view(torch.int32)), "ring contents differ between two runs"


def test_policy_input_scale_reaches_both_mlp_kernels():
    """SitPolicy.obs_scale: the network sees obs * scale; the replay row keeps the raw observation.  A policy with scale s and
    layer-0 weights w must act exactly like a policy without scale and weights diag(s) w (fp32 CUDA-core MLP, same sums up to
    one rounding), on both MLP kernels."""
    from sac_maritime_ast_b200 import rollout as ro
    for tc in (False, True):
        a = ro.GaussianPolicy(hidden_size=64, device="cuda:0", seed=9, mode="mean", tensor_cores=tc, obs_scale="space")
        b = ro.GaussianPolicy(hidden_size=64, device="cuda:0", seed=9, mode="mean", tensor_cores=tc)
        b.w0.mul_(a.obs_scale[:, None])
        outs = []
        for pol in (a, b):
            smp = _make(512, "fp32", pol)
            smp.begin(); smp.policy_step(); smp.env_steps(1)
            torch.cuda.synchronize()
            rows = smp.transitions()
            assert rows.shape[0] == 512 and float(rows[:, 0].min()) >= 400.0, "replay state is the raw observation"
            outs.append(rows[:, 10].clone())
        assert float((outs[0] - outs[1]).abs().max()) <= (2e-3 if tc else 2e-5)
        assert float(outs[0].abs().max()) < 0.999, "scaled inputs keep tanh(mu) off saturation"
