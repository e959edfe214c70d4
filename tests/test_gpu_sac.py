"""GPU numerics of the SAC learner kernels (sit_sac_update) against a plain PyTorch fp32 autograd restatement of the same
update -- standard twin-Q soft actor-critic with the hyper-parameters of main_ast.py:41-68 -- driven by the SAME batch
indices and normal draws (read back from the learner's workspace).  The reference's own agent (ast_sac) is absent, so this
restatement is the definition both sides are held to: parity unpinned.  Tolerance: 2e-4 relative on every parameter
tensor after 1 and after 5 updates (fp32 sums in different orders), losses 1e-3."""
import copy
import math

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TOL = 2e-4


def _fake_replay(n, seed=0):
    g = torch.Generator().manual_seed(seed)
    rows = torch.zeros(n, 24)
    rows[:, 0:10] = torch.randn(n, 10, generator=g)
    rows[:, 10] = torch.rand(n, generator=g) * 2 - 1
    rows[:, 11] = torch.randn(n, generator=g) * 0.5
    rows[:, 12:22] = rows[:, 0:10] + 0.1 * torch.randn(n, 10, generator=g)
    rows[:, 22] = (torch.rand(n, generator=g) > 0.2).float()
    return rows


class TorchSAC:
    """The restatement: torch.nn-free autograd + torch.optim.Adam on clones of the learner's initial parameters."""

    def __init__(self, agent):
        self.cfg = agent.cfg
        self.obs_scale = agent.f["obs_scale"].clone()       # SitSacConfig.obs_scale: every network sees s * obs_scale
        self.nets = {}
        for net in ("policy", "q1", "q2", "q1_target", "q2_target"):
            self.nets[net] = [(w.clone().requires_grad_(not net.endswith("target")), b.clone().requires_grad_(not net.endswith("target")))
                              for w, b in agent.layers(net)]
        self.log_alpha = torch.zeros(1, device=DEV, requires_grad=True)
        self.alpha = float(self.cfg.alpha)
        flat = lambda net: [t for wb in self.nets[net] for t in wb]
        self.opt_q = torch.optim.Adam(flat("q1") + flat("q2"), lr=self.cfg.lr)
        self.opt_pi = torch.optim.Adam(flat("policy"), lr=self.cfg.lr)
        self.opt_a = torch.optim.Adam([self.log_alpha], lr=self.cfg.lr)
        self.updates = 0

    @staticmethod
    def mlp(layers, x):
        (w0, b0), (w1, b1), (w2, b2) = layers
        h = torch.relu(x @ w0 + b0)
        h = torch.relu(h @ w1 + b1)
        return h @ w2 + b2

    def pi(self, s, eps):
        out = self.mlp(self.nets["policy"], s)
        mu, ls = out[:, 0], out[:, 1].clamp(-20.0, 2.0)
        x = mu + ls.exp() * eps
        a = torch.tanh(x)
        logp = -0.5 * eps * eps - ls - 0.5 * math.log(2 * math.pi) - torch.log(1 - a * a + 1e-6)
        return a, logp

    def q(self, which, s, a):
        return self.mlp(self.nets[which], torch.cat([s, a[:, None]], 1))[:, 0]

    def update(self, rows, eps_next, eps_cur):
        c = self.cfg
        s, a, r, s2, m = rows[:, 0:10] * self.obs_scale, rows[:, 10], rows[:, 11], rows[:, 12:22] * self.obs_scale, rows[:, 22]
        with torch.no_grad():
            a2, lp2 = self.pi(s2, eps_next)
            qt = torch.min(self.q("q1_target", s2, a2), self.q("q2_target", s2, a2)) - self.alpha * lp2
            y = r + m * c.gamma * qt
        l1 = ((self.q("q1", s, a) - y) ** 2).mean()
        l2 = ((self.q("q2", s, a) - y) ** 2).mean()
        self.opt_q.zero_grad(); (l1 + l2).backward(); self.opt_q.step()
        ap, lp = self.pi(s, eps_cur)
        pl = (self.alpha * lp - torch.min(self.q("q1", s, ap), self.q("q2", s, ap))).mean()
        self.opt_pi.zero_grad(); pl.backward(); self.opt_pi.step()
        al = torch.zeros(())
        if c.automatic_entropy_tuning:
            al = -(self.log_alpha * (lp.detach() + (-1.0))).mean()
            self.opt_a.zero_grad(); al.backward(); self.opt_a.step()
            self.alpha = float(self.log_alpha.exp().item())
        if self.updates % c.target_update_interval == 0:
            with torch.no_grad():
                for src, dst in (("q1", "q1_target"), ("q2", "q2_target")):
                    for (w, b), (wt, bt) in zip(self.nets[src], self.nets[dst]):
                        wt.mul_(1 - c.tau).add_(c.tau * w); bt.mul_(1 - c.tau).add_(c.tau * b)
        self.updates += 1
        return float(l1.detach()), float(l2.detach()), float(pl.detach()), float(al.detach()), self.alpha


def _rel(a, b):
    return float((a - b).abs().max() / (b.abs().max() + 1e-6))


@pytest.mark.parametrize("hidden,batch,graph", [(64, 64, False), (256, 64, True), (128, 300, True)])
def test_update_matches_torch_autograd(hidden, batch, graph):
    from sac_maritime_ast_b200 import sac
    mem = sac.ReplayMemory(4096, device=DEV)
    mem.extend(_fake_replay(1000).to(DEV))
    # the fake replay holds unit-scale observations: no input scale, except one case with a non-trivial per-channel scale
    scale = tuple(0.5 + 0.1 * c for c in range(10)) if batch == 300 else None
    agent = sac.SAC(10, None, None, device=DEV, hidden_size=hidden, batch_size=batch, use_graph=graph, seed=11, obs_scale=scale)
    assert torch.allclose(agent.f["obs_scale"], torch.tensor(scale if scale else [1.0] * 10, device=DEV))
    ref = TorchSAC(agent)
    assert len(mem) == 1000
    for u in range(5):
        out = agent.update_parameters(mem, batch, u)
        torch.cuda.synchronize()
        idx = agent.f["idx"].long()
        assert int(idx.min()) >= 0 and int(idx.max()) < 1000
        want = ref.update(mem.buffer[idx], agent.f["eps_next"].clone(), agent.f["eps_cur"].clone())
        for k in range(3):
            assert abs(out[k] - want[k]) <= 1e-3 * max(1.0, abs(want[k])), (u, k, out, want)
        assert abs(out[3] - want[3]) <= 1e-3 and abs(out[4] - want[4]) <= 1e-5, (u, out, want)
        if u in (0, 4):
            for net in ("policy", "q1", "q2", "q1_target", "q2_target"):
                for li, ((w, b), (wr, br)) in enumerate(zip(agent.layers(net), ref.nets[net])):
                    assert _rel(w, wr.detach()) <= TOL, (u, net, li, "w", _rel(w, wr.detach()))
                    assert (b - br.detach()).abs().max() <= TOL * max(1.0, float(br.abs().max())), (u, net, li, "b")
    assert agent.updates == 5
    eps = torch.cat([agent.f["eps_next"], agent.f["eps_cur"]])
    assert torch.isfinite(eps).all()


def test_initialisation_and_determinism():
    """Glorot-uniform weights inside the limit, zero biases, targets == critics; two learners with one seed stay
    bit-identical over 20 graph-replayed updates; a different seed differs."""
    from sac_maritime_ast_b200 import sac
    mem = sac.ReplayMemory(2048, device=DEV)
    mem.extend(_fake_replay(2048, seed=3).to(DEV))
    a = sac.SAC(10, None, None, device=DEV, hidden_size=128, batch_size=64, seed=5, obs_scale=None)
    b = sac.SAC(10, None, None, device=DEV, hidden_size=128, batch_size=64, seed=5, obs_scale=None)
    c = sac.SAC(10, None, None, device=DEV, hidden_size=128, batch_size=64, seed=6, obs_scale=None)
    for net, kin, nout in (("policy", 10, 2), ("q1", 11, 1), ("q2", 11, 1)):
        for (w, bias), (K, N) in zip(a.layers(net), ((kin, 128), (128, 128), (128, nout))):
            lim = math.sqrt(6.0 / (K + N))
            assert w.shape == (K, N) and float(w.abs().max()) <= lim and float(w.abs().max()) > 0.8 * lim
            assert abs(float(w.mean())) < 0.1 * lim and float(bias.abs().max()) == 0.0
    assert torch.equal(a.f["q1"], a.f["q1_target"]) and torch.equal(a.f["q2"], a.f["q2_target"])
    assert not torch.equal(a.f["q1"], a.f["q2"])
    for ag in (a, b, c):
        ag.update(mem, 20)
    torch.cuda.synchronize()
    assert a.updates == 20
    for k in ("policy", "q1", "q2", "q1_target", "q2_target", "log_alpha", "adam_m", "adam_v"):
        assert torch.equal(a.f[k], b.f[k]), k
        assert torch.isfinite(a.f[k]).all()
    assert not torch.equal(a.f["policy"], c.f["policy"])


def test_learner_policy_drives_the_rollout():
    """The SitPolicy handed to the sampler points into the learner's workspace: rollout -> replay ring -> updates ->
    rollout again sees the new weights, nothing crosses to the host."""
    import gpu_helpers as G
    import scenarios as sc
    from sac_maritime_ast_b200 import rollout as ro, sac
    env = G.make_env(sc.ENV_SCENARIOS["default"], n_envs=256, precision="fp32")
    agent = sac.SAC(10, None, None, device=DEV, hidden_size=64, batch_size=64, seed=1)
    # raw observations are ~1e4 m: the learner's default obs_scale (1 / observation_space.high) brings them into [-1, 1]
    assert float(agent.f["obs_scale"][0]) == pytest.approx(1e-4)
    smp = ro.IntermediateWaypointSampler(env, agent.policy(), capacity=4096, num_steps_episode=200, auto_reset="hard",
                                         truncate_episodes=True)
    mem = sac.ReplayMemory.from_sampler(smp)
    smp.begin()
    smp.collect(30, 16)
    torch.cuda.synchronize()
    n0 = len(mem)
    assert n0 >= 256
    before = agent.f["policy"].clone()
    ls = agent.update(mem, 10)
    smp.collect(30, 16)
    torch.cuda.synchronize()
    assert len(mem) > n0 and torch.isfinite(ls).all()
    assert not torch.equal(before, agent.f["policy"])
    rows = smp.transitions()
    assert torch.isfinite(rows[:, :23]).all() and float(rows[:, 10].abs().max()) <= 1.0
    # the policy is not saturated: a non-trivial share of the actions lies strictly inside (-1, 1) (with raw
    # observations tanh(mu) is exactly +-1 and every policy gradient is 0)
    assert float((rows[:, 10].abs() < 0.999).float().mean()) > 0.5


def _grad_blocks(agent):
    """name -> gradient tensor of every layer block (weights + bias row) as the last update left them."""
    H, out, o = agent.hidden, {}, 0
    g = agent.f["grad"]
    for net, kin, nout in (("policy", 10, 2), ("q1", 11, 1), ("q2", 11, 1)):
        for li, (K, N) in enumerate(((kin, H), (H, H), (H, nout))):
            out[f"{net}.l{li}"] = g[o:o + (K + 1) * N].view(K + 1, N)
            o += (K + 1) * N
    return out


@pytest.mark.parametrize("hidden,batch,fused", [(256, 1024, True), (128, 4096, True), (256, 1500, True), (256, 1024, False),
                                                 (160, 1024, False), (256, 64, True), (128, 300, True), (256, 200, False), (256, 2048, True), (256, 2100, True)])
def test_tensor_core_update_matches_fp32_update(hidden, batch, fused, monkeypatch):
    """Large-batch learner: the hidden-layer products and every contraction over the batch run on tcgen05 (bf16 operands,
    fp32 accumulate in TMEM, split-k partials summed in a fixed order).  Against the fp32 CUDA-core update with the same
    seed, batch indices and normal draws: every layer's gradient within 4e-2 relative (L2), losses within 1e-2 -- the
    bf16 tolerance (operands carry 8 significant bits = 0.4 % per element, sums over >= 128 terms average the rounding
    out; the layer-0 gradient sits behind three bf16 products in a row and shows the largest error, 2.4e-2); the same
    update twice is bit-identical."""
    from sac_maritime_ast_b200 import sac
    # fused: a network's forward and its input-gradient chain as one kernel each (hidden a multiple of 64; 30 launches per
    # update); otherwise every product is its own launch (43) -- by width (160) or switched off through the environment
    if not fused:
        monkeypatch.setenv("SIT_SAC_FUSED_FWD", "0")
    mem = sac.ReplayMemory(8192, device=DEV)
    mem.extend(_fake_replay(8192, seed=2).to(DEV))
    kw = dict(device=DEV, hidden_size=hidden, batch_size=batch, seed=21, obs_scale=None)
    want = True if batch >= 1024 else 2          # 2: the tensor-core path below batch 1024 too (opt-in)
    tc, tc2, ref = (sac.SAC(10, None, None, tensor_cores=want, **kw), sac.SAC(10, None, None, tensor_cores=want, **kw),
                    sac.SAC(10, None, None, tensor_cores=False, **kw))
    assert torch.equal(tc.f["policy"], ref.f["policy"]) and torch.equal(tc.f["q1"], ref.f["q1"])
    assert tc.kernels_per_update == (22 if fused else 37) + (6 if batch >= 1024 else (4 if (fused and batch >= 256) else 0)) and ref.kernels_per_update == 37
    outs = [a.update_parameters(mem, batch, 0) for a in (tc, tc2, ref)]
    torch.cuda.synchronize()
    assert torch.equal(tc.f["idx"], ref.f["idx"]) and torch.equal(tc.f["eps_cur"], ref.f["eps_cur"])
    for k in range(3):
        assert abs(outs[0][k] - outs[2][k]) <= 1e-2 * max(1.0, abs(outs[2][k])), (k, outs[0], outs[2])
    gt, gr = _grad_blocks(tc), _grad_blocks(ref)
    for name in gr:
        rel = float((gt[name] - gr[name]).norm() / (gr[name].norm() + 1e-12))
        # (below batch 1024 the opt-in path sums the bf16 rounding over fewer rows: 5.3e-2 measured for q1.l0 at batch 300)
        assert rel <= (4e-2 if batch >= 1024 else 8e-2), (name, rel)
        assert float(gr[name].norm()) > 0
    for a in (tc, tc2, ref):
        a.update(mem, 6)
    torch.cuda.synchronize()
    for k in ("policy", "q1", "q2", "q1_target", "q2_target", "adam_m", "adam_v", "log_alpha"):
        assert torch.equal(tc.f[k], tc2.f[k]), f"{k}: two tensor-core learners with one seed diverged"
        assert torch.isfinite(tc.f[k]).all()
    # seven Adam steps of 3e-4: parameters of the two arithmetic paths stay within a few steps of each other
    assert float((tc.f["policy"] - ref.f["policy"]).abs().max()) <= 7 * 2 * 3e-4 + 1e-6
    assert tc.updates == ref.updates == 7


def test_tensor_core_learner_below_the_batch_threshold_is_the_fp32_path():
    from sac_maritime_ast_b200 import sac
    mem = sac.ReplayMemory(2048, device=DEV)
    mem.extend(_fake_replay(2048, seed=4).to(DEV))
    a = sac.SAC(10, None, None, device=DEV, hidden_size=128, batch_size=512, seed=3, tensor_cores=True, obs_scale=None)
    b = sac.SAC(10, None, None, device=DEV, hidden_size=128, batch_size=512, seed=3, tensor_cores=False, obs_scale=None)
    a.update(mem, 3); b.update(mem, 3)
    torch.cuda.synchronize()
    assert torch.equal(a.f["policy"], b.f["policy"]) and torch.equal(a.f["q1"], b.f["q1"])
