"""CPU restatement of the device-resident rollout loop -- TEST INFRASTRUCTURE.

Follows the loop contract of the reference's ``test_beds/main_ast.py:310-412`` (reset, init_step, per-step
select_action / convert / step, push on decision steps only with the mask rule of :387) around the C oracle's
``orc_env_step``.  The agent-side pieces that the reference imports from its absent ``ast_sac`` package
(decision schedule, scoping-angle -> waypoint conversion, validity resampling) are restated from the definitions
in ``include/shipsim.h`` (SitRollout) -- PARITY UNPINNED for those; the env step itself is the pinned oracle.
Random draws: Philox4x32-10 keyed by the seed, counter = (global env id lo, hi, decision serial, resample index).
"""
from __future__ import annotations

import ctypes as C
import math

import numpy as np

from oracle import cport

M0, M1 = 0xD2511F53, 0xCD9E8D57
W0, W1 = 0x9E3779B9, 0xBB67AE85
MASK = 0xFFFFFFFF


def philox4x32_10(c, k):
    """Salmon et al. 2011, Philox-4x32 with 10 rounds.  c: 4 uint32, k: 2 uint32 -> 4 uint32."""
    c0, c1, c2, c3 = [int(x) & MASK for x in c]
    k0, k1 = [int(x) & MASK for x in k]
    for _ in range(10):
        p0, p1 = M0 * c0, M1 * c2
        hi0, lo0, hi1, lo1 = p0 >> 32, p0 & MASK, p1 >> 32, p1 & MASK
        c0, c1, c2, c3 = hi1 ^ c1 ^ k0, lo1, hi0 ^ c3 ^ k1, lo0
        k0, k1 = (k0 + W0) & MASK, (k1 + W1) & MASK
    return c0, c1, c2, c3


def draw(seed, gid, serial, tries):
    x = philox4x32_10((gid & MASK, (gid >> 32) & MASK, serial, tries), (seed & MASK, (seed >> 32) & MASK))
    return (x[0] + 0.5) / 4294967296.0, (x[1] + 0.5) / 4294967296.0


def policy_mu_sig(weights, obs10):
    """float32 forward of mlp(obs; H, H, 2) with ReLU hidden layers (nn_models/mlp.py:95-148) and the log-sigma clip
    of distributions/normal.py:86-87."""
    w0, b0, w1, b1, w2, b2 = [np.asarray(w, np.float32) for w in weights]
    x = np.asarray(obs10, np.float32)
    h = np.maximum(x @ w0 + b0, 0).astype(np.float32)
    h = np.maximum(h @ w1 + b1, 0).astype(np.float32)
    out = (h @ w2 + b2).astype(np.float32)
    return float(out[0]), float(np.exp(np.clip(out[1], -20.0, 2.0)))


class RolloutRef:
    """One env.  mode: 'uniform' | 'gaussian' | 'mean'."""

    def __init__(self, p, m, e, gid, seed=25450, mode="uniform", weights=None, num_steps_episode=600,
                 max_route_resampling=1000, sampling_frequency=7, hard_reset=False, truncate=False):
        self.L = cport.lib()
        self.p, self.m, self.e = p, m, e
        self.gid, self.seed, self.mode, self.weights = gid, seed, mode, weights
        self.cap, self.max_resample, self.sf = num_steps_episode, max_route_resampling, sampling_frequency
        self.hard = hard_reset
        self.truncate = truncate     # False: the reference's loop (`while not done`, main_ast.py:322): num_steps_episode only masks
        self.serial = 0
        self.transitions = []          # (state[10], a, r, next_state[10], mask)
        self.decisions = []            # (global step, a, wp_n, wp_e, tries)
        self.steps = 0
        self._begin_episode()

    def _initial_state(self):
        t, o = self.e.test, self.e.obs     # MSRL_env_ex.py:86-90 (float32)
        f = np.float32
        return [float(f(t.init_north)), float(f(t.init_east)), float(f(t.init_psi)), 0.0, 0.0, 0.0,
                float(f(o.init_north)), float(f(o.init_east)), float(f(o.init_psi)), 0.0]

    def _begin_episode(self):
        self.L.orc_env_reset(C.byref(self.e))
        self.L.orc_env_init_step(C.byref(self.p), C.byref(self.p), C.byref(self.e))
        self.state = self._initial_state()
        self.pending, self.n_dec, self.ep_steps, self.fresh = True, 0, 0, True
        self.act = (0.0, 0.0)

    def _decide(self):
        e = self.e
        base = (e.obs.wp0_n[0], e.obs.wp0_e[0]) if self.n_dec == 0 else self.act
        if self.mode != "uniform":
            mu, sig = policy_mu_sig(self.weights, self.state)
        tries, bad = 0, True
        while bad and tries < 1 + self.max_resample:
            u1, u2 = draw(self.seed, self.gid, self.serial, tries)
            if self.mode == "uniform":
                a = float(np.float32(2.0 * u1 - 1.0))
            elif self.mode == "mean":
                a = float(np.tanh(np.float32(mu)))
            else:
                eps = np.float32(math.sqrt(-2.0 * math.log(u1)) * math.cos(2.0 * math.pi * u2))
                a = float(np.tanh(np.float32(np.float32(sig) * eps + np.float32(mu))))
            ang = e.ab_alpha + a * 0.52359877559829887308
            wn, we = base[0] + e.ab_segment_length * math.cos(ang), base[1] + e.ab_segment_length * math.sin(ang)
            m = self.m
            bad = (wn < m.min_north or wn > m.max_north or we < m.min_east or we > m.max_east) or \
                bool(self.L.orc_map_contains(C.byref(m), wn, we))
            tries += 1
            if self.mode == "mean":
                break
        self.decisions.append((self.steps, a, wn, we, tries))
        self.act, self.n_dec, self.serial = (wn, we), self.n_dec + 1, self.serial + 1
        return a

    def step(self):
        """One env step (taking a decision first when one is due)."""
        e = self.e
        su, a = False, None
        if self.pending:
            a = self._decide()
            su, self.pending = True, False
        o10 = (C.c_double * 10)(); r = C.c_double(); d = C.c_int(); st = C.c_int()
        self.L.orc_env_step(C.byref(self.p), C.byref(self.p), C.byref(self.m), C.byref(e), self.act[0], self.act[1],
                            int(su), int(self.fresh), o10, C.byref(r), C.byref(d), C.byref(st))
        self.fresh = False
        self.ep_steps += 1
        self.steps += 1
        nxt, done = list(o10), bool(d.value)
        timeup = self.cap > 0 and self.ep_steps + 1 == self.cap      # episode_steps == num_steps_episode, main_ast.py:387
        if su:
            mask = 1.0 if timeup else float(not done)      # main_ast.py:387
            self.transitions.append((list(self.state), a, r.value, nxt, mask))
        out = dict(obs=nxt, reward=r.value, done=done, status=st.value)
        if done or (timeup and self.truncate):
            if self.hard:
                raise NotImplementedError("hard reset is a product extension; the oracle keeps the reference's reset")
            self._begin_episode()
        else:
            self.state = nxt
            ob = e.obs
            if (not e.stop_obs) and self.n_dec < self.sf and ob.k < ob.n_wp - 1:
                dn, de = ob.wp_n[ob.k] - ob.north, ob.wp_e[ob.k] - ob.east
                if dn * dn + de * de <= self.p.ra * self.p.ra:
                    self.pending = True
        return out
