"""Device-resident SAC-AST rollout: the role of the reference's (empty)
``ast_core/samplers/intermediate_waypoint_sampler.py`` with the loop contract of
``test_beds/main_ast.py:310-412``.

``IntermediateWaypointSampler`` drives a batched ``MultiShipRLEnv`` entirely on the GPU: between step
launches ``sit_rollout_policy`` gives every env that reached a decision event an action (uniform warm-up
policy or the tanh-Gaussian MLP policy of ``ast_core/policies/gaussian_policy.py``), converts the scoping
angle to an intermediate waypoint, validates it against the map and opens a replay row;
``sit_rollout_step_k`` advances the envs, closes the rows and restarts finished episodes.  Nothing returns
to the host inside the loop.

What the reference does NOT pin (its ``ast_sac`` package is absent: ``agent.select_action``,
``agent.convert_action_to_simu_input``, main_ast.py:337-347,367) is defined in ``include/shipsim.h`` next
to ``SitRollout`` and restated for the tests in ``oracle/rollout_ref.py`` -- parity unpinned there.

``gather_replay`` / ``reduce_stats`` are the only collectives of the framework (NCCL on GPUs, gloo in the
CPU tests): they run after a rollout segment, never on the step path.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional, Sequence

import numpy as np
import torch

from . import _lib
from .core import nvtx_range, ptr
from .env import MultiShipRLEnv

N_OBS, TR_FLOATS = _lib.N_OBS, _lib.TR_FLOATS
LOG_SIG_CAP_MAX, LOG_SIG_CAP_MIN = 2.0, -20.0      # distributions/normal.py:14-15
EPS = 1e-6                                         # policies/gaussian_policy.py:16
# observation_space.high (MSRL_env_ex.py:67-78): dividing by it brings every channel of the raw observation (metres, rpm, kW)
# into [-1, 1]; unscaled, |mu| runs into the hundreds, tanh saturates to exactly +-1 in fp32 and every policy gradient is 0
OBS_HIGH = (10000.0, 20000.0, float(np.pi), 3000.0, 1000.0, 2000.0, 10000.0, 20000.0, float(np.pi), 1000.0)


def default_obs_scale():
    return tuple(1.0 / h for h in OBS_HIGH)


class GaussianPolicy:
    """Weights of the reference's Gaussian policy network: ``mlp(obs; hidden_layer_sizes, ReLU, linear 2*Da)``
    (nn_models/mlp.py:95-148, distributions/normal.py:78-83), Da = 1.  Glorot-uniform weights and zero biases
    like ``_weight_variable`` / ``_bias_variable`` (mlp.py:13-51).  Both hidden layers share one width H
    (multiple of 32, <= 256: main_ast.py:67 uses 256, the class default is (100, 100) -> use 128 or 96)."""

    def __init__(self, hidden_size: int = 256, device=None, seed: int = 25450, mode: str = "gaussian",
                 tensor_cores: bool = False, obs_scale=None):
        if hidden_size % 32 or not 32 <= hidden_size <= 256:
            raise ValueError("hidden_size must be a multiple of 32 in [32, 256]")
        self.hidden, self.device, self.mode = hidden_size, torch.device(device or "cuda"), mode
        # tensor_cores: hidden layer as tcgen05.mma with bf16 operands / fp32 accumulate (SIT_POLICY_TC_BF16)
        self.tensor_cores = bool(tensor_cores)
        g = torch.Generator().manual_seed(seed)

        def glorot(fan_in, fan_out):
            lim = float(np.sqrt(6.0 / (fan_in + fan_out)))
            return ((torch.rand(fan_in, fan_out, generator=g) * 2 - 1) * lim).to(self.device).contiguous()
        H = hidden_size
        self.w0, self.b0 = glorot(N_OBS, H), torch.zeros(H, device=self.device)
        self.w1, self.b1 = glorot(H, H), torch.zeros(H, device=self.device)
        self.w2, self.b2 = glorot(H, 2), torch.zeros(2, device=self.device)
        # per-channel input scale (None: the network sees the raw observation; "space": 1 / observation_space.high)
        if isinstance(obs_scale, str):
            obs_scale = default_obs_scale()
        self.obs_scale = None if obs_scale is None else torch.tensor(obs_scale, dtype=torch.float32, device=self.device)

    def parameters(self):
        return [self.w0, self.b0, self.w1, self.b1, self.w2, self.b2]

    def struct(self) -> "_lib.SitPolicy":
        p = _lib.SitPolicy()
        p.mode = {"uniform": _lib.POLICY_UNIFORM, "gaussian": _lib.POLICY_GAUSSIAN, "mean": _lib.POLICY_MEAN}[self.mode]
        if self.tensor_cores and self.mode != "uniform":
            p.mode |= _lib.POLICY_TC_BF16
        p.hidden = self.hidden
        for n in ("w0", "b0", "w1", "b1", "w2", "b2"):
            t = getattr(self, n)
            assert t.dtype == torch.float32 and t.is_contiguous() and t.device.type == "cuda"
            setattr(p, n, t.data_ptr())
        p.obs_scale = None if self.obs_scale is None else self.obs_scale.data_ptr()
        return p

    # plain torch restatement (used by the tests as the fp32 reference of the policy kernel)
    def mu_logsig(self, obs: torch.Tensor):
        x = obs.float() if self.obs_scale is None else obs.float() * self.obs_scale
        h = torch.relu(x @ self.w0 + self.b0)
        if self.tensor_cores:     # the kernel rounds both operands of the hidden-layer product to bf16 (fp32 accumulate)
            h = torch.relu(h.bfloat16().float() @ self.w1.bfloat16().float() + self.b1)
        else:
            h = torch.relu(h @ self.w1 + self.b1)
        out = h @ self.w2 + self.b2
        return out[:, 0], out[:, 1].clamp(LOG_SIG_CAP_MIN, LOG_SIG_CAP_MAX)


class UniformPolicy:
    """policies/uniform_policy.py:7-22: actions uniform in [-1, 1] (the start_steps warm-up, main_ast.py:335-341)."""
    mode = "uniform"

    def struct(self):
        p = _lib.SitPolicy()
        p.mode, p.hidden = _lib.POLICY_UNIFORM, 0
        return p


class IntermediateWaypointSampler:
    """Rollout collection for a batched MultiShipRLEnv (device resident).

    replay rows are float32 [capacity, 24]: state[0:10], action[10], reward[11], next_state[12:22], mask[22],
    global env id (int32 bits) [23] -- ``memory.push(state, action, reward, next_state, mask)`` of
    main_ast.py:396, pushed on decision steps only (:395)."""

    def __init__(self, env: MultiShipRLEnv, policy, capacity: Optional[int] = None, seed: int = 25450,
                 env_id_offset: int = 0, num_steps_episode: int = 600, max_route_resampling: int = 1000,
                 auto_reset="reference", truncate_episodes: bool = False):
        """num_steps_episode follows the reference (main_ast.py:322,387): the row that closes at that episode step gets
        mask = 1 and the episode goes on until a terminal state.  truncate_episodes=True is an extension that also ends
        (resets) the episode there."""
        self.env, self.policy = env, policy
        self.capacity = int(capacity if capacity is not None else max(env.n_envs * 8, 1024))
        if self.capacity < env.n_envs:
            raise ValueError("replay capacity must be >= n_envs")
        self.replay = torch.zeros((self.capacity, TR_FLOATS), dtype=torch.float32, device=env.device)
        self.counters = torch.zeros(_lib.RC_N, dtype=torch.int64, device=env.device)
        self.ro = _lib.SitRollout()
        self.ro.seed, self.ro.env_id_offset = int(seed), int(env_id_offset)
        self.ro.max_episode_steps, self.ro.max_route_resampling = int(num_steps_episode), int(max_route_resampling)
        self.ro.replay, self.ro.capacity, self.ro.counters = self.replay.data_ptr(), self.capacity, self.counters.data_ptr()
        self.ro.truncate_episodes = int(bool(truncate_episodes))
        self.io = _lib.SitEnvIO()
        self.io.auto_reset = 2 if auto_reset == "hard" else 1
        self.io.rec_steps = 1
        self.io.stats = env.stats.data_ptr()
        self.launches = 0
        self._side = None
        self._pol = policy.struct()

    def set_policy(self, policy):
        self.policy, self._pol = policy, policy.struct()

    def begin(self):
        """reset() + init_step() for every env and arm the first decision (main_ast.py:314,325-329)."""
        e = self.env
        e.reset()
        e.init_step()
        _lib.check(e.L.sit_rollout_begin(e.h, e.state_ptr(), e.stream()))
        self.launches += 1

    def policy_step(self):
        e = self.env
        _lib.check(e.L.sit_rollout_policy(e.h, e.state_ptr(), C.byref(self._pol), C.byref(self.ro), e.stream()))
        self.launches += 1

    def env_steps(self, K: int):
        e = self.env
        _lib.check(e.L.sit_rollout_step_k(e.h, e.state_ptr(), int(K), C.byref(self.ro), C.byref(self.io), e.stream()))
        self.launches += 1

    @nvtx_range("sit.rollout.collect")
    def collect(self, iterations: int, K: int = 16, graph: bool = True):
        """`iterations` x (policy on the pending envs, then K env steps).  Enqueues only.  With ``graph`` the loop runs
        inside libshipsim (sit_rollout_collect): one iteration is captured into a CUDA graph on a side stream (the
        legacy default stream cannot be captured) and replayed."""
        e = self.env
        if iterations <= 0:
            return
        if not graph:
            for _ in range(iterations):
                self.policy_step()
                self.env_steps(K)
            return
        cur = torch.cuda.current_stream(e.device)
        if self._side is None:
            self._side = torch.cuda.Stream(device=e.device)
        self._side.wait_stream(cur)
        _lib.check(e.L.sit_rollout_collect(e.h, e.state_ptr(), int(iterations), int(K), C.byref(self._pol), C.byref(self.ro),
                                           C.byref(self.io), 1, self._side.cuda_stream))
        cur.wait_stream(self._side)
        # per env group: gather + mlp + sample + resample + step kernels; plus one w1 pack per iteration (tensor cores)
        groups = C.c_int32(1)
        _lib.check(e.L.sit_rollout_group_count(e.h, C.byref(groups)))
        tc = bool(getattr(self.policy, 'tensor_cores', False))
        per_group = 5 + (0 if isinstance(self.policy, UniformPolicy) else 1)   # count, gather, [mlp], sample, resample, step
        # (several groups: one pack per call, the weights do not change inside it)
        self.launches += per_group * groups.value * iterations + (1 if tc else 0) * (iterations if groups.value == 1 else 1)

    # ---- results ----------------------------------------------------------------------------------------------
    def n_transitions(self) -> int:
        """Completed replay rows (SIT_RC_CLOSED: reward, next_state and mask written).  Between sit_rollout_policy and the
        next step launch the rows just opened are not counted."""
        return int(self.counters[_lib.RC_CLOSED].item())

    def n_opened(self) -> int:
        return int(self.counters[_lib.RC_TRANSITIONS].item())

    def transitions(self) -> torch.Tensor:
        """Completed replay rows pushed so far (ring order), [n, 24] float32."""
        n = min(self.n_transitions(), self.capacity)
        return self.replay[:n]

    @staticmethod
    def split(rows: torch.Tensor) -> Dict[str, torch.Tensor]:
        return {"state": rows[:, 0:10], "action": rows[:, 10:11], "reward": rows[:, 11], "next_state": rows[:, 12:22],
                "mask": rows[:, 22], "env_id": rows[:, 23].contiguous().view(torch.int32)}


# ---- collectives: after a rollout segment, never on the step path ------------------------------------------------
def gather_replay(rows: torch.Tensor, group=None) -> torch.Tensor:
    """All-gather of the compacted transitions of every rank: [n_r, 24] per rank -> [sum n_r, 24] on every rank
    (rank order).  Counts are exchanged first, rows travel in one padded all_gather."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return rows
    world = dist.get_world_size(group)
    n = torch.tensor([rows.shape[0]], dtype=torch.int64, device=rows.device)
    counts = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(counts, n, group=group)
    counts = [int(c.item()) for c in counts]
    cap = max(max(counts), 1)
    padded = torch.zeros((cap, rows.shape[1]), dtype=rows.dtype, device=rows.device)
    padded[:rows.shape[0]] = rows
    out = torch.empty((world * cap, rows.shape[1]), dtype=rows.dtype, device=rows.device)
    dist.all_gather_into_tensor(out, padded, group=group)
    return torch.cat([out[r * cap:r * cap + counts[r]] for r in range(world)], dim=0)


class ReplayExchange:
    """The post-segment exchange with NOTHING on the critical path: fixed-capacity blocks, no count exchange, no host
    synchronisation, collective on a side stream.

    A rollout segment of ``iterations`` policy calls opens at most ``iterations * n_envs`` rows per rank (one decision per env
    and policy call), so every rank ships one block of exactly ``1 + block_rows`` rows (header row with the count, then the
    rows; layout in include/shipsim.h next to sit_replay_pack).  ``pack()`` copies the segment's rows out of the sampler ring
    on the caller's stream (sit_replay_pack: the next segment may then overwrite the ring); ``exchange()`` runs the all-gather
    and the append (sit_replay_append: block order = rank order, row order) on a side stream and returns an event: the next
    ``sampler.collect`` overlaps it and only the consumer of the memory (the learner) waits.  With one rank the block is
    appended directly.  CPU tensors (the world-size-2 gloo tests of the host logic) take the same steps with torch ops."""

    def __init__(self, sampler: "IntermediateWaypointSampler", block_rows: int, group=None):
        self.smp, self.block, self.group = sampler, int(block_rows), group
        dev = sampler.replay.device
        self.cuda = dev.type == "cuda"
        import torch.distributed as dist
        self.world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
        self.seen = torch.zeros(1, dtype=torch.int64, device=dev)        # SIT_RC_CLOSED at the previous pack (device)
        self.dropped = torch.zeros(1, dtype=torch.int64, device=dev)     # rows a too-small block had to leave behind
        self.send = torch.zeros((1 + self.block, TR_FLOATS), dtype=torch.float32, device=dev)
        self.recv = (torch.zeros((self.world * (1 + self.block), TR_FLOATS), dtype=torch.float32, device=dev)
                     if self.world > 1 else self.send)
        self.side = torch.cuda.Stream(device=dev) if self.cuda else None
        self.done = None
        self.L = _lib.lib() if self.cuda else None

    @nvtx_range("sit.replay.pack")
    def pack(self):
        """Copy the rows closed since the previous pack into the send block (caller's stream, no host sync)."""
        smp = self.smp
        if self.cuda:
            cur = torch.cuda.current_stream(self.send.device)
            if self.done is not None:
                cur.wait_event(self.done)                                   # the previous exchange has read `send`
            _lib.check(self.L.sit_replay_pack(C.byref(smp.ro), self.seen.data_ptr(), self.dropped.data_ptr(),
                                              self.send.data_ptr(), self.block, cur.cuda_stream))
            return
        n_now = int(smp.counters[_lib.RC_CLOSED])
        fresh = n_now - int(self.seen)
        n = min(fresh, self.block)
        idx = (torch.arange(n_now - n, n_now)) % smp.capacity
        self.send[1:1 + n] = smp.replay[idx]
        self.send[0].view(torch.int64)[0] = n
        self.dropped += max(fresh - self.block, 0)
        self.seen.fill_(n_now)

    @nvtx_range("sit.replay.exchange")
    def exchange(self, memory) -> Optional["torch.cuda.Event"]:
        """All-gather the packed blocks of every rank and append the valid rows to ``memory`` (rank order, row order)."""
        import torch.distributed as dist
        if not self.cuda:
            if self.world > 1:
                dist.all_gather_into_tensor(self.recv, self.send, group=self.group)
            memory.append_blocks(self.recv, self.world, self.block)
            return None
        cur = torch.cuda.current_stream(self.send.device)
        self.side.wait_stream(cur)
        with torch.cuda.stream(self.side):
            if self.world > 1:
                dist.all_gather_into_tensor(self.recv, self.send, group=self.group)
            memory.append_blocks(self.recv, self.world, self.block)
            self.done = self.side.record_event()
        return self.done

    def wait(self):
        if self.cuda and self.done is not None:
            torch.cuda.current_stream(self.send.device).wait_event(self.done)

    def rows(self) -> torch.Tensor:
        """The valid rows of this rank's last packed block (host sync: for logging only)."""
        return self.send[1:1 + int(self.send[0].view(torch.int64)[0].item())]


def reduce_stats(stats: torch.Tensor, group=None) -> torch.Tensor:
    """Sum of the 16 episode counters (12 status bits, done steps, env steps, ...) over ranks."""
    import torch.distributed as dist
    out = stats.clone()
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(out, op=dist.ReduceOp.SUM, group=group)
    return out


def shard_range(n_total: int, rank: int, world: int):
    """Contiguous env block [lo, hi) of `rank`; the block start is the sampler's env_id_offset so that
    every env draws the same random numbers whatever the number of shards."""
    base, rem = divmod(n_total, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)
