"""SAC learner on the device replay ring -- the step after the rollout path.

The reference drives an agent object from its absent ``ast_sac`` package (``ast_core/algorithms/sac.py`` is empty):

    memory = ReplayMemory(args.replay_size, args.seed)                       # main_ast.py (setup)
    memory.push(state, action, reward, next_state, mask)                     # main_ast.py:396
    critic_1_loss, critic_2_loss, policy_loss, ent_loss, alpha = \
        agent.update_parameters(memory, args.batch_size, updates)            # main_ast.py:354

``SAC`` and ``ReplayMemory`` keep those names and signatures; the update rule (standard twin-Q soft actor-critic, the
hyper-parameters of main_ast.py:41-68) is defined in ``include/shipsim.h`` next to ``sit_sac_update`` -- parity unpinned,
checked against a plain PyTorch autograd restatement (tests/test_gpu_sac.py).  Everything runs in libshipsim's kernels on
one caller-owned workspace tensor; there is no torch.nn / torch.optim on this path and no CPU fallback.

Multi-GPU: the learner is *replicated*, not sharded.  Every rank all-gathers the transitions of a rollout segment
(``rollout.gather_replay``) into the same replay ring and runs the same Philox-keyed updates, so the 70 k-parameter
networks stay bit-identical on all ranks without a gradient all-reduce.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, Optional

import torch

from . import _lib
from .core import nvtx_range, ptr

TR_FLOATS = _lib.TR_FLOATS


class ReplayMemory:
    """Device replay ring with the reference's call names (``push``, ``sample``, ``__len__``).  Rows are the 24-float
    transitions of ``include/shipsim.h`` (SIT_TR_FLOATS).  ``IntermediateWaypointSampler`` writes its ring directly; this
    class wraps that ring (``from_sampler``) or owns one that is filled with ``push`` / ``extend``."""

    def __init__(self, capacity: int, seed: int = 25450, device="cuda"):
        self.capacity, self.seed, self.device = int(capacity), int(seed), torch.device(device)
        self.buffer = torch.zeros((self.capacity, TR_FLOATS), dtype=torch.float32, device=self.device)
        self.count = torch.zeros(1, dtype=torch.int64, device=self.device)      # rows ever pushed (device counter)
        self._gen = torch.Generator(device=self.device).manual_seed(self.seed)

    @classmethod
    def from_sampler(cls, sampler) -> "ReplayMemory":
        m = cls.__new__(cls)
        m.capacity, m.seed, m.device = sampler.capacity, int(sampler.ro.seed), sampler.replay.device
        # completed rows only (SIT_RC_CLOSED): a row opened by the policy call is not drawn before its step has closed it
        m.buffer, m.count = sampler.replay, sampler.counters[_lib.RC_CLOSED:_lib.RC_CLOSED + 1]
        m._gen = torch.Generator(device=m.device).manual_seed(m.seed)
        return m

    def __len__(self) -> int:
        return int(min(int(self.count.item()), self.capacity))

    def extend(self, rows: torch.Tensor):
        """Append [n, 24] transitions (e.g. the rows gathered from all ranks) in ring order."""
        n = rows.shape[0]
        if n == 0:
            return
        start = int(self.count.item())
        pos = (start + torch.arange(n, device=self.device)) % self.capacity
        self.buffer[pos] = rows.to(self.device, torch.float32)
        self.count += n

    def append_blocks(self, blocks: torch.Tensor, n_blocks: int, block_rows: int):
        """Append the valid rows of ``n_blocks`` exchange blocks ([1 + block_rows][24] each: header row with the int64 count,
        then the rows; include/shipsim.h sit_replay_append) in block order, row order, WITHOUT a host synchronisation on the
        device path.  The fixed-capacity exchange of ``rollout.ReplayExchange`` ends here."""
        if self.device.type == "cuda":
            _lib.check(_lib.lib().sit_replay_append(self.buffer.data_ptr(), self.capacity, self.count.data_ptr(),
                                                    blocks.data_ptr(), int(n_blocks), int(block_rows),
                                                    torch.cuda.current_stream(self.device).cuda_stream))
            return
        b = blocks.view(n_blocks, 1 + block_rows, TR_FLOATS)           # host logic of the gloo tests
        for r in range(n_blocks):
            n = max(0, min(int(b[r, 0].view(torch.int64)[0]), block_rows))
            self.extend(b[r, 1:1 + n])

    def push(self, state, action, reward, next_state, mask):
        row = torch.zeros(TR_FLOATS, dtype=torch.float32, device=self.device)
        row[0:10] = torch.as_tensor(state, dtype=torch.float32, device=self.device)
        row[10] = float(torch.as_tensor(action).reshape(-1)[0])
        row[11] = float(reward)
        row[12:22] = torch.as_tensor(next_state, dtype=torch.float32, device=self.device)
        row[22] = float(mask)
        self.extend(row[None])

    def sample(self, batch_size: int):
        n = len(self)
        idx = torch.randint(0, max(n, 1), (batch_size,), generator=self._gen, device=self.device)
        r = self.buffer[idx]
        return r[:, 0:10], r[:, 10:11], r[:, 11], r[:, 12:22], r[:, 22]


class SAC:
    """``agent = SAC(num_inputs, action_space, args)``; ``agent.update_parameters(memory, batch_size, updates)``."""

    FIELDS = ("policy", "q1", "q2", "q1_target", "q2_target", "log_alpha", "alpha", "grad", "adam_m", "adam_v", "idx",
              "eps_next", "eps_cur", "losses", "updates", "obs_scale")

    def __init__(self, num_inputs: int = 10, action_space=None, args=None, device="cuda", batch_size: Optional[int] = None,
                 use_graph: bool = True, **kw):
        def arg(name, default):
            return kw.get(name, getattr(args, name, default) if args is not None else default)
        if num_inputs != _lib.N_OBS:
            raise ValueError("the AST observation has 10 entries (MSRL_env_ex.py:67-78)")
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise _lib.ShipSimError("the SAC learner runs on a CUDA device only (no CPU fallback)")
        self.L = _lib.lib()
        cfg = _lib.SitSacConfig()
        cfg.hidden = int(arg("hidden_size", 256))
        cfg.batch_size = int(batch_size if batch_size is not None else arg("batch_size", 64))
        cfg.gamma, cfg.tau, cfg.lr, cfg.alpha = float(arg("gamma", 0.99)), float(arg("tau", 0.005)), float(arg("lr", 3e-4)), float(arg("alpha", 0.2))
        cfg.automatic_entropy_tuning = int(bool(arg("automatic_entropy_tuning", True)))
        cfg.target_update_interval = int(arg("target_update_interval", 1))
        cfg.seed = int(arg("seed", 25450))
        # input scale of all networks: default 1 / observation_space.high (MSRL_env_ex.py:67-78); None = raw observations
        from .rollout import default_obs_scale
        obs_scale = kw.get("obs_scale", "space")
        if isinstance(obs_scale, str):
            obs_scale = default_obs_scale()
        for c in range(_lib.N_OBS):
            cfg.obs_scale[c] = 1.0 if obs_scale is None else float(obs_scale[c])
        tcv = arg("tensor_cores", True)
        cfg.tensor_cores = 2 if (tcv == 2 and tcv is not True) else int(bool(tcv))   # 2: also below batch 1024
        self.cfg, self.hidden, self.batch_size, self.use_graph = cfg, cfg.hidden, cfg.batch_size, bool(use_graph)
        self.h = C.c_void_p()
        idx = self.device.index if self.device.index is not None else torch.cuda.current_device()
        _lib.check(self.L.sit_sac_create(C.byref(cfg), idx, C.byref(self.h)))
        nbytes = C.c_size_t()
        _lib.check(self.L.sit_sac_workspace_bytes(self.h, C.byref(nbytes)))
        self.ws = torch.zeros(nbytes.value // 4, dtype=torch.float32, device=self.device)
        self.f: Dict[str, torch.Tensor] = {}
        for name in self.FIELDS:
            off, cnt = C.c_int64(), C.c_int64()
            _lib.check(self.L.sit_sac_field(self.h, name.encode(), C.byref(off), C.byref(cnt)))
            self.f[name] = self.ws[off.value // 4: off.value // 4 + cnt.value]
        self.f["idx"] = self.f["idx"].view(torch.int32)
        self.f["updates"] = self.f["updates"].view(torch.int64)
        self.losses = torch.zeros(8, dtype=torch.float32, device=self.device)
        # kernels one update enqueues: 28 products + 9 element-wise = 37.  On the tensor-core path (batch >= 1024, hidden >= 128)
        # the 6 split-k products (weight gradients: contraction over the batch) are followed by a reduce kernel each (43); with
        # hidden a multiple of 128 a network's forward and its input-gradient chain are one kernel each (5 forwards, 3
        # backwards instead of 15 + 7 products, the policy's tanh-Gaussian head in its forward kernel) and one launch packs the
        # bf16 images of W1: 28 (22 below batch 1024, where no product is split)
        self.tensor_cores = bool(cfg.tensor_cores) and (cfg.batch_size >= 1024 or cfg.tensor_cores >= 2) and cfg.hidden >= 128
        fused = self.tensor_cores and cfg.hidden % 128 == 0 and os.environ.get("SIT_SAC_FUSED_FWD", "1") != "0"
        splitk = 6 if (self.tensor_cores and cfg.batch_size >= 1024) else 0          # reduce kernels of the split-k weight gradients
        # batch 256 .. 8192: the weight gradients of layers 0 and 1 come from bf16 operand images (product + reduce each, also
        # below batch 1024 where the fp32 products have no reduce)
        dw_img = fused and 256 <= cfg.batch_size <= 8192 and os.environ.get("SIT_SAC_DW_IMG", "1") != "0"
        self.kernels_per_update = (22 if fused else 37) + splitk + (4 if (dw_img and cfg.batch_size < 1024) else 0)
        self._side = None
        self.launches = 0
        _lib.check(self.L.sit_sac_init(self.h, self.ws.data_ptr(), torch.cuda.current_stream(self.device).cuda_stream))

    # ---- network views (augmented layers [K+1][N]: weight rows then the bias row) ------------------------------------
    def layers(self, net: str):
        H, kin, nout = self.hidden, (10 if net == "policy" else 11), (2 if net == "policy" else 1)
        p, o, out = self.f[net], 0, []
        for K, N in ((kin, H), (H, H), (H, nout)):
            blk = p[o:o + (K + 1) * N].view(K + 1, N)
            out.append((blk[:K], blk[K]))
            o += (K + 1) * N
        return out

    def policy_struct(self, mode: str = "gaussian", tensor_cores: bool = False) -> "_lib.SitPolicy":
        """SitPolicy pointing INTO the workspace: the rollout kernels read the weights the learner updates."""
        p = _lib.SitPolicy()
        _lib.check(self.L.sit_sac_policy(self.h, self.ws.data_ptr(), C.byref(p)))
        p.mode = {"gaussian": _lib.POLICY_GAUSSIAN, "mean": _lib.POLICY_MEAN}[mode]
        if tensor_cores:
            p.mode |= _lib.POLICY_TC_BF16
        return p

    def policy(self, mode: str = "gaussian", tensor_cores: bool = False):
        agent = self

        class _LivePolicy:
            hidden = agent.hidden

            def __init__(self):
                self.mode, self.tensor_cores = mode, tensor_cores

            def struct(self):
                return agent.policy_struct(self.mode, self.tensor_cores)
        return _LivePolicy()

    @property
    def alpha(self) -> float:
        return float(self.f["alpha"].item())

    @property
    def updates(self) -> int:
        return int(self.f["updates"][0].item())

    # ---- the update ------------------------------------------------------------------------------------------------
    @nvtx_range("sit.sac.update")
    def update(self, memory: ReplayMemory, n_updates: int = 1):
        """Enqueue ``n_updates`` x update_parameters; returns the device tensor that will hold the losses of the last one."""
        if n_updates <= 0:
            return self.losses
        cur = torch.cuda.current_stream(self.device)
        if self.use_graph:
            if self._side is None:
                self._side = torch.cuda.Stream(device=self.device)
            self._side.wait_stream(cur)
            st = self._side.cuda_stream
        else:
            st = cur.cuda_stream
        _lib.check(self.L.sit_sac_update(self.h, self.ws.data_ptr(), memory.buffer.data_ptr(), int(memory.capacity),
                                         memory.count.data_ptr(), int(n_updates), self.losses.data_ptr(), int(self.use_graph), st))
        if self.use_graph:
            cur.wait_stream(self._side)
        self.launches += self.kernels_per_update * n_updates
        return self.losses

    def update_parameters(self, memory: ReplayMemory, batch_size: int, updates: int):
        """main_ast.py:354.  ``updates`` is the caller's running count; the learner keeps its own on the device (they agree
        when every update goes through this object)."""
        if batch_size != self.batch_size:
            raise ValueError(f"this learner was built for batch_size {self.batch_size}")
        ls = self.update(memory, 1).tolist()
        return ls[0], ls[1], ls[2], ls[3], ls[4]

    def close(self):
        if getattr(self, "h", None) is not None and self.h:
            self.L.sit_sac_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
