"""Handle + caller-owned state buffer shared by ShipFleet and MultiShipRLEnv.

PyTorch is plumbing here: it owns the device memory (one uint8 buffer carved into named SoA views
according to ``sit_layout``) and supplies the CUDA stream; all simulation work is done by the
kernels in libshipsim.so.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict

import numpy as np
import torch

from . import _lib

_TORCH_DT = {_lib.DT_F64: torch.float64, _lib.DT_F32: torch.float32, _lib.DT_I32: torch.int32}


def _precision_code(precision) -> int:
    if precision in ("fp64", "f64", "float64", torch.float64, _lib.F64):
        return _lib.F64
    if precision in ("fp32", "f32", "float32", torch.float32, _lib.F32):
        return _lib.F32
    raise ValueError(f"precision must be 'fp64' or 'fp32', got {precision!r}")


def ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


# ---- tracing: NVTX ranges around the entry points of the wrapper -------------------------------------
# SIT_NVTX=1 brackets every launch-enqueueing call (ShipFleet.step / step_host, MultiShipRLEnv.step_k / step_host,
# IntermediateWaypointSampler.collect, ReplayExchange.pack / exchange, SAC.update) in an NVTX range, so a timeline (nsys, or
# ncu --nvtx --nvtx-include) shows which call a kernel belongs to.  Off by default: a range costs ~1 us of host time.
import functools
import os

NVTX = os.environ.get("SIT_NVTX", "0") not in ("", "0")


def nvtx_range(name):
    def deco(fn):
        if not NVTX:
            return fn

        @functools.wraps(fn)
        def wrapped(*a, **kw):
            torch.cuda.nvtx.range_push(name)
            try:
                return fn(*a, **kw)
            finally:
                torch.cuda.nvtx.range_pop()
        return wrapped
    return deco


class SimCore:
    """One libshipsim handle and its state buffer on one CUDA device."""

    def __init__(self, cfg: "_lib.SitConfig", n_envs: int, device):
        self.L = _lib.lib()
        if not torch.cuda.is_available():
            raise _lib.ShipSimError("no CUDA device: sac_maritime_ast_b200 has no CPU fallback")
        self.device = torch.device(device if device is not None else "cuda")
        if self.device.type != "cuda":
            raise _lib.ShipSimError("device must be a CUDA device (no CPU fallback)")
        index = self.device.index if self.device.index is not None else torch.cuda.current_device()
        self.device = torch.device("cuda", index)
        self.cfg = cfg
        self.n_envs = int(n_envs)
        self.real = torch.float64 if cfg.precision == _lib.F64 else torch.float32
        self.np_real = np.float64 if cfg.precision == _lib.F64 else np.float32
        h = C.c_void_p()
        _lib.check(self.L.sit_create(C.byref(cfg), self.n_envs, index, C.byref(h)))
        self.h = h
        self.n_ships = int(self.L.sit_n_ships(h))
        nbytes = C.c_size_t()
        _lib.check(self.L.sit_state_bytes(h, C.byref(nbytes)))
        self.state = torch.zeros(nbytes.value, dtype=torch.uint8, device=self.device)
        n = C.c_int32()
        _lib.check(self.L.sit_layout(h, None, 0, C.byref(n)))
        fields = (_lib.SitField * n.value)()
        _lib.check(self.L.sit_layout(h, fields, n.value, C.byref(n)))
        self.f: Dict[str, torch.Tensor] = {}
        for fd in fields:
            dt = _TORCH_DT[fd.dtype]
            esz = 8 if fd.dtype == _lib.DT_F64 else 4
            nb = esz * fd.rows * fd.count
            v = self.state[fd.offset:fd.offset + nb].view(dt)
            self.f[fd.name.decode()] = v.view(fd.rows, fd.count) if fd.rows > 1 else v
        self.launches = 0   # kernels of libshipsim enqueued through this handle

    # ---- helpers ---------------------------------------------------------------------------------
    def stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def state_ptr(self):
        return C.c_void_p(self.state.data_ptr())

    def capture_init(self):
        _lib.check(self.L.sit_capture_init(self.h, self.state_ptr(), self.stream()))
        self.launches += 1

    def reset(self, mask=None, hard=False):
        if mask is not None:
            mask = mask.to(device=self.device, dtype=torch.uint8).contiguous()
            assert mask.numel() == self.n_envs
        _lib.check(self.L.sit_reset(self.h, self.state_ptr(), ptr(mask), int(bool(hard)), self.stream()))
        self.launches += 1

    def close(self):
        if getattr(self, "h", None):
            self.L.sit_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- state_dict-style checkpointing of the whole simulation state ---------------------------
    def state_dict(self):
        return {"state": self.state.clone()}

    def load_state_dict(self, d):
        self.state.copy_(d["state"])
