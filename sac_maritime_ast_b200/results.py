"""Columnar simulation log: the reference's ``ship_model.simulation_results`` -- a dict of 27 Python
lists appended once per ``store_simulation_data`` call (ship_model.py:645-684) and turned into a DataFrame at the end of a
run (``pd.DataFrame().from_dict(ship_model.simulation_results)``, main_ast.py:634) -- as a device buffer
``[steps][27][n_ships]`` filled by the K1 kernel's log channel (``sit_fleet_step_k`` with ``log27``), with exporters that
rebuild the reference's dict / DataFrame for any ship of the fleet.
"""
from __future__ import annotations

from typing import Dict, List, Optional

import numpy as np
import torch

# key order of store_simulation_data, ship_model.py:645-684 (== channel order of SIT_N_LOG)
LOG_KEYS = (
    'time [s]', 'north position [m]', 'east position [m]', 'yaw angle [deg]', 'rudder angle [deg]',
    'forward speed [m/s]', 'sideways speed [m/s]', 'yaw rate [deg/sec]', 'propeller shaft speed [rpm]',
    'commanded load fraction me [-]', 'commanded load fraction hsg [-]', 'power me [kw]',
    'available power me [kw]', 'power electrical [kw]', 'available power electrical [kw]', 'power [kw]',
    'propulsion power [kw]', 'fuel rate me [kg/s]', 'fuel rate hsg [kg/s]', 'fuel rate [kg/s]',
    'fuel consumption me [kg]', 'fuel consumption hsg [kg]', 'fuel consumption [kg]',
    'motor torque [Nm]', 'thrust force [kN]', 'cross track error [m]', 'heading error [deg]')


class SimulationResults:
    """Recorder around a ``ShipFleet``: ``step(K)`` advances the fleet K steps in one launch with the 27-channel log
    on and keeps the chunk on the device; nothing is copied to the host until an exporter is called."""

    def __init__(self, fleet, ships: Optional[torch.Tensor] = None):
        """ships: optional index tensor -- keep only these ships' columns of every chunk (a full log is
        27 x 8 B x n_ships per step: 216 MB per step at 1 M fp64 ships)."""
        self.fleet = fleet
        self.ships = None if ships is None else torch.as_tensor(ships, dtype=torch.long, device=fleet.device)
        self.chunks: List[torch.Tensor] = []
        self._scratch: Dict[str, torch.Tensor] = {}

    def step(self, K: int = 1, **kw):
        out = self.fleet.step(K, log27=True, out=self._scratch, **kw)
        lg = out["log27"]
        self.chunks.append(lg.clone() if self.ships is None else lg.index_select(2, self.ships))
        return out

    def __len__(self) -> int:
        return sum(c.shape[0] for c in self.chunks)

    def columnar(self) -> torch.Tensor:
        """[steps, 27, n_kept_ships] device tensor (channel-major: one channel of one step is contiguous)."""
        if not self.chunks:
            n = self.fleet.n_ships if self.ships is None else self.ships.numel()
            return torch.empty((0, len(LOG_KEYS), n), dtype=self.fleet.real, device=self.fleet.device)
        if len(self.chunks) > 1:
            self.chunks = [torch.cat(self.chunks, dim=0)]
        return self.chunks[0]

    def to_dict(self, ship: int = 0) -> Dict[str, List[float]]:
        """The reference's ``simulation_results`` of one (kept) ship: 27 keys -> list of floats, one per stored step."""
        col = self.columnar()[:, :, ship].double().cpu().numpy()
        return {k: col[:, c].tolist() for c, k in enumerate(LOG_KEYS)}

    def to_dataframe(self, ship: int = 0):
        import pandas as pd
        return pd.DataFrame().from_dict(self.to_dict(ship))      # main_ast.py:634

    def save_npz(self, path: str):
        """Whole log as one compressed .npz: ``log`` [steps, 27, ships], ``keys``, ``ships`` (kept indices)."""
        ships = np.arange(self.fleet.n_ships) if self.ships is None else self.ships.cpu().numpy()
        np.savez_compressed(path, log=self.columnar().cpu().numpy(), keys=np.array(LOG_KEYS), ships=ships)

    def reset(self):
        self.chunks = []


# (ship, key) order of MultiShipRLEnv.reward_results, MSRL_env_ex.py:133-141 (== channel order of SIT_N_REWARD_PARTS)
REWARD_KEYS = (("test_ship", "reward_e_ct"), ("test_ship", "reward_near_col"), ("test_ship", "total_non_terminal"),
               ("obs_ship", "reward_base"), ("obs_ship", "reward_e_ct"), ("obs_ship", "reward_near_col"),
               ("obs_ship", "total_non_terminal"), ("shared", "total_non_terminal"))


class RewardResults:
    """Recorder around a ``MultiShipRLEnv``: the reference's ``reward_results`` -- cumulative sums of the non-terminal reward
    terms, one entry appended per ``step`` (reward_function, MSRL_env_ex.py:924-964), lists re-created by ``reset()``
    (through reward_function_params, :174) -- as a device buffer of per-step increments ``[steps][8][n_envs]`` written by the
    env-step kernel; the running sums are formed when an exporter is called.  ``reset()`` here resets the env and clears the
    record, like the reference's ``env.reset()`` does."""

    def __init__(self, env, envs: Optional[torch.Tensor] = None):
        self.env = env
        self.envs = None if envs is None else torch.as_tensor(envs, dtype=torch.long, device=env.device)
        self.chunks: List[torch.Tensor] = []

    def step_k(self, K, converted_action, SAC_update=False, **kw):
        out = self.env.step_k(K, converted_action, SAC_update, record_all=True, reward_parts=True, **kw)
        rp = out["reward_parts"]
        self.chunks.append(rp.clone() if self.envs is None else rp.index_select(2, self.envs))
        return out

    def reset(self, **kw):
        self.chunks = []
        return self.env.reset(**kw)

    def __len__(self) -> int:
        return sum(c.shape[0] for c in self.chunks)

    def increments(self) -> torch.Tensor:
        """[steps, 8, n_kept_envs] device tensor of per-step increments."""
        if not self.chunks:
            n = self.env.n_envs if self.envs is None else self.envs.numel()
            return torch.empty((0, len(REWARD_KEYS), n), dtype=self.env.real, device=self.env.device)
        if len(self.chunks) > 1:
            self.chunks = [torch.cat(self.chunks, dim=0)]
        return self.chunks[0]

    def to_dict(self, env: int = 0) -> Dict[str, Dict[str, List[float]]]:
        """The reference's nested ``reward_results`` dict of one (kept) env: running sums in float64, added in step order
        exactly like ``prev + r`` (numpy's cumsum is sequential)."""
        inc = self.increments()[:, :, env].double().cpu().numpy()
        cum = np.cumsum(inc, axis=0)
        res: Dict[str, Dict[str, List[float]]] = {"test_ship": {}, "obs_ship": {}, "shared": {}}
        for c, (ship, key) in enumerate(REWARD_KEYS):
            res[ship][key] = cum[:, c].tolist()
        return res
