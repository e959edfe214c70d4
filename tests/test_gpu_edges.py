"""Edge cases and error behaviour of the C ABI on a GPU: ragged sizes (1 ship, a partial warp, a partial CTA), maximum
route capacity, end-of-route hold, wrong-kind handles, invalid K / step counts -- the ABI returns negative codes with a
message, never throws, never falls back."""
import ctypes as C

import numpy as np
import pytest
import torch

import helpers as H
import scenarios as sc

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n", [1, 33, 257])
def test_ragged_fleet_sizes_match_a_full_fleet(n):
    """A fleet of n ships (not a multiple of the warp / CTA size) gives every ship the trajectory the same ship has in
    a larger fleet."""
    import gpu_helpers as G
    s = sc.SHIP_SCENARIOS["short_segments"]
    big, small = G.make_fleet(s, n_ships=512, precision="fp64"), G.make_fleet(s, n_ships=n, precision="fp64")
    usp = 4 + 5 * torch.rand(512, device="cuda", dtype=torch.float64, generator=torch.Generator("cuda").manual_seed(9))
    big.f["u_sp"].copy_(usp); small.f["u_sp"].copy_(usp[:n])
    a = big.step(300, traj=True)["traj"]
    b = small.step(300, traj=True)["traj"]
    assert torch.equal(a[:, :, :n], b)
    assert (small.f["steps"] == 300).all()


def test_maximum_route_capacity_and_end_of_route_hold():
    """16-waypoint route (SIT_MAX_WP) against the oracle; after the last waypoint the index holds (LOS_guidance.py:99-103)."""
    import gpu_helpers as G
    import sac_maritime_ast_b200 as sit
    from sac_maritime_ast_b200 import _lib
    rng = np.random.default_rng(5)
    pts, ang, p = [(0.0, 0.0)], 0.3, np.zeros(2)
    for _ in range(_lib.MAX_WP - 1):
        ang += rng.uniform(-0.3, 0.3)
        p = p + 450.0 * np.array([np.cos(ang), np.sin(ang)])
        pts.append((float(p[0]), float(p[1])))
    s = dict(route=pts, psi0=0.3, u_sp=9.0, nsteps=3200)
    out = G.run_fleet_scenario(s, n_ships=2, precision="fp64", log27=False)
    ref = H.run_oracle_ship(s)
    assert np.array_equal(out["traj"][:, 7, 0].astype(np.int8), ref["k"])
    assert ref["k"].max() == _lib.MAX_WP - 1, "the route must be sailed to its last waypoint"
    err = H.rel_err(out["traj"][:, :7, 0], ref["state"][:, :7], np.array([1.0, 1.0, 1.0, 1.0, 1e-3, 1e-3, 1.0]))
    assert err.max() <= 1e-9
    sm, tc, ap = sit.default_ship(pts + [(9e3, 9e3)], 0.3)
    with pytest.raises((ValueError, sit.ShipSimError)):
        sit.ShipFleet(sm, tc, ap, 9.0, n_ships=2, max_wp=_lib.MAX_WP).set_route(pts + [(9e3, 9e3)])


def test_abi_error_codes_on_device():
    import gpu_helpers as G
    from sac_maritime_ast_b200 import _lib
    L = _lib.lib()
    fleet = G.make_fleet(sc.SHIP_SCENARIOS["default"], n_ships=8, precision="fp32")
    env = G.make_env(sc.ENV_SCENARIOS["default"], n_envs=4, precision="fp32")
    io = _lib.SitFleetIO()
    assert L.sit_fleet_step_k(fleet.h, fleet.state_ptr(), 0, C.byref(io), None) == -1 and b"K" in L.sit_last_error()
    assert L.sit_fleet_step_k(None, fleet.state_ptr(), 1, C.byref(io), None) == -1
    assert L.sit_fleet_step_k(env.h, env.state_ptr(), 1, C.byref(io), None) < 0 and b"FLEET" in L.sit_last_error()
    assert L.sit_env_step_k(fleet.h, fleet.state_ptr(), 1, None, None) < 0 and b"ENV" in L.sit_last_error()
    d = torch.zeros((3, 4, 8), dtype=torch.float32, device="cuda")
    io.dist, io.dist_steps = d.data_ptr(), 3
    assert L.sit_fleet_step_k(fleet.h, fleet.state_ptr(), 5, C.byref(io), None) == -1 and b"dist_steps" in L.sit_last_error()
    eio = _lib.SitEnvIO()
    eio.act_steps, eio.rec_steps = 3, 1
    assert L.sit_env_step_k(env.h, env.state_ptr(), 5, C.byref(eio), None) == -1 and b"act_steps" in L.sit_last_error()
    g = C.c_int32(0)
    assert L.sit_rollout_group_count(fleet.h, C.byref(g)) < 0
    assert L.sit_rollout_group_count(env.h, C.byref(g)) == 0 and g.value >= 1
    # nothing above launched anything: the state is untouched and a valid call still works
    assert (fleet.f["steps"] == 0).all()
    fleet.step(2)
    assert (fleet.f["steps"] == 2).all()


def test_single_env_is_call_compatible_with_the_reference():
    """B = 1: reset() -> [1, 10] state, step(...) -> 4-tuple whose .item() values are the reference's scalars."""
    import gpu_helpers as G
    import sac_maritime_ast_b200 as sit
    s = sc.ENV_SCENARIOS["default"]
    g = H.golden("env_default")
    env = G.make_env(s, n_envs=1, precision="fp64")
    st0 = env.reset()
    assert tuple(st0.shape) == (1, 10)
    assert np.allclose(st0[0].cpu().numpy(), g["reset_state"], rtol=1e-6)
    env.init_step()
    a_n, a_e, sac = sc.action_at(s["actions"], 0)
    ns, r, d, st = env.step((a_n, a_e), sac, True)
    assert H.rel_err(ns[0].cpu().numpy(), g["ep1_obs"][0], 1e-3).max() <= 1e-9
    assert abs(r.item() - g["ep1_reward"][0]) <= 1e-9 * max(1.0, abs(g["ep1_reward"][0]))
    assert bool(d.item()) == bool(g["ep1_done"][0]) and int(st.item()) == int(g["ep1_status"][0])
    assert sit.MultiShipRLEnv.decode_status(int(st.item())).startswith(" |Test ship not in terminal state|")


@pytest.mark.parametrize("precision", ["fp32", "fp64"])
def test_outputs_stay_inside_their_buffers(precision):
    """Guard bands around every output of a ragged fleet launch and a ragged env launch stay untouched (compute-sanitizer
    is not available on the pool, so the bounds are checked with sentinels)."""
    import gpu_helpers as G
    real = torch.float32 if precision == "fp32" else torch.float64
    n, K, pad = 301, 37, 4096
    fl = G.make_fleet(sc.SHIP_SCENARIOS["short_segments"], n_ships=n, precision=precision)

    def banded(shape):
        numel = int(np.prod(shape))
        buf = torch.full((numel + 2 * pad,), -12345.0, dtype=real, device="cuda")
        return buf, buf[pad:pad + numel].view(shape)
    bt, traj = banded((K, 8, n)); bl, log = banded((K, 27, n)); bc, cmd = banded((K, 2, n))
    fl.step(K, traj=True, log27=True, cmd=True, out={"traj": traj, "log27": log, "cmd": cmd})
    bt2, traj2 = banded((K, 8, n))
    fl.step(K, traj=True, out={"traj": traj2})          # the specialised every-step trajectory path (running row pointer)
    torch.cuda.synchronize()
    for buf, view in ((bt, traj), (bl, log), (bc, cmd), (bt2, traj2)):
        assert (buf[:pad] == -12345.0).all() and (buf[-pad:] == -12345.0).all()
        assert (view != -12345.0).all(), "every element of the output is written"
    B = 45
    env = G.make_env(sc.ENV_SCENARIOS["default"], n_envs=B, precision=precision)
    env.reset(); env.init_step()
    bo, obs = banded((K, 10, B)); br, rew = banded((K, B)); bp, parts = banded((K, 8, B))
    env.step_k(K, (8300.0, 8600.0), True, init=True, record_all=True, reward_parts=True,
               out={"obs": obs, "reward": rew, "reward_parts": parts})
    torch.cuda.synchronize()
    for buf, view in ((bo, obs), (br, rew), (bp, parts)):
        assert (buf[:pad] == -12345.0).all() and (buf[-pad:] == -12345.0).all()
        assert (view != -12345.0).all()


def test_nvtx_ranges_do_not_change_results():
    """SIT_NVTX=1 wraps the wrapper's entry points in NVTX ranges (tracing): a fresh process with the
    ranges on gives the same fleet state as this process without them."""
    import os
    import subprocess
    import sys
    import gpu_helpers as G
    import scenarios as sc
    code = ("import sys; sys.path[:0] = [%r, %r]\n"
            "import gpu_helpers as G, scenarios as sc, sac_maritime_ast_b200.core as c\n"
            "assert c.NVTX\n"
            "f = G.make_fleet(sc.SHIP_SCENARIOS['default'], n_ships=8, precision='fp64'); f.step(50)\n"
            "print(repr(float(f.f['north'][3])), repr(float(f.f['psi'][3])))\n") % (
                os.path.dirname(os.path.dirname(os.path.abspath(__file__))), os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, SIT_NVTX="1"), capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-800:]
    f = G.make_fleet(sc.SHIP_SCENARIOS["default"], n_ships=8, precision="fp64"); f.step(50)
    assert out.stdout.split() == [repr(float(f.f["north"][3])), repr(float(f.f["psi"][3]))]
