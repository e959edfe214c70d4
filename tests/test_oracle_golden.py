"""The C oracle against the golden fixtures frozen from the REAL reference
(tests/golden/make_golden.py).  This is what pins the oracle (CPU, no GPU needed)."""
import ctypes as C

import numpy as np
import pytest

import helpers as H
import scenarios as sc
from oracle import cport

TOL = 1e-9   # north-star: fp64 per-step state within 1e-9 relative


@pytest.mark.parametrize("name", list(sc.SHIP_SCENARIOS))
def test_ship_step_matches_reference(name):
    s = sc.SHIP_SCENARIOS[name]
    g = H.golden("ship_" + name)
    out = H.run_oracle_ship(s)
    assert np.array_equal(out["k"], g["k"]), "waypoint index must be bit-exact at every step"
    rows = g["rows"]
    err = H.rel_err(out["state"][rows], g["state"], H.STATE_FLOORS)
    assert err.max() <= TOL, f"state rel err {err.max():.3e} at {np.unravel_index(err.argmax(), err.shape)}"
    cerr = H.rel_err(out["cmd"][rows], g["cmd"], 1e-3)
    assert cerr.max() <= TOL
    lerr = H.rel_err(out["log27"][g["log_rows"]], g["log27"], 1e-6)
    assert lerr.max() <= TOL, f"27-channel log rel err {lerr.max():.3e} ch {lerr.argmax(axis=1)}"
    assert H.rel_err(out["fuel3"], g["fuel3"], 1e-6).max() <= TOL
    assert out["time"] == float(g["time"])


@pytest.mark.parametrize("name", list(sc.ENV_SCENARIOS))
def test_env_step_matches_reference(name):
    s = sc.ENV_SCENARIOS[name]
    g = H.golden("env_" + name)
    p, m, e = H.oracle_env(s)
    L = cport.lib()
    L.orc_env_reset(C.byref(e))
    state0 = np.array([e.test.north, e.test.east, e.test.psi, 0, 0, 0, e.obs.north, e.obs.east, e.obs.psi, 0])
    assert np.allclose(state0, g["reset_state"], rtol=1e-6)      # reference returns float32
    L.orc_env_init_step(C.byref(p), C.byref(p), C.byref(e))
    episodes = [("ep1", s["actions"], s["nsteps"])]
    if "second_episode" in s:
        episodes.append(("ep2", s["second_episode"]["actions"], s["second_episode"]["nsteps"]))
    for i, (tag, actions, nsteps) in enumerate(episodes):
        if i > 0:
            L.orc_env_reset(C.byref(e))
            L.orc_env_init_step(C.byref(p), C.byref(p), C.byref(e))
        out = H.run_oracle_env_episode(p, m, e, actions, nsteps)
        assert np.array_equal(out["done"], g[tag + "_done"]), "done must be bit-exact"
        assert np.array_equal(out["status"], g[tag + "_status"]), "status bits must be bit-exact"
        assert np.array_equal(out["stop"], g[tag + "_stop"])
        assert H.rel_err(out["obs"], g[tag + "_obs"], 1e-3).max() <= TOL
        assert H.rel_err(out["reward"], g[tag + "_reward"], 1.0).max() <= TOL
        assert H.rel_err(out["dist"], g[tag + "_dist"], 1.0).max() <= TOL


def test_appendix_d_known_answers():
    """values printed by the live reference at survey time."""
    out = H.run_oracle_ship(sc.SHIP_SCENARIOS["default"])
    st = out["state"]
    assert st[0, 3] == pytest.approx(0.01067042564034778, rel=1e-12)
    assert st[0, 6] == pytest.approx(42.44934212602637, rel=1e-12)
    assert st[999, 0] == pytest.approx(1230.000457601091, rel=1e-10)
    assert st[1999, 1] == pytest.approx(2843.5270585339345, rel=1e-10)
    assert st[1999, 7] == 2
    assert out["cmd"][1999, 1] == pytest.approx(1578.8145807037772, rel=1e-10)
