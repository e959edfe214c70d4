"""CPU: the C oracle's SimplifiedMachineryModel branch against the fixture frozen from the live reference classes
(tests/golden/make_golden_variants.py), and -- where /root/reference exists -- the fixtures against a fresh run."""
import numpy as np

import helpers as H
import variants as V

FLOORS7 = np.array([1.0, 1.0, 1.0, 1.0, 1e-3, 1e-3, 1.0])


def test_oracle_simplified_machinery_matches_reference_golden():
    g = H.golden("variant_simplified")
    o = V.oracle_simplified()
    assert np.array_equal(o["k"], g["k"]), "waypoint index every step"
    assert H.rel_err(o["state"], g["state"], FLOORS7).max() <= 1e-11
    assert H.rel_err(o["cmd"], g["cmd"], 1e-2).max() <= 1e-9      # the heading PID differentiates: kd / dt = 180 amplifies 1e-13
    assert H.rel_err(o["fuel"], g["fuel"], 1e-6).max() <= 1e-11
    assert H.rel_err(o["ei_speed"], g["ei_speed"], 1.0).max() <= 1e-11
    # the logged machinery channels of the variant: power me / electrical [kW] and the load fractions are the base-class
    # distribute_load (ship_engine.py:46-76); thrust [kN] is the machinery state; rpm and torque do not exist (0)
    lg, ld = o["log27"], g["load"]
    assert H.rel_err(lg[:, 11], ld[:, 0] / 1000, 1e-6).max() <= 1e-11 and H.rel_err(lg[:, 13], ld[:, 1] / 1000, 1e-6).max() <= 1e-11
    assert H.rel_err(lg[:, 9], ld[:, 2], 1e-9).max() <= 1e-11 and H.rel_err(lg[:, 10], ld[:, 3], 1e-9).max() <= 1e-11
    assert (lg[:, 8] == 0).all() and (lg[:, 23] == 0).all()
    assert H.rel_err(lg[1:, 24], g["state"][:-1, 6] / 1000, 1e-6).max() <= 1e-11     # pre-step thrust of step t = post-step of t-1
    # the scenario exercises the saturation at both ends of its range and the run-time mode switch
    thr = g["cmd"][:, 1]
    assert thr.max() == 1.1 and (thr < 1.1).any() and g["k"][-1] == 3


def test_variant_fixtures_regenerate_from_the_live_reference(tmp_path):
    import pytest
    from oracle import ref_harness as rh
    if not rh.available():
        pytest.skip("reference tree not present")
    import importlib.util, os
    spec = importlib.util.spec_from_file_location("mgv", os.path.join(H.GOLDEN, "make_golden_variants.py"))
    mgv = importlib.util.module_from_spec(spec); spec.loader.exec_module(mgv)
    mgv.OUT = str(tmp_path)
    mgv.make_circles(); mgv.make_simplified()
    for name in ("variant_circles", "variant_simplified"):
        a, b = np.load(os.path.join(tmp_path, name + ".npz")), H.golden(name)
        for k in b.files:
            assert np.array_equal(a[k], b[k]), (name, k)
