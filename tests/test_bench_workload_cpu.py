"""Why bench.py draws the speed set points from U(9.5, 12) m/s and not from SURVEY 8(d)'s U(4, 9) -- shown on the float64 C
oracle (the pinned restatement of the reference step), so this is a property of the REFERENCE's controller, not of the kernels.

The reference's throttle cascade is unsaturated (controllers.py:138-143: both `sat` lines are commented out) and the torque
is power / (omega + 0.1) (ship_engine.py:368-375).  With a set point the PTI machinery can exceed, the speed PI winds the
commanded load negative, the shaft decelerates through omega = -0.1 rad/s, the torque term blows up and the state goes
non-finite.  At 9.5 .. 12 m/s the hull cannot reach the set point on 820 kW, the load stays on the torque cap and every ship
stays finite for the whole run -- which is what the headline configuration measures."""
import numpy as np

import bench_case as bc


def _run(u_sp_range, n=2048, launches=16):
    inp = bc.make_inputs(n, 25450, n_launches=launches, u_sp_range=u_sp_range)
    p, ships, arr = bc.oracle_ships(inp)
    omega_min = np.full(n, np.inf)
    for l in range(launches):
        bc.oracle_run(p, ships, arr, inp["dists"][l], 256)
        omega_min = np.minimum(omega_min, np.where(np.isfinite(arr["omega"]), arr["omega"], -np.inf))
    pose = bc.oracle_pose(arr)
    return np.isfinite(pose).all(axis=0), omega_min


def test_survey_set_points_drive_the_reference_controller_through_its_pole():
    finite, omega_min = _run((4.0, 9.0))
    bad = float((~finite).mean())
    print(f"\n[workload] u_sp ~ U(4, 9): {100 * bad:.1f} % of 2048 ships non-finite after 4096 steps (float64 oracle)")
    assert bad > 0.10, "the U(4, 9) workload should leave a large share of the reference's ships non-finite"
    assert (omega_min[~finite] < 0).all(), "every non-finite ship went through a negative shaft speed first"


def test_bench_set_points_keep_every_ship_finite():
    finite, omega_min = _run((9.5, 12.0))
    assert finite.all()
    assert omega_min.min() > 1.0, "the shaft never comes near the omega = -0.1 pole"
