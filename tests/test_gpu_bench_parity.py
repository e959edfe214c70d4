"""Parity of the BENCHMARKED configuration and of every instantiation of the fast fleet kernels against the C oracle.

  * the exact workload of bench.py (bench.make_routes / bench.make_dist: random 2..8-waypoint routes, per-ship set
    points, per-ship current / wind re-drawn per 256-step launch) on a 4096-ship sample, through ShipFleet.step
    (device-resident path, last-row trajectory = the kernel the headline number times) and through step_host;
  * AVOID x SINGLE_MODE x TRAJ x {fp32, fp64}: fp32 runs the packed two-ships-per-thread kernel
    (fleet_step_x2_kernel), fp64 runs fleet_step_fast_kernel<double, ...>;
  * fp32 waypoint decisions: the measured flip rate and the |step shift| histogram are printed, every shift is at most
    one step, and flipped ships stay inside a stated pose bound instead of being dropped.
fp64: per-step state within 1e-9 relative, waypoint index exact every step.  fp32: pose and speed within 1e-4 relative
after 1024 steps (ships whose switch moved by a step: within 1e-2)."""
import numpy as np
import pytest
import torch

import bench_case as bc
import helpers as H

pytestmark = pytest.mark.gpu

K_LAUNCH = 256
TOL64, TOL32, TOL32_FLIPPED = 1e-9, 1e-4, 1e-2
POSE_FLOORS = np.array([100.0, 100.0, 1.0, 1.0])[:, None]            # N, E [m], psi [rad], u [m/s]
TRAJ_FLOORS = np.array([1.0, 1.0, 1.0, 1.0, 1e-3, 1e-3, 1.0])[None, :, None]
POSE_KEYS = ("north", "east", "psi", "u", "v", "r", "omega", "k")


def fleet_pose(fleet):
    return np.stack([fleet.f[k].double().cpu().numpy() for k in POSE_KEYS])


def switch_steps(k_traj):
    """[T, n] waypoint index per step -> dict ship -> list of the steps at which the index advanced."""
    t, ship = np.nonzero(np.diff(k_traj, axis=0) != 0)
    out = {}
    for a, b in zip(ship, t):
        out.setdefault(int(a), []).append(int(b) + 1)
    return out


def flip_report(k_got, k_ref):
    """Ships whose waypoint-switch steps differ from the oracle's, and the histogram of |step shift|."""
    sg, sr = switch_steps(k_got), switch_steps(k_ref)
    flipped, hist, n_sw = set(), {}, 0
    for ship in set(sg) | set(sr):
        a, b = sg.get(ship, []), sr.get(ship, [])
        n_sw += len(b)
        if a == b:
            continue
        flipped.add(ship)
        if len(a) != len(b):
            hist["count"] = hist.get("count", 0) + 1
            continue
        for x, y in zip(a, b):
            if x != y:
                hist[abs(x - y)] = hist.get(abs(x - y), 0) + 1
    return np.array(sorted(flipped), dtype=np.int64), hist, n_sw


@pytest.mark.parametrize("path", ["device", "host"])
def test_fp32_benchmark_workload_matches_oracle(path):
    """The headline configuration: fp32, AVOID = 0, one machinery mode, last-row trajectory, disturbances re-drawn per
    launch.  Per-launch waypoint index and final pose / speed against the float64 oracle."""
    n, launches = 4096, 4
    inp = bc.make_inputs(n, 25450, n_launches=launches)
    p, ships, arr = bc.oracle_ships(inp)
    fleet = bc.make_fleet(inp, "fp32")
    pose_h = torch.empty((8, n), dtype=torch.float32).pin_memory()
    k_mismatch = []
    for l in range(launches):
        bc.oracle_run(p, ships, arr, inp["dists"][l], K_LAUNCH)
        d = torch.from_numpy(inp["dists"][l])
        if path == "device":
            fleet.step(K_LAUNCH, dist=d.cuda())
        else:
            fleet.step_host(K_LAUNCH, d.pin_memory(), pose_h)
        k_mismatch.append(int((fleet.f["k"].cpu().numpy() != arr["k"]).sum()))
    got, ref = fleet_pose(fleet), bc.oracle_pose(arr)
    assert np.isfinite(ref).all() and np.isfinite(got).all()
    if path == "host":
        assert np.array_equal(pose_h.numpy().astype(np.float64), got.astype(np.float32).astype(np.float64)), \
            "host pose rows are the state the device holds"
    same_k = got[7] == ref[7]
    err = H.rel_err(got[:4], ref[:4], POSE_FLOORS)
    print(f"\n[bench parity fp32 {path}] ships {n}, launches {launches} x {K_LAUNCH}: waypoint-index mismatches per launch "
          f"{k_mismatch}; pose/speed rel err max {err[:, same_k].max():.3e} (index-equal ships), {err.max():.3e} (all)")
    assert err[:, same_k].max() <= TOL32
    assert err.max() <= TOL32_FLIPPED, "a ship whose switch moved by a step stays within 1e-2"
    assert max(k_mismatch) <= 0.002 * n, f"waypoint index differs at a launch boundary for {k_mismatch} of {n} ships"
    assert (fleet.f["steps"] == launches * K_LAUNCH).all()


def test_fp64_benchmark_workload_matches_oracle():
    n, launches = 4096, 4
    inp = bc.make_inputs(n, 25450, n_launches=launches)
    p, ships, arr = bc.oracle_ships(inp)
    fleet = bc.make_fleet(inp, "fp64")
    for l in range(launches):
        bc.oracle_run(p, ships, arr, inp["dists"][l], K_LAUNCH)
        fleet.step(K_LAUNCH, dist=torch.from_numpy(inp["dists"][l]).cuda())
        assert np.array_equal(fleet.f["k"].cpu().numpy(), arr["k"]), "waypoint index must be exact at every launch boundary"
    got, ref = fleet_pose(fleet), bc.oracle_pose(arr)
    err = H.rel_err(got[:7], ref[:7], TRAJ_FLOORS[0])
    assert err.max() <= TOL64, f"rel err {err.max():.3e}"


@pytest.mark.parametrize("precision", ["fp32", "fp64"])
@pytest.mark.parametrize("traj", [False, True])
@pytest.mark.parametrize("multi_mode", [False, True])
@pytest.mark.parametrize("avoid", [False, True])
def test_fast_kernel_instantiations_match_oracle(avoid, multi_mode, traj, precision):
    """AVOID x SINGLE_MODE x TRAJ x precision on bench-style random fleets (per-ship machinery mode when several modes)."""
    n, launches = 1024 + 37, 3          # odd: the packed kernel's unpaired tail ship and unaligned rows are exercised
    # several modes without the avoidance clamp: set points the hull cannot reach in MEC either, so the reference's
    # unsaturated throttle PI never unwinds through the omega = -0.1 pole (ship_engine.py:374)
    inp = bc.make_inputs(n, 77 + 2 * avoid + multi_mode, n_launches=launches, multi_mode=multi_mode,
                         u_sp_range=(18.0, 20.0) if (multi_mode and not avoid) else None)
    p, ships, arr = bc.oracle_ships(inp)
    fleet = bc.make_fleet(inp, precision)
    k_ref, k_got = [], []
    for l in range(launches):
        ref = bc.oracle_run(p, ships, arr, inp["dists"][l], K_LAUNCH, avoid=avoid, traj=traj)
        out = fleet.step(K_LAUNCH, dist=torch.from_numpy(inp["dists"][l]).cuda(), avoid=avoid, traj=traj)
        if traj:
            got = out["traj"].double().cpu().numpy()
            k_ref.append(ref[:, 7]); k_got.append(got[:, 7])
            if precision == "fp64":
                assert np.array_equal(got[:, 7], ref[:, 7]), "waypoint index must be exact at every step"
                err = H.rel_err(got[:, :7], ref[:, :7], TRAJ_FLOORS)
                assert err.max() <= TOL64, f"launch {l}: per-step rel err {err.max():.3e}"
            else:
                assert np.array_equal(got[-1], fleet_pose(fleet).astype(np.float32).astype(np.float64)), \
                    "last trajectory row == stored state"
        elif precision == "fp64":
            assert np.array_equal(fleet.f["k"].cpu().numpy(), arr["k"])
    got, ref = fleet_pose(fleet), bc.oracle_pose(arr)
    assert np.isfinite(ref).all() and np.isfinite(got).all()
    if precision == "fp64":
        assert H.rel_err(got[:7], ref[:7], TRAJ_FLOORS[0]).max() <= TOL64
        fuel = np.stack([fleet.f[k].cpu().numpy() for k in ("fuel_me", "fuel_el", "fuel")])
        fuel_ref = np.stack([arr["fuel_me"], arr["fuel_el"], arr["fuel"]])
        assert H.rel_err(fuel, fuel_ref, 1e-6).max() <= TOL64
        return
    err = H.rel_err(got[:4], ref[:4], POSE_FLOORS)
    same_k = got[7] == ref[7]
    fuel = np.stack([fleet.f[k].double().cpu().numpy() for k in ("fuel_me", "fuel_el", "fuel")])
    fuel_ref = np.stack([arr["fuel_me"], arr["fuel_el"], arr["fuel"]])
    assert H.rel_err(fuel, fuel_ref, 1e-6).max() <= 1e-4, "fuel accumulators (fp32)"
    # controller integrators: not on north_star's tolerance list (a switch that moved by a step in mid-run shifts the LOS
    # integral by one e / delta term); checked loosely so a transcription error cannot hide behind the pose
    for key in ("e_ct_int", "ei_speed", "ei_shaft"):
        e = H.rel_err(fleet.f[key].double().cpu().numpy()[same_k], arr[key][same_k], 1.0)
        assert e.max() <= 1e-2, f"{key}: {e.max():.3e}"
    if traj:
        flipped, hist, n_sw = flip_report(np.concatenate(k_got), np.concatenate(k_ref))
        print(f"\n[fp32 avoid={avoid} multi_mode={multi_mode}] {n_sw} waypoint switches, {len(flipped)} ships with a moved "
              f"switch, |step shift| histogram {hist}")
        assert set(hist) <= {1}, f"a waypoint switch moved by more than one step: {hist}"
        assert len(flipped) <= 0.01 * n
    assert err[:, same_k].max() <= TOL32, f"fp32 pose/speed rel err {err[:, same_k].max():.3e}"
    assert err.max() <= TOL32_FLIPPED, f"flipped ships: {err.max():.3e}"


def test_fp32_flip_rate_on_benchmark_routes():
    """north_star: waypoint-index decisions bit-exact.  In fp32 that cannot hold for every ship (the ship closes on the
    acceptance circle at ~4 m per step with a pose error of millimetres): measure it on the benchmark's routes --
    how many switches move, by how many steps, and how far a ship with a moved switch ends from the oracle."""
    n, launches = 4096, 4
    inp = bc.make_inputs(n, 4242, n_launches=launches)
    p, ships, arr = bc.oracle_ships(inp)
    fleet = bc.make_fleet(inp, "fp32")
    k_ref, k_got = [], []
    for l in range(launches):
        ref = bc.oracle_run(p, ships, arr, inp["dists"][l], K_LAUNCH, traj=True)
        got = fleet.step(K_LAUNCH, dist=torch.from_numpy(inp["dists"][l]).cuda(), traj=True)["traj"].double().cpu().numpy()
        k_ref.append(ref[:, 7]); k_got.append(got[:, 7])
    flipped, hist, n_sw = flip_report(np.concatenate(k_got), np.concatenate(k_ref))
    got, ref = fleet_pose(fleet), bc.oracle_pose(arr)
    err = H.rel_err(got[:4], ref[:4], POSE_FLOORS)
    ok = np.ones(n, bool); ok[flipped] = False
    rate = len(flipped) / max(n_sw, 1)
    print(f"\n[fp32 flip rate] {n} ships x {launches * K_LAUNCH} steps: {n_sw} waypoint switches, {len(flipped)} moved "
          f"({100 * rate:.3f} % of switches), |step shift| histogram {hist}; pose/speed rel err: unflipped max "
          f"{err[:, ok].max():.3e}, flipped max {err[:, ~ok].max() if len(flipped) else 0.0:.3e}")
    assert set(hist) <= {1}, f"a switch moved by more than one step: {hist}"
    assert rate <= 0.01
    assert err[:, ok].max() <= TOL32
    assert err.max() <= TOL32_FLIPPED
