"""Golden fixtures of the model variants and snapshot data, frozen from the LIVE reference classes
(imported from /root/reference behind the shims of oracle/ref_harness.py).  Build-container only.

    python tests/golden/make_golden_variants.py

  variant_simplified.npz   SimplifiedMachineryModel (ship_engine.py:398-433) + ThrottleFromSpeedSetPointSimplifiedPropulsion
                           (controllers.py:154-172) driving the hull of ShipModelAST (three_dof_kinematics,
                           three_dof_kinetics(thrust_force=..., rudder_angle=...), ship_model.py:233-250,576-606) under the
                           LOS auto-pilot.  The reference has no ship class that joins these (its ShipModel* classes build a
                           ShipMachineryModel), so the loop below joins them in the order of ShipModel.update_differentials
                           (ship_model.py:484-490): every number comes out of the reference's own methods.
  variant_circles.npz      StaticObstacle (obstacle.py:8-90): obstacles_distance / if_ship_inside_obstacles on random, centre,
                           on-circle and near-circle points.
  variant_snapshots.npz    ship_draw = True in the real MultiShipRLEnv (MSRL_env_ex.py:416-421): the steps at which
                           ship_snap_shot() fired and the outline rows it appended (ship_model.py:326-338, ShipDraw
                           utils.py:56-101) for both ships, default scenario, 400 steps, timer started at 30 and at 0.
"""
from __future__ import annotations

import os
import sys
from types import SimpleNamespace

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import ref_harness as rh  # noqa: E402
import scenarios as sc  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))

SIMPLIFIED = dict(route=[(0.0, 0.0), (2500.0, 1500.0), (3000.0, 5000.0), (6000.0, 7000.0)], psi0=0.4, u_sp=7.0, nsteps=2400,
                  modes=("MEC", "PTO"), mode=0, mode_schedule=[(1200, 1)], thrust_time_constant=30.0, kp=0.25, ki=0.012,
                  initial_thrust_force=150e3)
CIRCLES = np.array([[2000.0, 3000.0, 500.0], [6000.0, 6500.0, 1200.0], [6100.0, 5000.0, 300.0], [9000.0, 1000.0, 50.0]])


def make_simplified():
    s = SIMPLIFIED
    m = rh.ref_modules()
    cfg = rh.default_config_dict()
    base = rh.build_ship(s["route"], s["psi0"], s["u_sp"], mode_names=s["modes"], mode_index=s["mode"])
    hull, ap = base.ship_model, base.auto_pilot
    se, ct = m.ship_engine, m.controllers
    full = hull.ship_machinery_model
    mc = se.SimplifiedPropulsionMachinerySystemConfiguration(
        hotel_load=cfg["machinery"]["hotel_load"], machinery_modes=full.machinery_modes, machinery_operating_mode=s["mode"],
        specific_fuel_consumption_coefficients_me=se.FuelConsumptionCoefficients(*cfg["sfc_me"]),
        specific_fuel_consumption_coefficients_dg=se.FuelConsumptionCoefficients(*cfg["sfc_dg"]),
        thrust_force_dynamic_time_constant=s["thrust_time_constant"],
        rudder_angle_to_sway_force_coefficient=cfg["machinery"]["rudder_angle_to_sway_force_coefficient"],
        rudder_angle_to_yaw_force_coefficient=cfg["machinery"]["rudder_angle_to_yaw_force_coefficient"],
        max_rudder_angle_degrees=cfg["machinery"]["max_rudder_angle_degrees"])
    mach = se.SimplifiedMachineryModel(machinery_config=mc, time_step=cfg["dt"], initial_thrust_force=s["initial_thrust_force"])
    hull.ship_machinery_model = mach          # the hull's rudder() reads c_rudder_v / c_rudder_r from here (ship_model.py:608-622)
    thr_ctl = ct.ThrottleFromSpeedSetPointSimplifiedPropulsion(kp=s["kp"], ki=s["ki"], time_step=cfg["dt"])
    sched = dict(s["mode_schedule"])
    n = s["nsteps"]
    state = np.zeros((n, 7)); cmd = np.zeros((n, 2)); k = np.zeros(n, np.int8); fuel = np.zeros((n, 3)); ei = np.zeros(n)
    load = np.zeros((n, 4))
    for t in range(n):
        if t in sched:
            mach.mode_selector(sched[t])
        rud = ap.rudder_angle_from_sampled_route(hull.north, hull.east, hull.yaw_angle)
        thr = thr_ctl.throttle(speed_set_point=s["u_sp"], measured_speed=hull.forward_speed)
        # what store_simulation_data does with the machinery (ship_model.py:645-684): load split and fuel accumulators
        pct_me, pct_el = mach.load_perc(thr)
        ld = mach.mode.distribute_load(load_perc=thr, hotel_load=mach.hotel_load)
        rate_me, rate_el, c_me, c_el, c_tot = mach.fuel_consumption(thr)
        load[t] = (ld.load_on_main_engine, ld.load_on_electrical, pct_me, pct_el)
        # ShipModel.update_differentials order (ship_model.py:484-490) with the simplified machinery
        hull.three_dof_kinematics()
        mach.update_thrust_force(thr)
        hull.three_dof_kinetics(thrust_force=mach.thrust, rudder_angle=rud)
        # BaseShipModel.integrate_differentials (ship_model.py:314-324) + the machinery's own Euler step
        hull.north = hull.int.integrate(x=hull.north, dx=hull.d_north)
        hull.east = hull.int.integrate(x=hull.east, dx=hull.d_east)
        hull.yaw_angle = hull.int.integrate(x=hull.yaw_angle, dx=hull.d_yaw)
        hull.forward_speed = hull.int.integrate(x=hull.forward_speed, dx=hull.d_forward_speed)
        hull.sideways_speed = hull.int.integrate(x=hull.sideways_speed, dx=hull.d_sideways_speed)
        hull.yaw_rate = hull.int.integrate(x=hull.yaw_rate, dx=hull.d_yaw_rate)
        mach.integrate_differentials()
        hull.int.next_time()
        state[t] = (hull.north, hull.east, hull.yaw_angle, hull.forward_speed, hull.sideways_speed, hull.yaw_rate, mach.thrust)
        cmd[t] = (rud, thr)
        k[t] = ap.next_wpt
        fuel[t] = (c_me, c_el, c_tot)
        ei[t] = thr_ctl.ship_speed_controller.error_i
    np.savez_compressed(os.path.join(OUT, "variant_simplified.npz"), state=state, cmd=cmd, k=k, fuel=fuel, ei_speed=ei, load=load)
    print("variant_simplified: final", state[-1], "k", k[-1], "thr range", cmd[:, 1].min(), cmd[:, 1].max())


def make_circles():
    m = rh.ref_modules()
    obs = m.obstacle.StaticObstacle(CIRCLES.copy())
    rng = np.random.default_rng(7)
    pts = [rng.uniform(0, 10000, size=(3000, 2))]
    for (cn, ce, r) in CIRCLES:                       # centres, points on the circle along the axes, and just in / out
        pts.append(np.array([[cn, ce], [cn + r, ce], [cn - r, ce], [cn, ce + r], [cn, ce - r]]))
        ang = rng.uniform(0, 2 * np.pi, 64)
        for f in (1.0, 1.0 - 1e-9, 1.0 + 1e-9, 0.5, 2.0):
            pts.append(np.stack([cn + f * r * np.cos(ang), ce + f * r * np.sin(ang)], 1))
    pts = np.concatenate(pts)
    dist = np.array([obs.obstacles_distance(n, e) for n, e in pts])
    inside = np.array([bool(obs.if_ship_inside_obstacles(n, e)) for n, e in pts])
    route_inside = np.array([bool(obs.if_route_inside_obstacles(n, e)) for n, e in pts])
    assert np.array_equal(inside, route_inside)
    np.savez_compressed(os.path.join(OUT, "variant_circles.npz"), circles=CIRCLES, points=pts, dist=dist, inside=inside)
    print("variant_circles:", len(pts), "points,", int(inside.sum()), "inside")


def make_snapshots():
    m = rh.ref_modules()
    s = sc.ENV_SCENARIOS["default"]
    out = {}
    for tag, t0 in (("t30", 30.0), ("t0", 0.0)):
        env = rh.build_env(s["test_route"], s["obs_route"], s["test_psi0"], s["obs_psi0"], s["test_u_sp"], s["obs_u_sp"])
        env.ship_draw, env.time_since_last_ship_drawing = True, t0
        from oracle.ref_harness import ref_modules
        for ship in (env.test.ship_model, env.obs.ship_model):
            ship.draw = m.utils.ShipDraw()
        env.reset(); env.init_step()
        nsteps = 400
        steps = []
        for t in range(nsteps):
            a_n, a_e, sac = sc.action_at(s["actions"], t)
            before = len(env.test.ship_model.ship_drawings[0])
            env.step((a_n, a_e), bool(sac), t == 0)
            if len(env.test.ship_model.ship_drawings[0]) > before:
                steps.append(t)
        out[f"{tag}_steps"] = np.array(steps, dtype=np.int32)
        for name, ship in (("test", env.test.ship_model), ("obs", env.obs.ship_model)):
            out[f"{tag}_{name}_x"] = np.array(ship.ship_drawings[0])
            out[f"{tag}_{name}_y"] = np.array(ship.ship_drawings[1])
        print("variant_snapshots", tag, "snap steps", steps[:4], "...", len(steps))
    np.savez_compressed(os.path.join(OUT, "variant_snapshots.npz"), **out)


if __name__ == "__main__":
    make_simplified()
    make_circles()
    make_snapshots()
