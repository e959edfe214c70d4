"""Time the REFERENCE'S OWN Python simulator on the host cores (BASELINE INFRASTRUCTURE: bench.py's cpu_baseline leg and
``--impl reference`` arm only).

What runs is the unmodified reference code -- ``ShipModelAST`` + ``EngineThrottleFromSpeedSetPoint`` +
``HeadingBySampledRouteController`` driven in the order of MSRL_env_ex.py:346-398 (ref_harness.bare_ship_step) -- from
/root/reference in the build container or from the copy staged under oracle/_ref/ on the GPU box (oracle/stage_ref.py).
The reference has no batching or threading of its own: P worker processes, each owning a few independent
ships of the benchmark workload (bench.make_routes / make_dist), all released together by a barrier; rate = sum of
ship-steps / (last finish - first start)."""
from __future__ import annotations

import multiprocessing as mp
import os
import sys
import time

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(_HERE)


def available() -> bool:
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    from oracle import ref_harness as rh
    return rh.available()


def _worker(rank, barrier, q, seed, ships_per_proc, k_inner, steps, warmup, max_wp):
    os.environ.setdefault("OMP_NUM_THREADS", "1")
    for p in (ROOT, os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import bench
    from oracle import ref_harness as rh
    rng = np.random.default_rng(seed + 7919 * rank)
    n_wp, wp_n, wp_e, psi0, u_sp = bench.make_routes(ships_per_proc, max_wp, rng)
    ships = []
    for j in range(ships_per_proc):
        route = [(float(wp_n[i, j]), float(wp_e[i, j])) for i in range(int(n_wp[j]))]
        ships.append(rh.build_ship(route, float(psi0[j]), float(u_sp[j])))

    def bench_step():
        d = bench.make_dist(ships_per_proc, rng).astype(np.float64)
        for j, s in enumerate(ships):
            sm = s.ship_model      # ship_model.py:154-156: plain attributes, assigned between steps
            sm.vel_c = np.array([d[0, j], d[1, j], 0.0]); sm.wind_speed = float(d[2, j]); sm.wind_dir = float(d[3, j])
            for _ in range(k_inner):
                rh.bare_ship_step(s)

    for _ in range(warmup):
        bench_step()
    barrier.wait()
    t0 = time.time()
    for _ in range(steps):
        bench_step()
    t1 = time.time()
    finite = all(np.isfinite(s.ship_model.north) and np.isfinite(s.ship_model.forward_speed) for s in ships)
    q.put((rank, t0, t1, ships_per_proc * k_inner * steps, finite))


def time_reference(procs=0, ships_per_proc=2, k_inner=256, steps=4, warmup=1, max_wp=8, seed=25450):
    """Returns dict(value [ship-steps/s], cores, wall_s, ship_steps, finite, python, numpy)."""
    if procs <= 0:
        procs = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    ctx = mp.get_context("spawn")      # the caller may hold a CUDA context: never fork it
    barrier, q = ctx.Barrier(procs), ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, barrier, q, seed, ships_per_proc, k_inner, steps, warmup, max_wp), daemon=True)
          for r in range(procs)]
    for p in ps:
        p.start()
    res = [q.get(timeout=1800) for _ in ps]
    for p in ps:
        p.join(timeout=60)
    t0, t1 = min(r[1] for r in res), max(r[2] for r in res)
    total = sum(r[3] for r in res)
    return {"value": total / (t1 - t0), "cores": procs, "wall_s": t1 - t0, "ship_steps": total,
            "finite": all(r[4] for r in res), "python": sys.version.split()[0], "numpy": np.__version__,
            "cpu_model": _cpu_model()}


def _cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


if __name__ == "__main__":
    print(time_reference())
