#!/usr/bin/env python
"""Benchmark of the hot path: ship env-steps/s (BASELINE.json metric) on N B200s of one node.

    python bench.py --gpus N --steps K --warmup W            # our arm (CUDA kernels via libshipsim.so)
    python bench.py --impl reference --gpus N --steps K ...  # CPU arm: the oracle port on all host cores

Workload (BASELINE config 3): per GPU 1,048,576 fp32 ships, random routes (2..8 waypoints),
per-ship set points, current / wind disturbances re-drawn every launch; one bench "step" = ONE launch of
the fused ship-step kernel advancing every ship K_INNER = 256 simulator steps.  Ships are sharded over the
GPUs with no data-path collective (weak scaling: the per-GPU fleet is fixed); NCCL only gathers the
rollout statistics after the timed region.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

SEED = 25450            # main_ast.py:61
F_SHIP = 250.0          # algorithmic FLOP per ship-step
TRAJ_BYTES_F64 = 64.0   # algorithmic bytes written per ship-step in trajectory mode: 8 channels x 8 B
METRIC = "ship env-steps/sec"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--ships-per-gpu", type=int, default=1 << 20)
    ap.add_argument("--k-inner", type=int, default=256)
    ap.add_argument("--max-wp", type=int, default=8)
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak (default): --ships-per-gpu ships on every GPU; strong: --ships-per-gpu ships IN TOTAL, sharded")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the e2e / HBM trajectory / peak legs (profiling runs)")
    return ap.parse_args()


# ---- synthetic workload (shared by both arms: generated with numpy so the CPU arm needs no GPU) ---------------
def make_routes(n, max_wp, rng):
    """Random routes: n_wp ~ U{2..max_wp}, start U([500, 9500]^2), segment length U(800, 3000) m,
    turn U(-60, 60) deg.  float32-representable values (the fp32 kernel and the fp64 oracle see the same)."""
    n_wp = rng.integers(2, max_wp + 1, n).astype(np.int32)
    wp_n = np.zeros((max_wp, n), np.float32); wp_e = np.zeros((max_wp, n), np.float32)
    wp_n[0] = rng.uniform(500, 9500, n); wp_e[0] = rng.uniform(500, 9500, n)
    heading = rng.uniform(-np.pi, np.pi, n)
    psi0 = heading.copy()
    for j in range(1, max_wp):
        seg = rng.uniform(800, 3000, n)
        wp_n[j] = wp_n[j - 1] + seg * np.cos(heading); wp_e[j] = wp_e[j - 1] + seg * np.sin(heading)
        heading = heading + np.deg2rad(rng.uniform(-60, 60, n))
    for j in range(2, max_wp):      # unused slots repeat the last waypoint (never read: n_wp bounds the index)
        unused = n_wp <= j
        wp_n[j, unused] = wp_n[j - 1, unused]; wp_e[j, unused] = wp_e[j - 1, unused]
    # set points above the ~9.1 m/s the default ship can reach with a 2.8 m/s following current: the reference's
    # unsaturated throttle PI (controllers.py:138-143) then never unwinds below zero, which would drive the shaft
    # through the omega = -0.1 singularity of ship_engine.py:374 (the float64 reference blows up there too)
    u_sp = rng.uniform(9.5, 12.0, n).astype(np.float32)
    return n_wp, wp_n, wp_e, psi0.astype(np.float32), u_sp


def make_dist(n, rng):
    return np.stack([rng.uniform(-2, 2, n), rng.uniform(-2, 2, n), rng.uniform(0, 15, n),
                     rng.uniform(-np.pi, np.pi, n)]).astype(np.float32)


# ---- CPU arm: the oracle port (plain C float64, OpenMP) on the host cores ----------------------------------------
def cpu_ship_steps_per_s(n, k_inner, steps, warmup, max_wp, seed, threads=0):
    """Times orc_run_ships on a bounded sample (n ships) of the same workload; returns (ship-steps/s, cores)."""
    import helpers as H
    from oracle import cport
    L = cport.lib()
    rng = np.random.default_rng(seed)
    n_wp, wp_n, wp_e, psi0, u_sp = make_routes(n, max_wp, rng)
    cfg = H.default_cfg()
    p = cport.make_params(cfg)
    ships = (cport.OrcShip * n)()
    proto = cport.OrcShip()
    cport.init_ship(proto, cfg, [(0, 0), (1, 1)], 0.0, 8.0)
    arr = np.ctypeslib.as_array(ships)
    arr[:] = np.ctypeslib.as_array((cport.OrcShip * 1)(proto))[0]
    arr["north"], arr["east"], arr["psi"], arr["u_sp"] = wp_n[0], wp_e[0], psi0, u_sp
    arr["wp_n"][:, :max_wp] = wp_n.T; arr["wp_e"][:, :max_wp] = wp_e.T
    arr["n_wp"] = n_wp
    if threads <= 0:
        # all host cores this process may run on (torchrun exports OMP_NUM_THREADS=1, which would otherwise make the
        # OpenMP port single-threaded)
        threads = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    cores = threads
    times = []
    for it in range(warmup + steps):
        d = np.ascontiguousarray(make_dist(n, rng).astype(np.float64))
        arr["vc_n"], arr["vc_e"], arr["wind_speed"], arr["wind_dir"] = d[0], d[1], d[2], d[3]
        t0 = time.perf_counter()
        L.orc_run_ships(C.byref(p), ships, n, k_inner, 0, None, None, threads)
        t1 = time.perf_counter()
        if it >= warmup:
            times.append(t1 - t0)
    total = float(np.sum(times))
    return n * k_inner * steps / total, cores, total


def calibrate_cpu_sample(k_inner, max_wp, per_step_s):
    """Pick a ship count so that one CPU bench step (k_inner simulator steps of every ship) takes about per_step_s."""
    rate, cores, _ = cpu_ship_steps_per_s(8192, k_inner, 1, 1, max_wp, 1)
    n = int(rate * per_step_s / k_inner)
    return max(1024, min(1 << 20, (n // 1024) * 1024))


def python_reference_rate(args, steps, warmup, ships_per_proc=4):
    """The reference's OWN Python simulator (oracle/_ref staged by oracle/stage_ref.py, or /root/reference) on all host
    cores: P processes x ships_per_proc independent ships of the benchmark workload.  None when the files are absent."""
    from oracle import ref_timing as rt
    if not rt.available():
        return None
    r = rt.time_reference(procs=0, ships_per_proc=ships_per_proc, k_inner=args.k_inner, steps=steps, warmup=warmup,
                          max_wp=args.max_wp, seed=SEED)
    r["sample"] = (f"{r['cores']} processes x {ships_per_proc} ships x {args.k_inner} steps per bench step x {steps} steps "
                   f"(bounded sample of the {args.ships_per_gpu}-ship fleet), the reference's own Python simulator "
                   f"(ShipModelAST + controllers, MSRL_env_ex.py:346-398 order), Python {r['python']}, numpy {r['numpy']}, "
                   f"{r['cpu_model']}")
    return r


def run_reference(args, rank, world):
    """CPU arm.  Times the reference's own Python simulator when its files are present (staged under oracle/_ref by
    __graft_entry__.build(); kind "reference"), else the float64 C port of the same step (kind "port").  The port's rate
    is reported beside it either way: it is ~10^3 x faster per core than the Python and is the conservative baseline."""
    if rank != 0:
        return
    wu = min(args.warmup, 2)
    pr = python_reference_rate(args, args.steps, max(1, wu))
    n = calibrate_cpu_sample(args.k_inner, args.max_wp, per_step_s=(20.0 if pr else 60.0) / (args.steps + wu))
    rate, cores, total = cpu_ship_steps_per_s(n, args.k_inner, args.steps, wu, args.max_wp, SEED)
    port = {"value": rate, "unit": "ship-steps/s", "cores": cores, "kind": "port", "ms_per_step": 1e3 * total / args.steps,
            "sample": f"{n} ships x {args.k_inner} steps per bench step (bounded sample of the {args.ships_per_gpu}-ship "
                      f"fleet), float64 C port of the reference step (oracle/ship_oracle.c), OpenMP"}
    if pr:
        value, ms = pr["value"], 1e3 * pr["wall_s"] / args.steps
        base = {"value": value, "unit": "ship-steps/s", "cores": pr["cores"], "kind": "reference", "sample": pr["sample"],
                "state_finite": pr["finite"], "port": port}
        note = ("reference arm = the reference's own Python simulator on every host core (one process per core: it has no "
                "batching or threading of its own); `cpu_baseline.port` is the float64 C restatement of the same step on the "
                "same cores")
    else:
        value, ms, base = rate, port["ms_per_step"], port
        note = ("oracle/_ref is absent (the reference's Python files were not staged): this arm times the plain-C float64 "
                "restatement (oracle/ship_oracle.c), ~10^3 x faster per core than the reference's own Python")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "ship-steps/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args), "cpu_baseline": base,
        "e2e": {"value": value, "unit": "ship-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "note": note,
    }
    emit(line)


def shard_range(n, rank, world):
    """Contiguous shard [lo, hi) of n units for `rank` of `world` (the partition of SURVEY 8e)."""
    per = (n + world - 1) // world
    return min(n, rank * per), min(n, (rank + 1) * per)


def workload_config(args):
    return {"workload": "BASELINE config 3: batched ship envs fp32, random routes (2..%d waypoints), current/wind "
                        "disturbances re-drawn every %d steps, sharded over GPUs" % (args.max_wp, args.k_inner),
            "ships_per_gpu": args.ships_per_gpu, "k_inner_steps_per_launch": args.k_inner, "max_wp": args.max_wp,
            "sharding": "ships split contiguously over ranks, no step-path collective",
            "l2": "state (~330 MB per GPU at 1M ships) and disturbance ring larger than the 126 MB L2"}


# ---- clocks during the timed region (B200_PROFILING.md recipe) -------------------------------------------------
class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), line.strip()))

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self, t0, t1):
        sm, mx, reasons = [], 0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ts, line in self.rows:
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                clk, cmax = float(parts[0]), float(parts[1])
            except ValueError:
                continue
            mx = max(mx, cmax)
            if t0 <= ts <= t1:
                sm.append(clk)
                for nm, v in zip(names, parts[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm)}


def run_b200(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    import sac_maritime_ast_b200 as sit
    from sac_maritime_ast_b200 import _lib
    from sac_maritime_ast_b200 import rollout as ro_mod

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device and no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # keep stdout to the one JSON line
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    args.warmup = max(args.warmup, 3)                      # timing rule: at least 3 untimed warm-up steps
    KI = args.k_inner
    POSE_CH = ("north", "east", "psi", "u")                # what the reference's caller reads between steps (MSRL_env_ex.py:255-278)

    def fleet_leg(S, seed, e2e=True):
        """Device-resident timed region + (optionally) the same through the host-buffer entry point for a fleet of S ships
        on this rank.  Returns a dict with max-over-ranks times."""
        rng = np.random.default_rng(seed)
        n_wp, wp_n, wp_e, psi0, u_sp = make_routes(S, args.max_wp, rng)
        sm, tc, ap = sit.default_ship([(0, 0), (1000, 1000)], 0.0)
        fleet = sit.ShipFleet(sm, tc, ap, 8.0, n_ships=S, device=dev, precision="fp32", max_wp=args.max_wp)
        fleet.set_routes(torch.from_numpy(wp_n).to(dev), torch.from_numpy(wp_e).to(dev), torch.from_numpy(n_wp).to(dev))
        fleet.f["north"].copy_(torch.from_numpy(wp_n[0].astype(np.float64))); fleet.f["east"].copy_(torch.from_numpy(wp_e[0].astype(np.float64)))
        fleet.f["psi"].copy_(torch.from_numpy(psi0)); fleet.f["u_sp"].copy_(torch.from_numpy(u_sp))
        fleet.capture_init()
        n_ring = 8                                         # 8 x 16 MB of disturbances, resident in HBM
        ring_h = [torch.from_numpy(make_dist(S, rng)).pin_memory() for _ in range(n_ring)]
        ring_d = [t.to(dev) for t in ring_h]
        pose_h = torch.empty((len(POSE_CH), S), dtype=torch.float32).pin_memory()
        k8_h = torch.empty(S, dtype=torch.int8).pin_memory()
        snapshot = fleet.state_dict()
        out = {"S": S}
        for w in range(args.warmup):
            fleet.step(KI, dist=ring_d[w % n_ring])
        barrier()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = fleet.launches
        t_wall0 = time.perf_counter()
        ev0.record()
        for k in range(args.steps):
            fleet.step(KI, dist=ring_d[(args.warmup + k) % n_ring])
        ev1.record()
        barrier()
        out["wall_s"] = time.perf_counter() - t_wall0
        out["launches"] = fleet.launches - l0
        tmax = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        out["ms"] = float(tmax.item())
        out["finite"] = bool(torch.isfinite(fleet.f["north"]).all() and torch.isfinite(fleet.f["u"]).all())
        out["stats"] = torch.stack([fleet.f["steps"].sum().double(), fleet.f["k"].double().mean(), fleet.f["u"].double().mean()])
        if e2e:
            # end to end through the C-ABI host-buffer entry point: pinned host disturbances in, the channels the caller
            # reads (N, E, psi, u) + the waypoint index as int8 out, every step, copies inside the timed region
            fleet.load_state_dict(snapshot)
            for w in range(min(args.warmup, 3)):
                fleet.step_host(KI, ring_h[w % n_ring], pose_h, channels=POSE_CH, k8_host=k8_h)
            barrier()
            ev0.record()
            for k in range(args.steps):
                fleet.step_host(KI, ring_h[(args.warmup + k) % n_ring], pose_h, channels=POSE_CH, k8_host=k8_h)
            ev1.record()
            barrier()
            e2e_ms = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)
            out["e2e_ms"] = float(e2e_ms.item())
            out["h2d"] = int(ring_h[0].numel() * 4)
            out["d2h"] = int(pose_h.numel() * 4 + k8_h.numel())
            out["e2e_finite"] = bool(torch.isfinite(pose_h).all())
        fleet.close()
        return out

    # ---- headline: weak scaling, ships_per_gpu ships on every rank ---------------------------------------------------
    S = args.ships_per_gpu
    strong_total = args.ships_per_gpu                      # BASELINE config 3: 1 048 576 ships IN TOTAL, sharded over the GPUs
    if args.scaling == "strong":
        lo, hi = shard_range(strong_total, rank, world)
        S = hi - lo
    clocks = ClockSampler(local_rank)
    clocks.start()
    t_load0 = time.perf_counter()
    main = fleet_leg(S, SEED + 1000 * rank, e2e=not args.no_extras)   # global ship ids differ per rank: independent shards
    t_load1 = time.perf_counter()
    clocks.stop()
    total_ships = strong_total if args.scaling == "strong" else world * S
    value = total_ships * KI * args.steps / (main["ms"] * 1e-3)
    ms_max, launches, finite = main["ms"], main["launches"], main["finite"]
    cfg = workload_config(args)
    if args.scaling == "strong":
        cfg["ships_total"] = strong_total; cfg["ships_per_gpu"] = S
    line = {"metric": METRIC, "value": value, "unit": "ship-steps/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": args.scaling,
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfg,
            "gpu_launches": launches, "state_finite": finite}
    if "e2e_ms" in main:
        line["e2e"] = {"value": total_ships * KI * args.steps / (main["e2e_ms"] * 1e-3), "unit": "ship-steps/s",
                       "h2d_bytes_per_step": main["h2d"], "d2h_bytes_per_step": main["d2h"],
                       "api": "ShipFleet.step_host(channels=N,E,psi,u + int8 waypoint index) -> sit_fleet_step_host_sel "
                              "(pinned host buffers, chunk pipeline)", "pose_finite": main["e2e_finite"]}
    # sampled over warm-up + timed region + the e2e loop (the same kernel under load throughout); the timed
    # region alone can be shorter than one nvidia-smi sampling period
    line["clocks"] = clocks.summary(t_load0, t_load1)
    line["clocks"]["window_s"] = t_load1 - t_load0
    line["clocks"]["timed_region_s"] = main["wall_s"]

    # ---- strong scaling beside it: the SAME 1 048 576-ship fleet cut into N shards (BASELINE config 3 as written) ----------
    if not args.no_extras and args.scaling == "weak":
        if world == 1:
            line["strong"] = {"ships_total": S, "ships_per_gpu": S, "value": value, "ms_per_step": ms_max / args.steps,
                              "e2e": line["e2e"]["value"], "note": "N = 1: the strong-scaling point is the headline itself"}
        else:
            lo, hi = shard_range(strong_total, rank, world)
            st = fleet_leg(hi - lo, SEED + 1000 * rank + 500, e2e=True)
            line["strong"] = {"ships_total": strong_total, "ships_per_gpu": hi - lo,
                              "value": strong_total * KI * args.steps / (st["ms"] * 1e-3), "ms_per_step": st["ms"] / args.steps,
                              "e2e": strong_total * KI * args.steps / (st["e2e_ms"] * 1e-3),
                              "efficiency_vs_one_gpu_holding_the_whole_fleet": (ms_max / args.steps) / (world * st["ms"] / args.steps)
                              if S == strong_total else None,
                              "state_finite": st["finite"],
                              "note": "one GPU advances the whole fleet in the headline's ms_per_step (same run, same "
                                      "kernel); N GPUs each advance a 1/N shard in strong.ms_per_step"}

    # ---- rollout statistics gathered over NCCL (outside the step path) ------------------------------------------------
    stats = main["stats"]
    if world > 1:
        gathered = [torch.empty_like(stats) for _ in range(world)]
        dist.all_gather(gathered, stats)
        stats = torch.stack(gathered).mean(0)
    line["rollout_stats"] = {"mean_waypoint_index": float(stats[1]), "mean_speed_mps": float(stats[2])}

    if rank == 0 and not args.no_extras:
        # ---- roofline of the dominant kernel (fleet_step_kernel<float>): CUDA-core FP32 issue -----------------------
        L = _lib.lib()
        tf, pms = C.c_double(), C.c_double()
        _lib.check(L.sit_measure_fma_peak(local_rank, _lib.F32, C.byref(tf), C.byref(pms)))
        tf2 = C.c_double()
        _lib.check(L.sit_measure_fma_peak(local_rank, 2, C.byref(tf2), None))      # the same chain as packed FFMA2
        per_gpu = value / world
        ach = per_gpu * F_SHIP / 1e12
        line["roofline"] = {"bound": "fp32", "achieved": ach, "peak": tf.value, "unit": "TFLOP/s", "frac": ach / tf.value,
                            "peak_ffma2": tf2.value,
                            "traffic": ncu_traffic("fleet32") if (S == 1 << 20 and KI == 256) else None,
                            "traffic_unit": "DRAM bytes per launch (ncu --set full, profiles/r02_traffic.json)",
                            "kernel": "fleet_step_x2_kernel<AVOID=0, SINGLE_MODE=1, TRAJ=0> (packed f32x2, two ships per thread)",
                            "note": "CUDA-core FP32 pipe bound (no tensor-core or HBM roofline applies to the K-step kernel): "
                                    "achieved = ship-steps/s/GPU x 250 algorithmic FLOP; peak = FFMA chain measured live on this GPU "
                                    "(sit_measure_fma_peak; peak_ffma2 = the same chain as packed FFMA2: the same FLOP rate, "
                                    "half the issue slots); kernel time = ms_per_step (one launch per step, CUDA events)"}
        # ---- trajectory write-out mode (BASELINE config 2 style, fp64): HBM-write bound ------------------------------
        line["roofline_hbm"] = traj_roofline(sit, dev)
        line["roofline_hbm_fp32"] = traj_roofline(sit, dev, "fp32")
        tf64 = C.c_double()
        _lib.check(L.sit_measure_fma_peak(local_rank, _lib.F64, C.byref(tf64), None))
        line["fp64_fma_peak_tflops"] = tf64.value
        # ---- two-ship AST env step (BASELINE configs 4/5): env-steps/s, last-step record and every-step record ----------
        ast = env_leg(sit, dev, B=1 << 18, K=64, steps=10, warmup=3, precision="fp32", record_all=False)
        # utilisation as the counters give it (ncu --set full of this leg's kernel, profiles/r02_env_util.json): the kernel is
        # bound by instruction issue and short dependent chains, not by a FLOP or byte roofline
        try:
            ast["utilisation_ncu"] = json.load(open(os.path.join(ROOT, "profiles", "r02_env_util.json")))
        except OSError:
            ast["utilisation_ncu"] = None
        ast["notional_flop_note"] = ("algorithmic_tflops uses SURVEY 8(d)'s notional 4 800 FLOP / env-step (every map query as a full "
                                     "39-edge scan); the kernel skips most of that through its nearest-ring cache, so the figure is a "
                                     "throughput in the survey's units, not a pipe utilisation")
        ast["notional_flop_over_fma_peak"] = ast["algorithmic_tflops"] / tf.value
        line["ast_env"] = ast
        ast_rec = env_leg(sit, dev, B=1 << 18, K=64, steps=10, warmup=3, precision="fp32", record_all=True)
        line["ast_env_recorded"] = {k: ast_rec[k] for k in ("value", "unit", "ms_per_launch", "write_gbs")}
        line["ast_env_e2e"] = env_e2e_leg(sit, dev)
    if not args.no_extras:
        # ---- SAC-AST rollout (BASELINE config 4): policy fused into the device loop, samples/s.  Every rank rolls out its
        # own shard of envs (global env ids rank*B ..), then the replay rows of the timed segment are all-gathered and the
        # episode counters all-reduced over NCCL -- the only collectives of the framework, after the segment ----------------
        B_roll = 65536
        rl = rollout_leg(sit, dev, B=B_roll, env_id_offset=rank * B_roll, world=world, barrier=barrier)
        smp_keep = rl.pop("sampler")
        mem_roll = rl.pop("memory")
        tot_stats = ro_mod.reduce_stats(smp_keep.env.stats)
        agg = torch.tensor([rl["samples"], rl["env_steps"], rl["resamples"], rl["episodes"], rl["launches"], rl["rows_left_behind"]],
                           dtype=torch.float64, device=dev)
        tm = torch.tensor([rl["ms_total"], rl["ms_rollout_only"], rl["ms_exposed_exchange"]], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(agg, op=dist.ReduceOp.SUM)
            dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        allrows = mem_roll.buffer[:len(mem_roll)]
        ids = allrows[:, 23].contiguous().view(torch.int32)
        rl.update({"n_gpus": world, "envs": world * B_roll, "envs_per_gpu": B_roll, "samples": int(agg[0].item()),
                   "env_steps": int(agg[1].item()), "resamples": int(agg[2].item()), "episodes": int(agg[3].item()),
                   "launches": int(agg[4].item()), "rows_left_behind": int(agg[5].item()),
                   "lane_utilisation": float(agg[1].item()) / (world * B_roll * rl["k_per_launch"] * rl["iterations"]),
                   "fields": "samples, env_steps, resamples, episodes, launches are sums over all ranks; times are max over ranks",
                   "ms_total": float(tm[0].item()), "ms_rollout_only": float(tm[1].item()),
                   "ms_exposed_exchange": float(tm[2].item()),
                   "gathered_rows": int(allrows.shape[0]), "gathered_bytes": int(allrows.numel() * 4),
                   "gather_complete": bool(allrows.shape[0] == int(agg[0].item()) and int(agg[5].item()) == 0
                                           and int(ids.min()) >= 0 and int(ids.max()) < world * B_roll
                                           and int(tot_stats[13].item()) > 0),
                   "value": float(agg[0].item()) / (float(tm[0].item()) * 1e-3),
                   "value_without_gather": float(agg[0].item()) / (float(tm[1].item()) * 1e-3),
                   "env_steps_per_s": float(agg[1].item()) / (float(tm[1].item()) * 1e-3),
                   "collective": ("NCCL all_gather_into_tensor of fixed-capacity row blocks on a side stream after every segment "
                                  "(no count exchange, no host sync; segment s+1 rolls out under the exchange of segment s) + "
                                  "one all_reduce of 16 counters") if world > 1 else "none (one rank): blocks appended on the side stream"})
        del allrows
        # ---- long horizon (BASELINE config 5): 10 000 steps, run-time machinery-mode switching, all checks on; every rank
        # runs its own envs, the termination histogram is summed over ranks with one all-reduce after the run ---------------
        barrier()
        lh = long_horizon_leg(sit, dev, rank)
        barrier()
        lh_ms = torch.tensor([lh["ms"]], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(lh_ms, op=dist.ReduceOp.MAX)
        lh_stats = ro_mod.reduce_stats(lh["stats"]).cpu().numpy()
        if rank == 0:
            line["long_horizon"] = {
                "metric": "AST env-steps/s over a 10 000-step horizon with machinery-mode switching (config 5)",
                "value": float(lh_stats[13]) / (float(lh_ms.item()) * 1e-3), "unit": "env-steps/s", "n_gpus": world,
                "envs_per_gpu": lh["envs"], "steps": lh["steps"], "launches_per_gpu": lh["launches"],
                "ms_total": float(lh_ms.item()), "mode_redraw_every": 512, "modes": ["PTO", "PTI", "MEC"],
                "env_steps_counted": int(lh_stats[13]), "episodes_done": int(lh_stats[12]),
                "status_counts": {i: int(lh_stats[i]) for i in range(12)}, "state_finite": lh["finite"],
                "collective": "one all_reduce of 16 counters after the run" if world > 1 else "none (one rank)"}
            line["rollout"] = rl
            line["learner"] = learner_leg(sit, dev, smp_keep)
    if rank == 0 and world == 1 and not args.no_cpu_baseline and not args.no_extras:
        # N = 1 only: with several ranks the others would sit in a barrier on the same host cores
        n = calibrate_cpu_sample(KI, args.max_wp, per_step_s=2.5)
        rate, cores, total = cpu_ship_steps_per_s(n, KI, 4, 1, args.max_wp, SEED)
        port = {"value": rate, "unit": "ship-steps/s", "cores": cores, "kind": "port",
                "sample": f"{n} ships x {KI} steps x 4 (oracle/ship_oracle.c float64, OpenMP, {total:.1f} s)"}
        pr = python_reference_rate(args, steps=4, warmup=1)
        if pr:
            line["cpu_baseline"] = {"value": pr["value"], "unit": "ship-steps/s", "cores": pr["cores"], "kind": "reference",
                                    "sample": pr["sample"], "port": port}
        else:
            line["cpu_baseline"] = port
    if rank == 0:
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


TRAFFIC_FILE = "r02_traffic.json" if os.path.exists(os.path.join(ROOT, "profiles", "r02_traffic.json")) else "r01_traffic.json"


def ncu_traffic(label):
    """DRAM bytes per launch of the kernel `label` from the committed `ncu --set full` capture of the same workload
    (profiles/r01_traffic.json, written by tools/ncu_traffic.py), or None."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", TRAFFIC_FILE)))[label]["dram_bytes_per_launch"]
    except Exception:
        return None


def traj_roofline(sit, dev, precision="fp64"):
    """Fleet with full trajectory write-out [K][8][S] (BASELINE config 2 layout): achieved HBM GB/s.  fp64: 64 B per
    ship-step (the config-2 kernel, co-bound by the FP64 pipe); fp32: 32 B per ship-step (the config-3 arithmetic)."""
    import torch
    f64 = precision == "fp64"
    S, K = 148 * 4096, (32 if f64 else 64)      # 148 SMs x 16 CTAs of 256 ships: whole waves at 2 or 4 CTAs per SM
    sm, tc, ap = sit.default_ship([(0, 0), (3000, 3000), (6000, 9000)], np.pi / 4)
    fl = sit.ShipFleet(sm, tc, ap, 8.0, n_ships=S, device=dev, precision=precision)
    fl.f["u_sp"].copy_(4 + 5 * torch.rand(S, device=dev, dtype=torch.float64 if f64 else torch.float32))
    out = {}
    for _ in range(3):
        fl.step(K, traj=True, out=out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    e0.record()
    for _ in range(reps):
        fl.step(K, traj=True, out=out)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    bytes_per_step = TRAJ_BYTES_F64 if f64 else TRAJ_BYTES_F64 // 2
    gbs = bytes_per_step * S * K / (ms * 1e-3) / 1e9
    peak = 6551.0
    src = "fallback"
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]; src = "MEASURED_PEAKS.json"
    except Exception:
        pass
    return {"bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak,
            "traffic": ncu_traffic("traj64" if f64 else "traj32"), "algorithmic_bytes_per_launch": bytes_per_step * S * K,
            "kernel": ("fleet_step_fast_kernel<double, TRAJ>" if f64 else "fleet_step_x2_kernel<TRAJ> (packed f32x2)") + " (every-step trajectory write-out)", "ships": S, "k": K,
            "ms_per_launch": ms, "ship_steps_per_s": S * K / (ms * 1e-3), "peak_source": src,
            "note": f"algorithmic bytes = {bytes_per_step} B written per ship-step (8 {precision} channels); output ~1.2 GB per launch > L2"}


DEFAULT_MAP = [
    [(0, 10000), (5500, 10000), (5300, 9000), (4800, 8500), (4200, 7300), (4000, 5700), (4300, 4900),
     (4900, 4400), (4400, 4000), (3200, 4100), (2000, 4500), (1000, 4000), (900, 3500), (500, 2600), (0, 2350)],
    [(10000, 0), (4000, 0), (4250, 250), (5000, 400), (6000, 900), (8000, 1100), (8500, 1500), (9000, 2250),
     (9500, 3500), (10000, 4000)],
    [(5500, 5500), (5700, 7000), (6200, 8100), (7500, 8000), (7800, 7000), (7600, 5500), (6900, 4700), (6000, 5000)],
    [(2000, 2000), (2500, 2300), (4000, 2500), (5000, 3000), (4200, 2100), (3400, 1900)],
]   # main_ast.py:200-205
F_ENV = 4800.0   # algorithmic FLOP per AST env-step on the default 39-edge map


def make_bench_env(sit, dev, B, precision, seed, modes=("PTO", "PTI", "MEC")):
    """BASELINE config 4/5 style two-ship envs on the default map: per-env set points and machinery modes so
    that lanes diverge like a real rollout batch does."""
    import torch
    from types import SimpleNamespace
    kw = dict(mode_names=modes, mode_index=1)
    t = sit.default_ship([(500, 500), (9500, 9500)], np.pi / 4, **kw)
    o = sit.default_ship([(9500, 9000), (1000, 1000)], -3 * np.pi / 4, **kw)
    assets = [sit.ShipAssets(*t, 8.0, [], [], "test_ship", False), sit.ShipAssets(*o, 8.0, [], [], "obs_ship", False)]
    env = sit.MultiShipRLEnv(assets=assets, map=sit.PolygonObstacle(DEFAULT_MAP), ship_draw=False,
                             time_since_last_ship_drawing=30, args=SimpleNamespace(sampling_frequency=7, theta=2),
                             n_envs=B, device=dev, precision=precision)
    g = torch.Generator(device=dev); g.manual_seed(seed)
    # obstacle-ship set points above the ~9.1 m/s the hull reaches with the following current (see make_routes): the
    # reference's unsaturated throttle PI then never unwinds through the omega = -0.1 singularity
    env.ship[0].u_sp.copy_(6.0 + 4.0 * torch.rand(B, generator=g, device=dev, dtype=env.real))
    env.ship[1].u_sp.copy_(9.5 + 2.5 * torch.rand(B, generator=g, device=dev, dtype=env.real))
    # the test ship stays in PTI (the reference default, main_ast.py:159-161): with the always-on avoidance branch its
    # throttle sits at 1.1, which is an immediate blackout failure in PTO / MEC; the obstacle ship mixes PTI / MEC
    env.ship[1].mode.copy_(torch.randint(1, 3, (B,), generator=g, device=dev, dtype=torch.int32))
    env.capture_init()
    env.reset(); env.init_step()
    return env, g


def draw_waypoints(env, g, dev):
    """Intermediate waypoints (north, east) for every env: a point on the obstacle ship's A->B line plus a lateral
    offset (the role of agent.convert_action_to_simu_input, main_ast.py:367; its code is absent from the reference)."""
    import torch
    B = env.n_envs
    s = torch.rand(B, generator=g, device=dev, dtype=env.real)
    lat = 1600.0 * (torch.rand(B, generator=g, device=dev, dtype=env.real) - 0.5)
    a = torch.empty((B, 2), dtype=env.real, device=dev)
    a[:, 0] = 9500.0 - 8500.0 * s + 0.686 * lat
    a[:, 1] = 9000.0 - 8000.0 * s - 0.728 * lat
    # validity resampling (main_ast.py:47-51 max_route_resampling): redraw the waypoints that fall inside an island or
    # off the map, checked by the product's own map kernel (sit_map_query)
    for _ in range(20):
        bad = env.map_query(a[:, 0], a[:, 1]) != 0
        nb = int(bad.sum())
        if nb == 0:
            break
        s2 = torch.rand(nb, generator=g, device=dev, dtype=env.real)
        l2 = 1600.0 * (torch.rand(nb, generator=g, device=dev, dtype=env.real) - 0.5)
        a[bad, 0] = 9500.0 - 8500.0 * s2 + 0.686 * l2
        a[bad, 1] = 9000.0 - 8000.0 * s2 - 0.728 * l2
    return a


def env_leg(sit, dev, B=1 << 18, K=64, steps=10, warmup=3, precision="fp32", record_all=False):
    """AST env-steps/s: K MultiShipRLEnv.step per launch for B two-ship envs, in-kernel auto-reset, a fresh
    intermediate waypoint (SAC_update) per launch."""
    import torch
    env, g = make_bench_env(sit, dev, B, precision, SEED)
    acts = [draw_waypoints(env, g, dev) for _ in range(8)]
    sac = torch.ones(B, dtype=torch.uint8, device=dev)
    out = {}
    for w in range(warmup):
        env.step_k(K, acts[w % 8], sac, init=None, auto_reset="hard", record_all=record_all, out=out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    env.stats.zero_()
    e0.record()
    for k in range(steps):
        env.step_k(K, acts[(warmup + k) % 8], sac, init=None, auto_reset="hard", record_all=record_all, out=out)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    st = env.stats.cpu().numpy()
    rate = B * K / (ms * 1e-3)
    rec_bytes = (10 * 4 + 4 + 1 + 4) if precision == "fp32" else (10 * 8 + 8 + 1 + 4)
    return {"metric": "AST env-steps/s (MultiShipRLEnv.step, 2 ships + reward + termination)", "value": rate,
            "unit": "env-steps/s", "envs": B, "k_per_launch": K, "ms_per_launch": ms, "dtype": precision,
            "record_every_step": bool(record_all),
            "algorithmic_tflops": rate * F_ENV / 1e12,
            "write_gbs": (rate * rec_bytes / 1e9) if record_all else None,
            "state_finite": bool(torch.isfinite(env.f["north"]).all() and torch.isfinite(env.f["omega"]).all()),
            "auto_reset": "hard (fresh construction; the reference's reset keeps integrators, quirk Q3)",
            "episodes_done": int(st[12]), "env_steps_counted": int(st[13]),
            "status_counts": {i: int(st[i]) for i in range(12)}}


def env_e2e_leg(sit, dev, B=1 << 18, K=64, steps=10, warmup=3):
    """The env-step leg through the host-buffer entry point of the drop-in (MultiShipRLEnv.step_host -> sit_env_step_host):
    pinned host action / SAC_update in, host obs / reward / done / status of the last step out, copies inside the timed
    region.  No auto-reset on this entry point (the reference's caller resets), so episodes run into their terminal states."""
    import torch
    env, g = make_bench_env(sit, dev, B, "fp32", SEED)
    acts = [draw_waypoints(env, g, dev).t().contiguous().cpu().pin_memory() for _ in range(4)]
    sac = torch.ones(B, dtype=torch.uint8).pin_memory()
    out_h = dict(obs=torch.empty((10, B), dtype=torch.float32).pin_memory(), reward=torch.empty(B, dtype=torch.float32).pin_memory(),
                 done=torch.empty(B, dtype=torch.uint8).pin_memory(), status=torch.empty(B, dtype=torch.int32).pin_memory())
    for w in range(warmup):
        env.step_host(K, acts[w % 4], sac, w == 0, out_h)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for k in range(steps):
        env.step_host(K, acts[(warmup + k) % 4], sac, False, out_h)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    return {"value": B * K / (ms * 1e-3), "unit": "env-steps/s", "envs": B, "k_per_call": K, "ms_per_call": ms,
            "h2d_bytes_per_call": int(acts[0].numel() * 4 + B), "d2h_bytes_per_call": int(B * (10 * 4 + 4 + 1 + 4)),
            "api": "MultiShipRLEnv.step_host -> sit_env_step_host (pinned host buffers, chunk pipeline)",
            "obs_finite": bool(torch.isfinite(out_h["obs"]).all())}


def long_horizon_leg(sit, dev, rank, B=65536, K=64, total_steps=10000, mode_every=512, precision="fp32"):
    """BASELINE config 5: 10 000-step horizon with every termination / collision check on, in-kernel auto-reset, and the
    obstacle ship's machinery mode re-drawn per env from {PTO, PTI, MEC} every `mode_every` steps (run-time mode switch =
    mode_selector, ship_engine.py:233-237; the test ship stays in PTI, see make_bench_env).  Returns this rank's timing and
    the 16 episode counters (the caller reduces them over ranks)."""
    import torch
    env, g = make_bench_env(sit, dev, B, precision, SEED + 17 * rank)
    acts = [draw_waypoints(env, g, dev) for _ in range(8)]
    sac = torch.ones(B, dtype=torch.uint8, device=dev)
    n_draws = (total_steps + mode_every - 1) // mode_every
    modes = torch.randint(0, 3, (n_draws, B), generator=g, device=dev, dtype=torch.int32)
    out = {}
    for w in range(3):
        env.step_k(K, acts[w % 8], sac, init=None, auto_reset="hard", record_all=False, out=out)
    torch.cuda.synchronize()
    env.stats.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches, done_steps = 0, 0
    e0.record()
    while done_steps < total_steps:
        if done_steps % mode_every == 0:
            env.ship[1].mode.copy_(modes[done_steps // mode_every])
        k = min(K, total_steps - done_steps, mode_every - done_steps % mode_every)
        env.step_k(k, acts[launches % 8], sac, init=None, auto_reset="hard", record_all=False, out=out)
        launches += 1; done_steps += k
    e1.record()
    torch.cuda.synchronize()
    return {"ms": e0.elapsed_time(e1), "launches": launches, "stats": env.stats.clone(), "envs": B, "steps": total_steps,
            "finite": bool(torch.isfinite(env.f["north"]).all() and torch.isfinite(env.f["omega"]).all())}


def rollout_leg(sit, dev, B=65536, K=80, iters=40, warm_iters=20, hidden=256, precision="fp32", mode="gaussian",
                env_id_offset=0, tc=1, world=1, barrier=None, segments=4):
    """BASELINE config 4: SAC-AST rollout, policy MLP 10 -> 256 -> 256 -> 2 with fixed seeded weights evaluated on the
    decision-event subset between step launches, transitions compacted into a device replay ring; after every segment of
    iters / segments iterations the segment's rows go to every rank's replay memory (rollout.ReplayExchange: fixed-capacity
    blocks, side stream).  The timed region is the whole of it: `segments` x (collect, pack, exchange) and the wait for the
    last exchange.  Returns samples/s (pushed transitions) and env-steps/s actually executed (pending envs do not step).
    Episodes follow the reference (`while not done`, num_steps_episode = 600 only masks the row closing at that step)."""
    import torch
    from sac_maritime_ast_b200 import rollout as ro
    from sac_maritime_ast_b200 import sac as sac_mod
    env, g = make_bench_env(sit, dev, B, precision, SEED)
    if mode == "uniform":
        pol = ro.UniformPolicy()
    else:
        # the network sees obs / observation_space.high (the raw observation is metres / rpm / kW up to 2e4)
        pol = ro.GaussianPolicy(hidden_size=hidden, device=dev, seed=SEED, mode=mode, tensor_cores=bool(tc), obs_scale="space")
    smp = ro.IntermediateWaypointSampler(env, pol, capacity=max(B * 4, 1 << 16), seed=SEED, env_id_offset=env_id_offset,
                                         num_steps_episode=600, max_route_resampling=1000, auto_reset="hard")
    per_seg = iters // segments
    smp.begin()
    for _ in range(max(1, warm_iters // per_seg)):       # in the timed call shape (the collect graph is keyed by its arguments)
        smp.collect(per_seg, K)
    warm_iters = max(1, warm_iters // per_seg) * per_seg
    torch.cuda.synchronize()
    # block size of the exchange: the row rate seen in the warm-up with head-room (a host-known bound; every rank uses the
    # same number; rows that do not fit would be counted in rows_left_behind and fail gather_complete)
    rows_per_iter = torch.tensor([float(smp.counters[1].item()) / warm_iters], dtype=torch.float64, device=dev)
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(rows_per_iter, op=dist.ReduceOp.MAX)
    # (factor 2: with 1.25 the first timed segment left 60 248 rows behind -- decisions come in bursts as the obstacle ships
    # of a synchronised batch reach their waypoints together)
    block = int(min(per_seg * B, 2.0 * float(rows_per_iter.item()) * per_seg + 4096))
    block = (block + 255) // 256 * 256
    mem = sac_mod.ReplayMemory(world * block * (segments + 2), seed=SEED, device=dev)
    xch = ro.ReplayExchange(smp, block)
    xch.pack(); xch.exchange(mem); xch.wait()            # untimed: the warm-up rows, NCCL channel / allocator set-up at this size
    torch.cuda.synchronize()
    mem.count.zero_()

    def timed(exchange):
        if barrier:
            barrier()
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        e0.record()
        for _ in range(segments):
            smp.collect(per_seg, K)
            if exchange:
                xch.pack(); xch.exchange(mem)
        e1.record()
        if exchange:
            xch.wait()
        e2.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e2), e1.elapsed_time(e2)
    c0, s0, l0 = smp.counters.clone(), env.stats.clone(), smp.launches
    ms, ms_tail = timed(True)
    dc, ds = (smp.counters - c0).cpu().numpy(), (env.stats - s0).cpu().numpy()
    launches = smp.launches - l0
    left_behind = int(xch.dropped[0].item())
    c1 = smp.counters.clone()
    ms_plain, _ = timed(False)                             # the same number of iterations without the exchange
    dc1 = (smp.counters - c1).cpu().numpy()
    rows = smp.replay[:min(smp.n_transitions(), smp.capacity)]
    iters = per_seg * segments
    return {"metric": "SAC-AST rollout samples/s", "value": float(dc[4]) / (ms * 1e-3), "unit": "samples/s",
            "env_steps_per_s": float(ds[13]) / (ms * 1e-3), "envs": B, "k_per_launch": K, "iterations": iters,
            "segments": segments, "exchange_block_rows": block,
            "ms_per_iteration": ms / iters, "ms_total": ms, "ms_exposed_exchange": ms_tail,
            "ms_rollout_only": ms_plain * (float(dc[4]) / max(float(dc1[4]), 1.0)),   # scaled to the same number of samples
            "policy": f"{mode} MLP 10-{hidden}-{hidden}-2" if mode != "uniform" else "uniform",
            "policy_hidden_layer": "tcgen05 bf16 x bf16 -> fp32 (TMEM)" if (tc and mode != "uniform") else "fp32 CUDA cores",
            "policy_input": "obs / observation_space.high (SitPolicy.obs_scale)",
            "episodes_end": "terminal states only (reference loop); num_steps_episode = 600 masks the row closing at that step",
            "dtype": precision, "samples": int(dc[4]), "env_steps": int(ds[13]), "resamples": int(dc[2]),
            "episodes": int(ds[12]), "lane_utilisation": float(ds[13]) / (B * K * iters),
            "launches": launches, "bytes_per_sample": 96, "rows_left_behind": left_behind,
            "replay_finite": bool(torch.isfinite(rows[:, :23]).all()), "sampler": smp, "memory": mem}


def learner_leg(sit, dev, sampler, batch=64, hidden=256, n=400, warm=50):
    """SAC update_parameters/s (main_ast.py:354) on the replay ring the rollout leg filled: twin-Q SAC, policy and critics
    10/11 -> 256 -> 256, one CUDA-graph replay per update.  Batch 64 (main_ast.py:63-67): fp32 CUDA-core products (latency-bound
    chain of 37 small kernels).  Batch 4096: hidden-layer products and every contraction over the batch on tcgen05 (bf16
    operands, fp32 accumulate in TMEM), split-k partials reduced in a fixed order; the fp32-only path is timed beside it."""
    import torch
    from sac_maritime_ast_b200 import sac
    mem = sac.ReplayMemory.from_sampler(sampler)
    out = {}

    def run(b, tc):
        agent = sac.SAC(10, None, None, device=dev, hidden_size=hidden, batch_size=b, seed=SEED, tensor_cores=tc)
        agent.update(mem, warm)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ls = agent.update(mem, n)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        flop = 3 * 2 * b * (2 * (11 * hidden + hidden * hidden + 2 * hidden) + 4 * (12 * hidden + hidden * hidden + hidden))
        res = {"updates_per_s": n / (ms * 1e-3), "us_per_update": ms * 1e3 / n, "samples_per_s": n * b / (ms * 1e-3),
               "approx_tflops": flop * n / (ms * 1e-3) / 1e12, "kernels_per_update": agent.kernels_per_update,
               "products": "tcgen05 bf16 x bf16 -> fp32 (TMEM): fused forward / input-gradient kernels per network, split-k weight gradients"
                           if ((tc and b >= 1024) or tc == 2) else "fp32 CUDA cores",
               "losses_finite": bool(torch.isfinite(ls).all()), "alpha": float(ls[4])}
        agent.close()
        return res
    out[f"batch_{batch}"] = run(batch, True)
    out[f"batch_{batch}_tc"] = run(batch, 2)          # opt-in: the tensor-core path at the reference's batch size
    out["batch_1024"] = run(1024, True)
    out["batch_4096"] = run(4096, True)
    out["batch_4096_fp32_only"] = run(4096, False)
    out["replay_rows"] = len(mem)
    out["metric"] = "SAC update_parameters/s (twin-Q, auto-alpha, Adam, soft target update), CUDA graph"
    return out


_STDOUT_FD = None


def guard_stdout():
    """The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version line to stdout whatever
    NCCL_DEBUG_FILE says on some builds): everything this process and its libraries print goes to stderr, the JSON line
    is written to the saved descriptor."""
    global _STDOUT_FD
    if _STDOUT_FD is None:
        sys.stdout.flush()
        _STDOUT_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    sys.stdout.flush()
    if _STDOUT_FD is None:
        os.write(1, data)
    else:
        os.write(_STDOUT_FD, data)


def main():
    guard_stdout()
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    run_b200(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
