"""Batched SAC-AST training loop on the B200 simulator: the loop of the reference's test_beds/main_ast.py:310-412
(reset -> init_step -> select_action -> convert_action_to_simu_input -> step -> push on decision -> update_parameters)
with every stage on the device and `n_envs` two-ship environments advanced together.

    python test_beds/main_ast_b200.py --n_envs 4096 --num_steps 200000
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 test_beds/main_ast_b200.py --n_envs 65536

Hyper-parameter names and defaults are the reference's (main_ast.py:36-100).  What is batched differently:
* `num_steps`, `start_steps` count simulator steps PER ENV, as in the reference's single-env loop;
* one segment = `collect_iterations` x (policy on the envs at a decision event, `k_steps` env steps), one CUDA graph launch per
  iteration (sit_rollout_collect); the learner then runs `update_per_step` x steps-per-env updates, replayed from its own graph;
* with several ranks every rank rolls out its own contiguous shard of global env ids, the new replay rows are all-gathered
  over NCCL after the segment, and every rank applies the same updates to its replica of the agent (same seed, same rows,
  same Philox batch indices -> bit-identical replicas, no gradient traffic).
Ship, controller, machinery and map parameters are main_ast.py:106-237 (`sit.default_ship`, `sit.DEFAULT_MAP`)."""
from __future__ import annotations

import argparse
import json
import os
import sys
import time
from types import SimpleNamespace

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import sac_maritime_ast_b200 as sit  # noqa: E402
from sac_maritime_ast_b200 import rollout as ro  # noqa: E402
from sac_maritime_ast_b200.sac import SAC, ReplayMemory  # noqa: E402


def parse_args(argv=None):
    p = argparse.ArgumentParser(description="Ship Transit Soft Actor-Critic Args (B200 batched)")
    # main_ast.py:36-100, same names and defaults
    p.add_argument("--policy", default="Gaussian")
    p.add_argument("--gamma", type=float, default=0.99)
    p.add_argument("--tau", type=float, default=0.005)
    p.add_argument("--theta", type=float, default=2)
    p.add_argument("--sampling_frequency", type=int, default=7)
    p.add_argument("--max_route_resampling", type=int, default=1000)
    p.add_argument("--lr", type=float, default=0.0003)
    p.add_argument("--alpha", type=float, default=0.2)
    p.add_argument("--automatic_entropy_tuning", type=lambda s: str(s).lower() not in ("0", "false"), default=True)
    p.add_argument("--seed", type=int, default=25450)
    p.add_argument("--batch_size", type=int, default=64)
    p.add_argument("--replay_size", type=int, default=1000)
    p.add_argument("--hidden_size", type=int, default=256)
    p.add_argument("--num_steps", type=int, default=100000)
    p.add_argument("--num_steps_episode", type=int, default=600)
    p.add_argument("--start_steps", type=int, default=10000)
    p.add_argument("--update_per_step", type=int, default=1)
    p.add_argument("--target_update_interval", type=int, default=1)
    # batching
    p.add_argument("--n_envs", type=int, default=4096, help="two-ship environments over all ranks")
    p.add_argument("--k_steps", type=int, default=80, help="env steps per launch")
    p.add_argument("--collect_iterations", type=int, default=8, help="(policy, k_steps) iterations per segment")
    p.add_argument("--precision", default="fp32", choices=["fp32", "fp64"])
    p.add_argument("--tensor_cores", type=int, default=1, help="1: rollout policy hidden layer and the learner from batch 1024 on tcgen05 (bf16 operands); 2: the learner at every batch size; 0: fp32 CUDA cores")
    p.add_argument("--max_updates_per_segment", type=int, default=0, help="0 = update_per_step x steps per env")
    p.add_argument("--truncate_episodes", type=int, default=0,
                   help="0 (reference, main_ast.py:322,387): num_steps_episode only sets mask = 1 at that step and the episode "
                        "runs on to a terminal state; 1: also reset the episode there")
    p.add_argument("--log_json", default="")
    return p.parse_args(argv)


def build_env(args, n_envs, device):
    test = sit.default_ship([(500, 500), (9500, 9500)], np.pi / 4)                  # main_ast.py:207-237
    obs = sit.default_ship([(9500, 9000), (1000, 1000)], -3 * np.pi / 4)
    assets = [sit.ShipAssets(*test, 8.0, [], [], "test_ship", False), sit.ShipAssets(*obs, 8.0, [], [], "obs_ship", False)]
    return sit.MultiShipRLEnv(assets=assets, map=sit.PolygonObstacle(sit.DEFAULT_MAP), ship_draw=False,
                              time_since_last_ship_drawing=30,
                              args=SimpleNamespace(sampling_frequency=args.sampling_frequency, theta=args.theta),
                              n_envs=n_envs, device=device, precision=args.precision)


def main(argv=None):
    args = parse_args(argv)
    import torch.distributed as dist
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("main_ast_b200.py needs a CUDA device (no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=dev)

    lo, hi = ro.shard_range(args.n_envs, rank, world)
    env = build_env(args, hi - lo, dev)
    agent = SAC(10, env.action_space, args, device=dev)                                # every rank: same seed -> same init
    # the reference's replay_size default (1000) is smaller than one batch of envs: keep at least 8 decisions per env
    capacity = max(args.replay_size, 8 * args.n_envs, 4 * args.batch_size)
    memory = ReplayMemory(capacity, seed=args.seed, device=dev)
    uniform, learned = ro.UniformPolicy(), agent.policy("gaussian" if args.policy == "Gaussian" else "mean",
                                                         tensor_cores=bool(args.tensor_cores))
    # a segment opens at most collect_iterations rows per env: the sampler ring holds at least two segments, so the rows of
    # a segment are packed long before they can be overwritten
    block = args.collect_iterations * (hi - lo)
    smp = ro.IntermediateWaypointSampler(env, uniform, capacity=max(2 * block, 8 * (hi - lo), 4096), seed=args.seed,
                                         env_id_offset=lo, num_steps_episode=args.num_steps_episode,
                                         max_route_resampling=args.max_route_resampling,
                                         truncate_episodes=bool(args.truncate_episodes))
    # fixed-capacity exchange (every rank has the same block size: the shards differ by at most one env)
    xch = ro.ReplayExchange(smp, args.collect_iterations * (-(-args.n_envs // world)))
    smp.begin()
    steps_per_segment = args.k_steps * args.collect_iterations
    total_numsteps, updates, segment = 0, 0, 0
    history = []
    t_start = time.time()
    while total_numsteps < args.num_steps:
        use_learned = total_numsteps >= args.start_steps
        if use_learned != (smp.policy is learned):
            smp.set_policy(learned if use_learned else uniform)                        # mode 0 / mode 1 of agent.select_action
        stats0, count0 = env.stats.clone(), smp.counters.clone()
        smp.collect(args.collect_iterations, args.k_steps)
        # ---- push: the rows of this segment (ring order) -> every rank's memory.  Fixed-capacity blocks, all-gather and
        # append on a side stream, no host sync: the learner below waits for it, the next segment's rollout does not -------
        xch.pack()
        xch.exchange(memory)
        xch.wait()                                   # the update draws from this segment's rows too (main_ast.py:350-357)
        total_numsteps += steps_per_segment
        # ---- update_parameters: update_per_step per simulator step of an env (main_ast.py:350-362) ----------------
        losses = None
        if len(memory) > args.batch_size:
            n_upd = args.update_per_step * steps_per_segment
            if args.max_updates_per_segment > 0:
                n_upd = min(n_upd, args.max_updates_per_segment)
            losses = agent.update(memory, n_upd)
            updates += n_upd
        # ---- log -------------------------------------------------------------------------------------------------
        d = ro.reduce_stats(env.stats - stats0).cpu().numpy()
        c = ro.reduce_stats(smp.counters - count0).cpu().numpy()        # transitions, decisions, re-draws, episodes ended
        segment += 1
        rec = {"segment": segment, "steps_per_env": total_numsteps, "policy": "learned" if use_learned else "uniform",
               "memory": len(memory), "updates": updates, "episodes": int(c[3]), "terminal_steps": int(d[12]),
               "env_steps": int(d[13]), "decisions": int(c[1]), "redraws": int(c[2]),
               "status_counts": [int(x) for x in d[:12]], "elapsed_s": time.time() - t_start}
        if losses is not None:
            ls = losses.tolist()
            rec.update(critic_1_loss=ls[0], critic_2_loss=ls[1], policy_loss=ls[2], ent_loss=ls[3], alpha=ls[4])
        new_rows = xch.rows()
        rec["mean_reward_of_new_rows"] = float(new_rows[:, 11].mean()) if new_rows.shape[0] else None
        rec["share_of_unsaturated_actions"] = float((new_rows[:, 10].abs() < 0.999).float().mean()) if new_rows.shape[0] else None
        rec["rows_left_behind"] = int(xch.dropped.item())
        history.append(rec)
        if rank == 0:
            msg = (f"segment {segment:4d}  steps/env {total_numsteps:7d}  {rec['policy']:8s} memory {rec['memory']:8d} "
                   f"updates {updates:7d} episodes {rec['episodes']:7d}")
            if losses is not None:
                msg += f"  q1 {rec['critic_1_loss']:.4g} q2 {rec['critic_2_loss']:.4g} pi {rec['policy_loss']:.4g} alpha {rec['alpha']:.4g}"
            print(msg, flush=True)
    torch.cuda.synchronize()
    if rank == 0 and args.log_json:
        with open(args.log_json, "w") as f:
            json.dump(history, f)
    if world > 1:
        # the replicas must have stayed bit-identical: compare a checksum of every rank's policy / critic parameters
        pol = torch.cat([agent.f[k].view(torch.int32).to(torch.int64).sum()[None] for k in ("policy", "q1", "q2")])
        allp = [torch.empty_like(pol) for _ in range(world)]
        dist.all_gather(allp, pol)
        same = all(torch.equal(a, allp[0]) for a in allp)
        if rank == 0:
            print(f"learner replicas identical over {world} ranks: {same}", flush=True)
        if not same:
            raise SystemExit("learner replicas diverged")
        dist.barrier()
    return history, agent, memory


if __name__ == "__main__":
    main()
    import torch.distributed as _dist
    if _dist.is_available() and _dist.is_initialized():
        _dist.destroy_process_group()
