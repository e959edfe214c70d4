"""World-size-2 `gloo` tests (CPU) of the only collectives of the framework: the post-segment gather of compacted
replay rows and the reduction of the episode counters (sac_maritime_ast_b200/rollout.py)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from sac_maritime_ast_b200.rollout import gather_replay, reduce_stats, shard_range
    n_total = 11
    lo, hi = shard_range(n_total, rank, world)
    # rank r holds a different number of rows; row content encodes (global env id, index)
    n = (hi - lo) * (rank + 1)
    rows = torch.zeros((n, 24), dtype=torch.float32)
    for j in range(n):
        rows[j, 0], rows[j, 1] = lo + j % (hi - lo), j
        rows[j, 23] = torch.tensor(lo + j % (hi - lo), dtype=torch.int32).view(torch.float32)
    allrows = gather_replay(rows)
    stats = torch.arange(16, dtype=torch.int64) * (rank + 1)
    tot = reduce_stats(stats)
    empty = gather_replay(torch.zeros((0, 24)) if rank == 0 else rows[:2])   # ragged incl. an empty rank
    # fixed-capacity exchange (ReplayExchange): no count exchange, every block carries its own count in a header row.
    # A stand-in sampler: a ring, its counters, its capacity.  Two segments; rank 1's second segment is empty.
    from types import SimpleNamespace
    from sac_maritime_ast_b200 import _lib
    from sac_maritime_ast_b200.rollout import ReplayExchange
    from sac_maritime_ast_b200.sac import ReplayMemory
    ring = torch.zeros((16, 24)); counters = torch.zeros(_lib.RC_N, dtype=torch.int64)
    fake = SimpleNamespace(replay=ring, counters=counters, capacity=16)
    xch = ReplayExchange(fake, block_rows=6)
    mem = ReplayMemory(64, device="cpu")
    sent = []
    for seg, n_new in enumerate([(3, 5), (4, 0)]):
        k = n_new[rank]
        start = int(counters[_lib.RC_CLOSED])
        for j in range(k):
            slot = (start + j) % 16
            ring[slot, 0], ring[slot, 1] = 100 * rank + start + j, seg
            ring[slot, 23] = torch.tensor(rank, dtype=torch.int32).view(torch.float32)
        counters[_lib.RC_CLOSED] += k; counters[_lib.RC_TRANSITIONS] += k
        xch.pack(); xch.exchange(mem)
    q.put((rank, allrows.clone(), tot.clone(), empty.shape[0], n, mem.buffer[:len(mem)].clone(), int(xch.dropped[0])))
    dist.barrier()
    dist.destroy_process_group()


def _run_world(world):
    port = _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    try:
        res = sorted([q.get(timeout=300) for _ in range(world)], key=lambda x: x[0])
        for p in procs:
            p.join(timeout=120)
        return res if all(p.exitcode == 0 for p in procs) else None
    except Exception:
        return None
    finally:
        for p in procs:
            if p.is_alive():
                p.terminate()


def test_gather_replay_and_reduce_stats_world2():
    world = 2
    res = _run_world(world) or _run_world(world)     # one retry on a fresh port: the probe-then-bind of the port can race
    assert res is not None, "gloo world-size-2 run failed twice"
    n_rows = [r[4] for r in res]
    for rank, allrows, tot, n_empty, _, xrows, dropped in res:
        # fixed-capacity path: segment 0 = rank 0's 3 rows then rank 1's 5, segment 1 = rank 0's 4 rows; same on every rank
        assert dropped == 0 and xrows.shape[0] == 12 and torch.equal(xrows, res[0][5])
        assert xrows[:, 0].tolist() == [0, 1, 2, 100, 101, 102, 103, 104, 3, 4, 5, 6]
        assert xrows[:, 1].tolist() == [0] * 8 + [1] * 4
        assert allrows.shape == (sum(n_rows), 24)
        assert torch.equal(allrows, res[0][1])                       # same on every rank
        assert torch.equal(tot, torch.arange(16, dtype=torch.int64) * 3)
        assert n_empty == 2
    # rank order preserved: first block is rank 0's rows
    a = res[0][1]
    assert torch.equal(a[:n_rows[0], 1], torch.arange(n_rows[0], dtype=torch.float32))
    ids = a[:, 23].contiguous().view(torch.int32)
    assert int(ids.min()) == 0 and int(ids.max()) == 10
