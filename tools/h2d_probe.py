"""Host<->device copy bandwidth with N ranks copying at once (torchrun): is the host-buffer path of bench.py bound by the
box's host memory / PCIe fabric?  Each rank: IN bytes pinned H2D + OUT bytes D2H per iteration on two streams, like one bench
step.  Defaults = the masked host path of round 2 (16 MiB of disturbances in; N, E, psi, u + int8 waypoint index = 17 MiB
out); H2D_PROBE_IN / H2D_PROBE_OUT (bytes) override (round 1 copied 16 MiB + 32 MiB)."""
import os, torch, torch.distributed as dist
NIN, NOUT = int(os.environ.get("H2D_PROBE_IN", 16777216)), int(os.environ.get("H2D_PROBE_OUT", 17825792))
rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
if world > 1:
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
hin, hout = torch.empty(NIN, dtype=torch.uint8).pin_memory(), torch.empty(NOUT, dtype=torch.uint8).pin_memory()
din, dout = torch.empty(NIN, dtype=torch.uint8, device="cuda"), torch.empty(NOUT, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def run(n):
    for _ in range(n):
        with torch.cuda.stream(s1): din.copy_(hin, non_blocking=True)
        with torch.cuda.stream(s2): hout.copy_(dout, non_blocking=True)
run(5); torch.cuda.synchronize()
if world > 1: dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize(); e0.record(); run(100); s1.synchronize(); s2.synchronize(); e1.record(); torch.cuda.synchronize()
ms = torch.tensor([e0.elapsed_time(e1) / 100], device="cuda", dtype=torch.float64)
if world > 1: dist.all_reduce(ms, op=dist.ReduceOp.MAX)
if rank == 0:
    gb = (NIN + NOUT) / 1e6
    print(f"ranks {world}: {ms.item():.3f} ms per ({NIN / 2**20:.1f} MiB in + {NOUT / 2**20:.1f} MiB out) -> {gb / ms.item():.1f} GB/s per rank, {world * gb / ms.item():.1f} GB/s box")
if world > 1: dist.barrier(); dist.destroy_process_group()
