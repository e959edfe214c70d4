"""Host<->device copy bandwidth with N ranks copying at once (torchrun): is the host-buffer path of bench.py bound by the
box's host memory / PCIe fabric?  Each rank: 16 MB pinned H2D + 32 MB D2H per iteration on two streams, like one bench step."""
import os, torch, torch.distributed as dist
rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
if world > 1:
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
hin, hout = torch.empty(16 << 20, dtype=torch.uint8).pin_memory(), torch.empty(32 << 20, dtype=torch.uint8).pin_memory()
din, dout = torch.empty(16 << 20, dtype=torch.uint8, device="cuda"), torch.empty(32 << 20, dtype=torch.uint8, device="cuda")
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
    print(f"ranks {world}: {ms.item():.3f} ms per (16 MB in + 32 MB out) -> {48 * 1.048576 / ms.item():.1f} GB/s per rank, {world * 48 * 1.048576 / ms.item():.1f} GB/s box")
if world > 1: dist.barrier(); dist.destroy_process_group()
