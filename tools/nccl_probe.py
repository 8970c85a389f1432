"""All-reduce (sum, f64) latency of the payloads a sharded LM iteration moves, as NCCL runs them on this box (tools only).
torchrun --nproc-per-node N tools/nccl_probe.py [label]"""
import os, sys, torch, torch.distributed as dist
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
label = sys.argv[1] if len(sys.argv) > 1 else "default"
sizes = {"step scalars (64 B)": 8, "linearisation message K4 (571 KB)": 1700 * 42 + 16, "CG vector K5 (624 KB)": 78000,
         "S blocks K4 (15.5 MB)": 53692 * 36, "S tile front K4 (27.5 MB)": 840 * 4096 + 10200}
for name, n in sizes.items():
    x = torch.ones(n, dtype=torch.float64, device="cuda")
    for _ in range(10):
        dist.all_reduce(x)
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 100
    e0.record()
    for _ in range(reps):
        dist.all_reduce(x)
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / reps], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if dist.get_rank() == 0:
        us = t.item() * 1e3
        print(f"[{label}] world {dist.get_world_size()}  {name:36s} {us:8.1f} us  algbw {8 * n / us / 1e3:7.1f} GB/s", flush=True)
dist.destroy_process_group()
