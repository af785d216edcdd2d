"""Host -> device copy bandwidth of pinned memory with all ranks copying at once (torchrun) or alone: the floor under the
end-to-end number of the pair-batch path (616 MB up, 20 MB down per step and GPU).  One JSON line on rank 0."""
import json, os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
try:
    from bench import bind_to_gpu_numa
    bind_to_gpu_numa(local)
except Exception:
    pass
torch.cuda.set_device(local)
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = 616_000_000
x = torch.empty(n, dtype=torch.uint8).pin_memory(); x.fill_(7)
y = torch.empty(n, dtype=torch.uint8, device="cuda")
z = torch.empty(20_000_000, dtype=torch.uint8).pin_memory()
for _ in range(2):
    y.copy_(x, non_blocking=True)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
torch.cuda.synchronize(); t = time.perf_counter()
for _ in range(5):
    y.copy_(x, non_blocking=True); z.copy_(y[:20_000_000], non_blocking=True)
torch.cuda.synchronize(); dt = (time.perf_counter() - t) / 5
v = torch.tensor([dt], dtype=torch.float64, device="cuda")
if world > 1:
    dist.all_reduce(v, op=dist.ReduceOp.MAX)
if rank == 0:
    print(json.dumps({"ranks": world, "h2d_mb": n / 1e6, "d2h_mb": 20, "ms_per_step_max_rank": v.item() * 1e3,
                      "h2d_gbs_per_rank": n / v.item() / 1e9, "h2d_gbs_aggregate": world * n / v.item() / 1e9}))
if world > 1:
    dist.destroy_process_group()
