import torch, time
x = torch.empty(600_000_000, dtype=torch.uint8).pin_memory()
y = torch.empty_like(x, device="cuda")
for n in (600_000_000, 150_000_000, 40_000_000):
    for _ in range(2):
        y[:n].copy_(x[:n], non_blocking=True)
    torch.cuda.synchronize(); t = time.perf_counter()
    for _ in range(5):
        y[:n].copy_(x[:n], non_blocking=True)
    torch.cuda.synchronize(); dt = (time.perf_counter() - t) / 5
    print("H2D pinned %d MB: %.2f ms  %.1f GB/s" % (n // 1_000_000, dt * 1e3, n / dt / 1e9))
z = torch.empty(16_000_000, dtype=torch.uint8).pin_memory()
torch.cuda.synchronize(); t = time.perf_counter(); z.copy_(y[:16_000_000]); torch.cuda.synchronize(); print("D2H 16MB %.2f ms" % ((time.perf_counter() - t) * 1e3))
