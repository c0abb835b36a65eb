import torch, time
x = torch.empty(134217728 // 4, dtype=torch.float32).pin_memory()
d = torch.empty_like(x, device="cuda")
for n in range(3):
    torch.cuda.synchronize(); t = time.perf_counter(); d.copy_(x, non_blocking=True); torch.cuda.synchronize(); dt = time.perf_counter() - t
    print(f"H2D pinned 134 MB: {0.134217728 / dt:.1f} GB/s")
y = torch.empty(134217728 // 4, dtype=torch.float32)
torch.cuda.synchronize(); t = time.perf_counter(); d.copy_(y); torch.cuda.synchronize(); dt = time.perf_counter() - t
print(f"H2D pageable: {0.134217728 / dt:.1f} GB/s")
torch.cuda.synchronize(); t = time.perf_counter(); x.copy_(d); torch.cuda.synchronize(); dt = time.perf_counter() - t
print(f"D2H pinned: {0.134217728 / dt:.1f} GB/s")
