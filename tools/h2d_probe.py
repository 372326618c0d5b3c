import torch, time
n = 4 * 1024**3
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda")
for _ in range(2):
    d.copy_(h, non_blocking=True); torch.cuda.synchronize()
t0 = time.time()
for _ in range(4):
    d.copy_(h, non_blocking=True)
torch.cuda.synchronize()
print("H2D pinned GB/s", 4 * n / (time.time() - t0) / 1e9)
