import torch
xs = []
try:
    for _ in range(42):
        xs.append(torch.full((1 << 30,), -1, dtype=torch.int32, device="cuda"))
except Exception as e:
    print("stopped at", len(xs), type(e).__name__)
torch.cuda.synchronize()
print("dirtied GiB:", 4 * len(xs))
