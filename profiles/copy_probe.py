import torch
a = torch.empty(1<<28, dtype=torch.int32, device="cuda"); b = torch.empty_like(a)
for _ in range(3): b.copy_(a)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); b.copy_(a); e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1); print(f"torch copy 2^28 int32: {ms:.3f} ms {8*(1<<28)/ms/1e6:.0f} GB/s")
