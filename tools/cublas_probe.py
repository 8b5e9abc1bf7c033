import torch
a = torch.randn(128, 8192, device="cuda", dtype=torch.bfloat16)
b = torch.randn(8192, 8192, device="cuda", dtype=torch.bfloat16)
for _ in range(5):
    c = a @ b.t()
torch.cuda.synchronize()
