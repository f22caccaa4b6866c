import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from houv_b200.chamfer import chamfer_3D
dev = torch.device("cuda"); B, N, M = 32, 2048, 2048
torch.manual_seed(0)
x1, x2 = torch.rand(B, N, 3, device=dev), torch.rand(B, M, 3, device=dev)
d1, d2 = torch.empty(B, N, device=dev), torch.empty(B, M, device=dev)
i1 = torch.empty(B, N, device=dev, dtype=torch.int32); i2 = torch.empty(B, M, device=dev, dtype=torch.int32)
for _ in range(5):
    chamfer_3D.forward(x1, x2, d1, d2, i1, i2)
torch.cuda.synchronize()
