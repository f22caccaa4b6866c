"""Small fixed K1 workload for ncu captures: a few forward launches at
B=1024, N=M=2048 (enough CTAs for ~7 waves)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from houv_b200.chamfer import chamfer_3D
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
torch.manual_seed(0)
x1 = torch.rand(B, 2048, 3, device="cuda"); x2 = torch.rand(B, 2048, 3, device="cuda")
d1 = torch.empty(B, 2048, device="cuda"); d2 = torch.empty(B, 2048, device="cuda")
i1 = torch.empty(B, 2048, device="cuda", dtype=torch.int32); i2 = torch.empty_like(i1)
for _ in range(4):
    chamfer_3D.forward(x1, x2, d1, d2, i1, i2)
torch.cuda.synchronize()
print("ok", float(d1.sum()))
