"""Small fixed K2 / cd_reduce workload for ncu captures: backward + calc_cd epilogue at the HOUV
batch shape (B=6400, N=M=2048)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from houv_b200.chamfer import chamfer_3D
from houv_b200.loss import calc_cd
B = int(sys.argv[1]) if len(sys.argv) > 1 else 6400
torch.manual_seed(0)
x1 = torch.rand(B, 2048, 3, device="cuda"); x2 = torch.rand(B, 2048, 3, device="cuda")
d1 = torch.empty(B, 2048, device="cuda"); d2 = torch.empty(B, 2048, device="cuda")
i1 = torch.empty(B, 2048, device="cuda", dtype=torch.int32); i2 = torch.empty_like(i1)
chamfer_3D.forward(x1, x2, d1, d2, i1, i2)
gd1, gd2 = torch.rand(B, 2048, device="cuda"), torch.rand(B, 2048, device="cuda")
for _ in range(3):
    g1, g2 = torch.zeros_like(x1), torch.zeros_like(x2)
    chamfer_3D.backward(x1, x2, g1, g2, gd1, gd2, i1, i2)
with torch.no_grad():
    for _ in range(3):
        out = calc_cd(x2, x1, calc_f1=True)
torch.cuda.synchronize()
print("ok", float(g1.abs().sum()), float(out[0].mean()))
