"""Fixed EMD workload for ncu captures: usage  python tools/ncu_emd.py B n eps iters"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from houv_b200.emd import emd_forward
B, n, eps, iters = int(sys.argv[1]), int(sys.argv[2]), float(sys.argv[3]), int(sys.argv[4])
torch.manual_seed(0)
x1, x2 = torch.rand(B, n, 3).cuda(), torch.rand(B, n, 3).cuda()
for _ in range(2):
    d, a, used = emd_forward(x1, x2, eps, iters, want_iters=True)
torch.cuda.synchronize()
print("ok", float(torch.sqrt(d).mean()), int(used.max()))
