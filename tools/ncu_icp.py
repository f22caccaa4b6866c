"""Fixed ICP workload for ncu captures: P MVP-shaped pairs, train_ICP settings."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from houv_b200 import icp, synth
P = int(sys.argv[1]) if len(sys.argv) > 1 else 148
s, t, tr = synth.make_pairs(0, P)
src, tgt = torch.from_numpy(s).cuda(), torch.from_numpy(t).cuda()
for _ in range(2):
    reg = icp.registration_icp(src, tgt, 0.02, icp.TRANS_INIT, 500)
torch.cuda.synchronize()
print("ok", float(reg.iterations.mean()), float(reg.iterations.max()))
