"""Small fixed K3 workload for ncu captures: 5 pairs x 64 starts (320 rows, one full
wave of 296 resident CTAs + tail) x 20 iterations at 2048 points."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from houv_b200 import solver, synth
P = int(sys.argv[1]) if len(sys.argv) > 1 else 5
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 20
s, t, _ = synth.make_pairs(0, P)
src, tgt = torch.from_numpy(s).cuda(), torch.from_numpy(t).cuda()
for _ in range(2):
    p = torch.from_numpy(solver.initial_parameters(P * 64)).cuda()
    out = solver.pose_optimise(src, tgt, p, 64, iters)
torch.cuda.synchronize()
print("ok", float(out["min_1"].min()))
