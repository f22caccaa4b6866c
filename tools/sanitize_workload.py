"""Tiny workload that touches every kernel once, for compute-sanitizer:
    compute-sanitizer --tool memcheck|racecheck|synccheck|initcheck python tools/sanitize_workload.py"""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from houv_b200 import _lib, icp, kabsch, solver, synth, train_utils
from houv_b200.chamfer import chamfer_3DDist

lib = _lib.load()
torch.manual_seed(0)
for variant in (-1, 10):                       # directed and symmetric K1, K2 behind them
    lib.houv_set_nn_variant(variant)
    for (B, N, M) in ((2, 300, 4500), (3, 2050, 70)):
        a = torch.rand(B, N, 3, device="cuda", requires_grad=True)
        b = torch.rand(B, M, 3, device="cuda", requires_grad=True)
        d1, d2, i1, i2 = chamfer_3DDist()(a, b)
        (d1.sum() + d2.sum()).backward()
lib.houv_set_nn_variant(-1)
s, t, tr = synth.make_pairs(0, 2, n_points=200)
src, tgt = torch.from_numpy(s).cuda(), torch.from_numpy(t).cuda()
for variant in (12, 0):                        # symmetric and two-scan K3
    lib.houv_set_pose_variant(variant)
    p = torch.from_numpy(solver.initial_parameters(2 * 32)).cuda()
    solver.pose_optimise(src, tgt, p, 32, 2, want_loss=True, want_trace=True)
    np.random.seed(0)
    train_utils.getPredict_angle(src, tgt, kernel=8, num_epochs=2)
lib.houv_set_pose_variant(-1)
icp.registration_icp(src, tgt, 0.05, icp.TRANS_INIT, 5)
x = torch.rand(3, 3, 100, device="cuda")
kabsch.SVDHead(None)(x, x + 0.01 * torch.rand_like(x))
torch.cuda.synchronize()
print("sanitize workload ok")
