"""Small workload over the kernels added in round 2, for compute-sanitizer:
    compute-sanitizer --tool racecheck|memcheck|synccheck python tools/sanitize_r2.py
pose_duo_kernel (both forms, incl. the one-row tail / unpaired pass), the K1 column-slice planes + unpack, the
deterministic K2 kernels (plain and TMA-pipelined, assign and accumulate), houv_chamfer3d_forward_static."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from houv_b200 import _lib, solver, synth
from houv_b200.chamfer import chamfer_3D

lib = _lib.load()
s, t, _ = synth.make_pairs(0, 3, n_points=200)
src, tgt = torch.from_numpy(s).cuda(), torch.from_numpy(t).cuda()
for variant in (20, 21):
    lib.houv_set_pose_variant(variant)
    p = torch.from_numpy(solver.initial_parameters(3 * 32)).cuda()
    solver.pose_optimise(src, tgt, p, 32, 2, want_loss=True, want_trace=True)
    rp = torch.tensor([0] * 5 + [1] * 8 + [2] * 4, dtype=torch.int32).cuda()       # unpaired rows + odd count
    p = torch.from_numpy(solver.initial_parameters(32))[:17].contiguous().cuda()
    solver.pose_optimise(src, tgt, p, 32, 2, row_pair=rp)
lib.houv_set_pose_variant(-1)
torch.manual_seed(0)
lib.houv_set_nn_variant(10)                                                         # symmetric K1 whatever the shape
for (B, N, M) in ((3, 700, 600), (2, 2100, 300)):
    a, b = torch.rand(B, N, 3, device="cuda"), torch.rand(B, M, 3, device="cuda")
    d1, d2 = torch.empty(B, N, device="cuda"), torch.empty(B, M, device="cuda")
    i1 = torch.empty(B, N, device="cuda", dtype=torch.int32); i2 = torch.empty(B, M, device="cuda", dtype=torch.int32)
    chamfer_3D.forward(a, b, d1, d2, i1, i2)
    rc = lib.houv_chamfer3d_forward_static(a.data_ptr(), b.data_ptr(), B, N, M, d1.data_ptr(), d2.data_ptr(),
                                           i1.data_ptr(), i2.data_ptr(), _lib.stream_ptr())
    assert rc == 0
lib.houv_set_nn_variant(-1)
torch.use_deterministic_algorithms(True)
for (B, N, M) in ((5, 256, 300), (3, 77, 50), (150, 64, 64)):                       # pipelined / plain gather
    a, b = torch.rand(B, N, 3, device="cuda"), torch.rand(B, M, 3, device="cuda")
    d1, d2 = torch.empty(B, N, device="cuda"), torch.empty(B, M, device="cuda")
    i1 = torch.empty(B, N, device="cuda", dtype=torch.int32); i2 = torch.empty(B, M, device="cuda", dtype=torch.int32)
    chamfer_3D.forward(a, b, d1, d2, i1, i2)
    g1, g2 = torch.zeros_like(a), torch.zeros_like(b)
    chamfer_3D.backward(a, b, g1, g2, torch.rand(B, N, device="cuda"), torch.rand(B, M, device="cuda"), i1, i2)
    chamfer_3D.backward(a, b, g1, g2, torch.rand(B, N, device="cuda"), torch.rand(B, M, device="cuda"), i1, i2,
                        assign=True)
torch.use_deterministic_algorithms(False)
torch.cuda.synchronize()
print("sanitize r2 workload ok")
