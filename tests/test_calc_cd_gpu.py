"""GPU: calc_cd (+ fscore) -- the chamfer_3DDist consumer of the completion pipeline
(completion/model_utils.py:69-80 = registration/model_utils_completion.py:69-80; VRCNet calls it
four times per step, completion/models/vrcnet.py:631-634, and with calc_f1 at :647) -- as torch
ops over K1/K2 (training) and as K1 + the fused reduction houv_cd_reduce (evaluation)."""
import numpy as np
import pytest
import torch

from oracle import ref_ext
from tests.util import load_npz

pytestmark = pytest.mark.gpu


def _rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-30)


def test_matches_reference_golden_values_and_gradient():
    from houv_b200.loss import calc_cd
    g = load_npz("calccd_ref.npz")
    gt = torch.from_numpy(g["gt"]).cuda()
    o = torch.from_numpy(g["output"]).cuda().requires_grad_()
    cd_p, cd_t, f1 = calc_cd(o, gt, calc_f1=True)           # autograd path
    (cd_p.mean() + cd_t.mean()).backward()
    assert _rel(cd_p.detach().cpu(), g["cd_p"]) < 1e-6 and _rel(cd_t.detach().cpu(), g["cd_t"]) < 1e-6
    assert np.array_equal(f1.cpu().numpy(), g["f1"])
    # coincident points (d = 0): sqrt's backward is 0 * inf = NaN in the reference, and here
    got, ref = o.grad.cpu().numpy(), g["grad_output"]
    assert np.isnan(ref).any() and np.array_equal(np.isnan(got), np.isnan(ref))
    assert _rel(np.nan_to_num(got), np.nan_to_num(ref)) < 1e-5
    with torch.no_grad():                                    # fused evaluation path
        p2, t2, f2 = calc_cd(o, gt, calc_f1=True)
        p3, t3 = calc_cd(o, gt)
    assert _rel(p2.cpu(), g["cd_p"]) < 1e-6 and _rel(t2.cpu(), g["cd_t"]) < 1e-6
    assert np.array_equal(f2.cpu().numpy(), g["f1"])
    assert torch.equal(p2, p3) and torch.equal(t2, t3)


def test_fscore_nan_becomes_zero():
    from houv_b200.loss import calc_cd
    a = torch.zeros(2, 64, 3, device="cuda")
    b = torch.ones(2, 32, 3, device="cuda")                  # nothing within the threshold: 0/0
    with torch.no_grad():
        _, _, f1 = calc_cd(a, b, calc_f1=True)
    assert float(f1.abs().max()) == 0.0


@pytest.mark.skipif(not ref_ext.available(), reason="oracle/_ref/chamfer_3D.so not built")
@pytest.mark.parametrize("shape", [(64, 2048, 16384), (32, 1024, 2048)],
                         ids=["C5_vrcnet_fine", "vrcnet_coarse"])
def test_full_size_against_reference_extension(shape):
    """BASELINE config 5 shape: the reference's calc_cd expression over the reference's own CUDA
    extension vs the fused path."""
    from houv_b200.loss import calc_cd
    B, n_out, n_gt = shape
    torch.manual_seed(3)
    gt = torch.rand(B, n_gt, 3, device="cuda")
    out = gt[:, :n_out] + 0.003 * torch.randn(B, n_out, 3, device="cuda")
    d1, d2, _, _ = ref_ext.chamfer_forward(gt, out)
    cd_p = (torch.sqrt(d1).mean(1) + torch.sqrt(d2).mean(1)) / 2
    cd_t = d1.mean(1) + d2.mean(1)
    p1 = (d1 < 1e-4).float().mean(1)
    p2 = (d2 < 1e-4).float().mean(1)
    f1 = 2 * p1 * p2 / (p1 + p2)
    f1[torch.isnan(f1)] = 0
    with torch.no_grad():
        a, b, c = calc_cd(out, gt, calc_f1=True)
    assert _rel(a.cpu(), cd_p.cpu()) < 2e-6 and _rel(b.cpu(), cd_t.cpu()) < 2e-6
    assert _rel(c.cpu(), f1.cpu()) < 1e-6 and float(c.min()) > 0
