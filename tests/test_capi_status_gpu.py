"""Status codes of the C ABI called directly through ctypes (no Python-layer checks in between): bad
arguments are refused before any launch with HOUV_E_BADARG / HOUV_E_TOOBIG, empty problems return 0."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu



def _codes():
    import re, os
    h = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "houv_b200.h")).read()
    return {k: int(v) for k, v in re.findall(r"#define\s+(HOUV_E_\w+)\s+\(?(-?\d+)\)?", h)}


def test_error_codes_of_the_new_entry_points():
    from houv_b200 import _lib
    lib = _lib.load()
    codes = _codes()
    bad, big = codes["HOUV_E_BADARG"], codes["HOUV_E_TOOBIG"]
    x = torch.rand(2, 64, 3, device="cuda")
    d = torch.empty(2, 64, device="cuda")
    a = torch.empty(2, 64, device="cuda", dtype=torch.int32)
    p = lambda t: t.data_ptr()
    # EMD: iteration cap, eps, null outputs, missing workspace for n > 4096
    assert lib.houv_emd_forward(p(x), p(x), 2, 64, 0.005, 0, p(d), p(a), 0, 0, 0, 0) == bad
    assert lib.houv_emd_forward(p(x), p(x), 2, 64, -1.0, 5, p(d), p(a), 0, 0, 0, 0) == bad
    assert lib.houv_emd_forward(p(x), p(x), 2, 64, 0.005, 5, 0, p(a), 0, 0, 0, 0) == bad
    assert lib.houv_emd_forward(p(x), p(x), 0, 64, 0.005, 5, p(d), p(a), 0, 0, 0, 0) == 0
    big_x = torch.rand(1, 5000, 3, device="cuda")
    assert lib.houv_emd_workspace_bytes(1, 5000) > 0 and lib.houv_emd_workspace_bytes(1, 4096) == 0
    assert lib.houv_emd_forward(p(big_x), p(big_x), 1, 5000, 0.005, 5, p(d), p(a), 0, 0, 0, 0) == bad
    assert lib.houv_emd_backward(p(x), p(x), 0, p(a), 2, 64, p(x), 0) == bad
    # IDAM step: size cap, R/t accumulators only together
    idx = torch.zeros(1, 9000, device="cuda", dtype=torch.int32)
    lg = torch.zeros(1, 9000, device="cuda")
    s9 = torch.rand(1, 3, 9000, device="cuda")
    assert lib.houv_idam_step(p(s9), p(s9), p(idx), 0, p(lg), 1, 9000, 9000, 0, 0, 0, 0, 0, 0, 0) == big
    s1 = torch.rand(1, 3, 64, device="cuda")
    R = torch.eye(3, device="cuda").reshape(1, 9).contiguous()
    assert lib.houv_idam_step(p(s1), p(s1), p(idx), 0, p(lg), 1, 64, 64, 0, p(R), 0, 0, 0, 0, 0) == bad
    # backward-assign keeps the argument checks of backward
    assert lib.houv_chamfer3d_backward_assign(p(x), p(x), 0, p(d), p(a), p(a), 2, 64, 64, p(x), p(x), 0) == bad
    assert lib.houv_chamfer3d_backward_assign(p(x), p(x), p(d), p(d), p(a), p(a), -1, 64, 64, p(x), p(x), 0) == bad
    torch.cuda.synchronize()
