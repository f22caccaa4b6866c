"""houv_chamfer3d_forward_static -- the entry point that allocates nothing (its scratch is a static
device array inside the library, include/houv_b200.h) -- called through the C ABI and compared
with the existing entry point houv_chamfer3d_forward_ws, bit for bit (the bar of every K1 test),
on calls in a row that share the array.  No counterpart in the reference.
This file is named to be collected LAST."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

E_SCRATCH = 100003  # include/houv_b200.h HOUV_E_SCRATCH


def _outputs(B, N, M):
    dev = "cuda"
    return (torch.full((B, N), -1.0, device=dev), torch.full((B, M), -1.0, device=dev),
            torch.full((B, N), -1, device=dev, dtype=torch.int32),
            torch.full((B, M), -1, device=dev, dtype=torch.int32))


def _call_static(lib, x1, x2, out, stream):
    B, N, _ = x1.shape
    M = x2.shape[1]
    return lib.houv_chamfer3d_forward_static(x1.data_ptr(), x2.data_ptr(), B, N, M, out[0].data_ptr(),
                                             out[1].data_ptr(), out[2].data_ptr(), out[3].data_ptr(),
                                             stream)


def _call_ws(lib, x1, x2, out, stream):
    B, N, _ = x1.shape
    M = x2.shape[1]
    need = lib.houv_chamfer3d_workspace_bytes(B, N, M)
    ws = torch.empty(max(need, 8) // 8, device="cuda", dtype=torch.int64)  # the CALLER's workspace
    rc = lib.houv_chamfer3d_forward_ws(x1.data_ptr(), x2.data_ptr(), B, N, M, out[0].data_ptr(),
                                       out[1].data_ptr(), out[2].data_ptr(), out[3].data_ptr(),
                                       ws.data_ptr() if need else 0, need, stream)
    return rc, need


def test_static_scratch_entry_point_matches_workspace_entry_point():
    from houv_b200 import _lib
    lib = _lib.load()
    cap = lib.houv_chamfer3d_static_scratch_bytes()
    assert cap == 8 << 20
    stream = _lib.stream_ptr()
    g = torch.Generator(device="cuda").manual_seed(5)
    # shapes whose symmetric kernel USES scratch: row-key planes (column slices of a small batch),
    # column keys + planes (several row chunks), column keys alone; then a small call that needs
    # none -- in a row, so every call finds the previous call's contents in the shared array
    shapes = [(16, 2048, 2048), (6, 4100, 2048), (16, 2048, 2048), (64, 8192, 2048), (3, 100, 70)]
    used = 0
    for k, (B, N, M) in enumerate(shapes):
        x1 = torch.rand(B, N, 3, device="cuda", generator=g)
        x2 = torch.rand(B, M, 3, device="cuda", generator=g)
        if k == 2:
            x2[:, 7] = x2[:, 3]  # exact ties: the first index must win in both entry points
        ref, got = _outputs(B, N, M), _outputs(B, N, M)
        rc, need = _call_ws(lib, x1, x2, ref, stream)
        assert rc == 0 and need <= cap
        used += need > 0
        assert _call_static(lib, x1, x2, got, stream) == 0
        torch.cuda.synchronize()
        for a, b in zip(got, ref):
            assert torch.equal(a, b), (B, N, M)
        assert int(got[2].min()) >= 0 and int(got[2].max()) < M
        # the same call again straight away: the array now holds THIS call's keys
        again = _outputs(B, N, M)
        assert _call_static(lib, x1, x2, again, stream) == 0
        torch.cuda.synchronize()
        for a, b in zip(again, ref):
            assert torch.equal(a, b), (B, N, M)
    assert used >= 4  # the scratch path was exercised, not only the directed kernel


def test_static_scratch_rejects_what_does_not_fit():
    from houv_b200 import _lib
    lib = _lib.load()
    B, N, M = 200, 8192, 8192  # 200 x 8192 column keys = 12.5 MiB > 8 MiB
    assert lib.houv_chamfer3d_workspace_bytes(B, N, M) > lib.houv_chamfer3d_static_scratch_bytes()
    x = torch.rand(B, N, 3, device="cuda")
    out = _outputs(B, N, M)
    assert _call_static(lib, x, x, out, _lib.stream_ptr()) == E_SCRATCH
    torch.cuda.synchronize()
    assert float(out[0].max()) == -1.0 and int(out[2].max()) == -1  # rejected before any launch
    with pytest.raises(RuntimeError):  # what the Python layer turns a status into
        _lib.check(E_SCRATCH, "houv_chamfer3d_forward_static")
    # bad arguments are still bad arguments
    assert lib.houv_chamfer3d_forward_static(0, 0, -1, 4, 4, 0, 0, 0, 0, 0) == 100001
