"""DEV-CONTAINER ONLY (needs /root/reference).  Imports the reference's HOUV
python files UNMODIFIED on a GPU-less host so that golden vectors can be made
from the reference's own code:

  registration/models/houv.py, registration/model_utils_completion.py,
  registration/train_utils.py, registration/model_utils.py (SVDHead)

What is stubbed, and only that:
  * `.cuda()` on tensors/modules becomes the identity (houv.py:64,67,76,
    model_utils_completion.py:159 hard-code it);
  * `metrics.cd` (the JIT CUDA extension, utils/metrics/CD/chamfer3D) is served
    by the reference's own kernel SOURCE compiled for the host
    (oracle/_ref/libref_chamfer.so) wrapped in the same autograd Function shape
    as dist_chamfer_3D.py:26-64; `metrics.fscore` is the reference's fscore.py;
  * absent third-party imports that the hot path never calls (mm3d_pn2/mmcv,
    emd, visu_utils/matplotlib/open3d, models.simple_icp) are empty modules.
"""
import importlib.util
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
REF = "/root/reference"
REG = os.path.join(REF, "registration")

_loaded = {}
FAST_CD = False      # serve metrics.cd from oracle/_build (threads + AVX2) instead of oracle/_ref
FAST_CD_CHECK = 40   # ... and compare with oracle/_ref every so many calls
_calls = [0]
BACKWARD_REVERSED = False  # see _RefChamferFunction.backward


class _RefChamferFunction(torch.autograd.Function):
    """dist_chamfer_3D.py:26-64 with chamfer_3D.forward/backward served by the
    host build of the reference kernels."""

    @staticmethod
    def forward(ctx, xyz1, xyz2):
        from oracle import chamfer as oc
        x1, x2 = xyz1.detach().numpy(), xyz2.detach().numpy()
        if FAST_CD:
            # full-size goldens: the thread-parallel C restatement (bit-equal to the reference
            # kernel source, tests/test_oracle_cpu.py), re-checked here against oracle/_ref on
            # the first rows of every FAST_CD_CHECK-th call
            d1, d2, i1, i2 = oc.chamfer_forward(x1, x2)
            _calls[0] += 1
            if _calls[0] % FAST_CD_CHECK == 1:
                r = oc.ref_chamfer_forward(x1[:2], x2[:2])
                for a, b in zip((d1, d2, i1, i2), r):
                    assert np.array_equal(a[:2], b, equal_nan=True), "C oracle != reference kernel"
        else:
            d1, d2, i1, i2 = oc.ref_chamfer_forward(x1, x2)
        dist1, dist2 = torch.from_numpy(d1), torch.from_numpy(d2)
        idx1, idx2 = torch.from_numpy(i1), torch.from_numpy(i2)
        ctx.save_for_backward(xyz1, xyz2, idx1, idx2)
        return dist1, dist2, idx1, idx2

    @staticmethod
    def backward(ctx, graddist1, graddist2, gradidx1, gradidx2):
        from oracle import chamfer as oc
        xyz1, xyz2, idx1, idx2 = ctx.saved_tensors
        x1, x2 = xyz1.detach().numpy(), xyz2.detach().numpy()
        gd1, gd2 = graddist1.contiguous().numpy(), graddist2.contiguous().numpy()
        i1, i2 = idx1.numpy(), idx2.numpy()
        if BACKWARD_REVERSED:
            # The reference adds gradients with float atomics in whatever order the GPU schedules
            # its threads (chamfer3D.cu:166-171; README.md:82-90 concedes the run-to-run drift).  Its
            # kernel body run by ONE host thread adds in ascending point order; handing it the
            # clouds with their points in descending order makes the same kernel add in the
            # opposite order -- a second, equally valid execution of the reference, used to measure
            # how far the reference drifts from ITSELF over 200 iterations.
            n, m = x1.shape[1], x2.shape[1]
            f = lambda a: np.ascontiguousarray(a[:, ::-1])
            g1, g2 = oc.ref_chamfer_backward(f(x1), f(x2), f(gd1), f(gd2),
                                             f((m - 1 - i1).astype(np.int32)),
                                             f((n - 1 - i2).astype(np.int32)))
            return torch.from_numpy(f(g1)), torch.from_numpy(f(g2))
        g1, g2 = oc.ref_chamfer_backward(x1, x2, gd1, gd2, i1, i2)
        return torch.from_numpy(g1), torch.from_numpy(g2)


class _RefChamferDist(torch.nn.Module):
    def forward(self, input1, input2):
        return _RefChamferFunction.apply(input1.contiguous().float(), input2.contiguous().float())


def _file_module(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


class _Anything(types.ModuleType):
    def __getattr__(self, k):
        if k.startswith("__"):
            raise AttributeError(k)
        return None


def load():
    """-> dict(houv=<module>, muc=<model_utils_completion>, tu=<train_utils>)."""
    if _loaded:
        return _loaded
    torch.Tensor.cuda = lambda self, *a, **k: self
    torch.nn.Module.cuda = lambda self, *a, **k: self

    fscore_mod = _file_module("ref_fscore", os.path.join(REF, "utils/metrics/CD/fscore.py"))
    metrics = types.ModuleType("metrics")
    metrics.cd = _RefChamferDist
    metrics.fscore = fscore_mod.fscore
    metrics.emd = None
    sys.modules["metrics"] = metrics
    for name in ("mm3d_pn2", "visu_utils", "models", "models.simple_icp", "open3d"):
        if name not in sys.modules:
            sys.modules[name] = _Anything(name)
    sys.path.insert(0, REG)
    cwd = os.getcwd()
    os.chdir(REG)
    try:
        muc = _file_module("model_utils_completion", os.path.join(REG, "model_utils_completion.py"))
        tu = _file_module("train_utils", os.path.join(REG, "train_utils.py"))
        houv = _file_module("ref_houv", os.path.join(REG, "models/houv.py"))
    finally:
        os.chdir(cwd)
    _loaded.update(houv=houv, muc=muc, tu=tu)
    return _loaded


def load_svdhead(rel="model_utils.py"):
    """registration/model_utils.py (and models/idam.py) import open3d & co. at the top; SVDHead
    only needs torch, so exec the class source alone (text is read, not copied)."""
    import ast
    src = open(os.path.join(REG, rel)).read()
    tree = ast.parse(src)
    node = [n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "SVDHead"][0]
    ns = {"torch": torch, "nn": torch.nn}
    exec(compile(ast.Module(body=[node], type_ignores=[]), "model_utils.py:SVDHead", "exec"), ns)
    return ns["SVDHead"]


def load_function(name, rel="model_utils.py"):
    """One top-level function of a reference file that cannot be imported whole (open3d & co. at
    the top): its source is read and executed as-is."""
    import ast
    tree = ast.parse(open(os.path.join(REG, rel)).read())
    node = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == name][0]
    ns = {"torch": torch, "np": np}
    exec(compile(ast.Module(body=[node], type_ignores=[]), rel + ":" + name, "exec"), ns)
    return ns[name]


if __name__ == "__main__":
    m = load()
    torch.manual_seed(0)
    np.random.seed(0)
    src = torch.rand(1, 64, 3) - 0.5
    tgt = torch.rand(1, 64, 3) - 0.5
    net = m["houv"].HOUV(32, 0)  # reset_weight needs >= 26 rows (houv.py:43-51)
    loss, R, T = m["houv"].predict_model(net, src, tgt, kernel=32, num_epochs=3)
    print(loss, R.shape, T.shape)
