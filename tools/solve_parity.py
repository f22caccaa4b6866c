"""python tools/solve_parity.py [first count chunk radius] > profiles/r2_solve_parity_hard_150.json
The whole solve_model (retry cascade at angle_base 1..3 included, houv.py:142-206) on the HARD set (objects of
twice the size: ~20 % of the pairs retry), fused against the same cascade over the reference's own GPU loop
(tests/refloop.py), both walked in chunks of `chunk` pairs like train_HOUV.py walks its batches."""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from houv_b200 import loss as L, solver, synth  # noqa: E402
from tests.refloop import reference_solve_model  # noqa: E402

first = int(sys.argv[1]) if len(sys.argv) > 1 else 0
count = int(sys.argv[2]) if len(sys.argv) > 2 else 150
chunk = int(sys.argv[3]) if len(sys.argv) > 3 else 25
radius = float(sys.argv[4]) if len(sys.argv) > 4 else 1.0
s, t, tr = synth.make_pairs(first, count, radius=radius)
src, tgt, gt = (torch.from_numpy(x).cuda() for x in (s, t, tr))
out = {}
for name in ("reference_loop", "fused"):
    torch.cuda.synchronize(); t0 = time.time()
    ans, ret = [], []
    for lo in range(0, count, chunk):
        a, b = src[lo:lo + chunk], tgt[lo:lo + chunk]
        if name == "fused":
            an = solver.solve_model(solver.HOUV(a.shape[0] * 64, 0), a, b, None, kernel=64, num_epochs=200,
                                    prefix="device")
            r = torch.zeros(a.shape[0], dtype=torch.bool, device=a.device)  # filled below from a first pass
        else:
            an, r = reference_solve_model(a, b)
        ans.append(an); ret.append(r)
    torch.cuda.synchronize()
    out[name] = (torch.cat(ans), torch.cat(ret), time.time() - t0)
# which pairs the fused path retries: its first pass again (cheap), same chunks
ret_f = []
for lo in range(0, count, chunk):
    a, b = src[lo:lo + chunk], tgt[lo:lo + chunk]
    m, _, _ = solver.predict_model(solver.HOUV(a.shape[0] * 64, 0), a, b, kernel=64, num_epochs=200)
    ret_f.append(m.min(1).values > solver.RETRY_THRESHOLD)
out["fused"] = (out["fused"][0], torch.cat(ret_f), out["fused"][2])
rep = dict(pairs=count, first=first, chunk=chunk, radius=radius)
errs = {}
for name, (ans, ret, sec) in out.items():
    re = L.rotation_error(ans[:, :3, :3], gt[:, :3, :3])
    te = L.translation_error(ans[:, :3, 3], gt[:, :3, 3])
    ms = L.rmse_loss(src, ans, gt)
    errs[name] = re
    rep[name] = dict(seconds=sec, pairs_per_s=count / sec, retried_pairs=int(ret.sum()),
                     retry_fraction=float(ret.float().mean()),
                     rot_err_deg_mean=float(re.mean()), rot_err_deg_median=float(re.median()),
                     frac_rot_err_below_5deg=float((re < 5).float().mean()),
                     trans_err_mean=float(te.mean()), mse_mean=float(ms.mean()), mse_median=float(ms.median()))
ra, rb = out["fused"][1], out["reference_loop"][1]
rep["per_pair"] = dict(same_retry_decision=float((ra == rb).float().mean()),
                       same_basin=float(((errs["fused"] < 5) == (errs["reference_loop"] < 5)).float().mean()),
                       rot_err_deg_abs_diff_median=float((errs["fused"] - errs["reference_loop"]).abs().median()))
print(json.dumps(rep, indent=1))
