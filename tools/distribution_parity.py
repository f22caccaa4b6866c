"""python tools/distribution_parity.py [first count chunk] > profiles/r2_distribution_parity_150.json
Distributional parity over a whole shard (report (iv)): the fused predict_model against the
reference's own GPU loop (torch forward / autograd / Adam over the reference's unmodified CUDA extension,
tests/refloop.py), angle_base 0, 64 starts x 200 iterations, both walked in the same chunks of `chunk` pairs
(initial values and the 1/rows gradient scale depend on the batch: houv.py:116,124)."""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from houv_b200 import loss as L, solver, synth  # noqa: E402
from tests.refloop import reference_predict_model  # noqa: E402

first = int(sys.argv[1]) if len(sys.argv) > 1 else 0
count = int(sys.argv[2]) if len(sys.argv) > 2 else 150
chunk = int(sys.argv[3]) if len(sys.argv) > 3 else 25
s, t, tr = synth.make_pairs(first, count)
src, tgt, gt = (torch.from_numpy(x).cuda() for x in (s, t, tr))
res = {"reference_loop": [], "fused": []}
secs = {"reference_loop": 0.0, "fused": 0.0}
for lo in range(0, count, chunk):
    a, b = src[lo:lo + chunk], tgt[lo:lo + chunk]
    for name, fn in (("reference_loop", lambda: reference_predict_model(a, b, num_epochs=200)),
                     ("fused", lambda: solver.predict_model(solver.HOUV(a.shape[0] * 64, 0), a, b, kernel=64,
                                                            num_epochs=200))):
        torch.cuda.synchronize(); t0 = time.time()
        m, R, T = fn()
        torch.cuda.synchronize(); secs[name] += time.time() - t0
        res[name].append((m, R, T))
rep = dict(pairs=count, first=first, chunk=chunk, iterations=200, points=int(src.shape[1]),
           seconds=secs, pairs_per_s={k: count / v for k, v in secs.items()})
side = {}
idx = torch.arange(count, device=src.device)
for name, parts in res.items():
    m, R, T = (torch.cat([p[k] for p in parts]) for k in range(3))
    sel = m.argmin(1)
    ans = torch.zeros(count, 4, 4, device=src.device)
    ans[:, :3, :3], ans[:, :3, 3] = R[idx, sel], T[idx, sel]
    re, te = L.rotation_error(ans[:, :3, :3], gt[:, :3, :3]), L.translation_error(ans[:, :3, 3], gt[:, :3, 3])
    ms = L.rmse_loss(src, ans, gt)
    side[name] = (m, sel, re)
    rep[name] = dict(rot_err_deg_mean=float(re.mean()), rot_err_deg_median=float(re.median()),
                     frac_rot_err_below_5deg=float((re < 5).float().mean()),
                     trans_err_mean=float(te.mean()), trans_err_median=float(te.median()),
                     mse_mean=float(ms.mean()), mse_median=float(ms.median()),
                     best_min_1_mean=float(m[idx, sel].mean()),
                     frac_best_min_1_above_0_030=float((m[idx, sel] > 0.030).float().mean()))
(ma, sa, ra), (mb, sb, rb) = side["fused"], side["reference_loop"]
rel = ((ma[idx, sa] - mb[idx, sb]).abs() / mb[idx, sb]).cpu().numpy()
rep["per_pair"] = dict(same_selected_start=float((sa == sb).float().mean()),
                       best_min_1_rel_median=float(np.median(rel)), best_min_1_rel_p90=float(np.percentile(rel, 90)),
                       starts_same_end_1e_3=float((((ma - mb).abs() / mb) <= 1e-3).float().mean()),
                       same_basin=float((((ra < 5) == (rb < 5))).float().mean()),
                       same_retry_decision=float(((ma[idx, sa] > 0.030) == (mb[idx, sb] > 0.030)).float().mean()))
print(json.dumps(rep, indent=1))
