"""Retry rate and quality of solve_model on a synthetic set: python tools/hard_probe.py shape first count"""
import json, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from houv_b200 import shard, solver, synth
shape, first, count = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
radius = float(sys.argv[4]) if len(sys.argv) > 4 else 0.5
s, t, tr = synth.make_pairs(first, count, shape=shape, radius=radius)
src, tgt, gt = (torch.from_numpy(x).cuda() for x in (s, t, tr))
for rep in range(2):
    torch.cuda.synchronize(); t0 = time.time()
    ans, met = shard.register_pairs(src, tgt, gt, batch_size=50)
    torch.cuda.synchronize(); dt = time.time() - t0
m = met.cpu().numpy()
ang = np.degrees(np.arccos(np.clip((np.trace(tr[:, :3, :3], axis1=1, axis2=2) - 1) / 2, -1, 1)))
big = ang > 45
print(json.dumps(dict(shape=shape, radius=radius, pairs=count, seconds=dt, pairs_per_s=count / dt, stats=solver.LAST_STATS,
                      retry_fraction=solver.LAST_STATS["retried_pairs"] / count,
                      rot_err_deg_mean=float(m[:, 0].mean()), rot_err_deg_median=float(np.median(m[:, 0])),
                      frac_rot_err_below_5deg=float((m[:, 0] < 5).mean()),
                      frac_big_rotation_pairs=float(big.mean()),
                      rot_err_mean_small_rot=float(m[~big, 0].mean()), rot_err_mean_big_rot=float(m[big, 0].mean()),
                      mse_mean=float(m[:, 2].mean()))))
