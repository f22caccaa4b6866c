"""CPU study for DESIGN section 8 item 1(a): how many (64-row x 64-column) blocks of the K3 sweep could be
skipped by bounding-box lower bounds in ALL four metrics (3-D and the three views), on MVP-shaped synthetic
pairs at an aligned pose (ground truth) and at random start poses.  numpy only; no GPU.
usage: python tools/prune_study.py [pairs]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from houv_b200 import synth

def morton(p):
    q = np.clip(((p - p.min(0)) / (np.ptp(p, 0).max() + 1e-9) * 1023).astype(np.int64), 0, 1023)
    def spread(v):
        v = (v | (v << 16)) & 0x030000FF; v = (v | (v << 8)) & 0x0300F00F
        v = (v | (v << 4)) & 0x030C30C3; v = (v | (v << 2)) & 0x09249249
        return v
    return spread(q[:, 0]) | (spread(q[:, 1]) << 1) | (spread(q[:, 2]) << 2)

MASKS = [np.array(m, bool) for m in ([1, 1, 1], [0, 1, 1], [1, 0, 1], [1, 1, 0])]

def study(y, t, blk=64):
    y = y[np.argsort(morton(y))]; t = t[np.argsort(morton(t))]
    nb = len(y) // blk
    d = y[:, None, :] - t[None, :, :]
    skip_all = np.ones((nb, nb), bool)
    for m in MASKS:
        dm = (d[..., m] ** 2).sum(-1)
        rmin, cmin = dm.min(1), dm.min(0)
        rmax = rmin.reshape(nb, blk).max(1); cmax = cmin.reshape(nb, blk).max(1)
        ylo = y[:, m].reshape(nb, blk, -1).min(1); yhi = y[:, m].reshape(nb, blk, -1).max(1)
        tlo = t[:, m].reshape(nb, blk, -1).min(1); thi = t[:, m].reshape(nb, blk, -1).max(1)
        gap = np.maximum(0, np.maximum(ylo[:, None] - thi[None], tlo[None] - yhi[:, None]))
        lb = (gap ** 2).sum(-1)
        # skippable with perfect knowledge of the final minima (upper bound on what a running bound achieves)
        skip_all &= (lb > rmax[:, None]) & (lb > cmax[None, :])
    return skip_all.mean()

def rand_rot(rng):
    q = rng.normal(size=4); q /= np.linalg.norm(q); w, x, yy, z = q
    return np.array([[1-2*(yy*yy+z*z), 2*(x*yy-z*w), 2*(x*z+yy*w)], [2*(x*yy+z*w), 1-2*(x*x+z*z), 2*(yy*z-x*w)],
                     [2*(x*z-yy*w), 2*(yy*z+x*w), 1-2*(x*x+yy*yy)]])

P = int(sys.argv[1]) if len(sys.argv) > 1 else 4
s, t, tr = synth.make_pairs(0, P)
rng = np.random.RandomState(0)
al, rd = [], []
for p in range(P):
    R, T = tr[p, :3, :3].astype(np.float64), tr[p, :3, 3].astype(np.float64)
    al.append(study(s[p] @ R.T + T, t[p]))
    for _ in range(3):
        rd.append(study(s[p] @ rand_rot(rng).T + rng.normal(size=3) * 0.05, t[p]))
print("blocks skippable in all four metrics (64x64 blocks, Morton order, oracle bounds): aligned pose %.1f %%, random poses %.1f %%"
      % (100 * np.mean(al), 100 * np.mean(rd)))
