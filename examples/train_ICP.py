#!/usr/bin/env python
"""Entry script with the shape of the reference's registration/train_ICP.py (:59-234): per batch,
point-to-point ICP of every pair (threshold 0.02, fixed initial rotation, <= 500 iterations,
:124-153) then RotE / transE / MSE (:187-190) -- one houv_icp launch per batch instead of a Python
loop over Open3D calls.  Plotting (:122, :222-234) is left out.

    python examples/train_ICP.py -c examples/cfgs/houv.yaml [--pairs N]"""
import argparse
import os
import sys
import time

import torch
import yaml

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from houv_b200 import icp, shard, synth  # noqa: E402
from houv_b200.train_utils import AverageValueMeter  # noqa: E402


def train(args):
    if args["data"] == "synthetic":
        s, t, tr = synth.make_pairs(0, int(args["pairs"]), n_points=int(args["num_points"]))
    else:
        s, t, tr = shard.load_mvp_pairs(args["data"])
    meters = {k: AverageValueMeter() for k in ("RotE", "transE", "MSE")}
    bs = int(args["batch_size"])
    for l in range(0, len(s), bs):
        src = torch.from_numpy(s[l:l + bs]).float().cuda()
        tgt = torch.from_numpy(t[l:l + bs]).float().cuda()
        gt = torch.from_numpy(tr[l:l + bs]).float().cuda()
        start = time.time()                                            # train_ICP.py:117-118
        ans, r_err, t_err, mse = icp.icp_batch(src, tgt, gt)
        torch.cuda.synchronize()
        print("time:", time.time() - start)                            # :192-193
        meters["RotE"].update(r_err.mean().item())
        meters["transE"].update(t_err.mean().item())
        meters["MSE"].update(mse.mean().item())
    print(meters["RotE"].avg)
    print(meters["transE"].avg)
    print(meters["MSE"].avg)


if __name__ == "__main__":
    ap = argparse.ArgumentParser(description="Train config file")
    ap.add_argument("-c", "--config", required=True)
    ap.add_argument("--pairs", type=int, default=None)
    a = ap.parse_args()
    cfg = yaml.safe_load(open(a.config))
    if a.pairs is not None:
        cfg["pairs"] = a.pairs
    train(cfg)
