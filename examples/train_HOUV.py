#!/usr/bin/env python
"""Entry script with the shape of the reference's registration/train_HOUV.py (:40-153): read the
YAML config, seed, walk the pairs in batches, solve_model per batch, running RotE / transE / MSE
logged every 10 batches and printed at the end -- on houv_b200 instead of models.houv.

    python examples/train_HOUV.py -c examples/cfgs/houv.yaml [--pairs N]

Data: the reference reads MVP_Test_RG.h5 through MVP_RG_rotated('val') (dataset.py:189-233), which
is not shipped; `data: synthetic` uses the MVP-shaped generator, a path reads the same arrays from
an MVP h5 file (h5py required)."""
import argparse
import logging
import os
import random
import sys

import numpy as np
import torch
import yaml

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from houv_b200 import shard, synth  # noqa: E402
from houv_b200.solver import HOUV, solve_model  # noqa: E402  (train_HOUV.py:38)
from houv_b200.train_utils import AverageValueMeter, rmse_loss  # noqa: E402


def train(args):
    logging.info(str(args))
    if args["data"] == "synthetic":
        s, t, tr = synth.make_pairs(0, int(args["pairs"]), n_points=int(args["num_points"]))
    else:
        s, t, tr = shard.load_mvp_pairs(args["data"])
    logging.info("Length of dataset:%d", len(s))
    seed = int(args["manual_seed"]) if args.get("manual_seed") else random.randint(1, 10000)
    logging.info("Random Seed: %d", seed)
    random.seed(seed); np.random.seed(seed); torch.manual_seed(seed)   # train_HOUV.py:74-82

    meters = {k: AverageValueMeter() for k in ("RotE", "transE", "MSE")}
    bs = int(args["batch_size"])
    net = HOUV(bs * int(args["kernel"]), 0)                            # train_HOUV.py:84-85
    net.cuda()
    for i, l in enumerate(range(0, len(s), bs)):
        src_rotated = torch.from_numpy(s[l:l + bs]).float().cuda()
        tgt_rotated = torch.from_numpy(t[l:l + bs]).float().cuda()
        transform = torch.from_numpy(tr[l:l + bs]).float().cuda()
        r_err, t_err, ans = solve_model(net, src_rotated, tgt_rotated, transform)   # :112
        mse = rmse_loss(src_rotated, ans, transform)
        meters["RotE"].update(r_err.mean().item())
        meters["transE"].update(t_err.mean().item())
        meters["MSE"].update(mse.mean().item())
        if i % 10 == 0:
            logging.info("RotE:{} TransE:{} MSE:{}".format(meters["RotE"].avg, meters["transE"].avg,
                                                           meters["MSE"].avg))
    print(meters["RotE"].avg)
    print(meters["transE"].avg)
    print(meters["MSE"].avg)


if __name__ == "__main__":
    ap = argparse.ArgumentParser(description="Train config file")
    ap.add_argument("-c", "--config", help="path to config file", required=True)
    ap.add_argument("--pairs", type=int, default=None, help="override `pairs` (synthetic data)")
    a = ap.parse_args()
    cfg = yaml.safe_load(open(a.config))
    if a.pairs is not None:
        cfg["pairs"] = a.pairs
    log_dir = os.path.join(cfg["work_dir"], "%s_%s" % (cfg["model_name"], cfg["flag"]))
    os.makedirs(log_dir, exist_ok=True)
    logging.basicConfig(level=logging.INFO, handlers=[
        logging.FileHandler(os.path.join(log_dir, "train.log")), logging.StreamHandler(sys.stdout)])
    train(cfg)
