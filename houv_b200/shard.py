"""Multi-GPU driver: independent pairs sharded by contiguous index range, one
process per GPU, no traffic during the optimisation, ONE all-gather of the
per-pair poses and metrics at the end.

Replaces the reference's 4-background-jobs + .npy files + `sleep 600` + combine
scheme (/root/reference/registration/run_test.sh:6,21-23;
test_mult.py:34,65-66,70-81; dataset.py:354-372 MVP_RG_rotated_bound(l, r)).

Batching follows the reference's DataLoader (cfgs/houv.yaml:1 batch_size,
train_HOUV.py:69,100-112): each rank walks its range in batches and calls
solve_model per batch.  A pair's result depends on the batch it sits in
(initial values are drawn per batch row, houv.py:116, and the loss is the mean
over the batch, :124), so results are independent of the number of GPUs exactly
when `batch_size` divides the per-rank range -- `check_layout` says so (bench.py prints it).
"""
import os

import torch

from . import solver
from .loss import rmse_loss, rotation_error, translation_error


def shard_range(total, rank, world):
    """Contiguous [l, r) of `total` pairs for `rank` (test_mult.py's -l/-r)."""
    per = (total + world - 1) // world
    l = min(rank * per, total)
    return l, min(l + per, total)


def init_distributed(backend=None):
    """torchrun environment -> (rank, world, local_rank); no-op for one process."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not torch.distributed.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        torch.distributed.init_process_group(backend=backend)
    return rank, world, local


def register_pairs(src, tgt, transform=None, batch_size=50, kernel=64, num_epochs=200,
                   solve=None, balance=None):
    """Full HOUV registration (solve_model semantics) of src[P,N,3] -> tgt[P,N,3] in
    batches.  -> ans[P,4,4] and, when the ground truth is given, metrics[P,3] =
    (rotation error deg, translation error, mse) as train_HOUV.py:112-118 logs them.
    `solve`: a solve_model-shaped callable to run batch by batch instead (the CPU tests pass one).
    `balance`: see solver.solve_model_batches (None = on under torch.distributed with > 1 rank)."""
    P = src.shape[0]
    if solve is None:
        # every batch's first pass in one launch, every batch's retries in a second one:
        # per-pair results identical to solve_model batch by batch (solver.solve_model_batches)
        ans = solver.solve_model_batches(src, tgt, batch_size=batch_size, kernel=kernel,
                                         num_epochs=num_epochs, balance=balance)
        met = None
        if transform is not None:
            met = torch.stack([rotation_error(ans[:, :3, :3], transform[:, :3, :3]),
                               translation_error(ans[:, :3, 3], transform[:, :3, 3]),
                               rmse_loss(src, ans, transform)], dim=1)
        return ans, met
    net = None
    ans_all, met_all = [], []
    for l in range(0, P, batch_size):
        r = min(l + batch_size, P)
        if net is None:
            net = solver.HOUV((r - l) * kernel, 0, device=src.device)
        ans = solve(net, src[l:r], tgt[l:r], None, kernel=kernel, num_epochs=num_epochs,
                    prefix='device')
        ans_all.append(ans)
        if transform is not None:
            gt = transform[l:r]
            met_all.append(torch.stack([
                rotation_error(ans[:, :3, :3], gt[:, :3, :3]),
                translation_error(ans[:, :3, 3], gt[:, :3, 3]),
                rmse_loss(src[l:r], ans, gt)], dim=1))
    ans = torch.cat(ans_all) if ans_all else torch.zeros(0, 4, 4, device=src.device)
    met = torch.cat(met_all) if met_all else None
    return ans, met


def gather_results(ans, metrics, world):
    """The only collective of the path: all_gather of [P_local, 16+3] floats
    (NCCL over NVLink on GPUs, gloo in the CPU tests).  Ranks may hold DIFFERENT counts
    (shard_range gives 3,3,3,1 for 10 pairs on 4 ranks, or none at all): the per-rank counts are
    gathered first, every rank pads to the largest, and the padding is dropped after the
    gather -- rows come back in rank order = pair order."""
    packed = torch.cat([ans.reshape(ans.shape[0], 16),
                        metrics if metrics is not None else ans.new_zeros(ans.shape[0], 3)], dim=1)
    packed = packed.contiguous()
    if world > 1:
        dist = torch.distributed
        n_l = packed.shape[0]
        cnt = torch.tensor([n_l], device=packed.device, dtype=torch.int64)
        cnts = torch.empty(world, device=packed.device, dtype=torch.int64)
        dist.all_gather_into_tensor(cnts, cnt)
        counts = cnts.tolist()
        nmax = max(counts)
        if nmax == 0:
            out = packed
        else:
            padded = packed.new_zeros(nmax, packed.shape[1])
            padded[:n_l] = packed
            g = packed.new_empty(world * nmax, packed.shape[1])
            dist.all_gather_into_tensor(g, padded)
            g = g.reshape(world, nmax, packed.shape[1])
            out = torch.cat([g[k, :counts[k]] for k in range(world)])
    else:
        out = packed
    return out[:, :16].reshape(-1, 4, 4), out[:, 16:]


def check_layout(total, world, batch_size):
    """Results are independent of the number of GPUs exactly when every rank's range is a whole
    number of batches (a pair's result depends on its position in its batch: houv.py:116,124).
    -> None when that holds, else a one-line warning text (bench.py and examples print it)."""
    per = (total + world - 1) // world
    if world > 1 and (per % batch_size != 0 or total % per != 0):
        return ("%d pairs on %d ranks in batches of %d: per-rank range %d is not a whole number of "
                "batches (or the last rank is short) -- poses will differ from a 1-GPU run"
                % (total, world, batch_size, per))
    return None


# ----------------------------------------------------------------- result files
# The reference's file-based merge, kept for drop-in use next to the all-gather:
# every process saves its range as `{l}_{r}.npy` (test_mult.py:65-66) and a --combine pass
# concatenates them into `results.h5`, dataset 'results' [P,4,4] float32 (test_mult.py:70-81;
# test.py:67-77 writes the same file for a single process).

def save_range(ans, log_dir, l, r):
    """test_mult.py:65-66: np.save(log_dir + '/{l}_{r}.npy', all_results)."""
    import numpy as np
    os.makedirs(log_dir, exist_ok=True)
    path = os.path.join(log_dir, '{}_{}.npy'.format(l, r))
    np.save(path, ans.detach().cpu().numpy().astype('float32') if torch.is_tensor(ans) else ans)
    return path


def combine(log_dir, step=500, num=4):
    """test_mult.py:70-81: concatenate the `num` range files of width `step` and write
    results.h5 (dataset 'results').  h5py is an optional dependency: without it the merged
    array is written as results.npy only.  A missing range file raises FileNotFoundError like
    the reference's np.load.  -> the merged [P,4,4] array."""
    import numpy as np
    parts = [np.load(os.path.join(log_dir, '{}_{}.npy'.format(step * i, step * i + step)))
             for i in range(num)]
    all_results = np.concatenate(parts, axis=0)
    np.save(os.path.join(log_dir, 'results.npy'), all_results)
    try:
        import h5py
    except ImportError:
        h5py = None
    if h5py is not None:
        with h5py.File(os.path.join(log_dir, 'results.h5'), 'w') as f:
            f.create_dataset('results', data=all_results)
    return all_results


def load_mvp_pairs(path, l=None, r=None):
    """The arrays MVP_RG_rotated(_bound) hands to the solver (dataset.py:189-233, 354-372):
    'rotated_src' / 'rotated_tgt' [P,2048,3] float32 and, where the file has them (the val
    split), 'transforms' [P,4,4].  Needs h5py (the MVP files are HDF5); -> (src, tgt, transforms
    or None) as numpy arrays for the index range [l, r)."""
    try:
        import h5py
    except ImportError as e:  # pragma: no cover
        raise RuntimeError("reading MVP_*_RG.h5 needs h5py, which this environment lacks") from e
    with h5py.File(path, 'r') as f:
        sl = slice(l, r)
        src = f['rotated_src'][sl].astype('float32')
        tgt = f['rotated_tgt'][sl].astype('float32')
        tr = f['transforms'][sl].astype('float32') if 'transforms' in f else None
    return src, tgt, tr
