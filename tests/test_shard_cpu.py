"""CPU: host logic of the multi-GPU driver with a world_size-2 gloo group -- contiguous
range sharding (test_mult.py -l/-r semantics), per-batch solve calls, and the ONE all-gather
of poses + metrics.  The GPU solve is replaced by a deterministic stand-in (no compute calls
without a GPU); the real path is covered by bench.py --gpus N on the box."""
import os
import socket

import numpy as np
import torch
import torch.multiprocessing as mp

from houv_b200 import shard


def test_shard_ranges_cover_and_are_contiguous():
    for total, world in [(1200, 8), (1200, 1), (10, 4), (3, 8), (0, 2)]:
        seen = []
        for r in range(world):
            l, rr = shard.shard_range(total, r, world)
            assert 0 <= l <= rr <= total
            seen += list(range(l, rr))
        assert seen == list(range(total))
    assert shard.shard_range(1200, 3, 8) == (450, 600)


def _fake_solve(net, src, tgt, pose, kernel=64, num_epochs=200, prefix='device'):
    # pose that depends only on the pair's data: rank/batch layout must not matter
    ans = torch.zeros(src.shape[0], 4, 4)
    ans[:, :3, :3] = torch.eye(3)
    ans[:, :3, 3] = tgt.mean(1) - src.mean(1)
    return ans


class _NoNet:
    def __init__(self, *a, **k):
        pass


def _worker(rank, world, port, total, batch, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    r, w, _ = shard.init_distributed("gloo")
    assert (r, w) == (rank, world)
    rng = np.random.RandomState(0)
    src = torch.from_numpy(rng.rand(total, 16, 3).astype(np.float32))
    tgt = torch.from_numpy(rng.rand(total, 16, 3).astype(np.float32))
    gt = torch.eye(4).repeat(total, 1, 1)
    l, rr = shard.shard_range(total, rank, world)
    shard.solver.HOUV = _NoNet
    ans, met = shard.register_pairs(src[l:rr], tgt[l:rr], gt[l:rr], batch_size=batch, kernel=4,
                                    num_epochs=1, solve=_fake_solve)
    ans_all, met_all = shard.gather_results(ans, met, world)
    q.put((rank, ans_all.numpy(), met_all.numpy()))
    torch.distributed.barrier()
    torch.distributed.destroy_process_group()


def test_gloo_world2_gather_matches_single_process():
    total, batch, world = 12, 3, 2
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, total, batch, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # single-process reference of the same computation
    rng = np.random.RandomState(0)
    src = torch.from_numpy(rng.rand(total, 16, 3).astype(np.float32))
    tgt = torch.from_numpy(rng.rand(total, 16, 3).astype(np.float32))
    gt = torch.eye(4).repeat(total, 1, 1)
    real = shard.solver.HOUV
    shard.solver.HOUV = _NoNet
    try:
        ans, met = shard.register_pairs(src, tgt, gt, batch_size=batch, kernel=4, num_epochs=1,
                                        solve=_fake_solve)
    finally:
        shard.solver.HOUV = real
    for rank, a, m in got:
        assert a.shape == (total, 4, 4) and m.shape == (total, 3)
        assert np.array_equal(a, ans.numpy()) and np.allclose(m, met.numpy())


def test_range_files_and_combine(tmp_path):
    """test_mult.py:65-81: per-range .npy files merged in index order; a missing range raises."""
    import numpy as np
    import pytest
    import torch
    from houv_b200 import shard
    full = np.arange(8 * 16, dtype=np.float32).reshape(8, 4, 4)
    for i in range(4):
        p = shard.save_range(torch.from_numpy(full[2 * i:2 * i + 2]), str(tmp_path), 2 * i, 2 * i + 2)
        assert p.endswith("%d_%d.npy" % (2 * i, 2 * i + 2))
    merged = shard.combine(str(tmp_path), step=2, num=4)
    assert merged.dtype == np.float32 and np.array_equal(merged, full)
    assert np.array_equal(np.load(str(tmp_path / "results.npy")), full)
    with pytest.raises(FileNotFoundError):
        shard.combine(str(tmp_path), step=2, num=5)


def _fake_pose_optimise(src, tgt, params, kernel, iters, row_pair=None, row_base=None, row_scale=None, **kw):
    # a row's result depends on its pair's clouds, its initial values, angle_base and gradient scale only
    p = row_pair.long()
    c = (src[p].sum((1, 2)) - 2.0 * tgt[p].sum((1, 2))).float()
    m = c + params.sum(1) + 10.0 * row_base.float() + 1000.0 * row_scale
    R = m[:, None, None] * torch.arange(9.0).reshape(1, 3, 3)
    T = m[:, None] + torch.arange(3.0)[None]
    return dict(min_1=m, R=R, T=T)


def _balance_worker(rank, world, port, n_fail, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    shard.init_distributed("gloo")
    from houv_b200 import solver
    solver.pose_optimise = _fake_pose_optimise
    kernel, n = 4, n_fail[rank]
    rng = np.random.RandomState(100 + rank)
    srcF = torch.from_numpy(rng.rand(n, 16, 3).astype(np.float32))
    tgtF = torch.from_numpy(rng.rand(n, 16, 3).astype(np.float32))
    init = torch.from_numpy(rng.rand(max(n, 1) * kernel, 8).astype(np.float32))
    p_rows, pair_rows, base_rows, scale_rows = [], [], [], []
    if n:
        local = torch.arange(n, dtype=torch.int32).repeat_interleave(kernel)
        for b in (1, 2, 3):
            p_rows.append(init[:n * kernel]); pair_rows.append(local)
            base_rows.append(torch.full((n * kernel,), b, dtype=torch.int32))
            scale_rows.append(torch.full((n * kernel,), 1.0 / (n * kernel)))
    got = solver._balanced_retry(srcF, tgtF, p_rows, pair_rows, base_rows, scale_rows, kernel, 1)
    if n:
        want = _fake_pose_optimise(srcF, tgtF, torch.cat(p_rows), kernel, 1, row_pair=torch.cat(pair_rows),
                                   row_base=torch.cat(base_rows), row_scale=torch.cat(scale_rows))
        same = all(torch.equal(got[k], want[k]) for k in ("min_1", "R", "T"))
    else:
        same = got["min_1"].numel() == 0 and tuple(got["R"].shape) == (0, 3, 3)
    q.put((rank, bool(same)))
    torch.distributed.barrier()
    torch.distributed.destroy_process_group()


def _run_balance(n_fail):
    world = len(n_fail)
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_balance_worker, args=(r, world, port, n_fail, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    return got


def test_gloo_world2_balanced_retry_returns_every_rank_its_own_rows():
    """solver._balanced_retry (the retry rows of all ranks dealt evenly, SURVEY 8e / DESIGN 5) with a stand-in for
    the K3 launch: uneven failed-pair counts, a rank with none, nobody failing -- each rank gets exactly the
    results of ITS rows, whichever rank computed them."""
    assert _run_balance([5, 1]) == {0: True, 1: True}
    assert _run_balance([0, 3]) == {0: True, 1: True}
    assert _run_balance([0, 0]) == {0: True, 1: True}


def _uneven_worker(rank, world, port, total, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    shard.init_distributed("gloo")
    l, r = shard.shard_range(total, rank, world)
    ans = torch.arange(l, r, dtype=torch.float32).reshape(-1, 1, 1).expand(-1, 4, 4).contiguous()
    met = torch.arange(l, r, dtype=torch.float32).reshape(-1, 1).expand(-1, 3).contiguous() * 10
    a, m = shard.gather_results(ans, met if rank != 1 else None, world)
    q.put((rank, a.numpy(), m.numpy()))
    torch.distributed.barrier()
    torch.distributed.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_gather_results_with_uneven_and_empty_shards():
    """total % world != 0: shard_range(5, r, 2) = 3 + 2 pairs, shard_range(1, r, 2) = 1 + 0; the
    gather pads to the largest count and trims, rows in pair order on every rank."""
    for total in (5, 1):
        world, port = 2, _free_port()
        ctx = mp.get_context("spawn")
        q = ctx.Queue()
        ps = [ctx.Process(target=_uneven_worker, args=(r, world, port, total, q)) for r in range(world)]
        for p in ps:
            p.start()
        res = [q.get(timeout=120) for _ in ps]
        for p in ps:
            p.join(timeout=60)
            assert p.exitcode == 0
        for rank, a, m in res:
            assert a.shape == (total, 4, 4) and m.shape == (total, 3)
            assert np.array_equal(a[:, 0, 0], np.arange(total, dtype=np.float32))
            l1, r1 = shard.shard_range(total, 1, 2)  # rank 1 passed metrics=None -> zeros
            want = np.arange(total, dtype=np.float32) * 10
            want[l1:r1] = 0
            assert np.array_equal(m[:, 0], want)


def test_check_layout():
    assert shard.check_layout(1200, 8, 50) is None and shard.check_layout(1200, 1, 50) is None
    assert shard.check_layout(1200, 8, 100) is not None  # 150 per rank, batches of 100
    assert shard.check_layout(10, 4, 1) is not None       # 3,3,3,1


def test_results_h5_round_trip(tmp_path):
    """test_mult.py:65-81 / test.py:67-77: range files -> results.h5 ('results' [P,4,4] float32);
    dataset.py:189-233: the MVP reader.  h5py is absent from this image ('unavailable
    on this box'), so this test runs wherever h5py exists and is skipped here; the .npy side of the
    same functions is covered above."""
    import pytest
    h5py = pytest.importorskip("h5py")
    rng = np.random.RandomState(3)
    parts = [rng.rand(5, 4, 4).astype(np.float32) for _ in range(2)]
    for i, p in enumerate(parts):
        shard.save_range(torch.from_numpy(p), str(tmp_path), 5 * i, 5 * i + 5)
    merged = shard.combine(str(tmp_path), step=5, num=2)
    with h5py.File(str(tmp_path / "results.h5"), "r") as f:
        assert f["results"].dtype == np.float32 and np.array_equal(f["results"][:], merged)
    src, tgt, tr = rng.rand(6, 8, 3), rng.rand(6, 8, 3), rng.rand(6, 4, 4)
    with h5py.File(str(tmp_path / "MVP_Test_RG.h5"), "w") as f:
        f.create_dataset("rotated_src", data=src)
        f.create_dataset("rotated_tgt", data=tgt)
        f.create_dataset("transforms", data=tr)
    a, b, c = shard.load_mvp_pairs(str(tmp_path / "MVP_Test_RG.h5"), 2, 5)
    assert a.dtype == np.float32 and np.array_equal(a, src[2:5].astype(np.float32))
    assert np.array_equal(b, tgt[2:5].astype(np.float32)) and np.array_equal(c, tr[2:5].astype(np.float32))
