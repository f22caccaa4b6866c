"""Generates the committed golden fixtures.  Runs ONLY in the dev container,
where /root/reference exists; the GPU box consumes the .npz files.

    python tests/golden/make_golden.py            # all fixtures
    python tests/golden/make_golden.py chamfer    # one family

chamfer_unit.npz
    the reference's own unit test case (utils/metrics/CD/unit_test.py:14-33:
    rand(4,100,3) vs rand(4,200,3)) with
      py_*   = outputs of the reference's chamfer_python.distChamfer, imported
               as-is from /root/reference/utils/metrics/CD/chamfer_python.py
      ref_*  = outputs of the reference's CUDA kernel SOURCE compiled for the
               host (oracle/_ref/libref_chamfer.so, see oracle/Makefile)
chamfer_cases.npz
    ragged / tie / tile-boundary cases, ref_* outputs as above.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
REF = "/root/reference"


def chamfer_cases():
    """name -> (xyz1, xyz2); shared with the tests through the fixture file."""
    rng = np.random.RandomState(1234)
    f32 = np.float32
    cases = {}
    cases["ragged_7x13"] = (rng.rand(3, 7, 3), rng.rand(3, 13, 3))
    cases["one_target"] = (rng.rand(2, 33, 3), rng.rand(2, 1, 3))
    cases["tile_edge_511_513"] = (rng.rand(2, 511, 3), rng.rand(2, 513, 3))
    cases["odd_1000_1537"] = (rng.rand(2, 1000, 3), rng.rand(2, 1537, 3))
    # exact ties: duplicated targets (first index must win), also across the
    # 512-tile boundary of the reference and the 8-wide groups of our kernel
    t = rng.rand(1, 600, 3)
    t[0, 300] = t[0, 5]
    t[0, 517] = t[0, 5]
    t[0, 599] = t[0, 510]
    q = np.concatenate([t[:, [5, 510, 7]] + 1e-3, rng.rand(1, 40, 3)], axis=1)
    cases["dup_targets"] = (q, t)
    # grid-quantised coordinates: many exactly equal distances
    cases["grid_ties"] = (rng.randint(0, 4, (2, 200, 3)) / 4.0, rng.randint(0, 4, (2, 300, 3)) / 4.0)
    # identical clouds: d == 0 everywhere, idx == own index unless duplicated
    same = rng.rand(1, 257, 3)
    cases["identical"] = (same, same.copy())
    # negative / larger-magnitude coordinates (MVP clouds are centred)
    cases["centred"] = (rng.randn(2, 300, 3) * 0.3, rng.randn(2, 260, 3) * 0.3)
    return {k: (a.astype(f32), b.astype(f32)) for k, (a, b) in cases.items()}


def make_chamfer():
    import torch
    from oracle import chamfer as oc
    sys.path.insert(0, os.path.join(REF, "utils/metrics/CD"))
    import chamfer_python  # the reference's own file, unmodified

    torch.manual_seed(0)
    p1 = torch.rand(4, 100, 3)
    p2 = torch.rand(4, 200, 3)
    py = chamfer_python.distChamfer(p1, p2)
    x1, x2 = p1.numpy(), p2.numpy()
    ref = oc.ref_chamfer_forward(x1, x2)
    rng = np.random.RandomState(7)
    gd1 = rng.randn(4, 100).astype(np.float32)
    gd2 = rng.randn(4, 200).astype(np.float32)
    g1, g2 = oc.ref_chamfer_backward(x1, x2, gd1, gd2, ref[2], ref[3])
    np.savez_compressed(
        os.path.join(HERE, "chamfer_unit.npz"), xyz1=x1, xyz2=x2,
        py_dist1=py[0].numpy(), py_dist2=py[1].numpy(), py_idx1=py[2].numpy(),
        py_idx2=py[3].numpy(), ref_dist1=ref[0], ref_dist2=ref[1], ref_idx1=ref[2],
        ref_idx2=ref[3], graddist1=gd1, graddist2=gd2, ref_gradxyz1=g1, ref_gradxyz2=g2)

    out = {}
    for name, (a, b) in chamfer_cases().items():
        d1, d2, i1, i2 = oc.ref_chamfer_forward(a, b)
        rs = np.random.RandomState(len(name))
        gd1 = rs.randn(*d1.shape).astype(np.float32)
        gd2 = rs.randn(*d2.shape).astype(np.float32)
        g1, g2 = oc.ref_chamfer_backward(a, b, gd1, gd2, i1, i2)
        for k, v in dict(xyz1=a, xyz2=b, dist1=d1, dist2=d2, idx1=i1, idx2=i2, gd1=gd1, gd2=gd2,
                         g1=g1, g2=g2).items():
            out[name + "/" + k] = v
    np.savez_compressed(os.path.join(HERE, "chamfer_cases.npz"), **out)
    print("chamfer fixtures written")


def make_houv():
    """houv_ref.npz: outputs of the reference's OWN python files (houv.py,
    model_utils_completion.py, train_utils.py, loaded unmodified by ref_loader) on small
    synthetic pairs: reset_weight draws, first-forward loss terms, predict_model after 3
    iterations at angle_base 0 and 2, solve_model (10 iterations) and the three metrics."""
    import torch
    import ref_loader
    from houv_b200 import synth
    m = ref_loader.load()
    H, tu = m["houv"], m["tu"]
    s, t, tr = synth.make_pairs(100, 2, n_points=128)
    src, tgt, gt = torch.from_numpy(s), torch.from_numpy(t), torch.from_numpy(tr)
    out = dict(src=s, tgt=t, transform=tr)
    net = H.HOUV(64, 0)
    net.reset_weight(64, 0, seed=2021)
    for k in ("V_c", "angle_c", "tran_c", "tran_s"):
        out["init_" + k] = getattr(net, k).detach().numpy().copy()
    rep = lambda x: x.unsqueeze(1).expand(-1, 32, -1, -1).reshape(-1, 128, 3)
    src_temp, R, T = net(rep(src))
    loss, min_1 = H.Predict_loss(src_temp, rep(tgt))
    out.update(fwd_src_temp=src_temp.detach().numpy(), fwd_R=R.detach().numpy(),
               fwd_T=T.detach().numpy(), fwd_loss=loss.detach().numpy(),
               fwd_min_1=min_1.detach().numpy())
    for base in (0, 2):
        l, R, T = H.predict_model(H.HOUV(64, 0), src, tgt, kernel=32, num_epochs=3, angle_base=base)
        out["pm%d_loss" % base] = l.detach().numpy()
        out["pm%d_R" % base] = R.detach().numpy()
        out["pm%d_T" % base] = T.detach().numpy()
    r_err, t_err, ans = H.solve_model(H.HOUV(64, 0), src, tgt, gt, kernel=32, num_epochs=10)
    out.update(solve_ans=ans.detach().numpy(), solve_r_err=r_err.detach().numpy(),
               solve_t_err=t_err.detach().numpy(),
               solve_mse=tu.rmse_loss(src, ans.detach(), gt).numpy())
    np.savez_compressed(os.path.join(HERE, "houv_ref.npz"), **out)
    print("houv fixtures written")


def make_solve():
    """solve_ref.npz: the reference's OWN train_utils.getPredict_angle / solve
    (registration/train_utils.py:359-572, unmodified) on small synthetic pairs with NumPy seeded
    just before each call.  solve() passes the literal 500 to getPredict_angle (:488, :503); the
    fixture intercepts that inner call to shorten the horizon -- solve's own code runs as is."""
    import torch
    import ref_loader
    from houv_b200 import synth
    tu = ref_loader.load()["tu"]
    s, t, tr = synth.make_pairs(200, 3, n_points=128)
    src, tgt, gt = torch.from_numpy(s), torch.from_numpy(t), torch.from_numpy(tr)
    out = dict(src=s, tgt=t, transform=tr)
    for base, its in ((0, 1), (0, 4), (2, 4)):
        np.random.seed(77)
        l, R, T, ts = tu.getPredict_angle(src, tgt, kernel=16, num_epochs=its, angle_base=base)
        k = "gpa_b%d_i%d_" % (base, its)
        out[k + "min_1"], out[k + "R"] = l.detach().numpy(), R.detach().numpy()
        out[k + "T"], out[k + "tran_s"] = T.detach().numpy(), ts.detach().numpy()
    orig = tu.getPredict_angle
    tu.getPredict_angle = lambda *a, **k: orig(*a, **dict(k, num_epochs=6))
    try:
        np.random.seed(78)
        # prefix='test' is how test.py:64 calls it (the 'train' branch plots pairs above 70 deg
        # through matplotlib, train_utils.py:556-562)
        ans = tu.solve(src, tgt, gt, kernel=16, prefix='test')
    finally:
        tu.getPredict_angle = orig
    out.update(solve_ans=ans.detach().numpy(),
               solve_r_err=tu.rotation_error(ans[:, :3, :3].detach(), gt[:, :3, :3]).numpy(),
               solve_t_err=tu.translation_error(ans[:, :3, 3].detach(), gt[:, :3, 3]).numpy())
    np.savez_compressed(os.path.join(HERE, "solve_ref.npz"), **out)
    print("solve fixtures written")


def make_solve_full():
    """solve_full.npz: the reference's OWN train_utils.getPredict_angle (train_utils.py:359-452,
    unmodified) at the shape test.py:64 runs it -- 2048 points x 64 starts x 500 Adam iterations at
    lr 0.1, float64 leaves -- on two well-conditioned pairs (synth.make_overlap_pair), NumPy seeded
    just before the call.  min_1 of every forward is recorded by wrapping (not editing) the module's
    calc_cd_percent name; metrics.cd as in make_houv_full."""
    import time
    import torch
    import ref_loader
    from houv_b200 import synth
    from oracle import chamfer as oc
    oc.set_threads(os.cpu_count() or 8)
    ref_loader.FAST_CD = True
    tu = ref_loader.load()["tu"]
    pairs = [synth.make_overlap_pair(10 + i) for i in range(2)]
    s, t, tr = (np.stack([q[k] for q in pairs]) for k in range(3))
    src, tgt = torch.from_numpy(s), torch.from_numpy(t)
    trace = []
    orig = tu.calc_cd_percent

    def rec(a, b, *k, **kw):
        cd_t, cd_p = orig(a, b, *k, **kw)
        trace.append(torch.min(cd_t, cd_p).detach().numpy().copy())
        return cd_t, cd_p
    tu.calc_cd_percent = rec
    try:
        np.random.seed(91)
        t0 = time.time()
        l, R, T, ts = tu.getPredict_angle(src, tgt, kernel=64, num_epochs=500, angle_base=0)
        print("getPredict_angle 2 pairs x 64 x 500: %.0f s" % (time.time() - t0), flush=True)
    finally:
        tu.calc_cd_percent = orig
    np.savez_compressed(os.path.join(HERE, "solve_full.npz"), src=s, tgt=t, transform=tr, seed=np.int32(91),
                        trace=np.stack(trace).astype(np.float32), min_1=l.detach().numpy(),
                        R=R.detach().numpy(), T=T.detach().numpy(), tran_s=ts.detach().numpy())
    print("solve_full written", np.stack(trace).shape, l.detach().numpy().min(1))


def make_calccd():
    """calccd_ref.npz: the reference's own calc_cd (model_utils_completion.py:69-80, identical to
    completion/model_utils.py:69-80) with calc_f1, and the gradient of mean(cd_p)+mean(cd_t) with
    respect to `output`, on a completion-shaped input (coarse output against a denser gt whose
    points partly coincide with it, so the F-score threshold 1e-4 is exercised)."""
    import torch
    import ref_loader
    muc = ref_loader.load()["muc"]
    rng = np.random.RandomState(11)
    gt = rng.rand(3, 500, 3).astype(np.float32)
    out = gt[:, rng.permutation(500)[:200]] + (rng.rand(3, 200, 3) < 0.5) * \
        rng.normal(0, 0.02, (3, 200, 3))
    out = out.astype(np.float32)
    o = torch.from_numpy(out).requires_grad_()
    cd_p, cd_t, f1 = muc.calc_cd(o, torch.from_numpy(gt), calc_f1=True)
    (cd_p.mean() + cd_t.mean()).backward()
    np.savez_compressed(os.path.join(HERE, "calccd_ref.npz"), output=out, gt=gt,
                        cd_p=cd_p.detach().numpy(), cd_t=cd_t.detach().numpy(), f1=f1.numpy(),
                        grad_output=o.grad.numpy())
    print("calc_cd fixtures written")


def make_small():
    """small_ref.npz: the remaining reference functions of the path that had no fixture --
    calc_cd_percent_aligned (model_utils_completion.py:103-116) at percent 0.5 and 1, and
    nearest_neighbor (model_utils.py:33-37, source executed as-is) on [3, points] inputs."""
    import torch
    import ref_loader
    muc = ref_loader.load()["muc"]
    rng = np.random.RandomState(21)
    out = {}
    o = (rng.rand(3, 96, 3) - 0.5).astype(np.float32)
    g = (rng.rand(3, 96, 3) - 0.5).astype(np.float32)
    out["al_output"], out["al_gt"] = o, g
    for tag, pc in (("50", 0.5), ("100", 1)):
        r = muc.calc_cd_percent_aligned(torch.from_numpy(o), torch.from_numpy(g), percent=pc)
        for k, v in zip(("cd_p", "cd_t", "idx1", "idx2", "idxx1", "idxx2"), r):
            out["al%s_%s" % (tag, k)] = v.numpy()
    nn = ref_loader.load_function("nearest_neighbor")
    a = (rng.rand(3, 150) - 0.5).astype(np.float32)
    b = (rng.rand(3, 170) - 0.5).astype(np.float32)
    d, i = nn(torch.from_numpy(a), torch.from_numpy(b))
    out.update(nn_src=a, nn_dst=b, nn_dist=d.numpy(), nn_idx=i.numpy())
    np.savez_compressed(os.path.join(HERE, "small_ref.npz"), **out)
    print("small fixtures written", d.shape, i.dtype)


def make_svd():
    """svdhead_ref.npz: the reference's SVDHead class (model_utils.py:213-255, source read
    and executed as-is) on noisy rigid correspondences, with and without weights, including
    a mirrored sample that exercises the det<0 branch."""
    import torch
    import ref_loader

    class Args:
        use_fpfh = False
        descriptor_size = 8
    head = ref_loader.load_svdhead()(Args())
    torch.manual_seed(11)
    B, N = 6, 200
    src = torch.randn(B, 3, N) * 0.3
    Rg = torch.linalg.qr(torch.randn(B, 3, 3))[0]
    Rg = Rg * torch.sign(torch.det(Rg)).reshape(B, 1, 1)
    corr = Rg @ src + torch.randn(B, 3, 1) * 0.2 + 0.005 * torch.randn(B, 3, N)
    corr[5] = corr[5] * torch.tensor([1.0, 1.0, -1.0]).reshape(3, 1)  # reflection -> det fix
    w = torch.rand(B, 1, N)
    R0, t0 = head(src, corr)
    R1, t1 = head(src, corr, w)
    np.savez_compressed(os.path.join(HERE, "svdhead_ref.npz"), src=src.numpy(), corr=corr.numpy(),
                        weights=w.numpy(), R=R0.numpy(), t=t0.numpy(), Rw=R1.numpy(),
                        tw=t1.numpy())
    print("svd fixtures written")


def make_idam():
    """idam_ref.npz: three chained refinement steps (idam.py:317-342) with the reference's own IDAM
    SVDHead class (models/idam.py:150-189, source read and executed as-is); correspondences are the
    true ones with a fraction scrambled, logits favour the true ones; even and odd N (median rule)."""
    import torch
    import ref_loader
    sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
    from oracle import houv as oh

    class Args:
        use_fpfh = False
        descriptor_size = 8
    head = ref_loader.load_svdhead("models/idam.py")(Args())
    out = {}
    for tag, (B, N, M) in {"even": (5, 128, 160), "odd": (3, 77, 77)}.items():
        torch.manual_seed(N)
        src = torch.randn(B, 3, N) * 0.3
        Rg = torch.linalg.qr(torch.randn(B, 3, 3))[0]
        Rg = Rg * torch.sign(torch.det(Rg)).reshape(B, 1, 1)
        tgt = torch.randn(B, 3, M) * 0.3
        perm = torch.stack([torch.randperm(M)[:N] for _ in range(B)])
        moved = Rg @ src + torch.randn(B, 3, 1) * 0.1 + 0.003 * torch.randn(B, 3, N)
        tgt[torch.arange(B)[:, None], :, perm] = moved.transpose(1, 2)
        R = torch.eye(3).repeat(B, 1, 1)
        t = torch.zeros(B, 3)
        cur = src
        for it in range(3):
            corr = perm.clone()
            bad = torch.rand(B, N) < 0.3
            corr[bad] = torch.randint(0, M, (int(bad.sum()),))
            logits = torch.randn(B, N) + 2.0 * (~bad).float()
            if it == 1:
                logits[:, : N // 3] = 0.25  # ties at and around the median
            out["%s_src%d" % (tag, it)] = cur.numpy()
            out["%s_corr%d" % (tag, it)] = corr.numpy()
            out["%s_logits%d" % (tag, it)] = logits.numpy()
            out["%s_Rin%d" % (tag, it)] = R.numpy()
            out["%s_tin%d" % (tag, it)] = t.numpy()
            cur, R, t, Rab, tab, w = oh.idam_step(cur, tgt, corr, logits, R, t, head=head)
            for k, v in dict(out=cur, R=R, t=t, Rab=Rab, tab=tab, w=w).items():
                out["%s_%s%d" % (tag, k, it)] = v.numpy()
        out[tag + "_tgt"] = tgt.numpy()
    np.savez_compressed(os.path.join(HERE, "idam_ref.npz"), **out)
    print("idam fixtures written")


def make_houv_full(n_easy=2, hard=(3,), tag="houv_full", reversed_atomics=False):
    """houv_full.npz: the reference's OWN solve_model / predict_model (houv.py:106-206, loaded
    unmodified) at BASELINE config 4's shape -- 2048 points x 64 starts x 200 iterations -- on
    `n_easy` well-conditioned pairs (synth.make_overlap_pair) and the synth.make_pair pairs in
    `hard` (partial overlap; whichever of them ends above 0.030 takes the reference through its
    retry cascade at angle_base 1..3).  metrics.cd is served by the thread-parallel C restatement,
    re-checked against the reference kernel source every 40th call (ref_loader.FAST_CD).
    Recorded by wrapping (not editing) the module's Predict_loss / predict_model names:
      pm{c}_base, pm{c}_npairs           which predict_model call (c = 0: all pairs, base 0;
                                          c = 1..3: the retry sub-batch at angle_base c)
      pm{c}_trace[its, rows] float32     min_1 of every forward (houv.py:123)
      pm{c}_loss[its, rows]              6*min_1 + views of every forward
      pm{c}_min1[B,K], pm{c}_R, pm{c}_T  what predict_model returned (:134-138)
      ans, r_err, t_err, mse             solve_model's result and the three metrics."""
    import time
    import torch
    import ref_loader
    from houv_b200 import synth
    from oracle import chamfer as oc
    oc.set_threads(os.cpu_count() or 8)
    ref_loader.FAST_CD = True
    ref_loader.BACKWARD_REVERSED = reversed_atomics
    m = ref_loader.load()
    H, tu = m["houv"], m["tu"]
    pairs = [synth.make_overlap_pair(i) for i in range(n_easy)] + [synth.make_pair(p) for p in hard]
    s, t, tr = (np.stack([q[k] for q in pairs]) for k in range(3))
    src, tgt, gt = torch.from_numpy(s), torch.from_numpy(t), torch.from_numpy(tr)
    out = dict(src=s, tgt=t, transform=tr)
    calls, cur = [], {}
    pl, pm = H.Predict_loss, H.predict_model

    def rec_loss(a, b, *k, **kw):
        loss, min_1 = pl(a, b, *k, **kw)
        cur["trace"].append(min_1.detach().numpy().copy())
        cur["loss"].append(loss.detach().numpy().copy())
        return loss, min_1

    def rec_pm(net, a, b, *k, **kw):
        cur.clear()
        cur.update(trace=[], loss=[], base=kw.get("angle_base", 0), npairs=a.shape[0])
        t0 = time.time()
        l, R, T = pm(net, a, b, *k, **kw)
        print("  predict_model base %d, %d pairs: %.0f s" % (cur["base"], a.shape[0], time.time() - t0),
              flush=True)
        calls.append(dict(cur, min1=l.detach().numpy().copy(), R=R.detach().numpy().copy(),
                          T=T.detach().numpy().copy()))
        return l, R, T

    H.Predict_loss, H.predict_model = rec_loss, rec_pm
    try:
        r_err, t_err, ans = H.solve_model(H.HOUV(64, 0), src, tgt, gt, kernel=64, num_epochs=200)
    finally:
        H.Predict_loss, H.predict_model = pl, pm
    for c, d in enumerate(calls):
        out["pm%d_base" % c], out["pm%d_npairs" % c] = np.int32(d["base"]), np.int32(d["npairs"])
        out["pm%d_trace" % c] = np.stack(d["trace"]).astype(np.float32)
        out["pm%d_loss" % c] = np.stack(d["loss"]).astype(np.float32)
        out["pm%d_min1" % c], out["pm%d_R" % c], out["pm%d_T" % c] = d["min1"], d["R"], d["T"]
    out.update(n_calls=np.int32(len(calls)), ans=ans.detach().numpy(), r_err=r_err.detach().numpy(),
               t_err=t_err.detach().numpy(), mse=tu.rmse_loss(src, ans.detach(), gt).numpy())
    np.savez_compressed(os.path.join(HERE, tag + ".npz"), **out)
    print(tag, "written: r_err", out["r_err"], "t_err", out["t_err"], "calls", len(calls))


FAMILIES = {"small": make_small, "solve_full": make_solve_full, "houv_full": make_houv_full,
            # the same run with the reference kernel's gradient adds in the opposite order: the
            # reference against ITSELF (ref_loader.BACKWARD_REVERSED)
            "houv_full_alt": lambda: make_houv_full(tag="houv_full_alt", reversed_atomics=True),
            "houv_full_retry": lambda: make_houv_full(0, (6, 26, 30), "houv_full_retry"),
            "idam": make_idam, "chamfer": make_chamfer, "houv": make_houv, "svd": make_svd, "solve": make_solve, "calccd": make_calccd}

if __name__ == "__main__":
    want = sys.argv[1:] or list(FAMILIES)
    for w in want:
        FAMILIES[w]()
