"""Kernel micro-benchmarks on one B200 (CUDA events, warm-up, L2 flush between
iterations).  Not the contract bench (that is bench.py); this is the tuning tool.

    python tools/microbench.py probe            # FP32 FFMA peak (scalar and packed)
    python tools/microbench.py chamfer          # K1/K2 at BASELINE configs C2 and C5
"""
import ctypes
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from houv_b200 import _lib  # noqa: E402


def timeit(fn, iters=20, warmup=5, flush=None):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.add_(1.0)  # 256 MB write: evicts the 126 MB L2
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2], ts[0]


def fp32_peak(packed, iters=4096):
    lib = _lib.load()
    scratch = torch.zeros(16, device="cuda")
    flop = ctypes.c_double(0)

    def run():
        _lib.check(lib.houv_fp32_probe(scratch.data_ptr(), iters, int(packed), ctypes.byref(flop),
                                       _lib.stream_ptr()), "probe")
    med, best = timeit(run, iters=10, warmup=3)
    return {"packed": bool(packed), "tflops_median": flop.value / med / 1e9,
            "tflops_best": flop.value / best / 1e9, "ms": med}


def cmd_probe():
    for p in (0, 1):
        print(json.dumps(fp32_peak(p)))


def cmd_chamfer():
    from houv_b200.chamfer import chamfer_3DDist, chamfer_3D
    lib = _lib.load()
    flush = torch.zeros(64 * 1024 * 1024, device="cuda")
    peak = fp32_peak(0)["tflops_best"]
    print(json.dumps({"fp32_peak_tflops": peak}))
    for name, (B, N, M) in {"C2": (32, 2048, 2048), "C5": (64, 16384, 2048),
                            "unit_timing": (32, 2000, 1000), "B1": (1, 2048, 2048), "rows592": (592, 2048, 2048),
                            "rows6400": (6400, 2048, 2048)}.items():
        torch.manual_seed(0)
        x1 = torch.rand(B, N, 3, device="cuda")
        x2 = torch.rand(B, M, 3, device="cuda")
        d1 = torch.empty(B, N, device="cuda")
        d2 = torch.empty(B, M, device="cuda")
        i1 = torch.empty(B, N, device="cuda", dtype=torch.int32)
        i2 = torch.empty(B, M, device="cuda", dtype=torch.int32)
        for variant in [int(v) for v in os.environ.get('NN_VARIANTS', '0,1,10').split(',')]:
            lib.houv_set_nn_variant(variant)
            med, best = timeit(lambda: chamfer_3D.forward(x1, x2, d1, d2, i1, i2), flush=flush)
            evals = 2.0 * B * N * M
            print(json.dumps({"case": name, "kernel": "K1 fwd", "variant": variant,
                              "ms_median": round(med, 4), "ms_best": round(best, 4),
                              "Gpts2_per_s": round(evals / med / 1e6, 1),
                              "tflops_8flop": round(8 * evals / med / 1e9, 2),
                              "frac_fp32_peak": round(8 * evals / med / 1e9 / peak, 3)}))
        lib.houv_set_nn_variant(-1)
        gd1, gd2 = torch.rand(B, N, device="cuda"), torch.rand(B, M, device="cuda")
        g1, g2 = torch.zeros_like(x1), torch.zeros_like(x2)
        med, best = timeit(lambda: chamfer_3D.backward(x1, x2, g1, g2, gd1, gd2, i1, i2),
                           flush=flush)
        byts = 32.0 * B * (N + M)
        print(json.dumps({"case": name, "kernel": "K2 bwd", "ms_median": round(med, 4),
                          "GBps": round(byts / med / 1e6, 1)}))
        # what autograd pays: zero-fill + accumulate (the reference's scheme) against the assign entry
        def zero_then_accumulate():
            g1.zero_(); g2.zero_()
            chamfer_3D.backward(x1, x2, g1, g2, gd1, gd2, i1, i2)
        med_z, _ = timeit(zero_then_accumulate, flush=flush)
        med_a, _ = timeit(lambda: chamfer_3D.backward(x1, x2, g1, g2, gd1, gd2, i1, i2, assign=True),
                          flush=flush)
        print(json.dumps({"case": name, "kernel": "K2 bwd", "zero_fill_plus_accumulate_ms": round(med_z, 4),
                          "assign_ms": round(med_a, 4), "assign_GBps": round(byts / med_a / 1e6, 1)}))


def cmd_refext():
    """The reference's own CUDA extension (oracle/_ref/chamfer_3D.so, unmodified, sm_100a) beside
    K1/K2 on the same GPU: forward, backward, and forward as the reference WRAPPER runs it
    (four CPU zero tensors copied to the device per call, dist_chamfer_3D.py:33-42)."""
    from oracle import ref_ext
    from houv_b200.chamfer import chamfer_3D
    ref = ref_ext.load()
    flush = torch.zeros(64 * 1024 * 1024, device="cuda")
    for name, (B, N, M) in {"C2": (32, 2048, 2048), "C5": (64, 16384, 2048),
                            "unit_timing": (32, 2000, 1000),
                            "rows6400": (6400, 2048, 2048)}.items():
        torch.manual_seed(0)
        x1 = torch.rand(B, N, 3, device="cuda")
        x2 = torch.rand(B, M, 3, device="cuda")
        d1 = torch.zeros(B, N, device="cuda"); d2 = torch.zeros(B, M, device="cuda")
        i1 = torch.zeros(B, N, device="cuda", dtype=torch.int32)
        i2 = torch.zeros(B, M, device="cuda", dtype=torch.int32)
        gd1, gd2 = torch.rand(B, N, device="cuda"), torch.rand(B, M, device="cuda")
        g1, g2 = torch.zeros_like(x1), torch.zeros_like(x2)
        n_it = 5 if B > 1000 else 20

        def ref_wrapper_fwd():
            a = torch.zeros(B, N).to("cuda"); b = torch.zeros(B, M).to("cuda")
            c = torch.zeros(B, N).type(torch.IntTensor).to("cuda")
            d = torch.zeros(B, M).type(torch.IntTensor).to("cuda")
            ref.forward(x1, x2, a, b, c, d)
        rf, _ = timeit(lambda: ref.forward(x1, x2, d1, d2, i1, i2), iters=n_it, flush=flush)
        rb, _ = timeit(lambda: ref.backward(x1, x2, g1, g2, gd1, gd2, i1, i2), iters=n_it, flush=flush)
        rw, _ = timeit(ref_wrapper_fwd, iters=n_it, flush=flush)
        of, _ = timeit(lambda: chamfer_3D.forward(x1, x2, d1, d2, i1, i2), iters=n_it, flush=flush)
        ob, _ = timeit(lambda: chamfer_3D.backward(x1, x2, g1, g2, gd1, gd2, i1, i2), iters=n_it, flush=flush)
        print(json.dumps({"case": name, "ref_fwd_ms": round(rf, 4), "ref_bwd_ms": round(rb, 4),
                          "ref_wrapper_fwd_ms": round(rw, 4), "ours_fwd_ms": round(of, 4),
                          "ours_bwd_ms": round(ob, 4), "fwd_speedup": round(rf / of, 2),
                          "bwd_speedup": round(rb / ob, 2),
                          "fwd_bwd_speedup": round((rf + rb) / (of + ob), 2),
                          "ref_Gpts2_per_s": round(2.0 * B * N * M / rf / 1e6, 1),
                          "ours_Gpts2_per_s": round(2.0 * B * N * M / of / 1e6, 1)}))


def cmd_solve():
    """train_utils.getPredict_angle shape: P pairs x 64 starts x `iters` iterations
    at lr 0.1 with the 3-D term only; one-metric sweep (default) vs the four-metric sweep."""
    from houv_b200 import solver, synth, train_utils as tu
    P = int(sys.argv[2]) if len(sys.argv) > 2 else 37
    iters = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    s, t, _ = synth.make_pairs(0, P)
    src, tgt = torch.from_numpy(s).cuda(), torch.from_numpy(t).cuda()
    np.random.seed(0)
    init = torch.from_numpy(tu.initial_variables(P * 64)).cuda()

    def run():
        solver.pose_optimise(src, tgt, init.clone(), 64, iters, lr=0.1, wview=0.0, tran_scale=1.0,
                             tran_offset=0.0)
    for v in (12, 13, 15):
        _lib.load().houv_set_pose_variant(v)
        med, best = timeit(run, iters=3, warmup=1)
        print(json.dumps({"variant": v, "ms": round(med, 2),
                          "us_per_row_iter_per_sm": round(med * 1e3 / (P * 64 * iters) * 148, 1),
                          "pairs_per_s_500its": round(P / (med * 500 / iters) * 1e3, 2)}))
    _lib.load().houv_set_pose_variant(-1)


def cmd_refhouv():
    """The reference's HOUV loop as the reference runs it on a GPU, for a GPU-side baseline of
    pairs/s: restated predict_model loop (houv.py:106-138: torch HOUV.forward, Predict_loss =
    4 chamfer calls + 8 top-k sorts, autograd backward, torch Adam) over the reference's OWN CUDA
    extension (oracle/_ref/chamfer_3D.so) with the reference wrapper's allocation pattern (CPU
    zeros copied to the device per call, dist_chamfer_3D.py:33-42,56-60).  B pairs x 64 starts,
    `n` timed iterations; pairs/s = B / (t_iter * 200), retries not added."""
    from torch.autograd import Function
    from houv_b200 import loss as L, solver, synth
    from oracle import ref_ext
    ref = ref_ext.load()
    faithful = os.environ.get("REF_ALLOC", "cpu") == "cpu"

    def zeros(shape, dev, dtype=torch.float32):
        if faithful:
            z = torch.zeros(*shape)
            return (z.type(torch.IntTensor) if dtype == torch.int32 else z).to(dev)
        return torch.zeros(*shape, device=dev, dtype=dtype)

    class RefFn(Function):  # dist_chamfer_3D.py:26-64
        @staticmethod
        def forward(ctx, xyz1, xyz2):
            B, n, _ = xyz1.shape
            m = xyz2.shape[1]
            dev = xyz1.device
            d1, d2 = zeros((B, n), dev), zeros((B, m), dev)
            i1, i2 = zeros((B, n), dev, torch.int32), zeros((B, m), dev, torch.int32)
            ref.forward(xyz1, xyz2, d1, d2, i1, i2)
            ctx.save_for_backward(xyz1, xyz2, i1, i2)
            return d1, d2, i1, i2

        @staticmethod
        def backward(ctx, g1, g2, gi1, gi2):
            xyz1, xyz2, i1, i2 = ctx.saved_tensors
            gx1, gx2 = zeros(tuple(xyz1.shape), xyz1.device), zeros(tuple(xyz2.shape), xyz2.device)
            ref.backward(xyz1, xyz2, gx1, gx2, g1.contiguous(), g2.contiguous(), i1, i2)
            return gx1, gx2

    class RefCd(torch.nn.Module):
        def forward(self, a, b):
            return RefFn.apply(a.contiguous(), b.contiguous())

    B = int(sys.argv[2]) if len(sys.argv) > 2 else 8
    n_it = int(sys.argv[3]) if len(sys.argv) > 3 else 5
    s, t, _ = synth.make_pairs(0, B)
    src = torch.from_numpy(s).cuda().unsqueeze(1).expand(-1, 64, -1, -1).reshape(-1, 2048, 3)
    tgt = torch.from_numpy(t).cuda().unsqueeze(1).expand(-1, 64, -1, -1).reshape(-1, 2048, 3)
    old_cd = L.cd
    L.cd = RefCd
    try:
        net = solver.HOUV(B * 64, 0)
        net.reset_weight(B * 64, 0)
        opt = torch.optim.Adam(net.parameters(), lr=0.01)

        def it():
            opt.zero_grad()
            src_temp, R, T = net(src)
            rmse, min_1 = L.Predict_loss(src_temp, tgt)
            rmse.mean().backward()
            opt.step()
        it(); it()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n_it):
            it()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n_it
    finally:
        L.cd = old_cd
    print(json.dumps({"what": "reference HOUV loop on the reference CUDA ext", "pairs": B,
                      "alloc": "cpu zeros + H2D (reference wrapper)" if faithful else "device zeros",
                      "ms_per_iteration": round(ms, 2),
                      "pairs_per_s_200its": round(B / (ms * 200 / 1e3), 4)}))


def cmd_icp():
    """BASELINE config 3: the ICP block of train_ICP.py over 1200 synthetic pairs (2048 points,
    threshold 0.02, <= 500 iterations, fixed initial rotation) in one launch; the CPU float64
    restatement timed on a few pairs beside it."""
    import time
    from houv_b200 import icp, synth
    from oracle import houv as oh
    P = int(sys.argv[2]) if len(sys.argv) > 2 else 1200
    s, t, tr = synth.make_pairs(0, P)
    src, tgt, gt = (torch.from_numpy(a).cuda() for a in (s, t, tr))
    med, best = timeit(lambda: icp.icp_batch(src, tgt, gt), iters=3, warmup=1)
    reg = icp.registration_icp(src, tgt, icp.THRESHOLD, icp.TRANS_INIT, icp.MAX_ITERATION)
    ans, r_err, t_err, mse = icp.icp_batch(src, tgt, gt)
    n_cpu = 3
    t0 = time.perf_counter()
    for i in range(n_cpu):
        oh.icp_point_to_point(s[i].astype(np.float64), t[i].astype(np.float64), 0.02,
                              icp.TRANS_INIT.astype(np.float64), 500)
    cpu = (time.perf_counter() - t0) / n_cpu
    print(json.dumps({"kernel": "ICP (houv_icp)", "pairs": P, "ms": round(med, 2),
                      "pairs_per_s": round(P / med * 1e3, 1),
                      "iterations_mean": float(reg.iterations.mean()),
                      "iterations_max": float(reg.iterations.max()),
                      "fitness_mean": float(reg.fitness.mean()),
                      "rot_err_deg_mean": float(r_err.mean()), "trans_err_mean": float(t_err.mean()),
                      "cpu_oracle_s_per_pair": round(cpu, 3),
                      "cpu_oracle_pairs_per_s": round(1.0 / cpu, 2)}))


def cmd_idam():
    """K6 (houv_idam_step) beside the same iteration tail in torch ops on the same GPU (the oracle's
    restatement of idam.py:317-342 run on CUDA tensors, i.e. the reference's own execution scheme:
    ~15 launches plus a per-sample torch.svd loop with a host-synchronising `if`)."""
    from houv_b200.kabsch import idam_step
    from oracle import houv as oh
    for B, N in ((32, 128), (32, 1024), (256, 128)):
        torch.manual_seed(0)
        src, tgt = torch.randn(B, 3, N).cuda() * 0.3, torch.randn(B, 3, N).cuda() * 0.3
        corr = torch.randint(0, N, (B, N)).cuda()
        logits = torch.randn(B, N).cuda()
        R0, t0 = torch.eye(3).repeat(B, 1, 1).cuda(), torch.zeros(B, 3).cuda()
        med, _ = timeit(lambda: idam_step(src, tgt, corr, logits, R0, t0), iters=20)
        head = oh.SVDHead().cuda()
        ref, _ = timeit(lambda: oh.idam_step(src, tgt, corr, logits, R0, t0, head=head), iters=3, warmup=1)
        print(json.dumps({"kernel": "K6 idam_step", "B": B, "N": N, "ms": round(med, 4),
                          "torch_ops_same_gpu_ms": round(ref, 3), "ratio": round(ref / med, 1)}))


def cmd_emd():
    """K7 (houv_emd_forward) beside the reference's own EMD extension (oracle/_ref/emd.so, unmodified,
    sm_100a) on the same GPU: training setting (eps 0.005, 50 iterations), evaluation setting
    (pcn.py:169: eps 0.004, 3000 iterations) and emd_module.py:95-108 `test_emd` (20 x 8192, eps 0.05, 3000)."""
    from houv_b200.emd import emd_forward
    from oracle import emd as oe
    for B, n, eps, iters in ((32, 2048, 0.005, 50), (32, 2048, 0.004, 3000), (148, 2048, 0.004, 3000),
                             (20, 8192, 0.05, 3000)):
        torch.manual_seed(0)
        x1, x2 = torch.rand(B, n, 3).cuda(), torch.rand(B, n, 3).cuda()
        med, _ = timeit(lambda: emd_forward(x1, x2, eps, iters), iters=3, warmup=1)
        d, a, used = emd_forward(x1, x2, eps, iters, want_iters=True)
        line = {"kernel": "K7 emd_forward", "B": B, "n": n, "eps": eps, "iters": iters, "ms": round(med, 3),
                "emd": round(float(torch.sqrt(d).mean()), 6), "iters_used_max": int(used.max()),
                "unique_frac": round(float(np.mean([len(set(r)) / n for r in a.cpu().numpy()])), 4)}
        if oe.ref_available():
            ref, _ = timeit(lambda: oe.ref_emd_forward(x1, x2, eps, iters), iters=2, warmup=1)
            rd, ra = oe.ref_emd_forward(x1, x2, eps, iters)
            line.update({"reference_ext_ms": round(ref, 2), "ratio": round(ref / med, 1),
                         "reference_emd": round(float(torch.sqrt(rd).mean()), 6),
                         "same_assignment_frac": round(float((ra == a).float().mean()), 4)})
        print(json.dumps(line), flush=True)


def cmd_pose():
    """K3 at the BASELINE shape: P pairs x 64 starts x `iters` iterations, 2048 points."""
    import numpy as np
    from houv_b200 import solver, synth
    P = int(sys.argv[2]) if len(sys.argv) > 2 else 37
    iters = int(sys.argv[3]) if len(sys.argv) > 3 else 200
    s, t, _ = synth.make_pairs(0, P)
    src, tgt = torch.from_numpy(s).cuda(), torch.from_numpy(t).cuda()
    init = torch.from_numpy(solver.initial_parameters(P * 64)).cuda()
    peak = fp32_peak(0)["tflops_best"]

    def run():
        solver.pose_optimise(src, tgt, init.clone(), 64, iters)
    variants = [int(v) for v in sys.argv[4].split(",")] if len(sys.argv) > 4 else [-1]
    ref = None
    for v in variants:
        _lib.load().houv_set_pose_variant(v)
        p0 = init.clone()
        o = solver.pose_optimise(src, tgt, p0, 64, 3)
        same = None if ref is None else bool(torch.equal(ref, p0))
        ref = p0 if ref is None else ref
        med, best = timeit(run, iters=3, warmup=1)
        print(json.dumps({"variant": v, "ms": round(med, 2), "identical_to_first": same,
                          "pairs_per_s": round(P / med * 1e3, 2)}))
    _lib.load().houv_set_pose_variant(-1)
    med, best = timeit(run, iters=3, warmup=1)
    evals = 2.0 * 2048 * 2048 * P * 64 * iters  # directed point pairs, each serving 4 metrics
    print(json.dumps({"kernel": "K3 pose_optimise", "pairs": P, "iters": iters,
                      "ms": round(med, 2), "pairs_per_s_base0": round(P / med * 1e3, 3),
                      "us_per_row_iter": round(med * 1e3 / (P * 64 * iters) * 148 * 2, 2),
                      "tflops_executed_26flop": round(13 * evals / med / 1e9, 2),
                      "tflops_ref_equiv_64flop": round(32 * evals / med / 1e9, 2),
                      "frac_fp32_peak_executed": round(13 * evals / med / 1e9 / peak, 3),
                      "fp32_peak": round(peak, 2)}))


if __name__ == "__main__":
    {"probe": cmd_probe, "chamfer": cmd_chamfer, "pose": cmd_pose, "refext": cmd_refext, "icp": cmd_icp, "solve": cmd_solve, "refhouv": cmd_refhouv, "idam": cmd_idam, "emd": cmd_emd}[sys.argv[1] if len(sys.argv) > 1 else "probe"]()
