"""K2 timing: python tools/k2_time.py  (HOUV_K2_VARIANT=0/1/2 to force a kernel)"""
import json, os, statistics, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from houv_b200.chamfer import chamfer_3D
dev = torch.device("cuda")
flush = torch.zeros(64 * 1024 * 1024, device=dev)
for name, (B, N, M) in (("C2", (32, 2048, 2048)), ("b6400", (6400, 2048, 2048)), ("b300", (300, 2048, 2048))):
    torch.manual_seed(0)
    x1, x2 = torch.rand(B, N, 3, device=dev), torch.rand(B, M, 3, device=dev)
    d1, d2 = torch.empty(B, N, device=dev), torch.empty(B, M, device=dev)
    i1 = torch.empty(B, N, device=dev, dtype=torch.int32); i2 = torch.empty(B, M, device=dev, dtype=torch.int32)
    chamfer_3D.forward(x1, x2, d1, d2, i1, i2)
    gd1 = torch.full((B, N), 1.0 / N, device=dev); gd2 = torch.full((B, M), 1.0 / M, device=dev)
    g1, g2 = torch.empty_like(x1), torch.empty_like(x2)
    ts = []
    for i in range(10):
        flush.add_(1.0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); chamfer_3D.backward(x1, x2, g1, g2, gd1, gd2, i1, i2, assign=True); e1.record()
        torch.cuda.synchronize()
        if i >= 3: ts.append(e0.elapsed_time(e1))
    us = statistics.median(ts) * 1e3
    print(json.dumps({"case": name, "variant": os.environ.get("HOUV_K2_VARIANT", "default"), "us": round(us, 1),
                      "GBps_algorithmic": round(32.0 * B * (N + M) / us / 1e3, 1)}))
