"""C2 / C5 forward timing (CUDA events, L2 flushed): python tools/c2_time.py"""
import json, os, statistics, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from houv_b200.chamfer import chamfer_3D
dev = torch.device("cuda")
flush = torch.zeros(64 * 1024 * 1024, device=dev)
for name, (B, N, M) in (("C2", (32, 2048, 2048)), ("b16", (16, 2048, 2048)), ("b64", (64, 2048, 2048)), ("b8_1024", (8, 2048, 1024)), ("C5", (64, 16384, 2048))):
    torch.manual_seed(0)
    x1, x2 = torch.rand(B, N, 3, device=dev), torch.rand(B, M, 3, device=dev)
    d1, d2 = torch.empty(B, N, device=dev), torch.empty(B, M, device=dev)
    i1 = torch.empty(B, N, device=dev, dtype=torch.int32); i2 = torch.empty(B, M, device=dev, dtype=torch.int32)
    ts = []
    for i in range(12):
        flush.add_(1.0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); chamfer_3D.forward(x1, x2, d1, d2, i1, i2); e1.record()
        torch.cuda.synchronize()
        if i >= 4: ts.append(e0.elapsed_time(e1))
    print(json.dumps({"case": name, "us": round(statistics.median(ts) * 1e3, 1), "min_us": round(min(ts) * 1e3, 1)}))
