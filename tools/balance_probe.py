"""torchrun script: every rank registers its own shard of the C4 set (150 pairs from rank*150) with the retry
rows kept local and with them dealt evenly over the ranks (solver.solve_model_batches balance=False / True);
prints per-rank times and whether the poses are identical bit for bit.
usage: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/balance_probe.py [pairs]"""
import os, sys, json
import torch
import torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from houv_b200 import solver, synth
rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
P = int(sys.argv[1]) if len(sys.argv) > 1 else 150
s, t, _ = synth.make_pairs(rank * P, P)
src, tgt = torch.from_numpy(s).cuda(), torch.from_numpy(t).cuda()
res = {}
for balance in (False, True):
    dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    ans = solver.solve_model_batches(src, tgt, balance=balance)
    e1.record(); torch.cuda.synchronize()
    res.setdefault(balance, ans)
    print(json.dumps({"rank": rank, "balance": balance, "ms": round(e0.elapsed_time(e1), 1)}), flush=True)
print(json.dumps({"rank": rank, "identical": bool(torch.equal(res[False], res[True]))}), flush=True)
dist.barrier()
dist.destroy_process_group()
