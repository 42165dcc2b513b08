"""Run under b200run -np N (CPU/gloo).  Writes one JSON result per rank."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import b200dp as hvd
from b200dp.parallel.collectives import adasum_tree

out_dir = sys.argv[1]
hvd.init(device="cpu")
r, n = hvd.rank(), hvd.size()
res = {"rank": r, "size": n, "local_rank": hvd.local_rank(), "local_size": hvd.local_size()}

g = torch.Generator().manual_seed(100 + r)
t = torch.randn(1001, generator=g)
all_t = [torch.randn(1001, generator=torch.Generator().manual_seed(100 + i)) for i in range(n)]
avg = hvd.allreduce(t, op=hvd.Average)
res["avg_ok"] = bool(torch.allclose(avg, torch.stack(all_t).mean(0), atol=1e-6))
res["sum_ok"] = bool(torch.allclose(hvd.allreduce(t, op=hvd.Sum), torch.stack(all_t).sum(0), atol=1e-5))
res["input_untouched"] = bool(torch.equal(t, all_t[r]))
ada = hvd.allreduce(t, op=hvd.Adasum)
res["adasum_ok"] = bool(torch.allclose(ada, adasum_tree(all_t), atol=1e-5))
b = torch.full((17,), float(r))
hvd.broadcast_(b, root_rank=n - 1)
res["bcast_ok"] = bool(torch.equal(b, torch.full((17,), float(n - 1))))
res["obj"] = hvd.broadcast_object({"from": r}, 0)["from"]
bf = torch.full((8,), float(r), dtype=torch.bfloat16)
res["bf16_avg"] = float(hvd.allreduce(bf)[0])
res["gather"] = hvd.allgather(torch.tensor([r])).tolist()

# generic DistributedOptimizer over a torch optimizer: replicas stay identical
torch.manual_seed(r)                       # different init per rank on purpose
lin = torch.nn.Linear(6, 3)
hvd.broadcast_parameters(lin.state_dict(), 0)
opt = hvd.DistributedOptimizer(torch.optim.Adam(lin.parameters(), lr=0.1), bucket_bytes=40)
for step in range(3):
    opt.zero_grad()
    x = torch.randn(4, 6, generator=torch.Generator().manual_seed(1000 * step + r))
    lin(x).pow(2).mean().backward()
    opt.step()
res["overlapped_launches"] = opt.overlapped_launches      # buckets reduced from the grad hooks, during backward
hvd.broadcast_optimizer_state(opt, 0)
flat = torch.cat([p.detach().flatten() for p in lin.parameters()])
ref = hvd.broadcast(flat, 0)
res["replicas_identical"] = bool(torch.equal(flat, ref))

# gradient accumulation: two backward passes per step reduce once, with the mean of the two gradients
lin2 = torch.nn.Linear(4, 2)
hvd.broadcast_parameters(lin2.state_dict(), 0)
opt2 = hvd.DistributedOptimizer(torch.optim.SGD(lin2.parameters(), lr=1.0), backward_passes_per_step=2)
w0 = lin2.weight.detach().clone()
xs2 = [torch.randn(3, 4, generator=torch.Generator().manual_seed(50 + 10 * k + r)) for k in range(2)]
for x in xs2:
    lin2(x).sum().backward()
opt2.step()
exp = torch.zeros_like(w0)
for rr in range(n):
    for k in range(2):
        exp += torch.randn(3, 4, generator=torch.Generator().manual_seed(50 + 10 * k + rr)).sum(0).expand(2, 4)
res["accumulation_ok"] = bool(torch.allclose(w0 - lin2.weight.detach(), exp / (2 * n), atol=1e-5))
json.dump(res, open(os.path.join(out_dir, f"r{r}.json"), "w"))
hvd.shutdown()
