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

# ---- the rest of Horovod's collective API (b200dp/parallel/extras.py) ----------------------------------------------
h = hvd.allreduce_async(t, op=hvd.Average)
h2 = hvd.broadcast_async(torch.full((5,), float(r)), root_rank=0)
res["async_avg_ok"] = bool(torch.allclose(hvd.synchronize(h), torch.stack(all_t).mean(0), atol=1e-6)) and hvd.poll(h)
res["async_bcast_ok"] = bool(torch.equal(hvd.synchronize(h2), torch.zeros(5)))
a, b2 = torch.full((3,), float(r + 1)), torch.full((2, 2), float(10 * (r + 1)), dtype=torch.float64)
ga, gb = hvd.grouped_allreduce([a, b2], op=hvd.Sum)
tot = n * (n + 1) / 2
res["grouped_ok"] = bool(torch.allclose(ga, torch.full((3,), tot)) and torch.allclose(gb, torch.full((2, 2), 10 * tot, dtype=torch.float64))
                         and gb.dtype == torch.float64 and torch.equal(a, torch.full((3,), float(r + 1))))
gad = hvd.grouped_allreduce([all_t[r][:100].clone(), all_t[r][100:300].clone()], op=hvd.Adasum)
res["grouped_adasum_ok"] = bool(torch.allclose(gad[0], adasum_tree([x[:100] for x in all_t]), atol=1e-5)
                                and torch.allclose(gad[1], adasum_tree([x[100:300] for x in all_t]), atol=1e-5))
# alltoall: rank r sends (dst + 1) rows tagged 100*r + dst to every dst
send = torch.cat([torch.full((d + 1, 2), float(100 * r + d)) for d in range(n)])
got, counts = hvd.alltoall(send, splits=[d + 1 for d in range(n)])
exp_rows = torch.cat([torch.full((r + 1, 2), float(100 * src + r)) for src in range(n)])
res["alltoall_ok"] = bool(torch.equal(got, exp_rows)) and counts == [r + 1] * n
even = hvd.alltoall(torch.arange(2 * n, dtype=torch.float32) + 1000 * r)
res["alltoall_even_ok"] = bool(torch.equal(even, torch.cat([torch.arange(2 * r, 2 * r + 2, dtype=torch.float32) + 1000 * src for src in range(n)])))
rs = hvd.reducescatter(torch.arange(2 * n + 1, dtype=torch.float32) * (r + 1), op=hvd.Sum)
full = torch.arange(2 * n + 1, dtype=torch.float32) * tot
lo = r * 2 + min(r, 1)
res["reducescatter_ok"] = bool(torch.equal(rs, full[lo:lo + 2 + (1 if r < 1 else 0)]))
res["minmax_ok"] = (float(hvd.allreduce(torch.tensor([float(r)]), op=hvd.Max)) == n - 1 and float(hvd.allreduce(torch.tensor([float(r)]), op=hvd.Min)) == 0.0
                    and int(hvd.allreduce(torch.tensor([2]), op=hvd.Product)) == 2 ** n)
rag = hvd.allgather(torch.full((r + 1, 2), float(r)))
res["ragged_gather_ok"] = bool(torch.equal(rag, torch.cat([torch.full((i + 1, 2), float(i)) for i in range(n)])))
res["join"] = hvd.join()
res["cross"] = [hvd.cross_rank(), hvd.cross_size()]
# wire compression: gradients rounded to fp16 for the reduction, replicas still identical
lin3 = torch.nn.Linear(5, 2)
hvd.broadcast_parameters(lin3.state_dict(), 0)
opt3 = hvd.DistributedOptimizer(torch.optim.SGD(lin3.parameters(), lr=0.5), compression=hvd.Compression.fp16)
lin3(torch.randn(4, 5, generator=torch.Generator().manual_seed(77 + r))).pow(2).sum().backward()
opt3.step()
f3 = torch.cat([p.detach().flatten() for p in lin3.parameters()])
res["compression_ok"] = bool(torch.equal(f3, hvd.broadcast(f3, 0))) and lin3.weight.grad.dtype == torch.float32
json.dump(res, open(os.path.join(out_dir, f"r{r}.json"), "w"))
hvd.shutdown()
