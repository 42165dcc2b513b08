"""Stand-in for the reference's Horovod+NCCL(+cuDNN/cuBLAS) stack, which cannot be installed offline
(SURVEY §0.2-3): the identical CNN / Adam / batch / LR rule in PyTorch eager + DistributedDataParallel
over NCCL, bf16 autocast, channels-last, fused Adam.  "A path that only calls NCCL" — the baseline.

Two forms: the eager loop (DistributedDataParallel, what a user writes first) and ``--graph``: the WHOLE step — forward,
backward, one NCCL all-reduce of a flat gradient buffer the parameters' .grad fields are views of, the 1/size scale and the
fused (capturable) Adam — captured once in a CUDA graph and replayed, which removes the host launch latency that dominates
the eager loop on a model this small.  Every ratio quoted in BASELINE.md uses the faster of the two."""
import json
import os

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F


class Net(nn.Module):
    def __init__(self, width=1.0):
        super().__init__()
        c1, c2, d = int(round(32 * width)), int(round(64 * width)), int(round(1024 * width))
        self.conv1 = nn.Conv2d(1, c1, 5, padding=2)
        self.conv2 = nn.Conv2d(c1, c2, 5, padding=2)
        self.fc1 = nn.Linear(7 * 7 * c2, d)
        self.drop = nn.Dropout(0.5)
        self.fc2 = nn.Linear(d, 10)

    def forward(self, x):
        x = x.view(-1, 1, 28, 28)
        x = F.max_pool2d(F.relu(self.conv1(x)), 2)
        x = F.max_pool2d(F.relu(self.conv2(x)), 2)
        x = x.permute(0, 2, 3, 1).reshape(x.shape[0], -1)
        return self.fc2(self.drop(F.relu(self.fc1(x))))


class KerasNet(nn.Module):
    """The Keras MNIST example shape (keras128 in b200dp.models.spec)."""

    def __init__(self, width=1.0):
        super().__init__()
        c1, c2, d = int(round(32 * width)), int(round(64 * width)), int(round(128 * width))
        self.conv1 = nn.Conv2d(1, c1, 3)
        self.conv2 = nn.Conv2d(c1, c2, 3)
        self.drop1 = nn.Dropout(0.25)
        self.fc1 = nn.Linear(12 * 12 * c2, d)
        self.drop2 = nn.Dropout(0.5)
        self.fc2 = nn.Linear(d, 10)

    def forward(self, x):
        x = x.view(-1, 1, 28, 28)
        x = F.relu(self.conv1(x))
        x = self.drop1(F.max_pool2d(F.relu(self.conv2(x)), 2))
        x = x.permute(0, 2, 3, 1).reshape(x.shape[0], -1)
        return self.fc2(self.drop2(F.relu(self.fc1(x))))


def run(args) -> int:
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    torch.backends.cudnn.benchmark = True
    B = args.batch
    model = (KerasNet if getattr(args, "arch", "mnist_cnn") == "keras128" else Net)(args.width).to(dev).to(memory_format=torch.channels_last)
    eager_ddp = world > 1 and not getattr(args, "graph", False)
    net = nn.parallel.DistributedDataParallel(model, device_ids=[local], bucket_cap_mb=64, gradient_as_bucket_view=True) if eager_ddp else model
    opt = torch.optim.Adam(model.parameters(), lr=0.001 * world, fused=True)
    g = torch.Generator(device="cpu").manual_seed(1234 + rank)
    pool_n = max(8, int(160e6 // (B * 784 * 4)) + 1)
    px = torch.rand(pool_n, B, 784, generator=g).to(dev)
    py = torch.randint(0, 10, (pool_n, B), generator=g).to(dev)

    def step(x, y):
        opt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            loss = F.cross_entropy(net(x).float(), y)
        loss.backward()
        opt.step()
        return loss

    graphed = bool(getattr(args, "graph", False))
    if graphed:
        # static buffers + one flat gradient buffer (each p.grad is a view of it: a single all-reduce, no copies)
        params = [p for p in model.parameters()]
        flat = torch.zeros(sum(p.numel() for p in params), device=dev)
        off = 0
        for p in params:
            # same memory layout as the parameter (conv weights are channels-last): the fused Adam insists on it
            p.grad = flat[off:off + p.numel()].as_strided(p.size(), p.stride())
            off += p.numel()
        opt = torch.optim.Adam(params, lr=0.001 * world, fused=True, capturable=True)
        sx = torch.zeros(B, 784, device=dev)
        sy = torch.zeros(B, dtype=torch.long, device=dev)
        sloss = torch.zeros((), device=dev)

        def body():
            flat.zero_()
            with torch.autocast("cuda", dtype=torch.bfloat16):
                loss = F.cross_entropy(model(sx).float(), sy)
            loss.backward()
            if world > 1:
                dist.all_reduce(flat)
                flat.mul_(1.0 / world)
            opt.step()
            sloss.copy_(loss.detach())

        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(11):
                body()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        cg = torch.cuda.CUDAGraph()
        with torch.cuda.graph(cg):
            body()

        def step(x, y):  # noqa: F811
            sx.copy_(x, non_blocking=True)
            sy.copy_(y, non_blocking=True)
            cg.replay()
            return sloss

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(max(args.warmup, 3)):
        step(px[i % pool_n], py[i % pool_n])
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        step(px[(args.warmup + i) % pool_n], py[(args.warmup + i) % pool_n])
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    # end to end: pinned host batch -> H2D -> step -> loss D2H
    hp = 64
    hx = torch.rand(hp, B, 784, generator=g).pin_memory()
    hy = torch.randint(0, 10, (hp, B), generator=g).pin_memory()
    host_loss = torch.zeros(args.steps, dtype=torch.float32).pin_memory()
    for i in range(3):
        step(hx[i].to(dev, non_blocking=True), hy[i].to(dev, non_blocking=True))
    barrier()
    e0.record()
    for i in range(args.steps):
        l = step(hx[i % hp].to(dev, non_blocking=True), hy[i % hp].to(dev, non_blocking=True))
        host_loss[i:i + 1].copy_(l.detach().view(1), non_blocking=True)
    e1.record()
    barrier()
    ems = e0.elapsed_time(e1)
    t = torch.tensor([ms, ems], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ems = float(t[0]), float(t[1])
    if rank == 0:
        print(json.dumps({
            "metric": "MNIST CNN training throughput (images/sec), device-timed, max over ranks",
            "impl": "standin-graph" if graphed else "standin", "cuda_graph": graphed,
            "note": "stand-in for Horovod+NCCL (reference not installable offline): PyTorch DDP + NCCL + cuDNN/cuBLAS, bf16 autocast",
            "value": round(args.steps * B * world / (ms / 1e3), 1), "unit": "images/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ms / args.steps, 5), "higher_is_better": True, "scaling": "weak",
            "dtype": "bf16", "data": "synthetic", "config": {"model": getattr(args, "arch", "mnist_cnn"), "global_batch": B * world, "parallelism": f"dp{world}",
                                                              "width": args.width},
            "e2e": {"value": round(args.steps * B * world / (ems / 1e3), 1), "unit": "images/s",
                    "h2d_bytes_per_step": B * 784 * 4 + B * 8, "d2h_bytes_per_step": 4}}))
    if graphed:
        # tearing NCCL down under a live captured graph that holds its kernels has hung at exit: the line is printed,
        # leave without the teardown
        import sys
        sys.stdout.flush()
        if world > 1:
            dist.barrier()
        os._exit(0)
    if world > 1:
        dist.destroy_process_group()
    return 0
