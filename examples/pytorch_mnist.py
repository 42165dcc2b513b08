#!/usr/bin/env python
"""Horovod's PyTorch recipe on b200dp: any ``torch.nn.Module`` + any ``torch.optim`` optimizer, data-parallel.

The reference image builds Horovod with PyTorch support next to TensorFlow (``horovod/Dockerfile:44-49,64-65``:
``HOROVOD_WITH_PYTORCH=1``) although its two scripts are TensorFlow programs.  This is the same five-step recipe those
scripts follow (init, pin device, scale LR, wrap optimizer, broadcast from rank 0), written against ``torch``:
gradients are bucketed from autograd hooks and reduced by the NVLink peer-memory kernels (CUDA) or gloo (CPU) while
backward is still running.

    bin/b200run -np 8 python examples/pytorch_mnist.py --synthetic --epochs 2
"""
import argparse
import os
import sys

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200dp as hvd
from b200dp.data import mnist

parser = argparse.ArgumentParser(description="PyTorch MNIST, Horovod style")
parser.add_argument("--batch-size", type=int, default=64)
parser.add_argument("--epochs", type=int, default=1)
parser.add_argument("--lr", type=float, default=0.01)
parser.add_argument("--momentum", type=float, default=0.5)
parser.add_argument("--use-adasum", action="store_true", default=False)
parser.add_argument("--fp16-allreduce", action="store_true", default=False, help="round gradients to fp16 on the wire")
parser.add_argument("--batches-per-allreduce", type=int, default=1)
parser.add_argument("--synthetic", action="store_true")
parser.add_argument("--train-samples", type=int, default=60000)
parser.add_argument("--test-samples", type=int, default=10000)
parser.add_argument("--seed", type=int, default=42)
args = parser.parse_args()


class Net(nn.Module):
    def __init__(self):
        super().__init__()
        self.conv1 = nn.Conv2d(1, 10, kernel_size=5)
        self.conv2 = nn.Conv2d(10, 20, kernel_size=5)
        self.drop = nn.Dropout2d()
        self.fc1 = nn.Linear(320, 50)
        self.fc2 = nn.Linear(50, 10)

    def forward(self, x):
        x = F.relu(F.max_pool2d(self.conv1(x), 2))
        x = F.relu(F.max_pool2d(self.drop(self.conv2(x)), 2))
        x = F.relu(self.fc1(x.flatten(1)))
        return self.fc2(F.dropout(x, training=self.training))


def main():
    hvd.init()
    torch.manual_seed(args.seed)
    dev = hvd.device()
    from b200dp.utils.logging import get_logger
    log = get_logger()

    if args.synthetic:
        (x, y), (xt, yt) = mnist.synthetic_mnist(args.train_samples, args.test_samples, seed=args.seed)
    else:
        (x, y), (xt, yt) = mnist.load_data("MNIST-data-%d" % hvd.rank())
    x, xt = mnist.preprocess(x).reshape(-1, 1, 28, 28), mnist.preprocess(xt).reshape(-1, 1, 28, 28)
    # DistributedSampler semantics: every rank sees its own 1/size of the data
    x, y = x[hvd.rank()::hvd.size()], y[hvd.rank()::hvd.size()]

    model = Net().to(dev)
    lr_scaler = hvd.lr_scaler(args.use_adasum) * args.batches_per_allreduce
    opt = torch.optim.SGD(model.parameters(), lr=args.lr * lr_scaler, momentum=args.momentum)
    hvd.broadcast_parameters(model.state_dict(), root_rank=0)
    hvd.broadcast_optimizer_state(opt, root_rank=0)
    opt = hvd.DistributedOptimizer(opt, named_parameters=model.named_parameters(),
                                   compression=hvd.Compression.fp16 if args.fp16_allreduce else hvd.Compression.none,
                                   op=hvd.Adasum if args.use_adasum else hvd.Average,
                                   backward_passes_per_step=args.batches_per_allreduce)

    def metric_average(val: float) -> float:
        return float(hvd.allreduce(torch.tensor([val], dtype=torch.float64), op=hvd.Average)[0])

    rng = np.random.RandomState(args.seed + hvd.rank())
    for epoch in range(1, args.epochs + 1):
        model.train()
        order = rng.permutation(len(x))
        nb = len(x) // args.batch_size
        for b in range(nb):
            idx = order[b * args.batch_size:(b + 1) * args.batch_size]
            xb = torch.from_numpy(x[idx]).to(dev)
            yb = torch.from_numpy(y[idx].astype(np.int64)).to(dev)
            if b % args.batches_per_allreduce == 0:
                opt.zero_grad()
            loss = F.cross_entropy(model(xb), yb)
            loss.backward()
            if (b + 1) % args.batches_per_allreduce == 0:
                opt.step()
            if b % 10 == 0:
                log.step(b + (epoch - 1) * nb, float(loss), epoch=epoch)
        model.eval()
        with torch.no_grad():
            n = min(len(xt), 2048)
            out = model(torch.from_numpy(xt[:n]).to(dev))
            tl = float(F.cross_entropy(out, torch.from_numpy(yt[:n].astype(np.int64)).to(dev)))
            acc = float((out.argmax(1).cpu().numpy() == yt[:n]).mean())
        tl, acc = metric_average(tl), metric_average(acc)
        if hvd.rank() == 0:
            print(f"Test set: Average loss: {tl:.4f}, Accuracy: {100.0 * acc:.2f}%")
    flat = torch.cat([p.detach().flatten().cpu() for p in model.parameters()])
    same = bool(torch.equal(flat, hvd.broadcast(flat, 0)))
    if hvd.rank() == 0:
        print(f"replicas identical: {same}")
    hvd.shutdown()
    return 0 if same else 1


if __name__ == "__main__":
    sys.exit(main())
