#!/usr/bin/env python
"""MNIST-shaped training with the Horovod-style API (workload of the reference's
``examples/torch/pytorch_mnist.py``).  Uses torchvision's MNIST when ``--data-dir`` holds it,
otherwise a synthetic stand-in with the same shapes (this sandbox has no network).

  torchrun --standalone --nproc-per-node 2 examples/torch/pytorch_mnist.py --epochs 2
"""
import argparse
import os
import sys

import torch
import torch.nn.functional as F
import torch.utils.data.distributed

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import grace_b200.torch as hvd  # noqa: E402
from grace_b200.models import MnistNet  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--batch-size", type=int, default=64)
ap.add_argument("--epochs", type=int, default=2)
ap.add_argument("--lr", type=float, default=0.01)
ap.add_argument("--momentum", type=float, default=0.5)
ap.add_argument("--data-dir", default="")
ap.add_argument("--compressor", default="topk")
ap.add_argument("--memory", default="residual")
ap.add_argument("--communicator", default="allgather")
ap.add_argument("--no-cuda", action="store_true")
# remaining flags of the reference example (pytorch_mnist.py:15-40)
ap.add_argument("--test-batch-size", type=int, default=1000)
ap.add_argument("--seed", type=int, default=42)
ap.add_argument("--log-interval", type=int, default=10)
ap.add_argument("--fp16-allreduce", action="store_true", help="use the fp16 codec (what Horovod's fp16 compression did)")
args = ap.parse_args()
if args.fp16_allreduce:
    args.compressor, args.memory, args.communicator = "fp16", "none", "allreduce"
torch.manual_seed(args.seed)
cuda = not args.no_cuda and torch.cuda.is_available()
hvd.init()
torch.manual_seed(42)


def dataset(train):
    try:
        from torchvision import datasets, transforms
        tf = transforms.Compose([transforms.ToTensor(), transforms.Normalize((0.1307,), (0.3081,))])
        return datasets.MNIST(args.data_dir, train=train, download=False, transform=tf)
    except Exception:
        g = torch.Generator().manual_seed(0 if train else 1)
        n = 4096 if train else 1024
        y = torch.randint(0, 10, (n,), generator=g)
        x = torch.randn(n, 1, 28, 28, generator=g) * 0.3
        for c in range(10):                                   # class-dependent blob so the task is learnable
            x[y == c, 0, 2 * c:2 * c + 6, 2 * c:2 * c + 6] += 2.0
        return torch.utils.data.TensorDataset(x, y)


train_set, test_set = dataset(True), dataset(False)
sampler = torch.utils.data.distributed.DistributedSampler(train_set, num_replicas=hvd.size(), rank=hvd.rank())
train_loader = torch.utils.data.DataLoader(train_set, batch_size=args.batch_size, sampler=sampler)
test_loader = torch.utils.data.DataLoader(test_set, batch_size=1000)

model = MnistNet()
if cuda:
    model.cuda()
optimizer = torch.optim.SGD(model.parameters(), lr=args.lr * hvd.size(), momentum=args.momentum)
hvd.broadcast_parameters(model.state_dict(), root_rank=0)
hvd.broadcast_optimizer_state(optimizer, root_rank=0)
grc = hvd.grace_from_params({"compressor": args.compressor, "memory": args.memory, "communicator": args.communicator})
optimizer = hvd.DistributedOptimizer(optimizer, grc, named_parameters=model.named_parameters())


def metric_average(val):
    return float(hvd.allreduce(torch.tensor([val], device="cuda" if cuda else "cpu"), average=True))


for epoch in range(1, args.epochs + 1):
    model.train()
    sampler.set_epoch(epoch)
    for i, (x, y) in enumerate(train_loader):
        if cuda:
            x, y = x.cuda(), y.cuda()
        optimizer.zero_grad()
        loss = F.nll_loss(model(x), y)
        loss.backward()
        optimizer.step()
        if i % 20 == 0 and hvd.rank() == 0:
            print("epoch %d batch %d loss %.4f" % (epoch, i, float(loss)), flush=True)
    model.eval()
    correct = total = 0
    with torch.no_grad():
        for x, y in test_loader:
            if cuda:
                x, y = x.cuda(), y.cuda()
            correct += int((model(x).argmax(1) == y).sum())
            total += y.numel()
    acc = metric_average(correct / total)
    if hvd.rank() == 0:
        print("epoch %d test accuracy %.2f%%" % (epoch, 100 * acc), flush=True)
hvd.shutdown()
