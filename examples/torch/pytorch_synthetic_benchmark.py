#!/usr/bin/env python
"""Synthetic img/s benchmark, Horovod-style API (workload of the reference's
``examples/torch/pytorch_synthetic_benchmark.py``: torchvision model, batch 32, fixed fake data,
10 warm-up batches, 10 iterations x 10 batches, img/sec +- 1.96 sigma).

  torchrun --standalone --nproc-per-node 8 examples/torch/pytorch_synthetic_benchmark.py \
      --compressor topk --memory residual --communicator allgather
"""
import argparse
import os
import sys
import timeit

import numpy as np
import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
import grace_b200.torch as hvd  # noqa: E402
from grace_b200.utils.metrics import CompressionStats  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="resnet50")
ap.add_argument("--batch-size", type=int, default=32)
ap.add_argument("--num-warmup-batches", type=int, default=10)
ap.add_argument("--num-batches-per-iter", type=int, default=10)
ap.add_argument("--num-iters", type=int, default=10)
ap.add_argument("--compressor", default="signum")
ap.add_argument("--memory", default="residual")
ap.add_argument("--communicator", default="allgather")
ap.add_argument("--compress-ratio", type=float, default=0.01)
ap.add_argument("--bf16", action="store_true")
ap.add_argument("--graph", action="store_true",
                help="replay the whole step (forward, backward, fused exchange, optimizer) as ONE CUDA graph and feed it "
                     "from pinned host memory with hvd.GraphedStep.feed/run")
ap.add_argument("--no-cuda", action="store_true")
# remaining flags of the reference harness (pytorch_synthetic_benchmark.py:42-72); --use-adasum has no equivalent
ap.add_argument("--fp16-allreduce", action="store_true", help="use the fp16 codec")
ap.add_argument("--use-adasum", action="store_true", help="accepted and ignored")
ap.add_argument("--seed", type=int, default=42)
ap.add_argument("--lr", type=float, default=0.01)
ap.add_argument("--momentum", type=float, default=0.5)
ap.add_argument("--epochs", type=int, default=1, help="accepted and ignored (the harness runs --num-iters x --num-batches-per-iter)")
ap.add_argument("--log-interval", type=int, default=10, help="accepted and ignored")
args = ap.parse_args()
if args.fp16_allreduce:
    args.compressor, args.memory, args.communicator = "fp16", "none", "allreduce"
torch.manual_seed(args.seed)
cuda = not args.no_cuda and torch.cuda.is_available()

hvd.init()
import torchvision  # noqa: E402
model = getattr(torchvision.models, args.model)()
if cuda:
    model = model.cuda().to(memory_format=torch.channels_last)
optimizer = torch.optim.SGD(model.parameters(), lr=args.lr * hvd.size(), momentum=args.momentum)
hvd.broadcast_parameters(model.state_dict(), root_rank=0)
hvd.broadcast_optimizer_state(optimizer, root_rank=0)
grc = hvd.grace_from_params({"compressor": args.compressor, "memory": args.memory, "communicator": args.communicator,
                             "compress_ratio": args.compress_ratio})
optimizer = hvd.DistributedOptimizer(optimizer, grc, named_parameters=model.named_parameters())
stats = CompressionStats(optimizer)

data = torch.randn(args.batch_size, 3, 224, 224)
target = torch.randint(0, 1000, (args.batch_size,))
if cuda:
    data, target = data.cuda().contiguous(memory_format=torch.channels_last), target.cuda()


def benchmark_step():
    optimizer.zero_grad()
    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=cuda and args.bf16):
        loss = F.cross_entropy(model(data), target)
    loss.backward()
    optimizer.step()
    stats.step()
    if cuda:
        torch.cuda.synchronize()


graphed = None
if args.graph and cuda:
    graphed = hvd.GraphedStep(model, optimizer, lambda m, x, y: F.cross_entropy(m(x), y), [data, target],
                              amp_dtype=torch.bfloat16 if args.bf16 else None)
    stats.graph_launches_per_step = graphed.native_launches_per_step
    host_data, host_target = data.cpu().contiguous().pin_memory(), target.cpu().pin_memory()   # "data loader" output
    graphed.feed(host_data, host_target)
    pending = []

    def benchmark_step():                       # noqa: F811
        pending.append(graphed.run())           # consumes the fed batch, replays, starts the loss read-back
        graphed.feed(host_data, host_target)    # next batch: copied on the copy stream while the replay runs
        if len(pending) > 1:
            pending.pop(0).item()               # loss of the previous step, already on the host
        stats.step()


def log(s):
    if hvd.rank() == 0:
        print(s, flush=True)


log("Model: %s  batch %d  GPUs %d  grace %s+%s+%s" % (args.model, args.batch_size, hvd.size(), args.compressor,
                                                       args.memory, args.communicator))
timeit.timeit(benchmark_step, number=args.num_warmup_batches)
rates = []
for i in range(args.num_iters):
    t = timeit.timeit(benchmark_step, number=args.num_batches_per_iter)
    rates.append(args.batch_size * args.num_batches_per_iter / t)
    log("Iter #%d: %.1f img/sec per GPU" % (i, rates[-1]))
mean, conf = np.mean(rates), 1.96 * np.std(rates)
log("Img/sec per GPU: %.1f +-%.1f" % (mean, conf))
log("Total img/sec on %d GPU(s): %.1f +-%.1f" % (hvd.size(), hvd.size() * mean, hvd.size() * conf))
log("compression: %s" % stats.summary())
hvd.shutdown()
