#!/usr/bin/env python
"""CIFAR10 DAWNBench-style ResNet-9 with the torch.distributed-style manual loop (workload of the
reference's ``examples/dist/CIFAR10-dawndist/{dawn,core}.py``): piecewise-linear LR, Nesterov SGD,
batch 512, Crop/FlipLR/Cutout augmentation, TableLogger + ``logs.tsv``, DAWN timing (evaluation excluded).  After
``backward()`` the gradients are exchanged through the torch.distributed-flavour API: ``grc.step_all(named_grads)``,
the bucketed form of the reference's ``for name, p: grc.step(p.grad, name)`` loop (``core.py:202-208``;
``--per-tensor`` runs that loop literally).  Synthetic CIFAR-shaped data unless ``--data-dir`` holds the real set.

  torchrun --standalone --nproc-per-node 2 examples/dist/cifar10_dawn.py --epochs 2
"""
import argparse
import os
import sys
import time

import torch
import torch.distributed as dist
import torch.nn.functional as F

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from grace_b200.dist import grace_from_params  # noqa: E402
from grace_b200.models import resnet9  # noqa: E402
from grace_b200.utils.augment import DawnAugment, normalise, pad_reflect  # noqa: E402
from grace_b200.utils.metrics import TableLogger, Timer, TSVLogger  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--epochs", type=int, default=24)
ap.add_argument("--batch-size", type=int, default=512)
ap.add_argument("--data-dir", "--data_dir", dest="data_dir", default="")
ap.add_argument("--backend", default="nccl" if torch.cuda.is_available() else "gloo")
ap.add_argument("--compressor", "--compression_method", dest="compressor", default="none")
ap.add_argument("--memory", nargs="?", const="residual", default="none",
                help="memory name; the reference's bare --memory flag (dawn.py:18) selects the residual memory")
ap.add_argument("--communicator", "--communication_method", dest="communicator", default="allreduce")
ap.add_argument("--compress-ratio", "--compress_ratio", dest="compress_ratio", type=float, default=0.01)
# flags of the reference's dawn.py:8-22, accepted so its command lines keep working (the reference parses most of them
# and then ignores them in favour of hard-coded params, dawn.py:125-127; here they are honoured)
ap.add_argument("--log_dir", default=".")
ap.add_argument("--per-tensor", action="store_true", help="exchange with one grc.step per parameter (the reference loop)")
ap.add_argument("--no-augment", action="store_true")
ap.add_argument("--graph", action="store_true",
                help="CUDA only: capture forward, backward, grc.step_all and the optimizer update as ONE CUDA graph "
                     "(grace_b200.dist.GraphedStep); the learning rate lives in a device tensor updated in place")
ap.add_argument("--master_address", "-m", default="127.0.0.1")
ap.add_argument("--rank", "-r", type=int, default=None)
ap.add_argument("--world_size", "-w", type=int, default=None)
ap.add_argument("--clipping", action="store_true")
ap.add_argument("--efsgdlr", type=float, default=0.1)
ap.add_argument("--momentum", type=float, default=0.9)
ap.add_argument("--threshold", type=float, default=0.01)
ap.add_argument("--qstates", type=int, default=256)
args = ap.parse_args()
if args.rank is not None and "RANK" not in os.environ:          # reference style: explicit rank / world size / master
    os.environ.update(RANK=str(args.rank), WORLD_SIZE=str(args.world_size or 2), MASTER_ADDR=args.master_address,
                      MASTER_PORT=os.environ.get("MASTER_PORT", "1111"), LOCAL_RANK=os.environ.get("LOCAL_RANK", str(args.rank)))

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
if world > 1:
    dist.init_process_group(args.backend)
device = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0))) if torch.cuda.is_available() else torch.device("cpu")
if device.type == "cuda":
    torch.cuda.set_device(device)
torch.manual_seed(0)


def load():
    """uint8 NHWC images + labels: the real CIFAR10 when --data-dir has it, else a synthetic set of the same shape."""
    try:
        import torchvision
        tr = torchvision.datasets.CIFAR10(args.data_dir, train=True, download=False)
        te = torchvision.datasets.CIFAR10(args.data_dir, train=False, download=False)
        f = lambda d: (torch.tensor(d.data), torch.tensor(d.targets))
        return f(tr), f(te)
    except Exception:
        g = torch.Generator().manual_seed(1)
        def fake(n):
            y = torch.randint(0, 10, (n,), generator=g)
            x = torch.rand(n, 32, 32, 3, generator=g) * 0.5
            for c in range(10):
                x[y == c, 3 * c:3 * c + 5, :, c % 3] += 0.5
            return (x * 255).to(torch.uint8), y
        return fake(8192), fake(2048)


timer = Timer(synch=torch.cuda.synchronize if device.type == "cuda" else None)
(train_x, train_y), (test_x, test_y) = load()
# reference preprocessing (core.py:50-62, dawn.py:105-108): normalise with the CIFAR10 statistics, reflect-pad the
# training images by 4, NCHW -- once, on the device; the augmentation then crops 32x32 windows out of the padded set
train_x = pad_reflect(normalise(train_x.to(device)), 4)
test_x = normalise(test_x.to(device))
train_y, test_y = train_y.to(device), test_y.to(device)
augment = DawnAugment(crop=(32, 32), cutout=(8, 8), flip=True,
                      generator=torch.Generator(device=device).manual_seed(1000 + int(os.environ.get("RANK", 0))))
if os.environ.get("RANK", "0") == "0":
    print("Preprocessing finished in %.2f seconds" % timer())
model = resnet9().to(device).to(memory_format=torch.channels_last)
if world > 1:
    for p in model.parameters():
        dist.broadcast(p.data, 0)
steps_per_epoch = len(train_x) // (args.batch_size * world)
lr_at = lambda e: max(0.0, (e / 5 if e < 5 else (args.epochs - e) / max(args.epochs - 5, 1))) * 0.4   # piecewise linear
use_graph = args.graph and device.type == "cuda" and not args.per_tensor
lr_t = torch.zeros((), device=device) if use_graph else None     # a Python float would be frozen into the graph
# (torch's foreach SGD reads a tensor learning rate back with .item(), which a capturing stream forbids; fused=True
#  takes the tensor as is)
opt = torch.optim.SGD(model.parameters(), lr=lr_t if use_graph else 0.0, momentum=0.9, nesterov=True,
                      weight_decay=5e-4 * args.batch_size, fused=True if use_graph else None)
grc = grace_from_params({"compressor": args.compressor, "memory": args.memory, "communicator": args.communicator,
                         "compress_ratio": args.compress_ratio, "world_size": world, "threshold": args.threshold,
                         "quantum_num": args.qstates, "lr": args.efsgdlr, "momentum": args.momentum,
                         "gradient_clipping": args.clipping})
table, tsv = TableLogger(rank=rank), TSVLogger()
named = [(n, p) for n, p in model.named_parameters() if p.requires_grad]
gstep = None
if use_graph:
    from grace_b200.dist import GraphedStep
    model.train()
    ex = train_x[:args.batch_size]
    ex = (ex[:, :, 4:36, 4:36] if args.no_augment else augment(ex)).contiguous(memory_format=torch.channels_last)
    # three warm-up steps at lr = 0 (parameters do not move), then the capture; every later step is one graph replay
    gstep = GraphedStep(model, opt, lambda m, x, y: F.cross_entropy(m(x), y, reduction="sum"), [ex, train_y[:args.batch_size]],
                        amp_dtype=torch.bfloat16, warmup=3, communicator=grc)
    for st in opt.state.values():                              # the warm-up gradients must not seed the momentum
        if "momentum_buffer" in st and st["momentum_buffer"] is not None:
            st["momentum_buffer"].zero_()
for epoch in range(args.epochs):
    model.train()
    perm = torch.randperm(len(train_x), generator=torch.Generator().manual_seed(epoch)).to(device)
    loss_sum = torch.zeros((), device=device)
    for s in range(steps_per_epoch):
        idx = perm[(s * world + rank) * args.batch_size:(s * world + rank + 1) * args.batch_size]
        x = train_x[idx]
        x = (x[:, :, 4:36, 4:36] if args.no_augment else augment(x)).contiguous(memory_format=torch.channels_last)
        y = train_y[idx]
        if gstep is not None:
            lr_t.fill_(lr_at(epoch + s / steps_per_epoch) / args.batch_size)
            loss_sum += gstep(x, y)
            continue
        for g in opt.param_groups:
            g["lr"] = lr_at(epoch + s / steps_per_epoch) / args.batch_size
        opt.zero_grad(set_to_none=False)
        with torch.autocast(device.type, dtype=torch.bfloat16, enabled=device.type == "cuda"):
            loss = F.cross_entropy(model(x), y, reduction="sum")
        loss.backward()
        if args.per_tensor:
            for name, p in named:                              # the reference's sequential exchange loop
                p.grad.copy_(grc.step(p.grad, name))
        else:
            grc.step_all([(n, p.grad) for n, p in named], inplace=True)      # same exchange, O(#buckets) launches
        opt.step()
        loss_sum += loss.detach()
    train_time = timer()                                       # DAWN timing: training counts ...
    model.eval()
    with torch.no_grad(), torch.autocast(device.type, dtype=torch.bfloat16, enabled=device.type == "cuda"):
        correct = sum(int((model(test_x[i:i + 1000].contiguous(memory_format=torch.channels_last)).argmax(1)
                           == test_y[i:i + 1000]).sum()) for i in range(0, len(test_x), 1000))
    test_time = timer(include_in_total=False)                  # ... evaluation does not (dawn.py:129 test_time_in_total=False)
    row = {"epoch": epoch + 1, "lr": lr_at(epoch + 1), "train_time": train_time,
           "train_loss": float(loss_sum) / (steps_per_epoch * args.batch_size), "test_time": test_time,
           "test_acc": correct / len(test_x), "total_time": timer.total_time}
    table.append(row)
    tsv.append(row)
if rank == 0:
    tsv.write(os.path.join(os.path.expanduser(args.log_dir), "logs.tsv"))
if hasattr(grc, "close"):
    grc.close()
if world > 1:
    dist.destroy_process_group()
