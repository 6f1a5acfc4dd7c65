#!/usr/bin/env python
"""Reference side of the codec micro-benchmark (BASELINE.json config 5).

Times the UNMODIFIED reference (``baseline/_ref``: ``grace_dl.dist`` over NCCL) doing what
``benchmarks/micro.py`` times for this repo: one compressed exchange of all gradient tensors of a
model -- i.e. the reference's per-parameter loop ``for name, g in grads: grc.step(g, name)``
(``examples/dist/CIFAR10-dawndist/core.py:202-208``) -- plus a plain NCCL all-reduce of the same
bytes.  Device-timed with CUDA events, max over ranks, L2 flushed between iterations.

  python benchmarks/micro_reference.py --shape resnet50
  torchrun --nproc-per-node 8 benchmarks/micro_reference.py --shape resnet50
"""
import argparse
import json
import os
import socket
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "baseline", "_ref"))      # the reference's grace_dl, not this repo's shim

import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from clocks import ClockSampler  # noqa: E402

CODECS = [("none", "none", "allreduce"), ("fp16", "none", "allreduce"), ("topk", "residual", "allgather"),
          ("randomk", "residual", "allreduce"), ("threshold", "residual", "allgather"), ("dgc", "dgc", "allgather"),
          ("qsgd", "none", "allgather"), ("terngrad", "residual", "allgather"), ("signsgd", "none", "allgather"),
          ("signum", "residual", "allgather"), ("efsignsgd", "residual", "allgather"), ("onebit", "residual", "allgather"),
          ("powersgd", "powersgd", "allreduce")]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="resnet50")
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--codecs", default="")
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    if "MASTER_ADDR" not in os.environ:
        with socket.socket() as s:
            s.bind(("127.0.0.1", 0))
            os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(s.getsockname()[1]))
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", torch.cuda.current_device()))
    import grace_dl
    assert "baseline/_ref" in os.path.realpath(grace_dl.__file__), grace_dl.__file__
    from grace_dl.dist.helper import grace_from_params
    if args.shape.startswith("flat:"):           # one tensor of N elements (BASELINE config 5 sizes); 2-D so that PowerSGD applies
        n_el = int(args.shape.split(":")[1])
        named = [("flat", (n_el // 4096, 4096) if n_el % 4096 == 0 else (n_el,))]
    else:
        import torchvision
        named = [(n, tuple(p.shape)) for n, p in getattr(torchvision.models, args.shape)().named_parameters()]
    sampler = ClockSampler(torch.cuda.current_device()) if rank == 0 else None
    numel = sum(int(torch.tensor(s).prod()) for _, s in named)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    want = [c for c in args.codecs.split(",") if c]
    rows = []

    def timed(fn):
        times = []
        for it in range(args.warmup + args.iters):
            flush.zero_()
            dist.barrier()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            if it >= args.warmup:
                times.append(a.elapsed_time(b))
        t = torch.tensor([sorted(times)[len(times) // 2]], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    flat = torch.randn(numel, device="cuda")
    t_b = time.time()
    ms = timed(lambda: dist.all_reduce(flat))
    rows.append(dict(codec="nccl-allreduce-flat", world=world, ms=ms, elements=numel,
                     clocks=sampler.window(t_b, time.time()) if sampler else None))
    if rank == 0:
        print("%-22s W=%d %9.3f ms  (one NCCL all-reduce of the %d MB flat gradient)" % ("nccl-allreduce-flat", world, ms, numel * 4 >> 20), flush=True)

    for comp, mem, comm in CODECS:
        if want and comp not in want:
            continue
        params = dict(compressor=comp, memory=mem, communicator=comm, compress_ratio=0.01, threshold=2.5, quantum_num=64,
                      gradient_clipping=False, momentum=0.9, lr=0.1, world_size=world, compress_rank=4)
        try:
            grc = grace_from_params(params)
            if comp == "powersgd":
                grc.compressor.rank, grc.compressor.world_size = 4, world
            grads = [(n, torch.randn(s, device="cuda")) for n, s in named]

            def exchange():
                for n, g in grads:
                    grc.step(g, n)
            t_b = time.time()
            ms = timed(exchange)
            rows.append(dict(codec=comp, memory=mem, communicator=comm, world=world, ms=ms, elements=numel,
                             clocks=sampler.window(t_b, time.time()) if sampler else None))
            del grads, grc
            torch.cuda.empty_cache()
            if rank == 0:
                print("%-10s %-9s %-9s W=%d %9.3f ms" % (comp, mem, comm, world, ms), flush=True)
        except Exception as exc:           # several reference paths are broken
            if rank == 0:
                print("%-10s %-9s %-9s W=%d  fails in the reference: %s" % (comp, mem, comm, world, str(exc)[:90]), flush=True)
    if sampler:
        sampler.close()
    if rank == 0 and args.out:
        with open(args.out, "w") as f:
            json.dump(rows, f, indent=1)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
