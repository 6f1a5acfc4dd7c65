#!/usr/bin/env python
"""Headline benchmark: ResNet-50 synthetic training throughput (img/s, whole job) with gradient
compression, the harness of the reference's ``examples/torch/pytorch_synthetic_benchmark.py``
(batch 32 per GPU, fixed fake data 3x224x224) on BASELINE.json config 2: bf16 compute,
TopK 1 % + ResidualMemory + Allgather.

  python bench.py --gpus N --steps K --warmup W                # this framework (fused engine)
  python bench.py --impl reference --gpus N --steps K ...      # UNMODIFIED reference from baseline/_ref

For N > 1 launch through ``python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N``.
Rank 0 prints ONE JSON line.  ``value`` is device-timed (CUDA events, max over ranks) with the
batch resident on the GPU like the reference harness; ``e2e`` repeats the measurement with a
per-step pinned-host -> device copy of the inputs and a device -> host read of the loss.
"""
import argparse
import json
import os
import socket
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "ddp"],
                    help="ours | reference (unmodified baseline/_ref) | ddp (library baseline: torch DDP + NCCL, no compression)")
    ap.add_argument("--model", default="resnet50", help="torchvision model name, or bert-large / bert-base")
    ap.add_argument("--seq-len", type=int, default=128, help="BERT only")
    ap.add_argument("--batch-size", type=int, default=32, help="per GPU (reference harness default)")
    ap.add_argument("--compressor", default="topk")
    ap.add_argument("--memory", default=None)
    ap.add_argument("--communicator", default=None)
    ap.add_argument("--compress-ratio", type=float, default=0.01)
    ap.add_argument("--quantum-num", type=int, default=64)
    ap.add_argument("--compress-rank", type=int, default=4)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="ours: run the step eagerly instead of as one CUDA graph")
    ap.add_argument("--memory-format", default="auto", choices=["auto", "channels_last", "contiguous"])
    return ap.parse_args()


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


class ClockSampler:
    """nvidia-smi clocks and throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [c.strip() for c in line.split(",")]))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for t, r in self.rows if t0 <= t <= t1 and len(r) >= 7] or [r for _, r in self.rows if len(r) >= 7]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm = sorted(float(r[0]) for r in rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(rows[0][1]), "samples": len(rows),
                "power_w_max": max(float(r[2]) for r in rows), "reasons": reasons}


def triple(args):
    comp = args.compressor
    mem = args.memory or {"topk": "residual", "qsgd": "none", "powersgd": "powersgd", "efsignsgd": "efsignsgd",
                          "dgc": "dgc", "none": "none"}.get(comp, "residual")
    comm = args.communicator or ("allreduce" if comp in ("powersgd", "none", "fp16") else "allgather")
    return comp, mem, comm


def build_model(args, device):
    import torch
    torch.manual_seed(0)
    if args.model.startswith("bert"):
        # BASELINE.json config 4: BERT pre-training (MLM + NSP heads), random init, synthetic token ids
        from transformers import BertConfig, BertForPreTraining
        large = args.model == "bert-large"
        cfg = BertConfig(hidden_size=1024 if large else 768, num_hidden_layers=24 if large else 12,
                         num_attention_heads=16 if large else 12, intermediate_size=4096 if large else 3072)
        args.memory_format = "contiguous"
        return BertForPreTraining(cfg).to(device)
    import torchvision
    model = getattr(torchvision.models, args.model)().to(device)
    fmt = args.memory_format
    if fmt == "auto":
        # the reference's PowerSGD does tensor.view(shape[0], -1) (compressor/powersgd.py:35), which
        # cannot take channels_last gradients; every other configuration runs channels_last
        fmt = "contiguous" if (args.impl == "reference" and args.compressor == "powersgd") else "channels_last"
    args.memory_format = fmt
    if fmt == "channels_last":
        model = model.to(memory_format=torch.channels_last)
    return model


def main():
    args = parse()
    comp, mem, comm = triple(args)
    if args.impl == "reference":
        ref_root = os.path.join(ROOT, "baseline", "_ref")
        if not os.path.isdir(os.path.join(ref_root, "grace_dl", "dist")):
            print(json.dumps({"impl": "reference", "unavailable": "baseline/_ref/grace_dl/dist missing: run baseline/install_reference.sh"}))
            return 0
        sys.path.insert(0, ref_root)          # the reference's grace_dl must shadow this repo's import shim
    else:
        sys.path.insert(0, ROOT)

    # keep the real stdout for the single JSON line: libraries (NCCL's version banner, torchrun
    # notices) write to fd 1, so everything else is diverted to stderr
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)

    import torch
    import torch.distributed as dist
    import torch.nn.functional as F

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d" % (args.gpus, world))
    if args.gpus > 1 and world == 1:
        # convenience: re-launch ourselves under torchrun
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.gpus),
               "--master-addr", "127.0.0.1", "--master-port", str(free_port()), os.path.abspath(__file__)] + sys.argv[1:]
        return subprocess.call(cmd, stdout=real_stdout)
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1 or args.impl == "reference":
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", str(free_port()))
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=device)

    torch.backends.cudnn.benchmark = True
    model = build_model(args, device)
    base_opt = torch.optim.SGD(model.parameters(), lr=0.01, momentum=0.9)
    n_params = sum(p.numel() for p in model.parameters())

    fused = None
    if args.impl == "ours":
        import grace_b200.torch as hvd
        hvd.broadcast_parameters(model.state_dict(), root_rank=0)
        grc = hvd.grace_from_params({"compressor": comp, "memory": mem, "communicator": comm,
                                     "compress_ratio": args.compress_ratio, "quantum_num": args.quantum_num,
                                     "compress_rank": args.compress_rank, "world_size": world})
        opt = hvd.DistributedOptimizer(base_opt, grc, named_parameters=model.named_parameters(), fused="auto")
        fused = opt._fused

        def exchange_and_step():
            opt.step()

        def zero():
            opt.zero_grad()
    elif args.impl == "ddp":
        # library baseline: uncompressed gradients through torch DDP + NCCL all-reduce
        if world > 1:
            from torch.nn.parallel import DistributedDataParallel as DDP
            model = DDP(model, device_ids=[local_rank], gradient_as_bucket_view=True)
        comp, mem, comm = "none (torch DDP)", "none", "nccl all-reduce"

        def exchange_and_step():
            base_opt.step()

        def zero():
            base_opt.zero_grad(set_to_none=False)
    else:
        # the reference's own public API and stock code path: grace_from_params + the manual
        # per-parameter loop of examples/dist/CIFAR10-dawndist/core.py:202-208
        import grace_dl
        assert os.path.realpath(grace_dl.__file__).startswith(os.path.realpath(os.path.join(ROOT, "baseline", "_ref")))
        from grace_dl.dist.helper import grace_from_params
        if world > 1:
            for p in model.parameters():
                dist.broadcast(p.data, 0)
        params = {"compressor": comp, "memory": mem, "communicator": comm, "compress_ratio": args.compress_ratio,
                  "quantum_num": args.quantum_num, "compress_rank": args.compress_rank, "world_size": world}
        grc = grace_from_params(params)
        if comp == "powersgd":       # the dist helper drops rank/world_size: set what it forgot
            grc.compressor.rank, grc.compressor.world_size = args.compress_rank, world
        named = list(model.named_parameters())

        def exchange_and_step():
            for name, p in named:
                new_grad = grc.step(p.grad, name)
                p.grad.copy_(new_grad)
            base_opt.step()

        def zero():
            base_opt.zero_grad(set_to_none=False)

    bs = args.batch_size
    is_bert = args.model.startswith("bert")
    mf = torch.channels_last if args.memory_format == "channels_last" else torch.contiguous_format
    if is_bert:
        S = args.seq_len
        host_x = torch.randint(0, 30522, (bs, S)).pin_memory()                  # token ids (also the MLM labels)
        host_y = torch.randint(0, 2, (bs,)).pin_memory()                        # next-sentence labels
        to_dev = lambda x: x.to(device, non_blocking=True)
        units_per_sample, unit = S, "tokens/s"
    else:
        host_x = torch.randn(bs, 3, 224, 224).pin_memory()
        host_y = torch.randint(0, 1000, (bs,)).pin_memory()
        to_dev = lambda x: x.to(device, non_blocking=True).contiguous(memory_format=mf)
        units_per_sample, unit = 1, "img/s"
    dev_x = to_dev(host_x)
    dev_y = host_y.to(device, non_blocking=True)

    def step(x, y):
        zero()
        with torch.autocast("cuda", dtype=torch.bfloat16):
            if is_bert:
                loss = model(input_ids=x, labels=x, next_sentence_label=y).loss
            else:
                loss = F.cross_entropy(model(x), y)
        loss.backward()
        exchange_and_step()
        return loss

    # ours: the whole step (forward, backward, fused exchange, optimizer) as ONE CUDA graph
    graphed, graph_note = None, "eager"
    if args.impl == "ours" and not args.no_graph:
        try:
            def loss_fn(m, x, y):
                return m(input_ids=x, labels=x, next_sentence_label=y).loss if is_bert else F.cross_entropy(m(x), y)
            graphed = hvd.GraphedStep(model, opt, loss_fn, [dev_x, dev_y], amp_dtype=torch.bfloat16, warmup=3)
            dev_x, dev_y = graphed.static_inputs
            graph_note = "cuda-graph (whole step)"
        except Exception as exc:                      # capture refused: fall back to the eager loop
            graphed, graph_note = None, "eager (graph capture failed: %r)" % (exc,)
            torch.cuda.synchronize()

    if graphed is not None:
        eager_step = step

        def graph_step(x, y):
            return graphed(x, y)

        def quick(fn, n=5):
            for _ in range(3):                        # re-warm (allocator pools differ between the two modes)
                fn(dev_x, dev_y)
            torch.cuda.synchronize()
            a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(n):
                fn(dev_x, dev_y)
            b_.record()
            torch.cuda.synchronize()
            return a.elapsed_time(b_) / n
        # keep whichever execution mode is faster for this model (all ranks must agree)
        t_graph, t_eager = quick(graph_step), quick(eager_step)
        pick = torch.tensor([1.0 if t_graph <= t_eager else 0.0], device=device)
        if world > 1:
            dist.all_reduce(pick, op=dist.ReduceOp.MIN)
        if float(pick) > 0:
            step = graph_step
            graph_note = "cuda-graph (whole step; probe %.2f ms vs %.2f ms eager)" % (t_graph, t_eager)
        else:
            graphed = None
            graph_note = "eager (whole-step graph probed slower: %.2f ms vs %.2f ms)" % (t_graph, t_eager)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(n, e2e):
        barrier()
        launches0 = fused.launches() if fused is not None else 0
        start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.time()
        start.record()
        last = None
        if e2e and args.impl == "ours":
            # The repo's host-feeding API (grace_b200.torch.GraphedStep.feed/run, PinnedPrefetcher, read_back):
            # every step's inputs are copied from pinned host memory inside the timed region -- on a copy stream,
            # overlapping the step in flight -- and every step's loss is read back on the host (one step late,
            # so the read never stalls the next launch).  n copies in, n losses out for n steps.
            feeder = None
            if graphed is not None:
                graphed.feed(host_x, host_y)
            else:
                feeder = hvd.PinnedPrefetcher([host_x, host_y])
                feeder.feed(host_x, host_y)
            prev = None
            for i in range(n):
                if graphed is not None:
                    pend = graphed.run()
                    if i + 1 < n:
                        graphed.feed(host_x, host_y)
                else:
                    x, y = feeder.get()
                    if not is_bert:
                        x = x.contiguous(memory_format=mf)
                    loss = step(x, y)
                    feeder.release()
                    pend = hvd.read_back(loss.detach())
                    if i + 1 < n:
                        feeder.feed(host_x, host_y)
                if prev is not None:
                    last = prev.item()                     # device -> host read of step i-1's loss
                prev = pend
            last = prev.item()
        else:
            for _ in range(n):
                if e2e:                                    # reference arm: the reference examples' own loop shape
                    x = to_dev(host_x)
                    y = host_y.to(device, non_blocking=True)
                    last = step(x, y).item()               # device -> host read of the step's result
                else:
                    last = step(dev_x, dev_y)
        stop.record()
        barrier()
        t1 = time.time()
        ms = start.elapsed_time(stop)
        if world > 1:
            t = torch.tensor([ms], device=device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t)
        launches = (fused.launches() - launches0) if fused is not None else 0
        if graphed is not None:                       # replays do not pass through the host-side counter
            launches = graphed.native_launches_per_step * n
        return ms, launches, (t0, t1), last

    for _ in range(max(args.warmup, 3)):
        step(dev_x, dev_y)
    sampler = ClockSampler(torch.cuda.current_device()) if rank == 0 else None
    ms, launches, window, _ = timed(args.steps, e2e=False)
    clocks = sampler.stop(*window) if sampler else None
    e2e = None
    if not args.no_e2e:
        step(dev_x, dev_y)
        ems, _, _, last = timed(args.steps, e2e=True)
        e2e = {"value": world * bs * units_per_sample * args.steps / (ems / 1e3), "unit": unit, "ms_per_step": ems / args.steps,
               "h2d_bytes_per_step": host_x.numel() * host_x.element_size() + host_y.numel() * 8, "d2h_bytes_per_step": 4,
               "last_loss": last,
               "feeding": ("pipelined through GraphedStep.feed/run / PinnedPrefetcher: each step's pinned-host -> device copy "
                           "is issued inside the timed region on a copy stream (overlapping the previous step), each step's "
                           "loss is read on the host one step late") if args.impl == "ours" else
                          "synchronous per step (copy, step, loss.item()) like the reference examples"}
    if fused is not None:
        fused.engine.check_status()

    if rank == 0:
        value = world * bs * units_per_sample * args.steps / (ms / 1e3)
        out = {
            "metric": "%s synthetic training throughput with %s gradient compression"
                      % ("ResNet-50" if args.model == "resnet50" else args.model, comp),
            "value": value, "unit": unit, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic (fixed random batch, random-init weights)", "impl": args.impl,
            "config": {"model": args.model, "global_batch": world * bs, "per_gpu_batch": bs, "seq_len": args.seq_len if is_bert else None,
                       "image": None if is_bert else "3x224x224", "parallelism": "dp%d" % world, "compressor": comp, "memory": mem,
                       "communicator": comm, "compress_ratio": args.compress_ratio, "params": n_params,
                       "grad_dtype": "fp32 master gradients, bf16 autocast compute", "memory_format": args.memory_format,
                       "optimizer": "SGD momentum 0.9 (inside the timed region)",
                       "l2": "per-step working set (activations, >1 GB) exceeds the 126 MB L2; no explicit flush"},
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches,
        }
        if fused is not None:
            out["config"]["buckets"] = len(fused.buckets)
            out["config"]["transport"] = fused.transport
            out["config"]["execution"] = graph_note
            out["config"]["wire_bytes_per_peer_per_step"] = fused.wire_bytes()
        real_stdout.write(json.dumps(out) + "\n")
        real_stdout.flush()
    if dist.is_initialized():
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
