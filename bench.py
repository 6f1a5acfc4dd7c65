#!/usr/bin/env python
"""Headline benchmark: ResNet-50 synthetic training throughput (img/s, whole job) with gradient
compression -- the harness of the reference's ``examples/torch/pytorch_synthetic_benchmark.py``
(batch 32 per GPU, fixed fake data 3x224x224), bf16 compute, for the three codecs BASELINE.json's
metric names: TopK 1 % + ResidualMemory + Allgather (the headline ``value``), QSGD + Allgather and
PowerSGD rank 4 + PowerSGDMemory + Allreduce (``sub_results``).

  python bench.py --gpus N --steps K --warmup W                # this framework: all three codecs
  python bench.py --impl reference --gpus N --steps K ...      # UNMODIFIED reference from baseline/_ref
  python bench.py --compressor qsgd ...                        # one codec only

For N > 1 launch through ``python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N``.
Rank 0 prints ONE JSON line.  ``value`` is device-timed (CUDA events, max over ranks) with the
batch resident on the GPU like the reference harness; ``e2e`` repeats the measurement with a
per-step pinned-host -> device copy of the inputs and a device -> host read of the loss.  Our arm
also reports how the gap to the reference decomposes (``decomposition``: the same engine run eagerly,
and the graph fed synchronously like the reference loop) and, after every timed region, proves that
what it timed was correct: all ranks hold bit-identical parameters (``replicas_identical``) and two
exchanges of rank-dependent gradients at the model's real geometry match the generic
torch.distributed implementation of the same triple (``matches_generic``) and are bit-identical on
all ranks.  A failed check makes the process exit non-zero.
"""
import argparse
import gc
import json
import os
import socket
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
HEADLINE_CODECS = ["topk", "qsgd", "powersgd"]


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
    ap.add_argument("--compressor", default="all", help="'all' = topk (headline) + qsgd + powersgd sub-results; or a comma list")
    ap.add_argument("--memory", default=None)
    ap.add_argument("--communicator", default=None)
    ap.add_argument("--compress-ratio", type=float, default=0.01)
    ap.add_argument("--quantum-num", type=int, default=64)
    ap.add_argument("--compress-rank", type=int, default=4)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="ours: run the step eagerly instead of as one CUDA graph")
    ap.add_argument("--no-check", action="store_true", help="ours: skip the replica / generic-path correctness checks")
    ap.add_argument("--no-decomposition", action="store_true", help="ours: skip the eager and synchronous-feed timings")
    ap.add_argument("--api", default="torch", choices=["torch", "dist"],
                    help="ours: torch = DistributedOptimizer (hooks, overlap); dist = the reference's manual loop shape "
                         "through grace_b200.dist (bucketed step_all)")
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


def triple(args, comp):
    mem = args.memory or {"topk": "residual", "qsgd": "none", "powersgd": "powersgd", "efsignsgd": "efsignsgd",
                          "dgc": "dgc", "none": "none"}.get(comp, "residual")
    comm = args.communicator or ("allreduce" if comp in ("powersgd", "none", "fp16") else "allgather")
    if args.impl == "reference" and mem == "efsignsgd":
        mem = "residual"      # the reference's dist helper has no 'efsignsgd' memory branch (helper.py:57-74); its own
                              # older EFSignSGD wires ResidualMemory(beta=1, gamma=lr) (compression.py:255-257)
    return comp, mem, comm


def build_model(args, comp, device):
    import torch
    torch.manual_seed(0)
    if args.model.startswith("bert"):
        # BASELINE.json config 4: BERT pre-training (MLM + NSP heads), random init, synthetic token ids
        from transformers import BertConfig, BertForPreTraining
        large = args.model == "bert-large"
        cfg = BertConfig(hidden_size=1024 if large else 768, num_hidden_layers=24 if large else 12,
                         num_attention_heads=16 if large else 12, intermediate_size=4096 if large else 3072)
        return BertForPreTraining(cfg).to(device), "contiguous"
    import torchvision
    model = getattr(torchvision.models, args.model)().to(device)
    fmt = args.memory_format
    if fmt == "auto":
        # the reference's PowerSGD does tensor.view(shape[0], -1) (compressor/powersgd.py:35), which
        # cannot take channels_last gradients; every other configuration runs channels_last
        fmt = "contiguous" if (args.impl == "reference" and comp == "powersgd") else "channels_last"
    if fmt == "channels_last":
        model = model.to(memory_format=torch.channels_last)
    return model, fmt


def param_digest(tensors):
    """Two 64-bit checksums of the bit patterns of a list of tensors (order-sensitive), as a device tensor."""
    import torch
    acc = torch.zeros(2, dtype=torch.int64, device=tensors[0].device)
    for i, t in enumerate(tensors):
        bits = t.detach().contiguous().view(-1)
        bits = bits.view(torch.int32) if bits.element_size() == 4 else bits.view(torch.int16).to(torch.int32)
        b64 = bits.to(torch.int64)
        acc[0] += b64.sum() * (2 * i + 1)
        acc[1] += (b64 * (torch.arange(b64.numel(), device=b64.device) % 8191 + 1)).sum()
    return acc


def same_on_all_ranks(digest, world):
    import torch
    import torch.distributed as dist
    if world == 1:
        return True
    got = [torch.zeros_like(digest) for _ in range(world)]
    dist.all_gather(got, digest)
    return all(bool(torch.equal(g, got[0])) for g in got)


def run_codec(args, comp_name, env):
    """Build the model + optimizer for one (compressor, memory, communicator) triple, time it, check it."""
    import torch
    import torch.distributed as dist
    import torch.nn.functional as F

    world, rank, local_rank, device = env["world"], env["rank"], env["local_rank"], env["device"]
    comp, mem, comm = triple(args, comp_name)
    model, memory_format = build_model(args, comp, device)
    base_opt = torch.optim.SGD(model.parameters(), lr=0.01, momentum=0.9)
    n_params = sum(p.numel() for p in model.parameters())
    gparams = {"compressor": comp, "memory": mem, "communicator": comm, "compress_ratio": args.compress_ratio,
               "quantum_num": args.quantum_num, "compress_rank": args.compress_rank, "world_size": world}

    fused, opt, hvd, dist_grc = None, None, None, None
    if args.impl == "ours" and args.api == "torch":
        import grace_b200.torch as hvd
        hvd.broadcast_parameters(model.state_dict(), root_rank=0)
        grc = hvd.grace_from_params(gparams)
        opt = hvd.DistributedOptimizer(base_opt, grc, named_parameters=model.named_parameters(), fused="auto")
        fused = opt._fused

        def exchange_and_step():
            opt.step()

        def zero():
            opt.zero_grad()
    elif args.impl == "ours":
        # the reference's integration for the torch.distributed flavour is a manual loop over the parameters
        # (examples/dist/CIFAR10-dawndist/core.py:202-208); grace_b200.dist offers the same call per tensor
        # and, used here, the bucketed form of that loop: one step_all over (name, grad) pairs
        import grace_b200.torch as hvd
        from grace_b200.dist import grace_from_params as dist_from_params
        hvd.broadcast_parameters(model.state_dict(), root_rank=0)
        dist_grc = dist_from_params(gparams)
        named = [(n, p) for n, p in model.named_parameters() if p.requires_grad]

        def exchange_and_step():
            dist_grc.step_all([(n, p.grad) for n, p in named], inplace=True)
            base_opt.step()

        def zero():
            base_opt.zero_grad(set_to_none=False)
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
        grc = grace_from_params(dict(gparams))
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
    mf = torch.channels_last if memory_format == "channels_last" else torch.contiguous_format
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

    def eager_step(x, y):
        zero()
        with torch.autocast("cuda", dtype=torch.bfloat16):
            if is_bert:
                loss = model(input_ids=x, labels=x, next_sentence_label=y).loss
            else:
                loss = F.cross_entropy(model(x), y)
        loss.backward()
        exchange_and_step()
        return loss

    step = eager_step
    # ours: the whole step (forward, backward, fused exchange, optimizer) as ONE CUDA graph
    graphed, graph_note = None, "eager"
    if args.impl == "ours" and (opt is not None or dist_grc is not None) and not args.no_graph:
        try:
            from grace_b200.utils.capture import unmasked_attention_when_capturing

            def loss_fn(m, x, y):
                if not is_bert:
                    return F.cross_entropy(m(x), y)
                # no attention_mask is passed: keep transformers from materialising an all-ones mask while the
                # stream captures (that mask sends SDPA to the fp32 math path; see grace_b200/utils/capture.py)
                with unmasked_attention_when_capturing():
                    return m(input_ids=x, labels=x, next_sentence_label=y).loss
            if opt is not None:
                graphed = hvd.GraphedStep(model, opt, loss_fn, [dev_x, dev_y], amp_dtype=torch.bfloat16, warmup=3)
            else:        # dist flavour: the manual loop, bucketed (step_all), captured with the plain optimizer
                graphed = hvd.GraphedStep(model, base_opt, loss_fn, [dev_x, dev_y], amp_dtype=torch.bfloat16, warmup=3,
                                          communicator=dist_grc)
            dev_x, dev_y = graphed.static_inputs
            graph_note = "cuda-graph (whole step)"
        except Exception as exc:                      # capture refused: fall back to the eager loop
            graphed, graph_note = None, "eager (graph capture failed: %r)" % (exc,)
            torch.cuda.synchronize()

    def quick(fn, n=5):
        for _ in range(3):                            # re-warm (allocator pools differ between the two modes)
            fn(dev_x, dev_y)
        torch.cuda.synchronize()
        a, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(n):
            fn(dev_x, dev_y)
        b_.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b_) / n

    if graphed is not None:
        def graph_step(x, y):
            return graphed(x, y)
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

    def native_launch_count():
        if fused is not None:
            return fused.launches()
        if dist_grc is not None and getattr(dist_grc, "_stepper", None) is not None:
            return dist_grc._stepper.launches()
        return 0

    def timed(n, mode, fn=None):
        """mode: 'device' (batch resident on the GPU), 'pipelined' (our host-feeding API), 'sync' (copy, step,
        loss.item() per step -- the loop shape of the reference examples)."""
        fn = fn or step
        use_graph = graphed is not None and fn is step
        barrier()
        launches0 = native_launch_count()
        start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.time()
        start.record()
        last = None
        if mode == "pipelined":
            # The repo's host-feeding API (grace_b200.torch.GraphedStep.feed/run, PinnedPrefetcher, read_back):
            # every step's inputs are copied from pinned host memory inside the timed region -- on a copy stream,
            # overlapping the step in flight -- and every step's loss is read back on the host (one step late,
            # so the read never stalls the next launch).  n copies in, n losses out for n steps.
            feeder = None
            if use_graph:
                graphed.feed(host_x, host_y)
            else:
                feeder = hvd.PinnedPrefetcher([host_x, host_y])
                feeder.feed(host_x, host_y)
            prev = None
            for i in range(n):
                if use_graph:
                    pend = graphed.run()
                    if i + 1 < n:
                        graphed.feed(host_x, host_y)
                else:
                    x, y = feeder.get()
                    if not is_bert:
                        x = x.contiguous(memory_format=mf)
                    loss = fn(x, y)
                    feeder.release()
                    pend = hvd.read_back(loss.detach())
                    if i + 1 < n:
                        feeder.feed(host_x, host_y)
                if prev is not None:
                    last = prev.item()                     # device -> host read of step i-1's loss
                prev = pend
            last = prev.item()
        elif mode == "sync":
            for _ in range(n):
                x = to_dev(host_x)
                y = host_y.to(device, non_blocking=True)
                last = fn(x, y).item()                     # device -> host read of the step's result
        else:
            for _ in range(n):
                last = fn(dev_x, dev_y)
        stop.record()
        barrier()
        t1 = time.time()
        ms = start.elapsed_time(stop)
        if world > 1:
            t = torch.tensor([ms], device=device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t)
        launches = native_launch_count() - launches0
        if use_graph:                                  # replays do not pass through the host-side counter
            launches = graphed.native_launches_per_step * n
        return ms, launches, (t0, t1), last

    def throughput(ms, n):
        return world * bs * units_per_sample * n / (ms / 1e3)

    for _ in range(max(args.warmup, 3)):
        step(dev_x, dev_y)
    sampler = ClockSampler(torch.cuda.current_device()) if rank == 0 else None
    ms, launches, window, _ = timed(args.steps, "device")
    clocks = sampler.stop(*window) if sampler else None
    e2e = None
    if not args.no_e2e:
        step(dev_x, dev_y)
        e2e_mode = "pipelined" if args.impl == "ours" else "sync"
        ems, _, _, last = timed(args.steps, e2e_mode)
        e2e = {"value": throughput(ems, args.steps), "unit": unit, "ms_per_step": ems / args.steps,
               "h2d_bytes_per_step": host_x.numel() * host_x.element_size() + host_y.numel() * 8, "d2h_bytes_per_step": 4,
               "last_loss": last,
               "feeding": ("pipelined through GraphedStep.feed/run / PinnedPrefetcher: each step's pinned-host -> device copy "
                           "is issued inside the timed region on a copy stream (overlapping the previous step), each step's "
                           "loss is read on the host one step late") if args.impl == "ours" else
                          "synchronous per step (copy, step, loss.item()) like the reference examples"}

    # how the gap to the reference arm decomposes: same engine without the graph, and the graph fed like the reference loop
    decomposition = None
    if args.impl == "ours" and not args.no_decomposition:
        decomposition = {}
        if graphed is not None:
            for _ in range(3):
                eager_step(dev_x, dev_y)
            dms, _, _, _ = timed(args.steps, "device", fn=eager_step)
            decomposition["eager"] = {"value": throughput(dms, args.steps), "ms_per_step": dms / args.steps,
                                      "what": "same fused engine and overlap, step issued eagerly (no CUDA graph), batch resident"}
            for _ in range(3):
                step(dev_x, dev_y)
        if not args.no_e2e:
            sms, _, _, _ = timed(args.steps, "sync")
            decomposition["sync_feed"] = {"value": throughput(sms, args.steps), "ms_per_step": sms / args.steps,
                                          "what": "e2e with the reference loop shape: blocking copy, step, loss.item() every step"}

    checks = None
    if args.impl == "ours" and not args.no_check:
        torch.cuda.synchronize()
        if fused is not None:
            fused.engine.check_status()
        checks = {"replicas_identical": same_on_all_ranks(param_digest(list(model.parameters())), world)}

    result = {
        "metric": "%s synthetic training throughput with %s gradient compression"
                  % ("ResNet-50" if args.model == "resnet50" else args.model, comp),
        "value": throughput(ms, args.steps), "unit": unit, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic (fixed random batch, random-init weights)", "impl": args.impl,
        "config": {"model": args.model, "global_batch": world * bs, "per_gpu_batch": bs, "seq_len": args.seq_len if is_bert else None,
                   "image": None if is_bert else "3x224x224", "parallelism": "dp%d" % world, "compressor": comp, "memory": mem,
                   "communicator": comm, "compress_ratio": args.compress_ratio, "quantum_num": args.quantum_num,
                   "compress_rank": args.compress_rank, "params": n_params,
                   "grad_dtype": "fp32 master gradients, bf16 autocast compute", "memory_format": memory_format,
                   "optimizer": "SGD momentum 0.9 (inside the timed region)",
                   "l2": "per-step working set (activations, >1 GB) exceeds the 126 MB L2; no explicit flush"},
        "clocks": clocks, "e2e": e2e, "gpu_launches": launches,
    }
    if args.impl == "ours":
        result["config"]["api"] = args.api
        result["config"]["execution"] = graph_note
        if fused is not None:
            result["config"]["buckets"] = len(fused.buckets)
            result["config"]["transport"] = fused.transport
            result["config"]["wire_bytes_per_peer_per_step"] = fused.wire_bytes()
        if decomposition is not None:
            result["decomposition"] = decomposition

    # ---- tear down, then (ours) two exchanges at the model's geometry against the generic implementation
    shapes = [(n, tuple(p.shape)) for n, p in model.named_parameters() if p.requires_grad]
    graphed = step = eager_step = None
    if opt is not None:
        opt.close()
    if dist_grc is not None and hasattr(dist_grc, "close"):
        dist_grc.close()
    opt = fused = dist_grc = model = base_opt = exchange_and_step = zero = None
    gc.collect()
    torch.cuda.empty_cache()
    if checks is not None:
        checks.update(exchange_check(gparams, shapes, memory_format, env))
        result.update(checks)
        result["checks_ok"] = bool(checks["replicas_identical"] and checks["matches_generic"]
                                   and checks["exchange_identical_on_all_ranks"])
    gc.collect()
    torch.cuda.empty_cache()
    return result


def exchange_check(gparams, shapes, memory_format, env):
    """Two exchanges of rank-dependent random gradients with the model's tensor shapes: the bucketed native path
    (step_all) against the generic torch.distributed implementation of the same triple, and across ranks.
    The stochastic codecs and PowerSGD are checked on contiguous gradients: their random streams / projections are
    indexed by memory order (native) and by view(-1) order (generic), which coincide only for contiguous tensors."""
    import torch
    from grace_b200.dist import grace_from_params
    world, rank = env["world"], env["rank"]
    fast, slow = grace_from_params(dict(gparams)), grace_from_params(dict(gparams))
    slow.fused = False
    gen = torch.Generator(device="cuda").manual_seed(4242 + rank)
    corrupt = os.environ.get("GRACE_B200_BENCH_CORRUPT")            # self-test of the checker: see BASELINE.md
    num = den = 0.0
    identical = True
    channels_last = memory_format == "channels_last" and gparams["compressor"] == "topk"
    for it in range(2):
        grads = []
        for name, shape in shapes:
            g = torch.randn(shape, device="cuda", generator=gen) * 0.01
            if len(shape) == 4 and channels_last:
                g = g.contiguous(memory_format=torch.channels_last)
            grads.append((name, g))
        outs = fast.step_all([(n, g.clone()) for n, g in grads])
        if corrupt is not None and rank == int(corrupt) % world:
            outs[0].view(-1)[0] += 1.0
        identical = identical and same_on_all_ranks(param_digest(outs), world)
        for (name, g), a in zip(grads, outs):
            b = slow.step(g.clone(), name)
            num += float((a.float() - b.float()).pow(2).sum())
            den += float(b.float().pow(2).sum())
    rel = (num / max(den, 1e-30)) ** 0.5
    tol = {"topk": 1e-6, "qsgd": 5e-3, "powersgd": 3e-2}.get(gparams["compressor"], 5e-3)
    native = getattr(fast, "_stepper", None) is not None and bool(fast._stepper.enabled)
    if hasattr(fast, "close"):
        fast.close()
    t = torch.tensor([rel], device="cuda")
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return {"matches_generic": bool(float(t) <= tol) and native, "generic_rel_err": float(t), "generic_tol": tol,
            "exchange_identical_on_all_ranks": identical, "exchange_check_native": native}


def main():
    args = parse()
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
    env = {"world": world, "rank": rank, "local_rank": local_rank, "device": device}

    if args.compressor == "all":
        codecs = HEADLINE_CODECS if not args.model.startswith("bert") else ["qsgd", "efsignsgd"]
    else:
        codecs = [c for c in args.compressor.split(",") if c]
    results = [run_codec(args, c, env) for c in codecs]
    out = results[0]
    if len(results) > 1:
        out["sub_results"] = {r["config"]["compressor"]: r for r in results[1:]}
    ok = all(r.get("checks_ok", True) for r in results)
    if rank == 0:
        real_stdout.write(json.dumps(out) + "\n")
        real_stdout.flush()
    if dist.is_initialized():
        dist.destroy_process_group()
    if not ok:
        sys.stderr.write("bench.py: correctness check FAILED: %s\n" % json.dumps(
            {r["config"]["compressor"]: {k: r.get(k) for k in ("replicas_identical", "matches_generic", "generic_rel_err",
                                                                "exchange_identical_on_all_ranks")} for r in results}))
        return 3
    return 0


if __name__ == "__main__":
    sys.exit(main())
