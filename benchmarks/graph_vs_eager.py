"""Why is a whole-step CUDA graph faster than the eager loop for ResNet-50 and slower for BERT-large?

Diagnostic only (kineto/CUPTI timeline; numbers taken under a profiler are never bench values): runs the same step
of ``bench.py`` -- model, autocast, fused exchange, optimizer -- eagerly and as a ``GraphedStep`` replay, records the
kernel timeline of a few steps of each, and prints per mode: kernels per step, summed kernel time, the wall span of
a step, idle time on the busiest stream, and the kernels whose duration changes most between the two modes.

    python benchmarks/graph_vs_eager.py --model bert-large --compressor qsgd --out gpurun_out/graph_vs_eager_bert.json
"""
import argparse
import collections
import json
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def kernels_of(trace_path):
    with open(trace_path) as f:
        tr = json.load(f)
    ks = []
    for e in tr.get("traceEvents", []):
        if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset") and "dur" in e:
            ks.append((float(e["ts"]), float(e["dur"]), e.get("name", "?"), e.get("args", {}).get("stream", 0), e["cat"]))
    ks.sort()
    return ks


def summarise(ks, steps):
    if not ks:
        return {}
    span = ks[-1][0] + ks[-1][1] - ks[0][0]
    by_stream = collections.defaultdict(float)
    for ts, dur, _, st, _ in ks:
        by_stream[st] += dur
    main = max(by_stream, key=by_stream.get)
    # union of busy intervals over all streams: time during which at least one kernel runs
    busy, end = 0.0, -1.0
    for ts, dur, _, _, _ in ks:
        if ts > end:
            busy += dur
            end = ts + dur
        elif ts + dur > end:
            busy += ts + dur - end
            end = ts + dur
    gaps = []
    prev_end = None
    for ts, dur, _, st, _ in ks:
        if st != main:
            continue
        if prev_end is not None:
            gaps.append(max(0.0, ts - prev_end))
        prev_end = ts + dur
    gaps.sort()
    return {"kernels_per_step": len(ks) / steps, "sum_kernel_ms_per_step": sum(k[1] for k in ks) / steps / 1e3,
            "span_ms_per_step": span / steps / 1e3, "any_kernel_running_ms_per_step": busy / steps / 1e3,
            "idle_ms_per_step": (span - busy) / steps / 1e3,
            "main_stream_busy_ms_per_step": by_stream[main] / steps / 1e3,
            "median_gap_us_main_stream": gaps[len(gaps) // 2] if gaps else None,
            "mean_gap_us_main_stream": sum(gaps) / len(gaps) if gaps else None,
            "streams": {str(k): v / steps / 1e3 for k, v in by_stream.items()}}


def by_name(ks, steps):
    d = collections.defaultdict(lambda: [0, 0.0])
    for _, dur, name, _, _ in ks:
        d[name][0] += 1
        d[name][1] += dur
    return {k: (v[0] / steps, v[1] / steps) for k, v in d.items()}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="bert-large")
    ap.add_argument("--compressor", default="qsgd")
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--out", default=None)
    ap.add_argument("--keep-mask", action="store_true", help="capture without grace_b200.utils.capture's mask hint")
    a = ap.parse_args()

    import torch
    import torch.nn.functional as F
    from torch.profiler import ProfilerActivity, profile

    import bench
    import grace_b200.torch as hvd

    sys.argv = ["bench.py", "--model", a.model, "--compressor", a.compressor]
    args = bench.parse()
    device = torch.device("cuda", 0)
    torch.cuda.set_device(device)
    comp, mem, comm = bench.triple(args, a.compressor)
    model, memory_format = bench.build_model(args, comp, device)
    base_opt = torch.optim.SGD(model.parameters(), lr=0.01, momentum=0.9)
    grc = hvd.grace_from_params({"compressor": comp, "memory": mem, "communicator": comm, "compress_ratio": args.compress_ratio,
                                 "quantum_num": args.quantum_num, "compress_rank": args.compress_rank, "world_size": 1})
    opt = hvd.DistributedOptimizer(base_opt, grc, named_parameters=model.named_parameters(), fused="auto")
    is_bert = a.model.startswith("bert")
    bs = args.batch_size
    if is_bert:
        x = torch.randint(0, 30522, (bs, args.seq_len), device=device)
        y = torch.randint(0, 2, (bs,), device=device)
    else:
        mf = torch.channels_last if memory_format == "channels_last" else torch.contiguous_format
        x = torch.randn(bs, 3, 224, 224, device=device).contiguous(memory_format=mf)
        y = torch.randint(0, 1000, (bs,), device=device)

    import contextlib
    from grace_b200.utils.capture import unmasked_attention_when_capturing
    hint = contextlib.nullcontext if a.keep_mask else unmasked_attention_when_capturing

    def loss_fn(m, x, y):
        if not is_bert:
            return F.cross_entropy(m(x), y)
        with hint():
            return m(input_ids=x, labels=x, next_sentence_label=y).loss

    def eager(x, y):
        opt.zero_grad()
        with torch.autocast("cuda", dtype=torch.bfloat16):
            loss = loss_fn(model, x, y)
        loss.backward()
        opt.step()
        return loss

    graphed = hvd.GraphedStep(model, opt, loss_fn, [x, y], amp_dtype=torch.bfloat16, warmup=3)
    x, y = graphed.static_inputs

    def timed(fn, n=5):
        for _ in range(3):
            fn(x, y)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            fn(x, y)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n

    out = {"model": a.model, "compressor": a.compressor, "mask_hint": not a.keep_mask,
           "unprofiled_ms_per_step": {"graph": timed(graphed), "eager": timed(eager)}}
    tmp = tempfile.mkdtemp()
    traces = {}
    for mode, fn in (("eager", eager), ("graph", graphed)):
        for _ in range(3):
            fn(x, y)
        torch.cuda.synchronize()
        with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
            for _ in range(a.steps):
                fn(x, y)
            torch.cuda.synchronize()
        path = os.path.join(tmp, mode + ".json")
        prof.export_chrome_trace(path)
        traces[mode] = kernels_of(path)
        out[mode] = summarise(traces[mode], a.steps)
    ne, ng = by_name(traces["eager"], a.steps), by_name(traces["graph"], a.steps)
    diff = []
    for name in set(ne) | set(ng):
        ce, te = ne.get(name, (0, 0.0))
        cg, tg = ng.get(name, (0, 0.0))
        diff.append((tg - te, name[:90], ce, te / 1e3, cg, tg / 1e3))
    diff.sort(reverse=True)
    out["slower_in_graph_top"] = [dict(name=n, eager_count=ce, eager_ms=te, graph_count=cg, graph_ms=tg, delta_ms=d / 1e3)
                                  for d, n, ce, te, cg, tg in diff[:12]]
    out["faster_in_graph_top"] = [dict(name=n, eager_count=ce, eager_ms=te, graph_count=cg, graph_ms=tg, delta_ms=d / 1e3)
                                  for d, n, ce, te, cg, tg in diff[-6:]]
    for mode, names in (("eager", ne), ("graph", ng)):
        top = sorted(names.items(), key=lambda kv: -kv[1][1])[:18]
        out["top_kernels_" + mode] = [dict(name=k[:110], count=c, ms=t / 1e3) for k, (c, t) in top]
    text = json.dumps(out, indent=1)
    print(text)
    if a.out:
        os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
        with open(a.out, "w") as f:
            f.write(text)


if __name__ == "__main__":
    main()
