#!/usr/bin/env python
"""Codec x communicator micro-benchmark (BASELINE.json config 5).

Times one compressed exchange (send + receive kernels) of a gradient bucket for every built-in
codec, device-timed with CUDA events (max over ranks), and reports the achieved fraction of the
roofline  max(HBM bytes / measured copy bandwidth, wire bytes per direction / NVLink bandwidth).

  python benchmarks/micro.py --shape resnet50 --iters 20                    # 1 GPU
  torchrun --nproc-per-node 8 benchmarks/micro.py --shape flat:268435456     # 8 GPUs, 2^28 elements
  python benchmarks/micro.py --reference   # also time the reference's op chain (baseline/_ref) per tensor
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from clocks import ClockSampler  # noqa: E402

CODECS = [("none", "none"), ("fp16", "none"), ("topk", "residual"), ("randomk", "residual"), ("threshold", "residual"),
          ("dgc", "dgc"), ("qsgd", "none"), ("terngrad", "residual"), ("natural", "residual"), ("u8bit", "none"),
          ("signsgd", "none"), ("signum", "residual"), ("efsignsgd", "efsignsgd"), ("onebit", "residual"),
          ("sketch", "residual"), ("adaq", "residual"), ("inceptionn", "residual")]


def shapes_of(spec):
    if spec.startswith("flat:"):
        return [("flat", int(spec.split(":")[1]))]
    import torchvision
    model = getattr(torchvision.models, spec)()
    return [(n, p.numel()) for n, p in model.named_parameters()]


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f)["hbm_gbs"], "measured"
    except Exception:
        return 6650.0, "fallback"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="resnet50")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--codecs", default="")
    ap.add_argument("--ratio", type=float, default=0.01)
    ap.add_argument("--out", default="")
    ap.add_argument("--graph", action="store_true", help="time the exchange as a CUDA-graph replay (how it runs inside GraphedStep)")
    ap.add_argument("--flush", default="write", choices=["write", "write+read"],
                    help="L2 flush between iterations: write 256 MB (leaves the L2 full of DIRTY lines whose write-back "
                         "overlaps the timed kernels) or write 256 MB and then read another 256 MB (cold, clean L2)")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    from grace_b200.parallel.engine import Engine
    from grace_b200.parallel.fused import codec_params
    from grace_b200.registry import build_compressor, build_memory

    named = shapes_of(args.shape)
    numel = sum(n for _, n in named)
    hbm, how = peaks()
    nvlink = 770.0
    eng = Engine.create(arena_bytes=int(os.environ.get("GRACE_B200_ARENA_MB", "4096")) << 20)
    flush_w = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")      # > 126 MB L2
    flush_r = torch.ones(64 << 20, dtype=torch.float32, device="cuda") if args.flush == "write+read" else None

    class _Flush:
        def zero_(self):
            flush_w.zero_()
            if flush_r is not None:
                flush_r.sum()              # evicts the dirty lines of the write before the timed region starts
    flush = _Flush()
    rows = []
    if rank == 0:
        print("engine: world %d, payload transport %s" % (world, "nvls-multimem-store" if eng.multicast else "p2p-store"), flush=True)
    sampler = ClockSampler(torch.cuda.current_device()) if rank == 0 else None
    want = [c for c in args.codecs.split(",") if c]
    for comp, mem in CODECS:
        if want and comp not in want:
            continue
        prm = dict(compressor=comp, memory=mem, compress_ratio=args.ratio, threshold=2.5, quantum_num=64,
                   gradient_clipping=False, momentum=0.9, lr=0.1, world_size=world)
        c = build_compressor(prm, "dist", world)
        m = build_memory(prm, "dist", world, c)
        cp = codec_params(c, m)
        if comp in ("dgc", "adaq"):
            cp.capacity = 0.1              # variable-size payload: slots hold up to 10 % of the bucket
        if comp == "threshold":
            cp.capacity = 0.25             # |x| > 2.5 sigma selects ~1.2 % of fresh noise; the residual raises that
        try:
            plan = eng.plan([n for _, n in named], [k for k, _ in named], cp)
        except Exception as exc:
            if rank == 0:
                print("%-10s skipped: %s" % (comp, str(exc)[:120]), flush=True)
            continue
        grad = torch.randn(plan.padded, device="cuda")
        out = torch.empty(plan.padded, device="cuda")
        times = []
        t_begin = time.time()
        graph = None
        if args.graph:
            # epochs and codec steps live in device memory, so one captured exchange can be replayed
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                plan.send(grad, stream=side); plan.recv(out, stream=side)          # warm-up outside the capture
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                plan.send(grad, stream=torch.cuda.current_stream())
                plan.recv(out, stream=torch.cuda.current_stream())
        for it in range(args.warmup + args.iters):
            if mem != "none":
                grad.normal_()             # fresh gradient every exchange so residuals do not pile up on one sample
            flush.zero_()                                                   # flush L2 between iterations
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            if graph is not None:
                graph.replay()
            else:
                plan.send(grad)
                plan.recv(out)
            b.record()
            torch.cuda.synchronize()
            if it >= args.warmup:
                times.append(a.elapsed_time(b))
        ms = sorted(times)[len(times) // 2]
        if world > 1:
            t = torch.tensor([ms], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t)
        wire = plan.wire_bytes
        # algorithmic HBM traffic: read grad (+state r/w), write payload to W slots, read W slots, write out
        state_passes = {"none": 0, "residual": 2, "dgc": 4, "efsignsgd": 2}[mem]
        hbm_bytes = numel * 4 * (1 + state_passes + 1) + wire * 2 * world
        t_hbm = hbm_bytes / (hbm * 1e9) * 1e3
        t_link = wire * max(world - 1, 0) / (nvlink * 1e9) * 1e3
        roof = max(t_hbm, t_link)
        rows.append(dict(codec=comp, memory=mem, world=world, elements=numel, ms=ms, launches_per_exchange=None,
                         wire_bytes_per_peer=wire, roofline_ms=roof, fraction_of_roofline=roof / ms,
                         grad_gbs=numel * 4 / ms / 1e6, peaks=how,
                         clocks=sampler.window(t_begin, time.time()) if sampler else None))
        n0 = plan.launches()
        plan.send(grad); plan.recv(out)
        rows[-1]["launches_per_exchange"] = plan.launches() - n0
        torch.cuda.synchronize()
        try:
            eng.check_status()             # e.g. threshold + residual on repeated noise outgrows a 10 % slot
        except Exception as exc:
            rows[-1]["status"] = str(exc)[:100]
            if rank == 0:
                print("%-10s status: %s" % (comp, str(exc)[:100]), flush=True)
        plan.destroy()
        if rank == 0:
            r = rows[-1]
            print("%-10s %-9s W=%d  %8.3f ms  %7.1f GB/s of gradient  wire %9d B/peer  roofline %6.3f ms (%.2f of %s)  launches %d"
                  % (comp, mem, world, r["ms"], r["grad_gbs"], wire, roof, r["fraction_of_roofline"], how,
                     r["launches_per_exchange"]), flush=True)
    if world > 1 and (not want or "nvls" in want):
        # dense all-reduce in place on an NVLS multicast-bound bucket (multimem.ld_reduce / multimem.st)
        try:
            import torch.distributed._symmetric_memory as symm
            from grace_b200.parallel.engine import NvlsPlan
            padded = (numel + 127) // 128 * 128
            buf = symm.empty(padded, dtype=torch.float32, device="cuda")
            hd = symm.rendezvous(buf, dist.group.WORLD)
            plan = NvlsPlan(eng, padded, hd.multicast_ptr + (buf.data_ptr() - hd.buffer_ptrs[rank]), True)
            buf.normal_()
            times = []
            t_nv = time.time()
            for it in range(args.warmup + args.iters):
                flush.zero_()
                dist.barrier()
                torch.cuda.synchronize()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                plan.signal(); plan.reduce()
                b.record()
                torch.cuda.synchronize()
                if it >= args.warmup:
                    times.append(a.elapsed_time(b))
            ms = sorted(times)[len(times) // 2]
            t = torch.tensor([ms], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t)
            eng.check_status()
            busbw = 2 * (world - 1) / world * numel * 4 / ms / 1e6
            rows.append(dict(codec="none-nvls", memory="none", world=world, elements=numel, ms=ms, launches_per_exchange=2,
                             bus_bandwidth_gbs=busbw, peaks=how, clocks=sampler.window(t_nv, time.time()) if sampler else None))
            if rank == 0:
                print("%-10s %-9s W=%d  %8.3f ms  all-reduce bus bandwidth %.0f GB/s (NCCL reference on this pool: 725 GB/s at 1 GiB)"
                      % ("none-nvls", "none", world, ms, busbw), flush=True)
        except Exception as exc:
            if rank == 0:
                print("none-nvls  unavailable: %r" % (exc,), flush=True)
    if world > 1 and (not want or "nccl" in want):
        # library baseline: one NCCL all-reduce (+ 1/W scale) of the same 102 MB buffer, timed the same way
        buf = torch.randn((numel + 127) // 128 * 128, device="cuda")
        times = []
        t_nc = time.time()
        for it in range(args.warmup + args.iters):
            flush.zero_()
            dist.barrier()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            dist.all_reduce(buf)
            buf.div_(world)
            b.record()
            torch.cuda.synchronize()
            if it >= args.warmup:
                times.append(a.elapsed_time(b))
        ms = sorted(times)[len(times) // 2]
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t)
        busbw = 2 * (world - 1) / world * numel * 4 / ms / 1e6
        rows.append(dict(codec="nccl-allreduce", memory="none", world=world, elements=numel, ms=ms, launches_per_exchange=2,
                         bus_bandwidth_gbs=busbw, peaks=how, clocks=sampler.window(t_nc, time.time()) if sampler else None))
        if rank == 0:
            print("%-10s %-9s W=%d  %8.3f ms  all-reduce bus bandwidth %.0f GB/s (library baseline: dist.all_reduce + div_)"
                  % ("nccl", "none", world, ms, busbw), flush=True)
    if not want or "powersgd" in want:
        # K3: all >=2-D tensors through the native PowerSGD plan (rank 4, with residual memory)
        from grace_b200.parallel.engine import PowerSGDPlan
        import torchvision
        mats = [(n, tuple(p.shape)) for n, p in getattr(torchvision.models, args.shape)().named_parameters() if p.dim() >= 2] \
            if not args.shape.startswith("flat:") else [("flat", (int(args.shape.split(":")[1]) // 4096, 4096))]
        shapes = [(sh[0], int(torch.tensor(sh[1:]).prod())) for _, sh in mats]
        for use_tc in (True, False):
            plan = PowerSGDPlan(eng, shapes, [n for n, _ in mats], 4, 1234, True, use_tc)
            grad = torch.randn(plan.padded, device="cuda")
            out = torch.empty(plan.padded, device="cuda")
            times = []
            t_ps = time.time()
            for it in range(args.warmup + args.iters):
                flush.zero_()
                if world > 1:
                    dist.barrier()
                torch.cuda.synchronize()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                plan.send(grad); plan.mid(); plan.recv(out)
                b.record()
                torch.cuda.synchronize()
                if it >= args.warmup:
                    times.append(a.elapsed_time(b))
            ms = sorted(times)[len(times) // 2]
            if world > 1:
                t = torch.tensor([ms], device="cuda")
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                ms = float(t)
            eng.check_status()
            n_el = sum(r * c for r, c in shapes)
            # prep reads grad and M and writes M (the residual is implicit), GEMM 1 and GEMM 2 read M, the final pass
            # writes the output: 6 passes over the bucket (8 before the residual buffer was dropped)
            hbm_bytes = n_el * 4 * 6
            roof = max(hbm_bytes / (hbm * 1e9) * 1e3, plan.wire_bytes * max(world - 1, 0) / (nvlink * 1e9) * 1e3)
            n0 = plan.launches(); plan.send(grad); plan.mid(); plan.recv(out); torch.cuda.synchronize()
            rows.append(dict(codec="powersgd-r4" + ("-tcgen05" if use_tc else "-simt"), memory="powersgd", world=world,
                             elements=n_el, ms=ms, launches_per_exchange=plan.launches() - n0,
                             wire_bytes_per_peer=plan.wire_bytes, roofline_ms=roof, fraction_of_roofline=roof / ms,
                             grad_gbs=n_el * 4 / ms / 1e6, peaks=how, tc_matrices=plan.tc_matrices,
                             clocks=sampler.window(t_ps, time.time()) if sampler else None))
            if rank == 0:
                r = rows[-1]
                print("%-22s W=%d  %8.3f ms  %7.1f GB/s of gradient  wire %9d B/peer  roofline %6.3f ms (%.2f of %s)  launches %d  tcgen05 matrices %d/%d"
                      % (r["codec"], world, ms, r["grad_gbs"], plan.wire_bytes, roof, r["fraction_of_roofline"], how,
                         r["launches_per_exchange"], plan.tc_matrices, len(shapes)), flush=True)
            plan.destroy()
    if sampler:
        sampler.close()
    if rank == 0 and args.out:
        with open(args.out, "w") as f:
            json.dump(rows, f, indent=1)
    eng.destroy()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
