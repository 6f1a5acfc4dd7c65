"""Receive-side kernels at W = 2/4/8 on ONE GPU (so that `ncu` may wrap the command).

`Engine.local(W)` puts W virtual ranks on one device; every rank's send kernels run first, then rank 0's receive
kernels are timed alone (their flag waits are already satisfied) -- the decode / merge work a real rank does at that
world size, without the NVLink transfer.  ResNet-50 gradient geometry (161 tensors, 102 MB).

    python benchmarks/local_world.py --world 8 --codecs topk,qsgd
    ncu --set full --import-source on -k regex:k1_merge -c 2 ... python benchmarks/local_world.py --world 8 --codecs topk --iters 2
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--world", type=int, default=8)
    ap.add_argument("--codecs", default="topk,qsgd,signsgd,natural")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--ratio", type=float, default=0.01)
    a = ap.parse_args()
    import torch
    import torchvision

    from grace_b200.ops.native import CODEC, MEMORY, CodecParams
    from grace_b200.parallel.engine import Engine

    shapes = [(n, p.numel()) for n, p in torchvision.models.resnet50().named_parameters()]
    names, sizes = [n for n, _ in shapes], [s for _, s in shapes]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for codec in a.codecs.split(","):
        engines = Engine.local(a.world, arena_bytes=(1 << 30))
        prm = CodecParams()
        prm.codec = CODEC[codec]
        prm.memory = MEMORY["dgc" if codec == "dgc" else "none" if codec in ("qsgd", "signsgd", "u8bit", "none", "fp16") else "residual"]
        prm.ratio, prm.beta, prm.gamma, prm.quantum_num, prm.seed, prm.average, prm.capacity = a.ratio, 1.0, 1.0, 64, 1234, 1, 1.0
        prm.threshold, prm.momentum, prm.lr = 2.5, 0.9, 0.1           # threshold: ~1.2 % of N(0,1) gradients pass
        plans = [e.plan(sizes, names, prm) for e in engines]
        grads = [torch.randn(plans[0].padded, device="cuda") for _ in range(a.world)]
        out = torch.empty(plans[0].padded, device="cuda")
        times = []
        for it in range(a.iters + 2):
            for r in range(a.world):
                plans[r].send(grads[r])
            for r in range(a.world):
                plans[r].mid()
            for r in range(1, a.world):
                plans[r].recv(out)
            flush.zero_()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            plans[0].recv(out)
            e1.record()
            torch.cuda.synchronize()
            if it >= 2:
                times.append(e0.elapsed_time(e1))
        for e in engines:
            e.check_status()
        times.sort()
        print("%-10s W=%d (one GPU)  receive side of rank 0: %.3f ms median, %.3f min" % (codec, a.world, times[len(times) // 2], times[0]),
              flush=True)
        for e in engines:
            e.destroy()


if __name__ == "__main__":
    main()
