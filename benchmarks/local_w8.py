"""Single-GPU emulation of a W-rank exchange (Engine.local): every virtual rank sends, then every rank receives.
Times one rank's send and one rank's receive kernel sequence separately -- the compute side of a W-rank exchange
without NVLink (all "peers" are the same device)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torchvision
from grace_b200.parallel.engine import Engine
from grace_b200.parallel.fused import codec_params
from grace_b200.registry import build_compressor, build_memory

W = int(sys.argv[1]) if len(sys.argv) > 1 else 8
codecs = sys.argv[2].split(",") if len(sys.argv) > 2 else ["signsgd", "efsignsgd", "qsgd", "topk"]
named = [(n, p.numel()) for n, p in torchvision.models.resnet50().named_parameters()]
engines = Engine.local(W, arena_bytes=6 << 30)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for comp in codecs:
    mem = {"topk": "residual", "efsignsgd": "efsignsgd"}.get(comp, "none")
    prm = dict(compressor=comp, memory=mem, compress_ratio=0.01, quantum_num=64, lr=0.1, momentum=0.9, world_size=W)
    c = build_compressor(prm, "dist", W); m = build_memory(prm, "dist", W, c)
    plans = [e.plan([n for _, n in named], [k for k, _ in named], codec_params(c, m)) for e in engines]
    grads = [torch.randn(plans[0].padded, device="cuda") for _ in range(W)]
    out = torch.empty(plans[0].padded, device="cuda")
    ts, tr = [], []
    for it in range(8):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        for r in range(1, W):
            plans[r].send(grads[r])
        flush.zero_()
        ev[0].record(); plans[0].send(grads[0]); ev[1].record()
        flush.zero_()
        ev[2].record(); plans[0].recv(out); ev[3].record()
        for r in range(1, W):
            plans[r].recv(out)
        torch.cuda.synchronize()
        if it >= 2:
            ts.append(ev[0].elapsed_time(ev[1])); tr.append(ev[2].elapsed_time(ev[3]))
    for e in engines:
        e.check_status()
    print("%-10s local W=%d  send %.3f ms  recv %.3f ms" % (comp, W, sorted(ts)[len(ts) // 2], sorted(tr)[len(tr) // 2]), flush=True)
    for p in plans:
        p.destroy()
