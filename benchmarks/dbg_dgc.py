"""Debug helper: per-iteration selected fraction and status of the native DGC plan on ResNet-50 shapes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torchvision
from grace_b200.parallel.engine import Engine
from grace_b200.parallel.fused import codec_params
from grace_b200.registry import build_compressor, build_memory

named = [(n, p.numel()) for n, p in torchvision.models.resnet50().named_parameters()]
eng = Engine.create(arena_bytes=2048 << 20)
prm = dict(compressor="dgc", memory="dgc", compress_ratio=0.01, gradient_clipping=False, momentum=0.9, world_size=1)
c = build_compressor(prm, "dist", 1); m = build_memory(prm, "dist", 1, c)
cp = codec_params(c, m); cp.capacity = 0.1
plan = eng.plan([n for _, n in named], [k for k, _ in named], cp)
grad = torch.randn(plan.padded, device="cuda"); out = torch.empty(plan.padded, device="cuda")
for it in range(12):
    grad.normal_()
    plan.send(grad); plan.recv(out); torch.cuda.synchronize()
    try:
        eng.check_status(); st = "ok"
    except Exception as e:
        st = str(e)[:60]
    nz = [(int((out[o:o + n] != 0).sum()), n) for o, n in zip(plan.offsets, plan.numels)]
    worst = sorted(((a / b, b) for a, b in nz if b > 100000), reverse=True)[:3]
    print(it, "selected %.4f" % (sum(a for a, _ in nz) / sum(b for _, b in nz)), st, worst, flush=True)
