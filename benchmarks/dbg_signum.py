"""Debug helper: Signum + residual through the fused DistributedOptimizer vs the manual generic loop, step by step."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch, torch.nn.functional as F
from test_fused_gpu import _data
from grace_b200.models import MnistNet
import grace_b200.torch as hvd
from grace_b200.dist import grace_from_params
import grace_b200.parallel.fused as fused_mod

params = dict(compressor="signum", memory="residual", communicator="allgather", momentum=0.9)
data = _data()

def run(mode):
    torch.manual_seed(0)
    model = MnistNet().cuda().eval()
    opt = torch.optim.SGD(model.parameters(), lr=0.05, momentum=0.5)
    outs = []
    if mode == "manual":
        grc = grace_from_params(dict(params, world_size=1)); grc.fused = False
    else:
        grc = hvd.grace_from_params(dict(params, world_size=1))
        opt = hvd.DistributedOptimizer(opt, grc, named_parameters=model.named_parameters(), fused=True, bucket_bytes=40000)
    for x, y in data:
        opt.zero_grad(); F.nll_loss(model(x), y).backward()
        if mode == "manual":
            for name, p in model.named_parameters():
                p.grad.copy_(grc.step(p.grad, name))
        else:
            opt.synchronize()
        outs.append(torch.cat([p.grad.detach().flatten().float() for p in model.parameters()]).clone())
        if mode == "manual":
            opt.step()
        else:
            with opt.skip_synchronize():
                opt.step()
    return outs

a = run("manual")
b = run("fused")
orig = fused_mod.codec_params
def patched(c, m):
    p = orig(c, m)
    if p is not None:
        p.one_minus_momentum = 0.0
    return p
fused_mod.codec_params = patched
c = run("fused")
for s in range(len(a)):
    print("step", s, "fused(om) mismatches", int((a[s] != b[s]).sum()), " fused(fallback) mismatches", int((a[s] != c[s]).sum()), "of", a[s].numel(),
          "| values", sorted(set(a[s].tolist()))[:4], sorted(set(b[s].tolist()))[:4], flush=True)
