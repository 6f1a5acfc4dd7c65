"""Is the ResNet-50 batch-32 step launch-bound?  Eager vs whole-step CUDA graph (no gradient exchange)."""
import torch, torchvision, time
import torch.nn.functional as F
torch.backends.cudnn.benchmark = True
m = torchvision.models.resnet50().cuda().to(memory_format=torch.channels_last)
opt = torch.optim.SGD(m.parameters(), lr=0.01, momentum=0.9)
x = torch.randn(32, 3, 224, 224, device="cuda").contiguous(memory_format=torch.channels_last)
y = torch.randint(0, 1000, (32,), device="cuda")
def step():
    opt.zero_grad(set_to_none=False)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        loss = F.cross_entropy(m(x), y)
    loss.backward()
    opt.step()
    return loss
for p in m.parameters():
    p.grad = torch.zeros_like(p)
def timeit(fn, n=30):
    for _ in range(5): fn()
    torch.cuda.synchronize(); a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n
print("eager ms/step", timeit(step))
s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    for _ in range(3): step()
torch.cuda.current_stream().wait_stream(s)
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    step()
print("graph ms/step", timeit(g.replay))
