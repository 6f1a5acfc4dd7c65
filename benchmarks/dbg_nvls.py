import os, sys
sys.path.insert(0, os.getcwd())
import torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
from grace_b200.parallel.fused import get_engine
from grace_b200.parallel.engine import NvlsPlan
eng = get_engine()
n = 128 * 1000
flat = symm.empty(n, dtype=torch.float32, device="cuda")
h = symm.rendezvous(flat, dist.group.WORLD)
print(rank, "data_ptr", hex(flat.data_ptr()), "buffer_ptrs", [hex(p) for p in h.buffer_ptrs], "mc", hex(h.multicast_ptr),
      "offset", getattr(h, "offset", None), "buffer_size", h.buffer_size, flush=True)
plan = NvlsPlan(eng, n, h.multicast_ptr + (flat.data_ptr() - h.buffer_ptrs[rank]), True)
for trial in range(3):
    flat.copy_(torch.arange(n, device="cuda", dtype=torch.float32) * (rank + 1) + trial)
    torch.cuda.synchronize(); dist.barrier()
    plan.signal(); plan.reduce(); torch.cuda.synchronize()
    want = torch.arange(n, device="cuda", dtype=torch.float32) * (sum(range(1, world + 1)) / world) + trial
    bad = (flat - want).abs() > 1e-3 * want.abs().clamp(min=1)
    print(rank, "trial", trial, "bad", int(bad.sum()), "first bad", bad.nonzero()[:4].flatten().tolist(),
          "vals", flat[bad][:4].tolist(), "want", want[bad][:4].tolist(), "status", eng.status(), flush=True)
dist.destroy_process_group()
