# NOTE: needs the debug hooks: rebuild with  GB_NVCC_EXTRA=-DGB_K3_DEBUG bash grace_b200/csrc/build.sh
import os, sys, ctypes as C
sys.path.insert(0, os.getcwd())
import torch, numpy as np
from grace_b200.ops import native
from grace_b200.parallel.engine import Engine, PowerSGDPlan
L = native.lib()
L.gb_psgd_debug_set.argtypes = [C.POINTER(C.c_int)]
L.gb_psgd_debug_dump.argtypes = [C.c_void_p]
def peek(plan, what, n):
    out = torch.empty(n, device="cuda")
    native.check(L.gb_psgd_peek(plan._h, what, C.c_void_p(out.data_ptr()), n, C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    torch.cuda.synchronize()
    return out
shape = (512, 256)     # mode 2: M-dim = 256 cols (2 tiles), K = 512 rows -> 2 chunks of 256
n, m = shape
g = (torch.arange(n * m, device="cuda").float().view(n, m) % 97) / 97.0 + torch.randn(n, m, device="cuda") * 0.01
first = True
for cfg in [(0, 0, 0, 0, 0, 0, 0), (1, 4096, 512, 1024, 1, 0, 1), (1, 512, 4096, 1024, 1, 0, 1), (1, 4096, 1024, 1024, 1, 0, 1),
            (1, 4096, 512, 1024, 1, 0, 2)]:
    arr = (C.c_int * 8)(*cfg, 0)
    native.check(L.gb_psgd_debug_set(arr))
    (e,) = Engine.local(1)
    plan = PowerSGDPlan(e, [shape], ["w"], 4, 7, False, True)
    buf = torch.zeros(plan.padded, device="cuda"); buf[:n*m] = g.reshape(-1)
    plan.send(buf); plan.mid(); torch.cuda.synchronize()
    ph = peek(plan, 4, n * 4).view(n, 4)
    qacc = peek(plan, 1, m * 4).view(m, 4)
    wantq = g.t() @ ph
    print("cfg(lbo,sbo,kstep,amajor,dump,layout)", cfg[1:], "|Q|", float(qacc.norm()), "|want|", float(wantq.norm()),
          "relerr", float((qacc - wantq).norm() / (wantq.norm() + 1e-9)), "status", e.status(), flush=True)
    if cfg[5]:
        host = np.zeros(8192, dtype=np.float32)
        native.check(L.gb_psgd_debug_dump(host.ctypes.data_as(C.c_void_p)))
        sA = host[:4096].reshape(4, 32, 32)     # [box b][k row][32 floats (swizzled 16B chunks)]
        gm = g.cpu().numpy()
        # un-swizzle: chunk c of row r holds logical chunk c ^ (r % 8)
        for b in range(2):
            for r in (0, 1, 9):
                row = sA[b, r].reshape(8, 4)
                logical = np.concatenate([row[c ^ (r % 8)] for c in range(8)])
                print("  box", b, "krow", r, "max|smem-gmem|", float(np.abs(logical - gm[r, b*32:(b+1)*32]).max()))
        sB = host[4096:4096+512].reshape(16, 32)
        pht = ph.t().cpu().numpy()
        for r in (0, 1, 3, 5):
            row = sB[r].reshape(8, 4)
            logical = np.concatenate([row[c ^ (r % 8)] for c in range(8)])
            print("  B row", r, "max|smem-pht|", float(np.abs(logical - (pht[r, :32] if r < 4 else 0)).max()), "|row|", float(np.abs(logical).max()))
    e.destroy()
