"""K3: native PowerSGD (tcgen05/TMEM GEMMs fed by TMA) against a plain fp32 torch restatement."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from grace_b200.compressor.lowrank import orthogonalize          # noqa: E402
from grace_b200.utils.philox import philox_normal, fnv1a, STREAM_NORMAL   # noqa: E402

SEED = 99
SHAPES = [(512, 4608), (256, 1152), (64, 147), (1000, 2048), (128, 64), (40, 36), (2048, 512), (300, 1000)]
NAMES = ["m%d.weight" % i for i in range(len(SHAPES))]


def oracle(world, rank_r, grads_per_step, memory):
    """grads_per_step[step][rank][i] -> outs[step][i] (identical on every rank)."""
    residual = [[torch.zeros(s, device="cuda") for s in SHAPES] for _ in range(world)]
    outs = []
    for step, grads in enumerate(grads_per_step):
        res = []
        for i, (shape, name) in enumerate(zip(SHAPES, NAMES)):
            n, m = shape
            r = min(n, m, rank_r)
            mats = [grads[w][i] + (residual[w][i] if (memory and step > 0) else 0) for w in range(world)]
            q0 = philox_normal(m * r, SEED, fnv1a(name), step, STREAM_NORMAL, device="cuda").view(m, r)
            q0 = orthogonalize(q0.clone())
            p = sum(mt @ q0 for mt in mats) / world
            p = orthogonalize(p)
            q = sum(mt.t() @ p for mt in mats) / world
            out = p @ q.t()
            for w in range(world):
                residual[w][i] = mats[w] - out
            res.append(out)
        outs.append(res)
    return outs, residual


def run_native(world, rank_r, grads_per_step, memory, use_tc):
    from grace_b200.parallel.engine import Engine, PowerSGDPlan
    engines = Engine.local(world)
    plans = [PowerSGDPlan(e, SHAPES, NAMES, rank_r, SEED, memory, use_tc) for e in engines]
    outs = []
    try:
        for grads in grads_per_step:
            bufs = []
            for w in range(world):
                buf = torch.zeros(plans[w].padded, device="cuda")
                for off, g in zip(plans[w].offsets, grads[w]):
                    buf[off:off + g.numel()] = g.reshape(-1)
                bufs.append(buf)
            for w in range(world):
                plans[w].send(bufs[w])
            for w in range(world):
                plans[w].mid()
            step_out = []
            for w in range(world):
                out = torch.full((plans[w].padded,), float("nan"), device="cuda")
                plans[w].recv(out)
                step_out.append([out[o:o + n * m].view(n, m).clone() for o, (n, m) in zip(plans[w].offsets, SHAPES)])
            torch.cuda.synchronize()
            for e in engines:
                e.check_status()
            outs.append(step_out)
        info = (plans[0].tc_matrices, plans[0].launches())
    finally:
        for e in engines:
            e.destroy()
    return outs, info


def make(world, steps, lowrank=False):
    g = torch.Generator(device="cuda").manual_seed(5)
    out = []
    for _ in range(steps):
        per_rank = []
        for w in range(world):
            ts = []
            for (n, m) in SHAPES:
                if lowrank:
                    a = torch.randn(n, 2, device="cuda", generator=g) @ torch.randn(2, m, device="cuda", generator=g)
                    ts.append(a + 0.01 * torch.randn(n, m, device="cuda", generator=g))
                else:
                    ts.append(torch.randn(n, m, device="cuda", generator=g))
            per_rank.append(ts)
        out.append(per_rank)
    return out


def compare(native, want, tol, what):
    for s, (ns, ws) in enumerate(zip(native, want)):
        for w, per_rank in enumerate(ns):
            for i, (a, b) in enumerate(zip(per_rank, ws)):
                err = float((a - b).norm() / (b.norm() + 1e-12))
                assert err < tol, "%s step %d rank %d matrix %d %s: rel err %g" % (what, s, w, i, SHAPES[i], err)
            for i, a in enumerate(per_rank):
                assert torch.equal(a, ns[0][i]), "ranks disagree on matrix %d" % i


@pytest.mark.parametrize("use_tc", [False, True])
@pytest.mark.parametrize("world,rank_r,memory", [(1, 4, False), (2, 4, True), (4, 2, True), (2, 1, False)])
def test_powersgd_matches_torch(use_tc, world, rank_r, memory):
    grads = make(world, 3, lowrank=True)
    native, (tc_mats, launches) = run_native(world, rank_r, grads, memory, use_tc)
    want, _ = oracle(world, rank_r, grads, memory)
    # CUDA-core path is fp32; the tensor-core path multiplies in tf32 (10-bit mantissa)
    compare(native, want, 2e-2 if use_tc else 2e-4, "tc" if use_tc else "simt")
    assert (tc_mats > 0) == use_tc and launches > 0


def test_tcgen05_gemm_agrees_with_cuda_core_gemm_on_full_rank_noise():
    grads = make(1, 1, lowrank=False)
    a, _ = run_native(1, 4, grads, False, True)
    b, _ = run_native(1, 4, grads, False, False)
    compare(a, b[0:1] and [b[0][0]], 2e-2, "tc-vs-simt")
