"""Numerics of the fused sm_100a kernels against the pure-torch codecs (fp32 oracle).

W virtual ranks live on one GPU (``Engine.local``): every rank's send kernels run first,
then every rank's receive kernels, so flag waits never block (B200_PROFILING.md: kernels that
wait on one another must not be separate launches on one GPU)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from grace_b200 import compressor as C          # noqa: E402
from grace_b200 import memory as M              # noqa: E402

SIZES = [1000, 64, 4096, 5001, 1, 12345, 70000, 3]
NAMES = ["seg%d.weight" % i for i in range(len(SIZES))]
SEED = 1234


def _params(codec, memory="none", **kw):
    from grace_b200.ops.native import CodecParams, CODEC, MEMORY
    p = CodecParams()
    p.codec, p.memory = CODEC[codec], MEMORY[memory]
    p.ratio = kw.get("ratio", 0.05)
    p.beta, p.gamma = kw.get("beta", 1.0), kw.get("gamma", 1.0)
    p.threshold = kw.get("threshold", 1.5)
    p.quantum_num = kw.get("quantum_num", 64)       # also SketchML's number of quantiles
    p.momentum = kw.get("momentum", 0.9)
    p.lr = kw.get("lr", 0.1)
    p.seed = SEED
    p.average = kw.get("average", 1)
    p.capacity = 1.0
    return p


def _flat(plan, tensors):
    buf = torch.zeros(plan.padded, device="cuda")
    for off, t in zip(plan.offsets, tensors):
        buf[off:off + t.numel()] = t
    return buf


def _split(plan, buf):
    return [buf[off:off + n].clone() for off, n in zip(plan.offsets, plan.numels)]


def run_fused(world, params, grads_per_step, sizes=None, names=None, arena_bytes=None):
    """grads_per_step[step][rank] = list of tensors -> outs[step][rank] = list of tensors."""
    from grace_b200.parallel.engine import Engine
    sizes, names = sizes or SIZES, names or NAMES
    engines = Engine.local(world, arena_bytes=arena_bytes)
    plans = [e.plan(sizes, names, params) for e in engines]
    outs = []
    try:
        for grads in grads_per_step:
            bufs = [_flat(plans[r], grads[r]) for r in range(world)]
            for r in range(world):
                plans[r].presend(bufs[r])           # only DGC with clipping has one (its send waits for the peers)
            for r in range(world):
                plans[r].send(bufs[r])
            for r in range(world):
                plans[r].mid()
            res = []
            for r in range(world):
                out = torch.full((plans[r].padded,), float("nan"), device="cuda")
                plans[r].recv(out)
                res.append(_split(plans[r], out))
            torch.cuda.synchronize()
            for e in engines:
                e.check_status()
            outs.append(res)
        states = [[p.get_state(0), p.get_state(1)] for p in plans]
        states = [[None if s is None else _split(plans[r], s) for s in st] for r, st in enumerate(states)]
    finally:
        for e in engines:
            e.destroy()
    return outs, states


def run_oracle(world, make_comp, make_mem, grads_per_step, names=None):
    names = names or NAMES
    comps = [make_comp() for _ in range(world)]
    mems = [make_mem() for _ in range(world)]
    outs = []
    for grads in grads_per_step:
        step_out = [[None] * len(names) for _ in range(world)]
        for i, name in enumerate(names):
            payloads, ctxs = [], []
            for r in range(world):
                x = mems[r].compensate(grads[r][i], name)
                pl, ctx = comps[r].compress(x, name)
                mems[r].update(x, name, comps[r], pl, ctx)
                payloads.append(pl)
                ctxs.append(ctx)
            for r in range(world):
                dec = [comps[r].decompress(pl, ctxs[r]) for pl in payloads]
                agg = comps[r].aggregate(dec)
                step_out[r][i] = agg / world if comps[r].average else agg
        outs.append(step_out)
    return outs, comps, mems


def make_grads(world, steps, scale=1.0, seed=0, sizes=None):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return [[[torch.randn(n, device="cuda", generator=g) * scale * (1 + 0.3 * r) for n in (sizes or SIZES)]
             for r in range(world)] for _ in range(steps)]


def assert_same(fused, oracle, exact=True, frac=0.0, rtol=1e-5, atol=1e-6, what=""):
    for s, (fs, os_) in enumerate(zip(fused, oracle)):
        for r, (fr, orr) in enumerate(zip(fs, os_)):
            for i, (a, b) in enumerate(zip(fr, orr)):
                b = b.reshape(-1).float()
                if exact:
                    bad = (a != b)
                else:
                    bad = (a - b).abs() > atol + rtol * b.abs()
                nbad = int(bad.sum())
                assert nbad <= frac * a.numel(), (
                    "%s step %d rank %d seg %d (n=%d): %d mismatches, max|diff| %g, first idx %s" %
                    (what, s, r, i, a.numel(), nbad, float((a - b).abs().max()), bad.nonzero()[:5].flatten().tolist()))


@pytest.mark.parametrize("world", [1, 2, 4])
@pytest.mark.parametrize("memory", ["none", "residual"])
def test_topk(world, memory):
    grads = make_grads(world, 3)
    prm = _params("topk", memory, ratio=0.05, beta=0.9, gamma=1.1)
    fused, states = run_fused(world, prm, grads)
    mk = (lambda: M.ResidualMemory(0.9, 1.1)) if memory == "residual" else M.NoneMemory
    oracle, _, mems = run_oracle(world, lambda: C.TopKCompressor(0.05), mk, grads)
    assert_same(fused, oracle, exact=True, what="topk")
    if memory == "residual":
        for r in range(world):
            for i, name in enumerate(NAMES):
                assert torch.equal(states[r][0][i], mems[r].residuals[name].reshape(-1)), (r, name)


def test_topk_ties_take_lowest_indices():
    grads = [[[torch.randint(-3, 4, (n,), device="cuda").float() for n in SIZES]]]
    fused, _ = run_fused(1, _params("topk", ratio=0.1), grads)
    for out, g, n in zip(fused[0][0], grads[0][0], SIZES):
        k = max(1, int(n * 0.1))
        assert int((out != 0).sum()) <= k
        kth = g.abs().sort(descending=True)[0][k - 1]
        assert torch.equal(out[g.abs() > kth], g[g.abs() > kth])          # everything above the threshold
        ties = (g.abs() == kth).nonzero().flatten()
        taken = k - int((g.abs() > kth).sum())
        if kth > 0:
            assert torch.equal(out[ties[:taken]], g[ties[:taken]]) and (out[ties[taken:]] == 0).all()


@pytest.mark.parametrize("world", [1, 3])
def test_threshold(world):
    grads = make_grads(world, 2)
    fused, _ = run_fused(world, _params("threshold", "residual", threshold=1.5), grads)
    oracle, _, _ = run_oracle(world, lambda: C.ThresholdCompressor(1.5), M.ResidualMemory, grads)
    assert_same(fused, oracle, exact=True, what="threshold")


@pytest.mark.parametrize("world", [1, 2])
@pytest.mark.parametrize("memory", ["none", "residual"])
def test_randomk(world, memory):
    grads = make_grads(world, 3)
    fused, _ = run_fused(world, _params("randomk", memory, ratio=0.05), grads)
    mk = M.ResidualMemory if memory == "residual" else M.NoneMemory
    oracle, _, _ = run_oracle(world, lambda: C.RandomKCompressor(0.05, seed=SEED), mk, grads)
    assert_same(fused, oracle, exact=True, what="randomk")


@pytest.mark.parametrize("world", [1, 2])
def test_dgc_with_dgc_memory(world):
    grads = make_grads(world, 3)
    fused, states = run_fused(world, _params("dgc", "dgc", ratio=0.05, momentum=0.9), grads)
    oracle, _, mems = run_oracle(world, lambda: C.DgcCompressor(0.05, seed=SEED),
                                 lambda: M.DgcMemory(0.9, False, world), grads)
    assert_same(fused, oracle, exact=True, what="dgc")
    for r in range(world):
        for i, name in enumerate(NAMES):
            assert torch.equal(states[r][0][i], mems[r].gradients[name].reshape(-1))
            assert torch.equal(states[r][1][i], mems[r].residuals[name].reshape(-1))


@pytest.mark.parametrize("world", [1, 2, 3])
def test_dgc_with_gradient_clipping(world):
    """DgcMemory(gradient_clipping=True) -- the dist helper's default (helper.py:60-62) -- stays native: the per-tensor
    sum(g^2) is all-reduced through the engine's flags (one float per tensor per rank), the clamp is fused into the
    momentum pass.  The kernel sums g^2 in double, torch in float: clip values agree to an ulp, so values are compared
    with a tolerance and a few threshold flips are allowed."""
    grads = make_grads(world, 3)
    prm = _params("dgc", "dgc", ratio=0.05, momentum=0.9)
    prm.clip = 1
    fused, states = run_fused(world, prm, grads)
    # the oracle's all-reduce: sum the ranks' scalars in rank order
    pending = {}

    def make_mem():
        mem = M.DgcMemory(0.9, True, world)
        return mem
    comps = [C.DgcCompressor(0.05, seed=SEED) for _ in range(world)]
    mems = [make_mem() for _ in range(world)]
    oracle = []
    for step in grads:
        out = [[None] * len(SIZES) for _ in range(world)]
        for i, name in enumerate(NAMES):
            total = sum(torch.sum(step[r][i] * step[r][i]) for r in range(world))
            payloads, ctxs = [], []
            for r in range(world):
                mems[r]._allreduce_sum = lambda t, total=total: total
                x = mems[r].compensate(step[r][i], name)
                pl, ctx = comps[r].compress(x, name)
                mems[r].update(x, name, comps[r], pl, ctx)
                payloads.append(pl); ctxs.append(ctx)
            for r in range(world):
                dec = [comps[r].decompress(pl, ctxs[r]) for pl in payloads]
                out[r][i] = comps[r].aggregate(dec) / world
        oracle.append(out)
    assert_same(fused, oracle, exact=False, frac=0.01, rtol=1e-5, atol=1e-6, what="dgc+clip")
    for s in fused:
        for r in range(1, world):
            for a, b in zip(s[0], s[r]):
                assert torch.equal(a, b)            # all ranks hold the same aggregate
    if world > 1:                                   # the clamp is active: the 1-element tensor of the largest rank is cut
        clip = torch.sqrt(sum(grads[0][r][4] ** 2 for r in range(world)) / world)
        assert any(float(grads[0][r][4].abs()) > float(clip) for r in range(world))


@pytest.mark.parametrize("world", [1, 2, 4])
@pytest.mark.parametrize("memory", ["none", "residual"])
def test_adaq(world, memory):
    """K1-adaq: same picked index sets as the torch oracle (sampling, k-th sample, <= 20 refinements per sign
    are all integer/fp32-exact); the two means per tensor are fixed-order double sums, so values agree to an ulp."""
    grads = make_grads(world, 3)
    fused, states = run_fused(world, _params("adaq", memory, ratio=0.05, beta=0.9, gamma=1.1), grads)
    mk = (lambda: M.ResidualMemory(0.9, 1.1)) if memory == "residual" else M.NoneMemory
    oracle, _, mems = run_oracle(world, lambda: C.AdaqCompressor(0.05, seed=SEED), mk, grads)
    for s in range(len(grads)):                       # support (which positions were kept) must match exactly
        for r in range(world):
            for i in range(len(SIZES)):
                a, b = fused[s][r][i], oracle[s][r][i].reshape(-1)
                if memory == "none" or s == 0:
                    assert torch.equal(a != 0, b != 0), (s, r, i)
    assert_same(fused, oracle, exact=False, frac=0.001 if memory == "residual" else 0.0, rtol=1e-5, what="adaq")
    if memory == "residual":
        for r in range(world):
            for i, name in enumerate(NAMES):
                a, b = states[r][0][i], mems[r].residuals[name].reshape(-1)
                assert int(((a - b).abs() > 1e-5 + 1e-5 * b.abs()).sum()) <= 0.001 * a.numel() + 1, (r, name)


@pytest.mark.parametrize("world", [1, 2, 4])
@pytest.mark.parametrize("memory", ["none", "residual"])
def test_inceptionn(world, memory):
    """K2-inceptionn: every step of the codec is exact fp32 / integer arithmetic -> bit-exact against the oracle.
    Gradients span all four encodings (0, 8, 16, 32 bits: |x| from 1e-5 to ~6)."""
    g = torch.Generator(device="cuda").manual_seed(7)
    grads = [[[torch.randn(n, device="cuda", generator=g) * torch.pow(10.0, torch.rand(n, device="cuda", generator=g) * 5 - 4.5)
               for n in SIZES] for r in range(world)] for _ in range(3)]
    comp = C.INCEPTIONNCompressor()
    prm = _params("inceptionn", memory, beta=0.9, gamma=1.1, quantum_num=comp.low_exp | (comp.mid_exp << 8))
    fused, states = run_fused(world, prm, grads)
    mk = (lambda: M.ResidualMemory(0.9, 1.1)) if memory == "residual" else M.NoneMemory
    oracle, _, mems = run_oracle(world, C.INCEPTIONNCompressor, mk, grads)
    assert_same(fused, oracle, exact=True, what="inceptionn")
    if memory == "residual":
        for r in range(world):
            for i, name in enumerate(NAMES):
                assert torch.equal(states[r][0][i], mems[r].residuals[name].reshape(-1)), (r, name)


@pytest.mark.parametrize("codec,make,exact,frac", [
    ("none", C.NoneCompressor, True, 0.0),
    ("fp16", C.FP16Compressor, True, 0.0),
    ("signsgd", C.SignSGDCompressor, True, 0.0),
    ("signum", lambda: C.SignumCompressor(0.9), True, 0.0),
    ("natural", lambda: C.NaturalCompressor(seed=SEED), True, 0.0),
    ("efsignsgd", lambda: C.EFSignSGDCompressor(0.1), False, 0.0),
    ("onebit", C.OneBitCompressor, False, 0.0),
    ("qsgd", lambda: C.QSGDCompressor(64, seed=SEED), False, 0.002),
    ("terngrad", lambda: C.TernGradCompressor(seed=SEED), False, 0.002),
    ("u8bit", C.U8bitCompressor, False, 0.002),
])
@pytest.mark.parametrize("world", [1, 2, 4])
@pytest.mark.parametrize("memory", ["none", "residual"])
def test_quantizers(codec, make, exact, frac, world, memory):
    grads = make_grads(world, 3, scale=0.02)
    avg = 0 if codec in ("signsgd", "signum", "efsignsgd") else 1
    fused, _ = run_fused(world, _params(codec, memory, average=avg, gamma=0.1 if codec == "efsignsgd" else 1.0), grads)
    if memory == "residual":
        mk = (lambda: M.EFSignSGDMemory(0.1)) if codec == "efsignsgd" else M.ResidualMemory
    else:
        mk = M.NoneMemory
    oracle, _, _ = run_oracle(world, make, mk, grads)
    assert_same(fused, oracle, exact=exact, frac=frac, rtol=2e-5, atol=1e-7, what=codec)


@pytest.mark.parametrize("world", [1, 2, 4])
@pytest.mark.parametrize("memory", ["none", "residual"])
def test_sketch(world, memory):
    """SketchML: same cut points and buckets as the torch restatement.  With a residual the bucket means
    (summed in a different order than torch's atomic index_add_) feed back into the next step's input, so a
    few values cross a cut point: later steps are compared in the L2 sense."""
    grads = make_grads(world, 3, scale=0.02)
    fused, _ = run_fused(world, _params("sketch", memory, quantum_num=64), grads)
    mk = M.ResidualMemory if memory == "residual" else M.NoneMemory
    oracle, _, _ = run_oracle(world, lambda: C.SketchCompressor(64), mk, grads)
    steps = 3 if memory == "none" else 1
    assert_same(fused[:steps], oracle[:steps], exact=False, frac=0.002, rtol=2e-5, atol=1e-7, what="sketch")
    for fs, os_ in zip(fused, oracle):
        for fr, orr in zip(fs, os_):
            for a, b in zip(fr, orr):
                b = b.reshape(-1)
                assert float((a - b).norm()) <= 0.05 * float(b.norm()) + 1e-6


@pytest.mark.parametrize("q", [256, 300, 1000])
def test_sketch_with_16_bit_bins(q):
    """>= 256 quantiles travel as uint16 bins (reference: tensorflow/compressor/sketch.py:26); the hand-written radix
    sort orders all tensors of the bucket in one go (no library sort on the path)."""
    grads = make_grads(2, 2, scale=0.02)
    fused, _ = run_fused(2, _params("sketch", "none", quantum_num=q), grads)
    oracle, _, _ = run_oracle(2, lambda: C.SketchCompressor(q), M.NoneMemory, grads)
    assert_same(fused, oracle, exact=False, frac=0.002, rtol=2e-5, atol=1e-7, what="sketch%d" % q)


@pytest.mark.parametrize("q,bits", [(4, 4), (7, 4), (8, 5), (15, 5), (16, 6), (31, 6), (32, 7), (63, 7), (64, 8), (256, 16)])
def test_qsgd_bit_widths(q, bits):
    grads = make_grads(2, 1, scale=0.5)
    fused, _ = run_fused(2, _params("qsgd", quantum_num=q), grads)
    from grace_b200.parallel.engine import Engine
    (e,) = Engine.local(1)
    try:
        pl = e.plan(SIZES, NAMES, _params("qsgd", quantum_num=q))
        assert pl.padded * bits // 8 <= pl.wire_bytes <= pl.padded * bits // 8 + 4096, (pl.wire_bytes, pl.padded, bits)
    finally:
        e.destroy()
    oracle, _, _ = run_oracle(2, lambda: C.QSGDCompressor(q, seed=SEED), M.NoneMemory, grads)
    assert_same(fused, oracle, exact=False, frac=0.002, rtol=2e-5, atol=1e-7, what="qsgd%d" % q)


def test_standalone_topk_matches_torch():
    import ctypes as ct
    from grace_b200.ops import native
    L = native.lib()
    gen = torch.Generator(device="cuda").manual_seed(4242)
    for n, k in [(1, 1), (777, 10), (4096, 40), (100000, 1000), (1 << 20, 10485)]:
        x = torch.randn(n, device="cuda", generator=gen)
        vals = torch.empty(k, device="cuda")
        idx = torch.empty(k, dtype=torch.int32, device="cuda")
        scratch = torch.empty(L.gb_topk_scratch_bytes(n), dtype=torch.uint8, device="cuda")
        native.check(L.gb_topk_compress(ct.c_void_p(x.data_ptr()), n, k, ct.c_void_p(vals.data_ptr()),
                                        ct.c_void_p(idx.data_ptr()), ct.c_void_p(scratch.data_ptr()), scratch.numel(),
                                        ct.c_void_p(torch.cuda.current_stream().cuda_stream)))
        torch.cuda.synchronize()
        # equal magnitudes at the k-th place are taken lowest index first (a stable sort states the same rule;
        # torch.topk leaves it open, and with 2^20 fp32 normals a tie there happens in ~1 % of the draws)
        ref = torch.argsort(x.abs(), descending=True, stable=True)[:k].sort()[0]
        assert torch.equal(idx.long(), ref) and torch.equal(vals, x[ref])


def test_state_survives_plan_roundtrip():
    grads = make_grads(1, 2)
    from grace_b200.parallel.engine import Engine
    (e,) = Engine.local(1)
    try:
        p = e.plan(SIZES, NAMES, _params("topk", "residual"))
        buf = _flat(p, grads[0][0])
        p.send(buf)
        out = torch.empty(p.padded, device="cuda")
        p.recv(out)
        torch.cuda.synchronize()
        saved, step = p.get_state(0), p.step
        q = e.plan(SIZES, NAMES, _params("topk", "residual"))
        q.set_state(0, saved)
        q.step = step
        buf2 = _flat(p, grads[1][0])
        o1, o2 = torch.empty(p.padded, device="cuda"), torch.empty(p.padded, device="cuda")
        p.send(buf2); p.recv(o1)
        q.send(buf2); q.recv(o2)
        torch.cuda.synchronize()
        assert step == 1
        for a, b in zip(_split(p, o1), _split(p, o2)):          # pad elements are never written
            assert torch.equal(a, b)
    finally:
        e.destroy()


def test_missing_peer_sets_timeout_status_instead_of_hanging():
    """Failure detection: a rank that never sends makes the receiver's bounded flag wait expire; the
    kernel returns, the host-mapped status word carries the reason and check_status() raises.  The next exchange
    works again once the late rank has caught up."""
    from grace_b200.ops.native import NativeError
    from grace_b200.parallel.engine import Engine
    engines = Engine.local(2)
    try:
        for e in engines:
            e.set_timeout_ms(30)
        plans = [e.plan(SIZES, NAMES, _params("signsgd")) for e in engines]
        bufs = [torch.randn(plans[0].padded, device="cuda") for _ in range(2)]
        outs = [torch.empty(plans[0].padded, device="cuda") for _ in range(2)]
        plans[0].send(bufs[0])                      # rank 1 stays silent
        plans[0].recv(outs[0])
        torch.cuda.synchronize()
        with pytest.raises(NativeError, match="timed out"):
            engines[0].check_status()
        engines[0].check_status()                   # reading the status clears it
        plans[1].send(bufs[1]); plans[1].recv(outs[1])      # the late rank catches up with epoch 1
        for r in range(2):
            plans[r].send(bufs[r])
        for r in range(2):
            plans[r].recv(outs[r])
        torch.cuda.synchronize()
        for e in engines:
            e.check_status()
        vote = torch.where(torch.sign(bufs[0]) + torch.sign(bufs[1]) + (bufs[0] == 0).float() + (bufs[1] == 0).float() >= 0, 1.0, -1.0)
        for off, n in zip(plans[0].offsets, plans[0].numels):
            assert torch.equal(outs[0][off:off + n], vote[off:off + n])
            assert torch.equal(outs[0][off:off + n], outs[1][off:off + n])     # pad elements are never written
    finally:
        for e in engines:
            e.destroy()


def test_variable_size_payload_overflow_is_reported():
    from grace_b200.ops.native import NativeError
    from grace_b200.parallel.engine import Engine
    (e,) = Engine.local(1)
    try:
        prm = _params("threshold", threshold=0.001)
        prm.capacity = 0.01                          # slots hold 1 % of the bucket, ~100 % of it is selected
        p = e.plan(SIZES, NAMES, prm)
        buf = torch.randn(p.padded, device="cuda")
        out = torch.empty(p.padded, device="cuda")
        p.send(buf); p.recv(out)
        torch.cuda.synchronize()
        with pytest.raises(NativeError, match="capacity"):
            e.check_status()
    finally:
        e.destroy()


def test_failed_plan_creation_is_clean():
    """A plan that cannot be built is refused with a message and released; the engine keeps working."""
    from grace_b200.ops.native import NativeError
    from grace_b200.parallel.engine import Engine
    (e,) = Engine.local(1, arena_bytes=8 << 20)
    try:
        with pytest.raises(NativeError, match="elements"):
            e.plan([100, 1 << 31], ["ok.weight", "huge.weight"], _params("qsgd"))      # refused inside the builder
        p = e.plan(SIZES, NAMES, _params("qsgd"))
        buf = torch.randn(p.padded, device="cuda")
        out = torch.empty(p.padded, device="cuda")
        p.send(buf); p.recv(out)
        torch.cuda.synchronize()
        e.check_status()
        assert float((out[:1000] - buf[:1000]).abs().max()) < 0.6        # one QSGD quantum = norm / 64 ~ 0.5
    finally:
        e.destroy()


@pytest.mark.parametrize("world", [1, 2])
def test_arena_grows_when_a_plan_does_not_fit(world):
    """The peer arena is no longer sized by the first caller: a plan whose slots do not fit makes every (virtual)
    rank add a region, and exchanges through the new region give the same result as the oracle."""
    from grace_b200.ops import native
    sizes, names = [3 << 20, 1000], ["huge.weight", "small.bias"]
    grads = make_grads(world, 2, scale=0.02, sizes=sizes)
    prm = _params("signsgd", "none", average=0)
    from grace_b200.parallel.engine import Engine
    engines = Engine.local(world, arena_bytes=2 << 20)          # flags + < 2 MB of slots
    try:
        small = [e.plan(SIZES, NAMES, prm) for e in engines]     # fits region 0
        big = [e.plan(sizes, names, _params("fp16", "none")) for e in engines]      # several MB of slots: needs a new region
        assert all(native.lib().gb_comm_regions(e._h) >= 2 for e in engines)
        for step in grads:
            bufs = [_flat(big[r], step[r]) for r in range(world)]
            for r in range(world):
                big[r].send(bufs[r])
            for r in range(world):
                big[r].mid()
            outs = []
            for r in range(world):
                out = torch.full((big[r].padded,), float("nan"), device="cuda")
                big[r].recv(out)
                outs.append(out)
            torch.cuda.synchronize()
            for e in engines:
                e.check_status()
            for r in range(1, world):
                assert torch.equal(outs[0][:sizes[0]], outs[r][:sizes[0]])
            want = sum(step[r][0] for r in range(world)) / world
            assert float((outs[0][:sizes[0]] - want).abs().max()) < 1e-4          # fp16 rounding of |x| < 0.2
        # the plan created before the growth still works
        g2 = make_grads(world, 1)
        bufs = [_flat(small[r], g2[0][r]) for r in range(world)]
        for r in range(world):
            small[r].send(bufs[r])
        for r in range(world):
            out = torch.empty(small[r].padded, device="cuda")
            small[r].recv(out)
        torch.cuda.synchronize()
        for e in engines:
            e.check_status()
    finally:
        for e in engines:
            e.destroy()


def _resnet50_shapes():
    import torchvision
    m = torchvision.models.resnet50()
    return [(n, tuple(p.shape)) for n, p in m.named_parameters()]


@pytest.mark.parametrize("codec", ["topk", "qsgd", "signsgd"])
def test_resnet50_geometry(codec):
    """One plan over the 161 gradient tensors of ResNet-50 (25.6 M elements, 64 ... 2.4 M per tensor, 6 300 tiles) on two
    virtual ranks against the per-tensor torch oracle: the shapes the headline benchmark runs, not toy segments."""
    shapes = _resnet50_shapes()
    names = [n for n, _ in shapes]
    sizes = [int(torch.Size(s).numel()) for _, s in shapes]
    assert len(sizes) == 161
    world = 2
    grads = make_grads(world, 2, scale=0.02, sizes=sizes, seed=11)
    if codec == "topk":
        prm, mk_c, mk_m, exact, frac = _params("topk", "residual", ratio=0.01), (lambda: C.TopKCompressor(0.01)), M.ResidualMemory, True, 0.0
    elif codec == "qsgd":
        prm, mk_c, mk_m, exact, frac = _params("qsgd", "none"), (lambda: C.QSGDCompressor(64, seed=SEED)), M.NoneMemory, False, 0.002
    else:
        prm, mk_c, mk_m, exact, frac = _params("signsgd", "none", average=0), C.SignSGDCompressor, M.NoneMemory, True, 0.0
    fused, _ = run_fused(world, prm, grads, sizes=sizes, names=names, arena_bytes=1 << 30)
    oracle, _, _ = run_oracle(world, mk_c, mk_m, grads, names=names)
    assert_same(fused, oracle, exact=exact, frac=frac, rtol=2e-5, atol=1e-7, what="resnet50/" + codec)


def test_topk_tensor_larger_than_2_pow_27():
    """Tile descriptors address tensors up to 2^31 elements (BASELINE config 5 goes to 1 B); 2^27 + 5 elements is the
    first size whose tile index needs more than 15 bits."""
    n = (1 << 27) + 5
    from grace_b200.parallel.engine import Engine
    (e,) = Engine.local(1, arena_bytes=64 << 20)
    try:
        ratio = 0.001
        p = e.plan([n], ["giant.weight"], _params("topk", "none", ratio=ratio))
        k = max(1, int(n * ratio))
        gen = torch.Generator(device="cuda").manual_seed(2027)
        for _ in range(8):        # 1.3e8 fp32 normals: a tie exactly at the k-th magnitude happens in ~10 % of the draws
            g = torch.randn(p.padded, device="cuda", generator=gen)
            g[n:] = 0
            x = g[:n]
            kth = torch.topk(x.abs(), k, sorted=True)[0][-1]
            want = torch.where(x.abs() >= kth, x, torch.zeros_like(x))
            if int((want != 0).sum()) == k:
                break
        assert int((want != 0).sum()) == k
        out = torch.empty(p.padded, device="cuda")
        p.send(g); p.recv(out)
        torch.cuda.synchronize()
        e.check_status()
        assert torch.equal(out[:n], want)
    finally:
        e.destroy()


@pytest.mark.parametrize("codec,memory", [
    ("topk", "residual"), ("randomk", "residual"), ("threshold", "residual"), ("dgc", "dgc"), ("qsgd", "none"),
    ("terngrad", "residual"), ("natural", "residual"), ("u8bit", "none"), ("signsgd", "none"), ("signum", "residual"),
    ("efsignsgd", "residual"), ("onebit", "residual"), ("sketch", "residual"), ("adaq", "residual"),
    ("inceptionn", "residual"), ("none", "none"), ("fp16", "none")])
def test_degenerate_gradients_stay_finite(codec, memory):
    """All-zero and constant gradients (dead layers, frozen inputs): zero norms, empty selections, k-way ties and
    single-bucket quantiles must neither produce NaN/Inf nor trip the overflow / timeout status."""
    comp = C.INCEPTIONNCompressor()
    prm = _params(codec, memory, ratio=0.05, quantum_num=(comp.low_exp | (comp.mid_exp << 8)) if codec == "inceptionn" else 64)
    zeros = [[[torch.zeros(n, device="cuda") for n in SIZES]] for _ in range(2)]
    outs, states = run_fused(1, prm, zeros)
    for step in outs:
        for t in step[0]:
            assert torch.isfinite(t).all(), codec
            if codec not in ("signsgd", "signum", "onebit"):
                assert float(t.abs().max()) == 0.0, (codec, float(t.abs().max()))
    const = [[[torch.full((n,), -0.75, device="cuda") for n in SIZES]] for _ in range(2)]
    outs, states = run_fused(1, prm, const)
    for step in outs:
        for t in step[0]:
            assert torch.isfinite(t).all(), codec
    for st in states[0]:
        if st is not None:
            assert all(torch.isfinite(s).all() for s in st), codec


@pytest.mark.parametrize("world", [1, 3])
def test_point_collectives_over_the_arena(world):
    """csrc/ops.cu, the native half of hvd.allreduce / allgather / broadcast (reference: Horovod core behind
    patch_files/horovod/torch/mpi_ops.py:157,216,362): gather, broadcast and rank-ordered sums for the four element
    types, several ops back to back on the same double-buffered slots."""
    import ctypes as ct
    from grace_b200.ops import native
    from grace_b200.parallel.engine import Engine
    L = native.lib()
    engines = Engine.local(world, arena_bytes=32 << 20)
    st = ct.c_void_p(torch.cuda.current_stream().cuda_stream)
    try:
        ops = []
        for e in engines:
            h = ct.c_void_p()
            native.check(L.gb_ops_create(e._h, 1 << 20, ct.byref(h)))
            ops.append(h)
        g = torch.Generator(device="cuda").manual_seed(3)

        def run(srcs, nbytes, first, count, stride, mode, scale, outs):
            for r in range(world):
                native.check(L.gb_ops_send(ops[r], ct.c_void_p(srcs[r].data_ptr() if srcs[r] is not None else 0), nbytes, st))
            for r in range(world):
                native.check(L.gb_ops_recv(ops[r], nbytes, first, count, stride, mode, ct.c_double(scale),
                                           ct.c_void_p(outs[r].data_ptr()), st))
            torch.cuda.synchronize()
            for e in engines:
                e.check_status()

        for it in range(3):
            n = 4096 + 4 * it
            xs = [torch.randn(n, device="cuda", generator=g) for _ in range(world)]
            outs = [torch.empty(n * world, device="cuda") for _ in range(world)]
            run(xs, n * 4, 0, world, n * 4, 0, 1.0, outs)                       # all-gather
            for o in outs:
                assert torch.equal(o, torch.cat(xs))
            root = it % world
            outs = [torch.empty(n, device="cuda") for _ in range(world)]
            run([xs[r] if r == root else None for r in range(world)], n * 4, root, 1, n * 4, 0, 1.0, outs)    # broadcast
            for o in outs:
                assert torch.equal(o, xs[root])
            outs = [torch.empty(n, device="cuda") for _ in range(world)]
            run(xs, n * 4, 0, world, n * 4, 1, 1.0 / world, outs)                # all-reduce (mean), rank order
            want = xs[0].clone()
            for x in xs[1:]:
                want = want + x
            want = want * torch.tensor(1.0 / world, device="cuda")
            for o in outs:
                assert torch.equal(o, want)
        for dtype, mode in ((torch.float64, 2), (torch.int64, 3), (torch.int32, 4)):
            xs = [(torch.randn(64, device="cuda", generator=g) * 100).to(dtype) for _ in range(world)]
            outs = [torch.empty(64, dtype=dtype, device="cuda") for _ in range(world)]
            run(xs, 64 * xs[0].element_size(), 0, world, 64 * xs[0].element_size(), mode, 1.0, outs)
            want = xs[0].clone()
            for x in xs[1:]:
                want = want + x
            for o in outs:
                assert torch.equal(o, want)
        for h in ops:
            L.gb_ops_destroy(h)
    finally:
        for e in engines:
            e.destroy()


def test_guard_bands_catch_a_stray_peer_store(monkeypatch):
    """compute-sanitizer is closed on the GPU pool; the substitute: with GRACE_B200_GUARD=1 every wire slot ends in a
    poisoned 256-byte band that no kernel may touch.  A full exchange of several codec families leaves all bands intact;
    one overwritten word is reported through the status word."""
    from grace_b200.ops import native
    from grace_b200.ops.native import NativeError
    from grace_b200.parallel.engine import Engine, PowerSGDPlan
    monkeypatch.setenv("GRACE_B200_GUARD", "1")
    engines = Engine.local(2, arena_bytes=64 << 20)
    try:
        grads = make_grads(2, 1)[0]
        for codec, memory in (("topk", "residual"), ("qsgd", "none"), ("signsgd", "none"), ("threshold", "none"), ("fp16", "none")):
            plans = [e.plan(SIZES, NAMES, _params(codec, memory)) for e in engines]
            bufs = [_flat(plans[r], grads[r]) for r in range(2)]
            for r in range(2):
                plans[r].send(bufs[r])
            for r in range(2):
                plans[r].mid()
            for r in range(2):
                plans[r].recv(torch.empty(plans[r].padded, device="cuda"))
        pp = [PowerSGDPlan(e, [(64, 96), (40, 36)], ["a.weight", "b.weight"], 4, 7, True, True) for e in engines]
        bufs = [torch.randn(pp[0].padded, device="cuda") for _ in range(2)]
        for r in range(2):
            pp[r].send(bufs[r])
        for r in range(2):
            pp[r].mid()
        for r in range(2):
            pp[r].recv(torch.empty(pp[r].padded, device="cuda"))
        torch.cuda.synchronize()
        for e in engines:
            assert e.check_guards() >= 6                      # every plan's channel is guarded, none was touched
        native.check(native.lib().gb_comm_guard_poke(engines[1]._h))
        with pytest.raises(NativeError, match="guard band"):
            engines[1].check_guards()
        engines[0].check_guards()                             # the other rank's arena is intact
    finally:
        monkeypatch.delenv("GRACE_B200_GUARD")
        for e in engines:
            e.destroy()
