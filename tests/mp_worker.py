"""Worker for the multi-GPU tests: one process per GPU under torchrun (NCCL only for bootstrap
and for the generic-path oracle).  Compares the native peer-memory path with the generic
torch.distributed path and checks that all ranks hold bit-identical results."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch
import torch.distributed as dist
import torch.nn.functional as F


def same_on_all_ranks(t):
    ref = t.clone()
    dist.broadcast(ref, 0)
    return bool(torch.equal(ref, t))


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
    from grace_b200.dist import grace_from_params
    failures = []
    triples = [
        ("topk", "residual", "allgather", True), ("threshold", "residual", "allgather", True),
        ("randomk", "residual", "allreduce", True), ("dgc", "dgc", "allgather", True),
        ("none", "none", "allreduce", True), ("fp16", "none", "allgather", True),
        ("signsgd", "none", "allgather", True), ("signum", "residual", "allgather", True),
        ("natural", "residual", "allgather", True), ("efsignsgd", "efsignsgd", "allgather", False),
        ("onebit", "residual", "allgather", False), ("qsgd", "none", "allgather", False),
        ("terngrad", "residual", "allgather", False), ("u8bit", "none", "allgather", False),
        ("sketch", "none", "allgather", False), ("adaq", "residual", "allgather", False),
        ("inceptionn", "residual", "allgather", True),
    ]
    shapes = {"conv.weight": (64, 33, 3, 3), "fc.weight": (100, 257), "bn.bias": (77,), "big": (300001,)}
    for comp, mem, comm, exact in triples:
        params = dict(compressor=comp, memory=mem, communicator=comm, world_size=world, compress_ratio=0.02,
                      threshold=1.2, quantum_num=64, gradient_clipping=False, momentum=0.9, lr=0.1)
        fast, slow = grace_from_params(params), grace_from_params(params)
        slow.fused = False
        gen = torch.Generator(device="cuda").manual_seed(1000 + rank)
        for step in range(3):
            for name, shape in shapes.items():
                g = torch.randn(shape, device="cuda", generator=gen) * (0.02 if not exact else 1.0)
                a = fast.step(g.clone(), name).clone()
                b = slow.step(g.clone(), name)
                if exact and (comm != "allreduce" or world <= 2):
                    ok = torch.equal(a, b)
                elif exact:
                    # the generic all-reduce path sums in NCCL's ring/tree order, the native path in
                    # rank order: equal up to rounding once more than two ranks contribute
                    ok = torch.allclose(a, b, rtol=1e-5, atol=1e-6)
                else:
                    ok = float(((a - b).abs() > 1e-6 + 2e-5 * b.abs()).float().mean()) < 0.003
                if not ok:
                    failures.append("%s/%s/%s step %d %s: max diff %g" % (comp, mem, comm, step, name,
                                                                       float((a - b).abs().max())))
                if not same_on_all_ranks(a):
                    failures.append("%s/%s/%s step %d %s: ranks disagree" % (comp, mem, comm, step, name))
        assert fast._stepper is not None and fast._stepper.enabled, comp

    # fused, overlapped DistributedOptimizer: replicas stay bit-identical and match the generic path
    import grace_b200.torch as hvd
    from grace_b200.models import MnistNet

    def train(fused):
        torch.manual_seed(0)
        model = MnistNet().cuda().eval()
        opt = torch.optim.SGD(model.parameters(), lr=0.05, momentum=0.5)
        hvd.broadcast_parameters(model.state_dict(), root_rank=0)
        grc = hvd.grace_from_params(dict(compressor="topk", memory="residual", communicator="allgather",
                                         compress_ratio=0.01))
        opt = hvd.DistributedOptimizer(opt, grc, named_parameters=model.named_parameters(), fused=fused,
                                       bucket_bytes=40000)
        gen = torch.Generator(device="cuda").manual_seed(5 + rank)
        for _ in range(5):
            x = torch.randn(16, 1, 28, 28, device="cuda", generator=gen)
            y = torch.randint(0, 10, (16,), device="cuda", generator=gen)
            opt.zero_grad()
            F.nll_loss(model(x), y).backward()
            opt.step()
        return torch.cat([p.detach().flatten() for p in model.parameters()])

    # flag-protocol stress: ranks arrive in random order with random device-side delays, many
    # back-to-back exchanges on the same double-buffered slots (SURVEY §5 race-detection row)
    import random
    rnd = random.Random(1234 + rank)
    stress = [
        ("topk", dict(compressor="topk", memory="residual", communicator="allgather", compress_ratio=0.02), (50000,)),
        ("qsgd", dict(compressor="qsgd", memory="none", communicator="allgather", quantum_num=64), (50000,)),
        ("signsgd", dict(compressor="signsgd", memory="none", communicator="allgather"), (50000,)),
        ("none", dict(compressor="none", memory="none", communicator="allreduce"), (50000,)),
        ("powersgd", dict(compressor="powersgd", memory="powersgd", communicator="allreduce", compress_rank=4), (200, 250)),
    ]
    for label, params, shape in stress:
        fast = grace_from_params(dict(params, world_size=world))
        gen = torch.Generator(device="cuda").manual_seed(77 + rank)
        n = 1
        for d in shape:
            n *= d
        for it in range(40):
            if rnd.random() < 0.5:
                torch.cuda._sleep(int(rnd.random() * 3e6))          # up to ~1.5 ms of skew on this rank
            g = torch.randn(shape, device="cuda", generator=gen)
            want = None
            if label == "none":
                want = g.clone()
                dist.all_reduce(want)
                want /= world
            if label == "powersgd":            # K3's native path is the bucketed entry
                a = torch.cat([o.flatten() for o in fast.step_all([("stress_a", g), ("stress_b", g[:64] * 2.0)])]).clone()
            else:
                a = fast.step(g, "stress_" + label).clone()
            if not same_on_all_ranks(a):
                failures.append("stress %s iteration %d: ranks disagree" % (label, it))
                break
            if not bool(torch.isfinite(a).all()):
                failures.append("stress %s iteration %d: non-finite output" % (label, it))
                break
            if label == "topk" and int((a != 0).sum()) > world * max(1, int(n * 0.02)):
                failures.append("stress iteration %d: %d non-zeros > W*k" % (it, int((a != 0).sum())))
                break
            if want is not None and not torch.allclose(a, want, rtol=1e-5, atol=1e-6):
                failures.append("stress none iteration %d: differs from all_reduce by %g" % (it, float((a - want).abs().max())))
                break
        assert fast._stepper is not None and (fast._stepper.enabled or label == "powersgd"), label
        if label == "powersgd":
            assert fast._stepper.launches() > 0, "PowerSGD stress ran on the generic path"
        fast.close()

    # dense buckets: NVLS multimem when the host has it, else the P2P two-shot plan -- same API
    def train_dense():
        import copy
        torch.manual_seed(0)
        model = MnistNet().cuda().eval()
        hvd.broadcast_parameters(model.state_dict(), root_rank=0)
        twin = copy.deepcopy(model)                    # plain replica: source of the expected gradients
        opt = torch.optim.SGD(model.parameters(), lr=0.0)
        grc = hvd.grace_from_params(dict(compressor="none", memory="none", communicator="allreduce"))
        opt = hvd.DistributedOptimizer(opt, grc, named_parameters=model.named_parameters(), fused=True)
        gen = torch.Generator(device="cuda").manual_seed(5 + rank)
        x = torch.randn(16, 1, 28, 28, device="cuda", generator=gen)
        y = torch.randint(0, 10, (16,), device="cuda", generator=gen)
        F.nll_loss(twin(x), y).backward()
        want = torch.cat([p.grad.detach().flatten() for p in twin.parameters()])
        dist.all_reduce(want)
        want /= world
        outs = []
        for _ in range(3):
            opt.zero_grad()
            F.nll_loss(model(x), y).backward()
            # (between backward and synchronize the bucket belongs to the exchange: with NVLS peers
            #  already write reduced shards into it, so gradients are only inspected afterwards)
            opt.synchronize()
            got = torch.cat([p.grad.detach().flatten() for p in model.parameters()])
            outs.append((got.clone(), want))
            with opt.skip_synchronize():
                opt.step()
        return outs, opt._fused.transport, opt._fused.nvls_error

    dense, transport, nvls_err = train_dense()
    if rank == 0:
        print("dense transport: %s (nvls error: %s)" % (transport, nvls_err), flush=True)
    for got, want in dense:
        if not torch.allclose(got, want, rtol=1e-5, atol=1e-7):
            failures.append("dense all-reduce (%s) differs by %g" % (transport, float((got - want).abs().max())))
        if not same_on_all_ranks(got):
            failures.append("dense all-reduce (%s): ranks disagree" % transport)

    # PowerSGD (K3): tile-wise peer pushes of P and Q from the GEMM epilogues -- fused DistributedOptimizer against the
    # generic cuBLAS + NCCL path (tf32 tolerance), replicas bit-identical; and the bucketed dist API (step_all)
    def train_psgd(fused):
        torch.manual_seed(0)
        model = MnistNet().cuda().eval()
        opt = torch.optim.SGD(model.parameters(), lr=0.05, momentum=0.5)
        hvd.broadcast_parameters(model.state_dict(), root_rank=0)
        grc = hvd.grace_from_params(dict(compressor="powersgd", memory="powersgd", communicator="allreduce", compress_rank=4))
        opt = hvd.DistributedOptimizer(opt, grc, named_parameters=model.named_parameters(), fused=fused, bucket_bytes=40000)
        gen = torch.Generator(device="cuda").manual_seed(5 + rank)
        for _ in range(4):
            x = torch.randn(16, 1, 28, 28, device="cuda", generator=gen)
            y = torch.randint(0, 10, (16,), device="cuda", generator=gen)
            opt.zero_grad()
            F.nll_loss(model(x), y).backward()
            opt.step()
        return torch.cat([p.detach().flatten() for p in model.parameters()])

    qf, qg = train_psgd(True), train_psgd(False)
    if not same_on_all_ranks(qf):
        failures.append("PowerSGD fused DistributedOptimizer: replicas diverged")
    rel = float((qf - qg).norm() / qg.norm())
    if rel > 3e-2:
        failures.append("PowerSGD fused vs generic differ by %g (relative)" % rel)

    # the dist helper's default DGC memory clips (gradient_clipping = 0.3 -> truthy): native scalar all-reduce
    prm = dict(compressor="dgc", memory="dgc", communicator="allgather", world_size=world, compress_ratio=0.05)
    fast, slow = grace_from_params(prm), grace_from_params(prm)
    slow.fused = False
    assert fast.memory.gradient_clipping
    gen = torch.Generator(device="cuda").manual_seed(19 + rank)
    for step in range(3):
        for name, shape in shapes.items():
            g = torch.randn(shape, device="cuda", generator=gen)
            a = fast.step(g.clone(), name)
            b = slow.step(g.clone(), name)
            bad = float(((a - b).abs() > 1e-6 + 1e-5 * b.abs()).float().mean())
            if bad > 0.01:
                failures.append("dgc+clip step %d %s: %.4f of the elements differ" % (step, name, bad))
            if not same_on_all_ranks(a):
                failures.append("dgc+clip step %d %s: ranks disagree" % (step, name))
    if not (fast._stepper is not None and fast._stepper.enabled):
        failures.append("dgc with clipping did not take the native path")
    fast.close()

    for comp, mem, comm, tol in (("topk", "residual", "allgather", 0.0), ("qsgd", "none", "allgather", 5e-3),
                                 ("powersgd", "powersgd", "allreduce", 5e-2), ("signsgd", "none", "allgather", 0.0)):
        prm = dict(compressor=comp, memory=mem, communicator=comm, world_size=world, compress_ratio=0.02, quantum_num=64,
                   compress_rank=4)
        fast, slow = grace_from_params(prm), grace_from_params(prm)
        slow.fused = False
        gen = torch.Generator(device="cuda").manual_seed(31 + rank)
        for step in range(2):
            named = [(n, torch.randn(sh, device="cuda", generator=gen) * 0.05) for n, sh in shapes.items()]
            outs = fast.step_all([(n, g.clone()) for n, g in named])
            for (n, g), a in zip(named, outs):
                b = slow.step(g.clone(), n)
                err = float((a - b).norm()) / max(float(b.norm()), 1e-12)
                if err > tol + 1e-6:
                    failures.append("step_all %s step %d %s: relative error %g" % (comp, step, n, err))
                if not same_on_all_ranks(a.contiguous()):
                    failures.append("step_all %s step %d %s: ranks disagree" % (comp, step, n))
        if not (fast._stepper is not None and fast._stepper.enabled):
            failures.append("step_all %s did not take the native path" % comp)
        fast.close()

    # whole step as one CUDA graph (forward + backward + overlapped exchange + optimizer): replicas stay identical
    def train_graphed():
        torch.manual_seed(0)
        model = MnistNet().cuda().eval()
        opt = torch.optim.SGD(model.parameters(), lr=0.05, momentum=0.5)
        hvd.broadcast_parameters(model.state_dict(), root_rank=0)
        grc = hvd.grace_from_params(dict(compressor="topk", memory="residual", communicator="allgather", compress_ratio=0.01))
        opt = hvd.DistributedOptimizer(opt, grc, named_parameters=model.named_parameters(), fused=True, bucket_bytes=40000)
        gen = torch.Generator(device="cuda").manual_seed(5 + rank)
        xs = [torch.randn(16, 1, 28, 28, device="cuda", generator=gen) for _ in range(5)]
        ys = [torch.randint(0, 10, (16,), device="cuda", generator=gen) for _ in range(5)]
        step = hvd.GraphedStep(model, opt, lambda m, x, y: F.nll_loss(m(x), y), [xs[0], ys[0]], amp_dtype=None, warmup=2)
        for x, y in zip(xs, ys):
            step(x, y)
        torch.cuda.synchronize()
        return torch.cat([p.detach().flatten() for p in model.parameters()])

    gp = train_graphed()
    if not same_on_all_ranks(gp):
        failures.append("GraphedStep: replicas diverged")

    # Horovod-style ops on CUDA tensors run on the engine (csrc/ops.cu), not through NCCL
    from grace_b200.torch import mpi_ops
    n0 = mpi_ops.native_op_launches()
    x = torch.arange(1000, device="cuda", dtype=torch.float32) * (rank + 1)
    got = hvd.allreduce(x, average=True)
    want = torch.arange(1000, device="cuda", dtype=torch.float32) * (sum(range(1, world + 1)) / world)
    if not torch.allclose(got, want, rtol=1e-6):
        failures.append("native hvd.allreduce differs by %g" % float((got - want).abs().max()))
    big = torch.full((3 << 20,), float(rank + 1), device="cuda")                    # > 4 MB: dense two-shot plan
    got = hvd.allreduce(big, average=False)
    if not torch.equal(got, torch.full_like(big, float(sum(range(1, world + 1))))):
        failures.append("native large hvd.allreduce wrong")
    rows = torch.full((rank + 1, 3), float(rank), device="cuda", dtype=torch.bfloat16)
    got = hvd.allgather(rows)
    want = torch.cat([torch.full((r + 1, 3), float(r), device="cuda", dtype=torch.bfloat16) for r in range(world)])
    if got.shape != want.shape or not torch.equal(got, want):
        failures.append("native hvd.allgather wrong: %s" % (tuple(got.shape),))
    z = torch.full((7,), float(rank), device="cuda", dtype=torch.float64)
    hvd.broadcast_(z, root_rank=world - 1)
    if not torch.equal(z, torch.full_like(z, float(world - 1))):
        failures.append("native hvd.broadcast_ wrong")
    h = hvd.allreduce_async_(torch.ones(5, device="cuda", dtype=torch.int64) * rank, average=False)
    while not hvd.poll(h):
        pass
    if int(hvd.synchronize(h)[0]) != sum(range(world)):
        failures.append("native int64 allreduce wrong")
    if mpi_ops.native_op_launches() <= n0:
        failures.append("hvd ops on CUDA tensors did not take the native path")

    pf, pg = train(True), train(False)
    if not same_on_all_ranks(pf):
        failures.append("fused DistributedOptimizer: replicas diverged")
    if not torch.allclose(pf, pg, rtol=1e-5, atol=1e-6):
        failures.append("fused vs generic DistributedOptimizer differ by %g" % float((pf - pg).abs().max()))
    from grace_b200.parallel.fused import get_engine
    torch.cuda.synchronize()
    get_engine().check_status()

    flag = torch.tensor([len(failures)], device="cuda")
    dist.all_reduce(flag)
    for f in failures:
        print("[rank %d] FAIL %s" % (rank, f), flush=True)
    if rank == 0:
        print("MULTIGPU %s world=%d" % ("OK" if int(flag) == 0 else "FAILED", world), flush=True)
    dist.destroy_process_group()
    sys.exit(1 if int(flag) else 0)


if __name__ == "__main__":
    main()
