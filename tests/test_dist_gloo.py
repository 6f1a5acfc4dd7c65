"""Two-process Gloo runs on CPU: the plumbing path and BASELINE.json config 1
(small CNN on synthetic 28x28, TopK 1% + ResidualMemory + Allgather, world_size=2)."""
import io
import os
import socket
import sys
import time

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _spawn(fn, world=2, *args):
    port = _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    procs = [ctx.Process(target=_entry, args=(fn, r, world, port, q) + args) for r in range(world)]
    for p in procs:
        p.start()
    out = {}
    deadline = time.time() + 240
    # drain the queue BEFORE joining: a child blocks in put() until its bytes are read
    while len(out) < world and time.time() < deadline:
        if not q.empty():
            r, v = q.get()
            out[r] = torch.load(io.BytesIO(v), weights_only=False)
        elif any(p.exitcode not in (None, 0) for p in procs):
            break
        else:
            time.sleep(0.05)
    for p in procs:
        p.join(20)
        if p.is_alive():
            p.kill()
    assert len(out) == world, "worker failed or timed out"
    return [out[r] for r in range(world)]


def _entry(fn, rank, world, port, q, *args):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        buf = io.BytesIO()
        torch.save(fn(rank, world, *args), buf)
        q.put((rank, buf.getvalue()))
    finally:
        dist.destroy_process_group()


TRIPLES = [
    ("none", "none", "allreduce"), ("none", "none", "allgather"), ("none", "none", "broadcast"),
    ("fp16", "none", "allreduce"), ("topk", "residual", "allgather"), ("topk", "residual", "broadcast"),
    ("randomk", "residual", "allreduce"), ("randomk", "residual", "allgather"),
    ("threshold", "residual", "allgather"), ("dgc", "dgc", "allgather"), ("adaq", "residual", "allgather"),
    ("qsgd", "none", "allgather"), ("terngrad", "residual", "allgather"), ("natural", "residual", "allgather"),
    ("u8bit", "none", "allgather"), ("sketch", "none", "allgather"), ("inceptionn", "none", "allgather"),
    ("signsgd", "none", "allgather"), ("signum", "residual", "allgather"),
    ("efsignsgd", "efsignsgd", "allgather"), ("onebit", "residual", "broadcast"),
    ("powersgd", "powersgd", "allreduce"), ("powersgd", "none", "allreduce"),
]


def _run_triples(rank, world, flavour):
    if flavour == "dist":
        from grace_b200.dist import grace_from_params
    else:
        from grace_b200.torch import grace_from_params
    results = {}
    for comp, mem, comm in TRIPLES:
        params = dict(compressor=comp, memory=mem, communicator=comm, world_size=world, compress_ratio=0.1,
                      threshold=0.5, quantum_num=64, gradient_clipping=(comp == "dgc"), compress_rank=2)
        grc = grace_from_params(params)
        gen = torch.Generator().manual_seed(100 + rank)
        outs = []
        for _ in range(3):
            w = torch.randn(64, 32, generator=gen)
            b = torch.randn(32, generator=gen)
            outs += [grc.step(w, "weight"), grc.step(b, "bias")]
        assert all(o.shape in (w.shape, b.shape) for o in outs)
        assert all(torch.isfinite(o).all() for o in outs)
        results["%s+%s+%s" % (comp, mem, comm)] = [o.clone() for o in outs]
    return results


@pytest.mark.parametrize("flavour", ["dist", "torch"])
def test_every_working_triple_runs_and_ranks_agree(flavour):
    r0, r1 = _spawn(_run_triples, 2, flavour)
    for key in r0:
        for a, b in zip(r0[key], r1[key]):
            assert torch.equal(a, b), key
    # the dense triples give the exact mean of the two ranks' inputs
    g0, g1 = torch.Generator().manual_seed(100), torch.Generator().manual_seed(101)
    mean_w = (torch.randn(64, 32, generator=g0) + torch.randn(64, 32, generator=g1)) / 2
    for key in ("none+none+allreduce", "none+none+allgather", "none+none+broadcast"):
        assert torch.allclose(r0[key][0], mean_w, atol=1e-6), key
    assert torch.allclose(r0["fp16+none+allreduce"][0], mean_w, atol=2e-3)
    # exactly low-rank input is reproduced by PowerSGD (mean formed)


def _powersgd_mean(rank, world):
    from grace_b200.dist import grace_from_params
    grc = grace_from_params(dict(compressor="powersgd", memory="none", communicator="allreduce",
                                 world_size=world, compress_rank=2))
    u = torch.randn(30, 1, generator=torch.Generator().manual_seed(1))
    v = torch.randn(1, 20, generator=torch.Generator().manual_seed(2))
    return grc.step((rank + 1) * (u @ v), "m"), 1.5 * (u @ v)


def test_powersgd_forms_the_mean():
    (out, exp), _ = _spawn(_powersgd_mean, 2)
    assert torch.allclose(out, exp, rtol=1e-4, atol=1e-5)


def _topk_allreduce_rejected(rank, world):
    from grace_b200.dist import grace_from_params
    grc = grace_from_params(dict(compressor="topk", memory="residual", communicator="allreduce", world_size=world))
    try:
        grc.step(torch.randn(8, 8), "w")
    except TypeError as e:
        return str(e)
    return ""


def test_topk_with_allreduce_is_rejected_clearly():
    msg, _ = _spawn(_topk_allreduce_rejected, 2)
    assert "Allgather" in msg


def _train_small_cnn(rank, world, flavour):
    """BASELINE config 1."""
    import torch.nn.functional as F
    from grace_b200.models import MnistNet
    torch.manual_seed(0)
    model = MnistNet()
    model.eval()                                          # no dropout: loss is a deterministic function
    gen = torch.Generator().manual_seed(7 + rank)
    data = torch.randn(4, 16, 1, 28, 28, generator=gen)
    target = torch.randint(0, 10, (4, 16), generator=gen)
    params = dict(compressor="topk", memory="residual", communicator="allgather", compress_ratio=0.01, world_size=world)
    losses = []
    if flavour == "dist":
        from grace_b200.dist import grace_from_params
        grc = grace_from_params(params)
        opt = torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.5)
        for step in range(40):
            opt.zero_grad()
            torch.manual_seed(step)                       # same dropout masks on both ranks is not required
            loss = F.nll_loss(model(data[step % 2]), target[step % 2])
            loss.backward()
            for name, p in model.named_parameters():       # the reference's manual loop (core.py:202-208)
                p.grad.copy_(grc.step(p.grad, name))
            opt.step()
            losses.append(float(loss.detach()))
    else:
        import grace_b200.torch as hvd
        grc = hvd.grace_from_params(params)
        opt = torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.5)
        hvd.broadcast_parameters(model.state_dict(), root_rank=0)
        hvd.broadcast_optimizer_state(opt, root_rank=0)
        opt = hvd.DistributedOptimizer(opt, grc, named_parameters=model.named_parameters())
        for step in range(40):
            opt.zero_grad()
            torch.manual_seed(step)
            loss = F.nll_loss(model(data[step % 2]), target[step % 2])
            loss.backward()
            opt.step()
            losses.append(float(loss.detach()))
    flat = torch.cat([p.detach().flatten() for p in model.parameters()])
    return flat, losses


@pytest.mark.parametrize("flavour", ["dist", "torch"])
def test_config1_small_cnn_topk_residual_allgather(flavour):
    (p0, l0), (p1, l1) = _spawn(_train_small_cnn, 2, flavour)
    assert torch.equal(p0, p1)                              # replicas stay bit-identical
    assert sum(l0[-4:]) < sum(l0[:4])                       # and the loss goes down


def _autograd_collectives(rank, world):
    import grace_b200.torch as hvd
    hvd.init()
    out = {}
    # allreduce (mean): d/dx sum(w * mean_r x_r) = mean_r w_r
    x = torch.full((3,), float(rank + 1), requires_grad=True)
    w = torch.arange(3.0) + 10 * rank
    (hvd.allreduce(x, average=True) * w).sum().backward()
    out["allreduce"] = x.grad.clone()
    # allgather with different first dimensions: every rank gets back the rows it contributed, summed over ranks
    y = torch.full((rank + 1, 2), float(rank), requires_grad=True)
    g = hvd.allgather(y)
    coef = torch.arange(float(g.numel())).view_as(g) * (rank + 1)
    (g * coef).sum().backward()
    out["allgather_shape"] = tuple(g.shape)
    out["allgather"] = y.grad.clone()
    # broadcast from rank 1: only the root gets a gradient (the sum over ranks)
    z = torch.full((2,), float(rank), requires_grad=True)
    (hvd.broadcast(z, root_rank=1) * (rank + 1)).sum().backward()
    out["broadcast"] = z.grad.clone()
    # plain tensors keep the non-differentiable path
    out["plain"] = hvd.allreduce(torch.ones(2) * rank, average=False)
    return out


def test_blocking_collectives_are_differentiable():
    """Reference: the autograd Functions of patch_files/horovod/torch/mpi_ops.py:116-127,237-255,319-333."""
    r0, r1 = _spawn(_autograd_collectives, 2)
    assert torch.equal(r0["allreduce"], torch.tensor([5.0, 6.0, 7.0])) and torch.equal(r1["allreduce"], r0["allreduce"])
    assert r0["allgather_shape"] == (3, 2)
    coef_sum = torch.arange(6.0).view(3, 2) * 3            # (1 + 2) x position index
    assert torch.equal(r0["allgather"], coef_sum[:1]) and torch.equal(r1["allgather"], coef_sum[1:])
    assert torch.equal(r0["broadcast"], torch.zeros(2)) and torch.equal(r1["broadcast"], torch.full((2,), 3.0))
    assert torch.equal(r0["plain"], torch.ones(2))


def _optimizer_state_from_root(rank, world):
    import grace_b200.torch as hvd
    hvd.init()
    torch.manual_seed(rank)
    model = torch.nn.Linear(4, 3)
    # a resumed rank 0 and a fresh rank 1: different hyper-parameters and different momentum buffers
    opt = torch.optim.SGD(model.parameters(), lr=0.5 if rank == 0 else 0.1, momentum=0.9 if rank == 0 else 0.5,
                          weight_decay=1e-3 if rank == 0 else 0.0)
    model(torch.randn(2, 4)).sum().backward()
    opt.step()
    hvd.broadcast_optimizer_state(opt, root_rank=0)
    g = opt.param_groups[0]
    bufs = torch.cat([opt.state[p]["momentum_buffer"].flatten() for p in model.parameters()])
    return (g["lr"], g["momentum"], g["weight_decay"]), bufs


def test_broadcast_optimizer_state_propagates_hyperparameters():
    """Reference: patch_files/horovod/torch/__init__.py:356-403 writes root's lr/momentum/... into every rank's
    param_groups; replicas that keep their own lr after a resume diverge silently."""
    (h0, b0), (h1, b1) = _spawn(_optimizer_state_from_root, 2)
    assert h0 == (0.5, 0.9, 1e-3)
    assert h1 == h0
    assert torch.equal(b0, b1)
