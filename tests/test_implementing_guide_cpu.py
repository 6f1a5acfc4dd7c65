"""The code of IMPLEMENTING.md, executed: a user compressor + memory through both flavours and the registry."""
import torch

from grace_b200.dist import Compressor, Memory


class ClipCompressor(Compressor):
    def __init__(self, c):
        super().__init__(average=True, tensors_size_are_same=True)
        self.c = c

    def compress(self, tensor, name):
        return [tensor.clamp(-self.c, self.c).half()], (tensor.dtype, tensor.size())

    def decompress(self, tensors, ctx):
        dtype, shape = ctx
        return tensors[0].to(dtype).view(shape)


class DecayedResidual(Memory):
    _state_fields = ("residuals",)

    def __init__(self, decay):
        self.decay, self.residuals = decay, {}

    def compensate(self, tensor, name):
        return tensor + self.decay * self.residuals[name] if name in self.residuals else tensor

    def update(self, tensor, name, compressor, tensor_compressed, ctx):
        self.residuals[name] = tensor - compressor.decompress(tensor_compressed, ctx)


def test_user_classes_run_on_the_generic_path_in_both_flavours():
    from grace_b200.dist.communicator.allgather import Allgather
    g = torch.tensor([[0.05, -0.5], [0.2, 0.0]])
    grc = Allgather(ClipCompressor(0.1), DecayedResidual(0.9), 1)
    out = grc.step(g.clone(), "w")
    assert out.shape == g.shape and float(out.abs().max()) <= 0.1 + 1e-3
    assert torch.allclose(grc.memory.residuals["w"], g - out, atol=1e-3)
    out2 = grc.step(g.clone(), "w")                                   # residual fed back
    assert torch.allclose(out2, (g + 0.9 * (g - out)).clamp(-0.1, 0.1), atol=1e-3)

    import grace_b200.torch as hvd
    from grace_b200.torch.communicator.allgather import Allgather as HAllgather
    hvd.init()
    model = torch.nn.Linear(4, 2)
    opt = hvd.DistributedOptimizer(torch.optim.SGD(model.parameters(), lr=0.1),
                                   HAllgather(ClipCompressor(0.1), DecayedResidual(0.9), hvd.size()),
                                   named_parameters=model.named_parameters())
    before = [p.detach().clone() for p in model.parameters()]
    opt.zero_grad()
    model(torch.randn(3, 4)).sum().backward()
    opt.step()                 # one process: like Horovod's wrapper, no exchange is set up when size() == 1
    assert any(not torch.equal(a, p.detach()) for a, p in zip(before, model.parameters()))
    h, ctx = opt.grace.send_step(torch.full((4,), 0.5), "x") if hasattr(opt, "grace") else (None, None)
    if h is not None:
        assert float(opt.grace.receive_step(h, ctx).abs().max()) <= 0.1 + 1e-3


def test_registry_makes_user_classes_available_by_name():
    from grace_b200.dist import grace_from_params
    from grace_b200.registry import register_compressor, register_memory
    register_compressor("clip", ClipCompressor, [("c", "clip_value", 0.1, 0.1)])
    register_memory("decayed", DecayedResidual, [("decay", "decay", 0.9, 0.9)])
    grc = grace_from_params({"compressor": "clip", "memory": "decayed", "communicator": "allgather", "world_size": 1,
                             "clip_value": 0.25})
    assert isinstance(grc.compressor, ClipCompressor) and grc.compressor.c == 0.25 and isinstance(grc.memory, DecayedResidual)
    state = grc.state_dict()
    grc.step(torch.ones(3), "w")
    assert "w" in grc.memory.residuals
    grc.load_state_dict(state)
    assert "w" not in grc.memory.residuals
