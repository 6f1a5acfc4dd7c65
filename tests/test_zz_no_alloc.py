"""gb_topk_compress_static: the allocation-free top-k entry point.

It exists for hosts that cannot supply scratch memory -- they have no allocator of their own
(all device memory already belongs to another pool) or they sit on a path where cudaMalloc's
device-wide synchronisation is unacceptable.  The check: through the C ABI, its output must be
identical to the caller-scratch entry point gb_topk_compress, on two calls in a row (the second
call re-uses the built-in scratch that still holds the first call's leftovers), and inputs
beyond the built-in limit must be refused with a clear message.

Collected last on purpose (file name): a stop-at-first-failure run has then already run
everything else.  It must also pass on its own in a fresh process.
"""
import ctypes as ct

import pytest
import torch

pytestmark = pytest.mark.gpu


def _ptr(t):
    return ct.c_void_p(t.data_ptr())


def _run(L, fn_static, x, k):
    from grace_b200.ops import native
    n = x.numel()
    stream = ct.c_void_p(torch.cuda.current_stream().cuda_stream)
    vals = torch.full((k,), float("nan"), device="cuda")
    idx = torch.full((k,), -1, dtype=torch.int32, device="cuda")
    if fn_static:
        native.check(L.gb_topk_compress_static(_ptr(x), n, k, _ptr(vals), _ptr(idx), stream))
    else:
        scratch = torch.empty(L.gb_topk_scratch_bytes(n), dtype=torch.uint8, device="cuda")
        native.check(L.gb_topk_compress(_ptr(x), n, k, _ptr(vals), _ptr(idx), _ptr(scratch), scratch.numel(), stream))
    torch.cuda.synchronize()
    return vals, idx


def test_static_scratch_entry_point_matches_caller_scratch_entry_point():
    from grace_b200.ops import native
    L = native.lib()
    g = torch.Generator(device="cuda").manual_seed(77)
    # two calls in a row with different sizes: the second sees the first one's leftovers
    for n, k in [(300000, 3000), (5001, 17), (1 << 20, 10485)]:
        x = torch.randn(n, device="cuda", generator=g)
        ref_vals, ref_idx = _run(L, False, x, k)
        for _ in range(2):
            vals, idx = _run(L, True, x, k)
            assert torch.equal(idx, ref_idx) and torch.equal(vals, ref_vals)
        # and against plain torch
        want = torch.topk(x.abs(), k)[1].sort()[0]
        assert torch.equal(ref_idx.long(), want)


def test_static_scratch_entry_point_rejects_oversize_input():
    from grace_b200.ops import native
    L = native.lib()
    limit = L.gb_topk_static_max_n()
    assert limit == 1 << 24
    x = torch.zeros(16, device="cuda")
    out_v, out_i = torch.zeros(4, device="cuda"), torch.zeros(4, dtype=torch.int32, device="cuda")
    rc = L.gb_topk_compress_static(_ptr(x), limit + 1, 4, _ptr(out_v), _ptr(out_i),
                                   ct.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc != 0
    msg = L.gb_last_error().decode()
    assert "exceeds the built-in scratch" in msg and str(limit) in msg
    # the largest accepted size works
    big = torch.randn(limit, device="cuda")
    vals, idx = _run(L, True, big, 1000)
    assert torch.equal(idx.long(), torch.topk(big.abs(), 1000)[1].sort()[0])
