"""`sn_linear_forward_noalloc` — the entry point for callers that may not allocate GPU memory
(their HBM belongs to another allocator, or they sit on a path that cannot take cudaMalloc's
device-wide stall).  It must give what `sn_linear_forward` gives while using only the scratch
array built into the library.  Collected last on purpose (pytest orders files by name), and
written so it also passes alone in a fresh process, where its first call is the one that looks up
the scratch address."""
import numpy as np
import pytest
import torch

from tests.helpers import rel_err

pytestmark = pytest.mark.gpu


def _case(B, K, N, seed):
    rng = np.random.RandomState(seed)
    x = (rng.rand(B, K) * 50 * (rng.rand(B, K) < 0.6)).astype(np.float32)
    W = (rng.uniform(-1, 1, (N, K)) / np.sqrt(K)).astype(np.float32)
    bias = rng.uniform(-1, 1, N).astype(np.float32)
    mask = (rng.rand(B, N) < 0.5).astype(np.uint8)
    return [torch.from_numpy(a).cuda() for a in (x, W, bias, mask)]


def test_noalloc_matches_linear_forward_twice_in_a_row():
    from splitlearning_b200 import _lib, engine as E
    lib = _lib.load()
    assert lib.sn_linear_forward_noalloc_max_scratch_bytes() == 4 << 20
    # calls back to back on the shared scratch, every one a split-K shape that really goes through
    # it: a large layer, then a smaller one with a ragged batch and a dropout mask (the array still
    # holds the first call's partial sums, in a different layout), then a 40-row batch (three row
    # tiles) and a narrow layer with a long reduction
    cases = [(16, 5408, 5000, True, False, 1), (7, 5000, 1000, False, True, 2), (40, 5000, 200, True, True, 3),
             (16, 40000, 10, False, False, 4)]
    # the very first call on a device looks up the array's address (documented set-up step); the
    # no-allocation assertions below apply to every call, this one included, but module loading
    # under CUDA's lazy mode is excluded by running each shape's kernels once beforehand
    for (B, K, N, relu, use_mask, seed) in cases:
        assert 0 < lib.sn_linear_forward_scratch_bytes(B, K, N) <= 4 << 20
        x, W, bias, mask = _case(B, K, N, seed)
        km, scale = (mask, 2.0) if use_mask else (None, 1.0)
        want = E.linear_forward(x, W, bias, relu=relu, keep_mask=km, mask_scale=scale)
        out = torch.full((B, N), float("nan"), device="cuda")
        torch.cuda.synchronize()
        # the call must not allocate: neither the driver's free-memory figure nor the number of
        # allocations torch's caching allocator has served may move across it (the output buffer
        # was made above; earlier results are dropped first so that only the call is measured)
        got = None
        count = lambda: (torch.cuda.memory_stats()["allocation.all.allocated"], torch.cuda.memory_stats()["segment.all.allocated"])
        free0, n0 = torch.cuda.mem_get_info()[0], count()
        got = E.linear_forward_noalloc(x, W, bias, relu=relu, keep_mask=km, mask_scale=scale, out=out)
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] == free0
        assert count() == n0
        assert got.data_ptr() == out.data_ptr()
        assert rel_err(got.cpu(), want.cpu()) <= 1e-6          # same kernels, fixed-order sums
        ref = torch.nn.functional.linear(x.cpu(), W.cpu(), bias.cpu())
        ref = torch.relu(ref) if relu else ref
        ref = ref * mask.cpu() * 2.0 if use_mask else ref
        assert rel_err(got.cpu(), ref) <= 1e-5                 # tolerance of tests/test_gpu_kernels.py


def test_noalloc_refuses_shapes_beyond_its_scratch():
    from splitlearning_b200 import _lib, engine as E
    lib = _lib.load()
    B, K, N = 64, 5408, 5000
    assert lib.sn_linear_forward_scratch_bytes(B, K, N) > lib.sn_linear_forward_noalloc_max_scratch_bytes()
    x = torch.zeros(B, K, device="cuda")
    W = torch.zeros(N, K, device="cuda")
    bias = torch.zeros(N, device="cuda")
    with pytest.raises(_lib.SplitNNError, match="built-in scratch holds"):
        E.linear_forward_noalloc(x, W, bias)


def test_noalloc_tensor_core_route_uses_the_same_array():
    """With the tcgen05 weight stream switched on, a big layer's partial sums take the other layout
    through the same built-in array; the result must still be sn_linear_forward's."""
    from splitlearning_b200 import _lib, engine as E
    lib = _lib.load()
    B, K, N = 16, 5408, 5000                       # 27 M weights: above the tensor route's threshold
    x, W, bias, _ = _case(B, K, N, 9)
    want = E.linear_forward(x, W, bias, relu=True)
    was = lib.sn_get_skinny_tensor_forward()
    try:
        _lib.check(lib.sn_set_skinny_tensor_forward(1))
        assert lib.sn_linear_forward_scratch_bytes(B, K, N) <= lib.sn_linear_forward_noalloc_max_scratch_bytes()
        got = E.linear_forward_noalloc(x, W, bias, relu=True)
        got2 = E.linear_forward_noalloc(x, W, bias, relu=True)
        torch.cuda.synchronize()
    finally:
        _lib.check(lib.sn_set_skinny_tensor_forward(was))
    assert rel_err(got.cpu(), want.cpu()) <= 1e-5              # 3xTF32 against fp32 FMA
    assert torch.equal(got, got2)
