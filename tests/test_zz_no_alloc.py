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
    # two calls back to back on the shared scratch: a large layer, then a smaller one with a
    # ragged batch and a dropout mask (the array still holds the first call's partial sums)
    for (B, K, N, relu, use_mask, seed) in [(16, 5408, 5000, True, False, 1), (7, 1000, 100, False, True, 2)]:
        x, W, bias, mask = _case(B, K, N, seed)
        km, scale = (mask, 2.0) if use_mask else (None, 1.0)
        want = E.linear_forward(x, W, bias, relu=relu, keep_mask=km, mask_scale=scale)
        out = torch.full((B, N), float("nan"), device="cuda")
        got = E.linear_forward_noalloc(x, W, bias, relu=relu, keep_mask=km, mask_scale=scale, out=out)
        torch.cuda.synchronize()
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
