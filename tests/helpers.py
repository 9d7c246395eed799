"""Shared test helpers: golden-fixture access and the norm-wise parity metric."""
import os

import numpy as np
import torch

from oracle.gen_golden import seeded_batch, seeded_masks, seeded_weights, sample_idx  # noqa: F401

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name), allow_pickle=False)


def rel_err(a, b):
    """||a-b||_F / ||b||_F with b the reference (float64 accumulate)."""
    a = np.asarray(a, np.float64).reshape(-1)
    b = np.asarray(b, np.float64).reshape(-1)
    den = np.linalg.norm(b)
    return float(np.linalg.norm(a - b) / den) if den > 0 else float(np.linalg.norm(a - b))


def check_summary(g, name, tensor, tol, what=""):
    """Compare `tensor` with a fixture entry written by gen_golden.summarize()."""
    t = tensor.detach().cpu().numpy() if torch.is_tensor(tensor) else np.asarray(tensor)
    flat = t.astype(np.float32).reshape(-1)
    if name in g.files:
        e = rel_err(flat, g[name].reshape(-1))
    else:
        idx = g[name + "__idx"]
        e = rel_err(flat[idx], g[name + "__val"])
    assert e <= tol, f"{what}{name}: rel err {e:.3e} > {tol:.1e}"
    l2 = float(np.sqrt((flat.astype(np.float64) ** 2).sum()))
    ref = float(g[name + "__l2"])
    assert abs(l2 - ref) <= max(tol * ref * 4, 1e-12), f"{what}{name}: l2 {l2} vs {ref}"
    return e


def check_adam_weights(got, ref, name="", frob=1e-3, frac_tol=1e-4):
    """Adam-updated weights.  One Adam step moves every weight by ~lr*g/(|g|+eps): where a gradient
    entry is within rounding of zero the step's sign depends on fp32 summation order, so a handful
    of entries legitimately differ by up to 2*lr between ANY two fp32 implementations (torch-CPU
    fp32 vs fp64-derived gradients already differ by 1.05e-4 norm-wise on fc1).
    Measured on a B200, one teacher-forced step: 1.3e-6 (fc1) and 1.4e-5 (fc2) of the entries are off
    by more than 1e-5, all with |g| < 1e-6.  Require: norm-wise 1e-3, and all but a 1e-4 fraction of
    entries within 1e-5 absolute."""
    a = np.asarray(got, np.float64).reshape(-1)
    b = np.asarray(ref, np.float64).reshape(-1)
    e = rel_err(a, b)
    assert e <= frob, f"{name}: rel err {e:.3e}"
    frac = float((np.abs(a - b) > 1e-5).mean())
    assert frac <= frac_tol, f"{name}: {frac:.2e} of entries differ by more than 1e-5"


def check_grad_through_relu(got, ref, fwd_eps, name=""):
    """Gradients that passed through ReLU gates computed by a forward with relative error
    `fwd_eps`: a pre-activation within fwd_eps of zero can land on the other side of the kink, which
    changes that gradient entry by O(1).  The flipped fraction is ~fwd_eps, so the norm-wise error is
    ~sqrt(fwd_eps) (measured: 3xTF32 forward 3.9e-5 at K=5408 -> 5e-3 on the cut gradient, while
    every single GEMM of the backward is exact to 1e-6; one flipped hidden unit changes a whole
    row of the layer below, so no per-entry criterion applies).  Require the norm-wise bound."""
    e = rel_err(got, ref)
    assert e <= 4.0 * np.sqrt(fwd_eps), f"{name}: rel err {e:.3e}"
