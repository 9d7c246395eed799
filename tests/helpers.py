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
