"""Plain-numpy statement of the operator formulas on the hot path, fp32, for SMALL cases
(TEST INFRASTRUCTURE — see oracle/__init__.py).

The reference delegates this arithmetic to PyTorch (third-party, not vendored in
`/root/reference`); these are the published definitions of those operators, written out so the
backward formulas the CUDA kernels implement are stated somewhere explicitly.  tests/ checks
them against `torch_ref.py` (autograd) before they are trusted.
"""
import numpy as np

f32 = np.float32


def conv_pool_forward(x, w, b):
    """conv3x3 valid (cross-correlation) + bias -> ReLU -> maxpool 2x2 stride 2 -> flatten.
    x [B,1,28,28], w [32,1,3,3], b [32] -> act [B,5408], plus the per-output winner position
    (first maximum in row-major window order, torch's CPU tie-break) for the backward.
    Reference call site: models.py:20-25,27-30."""
    B = x.shape[0]
    conv = np.zeros((B, 32, 26, 26), f32)
    for u in range(3):
        for v in range(3):
            conv += w[None, :, 0, u, v, None, None] * x[:, 0, None, u:u + 26, v:v + 26]
    conv += b[None, :, None, None]
    win = conv.reshape(B, 32, 13, 2, 13, 2).transpose(0, 1, 2, 4, 3, 5).reshape(B, 32, 13, 13, 4)
    arg = win.argmax(-1)                       # first max wins
    pooled = np.maximum(win.max(-1), 0)        # relu(max) == max(relu)
    return pooled.reshape(B, 5408).astype(f32), arg, conv


def conv_pool_backward(x, w, b, dact):
    """Gradients of the block above w.r.t. w and b (no dX: the input is data)."""
    B = x.shape[0]
    _, arg, conv = conv_pool_forward(x, w, b)
    d = dact.reshape(B, 32, 13, 13)
    dconv = np.zeros((B, 32, 26, 26), f32)
    bi, ci, ii, jj = np.meshgrid(np.arange(B), np.arange(32), np.arange(13), np.arange(13), indexing="ij")
    hh = 2 * ii + arg // 2
    ww = 2 * jj + arg % 2
    live = conv[bi, ci, hh, ww] > 0            # ReLU gate
    dconv[bi, ci, hh, ww] = np.where(live, d, 0)
    dw = np.zeros((32, 1, 3, 3), f32)
    for u in range(3):
        for v in range(3):
            dw[:, 0, u, v] = np.einsum("bchw,bhw->c", dconv, x[:, 0, u:u + 26, v:v + 26])
    db = dconv.sum((0, 2, 3))
    return dw.astype(f32), db.astype(f32)


def mlp_forward(act, params, relu_last, dropout_p, masks):
    """Linear stack: relu on hidden layers (+ last if `relu_last`), inverted dropout with given
    keep-masks on hidden layers.  Returns logits and the per-layer inputs (post-dropout)."""
    h = act.astype(f32)
    L = len(params) // 2
    inputs = []
    for l in range(L):
        inputs.append(h)
        z = h @ params[2 * l].T + params[2 * l + 1]
        last = l == L - 1
        if not last or relu_last:
            z = np.maximum(z, 0)
        if not last and dropout_p > 0 and masks is not None:
            z = z * (masks[l].astype(f32) / f32(1.0 - dropout_p))
        h = z.astype(f32)
    return h, inputs


def cross_entropy(logits, labels):
    """mean_b( logsumexp(z_b) - z_b[y_b] ) and its gradient (softmax - onehot)/B."""
    z = logits.astype(f32)
    zmax = z.max(1, keepdims=True)
    e = np.exp(z - zmax)
    s = e.sum(1, keepdims=True)
    B = z.shape[0]
    loss = (np.log(s[:, 0]) + zmax[:, 0] - z[np.arange(B), labels]).mean()
    g = e / s
    g[np.arange(B), labels] -= 1
    return f32(loss), (g / B).astype(f32)


def mlp_backward(dout, out, inputs, params, relu_last, dropout_p, masks, need_dx=True):
    """Backward of `mlp_forward`.  `dout` is dL/d(output of the stack), `out` the stack output
    (needed only when `relu_last`).  Returns ([dW1,db1,...], dX or None)."""
    L = len(params) // 2
    grads = [None] * (2 * L)
    dz = dout.astype(f32)
    if relu_last:
        dz = dz * (out > 0)
    for l in reversed(range(L)):
        a = inputs[l]
        grads[2 * l] = (dz.T @ a).astype(f32)
        grads[2 * l + 1] = dz.sum(0).astype(f32)
        if l == 0 and not need_dx:
            return grads, None
        da = dz @ params[2 * l]
        if l > 0:
            # a = relu(z_prev) * mask / (1-p): gate by (a > 0) and rescale by the same factor
            if dropout_p > 0 and masks is not None:
                da = da * (masks[l - 1].astype(f32) / f32(1.0 - dropout_p))
                da = da * (a > 0)
            else:
                da = da * (a > 0)
        dz = da.astype(f32)
    return grads, dz


def sgd_momentum(p, g, buf, lr, momentum, first):
    """`F.sgd`: buf <- g on the first step else momentum*buf + g ; p <- p - lr*buf."""
    buf = g.copy() if first else (f32(momentum) * buf + g).astype(f32)
    return (p - f32(lr) * buf).astype(f32), buf


def adam(p, g, m, v, t, lr, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0):
    """`_single_tensor_adam`: coupled L2, lerp first moment, bias corrections in double."""
    if weight_decay != 0:
        g = (g + f32(weight_decay) * p).astype(f32)
    m = (m + f32(1 - beta1) * (g - m)).astype(f32)
    v = (v * f32(beta2) + f32(1 - beta2) * g * g).astype(f32)
    bc1 = 1 - beta1 ** t
    bc2 = 1 - beta2 ** t
    step_size = f32(lr / bc1)
    denom = (np.sqrt(v) / f32(bc2 ** 0.5) + f32(eps)).astype(f32)
    return (p - step_size * (m / denom)).astype(f32), m, v
