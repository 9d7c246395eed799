"""Thin host wrappers over the C-ABI handles.  torch is used for device memory and streams only;
all arithmetic happens inside libsplitnn.so.  Every call passes raw device pointers
(`tensor.data_ptr()`) and torch's current CUDA stream.
"""
import ctypes as C

import torch

from . import _lib
from ._lib import SN_OPT_ADAM, SN_OPT_NONE, SN_OPT_SGD_MOMENTUM, check, ptr, ptr_array  # noqa: F401

D0 = 5408


def set_matmul_precision(mode):
    """Large-batch (B > 64) products: 0 = 3xTF32 tcgen05 (default, fp32-accurate), 1 = TF32 tcgen05
    (perf mode), 2 = fp32 SIMT fallback."""
    check(_lib.load().sn_set_matmul_precision(int(mode)))


def set_skinny_tensor_forward(on):
    """Batch <= 16 forward of big layers on tensor cores (3xTF32, weights in tensor memory) instead of
    the FMA kernel.  fp32-accurate but slower at this batch size (see include/splitnn.h); off by default."""
    check(_lib.load().sn_set_skinny_tensor_forward(1 if on else 0))


def _stream(device=None):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _f32c(t, device=None):
    if t.dtype != torch.float32 or not t.is_contiguous() or not t.is_cuda:
        raise _lib.SplitNNError("libsplitnn expects contiguous float32 CUDA tensors "
                                f"(got {t.dtype}, cuda={t.is_cuda}, contiguous={t.is_contiguous()})")
    return t


def _opt_args(optimizer, lr, momentum, betas, eps, weight_decay):
    return (int(optimizer), float(lr), float(momentum), float(betas[0]), float(betas[1]), float(eps),
            float(weight_decay))


class BobEngine:
    """`sn_bob`: a server MLP bound to caller-owned weight tensors (updated in place)."""

    def __init__(self, dims, weights, biases, *, relu_last=False, dropout_p=0.5, optimizer=SN_OPT_SGD_MOMENTUM,
                 lr=1e-3, momentum=0.9, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, n_opt_slots=1,
                 max_batch=16):
        self.lib = _lib.load()
        self.dims = [int(d) for d in dims]
        self.L = len(self.dims) - 1
        assert len(weights) == self.L and len(biases) == self.L
        self.device = weights[0].device
        for l, (w, b) in enumerate(zip(weights, biases)):
            _f32c(w), _f32c(b)
            assert tuple(w.shape) == (self.dims[l + 1], self.dims[l]) and tuple(b.shape) == (self.dims[l + 1],)
        self.weights, self.biases = list(weights), list(biases)   # keep the storage alive
        self.max_batch = int(max_batch)
        self.n_opt_slots = int(n_opt_slots)
        self.dropout_p = float(dropout_p)
        self.optimizer = int(optimizer)
        h = C.c_void_p()
        cdims = (C.c_int * len(self.dims))(*self.dims)
        check(self.lib.sn_bob_create(C.byref(h), self.device.index or 0, cdims, self.L, int(bool(relu_last)),
                                     float(dropout_p), *_opt_args(optimizer, lr, momentum, betas, eps, weight_decay),
                                     self.n_opt_slots, self.max_batch, ptr_array(self.weights), ptr_array(self.biases)))
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.lib.sn_bob_destroy(self.h)
            self.h = None

    __del__ = close

    def _masks(self, masks):
        if masks is None:
            return None, None
        keep = [m if m.dtype == torch.uint8 else m.to(torch.uint8) for m in masks]
        keep = [m.contiguous() for m in keep]
        return ptr_array(keep), keep

    def train_step(self, act, labels, slot=0, masks=None, seed=0, offset=0, want_cut_grad=True,
                   want_logits=False, loss_out=None, cut_out=None):
        """Returns (loss[device scalar], cut_grad or None, logits or None).  `cut_out`: where the
        cut-layer gradient goes (e.g. a peer GPU's mailbox, peer.py); allocated when omitted."""
        B = act.shape[0]
        _f32c(act)
        cut = cut_out if cut_out is not None else (torch.empty_like(act) if want_cut_grad else None)
        logits = torch.empty(B, self.dims[-1], device=act.device) if want_logits else None
        loss = loss_out if loss_out is not None else torch.empty((), device=act.device)
        marr, keep = self._masks(masks)
        check(self.lib.sn_bob_train_step(self.h, slot, ptr(act), ptr(labels), B, marr, seed, offset,
                                         ptr(cut), ptr(logits), ptr(loss), _stream(act.device)))
        return loss, cut, logits

    def forward(self, act, train_mode=False, masks=None, seed=0, offset=0, out=None):
        B = act.shape[0]
        _f32c(act)
        out = out if out is not None else torch.empty(B, self.dims[-1], device=act.device)
        marr, keep = self._masks(masks)
        check(self.lib.sn_bob_forward(self.h, ptr(act), B, int(bool(train_mode)), marr, seed, offset, ptr(out),
                                      _stream(act.device)))
        return out

    def backward_update(self, act, dout, slot=0, want_cut_grad=True):
        B = act.shape[0]
        _f32c(act), _f32c(dout)
        cut = torch.empty_like(act) if want_cut_grad else None
        check(self.lib.sn_bob_backward_update(self.h, slot, ptr(act), ptr(dout), B, ptr(cut), _stream(act.device)))
        return cut

    def server_epoch(self, cache, labels, batch_offsets, slot=0, masks=None, seed=0, offset=0, loss_per_step=None):
        _f32c(cache)
        offs = (C.c_int64 * len(batch_offsets))(*[int(o) for o in batch_offsets])
        marr, keep = self._masks(masks)
        check(self.lib.sn_bob_server_epoch(self.h, slot, ptr(cache), ptr(labels), offs, len(batch_offsets) - 1,
                                           marr, seed, offset, ptr(loss_per_step), _stream(cache.device)))

    def reset_opt_slot(self, slot=0):
        check(self.lib.sn_bob_reset_opt_slot(self.h, slot, _stream(self.device)))

    def opt_state(self, slot, which, param):
        """A torch view (no copy) of an optimizer-state buffer."""
        p, n = C.c_void_p(), C.c_size_t()
        check(self.lib.sn_bob_opt_state(self.h, slot, which, param, C.byref(p), C.byref(n)))
        return _view_f32(p.value, n.value, self.device)

    def get_opt_step(self, slot=0):
        s = C.c_int64()
        check(self.lib.sn_bob_get_opt_step(self.h, slot, C.byref(s)))
        return s.value

    def set_opt_step(self, slot, step):
        check(self.lib.sn_bob_set_opt_step(self.h, slot, int(step)))

    def set_lr(self, lr):
        check(self.lib.sn_bob_set_lr(self.h, float(lr)))

    def profile(self, enable):
        check(self.lib.sn_bob_profile(self.h, int(bool(enable))))

    def profile_read(self, kind, layer):
        """(total ms, launches) of kind 0 = forward / 1 = backward+update kernels of `layer`."""
        t, n = C.c_double(), C.c_int()
        check(self.lib.sn_bob_profile_read(self.h, kind, layer, C.byref(t), C.byref(n)))
        return t.value, n.value

    def set_ce_heads(self, heads):
        check(self.lib.sn_bob_set_ce_heads(self.h, int(heads)))

    def set_stream_rows(self, rows, max_scratch_bytes=1 << 30):
        """Let eval-mode `forward` calls of up to `rows` rows run the exact fp32 weight-stream
        kernels (row tiles of 16 in one launch per layer); returns the row count granted."""
        got = C.c_int(0)
        check(self.lib.sn_bob_set_stream_rows(self.h, int(rows), int(max_scratch_bytes), C.byref(got)))
        self.stream_rows = got.value
        return got.value


class ShardEngine:
    """`sn_shard`: rows [n_off, n_off+n_rows) of a Linear(+ReLU, dropout) layer with that slice's
    optimizer state - the unit of the tensor-parallel first layer (tp.py).  Borrows `W_slice`
    [n_rows,K] and `b_slice` [n_rows] (updated in place)."""

    def __init__(self, W_slice, b_slice, n_total, n_off, *, relu=True, dropout_p=0.5, optimizer=SN_OPT_ADAM, lr=1e-3,
                 momentum=0.0, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0):
        self.lib = _lib.load()
        self.W, self.b = _f32c(W_slice), _f32c(b_slice)
        self.device = W_slice.device
        self.n_rows, self.K = W_slice.shape
        self.n_total, self.n_off = int(n_total), int(n_off)
        h = C.c_void_p()
        dev = self.device.index if self.device.index is not None else torch.cuda.current_device()
        check(self.lib.sn_shard_create(C.byref(h), dev, self.K, self.n_total, self.n_off, self.n_rows, int(bool(relu)),
                                       float(dropout_p), *_opt_args(optimizer, lr, momentum, betas, eps, weight_decay),
                                       ptr(self.W), ptr(self.b)))
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.lib.sn_shard_destroy(self.h)
            self.h = None

    __del__ = close

    def forward(self, X, H_out, ld_out, train=True, mask=None, seed=0, offset=0):
        """`H_out`: the full-width [B, n_total] activation (tensor or raw peer address holder)."""
        _f32c(X)
        check(self.lib.sn_shard_forward(self.h, ptr(X), X.shape[0], int(bool(train)), ptr(mask), int(seed), int(offset),
                                        ptr(H_out), int(ld_out), _stream(self.device)))

    def backward_update(self, X, dH, ld_dh, X_next=None):
        _f32c(X)
        check(self.lib.sn_shard_backward_update(self.h, ptr(X), ptr(dH), int(ld_dh), X.shape[0], ptr(X_next),
                                                0 if X_next is None else X_next.shape[0], _stream(self.device)))

    def reset_opt(self):
        check(self.lib.sn_shard_reset_opt(self.h, _stream(self.device)))

    def opt_state(self, which, param):
        """A torch view (no copy) of the slice's optimizer state: which 0/1 = first/second moment,
        param 0/1 = weight slice / bias slice."""
        p, n = C.c_void_p(), C.c_size_t()
        check(self.lib.sn_shard_opt_state(self.h, which, param, C.byref(p), C.byref(n)))
        return _view_f32(p.value, n.value, self.device)

    def get_opt_step(self):
        s = C.c_int64()
        check(self.lib.sn_shard_get_opt_step(self.h, C.byref(s)))
        return s.value

    def set_opt_step(self, step):
        check(self.lib.sn_shard_set_opt_step(self.h, int(step)))


class AliceEngine:
    """`sn_alice`: the client conv block bound to caller-owned conv weight/bias tensors."""

    def __init__(self, w, b, *, optimizer=SN_OPT_SGD_MOMENTUM, lr=1e-3, momentum=0.9, betas=(0.9, 0.999), eps=1e-8,
                 weight_decay=0.0, max_batch=16):
        self.lib = _lib.load()
        _f32c(w), _f32c(b)
        assert tuple(w.shape) == (32, 1, 3, 3) and tuple(b.shape) == (32,)
        self.w, self.b = w, b
        self.device = w.device
        self.max_batch = int(max_batch)
        h = C.c_void_p()
        check(self.lib.sn_alice_create(C.byref(h), self.device.index or 0,
                                       *_opt_args(optimizer, lr, momentum, betas, eps, weight_decay),
                                       self.max_batch, ptr(w), ptr(b)))
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.lib.sn_alice_destroy(self.h)
            self.h = None

    __del__ = close

    def forward(self, x, out=None, ld=None):
        """act = conv block(x).  `out`: where to write (a tensor, or a raw peer address wrapper);
        `ld`: row stride of `out` in floats when the 5408 activations of a sample are one slot of
        a wider row (the --concat server's cache) - `out` then points at the slot's first column."""
        B = x.shape[0]
        _f32c(x)
        if ld is None:
            out = out if out is not None else torch.empty(B, D0, device=x.device)
            check(self.lib.sn_alice_forward(self.h, ptr(x), B, ptr(out), _stream(x.device)))
        else:
            if out is None or int(ld) < D0:
                raise ValueError("forward(ld=...) needs an output address and ld >= 5408")
            check(self.lib.sn_alice_forward_ld(self.h, ptr(x), B, C.c_void_p(out.data_ptr()), int(ld), _stream(x.device)))
        return out

    def forward_host(self, host_x, host_y, x_dev, y_dev, out=None):
        """`sn_alice_forward_host`: the batch is still in PINNED host memory; the conv kernel reads it over
        PCIe in its prologue, writes the activations to `out` (a tensor or a raw peer address) and leaves
        device copies in `x_dev` / `y_dev` (alternate between two pairs from step to step)."""
        if not host_x.is_pinned() or (host_y is not None and not host_y.is_pinned()):
            raise ValueError("forward_host needs pinned host tensors (tensor.pin_memory())")
        B = host_x.shape[0]
        out = out if out is not None else torch.empty(B, D0, device=self.device)
        with torch.cuda.device(self.device):
            check(self.lib.sn_alice_forward_host(self.h, ptr(host_x), ptr(host_y), B, C.c_void_p(out.data_ptr()), ptr(x_dev),
                                                 ptr(y_dev) if host_y is not None else None, _stream(self.device)))
        return out

    def backward_update(self, x, cut_grad):
        _f32c(x), _f32c(cut_grad)
        check(self.lib.sn_alice_backward_update(self.h, ptr(x), ptr(cut_grad), x.shape[0], _stream(x.device)))

    def local_step(self, x, labels, loss_out=None):
        _f32c(x)
        loss = loss_out if loss_out is not None else torch.empty((), device=x.device)
        check(self.lib.sn_alice_local_step(self.h, ptr(x), ptr(labels), x.shape[0], ptr(loss), _stream(x.device)))
        return loss

    def local_epoch(self, x, labels, batch, loss_per_step=None):
        """All batches of one pass over (x [n,1,28,28], labels [n]) in a single library call."""
        _f32c(x)
        check(self.lib.sn_alice_local_epoch(self.h, ptr(x), ptr(labels), int(x.shape[0]), int(batch),
                                            ptr(loss_per_step), _stream(x.device)))

    def reset_opt(self):
        check(self.lib.sn_alice_reset_opt(self.h, _stream(self.device)))

    def opt_state(self, which, param):
        p, n = C.c_void_p(), C.c_size_t()
        check(self.lib.sn_alice_opt_state(self.h, which, param, C.byref(p), C.byref(n)))
        return _view_f32(p.value, n.value, self.device)

    def get_opt_step(self):
        s = C.c_int64()
        check(self.lib.sn_alice_get_opt_step(self.h, C.byref(s)))
        return s.value

    def set_opt_step(self, step):
        check(self.lib.sn_alice_set_opt_step(self.h, int(step)))


def vanilla_step(alice, bob, x, labels, slot, seed, offset, act, cut, loss):
    """`sn_vanilla_step`: Alice forward, Bob train step, Alice backward+update in one library call
    (all on one device; `act`/`cut` are preallocated [B,5408] buffers)."""
    check(bob.lib.sn_vanilla_step(alice.h, bob.h, int(slot), ptr(x), ptr(labels), int(x.shape[0]), int(seed), int(offset),
                                  ptr(act), ptr(cut), ptr(loss), _stream(x.device)))


def vanilla_step_host(alice, bob, host_x, host_y, slot, seed, offset, x_dev, y_dev, act, cut, loss):
    """`sn_vanilla_step_host`: the same step with the batch still in PINNED host memory - the conv kernel
    pulls it over PCIe in its prologue and leaves device copies in `x_dev` / `y_dev` (alternate between
    two pairs from step to step)."""
    if not (host_x.is_pinned() and host_y.is_pinned()):
        raise ValueError("vanilla_step_host needs pinned host tensors (tensor.pin_memory())")
    dev = act.device
    with torch.cuda.device(dev):
        check(bob.lib.sn_vanilla_step_host(alice.h, bob.h, int(slot), ptr(host_x), ptr(host_y), int(host_x.shape[0]), int(seed),
                                           int(offset), ptr(x_dev), ptr(y_dev), ptr(act), ptr(cut), ptr(loss), _stream(dev)))


class _RawCuda:
    """Minimal __cuda_array_interface__ carrier so torch can view library-owned memory."""

    def __init__(self, address, count):
        self.__cuda_array_interface__ = {"shape": (count,), "typestr": "<f4", "data": (address, False), "version": 2}


def _view_f32(address, count, device):
    with torch.cuda.device(device):
        return torch.as_tensor(_RawCuda(address, count), device=device)


# ------------------------------------------------------------------------- stateless operators
def conv_pool_forward(x, w, b):
    lib = _lib.load()
    out = torch.empty(x.shape[0], D0, device=x.device)
    with torch.cuda.device(x.device):          # the stateless entry points launch on the CURRENT device
        check(lib.sn_conv_pool_forward(ptr(_f32c(x)), ptr(_f32c(w)), ptr(_f32c(b)), x.shape[0], ptr(out), _stream(x.device)))
    return out


def conv_pool_backward(x, w, b, dact):
    lib = _lib.load()
    B = x.shape[0]
    dw = torch.empty(32, 1, 3, 3, device=x.device)
    db = torch.empty(32, device=x.device)
    with torch.cuda.device(x.device):
        scratch = torch.empty(lib.sn_conv_pool_backward_scratch_bytes(B), dtype=torch.uint8, device=x.device)
        check(lib.sn_conv_pool_backward(ptr(_f32c(x)), ptr(_f32c(w)), ptr(_f32c(b)), ptr(_f32c(dact)), B, ptr(dw), ptr(db),
                                        ptr(scratch), _stream(x.device)))
    return dw, db


def linear_forward(x, W, bias, relu=False, keep_mask=None, mask_scale=1.0):
    lib = _lib.load()
    B, K = x.shape
    N = W.shape[0]
    y = torch.empty(B, N, device=x.device)
    with torch.cuda.device(x.device):
        nbytes = lib.sn_linear_forward_scratch_bytes(B, K, N)
        scratch = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
        check(lib.sn_linear_forward(ptr(_f32c(x)), ptr(_f32c(W)), ptr(_f32c(bias)), B, K, N, int(bool(relu)),
                                    ptr(keep_mask), float(mask_scale), ptr(y), ptr(scratch), nbytes, _stream(x.device)))
    return y


def conv_weights_copy(dst_w, dst_b, src_w, src_b):
    """`sn_conv_weights_copy`: dst <- src for a conv block's weight and bias, enqueued on dst's stream."""
    with torch.cuda.device(dst_w.device):
        check(_lib.load().sn_conv_weights_copy(ptr(_f32c(dst_w)), ptr(_f32c(dst_b)), ptr(_f32c(src_w)), ptr(_f32c(src_b)),
                                               _stream(dst_w.device)))


def linear_forward_noalloc(x, W, bias, relu=False, keep_mask=None, mask_scale=1.0, out=None):
    """`sn_linear_forward_noalloc`: same result as `linear_forward`, but the library uses its own
    built-in 4 MiB scratch and allocates nothing.  Pass `out` to keep this wrapper allocation-free
    as well (otherwise torch allocates the result tensor)."""
    lib = _lib.load()
    B, K = x.shape
    N = W.shape[0]
    y = out if out is not None else torch.empty(B, N, device=x.device)
    with torch.cuda.device(x.device):
        check(lib.sn_linear_forward_noalloc(ptr(_f32c(x)), ptr(_f32c(W)), ptr(_f32c(bias)), B, K, N, int(bool(relu)),
                                            ptr(keep_mask), float(mask_scale), ptr(_f32c(y)), _stream(x.device)))
    return y


def cross_entropy(logits, labels, want_grad=True):
    lib = _lib.load()
    B, Cn = logits.shape
    loss = torch.empty((), device=logits.device)
    d = torch.empty_like(logits) if want_grad else None
    with torch.cuda.device(logits.device):
        check(lib.sn_cross_entropy(ptr(_f32c(logits)), ptr(labels), B, Cn, ptr(loss), ptr(d), _stream(logits.device)))
    return loss, d


def eval_counts(logits, labels, omit_label, counters=None):
    lib = _lib.load()
    B, Cn = logits.shape
    if counters is None:
        counters = torch.zeros(6, dtype=torch.int64, device=logits.device)
    with torch.cuda.device(logits.device):
        check(lib.sn_eval_counts(ptr(_f32c(logits)), ptr(labels), B, Cn, int(omit_label), ptr(counters),
                                 _stream(logits.device)))
    return counters
