"""Network definitions with the reference's class names, constructor arguments and `state_dict`
keys/shapes (`/root/reference/models.py:5-94`), whose `forward` runs the sm_100a kernels in
libsplitnn.so instead of PyTorch operators.

The `nn.Conv2d` / `nn.Linear` children are parameter containers only: they give the same
initialisation draws as the reference (same constructors in the same order, on the CPU generator),
the same `state_dict` layout (`conv_layers.0.*`, `conv1.*`, `fc1/fc2/fc3.*`), and are never called.
Parameters live on the CUDA device; training engines (`make_engine`) bind them by pointer and
update them in place, so `state_dict()` / `load_state_dict()` always see current values.
"""
import torch
import torch.nn as nn

from . import engine as E


def default_device():
    if not torch.cuda.is_available():
        raise E._lib.SplitNNError("splitlearning_b200 needs a CUDA device (B200); there is no CPU path")
    return torch.device("cuda", torch.cuda.current_device())


class PhiloxMasks:
    """Perf-mode dropout: masks are generated inside the kernels from (seed, step counter)."""

    def __init__(self, seed=0):
        self.seed, self.offset = int(seed), 0

    def next(self):
        self.offset += 1
        return None, self.seed, self.offset


class TorchCpuMasks:
    """Parity-mode dropout: the keep-masks torch's CPU `F.dropout` would draw from `generator`
    (None = the global CPU generator), in layer order, shipped to the device as bytes.  With the
    same seed this reproduces the reference's mask stream bit for bit."""

    def __init__(self, widths, p, generator=None):
        self.widths, self.p, self.generator = list(widths), float(p), generator

    def draw(self, batch, device):
        return [torch.empty(batch, n).bernoulli_(1.0 - self.p, generator=self.generator).to(torch.uint8)
                .to(device, non_blocking=True) for n in self.widths]


class _ConvBlock(nn.Module):
    """conv3x3(1->32) + ReLU + maxpool2x2 [+ flatten]; output memory is [B,5408] either way."""
    flatten_out = True

    def _wb(self):
        raise NotImplementedError

    def forward(self, x):
        w, b = self._wb()
        x = x.to(device=w.device, dtype=torch.float32).reshape(-1, 1, 28, 28).contiguous()
        act = E.conv_pool_forward(x, w.data, b.data)
        return act if self.flatten_out else act.view(-1, 32, 13, 13)

    def make_engine(self, **opt):
        w, b = self._wb()
        return E.AliceEngine(w.data, b.data, **opt)

    def copy_weights_from(self, other):
        """The Alice -> Alice weight hand-off (`give_weights` + `load_state_dict`,
        `data_entities_vanilla.py:59-64`) as a stream-ordered device copy of 288 + 32 floats: no
        host-side state_dict clone; across GPUs two peer copies."""
        (w, b), (ow, ob) = self._wb(), other._wb()
        if w.device == ow.device:
            E.conv_weights_copy(w.data, b.data, ow.data, ob.data)      # one small launch in the stream's chain
        else:
            with torch.no_grad():
                w.copy_(ow, non_blocking=True)
                b.copy_(ob, non_blocking=True)


class model1(_ConvBlock):
    """U-shape client head (`models.py:5-14`): keys `conv1.weight`, `conv1.bias`."""
    flatten_out = False

    def __init__(self, device=None):
        super().__init__()
        self.conv1 = nn.Conv2d(1, 32, 3)
        self.pool = nn.MaxPool2d(2, 2)
        self.to(device or default_device())

    def _wb(self):
        return self.conv1.weight, self.conv1.bias


class model1_sisa(_ConvBlock):
    """Vanilla/SISA client block (`models.py:16-30`): keys `conv_layers.0.weight/bias`."""

    def __init__(self, device=None):
        super().__init__()
        self.conv_layers = nn.Sequential(nn.Conv2d(in_channels=1, out_channels=32, kernel_size=3), nn.ReLU(),
                                         nn.MaxPool2d(2, 2))
        self.flatten = nn.Flatten()
        self.to(device or default_device())

    def _wb(self):
        return self.conv_layers[0].weight, self.conv_layers[0].bias


class _ServerMLP(nn.Module):
    """Linear stack; `_layers` names the nn.Linear children in order."""
    relu_last = False
    dropout_p = 0.0
    _layers = ()

    def _finish(self, device):
        self.to(device or default_device())
        self._infer = None
        self.mask_source = PhiloxMasks()      # replace with TorchCpuMasks for reference-stream parity

    @property
    def dims(self):
        ls = [getattr(self, n) for n in self._layers]
        return [ls[0].in_features] + [l.out_features for l in ls]

    def weight_list(self):
        return [getattr(self, n).weight.data for n in self._layers]

    def bias_list(self):
        return [getattr(self, n).bias.data for n in self._layers]

    def make_engine(self, **kw):
        kw.setdefault("relu_last", self.relu_last)
        kw.setdefault("dropout_p", self.dropout_p)
        return E.BobEngine(self.dims, self.weight_list(), self.bias_list(), **kw)

    def dropout_args(self, batch, device):
        """(masks, seed, offset) for one train-mode forward."""
        if self.dropout_p <= 0:
            return None, 0, 0
        if isinstance(self.mask_source, TorchCpuMasks):
            return self.mask_source.draw(batch, device), 0, 0
        return self.mask_source.next()

    # Inference over more than 64 rows (a whole test shard at once): chunks of up to this many rows go
    # through the exact fp32 weight-stream kernels in one launch per layer (sn_bob_set_stream_rows), so
    # every row gets the numbers a batch-of-16 loop would give it.
    max_infer_rows = 4096
    # rows per dropout draw in train-mode inference with TorchCpuMasks: the reference draws one mask per
    # loader batch, so a fused pass must consume the CPU stream batch by batch (None: one draw per call)
    eval_batch = None

    def _chunk_masks(self, rows, device):
        """(masks, seed, offset) for a train-mode inference chunk of `rows` rows."""
        bs = self.eval_batch
        if bs and isinstance(self.mask_source, TorchCpuMasks) and self.dropout_p > 0 and rows > bs:
            per_layer = [[] for _ in self.mask_source.widths]
            for s in range(0, rows, bs):
                for l, m in enumerate(self.mask_source.draw(min(bs, rows - s), device)):
                    per_layer[l].append(m)
            return [torch.cat(ms, 0) for ms in per_layer], 0, 0
        return self.dropout_args(rows, device)

    def forward(self, x):
        w0 = self.weight_list()[0]
        x = torch.flatten(x.to(device=w0.device, dtype=torch.float32), 1).contiguous()
        B = x.shape[0]
        want = B if B <= 64 else min(B, self.max_infer_rows)
        if self._infer is None or self._infer.max_batch < want or self._infer.weights[0].data_ptr() != w0.data_ptr():
            self._infer = self.make_engine(optimizer=E.SN_OPT_NONE, max_batch=max(want, 16))
            self._infer_rows = self._infer.set_stream_rows(want) if want > 64 else self._infer.max_batch
        rows = max(1, min(self._infer_rows, self._infer.max_batch))
        if B <= rows:
            masks, seed, offset = self._chunk_masks(B, x.device) if self.training else (None, 0, 0)
            return self._infer.forward(x, train_mode=self.training, masks=masks, seed=seed, offset=offset)
        out = torch.empty(B, self.dims[-1], device=x.device)
        for s in range(0, B, rows):
            n = min(rows, B - s)
            masks, seed, offset = self._chunk_masks(n, x.device) if self.training else (None, 0, 0)
            self._infer.forward(x[s:s + n], train_mode=self.training, masks=masks, seed=seed, offset=offset, out=out[s:s + n])
        return out


class model2(_ServerMLP):
    """U-shape server (`models.py:33-44`): relu(fc1 5408->1000), relu(fc2 1000->100)."""
    relu_last = True
    _layers = ("fc1", "fc2")

    def __init__(self, device=None):
        super().__init__()
        self.fc1 = nn.Linear(32 * 13 * 13, 1000)
        self.fc2 = nn.Linear(1000, 100)
        self._finish(device)


class model2_sisa(_ServerMLP):
    """Vanilla/SISA server (`models.py:46-63`): 5408->5000->1000->100, dropout 0.5 after fc1, fc2."""
    dropout_p = 0.5
    _layers = ("fc1", "fc2", "fc3")

    def __init__(self, device=None):
        super().__init__()
        self.fc1 = nn.Linear(32 * 13 * 13, 5000)
        self.dropout1 = nn.Dropout(0.5)
        self.fc2 = nn.Linear(5000, 1000)
        self.dropout2 = nn.Dropout(0.5)
        self.fc3 = nn.Linear(1000, 100)
        self._finish(device)


class model2_sisa_concat(_ServerMLP):
    """Concat server (`models.py:66-82`): fc1 takes 5408*num_clients, fc3 gives 100*num_clients."""
    dropout_p = 0.5
    _layers = ("fc1", "fc2", "fc3")

    def __init__(self, num_clients, device=None):
        super().__init__()
        self.num_clients = num_clients
        self.fc1 = nn.Linear(32 * 13 * 13 * num_clients, 5000)
        self.dropout1 = nn.Dropout(0.5)
        self.fc2 = nn.Linear(5000, 1000)
        self.dropout2 = nn.Dropout(0.5)
        self.fc3 = nn.Linear(1000, 100 * num_clients)
        self._finish(device)


class model3(_ServerMLP):
    """U-shape client tail (`models.py:87-94`): fc 100->10."""
    _layers = ("fc3",)

    def __init__(self, device=None):
        super().__init__()
        self.fc3 = nn.Linear(100, 10)
        self._finish(device)
