"""Torch-CPU restatement of the reference's hot path (TEST INFRASTRUCTURE — see oracle/__init__.py).

Every function calls the same PyTorch CPU operators the reference calls, with dropout masks and
optimizer state made explicit so that one step can be teacher-forced.  Citations are into
`/root/reference`.
"""
from dataclasses import dataclass, field
from typing import List, Optional, Sequence

import torch
import torch.nn.functional as F


# ----------------------------------------------------------------------------- model shapes
@dataclass(frozen=True)
class BobSpec:
    """Server MLP shape. `relu_last`: U-shape `model2` applies ReLU to its last layer too."""
    dims: tuple
    relu_last: bool = False
    dropout_p: float = 0.5

    @property
    def n_layers(self):
        return len(self.dims) - 1


MODEL2_SISA = BobSpec((5408, 5000, 1000, 100), False, 0.5)      # models.py:46-63
MODEL2 = BobSpec((5408, 1000, 100), True, 0.0)                  # models.py:33-44


def model2_sisa_concat_spec(num_clients):                       # models.py:66-82
    return BobSpec((5408 * num_clients, 5000, 1000, 100 * num_clients), False, 0.5)


# ----------------------------------------------------------------------------- forward pieces
def alice_forward(x, conv_w, conv_b, flatten=True):
    """`model1_sisa.forward` (models.py:27-30) / `model1.forward` (:11-14, flatten=False):
    conv3x3(1->32) -> ReLU -> maxpool 2x2 (-> flatten, channel-major c*169+h*13+w)."""
    y = F.max_pool2d(F.relu(F.conv2d(x, conv_w, conv_b)), 2, 2)
    return torch.flatten(y, 1) if flatten else y


def bob_forward(act, params: Sequence[torch.Tensor], spec: BobSpec, masks=None):
    """`model2_sisa.forward` (models.py:55-63), `model2.forward` (:39-44),
    `model2_sisa_concat.forward` (:75-82).

    `params` = [W1,b1,W2,b2,...] with W as [out,in] (state_dict order).  `masks` = one {0,1}
    keep-mask per hidden layer (train mode, `F.dropout` == x*mask/(1-p)) or None
    (eval mode / no dropout)."""
    h = torch.flatten(act, 1)
    L = spec.n_layers
    for l in range(L):
        h = F.linear(h, params[2 * l], params[2 * l + 1])
        last = l == L - 1
        if not last or spec.relu_last:
            h = F.relu(h)
        if not last and spec.dropout_p > 0 and masks is not None:
            h = h * (masks[l].to(h.dtype) / (1.0 - spec.dropout_p))
    return h


def cross_entropy(logits, labels):
    """`nn.CrossEntropyLoss()` defaults: mean over the batch of -log_softmax(z)[y]
    (data_entities_vanilla.py:224,230; data_entities_vanilla_sisa.py:45,65,265,308;
    data_entities.py:34,76)."""
    return F.cross_entropy(logits, labels)


def draw_dropout_masks(batch, spec: BobSpec, generator=None):
    """The keep-masks torch's CPU `F.dropout` would draw, in layer order, from `generator`
    (None = global CPU generator): `empty_like(x).bernoulli_(1-p)`."""
    masks = []
    for l in range(spec.n_layers - 1):
        m = torch.empty(batch, spec.dims[l + 1]).bernoulli_(1.0 - spec.dropout_p, generator=generator)
        masks.append(m)
    return masks


# ----------------------------------------------------------------------------- optimizers
@dataclass
class SGDState:
    """`torch.optim.SGD(momentum=0.9)` as `_FunctionalSGD` applies it
    (data_entities_vanilla.py:37-42; torch optim/sgd.py single-tensor path)."""
    bufs: List[Optional[torch.Tensor]] = field(default_factory=list)


def sgd_momentum_(params, grads, state: SGDState, lr, momentum=0.9):
    if not state.bufs:
        state.bufs = [None] * len(params)
    for i, (p, g) in enumerate(zip(params, grads)):
        if g is None:
            continue
        if state.bufs[i] is None:
            state.bufs[i] = g.clone()                   # first step: buf <- g
        else:
            state.bufs[i].mul_(momentum).add_(g)        # buf <- mom*buf + g (dampening 0)
        p.add_(state.bufs[i], alpha=-lr)                # p <- p - lr*buf


@dataclass
class AdamState:
    """`torch.optim.Adam` single-tensor CPU path (torch optim/adam.py `_single_tensor_adam`),
    coupled L2 `weight_decay` (data_entities_vanilla_sisa.py:48,266; U-shape wd=0:
    data_entities.py:43-47)."""
    step: int = 0
    m: List[torch.Tensor] = field(default_factory=list)
    v: List[torch.Tensor] = field(default_factory=list)


def adam_(params, grads, state: AdamState, lr, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0):
    if not state.m:
        state.m = [torch.zeros_like(p) for p in params]
        state.v = [torch.zeros_like(p) for p in params]
    state.step += 1
    t = state.step
    bc1 = 1 - beta1 ** t                                  # python doubles, as in torch
    bc2 = 1 - beta2 ** t
    step_size = lr / bc1
    bc2_sqrt = bc2 ** 0.5
    for p, g, m, v in zip(params, grads, state.m, state.v):
        if g is None:
            continue
        if weight_decay != 0:
            g = g.add(p, alpha=weight_decay)
        m.lerp_(g, 1 - beta1)
        v.mul_(beta2).addcmul_(g, g, value=1 - beta2)
        denom = (v.sqrt() / bc2_sqrt).add_(eps)
        p.addcdiv_(m, denom, value=-step_size)


# ----------------------------------------------------------------------------- whole steps
def _leaf(ts):
    return [t.detach().clone().requires_grad_(True) for t in ts]


def vanilla_step(alice_params, bob_params, x, labels, masks, spec: BobSpec = MODEL2_SISA):
    """One `--vanilla` batch up to (not including) the optimizer step
    (data_entities_vanilla.py:70-73 on Alice, :229-231 on Bob).  Returns every tensor that
    crosses the cut plus all gradients."""
    a = _leaf(alice_params)
    b = _leaf(bob_params)
    act = alice_forward(x, a[0], a[1])
    act.retain_grad()
    logits = bob_forward(act, b, spec, masks)
    loss = cross_entropy(logits, labels)
    loss.backward()
    return dict(act=act.detach(), logits=logits.detach(), loss=loss.detach(),
                cut_grad=act.grad.detach(), bob_grads=[p.grad for p in b],
                alice_grads=[p.grad for p in a])


def sisa_server_step(bob_params, act, labels, masks, spec: BobSpec = MODEL2_SISA):
    """One batch of `bob.train_and_backward` in SISA up to the Adam step
    (data_entities_vanilla_sisa.py:305-311): input detached, so no cut-layer gradient."""
    b = _leaf(bob_params)
    logits = bob_forward(act.detach(), b, spec, masks)
    loss = cross_entropy(logits, labels)
    loss.backward()
    return dict(logits=logits.detach(), loss=loss.detach(), bob_grads=[p.grad for p in b])


def concat_server_step(bob_params, act_cat, labels_bk, masks, num_clients):
    """One `--sisa --concat` server batch up to the Adam step.  Forward/backward are the
    reference module's (`model2_sisa_concat`, models.py:66-82) on the feature-concatenated
    activations [B, 5408*K]; the loss is the one DEFINED BY THIS BUILD (the reference's loop crashes
    before it forms one, data_entities_sisa_concat.py:270-276): the output [B, 100*K] read as K heads
    of 100 logits, mean CE over the B*K (row, head) pairs with labels [B, K]."""
    spec = model2_sisa_concat_spec(num_clients)
    b = _leaf(bob_params)
    logits = bob_forward(act_cat.detach(), b, spec, masks)
    loss = cross_entropy(logits.reshape(-1, 100), labels_bk.reshape(-1))
    loss.backward()
    return dict(logits=logits.detach(), loss=loss.detach(), bob_grads=[p.grad for p in b])


def sisa_alice_step(alice_params, x, labels):
    """One batch of SISA `alice.train` up to the Adam step
    (data_entities_vanilla_sisa.py:62-68): CE taken directly over the 5408 cut activations."""
    a = _leaf(alice_params)
    act = alice_forward(x, a[0], a[1])
    act.retain_grad()
    loss = cross_entropy(act, labels)
    loss.backward()
    return dict(act=act.detach(), loss=loss.detach(), dact=act.grad.detach(),
                alice_grads=[p.grad for p in a])


def ushape_step(alice_params, tail_params, bob_params, x, labels, spec: BobSpec = MODEL2):
    """One default-mode (U-shape) batch up to the optimizer step (data_entities.py:70-79):
    Alice `model1` -> Bob `model2` -> Alice `model3` (fc 100->10, models.py:87-94) -> CE."""
    a = _leaf(alice_params)
    t = _leaf(tail_params)
    b = _leaf(bob_params)
    act = alice_forward(x, a[0], a[1], flatten=False)
    act.retain_grad()
    hb = bob_forward(act, b, spec, None)
    hb.retain_grad()
    out = F.linear(hb, t[0], t[1])
    loss = cross_entropy(out, labels)
    loss.backward()
    return dict(act=act.detach(), bob_out=hb.detach(), logits=out.detach(), loss=loss.detach(),
                bob_out_grad=hb.grad.detach(), cut_grad=act.grad.detach(),
                alice_grads=[p.grad for p in a], tail_grads=[p.grad for p in t],
                bob_grads=[p.grad for p in b])


def eval_counts(logits, labels, omit_label):
    """The six counters of `alice.eval_breakdown` (data_entities_vanilla.py:181-194)."""
    _, pred = torch.max(logits, 1)
    hit = pred == labels
    unl = labels == omit_label
    return [int(hit.sum()), int(labels.numel()), int((hit & unl).sum()), int(unl.sum()),
            int((hit & ~unl).sum()), int((~unl).sum())]
