"""Single-process restatement of who-calls-whom in the reference's entity code
(TEST INFRASTRUCTURE — see oracle/__init__.py).

The reference runs Bob and each Alice in separate processes joined by torch.distributed.rpc;
every process has its own torch CPU RNG stream.  Here each logical process owns an `RngStream`
that is swapped into the global generator whenever "its" code runs (model construction,
DataLoader iteration, dropout), which SURVEY.md A.3 showed is bit-exact with the RPC run.
tests/ pins this file against trajectories recorded from the reference's own RPC code
(`tests/golden/rpc_*.npz`, written by `gen_golden.py`).
"""
import contextlib
from copy import deepcopy

import torch
import torch.nn as nn

from . import torch_ref as R


class RngStream:
    """The torch CPU generator state of one logical process."""

    def __init__(self, seed):
        keep = torch.get_rng_state()
        torch.manual_seed(seed)
        self.state = torch.get_rng_state()
        torch.set_rng_state(keep)

    @contextlib.contextmanager
    def active(self):
        keep = torch.get_rng_state()
        torch.set_rng_state(self.state)
        try:
            yield
        finally:
            self.state = torch.get_rng_state()
            torch.set_rng_state(keep)


def _linear_init(n_in, n_out):
    m = nn.Linear(n_in, n_out)          # same init draws as the reference's nn.Linear
    return [m.weight.detach().clone(), m.bias.detach().clone()]


def _conv_init():
    m = nn.Conv2d(1, 32, 3)
    return [m.weight.detach().clone(), m.bias.detach().clone()]


def init_bob(spec: R.BobSpec):
    """Parameters in construction order fc1, fc2, (fc3) — models.py:36-37,49-53,69-73."""
    params = []
    for l in range(spec.n_layers):
        params += _linear_init(spec.dims[l], spec.dims[l + 1])
    return params


class Recorder:
    def __init__(self, keep_logits=True):
        self.steps = []
        self.keep_logits = keep_logits

    def add(self, **kw):
        if not self.keep_logits:
            kw.pop("logits", None)
        self.steps.append({k: (v.clone() if torch.is_tensor(v) else v) for k, v in kw.items()})


# ----------------------------------------------------------------------------- vanilla
def run_vanilla(train_loaders, lr, epochs, iterations, bob_seed=100, alice_seed=lambda r: 200 + r,
                recorder=None, spec=R.MODEL2_SISA, max_steps_per_request=None):
    """`split_nn.py:49-52` driving `data_entities_vanilla.py`: round-robin Alices, weight
    hand-off (:59-64), per-step Alice fwd -> Bob fwd/CE/bwd (:226-232) -> one SGD-momentum
    optimizer PER ALICE over [Bob params + her params] (:37-42,76)."""
    K = len(train_loaders)
    bob_rng = RngStream(bob_seed)
    with bob_rng.active():
        bob_params = init_bob(spec)
    alice_rng, alice_params, opt = {}, {}, {}
    for c in range(1, K + 1):
        alice_rng[c] = RngStream(alice_seed(c))
        with alice_rng[c].active():
            alice_params[c] = _conv_init()
        opt[c] = R.SGDState()
    last = None
    for _ in range(iterations):
        for c in range(1, K + 1):
            if last is not None:
                alice_params[c] = deepcopy(alice_params[last])          # give_weights (:79-80)
            for _ in range(epochs):
                with alice_rng[c].active():
                    it = iter(train_loaders[c])
                n = 0
                while True:
                    with alice_rng[c].active():
                        batch = next(it, None)
                    if batch is None:
                        break
                    x, y = batch
                    with bob_rng.active():
                        masks = R.draw_dropout_masks(x.shape[0], spec)
                    out = R.vanilla_step(alice_params[c], bob_params, x, y, masks, spec)
                    if recorder is not None:
                        recorder.add(client=c, x=x, labels=y, logits=out["logits"], loss=out["loss"])
                    R.sgd_momentum_(bob_params + alice_params[c], out["bob_grads"] + out["alice_grads"],
                                    opt[c], lr, 0.9)
                    n += 1
                    if max_steps_per_request and n >= max_steps_per_request:
                        break
            last = c
    return bob_params, alice_params


# ----------------------------------------------------------------------------- U-shape
def run_ushape(train_loaders, lr, epochs, iterations, bob_seed=100, alice_seed=lambda r: 200 + r,
               recorder=None, spec=R.MODEL2):
    """`data_entities.py:52-81`: Alice model1 -> Bob model2 -> Alice model3 -> CE on Alice;
    dist Adam per Alice over [model3, Bob, model1] (:43-47); hand-off of both Alice models."""
    K = len(train_loaders)
    bob_rng = RngStream(bob_seed)
    with bob_rng.active():
        bob_params = init_bob(spec)
    alice_rng, head, tail, opt = {}, {}, {}, {}
    for c in range(1, K + 1):
        alice_rng[c] = RngStream(alice_seed(c))
        with alice_rng[c].active():
            head[c] = _conv_init()                   # model1 first (:31)
            tail[c] = _linear_init(100, 10)          # then model3 (:32)
        opt[c] = R.AdamState()
    last = None
    for _ in range(iterations):
        for c in range(1, K + 1):
            if last is not None:
                head[c], tail[c] = deepcopy(head[last]), deepcopy(tail[last])
            for _ in range(epochs):
                with alice_rng[c].active():
                    it = iter(train_loaders[c])
                while True:
                    with alice_rng[c].active():
                        batch = next(it, None)
                    if batch is None:
                        break
                    x, y = batch
                    out = R.ushape_step(head[c], tail[c], bob_params, x, y, spec)
                    if recorder is not None:
                        recorder.add(client=c, x=x, labels=y, logits=out["logits"], loss=out["loss"])
                    R.adam_(tail[c] + bob_params + head[c],
                            out["tail_grads"] + out["bob_grads"] + out["alice_grads"], opt[c], lr)
            last = c
    return bob_params, head, tail


# ----------------------------------------------------------------------------- SISA
def sisa_alice_train(params, loader, rng, opt, lr, epochs, recorder=None, client=None):
    """SISA `alice.train` (data_entities_vanilla_sisa.py:55-70) and the retrain loop of
    `unlearn` (:157-168): local CE over the 5408 cut activations, Adam wd=1e-5."""
    for _ in range(epochs):
        with rng.active():
            it = iter(loader)
        while True:
            with rng.active():
                batch = next(it, None)
            if batch is None:
                break
            x, y = batch
            out = R.sisa_alice_step(params, x, y)
            if recorder is not None:
                recorder.add(client=client, phase="alice", loss=out["loss"])
            R.adam_(params, out["alice_grads"], opt, lr, weight_decay=1e-5)


def sisa_extract(params, loader, rng):
    """`alice.give_activation_and_labels` (:196-211): one pass over her loader, frozen."""
    acts, labels = [], []
    with rng.active():
        it = iter(loader)
    while True:
        with rng.active():
            batch = next(it, None)
        if batch is None:
            break
        x, y = batch
        acts.append(R.alice_forward(x, params[0], params[1]).detach())
        labels.append(y)
    return acts, labels


def sisa_server_epochs(bob_params, opt, cache_lists, bob_rng, lr, server_epochs, recorder=None,
                       spec=R.MODEL2_SISA):
    """`bob.train_and_backward` (:294-315): for each epoch, for each client in order, for each
    cached batch: zero_grad, fwd (train mode), CE, backward, Adam(wd=1e-5)."""
    for _ in range(server_epochs):
        for c, (acts, labels) in enumerate(cache_lists, start=1):
            for a, y in zip(acts, labels):
                with bob_rng.active():
                    masks = R.draw_dropout_masks(a.shape[0], spec)
                out = R.sisa_server_step(bob_params, a, y, masks, spec)
                if recorder is not None:
                    recorder.add(client=c, phase="bob", labels=y, logits=out["logits"], loss=out["loss"])
                R.adam_(bob_params, out["bob_grads"], opt, lr, weight_decay=1e-5)


def unlearn_filter(loader, omit_label, batch_size):
    """`alice.unlearn` data filter (:149-152): iterate the (shuffling) train loader once, drop
    label==omit, re-batch in that order without shuffling."""
    kept = [(xi, yi) for xb, yb in loader for xi, yi in zip(xb, yb) if yi != omit_label]
    return torch.utils.data.DataLoader(kept, batch_size=batch_size)


def run_sisa(train_loaders, lr, epochs, server_epochs, batch_size, omit_label=9, unlearn_client=1,
             bob_seed=100, alice_seed=lambda r: 200 + r, recorder=None, do_unlearn=True):
    """`split_nn.py:74-110` driving `data_entities_vanilla_sisa.py`."""
    spec = R.MODEL2_SISA
    K = len(train_loaders)
    bob_rng = RngStream(bob_seed)
    with bob_rng.active():
        bob_params = init_bob(spec)
    bob_opt = R.AdamState()
    a_rng, a_params, a_opt = {}, {}, {}
    for c in range(1, K + 1):
        a_rng[c] = RngStream(alice_seed(c))
        with a_rng[c].active():
            a_params[c] = _conv_init()
        a_opt[c] = R.AdamState()
    for c in range(1, K + 1):                                    # train_request_parallel
        sisa_alice_train(a_params[c], train_loaders[c], a_rng[c], a_opt[c], lr, epochs, recorder, c)
    cache = {c: sisa_extract(a_params[c], train_loaders[c], a_rng[c]) for c in range(1, K + 1)}
    sisa_server_epochs(bob_params, bob_opt, [cache[c] for c in range(1, K + 1)], bob_rng, lr,
                       server_epochs, recorder)
    if do_unlearn:
        c = unlearn_client
        a_opt[c] = R.AdamState()                                 # fresh Adam; reset_model no-op
        with a_rng[c].active():
            unl_loader = unlearn_filter(train_loaders[c], omit_label, batch_size)
        sisa_alice_train(a_params[c], unl_loader, a_rng[c], a_opt[c], lr, epochs, recorder, c)
        cache[c] = sisa_extract(a_params[c], unl_loader, a_rng[c])
        sisa_server_epochs(bob_params, bob_opt, [cache[k] for k in range(1, K + 1)], bob_rng, lr,
                           server_epochs, recorder)
    return bob_params, a_params
