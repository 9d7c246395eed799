"""Parity of the CUDA path (through the C ABI) against the oracle and the committed fixtures.

Tolerance: BASELINE.json asks for logits/losses within 1e-4 relative per step; measured
norm-wise, ||cuda - ref||_F / ||ref||_F <= 1e-4 (per-element relative error is
ill-posed, fp32 itself misses it near zero).  Integer results (eval counters) are exact.
"""
import os

import numpy as np
import pytest
import torch

from oracle import numpy_ref as NR
from oracle import torch_ref as R
from tests.helpers import check_adam_weights, check_grad_through_relu, check_summary, load_golden, rel_err, seeded_batch, seeded_masks, seeded_weights

pytestmark = pytest.mark.gpu
TOL = 1e-4
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

A_SHAPES = [(32, 1, 3, 3), (32,)]
B_SHAPES = [(5000, 5408), (5000,), (1000, 5000), (1000,), (100, 1000), (100,)]
B_NAMES = ["fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias", "fc3.weight", "fc3.bias"]


@pytest.fixture(scope="module")
def eng():
    from splitlearning_b200 import engine
    assert torch.cuda.is_available()
    return engine


def cu(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def T(arrs):
    return [torch.from_numpy(a.copy()) for a in arrs]


# --------------------------------------------------------------------------- stateless operators
@pytest.mark.parametrize("B", [1, 5, 16, 33])
def test_conv_pool_forward(eng, B):
    w, b = seeded_weights(A_SHAPES, 11)
    x, _ = seeded_batch(B, 900 + B)
    ref = R.alice_forward(torch.from_numpy(x), torch.from_numpy(w), torch.from_numpy(b))
    got = eng.conv_pool_forward(cu(x), cu(w), cu(b)).cpu()
    assert rel_err(got, ref) <= 1e-6
    ref_np, _, _ = NR.conv_pool_forward(x, w, b)
    assert rel_err(got, ref_np) <= 1e-6


@pytest.mark.parametrize("B", [1, 16, 40])
def test_conv_pool_backward(eng, B):
    w, b = seeded_weights(A_SHAPES, 11)
    x, _ = seeded_batch(B, 910 + B)
    rng = np.random.RandomState(5)
    dact = (rng.randn(B, 5408) * 1e-3).astype(np.float32)
    wt, bt = torch.from_numpy(w).requires_grad_(True), torch.from_numpy(b).requires_grad_(True)
    R.alice_forward(torch.from_numpy(x), wt, bt).backward(torch.from_numpy(dact))
    dw, db = eng.conv_pool_backward(cu(x), cu(w), cu(b), cu(dact))
    assert rel_err(dw.cpu(), wt.grad) <= 1e-5
    assert rel_err(db.cpu(), bt.grad) <= 1e-5


@pytest.mark.parametrize("B,K,N,relu", [(16, 5408, 5000, True), (16, 5000, 1000, True), (16, 1000, 100, False),
                                        (3, 5408, 1000, True), (16, 100, 10, False), (40, 1000, 100, True),
                                        (100, 512, 104, False), (7, 52, 12, True),
                                        # narrow layers with a long reduction: fewer row groups than warps
                                        # (a group cut into 3+ pieces used to lose partial sums)
                                        (16, 40000, 10, False), (16, 65536, 40, True), (5, 39940, 24, False),
                                        (16, 2112, 4, False)])
def test_linear_forward(eng, B, K, N, relu):
    rng = np.random.RandomState(B + K + N)
    x = (rng.rand(B, K) * 50 * (rng.rand(B, K) < 0.6)).astype(np.float32)
    W = (rng.uniform(-1, 1, (N, K)) / np.sqrt(K)).astype(np.float32)
    bias = rng.uniform(-1, 1, N).astype(np.float32)
    mask = (rng.rand(B, N) < 0.5).astype(np.uint8)
    ref = torch.nn.functional.linear(torch.from_numpy(x), torch.from_numpy(W), torch.from_numpy(bias))
    if relu:
        ref = torch.relu(ref)
    got = eng.linear_forward(cu(x), cu(W), cu(bias), relu=relu).cpu()
    assert rel_err(got, ref) <= 1e-5
    got = eng.linear_forward(cu(x), cu(W), cu(bias), relu=relu, keep_mask=cu(mask), mask_scale=2.0).cpu()
    assert rel_err(got, ref * torch.from_numpy(mask) * 2.0) <= 1e-5


# row forms: registers (<= 128 classes), one CTA per row (129..6144), streaming warp per row (beyond)
@pytest.mark.parametrize("B,Cn", [(16, 100), (1, 10), (16, 5408), (37, 100), (300, 10), (16, 128), (5, 129), (3, 6144),
                                  (2, 6145), (9, 1000)])
def test_cross_entropy(eng, B, Cn):
    rng = np.random.RandomState(B * 7 + Cn)
    z = (rng.randn(B, Cn) * 8).astype(np.float32)
    y = rng.randint(0, min(Cn, 10), size=B).astype(np.int64)
    zt = torch.from_numpy(z).requires_grad_(True)
    ref = torch.nn.functional.cross_entropy(zt, torch.from_numpy(y))
    ref.backward()
    loss, dz = eng.cross_entropy(cu(z), cu(y))
    # fp64 truth within 1e-5; torch's own fp32 result within the 1e-4 contract (when one class
    # dominates, torch-fp32 itself is ~1e-5 off the fp64 value)
    ref64 = torch.nn.functional.cross_entropy(torch.from_numpy(z).double(), torch.from_numpy(y)).item()
    assert abs(loss.item() - ref64) <= 1e-5 * abs(ref64)
    assert abs(loss.item() - ref.item()) <= 1e-4 * abs(ref.item())
    assert rel_err(dz.cpu(), zt.grad) <= 1e-5
    l2, g2 = NR.cross_entropy(z, y)
    assert abs(loss.item() - float(l2)) <= 1e-4 * abs(float(l2))
    assert rel_err(dz.cpu(), g2) <= 1e-4     # numpy's fp32 exp rounds differently where p-1 cancels
    # ties: the first maximum is the one left out of the log1p sum (torch's argmax convention)
    zt2 = z.copy()
    zt2[0, :] = -3.0
    zt2[0, [Cn // 3, Cn - 1]] = 5.0
    loss2, _ = eng.cross_entropy(cu(zt2), cu(y))
    ref2 = torch.nn.functional.cross_entropy(torch.from_numpy(zt2).double(), torch.from_numpy(y)).item()
    assert abs(loss2.item() - ref2) <= 1e-5 * abs(ref2)


def test_eval_counts_exact(eng):
    rng = np.random.RandomState(3)
    z = rng.randn(301, 100).astype(np.float32)
    z[5, 7] = z[5, 3] = 9.0          # tie: first index must win
    y = rng.randint(0, 10, size=301).astype(np.int64)
    y[5] = 3
    ref = R.eval_counts(torch.from_numpy(z), torch.from_numpy(y), 9)
    c = eng.eval_counts(cu(z), cu(y), 9)
    c = eng.eval_counts(cu(z), cu(y), 9, counters=c)       # accumulates
    assert c.cpu().tolist() == [2 * v for v in ref]


# --------------------------------------------------------------------------- whole steps
def _bob(eng, params, **kw):
    gp = [cu(p.numpy()) for p in params]
    e = eng.BobEngine([5408, 5000, 1000, 100], gp[0::2], gp[1::2], **kw)
    return e, gp


@pytest.mark.parametrize("B", [16, 7])
def test_vanilla_step_vs_oracle_and_golden(eng, B):
    """Teacher-forced: identical weights / batch / masks; two consecutive steps (the second
    exercises the momentum buffers).  Checked against the oracle and, for B=16, the fixture
    generated from the reference's own models.py."""
    g = load_golden("modules_vanilla.npz") if B == 16 else None
    a, b = T(seeded_weights(A_SHAPES, 11)), T(seeded_weights(B_SHAPES, 12))
    bob, gp = _bob(eng, b, optimizer=eng.SN_OPT_SGD_MOMENTUM, lr=1e-3, momentum=0.9)
    gw, gb = cu(a[0].numpy()), cu(a[1].numpy())
    alice = eng.AliceEngine(gw, gb, optimizer=eng.SN_OPT_SGD_MOMENTUM, lr=1e-3, momentum=0.9)
    opt = R.SGDState()
    for step in range(2):
        x, y = seeded_batch(B, 100 + step)
        masks = seeded_masks(B, (5000, 1000), 200 + step)
        ref = R.vanilla_step(a, b, torch.from_numpy(x), torch.from_numpy(y), T(masks))
        R.sgd_momentum_(b + a, ref["bob_grads"] + ref["alice_grads"], opt, 1e-3, 0.9)
        xg, yg = cu(x), cu(y)
        act = alice.forward(xg)
        loss, cut, logits = bob.train_step(act, yg, masks=[cu(m) for m in masks], want_logits=True)
        alice.backward_update(xg, cut)
        torch.cuda.synchronize()
        assert rel_err(act.cpu(), ref["act"]) <= TOL
        assert rel_err(logits.cpu(), ref["logits"]) <= TOL
        assert abs(loss.item() - ref["loss"].item()) <= TOL * abs(ref["loss"].item())
        assert rel_err(cut.cpu(), ref["cut_grad"]) <= TOL
        for n, p_ref, p in zip(B_NAMES, b, gp):
            assert rel_err(p.cpu(), p_ref) <= TOL, n
        assert rel_err(gw.cpu(), a[0]) <= TOL and rel_err(gb.cpu(), a[1]) <= TOL
        if g is not None:
            check_summary(g, f"s{step}_logits", logits, TOL)
            check_summary(g, f"s{step}_cut_grad", cut, TOL)
            assert abs(loss.item() - float(g[f"s{step}_loss"])) <= TOL * abs(float(g[f"s{step}_loss"]))
            for n, p in zip(B_NAMES, gp):
                check_summary(g, f"s{step}_new_{n}", p, TOL)
            check_summary(g, f"s{step}_new_alice.conv_layers.0.weight", gw, TOL)
            check_summary(g, f"s{step}_new_alice.conv_layers.0.bias", gb, TOL)
    # the update itself (not just W ~= W_old) must match: compare the applied delta
    w0 = seeded_weights(B_SHAPES, 12)
    for n, p_ref, p, p0 in zip(B_NAMES, b, gp, w0):
        assert rel_err(p.cpu().numpy() - p0, p_ref.numpy() - p0) <= 2e-3, n


@pytest.mark.parametrize("B", [16, 9])
def test_single_call_vanilla_step_equals_the_three_calls(eng, B):
    """sn_vanilla_step (one call: Alice forward, Bob's train step, Alice backward) against the three
    separate calls: after three steps every weight on both sides and every loss must be EQUAL."""
    a0, b0 = seeded_weights(A_SHAPES, 11), seeded_weights(B_SHAPES, 12)
    kw = dict(optimizer=eng.SN_OPT_SGD_MOMENTUM, lr=1e-3, momentum=0.9)
    finals = []
    for single in (False, True):
        gw, gb = cu(a0[0]), cu(a0[1])
        bob, gp = _bob(eng, T(b0), **kw)
        alice = eng.AliceEngine(gw, gb, **kw)
        act, cut, loss = torch.empty(B, 5408, device="cuda"), torch.empty(B, 5408, device="cuda"), torch.zeros((), device="cuda")
        losses = []
        for step in range(3):
            x, y = seeded_batch(B, 100 + step)
            xg, yg = cu(x), cu(y)
            if single:
                eng.vanilla_step(alice, bob, xg, yg, 0, 5, step + 1, act, cut, loss)
            else:
                alice.forward(xg, out=act)
                bob.train_step(act, yg, slot=0, masks=None, seed=5, offset=step + 1, loss_out=loss, cut_out=cut)
                alice.backward_update(xg, cut)
            losses.append(loss.clone())
        torch.cuda.synchronize()
        finals.append([gw.clone(), gb.clone()] + [p.clone() for p in gp] + losses)
    for u, v in zip(*finals):
        assert torch.equal(u, v)


@pytest.mark.parametrize("B", [16, 11])
def test_host_input_step_equals_the_device_input_step(eng, B):
    """sn_vanilla_step_host: the batch stays in pinned host memory, the conv kernel reads it over PCIe in its
    prologue and leaves device copies for the backward and the loss.  Three steps (alternating the two
    device-copy pairs) must EQUAL sn_vanilla_step on a device-resident batch: weights, losses, and the
    copies themselves."""
    a0, b0 = seeded_weights(A_SHAPES, 11), seeded_weights(B_SHAPES, 12)
    kw = dict(optimizer=eng.SN_OPT_SGD_MOMENTUM, lr=1e-3, momentum=0.9)
    finals = []
    for host in (False, True):
        gw, gb = cu(a0[0]), cu(a0[1])
        bob, gp = _bob(eng, T(b0), **kw)
        alice = eng.AliceEngine(gw, gb, **kw)
        act, cut = torch.empty(B, 5408, device="cuda"), torch.empty(B, 5408, device="cuda")
        loss = torch.zeros(3, dtype=torch.float32).pin_memory() if host else torch.zeros(3, device="cuda")
        xd = [torch.empty(B, 1, 28, 28, device="cuda") for _ in range(2)]
        yd = [torch.empty(B, dtype=torch.int64, device="cuda") for _ in range(2)]
        for step in range(3):
            x, y = seeded_batch(B, 100 + step)
            if host:
                hx, hy = torch.from_numpy(x).pin_memory(), torch.from_numpy(y).pin_memory()
                eng.vanilla_step_host(alice, bob, hx, hy, 0, 5, step + 1, xd[step & 1], yd[step & 1], act, cut, loss[step])
                torch.cuda.synchronize()                      # hx / hy go out of scope below
                assert torch.equal(xd[step & 1].cpu(), torch.from_numpy(x)) and torch.equal(yd[step & 1].cpu(), torch.from_numpy(y))
            else:
                eng.vanilla_step(alice, bob, cu(x), cu(y), 0, 5, step + 1, act, cut, loss[step])
        torch.cuda.synchronize()
        finals.append([gw.clone(), gb.clone()] + [p.clone() for p in gp] + [loss.clone().cuda()])
    for u, v in zip(*finals):
        assert torch.equal(u, v)


@pytest.mark.parametrize("opt", ["sgd", "adam"])
def test_persistent_tail_kernel_is_bit_identical_to_the_launch_chain(eng, opt):
    """sn_set_persistent_tail(1): the middle of the step (fc1 epilogue .. fc2 backward) runs as one
    persistent kernel with grid barriers instead of seven launches.  Same kernel bodies, same order of
    every sum: logits, loss, cut gradient and all updated weights must be EQUAL bit for bit, over
    three steps, with given masks and with in-kernel Philox dropout, full and ragged batches.  (Both
    forms are separately checked against the oracle by the tests around this one.)"""
    from splitlearning_b200 import _lib
    lib = _lib.load()
    kw = (dict(optimizer=eng.SN_OPT_SGD_MOMENTUM, lr=1e-3, momentum=0.9) if opt == "sgd"
          else dict(optimizer=eng.SN_OPT_ADAM, lr=1e-3, weight_decay=1e-5))
    a = T(seeded_weights(A_SHAPES, 11))
    results = []
    was = lib.sn_get_persistent_tail()
    try:
        for tail in (0, 1):
            _lib.check(lib.sn_set_persistent_tail(tail))
            assert lib.sn_get_persistent_tail() == tail
            bob, gp = _bob(eng, T(seeded_weights(B_SHAPES, 12)), **kw)
            out = []
            n0 = lib.sn_launch_count()
            for step, B in enumerate((16, 16, 5)):
                x, y = seeded_batch(B, 100 + step)
                act = cu(R.alice_forward(torch.from_numpy(x), a[0], a[1]).numpy())
                masks = [cu(m) for m in seeded_masks(B, (5000, 1000), 200 + step)] if step != 1 else None
                loss, cut, logits = bob.train_step(act, cu(y), masks=masks, seed=3, offset=step + 1, want_logits=True,
                                                   want_cut_grad=opt == "sgd")
                out += [loss.clone(), logits.clone()] + ([cut.clone()] if cut is not None else [])
            torch.cuda.synchronize()
            out += [p.clone() for p in gp]
            results.append((out, lib.sn_launch_count() - n0))
    finally:
        _lib.check(lib.sn_set_persistent_tail(was))
    (chain, n_chain), (tail, n_tail) = results
    assert n_tail < n_chain                               # six launches per step fewer
    assert len(chain) == len(tail)
    for u, v in zip(chain, tail):
        assert torch.equal(u, v)


def test_sisa_server_steps_vs_oracle_and_golden(eng):
    g = load_golden("modules_sisa_server.npz")
    a, b = T(seeded_weights(A_SHAPES, 11)), T(seeded_weights(B_SHAPES, 12))
    bob, gp = _bob(eng, b, optimizer=eng.SN_OPT_ADAM, lr=1e-3, weight_decay=1e-5)
    opt = R.AdamState()
    for step in range(2):
        x, y = seeded_batch(16, 300 + step)
        masks = seeded_masks(16, (5000, 1000), 400 + step)
        act = R.alice_forward(torch.from_numpy(x), a[0], a[1])
        ref = R.sisa_server_step(b, act, torch.from_numpy(y), T(masks))
        R.adam_(b, ref["bob_grads"], opt, 1e-3, weight_decay=1e-5)
        loss, cut, logits = bob.train_step(cu(act.numpy()), cu(y), masks=[cu(m) for m in masks],
                                           want_cut_grad=False, want_logits=True)
        assert cut is None
        assert rel_err(logits.cpu(), ref["logits"]) <= TOL
        assert abs(loss.item() - ref["loss"].item()) <= TOL * abs(ref["loss"].item())
        check_summary(g, f"s{step}_logits", logits, TOL)
        for n, p_ref, p in zip(B_NAMES, b, gp):
            check_adam_weights(p.cpu(), p_ref, n)
            check_summary(g, f"s{step}_new_{n}", p, 1e-3)
        assert bob.get_opt_step(0) == opt.step
        for i, (p_ref, p) in enumerate(zip(b, gp)):      # teacher forcing for the next step
            p.copy_(p_ref)
            bob.opt_state(0, 0, i).copy_(opt.m[i].reshape(-1))
            bob.opt_state(0, 1, i).copy_(opt.v[i].reshape(-1))
    # eval-mode inference with the updated weights
    x, _ = seeded_batch(16, 500)
    act = R.alice_forward(torch.from_numpy(x), a[0], a[1])
    out = bob.forward(cu(act.numpy()), train_mode=False)
    check_summary(g, "eval_logits", out, TOL)


def test_concat7_server_steps_vs_oracle_and_golden(eng):
    """BASELINE.json config 4: the --sisa --concat server at K = 7 clients (fc1 5000 x 37856 = 757 MB,
    fc3 1000 -> 700, seven CE heads), two teacher-forced Adam(wd 1e-5) steps against the oracle and
    against the fixture written from the reference's model2_sisa_concat module."""
    from oracle.gen_golden import concat7_inputs
    K = 7
    g = load_golden("modules_concat7.npz")
    w, xs, ys, masks = concat7_inputs()
    b = T(w)
    gp = [cu(p.numpy()) for p in b]
    bob = eng.BobEngine([5408 * K, 5000, 1000, 100 * K], gp[0::2], gp[1::2], optimizer=eng.SN_OPT_ADAM, lr=1e-3,
                        weight_decay=1e-5)
    bob.set_ce_heads(K)
    opt = R.AdamState()
    for step in range(2):
        ref = R.concat_server_step(b, torch.from_numpy(xs[step]), torch.from_numpy(ys[step]), T(masks[step]), K)
        R.adam_(b, ref["bob_grads"], opt, 1e-3, weight_decay=1e-5)
        loss, cut, logits = bob.train_step(cu(xs[step]), cu(ys[step]), masks=[cu(m) for m in masks[step]],
                                           want_cut_grad=False, want_logits=True)
        assert cut is None and logits.shape == (16, 100 * K)
        assert rel_err(logits.cpu(), ref["logits"]) <= TOL
        assert abs(loss.item() - ref["loss"].item()) <= TOL * abs(ref["loss"].item())
        check_summary(g, f"s{step}_logits", logits, TOL)
        assert abs(loss.item() - float(g[f"s{step}_loss"])) <= TOL * abs(float(g[f"s{step}_loss"]))
        for n, p_ref, p in zip(B_NAMES, b, gp):
            check_adam_weights(p.cpu(), p_ref, n)
            check_summary(g, f"s{step}_new_{n}", p, 1e-3)
        for i, (p_ref, p) in enumerate(zip(b, gp)):      # teacher forcing for the next step
            p.copy_(p_ref)
            bob.opt_state(0, 0, i).copy_(opt.m[i].reshape(-1))
            bob.opt_state(0, 1, i).copy_(opt.v[i].reshape(-1))
    # the server-epoch form (first layer's next forward fused into the Adam traversal) on the same two
    # batches must give the step-wise losses
    for i, p in enumerate(T(w)):
        gp[i].copy_(p)
    bob.reset_opt_slot(0)
    acts = cu(np.concatenate(xs, 0))
    labels = cu(np.concatenate(ys, 0))
    mm = [cu(np.concatenate([masks[0][l], masks[1][l]], 0)) for l in range(2)]
    losses = torch.empty(2, device="cuda")
    bob.server_epoch(acts, labels, [0, 16, 32], slot=0, masks=mm, loss_per_step=losses)
    torch.cuda.synchronize()
    b2 = T(w)
    opt2 = R.AdamState()
    for step in range(2):
        ref = R.concat_server_step(b2, torch.from_numpy(xs[step]), torch.from_numpy(ys[step]), T(masks[step]), K)
        R.adam_(b2, ref["bob_grads"], opt2, 1e-3, weight_decay=1e-5)
        assert abs(losses[step].item() - ref["loss"].item()) <= (TOL if step == 0 else 2e-3) * abs(ref["loss"].item())


def test_sisa_alice_local_steps(eng):
    g = load_golden("modules_sisa_alice.npz")
    a = T(seeded_weights(A_SHAPES, 11))
    gw, gb = cu(a[0].numpy()), cu(a[1].numpy())
    alice = eng.AliceEngine(gw, gb, optimizer=eng.SN_OPT_ADAM, lr=1e-3, weight_decay=1e-5)
    opt = R.AdamState()
    for step in range(3):
        x, y = seeded_batch(16, 600 + step)
        ref = R.sisa_alice_step(a, torch.from_numpy(x), torch.from_numpy(y))
        R.adam_(a, ref["alice_grads"], opt, 1e-3, weight_decay=1e-5)
        loss = alice.local_step(cu(x), cu(y))
        assert abs(loss.item() - ref["loss"].item()) <= TOL * abs(ref["loss"].item())
        assert abs(loss.item() - float(g[f"s{step}_loss"])) <= TOL * abs(float(g[f"s{step}_loss"]))
        assert rel_err(gw.cpu(), a[0]) <= TOL and rel_err(gb.cpu(), a[1]) <= TOL
        check_summary(g, f"s{step}_new_conv_layers.0.weight", gw, TOL)
        check_summary(g, f"s{step}_new_conv_layers.0.bias", gb, TOL)


def test_ushape_steps(eng):
    g = load_golden("modules_ushape.npz")
    a = T(seeded_weights(A_SHAPES, 11))
    b = T(seeded_weights([(1000, 5408), (1000,), (100, 1000), (100,)], 13))
    t = T(seeded_weights([(10, 100), (10,)], 14))
    gp = [cu(p.numpy()) for p in b]
    gt = [cu(p.numpy()) for p in t]
    gw, gb = cu(a[0].numpy()), cu(a[1].numpy())
    adam = dict(optimizer=eng.SN_OPT_ADAM, lr=1e-3)
    bob = eng.BobEngine([5408, 1000, 100], gp[0::2], gp[1::2], relu_last=True, dropout_p=0.0, **adam)
    tail = eng.BobEngine([100, 10], [gt[0]], [gt[1]], dropout_p=0.0, **adam)
    alice = eng.AliceEngine(gw, gb, **adam)
    opt = R.AdamState()
    for step in range(2):
        x, y = seeded_batch(16, 700 + step)
        ref = R.ushape_step(a, t, b, torch.from_numpy(x), torch.from_numpy(y))
        R.adam_(t + b + a, ref["tail_grads"] + ref["bob_grads"] + ref["alice_grads"], opt, 1e-3)
        xg, yg = cu(x), cu(y)
        act = alice.forward(xg)
        hb = bob.forward(act, train_mode=True)
        loss, dhb, logits = tail.train_step(hb, yg, want_logits=True)
        cut = bob.backward_update(act, dhb)
        alice.backward_update(xg, cut)
        assert rel_err(hb.cpu(), ref["bob_out"]) <= TOL
        assert rel_err(logits.cpu(), ref["logits"]) <= TOL
        assert rel_err(dhb.cpu(), ref["bob_out_grad"]) <= TOL
        assert rel_err(cut.cpu(), ref["cut_grad"].flatten(1)) <= TOL
        check_summary(g, f"s{step}_bob_out", hb, TOL)
        check_summary(g, f"s{step}_cut_grad", cut, TOL)
        for n, p in zip(["fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias"], gp):
            check_summary(g, f"s{step}_new_m2.{n}", p, TOL)
        check_summary(g, f"s{step}_new_m3.fc3.weight", gt[0], TOL)
        check_summary(g, f"s{step}_new_m1.conv1.weight", gw, TOL)


@pytest.mark.parametrize("mode,width", [(0, 256), (2, 256), (0, 1024)])
def test_general_batch_path_matches_oracle(eng, mode, width):
    """B > 16 takes the general-batch kernels - 3xTF32 tcgen05 GEMMs by default (mode 0), the fp32
    SIMT fallback in mode 2; same oracle, same 1e-4 bound, small widths.  At these sizes the forward
    GEMMs are split along K (few output tiles); width 1024 also splits the input-gradient GEMM."""
    eng.set_matmul_precision(mode)
    try:
        _general_batch_body(eng, mode, width)
    finally:
        eng.set_matmul_precision(0)


def test_general_batch_path_on_cta_pairs():
    """The same oracle check with every unsplit 3xTF32 forward / input-gradient GEMM forced onto CTA
    pairs (tcgen05 cta_group::2, 256 x 256 per pair): SN_X3_PAIR is read once per process, so this
    runs in a child."""
    import subprocess
    import sys
    env = dict(os.environ, SN_X3_PAIR="1")
    code = ("import sys; sys.path.insert(0, %r); sys.path.insert(0, %r)\n"
            "import tests.test_gpu_kernels as t\n"
            "from splitlearning_b200 import engine as E\n"
            "t._general_batch_body(E, 0, 256); t._general_batch_body(E, 0, 1024); print('PAIR_OK')\n") % (ROOT, os.path.join(ROOT, "tests"))
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "PAIR_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


def test_large_batch_step_on_one_pass_cta_pairs():
    """The one-pass TF32 step (forward, dX, dW + optimizer in the epilogue) with every unsplit GEMM forced
    onto CTA pairs (SS operands, both CTAs' TMA bytes counted on the leader's barrier): same oracle
    bounds as the single-CTA form.  SN_TF32_PAIR is read once per process, so this runs in a child."""
    import subprocess
    import sys
    env = dict(os.environ, SN_TF32_PAIR="1")
    code = ("import sys; sys.path.insert(0, %r)\n"
            "import tests.test_gpu_kernels as t\n"
            "from splitlearning_b200 import engine as E\n"
            "t.test_tcgen05_large_batch_train_step(E, 'sgd'); t.test_tcgen05_large_batch_train_step(E, 'adam'); print('PAIR_OK')\n") % ROOT
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "PAIR_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]


def _general_batch_body(eng, mode, width):
    rng = np.random.RandomState(77)
    dims = [5408, width, 52, 100]
    shapes = [(width, 5408), (width,), (52, width), (52,), (100, 52), (100,)]
    b = T(seeded_weights(shapes, 21))
    spec = R.BobSpec(tuple(dims), False, 0.5)
    gp = [cu(p.numpy()) for p in b]
    bob = eng.BobEngine(dims, gp[0::2], gp[1::2], optimizer=eng.SN_OPT_ADAM, lr=1e-3, weight_decay=1e-5, max_batch=80)
    opt = R.AdamState()
    for step, B in enumerate([80, 48]):
        act = (rng.rand(B, 5408) * 40 * (rng.rand(B, 5408) < 0.6)).astype(np.float32)
        y = rng.randint(0, 10, size=B).astype(np.int64)
        masks = seeded_masks(B, (width, 52), 50 + step)
        bl = [p.detach().clone().requires_grad_(True) for p in b]
        at = torch.from_numpy(act).requires_grad_(True)
        logits_ref = R.bob_forward(at, bl, spec, T(masks))
        loss_ref = R.cross_entropy(logits_ref, torch.from_numpy(y))
        loss_ref.backward()
        R.adam_(b, [p.grad for p in bl], opt, 1e-3, weight_decay=1e-5)
        loss, cut, logits = bob.train_step(cu(act), cu(y), masks=[cu(m) for m in masks], want_logits=True)
        assert rel_err(logits.cpu(), logits_ref.detach()) <= TOL
        assert abs(loss.item() - loss_ref.item()) <= TOL * abs(loss_ref.item())
        if mode == 2 or B <= 64:     # fp32 forward: no ReLU gate can flip
            assert rel_err(cut.cpu(), at.grad) <= TOL
        else:                        # 3xTF32 forward (4e-5 at K=5408): see check_grad_through_relu
            check_grad_through_relu(cut.cpu(), at.grad, 4e-5, "cut gradient")
        for p_ref, p in zip(b, gp):
            if mode == 2 or B <= 64:
                check_adam_weights(p.cpu(), p_ref)
            else:                    # gradients carry the gate flips above; Adam turns them into +-lr
                check_adam_weights(p.cpu(), p_ref, frob=5e-3, frac_tol=2e-2)
        # teacher forcing: the next step starts from the oracle's weights and Adam state on both sides
        assert bob.get_opt_step(0) == opt.step
        for i, (p_ref, p) in enumerate(zip(b, gp)):
            p.copy_(p_ref)
            bob.opt_state(0, 0, i).copy_(opt.m[i].reshape(-1))
            bob.opt_state(0, 1, i).copy_(opt.v[i].reshape(-1))


def test_philox_dropout_statistics_and_determinism(eng):
    b = T(seeded_weights(B_SHAPES, 12))
    bob, gp = _bob(eng, b, optimizer=eng.SN_OPT_SGD_MOMENTUM, lr=0.0, momentum=0.0)
    x, _ = seeded_batch(16, 1)
    a = T(seeded_weights(A_SHAPES, 11))
    act = cu(R.alice_forward(torch.from_numpy(x), a[0], a[1]).numpy())
    o1 = bob.forward(act, train_mode=True, seed=5, offset=1).cpu()
    o2 = bob.forward(act, train_mode=True, seed=5, offset=1).cpu()
    o3 = bob.forward(act, train_mode=True, seed=5, offset=2).cpu()
    assert torch.equal(o1, o2) and not torch.equal(o1, o3)
    # one hidden layer, linear head of ones: output = sum of kept, doubled activations
    ones_w, ones_b = torch.ones(1, 5000, device="cuda"), torch.zeros(1, device="cuda")
    net = eng.BobEngine([5408, 5000, 1], [gp[0], ones_w], [gp[1], ones_b], dropout_p=0.5, optimizer=eng.SN_OPT_NONE)
    full = net.forward(act, train_mode=False).cpu()
    drops = torch.stack([net.forward(act, train_mode=True, seed=3, offset=i).cpu() for i in range(8)])
    ratio = (drops.mean(0) / full).mean().item()          # E[keep*2] = 1
    assert 0.9 < ratio < 1.1, ratio


# --------------------------------------------------------------------------- tcgen05 TF32 path
def _tf32_trunc(a):
    return (a.view(np.uint32) & np.uint32(0xFFFFE000)).view(np.float32)


@pytest.mark.parametrize("B,K,N,relu", [(128, 32, 128, False), (256, 512, 256, True), (200, 1000, 100, True),
                                        (65, 100, 12, False), (600, 5408, 1000, True),
                                        (2500, 300, 2100, True)])      # 20 x 9 tiles of 128x256: the wide-tile kernel
def test_tcgen05_tf32_linear_forward(eng, B, K, N, relu):
    """Large-batch perf mode: tcgen05.mma kind::tf32 reads the top 19 bits of each fp32 operand and
    accumulates in fp32.  Checked tightly against an emulation of exactly that (truncated operands,
    fp64 accumulate) and loosely (TF32's own error) against the fp32 result."""
    rng = np.random.RandomState(B + K + N)
    x = (rng.rand(B, K) * 4 - 1).astype(np.float32)
    W = (rng.uniform(-1, 1, (N, K)) / np.sqrt(K)).astype(np.float32)
    bias = rng.uniform(-1, 1, N).astype(np.float32)
    mask = (rng.rand(B, N) < 0.5).astype(np.uint8)
    eng.set_matmul_precision(1)
    try:
        got = eng.linear_forward(cu(x), cu(W), cu(bias), relu=relu, keep_mask=cu(mask), mask_scale=2.0).cpu().numpy()
    finally:
        eng.set_matmul_precision(0)
    emu = _tf32_trunc(x).astype(np.float64) @ _tf32_trunc(W).astype(np.float64).T + bias
    ref = x.astype(np.float64) @ W.astype(np.float64).T + bias
    if relu:
        emu, ref = np.maximum(emu, 0), np.maximum(ref, 0)
    emu, ref = emu * mask * 2.0, ref * mask * 2.0
    assert rel_err(got, emu) <= 5e-5          # the kernel computes what TF32 tensor cores define
    assert rel_err(got, ref) <= 3e-3          # TF32 itself: ~6e-4 here, documented as perf mode only
    # default precision for B > 64 is 3xTF32 on the same kernel: fp32-accurate (parity mode)
    exact = eng.linear_forward(cu(x), cu(W), cu(bias), relu=relu, keep_mask=cu(mask), mask_scale=2.0).cpu().numpy()
    assert rel_err(exact, ref) <= 6e-5        # tensor-core fp32 accumulation over K=5408 is the floor
    eng.set_matmul_precision(2)               # and the plain fp32 SIMT fallback still agrees
    try:
        simt = eng.linear_forward(cu(x), cu(W), cu(bias), relu=relu, keep_mask=cu(mask), mask_scale=2.0).cpu().numpy()
    finally:
        eng.set_matmul_precision(0)
    assert rel_err(simt, ref) <= 1e-5


@pytest.mark.parametrize("opt", ["sgd", "adam"])
def test_tcgen05_large_batch_train_step(eng, opt):
    """Whole large-batch step on tensor cores (forward, dX with MN-major W, dW fused with the
    optimizer with both operands MN-major) against the fp32 oracle, at TF32 accuracy."""
    rng = np.random.RandomState(5)
    dims = [5408, 512, 104, 100]
    shapes = [(512, 5408), (512,), (104, 512), (104,), (100, 104), (100,)]
    b = T(seeded_weights(shapes, 31))
    spec = R.BobSpec(tuple(dims), False, 0.5)
    gp = [cu(p.numpy()) for p in b]
    kw = dict(optimizer=eng.SN_OPT_ADAM, lr=1e-3, weight_decay=1e-5) if opt == "adam" else \
        dict(optimizer=eng.SN_OPT_SGD_MOMENTUM, lr=1e-3, momentum=0.9)
    bob = eng.BobEngine(dims, gp[0::2], gp[1::2], max_batch=300, **kw)
    B = 300
    act = (rng.rand(B, 5408) * 4 * (rng.rand(B, 5408) < 0.6)).astype(np.float32)
    y = rng.randint(0, 10, size=B).astype(np.int64)
    masks = seeded_masks(B, (512, 104), 60)
    bl = [p.detach().clone().requires_grad_(True) for p in b]
    at = torch.from_numpy(act).requires_grad_(True)
    logits_ref = R.bob_forward(at, bl, spec, T(masks))
    loss_ref = R.cross_entropy(logits_ref, torch.from_numpy(y))
    loss_ref.backward()
    w_old = [p.clone() for p in b]
    if opt == "adam":
        R.adam_(b, [p.grad for p in bl], R.AdamState(), 1e-3, weight_decay=1e-5)
    else:
        R.sgd_momentum_(b, [p.grad for p in bl], R.SGDState(), 1e-3, 0.9)
    eng.set_matmul_precision(1)
    try:
        loss, cut, logits = bob.train_step(cu(act), cu(y), masks=[cu(m) for m in masks], want_logits=True)
        torch.cuda.synchronize()
    finally:
        eng.set_matmul_precision(0)
    assert rel_err(logits.cpu(), logits_ref.detach()) <= 5e-3
    assert abs(loss.item() - loss_ref.item()) <= 5e-3 * abs(loss_ref.item())
    # TF32 truncates operands (10-bit mantissa, biased toward zero); six chained GEMMs and the
    # softmax gradient compound it to a few percent on the cut gradient.  Perf mode only.
    assert rel_err(cut.cpu(), at.grad) <= 5e-2
    for p_ref, p, p0 in zip(b, gp, w_old):
        # Adam's step is ~lr*sign(g): TF32 gradient error flips it wherever |g| is small
        assert rel_err(p.cpu(), p_ref) <= (2e-2 if opt == "adam" else 2e-3)
        if opt == "sgd":        # the applied step itself (lr * grad) at TF32 accuracy
            assert rel_err(p.cpu() - p0, p_ref - p0) <= 5e-2


def test_server_epoch_fused_forward_matches_stepwise_and_oracle(eng):
    """sn_bob_server_epoch fuses step t+1's first-layer forward into step t's weight traversal.
    Same batches through (a) the fused epoch call, (b) one sn_bob_train_step per batch (never
    fused), (c) the oracle: per-step losses and final weights must agree."""
    rng = np.random.RandomState(11)
    b = T(seeded_weights(B_SHAPES, 12))
    sizes = [16, 16, 16, 16, 9]                          # ragged last batch
    offs = np.concatenate([[0], np.cumsum(sizes)]).tolist()
    n = offs[-1]
    acts = (rng.rand(n, 5408) * 30 * (rng.rand(n, 5408) < 0.6)).astype(np.float32)
    y = rng.randint(0, 10, size=n).astype(np.int64)
    m1 = (rng.rand(n, 5000) < 0.5).astype(np.uint8)
    m2 = (rng.rand(n, 1000) < 0.5).astype(np.uint8)
    kw = dict(optimizer=eng.SN_OPT_ADAM, lr=1e-3, weight_decay=1e-5)
    bob_f, gp_f = _bob(eng, b, **kw)
    bob_s, gp_s = _bob(eng, b, **kw)
    acts_g, y_g, m1_g, m2_g = cu(acts), cu(y), cu(m1), cu(m2)
    loss_f = torch.zeros(len(sizes), device="cuda")
    bob_f.server_epoch(acts_g, y_g, offs, masks=[m1_g, m2_g], loss_per_step=loss_f)
    loss_s, loss_o = [], []
    opt = R.AdamState()
    for i, (s0, s1) in enumerate(zip(offs[:-1], offs[1:])):
        l, _, _ = bob_s.train_step(acts_g[s0:s1], y_g[s0:s1], masks=[m1_g[s0:s1], m2_g[s0:s1]], want_cut_grad=False)
        loss_s.append(l.item())
        ref = R.sisa_server_step(b, torch.from_numpy(acts[s0:s1]), torch.from_numpy(y[s0:s1]),
                                 [torch.from_numpy(m1[s0:s1]), torch.from_numpy(m2[s0:s1])])
        R.adam_(b, ref["bob_grads"], opt, 1e-3, weight_decay=1e-5)
        loss_o.append(ref["loss"].item())
    torch.cuda.synchronize()
    lf = loss_f.cpu().numpy()
    np.testing.assert_allclose(lf, np.array(loss_s), rtol=2e-4)     # fused vs unfused: summation order only
    np.testing.assert_allclose(lf[:1], np.array(loss_o[:1]), rtol=1e-4)
    np.testing.assert_allclose(lf, np.array(loss_o), rtol=5e-3)     # free-running Adam, 5 steps
    assert bob_f.get_opt_step(0) == len(sizes) == bob_s.get_opt_step(0)
    for n_, pf, ps, po in zip(B_NAMES, gp_f, gp_s, b):
        check_adam_weights(pf.cpu(), ps.cpu(), n_ + " fused vs stepwise")
        assert rel_err(pf.cpu(), po) <= 2e-3, n_


# --------------------------------------------------------------------------- edge cases / errors
def test_error_paths_and_edge_shapes(eng):
    from splitlearning_b200 import _lib
    b = T(seeded_weights([(12, 52), (12,), (10, 12), (10,)], 3))
    gp = [cu(p.numpy()) for p in b]
    bob = eng.BobEngine([52, 12, 10], gp[0::2], gp[1::2], dropout_p=0.0, optimizer=eng.SN_OPT_SGD_MOMENTUM, lr=0.1,
                        momentum=0.0, max_batch=8)
    x = torch.rand(9, 52, device="cuda")
    y = torch.zeros(9, dtype=torch.int64, device="cuda")
    with pytest.raises(_lib.SplitNNError, match="max_batch"):
        bob.train_step(x, y)
    with pytest.raises(_lib.SplitNNError, match="slot"):
        bob.train_step(x[:4], y[:4], slot=3)
    with pytest.raises(_lib.SplitNNError, match="no matching"):
        bob.backward_update(x[:4], torch.zeros(4, 10, device="cuda"))
    with pytest.raises(_lib.SplitNNError):
        eng.BobEngine([52, 12, 10], gp[0::2], gp[1::2], dropout_p=1.5)
    with pytest.raises(_lib.SplitNNError, match="float32"):
        bob.forward(x[:4].double())
    # B = 1, width not a multiple of 4 anywhere except the input: the small-layer kernel's scalar path
    b2 = T(seeded_weights([(7, 50), (7,), (3, 7), (3,)], 4))
    gp2 = [cu(p.numpy()) for p in b2]
    net = eng.BobEngine([50, 7, 3], gp2[0::2], gp2[1::2], dropout_p=0.0, optimizer=eng.SN_OPT_SGD_MOMENTUM, lr=0.1,
                        momentum=0.9, max_batch=16)
    spec = R.BobSpec((50, 7, 3), False, 0.0)
    opt = R.SGDState()
    for B_ in (1, 5):
        xx = torch.rand(B_, 50)
        yy = torch.randint(0, 3, (B_,))
        bl = [p.detach().clone().requires_grad_(True) for p in b2]
        xt = xx.clone().requires_grad_(True)
        loss_ref = R.cross_entropy(R.bob_forward(xt, bl, spec, None), yy)
        loss_ref.backward()
        R.sgd_momentum_(b2, [p.grad for p in bl], opt, 0.1, 0.9)
        loss, cut, _ = net.train_step(xx.cuda(), yy.cuda())
        assert abs(loss.item() - loss_ref.item()) <= 1e-5 * abs(loss_ref.item())
        assert rel_err(cut.cpu(), xt.grad) <= 1e-5
        for p_ref, p in zip(b2, gp2):
            assert rel_err(p.cpu(), p_ref) <= 1e-5
    # labels outside [0, C): torch raises; here the loss becomes NaN instead of reading out of bounds
    bad = torch.full((4,), 99, dtype=torch.int64, device="cuda")
    loss, _ = eng.cross_entropy(torch.rand(4, 10, device="cuda"), bad)
    assert torch.isnan(loss).item()
    # a zero-length SISA epoch is a no-op
    bob.server_epoch(torch.zeros(0, 52, device="cuda"), torch.zeros(0, dtype=torch.int64, device="cuda"), [0])
    torch.cuda.synchronize()


def test_skinny_forward_on_tensor_cores_matches_fma_kernel(eng):
    """The selectable tcgen05 form of the batch<=16 weight stream (3xTF32, weight operand in tensor
    memory, 12 independent TMEM accumulators): same inputs through both kernels and fp64."""
    rng = np.random.RandomState(77)
    for B, K, N in ((16, 5408, 5000), (9, 5408, 5000), (16, 4100, 4200)):
        x = (rng.rand(B, K) * 30 * (rng.rand(B, K) < 0.6)).astype(np.float32)
        W = (rng.uniform(-1, 1, (N, K)) / np.sqrt(K)).astype(np.float32)
        bias = rng.uniform(-1, 1, N).astype(np.float32)
        mask = (rng.rand(B, N) < 0.5).astype(np.uint8)
        ref = np.maximum(x.astype(np.float64) @ W.astype(np.float64).T + bias, 0) * mask * 2.0
        fma = eng.linear_forward(cu(x), cu(W), cu(bias), relu=True, keep_mask=cu(mask), mask_scale=2.0).cpu().numpy()
        eng.set_skinny_tensor_forward(True)
        try:
            tc = eng.linear_forward(cu(x), cu(W), cu(bias), relu=True, keep_mask=cu(mask), mask_scale=2.0).cpu().numpy()
            tc2 = eng.linear_forward(cu(x), cu(W), cu(bias), relu=True, keep_mask=cu(mask), mask_scale=2.0).cpu().numpy()
        finally:
            eng.set_skinny_tensor_forward(False)
        assert np.array_equal(tc, tc2)                  # deterministic
        assert rel_err(fma, ref) <= 2e-6
        assert rel_err(tc, ref) <= 1e-5                 # 3xTF32 + fp32 tensor-core accumulation


@pytest.mark.parametrize("tiled", [False, True])
@pytest.mark.parametrize("B", [17, 33, 64])
def test_mid_batch_backward_sgd(eng, B, tiled):
    """17..64 rows: the forward stays on the fp32 weight-stream kernels (row tiles of 16); the backward
    runs as tcgen05 GEMMs with the optimizer in the dW GEMM's epilogue (default), or - tiled - as fp32
    traversals in tiles of 16 with the per-tile dW added into a gradient buffer (OPT=3) and one optimizer
    pass at the end.  One SGD-momentum step (momentum buffers pre-loaded) against the oracle, including
    a ragged last tile and the cut gradient; then a second step to see the buffers are re-armed."""
    from splitlearning_b200 import _lib
    lib = _lib.load()
    was = lib.sn_get_bwd_tiled_max()
    _lib.check(lib.sn_set_bwd_tiled_max(64 if tiled else 16))
    try:
        _mid_batch_body(eng, B)
    finally:
        _lib.check(lib.sn_set_bwd_tiled_max(was))


def _mid_batch_body(eng, B):
    rng = np.random.RandomState(B)
    dims = [5408, 320, 60, 100]
    shapes = [(320, 5408), (320,), (60, 320), (60,), (100, 60), (100,)]
    b = T(seeded_weights(shapes, 33))
    spec = R.BobSpec(tuple(dims), False, 0.5)
    gp = [cu(p.numpy()) for p in b]
    bob = eng.BobEngine(dims, gp[0::2], gp[1::2], optimizer=eng.SN_OPT_SGD_MOMENTUM, lr=1e-3, momentum=0.9, max_batch=64)
    opt = R.SGDState()
    for step in range(2):
        act = (rng.rand(B, 5408) * 40 * (rng.rand(B, 5408) < 0.6)).astype(np.float32)
        y = rng.randint(0, 10, size=B).astype(np.int64)
        masks = seeded_masks(B, (320, 60), 70 + step)
        bl = [p.detach().clone().requires_grad_(True) for p in b]
        at = torch.from_numpy(act).requires_grad_(True)
        loss_ref = R.cross_entropy(R.bob_forward(at, bl, spec, T(masks)), torch.from_numpy(y))
        loss_ref.backward()
        w_old = [p.clone() for p in b]
        R.sgd_momentum_(b, [p.grad for p in bl], opt, 1e-3, 0.9)
        loss, cut, _ = bob.train_step(cu(act), cu(y), masks=[cu(m) for m in masks])
        assert abs(loss.item() - loss_ref.item()) <= TOL * abs(loss_ref.item())
        assert rel_err(cut.cpu(), at.grad) <= TOL
        for p_ref, p, p0 in zip(b, gp, w_old):
            assert rel_err(p.cpu(), p_ref) <= 1e-5
            assert rel_err(p.cpu() - p0, p_ref - p0) <= TOL          # the applied step itself


@pytest.mark.parametrize("mode", [0, 1])
def test_philox_dropout_is_the_same_function_on_every_forward_path(eng, mode):
    """The keep decision of an element depends on (seed, step, layer, element index) only: a 96-row
    training-mode forward through the tcgen05 GEMMs (Philox words shared inside quads of lanes in the
    epilogue; also split along K where the output has few tiles) keeps exactly the elements the fp32
    weight-stream kernels keep (row tiles of 16, one Philox evaluation per element), for hidden widths
    that are and are not whole 128-column tiles."""
    rng = np.random.RandomState(11)
    dims = [5408, 320, 200, 100]
    shapes = [(320, 5408), (320,), (200, 320), (200,), (100, 200), (100,)]
    w = [cu(a) for a in seeded_weights(shapes, 91)]
    # positive biases keep most pre-activations away from the ReLU kink, so zeros mark dropped elements
    w[1] += 3.0
    w[3] += 3.0
    B = 96
    x = cu((rng.rand(B, 5408) * 2).astype(np.float32))
    a = eng.BobEngine(dims, w[0::2], w[1::2], optimizer=eng.SN_OPT_NONE, max_batch=B, dropout_p=0.5, relu_last=True)
    b = eng.BobEngine(dims, w[0::2], w[1::2], optimizer=eng.SN_OPT_NONE, max_batch=B, dropout_p=0.5, relu_last=True)
    assert a.set_stream_rows(B) >= B                     # a: fp32 weight-stream kernels; b: GEMMs (B > 64)
    eng.set_matmul_precision(mode)
    try:
        for off in (0, 5):
            ya = a.forward(x, train_mode=True, seed=1234, offset=off).cpu()
            yb = b.forward(x, train_mode=True, seed=1234, offset=off).cpu()
            assert rel_err(yb, ya) <= (TOL if mode == 0 else 5e-3)
    finally:
        eng.set_matmul_precision(0)


_SPLITK_CHILD = r"""
import sys
sys.path.insert(0, %r)
import numpy as np, torch
from splitlearning_b200 import engine as E
from tests.helpers import seeded_weights, seeded_masks
rng = np.random.RandomState(3)
dims = [5408, 128, 104, 100]
shapes = [(128, 5408), (128,), (104, 128), (104,), (100, 104), (100,)]
gp = [torch.from_numpy(a.copy()).cuda() for a in seeded_weights(shapes, 44)]
B = 1200
bob = E.BobEngine(dims, gp[0::2], gp[1::2], optimizer=E.SN_OPT_SGD_MOMENTUM, lr=1e-3, momentum=0.9, max_batch=B)
act = torch.from_numpy((rng.rand(B, 5408) * 4 * (rng.rand(B, 5408) < 0.6)).astype(np.float32)).cuda()
y = torch.from_numpy(rng.randint(0, 10, size=B).astype(np.int64)).cuda()
masks = [torch.from_numpy(m).cuda() for m in seeded_masks(B, (128, 104), 9)]
for step in range(2):
    loss, cut, _ = bob.train_step(act, y, masks=masks)
torch.cuda.synchronize()
np.savez(sys.argv[1], loss=loss.cpu().numpy(), cut=cut.cpu().numpy(), **{"p%%d" %% i: p.cpu().numpy() for i, p in enumerate(gp)})
"""


def test_split_k_gemms_equal_the_unsplit_ones(tmp_path):
    """Narrow layers at a large batch split the reduction of all three GEMM forms (forward, input
    gradient, dW + optimizer through a slice buffer and `splitk_opt_kernel`): two SGD-momentum steps
    with and without the split (SN_NO_SPLITK=1, read once per process, hence children) agree to the
    summation-order level."""
    import subprocess
    import sys
    out = {}
    for tag, env in (("split", {}), ("whole", {"SN_NO_SPLITK": "1"})):
        path = str(tmp_path / (tag + ".npz"))
        r = subprocess.run([sys.executable, "-c", _SPLITK_CHILD % ROOT, path], env=dict(os.environ, **env), capture_output=True,
                           text=True, timeout=300)
        assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
        out[tag] = np.load(path)
    a, b = out["split"], out["whole"]
    assert abs(float(a["loss"]) - float(b["loss"])) <= 1e-5 * abs(float(b["loss"]))
    assert rel_err(a["cut"], b["cut"]) <= 1e-4
    w0 = seeded_weights([(128, 5408), (128,), (104, 128), (104,), (100, 104), (100,)], 44)
    for i in range(6):
        assert rel_err(a["p%d" % i] - w0[i], b["p%d" % i] - w0[i]) <= 1e-4, i
