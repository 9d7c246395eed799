"""The drop-in entity classes on the GPU against trajectories recorded from the REFERENCE's own
RPC run (tests/golden/rpc_*_ws3.npz: Bob + 2 Alices in 3 processes, same synthetic loaders, torch
seeds Bob 100 / Alice r 200+r).  Free-running (no teacher forcing): per-entity RNG streams
reproduce the reference's initial weights, shuffles and dropout masks, so the only difference is
fp32 rounding, compared over the short horizon the fixtures cover."""
import os
import random
from argparse import Namespace

import numpy as np
import pytest
import torch

from tests.helpers import check_summary, load_golden, rel_err

pytestmark = pytest.mark.gpu
B_NAMES = ["fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias", "fc3.weight", "fc3.bias"]


def _world(tmp_path, monkeypatch, n=300, K=2, **flags):
    from splitlearning_b200.data.relational_table_preprocessor import image_preprocess_dl
    from splitlearning_b200.data.synthetic import make_synthetic_mnist
    monkeypatch.chdir(tmp_path)
    os.makedirs("data/mnist_flat", exist_ok=True)
    data, labels = make_synthetic_mnist(n, seed=1234)
    args = Namespace(world_size=K + 1, epochs=1, iterations=1, batch_size=16, partition_alpha=0.5,
                     datapath="data/mnist_flat", lr=1e-3, server_epochs=2, vanilla=False, sisa=False, concat=False,
                     control=False, client_num_in_total=K, class_num=10,
                     torch_seeds={"bob": 100, "alice_base": 200}, dropout_parity=True)
    for k, v in flags.items():
        setattr(args, k, v)
    np.random.seed(0)
    random.seed(0)
    parts = image_preprocess_dl(args, data.astype(np.float64), labels.tolist(), 0.2)
    for k in parts[5]:
        torch.save(parts[5][k], f"data/mnist_flat/data_worker{k + 1}_train.pt")
        torch.save(parts[6][k], f"data/mnist_flat/data_worker{k + 1}_test.pt")
    return args


def test_vanilla_matches_reference_rpc_run(tmp_path, monkeypatch):
    from splitlearning_b200 import data_entities_vanilla as ent
    g = load_golden("rpc_vanilla_ws3.npz")
    args = _world(tmp_path, monkeypatch, vanilla=True)
    BOB = ent.bob(args)
    losses, labels = [], []
    orig = BOB.train_and_backward

    def rec(x, y, ctx):
        loss = orig(x, y, ctx)
        losses.append(loss.clone())
        labels.append(y.clone())
        return loss
    BOB.train_and_backward = rec
    for c in (1, 2):
        BOB.train_request(c)
    torch.cuda.synchronize()
    assert np.array_equal(torch.cat(labels).cpu().numpy(), g["labels"])       # same shuffles, same batches
    got = torch.stack(losses).cpu().numpy()
    np.testing.assert_allclose(got, g["loss"], rtol=2e-3)
    assert abs(got[0] - g["loss"][0]) <= 1e-4 * abs(g["loss"][0])              # first step: pure fp32 parity
    for n, p in zip(B_NAMES, BOB.model.parameters()):
        check_summary(g, "final_" + n, p, 1e-3)
    for c in (1, 2):
        w = BOB.alices[c].give_weights()["conv_layers.0.weight"].cpu().numpy()
        assert rel_err(w, g[f"final_alice{c}.conv_layers.0.weight"]) <= 1e-3
    # logs use the reference's line formats
    BOB.eval_request_breakdown(9)
    text = open("logs/bob.log").read()
    assert "Bob Started Getting Tipsy" in text and "Train Request for Alice-2" in text
    assert "Accuracy over all data: " in text and "Accuracy over remaining data: " in text
    assert "Alice-1 receiving weights" not in open("logs/alice1.log").read()
    assert "Alice-2 receiving weights from Alice-1" in open("logs/alice2.log").read()


def test_sisa_matches_reference_rpc_run(tmp_path, monkeypatch):
    from splitlearning_b200 import data_entities_vanilla_sisa as ent
    g = load_golden("rpc_sisa_ws3.npz")
    args = _world(tmp_path, monkeypatch, sisa=True)
    BOB = ent.bob(args)
    BOB.record_loss = True
    BOB.train_request_parallel()
    BOB.freeze_alice_weights(range(1, 3))
    BOB.train_and_backward(unlearn_request_from_alices=[], unlearn_id=None)
    n_first = sum(int(l.numel()) for l in BOB.loss_history)
    BOB.unfreeze_alice_weights([1])
    BOB.unlearn_request(client_id=1, omit_label=9)
    BOB.freeze_alice_weights([1])
    BOB.train_and_backward([1], unlearn_id=9)
    torch.cuda.synchronize()
    got = torch.cat(BOB.loss_history).cpu().numpy()
    assert got.shape == g["loss"].shape
    # first server phase (2 epochs): fp32-level agreement that loosens as Adam's chaos accumulates
    # un-normalised 0..255 inputs push early losses into the hundreds and the trajectory is chaotic
    # (SURVEY.md 0.8): fp32 rounding differences grow ~10x per few steps, so only a short horizon is
    # comparable in a free-running run; per-step parity is the teacher-forced tests' job
    np.testing.assert_allclose(got[:4], g["loss"][:4], rtol=1e-3)
    np.testing.assert_allclose(got[:6], g["loss"][:6], rtol=2e-2)
    np.testing.assert_allclose(got[:10], g["loss"][:10], rtol=0.25)
    assert np.isfinite(got).all() and 0 < n_first < len(got)
    # cache replay: every epoch of a phase sees the same label sequence
    labs = np.concatenate([c.labels.cpu().numpy() for k, c in sorted(BOB.activation_and_labels_cache.items(), key=lambda kv: (kv[0][1], kv[0][0])) if not k[1]])
    assert np.array_equal(np.tile(labs, 2), g["labels"][:2 * len(labs)])
    acc = BOB.eval_request_breakdown(9)
    assert 0.0 <= acc[0] <= 1.0
    for c in (1, 2):
        w = BOB.alices[c].give_weights()["conv_layers.0.weight"].cpu().numpy()
        assert rel_err(w, g[f"final_alice{c}.conv_layers.0.weight"]) <= 5e-3
    assert (1, True, 9) in BOB.activation_and_labels_cache and (2, False, None) in BOB.activation_and_labels_cache


def test_ushape_matches_reference_rpc_run(tmp_path, monkeypatch):
    from splitlearning_b200 import data_entities as ent
    g = load_golden("rpc_ushape_ws3.npz")
    args = _world(tmp_path, monkeypatch)
    BOB = ent.bob(args)
    outs = []
    orig = BOB.inference_train

    def rec(x):
        out = orig(x)
        if len(outs) < 3:
            outs.append(out.clone())
        return out
    BOB.inference_train = rec
    for c in (1, 2):
        BOB.train_request(c)
    torch.cuda.synchronize()
    # per-step quantity: Bob's output of the first steps against the recorded RPC run (step 0 depends on
    # the initial weights and the first batch only: 1e-4; then the free-running drift of Adam sets in)
    rows = g["logit_rows"]
    ref0 = g["logits"][:rows[0] * 100].reshape(rows[0], 100)
    assert rel_err(outs[0].cpu().numpy(), ref0) <= 1e-4
    ref1 = g["logits"][rows[0] * 100:(rows[0] + rows[1]) * 100].reshape(rows[1], 100)
    assert rel_err(outs[1].cpu().numpy(), ref1) <= 5e-3
    for n, p in zip(["fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias"], BOB.model.parameters()):
        check_summary(g, "final_" + n, p, 5e-3)
    for c in (1, 2):
        m1, m3 = BOB.alices[c].give_weights()
        assert rel_err(m1["conv1.weight"].cpu().numpy(), g[f"final_alice{c}_m0.conv1.weight"]) <= 5e-3
        assert rel_err(m3["fc3.weight"].cpu().numpy(), g[f"final_alice{c}_m1.fc3.weight"]) <= 5e-3
    # the methods the reference's CLI needs but its module lacks
    acc = BOB.eval_request_breakdown(9)
    assert 0.0 <= acc[0] <= 1.0


def test_concat_defined_semantics(tmp_path, monkeypatch):
    """--concat is unpinned by the reference (its loop crashes); check the semantics this build
    defines against the oracle's forward/CE on the concatenated tensors."""
    from oracle import torch_ref as R
    from splitlearning_b200 import data_entities_sisa_concat as ent
    args = _world(tmp_path, monkeypatch, sisa=True, concat=True)
    BOB = ent.bob(args)
    BOB.record_loss = True
    BOB.server_epochs = 1
    BOB.train_request_parallel()
    BOB.freeze_alice_weights(range(1, 3))
    params0 = [p.detach().cpu().clone() for p in BOB.model.parameters()]
    # replay the masks the server epoch will draw
    state = BOB.rng.state.clone()
    BOB.train_and_backward(unlearn_request_from_alices=[], unlearn_id=None)
    torch.cuda.synchronize()
    acts, labels, offsets = BOB._concat
    keep = torch.get_rng_state()
    torch.set_rng_state(state)
    spec = R.model2_sisa_concat_spec(2)
    masks = R.draw_dropout_masks(16, spec)
    torch.set_rng_state(keep)
    x0 = acts[:16].cpu()
    logits = R.bob_forward(x0, params0, spec, masks)
    ref = R.cross_entropy(logits.reshape(32, 100), labels[:16].cpu().reshape(32))
    got = float(BOB.loss_history[0][0].item())
    assert abs(got - float(ref)) <= 1e-4 * abs(float(ref))
    assert acts.shape[1] == 5408 * 2 and labels.shape[1] == 2
    acc = BOB.eval_request_breakdown(9)
    assert 0.0 <= acc[0] <= 1.0
    BOB.unfreeze_alice_weights([1])
    BOB.unlearn_request(client_id=1, omit_label=9)
    BOB.freeze_alice_weights([1])
    BOB.train_and_backward([1], unlearn_id=9)
    torch.cuda.synchronize()
    assert BOB._concat_key[0] == (1, True, 9)


def test_cli_runs_vanilla_and_sisa(tmp_path, monkeypatch):
    from splitlearning_b200 import split_nn
    monkeypatch.chdir(tmp_path)
    BOB = split_nn.main(["--vanilla", "--world_size", "3", "--iterations", "1", "--synthetic_samples", "400", "--seed", "0"])
    assert os.path.exists("logs/bob.log") and os.path.exists("data/mnist_flat/data_worker2_test.pt")
    assert open("logs/bob.log").read().count("Accuracy over all data") == 2
    BOB = split_nn.main(["--sisa", "--world_size", "3", "--server_epochs", "1", "--synthetic_samples", "400", "--seed", "0"])
    assert open("logs/bob.log").read().count("Global training completed.") == 2


def test_eval_accuracy_parity_with_reference_weights(tmp_path, monkeypatch):
    """Accuracy parity (BASELINE.json: within 0.3 pp).  Free-running training is chaotic on this
    model (un-normalised inputs; measured on a B200: the SAME GPU run with three dropout seeds ends
    at 60 % / 69 % / 91 % test accuracy after 4 server epochs, profiles/accuracy_parity.py), so a
    trained-vs-trained comparison measures seed noise, not the implementation.  What can be pinned:
    weights trained by the CPU oracle, loaded through state_dict into the GPU modules, must give the
    reference's evaluation counters on the same test split."""
    from oracle import protocol as P
    from oracle import torch_ref as R
    from splitlearning_b200 import data_entities_vanilla_sisa as ent
    args = _world(tmp_path, monkeypatch, n=1500, sisa=True)
    BOB = ent.bob(args)
    train = {c: BOB.alices[c].train_dataloader for c in (1, 2)}
    bob_p, alice_p = P.run_sisa(train, lr=1e-3, epochs=1, server_epochs=1, batch_size=16, do_unlearn=False)
    names = ["fc1.weight", "fc1.bias", "fc2.weight", "fc2.bias", "fc3.weight", "fc3.bias"]
    BOB.model.load_state_dict({k: v for k, v in zip(names, bob_p)})
    ref_counts = [0] * 6
    for c in (1, 2):
        BOB.alices[c].model.load_state_dict({"conv_layers.0.weight": alice_p[c][0], "conv_layers.0.bias": alice_p[c][1]})
        x, y = BOB.alices[c].test_dataloader.dataset.tensors
        with torch.no_grad():
            out = R.bob_forward(R.alice_forward(x, alice_p[c][0], alice_p[c][1]), bob_p, R.MODEL2_SISA, None)
        for i, v in enumerate(R.eval_counts(out, y, 9)):
            ref_counts[i] += v
    got = [0] * 6
    for c in (1, 2):
        for i, v in enumerate(BOB.alices[c].eval_breakdown(9)):
            got[i] += v
    assert got[1] == ref_counts[1] and got[3] == ref_counts[3] and got[5] == ref_counts[5]     # same split
    acc_ref, acc_got = ref_counts[0] / ref_counts[1], got[0] / got[1]
    assert abs(acc_ref - acc_got) <= 0.003, (acc_ref, acc_got)
    assert abs(got[0] - ref_counts[0]) <= 1          # at most a near-tie flips

    # the same weights on a 12,000-row test set (SURVEY.md 8c asks for >= 10k): now 0.3 pp is 36 rows
    from splitlearning_b200._entity_common import DeviceDataset
    from splitlearning_b200.data.synthetic import make_synthetic_mnist
    data, labels = make_synthetic_mnist(12000, seed=4321)
    xb, yb = torch.from_numpy(data.astype(np.float32)).reshape(-1, 1, 28, 28), torch.tensor(labels, dtype=torch.int64)
    with torch.no_grad():
        out = R.bob_forward(R.alice_forward(xb, alice_p[1][0], alice_p[1][1]), bob_p, R.MODEL2_SISA, None)
    big_ref = R.eval_counts(out, yb, 9)
    A = BOB.alices[1]
    A.test_data = DeviceDataset.from_tensors(xb, yb, 16, False, A.device)
    big_got = list(A.eval_breakdown(9))
    assert big_got[1] == 12000 and [big_got[i] for i in (1, 3, 5)] == [big_ref[i] for i in (1, 3, 5)]
    assert abs(big_got[0] - big_ref[0]) / 12000 <= 0.003, (big_got, big_ref)
    assert abs(big_got[0] - big_ref[0]) <= 3, (big_got, big_ref)       # only near-tie rows may flip


@pytest.mark.parametrize("mode", ["sisa", "vanilla"])
def test_fused_evaluation_gives_the_per_batch_counters(tmp_path, monkeypatch, mode):
    """Evaluation over a whole shard in a few launches (eval_loop's chunked pass) against the
    reference-shaped loop of 16-sample batches: the six counters must be IDENTICAL on a 10,007-row test
    set - in SISA's eval mode, and in --vanilla where dropout stays live during evaluation
    (`data_entities_vanilla.py:234-235`) with the mask stream drawn batch by batch from the CPU generator."""
    from splitlearning_b200 import _lib
    from splitlearning_b200._entity_common import DeviceDataset
    from splitlearning_b200.data.synthetic import make_synthetic_mnist
    if mode == "sisa":
        from splitlearning_b200 import data_entities_vanilla_sisa as ent
        args = _world(tmp_path, monkeypatch, sisa=True)
    else:
        from splitlearning_b200 import data_entities_vanilla as ent
        args = _world(tmp_path, monkeypatch, vanilla=True)
    BOB = ent.bob(args)
    BOB.train_request(1)                      # a few real steps so that the logits are not degenerate
    data, labels = make_synthetic_mnist(10007, seed=99)            # ragged last batch (10007 = 625*16 + 7)
    xb, yb = torch.from_numpy(data.astype(np.float32)).reshape(-1, 1, 28, 28), torch.tensor(labels, dtype=torch.int64)
    A = BOB.alices[1]
    A.test_data = DeviceDataset.from_tensors(xb, yb, 16, False, A.device)
    lib = _lib.load()

    def run(chunk):
        monkeypatch.setenv("SN_EVAL_CHUNK", str(chunk))
        state_a, state_b = A.rng.state.clone(), BOB.rng.state.clone()
        n0 = lib.sn_launch_count()
        c = A.eval_breakdown(9)
        launches = lib.sn_launch_count() - n0
        A.rng.state, BOB.rng.state = state_a, state_b              # both passes see the same generator streams
        return c, launches
    loop, n_loop = run(0)
    fused, n_fused = run(4096)
    assert loop[1] == 10007 and fused == loop, (fused, loop)
    assert n_fused <= 30 and n_loop >= 626 * 5, (n_fused, n_loop)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
@pytest.mark.parametrize("mode", [["--vanilla"], ["--sisa", "--server_epochs", "1"], []])
def test_cli_entities_on_two_gpus_match_one_gpu(tmp_path, monkeypatch, mode):
    """`place_entities`: Bob on cuda:0, the Alices on cuda:1, one process.  The drop-in classes then
    let Alice's kernel store into Bob's GPU (and back) with events between the streams; the run must
    produce exactly the one-GPU run's weights (same kernels, same order, same Philox streams)."""
    from splitlearning_b200 import split_nn
    monkeypatch.chdir(tmp_path)
    common = mode + ["--world_size", "3", "--iterations", "1", "--synthetic_samples", "400", "--seed", "0"]
    one = split_nn.main(common + ["--n_gpus", "1"])
    w_one = [p.detach().cpu().clone() for p in one.model.parameters()]
    a_one = [p.detach().cpu().clone() for c in (1, 2) for p in (one.alices[c].model.parameters() if hasattr(one.alices[c], "model")
                                                                else one.alices[c].model1.parameters())]
    two = split_nn.main(common + ["--n_gpus", "2"])
    torch.cuda.synchronize(0)
    torch.cuda.synchronize(1)
    assert two.device == torch.device("cuda", 0) and two.alices[1].device == torch.device("cuda", 1)
    assert "Alice-1 on cuda:1" in open("logs/alice1.log").read() and "Bob on cuda:0" in open("logs/bob.log").read()
    for a, b in zip(w_one, two.model.parameters()):
        assert torch.equal(a, b.detach().cpu())
    a_two = [p.detach().cpu() for c in (1, 2) for p in (two.alices[c].model.parameters() if hasattr(two.alices[c], "model")
                                                        else two.alices[c].model1.parameters())]
    for a, b in zip(a_one, a_two):
        assert torch.equal(a, b)
