"""Tensor-parallel first layer (sn_shard_*, splitlearning_b200/tp.py) against the unsharded SISA
server epoch (sn_bob_server_epoch): same cached batches, same dropout masks, same Adam - losses per
step and final weights must agree (the sharding only regroups fc1's partial sums, ~1e-6).
  * one GPU: the row slices of fc1 as two sn_shard handles driven by hand (no transport);
  * two GPUs (skipped on a single-GPU box): tp.TPServer over CUDA IPC mailboxes and gate flags."""
import os
import tempfile

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

B = 16
DIMS = [5408, 5000, 1000, 100]
SIZES = [16, 16, 16, 16, 16, 11]          # ragged last batch


def _weights(device, seed=21):
    g = torch.Generator().manual_seed(seed)
    Ws = [(torch.randn(DIMS[i + 1], DIMS[i], generator=g) / DIMS[i] ** 0.5).to(device) for i in range(3)]
    bs = [(torch.randn(DIMS[i + 1], generator=g) * 0.01).to(device) for i in range(3)]
    return Ws, bs


def _data(device):
    g = torch.Generator().manual_seed(5)
    n = sum(SIZES)
    acts = (torch.rand(n, 5408, generator=g) * 30 * (torch.rand(n, 5408, generator=g) < 0.6)).to(device)
    y = torch.randint(0, 10, (n,), generator=g).to(device)
    m1 = (torch.rand(n, 5000, generator=g) < 0.5).to(torch.uint8).to(device)
    m2 = (torch.rand(n, 1000, generator=g) < 0.5).to(torch.uint8).to(device)
    offs = np.concatenate([[0], np.cumsum(SIZES)]).tolist()
    return acts, y, m1, m2, offs


def _reference(device):
    from splitlearning_b200 import engine as E
    Ws, bs = _weights(device)
    acts, y, m1, m2, offs = _data(device)
    bob = E.BobEngine(DIMS, Ws, bs, optimizer=E.SN_OPT_ADAM, lr=1e-3, weight_decay=1e-5, n_opt_slots=1, max_batch=B)
    losses = torch.zeros(2, len(SIZES), device=device)
    bob.server_epoch(acts, y, offs, masks=[m1, m2], loss_per_step=losses[0])
    bob.server_epoch(acts, y, offs, masks=[m1, m2], loss_per_step=losses[1])  # second epoch: Adam state carried over
    torch.cuda.synchronize()
    return [w.cpu().numpy() for w in Ws], [b.cpu().numpy() for b in bs], losses.cpu().numpy()


def _check(losses, Ws, bs, rloss, rW, rb):
    """Free-running Adam on un-normalised activations amplifies rounding differences step by step
    (numerical notes): the first steps pin the arithmetic, the rest the trajectory."""
    assert np.allclose(losses[0][:3], rloss[0][:3], rtol=2e-5), (losses, rloss)
    assert np.allclose(losses[0], rloss[0], rtol=5e-4), (losses, rloss)
    assert np.allclose(losses[1], rloss[1], rtol=2e-2), (losses, rloss)
    for got, ref in zip(Ws, rW):
        assert _rel(got, ref) <= 2e-3
    for got, ref in zip(bs, rb):
        assert _rel(got, ref) <= 2e-2


def _rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def test_two_shards_on_one_gpu_match_unsharded_epoch():
    from splitlearning_b200 import engine as E
    dev = torch.device("cuda", 0)
    rW, rb, rloss = _reference(dev)
    Ws, bs = _weights(dev)
    acts, y, m1, m2, offs = _data(dev)
    cuts = [0, 2600, 5000]                   # uneven slices
    kw = dict(optimizer=E.SN_OPT_ADAM, lr=1e-3, weight_decay=1e-5)
    shards = [E.ShardEngine(Ws[0][lo:hi], bs[0][lo:hi], 5000, lo, dropout_p=0.5, **kw) for lo, hi in zip(cuts[:-1], cuts[1:])]
    head = E.BobEngine(DIMS[1:], Ws[1:], bs[1:], dropout_p=0.5, n_opt_slots=1, max_batch=B, **kw)
    H1 = torch.zeros(B, 5000, device=dev)
    dH1 = torch.zeros(B, 5000, device=dev)
    losses = torch.zeros(2, len(SIZES), device=dev)
    for ep in range(2):
        for i in range(len(SIZES)):
            s, e = offs[i], offs[i + 1]
            X = acts[s:e]
            Xn = acts[offs[i + 1]:offs[i + 2]] if i + 1 < len(SIZES) else None
            for sh in shards:
                sh.forward(X, H1, 5000, train=True, mask=m1[s:e])
            head.train_step(H1[:e - s], y[s:e], masks=[m2[s:e]], loss_out=losses[ep][i], cut_out=dH1[:e - s])
            for sh in shards:
                sh.backward_update(X, dH1, 5000, X_next=Xn)
    torch.cuda.synchronize()
    _check(losses.cpu().numpy(), [w.cpu().numpy() for w in Ws], [b.cpu().numpy() for b in bs], rloss, rW, rb)
    with pytest.raises(Exception):           # backward without a matching train-mode forward
        shards[0].backward_update(acts[:16], dH1, 5000)


def test_sharded_first_layer_steps_vs_oracle():
    """Teacher-forced against the CPU oracle (not against this library's own unsharded path): three
    SISA server steps with fc1 split into two uneven sn_shard slices + a head over fc2/fc3 on one GPU;
    logits and loss at 1e-4, Adam-updated weights by check_adam_weights, every step restarted from the
    oracle's weights and moments (`data_entities_vanilla_sisa.py:305-313`, torch Adam wd 1e-5)."""
    from oracle import torch_ref as R
    from splitlearning_b200 import engine as E
    from tests.helpers import check_adam_weights, rel_err
    dev = torch.device("cuda", 0)
    Ws, bs = _weights(dev)
    acts, y, m1, m2, offs = _data(dev)
    ref_p = []
    for w, b in zip(Ws, bs):
        ref_p += [w.cpu().clone(), b.cpu().clone()]
    cuts = [0, 2600, 5000]
    kw = dict(optimizer=E.SN_OPT_ADAM, lr=1e-3, weight_decay=1e-5)
    shards = [E.ShardEngine(Ws[0][lo:hi], bs[0][lo:hi], 5000, lo, dropout_p=0.5, **kw) for lo, hi in zip(cuts[:-1], cuts[1:])]
    head = E.BobEngine(DIMS[1:], Ws[1:], bs[1:], dropout_p=0.5, n_opt_slots=1, max_batch=B, **kw)
    H1 = torch.zeros(B, 5000, device=dev)
    dH1 = torch.zeros(B, 5000, device=dev)
    opt = R.AdamState()
    for i in (0, 1, 5):                                    # two full batches and the ragged last one
        s, e = offs[i], offs[i + 1]
        masks = [m1[s:e].cpu(), m2[s:e].cpu()]
        ref = R.sisa_server_step(ref_p, acts[s:e].cpu(), y[s:e].cpu(), masks)
        R.adam_(ref_p, ref["bob_grads"], opt, 1e-3, weight_decay=1e-5)
        for sh in shards:
            sh.forward(acts[s:e], H1, 5000, train=True, mask=m1[s:e])
        loss, _, logits = head.train_step(H1[:e - s], y[s:e], masks=[m2[s:e]], cut_out=dH1[:e - s], want_logits=True)
        for sh in shards:
            sh.backward_update(acts[s:e], dH1, 5000)
        torch.cuda.synchronize()
        assert rel_err(logits.cpu(), ref["logits"]) <= 1e-4
        assert abs(loss.item() - ref["loss"].item()) <= 1e-4 * abs(ref["loss"].item())
        for j, (w, b) in enumerate(zip(Ws, bs)):
            check_adam_weights(w.cpu(), ref_p[2 * j], f"fc{j + 1}.weight")
            check_adam_weights(b.cpu(), ref_p[2 * j + 1], f"fc{j + 1}.bias")
        # teacher forcing: the next step starts from the oracle's weights, moments and step count
        for j, (w, b) in enumerate(zip(Ws, bs)):
            w.copy_(ref_p[2 * j])
            b.copy_(ref_p[2 * j + 1])
        for sh, lo, hi in zip(shards, cuts[:-1], cuts[1:]):
            sh.opt_state(0, 0).copy_(opt.m[0][lo:hi].reshape(-1))
            sh.opt_state(1, 0).copy_(opt.v[0][lo:hi].reshape(-1))
            sh.opt_state(0, 1).copy_(opt.m[1][lo:hi])
            sh.opt_state(1, 1).copy_(opt.v[1][lo:hi])
            sh.set_opt_step(opt.step)
            assert sh.get_opt_step() == opt.step
        for j in range(2, 6):
            head.opt_state(0, 0, j - 2).copy_(opt.m[j].reshape(-1))
            head.opt_state(0, 1, j - 2).copy_(opt.v[j].reshape(-1))
        assert head.get_opt_step(0) == opt.step


def _tp_rank(rank, world, out_dir, port):
    import torch.distributed as dist
    from splitlearning_b200 import tp
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    device = torch.device("cuda", rank)
    torch.cuda.set_device(device)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=device)
    Ws, bs = _weights(device)
    acts, y, m1, m2, offs = _data(device)
    srv = tp.TPServer(rank, world, device, Ws, bs, B=B, lr=1e-3, weight_decay=1e-5, dropout_p=0.5)
    losses = torch.zeros(2, len(SIZES), device=device)
    for ep in range(2):
        srv.server_epoch(acts, y, offs, masks=[m1, m2], loss_per_step=losses[ep] if rank == 0 else None)
    torch.cuda.synchronize()
    srv.gather_fc1(Ws[0], bs[0])
    torch.cuda.synchronize()
    if rank == 0:
        np.savez(os.path.join(out_dir, "tp.npz"), losses=losses.cpu().numpy(),
                 **{f"W{i}": w.cpu().numpy() for i, w in enumerate(Ws)}, **{f"b{i}": b.cpu().numpy() for i, b in enumerate(bs)})
    srv.close()
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_tensor_parallel_server_two_gpus_matches_one():
    import torch.multiprocessing as mp
    rW, rb, rloss = _reference(torch.device("cuda", 0))
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(_tp_rank, args=(2, d, 29841), nprocs=2, join=True)
        with np.load(os.path.join(d, "tp.npz")) as f:
            got = {k: f[k] for k in f.files}
    _check(got["losses"], [got[f"W{i}"] for i in range(3)], [got[f"b{i}"] for i in range(3)], rloss, rW, rb)
