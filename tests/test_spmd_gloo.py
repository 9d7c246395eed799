"""The one-process-per-GPU protocol (splitlearning_b200/spmd.py) on CPU: world_size 2 and 3 over
gloo with oracle-backed stand-in engines, against the same schedule run in a single process.
Checks the message flow (activations+labels one way, cut gradient back), the per-Alice optimizer
slots on Bob and the Alice->Alice weight hand-off across ranks: results must be bit-identical."""
import os
import tempfile

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import torch_ref as R
from splitlearning_b200 import spmd

B, N_ALICES, TURN, STEPS = 4, 3, 2, 13
SPEC = R.BobSpec((5408, 24, 10), False, 0.0)


class FakeAlice:
    def __init__(self, seed):
        g = torch.Generator().manual_seed(seed)
        self.p = [torch.randn(32, 1, 3, 3, generator=g) * 0.1, torch.randn(32, generator=g) * 0.1]
        self.opt = R.SGDState()

    def forward(self, x, out=None):
        act = R.alice_forward(x, self.p[0], self.p[1])
        if out is not None:
            out.copy_(act)
            return out
        return act

    def backward_update(self, x, cut):
        w, b = [t.detach().clone().requires_grad_(True) for t in self.p]
        R.alice_forward(x, w, b).backward(cut)
        R.sgd_momentum_(self.p, [w.grad, b.grad], self.opt, 1e-3, 0.9)

    def get_weights(self):
        return torch.cat([self.p[0].reshape(-1), self.p[1].reshape(-1)])

    def set_weights(self, flat):
        self.p[0].copy_(flat[:288].view(32, 1, 3, 3))
        self.p[1].copy_(flat[288:320])


class FakeBob:
    def __init__(self):
        g = torch.Generator().manual_seed(5)
        self.p = [torch.randn(24, 5408, generator=g) * 0.01, torch.zeros(24), torch.randn(10, 24, generator=g) * 0.1,
                  torch.zeros(10)]
        self.opts = [R.SGDState() for _ in range(N_ALICES)]
        self.slots_seen = []

    def train_step(self, act, y, slot):
        a = act.detach().clone().requires_grad_(True)
        leaf = [t.detach().clone().requires_grad_(True) for t in self.p]
        loss = R.cross_entropy(R.bob_forward(a, leaf, SPEC, None), y)
        loss.backward()
        R.sgd_momentum_(self.p, [t.grad for t in leaf], self.opts[slot], 1e-3, 0.9)
        self.slots_seen.append(slot)
        return loss.detach(), a.grad, None


def _batches():
    g = torch.Generator().manual_seed(99)
    x = torch.randint(0, 256, (STEPS, B, 1, 28, 28), generator=g).float()
    y = torch.randint(0, 10, (STEPS, B), generator=g)
    return x, y


def _rows(i):
    """A short batch every third step (the reference's loaders have no drop_last)."""
    return B if i % 3 else B - 1 - (i % 2)


def _run(rank, world, out_dir, port, ragged=False):
    if world > 1:
        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(1)
    x, y = _batches()
    bob = FakeBob() if rank == 0 else None
    mine = {k: FakeAlice(100 + k) for k in range(N_ALICES) if spmd.owner_rank(k, world) == rank}
    losses = {}
    if ragged:
        spmd.run_vanilla_schedule(rank, world, bob, mine, lambda i: (x[i][:_rows(i)], y[i][:_rows(i)]), N_ALICES, TURN, 0,
                                  STEPS, B, torch.device("cpu"), lambda i, l: losses.__setitem__(i, float(l)), rows_fn=_rows)
    else:
        spmd.run_vanilla_schedule(rank, world, bob, mine, lambda i: (x[i], y[i]), N_ALICES, TURN, 0, STEPS, B,
                                  torch.device("cpu"), lambda i, l: losses.__setitem__(i, float(l)))
    out = {f"alice{k}": a.get_weights().numpy() for k, a in mine.items()}
    if bob is not None:
        out.update({f"bob{i}": p.numpy() for i, p in enumerate(bob.p)})
        out["losses"] = np.array([losses[i] for i in range(STEPS)])
        out["slots"] = np.array(bob.slots_seen)
    np.savez(os.path.join(out_dir, f"w{world}_r{rank}.npz"), **out)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def _collect(out_dir, world):
    merged = {}
    for r in range(world):
        with np.load(os.path.join(out_dir, f"w{world}_r{r}.npz")) as f:
            merged.update({k: f[k] for k in f.files})
    return merged


@pytest.mark.parametrize("world", [2, 3])
def test_spmd_protocol_matches_single_process(world):
    with tempfile.TemporaryDirectory() as d:
        _run(0, 1, d, 0)
        ref = _collect(d, 1)
        mp.spawn(_run, args=(world, d, 29700 + world), nprocs=world, join=True)
        got = _collect(d, world)
    assert set(got) == set(ref)
    assert got["slots"].tolist() == [(i // TURN) % N_ALICES for i in range(STEPS)]
    for k in ref:
        assert np.array_equal(got[k], ref[k]), k


def test_spmd_ragged_batches_reach_bob_with_their_row_count():
    """Short batches: Bob must train on exactly the rows Alice sent, not on the stale tail of the
    B-row message buffer (and a batch_fn that disagrees with rows_fn is refused)."""
    with tempfile.TemporaryDirectory() as d:
        _run(0, 1, d, 0, True)
        ref = _collect(d, 1)
        mp.spawn(_run, args=(2, d, 29711, True), nprocs=2, join=True)
        got = _collect(d, 2)
    for k in ref:
        assert np.array_equal(got[k], ref[k]), k
    x, y = _batches()
    with pytest.raises(ValueError, match="rows_fn"):
        spmd.run_vanilla_schedule(0, 1, FakeBob(), {k: FakeAlice(k) for k in range(N_ALICES)},
                                  lambda i: (x[i][:2], y[i][:2]), N_ALICES, TURN, 0, 1, B, torch.device("cpu"))


def test_owner_placement():
    assert [spmd.owner_rank(k, 1) for k in range(4)] == [0, 0, 0, 0]
    assert [spmd.owner_rank(k, 2) for k in range(4)] == [1, 1, 1, 1]
    assert [spmd.owner_rank(k, 8) for k in range(7)] == [1, 2, 3, 4, 5, 6, 7]
    assert [spmd.owner_rank(k, 4) for k in range(4)] == [1, 2, 3, 1]
    msg, act, lab = spmd.pack_message(16, torch.device("cpu"))
    assert act.shape == (16, 5408) and lab.shape == (16,) and lab.dtype == torch.int64
    lab.copy_(torch.arange(16))
    act.fill_(1.5)
    assert lab.tolist() == list(range(16)) and float(msg[0]) == 1.5


# ----------------------------------------------------------------------------- SISA over ranks
class FakeSisaAlice(FakeAlice):
    def __init__(self, seed):
        super().__init__(seed)
        self.adam = R.AdamState()

    def local_step(self, x, y):
        out = R.sisa_alice_step(self.p, x, y)
        R.adam_(self.p, out["alice_grads"], self.adam, 1e-3, weight_decay=1e-5)

    def reset_opt(self):
        self.adam = R.AdamState()


class FakeSisaBob:
    def __init__(self):
        g = torch.Generator().manual_seed(5)
        self.p = [torch.randn(24, 5408, generator=g) * 0.01, torch.zeros(24), torch.randn(10, 24, generator=g) * 0.1,
                  torch.zeros(10)]
        self.adam = R.AdamState()
        self.batches = []

    def server_epoch(self, acts, labels, offsets):
        for s, e in zip(offsets[:-1], offsets[1:]):
            out = R.sisa_server_step(self.p, acts[s:e], labels[s:e], None, SPEC)
            R.adam_(self.p, out["bob_grads"], self.adam, 1e-3, weight_decay=1e-5)
            self.batches.append(e - s)


SISA_SIZES = [10, 7, 9]


def _sisa_data():
    g = torch.Generator().manual_seed(77)
    out = {}
    for k, n in enumerate(SISA_SIZES):
        out[k] = (torch.randint(0, 256, (n, 1, 28, 28), generator=g).float(), torch.randint(8, 10, (n,), generator=g))
    return out


def _run_sisa(rank, world, out_dir, port):
    if world > 1:
        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(1)
    full = _sisa_data()
    mine = {k: FakeSisaAlice(300 + k) for k in range(len(SISA_SIZES)) if spmd.owner_rank(k, world) == rank}
    data = {k: full[k] for k in mine}
    bob = FakeSisaBob() if rank == 0 else None
    cache = spmd.run_sisa_schedule(rank, world, bob, mine, data, SISA_SIZES, B, 1, 2, torch.device("cpu"), unlearn=(0, 9))
    out = {f"alice{k}": a.get_weights().numpy() for k, a in mine.items()}
    if rank == 0:
        out.update({f"bob{i}": p.numpy() for i, p in enumerate(bob.p)})
        out["batches"] = np.array(bob.batches)
        out["cache0_n"] = np.array(cache[0][0].shape[0])
        out["cache0_labels"] = cache[0][1].numpy()
    np.savez(os.path.join(out_dir, f"s{world}_r{rank}.npz"), **out)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sisa_schedule_matches_single_process(world):
    def collect(d, w):
        merged = {}
        for r in range(w):
            with np.load(os.path.join(d, f"s{w}_r{r}.npz")) as f:
                merged.update({k: f[k] for k in f.files})
        return merged
    with tempfile.TemporaryDirectory() as d:
        _run_sisa(0, 1, d, 0)
        ref = collect(d, 1)
        mp.spawn(_run_sisa, args=(world, d, 29720 + world), nprocs=world, join=True)
        got = collect(d, world)
    assert set(got) == set(ref)
    assert (ref["cache0_labels"] != 9).all() and int(ref["cache0_n"]) < SISA_SIZES[0]      # the unlearned shard was re-shipped
    for k in ref:
        assert np.array_equal(got[k], ref[k]), k


# ----------------------------------------------------------------------------- --sisa --concat over ranks
CONCAT_SIZES = [13, 9, 11]           # rows = full batches of the smallest shard = 8 at B = 4
CONCAT_SPEC = R.BobSpec((5408 * 3, 24, 10 * 3), False, 0.0)


class FakeConcatBob:
    """K-head server on the feature-concatenated rows (semantics of data_entities_sisa_concat.py)."""

    def __init__(self):
        g = torch.Generator().manual_seed(6)
        self.p = [torch.randn(24, 5408 * 3, generator=g) * 0.01, torch.zeros(24), torch.randn(30, 24, generator=g) * 0.1,
                  torch.zeros(30)]
        self.adam = R.AdamState()
        self.seen = []

    def server_epoch(self, acts, labels, offsets):
        assert acts.shape[1] == 5408 * 3 and labels.shape[1] == 3
        for s, e in zip(offsets[:-1], offsets[1:]):
            leaf = [t.detach().clone().requires_grad_(True) for t in self.p]
            logits = R.bob_forward(acts[s:e], leaf, CONCAT_SPEC, None)
            R.cross_entropy(logits.reshape(-1, 10), labels[s:e].reshape(-1)).backward()
            R.adam_(self.p, [t.grad for t in leaf], self.adam, 1e-3, weight_decay=1e-5)
            self.seen.append(e - s)


def _concat_data():
    g = torch.Generator().manual_seed(78)
    return {k: (torch.randint(0, 256, (n, 1, 28, 28), generator=g).float(), torch.randint(7, 10, (n,), generator=g))
            for k, n in enumerate(CONCAT_SIZES)}


def _run_concat(rank, world, out_dir, port):
    if world > 1:
        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(1)
    full = _concat_data()
    mine = {k: FakeSisaAlice(400 + k) for k in range(len(CONCAT_SIZES)) if spmd.owner_rank(k, world) == rank}
    data = {k: full[k] for k in mine}
    bob = FakeConcatBob() if rank == 0 else None
    res = spmd.run_sisa_concat_schedule(rank, world, bob, mine, data, CONCAT_SIZES, B, 1, 2, torch.device("cpu"), unlearn=(0, 9))
    out = {f"alice{k}": a.get_weights().numpy() for k, a in mine.items()}
    if rank == 0:
        out.update({f"bob{i}": p.numpy() for i, p in enumerate(bob.p)})
        out["seen"] = np.array(bob.seen)
        out["acts"] = res[0].numpy()
        out["labels"] = res[1].numpy()
    np.savez(os.path.join(out_dir, f"c{world}_r{rank}.npz"), **out)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [3, 4])
def test_sisa_concat_schedule_matches_single_process(world):
    """World 3: two Alices share rank 1; world 4: one Alice per rank (config 4's placement in small).
    The concatenated cache, the batch sequence and every weight must equal the one-process run, and
    the cache must be what the entity module defines: row j = cat_k(act_k[j]), labels [rows, K]."""
    def collect(d, w):
        merged = {}
        for r in range(w):
            with np.load(os.path.join(d, f"c{w}_r{r}.npz")) as f:
                merged.update({k: f[k] for k in f.files})
        return merged
    with tempfile.TemporaryDirectory() as d:
        _run_concat(0, 1, d, 0)
        ref = collect(d, 1)
        mp.spawn(_run_concat, args=(world, d, 29740 + world), nprocs=world, join=True)
        got = collect(d, world)
    assert set(got) == set(ref)
    for k in ref:
        assert np.array_equal(got[k], ref[k]), k
    # 8 rows before the unlearning (2 full batches of the 9-sample shard), then Alice 0 loses her 9s
    full = _concat_data()
    n0 = int((full[0][1] != 9).sum())
    rows_after = min(8, n0 // B * B)
    assert ref["seen"].tolist() == [B] * (2 * 2) + [B] * (2 * rows_after // B)
    assert ref["acts"].shape == (rows_after, 5408 * 3) and ref["labels"].shape == (rows_after, 3)
    assert (ref["labels"][:, 0] != 9).all()
    assert np.array_equal(ref["labels"][:, 1], full[1][1][:rows_after].numpy())
    a1 = FakeSisaAlice(401)                      # Alice 1 never retrains after phase 1: her slot can be recomputed
    for s in range(0, CONCAT_SIZES[1], B):
        a1.local_step(full[1][0][s:s + B], full[1][1][s:s + B])
    assert np.array_equal(ref["acts"][:, 5408:2 * 5408], a1.forward(full[1][0][:rows_after]).numpy())
    assert spmd.concat_rows([16271, 21685, 15329], 16) == 15328
