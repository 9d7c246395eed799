"""One-process-per-GPU execution of the vanilla split protocol (torchrun / torch.distributed).

Placement (BASELINE.json north_star): Bob on rank 0 (GPU 0), Alice k on rank 1 + (k mod (W-1)).
The reference moves cut-layer activations, labels, cut-layer gradients and the Alice->Alice weight
hand-off through torch.distributed.rpc / dist_autograd on CPU tensors
(`data_entities_vanilla.py:63,73,231`); here they are point-to-point `send`/`recv` of device
buffers (NCCL over NVLink; gloo in the CPU tests), enqueued on the compute stream so no host
synchronisation happens inside a step:

    Alice rank                          Bob rank
    act = forward(x)
    send [act | labels]  ------------>  recv
                                        train_step(slot k): fwd, CE, bwd, per-Alice SGD, cut_grad
    recv cut_grad        <------------  send cut_grad
    backward_update(x, cut_grad)

The protocol is strictly sequential by construction (one Alice active at a time; step t+1's
forward needs step t's update: SURVEY.md 8e), so extra GPUs add NVLink hops, not throughput: this
path shards nothing and bench.py reports `"scaling": "strong"`.  Every rank walks the same global
schedule and only acts on its own turns, so no control messages are needed.

The engines are duck-typed (`forward`, `backward_update`, `train_step`, `get_weights`,
`set_weights`) so tests/test_spmd_gloo.py can drive the same code with CPU stand-ins.
"""
import os

import torch
import torch.distributed as dist

D0 = 5408


def owner_rank(alice_index, world):
    """Rank hosting Alice `alice_index` (0-based): everything on rank 0 when there is one rank."""
    return 0 if world == 1 else 1 + alice_index % (world - 1)


def pack_message(B, device):
    """One buffer carries the activations [B,5408] fp32 and the labels [B] int64 (8-byte aligned
    tail), so a step costs one message each way."""
    msg = torch.empty(B * D0 + 2 * B, dtype=torch.float32, device=device)
    return msg, msg[:B * D0].view(B, D0), msg[B * D0:].view(torch.int64)


def run_vanilla_schedule(rank, world, bob, alices, batch_fn, n_alices, turn, first_step, n_steps, B, device,
                         loss_sink=None, link=None, rows_fn=None):
    """Walk steps [first_step, first_step+n_steps) of the round-robin schedule.

    bob: engine or None (ranks != 0); alices: {alice_index: engine} for the Alices this rank hosts;
    batch_fn(i) -> (x, y) device tensors for step i (only called on the owning rank).
    rows_fn(i) -> rows of step i's batch (1..B), evaluated on EVERY rank: the mailboxes and messages
    are B rows deep, and Bob - who never sees Alice's loader - must know how many of them belong to
    a short last batch (the reference's DataLoader has no drop_last).  None = every batch has B rows,
    and a batch_fn that returns anything else is refused.
    link: a peer.PeerLink (CUDA IPC mailboxes: the kernels write straight into the consumer's GPU and
    4-byte flags order the streams) or None for send/recv messages (NCCL; gloo in the CPU tests).

    Calling it again on the same link: inside one call an Alice on another rank is ordered behind her
    predecessor by the weight hand-off in front of her turn.  A call whose FIRST step has no hand-off
    in front of it (step 0, or a step inside a turn) has no such edge to the previous call's last
    Alice: drain first (`torch.cuda.synchronize(); dist.barrier()` on every rank), or her batch lands in
    Bob's mailbox - one buffer, one flag - while Bob still works on the previous call's steps."""
    msg, act_view, lab_view = pack_message(B, device)
    cut = torch.empty(B, D0, dtype=torch.float32, device=device)
    wbuf = torch.empty(320, dtype=torch.float32, device=device)
    for i in range(first_step, first_step + n_steps):
        k = (i // turn) % n_alices
        own = owner_rank(k, world)
        if i % turn == 0 and i > 0:                       # weight hand-off from the previous Alice (:59-64)
            kp = (k - 1) % n_alices
            src = owner_rank(kp, world)
            if src == own:
                if rank == own:
                    alices[k].set_weights(alices[kp].get_weights())
            else:
                if rank == src:
                    dist.send(alices[kp].get_weights().contiguous(), dst=own)
                elif rank == own:
                    dist.recv(wbuf, src=src)
                    alices[k].set_weights(wbuf)
        if link is not None:
            link.next_step()
        n = B if rows_fn is None else int(rows_fn(i))
        if not 1 <= n <= B:
            raise ValueError(f"step {i}: rows_fn gave {n} rows, mailboxes hold 1..{B}")
        host_in = False
        if rank == own or world == 1:
            x, y = batch_fn(i)
            if x.shape[0] != n:
                raise ValueError(f"step {i}: batch_fn returned {x.shape[0]} rows where every rank expects {n} "
                                 "(pass rows_fn for ragged batches)")
            # a batch still in pinned host memory: Alice's conv kernel reads it over PCIe itself and
            # leaves device copies for her backward and for the labels' way to Bob (no copy op, no event)
            host_in = x.device.type == "cpu" and hasattr(alices[k], "forward_host")
        if world == 1:
            if host_in:
                act, x, y = alices[k].forward_host(x, y, i)
            else:
                act = alices[k].forward(x)
            loss, g, _ = bob.train_step(act, y, slot=k)
            alices[k].backward_update(x, g)
        elif link is not None and rank == own:
            if host_in:
                _, x, y = alices[k].forward_host(x, y, i, out=link.alice_targets())
            else:
                alices[k].forward(x, out=link.alice_targets())   # the conv kernel stores into Bob's GPU
            link.alice_publish(y, n)
            alices[k].backward_update(x, link.alice_wait_cut()[:n])
        elif link is not None and rank == 0:
            act, lab = link.bob_wait_batch()
            # ... and Bob's last kernel stores the n cut-gradient rows into Alice's GPU
            loss, g, _ = bob.train_step(act[:n], lab[:n], slot=k, cut_out=link.bob_cut_target(own))
            link.bob_publish_cut(own)
        elif rank == own:
            if host_in:
                _, x, y = alices[k].forward_host(x, y, i, out=act_view[:n])
            else:
                alices[k].forward(x, out=act_view[:n])
            lab_view[:n].copy_(y)
            dist.send(msg, dst=0)
            dist.recv(cut, src=0)
            alices[k].backward_update(x, cut[:n])
        elif rank == 0:
            dist.recv(msg, src=own)
            loss, g, _ = bob.train_step(act_view[:n], lab_view[:n], slot=k)
            if n < B:                                            # the reply is always B rows deep
                cut[:n].copy_(g)
                g = cut
            dist.send(g, dst=own)
        else:
            continue
        if loss_sink is not None and (rank == 0):
            loss_sink(i, loss)


# ------------------------------------------------------------------------------- SISA
def run_sisa_schedule(rank, world, bob, alices, data, sizes, B, local_epochs, server_epochs, device,
                      unlearn=None, timers=None):
    """SISA over ranks (`split_nn.py:74-110` driving `data_entities_vanilla_sisa.py`).

    This is the part of the reference that DOES shard: every Alice owns her model and her data.
      phase 1  each rank trains its Alices locally, no communication (:55-70, :326-334);
      phase 2  each Alice runs her frozen model over her whole shard in one launch and ships the
               [n_k,5408] block + labels to rank 0 in one message each - the bulk transfer that
               replaces a pickled list of per-batch tensors over RPC (:196-211,:279);
      phase 3  rank 0 runs the server epochs over the cached blocks, client by client (:298-313);
      phase 4  (unlearn=(k, omit_label)) Alice k drops the label, retrains with a fresh optimizer,
               re-ships her block; rank 0 replaces that cache entry and trains again (:139-168).
    alices: {k: engine with local_step/forward/reset_opt}; data: {k: (x[n,1,28,28], y[n])} for
    the Alices this rank hosts; sizes[k] = n_k for every Alice (all ranks know the partition).
    Returns rank 0's cache {k: (acts, labels)} (None elsewhere)."""
    n_alices = len(sizes)
    tick = timers if timers is not None else (lambda name: None)

    def train_local(k, x, y):
        for _ in range(local_epochs):
            if hasattr(alices[k], "local_epoch"):
                alices[k].local_epoch(x, y, B)
            else:
                for s in range(0, x.shape[0], B):
                    alices[k].local_step(x[s:s + B], y[s:s + B])

    def ship(k, x, y, cache):
        own = owner_rank(k, world)
        n = x.shape[0] if rank == own else None
        if rank == own:
            acts = alices[k].forward(x)
            if own == 0:
                cache[k] = (acts, y)
            else:
                dist.send(acts, dst=0)
                dist.send(y, dst=0)
        elif rank == 0:
            acts = torch.empty(sizes[k], D0, dtype=torch.float32, device=device)
            lab = torch.empty(sizes[k], dtype=torch.int64, device=device)
            dist.recv(acts, src=own)
            dist.recv(lab, src=own)
            cache[k] = (acts, lab)
        return n

    def server(cache):
        for _ in range(server_epochs):
            for k in range(n_alices):
                acts, lab = cache[k]
                n = acts.shape[0]
                if n:
                    bob.server_epoch(acts, lab, list(range(0, n, B)) + [n])

    tick("start")
    for k in sorted(alices):
        train_local(k, *data[k])
    tick("local_training")
    cache = {} if rank == 0 else None
    for k in range(n_alices):
        ship(k, *(data[k] if k in alices else (None, None)), cache)
    tick("activation_gather")
    if rank == 0:
        server(cache)
    tick("server_epochs")
    if unlearn is not None:
        k, omit = unlearn
        if k in alices:
            x, y = data[k]
            keep = y != omit
            xu, yu = x[keep].contiguous(), y[keep].contiguous()
            alices[k].reset_opt()
            train_local(k, xu, yu)
            data = dict(data)
            data[k] = (xu, yu)
        # every rank must know the new shard size: it is a function of the (shared) partition
        if world > 1:
            n_new = torch.tensor([data[k][0].shape[0] if k in alices else 0], device=device, dtype=torch.int64)
            dist.broadcast(n_new, src=owner_rank(k, world))
            sizes = list(sizes)
            sizes[k] = int(n_new.item())
        tick("unlearn_retrain")
        ship(k, *(data[k] if k in alices else (None, None)), cache)
        if rank == 0:
            server(cache)
        tick("unlearn_server_epochs")
    return cache


def concat_rows(sizes, B):
    """Rows of the --concat server's cache: full batches only, as many as the smallest shard has
    (`data_entities_sisa_concat.py` here: steps/epoch = min_k(number of FULL batches of Alice k))."""
    return min(int(n) // B for n in sizes) * B


def run_sisa_concat_schedule(rank, world, bob, alices, data, sizes, B, local_epochs, server_epochs, device,
                             unlearn=None, timers=None, cache=None):
    """`--sisa --concat` over ranks (BASELINE.json config 4: Bob on rank 0, K Alices on ranks 1..W-1;
    CLI branch `split_nn.py:64-72`, unlearning order `split_nn.py:95-110`, server model
    `models.py:66-82`).  The reference's own concat loop crashes (`data_entities_sisa_concat.py:270-276`),
    so the semantics are the ones `splitlearning_b200/data_entities_sisa_concat.py` defines: row j of
    Bob's cache is cat_k(act_k[j]) -> [rows, 5408*K] with labels [rows, K], rows = full batches of the
    smallest shard, K-head CE.
      phase 1  each rank trains its Alices locally (as --sisa);
      phase 2  Alice k runs her frozen model over her first `rows` samples and fills COLUMN SLOT k of
               Bob's cache: with `cache` (a peer.ConcatCache) her conv kernel stores straight into
               Bob's GPU over NVLink, row stride 5408*K, no staging copy and no torch.cat; without
               it (gloo tests) she sends the [rows,5408] block and rank 0 copies it into the slot;
      phase 3  rank 0 runs the server epochs over the concatenated rows;
      phase 4  (unlearn=(k, omit_label)) Alice k drops the label, retrains with a fresh optimizer and
               rebuilds ONLY her slot (the row count can only shrink); Bob trains again.
    Returns rank 0's (acts [rows,5408K], labels [rows,K]) (None elsewhere)."""
    K = len(sizes)
    tick = timers if timers is not None else (lambda name: None)
    rows0 = concat_rows(sizes, B)
    if rank == 0:
        if cache is not None:
            acts, labels = cache.acts, cache.labels
        else:
            acts = torch.zeros(rows0, K * D0, dtype=torch.float32, device=device)
            labels = torch.zeros(rows0, K, dtype=torch.int64, device=device)
    else:
        acts = labels = None

    def train_local(k, x, y):
        for _ in range(local_epochs):
            if hasattr(alices[k], "local_epoch"):
                alices[k].local_epoch(x, y, B)
            else:
                for s in range(0, x.shape[0], B):
                    alices[k].local_step(x[s:s + B], y[s:s + B])

    def fill_slot(k, rows):
        own = owner_rank(k, world)
        if rank == own:
            x, y = data[k]
            if cache is not None:
                alices[k].forward(x[:rows], out=cache.slot(k), ld=K * D0)
                if own == 0:
                    labels[:rows, k] = y[:rows]
                else:
                    dist.send(y[:rows].contiguous(), dst=0)
            else:
                block = alices[k].forward(x[:rows])
                if own == 0:
                    acts[:rows, k * D0:(k + 1) * D0] = block
                    labels[:rows, k] = y[:rows]
                else:
                    dist.send(block, dst=0)
                    dist.send(y[:rows].contiguous(), dst=0)
        elif rank == 0:
            if cache is None:
                block = torch.empty(rows, D0, dtype=torch.float32, device=device)
                dist.recv(block, src=own)
                acts[:rows, k * D0:(k + 1) * D0] = block
            lab = torch.empty(rows, dtype=torch.int64, device=device)
            dist.recv(lab, src=own)
            labels[:rows, k] = lab

    def fence():
        # the slot writes are plain kernels on the Alice GPUs: Bob may read once they have finished
        if cache is not None:
            torch.cuda.synchronize(device)
        if world > 1:
            dist.barrier()

    def server(rows):
        if rows:
            for _ in range(server_epochs):
                bob.server_epoch(acts[:rows], labels[:rows], list(range(0, rows + 1, B)))

    tick("start")
    for k in sorted(alices):
        train_local(k, *data[k])
    tick("local_training")
    for k in range(K):
        fill_slot(k, rows0)
    fence()
    tick("activation_gather")
    if rank == 0:
        server(rows0)
    tick("server_epochs")
    rows = rows0
    if unlearn is not None:
        k, omit = unlearn
        if k in alices:
            x, y = data[k]
            keep = y != omit
            xu, yu = x[keep].contiguous(), y[keep].contiguous()
            alices[k].reset_opt()
            train_local(k, xu, yu)
            data = dict(data)
            data[k] = (xu, yu)
        if world > 1:
            n_new = torch.tensor([data[k][0].shape[0] if k in alices else 0], device=device, dtype=torch.int64)
            dist.broadcast(n_new, src=owner_rank(k, world))
            n_k = int(n_new.item())
        else:
            n_k = data[k][0].shape[0]
        rows = min(rows0, (n_k // B) * B)
        tick("unlearn_retrain")
        fill_slot(k, rows)
        fence()
        tick("unlearn_gather")
        if rank == 0:
            server(rows)
        tick("unlearn_server_epochs")
    return (acts[:rows], labels[:rows]) if rank == 0 else None


# ------------------------------------------------------------------------------- CUDA engines
class _CudaAlice:
    def __init__(self, device, lr):
        from . import engine as E
        from . import models as M
        self.model = M.model1_sisa(device)
        self.eng = self.model.make_engine(optimizer=E.SN_OPT_SGD_MOMENTUM, lr=lr, momentum=0.9, max_batch=64)
        self.w, self.b = self.model._wb()
        self._mirrors = None

    def forward(self, x, out=None):
        return self.eng.forward(x, out=out)

    def forward_host(self, hx, hy, step, out=None):
        """(act, x_dev, y_dev) for a pinned host batch; the device copies alternate between two pairs."""
        if self._mirrors is None:
            dev, mb = self.eng.device, self.eng.max_batch
            self._mirrors = [(torch.empty(mb, 1, 28, 28, device=dev), torch.empty(mb, dtype=torch.int64, device=dev)) for _ in range(2)]
        n = hx.shape[0]
        xd, yd = self._mirrors[step & 1]
        act = self.eng.forward_host(hx, hy, xd, yd, out=out)
        return act, xd[:n], yd[:n]

    def backward_update(self, x, cut):
        self.eng.backward_update(x, cut)

    def get_weights(self):
        return torch.cat([self.w.data.reshape(-1), self.b.data.reshape(-1)])

    def set_weights(self, flat):
        self.w.data.copy_(flat[:288].view(32, 1, 3, 3))
        self.b.data.copy_(flat[288:320])


class _CudaSisaAlice:
    """SISA Alice: local Adam(wd 1e-5) steps (`data_entities_vanilla_sisa.py:48,62-70`)."""

    def __init__(self, device, lr, B):
        from . import engine as E
        from . import models as M
        self.model = M.model1_sisa(device)
        self.kw = dict(optimizer=E.SN_OPT_ADAM, lr=lr, weight_decay=1e-5, max_batch=B)
        self.eng = self.model.make_engine(**self.kw)
        self.loss = torch.zeros((), device=device)

    def local_step(self, x, y):
        self.eng.local_step(x, y, loss_out=self.loss)

    def local_epoch(self, x, y, B):
        self.eng.local_epoch(x, y, B)

    def forward(self, x, out=None, ld=None):
        return self.eng.forward(x, out=out, ld=ld)

    def reset_opt(self):
        self.eng.reset_opt()


class _CudaSisaBob:
    def __init__(self, device, lr, B):
        from . import engine as E
        from . import models as M
        self.model = M.model2_sisa(device)
        self.eng = self.model.make_engine(optimizer=E.SN_OPT_ADAM, lr=lr, weight_decay=1e-5, n_opt_slots=1, max_batch=B)
        self.offset = 0

    def server_epoch(self, acts, labels, offsets):
        self.eng.server_epoch(acts, labels, offsets, slot=0, masks=None, seed=7, offset=self.offset)
        self.offset += len(offsets)


class _CudaBob:
    def __init__(self, device, lr, n_slots, B):
        from . import engine as E
        from . import models as M
        self.model = M.model2_sisa(device)
        self.eng = self.model.make_engine(optimizer=E.SN_OPT_SGD_MOMENTUM, lr=lr, momentum=0.9, n_opt_slots=n_slots,
                                          max_batch=B)
        self.offset = 0
        self.loss = torch.zeros((), device=device)
        self.loss_host, self.loss_idx = None, 0      # pinned host array: the loss kernel stores there directly

    def train_step(self, act, y, slot, cut_out=None):
        self.offset += 1
        loss = self.loss
        if self.loss_host is not None:
            loss = self.loss_host[self.loss_idx % self.loss_host.shape[0]]
            self.loss_idx += 1
        return self.eng.train_step(act, y, slot=slot, masks=None, seed=1, offset=self.offset, want_cut_grad=True,
                                   loss_out=loss, cut_out=cut_out)


def bench_sisa(rank, world, device, B, n_alices):
    """SISA on the same ranks (BASELINE.json configs[2]/[3] shape: 4 Alices, 55,999 train samples,
    1 local epoch, 1 server epoch, then Alice 1 unlearns label 9 and Bob retrains).  Phase times are
    max over ranks (CUDA events + barriers)."""
    sizes = [16271 * 8 // 10, 21685 * 8 // 10, 15329 * 8 // 10, 16715 * 8 // 10][:n_alices]
    g = torch.Generator().manual_seed(5)
    data, mine = {}, {}
    for k, n in enumerate(sizes):
        x = torch.randint(0, 256, (n, 1, 28, 28), generator=g).float() * (torch.rand(n, 1, 28, 28, generator=g) < 0.4)
        y = torch.randint(0, 10, (n,), generator=g)
        if owner_rank(k, world) == rank:
            data[k] = (x.to(device), y.to(device))
            mine[k] = _CudaSisaAlice(device, 1e-3, B)
    bob = _CudaSisaBob(device, 1e-3, B) if rank == 0 else None
    marks = []
    if world > 1:
        # NCCL opens its point-to-point connections lazily (~100 ms each): do that outside the timed
        # phases (the vanilla schedule above no longer uses NCCL for its messages)
        tiny = torch.zeros(4, device=device)
        for r in range(1, world):
            if rank == 0:
                dist.recv(tiny, src=r)
            elif rank == r:
                dist.send(tiny, dst=0)
        torch.cuda.synchronize()

    def tick(name):
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        ev = torch.cuda.Event(enable_timing=True)
        ev.record()
        marks.append((name, ev))
    run_sisa_schedule(rank, world, bob, mine, data, sizes, B, 1, 1, device, unlearn=(0, 9), timers=tick)
    torch.cuda.synchronize()
    out = {"samples": sum(sizes), "alices": n_alices, "placement": [owner_rank(k, world) for k in range(n_alices)]}
    for (n0, e0), (n1, e1) in zip(marks[:-1], marks[1:]):
        t = torch.tensor([e0.elapsed_time(e1)], device=device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out[n1 + "_ms"] = round(float(t.item()), 3)
    out["gather_bytes"] = sum(sizes[k] * (D0 * 4 + 8) for k in range(n_alices) if owner_rank(k, world) != 0)
    if world > 1 and os.environ.get("SN_NO_TP") is None:
        out["tensor_parallel"] = bench_tp_epoch(rank, world, device, B, sizes)
    return out


class _CudaConcatBob:
    """The --concat server: model2_sisa_concat(K) with K CE heads, Adam wd 1e-5 (`models.py:66-82`)."""

    def __init__(self, device, lr, B, K):
        from . import engine as E
        from . import models as M
        self.model = M.model2_sisa_concat(num_clients=K, device=device)
        self.eng = self.model.make_engine(optimizer=E.SN_OPT_ADAM, lr=lr, weight_decay=1e-5, n_opt_slots=1, max_batch=B)
        self.eng.set_ce_heads(K)
        self.offset = 0
        self.losses = None

    def server_epoch(self, acts, labels, offsets):
        self.eng.server_epoch(acts, labels, offsets, slot=0, masks=None, seed=7, offset=self.offset, loss_per_step=self.losses)
        self.offset += len(offsets)


CONCAT_K = 7
# LDA K=7 client sizes of the 70,000-sample set, 80 % train
CONCAT_SIZES = [n * 8 // 10 for n in (10448, 11355, 10069, 6662, 8509, 9968, 12989)]


def bench_sisa_concat(rank, world, device, B, peak):
    """BASELINE.json configs[3]: `--sisa --concat`, world_size 8 (Bob + 7 Alices), one local epoch,
    one server epoch over the concatenated cache, then Alice 1 unlearns label 9 and Bob retrains
    (`split_nn.py:64-72,95-110`).  At N GPUs the 7 Alices sit on ranks 1..N-1 (all on rank 0 when
    N = 1) and each one's conv kernel fills her column slot of Bob's cache over NVLink.  Phase times
    are max over ranks."""
    from .peer import ConcatCache
    K, sizes = CONCAT_K, CONCAT_SIZES
    g = torch.Generator().manual_seed(9)
    data, mine = {}, {}
    for k, n in enumerate(sizes):
        x = torch.randint(0, 256, (n, 1, 28, 28), generator=g).float() * (torch.rand(n, 1, 28, 28, generator=g) < 0.4)
        y = torch.randint(0, 10, (n,), generator=g)
        if owner_rank(k, world) == rank:
            data[k] = (x.to(device), y.to(device))
            mine[k] = _CudaSisaAlice(device, 1e-3, B)
    torch.manual_seed(3)
    bob = _CudaConcatBob(device, 1e-3, B, K) if rank == 0 else None
    rows = concat_rows(sizes, B)
    cache = ConcatCache(rank, world, device, rows, K)
    if world > 1:
        tiny = torch.zeros(4, device=device)              # open the lazy NCCL point-to-point connections untimed
        for r in range(1, world):
            if rank == 0:
                dist.recv(tiny, src=r)
            elif rank == r:
                dist.send(tiny, dst=0)
    if bob is not None:                                    # warm-up: first-use costs of the big kernels stay out of the phases
        bob.server_epoch(cache.acts[:4 * B], cache.labels[:4 * B], list(range(0, 4 * B + 1, B)))
        for p in bob.model.parameters():
            pass
    torch.cuda.synchronize()
    marks = []

    def tick(name):
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        ev = torch.cuda.Event(enable_timing=True)
        ev.record()
        marks.append((name, ev))
    run_sisa_concat_schedule(rank, world, bob, mine, data, sizes, B, 1, 1, device, unlearn=(0, 9), timers=tick, cache=cache)
    torch.cuda.synchronize()
    P = sum(bob.model.dims[i] * bob.model.dims[i + 1] + bob.model.dims[i + 1] for i in range(3)) if bob is not None else 0
    out = {"alices": K, "placement": [owner_rank(k, world) for k in range(K)], "rows": rows, "steps": rows // B,
           "cache_bytes": rows * K * D0 * 4,
           "semantics": "feature-concat, K CE heads: defined by this build, reference unpinned (its loop crashes)"}
    for (n0, e0), (n1, e1) in zip(marks[:-1], marks[1:]):
        t = torch.tensor([e0.elapsed_time(e1)], device=device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out[n1 + "_ms"] = round(float(t.item()), 3)
    if rank == 0:
        steps = rows // B
        step_bytes = 24.0 * P + 4.0 * B * K * D0
        us = 1e3 * out["server_epochs_ms"] / max(steps, 1)
        out.update({"server_epoch_ms": out["server_epochs_ms"], "us_per_step": round(us, 2), "gather_ms": out["activation_gather_ms"],
                    "params": P, "algorithmic_bytes_per_step": step_bytes, "formula": "24*P + 4*B*K*d0",
                    "achieved_gbs": round(step_bytes / (us * 1e-6) / 1e9, 1),
                    "frac_of_24P": round(step_bytes / (us * 1e-6) / 1e9 / peak, 4)})
    cache.close()
    return out


def bench_tp_epoch(rank, world, device, B, sizes):
    """One SISA server epoch with the first layer's rows split over all ranks (tp.TPServer): the
    activation cache is replicated on every rank (timed separately), then one pass over it, timed
    as the max over ranks after a warm-up pass over the first 64 batches."""
    from . import models as M
    from . import tp
    n = sum(sizes)
    g = torch.Generator(device="cpu").manual_seed(17)
    if rank == 0:
        acts = (torch.rand(n, D0, generator=g) * 30 * (torch.rand(n, D0, generator=g) < 0.6)).to(device)
        labels = torch.randint(0, 10, (n,), generator=g).to(device)
    else:
        acts = torch.empty(n, D0, device=device)
        labels = None
    offs = []
    base = 0
    for nk in sizes:                                   # client by client, as the reference's epoch body
        offs += list(range(base, base + nk, B))
        base += nk
    offs.append(n)

    def timed(fn):
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    rep_ms = timed(lambda: dist.broadcast(acts, src=0))
    torch.manual_seed(1)                               # the same initial server model on every rank
    model = M.model2_sisa(device)
    Ws = [model.fc1.weight.data, model.fc2.weight.data, model.fc3.weight.data]
    bs = [model.fc1.bias.data, model.fc2.bias.data, model.fc3.bias.data]
    srv = tp.TPServer(rank, world, device, Ws, bs, B=B, lr=1e-3, weight_decay=1e-5, dropout_p=0.5)
    srv.server_epoch(acts, labels, offs[:65])          # warm-up
    ms = timed(lambda: srv.server_epoch(acts, labels, offs))
    steps = len(offs) - 1
    srv.close()
    return {"ranks": world, "server_epoch_ms": round(ms, 3), "steps": steps, "us_per_step": round(1e3 * ms / steps, 2),
            "replicate_cache_ms": round(rep_ms, 3), "cache_bytes": n * D0 * 4,
            "layout": "fc1 rows split over all ranks (1/G of W1 + Adam state per GPU); fc2, fc3, loss on rank 0"}


def bench_parity(rank, world, device, B, link):
    """Outside every timed region, on the ranks the bench already has: do the multi-GPU paths
    compute what the one-GPU path computes?  (The driver's GPU test box has one GPU, so the two-GPU
    tests in tests/ are skipped there; this block puts the same comparison into the bench line.)
      vanilla   first 3 losses with Alice on rank 1 and Bob on rank 0 over the peer mailboxes, against
                the same model, data and Philox stream stepped on rank 0 alone;
      tp        first 3 losses of sn_tp_server_epoch over all ranks against sn_bob_server_epoch on
                rank 0, same cached batches and dropout masks.
    Returns {"vanilla_rel": [...], "tp_rel": [...]} on rank 0."""
    from . import engine as E
    from . import models as M
    from . import tp
    steps = 3
    g = torch.Generator().manual_seed(77)
    x = (torch.randint(0, 256, (steps, B, 1, 28, 28), generator=g).float() * (torch.rand(steps, B, 1, 28, 28, generator=g) < 0.4)).to(device)
    y = torch.randint(0, 10, (steps, B), generator=g).to(device)
    out = {}

    def fresh(with_bob, with_alice):
        torch.manual_seed(11)
        bob = _CudaBob(device, 1e-3, 1, B) if with_bob else None
        torch.manual_seed(12)
        alice = _CudaAlice(device, 1e-3) if with_alice else None
        return bob, alice

    def run(r, w, bob, alice, lk):
        got = []
        run_vanilla_schedule(r, w, bob, {0: alice} if alice is not None else {}, lambda i: (x[i], y[i]), 1, 64, 0, steps, B,
                             device, lambda i, l: got.append(l.clone()), lk)
        torch.cuda.synchronize()
        return [float(v.item()) for v in got]

    if link is not None:
        bob, alice = fresh(rank == 0, rank == owner_rank(0, world))
        over_link = run(rank, world, bob, alice, link)
        dist.barrier()
        if rank == 0:
            bob1, alice1 = fresh(True, True)
            alone = run(0, 1, bob1, alice1, None)
            out["vanilla_losses"] = over_link
            out["vanilla_rel"] = [abs(a - b) / max(abs(b), 1e-30) for a, b in zip(over_link, alone)]

    # tensor-parallel server against the unsharded epoch
    torch.manual_seed(5)
    model = M.model2_sisa(device)
    Ws = [model.fc1.weight.data, model.fc2.weight.data, model.fc3.weight.data]
    bs = [model.fc1.bias.data, model.fc2.bias.data, model.fc3.bias.data]
    ref_W, ref_b = [w.clone() for w in Ws], [b.clone() for b in bs]
    n = steps * B
    acts = (torch.rand(n, D0, generator=g) * 30 * (torch.rand(n, D0, generator=g) < 0.6)).to(device)
    labels = torch.randint(0, 10, (n,), generator=g).to(device)
    m1 = (torch.rand(n, 5000, generator=g) < 0.5).to(torch.uint8).to(device)
    m2 = (torch.rand(n, 1000, generator=g) < 0.5).to(torch.uint8).to(device)
    offs = list(range(0, n + 1, B))
    srv = tp.TPServer(rank, world, device, Ws, bs, B=B, lr=1e-3, weight_decay=1e-5, dropout_p=0.5)
    l_tp = torch.zeros(steps, device=device)
    srv.server_epoch(acts, labels, offs, masks=[m1, m2], loss_per_step=l_tp if rank == 0 else None)
    torch.cuda.synchronize()
    srv.close()
    if rank == 0:
        bob = E.BobEngine([D0, 5000, 1000, 100], ref_W, ref_b, optimizer=E.SN_OPT_ADAM, lr=1e-3, weight_decay=1e-5,
                          n_opt_slots=1, max_batch=B)
        l_ref = torch.zeros(steps, device=device)
        bob.server_epoch(acts, labels, offs, masks=[m1, m2], loss_per_step=l_ref)
        torch.cuda.synchronize()
        out["tp_losses"] = [float(v) for v in l_tp.tolist()]
        out["tp_rel"] = [abs(a - b) / max(abs(b), 1e-30) for a, b in zip(l_tp.tolist(), l_ref.tolist())]
        out["ok"] = bool(all(v <= 1e-4 for v in out["tp_rel"]) and all(v <= 1e-4 for v in out.get("vanilla_rel", [0.0])))
        bob.close()
    return out


def bench_vanilla(args, B, n_alices, turn, config, peak):
    """bench.py's N>1 arm: barrier + synchronize on both sides, CUDA events on every rank, max over
    ranks; rank 0 prints the JSON line."""
    import json
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    device = torch.device("cuda", local)
    torch.cuda.set_device(device)
    os.environ.setdefault("TORCH_NCCL_SHOW_EAGER_INIT_P2P_SERIALIZATION_WARNING", "false")
    dist.init_process_group("nccl", device_id=device)
    from . import _lib
    lib = _lib.load()
    torch.manual_seed(0)
    bob = _CudaBob(device, 1e-3, n_alices, B) if rank == 0 else None
    mine = {k: _CudaAlice(device, 1e-3) for k in range(n_alices) if owner_rank(k, world) == rank}
    nb = 256
    g = torch.Generator().manual_seed(1234)
    hx = (torch.randint(0, 256, (nb, B, 1, 28, 28), generator=g).float() * (torch.rand(nb, B, 1, 28, 28, generator=g) < 0.4))
    hy = torch.randint(0, 10, (nb, B), generator=g)
    xs, ys = (hx.to(device), hy.to(device)) if mine else (None, None)
    K, W = args.steps, args.warmup
    # NCCL opens point-to-point connections lazily (~100 ms each): open the ones the Alice->Alice
    # weight hand-off will use before anything is timed
    tiny = torch.zeros(4, device=device)
    for k in range(n_alices):
        src, dst = owner_rank(k, world), owner_rank((k + 1) % n_alices, world)
        if src != dst:
            if rank == src:
                dist.send(tiny, dst=dst)
            elif rank == dst:
                dist.recv(tiny, src=src)
    torch.cuda.synchronize()
    # transport: peer-memory mailboxes (default) or NCCL send/recv (SN_LINK=nccl, for A/B timing)
    link = None
    mode = os.environ.get("SN_LINK", "peer")          # peer | peer-stream | nccl
    if mode != "nccl":
        from .peer import PeerLink
        link = PeerLink(rank, world, device, B, flags="stream" if mode == "peer-stream" else "kernel")
    config = dict(config, transport={"peer": "peer-memory mailboxes over NVLink (CUDA IPC), gate-kernel flags",
                                     "peer-stream": "peer-memory mailboxes over NVLink (CUDA IPC), stream-ordered flags",
                                     "nccl": "NCCL send/recv"}.get(mode, mode))

    def timed(batch_fn, first, sink=None):
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        run_vanilla_schedule(rank, world, bob, mine, batch_fn, n_alices, turn, first, K, B, device, sink, link)
        e1.record()
        torch.cuda.synchronize()
        dist.barrier()
        t = torch.tensor([e0.elapsed_time(e1)], device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # pre-heat (untimed; bench.PREHEAT, disclosed in config) + warm-up with the shortest turn that still
    # fits: every Alice (so every rank: kernels are loaded lazily, per process, on first launch) and every
    # hand-off route has run before the timed steps
    import bench as bench_mod
    run_vanilla_schedule(rank, world, bob, mine, lambda i: (xs[i % nb], ys[i % nb]), n_alices, 1, 0, bench_mod.PREHEAT, B,
                         device, None, link)
    # A schedule that starts again at step 0 has no hand-off in front of its first step, i.e. nothing that
    # orders its first Alice behind the previous schedule's LAST Alice on another rank: she would store
    # her batch into Bob's one mailbox (and raise its flag) while Bob still works through the others'
    # steps.  Drain every rank before restarting.  (timed() does the same in front of the timed steps.)
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    run_vanilla_schedule(rank, world, bob, mine, lambda i: (xs[i % nb], ys[i % nb]), n_alices,
                         max(1, min(turn, W // (n_alices + 1))), 0, W, B, device, None, link)
    l0 = lib.sn_launch_count()
    with bench_mod.ClockSampler(local) as clk:
        ms = timed(lambda i: (xs[i % nb], ys[i % nb]), W)
    launches = torch.tensor([lib.sn_launch_count() - l0], device=device, dtype=torch.int64)
    dist.all_reduce(launches)

    # end to end: inputs start in pinned host memory on the Alice ranks and are copied one owned step
    # ahead on a side stream (io.HostBatchPrefetcher); Bob's loss kernel stores each step's loss
    # straight into pinned host memory on rank 0 (4 B over PCIe, no D2H op in the stream)
    from .io import HostBatchPrefetcher
    phx, phy = (hx.pin_memory(), hy.pin_memory()) if mine else (None, None)
    hloss = torch.zeros(K + 1, dtype=torch.float32).pin_memory()
    pf = HostBatchPrefetcher(device, B) if mine else None
    first, last = W + K, W + 2 * K
    owned = [i for i in range(first, last) if owner_rank((i // turn) % n_alices, world) == rank]
    pos = {i: j for j, i in enumerate(owned)}

    zero_copy = os.environ.get("SN_E2E", "zerocopy") != "copy"

    def host_batch(i):
        if zero_copy:                                     # pinned tensors straight into the schedule
            return phx[i % nb], phy[i % nb]
        j = pos[i]
        if j == 0:
            pf.submit(0, phx[i % nb], phy[i % nb])
        else:
            pf.release((j - 1) % 2)                       # the previous owned step's kernels are all enqueued
        if j + 1 < len(owned):
            nxt = owned[j + 1]
            pf.submit((j + 1) % 2, phx[nxt % nb], phy[nxt % nb])
        return pf.acquire(j % 2)

    if bob is not None:
        bob.loss_host, bob.loss_idx = hloss, 0
    sink = None
    e2e_ms = timed(host_batch, W + K, sink)
    if bob is not None:
        bob.loss_host = None
    sisa = bench_sisa(rank, world, device, B, n_alices)
    sisa_concat = bench_sisa_concat(rank, world, device, B, peak[0])
    parity = bench_parity(rank, world, device, B, link)

    if rank == 0:
        value = K * B / (ms * 1e-3)
        step_bytes = 16.0 * bench_mod.P_BOB + 8.0 * B * D0
        line = {
            "metric": "split_nn_train_samples_per_sec", "value": round(value, 1), "unit": "samples/s", "n_gpus": world,
            "steps": K, "warmup": W, "ms_per_step": round(ms / K, 5), "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config, "clocks": clk.summary(),
            "gpu_launches": int(launches.item()),
            "e2e": {"value": round(K * B / (e2e_ms * 1e-3), 1), "unit": "samples/s",
                    "h2d_bytes_per_step": B * 784 * 4 + B * 8, "d2h_bytes_per_step": 4, "ms_per_step": round(e2e_ms / K, 5)},
            "roofline": {"bound": "hbm", "kernel": "whole step on Bob's GPU (sequential protocol: one GPU streams weights at a time)",
                         "achieved": round(step_bytes / (ms * 1e-3 / K) / 1e9, 1), "peak": peak[0], "unit": "GB/s",
                         "frac": round(step_bytes / (ms * 1e-3 / K) / 1e9 / peak[0], 4), "traffic": None,
                         "peak_source": peak[1]},
            "nvlink_bytes_per_step": 2 * B * D0 * 4 + B * 8,
            "sisa": sisa, "sisa_concat": sisa_concat, "parity": parity,
            "note": "vanilla split learning is a sequential chain (SURVEY.md 8e): extra GPUs host Alices and add "
                    "two NVLink messages per step; they cannot add throughput without changing results",
        }
        print(json.dumps(line))
    if link is not None:
        link.close()
    dist.barrier()
    dist.destroy_process_group()
