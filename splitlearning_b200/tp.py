"""Tensor-parallel SISA server epochs over the ranks of one box.

The server step at batch 16 is a weight stream, and the first layer (5408 -> 5000) is 84 % of its
bytes: 27 M weights plus two Adam moments, read and written every step.  Here its ROWS are split
over the G ranks - each GPU streams 1/G of W1 and of its optimizer state - while the small tail
(fc2, fc3, loss: 5 M weights) stays on rank 0, the "head".  Per step:

    every rank r                                          rank 0 (head)
    slice forward (or just its epilogue, when the
    previous traversal already multiplied this batch
    out) -> stores H1[:, rows_r] straight into the
    head's mailbox over NVLink; signal h_ready[r]
                                                          wait for all h_ready; ordinary 2-layer Bob step
                                                          on H1 (fc2, fc3, CE, backward, Adam); its cut
                                                          gradient dH1 lands in the mailbox; signal dh_ready
    wait dh_ready (polled over NVLink); slice backward:
    reads dH1[:, rows_r] from the head's mailbox, gates
    it with its own activation, dW + Adam in one
    traversal, fused with the next batch's forward

The cached activations are replicated on every rank (1.2 GB); labels live on the head.  The dropout
decision of every element is that of the unsharded layer (global element index), so the sharded run
computes the same function as `sn_bob_server_epoch` on one GPU - same per-row arithmetic, different
grouping of the first layer's partial sums (~1e-6).  Reference semantics: one pass of
`bob.train_and_backward`'s epoch body (`data_entities_vanilla_sisa.py:298-313`) per call.
"""
import ctypes as C

import torch
import torch.distributed as dist

from . import _lib
from . import engine as E
from ._lib import check
from .peer import _Raw, _View

D0 = 5408


class TPServer:
    """Collective object: construct on every rank with the SAME full weights (e.g. a model built
    under the same seed, or broadcast from rank 0).  `weights`/`biases`: fc1, fc2, fc3 tensors on this
    rank's device; rank r trains rows [r*n1/G, (r+1)*n1/G) of fc1 in place, rank 0 also fc2, fc3."""

    def __init__(self, rank, world, device, weights, biases, B=16, lr=1e-3, weight_decay=1e-5, dropout_p=0.5,
                 optimizer=E.SN_OPT_ADAM, momentum=0.0):
        self.lib = _lib.load()
        self.rank, self.world, self.device, self.B = rank, world, device, B
        W1, b1 = weights[0], biases[0]
        n1, K = W1.shape
        self.n1 = n1
        per = (n1 + world - 1) // world
        self.lo, self.hi = min(n1, rank * per), min(n1, (rank + 1) * per)
        kw = dict(optimizer=optimizer, lr=lr, momentum=momentum, weight_decay=weight_decay)
        self.W1s, self.b1s = W1[self.lo:self.hi], b1[self.lo:self.hi]          # contiguous row slices (views)
        self.shard = E.ShardEngine(self.W1s, self.b1s, n1, self.lo, relu=True, dropout_p=dropout_p, **kw)
        self.head = None
        if rank == 0:
            dims = [n1] + [w.shape[0] for w in weights[1:]]
            self.head = E.BobEngine(dims, list(weights[1:]), list(biases[1:]), dropout_p=dropout_p, n_opt_slots=1, max_batch=B,
                                    **kw)
        # mailbox on the head: H1 [B,n1] | dH1 [B,n1] | h_ready[world] (128 B apart) | dh_ready
        self.off_h, self.off_dh = 0, B * n1 * 4
        self.off_flags = 2 * B * n1 * 4
        self.off_dh_ready = self.off_flags + 128 * world
        nbytes = self.off_dh_ready + 128
        handle = (C.c_ubyte * 64)()
        base = C.c_void_p()
        if rank == 0:
            with torch.cuda.device(device):
                check(self.lib.sn_ipc_alloc(nbytes, C.byref(base), handle))
        t = torch.tensor(list(handle), dtype=torch.uint8, device=device)
        if world > 1:
            dist.broadcast(t, src=0)
        if rank == 0:
            self.base, self.mapped = base.value, False
        else:
            raw = (C.c_ubyte * 64)(*t.cpu().tolist())
            with torch.cuda.device(device):
                check(self.lib.sn_ipc_open(raw, C.byref(base)))
            self.base, self.mapped = base.value, True
        self.H1 = _Raw(self.base + self.off_h, (B, n1), device)
        self.dH1 = _Raw(self.base + self.off_dh, (B, n1), device)
        if rank == 0:
            with torch.cuda.device(device):
                self.H1_t = torch.as_tensor(_View(self.base + self.off_h, B * n1, "<f4"), device=device).view(B, n1)
                self.dH1_t = torch.as_tensor(_View(self.base + self.off_dh, B * n1, "<f4"), device=device).view(B, n1)
        self.seq = 0
        self.offset = 0
        self.loss = torch.zeros((), device=device)
        if world > 1:
            dist.barrier()

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def server_epoch(self, acts, labels, offsets, masks=None, seed=7, loss_per_step=None):
        """One pass over the cached batches acts[offsets[i]:offsets[i+1]] (replicated on every
        rank); labels only on rank 0.  masks: optional [m1 (n,n1), m2 (n,n2)] keep bytes for parity
        runs (m1 on every rank, m2 on rank 0); otherwise Philox keyed by (seed, step)."""
        lib, st = self.lib, self._stream
        nb = len(offsets) - 1
        for i in range(nb):
            s, e = offsets[i], offsets[i + 1]
            X = acts[s:e]
            Xn = acts[offsets[i + 1]:offsets[i + 2]] if i + 1 < nb else None
            self.seq += 1
            self.offset += 1
            m1 = masks[0][s:e] if masks is not None else None
            self.shard.forward(X, self.H1, self.n1, train=True, mask=m1, seed=seed, offset=self.offset)
            if self.rank != 0:
                check(lib.sn_flag_signal(C.c_void_p(self.base + self.off_flags + 128 * self.rank), self.seq, None, None, 0, st()))
                check(lib.sn_flag_wait(C.c_void_p(self.base + self.off_dh_ready), self.seq, st()))
            else:
                if self.world > 1:
                    check(lib.sn_flag_wait_many(C.c_void_p(self.base + self.off_flags + 128), self.world - 1, 128, self.seq, st()))
                n = e - s
                loss = loss_per_step[i] if loss_per_step is not None else self.loss
                self.head.train_step(self.H1_t[:n], labels[s:e], slot=0, masks=None if masks is None else [masks[1][s:e]],
                                     seed=seed + 1, offset=self.offset, loss_out=loss, cut_out=self.dH1_t[:n])
                if self.world > 1:
                    check(lib.sn_flag_signal(C.c_void_p(self.base + self.off_dh_ready), self.seq, None, None, 0, st()))
            self.shard.backward_update(X, self.dH1, self.n1, X_next=Xn)

    def gather_fc1(self, W1_full, b1_full):
        """Collect the trained slices into rank 0's full fc1 tensors (state_dict interchange)."""
        if self.world == 1:
            return
        per = (self.n1 + self.world - 1) // self.world
        for r in range(1, self.world):
            lo, hi = min(self.n1, r * per), min(self.n1, (r + 1) * per)
            if self.rank == r:
                dist.send(self.W1s.contiguous(), dst=0)
                dist.send(self.b1s.contiguous(), dst=0)
            elif self.rank == 0:
                dist.recv(W1_full[lo:hi], src=r)
                dist.recv(b1_full[lo:hi], src=r)

    def close(self):
        torch.cuda.synchronize(self.device)
        if self.world > 1:
            dist.barrier()
        self.shard.close()
        if self.head is not None:
            self.head.close()
        with torch.cuda.device(self.device):
            if self.mapped:
                self.lib.sn_ipc_close(C.c_void_p(self.base))
            elif self.base:
                self.lib.sn_ipc_free(C.c_void_p(self.base))
        self.base = 0
