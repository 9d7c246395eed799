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
                                                          gradient dH1 is pushed, column range by column
                                                          range, into each rank's own mailbox, and each
                                                          rank's dh_ready flag is set
    wait dh_ready (own memory); slice backward: gates
    dH1[:, rows_r] with its own activation, dW + Adam in
    one traversal, fused with the next batch's forward

(SN_TP_EARLY_CUT=1 hands dH1 out before fc2's own update - a read-only pass over fc2, then the update -
and `row_split(head_rows=...)` can give rank 0 fewer rows, or none: both built and measured, neither a
win on this box; see `row_split` and profiles/r2h_tp_trace_2gpu.txt, r2i_tp_trace_8gpu.txt.)

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


def row_split(n1, K, head_first_layer_elems, world, head_rows=None, min_rows=64):
    """Row ranges [bounds[r], bounds[r+1]) of the sharded first layer: rank 0 holds `head_rows` rows
    fewer than the others (none at all when that leaves it under `min_rows`), ranks 1.. equal ranges.
    Default head_rows = 0, an even split.  Rank 0 also runs the head, whose own update traversal (fc2:
    as many bytes as head_first_layer_elems / K rows of fc1) overlaps the shards' since the cut gradient
    goes out first - but the other ranks pay about the same again in NVLink stores, flag launches and
    flag latency before the head can start (profiles/r2h_tp_trace_2gpu.txt: rank 0 18 us of fc2 update
    against 8 + 6 + ~6 us on rank 1), so taking rows away from rank 0 made the 2-GPU epoch SLOWER
    (505 -> 532 ms with the late hand-out, and no better with the early one)."""
    if world == 1:
        return [0, n1]
    if head_rows is None:
        head_rows = 0.0
    other = (n1 + head_rows) / world
    r0 = (int(other - head_rows) + 7) // 8 * 8
    if r0 < min_rows:
        r0 = 0
    r0 = min(r0, n1)
    per = -(-(n1 - r0) // (world - 1))
    bounds = [0, r0]
    for r in range(1, world):
        bounds.append(min(n1, r0 + r * per))
    return bounds


class TPServer:
    """Collective object: construct on every rank with the SAME full weights (e.g. a model built
    under the same seed, or broadcast from rank 0).  `weights`/`biases`: fc1, fc2, fc3 tensors on this
    rank's device; rank r trains rows [bounds[r], bounds[r+1]) of fc1 in place (`row_split`: an even
    split by default), rank 0 also fc2 and fc3."""

    def __init__(self, rank, world, device, weights, biases, B=16, lr=1e-3, weight_decay=1e-5, dropout_p=0.5,
                 optimizer=E.SN_OPT_ADAM, momentum=0.0):
        self.lib = _lib.load()
        self.rank, self.world, self.device, self.B = rank, world, device, B
        W1, b1 = weights[0], biases[0]
        n1, K = W1.shape
        self.n1 = n1
        self.bounds = row_split(n1, K, weights[1].shape[0] * weights[1].shape[1], world)
        self.lo, self.hi = self.bounds[rank], self.bounds[rank + 1]
        kw = dict(optimizer=optimizer, lr=lr, momentum=momentum, weight_decay=weight_decay)
        self.W1s, self.b1s = W1[self.lo:self.hi], b1[self.lo:self.hi]          # contiguous row slices (views)
        self.shard = None
        if self.hi > self.lo:
            self.shard = E.ShardEngine(self.W1s, self.b1s, n1, self.lo, relu=True, dropout_p=dropout_p, **kw)
        self.head = None
        if rank == 0:
            dims = [n1] + [w.shape[0] for w in weights[1:]]
            self.head = E.BobEngine(dims, list(weights[1:]), list(biases[1:]), dropout_p=dropout_p, n_opt_slots=1, max_batch=B,
                                    **kw)
        # every rank owns a mailbox: H1 [B,n1] | dH1 [B,n1] | h_ready[world] (128 B apart) | dh_ready.
        # H1 and h_ready are used on the head only (the shards store their slices there); dH1 and
        # dh_ready on every rank (the head pushes rank r's column range and flag into rank r's).
        self.off_h, self.off_dh = 0, B * n1 * 4
        self.off_flags = 2 * B * n1 * 4
        self.off_dh_ready = self.off_flags + 128 * world
        nbytes = self.off_dh_ready + 128
        handle = (C.c_ubyte * 64)()
        base = C.c_void_p()
        with torch.cuda.device(device):
            check(self.lib.sn_ipc_alloc(nbytes, C.byref(base), handle))
        self.base = base.value
        self.peer = {rank: self.base}
        if world > 1:
            mine = torch.tensor(list(handle), dtype=torch.uint8, device=device)
            every = [torch.empty(64, dtype=torch.uint8, device=device) for _ in range(world)]
            dist.all_gather(every, mine)
            for r in (range(1, world) if rank == 0 else [0]):
                raw = (C.c_ubyte * 64)(*every[r].cpu().tolist())
                p = C.c_void_p()
                with torch.cuda.device(device):
                    check(self.lib.sn_ipc_open(raw, C.byref(p)))
                self.peer[r] = p.value
        self.head_base = self.peer[0]
        self.dh_table = self.flag_table = None
        if rank == 0:
            with torch.cuda.device(device):
                self.H1_t = torch.as_tensor(_View(self.base + self.off_h, B * n1, "<f4"), device=device).view(B, n1)
            if world > 1:       # device arrays of addresses for the head's cut-gradient push and its signal
                self.dh_table = torch.tensor([self.peer[r] + self.off_dh for r in range(world)], dtype=torch.int64, device=device)
                self.flag_table = torch.tensor([self.peer[r] + self.off_dh_ready for r in range(1, world)], dtype=torch.int64,
                                               device=device)
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
        nb = len(offsets) - 1
        if nb <= 0:
            return
        offs = (C.c_int64 * (nb + 1))(*[int(o) for o in offsets])
        marr, keep = None, None
        if masks is not None:
            keep = [m if m.dtype == torch.uint8 else m.to(torch.uint8) for m in masks]
            if self.rank != 0:
                keep = keep[:1]
            keep = [m.contiguous() for m in keep]
            self._keep = keep                                  # alive until the queued kernels have run
            marr = _lib.ptr_array(keep)
        # the loop itself runs in the library (sn_tp_server_epoch): from Python the six calls per step
        # made rank 0 host-bound (165 us/step at two GPUs)
        check(self.lib.sn_tp_server_epoch(self.shard.h if self.shard is not None else None,
                                          self.head.h if self.head is not None else None, self.rank, self.world,
                                          _lib.ptr(acts), _lib.ptr(labels) if self.rank == 0 else None, offs, nb, marr,
                                          int(seed), int(self.offset), C.c_void_p(self.head_base + self.off_h),
                                          C.c_void_p(self.base + self.off_dh), C.c_void_p(self.head_base + self.off_flags),
                                          C.c_void_p(self.base + self.off_dh_ready), _lib.ptr(self.dh_table),
                                          _lib.ptr(self.flag_table), self.seq, _lib.ptr(loss_per_step), self._stream()))
        self.seq += nb
        self.offset += nb

    def gather_fc1(self, W1_full, b1_full):
        """Collect the trained slices into rank 0's full fc1 tensors (state_dict interchange)."""
        if self.world == 1:
            return
        for r in range(1, self.world):
            lo, hi = self.bounds[r], self.bounds[r + 1]
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
        if self.shard is not None:
            self.shard.close()
        if self.head is not None:
            self.head.close()
        with torch.cuda.device(self.device):
            for r, p in self.peer.items():
                if r != self.rank:
                    self.lib.sn_ipc_close(C.c_void_p(p))
            if self.base:
                self.lib.sn_ipc_free(C.c_void_p(self.base))
        self.peer = {}
        self.base = 0
