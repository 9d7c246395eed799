"""Peer-memory transport for the one-process-per-GPU split protocol (NVLink, CUDA IPC).

The reference carries each step's cut-layer activations + labels to Bob and the cut-layer gradient
back through `rpc_sync` / `dist_autograd` on CPU tensors (`data_entities_vanilla.py:63,73,231`).
`spmd.py`'s first transport replaced that with one NCCL send/recv pair each way; at batch 16 those
two messages cost ~40 us of a 200 us step although they move 0.7 MB.  Here nothing is "sent":

    Alice rank (GPU a)                                 Bob rank (GPU 0)
    conv kernel writes act   ----- NVLink stores ----> Bob's mailbox.act
    labels                   ----- peer memcpy ------> Bob's mailbox.labels
    signal(Bob.flag_act, s)                            wait(flag_act >= s)
                                                       train step reads its own mailbox;
    Alice's mailbox.cut      <---- NVLink stores ----- its last kernel writes cut_grad
    wait(flag_cut >= s)                                signal(Alice.flag_cut, s)
    conv backward reads its own mailbox

Every rank owns one mailbox (`sn_ipc_alloc`), maps the peers' (`sn_ipc_open`), and orders the two
streams with 4-byte flags: by default one-CTA gate kernels (`sn_flag_signal/wait`: the waiter polls
its own GPU's memory, ~3 us per hop, and the kernel queued behind it overlaps its prologue with the
wait), or stream-ordered value ops (`sn_stream_signal/wait`, ~20 us per hop, for comparison).  No
host synchronisation, no collective.  The step counter `s` is the number of
steps the link has carried, identical on all ranks because every rank walks the whole schedule.

Measured at 2 GPUs (us per vanilla step; 162 on one GPU): NCCL send/recv 199, stream-ordered flags
202, gate kernels 178-181.  A third form - the flags folded into the step's own kernels (producers
signalling from their last CTA after fence.sys, consumers polling with ld.acquire.sys in every CTA) -
was built, matched bit for bit, and ran at 186: slower than the gates (system-scope fences in
hundreds of CTAs, and no prologue overlap), so it was dropped.
"""
import ctypes as C

import torch
import torch.distributed as dist

from . import _lib
from ._lib import check

D0 = 5408


class _Raw:
    """A device address with just enough of the tensor interface for `engine` (`data_ptr`, dtype,
    contiguity, the LOCAL device for stream selection): peer memory must not become a torch tensor -
    torch would place it on the owning device and synchronise across devices on every use."""

    is_cuda = True
    dtype = torch.float32

    def __init__(self, address, shape, device):
        self._a, self.shape, self.device = address, tuple(shape), device

    def data_ptr(self):
        return self._a

    def is_contiguous(self):
        return True


class _View:
    """torch view of LOCAL library-owned memory through __cuda_array_interface__."""

    def __init__(self, address, count, typestr):
        self.__cuda_array_interface__ = {"shape": (count,), "typestr": typestr, "data": (address, False), "version": 2}


class Mailbox:
    """Layout of one rank's mailbox: act [B,5408] f32 | cut [B,5408] f32 | labels [B] i64 | flags."""

    def __init__(self, B):
        self.B = B
        self.off_act = 0
        self.off_cut = B * D0 * 4
        self.off_lab = 2 * B * D0 * 4
        self.off_flag_act = self.off_lab + B * 8
        self.off_flag_cut = self.off_flag_act + 128
        self.bytes = self.off_flag_cut + 128


class PeerLink:
    """The vanilla schedule's transport over peer memory (see module docstring).  Collective set-up:
    every rank must construct it."""

    def __init__(self, rank, world, device, B, flags="kernel"):
        """flags: "kernel" = one-CTA gate kernels (sn_flag_signal/wait, ~3 us per hop), "stream" =
        stream-ordered value ops (sn_stream_signal/wait, ~20 us per hop; kept for A/B timing)."""
        self.lib = _lib.load()
        self.kernel_flags = flags == "kernel"
        self.rank, self.world, self.device, self.B = rank, world, device, B
        self.box = Mailbox(B)
        self.seq = 0
        base, handle = C.c_void_p(), (C.c_ubyte * 64)()
        with torch.cuda.device(device):
            check(self.lib.sn_ipc_alloc(self.box.bytes, C.byref(base), handle))
        self.base = base.value
        mine = torch.tensor(list(handle), dtype=torch.uint8, device=device)
        every = [torch.empty(64, dtype=torch.uint8, device=device) for _ in range(world)]
        dist.all_gather(every, mine)
        self.peer = {}
        # Bob maps every Alice rank's mailbox, Alice ranks map Bob's
        wanted = range(1, world) if rank == 0 else [0]
        for r in wanted:
            raw = (C.c_ubyte * 64)(*every[r].cpu().tolist())
            p = C.c_void_p()
            with torch.cuda.device(device):
                check(self.lib.sn_ipc_open(raw, C.byref(p)))
            self.peer[r] = p.value
        dist.barrier()
        with torch.cuda.device(device):
            self.act = torch.as_tensor(_View(self.base + self.box.off_act, B * D0, "<f4"), device=device).view(B, D0)
            self.cut = torch.as_tensor(_View(self.base + self.box.off_cut, B * D0, "<f4"), device=device).view(B, D0)
            self.labels = torch.as_tensor(_View(self.base + self.box.off_lab, B, "<i8"), device=device)

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def next_step(self):
        """Called once per schedule step on EVERY rank (keeps the flag values in lock step)."""
        self.seq += 1

    # ---- Alice side
    def alice_targets(self):
        """Where the conv kernel writes: Bob's mailbox."""
        return _Raw(self.peer[0] + self.box.off_act, (self.B, D0), self.device)

    def alice_publish(self, y, n):
        bob = self.peer[0]
        if self.kernel_flags:          # the gate kernel carries the labels too
            check(self.lib.sn_flag_signal(C.c_void_p(bob + self.box.off_flag_act), self.seq, C.c_void_p(bob + self.box.off_lab),
                                          C.c_void_p(y.data_ptr()), n * 8, self._stream()))
            return
        check(self.lib.sn_p2p_copy(C.c_void_p(bob + self.box.off_lab), 0, C.c_void_p(y.data_ptr()), self.device.index, n * 8,
                                   self._stream()))
        check(self.lib.sn_stream_signal(C.c_void_p(bob + self.box.off_flag_act), self.seq, self._stream()))

    def _wait(self, flag):
        fn = self.lib.sn_flag_wait if self.kernel_flags else self.lib.sn_stream_wait
        check(fn(C.c_void_p(flag), self.seq, self._stream()))

    def alice_wait_cut(self):
        self._wait(self.base + self.box.off_flag_cut)
        return self.cut

    # ---- Bob side
    def bob_wait_batch(self):
        self._wait(self.base + self.box.off_flag_act)
        return self.act, self.labels

    def bob_cut_target(self, alice_rank):
        return _Raw(self.peer[alice_rank] + self.box.off_cut, (self.B, D0), self.device)

    def bob_publish_cut(self, alice_rank):
        flag = C.c_void_p(self.peer[alice_rank] + self.box.off_flag_cut)
        if self.kernel_flags:
            check(self.lib.sn_flag_signal(flag, self.seq, None, None, 0, self._stream()))
        else:
            check(self.lib.sn_stream_signal(flag, self.seq, self._stream()))

    def close(self):
        torch.cuda.synchronize(self.device)
        if dist.is_initialized():
            dist.barrier()
        with torch.cuda.device(self.device):
            for p in self.peer.values():
                self.lib.sn_ipc_close(C.c_void_p(p))
            self.peer = {}
            if self.base:
                self.lib.sn_ipc_free(C.c_void_p(self.base))
                self.base = 0


class ConcatCache:
    """Bob's activation cache for `--sisa --concat` as ONE peer-visible allocation on rank 0:
    acts [rows, 5408*K] fp32 followed by labels [rows, K] int64.  Every Alice rank maps it (CUDA IPC)
    and her conv kernel stores her 5408-column slot of every row straight into it over NVLink
    (`engine.AliceEngine.forward(out=cache.slot(k), ld=5408*K)`), which replaces both the per-batch
    activation lists of `bob.get_activation_and_labels` (`data_entities_vanilla_sisa.py:272-292`) and
    the torch.cat a single-process concat needs.  Collective set-up: every rank constructs it."""

    def __init__(self, rank, world, device, rows, K):
        self.lib = _lib.load()
        self.rank, self.world, self.device, self.rows, self.K = rank, world, device, int(rows), int(K)
        self.act_bytes = self.rows * self.K * D0 * 4
        total = self.act_bytes + self.rows * self.K * 8
        handle = torch.zeros(64, dtype=torch.uint8, device=device)
        self.base, self.mapped = 0, 0
        if rank == 0:
            base, raw = C.c_void_p(), (C.c_ubyte * 64)()
            with torch.cuda.device(device):
                check(self.lib.sn_ipc_alloc(max(total, 256), C.byref(base), raw))
            self.base = base.value
            handle.copy_(torch.tensor(list(raw), dtype=torch.uint8))
        if world > 1:
            dist.broadcast(handle, src=0)
        if rank != 0:
            raw = (C.c_ubyte * 64)(*handle.cpu().tolist())
            p = C.c_void_p()
            with torch.cuda.device(device):
                check(self.lib.sn_ipc_open(raw, C.byref(p)))
            self.mapped = p.value
        if world > 1:
            dist.barrier()
        self.acts = self.labels = None
        if rank == 0:
            with torch.cuda.device(device):
                self.acts = torch.as_tensor(_View(self.base, self.rows * self.K * D0, "<f4"), device=device).view(self.rows, self.K * D0)
                self.labels = torch.as_tensor(_View(self.base + self.act_bytes, self.rows * self.K, "<i8"),
                                              device=device).view(self.rows, self.K)

    def slot(self, k):
        """Address of column 5408*k of row 0, as seen from THIS rank (Bob's own memory on rank 0)."""
        origin = self.base if self.rank == 0 else self.mapped
        return _Raw(origin + k * D0 * 4, (self.rows, D0), self.device)

    def close(self):
        torch.cuda.synchronize(self.device)
        if self.world > 1 and dist.is_initialized():
            dist.barrier()
        with torch.cuda.device(self.device):
            if self.mapped:
                self.lib.sn_ipc_close(C.c_void_p(self.mapped))
                self.mapped = 0
            if self.base:
                self.acts = self.labels = None
                self.lib.sn_ipc_free(C.c_void_p(self.base))
                self.base = 0
