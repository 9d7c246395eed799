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
                         loss_sink=None):
    """Walk steps [first_step, first_step+n_steps) of the round-robin schedule.

    bob: engine or None (ranks != 0); alices: {alice_index: engine} for the Alices this rank hosts;
    batch_fn(i) -> (x, y) device tensors for step i (only called on the owning rank)."""
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
        if world == 1:
            x, y = batch_fn(i)
            act = alices[k].forward(x)
            loss, g, _ = bob.train_step(act, y, slot=k)
            alices[k].backward_update(x, g)
        elif rank == own:
            x, y = batch_fn(i)
            alices[k].forward(x, out=act_view)
            lab_view.copy_(y)
            dist.send(msg, dst=0)
            dist.recv(cut, src=0)
            alices[k].backward_update(x, cut)
        elif rank == 0:
            dist.recv(msg, src=own)
            loss, g, _ = bob.train_step(act_view, lab_view, slot=k)
            dist.send(g, dst=own)
        else:
            continue
        if loss_sink is not None and (rank == 0):
            loss_sink(i, loss)


# ------------------------------------------------------------------------------- CUDA engines
class _CudaAlice:
    def __init__(self, device, lr):
        from . import engine as E
        from . import models as M
        self.model = M.model1_sisa(device)
        self.eng = self.model.make_engine(optimizer=E.SN_OPT_SGD_MOMENTUM, lr=lr, momentum=0.9, max_batch=64)
        self.w, self.b = self.model._wb()

    def forward(self, x, out=None):
        return self.eng.forward(x, out=out)

    def backward_update(self, x, cut):
        self.eng.backward_update(x, cut)

    def get_weights(self):
        return torch.cat([self.w.data.reshape(-1), self.b.data.reshape(-1)])

    def set_weights(self, flat):
        self.w.data.copy_(flat[:288].view(32, 1, 3, 3))
        self.b.data.copy_(flat[288:320])


class _CudaBob:
    def __init__(self, device, lr, n_slots, B):
        from . import engine as E
        from . import models as M
        self.model = M.model2_sisa(device)
        self.eng = self.model.make_engine(optimizer=E.SN_OPT_SGD_MOMENTUM, lr=lr, momentum=0.9, n_opt_slots=n_slots,
                                          max_batch=B)
        self.offset = 0
        self.loss = torch.zeros((), device=device)

    def train_step(self, act, y, slot):
        self.offset += 1
        return self.eng.train_step(act, y, slot=slot, masks=None, seed=1, offset=self.offset, want_cut_grad=True,
                                   loss_out=self.loss)


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

    def timed(batch_fn, first, sink=None):
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        run_vanilla_schedule(rank, world, bob, mine, batch_fn, n_alices, turn, first, K, B, device, sink)
        e1.record()
        torch.cuda.synchronize()
        dist.barrier()
        t = torch.tensor([e0.elapsed_time(e1)], device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    run_vanilla_schedule(rank, world, bob, mine, lambda i: (xs[i % nb], ys[i % nb]), n_alices, turn, 0, W, B, device)
    l0 = lib.sn_launch_count()
    import bench as bench_mod
    with bench_mod.ClockSampler(local) as clk:
        ms = timed(lambda i: (xs[i % nb], ys[i % nb]), W)
    launches = torch.tensor([lib.sn_launch_count() - l0], device=device, dtype=torch.int64)
    dist.all_reduce(launches)

    # end to end: inputs start in pinned host memory on the Alice ranks, loss read back on rank 0
    phx, phy = (hx.pin_memory(), hy.pin_memory()) if mine else (None, None)
    dx = torch.empty(B, 1, 28, 28, device=device)
    dy = torch.empty(B, dtype=torch.int64, device=device)
    hloss = torch.zeros(K + 2 * W + K, dtype=torch.float32).pin_memory()

    def host_batch(i):
        dx.copy_(phx[i % nb], non_blocking=True)
        dy.copy_(phy[i % nb], non_blocking=True)
        return dx, dy

    def sink(i, loss):
        hloss[i - (W + K)].copy_(loss, non_blocking=True)
    e2e_ms = timed(host_batch, W + K, sink)

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
            "note": "vanilla split learning is a sequential chain (SURVEY.md 8e): extra GPUs host Alices and add "
                    "two NVLink messages per step; they cannot add throughput without changing results",
        }
        print(json.dumps(line))
    dist.barrier()
    dist.destroy_process_group()
