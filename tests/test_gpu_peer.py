"""Two real GPUs: the peer-memory transport (splitlearning_b200/peer.py - CUDA IPC mailboxes, stream
flags) must carry the vanilla schedule to exactly the same weights and losses as the NCCL send/recv
transport.  Needs 2 GPUs (`gpurun --gpus 2`); skipped on a single-GPU box."""
import os
import tempfile

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

B, N_ALICES, TURN, STEPS = 16, 3, 4, 26


def _run(rank, world, out_dir, port, use_link, flags="kernel"):
    import torch.distributed as dist
    from splitlearning_b200 import spmd
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    device = torch.device("cuda", rank)
    torch.cuda.set_device(device)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=device)
    torch.manual_seed(3)                      # same initial weights on every rank / in both runs
    bob = spmd._CudaBob(device, 1e-3, N_ALICES, B) if rank == 0 else None
    torch.manual_seed(4)
    mine = {}
    for k in range(N_ALICES):
        a = spmd._CudaAlice(device, 1e-3)     # every rank draws every Alice so the streams stay aligned
        if spmd.owner_rank(k, world) == rank:
            mine[k] = a
    g = torch.Generator().manual_seed(9)
    xs = (torch.randint(0, 256, (STEPS, B, 1, 28, 28), generator=g).float() * (torch.rand(STEPS, B, 1, 28, 28, generator=g) < 0.4)).to(device)
    ys = torch.randint(0, 10, (STEPS, B), generator=g).to(device)
    link = None
    if use_link:
        from splitlearning_b200.peer import PeerLink
        link = PeerLink(rank, world, device, B, flags=flags)
    losses = {}
    spmd.run_vanilla_schedule(rank, world, bob, mine, lambda i: (xs[i], ys[i]), N_ALICES, TURN, 0, STEPS, B, device,
                              lambda i, l: losses.__setitem__(i, l.clone()), link)
    torch.cuda.synchronize()
    out = {f"alice{k}": a.get_weights().cpu().numpy() for k, a in mine.items()}
    if rank == 0:
        out["losses"] = np.array([float(losses[i]) for i in range(STEPS)])
        for i, p in enumerate(bob.model.parameters()):
            out[f"bob_p{i}"] = p.detach().cpu().numpy()
    np.savez(os.path.join(out_dir, f"{('peer-' + flags) if use_link else 'nccl'}_r{rank}.npz"), **out)
    if link is not None:
        link.close()
    dist.barrier()
    dist.destroy_process_group()


def _collect(d, tag, world):
    merged = {}
    for r in range(world):
        with np.load(os.path.join(d, f"{tag}_r{r}.npz")) as f:
            merged.update({k: f[k] for k in f.files})
    return merged


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_peer_mailboxes_match_nccl_transport():
    import torch.multiprocessing as mp
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(_run, args=(2, d, 29811, False), nprocs=2, join=True)
        mp.spawn(_run, args=(2, d, 29812, True, "kernel"), nprocs=2, join=True)
        mp.spawn(_run, args=(2, d, 29813, True, "stream"), nprocs=2, join=True)
        ref = _collect(d, "nccl", 2)
        got = {m: _collect(d, "peer-" + m, 2) for m in ("kernel", "stream")}
    assert "losses" in ref and "bob_p0" in ref and np.isfinite(ref["losses"]).all()
    for m, g in got.items():
        assert set(ref) == set(g), m
        for k in ref:
            assert np.array_equal(ref[k], g[k]), (m, k)
