"""Runs each kernel whose roofline claim DESIGN.md quotes a few times, so that ONE
`ncu --set full -k regex:...` pass over this script captures them all (profiles/README.md):

    ncu --set full --clock-control none --import-source on -k regex:'bwd_update|conv_pool_stream|skinny_fwd_x3|gemm_tf32|fwd_skinny' \
        -o gpurun_out/r2_targets python profiles/micro/ncu_targets.py

Sections (pick with argv): sisa (Adam traversal with the fused next forward), vanilla (SGD traversal
with dA), fill (conv_pool_stream_kernel over a whole shard), sktc (tcgen05 skinny forward), gemm
(tcgen05 fwd / dX / dW at batch 8192, TF32 and 3xTF32), shard (two row slices on one GPU), concat
(K=7 first layer).
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from splitlearning_b200 import _lib, engine as E, models as M  # noqa: E402

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
lib = _lib.load()
what = set(sys.argv[1:]) or {"sisa", "vanilla", "fill", "sktc", "gemm", "shard", "concat"}
B = 16
g = torch.Generator().manual_seed(3)


def acts(n, d=5408):
    return (torch.rand(n, d, generator=g) * 30 * (torch.rand(n, d, generator=g) < 0.6)).to(dev)


if "sisa" in what:
    torch.manual_seed(1)
    bob = M.model2_sisa(dev).make_engine(optimizer=E.SN_OPT_ADAM, lr=1e-3, weight_decay=1e-5, n_opt_slots=1, max_batch=B)
    a, y = acts(6 * B), torch.randint(0, 10, (6 * B,), generator=g).to(dev)
    bob.server_epoch(a, y, list(range(0, 6 * B + 1, B)), seed=7, offset=0)
    torch.cuda.synchronize()

if "vanilla" in what:
    torch.manual_seed(2)
    bob = M.model2_sisa(dev).make_engine(optimizer=E.SN_OPT_SGD_MOMENTUM, lr=1e-3, momentum=0.9, n_opt_slots=1, max_batch=B)
    am = M.model1_sisa(dev)
    alice = am.make_engine(optimizer=E.SN_OPT_SGD_MOMENTUM, lr=1e-3, momentum=0.9, max_batch=B)
    x = (torch.randint(0, 256, (B, 1, 28, 28), generator=g).float()).to(dev)
    y = torch.randint(0, 10, (B,), generator=g).to(dev)
    act, cut, loss = torch.empty(B, 5408, device=dev), torch.empty(B, 5408, device=dev), torch.zeros((), device=dev)
    for i in range(4):
        E.vanilla_step(alice, bob, x, y, 0, 1, i + 1, act, cut, loss)
    torch.cuda.synchronize()

if "fill" in what:
    am = M.model1_sisa(dev)
    alice = am.make_engine(optimizer=E.SN_OPT_NONE, max_batch=B)
    n = 17348
    x = (torch.randint(0, 256, (n, 1, 28, 28), generator=g).float()).to(dev)
    out = torch.empty(n, 5408, device=dev)
    for _ in range(3):
        alice.forward(x, out=out)
    torch.cuda.synchronize()

if "sktc" in what:
    lib.sn_set_skinny_tensor_forward(1)
    W = (torch.randn(5000, 5408, generator=g) / 73.0).to(dev)
    b = torch.zeros(5000, device=dev)
    xx = acts(B)
    for _ in range(3):
        E.linear_forward(xx, W, b, relu=True)
    torch.cuda.synchronize()
    lib.sn_set_skinny_tensor_forward(0)

if "gemm" in what:
    for mode in (1, 0):
        lib.sn_set_matmul_precision(mode)
        torch.manual_seed(4)
        BB = 8192
        bob = M.model2_sisa(dev).make_engine(optimizer=E.SN_OPT_SGD_MOMENTUM, lr=1e-3, momentum=0.9, n_opt_slots=1, max_batch=BB)
        a, y = acts(BB), torch.randint(0, 10, (BB,), generator=g).to(dev)
        for i in range(2):
            bob.train_step(a, y, seed=1, offset=i + 1)
        torch.cuda.synchronize()
        bob.close()
    lib.sn_set_matmul_precision(0)

if "shard" in what:
    torch.manual_seed(5)
    m = M.model2_sisa(dev)
    W1, b1 = m.fc1.weight.data, m.fc1.bias.data
    kw = dict(optimizer=E.SN_OPT_ADAM, lr=1e-3, weight_decay=1e-5)
    shards = [E.ShardEngine(W1[lo:hi], b1[lo:hi], 5000, lo, dropout_p=0.5, **kw) for lo, hi in ((0, 625), (625, 1250))]
    a = acts(3 * B)
    H1, dH1 = torch.zeros(B, 5000, device=dev), (torch.randn(B, 5000, generator=g) * 1e-3).to(dev)
    for i in range(2):
        for sh in shards:
            sh.forward(a[i * B:(i + 1) * B], H1, 5000, train=True, seed=1, offset=i)
        for sh in shards:
            sh.backward_update(a[i * B:(i + 1) * B], dH1, 5000, X_next=a[(i + 1) * B:(i + 2) * B])
    torch.cuda.synchronize()

if "concat" in what:
    torch.manual_seed(6)
    K = 7
    mc = M.model2_sisa_concat(num_clients=K, device=dev)
    bob = mc.make_engine(optimizer=E.SN_OPT_ADAM, lr=1e-3, weight_decay=1e-5, n_opt_slots=1, max_batch=B)
    bob.set_ce_heads(K)
    a = acts(4 * B, 5408 * K)
    y = torch.randint(0, 10, (4 * B, K), generator=g).to(dev)
    bob.server_epoch(a, y, list(range(0, 4 * B + 1, B)), seed=7, offset=0)
    torch.cuda.synchronize()
print("ncu_targets: done", sorted(what))
