import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from splitlearning_b200 import engine as E
from oracle import torch_ref as R
def rel(a, b): return float((a.double()-b.double()).norm()/b.double().norm())
for p in (0.0, 0.5):
  for dims in ([512, 104, 100], [5408, 512, 104, 100]):
    torch.manual_seed(0)
    L = len(dims) - 1
    Ws = [(torch.randn(dims[i+1], dims[i]) / dims[i] ** 0.5) for i in range(L)]
    bs = [torch.randn(dims[i+1]) * 0.1 for i in range(L)]
    B = 300
    x = torch.rand(B, dims[0]) * 2
    y = torch.randint(0, 10, (B,))
    masks = [(torch.rand(B, dims[i+1]) < 0.5).to(torch.uint8) for i in range(L-1)] if p > 0 else None
    spec = R.BobSpec(tuple(dims), False, p)
    params = []
    for w, b in zip(Ws, bs): params += [w.clone().requires_grad_(True), b.clone().requires_grad_(True)]
    xt = x.clone().requires_grad_(True)
    logits_ref = R.bob_forward(xt, params, spec, [m.float() for m in masks] if masks else None)
    loss_ref = R.cross_entropy(logits_ref, y); loss_ref.backward()
    for mode in (2, 0):
        gw = [w.clone().cuda() for w in Ws]; gb = [b.clone().cuda() for b in bs]
        bob = E.BobEngine(dims, gw, gb, dropout_p=p, optimizer=E.SN_OPT_SGD_MOMENTUM, lr=1.0, momentum=0.0, max_batch=B)
        E.set_matmul_precision(mode)
        loss, cut, logits = bob.train_step(x.cuda(), y.cuda(), masks=[m.cuda() for m in masks] if masks else None, want_logits=True)
        torch.cuda.synchronize(); E.set_matmul_precision(0)
        dws = [rel(Ws[i].cuda() - gw[i], params[2*i].grad.cuda()) for i in range(L)]
        dbs = [rel(bs[i].cuda() - gb[i], params[2*i+1].grad.cuda()) for i in range(L)]
        print(f"p={p} dims={dims} mode={mode}: logits {rel(logits.cpu(), logits_ref):.1e} cut {rel(cut.cpu(), xt.grad):.1e} dW {['%.1e' % v for v in dws]} db {['%.1e' % v for v in dbs]}", flush=True)
