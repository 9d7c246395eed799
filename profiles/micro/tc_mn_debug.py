import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from splitlearning_b200 import engine as E
def rel(a, b): return float((a.double()-b.double()).norm()/b.double().norm())
for mode in (1, 0):
    for (K, N, B) in [(256, 128, 128), (5408, 100, 80), (256, 52, 80), (52, 100, 48), (5408, 256, 80), (512, 104, 300), (104, 100, 300)]:
        torch.manual_seed(0)
        W = (torch.randn(N, K) / K ** 0.5).cuda(); b = torch.zeros(N).cuda()
        W0 = W.clone()
        bob = E.BobEngine([K, N], [W], [b], dropout_p=0.0, optimizer=E.SN_OPT_SGD_MOMENTUM, lr=1.0, momentum=0.0, max_batch=B)
        x = torch.randn(B, K).cuda(); dout = torch.randn(B, N).cuda()
        E.set_matmul_precision(mode)
        out = bob.forward(x, train_mode=True)
        cut = bob.backward_update(x, dout)
        torch.cuda.synchronize()
        E.set_matmul_precision(0)
        print(f"mode {mode} K={K:5d} N={N:4d} B={B:4d}: fwd {rel(out, x.double() @ W0.double().t()):.2e}  cut {rel(cut, dout.double() @ W0.double()):.2e}  dW {rel(W0 - W, dout.double().t() @ x.double()):.2e}", flush=True)
