import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from splitlearning_b200 import engine as E
torch.manual_seed(0)
K, N, B = 256, 128, 128
W = (torch.randn(N, K) / 16).cuda(); b = torch.zeros(N).cuda()
W0 = W.clone()
bob = E.BobEngine([K, N], [W], [b], dropout_p=0.0, optimizer=E.SN_OPT_SGD_MOMENTUM, lr=1.0, momentum=0.0, max_batch=B)
x = torch.randn(B, K).cuda()
dout = torch.randn(B, N).cuda()
E.set_matmul_precision(1)
out = bob.forward(x, train_mode=True)
cut = bob.backward_update(x, dout)
torch.cuda.synchronize()
E.set_matmul_precision(0)
ref_cut = dout @ W0
ref_dW = dout.t() @ x
got_dW = (W0 - W)          # lr=1, momentum 0: W_new = W - g
def rel(a, b): return float((a-b).norm()/b.norm())
print("fwd rel", rel(out, x @ W0.t()))
print("cut rel", rel(cut, ref_cut), "norms", float(cut.norm()), float(ref_cut.norm()))
print("dW  rel", rel(got_dW, ref_dW), "norms", float(got_dW.norm()), float(ref_dW.norm()))
# structure of the error: per 32-col block of cut
for c in range(0, K, 32):
    print("cut cols", c, rel(cut[:, c:c+32], ref_cut[:, c:c+32]), end=" | ")
print()
# try hypotheses: partial reductions
for nn in (8, 32, 64):
    print("cut vs partial n<%d" % nn, rel(cut, dout[:, :nn] @ W0[:nn]))
print("cut[0,:8]", cut[0,:8].cpu().numpy(), "\nref[0,:8]", ref_cut[0,:8].cpu().numpy())
print("dW[0,:8]", got_dW[0,:8].cpu().numpy(), "\nref[0,:8]", ref_dW[0,:8].cpu().numpy())
