"""BASELINE.json config 4 shape (--sisa --concat, world_size 8: 7 Alices): server steps of the concat
Bob (fc1 5000 x 37856 = 757 MB of weights, Adam) on one GPU.  Prints us/step and the fraction of the
24*P HBM floor; compares the first step with a torch fp32 reference of the same step."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch

from splitlearning_b200 import engine as E, models as M

K, B, D0 = 7, 16, 5408
dev = torch.device("cuda", 0)
torch.manual_seed(0)
model = M.model2_sisa_concat(K, dev)
bob = model.make_engine(optimizer=E.SN_OPT_ADAM, lr=1e-3, weight_decay=1e-5, n_opt_slots=1, max_batch=B)
bob.set_ce_heads(K)
nbat = 40
acts = (torch.rand(nbat * B, D0 * K, device=dev) * 30) * (torch.rand(nbat * B, D0 * K, device=dev) < 0.6)
labels = torch.randint(0, 10, (nbat * B, K), device=dev)
offs = list(range(0, nbat * B + 1, B))
# reference for the first step (eval-style: no dropout) through torch on the GPU
with torch.no_grad():
    W = [p.detach().clone() for p in (model.fc1.weight, model.fc2.weight, model.fc3.weight)]
    bs = [p.detach().clone() for p in (model.fc1.bias, model.fc2.bias, model.fc3.bias)]
x0 = acts[:B]
logits = bob.forward(x0, train_mode=False)
ref = torch.relu(torch.relu(x0 @ W[0].T + bs[0]) @ W[1].T + bs[1]) @ W[2].T + bs[2]
err = float((logits - ref).norm() / ref.norm())
print(f"concat K={K}: forward rel err vs torch fp32 (TF32 off) {err:.2e}")
losses = torch.zeros(nbat, device=dev)
bob.server_epoch(acts, labels.reshape(-1), offs, seed=3, loss_per_step=losses)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
bob.server_epoch(acts, labels.reshape(-1), offs, seed=4, loss_per_step=losses)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
P = sum(p.numel() for p in model.parameters())
us = 1e3 * ms / nbat
print(f"P = {P / 1e6:.1f} M params; {us:.1f} us/step; 24*P = {24 * P / 1e9:.2f} GB -> {24 * P / (us * 1e-6) / 1e9:.0f} GB/s "
      f"({24 * P / (us * 1e-6) / 1e9 / 6454:.2f} of measured HBM peak); losses {losses[:3].tolist()}")
