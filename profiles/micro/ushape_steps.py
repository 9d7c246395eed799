"""BASELINE.json configs[0] (default U-shape mode): Alice conv -> Bob 5408-1000-100 (ReLU last) -> Alice
tail 100-10 + CE -> backward through all three, Adam per entity.  us/step on one GPU, data resident."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch

from splitlearning_b200 import engine as E

dev = torch.device("cuda", 0)
g = torch.Generator().manual_seed(0)
rn = lambda *s: torch.randn(*s, generator=g)
W = [(rn(1000, 5408) / 5408 ** 0.5).to(dev), (rn(100, 1000) / 1000 ** 0.5).to(dev)]
b = [torch.zeros(1000, device=dev), torch.zeros(100, device=dev)]
Wt, bt = [(rn(10, 100) / 10).to(dev)], [torch.zeros(10, device=dev)]
cw, cb = (rn(32, 1, 3, 3) * 0.1).to(dev), torch.zeros(32, device=dev)
adam = dict(optimizer=E.SN_OPT_ADAM, lr=1e-3)
bob = E.BobEngine([5408, 1000, 100], W, b, relu_last=True, dropout_p=0.0, **adam)
tail = E.BobEngine([100, 10], Wt, bt, dropout_p=0.0, **adam)
alice = E.AliceEngine(cw, cb, **adam)
xs = (torch.randint(0, 256, (64, 16, 1, 28, 28), generator=g).float() * (torch.rand(64, 16, 1, 28, 28, generator=g) < 0.4)).to(dev)
ys = torch.randint(0, 10, (64, 16), generator=g).to(dev)
loss = torch.zeros((), device=dev)


def step(i):
    x, y = xs[i % 64], ys[i % 64]
    act = alice.forward(x)
    hb = bob.forward(act, train_mode=True)
    _, dhb, _ = tail.train_step(hb, y, loss_out=loss)
    cut = bob.backward_update(act, dhb)
    alice.backward_update(x, cut)


for i in range(100):
    step(i)
torch.cuda.synchronize()
K = 2000
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(K):
    step(i)
e1.record()
torch.cuda.synchronize()
us = 1e3 * e0.elapsed_time(e1) / K
P = 5408 * 1000 + 1000 * 100 + 100 * 10 + 1100 + 10 + 320
print(f"U-shape step: {us:.1f} us/step = {16 / us * 1e6:.0f} samples/s; 24*P = {24 * P / 1e6:.0f} MB -> {24 * P / (us * 1e-6) / 1e9:.0f} GB/s; loss {loss.item():.4f}")
