"""Where does a large-batch step go?  Per-launch CUDA-event times of the tcgen05 GEMMs of one training
step (forward, dX, dW+optimizer per layer) at the sweep's shapes.  usage: gemm_breakdown.py B H [mode]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from splitlearning_b200 import engine as E

B, H = int(sys.argv[1]), int(sys.argv[2])
mode = int(sys.argv[3]) if len(sys.argv) > 3 else 0
dims = [5408, H, max(4, (H // 5) // 4 * 4), 100]
dev = torch.device("cuda", 0)
g = torch.Generator().manual_seed(1)
Ws = [(torch.randn(dims[i + 1], dims[i], generator=g) / dims[i] ** 0.5).to(dev) for i in range(3)]
bs = [torch.zeros(dims[i + 1], device=dev) for i in range(3)]
bob = E.BobEngine(dims, Ws, bs, optimizer=E.SN_OPT_SGD_MOMENTUM, lr=1e-4, momentum=0.9, max_batch=B,
                  dropout_p=float(os.environ.get("DROPOUT", "0.5")))
x = (torch.rand(B, 5408, device=dev) * 100) * (torch.rand(B, 5408, device=dev) < 0.6)
y = torch.randint(0, 10, (B,), device=dev)
E.set_matmul_precision(mode)
for i in range(3):
    bob.train_step(x, y, seed=1, offset=i)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(10):
    bob.train_step(x, y, seed=1, offset=10 + i)
e1.record()
torch.cuda.synchronize()
print(f"B={B} H={H} mode={mode}: {e0.elapsed_time(e1) / 10:.3f} ms/step")
bob.profile(True)
for i in range(10):
    bob.train_step(x, y, seed=1, offset=30 + i)
torch.cuda.synchronize()
names = {0: "fwd", 1: "dW+opt", 2: "dX"}
tot = 0.0
for kind in (0, 2, 1):
    for l in range(3):
        t, n = bob.profile_read(kind, l)
        if n:
            K, N = dims[l], dims[l + 1]
            fl = 2.0 * B * K * N
            us = 1e3 * t / n
            tot += us
            extra = f" {16.0 * K * N / us / 1e3:7.0f} GB/s of the 16 B/weight stream" if kind == 1 else ""
            print(f"  {names[kind]:7s} layer {l} [{N}x{K}]: {us:8.1f} us  {fl / us / 1e6:7.1f} TFLOP/s{extra}")
print(f"  sum of bracketed launches {tot:.1f} us")
E.set_matmul_precision(0)
