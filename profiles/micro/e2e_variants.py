"""Where do the ~12 us/step between bench.py's device-resident `value` and its host-buffer `e2e`
go?  Times the same 2000 vanilla steps with (A) the bench's e2e loop, (B) no per-step loss copy,
(C) the loss written by ce_kernel straight into pinned host memory (UVA zero-copy store), (D) C with
device-resident inputs; prints GPU time, and how long the Python enqueue loop itself took (if that
is close to the GPU time the loop is CPU-bound)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch

import bench
from splitlearning_b200 import engine as E
from splitlearning_b200.io import iterate_host_batches

B, D0, N_ALICES, TURN = bench.B, bench.D0, bench.N_ALICES, 64
dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
bob_model, bob, alices = bench.make_world(torch, dev)
nb = 256
xs, ys = bench.synthetic_batches(torch, nb, dev)
hx, hy = bench.synthetic_batches(torch, nb, pin=True)
act_buf = torch.empty(B, D0, device=dev)
cut_buf = torch.empty(B, D0, device=dev)
K, W = 2000, 100
hloss = torch.zeros(K + W, dtype=torch.float32).pin_memory()
dl = torch.zeros((), device=dev)
seed = [0]


def step(i, x, y, loss_out):
    k = (i // TURN) % N_ALICES
    if i % TURN == 0 and i > 0:
        alices[k][0].load_state_dict(alices[(k - 1) % N_ALICES][0].state_dict())
    seed[0] += 1
    E.vanilla_step(alices[k][1], bob, x, y, k, 1, seed[0], act_buf, cut_buf, loss_out)


def run(kind, first, count):
    if kind == "D":
        for j in range(count):
            i = first + j
            step(i, xs[i % nb], ys[i % nb], hloss[i])
        return
    host = ((hx[i % nb], hy[i % nb]) for i in range(first, first + count))
    for j, (dx, dy) in enumerate(iterate_host_batches(host, dev, B)):
        i = first + j
        if kind == "A":
            step(i, dx, dy, dl)
            hloss[i].copy_(dl, non_blocking=True)
        elif kind == "B":
            step(i, dx, dy, dl)
        else:
            step(i, dx, dy, hloss[i])


for kind in ("A", "B", "C", "D", "A", "C"):
    run(kind, 0, W)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    run(kind, W, K)
    e1.record()
    t_enq = time.perf_counter() - t0
    torch.cuda.synchronize()
    t_all = time.perf_counter() - t0
    print(f"{kind}: gpu {1e3 * e0.elapsed_time(e1) / K:.1f} us/step, enqueue loop {1e6 * t_enq / K:.1f} us/step, "
          f"wall {1e6 * t_all / K:.1f} us/step, last loss {float(hloss[W + K - 1]):.4f}", flush=True)
