"""Quick check of the tcgen05 TF32 GEMM path (sn_linear_forward with matmul precision 1)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from splitlearning_b200 import engine as E

def tf32_trunc(a):
    return (a.view(np.uint32) & np.uint32(0xFFFFE000)).view(np.float32)

def rel(a, b):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / np.linalg.norm(b))

E.set_matmul_precision(1)
for (B, K, N) in [(128, 32, 128), (128, 64, 128), (256, 512, 256), (200, 1000, 100), (1000, 5408, 5000), (65, 100, 12)]:
    rng = np.random.RandomState(B + K + N)
    x = (rng.rand(B, K) * 4 - 1).astype(np.float32)
    W = (rng.uniform(-1, 1, (N, K)) / np.sqrt(K)).astype(np.float32)
    bias = rng.uniform(-1, 1, N).astype(np.float32)
    got = E.linear_forward(torch.from_numpy(x).cuda(), torch.from_numpy(W).cuda(), torch.from_numpy(bias).cuda(), relu=True)
    torch.cuda.synchronize()
    got = got.cpu().numpy()
    emu = np.maximum(tf32_trunc(x).astype(np.float64) @ tf32_trunc(W).astype(np.float64).T + bias, 0)
    ref = np.maximum(x.astype(np.float64) @ W.astype(np.float64).T + bias, 0)
    print(f"B={B:5d} K={K:5d} N={N:5d}  vs tf32-truncated emulation {rel(got, emu):.2e}   vs fp64 {rel(got, ref):.2e}", flush=True)
E.set_matmul_precision(0)
for (B, K, N) in [(256, 512, 256), (1000, 5408, 5000)]:
    rng = np.random.RandomState(B + K + N)
    x = (rng.rand(B, K) * 4 - 1).astype(np.float32)
    W = (rng.uniform(-1, 1, (N, K)) / np.sqrt(K)).astype(np.float32)
    bias = rng.uniform(-1, 1, N).astype(np.float32)
    got = E.linear_forward(torch.from_numpy(x).cuda(), torch.from_numpy(W).cuda(), torch.from_numpy(bias).cuda(), relu=True).cpu().numpy()
    ref = np.maximum(x.astype(np.float64) @ W.astype(np.float64).T + bias, 0)
    f32 = np.maximum(x @ W.T + bias, 0)
    print(f"3xTF32 B={B} K={K} N={N}: vs fp64 {rel(got, ref):.2e}   (numpy fp32 vs fp64: {rel(f32, ref):.2e})", flush=True)
# timing at a tensor-bound shape
B, K, N = 8192, 5408, 5000
x = torch.rand(B, K, device="cuda"); W = torch.rand(N, K, device="cuda") / K; bias = torch.zeros(N, device="cuda")
E.set_matmul_precision(1)
for _ in range(2): y = E.linear_forward(x, W, bias, relu=True)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5): y = E.linear_forward(x, W, bias, relu=True)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print(f"tcgen05 TF32 GEMM {B}x{K}x{N}: {ms:.3f} ms  {2.0*B*K*N/ms/1e9:.1f} TFLOP/s")
E.set_matmul_precision(0)
for _ in range(2): y3 = E.linear_forward(x, W, bias, relu=True)
torch.cuda.synchronize(); e0.record()
for _ in range(5): y3 = E.linear_forward(x, W, bias, relu=True)
e1.record(); torch.cuda.synchronize()
ms3 = e0.elapsed_time(e1) / 5
print(f"tcgen05 3xTF32 GEMM same shape: {ms3:.3f} ms  {2.0*B*K*N/ms3/1e9:.1f} TFLOP/s (fp32-equivalent)")
E.set_matmul_precision(1)
torch.backends.cuda.matmul.allow_tf32 = True
for _ in range(2): y2 = torch.relu(torch.nn.functional.linear(x, W, bias))
torch.cuda.synchronize(); e0.record()
for _ in range(5): y2 = torch.relu(torch.nn.functional.linear(x, W, bias))
e1.record(); torch.cuda.synchronize()
ms2 = e0.elapsed_time(e1) / 5
print(f"cuBLAS TF32 (torch) same shape: {ms2:.3f} ms  {2.0*B*K*N/ms2/1e9:.1f} TFLOP/s   rel diff {float((y-y2).norm()/y2.norm()):.2e}")
