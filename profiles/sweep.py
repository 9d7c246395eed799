"""BASELINE.json config 5: synthetic scaling sweep of Bob's training step (forward, CE, backward,
optimizer) over batch and hidden width, Bob dims [5408, H, H/5 (rounded to 4), 100], on one B200.
fp32 mode for every batch (parity path), TF32 tcgen05 mode for B >= 128.  Prints one JSON line per
point with time per step, samples/s, model TFLOP/s (6*P_w*B) and the fraction of the
relevant roofline: HBM (16*P bytes per step, SGD) for the weight-streaming regime, TF32 tensor
peak (measured bf16 peak / 2) for the tensor regime.

    python profiles/sweep.py [--quick] > profiles/r1_sweep.jsonl
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from splitlearning_b200 import engine as E  # noqa: E402


def peaks():
    try:
        d = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return d["hbm_gbs"], d["bf16_tflops"] / 2.0, "measured"
    except Exception:
        return 6650.0, 1590.0 / 2.0, "fallback"


def run_point(B, H, mode, steps):
    dims = [5408, H, max(4, (H // 5) // 4 * 4), 100]
    dev = torch.device("cuda", 0)
    g = torch.Generator(device="cpu").manual_seed(H)
    Ws = [(torch.randn(dims[i + 1], dims[i], generator=g) / dims[i] ** 0.5).to(dev) for i in range(3)]
    bs = [torch.zeros(dims[i + 1], device=dev) for i in range(3)]
    bob = E.BobEngine(dims, Ws, bs, optimizer=E.SN_OPT_SGD_MOMENTUM, lr=1e-4, momentum=0.9, max_batch=B)
    x = (torch.rand(B, 5408, device=dev) * 100) * (torch.rand(B, 5408, device=dev) < 0.6)
    y = torch.randint(0, 10, (B,), device=dev)
    loss = torch.zeros((), device=dev)
    E.set_matmul_precision(mode)
    try:
        for i in range(3):
            bob.train_step(x, y, seed=1, offset=i, want_cut_grad=True, loss_out=loss)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            bob.train_step(x, y, seed=1, offset=10 + i, want_cut_grad=True, loss_out=loss)
        e1.record()
        torch.cuda.synchronize()
    finally:
        E.set_matmul_precision(0)
    ms = e0.elapsed_time(e1) / steps
    pw = sum(dims[i] * dims[i + 1] for i in range(3))
    bob.close()
    return dims, pw, ms, float(loss.item())


def main():
    quick = "--quick" in sys.argv
    hbm, tf32_peak, src = peaks()
    Hs = [256, 1024, 4096, 8192] if quick else [256, 512, 1024, 2048, 4096, 8192]
    Bs = [16, 64, 256, 1024, 4096] if quick else [16, 64, 256, 1024, 4096, 16384, 65536]
    for H in Hs:
        for B in Bs:
            for mode in ([0] if B < 128 else [1]):
                if B * max(H, 5408) * 4 > 6e9:
                    continue
                steps = max(2, min(50, int(2e11 / (6.0 * 5408 * H * B + 1e9))))
                try:
                    dims, pw, ms, loss = run_point(B, H, mode, steps)
                except Exception as e:      # report, keep sweeping
                    print(json.dumps({"B": B, "H": H, "mode": mode, "error": str(e)[:200]}), flush=True)
                    continue
                flops = 6.0 * pw * B
                tfl = flops / (ms * 1e-3) / 1e12
                gbs = 16.0 * pw / (ms * 1e-3) / 1e9
                print(json.dumps({"B": B, "H": H, "dims": dims, "precision": "tf32-tcgen05" if mode else "fp32",
                                  "ms_per_step": round(ms, 4), "samples_per_s": round(B / (ms * 1e-3), 1),
                                  "tflops": round(tfl, 2), "frac_tf32_peak": round(tfl / tf32_peak, 4),
                                  "weight_stream_gbs": round(gbs, 1), "frac_hbm": round(gbs / hbm, 4),
                                  "peaks": src, "loss": loss}), flush=True)
                torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
