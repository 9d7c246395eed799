"""BASELINE.json config 5: synthetic scaling sweep of Bob's training step (forward, CE, backward,
optimizer) over batch and hidden width, Bob dims [5408, H, H/5 (rounded to 4), 100].

One GPU:   python profiles/sweep.py [--quick] > profiles/r2_sweep.jsonl
    Every batch size is run at REFERENCE PRECISION - the credited column ("credited": true): the fp32
    weight-stream kernels up to 16 rows; 17..64 rows the fp32 forward and 3xTF32 tcgen05 GEMMs for the
    backward; 3xTF32 on tcgen05 throughout above that (fp32-accurate, inside the 1e-4 parity bound).  Batches >= 128 are also run in one-pass TF32 ("credited": false: 5e-4 on the
    logits, outside the bound; shown as the tensor-pipe ceiling only).
    Per point: time per step, samples/s, model TFLOP/s (6*P_w*B) and the fraction of the
    relevant roofline: HBM (16*P bytes per step, SGD) for the weight-streaming regime, TF32 tensor
    peak (measured bf16 peak / 2) for the tensor regime.
world_size axis (2/4/8 GPUs; the only part of the step that shards is the SISA server step, B = 16):
    python -m torch.distributed.run --nproc-per-node G ... profiles/sweep.py --tp >> profiles/r2_sweep_tp.jsonl
    runs tp.TPServer (first layer's rows split over the ranks) for H in 2048/4096/8192.
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


def precision_name(B, mode):
    if mode:
        return "tf32-tcgen05"
    if B <= 16:
        return "fp32"
    return "fp32-forward+3xtf32-backward" if B <= 64 else "3xtf32-tcgen05"


def ceiling_frac(B, mode, tfl, tf32_peak):
    """3xTF32 executes three TF32 products per useful one: its ceiling is a third of the TF32 peak."""
    return round(tfl / (tf32_peak / 3.0), 4) if (mode == 0 and B > 64) else None


def main():
    quick = "--quick" in sys.argv
    hbm, tf32_peak, src = peaks()
    Hs = [256, 1024, 4096, 8192] if quick else [256, 512, 1024, 2048, 4096, 8192]
    Bs = [16, 64, 256, 1024, 4096] if quick else [16, 32, 64, 256, 1024, 4096, 16384, 65536]
    for H in Hs:
        for B in Bs:
            for mode in ([0] if B < 128 else [0, 1]):
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
                prec = precision_name(B, mode)
                print(json.dumps({"B": B, "H": H, "dims": dims, "precision": prec, "credited": mode == 0,
                                  "ms_per_step": round(ms, 4), "samples_per_s": round(B / (ms * 1e-3), 1),
                                  "tflops": round(tfl, 2), "frac_tf32_peak": round(tfl / tf32_peak, 4),
                                  "frac_3xtf32_ceiling": ceiling_frac(B, mode, tfl, tf32_peak),
                                  "weight_stream_gbs": round(gbs, 1), "frac_hbm": round(gbs / hbm, 4),
                                  "peaks": src, "loss": loss}), flush=True)
                torch.cuda.empty_cache()


def point_record(B, H, mode, steps=None):
    """One sweep point as a dict (bench.py's `sweep` extra uses this)."""
    hbm, tf32_peak, src = peaks()
    steps = steps or max(2, min(50, int(2e11 / (6.0 * 5408 * H * B + 1e9))))
    dims, pw, ms, loss = run_point(B, H, mode, steps)
    tfl = 6.0 * pw * B / (ms * 1e-3) / 1e12
    gbs = 16.0 * pw / (ms * 1e-3) / 1e9
    return {"B": B, "H": H, "precision": precision_name(B, mode),
            "credited": mode == 0, "ms_per_step": round(ms, 4), "samples_per_s": round(B / (ms * 1e-3), 1),
            "tflops": round(tfl, 2), "frac_tf32_peak": round(tfl / tf32_peak, 4),
            "frac_3xtf32_ceiling": ceiling_frac(B, mode, tfl, tf32_peak), "frac_hbm": round(gbs / hbm, 4)}


def main_tp():
    """world_size axis: SISA server steps (B = 16, Adam) with the first layer sharded over the ranks."""
    import torch.distributed as dist
    from splitlearning_b200 import tp
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    hbm, _, src = peaks()
    B, nbat = 16, 400
    for H in (2048, 4096, 8192):
        dims = [5408, H, max(4, (H // 5) // 4 * 4), 100]
        g = torch.Generator(device="cpu").manual_seed(H)
        Ws = [(torch.randn(dims[i + 1], dims[i], generator=g) / dims[i] ** 0.5).to(dev) for i in range(3)]
        bs = [torch.zeros(dims[i + 1], device=dev) for i in range(3)]
        acts = ((torch.rand(nbat * B, 5408, generator=g) * 100) * (torch.rand(nbat * B, 5408, generator=g) < 0.6)).to(dev)
        labels = torch.randint(0, 10, (nbat * B,), generator=g).to(dev)
        offs = list(range(0, nbat * B + 1, B))
        srv = tp.TPServer(rank, world, dev, Ws, bs, B=B, lr=1e-4, weight_decay=1e-5, dropout_p=0.5)
        srv.server_epoch(acts, labels, offs[:41])
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        srv.server_epoch(acts, labels, offs)
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        us = 1e3 * float(t.item()) / nbat
        pw = sum(dims[i] * dims[i + 1] for i in range(3))
        if rank == 0:
            print(json.dumps({"B": B, "H": H, "dims": dims, "world_size": world, "precision": "fp32", "credited": True,
                              "workload": "SISA server step, Adam, first layer's rows split over the ranks",
                              "us_per_step": round(us, 2), "samples_per_s": round(B / (us * 1e-6), 1),
                              "aggregate_weight_stream_gbs": round(24.0 * pw / (us * 1e-6) / 1e9, 1),
                              "frac_hbm_one_gpu": round(24.0 * pw / (us * 1e-6) / 1e9 / hbm, 4), "peaks": src}), flush=True)
        srv.close()
        del Ws, bs, acts
        torch.cuda.empty_cache()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    if "--tp" in sys.argv:
        main_tp()
    else:
        main()
