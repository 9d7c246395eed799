#!/usr/bin/env python
"""Benchmark of the split-NN hot path (BASELINE.json metric: split-NN train samples/s; SISA
server-epoch ms alongside; reference CPU path timed on the same box).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

One "step" = one batch (B=16, the reference's default) through the whole vanilla split step:
Alice conv forward -> Bob forward / cross-entropy / backward / SGD-momentum (per-Alice state)
-> cut-layer gradient -> Alice conv backward + SGD-momentum.  Workload at N=1 is
BASELINE.json configs[1]: `--vanilla`, world_size=5 (1 Bob + 4 Alices taking turns), all on one
B200.  Under torchrun (N>1): Bob on rank 0, the 4 Alices spread over ranks 1..N-1, activations and
cut-layer gradients written by the producing kernels straight into the consumer GPU's mailbox over
NVLink (peer.PeerLink; SN_LINK=nccl for send/recv); the protocol is strictly sequential
(SURVEY.md 8e), so total work is fixed: scaling = "strong".

`--impl reference`: the UNMODIFIED reference (oracle/_ref) through its torch.distributed.rpc path on
the host cores (oracle/ref_runner.py); under torchrun rank 0 alone runs it.

Prints ONE JSON line (rank 0).  Besides the contract's fields: `roofline` (dominant kernel + whole
step), `cpu_baseline` (reference over RPC on a bounded sample, the single-process port beside it),
`sisa` (server epoch, cache fill, tensor-parallel epoch at N>1), `sisa_concat` (BASELINE.json
configs[3]), `sweep` (five points of configs[4] at reference precision), `parity` (N>1: multi-GPU
paths against the one-GPU path).  See DESIGN.md "Measurement".
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

B = 16
D0 = 5408
DIMS = [5408, 5000, 1000, 100]
N_ALICES = 4
PREHEAT = 40                  # untimed steps before the W warm-up steps (see run_ours_single)
P_BOB = sum(DIMS[i] * DIMS[i + 1] + DIMS[i + 1] for i in range(3))        # 32,146,100
P_FC1 = DIMS[0] * DIMS[1]
L2_MB = 126


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """SM clock + throttle reasons sampled DURING the timed region (NVML, 20 ms period)."""
    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting"}

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self.period = float(os.environ.get("SN_CLOCK_PERIOD", "0.02"))
        self._stop = threading.Event()
        self._t = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        while not self._stop.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                mask = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                for bit, name in self.REASONS.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(self.period)

    def __enter__(self):
        if self.nv:
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._t:
            self._t.join()

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


# ----------------------------------------------------------------------------------- reference arm
REF_MAX_STEPS = 300           # the RPC reference runs ~150 ms per step: keep the arm within a few minutes


def turn_for(steps):
    """Steps an Alice trains before the next one takes over.  The driver's short run (20 steps) must
    still see every Alice train and time Alice->Alice hand-offs; long runs use 64.  (SN_TURN overrides,
    for A/B timing of the hand-off cost.)"""
    if os.environ.get("SN_TURN"):
        return max(1, int(os.environ["SN_TURN"]))
    return max(1, min(64, steps // N_ALICES))


def run_reference(args):
    """`--impl reference`: the reference's own CPU path on the box's host cores.  When the copy of
    the unmodified reference is present (oracle/_ref, made by oracle/fetch_ref.sh in the build
    container; it travels with the snapshot) this runs its alice/bob entity classes under
    torch.distributed.rpc in 5 processes (oracle/ref_runner.py) - kind "reference".  Without it, the
    torch-CPU port of the same step (oracle/cpu_baseline.py; no RPC, so faster) - kind "port"."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    from oracle import ref_runner
    ncpu = ref_runner.host_threads()
    extra = {}
    r = None
    if ref_runner.available():
        k = min(args.steps, REF_MAX_STEPS)
        try:
            r = ref_runner.time_vanilla_rpc(k, args.warmup, world=N_ALICES + 1, batch=B, threads=ncpu)
        except Exception as e:                      # the RPC harness could not run here: say so, time the port
            extra = {"reference_error": f"{type(e).__name__}: {e}"[:300]}
    if r is not None:
        value, dt, steps, threads = r["samples_per_s"], r["seconds"], r["steps"], r["threads"]
        kind = "reference"
        sample = (f"{steps} vanilla split steps of B={B} after {r['warmup']} warm-up steps through the UNMODIFIED reference: "
                  f"data_entities_vanilla.alice/bob in {N_ALICES + 1} processes over torch.distributed.rpc + dist_autograd + "
                  f"DistributedOptimizer, {threads} intra-op threads per process (torch default = all cores), "
                  f"LDA-partitioned synthetic shards {r['shard_train_sizes']}")
        if steps < args.steps:
            sample += f"; capped from --steps {args.steps} to keep the arm within minutes"
        extra = {"world_size": N_ALICES + 1}
    else:
        from oracle import cpu_baseline as cb
        w = cb.VanillaCPU(n_alices=N_ALICES, batch=B)
        # torchrun pins OMP_NUM_THREADS=1; the reference as shipped lets torch use the host's cores.
        best = None
        for nt in sorted({ncpu, max(1, ncpu * 2 // 3), max(1, ncpu // 2)}):
            torch.set_num_threads(nt)
            t = cb.time_steps(w, 3, 1)
            if best is None or t < best[0]:
                best = (t, nt)
        torch.set_num_threads(best[1])
        steps = args.steps
        dt = cb.time_steps(w, steps, args.warmup)
        value, threads, kind = steps * B / dt, best[1], "port"
        sample = (f"{steps} vanilla split steps of B={B} on torch-CPU, single process, no RPC "
                  f"({'the RPC harness failed' if extra else 'oracle/_ref absent'})")
    line = {
        "impl": "reference", "metric": "split_nn_train_samples_per_sec", "value": value, "unit": "samples/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.gpus, turn_for(args.steps)),
        "cpu_baseline": dict({"value": value, "unit": "samples/s", "cores": threads, "kind": kind, "sample": sample}, **extra),
        "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "host": {"cpu_count": os.cpu_count(), "usable_threads": ncpu},
    }
    print(json.dumps(line))


def workload_config(n_gpus, turn=64):
    return {"workload": "--vanilla split NN, world_size=5 (1 Bob + 4 Alices round-robin), batch 16, "
                        "model1_sisa + model2_sisa 5408-5000-1000-100, SGD momentum 0.9 per Alice, lr 1e-3; "
                        f"each Alice trains {turn} steps, then hands her weights to the next (every Alice trains and "
                        "hand-offs fall inside the timed window)",
            "global_batch": B, "parallelism": "single GPU" if n_gpus == 1 else f"Bob on rank 0 + Alices on ranks 1..{n_gpus - 1}",
            "l2": f"per-step working set {20 * P_BOB / 1e6:.0f} MB (weights + momentum) > {L2_MB} MB L2: no flush needed",
            "dropout": "in-kernel Philox (perf mode)",
            "preheat_steps": PREHEAT}


# ----------------------------------------------------------------------------------- our arm, N = 1
def make_world(torch, device, lr=1e-3):
    """Bob + 4 Alices through the public classes' engines, synthetic MNIST-shaped data in HBM."""
    from splitlearning_b200 import engine as E
    from splitlearning_b200 import models as M
    torch.manual_seed(0)
    bob_model = M.model2_sisa(device)
    bob = bob_model.make_engine(optimizer=E.SN_OPT_SGD_MOMENTUM, lr=lr, momentum=0.9, n_opt_slots=N_ALICES, max_batch=B)
    alices = []
    for _ in range(N_ALICES):
        m = M.model1_sisa(device)
        alices.append((m, m.make_engine(optimizer=E.SN_OPT_SGD_MOMENTUM, lr=lr, momentum=0.9, max_batch=B)))
    return bob_model, bob, alices


def synthetic_batches(torch, n, device=None, pin=False):
    g = torch.Generator().manual_seed(1234)
    x = torch.randint(0, 256, (n, B, 1, 28, 28), generator=g).float() * (torch.rand(n, B, 1, 28, 28, generator=g) < 0.4)
    y = torch.randint(0, 10, (n, B), generator=g)
    if pin:
        return x.pin_memory(), y.pin_memory()
    return x.to(device), y.to(device)


def run_ours_single(args):
    import torch
    from splitlearning_b200 import _lib
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    lib = _lib.load()
    bob_model, bob, alices = make_world(torch, dev)
    nb = 256
    xs, ys = synthetic_batches(torch, nb, dev)
    seed_ctr = [0]

    def hand_off(src, dst):      # the Alice -> Alice weight hand-off, as alice.train does it
        dst[0].copy_weights_from(src[0])

    from splitlearning_b200 import engine as E
    act_buf = torch.empty(B, D0, device=dev)
    cut_buf = torch.empty(B, D0, device=dev)

    TURN = turn_for(args.steps)

    def step(i, x, y, loss_out, turn=TURN):
        k = (i // turn) % N_ALICES
        if i % turn == 0 and i > 0:
            hand_off(alices[(k - 1) % N_ALICES], alices[k])
        seed_ctr[0] += 1
        # the three calls of alice.train's batch body (forward, Bob's train step, backward+update)
        # through the library's single-call form
        E.vanilla_step(alices[k][1], bob, x, y, k, 1, seed_ctr[0], act_buf, cut_buf, loss_out)

    loss_dev = torch.zeros((), device=dev)
    K, W = args.steps, args.warmup
    # ---- device-resident timing ("value")
    # pre-heat (untimed, disclosed in config): the driver's short run is 25 steps = 4 ms in all, too short
    # for clocks, L2 and lazily loaded kernels to reach the state a training job runs in
    for i in range(PREHEAT):
        step(i, xs[i % nb], ys[i % nb], loss_dev, turn=1)
    # warm-up with a shorter turn, so that the hand-off path (its kernel is loaded lazily on first use)
    # has run before anything is timed
    for i in range(W):
        step(i, xs[i % nb], ys[i % nb], loss_dev, turn=max(1, min(TURN, W // (N_ALICES + 1))))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = lib.sn_launch_count()
    with ClockSampler(0) as clk:
        torch.cuda.synchronize()
        e0.record()
        for i in range(K):
            step(W + i, xs[(W + i) % nb], ys[(W + i) % nb], loss_dev)
        e1.record()
        torch.cuda.synchronize()
    launches = lib.sn_launch_count() - l0
    ms = e0.elapsed_time(e1)
    value = K * B / (ms * 1e-3)
    final_loss = float(loss_dev.item())

    # ---- end-to-end through the same calls with HOST buffers ("e2e")
    # Inputs start in pinned host memory; splitlearning_b200.io copies step t+1's batch on a side stream
    # while step t computes (every step still does its own H2D copy and its own D2H loss read).
    # The loss read: ce_kernel stores each step's loss straight into pinned host memory (UVA-mapped,
    # 4 B over PCIe per step); a cudaMemcpyAsync D2H on the compute stream instead costs 7 us/step
    # (profiles/micro/e2e_variants.py: 193.6 -> 186.6 us/step).
    from splitlearning_b200.io import HostBatchPrefetcher, iterate_host_batches
    hx, hy = synthetic_batches(torch, nb, pin=True)
    hloss = torch.zeros(K + W, dtype=torch.float32).pin_memory()
    prefetcher = HostBatchPrefetcher(dev, B)          # side stream + staging buffers: set up once, like a DataLoader

    # Default: no copy op at all - the step's first kernel reads the pinned batch over PCIe itself, in its
    # prologue (sn_vanilla_step_host), and leaves device copies for the kernels that need the batch again;
    # the launch chain is never broken by an event wait.  SN_E2E=copy: the side-stream prefetcher instead.
    zero_copy = os.environ.get("SN_E2E", "zerocopy") != "copy"
    xd = [torch.empty(B, 1, 28, 28, device=dev) for _ in range(2)]
    yd = [torch.empty(B, dtype=torch.int64, device=dev) for _ in range(2)]

    def e2e_run(first, count):
        if zero_copy:
            for i in range(first, first + count):
                k = (i // TURN) % N_ALICES
                if i % TURN == 0 and i > 0:
                    hand_off(alices[(k - 1) % N_ALICES], alices[k])
                seed_ctr[0] += 1
                E.vanilla_step_host(alices[k][1], bob, hx[i % nb], hy[i % nb], k, 1, seed_ctr[0], xd[i & 1], yd[i & 1], act_buf,
                                    cut_buf, hloss[i])
            return
        host = ((hx[i % nb], hy[i % nb]) for i in range(first, first + count))
        for j, (dx, dy) in enumerate(iterate_host_batches(host, dev, B, prefetcher)):
            step(first + j, dx, dy, hloss[first + j])

    e2e_run(0, W)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    e2e_run(W, K)
    f1.record()
    torch.cuda.synchronize()
    e2e_wall = time.perf_counter() - t0
    e2e_ms = max(f0.elapsed_time(f1), e2e_wall * 1e3)
    e2e_value = K * B / (e2e_ms * 1e-3)

    # ---- per-kernel roofline pass (event-bracketed launches; separate from the headline timing)
    peak, peak_src = measured_peak()
    bob.profile(True)
    torch.cuda.synchronize()
    PK = min(K, 200)
    for i in range(PK):
        step(i, xs[i % nb], ys[i % nb], loss_dev)
    torch.cuda.synchronize()
    t_bwd1, n_bwd1 = bob.profile_read(1, 0)
    t_fwd1, n_fwd1 = bob.profile_read(0, 0)
    kern = {}
    for kind, name in ((0, "fwd"), (1, "bwd_update")):
        for l in range(3):
            t, n = bob.profile_read(kind, l)
            if n:
                kern[f"{name}_fc{l + 1}_us"] = round(1e3 * t / n, 2)
    bob.profile(False)
    dom_bytes = 16.0 * P_FC1 + 4.0 * B * (DIMS[1] + 2 * DIMS[0])     # W r/w + momentum r/w (+ dZ, A in, dA out)
    dom_s = (t_bwd1 / max(n_bwd1, 1)) * 1e-3
    achieved = dom_bytes / dom_s / 1e9
    step_bytes = 16.0 * P_BOB + 8.0 * B * D0
    roofline = {"bound": "hbm", "kernel": "bwd_update_kernel<SGD,dA> fc1 (5000x5408): dW+dX+momentum update in one traversal",
                "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 4),
                # dram bytes of this kernel per launch from the committed `ncu --set full` capture; null
                # when the kernel sources changed after it was taken (profiles/ncu_traffic.json)
                "traffic": ncu_traffic("bwd_update_sgd_da_fc1"), "peak_source": peak_src, "bytes_per_launch": dom_bytes,
                "us_per_launch": round(dom_s * 1e6, 2), "kernel_us": kern,
                "whole_step": {"algorithmic_bytes": step_bytes, "formula": "16*P + 8*B*d0 (SURVEY 8d)",
                               "achieved": round(step_bytes / (ms * 1e-3 / K) / 1e9, 1),
                               "frac": round(step_bytes / (ms * 1e-3 / K) / 1e9 / peak, 4)}}

    # ---- SISA server epoch (BASELINE.json's second number), same process
    sisa = run_sisa_epoch(torch, dev, peak)

    # ---- SISA --concat server (BASELINE.json configs[3] shape) on this GPU
    from splitlearning_b200 import spmd
    sisa_concat = spmd.bench_sisa_concat(0, 1, dev, B, peak)

    # ---- BASELINE.json configs[4]: a few representative points of the batch x width sweep (the full
    # sweep is profiles/sweep.py -> profiles/r2_sweep.jsonl).  Reference precision unless marked.
    sweep = run_sweep_points(torch)

    # ---- CPU baseline on the host cores (bounded sample)
    cpu, sisa_cpu = cpu_baselines()
    sisa.update(sisa_cpu)

    line = {
        "metric": "split_nn_train_samples_per_sec", "value": round(value, 1), "unit": "samples/s", "n_gpus": 1,
        "steps": K, "warmup": W, "ms_per_step": round(ms / K, 5), "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload_config(1, TURN),
        "clocks": clk.summary(), "gpu_launches": int(launches),
        "e2e": {"value": round(e2e_value, 1), "unit": "samples/s", "h2d_bytes_per_step": B * 784 * 4 + B * 8,
                "d2h_bytes_per_step": 4, "ms_per_step": round(e2e_ms / K, 5),
                "how": ("inputs read from pinned host memory by the step's first kernel (4 channel-group CTAs per image: "
                        f"{4 * B * 784 * 4 + B * 8} B cross PCIe per step), loss stored to pinned host memory by the loss kernel"
                        if zero_copy else "inputs copied from pinned host memory on a side stream one step ahead, loss "
                        "stored to pinned host memory by the loss kernel")},
        "roofline": roofline, "cpu_baseline": cpu, "sisa": sisa, "sisa_concat": sisa_concat, "sweep": sweep, "final_loss": final_loss,
    }
    print(json.dumps(line))


def run_sweep_points(torch):
    import importlib.util
    spec = importlib.util.spec_from_file_location("sn_sweep", os.path.join(ROOT, "profiles", "sweep.py"))
    sw = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(sw)
    out = []
    for (b, h, mode) in [(16, 8192, 0), (64, 8192, 0), (256, 8192, 0), (4096, 4096, 0), (4096, 4096, 1)]:
        try:
            out.append(sw.point_record(b, h, mode, steps=8))
        except Exception as e:
            out.append({"B": b, "H": h, "mode": mode, "error": f"{type(e).__name__}: {e}"[:200]})
        torch.cuda.empty_cache()
    return out


def cpu_baselines():
    """The timed CPU baseline next to the GPU numbers (rank 0, N=1 only; ~10 s + ~10 s of CPU work
    plus process start-up).  With oracle/_ref present: the UNMODIFIED reference through its RPC path
    (kind "reference"), and the single-process torch-CPU port as a second, labelled figure."""
    from oracle import cpu_baseline as cb
    from oracle import ref_runner
    v, n, dt, thr = cb.bounded(cb.VanillaCPU, budget_s=6.0, n_alices=N_ALICES, batch=B)
    port = {"value": round(v, 1), "unit": "samples/s", "cores": thr, "kind": "port",
            "sample": f"{n} vanilla split steps of B={B} ({dt:.1f} s) on torch-CPU, single process, no RPC"}
    v2, n2, dt2, _ = cb.bounded(cb.SisaServerCPU, budget_s=5.0, batch=B)
    sisa_cpu = {"cpu_baseline_ms_per_step": round(1e3 * B / v2, 2), "cpu_kind": "port",
                "cpu_sample": f"{n2} SISA server steps of B={B} ({dt2:.1f} s), Adam wd 1e-5, torch-CPU port"}
    if not ref_runner.available():
        return port, sisa_cpu
    try:
        r = ref_runner.time_vanilla_rpc(64, 4, world=N_ALICES + 1, batch=B)
        cpu = {"value": round(r["samples_per_s"], 1), "unit": "samples/s", "cores": r["threads"], "kind": "reference",
               "world_size": N_ALICES + 1,
               "sample": f"{r['steps']} vanilla split steps of B={B} ({r['seconds']:.1f} s) through the unmodified reference "
                         f"(data_entities_vanilla.py over torch.distributed.rpc, {N_ALICES + 1} processes, "
                         f"{r['threads']} intra-op threads each)",
               "port": port}
        q = ref_runner.time_sisa_server(40, batch=B)
        sisa_cpu = {"cpu_baseline_ms_per_step": round(q["ms_per_step"], 2), "cpu_kind": "reference",
                    "cpu_sample": f"{q['steps']} steps of the reference's own bob.train_and_backward "
                                  f"(data_entities_vanilla_sisa.py:294-315) on a pre-filled cache ({q['seconds']:.1f} s, "
                                  f"{q['threads']} threads)",
                    "cpu_port_ms_per_step": sisa_cpu["cpu_baseline_ms_per_step"]}
        return cpu, sisa_cpu
    except Exception as e:                                  # the RPC harness could not start here: say so, keep the port
        port["reference_error"] = f"{type(e).__name__}: {e}"[:300]
        return port, sisa_cpu


def ncu_traffic(kernel_key):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel_key` from the committed ncu
    capture (profiles/ncu_traffic.json, written by profiles/summarize.py from an `ncu --set full`
    run), or None when that kernel's source files have changed since the capture."""
    import hashlib
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            ent = json.load(f)["kernels"][kernel_key]
        h = hashlib.sha256()
        for name in ent["sources"]:
            with open(os.path.join(ROOT, name), "rb") as f:
                h.update(f.read())
        if h.hexdigest()[:16] != ent["source_sha16"]:
            return None
        return float(ent["dram_bytes_per_launch"])
    except Exception:
        return None


def run_sisa_epoch(torch, dev, peak):
    """SISA server-only epoch over frozen activations cached in HBM (BASELINE.json configs[2]:
    ws=5, 56,000 train samples -> 3,500 batches of 16), Adam wd 1e-5, one sn_bob_server_epoch per
    Alice per epoch.  Reports epoch ms (steady state) and the cache-stream / weight-stream rates."""
    from splitlearning_b200 import engine as E
    from splitlearning_b200 import models as M
    torch.manual_seed(1)
    model = M.model2_sisa(dev)
    bob = model.make_engine(optimizer=E.SN_OPT_ADAM, lr=1e-3, weight_decay=1e-5, n_opt_slots=1, max_batch=B)
    amodel = M.model1_sisa(dev)
    alice = amodel.make_engine(optimizer=E.SN_OPT_NONE, max_batch=B)
    sizes = [16271 * 8 // 10, 21685 * 8 // 10, 15329 * 8 // 10, 16715 * 8 // 10]      # LDA K=4 split, 80 % train
    caches = []
    g = torch.Generator().manual_seed(5)
    t0 = torch.cuda.Event(enable_timing=True)
    t1 = torch.cuda.Event(enable_timing=True)
    t0.record()
    for n in sizes:
        x = (torch.randint(0, 256, (n, 1, 28, 28), generator=g).float() * (torch.rand(n, 1, 28, 28, generator=g) < 0.4)).to(dev)
        y = torch.randint(0, 10, (n,), generator=g).to(dev)
        acts = alice.forward(x)                      # frozen Alice, whole shard in one launch
        caches.append((acts, y, list(range(0, n, B)) + [n]))
    t1.record()
    torch.cuda.synchronize()
    # activation cache fill in isolation: one frozen-Alice launch writes a whole shard into HBM
    xs_big = (torch.randint(0, 256, (sizes[1], 1, 28, 28), generator=g).float()).to(dev)
    dst = torch.empty(sizes[1], D0, device=dev)
    alice.forward(xs_big, out=dst)
    torch.cuda.synchronize()
    c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    FILL_REPS = 5                                     # 375 MB written per launch: nothing survives in L2 between them
    c0.record()
    for _ in range(FILL_REPS):
        alice.forward(xs_big, out=dst)
    c1.record()
    torch.cuda.synchronize()
    fill_ms = c0.elapsed_time(c1) / FILL_REPS
    fill_bytes = sizes[1] * (D0 * 4 + 784 * 4)
    fill_fma = sizes[1] * 32 * 169 * 36                # multiply-adds: 36 per output
    fp32_peak_fma = 148 * 128 * 1.965e9                # FFMA and FFMA2 both top out at 128 FMA/clk/SM (profiles/r2_ffma_rate.txt)
    del xs_big, dst
    n_steps = sum(len(c[2]) - 1 for c in caches)
    n_samples = sum(sizes)

    def epoch(seed_off):
        off = seed_off
        for acts, y, offs in caches:
            bob.server_epoch(acts, y, offs, slot=0, masks=None, seed=7, offset=off)
            off += len(offs)
    epoch(0)                                          # warm-up epoch (also epoch 0 of the reference flow)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    epoch(10000)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    # per-kernel times of the Adam traversal (event-bracketed, first cache shard only)
    bob.profile(True)
    acts, y, offs = caches[0]
    bob.server_epoch(acts, y, offs[:201], slot=0, masks=None, seed=7, offset=20000)
    torch.cuda.synchronize()
    kern = {}
    for kind, name in ((0, "fwd"), (1, "bwd_update")):
        for l in range(3):
            t, n = bob.profile_read(kind, l)
            if n:
                kern[f"{name}_fc{l + 1}_us"] = round(1e3 * t / n, 2)
    bob.profile(False)
    step_bytes = 24.0 * P_BOB + 4.0 * B * D0
    adam_fc1 = 24.0 * P_FC1 + 4.0 * B * (DIMS[1] + DIMS[0])
    return {"cache_fill": {"kernel": "conv_pool_stream_warp_kernel over one Alice's whole shard (frozen model -> HBM-resident cache)",
                           "samples": sizes[1], "ms": round(fill_ms, 3), "bytes": fill_bytes,
                           "achieved": round(fill_bytes / (fill_ms * 1e-3) / 1e9, 1), "unit": "GB/s",
                           "frac": round(fill_bytes / (fill_ms * 1e-3) / 1e9 / peak, 4),
                           "bound": "fp32 pipe (36 FMA per 4-byte output)",
                           "fp32_frac": round(fill_fma / (fill_ms * 1e-3) / fp32_peak_fma, 4)},
            "kernel_us": kern,
            "dominant": {"kernel": "bwd_update_kernel<Adam,no dA> fc1", "bytes_per_launch": adam_fc1,
                         "achieved": round(adam_fc1 / (kern.get("bwd_update_fc1_us", 1e9) * 1e-6) / 1e9, 1),
                         "frac": round(adam_fc1 / (kern.get("bwd_update_fc1_us", 1e9) * 1e-6) / 1e9 / peak, 4)},
            "server_epoch_ms": round(ms, 2), "steps": n_steps, "samples": n_samples,
            "us_per_step": round(1e3 * ms / n_steps, 2), "samples_per_sec": round(n_samples / (ms * 1e-3), 1),
            "cache_bytes": n_samples * D0 * 4,
            "roofline": {"bound": "hbm", "algorithmic_bytes_per_step": step_bytes, "formula": "24*P + 4*B*d0 (SURVEY 8d)",
                         "achieved": round(step_bytes * n_steps / (ms * 1e-3) / 1e9, 1), "peak": peak, "unit": "GB/s",
                         "frac": round(step_bytes * n_steps / (ms * 1e-3) / 1e9 / peak, 4)}}


# ----------------------------------------------------------------------------------- our arm, N > 1
def run_ours_multi(args):
    from splitlearning_b200 import spmd
    turn = turn_for(args.steps)
    spmd.bench_vanilla(args, B, N_ALICES, turn, workload_config(args.gpus, turn), measured_peak())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=100)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus > 1 or world > 1:
        return run_ours_multi(args)
    return run_ours_single(args)


if __name__ == "__main__":
    main()
