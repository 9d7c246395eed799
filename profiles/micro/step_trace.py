"""In-chain cost of every kernel of the vanilla step.  Needs the debug build
    SN_OUT=profiles/micro/libsplitnn_trace.so SN_NVCC_EXTRA=-DSN_TRACE sh splitlearning_b200/csrc/build.sh
in which CTA 0 of each kernel stamps %globaltimer right after its PDL dependency wait (= when the
previous kernel has completed).  Differences of consecutive stamps are what each kernel costs the
step with launch overlap as it really is - ncu serialises launches and cannot show that.
usage: python profiles/micro/step_trace.py [vanilla|sisa|local]
       python -m torch.distributed.run --nproc-per-node G ... profiles/micro/step_trace.py tp   (tensor-parallel SISA steps)"""
import ctypes as C
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import numpy as np
import torch

from splitlearning_b200 import _lib

_lib.LIB_PATH = os.environ.get("SN_TRACE_LIB", os.path.join(HERE, "libsplitnn_trace.so"))
import bench
from splitlearning_b200 import engine as E

mode = sys.argv[1] if len(sys.argv) > 1 else "vanilla"
lib = _lib.load()
lib.sn_trace_read.restype = C.c_int
lib.sn_trace_read.argtypes = [C.c_void_p, C.c_void_p]
B, D0, N_ALICES = bench.B, bench.D0, bench.N_ALICES
dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)


def read_trace():
    buf = np.zeros(8192, dtype=np.uint64)
    n = C.c_uint(0)
    assert lib.sn_trace_read(buf.ctypes.data, C.byref(n)) == 0
    return buf[:min(n.value, 8192)].astype(np.int64)


if mode == "tp":
    import torch.distributed as dist
    from splitlearning_b200 import models as M, tp
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    torch.manual_seed(1)
    model = M.model2_sisa(dev)
    srv = tp.TPServer(rank, world, dev, [model.fc1.weight.data, model.fc2.weight.data, model.fc3.weight.data],
                      [model.fc1.bias.data, model.fc2.bias.data, model.fc3.bias.data], B=B)
    nbat = 140
    g = torch.Generator().manual_seed(3)
    acts = torch.rand(nbat * B, D0, generator=g).to(dev)
    labels = torch.randint(0, 10, (nbat * B,), generator=g).to(dev)
    offs = list(range(0, nbat * B + 1, B))
    srv.server_epoch(acts, labels, offs)
    torch.cuda.synchronize(); dist.barrier()
    read_trace()
    srv.server_epoch(acts, labels, offs)
    torch.cuda.synchronize(); dist.barrier()
    t = read_trace()
    L = len(t) // nbat
    d = np.diff(t[: (len(t) // L) * L])
    d = d[: (len(d) // L) * L].reshape(-1, L)[5:-2]
    med = np.median(d, axis=0) / 1e3
    for r in range(world):
        if r == rank:
            print(f"rank {rank}: {L} launches/step, step {med.sum():.1f} us:", " ".join(f"{m:.1f}" for m in med), flush=True)
        dist.barrier()
    srv.close()
    dist.destroy_process_group()
    sys.exit(0)
if mode == "vanilla":
    bob_model, bob, alices = bench.make_world(torch, dev)
    xs, ys = bench.synthetic_batches(torch, 64, dev)
    act, cut = torch.empty(B, D0, device=dev), torch.empty(B, D0, device=dev)
    loss = torch.zeros((), device=dev)
    names = None      # one name per launch of a step; filled from the stamp count below

    def run(n):
        for i in range(n):
            E.vanilla_step(alices[0][1], bob, xs[i % 64], ys[i % 64], 0, 1, i + 1, act, cut, loss)
elif mode == "concat":
    # BASELINE.json config 4: the concat server (fc1 5000 x 37856), Adam, fused server steps
    from splitlearning_b200 import models as M
    torch.manual_seed(0)
    cm = M.model2_sisa_concat(7, dev)
    bob = cm.make_engine(optimizer=E.SN_OPT_ADAM, lr=1e-3, weight_decay=1e-5, n_opt_slots=1, max_batch=B)
    bob.set_ce_heads(7)
    nbat = 140
    acts = torch.rand(nbat * B, D0 * 7, device=dev)
    labels = torch.randint(0, 10, (nbat * B * 7,), device=dev)
    offs = list(range(0, nbat * B + 1, B))
    names = None

    def run(n):
        bob.server_epoch(acts, labels, offs, seed=3)
elif mode == "local":
    # SISA phase 1: an Alice's local epoch (conv fwd -> CE over the 5408 activations -> conv bwd + Adam)
    from splitlearning_b200 import models as M
    torch.manual_seed(0)
    am = M.model1_sisa(dev)
    alice = am.make_engine(optimizer=E.SN_OPT_ADAM, lr=1e-3, weight_decay=1e-5, max_batch=B)
    xl = (torch.randint(0, 256, (140 * B, 1, 28, 28)).float() * (torch.rand(140 * B, 1, 28, 28) < 0.4)).to(dev)
    yl = torch.randint(0, 10, (140 * B,)).to(dev)
    names = None

    def run(n):
        alice.local_epoch(xl, yl, B)
else:
    from splitlearning_b200 import models as M
    torch.manual_seed(1)
    model = M.model2_sisa(dev)
    bob = model.make_engine(optimizer=E.SN_OPT_ADAM, lr=1e-3, weight_decay=1e-5, n_opt_slots=1, max_batch=B)
    nbat = 140
    acts = torch.rand(nbat * B, D0, device=dev)
    labels = torch.randint(0, 10, (nbat * B,), device=dev)
    offs = list(range(0, nbat * B + 1, B))
    names = None

    def run(n):
        bob.server_epoch(acts, labels, offs, seed=3)

run(60)
read_trace()
run(140)
t = read_trace()
L = len(t) // 140                               # launches per step
KNOWN = {6: ["conv_fwd", "fwd_fc1", "tail(epi1,fc2,epi2,fc3,ce+bwd3,bwd2,bepi2)", "bwd_fc1", "bepi_fc1", "conv_bwd"],
         12: ["conv_fwd", "fwd_fc1", "tail:epi_fc1", "tail:fwd_fc2", "tail:epi_fc2", "tail:fwd_fc3", "tail:ce+bwd_fc3", "tail:bwd_fc2",
              "tail:bepi_fc2", "bwd_fc1", "bepi_fc1", "conv_bwd"] if os.environ.get("SN_TAIL") else
             ["conv_fwd", "fwd_fc1", "epi_fc1", "fwd_fc2", "epi_fc2", "fc3+ce", "bwd_fc3", "bwd_fc2", "bepi_fc2",
              "bwd_fc1", "bepi_fc1", "conv_bwd"],
         11: ["conv_fwd", "fwd_fc1", "epi_fc1", "fwd_fc2", "epi_fc2", "fc3+ce", "bwd_fc3", "bwd_fc2", "bepi_fc2", "bwd_fc1",
              "conv_bwd(+cut partial sums)"],
         10: ["conv_fwd", "fwd_fc1+epi", "fwd_fc2+epi", "fc3+ce", "bwd_fc3", "bwd_fc2", "bepi_fc2", "bwd_fc1", "bepi_fc1",
              "conv_bwd"],
         14: ["conv_fwd", "fwd_fc1", "epi_fc1", "fwd_fc2", "epi_fc2", "fwd_fc3", "ce", "bwd_fc3", "bepi_fc3", "bwd_fc2",
              "bepi_fc2", "bwd_fc1", "bepi_fc1", "conv_bwd"]}
names = KNOWN.get(L) if mode == "vanilla" else None
if names is None:
    names = [f"k{i}" for i in range(L)]
print("stamps", len(t), "per step", len(t) / 140)
t = t[: (len(t) // L) * L]
d = np.diff(t)                                  # d[j] = cost of kernel j (stamp j -> stamp j+1)
d = d[: (len(d) // L) * L].reshape(-1, L)[5:]   # drop the first steps
print("globaltimer granularity (smallest non-zero diff): %d ns" % d[d > 0].min())
med = np.median(d, axis=0) / 1e3
for nm, m in zip(names, med):
    print(f"  {nm:10s} {m:7.2f} us")
print("  sum       %7.2f us" % med.sum())
if hasattr(lib, "sn_trace_phase_read"):
    ph = np.zeros(16, dtype=np.int64)
    lib.sn_trace_phase_read.argtypes = [C.c_void_p]
    if lib.sn_trace_phase_read(ph.ctypes.data) == 0 and ph[0]:
        nz = ph[ph > 0]
        print("  phase stamps of the instrumented kernel (clock64 cycles from entry):", (nz - nz[0]).tolist())
        print("  raw phase slots:", ph.tolist())
