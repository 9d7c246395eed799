import os, sys, torch
sys.path.insert(0, os.environ.get("GRAFT_REPO_ROOT", "/root/repo"))
from splitlearning_b200 import engine as E, models as M
dev = torch.device("cuda", 0)
am = M.model1_sisa(dev)
alice = am.make_engine(optimizer=E.SN_OPT_NONE, max_batch=16)
n = 17348
g = torch.Generator().manual_seed(1)
x = (torch.randint(0, 256, (n, 1, 28, 28), generator=g).float()).to(dev)
out = torch.empty(n, 5408, device=dev)
for _ in range(3): alice.forward(x, out=out)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): alice.forward(x, out=out)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 20
small = torch.cat([alice.forward(x[s:s + 16]) for s in range(0, 640, 16)])
print(os.environ.get("SN_CONV_STREAM", "warp"), os.environ.get("SN_CONV_WCTAS", "-"), os.environ.get("SN_CONV_CTAS", "-"), "ms", round(ms, 4), "GB/s", round(n * (5408 + 784) * 4 / ms / 1e6, 1), "bit-equal to the 16-sample kernel:", torch.equal(small, out[:640]))
