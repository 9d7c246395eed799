"""A/B of sn_alice_local_epoch (SISA phase 1: conv fwd -> CE over 5408 'classes' -> conv bwd + Adam)
between two builds of the library.  usage: python local_epoch_ab.py <lib.so> ..."""
import os
import sys
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
if len(sys.argv) > 2:
    for lib in sys.argv[1:]:
        subprocess.run([sys.executable, __file__, lib])
    sys.exit(0)
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import torch
from splitlearning_b200 import _lib
_lib.LIB_PATH = os.path.abspath(sys.argv[1])
_lib.SIGNATURES = {k: v for k, v in _lib.SIGNATURES.items() if "skinny_tensor" not in k}
from splitlearning_b200 import engine as E, models as M
dev = torch.device("cuda", 0)
torch.manual_seed(0)
m = M.model1_sisa(dev)
a = m.make_engine(optimizer=E.SN_OPT_ADAM, lr=1e-3, weight_decay=1e-5, max_batch=16)
n = 14000
x = (torch.randint(0, 256, (n, 1, 28, 28)).float() * (torch.rand(n, 1, 28, 28) < 0.4)).to(dev)
y = torch.randint(0, 10, (n,)).to(dev)
a.local_epoch(x, y, 16)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
a.local_epoch(x, y, 16)
e1.record()
torch.cuda.synchronize()
print(sys.argv[1], "local epoch %.2f ms, %.1f us/step" % (e0.elapsed_time(e1), 1e3 * e0.elapsed_time(e1) / (n / 16)))
