"""A few SISA server steps (Adam, fused next-step forward) with the product library, as an ncu target:
ncu --set full -k regex:bwd_update -s 10 -c 3 python profiles/micro/sisa_steps.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch

from splitlearning_b200 import engine as E, models as M

B, D0 = 16, 5408
dev = torch.device("cuda", 0)
torch.manual_seed(1)
bob = M.model2_sisa(dev).make_engine(optimizer=E.SN_OPT_ADAM, lr=1e-3, weight_decay=1e-5, n_opt_slots=1, max_batch=B)
nbat = 24
acts = torch.rand(nbat * B, D0, device=dev)
labels = torch.randint(0, 10, (nbat * B,), device=dev)
bob.server_epoch(acts, labels, list(range(0, nbat * B + 1, B)), seed=3)
torch.cuda.synchronize()
print("ok")
