"""Timed CPU baseline: a port of the reference's hot path onto plain torch-CPU modules
(TEST/BENCH INFRASTRUCTURE — see oracle/__init__.py; `bench.py` is the only caller).

The reference itself (`/root/reference`) cannot travel to the GPU box, so its step is restated
here with the SAME torch operators its entity code calls — nn.Conv2d/ReLU/MaxPool2d/Flatten,
nn.Linear/F.relu/nn.Dropout, nn.CrossEntropyLoss, loss.backward(), torch.optim.SGD(momentum=0.9)
or torch.optim.Adam(weight_decay=1e-5) — in ONE process, i.e. WITHOUT the reference's
torch.distributed.rpc / dist_autograd / DistributedOptimizer round trips, which SURVEY.md A.4
measured at ~60 % of its wall time.  The port is therefore FASTER than the reference it stands in
for; speed-ups quoted against it are conservative.  kind = "port".
"""
import time

import torch
import torch.nn as nn
import torch.nn.functional as F


class _Alice(nn.Module):                       # models.py:16-30
    def __init__(self):
        super().__init__()
        self.conv_layers = nn.Sequential(nn.Conv2d(1, 32, 3), nn.ReLU(), nn.MaxPool2d(2, 2))
        self.flatten = nn.Flatten()

    def forward(self, x):
        return self.flatten(self.conv_layers(x))


class _Bob(nn.Module):                         # models.py:46-63
    def __init__(self):
        super().__init__()
        self.fc1, self.fc2, self.fc3 = nn.Linear(5408, 5000), nn.Linear(5000, 1000), nn.Linear(1000, 100)
        self.dropout1, self.dropout2 = nn.Dropout(0.5), nn.Dropout(0.5)

    def forward(self, x):
        x = self.dropout1(F.relu(self.fc1(torch.flatten(x, 1))))
        x = self.dropout2(F.relu(self.fc2(x)))
        return self.fc3(x)


def _synthetic(n, batch, seed=0):
    g = torch.Generator().manual_seed(seed)
    x = torch.randint(0, 256, (n, batch, 1, 28, 28), generator=g).float() * (torch.rand(n, batch, 1, 28, 28, generator=g) < 0.4)
    y = torch.randint(0, 10, (n, batch), generator=g)
    return x, y


class VanillaCPU:
    """data_entities_vanilla.py:67-76 + :226-232, one SGD-momentum optimizer per Alice."""

    def __init__(self, n_alices=4, batch=16, lr=1e-3, n_batches=8):
        torch.manual_seed(0)
        self.bob = _Bob()
        self.alices = [_Alice() for _ in range(n_alices)]
        self.opts = [torch.optim.SGD(list(self.bob.parameters()) + list(a.parameters()), lr=lr, momentum=0.9)
                     for a in self.alices]
        self.crit = nn.CrossEntropyLoss()
        self.x, self.y = _synthetic(n_batches, batch)
        self.i = 0
        self.batch = batch

    def step(self):
        k = (self.i // 8) % len(self.alices)
        j = self.i % self.x.shape[0]
        self.i += 1
        opt = self.opts[k]
        opt.zero_grad()
        act = self.alices[k](self.x[j])
        loss = self.crit(self.bob(act), self.y[j])
        loss.backward()
        opt.step()
        return loss


class SisaServerCPU:
    """data_entities_vanilla_sisa.py:304-313: Adam(wd 1e-5) over cached detached activations."""

    def __init__(self, batch=16, lr=1e-3, n_batches=8):
        torch.manual_seed(0)
        self.bob = _Bob()
        self.opt = torch.optim.Adam(self.bob.parameters(), lr=lr, weight_decay=1e-5)
        self.crit = nn.CrossEntropyLoss()
        g = torch.Generator().manual_seed(1)
        self.acts = (torch.rand(n_batches, batch, 5408, generator=g) * 100) * (torch.rand(n_batches, batch, 5408, generator=g) < 0.6)
        self.y = torch.randint(0, 10, (n_batches, batch), generator=g)
        self.i = 0
        self.batch = batch

    def step(self):
        j = self.i % self.acts.shape[0]
        self.i += 1
        self.opt.zero_grad()
        loss = self.crit(self.bob(self.acts[j].detach()), self.y[j])
        loss.backward()
        self.opt.step()
        return loss


def time_steps(workload, steps, warmup):
    for _ in range(warmup):
        workload.step()
    t0 = time.perf_counter()
    for _ in range(steps):
        workload.step()
    return time.perf_counter() - t0


def bounded(workload_cls, budget_s=15.0, warmup=3, max_steps=2000, **kw):
    """Run about `budget_s` seconds of CPU steps; returns (samples_per_s, steps, seconds, threads)."""
    w = workload_cls(**kw)
    probe = time_steps(w, 3, warmup)
    steps = int(max(5, min(max_steps, budget_s / max(probe / 3, 1e-4))))
    dt = time_steps(w, steps, 0)
    return steps * w.batch / dt, steps, dt, torch.get_num_threads()
