"""Runs the UNMODIFIED reference (a byte-for-byte copy under oracle/_ref, see oracle/fetch_ref.sh)
on the host cores and times it.  TEST/BENCH INFRASTRUCTURE: only bench.py calls this (its
`--impl reference` arm and its `cpu_baseline` leg); nothing under splitlearning_b200/ does.

What is timed is the reference's own CPU RPC path: its `alice` / `bob` entity classes
(`data_entities_vanilla.py`) in world_size processes under `torch.distributed.rpc` (TensorPipe),
`dist_autograd` and `DistributedOptimizer`, driven in the call order of `split_nn.example`'s
vanilla branch (`split_nn.py:47-52`: `BOB.train_request(client_id)` for each client).  `example()`
itself is not entered: it pins MASTER_ADDR to "localhost" with a fixed port, fetches MNIST from
OpenML at import (`split_nn.py:5`, no network here) and goes on to the evaluation / unlearning
phases, which the samples/s metric excludes.  Harness shims, reference untouched
(SURVEY.md 8c): synthetic `.pt` loaders written with the reference's own `image_preprocess_dl`;
TORCH_FORCE_NO_WEIGHTS_ONLY_LOAD=1 for its pickled DataLoaders; a time stamp taken on entry to
`bob.train_and_backward` (the one point every step passes) by wrapping the bound method.

SISA's server epoch is the reference's own `bob.train_and_backward`
(`data_entities_vanilla_sisa.py:294-315`) run in one process with a pre-filled activation cache
(SURVEY.md A.5).
"""
import json
import os
import sys
import tempfile
import time
from argparse import Namespace

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "_ref")
REPO = os.path.dirname(HERE)


def available():
    return os.path.isfile(os.path.join(REF, "data_entities_vanilla.py"))


def host_threads():
    return len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)


def _free_port():
    import socket
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


class _CleanLaunchEnv:
    """The reference's processes must rendezvous among themselves: when bench.py itself runs under
    torchrun, the launcher's variables (TORCHELASTIC_USE_AGENT_STORE makes env:// rendezvous a CLIENT of
    the launcher's store; RANK / WORLD_SIZE / MASTER_*; OMP_NUM_THREADS=1) must not leak into them."""
    NAMES = ("RANK", "LOCAL_RANK", "WORLD_SIZE", "LOCAL_WORLD_SIZE", "GROUP_RANK", "GROUP_WORLD_SIZE", "ROLE_RANK",
             "ROLE_WORLD_SIZE", "ROLE_NAME", "MASTER_ADDR", "MASTER_PORT", "OMP_NUM_THREADS", "NCCL_ASYNC_ERROR_HANDLING",
             "TORCH_NCCL_ASYNC_ERROR_HANDLING")

    def __enter__(self):
        self.saved = {k: os.environ.pop(k) for k in list(os.environ) if k in self.NAMES or k.startswith("TORCHELASTIC_")}
        return self

    def __exit__(self, *exc):
        os.environ.update(self.saved)
        return False


def _args(world, batch, epochs=1, server_epochs=1):
    return Namespace(world_size=world, epochs=epochs, iterations=1, batch_size=batch, partition_alpha=0.5,
                     datapath="data/mnist_flat", lr=1e-3, server_epochs=server_epochs, vanilla=True, sisa=False,
                     concat=False, control=False, client_num_in_total=world - 1, class_num=10)


def write_loaders(workdir, n_samples, n_clients, batch):
    """Synthetic MNIST-shaped `.pt` loaders through the REFERENCE's `image_preprocess_dl`
    (LDA partition alpha 0.5, 20 % test), seeds as SURVEY.md 8d.  Returns per-client train sizes."""
    import random

    import numpy as np
    import torch
    sys.path.insert(0, REF)
    sys.path.insert(0, REPO)
    from data.relational_table_preprocessor import image_preprocess_dl          # the reference's
    from splitlearning_b200.data.synthetic import make_synthetic_mnist
    data, labels = make_synthetic_mnist(n_samples, seed=1234)
    a = Namespace(client_num_in_total=n_clients, class_num=10, partition_alpha=0.5, batch_size=batch)
    np.random.seed(0)
    random.seed(0)
    parts = image_preprocess_dl(a, data.astype(np.float64), labels.tolist(), 0.2)
    os.makedirs(os.path.join(workdir, "data/mnist_flat"), exist_ok=True)
    os.makedirs(os.path.join(workdir, "logs"), exist_ok=True)
    sizes = []
    for k in sorted(parts[5]):
        torch.save(parts[5][k], os.path.join(workdir, f"data/mnist_flat/data_worker{k + 1}_train.pt"))
        torch.save(parts[6][k], os.path.join(workdir, f"data/mnist_flat/data_worker{k + 1}_test.pt"))
        sizes.append(len(parts[5][k].dataset))
    return sizes


def _vanilla_worker(rank, world, workdir, port, batch, threads, out_path):
    sys.stdout = sys.stderr                          # bench.py's stdout carries exactly one JSON line
    os.chdir(workdir)
    sys.path.insert(0, REF)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    os.environ["TORCH_FORCE_NO_WEIGHTS_ONLY_LOAD"] = "1"
    import torch
    import torch.distributed.rpc as rpc
    # the reference as shipped lets every process use torch's default intra-op threads (= all cores;
    # only one process computes at a time in this protocol); torchrun's OMP_NUM_THREADS=1 must not leak in
    torch.set_num_threads(threads)
    torch.manual_seed(100 if rank == 0 else 200 + rank)
    args = _args(world, batch)
    if rank != 0:
        rpc.init_rpc(f"alice{rank}", rank=rank, world_size=world)
        rpc.shutdown()
        return
    rpc.init_rpc("bob", rank=0, world_size=world)
    import data_entities_vanilla as ent                                        # the reference's
    BOB = ent.bob(args)
    stamps = []
    orig = BOB.train_and_backward

    def stamped(x, labels, ctx):
        stamps.append(time.perf_counter())
        return orig(x, labels, ctx)
    BOB.train_and_backward = stamped
    for client_id in range(1, world):                                         # split_nn.py:49-51
        BOB.train_request(client_id)
    stamps.append(time.perf_counter())
    with open(out_path, "w") as f:
        json.dump(stamps, f)
    rpc.shutdown()


def time_vanilla_rpc(steps, warmup, world=5, batch=16, threads=None):
    """K = `steps` timed vanilla split steps after `warmup` untimed ones through the reference's RPC
    path.  Returns dict(seconds, steps, warmup, samples_per_s, threads, world, total_steps)."""
    import torch.multiprocessing as mp
    threads = threads or host_threads()
    need = steps + warmup + 1
    # LDA shards are uneven; every client rounds its last batch up, so ~need*batch*1.25/0.8 samples
    # give at least `need` steps with margin (the run is cut to exactly `steps` by the time stamps)
    n_samples = max(400, int(need * batch / 0.8 * 1.15) + 40 * (world - 1))
    with tempfile.TemporaryDirectory() as d:
        sizes = write_loaders(d, n_samples, world - 1, batch)
        total = sum((n + batch - 1) // batch for n in sizes)
        out = os.path.join(d, "stamps.json")
        with _CleanLaunchEnv():
            mp.spawn(_vanilla_worker, args=(world, d, _free_port(), batch, threads, out), nprocs=world, join=True)
        with open(out) as f:
            stamps = json.load(f)
    n_steps = len(stamps) - 1                       # the last stamp closes the last step
    warmup = min(warmup, max(0, n_steps - 1))
    steps = min(steps, n_steps - warmup)
    dt = stamps[warmup + steps] - stamps[warmup]
    return {"seconds": dt, "steps": steps, "warmup": warmup, "samples_per_s": steps * batch / dt, "threads": threads,
            "world": world, "total_steps": n_steps, "shard_train_sizes": sizes, "expected_steps": total}


def _sisa_worker(rank, n_batches, batch, threads, port, out_path):
    sys.stdout = sys.stderr
    sys.path.insert(0, REF)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    wd = tempfile.mkdtemp()
    os.makedirs(os.path.join(wd, "logs"), exist_ok=True)
    os.chdir(wd)
    import torch
    import torch.distributed.rpc as rpc
    torch.set_num_threads(threads)
    torch.manual_seed(100)
    rpc.init_rpc("bob", rank=0, world_size=1)
    import data_entities_vanilla_sisa as ent                                   # the reference's
    BOB = ent.bob(Namespace(server_epochs=1, client_num_in_total=0, lr=1e-3))
    BOB.client_num_in_total = 1
    g = torch.Generator().manual_seed(1)
    acts = [(torch.rand(batch, 5408, generator=g) * 100) * (torch.rand(batch, 5408, generator=g) < 0.6) for _ in range(n_batches)]
    labs = [torch.randint(0, 10, (batch,), generator=g) for _ in range(n_batches)]
    BOB.activation_and_labels_cache[(1, False, None)] = (acts[:3], labs[:3])
    BOB.train_and_backward([], None)                                          # warm-up: 3 steps
    BOB.activation_and_labels_cache[(1, False, None)] = (acts, labs)
    t0 = time.perf_counter()
    BOB.train_and_backward([], None)
    dt = time.perf_counter() - t0
    with open(out_path, "w") as f:
        json.dump({"seconds": dt, "steps": n_batches}, f)
    rpc.shutdown()


def time_sisa_server(n_batches, batch=16, threads=None):
    """One pass of the reference's `bob.train_and_backward` over `n_batches` cached batches."""
    import torch.multiprocessing as mp
    threads = threads or host_threads()
    with tempfile.TemporaryDirectory() as d:
        out = os.path.join(d, "sisa.json")
        with _CleanLaunchEnv():
            mp.spawn(_sisa_worker, args=(n_batches, batch, threads, _free_port(), out), nprocs=1, join=True)
        with open(out) as f:
            r = json.load(f)
    r["ms_per_step"] = 1e3 * r["seconds"] / r["steps"]
    r["threads"] = threads
    return r


if __name__ == "__main__":
    k = int(sys.argv[1]) if len(sys.argv) > 1 else 20
    print(json.dumps(time_vanilla_rpc(k, 5)))
    print(json.dumps(time_sisa_server(20)))
