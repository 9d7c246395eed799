"""Pieces shared by the four `data_entities*` modules: log files with the reference's line
formats, the legacy `.pt` loader files, device-resident batch iteration that consumes the torch
CPU generator exactly like the reference's DataLoader, and the evaluation loop.
"""
import logging
import os
from collections import Counter

import torch

from . import engine as E


class RngStream:
    """The torch CPU generator state of one logical process.  The reference runs Bob and every
    Alice in separate processes, each with its own generator; here they share a process, so an
    entity built with a seed (`args.torch_seeds = {"bob": 100, "alice_base": 200}`) swaps its own
    state in around everything that draws (model init, loader shuffles, parity-mode dropout
    masks).  Without seeds this is a no-op and all entities draw from the global generator."""

    def __init__(self, seed=None):
        self.state = None
        if seed is not None:
            keep = torch.get_rng_state()
            torch.manual_seed(int(seed))
            self.state = torch.get_rng_state()
            torch.set_rng_state(keep)

    def __enter__(self):
        if self.state is not None:
            self._keep = torch.get_rng_state()
            torch.set_rng_state(self.state)
        return self

    def __exit__(self, *exc):
        if self.state is not None:
            self.state = torch.get_rng_state()
            torch.set_rng_state(self._keep)
        return False


def entity_device(args, rank):
    """Where an entity lives: rank 0 = Bob on `args.bob_device`; Alice r on `args.alice_devices[r]`
    when `split_nn.place_entities` spread the Alices over the GPUs of the box (one process drives
    them all), else on `args.alice_device`, else on the default device."""
    from .models import default_device
    if rank == 0:
        return getattr(args, "bob_device", None) or default_device()
    devs = getattr(args, "alice_devices", None)
    if devs and rank in devs:
        return devs[rank]
    return getattr(args, "alice_device", None) or default_device()


def handoff(src_device, dst_device):
    """Order two GPUs' streams inside one process: everything enqueued so far on `src_device`'s
    current stream happens before whatever is enqueued next on `dst_device`'s.  A no-op on one GPU.
    (Stands where the reference has an RPC boundary: `rpc_sync` returns only when the callee is done.)"""
    if src_device == dst_device:
        return
    ev = torch.cuda.Event()
    ev.record(torch.cuda.current_stream(src_device))
    torch.cuda.current_stream(dst_device).wait_event(ev)


def move(t, device):
    """`t` on `device` (a stream-ordered peer copy when it lives elsewhere)."""
    return t if t.device == device else t.to(device, non_blocking=True)


def entity_seed(args, rank):
    """rank 0 = Bob, r >= 1 = Alice r; None when the caller did not ask for per-entity streams."""
    seeds = getattr(args, "torch_seeds", None)
    if not seeds:
        return None
    return seeds["bob"] if rank == 0 else seeds["alice_base"] + rank


def make_logger(name, filename, first_line):
    """`start_logger` of every entity (e.g. `data_entities_vanilla.py:110-122,259-270`):
    FileHandler mode 'w', format '%(asctime)s: %(message)s', under ./logs (created if absent —
    the reference assumes it exists)."""
    os.makedirs(os.path.dirname(filename) or ".", exist_ok=True)
    logger = logging.getLogger(name)
    logger.setLevel(logging.INFO)
    for h in list(logger.handlers):
        logger.removeHandler(h)
        h.close()
    fh = logging.FileHandler(filename=filename, mode='w')
    fh.setFormatter(logging.Formatter("%(asctime)s: %(message)s"))
    fh.setLevel(logging.INFO)
    logger.addHandler(fh)
    logger.propagate = False
    logger.info(first_line)
    return logger


def load_loader(path):
    """The reference stores whole pickled DataLoaders (`mnist_flat_generator.py:38-41`) and reads
    them back with a bare `torch.load` (`data_entities_vanilla.py:101-102`), which torch >= 2.6
    refuses unless weights_only is off."""
    return torch.load(path, weights_only=False)


class DeviceDataset:
    """A client's TensorDataset staged once in HBM ([n,1,28,28] fp32 + [n] int64) with batch
    iteration that draws from the torch CPU generator exactly as the reference's
    `for data in dataloader` does:
      * every `iter(DataLoader)` draws one int64 (`_BaseDataLoaderIter` base seed);
      * a shuffling loader then draws one more int64 to seed `RandomSampler`'s private generator
        and takes `torch.randperm(n, generator=that)`.
    tests/test_host_logic.py checks this against a real DataLoader."""

    def __init__(self, loader, device):
        ds = loader.dataset
        if isinstance(ds, torch.utils.data.TensorDataset):
            x, y = ds.tensors
        else:                                   # list of (x, y) pairs (the unlearn loaders)
            x = torch.stack([p[0] for p in ds]) if len(ds) else torch.empty(0, 1, 28, 28)
            y = torch.stack([torch.as_tensor(p[1]) for p in ds]) if len(ds) else torch.empty(0, dtype=torch.int64)
        self.x = x.to(device=device, dtype=torch.float32).contiguous()
        self.y = y.to(device=device, dtype=torch.int64).contiguous()
        self.n = int(self.x.shape[0])
        self.batch_size = int(loader.batch_size)
        self.shuffle = isinstance(loader.sampler, torch.utils.data.RandomSampler)
        self.device = device

    @classmethod
    def from_tensors(cls, x, y, batch_size, shuffle, device):
        self = cls.__new__(cls)
        self.x = x.to(device=device, dtype=torch.float32).contiguous()
        self.y = y.to(device=device, dtype=torch.int64).contiguous()
        self.n, self.batch_size, self.shuffle, self.device = int(x.shape[0]), int(batch_size), bool(shuffle), device
        return self

    def epoch_order(self):
        """Consume the generator like one `iter(loader)` + full pass; returns the sample order
        (None = natural order)."""
        torch.empty((), dtype=torch.int64).random_()                 # DataLoader iterator base seed
        if not self.shuffle:
            return None
        seed = int(torch.empty((), dtype=torch.int64).random_().item())
        g = torch.Generator()
        g.manual_seed(seed)
        return torch.randperm(self.n, generator=g)

    def epoch(self, rng=None):
        """(x_epoch, y_epoch) laid out in this epoch's batch order: one gather, then every batch
        is a contiguous slice."""
        if rng is None:
            order = self.epoch_order()
        else:
            with rng:
                order = self.epoch_order()
        if order is None:
            return self.x, self.y
        idx = order.to(self.device)
        return self.x.index_select(0, idx), self.y.index_select(0, idx)

    def batches(self, rng=None):
        x, y = self.epoch(rng)
        for s in range(0, self.n, self.batch_size):
            yield x[s:s + self.batch_size], y[s:s + self.batch_size]

    def labels_list(self):
        return self.y.cpu().tolist()


def log_data_stats(logger, train, test):
    """`load_data` log lines (`data_entities_vanilla.py:104-108`)."""
    logger.info("Local Data Statistics:")
    logger.info("Dataset Size: {:.2f}".format(train.n))
    logger.info("Training dataset: {}".format(dict(Counter(train.labels_list()))))
    logger.info("Test dataset: {}".format(dict(Counter(test.labels_list()))))


def eval_chunk_rows():
    """Rows per evaluation pass (SN_EVAL_CHUNK; 0 = the reference's per-batch loop, kept for A/B runs)."""
    return int(os.environ.get("SN_EVAL_CHUNK", "4096"))


def eval_loop(forward_fn, test, omit_label, rng=None, chunk_rows=None):
    """Shared body of `alice.eval` / `alice.eval_breakdown` (`data_entities_vanilla.py:82-98,
    159-202`): activations -> Bob -> first-argmax -> six counters, accumulated on the device and
    read back once.  The reference walks the test loader 16 samples at a time (an RPC per batch);
    here a whole chunk of the shard (`chunk_rows`, default 4096) goes through in one pass: one conv
    launch, one launch (+ its split-K reduction) per server layer, one counter launch.  Every row
    still gets exactly the numbers the batch-of-16 loop gives it (the server's forward runs the same
    fp32 weight-stream kernels in row tiles of 16), so the counters are identical;
    tests/test_gpu_entities.py checks that.  The generator is consumed as by one pass over the loader."""
    chunk = eval_chunk_rows() if chunk_rows is None else int(chunk_rows)
    counters = torch.zeros(6, dtype=torch.int64, device=test.device)
    omit = -1 if omit_label is None else omit_label
    if chunk <= test.batch_size:
        for x, y in test.batches(rng):
            E.eval_counts(forward_fn(x), y, omit, counters)
    else:
        x, y = test.epoch(rng)
        chunk -= chunk % test.batch_size                      # whole loader batches per pass
        for s in range(0, test.n, chunk):
            E.eval_counts(forward_fn(x[s:s + chunk]), y[s:s + chunk], omit, counters)
    return [int(v) for v in counters.cpu().tolist()]


def log_breakdown(logger, client_id, omit_label, c, vanilla_style):
    correct, total, cu, tu, cr, tr = c
    logger.info(f"Alice-{client_id} Evaluating Data: {round(correct / total, 3)}")
    unl = round(cu / tu if tu else 0, 3)
    if vanilla_style:     # data_entities_vanilla.py:199 adds the raw counts on a second line
        logger.info(f"Alice-{client_id} Evaluating Unlearned label-{omit_label}: {unl} \n correct_unlearned: {cu}, total_unlearned: {tu}")
    else:                 # data_entities_vanilla_sisa.py:134
        logger.info(f"Alice-{client_id} Evaluating Unlearned label-{omit_label}: {unl}")
    logger.info(f"Alice-{client_id} Evaluating Remaining labels: {round(cr / tr if tr else 0, 3)}")


def filter_label(train, omit_label, rng=None):
    """The unlearn/control data filter (`data_entities_vanilla.py:140-143`): one pass over the
    (shuffling) train loader, drop label == omit, re-batch in that order, no shuffle."""
    x, y = train.epoch(rng)
    keep = y != omit_label
    return DeviceDataset.from_tensors(x[keep], y[keep], train.batch_size, False, train.device)


def bob_log_accuracy(logger, results):
    """`bob.eval_request_breakdown` summary lines (`data_entities_vanilla.py:299-301`);
    same ZeroDivisionError as the reference when no client holds the omitted label."""
    corr, tot, cu, tu, cr, tr = (sum(r[i] for r in results) for i in range(6))
    logger.info("Accuracy over all data: {:.3f}".format(corr / tot))
    logger.info("Accuracy over unlearned data: {:.3f}".format(cu / tu))
    logger.info("Accuracy over remaining data: {:.3f}".format(cr / tr))
    return corr / tot, cu / tu, cr / tr
