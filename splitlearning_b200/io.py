"""Host -> device input pipeline: batches that live in pinned host memory are copied to the GPU
on a side stream one step ahead of the compute stream (double buffered), so the H2D copy of step
t+1 overlaps step t's kernels.  The reference reads each batch from a CPU DataLoader and hands CPU
tensors to the model (`data_entities_vanilla.py:67-72`); this is the device-side equivalent of that
hand-off.  PyTorch is used for streams/events/memory only.
"""
import torch


class HostBatchPrefetcher:
    def __init__(self, device, batch, depth=2):
        self.device = device
        self.depth = depth
        self.copy_stream = torch.cuda.Stream(device)
        self.dx = [torch.empty(batch, 1, 28, 28, device=device) for _ in range(depth)]
        self.dy = [torch.empty(batch, dtype=torch.int64, device=device) for _ in range(depth)]
        self.ready = [torch.cuda.Event() for _ in range(depth)]
        self.free = [torch.cuda.Event() for _ in range(depth)]
        cur = torch.cuda.current_stream(device)
        for e in self.free:
            e.record(cur)

    def submit(self, slot, hx, hy):
        """Enqueue the copy of one pinned host batch into buffer `slot` (waits until the compute
        stream has released that buffer)."""
        n = hx.shape[0]
        with torch.cuda.stream(self.copy_stream):
            self.copy_stream.wait_event(self.free[slot])
            self.dx[slot][:n].copy_(hx, non_blocking=True)
            self.dy[slot][:n].copy_(hy, non_blocking=True)
            self.ready[slot].record(self.copy_stream)
        return n

    def acquire(self, slot, n=None):
        torch.cuda.current_stream(self.device).wait_event(self.ready[slot])
        if n is None:
            return self.dx[slot], self.dy[slot]
        return self.dx[slot][:n], self.dy[slot][:n]

    def release(self, slot):
        self.free[slot].record(torch.cuda.current_stream(self.device))


def iterate_host_batches(batches, device, batch_size, prefetcher=None):
    """Yield (x_dev, y_dev) for an iterable of pinned (hx, hy) host batches, prefetching one ahead.
    `prefetcher`: a HostBatchPrefetcher to reuse (its stream, events and staging buffers are set-up
    costs; build it once per loader, not once per epoch)."""
    pf = prefetcher or HostBatchPrefetcher(device, batch_size)
    it = iter(batches)
    nxt = next(it, None)
    slot = 0
    if nxt is not None:
        n_next = pf.submit(slot, *nxt)
    while nxt is not None:
        cur_slot, n_cur = slot, n_next
        nxt = next(it, None)
        if nxt is not None:
            slot = (slot + 1) % pf.depth
            n_next = pf.submit(slot, *nxt)
        yield pf.acquire(cur_slot, n_cur)
        pf.release(cur_slot)
