"""LDA (Dirichlet) non-IID partition, host side, bit-exact with the reference for the same seed.

Mirrors `/root/reference/data/noniid_partition.py:6-73` (driver) and `:76-91` (per-class split).
Only the classification branch is implemented: the reference's caller always passes
`task='task'` (`relational_table_preprocessor.py:62-66`), and anything other than
`'segmentation'` takes the classification branch (`noniid_partition.py:40,47,62`).

Bit-exactness depends on consuming numpy's *legacy global* MT19937 stream in exactly this order:
    for every restart of the min-size loop:
        for class k in 0..classes-1:  shuffle(idx_k ndarray) ; dirichlet([alpha]*client_num)
    for client i in 0..client_num-1:  shuffle(python list)
Shuffling an ndarray and shuffling a python list of the same length draw different numbers of
variates, so the container types below are part of the contract.

Two implementations, equal bit for bit (tests/test_host_logic.py): the numpy one below, and the
native routine `sn_lda_partition` (csrc/lda_partition.cpp) that takes numpy's MT19937 state in and
hands it back advanced.  `native=None` picks the native one when libsplitnn.so is there.
"""
import ctypes as C

import numpy as np


def _native_lib():
    from .. import _lib
    try:
        return _lib.load()
    except (_lib.SplitNNError, OSError):
        return None


def _partition_native(lib, labels, client_num, classes, alpha, rng):
    """The whole partition in one library call; numpy's generator state goes in and comes back."""
    name, key, pos, has_gauss, cached = rng.get_state()
    key = np.ascontiguousarray(key, dtype=np.uint32).copy()
    labels = np.ascontiguousarray(labels, dtype=np.int64)
    n = labels.shape[0]
    ids = np.empty(n, dtype=np.int64)
    offsets = np.empty(client_num + 1, dtype=np.int64)
    c_pos, c_has, c_gauss = C.c_int(int(pos)), C.c_int(int(has_gauss)), C.c_double(float(cached))
    rc = lib.sn_lda_partition(labels.ctypes.data, n, int(client_num), int(classes), float(alpha), key.ctypes.data,
                              C.byref(c_pos), C.byref(c_has), C.byref(c_gauss), ids.ctypes.data, offsets.ctypes.data)
    if rc != 0:
        raise ValueError("sn_lda_partition: invalid argument")
    rng.set_state((name, key, c_pos.value, c_has.value, c_gauss.value))
    return {i: ids[offsets[i]:offsets[i + 1]].tolist() for i in range(client_num)}


def _split_one_class(n_total, alpha, client_num, shards, class_idx, rng):
    """One class's samples spread over the clients (reference `noniid_partition.py:76-91`)."""
    rng.shuffle(class_idx)                                   # ndarray shuffle (:77)
    share = rng.dirichlet(np.repeat(alpha, client_num))      # (:80)
    # a client that already holds >= N/client_num samples gets nothing from this class (:83)
    cap = n_total / client_num
    share = np.array([s * (len(held) < cap) for s, held in zip(share, shards)])
    share = share / share.sum()                              # (:84)
    cuts = (np.cumsum(share) * len(class_idx)).astype(int)[:-1]   # (:85)
    pieces = np.split(class_idx, cuts)                       # (:88)
    shards = [held + piece.tolist() for held, piece in zip(shards, pieces)]
    return shards, min(len(s) for s in shards)


def non_iid_partition_with_dirichlet_distribution(label_list, client_num, classes, alpha,
                                                  task='classification', rng=None, native=None):
    """Return {client: [sample ids]}; same name/arguments as the reference (`:6-10`).

    `rng` defaults to numpy's global legacy generator (what the reference uses); pass a
    `np.random.RandomState` to keep the global stream untouched.  `native`: True = the library
    routine, False = numpy, None = the library when it loads.
    """
    if task == 'segmentation':
        raise NotImplementedError("segmentation partition is outside the split-NN hot path")
    rng = np.random if rng is None else rng
    labels = np.asarray(label_list)
    n_total = labels.shape[0]
    integral = n_total > 0 and labels.ndim == 1 and np.issubdtype(labels.dtype, np.integer) and hasattr(rng, "get_state")
    if native is not False and integral:
        lib = _native_lib()
        if lib is not None:
            return _partition_native(lib, labels, client_num, classes, alpha, rng)
    if native:
        raise RuntimeError("native LDA partition needs libsplitnn.so, 1-D integer labels and a legacy numpy generator")

    smallest = 0
    while smallest < 10:                                     # min-size guarantee (:43-44)
        shards = [[] for _ in range(client_num)]
        for k in range(classes):                             # (:64-68)
            class_idx = np.where(labels == k)[0]
            shards, smallest = _split_one_class(n_total, alpha, client_num, shards, class_idx, rng)

    out = {}
    for i in range(client_num):                              # list shuffle (:69-71)
        rng.shuffle(shards[i])
        out[i] = shards[i]
    return out


def record_data_stats(y_train, net_dataidx_map, task='classification'):
    """Per-client class histogram (reference `:94-103`, unused by the hot path)."""
    y = np.asarray(y_train)
    stats = {}
    for client, ids in net_dataidx_map.items():
        values, counts = np.unique(y[ids], return_counts=True)
        stats[client] = {int(v): int(c) for v, c in zip(values, counts)}
    return stats
