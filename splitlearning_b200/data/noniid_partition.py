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
"""
import numpy as np


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
                                                  task='classification', rng=None):
    """Return {client: [sample ids]}; same name/arguments as the reference (`:6-10`).

    `rng` defaults to numpy's global legacy generator (what the reference uses); pass a
    `np.random.RandomState` to keep the global stream untouched.
    """
    if task == 'segmentation':
        raise NotImplementedError("segmentation partition is outside the split-NN hot path")
    rng = np.random if rng is None else rng
    labels = np.asarray(label_list)
    n_total = labels.shape[0]

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
