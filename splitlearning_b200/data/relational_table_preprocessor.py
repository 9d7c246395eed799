"""Partition -> 80/20 split -> TensorDataset -> DataLoader per client (host side).

Mirrors `/root/reference/data/relational_table_preprocessor.py:57-104` (`image_preprocess_dl`)
and `:8-55` (`relational_table_preprocess_dl`, the flat-784 twin).  The two differ only in how
rows are indexed, so both go through `_split_clients`.  RNG call order that must be preserved
for a bit-exact split: the LDA partition first (numpy legacy global stream), then one
`random.shuffle(list)` per client in ascending client order (`:83`), `N_train = int(N*(1-test))`
(`:80`).  The returned list has the reference's 8 positions (`:103-104`).  Both shuffles run in the
library when it is there (`sn_lda_partition`, `sn_py_random_shuffle`: the generators' states go in
and come back advanced exactly as numpy / `random` would have left them).
"""
import ctypes as C
import random

import numpy as np
import torch
from torch.utils.data import ConcatDataset, DataLoader, TensorDataset

from .noniid_partition import _native_lib, non_iid_partition_with_dirichlet_distribution


def py_random_shuffle(ids, native=None):
    """`random.shuffle(ids)` on a list of sample ids, in place, on Python's global generator."""
    lib = _native_lib() if native is not False else None
    if lib is None:
        if native:
            raise RuntimeError("native shuffle needs libsplitnn.so")
        random.shuffle(ids)
        return
    version, internal, gauss_next = random.getstate()
    key = np.array(internal[:624], dtype=np.uint32)
    pos = C.c_int(internal[624])
    arr = np.asarray(ids, dtype=np.int64)
    if lib.sn_py_random_shuffle(arr.ctypes.data, arr.shape[0], key.ctypes.data, C.byref(pos)) != 0:
        raise ValueError("sn_py_random_shuffle: invalid argument")
    random.setstate((version, tuple(key.tolist()) + (pos.value,), gauss_next))
    ids[:] = arr.tolist()


def _split_clients(args, data, label_list, test_partition):
    labels = torch.LongTensor(label_list)
    data = torch.FloatTensor(data)
    owner = non_iid_partition_with_dirichlet_distribution(
        label_list=labels, client_num=args.client_num_in_total, classes=args.class_num,
        alpha=args.partition_alpha, task='task')

    train_loaders, test_loaders, train_sizes = {}, {}, {}
    train_sets, test_sets = [], []
    n_train_total = n_test_total = 0
    for client, ids in owner.items():
        n_client = len(ids)
        n_train = int(n_client * (1 - test_partition))
        train_sizes[client] = n_train
        py_random_shuffle(ids)
        head, tail = ids[:n_train], ids[n_train:]
        train_set = TensorDataset(data[head], labels[head])
        test_set = TensorDataset(data[tail], labels[tail])
        train_loaders[client] = DataLoader(train_set, batch_size=args.batch_size, shuffle=True)
        test_loaders[client] = DataLoader(test_set, batch_size=args.batch_size)
        train_sets.append(train_set)
        test_sets.append(test_set)
        n_train_total += n_train
        n_test_total += n_client - n_train

    train_global = DataLoader(ConcatDataset(train_sets), batch_size=args.batch_size)
    test_global = DataLoader(ConcatDataset(test_sets), batch_size=args.batch_size)
    return [n_train_total, n_test_total, train_global, test_global,
            train_sizes, train_loaders, test_loaders, args.class_num]


def image_preprocess_dl(args, data, label_list, test_partition=0.2):
    """`data` is [N,1,28,28]; same signature/return as the reference (`:57`)."""
    return _split_clients(args, data, label_list, test_partition)


def relational_table_preprocess_dl(args, data, label_list, test_partition=0.2):
    """`data` is [N,784]; same signature/return as the reference (`:8`)."""
    return _split_clients(args, data, label_list, test_partition)
