"""Writes `data_worker{k}_{train,test}.pt` (pickled DataLoaders, the reference's on-disk format).

Mirrors `/root/reference/data/mnist_flat/mnist_flat_generator.py:26-41` (`load_mnist_image`).
The reference fetches `mnist_784` from OpenML (`:27`); there is no network here, so the pixels
come from `fetch_mnist()`: a local `mnist_784.npz` under `args.datapath` if one exists (keys
`data` [N,784] or [N,1,28,28], `target` [N]), else the synthetic MNIST-shaped set.
"""
import os
from collections import Counter

import numpy as np
import torch

from ..relational_table_preprocessor import image_preprocess_dl, relational_table_preprocess_dl
from ..synthetic import make_synthetic_mnist


def fetch_mnist(args):
    local = os.path.join(args.datapath, "mnist_784.npz")
    if os.path.exists(local):
        blob = np.load(local)
        return blob["data"].reshape(-1, 1, 28, 28).astype(np.float32), blob["target"].astype(int).tolist()
    n = int(getattr(args, "synthetic_samples", 70000))
    data, labels = make_synthetic_mnist(n)
    return data, labels.tolist()


def _save_loaders(args, train_loaders, test_loaders):
    os.makedirs(args.datapath, exist_ok=True)
    for key in train_loaders.keys():
        torch.save(train_loaders[key], os.path.join(args.datapath, f"data_worker{key + 1}_train.pt"))
        hist = Counter(train_loaders[key].dataset[:][1].numpy().tolist())
        print(dict(sorted(dict(hist).items())))
        torch.save(test_loaders[key], os.path.join(args.datapath, f"data_worker{key + 1}_test.pt"))


def load_mnist_image(args):
    data, label_list = fetch_mnist(args)
    args.class_num = len(np.unique(label_list))
    parts = image_preprocess_dl(args, data, label_list, test_partition=0.2)
    args.class_num = parts[7]
    _save_loaders(args, parts[5], parts[6])


def load_mnist_flat(args):
    data, label_list = fetch_mnist(args)
    args.class_num = len(np.unique(label_list))
    parts = relational_table_preprocess_dl(args, data.reshape(len(data), -1), label_list, test_partition=0.2)
    args.class_num = parts[7]
    _save_loaders(args, parts[5], parts[6])
