"""Synthetic MNIST-shaped data: the real `mnist_784` needs a network fetch
(`/root/reference/data/mnist_flat/mnist_flat_generator.py:27`), which neither this container nor
the GPU box has.  Images are integer-valued float32 in [0,255] like the reference's
(`mnist_flat_generator.py:29` reshapes raw pixel values, un-normalised); labels are 0..9.
Each image is a fixed per-class prototype blended with uniform noise, so the set is learnable
and accuracy parity is meaningful.  numpy `RandomState` only: version-stable.
"""
import numpy as np


def make_synthetic_mnist(n=70000, seed=1234, noise=0.5):
    rng = np.random.RandomState(seed)
    labels = rng.randint(0, 10, size=n)
    protos = rng.randint(0, 256, size=(10, 28, 28)).astype(np.float32)
    # sparse-ish prototypes (MNIST is mostly background)
    protos *= (rng.rand(10, 28, 28) < 0.35)
    out = np.empty((n, 1, 28, 28), dtype=np.float32)
    chunk = 8192
    for s in range(0, n, chunk):
        lab = labels[s:s + chunk]
        u = rng.rand(len(lab), 28, 28).astype(np.float32) * 255.0
        img = (1.0 - noise) * protos[lab] + noise * u * (rng.rand(len(lab), 28, 28) < 0.5)
        out[s:s + chunk, 0] = np.clip(np.rint(img), 0, 255)
    return out, labels.astype(np.int64)
