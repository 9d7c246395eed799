"""Final-accuracy parity (BASELINE.json: within 0.3 pp): the same SISA run - same synthetic data,
same LDA partition, same per-entity torch seeds, same dropout mask stream - through the CPU oracle
(oracle/protocol.py, torch-CPU fp32) and through the GPU entity classes, then eval-mode test
accuracy over every client's test split.  usage: python profiles/accuracy_parity.py [n_samples]"""
import json
import os
import random
import sys
import tempfile
import time
from argparse import Namespace

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from oracle import protocol as P  # noqa: E402
from oracle import torch_ref as R  # noqa: E402
from splitlearning_b200.data.relational_table_preprocessor import image_preprocess_dl  # noqa: E402
from splitlearning_b200.data.synthetic import make_synthetic_mnist  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 12000
    K, epochs = 2, 1
    server_epochs = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    data, labels = make_synthetic_mnist(n, seed=1234)
    wd = tempfile.mkdtemp()
    os.chdir(wd)
    os.makedirs("data/mnist_flat")
    args = Namespace(world_size=K + 1, epochs=epochs, iterations=1, batch_size=16, partition_alpha=0.5,
                     datapath="data/mnist_flat", lr=1e-3, server_epochs=server_epochs, vanilla=False, sisa=True,
                     concat=False, control=False, client_num_in_total=K, class_num=10,
                     torch_seeds={"bob": 100, "alice_base": 200}, dropout_parity=True)
    np.random.seed(0)
    random.seed(0)
    parts = image_preprocess_dl(args, data.astype(np.float64), labels.tolist(), 0.2)
    for k in parts[5]:
        torch.save(parts[5][k], f"data/mnist_flat/data_worker{k + 1}_train.pt")
        torch.save(parts[6][k], f"data/mnist_flat/data_worker{k + 1}_test.pt")
    train = {k + 1: parts[5][k] for k in parts[5]}
    test = {k + 1: parts[6][k] for k in parts[6]}
    n_test = sum(len(t.dataset) for t in test.values())

    t0 = time.time()
    bob_p, alice_p = P.run_sisa(train, lr=1e-3, epochs=epochs, server_epochs=server_epochs, batch_size=16, do_unlearn=False)
    hit = tot = 0
    for c in test:
        x, y = test[c].dataset.tensors
        with torch.no_grad():
            out = R.bob_forward(R.alice_forward(x, alice_p[c][0], alice_p[c][1]), bob_p, R.MODEL2_SISA, None)
        hit += int((out.argmax(1) == y).sum())
        tot += len(y)
    acc_cpu = hit / tot
    t_cpu = time.time() - t0

    from splitlearning_b200 import data_entities_vanilla_sisa as ent
    t0 = time.time()
    BOB = ent.bob(args)
    BOB.train_request_parallel()
    BOB.freeze_alice_weights(range(1, K + 1))
    BOB.train_and_backward(unlearn_request_from_alices=[], unlearn_id=None)
    acc_gpu = BOB.eval_request()
    torch.cuda.synchronize()
    t_gpu = time.time() - t0
    # natural spread: the same GPU run with in-kernel Philox dropout and two other seeds
    spread = []
    for seed in (1, 2):
        a2 = Namespace(**vars(args))
        a2.dropout_parity = False
        B2 = ent.bob(a2)
        B2.model.mask_source.seed = seed
        B2.train_request_parallel()
        B2.freeze_alice_weights(range(1, K + 1))
        B2.train_and_backward(unlearn_request_from_alices=[], unlearn_id=None)
        spread.append(round(B2.eval_request(), 5))
    print(json.dumps({"server_epochs": server_epochs, "acc_gpu_other_dropout_seeds": spread, "samples": n, "test_samples": n_test, "acc_cpu_oracle": round(acc_cpu, 5), "acc_gpu": round(acc_gpu, 5),
                      "delta_pp": round(100 * abs(acc_cpu - acc_gpu), 3), "cpu_seconds": round(t_cpu, 1),
                      "gpu_seconds": round(t_gpu, 1)}))


if __name__ == "__main__":
    main()
