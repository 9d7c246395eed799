"""CPU-only checks of the host side: the bit-exact LDA partition and split (known answers from
SURVEY.md Appendix B), the loader emulation's RNG consumption, CLI validation, and that the C-ABI
library loads and exports every symbol the header declares (no compute calls without a GPU)."""
import ctypes
import hashlib
import os
import random
import re
from argparse import Namespace

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


# ----------------------------------------------------------------------------- LDA partition KATs
KAT = {
    1: ([70000], "b9bf2de0967c56b26b3f04bf55eab9bf7b1161a8b0e4038ff64d8e13c7e04b6c"),
    3: ([21860, 24741, 23399], "02238fcbe265bb5c407572643f7807cb03d7ed73b457e989a6da3498d997e08b"),
    4: ([16271, 21685, 15329, 16715], "3e07ba899ea463119826015e871d67e3de0ff9e1a660c3e5f942fa3ad50fbeed"),
    7: ([10448, 11355, 10069, 6662, 8509, 9968, 12989], "a2082ffb69b9a65602b2a641ddc1c702f90dbfd7bca0d5339e3408750ccfe626"),
}


@pytest.mark.parametrize("native", [False, True])
@pytest.mark.parametrize("K", [1, 3, 4, 7])
def test_lda_partition_known_answers(K, native):
    from splitlearning_b200.data.noniid_partition import non_iid_partition_with_dirichlet_distribution as part
    labels = np.random.RandomState(1234).randint(0, 10, size=70000)
    assert hashlib.sha256(labels.astype(np.int64).tobytes()).hexdigest() == \
        "51cfabb574dbfd600bb2efc08db687688c119041674d97a10ad1378dc9e75b0a"
    np.random.seed(0)
    m = part(torch.LongTensor(labels.tolist()), K, 10, 0.5, task='task', native=native)
    h = hashlib.sha256()
    for k in range(K):
        h.update(np.asarray(m[k], np.int64).tobytes())
    assert [len(m[k]) for k in range(K)] == KAT[K][0]
    assert h.hexdigest() == KAT[K][1]
    # ndarray labels and an explicit RandomState give the same answer
    m2 = part(labels, K, 10, 0.5, task='task', rng=np.random.RandomState(0))
    assert all(m2[k] == m[k] for k in range(K))


def test_lda_partition_edge_cases():
    from splitlearning_b200.data.noniid_partition import non_iid_partition_with_dirichlet_distribution as part
    with pytest.raises(NotImplementedError):
        part(np.zeros(10, int), 2, 10, 0.5, task='segmentation')
    # tiny, very skewed: the min-size loop must restart until every client holds >= 10 samples
    labels = np.random.RandomState(3).randint(0, 3, size=120)
    m = part(labels, 4, 3, 0.1, rng=np.random.RandomState(1))
    assert sorted(sum((m[k] for k in range(4)), [])) == list(range(120))
    assert min(len(m[k]) for k in range(4)) >= 10


@pytest.mark.parametrize("n,K,classes,alpha,warm", [
    (5000, 10, 10, 0.1, 0),       # skewed: some clients fill up and are masked out
    (600, 5, 3, 0.3, 0),          # the min-size loop restarts
    (3000, 16, 10, 2.5, 1),       # alpha > 1: Marsaglia-Tsang gammas, starting with a cached gaussian
    (2000, 9, 7, 1.0, 0),         # alpha == 1: exponentials
    (40000, 150, 10, 5.0, 2),     # > 128 clients: the renormalising sum recurses pairwise
])
def test_native_lda_partition_is_numpys(n, K, classes, alpha, warm):
    """csrc/lda_partition.cpp against the numpy implementation: the same shards, and the generator is
    left at the same position (the next uniform and the next gaussian agree)."""
    from splitlearning_b200.data.noniid_partition import non_iid_partition_with_dirichlet_distribution as part
    labels = np.random.RandomState(100 + n).randint(0, classes, n)
    a, b = np.random.RandomState(n), np.random.RandomState(n)
    for _ in range(warm):
        a.normal(), b.normal()
    ra = part(labels, K, classes, alpha, rng=a, native=False)
    rb = part(labels, K, classes, alpha, rng=b, native=True)
    assert ra == rb
    assert a.random_sample() == b.random_sample() and a.normal() == b.normal()
    sa, sb = a.get_state(), b.get_state()
    assert np.array_equal(sa[1], sb[1]) and sa[2:] == sb[2:]


@pytest.mark.parametrize("n", [0, 1, 2, 17, 1000, 48000])
def test_native_py_shuffle_is_random_shuffle(n):
    from splitlearning_b200.data.relational_table_preprocessor import py_random_shuffle
    random.seed(n + 3)
    want = list(range(n))
    random.shuffle(want)
    after = random.getstate()
    random.seed(n + 3)
    got = list(range(n))
    py_random_shuffle(got, native=True)
    assert got == want and random.getstate() == after


@pytest.mark.parametrize("native", [False, True])
def test_train_test_split_known_answer(native, monkeypatch):
    from splitlearning_b200.data import noniid_partition, relational_table_preprocessor
    from splitlearning_b200.data.relational_table_preprocessor import image_preprocess_dl
    if not native:          # the numpy / `random` path: as if the library were not there
        monkeypatch.setattr(noniid_partition, "_native_lib", lambda: None)
        monkeypatch.setattr(relational_table_preprocessor, "_native_lib", lambda: None)
    N = 4000
    lab = np.random.RandomState(1234).randint(0, 10, size=N)
    data = np.arange(N, dtype=np.float64).reshape(N, 1, 1, 1)
    args = Namespace(client_num_in_total=4, class_num=10, partition_alpha=0.5, batch_size=16)
    np.random.seed(0)
    random.seed(0)
    r = image_preprocess_dl(args, data, lab.tolist(), 0.2)
    assert r[0] == 3198 and r[1] == 802 and r[7] == 10
    h = hashlib.sha256()
    sizes = []
    for k in range(4):
        tr = r[5][k].dataset[:][0].reshape(-1).numpy().astype(np.int64)
        te = r[6][k].dataset[:][0].reshape(-1).numpy().astype(np.int64)
        sizes.append((len(tr), len(te)))
        h.update(tr.tobytes())
        h.update(te.tobytes())
    assert sizes == [(809, 203), (1007, 252), (573, 144), (809, 203)]
    assert h.hexdigest() == "c3c7044d91d1f9325f9084aa2d15de3d727d9a5bb97572754c4a7167d691af0e"


# ----------------------------------------------------------------------------- loader emulation
def _toy_loader(n, bs, shuffle):
    x = torch.arange(n, dtype=torch.float32).reshape(n, 1, 1, 1).expand(n, 1, 28, 28).contiguous()
    y = torch.arange(n) % 10
    return torch.utils.data.DataLoader(torch.utils.data.TensorDataset(x, y), batch_size=bs, shuffle=shuffle)


@pytest.mark.parametrize("shuffle", [True, False])
@pytest.mark.parametrize("n", [50, 64, 1])
def test_device_dataset_matches_dataloader_rng(shuffle, n):
    """Same batches, same order, and the same generator state afterwards as iterating the real
    DataLoader (two epochs, ragged last batch)."""
    from splitlearning_b200._entity_common import DeviceDataset
    loader = _toy_loader(n, 16, shuffle)
    torch.manual_seed(7)
    ref = [[(xb[:, 0, 0, 0].clone(), yb.clone()) for xb, yb in loader] for _ in range(2)]
    ref_state = torch.get_rng_state()
    torch.manual_seed(7)
    ds = DeviceDataset(loader, torch.device("cpu"))
    got = [[(xb[:, 0, 0, 0].clone(), yb.clone()) for xb, yb in ds.batches()] for _ in range(2)]
    assert torch.equal(torch.get_rng_state(), ref_state)
    for e in range(2):
        assert len(got[e]) == len(ref[e])
        for (gx, gy), (rx, ry) in zip(got[e], ref[e]):
            assert torch.equal(gx, rx) and torch.equal(gy, ry)


def test_filter_label_matches_reference_recipe():
    from splitlearning_b200._entity_common import DeviceDataset, filter_label
    loader = _toy_loader(70, 16, True)
    torch.manual_seed(3)
    kept = [(xi, yi) for xb, yb in loader for xi, yi in zip(xb, yb) if yi != 9]      # reference recipe
    ref_loader = torch.utils.data.DataLoader(kept, batch_size=16)
    ref = [(xb[:, 0, 0, 0].clone(), yb.clone()) for xb, yb in ref_loader]
    ref_state = torch.get_rng_state()
    torch.manual_seed(3)
    ds = DeviceDataset(loader, torch.device("cpu"))
    unl = filter_label(ds, 9)
    got = [(xb[:, 0, 0, 0].clone(), yb.clone()) for xb, yb in unl.batches()]
    assert torch.equal(torch.get_rng_state(), ref_state)
    assert len(got) == len(ref)
    for (gx, gy), (rx, ry) in zip(got, ref):
        assert torch.equal(gx, rx) and torch.equal(gy, ry)
    assert all((gy != 9).all() for _, gy in got)


def test_rng_stream_isolated():
    from splitlearning_b200._entity_common import RngStream
    torch.manual_seed(1)
    outer_before = torch.get_rng_state()
    s = RngStream(100)
    with s:
        a = torch.rand(3)
    with s:
        b = torch.rand(3)
    assert torch.equal(torch.get_rng_state(), outer_before)
    torch.manual_seed(100)
    assert torch.equal(torch.rand(3), a) and torch.equal(torch.rand(3), b)
    noop = RngStream(None)
    torch.manual_seed(5)
    with noop:
        c = torch.rand(1)
    torch.manual_seed(5)
    assert torch.equal(torch.rand(1), c)


# ----------------------------------------------------------------------------- CLI surface
def test_cli_flags_defaults_and_validation():
    from splitlearning_b200.split_nn import build_parser, get_alice_and_bob, validate
    a = validate(build_parser().parse_args([]))
    assert (a.world_size, a.epochs, a.iterations, a.batch_size, a.partition_alpha, a.datapath, a.lr,
            a.server_epochs) == (3, 1, 5, 16, 0.5, "data/mnist_flat", 0.001, 3)
    assert a.client_num_in_total == 2 and not (a.vanilla or a.sisa or a.concat or a.control)
    for bad in (["--concat"], ["--vanilla", "--sisa"], ["--vanilla", "--sisa", "--concat"], ["--control", "--vanilla"],
                ["--control", "--sisa"]):
        with pytest.raises(ValueError):
            validate(build_parser().parse_args(bad))
    assert validate(build_parser().parse_args(["--sisa", "--concat", "--world_size", "8"])).client_num_in_total == 7
    for flags, mod in ((["--vanilla"], "data_entities_vanilla"), (["--sisa"], "data_entities_vanilla_sisa"),
                       (["--sisa", "--concat"], "data_entities_sisa_concat"), ([], "data_entities")):
        al, bo = get_alice_and_bob(validate(build_parser().parse_args(flags)))
        assert al.__module__.endswith(mod) and bo.__module__.endswith(mod)


def test_entity_method_surface():
    """Every method name the reference's entity classes expose."""
    from splitlearning_b200 import (data_entities, data_entities_sisa_concat, data_entities_vanilla,
                                    data_entities_vanilla_sisa)
    van_a = {"train", "give_weights", "eval", "load_data", "start_logger", "unlearn", "eval_breakdown", "reset_model"}
    van_b = {"train_and_backward", "inference", "train_request", "eval_request", "start_logger", "unlearn_request",
             "eval_request_breakdown"}
    sisa_a = {"train", "eval", "eval_breakdown", "unlearn", "load_data", "start_logger", "give_activation_and_labels",
              "give_weights", "freeze_weights", "unfreeze_weights", "reset_model", "train_control"}
    sisa_b = {"get_activation_and_labels", "train_and_backward", "inference", "train_request", "train_request_parallel",
              "eval_request", "unlearn_request", "eval_request_breakdown", "freeze_alice_weights",
              "unfreeze_alice_weights", "switch_mode_to_train", "switch_mode_to_eval", "start_logger",
              "train_request_control"}
    u_a = {"train", "give_weights", "eval", "load_data", "start_logger"}
    u_b = {"train_request", "eval_request", "inference", "start_logger"}
    for mod, a, b in ((data_entities_vanilla, van_a, van_b), (data_entities_vanilla_sisa, sisa_a, sisa_b),
                      (data_entities_sisa_concat, sisa_a, sisa_b), (data_entities, u_a, u_b)):
        assert a <= set(dir(mod.alice)), a - set(dir(mod.alice))
        assert b <= set(dir(mod.bob)), b - set(dir(mod.bob))


def test_model_state_dict_layout():
    """Keys and shapes of the reference's modules, built on the CPU for the check."""
    from splitlearning_b200 import models as M
    cpu = torch.device("cpu")
    sd = M.model1_sisa(cpu).state_dict()
    assert {k: tuple(v.shape) for k, v in sd.items()} == {"conv_layers.0.weight": (32, 1, 3, 3), "conv_layers.0.bias": (32,)}
    sd = M.model1(cpu).state_dict()
    assert {k: tuple(v.shape) for k, v in sd.items()} == {"conv1.weight": (32, 1, 3, 3), "conv1.bias": (32,)}
    sd = M.model2(cpu).state_dict()
    assert {k: tuple(v.shape) for k, v in sd.items()} == {"fc1.weight": (1000, 5408), "fc1.bias": (1000,),
                                                         "fc2.weight": (100, 1000), "fc2.bias": (100,)}
    m = M.model2_sisa_concat(2, cpu)
    assert m.dims == [10816, 5000, 1000, 200] and tuple(m.state_dict()["fc3.weight"].shape) == (200, 1000)
    assert M.model3(cpu).dims == [100, 10]
    # same constructors in the same order => same init draws as torch's own layers
    torch.manual_seed(9)
    ours = M.model2_sisa(cpu).state_dict()
    torch.manual_seed(9)
    ref = [torch.nn.Linear(5408, 5000), torch.nn.Linear(5000, 1000), torch.nn.Linear(1000, 100)]
    assert torch.equal(ours["fc1.weight"], ref[0].weight) and torch.equal(ours["fc3.bias"], ref[2].bias)


# ----------------------------------------------------------------------------- C-ABI library
def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "splitnn.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(sn_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from splitlearning_b200 import _lib
    names = _declared_symbols()
    assert len(names) >= 30
    assert set(names) == set(_lib.SIGNATURES), set(names) ^ set(_lib.SIGNATURES)
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), n
    loaded = _lib.load()
    assert loaded.sn_version() >= 1
    assert loaded.sn_last_error() is not None


def test_missing_gpu_fails_loudly():
    """No CPU fallback: without a CUDA device the product path raises instead of computing."""
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from splitlearning_b200 import _lib, models as M
    with pytest.raises(_lib.SplitNNError):
        M.default_device()
    with pytest.raises(Exception):
        M.model2_sisa(torch.device("cpu"))(torch.zeros(2, 5408))
