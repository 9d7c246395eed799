"""Writes tests/golden/* from the UNMODIFIED reference in /root/reference (run in the build
container only; the GPU box has no /root/reference and only reads the committed fixtures).

    TORCH_FORCE_NO_WEIGHTS_ONLY_LOAD=1 python -m oracle.gen_golden [modules] [rpc] [concat7]

Two kinds of fixture:
  modules_*.npz  module-level vectors: the reference's own `models.py` classes loaded with
                 numpy-seeded weights (version-stable), one fwd/bwd/optimizer step through
                 torch autograd + torch.optim exactly as the reference's entity code calls them.
                 Inputs/weights are regenerated from the stored seeds; outputs are stored in full
                 where small and as sampled entries + sums where large.
  rpc_*.npz      trajectories recorded from the reference's RPC entity code (ws=3: Bob + 2
                 Alices in 3 processes, TensorPipe), via recorders hooked onto Bob's model.
Harness shims (reference untouched): synthetic .pt loaders written with the reference's own
`image_preprocess_dl`; per-process torch seeds; a forward hook on `bob.model`.
"""
import os
import sys
import tempfile
from argparse import Namespace

import numpy as np
import torch

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
BOB_SEED = 100


def alice_seed(r):
    return 200 + r


# ----------------------------------------------------------------------------- shared recipes
def seeded_weights(shapes, seed):
    """Version-stable pseudo-weights: uniform(-1/sqrt(fan_in), +) from numpy RandomState."""
    rng = np.random.RandomState(seed)
    out = []
    for shp in shapes:
        fan_in = int(np.prod(shp[1:])) if len(shp) > 1 else None
        out.append((shp, fan_in))
    ws, last_fan = [], 1
    for shp, fan_in in out:
        if fan_in is not None:
            last_fan = fan_in
        bound = 1.0 / np.sqrt(last_fan)
        ws.append(rng.uniform(-bound, bound, size=shp).astype(np.float32))
    return ws


def seeded_batch(B, seed, n_classes=10):
    rng = np.random.RandomState(seed)
    x = np.rint(rng.uniform(0, 255, size=(B, 1, 28, 28)) * (rng.rand(B, 1, 28, 28) < 0.4)).astype(np.float32)
    y = rng.randint(0, n_classes, size=B).astype(np.int64)
    return x, y


def seeded_masks(B, dims, seed):
    rng = np.random.RandomState(seed)
    return [(rng.rand(B, d) < 0.5).astype(np.uint8) for d in dims]


def sample_idx(n, k, seed):
    return np.random.RandomState(seed).randint(0, n, size=min(k, n)).astype(np.int64)


def summarize(name, t, store, k=4096, seed=7):
    a = t.detach().cpu().numpy().astype(np.float32).reshape(-1)
    if a.size <= 65536:
        store[name] = a.reshape(tuple(t.shape))
    else:
        idx = sample_idx(a.size, k, seed)
        store[name + "__idx"] = idx
        store[name + "__val"] = a[idx]
    store[name + "__sum"] = np.float64(a.astype(np.float64).sum())
    store[name + "__l2"] = np.float64(np.sqrt((a.astype(np.float64) ** 2).sum()))


# ----------------------------------------------------------------------------- module-level
def gen_modules():
    sys.path.insert(0, REF)
    import models as ref_models                      # the reference's models.py, unmodified
    import torch.nn as nn

    os.makedirs(OUT, exist_ok=True)
    B = 16

    def load(model, arrays):
        sd = model.state_dict()
        assert len(sd) == len(arrays)
        model.load_state_dict({k: torch.from_numpy(a) for k, a in zip(sd.keys(), arrays)})

    def forced_dropout(model, masks):
        """Teacher-force the dropout masks: replace nn.Dropout modules by fixed-mask multiplies
        (x * mask / (1-p) is what F.dropout computes)."""
        class Fixed(nn.Module):
            def __init__(self, m):
                super().__init__()
                self.m = torch.from_numpy(m.astype(np.float32)) / 0.5

            def forward(self, x):
                return x * self.m
        model.dropout1, model.dropout2 = Fixed(masks[0]), Fixed(masks[1])

    # ---- vanilla: model1_sisa + model2_sisa, CE on Bob, SGD momentum (two consecutive steps)
    st = {}
    a_w = seeded_weights([(32, 1, 3, 3), (32,)], 11)
    b_w = seeded_weights([(5000, 5408), (5000,), (1000, 5000), (1000,), (100, 1000), (100,)], 12)
    alice, bob = ref_models.model1_sisa(), ref_models.model2_sisa()
    load(alice, a_w)
    load(bob, b_w)
    params = list(bob.parameters()) + list(alice.parameters())
    opt = torch.optim.SGD(params, lr=1e-3, momentum=0.9)       # what _FunctionalSGD applies
    crit = nn.CrossEntropyLoss()
    for step in range(2):
        x, y = seeded_batch(B, 100 + step)
        masks = seeded_masks(B, (5000, 1000), 200 + step)
        forced_dropout(bob, masks)
        opt.zero_grad()
        act = alice(torch.from_numpy(x))
        act.retain_grad()
        logits = bob(act)
        loss = crit(logits, torch.from_numpy(y))
        loss.backward()
        summarize(f"s{step}_act", act, st)
        summarize(f"s{step}_logits", logits, st)
        st[f"s{step}_loss"] = np.float32(loss.item())
        summarize(f"s{step}_cut_grad", act.grad, st)
        for n, p in list(bob.named_parameters()) + [("alice." + k, v) for k, v in alice.named_parameters()]:
            summarize(f"s{step}_grad_{n}", p.grad, st)
        opt.step()
        for n, p in list(bob.named_parameters()) + [("alice." + k, v) for k, v in alice.named_parameters()]:
            summarize(f"s{step}_new_{n}", p, st)
    np.savez_compressed(os.path.join(OUT, "modules_vanilla.npz"), **st)

    # ---- SISA server: model2_sisa on detached activations, Adam wd=1e-5 (two steps)
    st = {}
    bob = ref_models.model2_sisa()
    load(bob, b_w)
    opt = torch.optim.Adam(bob.parameters(), lr=1e-3, weight_decay=1e-5)
    alice = ref_models.model1_sisa()
    load(alice, a_w)
    for step in range(2):
        x, y = seeded_batch(B, 300 + step)
        masks = seeded_masks(B, (5000, 1000), 400 + step)
        forced_dropout(bob, masks)
        act = alice(torch.from_numpy(x)).detach()
        opt.zero_grad()
        logits = bob(act.detach())
        loss = crit(logits, torch.from_numpy(y))
        loss.backward()
        opt.step()
        summarize(f"s{step}_logits", logits, st)
        st[f"s{step}_loss"] = np.float32(loss.item())
        for n, p in bob.named_parameters():
            summarize(f"s{step}_grad_{n}", p.grad, st)
            summarize(f"s{step}_new_{n}", p, st)
    # eval-mode inference (no dropout) with the twice-updated weights
    bob.dropout1, bob.dropout2 = nn.Dropout(0.5), nn.Dropout(0.5)
    bob.eval()
    x, y = seeded_batch(B, 500)
    with torch.no_grad():
        out = bob(alice(torch.from_numpy(x)))
    summarize("eval_logits", out, st)
    np.savez_compressed(os.path.join(OUT, "modules_sisa_server.npz"), **st)

    # ---- SISA Alice local: CE over the 5408 activations, Adam wd=1e-5 (three steps)
    st = {}
    alice = ref_models.model1_sisa()
    load(alice, a_w)
    opt = torch.optim.Adam(alice.parameters(), lr=1e-3, weight_decay=1e-5)
    for step in range(3):
        x, y = seeded_batch(B, 600 + step)
        opt.zero_grad()
        act = alice(torch.from_numpy(x))
        loss = crit(act, torch.from_numpy(y))
        loss.backward()
        st[f"s{step}_loss"] = np.float32(loss.item())
        for n, p in alice.named_parameters():
            summarize(f"s{step}_grad_{n}", p.grad, st)
        opt.step()
        for n, p in alice.named_parameters():
            summarize(f"s{step}_new_{n}", p, st)
    np.savez_compressed(os.path.join(OUT, "modules_sisa_alice.npz"), **st)

    # ---- U-shape: model1 -> model2 -> model3, Adam (two steps)
    st = {}
    u_b = seeded_weights([(1000, 5408), (1000,), (100, 1000), (100,)], 13)
    u_t = seeded_weights([(10, 100), (10,)], 14)
    m1, m2, m3 = ref_models.model1(), ref_models.model2(), ref_models.model3()
    load(m1, a_w)
    load(m2, u_b)
    load(m3, u_t)
    opt = torch.optim.Adam(list(m3.parameters()) + list(m2.parameters()) + list(m1.parameters()), lr=1e-3)
    for step in range(2):
        x, y = seeded_batch(B, 700 + step)
        opt.zero_grad()
        a1 = m1(torch.from_numpy(x))
        a1.retain_grad()
        hb = m2(a1)
        hb.retain_grad()
        out = m3(hb)
        loss = crit(out, torch.from_numpy(y))
        loss.backward()
        summarize(f"s{step}_bob_out", hb, st)
        summarize(f"s{step}_logits", out, st)
        st[f"s{step}_loss"] = np.float32(loss.item())
        summarize(f"s{step}_bob_out_grad", hb.grad, st)
        summarize(f"s{step}_cut_grad", a1.grad, st)
        opt.step()
        for tag, m in (("m1", m1), ("m2", m2), ("m3", m3)):
            for n, p in m.named_parameters():
                summarize(f"s{step}_grad_{tag}.{n}", p.grad, st)
                summarize(f"s{step}_new_{tag}.{n}", p, st)
    np.savez_compressed(os.path.join(OUT, "modules_ushape.npz"), **st)

    # ---- concat model shapes (reference entity code crashes; only the module is pinned)
    st = {}
    K = 2
    c_w = seeded_weights([(5000, 5408 * K), (5000,), (1000, 5000), (1000,), (100 * K, 1000), (100 * K,)], 15)
    mc = ref_models.model2_sisa_concat(K)
    load(mc, c_w)
    mc.eval()
    rng = np.random.RandomState(16)
    xin = (rng.rand(4, 5408 * K) * 50 * (rng.rand(4, 5408 * K) < 0.6)).astype(np.float32)
    with torch.no_grad():
        out = mc(torch.from_numpy(xin))
    summarize("eval_logits", out, st)
    np.savez_compressed(os.path.join(OUT, "modules_concat.npz"), **st)
    print("module fixtures written to", OUT)


def concat7_inputs(K=7, B=16):
    """Seeded weights / batch / masks / labels of the K=7 concat fixture (shared with the tests)."""
    w = seeded_weights([(5000, 5408 * K), (5000,), (1000, 5000), (1000,), (100 * K, 1000), (100 * K,)], 21)
    rng = np.random.RandomState(22)
    # sparse non-negative activations like the real cut tensor (63 % non-zero, mean 45)
    xs = [(rng.rand(B, 5408 * K) * 90 * (rng.rand(B, 5408 * K) < 0.63)).astype(np.float32) for _ in range(2)]
    ys = [rng.randint(0, 10, size=(B, K)).astype(np.int64) for _ in range(2)]
    masks = [seeded_masks(B, (5000, 1000), 23 + s) for s in range(2)]
    return w, xs, ys, masks


def gen_concat7():
    """modules_concat7.npz: BASELINE.json config 4's server (`model2_sisa_concat(7)`, fc1 5000 x 37856)
    through the reference's own module + torch autograd + torch.optim.Adam(wd 1e-5), two steps.  The
    7-head loss is this build's definition (see oracle/torch_ref.concat_server_step)."""
    sys.path.insert(0, REF)
    import models as ref_models
    import torch.nn as nn
    K, B = 7, 16
    w, xs, ys, masks = concat7_inputs(K, B)
    mc = ref_models.model2_sisa_concat(K)
    sd = mc.state_dict()
    mc.load_state_dict({k: torch.from_numpy(a) for k, a in zip(sd.keys(), w)})
    opt = torch.optim.Adam(mc.parameters(), lr=1e-3, weight_decay=1e-5)
    crit = nn.CrossEntropyLoss()

    class Fixed(nn.Module):
        def __init__(self, m):
            super().__init__()
            self.m = torch.from_numpy(m.astype(np.float32)) / 0.5

        def forward(self, x):
            return x * self.m
    st = {}
    for step in range(2):
        mc.dropout1, mc.dropout2 = Fixed(masks[step][0]), Fixed(masks[step][1])
        opt.zero_grad()
        logits = mc(torch.from_numpy(xs[step]))
        loss = crit(logits.reshape(B * K, 100), torch.from_numpy(ys[step]).reshape(B * K))
        loss.backward()
        summarize(f"s{step}_logits", logits, st)
        st[f"s{step}_loss"] = np.float32(loss.item())
        for n, p in mc.named_parameters():
            summarize(f"s{step}_grad_{n}", p.grad, st)
        opt.step()
        for n, p in mc.named_parameters():
            summarize(f"s{step}_new_{n}", p, st)
    np.savez_compressed(os.path.join(OUT, "modules_concat7.npz"), **st)
    print("modules_concat7.npz written")


# ----------------------------------------------------------------------------- RPC trajectories
def _write_loaders(workdir, n, K, batch_size):
    """Synthetic .pt loaders through the REFERENCE's own image_preprocess_dl."""
    import random
    sys.path.insert(0, REF)
    from data.relational_table_preprocessor import image_preprocess_dl
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, repo)
    from splitlearning_b200.data.synthetic import make_synthetic_mnist
    data, labels = make_synthetic_mnist(n, seed=1234)
    args = Namespace(client_num_in_total=K, class_num=10, partition_alpha=0.5, batch_size=batch_size)
    np.random.seed(0)
    random.seed(0)
    parts = image_preprocess_dl(args, data.astype(np.float64), labels.tolist(), 0.2)
    os.makedirs(os.path.join(workdir, "data/mnist_flat"), exist_ok=True)
    os.makedirs(os.path.join(workdir, "logs"), exist_ok=True)
    for k in parts[5]:
        torch.save(parts[5][k], os.path.join(workdir, f"data/mnist_flat/data_worker{k + 1}_train.pt"))
        torch.save(parts[6][k], os.path.join(workdir, f"data/mnist_flat/data_worker{k + 1}_test.pt"))


def _rpc_worker(rank, world_size, mode, workdir, port, out_path):
    os.chdir(workdir)
    sys.path.insert(0, REF)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    os.environ["TORCH_FORCE_NO_WEIGHTS_ONLY_LOAD"] = "1"
    import torch.distributed.rpc as rpc
    torch.manual_seed(BOB_SEED if rank == 0 else alice_seed(rank))
    args = Namespace(world_size=world_size, epochs=1, iterations=1, batch_size=16, partition_alpha=0.5,
                     datapath="data/mnist_flat", lr=1e-3, server_epochs=2, vanilla=mode == "vanilla",
                     sisa=mode == "sisa", concat=False, control=False,
                     client_num_in_total=world_size - 1, class_num=10)
    if rank != 0:
        rpc.init_rpc(f"alice{rank}", rank=rank, world_size=world_size)
        rpc.shutdown()
        return
    rpc.init_rpc("bob", rank=0, world_size=world_size)
    if mode == "vanilla":
        import data_entities_vanilla as ent
    elif mode == "sisa":
        import data_entities_vanilla_sisa as ent
    else:
        import data_entities as ent
    BOB = ent.bob(args)
    rec = {"logits": [], "labels": [], "loss": []}
    BOB.model.register_forward_hook(lambda m, i, o: rec["logits"].append(o.detach().clone()))
    if mode == "vanilla":
        orig = BOB.train_and_backward

        def wrapped(x, labels, ctx):
            rec["labels"].append(labels.clone())
            loss = orig(x, labels, ctx)
            rec["loss"].append(loss.detach().clone())
            return loss
        BOB.train_and_backward = wrapped
        for c in range(1, world_size):
            BOB.train_request(c)
    elif mode == "ushape":
        for c in range(1, world_size):
            BOB.train_request(c)
    else:
        crit = BOB.criterion

        class Crit:
            def __call__(self, out, lab):
                loss = crit(out, lab)
                rec["labels"].append(lab.clone())
                rec["loss"].append(loss.detach().clone())
                return loss
        BOB.criterion = Crit()
        BOB.train_request_parallel()
        BOB.freeze_alice_weights(range(1, args.client_num_in_total + 1))
        BOB.train_and_backward(unlearn_request_from_alices=[], unlearn_id=None)
        BOB.unfreeze_alice_weights([1])
        BOB.unlearn_request(client_id=1, omit_label=9)
        BOB.freeze_alice_weights([1])
        BOB.train_and_backward([1], unlearn_id=9)
    store = {"logits": torch.cat([l.reshape(-1) for l in rec["logits"]]).numpy(),
             "logit_rows": np.array([l.shape[0] for l in rec["logits"]], np.int64)}
    if rec["loss"]:
        store["loss"] = torch.stack(rec["loss"]).numpy()
        store["labels"] = torch.cat(rec["labels"]).numpy()
    for n, p in BOB.model.named_parameters():
        summarize("final_" + n, p, store)
    for c in range(1, world_size):
        w = BOB.alices[c].rpc_sync().give_weights()
        if isinstance(w, list):
            for j, sd in enumerate(w):
                for k, v in sd.items():
                    store[f"final_alice{c}_m{j}.{k}"] = v.numpy()
        else:
            for k, v in w.items():
                store[f"final_alice{c}.{k}"] = v.numpy()
    store["torch_version"] = np.array(torch.__version__)
    np.savez_compressed(out_path, **store)
    rpc.shutdown()


def gen_rpc():
    import torch.multiprocessing as mp
    os.makedirs(OUT, exist_ok=True)
    for port, (mode, n) in enumerate([("vanilla", 300), ("sisa", 300), ("ushape", 300)], start=29611):
        with tempfile.TemporaryDirectory() as wd:
            _write_loaders(wd, n, 2, 16)
            out_path = os.path.join(OUT, f"rpc_{mode}_ws3.npz")
            mp.spawn(_rpc_worker, args=(3, mode, wd, port, out_path), nprocs=3, join=True)
            print("wrote", out_path)


if __name__ == "__main__":
    what = sys.argv[1:] or ["modules", "rpc"]
    if "modules" in what:
        gen_modules()
    if "rpc" in what:
        gen_rpc()
    if "concat7" in what:
        gen_concat7()
