"""`--sisa`: Alices train locally, freeze, then Bob runs server-only epochs over their cached
cut-layer activations; one Alice can then unlearn a label and Bob retrains.

Same classes / methods / log lines as `/root/reference/data_entities_vanilla_sisa.py`.  What
changed underneath:

  alice.train / unlearn / train_control (:55-70,139-168,230-250)
        per batch zero_grad/forward/CE over the 5408 activations/backward/Adam(wd 1e-5)
        -> one sn_alice_local_step per batch
  alice.give_activation_and_labels (:196-211)
        a python list of per-batch tensors pickled over RPC
        -> ONE contiguous [n,5408] fp32 block written by sn_alice_forward straight into the HBM
           cache Bob trains from (+ labels + batch offsets), kept resident across epochs
  bob.train_and_backward (:294-315)
        per cached batch zero_grad/forward/CE/backward/Adam(wd 1e-5)
        -> one sn_bob_server_epoch per (epoch, client): the host loop over batches runs inside the
           library, each batch is streamed from the cache with 128-bit loads

Kept because it changes numbers: Adam with coupled L2 on both sides (:48,266), batch composition
frozen at first fetch and replayed in the same order every epoch (:291), `reset_model()` no-op
(:224-227), cache key (client, unlearned, unlearn_id) (:274), Bob in eval mode during evaluation
(:77,106) and back in train mode for training (:297).
"""
import os
from collections import Counter
from copy import deepcopy

import torch

from . import engine as E
from ._entity_common import (DeviceDataset, RngStream, bob_log_accuracy, entity_device, entity_seed, eval_loop,
                             filter_label, load_loader, log_breakdown, log_data_stats, make_logger, move)
from .models import TorchCpuMasks, default_device, model1_sisa, model2_sisa


class ActivationCache:
    """Frozen-Alice activations for one cache key, resident in Bob's HBM.  Unpacks like the
    reference's `(activations, labels)` pair of per-batch lists (views, no copies)."""

    def __init__(self, acts, labels, batch_size):
        self.acts, self.labels, self.batch_size = acts, labels, batch_size
        n = acts.shape[0]
        self.offsets = list(range(0, n, batch_size)) + [n]

    def __iter__(self):
        bs = self.batch_size
        yield [self.acts[s:s + bs] for s in self.offsets[:-1]]
        yield [self.labels[s:s + bs] for s in self.offsets[:-1]]

    @property
    def n_batches(self):
        return len(self.offsets) - 1


class alice(object):
    def __init__(self, server, bob_model_rrefs, rank, args):
        self.client_id = rank
        self.local_epochs = args.epochs
        self.server_epochs = args.server_epochs
        self.start_logger()
        self.bob = server
        self.device = entity_device(args, rank)
        self.logger.info(f"Alice-{rank} on {self.device}")
        self.rng = RngStream(entity_seed(args, rank))
        with self.rng:
            self.model = model1_sisa(self.device)
        self.lr = args.lr
        self.batch_size = args.batch_size
        self.frozen = False
        self._new_optimizer()
        self.load_data(args)

    def _new_optimizer(self):
        # torch.optim.Adam(lr, weight_decay=1e-5) (:48,145)
        self.engine = self.model.make_engine(optimizer=E.SN_OPT_ADAM, lr=self.lr, weight_decay=1e-5,
                                             max_batch=self.batch_size)

    def _local_epochs(self, dataset):
        for epoch in range(self.local_epochs):
            x, y = dataset.epoch(self.rng)          # this epoch's batch order, one gather
            if x.shape[0]:
                self.engine.local_epoch(x, y, dataset.batch_size)

    def train(self):
        self.logger.info("Local Training")
        assert not self.frozen, "weights are frozen"
        self._local_epochs(self.train_data)

    def _predict(self, x):
        """Test batch -> Bob's logits, back on this Alice's GPU for the counters."""
        return move(self.bob.inference(move(self.engine.forward(x), self.bob.device)), self.device)

    def eval(self):
        self.bob.switch_mode_to_eval()
        c = eval_loop(self._predict, self.test_data, None, self.rng)
        self.logger.info(f"Alice-{self.client_id} Evaluating Data: {round(c[0] / c[1], 3)}")
        return c[0], c[1]

    def eval_breakdown(self, omit_label):
        self.bob.switch_mode_to_eval()
        c = eval_loop(self._predict, self.test_data, omit_label, self.rng)
        log_breakdown(self.logger, self.client_id, omit_label, c, vanilla_style=False)
        return tuple(c)

    def unlearn(self, omit_label):
        self.logger.info(f"Unlearning label: {omit_label}, and reset the model")
        self.reset_model()
        self._new_optimizer()
        self.unlearn_data = filter_label(self.train_data, omit_label, self.rng)
        with self.rng:
            self.unlearn_data.epoch_order()          # the reference's logging pass over the loader
        self.logger.info("Retraining dataset: {}".format(dict(Counter(self.unlearn_data.labels_list()))))
        self.logger.info("Test dataset (retraining): {}".format(dict(Counter(self.test_data.labels_list()))))
        self._local_epochs(self.unlearn_data)

    def load_data(self, args):
        self.train_dataloader = load_loader(os.path.join(args.datapath, f"data_worker{self.client_id}_train.pt"))
        self.test_dataloader = load_loader(os.path.join(args.datapath, f"data_worker{self.client_id}_test.pt"))
        self.train_data = DeviceDataset(self.train_dataloader, self.device)
        self.test_data = DeviceDataset(self.test_dataloader, self.device)
        self.n_train = self.train_data.n
        log_data_stats(self.logger, self.train_data, self.test_data)

    def start_logger(self):
        self.logger = make_logger(f"alice{self.client_id}", f"logs/alice{self.client_id}.log", "Alice is going insane!")

    def give_activation_and_labels(self, unlearned=False, out_device=None):
        """One pass over her loader with frozen weights; the activations of the whole pass are
        written by a single kernel launch into one contiguous block on Bob's device."""
        data = self.train_data if not unlearned else self.unlearn_data
        x, y = data.epoch(self.rng)
        out_device = out_device or self.device
        if out_device == self.device:
            acts = self.engine.forward(x)
        else:
            # Alice on another GPU of the box: her kernel stores the whole block straight into Bob's
            # HBM over NVLink (no local staging copy); Bob's stream then waits for hers
            from ._entity_common import handoff
            acts = torch.empty(x.shape[0], 5408, device=out_device)
            if x.shape[0]:
                self.engine.forward(x, out=acts)
            y = y.to(out_device, non_blocking=True)
            handoff(self.device, out_device)
        return ActivationCache(acts, y, data.batch_size)

    def give_weights(self):
        return deepcopy(self.model.state_dict())

    def freeze_weights(self):
        self.frozen = True
        for param in self.model.parameters():
            param.requires_grad = False

    def unfreeze_weights(self):
        self.frozen = False
        for param in self.model.parameters():
            param.requires_grad = True

    def reset_model(self):
        for layer in self.model.children():        # no child has reset_parameters: a no-op (:224-227)
            if hasattr(layer, 'reset_parameters'):
                layer.reset_parameters()

    # ------------- Control group function ------------ #
    def train_control(self, omit_label):
        self.unlearn_data = filter_label(self.train_data, omit_label, self.rng)
        with self.rng:
            self.unlearn_data.epoch_order()
        self.logger.info("Filtered dataset: {}".format(dict(Counter(self.unlearn_data.labels_list()))))
        self._local_epochs(self.unlearn_data)


class bob(object):
    def __init__(self, args):
        self.server = self
        self.device = entity_device(args, 0)
        self.rng = RngStream(entity_seed(args, 0))
        with self.rng:
            self.model = self._make_model(args)
        self.model.eval_batch = args.batch_size
        if getattr(args, "dropout_parity", False):
            self.model.mask_source = TorchCpuMasks(self.model.dims[1:-1], self.model.dropout_p)
        self.server_epochs = args.server_epochs
        self.client_num_in_total = args.client_num_in_total
        self.batch_size = args.batch_size
        self.start_logger()
        self.logger.info(f"Bob on {self.device}")
        # torch.optim.Adam(lr, weight_decay=1e-5) over Bob's parameters (:266)
        self.engine = self.model.make_engine(optimizer=E.SN_OPT_ADAM, lr=args.lr, weight_decay=1e-5,
                                             n_opt_slots=1, max_batch=args.batch_size)
        alice_cls = getattr(self, "_alice_cls", None) or alice
        self.alices = {rank + 1: alice_cls(self.server, list(self.model.parameters()), rank + 1, args)
                       for rank in range(args.client_num_in_total)}
        self.activation_and_labels_cache = {}
        self.loss_history = []

    def _make_model(self, args):
        return model2_sisa(self.device)

    def get_activation_and_labels(self, client_id, unlearned=False, unlearn_id=None):
        cache_key = (client_id, unlearned, unlearn_id)
        if cache_key in self.activation_and_labels_cache:
            return self.activation_and_labels_cache[cache_key]
        entry = self.alices[client_id].give_activation_and_labels(unlearned, out_device=self.device)
        self.activation_and_labels_cache[cache_key] = entry
        return entry

    def _epoch_masks(self, cache):
        """Parity mode: draw every batch's masks from the CPU stream in step order."""
        if not isinstance(self.model.mask_source, TorchCpuMasks) or self.model.dropout_p <= 0:
            return None
        per_layer = [[] for _ in self.model.mask_source.widths]
        with self.rng:
            for i in range(cache.n_batches):
                b = cache.offsets[i + 1] - cache.offsets[i]
                for l, m in enumerate(self.model.mask_source.draw(b, self.device)):
                    per_layer[l].append(m)
        return [torch.cat(ms, 0) for ms in per_layer]

    def train_and_backward(self, unlearn_request_from_alices, unlearn_id):
        self.logger.info("Global Training")
        self.switch_mode_to_train()
        for epoch in range(self.server_epochs):
            for client_id in range(1, self.client_num_in_total + 1):
                cache = self.get_activation_and_labels(client_id, unlearned=True, unlearn_id=unlearn_id) \
                    if client_id in unlearn_request_from_alices else \
                    self.get_activation_and_labels(client_id, unlearned=False)
                if cache.n_batches == 0:
                    continue
                masks = self._epoch_masks(cache)
                losses = torch.empty(cache.n_batches, device=self.device) if getattr(self, "record_loss", False) else None
                seed, offset = 0, 0
                if masks is None and self.model.dropout_p > 0:
                    seed = self.model.mask_source.seed
                    offset = self.model.mask_source.offset + 1
                    self.model.mask_source.offset += cache.n_batches
                self.engine.server_epoch(cache.acts, cache.labels, cache.offsets, slot=0, masks=masks, seed=seed,
                                         offset=offset, loss_per_step=losses)
                if losses is not None:
                    self.loss_history.append(losses)
        self.logger.info("Global training completed.")

    def inference(self, x):
        with self.rng:
            return self.model(x)

    def train_request(self, client_id):
        self.logger.info(f"Train Request for Alice-{client_id}")
        self.alices[client_id].train()

    def train_request_parallel(self):
        self.logger.info(f"Train all Alices in parallel")
        # independent models and data: on several GPUs each Alice's stream runs concurrently
        for client_id in range(1, self.client_num_in_total + 1):
            self.alices[client_id].train()
        for client_id in range(1, self.client_num_in_total + 1):
            torch.cuda.synchronize(self.alices[client_id].device)

    def eval_request(self):
        self.logger.info("Initializing Evaluation of all Alices")
        res = [self.alices[c].eval() for c in range(1, self.client_num_in_total + 1)]
        acc = sum(r[0] for r in res) / sum(r[1] for r in res)
        self.logger.info("Accuracy over all data: {:.3f}".format(acc))
        return acc

    def unlearn_request(self, client_id, omit_label):
        self.logger.info(f"Unlearn Request for Alice-{client_id} upon the label-{omit_label}")
        self.alices[client_id].unlearn(omit_label)

    def eval_request_breakdown(self, omit_label):
        self.logger.info("Initializing Evaluation of all Alices. Breaking down to the unlearned and the remaining data")
        res = [self.alices[c].eval_breakdown(omit_label) for c in range(1, self.client_num_in_total + 1)]
        return bob_log_accuracy(self.logger, res)

    def freeze_alice_weights(self, client_ids):
        for client_id in client_ids:
            self.logger.info("Server training starts. Freezing weights for Alices-{}.".format(client_id))
            self.alices[client_id].freeze_weights()

    def unfreeze_alice_weights(self, client_ids):
        for client_id in client_ids:
            self.logger.info("Unfreezing weights for Alices-{}.".format(client_id))
            self.alices[client_id].unfreeze_weights()

    def switch_mode_to_train(self):
        self.model.train()

    def switch_mode_to_eval(self):
        self.model.eval()

    def start_logger(self):
        self.logger = make_logger("bob", "logs/bob.log", "Bob Started Getting Tipsy")

    # ------------- Control group function ------------ #
    def train_request_control(self, client_id, omit_label):
        self.logger.info(f"Train Request for Alice-{client_id}")
        self.alices[client_id].train_control(omit_label)
