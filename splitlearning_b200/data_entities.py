"""Default ("U-shape") split learning: Alice keeps the first AND the last layers, so labels never
leave her.  Same classes / methods / log lines as `/root/reference/data_entities.py`.

Per batch (reference :70-81 -> here):
  a1 = model1(inputs)                         -> sn_alice_forward                    [B,32,13,13]
  hb = bob.rpc_sync().inference(a1)           -> sn_bob_forward (train mode: keeps activations)
  out = model3(hb); loss = CE(out, labels)    -> sn_bob_train_step on the 100->10 tail, which also
  dist_autograd.backward + Adam on model3        returns dL/dhb and applies Adam to model3
  (backward through Bob, Adam on Bob)         -> sn_bob_backward_update (slot of THIS Alice)
  (backward through model1, Adam on model1)   -> sn_alice_backward_update
Four boundary crossings per step, as in the reference.  One Adam state per Alice on Bob: each
Alice builds her own DistributedOptimizer(Adam) over [model3, Bob, model1] (:43-47).

The reference's CLI calls `eval_request_breakdown` / `unlearn_request` on this module's bob, which
do not exist there (AttributeError after training, split_nn.py:54 — SURVEY.md 0.6).  They are
provided here, modelled on the vanilla module, so the default command line runs to completion.
"""
import os
from collections import Counter
from copy import deepcopy

import torch

from . import engine as E
from ._entity_common import (DeviceDataset, RngStream, bob_log_accuracy, entity_device, entity_seed, eval_loop,
                             filter_label, load_loader, log_breakdown, log_data_stats, make_logger, move)
from .models import default_device, model1, model2, model3


class alice(object):
    def __init__(self, server, bob_model_rrefs, rank, args):
        self.client_id = rank
        self.epochs = args.epochs
        self.start_logger()
        self.bob = server
        self.device = entity_device(args, rank)
        self.logger.info(f"Alice-{rank} on {self.device}")
        self.rng = RngStream(entity_seed(args, rank))
        with self.rng:
            self.model1 = model1(self.device)       # construction order as in the reference (:31-32)
            self.model2 = model3(self.device)
        self.lr = args.lr
        self.batch_size = args.batch_size
        self.slot = rank - 1
        self._new_optimizer()
        self.load_data(args)

    def _new_optimizer(self):
        adam = dict(optimizer=E.SN_OPT_ADAM, lr=self.lr, max_batch=self.batch_size)
        self.head = self.model1.make_engine(**adam)
        self.tail = self.model2.make_engine(**adam)

    def _run_batches(self, dataset):
        for inputs, labels in dataset.batches(self.rng):
            # the four hops of the U-shape step (:70-79); each `move` is a stream-ordered NVLink copy
            # when Bob sits on another GPU of the box, and nothing at all on one GPU
            a1 = move(self.head.forward(inputs), self.bob.device)
            hb = move(self.bob.inference_train(a1), self.device)
            _, dhb, _ = self.tail.train_step(hb, labels, want_cut_grad=True)
            cut = move(self.bob.backward(a1, move(dhb, self.bob.device), self.slot), self.device)
            self.head.backward_update(inputs, cut)

    def train(self, last_alice_rref, last_alice_id):
        self.logger.info("Training")
        if last_alice_rref is None:
            self.logger.info(f"Alice-{self.client_id} is first client to train")
        else:
            self.logger.info(f"Alice-{self.client_id} receiving weights from Alice-{last_alice_id}")
            model1_weights, model2_weights = last_alice_rref.give_weights()
            self.model1.load_state_dict(model1_weights)
            self.model2.load_state_dict(model2_weights)
        for epoch in range(self.epochs):
            self._run_batches(self.train_data)

    def give_weights(self):
        return [deepcopy(self.model1.state_dict()), deepcopy(self.model2.state_dict())]

    def _predict(self, x):
        return self.model2(move(self.bob.inference(move(self.head.forward(x), self.bob.device)), self.device))

    def eval(self):
        c = eval_loop(self._predict, self.test_data, None, self.rng)
        self.logger.info(f"Alice-{self.client_id} Evaluating Data: {round(c[0] / c[1], 3)}")
        return c[0], c[1]

    def eval_breakdown(self, omit_label):
        c = eval_loop(self._predict, self.test_data, omit_label, self.rng)
        log_breakdown(self.logger, self.client_id, omit_label, c, vanilla_style=True)
        return tuple(c)

    def unlearn(self, omit_label):
        self.logger.info("Retraining")
        self.bob.reset_optimizer_slot(self.slot)
        self._new_optimizer()
        self.unlearn_data = filter_label(self.train_data, omit_label, self.rng)
        with self.rng:
            self.unlearn_data.epoch_order()
        self.logger.info("Retraining dataset: {}".format(dict(Counter(self.unlearn_data.labels_list()))))
        for epoch in range(self.epochs):
            self._run_batches(self.unlearn_data)

    def load_data(self, args):
        self.train_dataloader = load_loader(os.path.join(args.datapath, f"data_worker{self.client_id}_train.pt"))
        self.test_dataloader = load_loader(os.path.join(args.datapath, f"data_worker{self.client_id}_test.pt"))
        self.train_data = DeviceDataset(self.train_dataloader, self.device)
        self.test_data = DeviceDataset(self.test_dataloader, self.device)
        self.n_train = self.train_data.n
        log_data_stats(self.logger, self.train_data, self.test_data)

    def start_logger(self):
        self.logger = make_logger(f"alice{self.client_id}", f"logs/alice{self.client_id}.log", "Alice is going insane!")


class bob(object):
    def __init__(self, args):
        self.server = self
        self.device = entity_device(args, 0)
        self.rng = RngStream(entity_seed(args, 0))
        with self.rng:
            self.model = model2(self.device)
        self.engine = self.model.make_engine(optimizer=E.SN_OPT_ADAM, lr=args.lr,
                                             n_opt_slots=max(1, args.client_num_in_total),
                                             max_batch=args.batch_size)
        model_rrefs = list(self.model.parameters())
        self.alices = {rank + 1: alice(self.server, model_rrefs, rank + 1, args)
                       for rank in range(args.client_num_in_total)}
        self.last_alice_id = None
        self.client_num_in_total = args.client_num_in_total
        self.start_logger()
        self.logger.info(f"Bob on {self.device}")

    def train_request(self, client_id):
        self.logger.info(f"Train Request for Alice-{client_id}")
        if self.last_alice_id is None:
            self.alices[client_id].train(None, None)
        else:
            self.alices[client_id].train(self.alices[self.last_alice_id], self.last_alice_id)
        self.last_alice_id = client_id

    def eval_request(self):
        self.logger.info("Initializing Evaluation of all Alices")
        res = [self.alices[c].eval() for c in range(1, self.client_num_in_total + 1)]
        acc = sum(r[0] for r in res) / sum(r[1] for r in res)
        self.logger.info("Accuracy over all data: {:.3f}".format(acc))
        return acc

    def inference(self, x):
        return self.model(x)

    def inference_train(self, x):
        """The autograd-recorded `inference` of a training step (:73): activations are kept for
        the `backward` that follows."""
        return self.engine.forward(torch.flatten(x, 1), train_mode=True)

    def backward(self, x, dout, slot):
        return self.engine.backward_update(torch.flatten(x, 1), dout, slot=slot, want_cut_grad=True)

    def reset_optimizer_slot(self, slot):
        self.engine.reset_opt_slot(slot)

    def start_logger(self):
        self.logger = make_logger("bob", "logs/bob.log", "Bob Started Getting Tipsy")

    # not in the reference's module (its CLI crashes without them) -------------------------------
    def unlearn_request(self, client_id, omit_label):
        self.logger.info(f"Unlearn Request for Alice-{client_id}")
        self.alices[client_id].unlearn(omit_label)

    def eval_request_breakdown(self, omit_label):
        self.logger.info("Initializing Evaluation of all Alices. Breaking down to the unlearned and the remaining data")
        res = [self.alices[c].eval_breakdown(omit_label) for c in range(1, self.client_num_in_total + 1)]
        return bob_log_accuracy(self.logger, res)
