"""`--vanilla` split learning: labels and loss on Bob, Alices take turns.

Same classes, constructor arguments, method names and log lines as
`/root/reference/data_entities_vanilla.py`; the torch.distributed.rpc / dist_autograd /
DistributedOptimizer round trips are replaced by direct calls into libsplitnn on CUDA streams:

  reference (per batch, :70-76)                      here
  act = model1_sisa(inputs)                    ->    sn_alice_forward
  bob.rpc_sync().train_and_backward(act,y,ctx) ->    sn_bob_train_step (fwd, CE, bwd, SGD-momentum of
    + dist_autograd.backward                          THIS Alice's optimizer slot, cut-layer gradient)
  dist_optimizer.step(ctx)                     ->    sn_alice_backward_update (conv wgrad + SGD-momentum)

Behaviour kept because it changes numbers: one optimizer state per Alice on Bob
(:37-42), `reset_model()` is a no-op (:204-207), Bob never leaves train mode so dropout is live
during evaluation (:234-235), 100-way CE with labels 0..9.
"""
import os
from copy import deepcopy

import torch
from tqdm import tqdm

from . import engine as E
from ._entity_common import (DeviceDataset, RngStream, bob_log_accuracy, entity_device, entity_seed, eval_loop,
                             filter_label, handoff, load_loader, log_breakdown, log_data_stats, make_logger, move)
from .models import default_device, model1_sisa, model2_sisa


class StepContext:
    """Stands in for the dist_autograd context id passed to `bob.train_and_backward`: names the
    calling Alice's optimizer slot on Bob and carries the cut-layer gradient back."""

    def __init__(self, slot):
        self.slot = slot
        self.cut_grad = None
        self.cut_out = None        # where Bob should write the cut-layer gradient (Alice on another GPU)


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
            self.model = model1_sisa(self.device)
        self.lr = args.lr
        self.batch_size = args.batch_size
        self.bob_model_rrefs = bob_model_rrefs
        self.context = StepContext(rank - 1)
        self._new_optimizer()
        self.load_data(args)

    def _new_optimizer(self):
        # Alice's half of DistributedOptimizer(torch.optim.SGD, lr, momentum=0.9) (:37-42)
        self.engine = self.model.make_engine(optimizer=E.SN_OPT_SGD_MOMENTUM, lr=self.lr, momentum=0.9,
                                             max_batch=self.batch_size)

    def _run_batches(self, dataset, desc="Batches"):
        for inputs, labels in tqdm(dataset.batches(self.rng), total=(dataset.n + dataset.batch_size - 1) // dataset.batch_size,
                                   desc=desc, ascii=" >=", disable=not getattr(self, "progress", False)):
            self.train_batch(inputs, labels)

    def train_batch(self, inputs, labels):
        """One split step (:70-76) on a device-resident batch; returns the loss (device scalar).
        With Bob on another GPU of the box nothing is copied: Alice's conv kernel stores its
        activations straight into a buffer in Bob's HBM over NVLink, Bob's last backward kernel
        stores the cut-layer gradient into Alice's, and two events order the two streams."""
        bob_dev = self.bob.device
        if bob_dev == self.device:
            activation_alice = self.engine.forward(inputs)
            loss = self.bob.train_and_backward(activation_alice, labels, self.context)
            self.engine.backward_update(inputs, self.context.cut_grad)
            return loss
        n = inputs.shape[0]
        if getattr(self, "_act_on_bob", None) is None:
            self._act_on_bob = torch.empty(self.batch_size, 5408, device=bob_dev)
            self._cut_here = torch.empty(self.batch_size, 5408, device=self.device)
        activation_alice = self.engine.forward(inputs, out=self._act_on_bob[:n])
        labels_bob = move(labels, bob_dev)
        handoff(self.device, bob_dev)
        self.context.cut_out = self._cut_here[:n]
        loss = self.bob.train_and_backward(activation_alice, labels_bob, self.context)
        handoff(bob_dev, self.device)
        self.engine.backward_update(inputs, self.context.cut_grad)
        return loss

    def _predict(self, x):
        """Test batch -> Bob's logits, back on this Alice's GPU for the counters."""
        return move(self.bob.inference(move(self.engine.forward(x), self.bob.device)), self.device)

    def train(self, last_alice_rref, last_alice_id):
        self.logger.info("Training")
        if last_alice_rref is None:
            self.logger.info(f"Alice-{self.client_id} is first client to train")
        else:
            self.logger.info(f"Alice-{self.client_id} receiving weights from Alice-{last_alice_id}")
            # same effect as load_state_dict(last_alice.give_weights()) without the host-side clone
            self.model.copy_weights_from(last_alice_rref.model)
        for epoch in range(self.epochs):
            self._run_batches(self.train_data)

    def give_weights(self):
        return deepcopy(self.model.state_dict())

    def eval(self):
        c = eval_loop(self._predict, self.test_data, None, self.rng)
        self.logger.info(f"Alice-{self.client_id} Evaluating Data: {round(c[0] / c[1], 3)}")
        return c[0], c[1]

    def load_data(self, args):
        self.train_dataloader = load_loader(os.path.join(args.datapath, f"data_worker{self.client_id}_train.pt"))
        self.test_dataloader = load_loader(os.path.join(args.datapath, f"data_worker{self.client_id}_test.pt"))
        self.train_data = DeviceDataset(self.train_dataloader, self.device)
        self.test_data = DeviceDataset(self.test_dataloader, self.device)
        self.n_train = self.train_data.n
        log_data_stats(self.logger, self.train_data, self.test_data)

    def start_logger(self):
        self.logger = make_logger(f"alice{self.client_id}", f"logs/alice{self.client_id}.log", "Alice is going insane!")

    # ------------- Control group function ------------ #
    def unlearn(self, omit_label):
        self.logger.info("Retraining")
        self.reset_model()
        self.bob.reset_optimizer_slot(self.context.slot)     # fresh DistributedOptimizer (:132-137)
        self._new_optimizer()
        self.unlearn_data = filter_label(self.train_data, omit_label, self.rng)
        with self.rng:
            self.unlearn_data.epoch_order()                   # the reference's logging pass draws once
        from collections import Counter
        self.logger.info("Retraining dataset: {}".format(dict(Counter(self.unlearn_data.labels_list()))))
        self.logger.info("Test dataset (retraining): {}".format(dict(Counter(self.test_data.labels_list()))))
        for epoch in range(self.epochs):
            self._run_batches(self.unlearn_data)

    def eval_breakdown(self, omit_label):
        c = eval_loop(self._predict, self.test_data, omit_label, self.rng)
        log_breakdown(self.logger, self.client_id, omit_label, c, vanilla_style=True)
        return tuple(c)

    def reset_model(self):
        # as in the reference (:204-207): children are nn.Sequential / nn.Flatten, neither has
        # reset_parameters, so nothing is reset and "unlearning" retrains from current weights
        for layer in self.model.children():
            if hasattr(layer, 'reset_parameters'):
                layer.reset_parameters()


class bob(object):
    def __init__(self, args):
        self.server = self
        self.device = entity_device(args, 0)
        self.rng = RngStream(entity_seed(args, 0))
        with self.rng:
            self.model = model2_sisa(self.device)
        self.model.eval_batch = args.batch_size
        if getattr(args, "dropout_parity", False):       # draw masks from torch's CPU stream like the reference
            from .models import TorchCpuMasks
            self.model.mask_source = TorchCpuMasks(self.model.dims[1:-1], self.model.dropout_p)
        self.client_num_in_total = args.client_num_in_total
        self.batch_size = args.batch_size
        # one SGD-momentum state per Alice: each Alice's DistributedOptimizer instantiates its own
        # _FunctionalSGD over Bob's parameters (:37-42)
        self.engine = self.model.make_engine(optimizer=E.SN_OPT_SGD_MOMENTUM, lr=args.lr, momentum=0.9,
                                             n_opt_slots=max(1, args.client_num_in_total),
                                             max_batch=args.batch_size)
        model_rrefs = list(self.model.parameters())
        self.alices = {rank + 1: alice(self.server, model_rrefs, rank + 1, args)
                       for rank in range(args.client_num_in_total)}
        self.last_alice_id = None
        self.start_logger()
        self.logger.info(f"Bob on {self.device}")
        self.last_loss = None

    def train_and_backward(self, x, labels, context_id):
        with self.rng:
            masks, seed, offset = self.model.dropout_args(x.shape[0], x.device)
        loss, cut, _ = self.engine.train_step(x, labels, slot=context_id.slot, masks=masks, seed=seed, offset=offset,
                                              want_cut_grad=True, cut_out=getattr(context_id, "cut_out", None))
        context_id.cut_grad = cut
        self.last_loss = loss
        return loss

    def inference(self, x):
        with self.rng:
            return self.model(x)    # train mode is never left: dropout stays live (:234-235)

    def reset_optimizer_slot(self, slot):
        self.engine.reset_opt_slot(slot)

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

    def start_logger(self):
        self.logger = make_logger("bob", "logs/bob.log", "Bob Started Getting Tipsy")

    # ------------- Control group function ------------ #
    def unlearn_request(self, client_id, omit_label):
        self.logger.info(f"Unlearn Request for Alice-{client_id}")
        self.alices[client_id].unlearn(omit_label)

    def eval_request_breakdown(self, omit_label):
        self.logger.info("Initializing Evaluation of all Alices. Breaking down to the unlearned and the remaining data")
        res = [self.alices[c].eval_breakdown(omit_label) for c in range(1, self.client_num_in_total + 1)]
        return bob_log_accuracy(self.logger, res)
