"""`--sisa --concat`: SISA with a server whose first layer takes every Alice's activations side by
side (`model2_sisa_concat`, `/root/reference/models.py:66-82`).

REFERENCE UNPINNED.  The reference's own concat loop cannot run: `train_and_backward`
row-concatenates (`dim=0`) and then iterates per SAMPLE into a model whose fc1 expects
5408*K features (`data_entities_sisa_concat.py:270-276` -> IndexError at `models.py:76`); its eval
has `# TODO: concat` (:112) and its CLI branch never reaches the unlearning phase
(`split_nn.py:64-72`).  Only the module shapes are pinned (tests/golden/modules_concat.npz).
The semantics below are DEFINED BY THIS BUILD ("D1: feature-concat, K heads"):

  * step j input  = cat_k(act_k[batch j], dim=1)            -> [B, 5408*K]
  * output        = [B, 100*K] read as K heads of 100 logits
  * loss          = mean over k of CE(head_k, labels_k[batch j])   (one CE over B*K rows)
  * steps/epoch   = min_k(number of FULL batches of Alice k)
  * eval, Alice k = her activations in slot k, zeros elsewhere, first argmax over head k
  * unlearning    = as `--sisa` (`split_nn.py:95-110`): Alice 1 retrains without the label, her
                    slot of the concatenated cache is rebuilt, Bob retrains.

Everything else (Alice-side training, freezing, cache keys, Adam wd 1e-5, log lines) is the
`--sisa` module's.
"""
import torch

from . import data_entities_vanilla_sisa as sisa
from ._entity_common import eval_loop, log_breakdown
from .models import model2_sisa_concat


class alice(sisa.alice):
    def _predict(self, x):
        out = self.bob.inference_slot(self.engine.forward(x), self.client_id)
        return out if out.device == self.device else out.to(self.device, non_blocking=True)

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


class bob(sisa.bob):
    _alice_cls = alice

    def __init__(self, args):
        super().__init__(args)
        self.engine.set_ce_heads(self.client_num_in_total)
        self._concat_key = None
        self._concat = None

    def _make_model(self, args):
        return model2_sisa_concat(num_clients=args.client_num_in_total, device=self.device)

    def _concat_cache(self, unlearn_request_from_alices, unlearn_id):
        K, bs = self.client_num_in_total, self.batch_size
        key = tuple((c, c in unlearn_request_from_alices, unlearn_id if c in unlearn_request_from_alices else None)
                    for c in range(1, K + 1))
        if key == self._concat_key:
            return self._concat
        parts = [self.get_activation_and_labels(c, unlearned=k[1], unlearn_id=k[2]) for c, k in zip(range(1, K + 1), key)]
        steps = min(p.acts.shape[0] // bs for p in parts)
        rows = steps * bs
        acts = torch.cat([p.acts[:rows] for p in parts], dim=1).contiguous()              # [rows, 5408K]
        labels = torch.stack([p.labels[:rows] for p in parts], dim=1).contiguous()        # [rows, K]
        self._concat_key, self._concat = key, (acts, labels, list(range(0, rows + 1, bs)))
        return self._concat

    def train_and_backward(self, unlearn_request_from_alices, unlearn_id):
        self.logger.info("Global Training")
        self.switch_mode_to_train()
        acts, labels, offsets = self._concat_cache(list(unlearn_request_from_alices), unlearn_id)
        n_batches = len(offsets) - 1
        for epoch in range(self.server_epochs):
            if n_batches == 0:
                break
            masks = None
            seed, offset = 0, 0
            if self.model.dropout_p > 0:
                from .models import TorchCpuMasks
                if isinstance(self.model.mask_source, TorchCpuMasks):
                    per_layer = [[] for _ in self.model.mask_source.widths]
                    with self.rng:
                        for i in range(n_batches):
                            for l, m in enumerate(self.model.mask_source.draw(self.batch_size, self.device)):
                                per_layer[l].append(m)
                    masks = [torch.cat(ms, 0) for ms in per_layer]
                else:
                    seed = self.model.mask_source.seed
                    offset = self.model.mask_source.offset + 1
                    self.model.mask_source.offset += n_batches
            losses = torch.empty(n_batches, device=self.device) if getattr(self, "record_loss", False) else None
            self.engine.server_epoch(acts, labels, offsets, slot=0, masks=masks, seed=seed, offset=offset,
                                     loss_per_step=losses)
            if losses is not None:
                self.loss_history.append(losses)
        self.logger.info("Global training completed.")

    def inference_slot(self, act, client_id):
        """Alice k's activations in slot k of an otherwise zero input; returns head k's logits."""
        K = self.client_num_in_total
        B = act.shape[0]
        x = torch.zeros(B, 5408 * K, device=self.device)
        x[:, (client_id - 1) * 5408:client_id * 5408] = act.to(self.device)
        out = self.inference(x)
        return out[:, (client_id - 1) * 100:client_id * 100].contiguous()
