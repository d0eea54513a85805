"""Training-loop bookkeeping on the device.

The reference's ``train_epoch`` / ``eval_epoch`` (src/train.py:153-173, 236-262) compute word accuracy and the entity /
action F1 counts per batch with ``cal_performance`` and ``calculate_f1`` (train.py:32-49) and read every count with
``.item()``: ~6 host synchronisations per video per step.  ``StepMetrics`` accumulates the same counts in one fp64 device
buffer with two kinds of launch and reads them ONCE, in ``result()``:

    metrics = StepMetrics(device)
    for batch in loader:
        loss, preds, ent_probs, ac_probs = model(...)
        metrics.update(loss, preds, labels, ent_probs, alignments, ac_probs, actions)     # no synchronisation
    loss_per_word, accuracy, ent_dict, ac_dict = metrics.result()                         # train.py:175-180

``labels[b]`` is the (S_b, max_t_len) int64 label block of video b (train.py:133-138), ``-1`` = ignore.
"""
from __future__ import annotations

import ctypes as C

import torch

from .cuda_backend import SvpcError
from .optim import _be, _p

IGNORE = -1          # RCDataset.IGNORE (src/rtransformer/dataset.py)


def compute_total_f1(n_correct, n_recall, n_precision):
    """train.py:51-67."""
    recall = 0 if n_recall == 0 else n_correct / n_recall
    precision = 0 if n_precision == 0 else n_correct / n_precision
    f1 = 0 if (recall == 0 and precision == 0) else 2 * (recall * precision) / (recall + precision)
    return {"recall": recall, "precision": precision, "f1": f1}


class StepMetrics:
    # acc layout: [n_correct, n_word, ent_correct, ent_recall, ent_precision, ac_correct, ac_recall, ac_precision, loss]
    def __init__(self, device):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise SvpcError("svpc_b200.metrics needs a CUDA device (no CPU path)")
        self.acc = torch.zeros(9, dtype=torch.float64, device=self.device)

    def reset(self):
        self.acc.zero_()

    def _f1(self, probs, golds, slot):
        if not probs:
            return
        p = torch.cat([x.detach().reshape(-1) for x in probs]).float()
        y = torch.cat([g.reshape(-1) for g in golds]).to(device=self.device, dtype=torch.float32)
        assert p.numel() == y.numel(), (p.numel(), y.numel())
        _be()._call("svpc_f1_counts", C.c_longlong(p.numel()), _p(p), _p(y), C.c_void_p(self.acc.data_ptr() + 8 * slot))

    @torch.no_grad()
    def update(self, loss, pred_scores_list, labels_list, entity_prob_list=(), alignments=(), action_prob_list=(),
               actions=()):
        dev = self.device
        if loss is not None:
            self.acc[8:9] += loss.detach().reshape(1).double()                       # total_loss += loss.item(), train.py:173
        preds = [p.detach() for p in pred_scores_list]
        golds = [g.to(dev).contiguous() for g in labels_list]
        for p, g in zip(preds, golds):
            if not (p.is_cuda and p.dtype == torch.float32 and p.is_contiguous() and p.dim() == 3):
                raise SvpcError("prediction scores must be contiguous fp32 CUDA tensors (S_b, L, V)")
            if g.dtype != torch.int64 or g.numel() != p.shape[0] * p.shape[1]:
                raise SvpcError("labels must be int64 with one entry per (clip, position)")
        if preds:
            rows = [p.shape[0] * p.shape[1] for p in preds]
            start = [0]
            for r in rows[:-1]:
                start.append(start[-1] + r)
            i64 = lambda v: torch.tensor(v, dtype=torch.int64).to(dev, non_blocking=True)
            i32 = lambda v: torch.tensor(v, dtype=torch.int32).to(dev, non_blocking=True)
            sp, gp = i64([p.data_ptr() for p in preds]), i64([g.data_ptr() for g in golds])
            rw, wd, rs = i32(rows), i32([p.shape[2] for p in preds]), i32(start)
            _be()._call("svpc_caption_accuracy", len(preds), _p(sp), _p(gp), _p(rw), _p(wd), _p(rs), int(sum(rows)),
                        C.c_longlong(IGNORE), _p(self.acc))
        self._f1(list(entity_prob_list), list(alignments), 2)
        self._f1(list(action_prob_list), list(actions), 5)

    def counts(self):
        """One synchronisation: the nine accumulated numbers as Python floats."""
        return [float(v) for v in self.acc.cpu()]

    def result(self):
        """(loss_per_word, accuracy, ent_dict, ac_dict) as train_epoch returns them (train.py:175-180)."""
        c = self.counts()
        n_word = c[1]
        return (1.0 * c[8] / n_word, 1.0 * c[0] / n_word, compute_total_f1(c[2], c[3], c[4]),
                compute_total_f1(c[5], c[6], c[7]))
