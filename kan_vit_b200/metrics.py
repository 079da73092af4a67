"""Per-epoch metric collection of the reference's training loop (train.py:27,42-46,74-78; utils.py:13-46) without
the per-step host round trips.

The reference copies the labels, ``argmax(y_hat)`` and ``softmax(y_hat)`` of every batch to the host and hands the
lists to sklearn at the end of the epoch.  Here the cross-entropy launch itself writes the probabilities and
predictions into device buffers sized for the epoch and counts (label, prediction) pairs in a confusion matrix;
the host reads them ONCE per epoch.  Accuracy, balanced accuracy and the weighted F1 score are exact functions
of the confusion matrix (same values as sklearn's, ``metrics_from_confusion``); ROC AUC needs the probabilities
and is computed by sklearn, like the reference, from the single device-to-host copy.
"""
from __future__ import annotations

import numpy as np
import torch


def metrics_from_confusion(conf: np.ndarray):
    """(accuracy, balanced_accuracy, weighted F1) of utils.py:38-40 from conf[true, pred] counts --
    sklearn's accuracy_score, balanced_accuracy_score (classes without samples are left out) and
    f1_score(average='weighted')."""
    conf = conf.astype(np.float64)
    total = conf.sum()
    tp = np.diag(conf)
    support = conf.sum(axis=1)
    predicted = conf.sum(axis=0)
    accuracy = tp.sum() / total if total else float("nan")
    seen = support > 0
    balanced = float(np.mean(tp[seen] / support[seen])) if seen.any() else float("nan")
    denom = support + predicted              # 2 tp + fp + fn
    f1 = np.where(denom > 0, 2.0 * tp / np.maximum(denom, 1.0), 0.0)
    weighted_f1 = float((f1 * support).sum() / total) if total else float("nan")
    return float(accuracy), balanced, weighted_f1


class EpochMetrics:
    """Device-side y_true / y_pred / y_pred_proba of one epoch (train.py:27) + the running loss."""

    def __init__(self, n_classes: int, capacity: int, device, keep_probs: bool = True):
        self.n_classes, self.capacity = n_classes, capacity
        self.confusion = torch.zeros((n_classes, n_classes), dtype=torch.int32, device=device)
        self.labels = torch.empty((capacity,), dtype=torch.int64, device=device)
        self.pred = torch.empty((capacity,), dtype=torch.int64, device=device)
        self.probs = torch.empty((capacity, n_classes), dtype=torch.float32, device=device) if keep_probs else None
        self.loss_sum = torch.zeros((), dtype=torch.float32, device=device)
        self.n = 0
        self.steps = 0

    def reset(self):
        self.confusion.zero_()
        self.loss_sum.zero_()
        self.n = self.steps = 0

    def loss_and_grad(self, logits, labels, need_grad=True):
        """Cross-entropy of one batch (ops.cross_entropy_with_grad) that also records the batch's metrics."""
        from . import ops
        B = logits.shape[0]
        if self.n + B > self.capacity:
            raise ValueError(f"EpochMetrics capacity {self.capacity} exceeded")
        sl = slice(self.n, self.n + B)
        self.labels[sl].copy_(labels)
        loss, dz = ops.cross_entropy_with_grad(logits, labels, need_grad=need_grad,
                                               probs_out=None if self.probs is None else self.probs[sl],
                                               pred_out=self.pred[sl], confusion=self.confusion)
        self.loss_sum += loss
        self.n += B
        self.steps += 1
        return loss, dz

    def compute(self, roc_auc: bool = True):
        """One device-to-host read; returns the dict the reference logs (train.py:48-53)."""
        conf = self.confusion.cpu().numpy()
        acc, bal, f1 = metrics_from_confusion(conf)
        out = {"loss": float(self.loss_sum) / max(self.steps, 1), "accuracy": acc, "balanced_accuracy": bal,
               "f1": f1, "roc_auc": None}
        if roc_auc and self.probs is not None:
            from sklearn.metrics import roc_auc_score          # utils.py:42-45
            y = self.labels[: self.n].cpu()
            onehot = torch.nn.functional.one_hot(y, num_classes=self.n_classes).numpy()
            out["roc_auc"] = float(roc_auc_score(onehot, self.probs[: self.n].cpu().numpy(), average="weighted",
                                                 multi_class="ovr"))
        return out
