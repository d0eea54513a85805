"""TEST INFRASTRUCTURE ONLY - CPU restatement of the reference's per-batch bookkeeping (src/train.py:32-67, 153-180).

Imported only by ``tests/``.  Pinned against the reference's own functions: ``tests/golden/make_metrics_golden.py``
extracts ``cal_performance`` / ``calculate_f1`` / ``compute_total_f1`` from /root/reference/src/train.py (the module
itself cannot be imported here: tensorboardX / h5py / spacy are absent), executes them unmodified on a seeded case and
commits the counts to ``tests/golden/metrics_counts.npz``."""
import torch

IGNORE = -1


def cal_performance(pred, gold):
    """train.py:32-38: number of labelled positions whose arg-max word equals the label."""
    pred = pred.max(2)[1].contiguous().view(-1)
    gold = gold.contiguous().view(-1)
    return int((pred.eq(gold) & gold.ne(IGNORE)).sum())


def n_word(gold):
    """train.py:156-157."""
    return int(gold.ne(IGNORE).sum())


def calculate_f1(prob_list, golds):
    """train.py:40-49."""
    c = r = q = 0.0
    for prob, gold in zip(prob_list, golds):
        pos = (prob > 0.5).flatten()
        gold = gold.flatten()
        c += float(gold[pos].sum())
        r += float(gold.sum())
        q += float(pos.sum())
    return c, r, q


def compute_total_f1(n_correct, n_recall, n_precision):
    """train.py:51-67."""
    recall = 0 if n_recall == 0 else n_correct / n_recall
    precision = 0 if n_precision == 0 else n_correct / n_precision
    f1 = 0 if (recall == 0 and precision == 0) else 2 * (recall * precision) / (recall + precision)
    return {"recall": recall, "precision": precision, "f1": f1}


def epoch(steps):
    """steps: iterable of (loss, preds, labels, ent_probs, alignments, ac_probs, actions) -> what train_epoch returns."""
    tot = nw = nc = 0.0
    e = [0.0, 0.0, 0.0]
    a = [0.0, 0.0, 0.0]
    for loss, preds, labels, ent, al, ac, acts in steps:
        for p, g in zip(preds, labels):
            nc += cal_performance(p, g)
            nw += n_word(g)
        for acc, (probs, gold) in ((e, (ent, al)), (a, (ac, acts))):
            c, r, q = calculate_f1(probs, gold)
            acc[0] += c; acc[1] += r; acc[2] += q
        tot += float(loss)
    return tot / nw, nc / nw, compute_total_f1(*e), compute_total_f1(*a), [nc, nw] + e + a + [tot]
