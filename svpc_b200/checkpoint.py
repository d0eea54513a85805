"""Checkpoint write / read in the reference's layout.

``src/train.py:402-406`` saves ``{"model": model.state_dict(), "model_cfg": model.config, "opt": opt, "epoch": epoch_i}``
while the EMA weights are assigned (:393-394) and ``src/translator.py:33-38`` reads ``model_cfg`` and ``model`` back.  The
``state_dict`` keys, shapes and dtypes of the drop-in model are the reference's (pinned by
tests/test_model_host.py), so files written here load into the reference model and the other way round.

``snapshot`` is what keeps the training stream busy: ``model.state_dict()`` aliases live parameter storage (with the in-place
``EMA.assign`` / ``resume`` of svpc_b200.optim, the storage is overwritten by the next ``resume``), so the values are staged
into ONE pinned host buffer by asynchronous D2H copies on the current stream and ``torch.save`` runs on host memory.
"""
from __future__ import annotations

import collections

import torch


def snapshot(state_dict):
    """Detached host copy of a ``state_dict``: one pinned arena, one async D2H per tensor, one synchronisation."""
    items = list(state_dict.items())
    cuda = [(k, v) for k, v in items if isinstance(v, torch.Tensor) and v.is_cuda]
    total, offs = 0, {}
    for k, v in cuda:
        offs[k] = total
        total += (v.numel() * v.element_size() + 255) // 256 * 256
    arena = torch.empty(max(total, 1), dtype=torch.uint8).pin_memory() if cuda else None
    out = collections.OrderedDict()
    for k, v in items:
        if isinstance(v, torch.Tensor) and v.is_cuda:
            h = arena[offs[k]:offs[k] + v.numel() * v.element_size()].view(v.dtype).view(v.shape)
            h.copy_(v.detach(), non_blocking=True)
            out[k] = h
        elif isinstance(v, torch.Tensor):
            out[k] = v.detach().clone()
        else:
            out[k] = v
    if cuda:
        torch.cuda.current_stream().synchronize()
        # every tensor is saved with its own storage (a view of the shared arena would serialise the whole arena per key)
        for k, _ in cuda:
            out[k] = out[k].clone()
    return out


def make_checkpoint(model, opt=None, epoch=0):
    """train.py:402-406 (call between ``ema.assign(model)`` and the next ``ema.resume(model)`` to save the EMA model)."""
    return {"model": snapshot(model.state_dict()), "model_cfg": model.config, "opt": opt, "epoch": epoch}


def save_checkpoint(path, model, opt=None, epoch=0):
    ckpt = make_checkpoint(model, opt, epoch)
    torch.save(ckpt, path)
    return ckpt


def load_checkpoint(path, model_cls=None, device=None):
    """translator.py:31-41: returns ``(checkpoint, model)``; ``model`` is built from ``checkpoint["model_cfg"]`` with
    ``model_cls`` (default: the drop-in ``StateAwareRecursiveTransformer``), loaded strictly, moved to ``device`` and put in
    eval mode."""
    ckpt = torch.load(path, map_location="cpu", weights_only=False)
    if model_cls is None:
        from .model import StateAwareRecursiveTransformer as model_cls
    model = model_cls(ckpt["model_cfg"])
    _prepare_embedding_layout(model, ckpt["model"])
    model.load_state_dict(ckpt["model"])
    if device is not None:
        model.to(device)
    model.eval()
    return ckpt, model


def _prepare_embedding_layout(model, sd):
    """A vivt / viv checkpoint holds ``reasoner.action_embeddings`` as a bare Parameter (the reference replaces the
    nn.Embedding in ``set_pretrained_embedding``, model.py:773-775); a freshly built model needs the same replacement before
    a strict load (translator.py relies on ``train.py`` having done it - SURVEY.md 8a-26)."""
    for prefix in ("reasoner", "recipe_reasoner"):
        mod = getattr(model, prefix, None)
        key = prefix + ".action_embeddings"
        if mod is not None and key in sd and hasattr(mod, "set_pretrained_embedding") and \
                not isinstance(getattr(mod, "action_embeddings", None), torch.nn.Parameter):
            mod.set_pretrained_embedding(sd[key].clone(), freeze=False)
