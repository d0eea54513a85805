"""Drop-in boundary, trainer side: the UNMODIFIED ``train_epoch`` of the reference (src/train.py:62-180)
- ``prepare_batch_inputs``, ``optimizer.zero_grad()``, ``model(*13 positionals, predict=False)``, ``loss.backward()``,
``nn.utils.clip_grad_norm_``, the reference ``BertAdam.step()`` over the NAME-BASED weight-decay groups of train.py:339-343,
``ema(model, niter)``, ``cal_performance`` / ``calculate_f1`` on the returned per-video tensors - drives
``svpc_b200.model.StateAwareRecursiveTransformer`` for several iterations and must leave the same statistics and the same
parameters as when it drives the reference model.  Runs where /root/reference is mounted, on the plain-torch operator
statements (host logic); the CUDA operators are covered by the GPU suites."""
import copy
import importlib
import sys
import types

import pytest
import torch

from oracle import refshim
from svpc_b200 import backend, synth
from svpc_b200.config import tiny_config
from svpc_b200.model import StateAwareRecursiveTransformer
from tests.torch_backend import TorchBackend

pytestmark = pytest.mark.skipif(not refshim.available(), reason="reference tree not mounted (GPU box)")


@pytest.fixture(autouse=True)
def _torch_backend():
    prev = backend._impl
    backend.set_backend(TorchBackend())
    yield
    backend.set_backend(prev)


def load_reference_trainer():
    refshim.load()
    if "tensorboardX" not in sys.modules:
        m = types.ModuleType("tensorboardX")

        class SummaryWriter:
            def __init__(self, *a, **k):
                self.rows = []

            def add_scalar(self, *a, **k):
                self.rows.append(a)

        m.SummaryWriter = SummaryWriter
        sys.modules["tensorboardX"] = m
    ds = sys.modules.get("src.rtransformer.recursive_caption_dataset")
    if ds is not None and not hasattr(ds, "caption_collate"):      # the translator test's constants-only stand-in
        del sys.modules["src.rtransformer.recursive_caption_dataset"]
        for k in ("src.translator", "src.translate", "src.train"):
            sys.modules.pop(k, None)
    return importlib.import_module("src.train")


def collated(batch, cfg):
    """A synthetic batch in the shape ``caption_collate`` hands to ``train_epoch`` (recursive_caption_dataset.py:528-576)."""
    B = len(batch["batch_step_num"])
    steps = []
    for t in range(len(batch["input_ids_list"])):
        steps.append(dict(name=[f"v{b}" for b in range(B)], input_ids=batch["input_ids_list"][t],
                          video_feature=batch["video_features_list"][t], input_mask=batch["input_masks_list"][t],
                          token_type_ids=batch["token_type_ids_list"][t], input_labels=batch["input_labels_list"][t]))
    ingr = [dict(ingr_ids=batch["ingr_input_ids"][b].tolist(), ingr_mask=batch["ingr_masks"][b].tolist(),
                 ingr_sep_mask=batch["ingr_sep_masks"][b].tolist(), ingr_id_dict=batch["ingr_id_dict"][b],
                 oov_word_dict=batch["oov_word_dict"][b]) for b in range(B)]
    return (steps, list(batch["batch_step_num"]), [None] * B, ingr, batch["alignments"], batch["actions"])


def run_trainer(tr, model, data, n_epoch=2):
    opt_mod = importlib.import_module("src.rtransformer.optimization")
    named = list(model.named_parameters())
    no_decay = ["bias", "LayerNorm.bias", "LayerNorm.weight"]                     # train.py:339-343
    groups = [{"params": [p for n, p in named if not any(nd in n for nd in no_decay)], "weight_decay": 0.01},
              {"params": [p for n, p in named if any(nd in n for nd in no_decay)], "weight_decay": 0.0}]
    ema = opt_mod.EMA(0.9)
    for n, p in named:
        if p.requires_grad:
            ema.register(n, p.data)
    optimizer = opt_mod.BertAdam(groups, lr=5e-3, warmup=0.1, t_total=len(data) * n_epoch, schedule="warmup_linear")
    opt = types.SimpleNamespace(pin_memory=False, debug=False, max_v_len=model.config.max_v_len, grad_clip=1.0)
    writer = sys.modules["tensorboardX"].SummaryWriter()
    out = []
    for epoch in range(n_epoch):
        out.append(tr.train_epoch(model, copy.deepcopy(data), optimizer, ema, torch.device("cpu"), opt, writer, epoch))
    return out, ema


@pytest.mark.parametrize("mode", ("video", "copy"))
def test_reference_train_epoch_drives_the_dropin(mode):
    tr = load_reference_trainer()
    # dropout probabilities 0: train_epoch switches the model to train mode, and the plain-torch statements are mask-free
    cfg = tiny_config(mode, hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)
    sd = synth.init_state_dict(cfg, seed=4)
    data = [collated(synth.make_batch(cfg, B=3, S=3, seed=30 + i, ragged=True, n_ingr=4, n_oov=1), cfg) for i in range(3)]
    ref = refshim.build_model(cfg, sd, train=True)
    mine = StateAwareRecursiveTransformer(cfg)
    mine.load_state_dict(sd)
    res_ref, ema_ref = run_trainer(tr, ref, data)
    res_mine, ema_mine = run_trainer(tr, mine, data)
    for a, b in zip(res_ref, res_mine):               # (loss per word, accuracy, entity F1 dict, action F1 dict) per epoch
        assert abs(a[0] - b[0]) <= 1e-4 * abs(a[0]), (a[0], b[0])
        assert abs(a[1] - b[1]) <= 1e-6, (a[1], b[1])
        assert a[2] == b[2] and a[3] == b[3]
    assert res_ref[0][0] != res_ref[1][0]             # the parameters moved between the epochs
    psd = dict(ref.named_parameters())
    for n, p in mine.named_parameters():
        q = psd[n]
        assert float((p - q).abs().max()) <= 2e-4 * max(float(q.abs().max()), 1e-3), n
        assert float((ema_mine.shadow[n] - ema_ref.shadow[n]).abs().max()) <= 2e-4 * max(float(q.abs().max()), 1e-3), n
