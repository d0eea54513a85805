"""Generates tests/golden/*.npz by running the UNMODIFIED reference (imported from /root/reference through
oracle/refshim.py) on seeded synthetic batches.  Only runs in the build container; the fixtures it writes are
committed and travel to the GPU box.

    python tests/golden/make_golden.py

Per model mode (v / vi / viv / vivt) on the tiny configuration (svpc_b200.config.tiny_config), ragged batch with
OOV copy words: total loss, prediction_scores / entity_probs / action_probs per video, every parameter gradient,
and the greedy captions of the reference Translator (LM head sharpened x20 so argmax margins are not at the
fp32 noise floor).  Inputs and weights are NOT stored: they are regenerated bit-identically
from the seeds by svpc_b200.synth (numpy RandomState).
"""
import copy
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from svpc_b200.config import tiny_config  # noqa: E402
from svpc_b200 import synth  # noqa: E402
from oracle import refshim  # noqa: E402

CASES = {
    # name: (weights seed, batch kwargs)
    "ragged": (1, dict(B=3, S=4, seed=5, ragged=True, n_ingr=4, n_oov=2)),
    "full_len": (2, dict(B=2, S=3, seed=7, ragged=False, n_ingr=3, n_oov=0)),
}
MODES = ("video", "copy", "reason_copy", "full")
GREEDY_SHARPEN = 20.0


def run_case(mode, case):
    wseed, bkw = CASES[case]
    cfg = tiny_config(mode)
    sd = synth.init_state_dict(cfg, seed=wseed)
    batch = synth.make_batch(cfg, **bkw)
    model = refshim.build_model(cfg, sd)
    loss, preds, ents, acts = refshim.run_forward(model, batch)
    loss.backward()
    out = {"loss": np.float64(loss.item())}
    for b, p in enumerate(preds):
        out[f"pred.{b}"] = p.detach().numpy()
    for b, (e, a) in enumerate(zip(ents, acts)):
        out[f"ent.{b}"] = e.detach().numpy()
        out[f"act.{b}"] = a.detach().numpy()
    for n, p in model.named_parameters():
        if p.grad is not None:
            out["grad." + n] = p.grad.numpy()
    # predict=True contract (dump_memories.py:60-73)
    mem, _, _ = refshim.run_forward(model, batch, predict=True)
    for b, d in enumerate(mem):
        for k, v in d.items():
            if k == "entity_vectors":
                out[f"mem.{b}.entity_vectors.0"] = v[0].detach().numpy()
                out[f"mem.{b}.entity_vectors.1"] = v[1].detach().numpy()
            else:
                out[f"mem.{b}.{k}"] = v.detach().numpy()
    # greedy captions
    sd2 = {k: v.clone() for k, v in sd.items()}
    sd2["decoder_classifier.decoder.weight"] *= GREEDY_SHARPEN
    model2 = refshim.build_model(cfg, sd2)
    tr = refshim.load_translator()
    from easydict import EasyDict
    T = tr.Translator(types.SimpleNamespace(cuda=False),
                      {"model_cfg": EasyDict(dict(cfg)), "model": model2.state_dict()}, model=model2)
    b2 = copy.deepcopy(batch)
    mi = (b2["input_ids_list"], b2["video_features_list"], b2["input_masks_list"], b2["token_type_ids_list"],
          b2["ingr_input_ids"].tolist(), b2["ingr_masks"].tolist(), b2["ingr_sep_masks"].tolist(),
          b2["ingr_id_dict"], b2["oov_word_dict"], b2["alignments"], b2["actions"], b2["batch_step_num"])
    seqs, _ = T.translate_batch(mi)
    for b, s in enumerate(seqs):
        out[f"greedy.{b}"] = s.numpy()
    return out


if __name__ == "__main__":
    assert refshim.available(), "needs /root/reference"
    here = os.path.dirname(os.path.abspath(__file__))
    for case in CASES:
        for mode in MODES:
            out = run_case(mode, case)
            path = os.path.join(here, f"{case}_{mode}.npz")
            np.savez_compressed(path, **out)
            print(path, os.path.getsize(path) // 1024, "KiB", "loss", out["loss"])
