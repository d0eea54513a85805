"""Model configuration for the SVPC hot path.

The reference passes an ``EasyDict`` as the only constructor argument of
``StateAwareRecursiveTransformer`` (reference src/train.py:657-686).  ``easydict`` is not a
dependency of this package, so ``SvpcConfig`` is a small dict-with-attribute-access that is
pickle/`state_dict` compatible with what the reference stores under ``checkpoint["model_cfg"]``.
"""
from __future__ import annotations

# special token ids: reference src/rtransformer/recursive_caption_dataset.py:27-34
PAD, CLS, SEP, VID, BOS, EOS, UNK, IGNORE = 0, 1, 2, 3, 4, 5, 6, -1

MODEL_MODES = ("video", "copy", "reason_copy", "full")            # v / vi / viv / vivt
MODE_ALIASES = {"v": "video", "vi": "copy", "viv": "reason_copy", "vivt": "full"}


class SvpcConfig(dict):
    """dict with attribute access (the part of EasyDict the model reads)."""

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:  # pragma: no cover - mirrors EasyDict behaviour
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v


def youcook2_config(model_mode: str = "full", **over) -> SvpcConfig:
    """Canonical hyper-parameters (reference src/train.py:463-608, scripts/train.sh:19-21)."""
    model_mode = MODE_ALIASES.get(model_mode, model_mode)
    assert model_mode in MODEL_MODES, model_mode
    cfg = SvpcConfig(
        hidden_size=768, intermediate_size=768, vocab_size=951, word_vec_size=300,
        action_vocab_size=384, lstm_hidden_size=768, video_feature_size=3072,
        max_position_embeddings=125, max_v_len=100, max_t_len=25, max_i_len=100,
        model_mode=model_mode, temperature=0.5, lambda_=0.5, type_vocab_size=2, unk_id=UNK,
        layer_norm_eps=1e-12, hidden_dropout_prob=0.1, num_hidden_layers=2,
        num_attention_heads=12, attention_probs_dropout_prob=0.1, n_memory_cells=1,
        memory_dropout_prob=0.1, initializer_range=0.02, label_smoothing=0.1,
        share_wd_cls_weight=False, use_asl=True, xl_grad=False,
    )
    cfg.update(over)
    return cfg


def tiny_config(model_mode: str = "full", **over) -> SvpcConfig:
    """A structurally identical but small configuration used by the golden fixtures and CPU tests."""
    cfg = youcook2_config(model_mode)
    cfg.update(hidden_size=64, intermediate_size=64, lstm_hidden_size=64, vocab_size=60,
               word_vec_size=24, action_vocab_size=16, video_feature_size=96,
               max_position_embeddings=20, max_v_len=12, max_t_len=8, max_i_len=16,
               num_attention_heads=4)
    cfg.update(over)
    return cfg
