"""TEST INFRASTRUCTURE ONLY - numpy restatement of the reference's clip-feature loading
(``src/rtransformer/recursive_caption_dataset.py``), used by tests/ as the checker of ``svpc_b200/features.py`` +
``svpc_features_gather``.  Pinned by ``tests/golden/features_clips.npz`` (made by ``tests/golden/make_features_golden.py``
from the UNMODIFIED reference methods) in ``tests/test_features.py``.  Never imported by the product."""
import math
import os

import numpy as np


def load_video_feature(video_feature_dir, video_name):
    """:187-189 - (T, 2048 + 1024) array in the files' dtype."""
    return np.concatenate([np.load(os.path.join(video_feature_dir, "{}_resnet.npy".format(video_name))),
                           np.load(os.path.join(video_feature_dir, "{}_bn.npy".format(video_name)))], axis=1)


def feat_index_st_ed(feat_len, timestamp, frm2sec):
    """:379-387"""
    st = int(math.floor(timestamp[0] / frm2sec))
    ed = int(math.ceil(timestamp[1] / frm2sec))
    ed = min(ed, feat_len - 1)
    st = min(st, ed - 1)
    assert st <= ed <= feat_len
    return st, ed


def load_indexed_video_feature(raw_feat, timestamp, frm2sec, max_v_len, max_t_len):
    """:389-416 followed by the ``.astype(np.float32)`` of :332 -> (feat (max_v_len + max_t_len, D) float32, tokens, mask)."""
    max_v_l = max_v_len - 2
    feat_len = len(raw_feat)
    st, ed = feat_index_st_ed(feat_len, timestamp, frm2sec)
    feat = np.zeros((max_v_len + max_t_len, raw_feat.shape[1]))                    # float64, as the reference
    if ed - st + 1 > max_v_l:
        idx = np.linspace(st, ed, max_v_l, endpoint=True).astype(int).tolist()
        assert max(idx) < feat_len
        feat[1:max_v_l + 1] = raw_feat[idx]
        tokens = ["[CLS]"] + ["[VID]"] * max_v_l + ["[SEP]"]
        mask = [1] * (max_v_l + 2)
    else:
        valid_l = ed - st + 1
        feat[1:valid_l + 1] = raw_feat[st:ed + 1]                                   # ValueError when the slice is short
        tokens = ["[CLS]"] + ["[VID]"] * valid_l + ["[SEP]"] + ["[PAD]"] * (max_v_l - valid_l)
        mask = [1] * (valid_l + 2) + [0] * (max_v_l - valid_l)
    return feat.astype(np.float32), tokens, mask
