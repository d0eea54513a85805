"""Generates tests/golden/features_clips.npz with the UNMODIFIED reference methods
``RecursiveCaptionDataset._load_indexed_video_feature`` / ``_convert_to_feat_index_st_ed``
(/root/reference/src/rtransformer/recursive_caption_dataset.py:379-416) on the files and clips of tests/features_case.py,
preceded by the reference's own ``np.concatenate([np.load(...), np.load(...)], axis=1)`` (:187-189) and followed by its
``.astype(np.float32)`` (:332).  Shims: the nltk / easydict stubs of oracle/refshim.py, and ``np.int = int`` (the alias the
reference uses at :403 was removed in numpy 1.24).

    python tests/golden/make_features_golden.py
"""
import importlib
import os
import sys
import tempfile
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import refshim  # noqa: E402
from tests import features_case as FC  # noqa: E402


def main():
    refshim.load()
    if not hasattr(np, "int"):
        np.int = int
    ds = importlib.import_module("src.rtransformer.recursive_caption_dataset").RecursiveCaptionDataset
    self = types.SimpleNamespace(max_v_len=FC.MAX_V_LEN, max_t_len=FC.MAX_T_LEN, CLS_TOKEN=ds.CLS_TOKEN, VID_TOKEN=ds.VID_TOKEN,
                                 SEP_TOKEN=ds.SEP_TOKEN, PAD_TOKEN=ds.PAD_TOKEN,
                                 _convert_to_feat_index_st_ed=ds._convert_to_feat_index_st_ed)
    out = {}
    with tempfile.TemporaryDirectory() as d:
        FC.write_files(d)
        raw = {}
        for name in FC.VIDEOS:
            raw[name] = np.concatenate([np.load(os.path.join(d, "{}_resnet.npy".format(name))),
                                        np.load(os.path.join(d, "{}_bn.npy".format(name)))], axis=1)      # :187-189
        for i, (name, ts) in enumerate(FC.CLIPS):
            feat, tokens, mask = ds._load_indexed_video_feature(self, raw[name], ts, FC.VIDEOS[name][2])
            out["feat_%d" % i] = feat.astype(np.float32)                                                   # :332
            out["tokens_%d" % i] = np.array(tokens)
            out["mask_%d" % i] = np.array(mask, dtype=np.int64)
            out["st_ed_%d" % i] = np.array(ds._convert_to_feat_index_st_ed(len(raw[name]), ts, FC.VIDEOS[name][2]))
        for i, (name, ts) in enumerate(FC.BAD_CLIPS):
            try:
                ds._load_indexed_video_feature(self, raw[name], ts, FC.VIDEOS[name][2])
                out["bad_%d" % i] = np.array("ok")
            except Exception as e:  # noqa: BLE001
                out["bad_%d" % i] = np.array(type(e).__name__)
    path = os.path.join(ROOT, "tests", "golden", "features_clips.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes;", [str(out["bad_%d" % i]) for i in range(len(FC.BAD_CLIPS))])


if __name__ == "__main__":
    main()
