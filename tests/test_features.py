"""Clip-feature reader: oracle vs the reference fixture, then ``svpc_b200.features`` + ``svpc_features_gather``
vs both, bit for bit (the copy is fp32 / a rounding conversion: no tolerance)."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from oracle import features_oracle as FO
from oracle import refshim
from svpc_b200 import cuda_backend
from svpc_b200.features import FeatureReader, clip_rows, convert_to_feat_index_st_ed
from tests import features_case as FC

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden", "features_clips.npz")


@pytest.fixture(scope="module")
def feat_dir(tmp_path_factory):
    d = str(tmp_path_factory.mktemp("feats"))
    FC.write_files(d)
    return d


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


def test_oracle_matches_reference_fixture(feat_dir, gold):
    for i, (name, ts) in enumerate(FC.CLIPS):
        raw = FO.load_video_feature(feat_dir, name)
        feat, tokens, mask = FO.load_indexed_video_feature(raw, ts, FC.VIDEOS[name][2], FC.MAX_V_LEN, FC.MAX_T_LEN)
        assert feat.dtype == np.float32 and np.array_equal(feat, gold["feat_%d" % i]), (i, name, ts)
        assert tokens == gold["tokens_%d" % i].tolist() and mask == gold["mask_%d" % i].tolist()
        assert list(FO.feat_index_st_ed(len(raw), ts, FC.VIDEOS[name][2])) == gold["st_ed_%d" % i].tolist()
    for i, (name, ts) in enumerate(FC.BAD_CLIPS):
        assert str(gold["bad_%d" % i]) == "ValueError"
        with pytest.raises(ValueError):
            FO.load_indexed_video_feature(FO.load_video_feature(feat_dir, name), ts, FC.VIDEOS[name][2], FC.MAX_V_LEN,
                                          FC.MAX_T_LEN)


@pytest.mark.skipif(not refshim.available(), reason="reference tree not present on this box")
def test_fixture_is_what_the_live_reference_produces(tmp_path, gold):
    env = dict(os.environ)
    subprocess.run([sys.executable, os.path.join(ROOT, "tests", "golden", "make_features_golden.py")], check=True, env=env,
                   capture_output=True)
    again = np.load(GOLD)
    assert sorted(again.files) == sorted(gold.files)
    for k in gold.files:
        assert np.array_equal(again[k], gold[k]), k


def test_reader_single_clips_bit_exact(feat_dir, gold):
    rd = FeatureReader(feat_dir, FC.MAX_V_LEN, FC.MAX_T_LEN, n_threads=3)
    for i, (name, ts) in enumerate(FC.CLIPS):
        frm2sec = FC.VIDEOS[name][2]
        feat, tokens, mask = rd.load_indexed_video_feature(name, ts, frm2sec)
        assert feat.dtype == torch.float32 and tuple(feat.shape) == (FC.MAX_V_LEN + FC.MAX_T_LEN, sum(FC.COLS))
        assert np.array_equal(feat.numpy(), gold["feat_%d" % i]), (i, name, ts)
        assert tokens == gold["tokens_%d" % i].tolist() and mask == gold["mask_%d" % i].tolist()
        assert list(convert_to_feat_index_st_ed(rd.info(name)[0], ts, frm2sec)) == gold["st_ed_%d" % i].tolist()


def test_reader_refuses_what_the_reference_refuses(feat_dir):
    rd = FeatureReader(feat_dir, FC.MAX_V_LEN, FC.MAX_T_LEN)
    for name, ts in FC.BAD_CLIPS:
        with pytest.raises(ValueError):
            rd.load_indexed_video_feature(name, ts, FC.VIDEOS[name][2])
    with pytest.raises(FileNotFoundError):
        rd.load_indexed_video_feature("no_such_video", [0.0, 1.0], 1.0)


def test_step_batch_into_a_strided_arena(feat_dir, gold):
    """A whole ragged batch in ONE library call, written into NaN-poisoned (B, L, D) step tensors: every row of every block is
    either a feature row or zero, padding slots hold clip (0, 0) as caption_collate pads (:562-567)."""
    rd = FeatureReader(feat_dir, FC.MAX_V_LEN, FC.MAX_T_LEN, n_threads=4)
    by_video = {}
    for i, (name, ts) in enumerate(FC.CLIPS):
        by_video.setdefault(name, []).append((i, ts))
    examples = [(name, [ts for _, ts in v], FC.VIDEOS[name][2]) for name, v in by_video.items()]
    B, S = len(examples), max(len(e[1]) for e in examples)
    L, D = FC.MAX_V_LEN + FC.MAX_T_LEN, sum(FC.COLS)
    views = [torch.full((B, L, D), float("nan")) for _ in range(S)]
    meta = rd.load_step_batch(examples, views)
    first = gold["feat_%d" % by_video[examples[0][0]][0][0]]
    for b, (name, tss, _) in enumerate(examples):
        assert len(meta[b]) == len(tss)
        for t in range(S):
            if t < len(tss):
                i = by_video[name][t][0]
                assert np.array_equal(views[t][b].numpy(), gold["feat_%d" % i]), (b, t)
                assert meta[b][t][0] == gold["tokens_%d" % i].tolist() and meta[b][t][1] == gold["mask_%d" % i].tolist()
            else:
                assert np.array_equal(views[t][b].numpy(), first), (b, t)


def test_bf16_blocks_are_the_rounded_fp32_blocks(feat_dir, gold):
    """bf16 destination (the half-size upload of the bf16 path): every value is the reference's fp32 value rounded to nearest
    even, the zero rows stay zero, in the same one-call gather into a strided arena."""
    rd = FeatureReader(feat_dir, FC.MAX_V_LEN, FC.MAX_T_LEN, n_threads=2)
    L, D = FC.MAX_V_LEN + FC.MAX_T_LEN, sum(FC.COLS)
    arena = torch.full((len(FC.CLIPS), L, D), float("nan"), dtype=torch.bfloat16)
    clips = [(name, ts, FC.VIDEOS[name][2]) for name, ts in FC.CLIPS]
    rd.load_clips(clips, [arena[i] for i in range(len(clips))])
    for i in range(len(clips)):
        want = torch.from_numpy(gold["feat_%d" % i]).to(torch.bfloat16)
        assert torch.equal(arena[i], want), i
    with pytest.raises(ValueError):
        rd.load_clips(clips[:1], [torch.zeros(L, D, dtype=torch.float16)])


def test_full_width_rows_and_downsampling_match_oracle(tmp_path):
    """YouCook2 widths (2048 + 1024), 100 video slots: long clips are down-sampled to 98 rows (:403-404)."""
    rng = np.random.default_rng(7)
    d = str(tmp_path)
    np.save(os.path.join(d, "v_resnet.npy"), rng.standard_normal((300, 2048)).astype(np.float32))
    np.save(os.path.join(d, "v_bn.npy"), rng.standard_normal((300, 1024)).astype(np.float32))
    rd = FeatureReader(d, 100, 25)
    raw = FO.load_video_feature(d, "v")
    for ts in ([0.0, 20.0], [13.7, 250.2], [100.0, 197.0], [100.0, 198.0], [290.0, 1000.0]):
        want, tokens, mask = FO.load_indexed_video_feature(raw, ts, 1.0, 100, 25)
        got, t2, m2 = rd.load_indexed_video_feature("v", ts, 1.0)
        assert np.array_equal(got.numpy(), want) and t2 == tokens and m2 == mask, ts
        rows, _, _ = clip_rows(300, ts, 1.0, 100)
        assert len(rows) <= 98


def test_npy_header_variants(tmp_path):
    d = str(tmp_path)
    rd = FeatureReader(d, 6, 2)
    a = np.arange(20, dtype=np.float32).reshape(5, 4)

    def put(name, res, bn):
        for suffix, arr in (("_resnet.npy", res), ("_bn.npy", bn)):
            if isinstance(arr, tuple):                                   # (array, format version)
                with open(os.path.join(d, name + suffix), "wb") as f:
                    np.lib.format.write_array(f, arr[0], version=arr[1])
            else:
                np.save(os.path.join(d, name + suffix), arr)

    put("v2", (a, (2, 0)), (a * 2, (3, 0)))
    got, _, _ = rd.load_indexed_video_feature("v2", [0.0, 2.0], 1.0)
    assert np.array_equal(got[1:4].numpy(), np.concatenate([a[0:3], 2 * a[0:3]], axis=1)) and float(got[0].abs().sum()) == 0
    put("fortran", np.asfortranarray(a), a)
    put("rank1", a[0].copy(), a)
    put("ints", a.astype(np.int32), a)
    put("rows", a[:4].copy(), a)
    for name in ("fortran", "rank1", "ints", "rows"):
        with pytest.raises(ValueError):
            rd.load_indexed_video_feature(name, [0.0, 2.0], 1.0)
    with open(os.path.join(d, "short_resnet.npy"), "wb") as f:
        f.write(open(os.path.join(d, "v2_resnet.npy"), "rb").read()[:-8])
    np.save(os.path.join(d, "short_bn.npy"), a)
    with pytest.raises(ValueError):
        rd.load_indexed_video_feature("short", [0.0, 2.0], 1.0)


def test_destination_layout_is_checked(feat_dir):
    rd = FeatureReader(feat_dir, FC.MAX_V_LEN, FC.MAX_T_LEN)
    L, D = FC.MAX_V_LEN + FC.MAX_T_LEN, sum(FC.COLS)
    for bad in (torch.empty(L, D, dtype=torch.float64), torch.empty(L - 1, D), torch.empty(L, D + 1), torch.empty(D, L).t()):
        with pytest.raises(ValueError):
            rd.load_clips([("vidA", [0.0, 5.0], 1.0)], [bad])


@pytest.mark.gpu
def test_reader_to_arena_to_device(feat_dir, gold):
    """file -> pinned arena (written in place by the library) -> pitched upload that skips the text rows -> device."""
    from svpc_b200.graph import InputArena
    dev = torch.device("cuda:0")
    names = ["vidA", "vidB", "vidE"]
    clips = {n: [(i, ts) for i, (m, ts) in enumerate(FC.CLIPS) if m == n][:3] for n in names}
    S, B = 3, len(names)
    L, D = FC.MAX_V_LEN + FC.MAX_T_LEN, sum(FC.COLS)
    arena = InputArena([torch.empty(B, L, D) for _ in range(S)], dev, used_rows={t: FC.MAX_V_LEN for t in range(S)})
    for v in arena.host_views():
        v.fill_(float("nan"))
    rd = FeatureReader(feat_dir, FC.MAX_V_LEN, FC.MAX_T_LEN)
    rd.load_step_batch([(n, [ts for _, ts in clips[n]], FC.VIDEOS[n][2]) for n in names], arena.host_views())
    arena.upload()
    torch.cuda.synchronize()
    got = [v.cpu() for v in arena._views(arena.stage)]
    for b, n in enumerate(names):
        for t, (i, _) in enumerate(clips[n]):
            assert np.array_equal(got[t][b, :FC.MAX_V_LEN].numpy(), gold["feat_%d" % i][:FC.MAX_V_LEN]), (n, t)
