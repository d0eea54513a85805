"""Clip features from the per-video ``.npy`` files straight into the pinned input arena.

Host mirror of the reference data layer for the one thing the hot path consumes from it, the ``video_feature`` block of a
clip (``src/rtransformer/recursive_caption_dataset.py``):

* ``convert_to_feat_index_st_ed``  = ``_convert_to_feat_index_st_ed``  (:379-387)
* ``clip_rows``                    = the row selection, tokens and mask of ``_load_indexed_video_feature`` (:389-416)
* ``FeatureReader.load_indexed_video_feature`` = ``np.load`` x 2 + ``np.concatenate`` (:187-189) + ``_load_indexed_video_feature``
  + ``.astype(np.float32)`` (:332) for one clip; ``FeatureReader.load_clips`` does a whole step's clips in one call of
  ``svpc_features_gather`` (include/svpc_b200.h), writing into caller-owned fp32 blocks - normally the host views of
  ``graph.InputArena``, so the rows go file -> pinned arena -> device with no intermediate array.

The row arithmetic is plain Python (a few integers per clip); the bytes are moved by ``libsvpc_b200.so``.  There is no
numpy fallback for the copy: a missing library raises.
"""
from __future__ import annotations

import ctypes as C
import math
import os

import numpy as np
import torch

from . import cuda_backend

PAD_TOKEN, CLS_TOKEN, SEP_TOKEN, VID_TOKEN = "[PAD]", "[CLS]", "[SEP]", "[VID]"      # recursive_caption_dataset.py:20-23


def convert_to_feat_index_st_ed(feat_len, timestamp, frm2sec):
    """Wall-clock (st, ed) seconds -> feature row indices (:379-387)."""
    st = int(math.floor(timestamp[0] / frm2sec))
    ed = int(math.ceil(timestamp[1] / frm2sec))
    ed = min(ed, feat_len - 1)
    st = min(st, ed - 1)
    assert st <= ed <= feat_len, "st {} <= ed {} <= feat_len {}".format(st, ed, feat_len)
    return st, ed


def clip_rows(feat_len, timestamp, frm2sec, max_v_len):
    """(rows, video_tokens, mask) of one clip (:389-416): ``rows[i]`` is the feature row that lands at position ``1 + i`` of
    the clip block (position 0 is [CLS] and stays zero).  More than ``max_v_len - 2`` rows are down-sampled with
    ``np.linspace(st, ed, max_v_len - 2).astype(int)`` (:403-404)."""
    max_v_l = max_v_len - 2
    st, ed = convert_to_feat_index_st_ed(feat_len, timestamp, frm2sec)
    if ed - st + 1 > max_v_l:
        rows = np.linspace(st, ed, max_v_l, endpoint=True).astype(np.int64)
        assert rows.max() < feat_len
        tokens = [CLS_TOKEN] + [VID_TOKEN] * max_v_l + [SEP_TOKEN]
        mask = [1] * (max_v_l + 2)
    else:
        valid_l = ed - st + 1
        # raw_feat[st:ed + 1] with Python slice semantics (:411): st is -1 when the clip ends in the first frame, the slice
        # is then empty (numpy refuses the assignment) unless the file has a single row (numpy broadcasts it)
        picked = list(range(feat_len))[st:ed + 1]
        if len(picked) == 1 and valid_l > 1:
            picked = picked * valid_l
        if len(picked) != valid_l:
            raise ValueError("could not broadcast input array from shape ({},D) into shape ({},D)".format(len(picked), valid_l))
        rows = np.asarray(picked, dtype=np.int64)
        tokens = [CLS_TOKEN] + [VID_TOKEN] * valid_l + [SEP_TOKEN] + [PAD_TOKEN] * (max_v_l - valid_l)
        mask = [1] * (valid_l + 2) + [0] * (max_v_l - valid_l)
    return rows, tokens, mask


class FeatureReader:
    """Reads ``<video>_resnet.npy`` and ``<video>_bn.npy`` (:187-188) rows into fp32 clip blocks."""

    SUFFIXES = ("_resnet.npy", "_bn.npy")

    def __init__(self, video_feature_dir, max_v_len, max_t_len, n_threads=None, fill_rows=None):
        """``fill_rows``: rows of each clip block the reader owns (features or zeros).  Default: all ``max_v_len + max_t_len``
        (what the reference produces); ``max_v_len`` leaves the text rows untouched - the model never reads them
        (model.py:889) and ``InputArena(used_rows=...)`` never uploads them."""
        self.fill_rows = fill_rows
        self.lib = cuda_backend.load_library()
        self.lib.svpc_npy_info.argtypes = [C.c_char_p, C.POINTER(C.c_longlong), C.POINTER(C.c_longlong), C.POINTER(C.c_int),
                                           C.POINTER(C.c_longlong)]
        self.dir = video_feature_dir
        self.max_v_len, self.max_t_len = max_v_len, max_t_len
        self.n_threads = n_threads or min(16, os.cpu_count() or 1)
        self._info = {}

    def paths(self, video_name):
        return [os.path.join(self.dir, video_name + s) for s in self.SUFFIXES]

    def info(self, video_name):
        """(rows, [columns of each file]) from the .npy headers; both files must have the same number of rows, as
        ``np.concatenate(axis=1)`` demands (:189)."""
        hit = self._info.get(video_name)
        if hit is None:
            rows, cols = [], []
            for p in self.paths(video_name):
                r, c, dt, off = C.c_longlong(), C.c_longlong(), C.c_int(), C.c_longlong()
                if self.lib.svpc_npy_info(p.encode(), C.byref(r), C.byref(c), C.byref(dt), C.byref(off)) != 0:
                    msg = self.lib.svpc_last_error().decode()
                    raise (FileNotFoundError if "cannot open" in msg else ValueError)(msg)
                rows.append(r.value)
                cols.append(c.value)
            if rows[0] != rows[1]:
                raise ValueError("all the input array dimensions except for the concatenation axis must match exactly, "
                                 "but along dimension 0 the arrays have sizes {} and {}".format(*rows))
            hit = self._info[video_name] = (rows[0], cols)
        return hit

    @property
    def block_rows(self):
        return self.max_v_len + self.max_t_len

    def load_clips(self, clips, dsts):
        """``clips``: list of ``(video_name, timestamp, frm2sec)``; ``dsts``: one fp32 (or bf16: the half-size upload of
        the bf16 path, values rounded to nearest even) CPU tensor ``(max_v_len + max_t_len, D)`` per clip with unit column
        stride and a common row pitch (views of a ``(B, L, D)`` arena tensor qualify).  Returns ``[(video_tokens, mask)]``.
        One library call for all clips."""
        assert len(clips) == len(dsts)
        if not clips:
            return []
        out, paths, row_start, rows, ptrs, col_off = [], [], [0], [], [], []
        pitch = dsts[0].stride(0)
        dt0 = dsts[0].dtype
        if dt0 not in (torch.float32, torch.bfloat16):
            raise ValueError("clip blocks must be fp32 or bf16, got {}".format(dt0))
        for (name, timestamp, frm2sec), d in zip(clips, dsts):
            feat_len, cols = self.info(name)
            r, tokens, mask = clip_rows(feat_len, timestamp, frm2sec, self.max_v_len)
            out.append((tokens, mask))
            if (d.dtype != dt0 or d.is_cuda or d.dim() != 2 or d.shape[0] != self.block_rows or
                    d.shape[1] != sum(cols) or d.stride(1) != 1 or d.stride(0) != pitch):
                raise ValueError("clip block must be a CPU fp32 / bf16 ({}, {}) tensor with unit column stride and row pitch {}, got "
                                 "{} {} strides {}".format(self.block_rows, sum(cols), pitch, d.dtype, tuple(d.shape), d.stride()))
            c0 = 0
            for p, c in zip(self.paths(name), cols):
                paths.append(p.encode())
                rows.extend(r.tolist())
                row_start.append(len(rows))
                ptrs.append(d.data_ptr())
                col_off.append(c0)
                c0 += c
        n = len(paths)
        rc = self.lib.svpc_features_gather(
            n, (C.c_char_p * n)(*paths), (C.c_longlong * (n + 1))(*row_start), (C.c_longlong * max(len(rows), 1))(*rows),
            (C.c_void_p * n)(*ptrs), 1 if dt0 == torch.bfloat16 else 0, C.c_longlong(pitch), (C.c_longlong * n)(*col_off), 1,
            int(self.fill_rows or self.block_rows), int(self.n_threads))
        if rc != 0:
            raise cuda_backend.SvpcError("svpc_features_gather failed ({}): {}".format(rc, self.lib.svpc_last_error().decode()))
        return out

    def load_indexed_video_feature(self, video_name, timestamp, frm2sec, out=None):
        """One clip: ``(feat (max_v_len + max_t_len, D) fp32, video_tokens, mask)`` - what ``clip_sentence_to_feature`` keeps
        of ``_load_indexed_video_feature`` (:295, :332)."""
        if out is None:
            out = torch.empty(self.block_rows, sum(self.info(video_name)[1]), dtype=torch.float32)
        (tokens, mask), = self.load_clips([(video_name, timestamp, frm2sec)], [out])
        return out, tokens, mask

    def load_step_batch(self, examples, step_views):
        """A collated batch: ``examples[b] = (video_name, [timestamps of its clips], frm2sec)``; ``step_views[t]`` is the
        ``(B, L, D)`` host tensor of recurrent step ``t`` (``video_features_list[t]``, train.py:98).  Videos with fewer clips
        than ``len(step_views)`` get the first clip of the first video in their padding slots, as ``caption_collate`` does
        (:562-567; the model never reads them past ``batch_step_num``).  Returns ``[[(tokens, mask)] * S_b]`` per video."""
        clips, dsts, owner = [], [], []
        pad_clip = (examples[0][0], examples[0][1][0], examples[0][2])
        for b, (name, timestamps, frm2sec) in enumerate(examples):
            for t, view in enumerate(step_views):
                real = t < len(timestamps)
                clips.append((name, timestamps[t], frm2sec) if real else pad_clip)
                dsts.append(view[b])
                owner.append(b if real else -1)
        res = self.load_clips(clips, dsts)
        out = [[] for _ in examples]
        for b, r in zip(owner, res):
            if b >= 0:
                out[b].append(r)
        return out
