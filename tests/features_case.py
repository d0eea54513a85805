"""Seeded clip-feature case shared by tests/golden/make_features_golden.py (reference side) and tests/test_features.py."""
import os

import numpy as np

MAX_V_LEN, MAX_T_LEN = 12, 5          # 10 feature slots per clip: short enough that long clips take the down-sampling branch
COLS = (16, 8)                        # columns of <v>_resnet.npy / <v>_bn.npy (2048 / 1024 in YouCook2)

# name -> (rows, dtype of the two files, frm2sec)
VIDEOS = {
    "vidA": (40, ("<f4", "<f4"), 1.0),
    "vidB": (23, ("<f8", "<f4"), 0.5),
    "vidC": (7, ("<f4", "<f8"), 2.5),
    "vidD": (1, ("<f4", "<f4"), 1.0),
    "vidE": (64, ("<f2", "<f2"), 0.25),
}

# (video, [st, ed] seconds): in-range, clipped at the end of the file, down-sampled (> 10 rows), single frame, fractional
# boundaries, a clip that ends in frame 0 of a one-row file (st = -1, numpy broadcasts the row), and zero-length clips
CLIPS = [
    ("vidA", [0.0, 5.0]), ("vidA", [3.2, 9.9]), ("vidA", [10.0, 35.0]), ("vidA", [38.0, 90.0]), ("vidA", [0.0, 39.0]),
    ("vidB", [0.0, 3.0]), ("vidB", [2.4, 4.1]), ("vidB", [5.0, 30.0]), ("vidB", [11.0, 11.2]),
    ("vidC", [0.0, 17.5]), ("vidC", [2.5, 5.0]), ("vidC", [16.0, 40.0]),
    ("vidD", [0.0, 0.0]), ("vidD", [0.0, 3.0]),
    ("vidE", [1.0, 2.0]), ("vidE", [0.3, 15.9]), ("vidE", [15.0, 16.0]),
]
# clips the reference refuses (:411 raises ValueError: empty slice into two rows)
BAD_CLIPS = [("vidA", [0.0, 0.0]), ("vidC", [0.0, 0.0])]


def write_files(directory, seed=2019):
    rng = np.random.default_rng(seed)
    for name, (rows, dts, _) in VIDEOS.items():
        for suffix, cols, dt in zip(("_resnet.npy", "_bn.npy"), COLS, dts):
            np.save(os.path.join(directory, name + suffix), rng.standard_normal((rows, cols)).astype(dt))
