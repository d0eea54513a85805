"""Host-side feature loading for one training step: the reference's sequence
(np.load x 2 + np.concatenate per video, float64 scratch + .astype(float32) per clip, default_collate per recurrent step;
recursive_caption_dataset.py:187-189, 389-416, 332, 519-527 - restated by oracle/features_oracle.py) against
svpc_b200.features.FeatureReader writing straight into a pinned arena.  Synthetic YouCook2-shaped files, page-cache warm.
Prints one JSON line."""
import json
import os
import sys
import tempfile
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import features_oracle as FO  # noqa: E402
from svpc_b200.features import FeatureReader  # noqa: E402

B, S, T, MAX_V, MAX_T = 16, 12, 1200, 100, 25


def main():
    rng = np.random.default_rng(0)
    reps = int(os.environ.get("REPS", "5"))
    with tempfile.TemporaryDirectory() as d:
        examples = []
        for b in range(B):
            name = "v%02d" % b
            np.save(os.path.join(d, name + "_resnet.npy"), rng.standard_normal((T, 2048)).astype(np.float32))
            np.save(os.path.join(d, name + "_bn.npy"), rng.standard_normal((T, 1024)).astype(np.float32))
            st = np.sort(rng.uniform(0, T - 100, S))
            examples.append((name, [[float(s), float(s + rng.uniform(40, 97))] for s in st], 1.0))
        L, D = MAX_V + MAX_T, 3072
        pinned = torch.cuda.is_available()
        views = [torch.empty(B, L, D).pin_memory() if pinned else torch.empty(B, L, D) for _ in range(S)]
        rd = FeatureReader(d, MAX_V, MAX_T)
        rd.load_step_batch(examples, views)                                   # warm the page cache and the header cache

        def reference_step():
            per_video = []
            for name, tss, f2s in examples:
                raw = FO.load_video_feature(d, name)
                per_video.append([FO.load_indexed_video_feature(raw, ts, f2s, MAX_V, MAX_T)[0] for ts in tss])
            return [torch.utils.data.dataloader.default_collate([per_video[b][t] for b in range(B)]) for t in range(S)]

        ref = reference_step()
        for t in range(S):
            assert torch.equal(ref[t], views[t]), t
        useful = sum(int((v[:, :MAX_V].abs().sum(-1) > 0).sum()) for v in views) * D * 4
        t_ref, t_new = [], []
        for _ in range(reps):
            t0 = time.perf_counter(); reference_step(); t_ref.append(time.perf_counter() - t0)
            t0 = time.perf_counter(); rd.load_step_batch(examples, views); t_new.append(time.perf_counter() - t0)
        rd2 = FeatureReader(d, MAX_V, MAX_T, fill_rows=MAX_V)
        t_new2 = []
        for _ in range(reps):
            t0 = time.perf_counter(); rd2.load_step_batch(examples, views); t_new2.append(time.perf_counter() - t0)
        for t in range(S):
            assert torch.equal(ref[t], views[t]), t
        src = torch.empty(B * S * L * D)
        dst = torch.empty_like(src)
        t0 = time.perf_counter(); dst.copy_(src); t_cp = time.perf_counter() - t0
        out = {"workload": "16 videos x 12 clips, files (1200, 2048) + (1200, 1024) fp32, clips of 41-98 rows",
               "feature_bytes_per_step": useful, "block_bytes_per_step": B * S * L * D * 4,
               "reference_ms": 1e3 * min(t_ref), "reader_ms": 1e3 * min(t_new), "speedup": min(t_ref) / min(t_new),
               "reader_video_rows_only_ms": 1e3 * min(t_new2),
               "reader_GBps_written": B * S * L * D * 4 / min(t_new) / 1e9, "torch_copy_GBps": src.numel() * 4 / t_cp / 1e9,
               "threads": rd.n_threads, "host_cores": os.cpu_count(), "pinned": pinned, "bit_identical": True}
        print(json.dumps(out))


if __name__ == "__main__":
    main()
