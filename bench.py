#!/usr/bin/env python
"""Benchmark of the SVPC hot path on B200 (BASELINE.json).

    python bench.py --gpus 1 --steps 10 --warmup 3                 # this framework, one rank per GPU (torchrun for N>1)
    python bench.py --impl reference --gpus 1 --steps 2 --warmup 1 # the reference algorithm on the host cores (oracle port)

Prints ONE JSON line (rank 0).  Headline (`value`): vivt ('full') training clips/s, forward + backward (+ gradient
all-reduce) of StateAwareRecursiveTransformer, 16 videos x 12 clips per GPU, bf16 path, train-mode dropout, batch resident
in HBM, CUDA-graph replay.  `e2e`: the same step with every input uploaded from pinned host memory each step and the loss
read back.  Beside it, in the same line:

    parity      eval-mode loss of the timed model(s) against the CPU oracle on the same batch (bf16 and fp32 paths)
    fp32_path   the same step on the fp32-accumulate (1e-4) path                                   (N = 1)
    modes       v / vi / viv training steps (BASELINE.json configs 0-2 on the GPU)                 (N = 1)
    ragged      SURVEY 8d secondary run: ragged batch, eager launches and graph replay             (N = 1)
    decode      greedy decode captions/s, batch sweep 16 / 64 / 256 / 512 videos sharded over the ranks (config 4)
    roofline    whole-step needed TFLOP/s against the measured tensor peak + the dominant tcgen05 GEMM timed alone
    cpu_baseline  the oracle on the host cores, full 192-clip batch                                (N = 1)
"""
import argparse
import copy
import gc
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

MODE_NAMES = {"video": "v", "copy": "vi", "reason_copy": "viv", "full": "vivt"}
# fwd+bwd GFLOP per clip of the reference as written (FlopCounterMode), and what this build needs:
# the last clip-encoder layer is evaluated for the [CLS] row only (-0.498 GF fwd, x3 for fwd+bwd) and, in eval mode, LN(3072)'s
# affine is folded into the projection so the 19200x768x3072 input-gradient GEMM is not needed (-0.472 GF; in train mode the
# dropout between LN and the projection forbids the folding, but the input gradient is still not needed: only dgamma / dbeta).
GF_PER_CLIP_AS_WRITTEN = {"video": 7.037, "copy": 7.097, "reason_copy": 7.131, "full": 7.189}
GF_SAVED_PER_CLIP = 3 * 0.498 + 2 * 100 * 3072 * 768 / 1e9
GF_PER_CAPTION_AS_WRITTEN, GF_PER_CAPTION_NEEDED = 12.06, 2.4
DECODE_SWEEP = (16, 64, 256, 512)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--mode", default="full", choices=list(MODE_NAMES))
    ap.add_argument("--videos", type=int, default=16, help="videos per GPU (reference batch size, train.py:543)")
    ap.add_argument("--clips", type=int, default=12)
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--task", default="all", choices=["all", "train", "decode"],
                    help="all: headline line with every block; train: headline step only; decode: greedy-decode line only")
    ap.add_argument("--decode-videos", type=int, nargs="*", default=list(DECODE_SWEEP), help="total videos of the decode sweep")
    ap.add_argument("--cpu-decode-videos", type=int, default=2, help="videos of the CPU greedy-decode sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-optimizer", action="store_true", help="skip the fwd+bwd+optimizer iteration measurement")
    ap.add_argument("--no-kernel-timing", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--eval-mode", action="store_true", help="run the training step with dropout off (model.eval())")
    ap.add_argument("--no-graph", action="store_true", help="enqueue every kernel from Python instead of replaying a CUDA graph")
    ap.add_argument("--ragged", action="store_true", help="headline step on the ragged batch of SURVEY 8d (train task)")
    ap.add_argument("--fp32-features", action="store_true",
                    help="bf16 path: keep the clip features fp32 in the input arena (default: the loader hands them over as bf16)")
    ap.add_argument("--trace", default=None, help="diagnostic: write the kernel timeline of 3 replays (start us, end us, name) to this file")
    ap.add_argument("--inject-noise", action="store_true",
                    help="upload the Gumbel noise of the re-simulator with the batch instead of drawing it on the device")
    ap.add_argument("--verify-ddp", action="store_true",
                    help="N >= 2: the NCCL data-parallel step on N shards against a 1-rank step on the concatenated batch")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------ helpers
def bind_to_gpu_numa(local_rank):
    """Pin this rank's host threads to the CPUs NVML reports as local to its GPU, BEFORE any pinned buffer is allocated:
    first-touch then places the input arena on the GPU's NUMA node.  Returns (description, previous affinity)."""
    prev = os.sched_getaffinity(0)
    try:
        import pynvml
        pynvml.nvmlInit()
        pr = torch.cuda.get_device_properties(local_rank)
        bus = "%08x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        h = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
        pynvml.nvmlDeviceSetCpuAffinity(h)
        now = os.sched_getaffinity(0)
        return f"gpu-local cpus via nvml ({len(now)} of {len(prev)})", prev
    except Exception as e:                                   # no NVML / no permission: leave the scheduler alone
        return f"unchanged ({type(e).__name__})", prev


def physical_cores():
    """Host threads for the CPU arm: physical cores available to this process (SMT siblings make torch's CPU GEMMs slower)."""
    avail = len(os.sched_getaffinity(0))
    try:
        import psutil
        phys = psutil.cpu_count(logical=False) or avail
    except Exception:
        phys = avail
    return max(1, min(avail, phys))


class ClockSampler(threading.Thread):
    """Samples SM clock / throttle reasons with nvidia-smi while the timed region runs."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self._stop_evt = index, [], threading.Event()

    def run(self):
        while not self._stop_evt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self._stop_evt.wait(0.2)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=6)
        sm = sorted(int(float(r[0])) for r in self.rows if r and r[0].replace(".", "").isdigit())
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        mx = [int(float(r[1])) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx[0] if mx else None,
                "reasons": sorted(reasons), "samples": len(self.rows)}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        with open(p) as f:
            d = json.load(f)
        return d.get("hbm_gbs", 6650.0), d.get("bf16_tflops_sustained", 1400.0), d.get("bf16_tflops", 1590.0), "measured (MEASURED_PEAKS.json)"
    return 6650.0, 1400.0, 1590.0, "fallback (B200_PROFILING.md)"


def workload_name(mode, videos, clips):
    return (f"{MODE_NAMES[mode]} training step fwd+bwd, {videos} videos x {clips} clips per GPU, "
            f"100 video + 25 text tokens, 3072-d features, Ni=10, hidden 768 x 2 layers")


def is_tlist(x):
    return isinstance(x, list) and len(x) > 0 and isinstance(x[0], torch.Tensor)


def tree(fn, x):
    if isinstance(x, torch.Tensor):
        return fn(x)
    if is_tlist(x):
        return [fn(t) for t in x]
    return x


def nbytes(x):
    if isinstance(x, torch.Tensor):
        return x.numel() * x.element_size()
    if is_tlist(x):
        return sum(t.numel() * t.element_size() for t in x)
    return 0


def free_cuda():
    gc.collect()
    torch.cuda.empty_cache()


class Timer:
    """K calls of fn between two CUDA events on the current stream, barrier + synchronize on both sides, max over ranks."""

    def __init__(self, dev, world):
        self.dev, self.world = dev, world

    def barrier(self):
        if self.world > 1:
            import torch.distributed as dist
            dist.barrier()
        torch.cuda.synchronize()

    def __call__(self, fn, steps):
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        self.barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=self.dev)
        if self.world > 1:
            import torch.distributed as dist
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms)


# ------------------------------------------------------------------------------------------------ reference arm (host cores)
def cpu_reference_train(cfg, sd, args, steps, warmup, cores):
    """The reference algorithm (oracle restatement, fp32) on the host cores, on the SAME batch the GPU arm times."""
    from oracle import svpc_oracle as O
    from svpc_b200 import synth
    torch.set_num_threads(cores)
    batch = synth.make_batch(cfg, B=args.videos, S=args.clips, seed=2019, ragged=False, n_ingr=10, n_oov=0)
    times = []
    for i in range(warmup + steps):
        b = copy.deepcopy(batch)               # (v mode rewrites labels in place) - outside the timed region
        t0 = time.perf_counter()
        O.loss_and_grads(sd, cfg, b)
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    per_step = sum(times) / len(times)
    return batch["n_clips"] / per_step, per_step, batch["n_clips"]


def cpu_reference_decode(cfg, sd, args, steps, cores, n_videos):
    from oracle import svpc_oracle as O
    from svpc_b200 import synth
    torch.set_num_threads(cores)
    batch = synth.make_batch(cfg, B=n_videos, S=args.clips, seed=2019, ragged=False, n_ingr=10, n_oov=0)
    ts = []
    for i in range(steps):
        b = copy.deepcopy(batch)
        t0 = time.perf_counter()
        O.greedy_decode(sd, cfg, b)
        ts.append(time.perf_counter() - t0)
    per = sum(ts) / len(ts)
    return batch["n_clips"] / per, per, batch["n_clips"]


def reference_arm(args, cfg, sd, rank):
    if rank != 0:
        return
    cores = physical_cores()
    if args.task == "decode":
        v, per, n = cpu_reference_decode(cfg, sd, args, max(1, args.steps), cores, args.cpu_decode_videos)
        sample = (f"oracle port of Translator.translate_batch_greedy (no KV cache, as the reference), fp32, {n} captions per step "
                  f"({args.cpu_decode_videos} videos x {args.clips} clips), torch CPU {cores} threads")
        metric, unit = "vivt greedy-decode captions/sec", "captions/s"
        workload = workload_name(args.mode, args.videos, args.clips).replace("training step fwd+bwd", "greedy decode")
    else:
        v, per, n = cpu_reference_train(cfg, sd, args, max(1, args.steps), max(0, min(args.warmup, 1)), cores)
        sample = (f"oracle port of model.py fwd+bwd (eval-mode arithmetic), fp32, the full batch of the GPU arm: {args.videos} videos x "
                  f"{args.clips} clips ({n} clips) per step, torch CPU {cores} threads (physical cores)")
        metric, unit = "vivt train clips/sec", "clips/s"
        workload = workload_name(args.mode, args.videos, args.clips)
    print(json.dumps({
        "impl": "reference", "metric": metric, "value": round(v, 3), "unit": unit, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(per * 1e3, 2), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
        "config": {"workload": workload, "sample": sample},
        "cpu_baseline": {"value": round(v, 3), "unit": unit, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": round(v, 3), "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


# ------------------------------------------------------------------------------------------------ training step
class TrainBench:
    """One model + one batch + (optionally) the captured step."""

    def __init__(self, mode, cd, args, dev, rank, world, train=True, ragged=False, graph=True, reducer_on=True):
        from svpc_b200 import synth, backend
        from svpc_b200.config import youcook2_config
        from svpc_b200.model import StateAwareRecursiveTransformer
        from svpc_b200.ddp import GradAllReducer
        self.mode, self.cd, self.dev, self.world = mode, cd, dev, world
        self.cfg = youcook2_config(mode)
        self.sd = synth.init_state_dict(self.cfg, seed=2019)
        self.model = StateAwareRecursiveTransformer(self.cfg, compute_dtype=cd)
        self.model.load_state_dict(self.sd)
        self.model.to(dev)
        self.model.train(train)
        self.reducer = GradAllReducer(self.model) if reducer_on else None
        self.be = backend.get()
        if ragged:      # SURVEY 8d secondary run: video length U{20..98}, words U{4..20}, S_b U{4..12}, Ni U{3..15}
            self.batch = synth.make_batch(self.cfg, B=args.videos, S=args.clips, seed=2019 + rank, ragged=True, n_ingr=10, n_oov=2)
        else:
            self.batch = synth.make_batch(self.cfg, B=args.videos, S=args.clips, seed=2019 + rank, ragged=False, n_ingr=10, n_oov=0)
        self.n_clips = self.batch["n_clips"]
        # bf16 path: the clip features are handed over as bf16 (what svpc_b200.features.FeatureReader writes into bf16 arena
        # views): half the upload; LN(3072), the first kernel to read them, computes its statistics in fp32 either way
        self.feature_dtype = torch.bfloat16 if (cd == torch.bfloat16 and not args.fp32_features) else torch.float32
        fa = synth.forward_args(self.batch)
        fa[1] = [t.to(self.feature_dtype) for t in fa[1]]
        self.host_args = [tree(lambda t: t.pin_memory(), a) for a in fa]
        # Gumbel noise: the reference draws it on the GPU inside the step (F.gumbel_softmax, model.py:1018) and so does the timed
        # step (svpc_gumbel_noise, seeded per step); --inject-noise uploads the batch's noise instead.  The parity forward always
        # gets the batch's noise: the oracle needs the same numbers.
        inject = getattr(args, "inject_noise", False)
        self.noise_host = [g.pin_memory() for g in self.batch["gumbel"]] if inject else None
        self.dev_args = [tree(lambda t: t.to(dev, non_blocking=True), a) for a in self.host_args]
        self.noise_dev = [g.to(dev, non_blocking=True) for g in self.noise_host] if inject else None
        self.parity_noise = [g.to(dev) for g in self.batch["gumbel"]]
        torch.cuda.synchronize()
        self.gstep, self.per_replay = None, 0
        if graph:
            from svpc_b200.graph import GraphedTrainStep
            self.be.reset_launch_count()
            self.gstep = GraphedTrainStep(self.model, self.dev_args, gumbel_noise=self.noise_dev,
                                          reducer=self.reducer if (world > 1 and reducer_on) else None, warmup=2)
            self.per_replay = self.be.launch_count() // 3          # 2 eager warm-ups + the captured step
        self.run = self.gstep.run if self.gstep is not None else self.eager_step

    def eager_step(self):
        self.model.zero_grad(set_to_none=True)
        loss = self.model(*self.dev_args, gumbel_noise=self.noise_dev)[0]
        loss.backward()
        if self.reducer is not None:
            self.reducer.finish()
        return loss

    def eval_loss(self):
        """Eval-mode forward of THIS model on its batch (dropout off: what the oracle states)."""
        was = self.model.training
        self.model.eval()
        with torch.no_grad():
            loss = float(self.model(*self.dev_args, gumbel_noise=self.parity_noise)[0])
        self.model.train(was)
        return loss

    def close(self):
        self.gstep = None
        self.run = None
        self.model = None
        self.reducer = None
        free_cuda()


def measure_train(mode, cd, args, dev, rank, world, timer, ragged=False, graph=True, want_eval_loss=False):
    tb = TrainBench(mode, cd, args, dev, rank, world, train=not args.eval_mode, ragged=ragged, graph=graph)
    for _ in range(max(args.warmup, 3)):
        loss = tb.run()
    torch.cuda.synchronize()
    assert torch.isfinite(loss).all(), f"non-finite loss ({mode}, {cd})"
    ms = timer(tb.run, args.steps)
    out = {"value": round(world * tb.n_clips * args.steps / (ms / 1e3), 2), "unit": "clips/s", "ms_per_step": round(ms / args.steps, 3),
           "clips_per_step_per_gpu": tb.n_clips}
    if want_eval_loss:
        out["_eval_loss"] = tb.eval_loss()
    tb.close()
    return out


def measure_ragged_stream(mode, cd, args, dev, n_sig=4, rounds=6):
    """A loader that buckets by batch signature: `n_sig` different ragged batches arriving round robin through
    GraphedStepCache (capture on the second sight, replay afterwards).  Steady-state clips/s once every signature is
    captured, and the cache statistics of the whole stream."""
    from svpc_b200 import synth
    from svpc_b200.config import youcook2_config
    from svpc_b200.graph import GraphedStepCache
    from svpc_b200.model import StateAwareRecursiveTransformer
    cfg = youcook2_config(mode)
    model = StateAwareRecursiveTransformer(cfg, compute_dtype=cd)
    model.load_state_dict(synth.init_state_dict(cfg, seed=2019))
    model.to(dev).train(not args.eval_mode)
    fdt = torch.bfloat16 if (cd == torch.bfloat16 and not args.fp32_features) else torch.float32
    batches = []
    for i in range(n_sig):
        b = synth.make_batch(cfg, B=args.videos, S=args.clips, seed=3000 + i, ragged=True, n_ingr=10, n_oov=2)
        fa = synth.forward_args(b)
        fa[1] = [t.to(fdt) for t in fa[1]]
        batches.append(([tree(lambda t: t.to(dev), a) for a in fa], [g.to(dev) for g in b["gumbel"]], b["n_clips"]))
    cache = GraphedStepCache(model, max_graphs=n_sig, capture_after=2)
    for r in range(2):                                    # first sight eager, second sight captures
        for a, g, _ in batches:
            cache.step(a, g)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    clips = 0
    e0.record()
    for r in range(rounds):
        for a, g, n in batches:
            loss = cache.step(a, g)
            clips += n
    e1.record()
    torch.cuda.synchronize()
    assert torch.isfinite(loss).all()
    ms = e0.elapsed_time(e1)
    out = {"value": round(clips / (ms / 1e3), 2), "unit": "clips/s", "ms_per_step": round(ms / (rounds * n_sig), 3),
           "what": f"{n_sig} ragged batch signatures round robin through GraphedStepCache, steady state (inputs resident, D2D load into the static buffers included)",
           "cache": cache.stats()}
    del cache, model
    free_cuda()
    return out


# ------------------------------------------------------------------------------------------------ greedy decode
def tile_decode_inputs(batch, n_videos):
    """The 12-tuple of Translator.translate_batch for `n_videos` videos, tiled from a 16-video synthetic batch (the
    decoder shares nothing between videos, so repeated content does not change the work)."""
    B = len(batch["batch_step_num"])
    idx = [i % B for i in range(n_videos)]
    it = torch.tensor(idx)
    pick_t = lambda t: t[it].contiguous()
    pick_l = lambda l: [l[i] for i in idx]
    return ([pick_t(t) for t in batch["input_ids_list"]], [pick_t(t) for t in batch["video_features_list"]],
            [pick_t(t) for t in batch["input_masks_list"]], [pick_t(t) for t in batch["token_type_ids_list"]],
            pick_t(batch["ingr_input_ids"]), pick_t(batch["ingr_masks"]), pick_t(batch["ingr_sep_masks"]),
            pick_l(batch["ingr_id_dict"]), pick_l(batch["oov_word_dict"]), pick_l(batch["alignments"]),
            pick_l(batch["actions"]), pick_l(batch["batch_step_num"]))


def measure_decode(model, batch, n_videos, args, dev, world, timer, be, e2e=True):
    """Greedy decode of `n_videos` videos on THIS rank (the sweep's total is n_videos * world): captions/s resident and,
    with pinned host inputs uploaded every step and the captions read back, end to end."""
    from svpc_b200.graph import GraphedDecoder
    host = tile_decode_inputs(batch, n_videos)
    n_caps = int(sum(host[11]))
    fdt = torch.bfloat16 if model.compute_dtype == torch.bfloat16 and not args.fp32_features else torch.float32
    resident = tuple(tree(lambda t: t.to(dev), x) for x in host)
    be.reset_launch_count()
    gdec = GraphedDecoder(model, resident, warmup=1, feature_dtype=fdt)
    per_replay = be.launch_count() // 2
    steps = max(2, min(args.steps, 10))
    with torch.no_grad():
        for _ in range(3):
            seqs = gdec.run()
        ms = timer(gdec.run, steps)
    out = {"videos_per_gpu": n_videos, "videos_total": n_videos * world, "captions_per_step": n_caps * world,
           "value": round(world * n_caps * steps / (ms / 1e3), 1), "unit": "captions/s", "ms_per_step": round(ms / steps, 3),
           "gpu_launches_per_step": per_replay}
    if e2e:
        # the loader's collate target: the pinned host twin of the decoder's input arena (features as `fdt`)
        hp = gdec.host_inputs()
        for i in GraphedDecoder.TENSOR_LISTS:
            for d_, s_ in zip(hp[i], host[i]):
                d_.copy_(s_)
        for i in GraphedDecoder.TENSORS:
            hp[i].copy_(torch.as_tensor(host[i]))
        h2d = gdec.arena.h2d_nbytes
        d2h = sum(s.numel() * 8 for s in seqs)

        def e2e_steps(k):
            gdec.prefetch(hp)                           # H2D of every input from pinned memory (copy stream)
            for i in range(k):
                s = gdec.run_prefetched()
                if i + 1 < k:
                    gdec.prefetch(hp)                   # next batch uploads while this one decodes
                outs = [t.cpu() for t in s]             # D2H of the captions
            return outs
        with torch.no_grad():
            e2e_steps(2)
            ms_e = timer(lambda: e2e_steps(steps), 1)
        out["e2e"] = {"value": round(world * n_caps * steps / (ms_e / 1e3), 1), "unit": "captions/s",
                      "ms_per_step": round(ms_e / steps, 3), "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                      "upload_dtype": gdec.arena.feature_dtype}
        del hp
    del gdec, resident, host
    free_cuda()
    return out


def decode_block(args, dev, rank, world, local_rank, timer, be, cd, with_cpu):
    """BASELINE.json configs[4]: vivt greedy decode, batch sweep over total videos, videos sharded over the ranks."""
    from svpc_b200 import synth
    from svpc_b200.config import youcook2_config
    from svpc_b200.model import StateAwareRecursiveTransformer
    cfg = youcook2_config("full")
    sd = synth.init_state_dict(cfg, seed=2019)
    model = StateAwareRecursiveTransformer(cfg, compute_dtype=cd)
    model.load_state_dict(sd)
    model.to(dev).eval()
    base = synth.make_batch(cfg, B=16, S=args.clips, seed=2019 + rank, ragged=False, n_ingr=10, n_oov=0)
    hbm, tf_sus, tf_burst, src = peaks()
    sweep = []
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    for total in args.decode_videos:
        per_rank = max(1, total // world)
        r = measure_decode(model, base, per_rank, args, dev, world, timer, be, e2e=not args.no_e2e)
        tf = r["value"] * GF_PER_CAPTION_NEEDED / 1e3
        r["roofline"] = {"bound": "tensor", "achieved": round(tf, 2), "peak": tf_sus, "unit": "TFLOP/s", "frac": round(tf / tf_sus, 4),
                         "what": f"captions/s x {GF_PER_CAPTION_NEEDED} GF needed per caption (KV cache; {GF_PER_CAPTION_AS_WRITTEN} GF as written)"}
        sweep.append(r)
    clocks = sampler.stop() if rank == 0 else None
    best = max(sweep, key=lambda r: r["value"])
    out = {"metric": "vivt greedy-decode captions/sec", "unit": "captions/s", "value": best["value"], "at_videos_total": best["videos_total"],
           "dtype": "bf16" if cd == torch.bfloat16 else "fp32",
           "config": {"workload": "vivt greedy decode, 12 clips per video, 24 positions, KV cache, all clips of all videos in lock-step",
                      "parallelism": f"videos sharded over {world} rank(s), no collective",
                      "launch": "cuda-graph replay of the whole decode", "gflop_per_caption_as_written": GF_PER_CAPTION_AS_WRITTEN,
                      "gflop_per_caption_needed": GF_PER_CAPTION_NEEDED},
           "sweep": sweep, "clocks": clocks}
    if with_cpu:
        cores = physical_cores()
        v, per, n = cpu_reference_decode(cfg, sd, args, 1, cores, args.cpu_decode_videos)
        out["cpu_baseline"] = {"value": round(v, 3), "unit": "captions/s", "cores": cores, "kind": "port",
                               "sample": f"oracle greedy decode (no KV cache, as the reference) of {n} captions ({args.cpu_decode_videos} videos), {cores} torch threads, {per:.2f} s"}
    del model
    free_cuda()
    return out


# ------------------------------------------------------------------------------------------------ dominant kernel
def dominant_gemm_roofline(tb, be, dev, ms_step, tf_burst, src):
    """The tcgen05 GEMM shape that carries the most algorithmic flops per step, timed ALONE: 20 back-to-back launches
    between two CUDA events on the launching stream, against the BURST peak (a kernel timed in isolation)."""
    was = None
    if tb.reducer is not None:
        was, tb.reducer.enabled = tb.reducer.enabled, False    # rank-0-only pass: no collective may be issued here
    be.start_timing()
    tb.eager_step()
    per = be.stop_timing()
    if was is not None:
        tb.reducer.enabled = was
    tc = []
    for k, (n, t) in per.items():
        if k.startswith("gemm_tc"):
            m, nn, kk = (int(x) for x in k.split(":")[1].split("x"))
            tc.append((2.0 * m * nn * kk * n, n, k))
    if not tc:
        return None
    _, n, k = max(tc, key=lambda x: (x[0], x[2].endswith(":NN")))
    M, N, K = (int(x) for x in k.split(":")[1].split("x"))
    lay = k.split(":")[2]
    a = torch.randn((K, M) if lay[0] == "T" else (M, K), device=dev).to(torch.bfloat16)
    b = torch.randn((K, N) if lay[1] == "T" else (N, K), device=dev).to(torch.bfloat16)
    a = a.t() if lay[0] == "T" else a
    b = b.t() if lay[1] == "T" else b
    out = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
    for _ in range(3):
        be.gemm(a, b, torch.bfloat16, out=out)
    reps = 20
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        be.gemm(a, b, torch.bfloat16, out=out)
    e1.record()
    torch.cuda.synchronize()
    avg_ms = e0.elapsed_time(e1) / reps
    fl = 2.0 * M * N * K
    ach = fl / (avg_ms / 1e3) / 1e12
    traffic, tsrc = None, None
    tpath = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if os.path.isfile(tpath):
        with open(tpath) as f:
            tj = json.load(f)
        traffic, tsrc = tj.get(k), tj.get("_capture")
    return {"kernel": k, "achieved": round(ach, 1), "peak": tf_burst, "unit": "TFLOP/s", "frac": round(ach / tf_burst, 4),
            "peak_source": f"{src}: bf16 burst (kernel timed alone)", "flop_per_launch": fl, "launches_per_step": n,
            "avg_ms": round(avg_ms, 4), "share_of_step": round(avg_ms * n / ms_step, 4), "traffic": traffic,
            "traffic_source": tsrc, "how": "20 back-to-back launches of this shape, CUDA events on the launching stream"}


def h2d_ceiling(dev, nbytes_=256 << 20):
    """Pinned host -> device copy bandwidth of this rank (GB/s): the ceiling of the end-to-end upload."""
    h = torch.empty(nbytes_, dtype=torch.uint8).pin_memory()
    d = torch.empty(nbytes_, dtype=torch.uint8, device=dev)
    d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(4):
        d.copy_(h, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    return 4 * nbytes_ / (e0.elapsed_time(e1) / 1e3) / 1e9


# ------------------------------------------------------------------------------------------------ main
def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    from svpc_b200 import synth
    from svpc_b200.config import youcook2_config
    cfg = youcook2_config(args.mode)
    sd = synth.init_state_dict(cfg, seed=2019)
    workload = workload_name(args.mode, args.videos, args.clips)

    if args.impl == "reference":
        return reference_arm(args, cfg, sd, rank)

    import torch.distributed as dist
    from svpc_b200 import backend
    assert torch.cuda.is_available(), "bench.py needs a GPU (the SVPC operators have no CPU path)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    affinity, full_affinity = bind_to_gpu_numa(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    cd = torch.bfloat16 if args.dtype == "bf16" else torch.float32
    be = backend.get()
    timer = Timer(dev, world)
    hbm, tf_sus, tf_burst, peak_src = peaks()

    if args.verify_ddp:
        from tools.verify_ddp import run as verify_run
        return verify_run(args, dev, rank, world)

    if args.task == "decode":
        blk = decode_block(args, dev, rank, world, local_rank, timer, be, cd, with_cpu=(world == 1 and not args.no_cpu_baseline))
        if rank == 0:
            best = max(blk["sweep"], key=lambda r: r["value"])
            print(json.dumps({"task": "decode", "metric": blk["metric"], "value": blk["value"], "unit": blk["unit"], "n_gpus": world,
                              "steps": args.steps, "warmup": 3, "ms_per_step": best["ms_per_step"], "higher_is_better": True,
                              "scaling": "strong", "vs_baseline": None, "dtype": blk["dtype"], "data": "synthetic", "config": blk["config"],
                              "e2e": best.get("e2e"), "gpu_launches": best["gpu_launches_per_step"] * args.steps, "clocks": blk["clocks"],
                              "roofline": best["roofline"], "cpu_baseline": blk.get("cpu_baseline"), "sweep": blk["sweep"]}))
        return finish(world)

    # ---------------------------------------------------------------- headline: training step
    tb = TrainBench(args.mode, cd, args, dev, rank, world, train=not args.eval_mode, ragged=args.ragged, graph=not args.no_graph)
    n_clips = tb.n_clips
    feat_name = "bf16 (rounded once by the loader)" if tb.feature_dtype == torch.bfloat16 else "fp32"
    for _ in range(max(args.warmup, 3)):
        loss = tb.run()
    torch.cuda.synchronize()
    assert torch.isfinite(loss).all(), "non-finite loss"

    if args.trace:
        # diagnostic: kernel timeline of 3 replays through the torch profiler (CUPTI), summarised per stream on rank 0
        from torch.profiler import profile, ProfilerActivity
        if world > 1:
            dist.barrier()
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            for _ in range(3):
                tb.run()
            torch.cuda.synchronize()
        if rank == 0:
            ev = [e for e in prof.events() if e.device_type.name == "CUDA" or "cuda" in str(e.device_type).lower()]
            rows = sorted(((e.time_range.start, e.time_range.end, e.name) for e in ev), key=lambda r: r[0])
            with open(args.trace, "w") as f:
                for a, b, n in rows:
                    f.write(f"{a:.3f}\t{b:.3f}\t{n}\n")
            print(json.dumps({"trace": args.trace, "events": len(rows)}))
        return finish(world)

    # ---- parity of the timed model: eval-mode loss on the timed batch against the CPU oracle (rank 0)
    parity = None
    if rank == 0 and world == 1 and not args.no_parity:      # (N > 1: the other ranks would spin at a barrier on the cores the oracle needs)
        from oracle import svpc_oracle as O
        torch.set_num_threads(physical_cores())
        t0 = time.perf_counter()
        ref = float(O.forward(tb.sd, tb.cfg, copy.deepcopy(tb.batch))[0])
        got = tb.eval_loss()
        tol = 1e-2 if cd == torch.bfloat16 else 1e-4
        parity = {"oracle_loss": round(ref, 4), "what": "eval-mode loss of the timed model on the timed batch vs oracle/svpc_oracle.py (fp32, CPU)",
                  args.dtype: {"loss": round(got, 4), "rel_err": round(abs(got - ref) / abs(ref), 8), "tol": tol,
                               "ok": bool(abs(got - ref) <= tol * abs(ref))},
                  "oracle_seconds": round(time.perf_counter() - t0, 2)}
        assert parity[args.dtype]["ok"], f"parity of the timed model failed: {parity}"
    if world > 1:
        dist.barrier()

    # ---- device-resident timing (`value`)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    be.reset_launch_count()
    ms = timer(tb.run, args.steps)
    launches = be.launch_count() if tb.gstep is None else tb.per_replay * args.steps
    clocks = sampler.stop() if rank == 0 else None
    value = world * n_clips * args.steps / (ms / 1e3)
    ms_step = ms / args.steps

    # ---- whole training iteration: step + global clip + BertAdam + EMA as three more launches
    iteration = None
    if not args.no_optimizer:
        from svpc_b200 import optim as SO
        named = [(n, p) for n, p in tb.model.named_parameters() if p.requires_grad]
        nod = ["bias", "LayerNorm.bias", "LayerNorm.weight"]                      # train.py:339-343
        groups = [{"params": [p for n, p in named if not any(x in n for x in nod)], "weight_decay": 0.01},
                  {"params": [p for n, p in named if any(x in n for x in nod)], "weight_decay": 0.0}]
        opt = SO.BertAdam(groups, lr=1e-4, warmup=0.1, t_total=100000, schedule="warmup_linear")   # train.py:353-357
        ema = SO.EMA(0.9999)                                                      # train.py:344-348, 529
        for n, p in named:
            ema.register(n, p.data)
        it = [0]

        def full_iteration():
            l_ = tb.run()
            opt.step(grad_clip=1.0, ema=ema, niter=it[0])                         # train.py:141-147
            it[0] += 1
            return l_
        for _ in range(3):
            full_iteration()
        ms_it = timer(full_iteration, args.steps)
        iteration = {"value": round(world * n_clips * args.steps / (ms_it / 1e3), 2), "unit": "clips/s",
                     "ms_per_step": round(ms_it / args.steps, 3),
                     "what": "fwd+bwd replay, then clip_grad_norm_(1.0) + BertAdam(lr 1e-4, warmup_linear) + EMA(0.9999) fused into 3 launches over 181 tensors"}
        assert torch.isfinite(tb.run()).all()

    # ---- end-to-end timing: pinned host inputs -> H2D every step, loss read back
    e2e = None
    if not args.no_e2e:
        h2d_bytes = sum(nbytes(a) for a in tb.host_args) + sum(nbytes(g) for g in (tb.noise_host or []))
        gstep = tb.gstep
        if gstep is not None:
            # the loader's collate target: pinned host views of the step's input arena
            arena_args, arena_noise = gstep.host_inputs()
            for dst, src in zip(arena_args, tb.host_args):
                if isinstance(dst, torch.Tensor):
                    dst.copy_(src)
                elif is_tlist(dst):
                    for d_, s_ in zip(dst, src):
                        d_.copy_(s_)
            for d_, s_ in zip(arena_noise or [], tb.noise_host or []):
                d_.copy_(s_)
            h2d_bytes = gstep.arena.h2d_nbytes     # the 25 never-read text rows of each (125, 3072) clip block are not uploaded

        def e2e_steps(k):
            """k steps, each with its own H2D of every input from pinned host memory and a D2H read of the loss.  With
            the graph, batch i+1 uploads on the copy stream while batch i computes (GraphedTrainStep.prefetch)."""
            if gstep is not None:
                gstep.prefetch(arena_args, arena_noise)
                pending, losses = None, []
                for i in range(k):
                    gstep.run_prefetched()                  # staged batch -> static buffers, graph replay (async)
                    h = gstep.loss_to_host()                # async D2H of THIS step's loss into a pinned ring slot
                    if i + 1 < k:
                        gstep.prefetch(arena_args, arena_noise)
                    if pending is not None:
                        losses.append(pending.get())        # host reads step i-1's loss while step i runs
                    pending = h
                losses.append(pending.get())
                assert len(losses) == k and all(v == v for v in losses)
                return
            for _ in range(k):
                a = [tree(lambda t: t.to(dev, non_blocking=True), x) for x in tb.host_args]
                nz = [g.to(dev, non_blocking=True) for g in tb.noise_host] if tb.noise_host is not None else None
                tb.model.zero_grad(set_to_none=True)
                l_ = tb.model(*a, gumbel_noise=nz)[0]
                l_.backward()
                if tb.reducer is not None:
                    tb.reducer.finish()
                float(l_)
        e2e_steps(2)
        ms_e2e = timer(lambda: e2e_steps(args.steps), 1)
        ceil = h2d_ceiling(dev)
        e2e = {"value": round(world * n_clips * args.steps / (ms_e2e / 1e3), 2), "unit": "clips/s",
               "h2d_bytes_per_step": int(h2d_bytes), "d2h_bytes_per_step": 4, "ms_per_step": round(ms_e2e / args.steps, 3),
               "h2d_ceiling_gbs": round(ceil, 1), "h2d_ms_at_ceiling": round(h2d_bytes / ceil / 1e6, 3),
               "upload_dtype": getattr(gstep.arena, "feature_dtype", "fp32") if gstep is not None else "fp32",
               "gumbel_noise": "uploaded with the batch (--inject-noise)" if tb.noise_host is not None else
                               "drawn on the device inside the step (svpc_gumbel_noise; the reference's F.gumbel_softmax draws it on the GPU too)",
               "pipeline": "every step uploads all of its inputs from pinned host memory; the upload of step i+1 runs on a copy stream under step i (first upload exposed); each step's loss goes to pinned host memory and is read by the host one step later" if gstep is not None else "serial"}

    # ---- roofline: the whole step against the tensor peak + the dominant tcgen05 GEMM alone
    gf_need = GF_PER_CLIP_AS_WRITTEN[args.mode] - GF_SAVED_PER_CLIP
    step_tf = gf_need * n_clips / ms_step          # per GPU: GFLOP needed per step / ms per step = TFLOP/s
    roofline = {"bound": "tensor", "achieved": round(step_tf, 1), "peak": tf_sus, "unit": "TFLOP/s", "frac": round(step_tf / tf_sus, 4),
                "what": "whole step per GPU (CUDA-graph replay): needed flops per step / step time, against the SUSTAINED bf16 peak",
                "peak_source": peak_src, "gflop_per_clip_needed": round(gf_need, 3), "traffic": None, "dominant_kernel": None}
    if not args.no_kernel_timing and rank == 0 and cd == torch.bfloat16:
        dk = dominant_gemm_roofline(tb, be, dev, ms_step, tf_burst, peak_src)
        roofline["dominant_kernel"] = dk
        if dk is not None:
            roofline["traffic"] = dk["traffic"]
    if world > 1:
        dist.barrier()
    eval_loss_main = None
    tb.close()
    del tb

    # ---- the other BASELINE.json configurations and the 1e-4 path (N = 1 only)
    fp32_path = modes = ragged = None
    if args.task == "all" and world == 1 and not args.ragged and not args.no_graph:
        other = torch.float32 if cd == torch.bfloat16 else torch.bfloat16
        oname = "fp32" if other == torch.float32 else "bf16"
        r = measure_train(args.mode, other, args, dev, rank, world, timer, want_eval_loss=parity is not None)
        if parity is not None:
            got = r.pop("_eval_loss")
            tol = 1e-2 if other == torch.bfloat16 else 1e-4
            ref = parity["oracle_loss"]
            parity[oname] = {"loss": round(got, 4), "rel_err": round(abs(got - ref) / abs(ref), 8), "tol": tol,
                             "ok": bool(abs(got - ref) <= tol * abs(ref))}
            assert parity[oname]["ok"], f"parity of the {oname} path failed: {parity}"
        r["dtype"] = oname
        r["what"] = "the same step on the fp32-storage, fp32-accumulate parity path (1e-4)" if other == torch.float32 else "bf16 path"
        r["ratio_to_headline"] = round(r["value"] / value, 4)
        fp32_path = r
        modes = {}
        for m in ("video", "copy", "reason_copy"):
            if m == args.mode:
                continue
            modes[MODE_NAMES[m]] = measure_train(m, cd, args, dev, rank, world, timer)
        rg = measure_train(args.mode, cd, args, dev, rank, world, timer, ragged=True, graph=True)
        re_ = measure_train(args.mode, cd, args, dev, rank, world, timer, ragged=True, graph=False)
        ragged = {"what": "SURVEY 8d secondary run: video length U{20..98}, words U{4..22}, S_b U{4..12}, Ni U{3..15}, 2 OOV words per video",
                  "graph_replay": rg, "eager_launches": re_, "signature_cache_stream": measure_ragged_stream(args.mode, cd, args, dev),
                  "note": "a captured step is keyed on (clips, ingredients, OOV width, copy tokens) per video; a loader that does not bucket by that signature runs the eager figure"}

    # ---- greedy decode sweep (all N; videos sharded over the ranks)
    decode = None
    if args.task == "all" and not args.ragged:
        decode = decode_block(args, dev, rank, world, local_rank, timer, be, cd, with_cpu=(world == 1 and not args.no_cpu_baseline))

    # ---- CPU baseline (rank 0, N = 1): the reference algorithm on the host cores, the same full batch
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        os.sched_setaffinity(0, full_affinity)              # the CPU baseline gets every host core back
        cores = physical_cores()
        v, per_step, n = cpu_reference_train(cfg, sd, args, 2, 1, cores)
        cpu = {"value": round(v, 3), "unit": "clips/s", "cores": cores, "kind": "port",
               "sample": f"oracle fwd+bwd fp32 on the full batch: {n} clips ({args.videos} videos x {args.clips}), {cores} torch threads (physical cores), {per_step:.2f} s/step"}

    if rank == 0:
        out = {
            "metric": "vivt train clips/sec", "value": round(value, 2), "unit": "clips/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": round(ms_step, 3), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
            "config": {"workload": workload if not args.ragged else workload + " (ragged)", "parallelism": f"dp{world}", "host_affinity": affinity,
                       "launch": "eager" if args.no_graph else "cuda-graph replay (fwd+bwd+weight refresh captured once)",
                       "global_videos": world * args.videos, "weights": "random-init (init_bert_weights scale), seed 2019",
                       "dropout": "off (model.eval())" if args.eval_mode else "train mode: hidden 0.1, attention 0.1, action selector 0.4 (reference defaults), counter-based masks recomputed in backward",
                       "l2": "inputs (features: 236 MB fp32 / 118 MB bf16 per step) exceed / fill the 126 MB L2; activations (~1 GB per step) evict them between steps",
                       "features": feat_name,
                       "gflop_per_clip_as_written": GF_PER_CLIP_AS_WRITTEN[args.mode], "gflop_per_clip_needed": round(gf_need, 3),
                       "step_tflops_needed": round(step_tf, 2)},
            "e2e": e2e, "iteration_with_optimizer": iteration, "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline,
            "cpu_baseline": cpu, "parity": parity, "fp32_path": fp32_path, "modes": modes, "ragged": ragged, "decode": decode,
        }
        print(json.dumps(out))
    finish(world)


def finish(world):
    if world > 1:
        import torch.distributed as dist
        # a captured graph holds NCCL kernels and autograd hooks: tearing the communicator down under it can block, so the
        # ranks meet once more, flush, and leave without running destructors
        dist.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


if __name__ == "__main__":
    main()
