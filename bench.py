#!/usr/bin/env python
"""bench.py -- throughput of the B200 GMM-HMM hot path (BASELINE.json metric), one JSON line on rank 0.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload em|decode|fb|connected|stress|fit|frontend]
                    [--scaling strong|weak] [--impl ours|reference] [--no-sub]

The default line is BASELINE config 2 (`configs[1]`, the configuration the EM metric is quoted on): one segmental-K-means
+ EM iteration over 100 k labelled 100-frame utterances (Viterbi alignment -> posteriors -> sufficient statistics ->
allreduce -> M-step).  At N = 1 the other four configurations and the full fit ride along as sub-lines, each a complete record of its
own (value, roofline with ncu traffic, cpu_baseline, e2e):
  decode     config 1 shape at benchmark scale: every 100-frame utterance scored against all 11 word models (11 x 5 x 4,
             D = 39) and Viterbi-decoded (core.py:79-87)
  fb         config 3: one Baum-Welch iteration, 8 states x 16 Gaussians per word, 1 M ragged utterances of 50-300 frames
  connected  config 4: connected digit strings, 7 layers x 11 words (393 states), 4 096 utterances x 500 frames, graph
             Viterbi with back-pointers, traceback and word strings
  stress     config 5: 2 000 states x 32 Gaussians x 39 dims on 10 M frames, best state per frame
  fit        config 2 in full: HMM.fit of all 11 words on the 100 k utterances (skmeans, split K-means and EM to the
             reference's own stop tests, re-alignment), float64 passes; strong scaling at N > 1
At N > 1 the EM line is STRONG scaling by default (BASELINE config 2: "100k synthetic utterances, sharded across
1/2/4/8"): the corpus is split over the ranks; `--scaling weak` gives every rank its own 100 k utterances.

A "step" is one pass of the workload over the rank's whole shard.  `value` is measured with the shard resident in HBM;
`e2e` goes through the host-facing API with pinned HOST buffers, H2D and D2H inside the timed region.  Every shard
(>= 0.19 GB) is larger than the 126 MB L2, so no flush is needed between steps.
"""
import argparse
import csv
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "speech-recognition_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402

W_, S_, M_, D_ = 11, 5, 4, 39
NCOMP = int(np.log(M_)) and 2 ** int(np.log(M_))   # hmm.py:104: n_splits = int(np.log(4)) = 1 -> 2 trained components
FMA_PER_FRAME_DECODE = W_ * S_ * M_ * 2 * D_          # 17 160 FMA per frame (all 11 words scored)

# stdout carries exactly ONE JSON line: the descriptor is saved here and fd 1 is pointed at stderr for the rest of the
# run, so that whatever a library prints (NCCL's version banner, for one) cannot land next to it.
_STDOUT_FD = os.dup(1)
os.dup2(2, 1)


def emit(line):
    os.write(_STDOUT_FD, (json.dumps(line) + "\n").encode())


def log(*a):
    print("[bench]", *a, file=sys.stderr, flush=True)


# ------------------------------------------------------------------------------------------ ncu traffic from profiles/
_UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def ncu_traffic(key, frames):
    """DRAM bytes of the dominant kernel per launch, re-derived at run time from the committed `ncu --set full` capture
    named in profiles/ncu_index.json: (dram__bytes_read.sum + dram__bytes_write.sum) of the captured launch, divided by
    the frames that launch processed, times the frames of a launch here."""
    try:
        idx = json.load(open(os.path.join(ROOT, "profiles", "ncu_index.json")))[key]
        rows = list(csv.reader(open(os.path.join(ROOT, idx["csv"]))))
        hdr, units = rows[0], rows[1]
        kn, rd, wr = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
        row = next(r for r in rows[2:] if idx["match"] in r[kn])
        total = float(row[rd]) * _UNIT[units[rd]] + float(row[wr]) * _UNIT[units[wr]]
        per = total / idx["frames"]
        return {"traffic": per * frames,
                "traffic_source": f"{per:.1f} B/frame (dram read+write of one {idx['frames']}-frame launch) x frames per "
                                  f"launch, {idx['csv']}"}
    except Exception as e:  # noqa: BLE001
        return {"traffic": None, "traffic_source": f"no ncu --set full capture indexed for {key} ({type(e).__name__})"}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default=os.environ.get("SRB_WORKLOAD", "em"),
                    choices=["decode", "em", "connected", "stress", "fb", "fit", "frontend"])
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--scaling", default=os.environ.get("SRB_SCALING", "strong"), choices=["strong", "weak"],
                    help="em at N > 1: split 100 k utterances over the ranks (strong) or 100 k per rank (weak)")
    ap.add_argument("--utts", type=int, default=100_000, help="utterances (em / decode: 100 frames each)")
    ap.add_argument("--frames", type=int, default=100)
    ap.add_argument("--cpu-utts", type=int, default=0, help="utterances of the CPU sample (0 = auto)")
    ap.add_argument("--no-sub", "--no-decode", dest="no_sub", action="store_true",
                    help="skip the sub-lines (configs 1, 3, 4, 5) of the default line")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline legs")
    ap.add_argument("--no-e2e", action="store_true", help="skip the end-to-end legs")
    ap.add_argument("--ffma", action="store_true", help="force the FP32-FMA kernels instead of the tcgen05 ones")
    ap.add_argument("--unfused", action="store_true", help="em: the separate K1/K2/K5 kernels instead of the fused pass")
    ap.add_argument("--fb-utts", type=int, default=1_000_000, help="config 3: utterances per GPU")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """Samples SM clock / throttle reasons of one GPU with NVML while the timed region runs."""

    def __init__(self, index):
        self.index, self.samples, self.reasons, self.stop_flag = index, [], set(), False
        self.max_mhz, self.thread = None, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self):
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
                 "hw_power_brake_slowdown": 0x80, "sync_boost": 0x10}
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.02)

    def start(self):
        if self.nv:
            self.thread = threading.Thread(target=self._loop, daemon=True)
            self.thread.start()
        return self

    def stop(self):
        self.stop_flag = True
        if self.thread:
            self.thread.join()
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 2250.0, "source": "fallback"}


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def timed_steps(fn, steps, warmup, torch, dist, world):
    """W warm-up + K timed steps, CUDA events per step on the current stream, barrier + synchronize on both sides.
    -> (total_ms max over ranks, per-step ms list of this rank)"""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    ev[0].record()
    for i in range(steps):
        fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    total = ev[0].elapsed_time(ev[steps])
    per = [ev[i].elapsed_time(ev[i + 1]) for i in range(steps)]
    if world > 1:
        t = torch.tensor([total], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total = float(t.item())
    return total, per


def sub_steps(args):
    """Steps / warm-up of the sub-lines: the timing rules' three warm-up steps, at most five timed ones."""
    return max(1, min(args.steps, 5)), 3


# ------------------------------------------------------------------------------------------ rooflines
def roof_fma(fma_pf, N, ms, peak_fma, kernel, note, key=None):
    flops = 2.0 * fma_pf * N / (ms * 1e-3)
    out = {"bound": "fp32_fma", "achieved": flops / 1e12, "peak": 2 * peak_fma / 1e12, "unit": "TFLOP/s",
           "frac": flops / (2 * peak_fma), "kernel": kernel, "kernel_ms": ms,
           "peak_source": "srb_fma_peak (FFMA2 probe) measured in this run", "algorithmic": note}
    out.update(ncu_traffic(key, N) if key else {"traffic": None})
    return out


def roof_tensor(fma_pf, N, ms, peak_fma, kernel, note, key, executed_flop_pf):
    """3xTF32 kernels: `achieved` is the ALGORITHMIC FP32-equivalent rate (SURVEY 8d's FMA per frame); `peak` is the rate
    an FP32-accurate tensor-core kernel can reach at best: the measured dense bf16 rate, of which TF32 gets half and the
    three TF32 products per FP32 product a third of that."""
    pk = peaks()
    bf16 = pk.get("bf16_tflops", 2250.0)
    flops = 2.0 * fma_pf * N / (ms * 1e-3)
    peak = bf16 / 2 / 3
    out = {"bound": "tensor", "achieved": flops / 1e12, "peak": peak, "unit": "TFLOP/s", "frac": flops / 1e12 / peak,
           "kernel": kernel, "kernel_ms": ms,
           "peak_source": ("MEASURED_PEAKS.json bf16_tflops (burst)" if "source" not in pk else "nominal 2250")
                          + " / 2 (TF32) / 3 (TF32 products per FP32 product)",
           "algorithmic": note,
           "tensor_pipe": {"executed_tflops": executed_flop_pf * N / (ms * 1e-3) / 1e12, "tf32_peak_tflops": bf16 / 2,
                           "frac": executed_flop_pf * N / (ms * 1e-3) / 1e12 / (bf16 / 2)},
           "fp32_fma": {"peak_tflops": 2 * peak_fma / 1e12, "frac": flops / (2 * peak_fma),
                        "note": "algorithmic rate over the measured FFMA2 peak; > 1 = above the CUDA-core roofline"}}
    out.update(ncu_traffic(key, N))
    return out


# ------------------------------------------------------------------------------------------ CPU legs (oracle port)
def _timed_cpu(fn, steps=1, warmup=0):
    for _ in range(warmup):
        fn()
    ts = []
    for _ in range(max(steps, 1)):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    return float(np.mean(ts))


def cpu_decode(models, base, n_utts, frames, steps=1, warmup=0, threads=None):
    """The float64 CPU restatement of the reference classifier (oracle/sr_oracle.c:orc_classify, OpenMP over utterances,
    linear-domain emissions like hmm_state.py:36-45) on a bounded sample.  -> (frames/s, cores, ms)"""
    from oracle import oracle as orc
    from srb200 import synth
    orc.set_num_threads(threads)
    X, off, _ = synth.make_utterances(2, base, n_utts, T=frames)
    X64, off64 = X.astype(np.float64), off.astype(np.int64)
    dt = _timed_cpu(lambda: orc.classify(X64, off64, models), steps, warmup)
    cores = orc.num_threads()
    orc.set_num_threads()
    return X.shape[0] / dt, cores, dt * 1e3


def cpu_em(models, base, n_utts, frames, steps=1, warmup=0, threads=None):
    """The CPU restatement of one training pass (oracle.train_iteration: align_gmm_states for every utterance, OpenMP over
    utterances, then one GMM.em iteration per tied state) on a bounded sample."""
    from oracle import oracle as orc
    from srb200 import synth
    orc.set_num_threads(threads)
    X, off, words = synth.make_utterances(2, base, n_utts, T=frames)
    X64, off64 = X.astype(np.float64), off.astype(np.int64)
    dt = _timed_cpu(lambda: orc.train_iteration(X64, off64, words, models, NCOMP), steps, warmup)
    cores = orc.num_threads()
    orc.set_num_threads()
    return X.shape[0] / dt, cores, dt * 1e3


def cpu_line(fn, models, base, n_utts, frames, what):
    """cpu_baseline object: all host threads on the sample, plus the 1-thread figure on a tenth of it."""
    v, cores, ms = fn(models, base, n_utts, frames)
    n1 = max(n_utts // 10, 64)
    v1, _, ms1 = fn(models, base, n1, frames, threads=1)
    return {"value": v, "unit": "frames/s", "cores": cores, "kind": "port",
            "sample": f"{n_utts} utterances x {frames} frames, {what} (oracle/ float64 C restatement, OpenMP), {ms:.0f} ms",
            "one_core": {"value": v1, "sample": f"{n1} utterances, 1 thread, {ms1:.0f} ms"},
            "note": "the reference itself is pure Python (about 3e2 frames/s/core on config 1, SURVEY section 6) and does "
                    "not travel to the GPU box; this C port is the stronger baseline"}


# ------------------------------------------------------------------------------------------ config 1: decode
def bench_decode(args, ctx, steps, warmup):
    """Config-1 shape: every utterance scored against all 11 word models + Viterbi (fused classify kernel)."""
    torch, dist, engine = ctx["torch"], ctx["dist"], ctx["engine"]
    X, off, words, pm, trans, world = ctx["X"], ctx["off"], ctx["words"], ctx["pm"], ctx["trans"], ctx["world"]
    N, B = int(X.shape[0]), len(off) - 1
    Xd, offd = engine.as_frames(X), engine.as_offsets(off)
    final = torch.empty((B, W_), dtype=torch.float32, device="cuda")
    label = torch.empty(B, dtype=torch.int32, device="cuda")
    tensor = pm.tensor_core_ok(W_, S_) and not args.ffma

    def step():
        engine.classify(Xd, offd, pm, trans, W_, S_, final, label, tensor=tensor)

    sampler = ClockSampler(ctx["local"]).start()
    total_ms, per = timed_steps(step, steps, warmup, torch, dist, world)
    clocks = sampler.stop()
    acc = float((label.cpu().numpy() == words).mean())
    kern_ms = float(np.mean(per))
    note = "17160 FMA/frame = 220 Gaussians x 2*39 (SURVEY 8d), FP32-equivalent; 156 B/frame of X read once"
    if tensor:
        # the contraction runs as 3 TF32 products per FP32 product on 128 x 224 x 80 tiles (N 220 -> 224, K 78 -> 80)
        roof = roof_tensor(FMA_PER_FRAME_DECODE, N, kern_ms, ctx["peak_fma"], "classify_tc_kernel<39>", note,
                           "classify_tc_kernel", 3 * 2.0 * 224 * 80)
    else:
        roof = roof_fma(FMA_PER_FRAME_DECODE, N, kern_ms, ctx["peak_fma"], "classify_kernel<39>", note, "classify_kernel")
    roof["hbm"] = {"achieved_gbs": (N * D_ * 4 + B * (W_ + 1) * 4) / (kern_ms * 1e-3) / 1e9,
                   "peak_gbs": peaks().get("hbm_gbs")}
    out = {"metric": "decoded frames/s (GMM scoring of all 11 word models + Viterbi)",
           "value": world * N * steps / (total_ms * 1e-3), "unit": "frames/s", "ms_per_step": total_ms / steps,
           "steps": steps, "warmup": warmup, "gpu_launches": steps, "roofline": roof, "accuracy": acc, "clocks": clocks,
           "config": {"workload": f"decode: {W_} words x {S_} states x {M_} diag Gaussians, D={D_}, {B} utterances x "
                                  f"{args.frames} frames per GPU, every utterance against every word"}}
    if not args.no_e2e:
        # pinned host buffers -> chunked H2D on two streams -> classify -> D2H labels
        Xh = torch.from_numpy(X).pin_memory()
        offh = torch.from_numpy(off.astype(np.int32))
        runner = engine.HostClassifier(pm, trans, W_, S_, max_utts=8192, max_frames=8192 * args.frames)
        out_labels = torch.empty(B, dtype=torch.int32).pin_memory()
        e2e_ms, _ = timed_steps(lambda: runner.run(Xh, offh, out_labels, sync=False), steps, warmup, torch, dist, world)
        assert np.array_equal(out_labels.numpy(), label.cpu().numpy()), "e2e labels differ from the resident run"
        out["e2e"] = {"value": world * N * steps / (e2e_ms * 1e-3), "unit": "frames/s",
                      "h2d_bytes_per_step": int(X.nbytes + off.nbytes), "d2h_bytes_per_step": int(B * 4),
                      "api": "srb200.engine.HostClassifier.run (host numpy/pinned buffers in, labels out)"}
        del Xh
    if ctx["rank"] == 0 and not args.no_cpu:
        out["cpu_baseline"] = cpu_line(cpu_decode, ctx["models"], ctx["base"], args.cpu_utts or min(B, 20_000),
                                       args.frames, "orc_classify")
    del Xd
    return out


# ------------------------------------------------------------------------------------------ config 2: em
def bench_em(args, ctx):
    """Config 2: one segmental-K-means + EM iteration over the rank's shard (align -> stats -> allreduce -> M-step)."""
    torch, dist, engine = ctx["torch"], ctx["dist"], ctx["engine"]
    from srb200.trainer import IsolatedTrainer
    X, off, words, world = ctx["X"], ctx["off"], ctx["words"], ctx["world"]
    N = int(X.shape[0])
    pm, _ = engine.PackedModel.from_words(ctx["models"])       # private copy: training moves it
    trans_np = np.stack([m[3] for m in ctx["models"]])
    tr = IsolatedTrainer(X, off, words, pm, trans_np, S_, NCOMP, fused=None if not args.unfused else False)
    sampler = ClockSampler(ctx["local"]).start()
    total_ms, per = timed_steps(tr.em_iteration, args.steps, args.warmup, torch, dist, world)
    clocks = sampler.stop()
    mean_cost = tr.mean_cost()
    pk = peaks()
    peak_fma = ctx["peak_fma"]
    Ntot = ctx["N_total"]
    if tr.fused:
        # phase breakdown (outside the timed region): CUDA events around the fused kernel call and the rest
        names = ["em_fused", "allreduce", "mstep"]
        acc = dict.fromkeys(names, 0.0)
        reps = 5
        for _ in range(reps):
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
            ev[0].record()
            tr.fused_pass(keep_slices=True)
            ev[1].record()
            tr.reduce()
            ev[2].record()
            tr.update()
            ev[3].record()
            torch.cuda.synchronize()
            for i, k in enumerate(names):
                acc[k] += ev[i].elapsed_time(ev[i + 1]) / reps
        # algorithmic work of the fused kernel per frame: scoring of the word's S*M Gaussians
        # (2*D FMA each) + accumulation of the ncomp trained components (2*D FMA each); 156 B of X read once
        fma_pf = S_ * M_ * 2 * D_ + NCOMP * 2 * D_
        kern_ms = acc["em_fused"]
        roof = roof_fma(fma_pf, N, kern_ms, peak_fma, "em_isolated_kernel<39,5,2>",
                        f"{fma_pf} FMA/frame = {S_ * M_} Gaussians x 2*39 scoring + {NCOMP} x 2*39 accumulation; "
                        f"{D_ * 4} B/frame of X read once", "em_isolated_kernel")
        gbs = (D_ * 4) * N / (kern_ms * 1e-3) / 1e9
        roof["hbm"] = {"achieved_gbs": gbs, "peak_gbs": pk.get("hbm_gbs"), "frac": gbs / pk.get("hbm_gbs")}
        roof["phase_ms"] = acc
        launches = tr.launches_per_iteration
    else:
        names = ["score", "viterbi", "stats", "allreduce", "mstep"]
        acc = dict.fromkeys(names, 0.0)
        reps = 3
        for _ in range(reps):
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
            ev[0].record()
            tr.score()
            ev[1].record()
            tr.plan.run(tr.E)
            ev[2].record()
            tr.accumulate()
            ev[3].record()
            tr.reduce()
            ev[4].record()
            tr.update()
            ev[5].record()
            torch.cuda.synchronize()
            for i, k in enumerate(names):
                acc[k] += ev[i].elapsed_time(ev[i + 1]) / reps
        bytes_pf = {"score": D_ * 4 + 4 * S_, "viterbi": 4 * S_ + 2 * S_ + 4, "stats": D_ * 4 + 4}
        dom = max(("score", "viterbi", "stats"), key=lambda k: acc[k])
        gbs = bytes_pf[dom] * N / (acc[dom] * 1e-3) / 1e9
        roof = {"bound": "hbm", "achieved": gbs, "peak": pk.get("hbm_gbs"), "unit": "GB/s",
                "frac": gbs / pk.get("hbm_gbs"), "traffic": None, "kernel": dom, "kernel_ms": acc[dom],
                "peak_source": "MEASURED_PEAKS.json hbm_gbs" if "source" not in pk else "fallback",
                "algorithmic": f"{bytes_pf[dom]} B/frame for the {dom} phase", "phase_ms": acc}
        launches = W_ + 6
    out = {
        "metric": "EM-iteration frames/s (segmental K-means alignment + GMM EM statistics + M-step)",
        "value": Ntot * args.steps / (total_ms * 1e-3), "unit": "frames/s", "ms_per_step": total_ms / args.steps,
        "gpu_launches": args.steps * launches, "roofline": roof,
        "mean_viterbi_cost": mean_cost, "clocks": clocks, "fused": bool(tr.fused),
    }
    if not args.no_e2e:
        # the shard comes from pinned host memory every step, the updated statistics / model go back
        Xh = torch.from_numpy(np.ascontiguousarray(tr.X.cpu().numpy())).pin_memory()
        host_model = torch.empty(tr.stats.numel(), dtype=torch.float64).pin_memory()

        def e2e_step():
            if tr.fused:
                tr.em_iteration_from_host(Xh)          # chunked H2D on a copy stream overlapped with the fused kernel
            else:
                tr.X.copy_(Xh, non_blocking=True)
                tr.em_iteration()
            host_model.copy_(tr.stats, non_blocking=True)

        e2e_ms, _ = timed_steps(e2e_step, args.steps, args.warmup, torch, dist, world)
        out["e2e"] = {"value": Ntot * args.steps / (e2e_ms * 1e-3), "unit": "frames/s",
                      "h2d_bytes_per_step": int(Xh.numel() * 4), "d2h_bytes_per_step": int(host_model.numel() * 8),
                      "api": "srb200.trainer.IsolatedTrainer.em_iteration_from_host (pinned host shard in, statistics / "
                             "model out)"}
        del Xh
    del tr
    return out


# ------------------------------------------------------------------------------------------ config 3: fb
def make_ragged_corpus_device(torch, base, B, lo, hi, seed, noise=0.7, chunk=1 << 22):
    """Config-3 corpus generated ON THE DEVICE (27 GB at 1 M utterances: a host generator would take minutes): word-sorted
    utterances of lo..hi frames walking left-to-right through their word's state means (synth.make_utterances's recipe).
    -> (X [N, D] float32 CUDA, off int32[B+1] numpy, words int64[B] numpy)"""
    rng = np.random.default_rng(seed)
    W, S, dim = base.shape
    words = np.sort(rng.integers(0, W, size=B))
    lens = rng.integers(lo, hi + 1, size=B)
    off = np.zeros(B + 1, np.int64)
    np.cumsum(lens, out=off[1:])
    N = int(off[-1])
    dev = torch.device("cuda")
    base_d = torch.from_numpy(base.astype(np.float32)).to(dev)
    X = torch.empty((N, dim), dtype=torch.float32, device=dev)
    lens_d, off_d = torch.from_numpy(lens).to(dev), torch.from_numpy(off).to(dev)
    words_d = torch.from_numpy(words).to(dev)
    gen = torch.Generator(device=dev).manual_seed(seed)
    u0 = 0
    while u0 < B:                                          # chunks of whole utterances of about `chunk` frames
        u1 = int(np.searchsorted(off, off[u0] + chunk, side="left"))
        u1 = min(max(u1, u0 + 1), B)
        f0, f1 = int(off[u0]), int(off[u1])
        u_of = torch.repeat_interleave(torch.arange(u0, u1, device=dev), lens_d[u0:u1])
        t_in = torch.arange(f0, f1, device=dev) - off_d[u_of]
        seg = torch.clamp((t_in * S) // lens_d[u_of], max=S - 1)
        X[f0:f1] = base_d[words_d[u_of], seg]
        X[f0:f1].add_(torch.randn((f1 - f0, dim), dtype=torch.float32, device=dev, generator=gen), alpha=noise)
        u0 = u1
    return X, off.astype(np.int32), words.astype(np.int64)


def cpu_fb(models, base, n_utts, S, M, ncomp):
    """CPU restatement of one Baum-Welch E-step + soft statistics on a bounded sample: emissions (orc_gmm_evaluate, OpenMP
    over frames), forward-backward per utterance (orc_forward_backward), gamma-weighted moments (numpy)."""
    from oracle import oracle as orc
    from srb200 import synth
    orc.set_num_threads()
    X, off, words = synth.make_utterances(3, base, n_utts, ragged=(50, 300))
    X64 = X.astype(np.float64)

    def run():
        G = len(models) * S
        S0, S1, S2 = np.zeros((G, ncomp)), np.zeros((G, ncomp, D_)), np.zeros((G, ncomp, D_))
        for u in range(len(words)):
            m = models[words[u]]
            x = X64[off[u]:off[u + 1]]
            comps = np.stack([orc.gmm_evaluate(x, m[0][s], m[1][s], m[2][s], log_domain=True)[1] for s in range(S)], axis=1)
            mx = comps.max(axis=2, keepdims=True)
            E = -(mx[:, :, 0] + np.log(np.exp(comps - mx).sum(axis=2)))
            _, _, _, gam = orc.forward_backward(E, m[3])
            lc = comps[:, :, :ncomp]
            p = np.exp(lc - lc.max(axis=2, keepdims=True))
            p = p / p.sum(axis=2, keepdims=True) * gam[:, :, None]            # [T, S, ncomp]
            g0 = words[u] * S
            S0[g0:g0 + S] += p.sum(axis=0)
            S1[g0:g0 + S] += np.einsum("tsm,td->smd", p, x)
            S2[g0:g0 + S] += np.einsum("tsm,td->smd", p, x * x)
        return S0

    dt = _timed_cpu(run)
    return X.shape[0] / dt, orc.num_threads(), dt * 1e3, X.shape[0]


def bench_fb(args, ctx, steps, warmup):
    """Config 3: one Baum-Welch iteration, 8 states x 16 Gaussians per word (int(log 16) = 2 splits -> 4 trained
    components, hmm.py:104), 1 M ragged 50-300 frame utterances: per-word K1 scoring + K4 forward-backward + soft-count
    statistics + allreduce + M-step + transition re-estimation.  No reference counterpart (the oracle is pinned by path
    enumeration and the (min,+) identity with decode_hmm_states)."""
    torch, dist, engine = ctx["torch"], ctx["dist"], ctx["engine"]
    from srb200 import synth
    from srb200.trainer import BaumWelchTrainer
    S, M = 8, 16
    ncomp = 2 ** int(np.log(M))
    models, base = synth.make_models(11, W=W_, S=S, M=M)
    B = args.fb_utts
    t0 = time.time()
    X, off, words = make_ragged_corpus_device(torch, base, B, 50, 300, 400 + ctx["rank"])
    torch.cuda.synchronize()
    log(f"fb: corpus of {B} utterances / {X.shape[0]} frames ({X.numel() * 4 / 1e9:.1f} GB) generated on the device in "
        f"{time.time() - t0:.1f} s")
    pm, _ = engine.PackedModel.from_words(models)
    tr = BaumWelchTrainer(X, off, words, pm, np.stack([m[3] for m in models]), S, ncomp)
    N = tr.N
    gb_in = X.numel() * 4 / 1e9
    del X
    sampler = ClockSampler(ctx["local"]).start()
    total_ms, per = timed_steps(tr.em_iteration, steps, warmup, torch, dist, ctx["world"])
    clocks = sampler.stop()
    names = ["score+forward_backward", "soft_stats", "allreduce", "mstep+transitions"]
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
    ev[0].record(); tr.e_step(); ev[1].record(); tr.accumulate(); ev[2].record(); tr.reduce(); ev[3].record()  # noqa: E702
    tr.update(); ev[4].record()  # noqa: E702
    torch.cuda.synchronize()
    phase = {k: ev[i].elapsed_time(ev[i + 1]) for i, k in enumerate(names)}
    ms = float(np.mean(per))
    # dominant kernel: the soft-count statistics (W launches); its algorithmic work is the posterior of the ncomp trained
    # components of every state above gmin (2*D FMA to score + 2*D to accumulate each), 156 B of X + 32 B of gamma / frame
    fma_soft = S * ncomp * 2 * D_ * 2
    roof = roof_fma(fma_soft, N, phase["soft_stats"], ctx["peak_fma"], f"gmm_stats_soft_kernel<{ncomp}> x{W_}",
                    f"{fma_soft} FMA/frame = {S} states x {ncomp} trained components x (2*39 scoring + 2*39 accumulation), "
                    f"an upper bound: states with gamma <= gmin are skipped; {D_ * 4 + S * 4} B/frame", "gmm_stats_soft_kernel")
    roof["phase_ms"] = phase
    fma_all = S * M * 2 * D_ + fma_soft
    roof["whole_iteration"] = {"fma_per_frame": fma_all,
                               "frac_of_fma_peak": 2.0 * fma_all * N / (ms * 1e-3) / (2 * ctx["peak_fma"])}
    out = {"metric": "Baum-Welch EM-iteration frames/s (8 states x 16 Gaussians, ragged 50-300 frames)",
           "value": ctx["world"] * N * steps / (total_ms * 1e-3), "unit": "frames/s", "ms_per_step": total_ms / steps,
           "steps": steps, "warmup": warmup, "gpu_launches": (4 * W_ + 8) * steps, "roofline": roof,
           "mean_neg_log_z_per_utt": tr.mean_neg_log_z(), "clocks": clocks,
           "config": {"workload": f"fb: {W_} words x {S} states x {M} Gaussians ({ncomp} trained), {B} utterances of 50-300 "
                                  f"frames per GPU ({N} frames, {gb_in:.1f} GB), ragged CSR batching"}}
    if not args.no_e2e:
        try:
            t0 = time.time()
            Xh = torch.empty((N, D_), dtype=torch.float32, pin_memory=True)
            Xh.copy_(tr.X)
            host_model = torch.empty(tr.stats.numel(), dtype=torch.float64).pin_memory()
            log(f"fb: {Xh.numel() * 4 / 1e9:.1f} GB pinned host copy of the corpus in {time.time() - t0:.1f} s")

            def e2e_step():
                tr.em_iteration_from_host(Xh)
                host_model.copy_(tr.stats, non_blocking=True)

            e_steps = max(1, min(steps, 3))
            e2e_ms, _ = timed_steps(e2e_step, e_steps, 1, torch, dist, ctx["world"])
            out["e2e"] = {"value": ctx["world"] * N * e_steps / (e2e_ms * 1e-3), "unit": "frames/s",
                          "h2d_bytes_per_step": int(Xh.numel() * 4), "d2h_bytes_per_step": int(host_model.numel() * 8),
                          "steps": e_steps,
                          "api": "srb200.trainer.BaumWelchTrainer.em_iteration_from_host (pinned host corpus in, word block "
                                 "by word block on a copy stream; statistics / model out)"}
            del Xh
        except Exception as e:  # noqa: BLE001
            out["e2e"] = {"value": None, "unavailable": f"{type(e).__name__}: {e}"}
    if ctx["rank"] == 0 and not args.no_cpu:
        n_cpu = args.cpu_utts or 1500
        v, cores, cms, nfr = cpu_fb(models, base, n_cpu, S, M, ncomp)
        out["cpu_baseline"] = {"value": v, "unit": "frames/s", "cores": cores, "kind": "port",
                               "sample": f"{n_cpu} utterances of 50-300 frames ({nfr} frames): emissions orc_gmm_evaluate "
                                         f"(OpenMP, {cores} threads) + orc_forward_backward and gamma-weighted moments "
                                         f"(1 thread), {cms:.0f} ms"}
    del tr
    torch.cuda.empty_cache()
    return out


# ------------------------------------------------------------------------------------------ config 2 in full: HMM.fit
def cpu_fit(base, n_per_word, frames):
    """oracle.hmm_fit (skmeans -> split K-means -> EM -> re-alignment, the reference's HMM.fit word by word) on a bounded
    sample: n_per_word utterances of each of the 11 words.  One thread (the per-state loops are numpy)."""
    from oracle import oracle as orc
    from srb200 import synth
    X, off, words = synth.make_utterances(2, base, n_per_word * W_, T=frames)
    X64 = X.astype(np.float64)
    ys = [[X64[off[u]:off[u + 1]] for u in np.flatnonzero(words == w)] for w in range(W_)]

    def run():
        for w in range(W_):
            np.random.seed(1000 + w)
            orc.hmm_fit(ys[w], S_, M_, use_em=True)

    dt = _timed_cpu(run)
    return X.shape[0] / dt, dt * 1e3, int(X.shape[0])


def bench_fit(args, ctx, steps, warmup):
    """BASELINE config 2 in full: `HMM(5).fit(ys, 4)` for the 11 words of the 100 k-utterance corpus (hmm.py:57-124:
    skmeans to convergence, binary-split K-means, per-state EM to convergence, Viterbi re-alignment), every phase batched
    over all 55 states (srb200.fit.WordFitter) and, at N > 1, the corpus split over the ranks with one fp64 allreduce of
    the statistics per iteration (strong scaling).  A step is one complete fit from the raw corpus."""
    torch, dist, engine = ctx["torch"], ctx["dist"], ctx["engine"]
    from srb200 import synth
    from srb200.fit import WordFitter
    from srb200.trainer import shard_range, sort_by_word
    rank, world = ctx["rank"], ctx["world"]
    X, off, words = synth.make_utterances(100, ctx["base"], args.utts, T=args.frames)
    lo, hi = shard_range(args.utts, rank, world)
    X, words = X[int(off[lo]):int(off[hi])], words[lo:hi]
    off = (off[lo:hi + 1] - off[lo]).astype(np.int64)
    Xs, offs, wordss, _ = sort_by_word(X, off, words)
    Xd = engine.as_frames(Xs)
    N_total = args.utts * args.frames
    seeds = [1000 + w for w in range(W_)]
    last = {}

    def step():
        ft = WordFitter(Xd, offs, wordss, W_, S_, M_, use_em=True, seeds=seeds,
                        verbose=bool(os.environ.get("SRB_FIT_VERBOSE")) and rank == 0)
        ft.fit()
        last["ft"] = ft

    steps, warmup = max(1, min(steps, 2)), 1
    sampler = ClockSampler(ctx["local"]).start()
    total_ms, per = timed_steps(step, steps, warmup, torch, dist, world)
    clocks = sampler.stop()
    ft = last["ft"]
    its = {"skmeans": int(ft.iterations["skmeans"].max()) + 1,
           "kmeans": [int(a.max()) + 1 for a in ft.iterations["kmeans"]],
           "em": [int(a.max()) + 1 for a in ft.iterations["em"]]}
    # phase split of one more fit, wall clock around synchronised phases
    ph = {}
    ft = WordFitter(Xd, offs, wordss, W_, S_, M_, use_em=True, seeds=seeds)
    for name, fn in (("skmeans", ft.skmeans), ("kmeans+em", ft.fit_states), ("align", ft.align)):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        ph[name] = (time.perf_counter() - t0) * 1e3
    # dominant kernel: srb_pool_stats, one launch per K-means / EM iteration over the rank's whole shard (X read once);
    # timed alone here with CUDA events on the launching stream, the call = pool-mask memset + kernel + slice reduction
    state = ft._labels_from_starts(ft.seg_starts)
    lc = engine.gauss_consts(ft.pm.var, ft.pm.w)
    n_loc = int(Xd.shape[0])
    calls = 20
    def call_ms(f64):
        engine.pool_em_stats(Xd, state, ft.pm, 2, lc=lc, f64=f64)
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        for _ in range(calls):
            engine.pool_em_stats(Xd, state, ft.pm, 2, lc=lc, f64=f64)
        ev1.record()
        torch.cuda.synchronize()
        return ev0.elapsed_time(ev1) / calls

    k_ms, k32_ms = call_ms(True), call_ms(False)
    pk = peaks()
    bytes_pf = 4 * D_ + 4
    gbs = n_loc * bytes_pf / (k_ms * 1e-3) / 1e9
    fma_pf = 2 * (2 * D_ + 2 * D_)                                   # 2 components x (scoring + moments), 2 FMA per dim each
    roof = {"bound": "hbm", "achieved": gbs, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": gbs / pk["hbm_gbs"],
            "kernel": "pool_stats64_kernel<2, 0> (srb_pool_stats_f64: pool-mask memset + kernel + slice reduction in one call)",
            "kernel_ms": k_ms,
            "peak_source": "MEASURED_PEAKS.json hbm_gbs" if "source" not in pk else "fallback 6650 GB/s",
            "algorithmic": f"{bytes_pf} B/frame = {D_} float32 features + one int32 label, read once per iteration; "
                           f"{fma_pf} FMA/frame",
            "fp64_pipe": {"ops_per_frame": 7 * D_ * 2, "frac": 7 * D_ * 2 * n_loc / (k_ms * 1e-3) / (148 * 64 * 1.965e9),
                          "note": "3 (scoring) + 4 (moments) double operations per dimension and component over the nominal "
                                  "148 SM x 64 FP64 lanes x 1.965 GHz; no measured FP64 peak in MEASURED_PEAKS.json"},
            "f32_kernel": {"kernel": "pool_stats_kernel<2, 0> (srb_pool_stats, precision='f32')", "kernel_ms": k32_ms,
                           "achieved": n_loc * bytes_pf / (k32_ms * 1e-3) / 1e9,
                           "frac": n_loc * bytes_pf / (k32_ms * 1e-3) / 1e9 / pk["hbm_gbs"],
                           "fp32_fma_frac": fma_pf * n_loc / (k32_ms * 1e-3) / ctx["peak_fma"]},
            "share_of_step": f"{sum(its['kmeans']) + sum(its['em'])} launches x {k_ms:.3f} ms of the {total_ms / steps:.0f} ms "
                             "step (the working set shrinks as states converge)"}
    roof.update(ncu_traffic("pool_stats64_kernel", n_loc))
    # the FP32 variant of the whole fit (precision="f32": faster passes, K-means may settle in a neighbouring fixed point)
    t32 = {}
    if not args.no_sub:
        def step32():
            WordFitter(Xd, offs, wordss, W_, S_, M_, use_em=True, seeds=seeds, precision="f32").fit()
        ms32, _ = timed_steps(step32, 1, 1, torch, dist, world)
        t32 = {"value": N_total / (ms32 * 1e-3), "unit": "frames/s", "ms_per_step": ms32, "steps": 1}
    out = {"metric": "full-training frames/s (HMM.fit of all 11 words: skmeans + split K-means + EM + re-alignment)",
           "value": N_total * steps / (total_ms * 1e-3), "unit": "frames/s", "ms_per_step": total_ms / steps,
           "steps": steps, "warmup": warmup, "scaling": "strong", "iterations_to_convergence": its, "phase_ms": ph,
           "clocks": clocks, "roofline": roof, "precision": "f64 passes (default)", "f32_variant": t32,
           "gpu_launches": (3 * its["skmeans"] + 4 * (sum(its["kmeans"]) + sum(its["em"])) + 2) * steps,
           "config": {"workload": f"fit: {W_} words x {S_} states x {M_} Gaussians, D={D_}, {args.utts} utterances x "
                                  f"{args.frames} frames in all, split over {world} GPU(s); data-dependent iteration "
                                  "counts (np.allclose stops of the reference), listed in iterations_to_convergence",
                      "l2": "corpus larger than L2"}}
    if not args.no_e2e:
        # the drop-in call: sr.recognition.fit_words on HOST lists of utterances (concatenation, H2D, training, models out)
        import sr.recognition as rec
        ys = [[X[off[u]:off[u + 1]] for u in np.flatnonzero(words == w)] for w in range(W_)]
        t0 = time.perf_counter()
        if world > 1:
            dist.barrier()
        models = rec.fit_words(ys, S_, M_, use_em=True, seeds=seeds)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        dt = time.perf_counter() - t0
        out["e2e"] = {"value": N_total / dt, "unit": "frames/s", "h2d_bytes_per_step": int(X.nbytes),
                      "d2h_bytes_per_step": int(sum((2 * D_ + 1) * M_ * 8 for m in models for _ in m.gmm_states)),
                      "steps": 1, "seconds": dt,
                      "api": "sr.recognition.fit_words(ys_per_word, 5, 4): host lists of [T, D] utterances in, trained HMM "
                             "objects (GMM states, transitions, segments) out; wall clock, barrier on both sides"}
    if rank == 0 and not args.no_cpu:
        npw = args.cpu_utts or 40
        v, cms, nfr = cpu_fit(ctx["base"], npw, args.frames)
        out["cpu_baseline"] = {"value": v, "unit": "frames/s", "cores": 1, "kind": "port",
                               "sample": f"{npw} utterances of each word ({nfr} frames), oracle.hmm_fit word by word "
                                         f"(numpy float64 restatement of hmm.py:57-124), {cms:.0f} ms; iteration counts "
                                         "differ with the corpus size"}
    return out


# ------------------------------------------------------------------------------------------ config 4: connected
def cpu_connected(models, base, n_utts, L, T):
    """orc_decode on the 393-state word-loop graph (decode.py:80-146: T*R*R candidate scans per utterance), one utterance
    per host thread (ctypes releases the GIL)."""
    from concurrent.futures import ThreadPoolExecutor
    from oracle import oracle as orc
    from srb200 import synth
    X, off, _ = synth.make_strings(5, base, n_utts, L, T)
    emit_, trans, ends = orc.build_state_sequences(S_, [m[3] for m in models], [list(range(W_))] * L)
    orc.set_num_threads(1)
    cores = os.cpu_count() or 1

    def one(u):
        x = X[off[u]:off[u + 1]].astype(np.float64)
        E = orc.emissions(x, models, emit_, log_domain=True)
        return orc.decode_hmm_states(E, trans, emit_ < 0, [[e, -1] for e in ends])[2]

    def run():
        with ThreadPoolExecutor(cores) as ex:
            list(ex.map(one, range(n_utts)))

    dt = _timed_cpu(run)
    orc.set_num_threads()
    return X.shape[0] / dt, min(cores, n_utts), dt * 1e3


def bench_connected(args, ctx, steps, warmup):
    """Config 4: connected-digit decoding, 7 layers x 11 words (393 states, 847 arcs), 500-frame utterances: K1 scoring of
    the 55 tied states + graph Viterbi with back-pointers and traceback + word strings (main.py:35-67)."""
    torch, dist, engine = ctx["torch"], ctx["dist"], ctx["engine"]
    from srb200 import synth
    from srb200.decoder import HostStringDecoder, StringDecoder
    B, T, L = 4096, 500, 7
    X, off, strings = synth.make_strings(300 + ctx["rank"], ctx["base"], B, L, T)
    pm = ctx["pm"]
    wt = np.stack([m[3] for m in ctx["models"]])
    dec = StringDecoder(pm, wt, L, off, tensor=None if not args.ffma else False)
    Xd = engine.as_frames(X)
    N = Xd.shape[0]
    sampler = ClockSampler(ctx["local"]).start()
    total_ms, per = timed_steps(lambda: dec.run(Xd), steps, warmup, torch, dist, ctx["world"])
    clocks = sampler.stop()
    words, nwords = dec.words.cpu().numpy(), dec.nwords.cpu().numpy()
    string_acc = float(np.mean((nwords == L) & (words == strings).all(axis=1)))
    full = nwords == L
    digit_acc = float(np.mean(words[full] == strings[full])) if full.any() else 0.0
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    ev[0].record(); dec.score(Xd); ev[1].record(); dec.decode(); ev[2].record()  # noqa: E702
    torch.cuda.synchronize()
    phase = {"score": ev[0].elapsed_time(ev[1]), "viterbi+words": ev[1].elapsed_time(ev[2])}
    ms = float(np.mean(per))
    pk = peaks()
    # dominant kernel: the one-graph banded Viterbi -- integer / compare work bounded by its HBM stream:
    # 393 B of back-pointers written + 220 B of cost row read per frame
    bytes_pf = 393 + 55 * 4
    gbs = bytes_pf * N / (phase["viterbi+words"] * 1e-3) / 1e9
    roof = {"bound": "hbm", "achieved": gbs, "peak": pk.get("hbm_gbs"), "unit": "GB/s", "frac": gbs / pk.get("hbm_gbs"),
            "kernel": "viterbi_banded1_kernel (one-graph banded form of srb_viterbi_banded) + path_words_kernel",
            "kernel_ms": phase["viterbi+words"],
            "peak_source": "MEASURED_PEAKS.json hbm_gbs" if "source" not in pk else "fallback",
            "algorithmic": f"{bytes_pf} B/frame: 393 back-pointer bytes written + 55 x 4 B of tied-state costs read; the "
                           "recursion itself is 847 arcs + 393 states of add / compare per frame (issue-bound, DESIGN 3)",
            "phase_ms": phase,
            "whole_step_fp32_fma": {"fma_per_frame": FMA_PER_FRAME_DECODE,
                                    "frac": 2.0 * FMA_PER_FRAME_DECODE * N / (ms * 1e-3) / (2 * ctx["peak_fma"])}}
    roof.update(ncu_traffic("viterbi_banded1_kernel", N))
    out = {"metric": "decoded frames/s (connected-digit grammar, GMM scoring + graph Viterbi with back-pointers)",
           "value": ctx["world"] * N * steps / (total_ms * 1e-3), "unit": "frames/s", "ms_per_step": total_ms / steps,
           "steps": steps, "warmup": warmup, "gpu_launches": 3 * steps, "roofline": roof, "string_accuracy": string_acc,
           "digit_accuracy": digit_acc, "clocks": clocks,
           "config": {"workload": f"connected: {L} layers x {W_} words (393 states, 847 arcs), {B} utterances x {T} frames "
                                  f"per GPU ({X.nbytes / 1e9:.2f} GB)"}}
    if not args.no_e2e:
        Xh = torch.from_numpy(X).pin_memory()
        hd = HostStringDecoder(pm, wt, L, off, chunks=4)
        wo = torch.empty((B, L), dtype=torch.int32).pin_memory()
        no = torch.empty(B, dtype=torch.int32).pin_memory()
        e2e_ms, _ = timed_steps(lambda: hd.run(Xh, wo, no), steps, warmup, torch, dist, ctx["world"])
        assert np.array_equal(no.numpy(), nwords) and np.array_equal(wo.numpy()[full], words[full]), "e2e strings differ"
        out["e2e"] = {"value": ctx["world"] * N * steps / (e2e_ms * 1e-3), "unit": "frames/s",
                      "h2d_bytes_per_step": int(X.nbytes), "d2h_bytes_per_step": int(B * (L + 1) * 4),
                      "api": "srb200.decoder.HostStringDecoder.run (pinned host frames in, digit strings out)"}
        del Xh, hd
    if ctx["rank"] == 0 and not args.no_cpu:
        n_cpu = args.cpu_utts or 2 * (os.cpu_count() or 8)
        v, cores, cms = cpu_connected(ctx["models"], ctx["base"], n_cpu, L, T)
        out["cpu_baseline"] = {"value": v, "unit": "frames/s", "cores": cores, "kind": "port",
                               "sample": f"{n_cpu} utterances x {T} frames, orc_decode on the 393-state graph, one utterance "
                                         f"per host thread, {cms:.0f} ms"}
    del Xd, dec
    return out


# ------------------------------------------------------------------------------------------ config 5: stress
def bench_stress(args, ctx, steps, warmup):
    """Config 5: likelihood-only sweep, 2000 tied states x 32 Gaussians x 39 dims on 10 M frames; per-frame best state
    streamed out (the 8 kB/frame cost matrix is never materialised)."""
    torch, dist, engine = ctx["torch"], ctx["dist"], ctx["engine"]
    R, M = 2000, 32
    rng = np.random.default_rng(7)
    mu = rng.normal(size=(R, 1, D_)) + 0.5 * rng.normal(size=(R, M, D_))
    var = rng.uniform(0.5, 1.5, size=(R, M, D_))
    w = rng.dirichlet(np.ones(M), size=R)
    pm = engine.PackedModel(mu, var, w)
    N = 10_000_000 if args.utts == 100_000 else args.utts * args.frames
    Xd = torch.randn((N, D_), dtype=torch.float32, device="cuda", generator=torch.Generator("cuda").manual_seed(9))
    bc = torch.empty(N, dtype=torch.float32, device="cuda")
    bs = torch.empty(N, dtype=torch.int32, device="cuda")
    L_ = ctx["lib"]
    use_tc = not args.ffma

    def step():
        if use_tc:
            engine.gmm_best_gemm(Xd, pm, bc, bs)
        else:
            L_.check(L_.load().srb_gmm_score(L_.ptr(Xd), N, D_, L_.ptr(pm.ab), L_.ptr(pm.c), L_.ptr(pm.shift), R, pm.Mp,
                                             None, R, L_.ptr(bc), L_.ptr(bs), None, L_.stream_ptr()))

    sampler = ClockSampler(ctx["local"]).start()
    total_ms, per = timed_steps(step, steps, warmup, torch, dist, ctx["world"])
    clocks = sampler.stop()
    ms = float(np.mean(per))
    fma_pf = R * M * 2 * D_
    note = "4.992 M FMA/frame = 64000 Gaussians x 2*39 (SURVEY 8d), FP32-equivalent; 156 B/frame of X read once"
    if use_tc:
        # executed: 3 TF32 products per FP32 product on 128 x 160 x 80 tiles (K 78 + constant -> 80), 400 blocks per tile
        roof = roof_tensor(fma_pf, N, ms, ctx["peak_fma"],
                           "gmm_best_gemm_kernel<39> (persistent tcgen05 3xTF32 GEMM, model streamed by TMA bulk copies)",
                           note, "gmm_best_gemm_kernel", 3 * 2.0 * R * M * 80)
    else:
        roof = roof_fma(fma_pf, N, ms, ctx["peak_fma"], "score_kernel<39>", note)
    out = {"metric": "GMM-scored frames/s (2000 states x 32 Gaussians, best state per frame)",
           "value": ctx["world"] * N * steps / (total_ms * 1e-3), "unit": "frames/s", "ms_per_step": total_ms / steps,
           "steps": steps, "warmup": warmup, "gpu_launches": steps, "roofline": roof, "clocks": clocks,
           "config": {"workload": f"stress: {R} states x {M} Gaussians, D={D_}, {N} frames per GPU ({N * D_ * 4 / 1e9:.2f} GB)"}}
    state_res = bs.cpu().numpy()
    if not args.no_e2e and use_tc:
        Xh = torch.empty((N, D_), dtype=torch.float32, pin_memory=True)
        Xh.copy_(Xd)
        co = torch.empty(N, dtype=torch.float32).pin_memory()
        so = torch.empty(N, dtype=torch.int32).pin_memory()
        runner = engine.HostBestState(pm, chunk_frames=1 << 20)
        e_steps = max(1, min(steps, 3))
        e2e_ms, _ = timed_steps(lambda: runner.run(Xh, co, so), e_steps, 1, torch, dist, ctx["world"])
        assert np.array_equal(so.numpy(), state_res), "e2e best states differ from the resident run"
        out["e2e"] = {"value": ctx["world"] * N * e_steps / (e2e_ms * 1e-3), "unit": "frames/s", "steps": e_steps,
                      "h2d_bytes_per_step": int(N * D_ * 4), "d2h_bytes_per_step": int(N * 8),
                      "api": "srb200.engine.HostBestState.run (pinned host frames in, best state / cost per frame out)"}
        del Xh, runner
    if ctx["rank"] == 0 and not args.no_cpu:
        from oracle import oracle as orc
        orc.set_num_threads()
        n_cpu = 2048
        idx = np.sort(np.random.default_rng(1).choice(N, n_cpu, replace=False))
        Xs = Xd[torch.from_numpy(idx).cuda()].cpu().numpy().astype(np.float64)
        cost = np.empty((n_cpu, R))

        def run():
            for r in range(R):
                cost[:, r] = orc.gmm_evaluate(Xs, mu[r], var[r], w[r], log_domain=True)[0]

        dt = _timed_cpu(run)
        # the sample doubles as a full-size parity check: the states the GPU named are the oracle's best
        so_ = np.sort(cost, axis=1)
        decided = (so_[:, 1] - so_[:, 0]) > 1e-4 * np.abs(so_[:, 0])
        agree = float(np.mean(cost.argmin(axis=1)[decided] == state_res[idx][decided]))
        out["cpu_baseline"] = {"value": n_cpu / dt, "unit": "frames/s", "cores": orc.num_threads(), "kind": "port",
                               "sample": f"{n_cpu} frames of the same corpus x 2000 states, orc_gmm_evaluate (float64, "
                                         f"OpenMP over frames), {dt * 1e3:.0f} ms",
                               "best_state_agreement_outside_margin": agree}
        assert agree == 1.0, "GPU best states differ from the oracle outside the 1e-4 margin"
    del Xd
    return out


# ------------------------------------------------------------------------------------------ front-end (f3)
def bench_frontend(args, ctx, steps, warmup):
    """SURVEY 8 f3 ("next" row): MFCC front-end, 1 s utterances at 16 kHz -> standardised [100, 39] features
    (srb_mfcc + srb_delta_standardize).  value: samples resident in HBM; e2e: host int16 arrays in, features out."""
    torch, dist, world = ctx["torch"], ctx["dist"], ctx["world"]
    from srb200.frontend import MfccFrontEnd
    B = min(args.utts, 5000)
    rng = np.random.default_rng(7 + ctx["rank"])
    t = np.arange(16000) / 16000.0
    sigs = [(8000 * np.sin(2 * np.pi * rng.uniform(100, 400) * t * (1 + 0.3 * t)) + 400 * rng.normal(size=16000)).astype(np.int16)
            for _ in range(B)]
    fe = MfccFrontEnd()
    st = fe.stage(sigs)
    out = fe.run(st)
    sampler = ClockSampler(ctx["local"]).start()
    total_ms, per = timed_steps(lambda: fe.run(st, out=out), steps, warmup, torch, dist, world)
    clocks = sampler.stop()
    F = st["F"]
    kern_ms = float(np.mean(per))
    # algorithmic bytes per frame: 160 new samples in, 13 cepstra out and back in, 39 features out, standardisation
    # re-reads them twice and rewrites once (L2-resident per utterance)
    bytes_pf = 160 * 4 + 13 * 4 * 2 + 39 * 4
    gbs = bytes_pf * F / (kern_ms * 1e-3) / 1e9
    pk = peaks()
    e_steps = max(1, steps // 3)
    e2e_ms, _ = timed_steps(lambda: fe.features(sigs)["features"].cpu(), e_steps, 1, torch, dist, world)
    from oracle import feature_oracle as fo
    t0 = time.perf_counter()
    n_cpu = 200
    for s in sigs[:n_cpu]:
        fo.load_features(s)
    cpu_s = time.perf_counter() - t0
    return {"metric": "MFCC feature frames/s (pre-emphasis, FFT-512, 40 mel filters, DCT, deltas, standardisation)",
            "value": world * F * steps / (total_ms * 1e-3), "unit": "frames/s", "ms_per_step": total_ms / steps,
            "steps": steps, "warmup": warmup, "gpu_launches": 2 * steps,
            "roofline": {"bound": "hbm", "achieved": gbs, "peak": pk.get("hbm_gbs"), "unit": "GB/s",
                         "frac": gbs / pk.get("hbm_gbs"), "traffic": None, "kernel": "mfcc_kernel", "kernel_ms": kern_ms,
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs" if "source" not in pk else "fallback",
                         "algorithmic": f"{bytes_pf} B/frame; the kernel is bound by the shared-memory FFT (one warp "
                                        f"per frame, 9 radix-2 stages), not by HBM"},
            "clocks": clocks,
            "e2e": {"value": world * F * e_steps / (e2e_ms * 1e-3), "unit": "frames/s",
                    "h2d_bytes_per_step": int(st["sig"].numel() * 2), "d2h_bytes_per_step": int(F * 39 * 4),
                    "api": "srb200.frontend.MfccFrontEnd.features (list of host int16 sample arrays in, features out; "
                           "includes the host-side concatenation of the arrays)"},
            "cpu_baseline": {"value": n_cpu * 100 / cpu_s, "unit": "frames/s", "cores": 1, "kind": "port",
                             "sample": f"{n_cpu} utterances of 1 s, oracle/feature_oracle.py (numpy float64), "
                                       f"{cpu_s * 1e3:.0f} ms"},
            "config": {"workload": f"frontend: {B} utterances x 16000 samples (1 s at 16 kHz) per GPU -> 100 frames x 39"}}


# ------------------------------------------------------------------------------------------ driver
def run_ours(args):
    import torch
    import torch.distributed as dist
    rank, local, world = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU path for the product arm")
    torch.cuda.set_device(local)
    from srb200 import numa
    cpus = numa.bind_to_gpu(local) if world > 1 else None      # pinned buffers then live on the GPU's NUMA node
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from srb200 import engine, synth
    from srb200 import lib as lib_
    from srb200.trainer import shard_range, sort_by_word

    models, base = synth.make_models(1, W=W_, S=S_, M=M_)
    pm, trans = engine.PackedModel.from_words(models)
    peak_fma = max(engine.fma_peak(True, 8192) for _ in range(3))      # FMA/s, measured now on this GPU
    ctx = {"torch": torch, "dist": dist, "engine": engine, "pm": pm, "trans": trans, "models": models, "base": base,
           "world": world, "local": local, "rank": rank, "peak_fma": peak_fma, "lib": lib_}
    common = {"n_gpus": world, "higher_is_better": True, "vs_baseline": None, "dtype": "f32", "data": "synthetic"}
    if world > 1:
        common["host_binding"] = (f"rank 0 bound to {len(cpus)} CPUs local to its GPU (NVML cpu affinity); every rank "
                                  "binds likewise before it allocates pinned buffers") if cpus else "unbound"
    sub_fns = {"connected": bench_connected, "stress": bench_stress, "fb": bench_fb, "frontend": bench_frontend,
               "fit": bench_fit}
    if args.workload in sub_fns:
        line = sub_fns[args.workload](args, ctx, args.steps, max(args.warmup, 3))
        line.update(common)
        line.setdefault("scaling", "weak")
        line["config"].update({"l2": "inputs larger than L2",
                               "parallelism": f"utterance-sharded x{world}, " + (
                                   "one fp64 allreduce of the statistics per iteration" if args.workload in ("fb", "fit")
                                   else "no collective")})
        if rank == 0:
            emit(line)
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- configs 1 / 2: 100-frame labelled utterances.  Strong scaling: the SAME 100 k-utterance corpus whatever N,
    # rank r keeps the contiguous range shard_range(B, r, N) of it; weak scaling: every rank draws its own corpus.
    em = args.workload == "em"
    strong = em and args.scaling == "strong"
    if strong:
        X, off, words = synth.make_utterances(100, base, args.utts, T=args.frames)
        lo, hi = shard_range(args.utts, rank, world)
        X, words = X[int(off[lo]):int(off[hi])], words[lo:hi]
        off = (off[lo:hi + 1] - off[lo]).astype(np.int32)
        N_total = args.utts * args.frames
    else:
        X, off, words = synth.make_utterances(100 + rank, base, args.utts, T=args.frames)
        N_total = world * args.utts * args.frames
    X, off, words, _ = sort_by_word(X, off, words)
    N, B = int(X.shape[0]), len(off) - 1
    ctx.update({"X": X, "off": off, "words": words, "N_total": N_total})
    if em:
        line = bench_em(args, ctx)
    else:
        line = bench_decode(args, ctx, args.steps, max(args.warmup, 3))
    line.update(common)
    line.update({"steps": args.steps, "warmup": args.warmup, "scaling": "strong" if strong else "weak"})
    line["config"] = {
        "workload": f"{args.workload}: {W_} words x {S_} states x {M_} diag Gaussians, D={D_}, "
                    + (f"{args.utts} utterances x {args.frames} frames in all, split over {world} GPU(s)" if strong else
                       f"{B} utterances x {args.frames} frames per GPU")
                    + (f", EM over the first {NCOMP} components (hmm.py:104)" if em else ""),
        "utterances_per_gpu": B, "frames_per_gpu": N, "l2": "shard (%.2f GB) larger than L2" % (X.nbytes / 1e9),
        "parallelism": f"utterance-sharded x{world}" + (", one fp64 allreduce of the statistics per step" if em
                                                        else ", no collective")}
    if em and rank == 0 and not args.no_cpu:
        line["cpu_baseline"] = cpu_line(cpu_em, models, base, args.cpu_utts or min(args.utts, 50_000), args.frames,
                                        "oracle.train_iteration")
    if em and world == 1 and not args.no_sub:
        # the other four configurations as complete sub-lines (BENCH_rNN.json then carries all five configs)
        steps, warmup = sub_steps(args)
        t0 = time.time()
        line["decode"] = bench_decode(args, ctx, steps, warmup)
        log(f"decode sub-line {time.time() - t0:.1f} s")
        for name in ("connected", "stress", "fb", "fit"):
            t0 = time.time()
            try:
                line[name] = sub_fns[name](args, ctx, steps, warmup)
            except Exception as e:  # noqa: BLE001  (a sub-line must not take the headline down with it)
                import traceback
                traceback.print_exc()
                line[name] = {"error": f"{type(e).__name__}: {e}"}
            torch.cuda.empty_cache()
            log(f"{name} sub-line {time.time() - t0:.1f} s")
    if em and world > 1 and strong and not args.no_sub:
        # the weak-scaling figure next to the strong one: every rank its own 100 k utterances
        from srb200.trainer import IsolatedTrainer
        Xw, offw, wordsw = synth.make_utterances(100 + rank, base, args.utts, T=args.frames)
        Xw, offw, wordsw, _ = sort_by_word(Xw, offw, wordsw)
        pmw, _ = engine.PackedModel.from_words(models)
        trw = IsolatedTrainer(Xw, offw, wordsw, pmw, np.stack([m[3] for m in models]), S_, NCOMP)
        steps, warmup = sub_steps(args)
        wms, _ = timed_steps(trw.em_iteration, steps, warmup, torch, dist, world)
        line["weak_scaling"] = {"value": world * Xw.shape[0] * steps / (wms * 1e-3), "unit": "frames/s",
                                "ms_per_step": wms / steps, "utterances_per_gpu": args.utts, "scaling": "weak"}
    if rank == 0:
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def run_reference(args):
    """Reference arm: the CPU restatement of the reference path (the reference itself is pure Python and does not
    travel to the GPU box), all host threads, same config/metric.  Rank 0 only."""
    rank, _, world = dist_env()
    if rank != 0:
        return
    from srb200 import synth
    models, base = synth.make_models(1, W=W_, S=S_, M=M_)
    em = args.workload == "em"
    n_cpu = args.cpu_utts or min(args.utts, 20_000 if em else 10_000)
    fn = cpu_em if em else cpu_decode
    v, cores, ms = fn(models, base, n_cpu, args.frames, steps=args.steps, warmup=min(args.warmup, 1))
    metric = ("EM-iteration frames/s (segmental K-means alignment + GMM EM statistics + M-step)" if em else
              "decoded frames/s (GMM scoring of all 11 word models + Viterbi)")
    line = {
        "impl": "reference",
        "metric": metric,
        "value": v, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms, "higher_is_better": True, "scaling": (args.scaling if em else "weak"), "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"{args.workload}: {W_} words x {S_} states x {M_} diag Gaussians, D={D_}, "
                               f"bounded sample of {n_cpu} utterances x {args.frames} frames per step"},
        "cpu_baseline": {"value": v, "unit": "frames/s", "cores": cores, "kind": "port",
                         "sample": f"{n_cpu} utterances x {args.frames} frames per step"},
        "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(line)


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
