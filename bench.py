#!/usr/bin/env python
"""bench.py -- throughput of the B200 GMM-HMM hot path (BASELINE.json metric), one JSON line on rank 0.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload decode|em] [--impl ours|reference]

Workloads (synthetic data of the shapes BASELINE.json names, seeded; see srb200/synth.py):
  decode  config 1 shape at benchmark scale: 11 words x 5 states x 4 Gaussians, D=39; every 100-frame utterance
          is scored against all 11 word models and Viterbi-decoded (core.py:79-87), `--utts` utterances per GPU.
  em      config 2: one segmental-K-means + EM iteration over `--utts` labelled utterances per GPU
          (Viterbi alignment -> posteriors -> sufficient statistics -> allreduce -> M-step).

A "step" is one pass of the workload over the rank's whole shard.  `value` is measured with the shard resident in
HBM; `e2e` goes through the host-facing API with pinned HOST buffers, H2D and D2H inside the timed region.
The shard (>= 1.5 GB) is larger than the 126 MB L2, so no flush is needed between steps.
"""
import argparse
import json
import os
import subprocess
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
# DRAM bytes per frame of the two headline kernels, from the committed `ncu --set full` captures
# (dram__bytes_read.sum + dram__bytes_write.sum over one 2 M-frame / 10 M-frame launch; algorithmic: 156 B/frame of X + outputs)
NCU_TRAFFIC_PER_FRAME = {"classify_kernel<39>": (318.694e6 / 2e6, "profiles/r01_ncu_classify_v3_raw.csv"),
                         "classify_tc_kernel<39>": ((312.401e6 + 4.543e6) / 2e6, "profiles/r01_ncu_classify_tc_v10_raw.csv"),
                         "em_isolated_kernel<39,5,2>": ((1565.19e6 + 4.78e6) / 1e7, "profiles/r01_ncu_emfused_v6_raw.csv")}


# stdout carries exactly ONE JSON line: the descriptor is saved here and fd 1 is pointed at stderr for the rest of the
# run, so that whatever a library prints (NCCL's version banner, for one) cannot land next to it.
_STDOUT_FD = os.dup(1)
os.dup2(2, 1)


def emit(line):
    os.write(_STDOUT_FD, (json.dumps(line) + "\n").encode())


def ncu_traffic(kernel, frames):
    if kernel not in NCU_TRAFFIC_PER_FRAME:
        return {"traffic": None, "traffic_source": "no ncu --set full capture committed for this kernel yet"}
    per, src = NCU_TRAFFIC_PER_FRAME[kernel]
    return {"traffic": per * frames, "traffic_source": f"{per:.1f} B/frame x frames per launch, {src}"}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default=os.environ.get("SRB_WORKLOAD", "em"), choices=["decode", "em", "connected", "stress", "fb", "frontend"])
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--utts", type=int, default=100_000, help="utterances per GPU (100 frames each)")
    ap.add_argument("--frames", type=int, default=100)
    ap.add_argument("--cpu-utts", type=int, default=0, help="utterances of the CPU sample (0 = auto)")
    ap.add_argument("--no-decode", action="store_true", help="skip the decode sub-benchmark of the em line")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--ffma", action="store_true", help="decode: force the FP32-FMA classifier instead of the tcgen05 one")
    ap.add_argument("--unfused", action="store_true", help="em: use the separate K1/K2/K5 kernels instead of the fused pass")
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

    def stop(self):
        self.stop_flag = True
        if self.thread:
            self.thread.join()
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------ CPU legs
def cpu_decode(models, base, n_utts, frames, steps=1, warmup=0):
    """Times the float64 CPU restatement of the reference classifier (oracle/sr_oracle.c:orc_classify, OpenMP over
    utterances, linear-domain emissions like hmm_state.py:36-45) on a bounded sample.  -> (frames/s, cores, ms)"""
    from oracle import oracle as orc
    from srb200 import synth
    orc.set_num_threads()
    X, off, _ = synth.make_utterances(2, base, n_utts, T=frames)
    X64, off64 = X.astype(np.float64), off.astype(np.int64)
    for _ in range(warmup):
        orc.classify(X64, off64, models)
    ts = []
    for _ in range(max(steps, 1)):
        t0 = time.perf_counter()
        orc.classify(X64, off64, models)
        ts.append(time.perf_counter() - t0)
    dt = float(np.mean(ts))
    return X.shape[0] / dt, orc.num_threads(), dt * 1e3


def cpu_em(models, base, n_utts, frames, steps=1, warmup=0):
    """Times the CPU restatement of one training pass (oracle.train_iteration: align_gmm_states for every utterance,
    OpenMP over utterances, then one GMM.em iteration per tied state) on a bounded sample."""
    from oracle import oracle as orc
    from srb200 import synth
    orc.set_num_threads()
    X, off, words = synth.make_utterances(2, base, n_utts, T=frames)
    X64, off64 = X.astype(np.float64), off.astype(np.int64)
    for _ in range(warmup):
        orc.train_iteration(X64, off64, words, models, NCOMP)
    ts = []
    for _ in range(max(steps, 1)):
        t0 = time.perf_counter()
        orc.train_iteration(X64, off64, words, models, NCOMP)
        ts.append(time.perf_counter() - t0)
    dt = float(np.mean(ts))
    return X.shape[0] / dt, orc.num_threads(), dt * 1e3


# ------------------------------------------------------------------------------------------ distributed plumbing
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


def bench_decode(args, ctx):
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

    sampler = ClockSampler(ctx["local"])
    sampler.start()
    total_ms, per = timed_steps(step, args.steps, args.warmup, torch, dist, world)
    clocks = sampler.stop()
    acc = float((label.cpu().numpy() == words).mean())
    kern_ms = float(np.mean(per))
    flops = 2.0 * FMA_PER_FRAME_DECODE * N / (kern_ms * 1e-3)
    peak_fma = ctx["peak_fma"]
    pk = peaks()
    hbm = {"achieved_gbs": (N * D_ * 4 + B * (W_ + 1) * 4) / (kern_ms * 1e-3) / 1e9, "peak_gbs": pk.get("hbm_gbs")}
    if tensor:
        # tcgen05 classifier: the contraction runs as 3 TF32 products per FP32 product on 128 x 224 x 80 tiles
        # (N 220 -> 224, K 78 -> 80).  `achieved` stays the ALGORITHMIC rate (17 160 FMA per frame);
        # `peak` is the measured dense bf16 rate, of which TF32 gets half and the 3-way split a third of that.
        bf16 = pk.get("bf16_tflops", 2250.0)
        executed = 3 * 2.0 * 224 * 80 * N / (kern_ms * 1e-3)
        roof = {"bound": "tensor", "achieved": flops / 1e12, "peak": bf16, "unit": "TFLOP/s", "frac": flops / 1e12 / bf16,
                **ncu_traffic("classify_tc_kernel<39>", N), "kernel": "classify_tc_kernel<39>", "kernel_ms": kern_ms,
                "peak_source": "MEASURED_PEAKS.json bf16_tflops (burst)" if "bf16_tflops" in pk else "nominal 2250",
                "algorithmic": "17160 FMA/frame = 220 Gaussians x 2*39 (SURVEY 8d), FP32-equivalent",
                "tensor_pipe": {"executed_tflops": executed / 1e12, "tf32_peak_tflops": bf16 / 2,
                                "frac": executed / 1e12 / (bf16 / 2),
                                "note": "executed = 3 TF32 MMAs per product on padded 128x224x80 tiles; ncu "
                                        "(profiles/r01_ncu_classify_tc_v10_raw.csv): sm__pipe_tensor_cycles_active "
                                        "43.4 % of elapsed"},
                "fp32_fma": {"peak_tflops": 2 * peak_fma / 1e12, "frac": flops / (2 * peak_fma),
                             "note": "algorithmic rate over the measured FFMA2 peak; > 1 = above the CUDA-core roofline"},
                "hbm": hbm}
    else:
        roof = {"bound": "fp32_fma", "achieved": flops / 1e12, "peak": 2 * peak_fma / 1e12,
                "unit": "TFLOP/s", "frac": flops / (2 * peak_fma), **ncu_traffic("classify_kernel<39>", N),
                "kernel": "classify_kernel<39>", "kernel_ms": kern_ms,
                "peak_source": "srb_fma_peak (FFMA2 probe) measured in this run",
                "algorithmic": "17160 FMA/frame = 220 Gaussians x 2*39 (SURVEY 8d)", "hbm": hbm}
    out = {
        "metric": "decoded frames/s (GMM scoring of all 11 word models + Viterbi)",
        "value": world * N * args.steps / (total_ms * 1e-3), "unit": "frames/s", "ms_per_step": total_ms / args.steps,
        "gpu_launches": args.steps,          # one fused classifier launch per step
        "roofline": roof, "accuracy": acc, "clocks": clocks,
    }
    # ---- end to end: pinned host buffers -> chunked H2D on two streams -> classify -> D2H labels
    Xh = torch.from_numpy(X).pin_memory()
    offh = torch.from_numpy(off.astype(np.int32))
    runner = engine.HostClassifier(pm, trans, W_, S_, max_utts=8192, max_frames=8192 * args.frames)
    out_labels = torch.empty(B, dtype=torch.int32).pin_memory()
    e2e_ms, _ = timed_steps(lambda: runner.run(Xh, offh, out_labels), args.steps, args.warmup, torch, dist, world)
    assert np.array_equal(out_labels.numpy(), label.cpu().numpy()), "e2e labels differ from the resident run"
    out["e2e"] = {"value": world * N * args.steps / (e2e_ms * 1e-3), "unit": "frames/s",
                  "h2d_bytes_per_step": int(X.nbytes + off.nbytes), "d2h_bytes_per_step": int(B * 4),
                  "api": "srb200.engine.HostClassifier.run (host numpy/pinned buffers in, labels out)"}
    del Xd, Xh
    return out


def bench_em(args, ctx):
    """Config 2: one segmental-K-means + EM iteration over the rank's shard (align -> stats -> allreduce -> M-step)."""
    torch, dist, engine = ctx["torch"], ctx["dist"], ctx["engine"]
    from srb200.trainer import IsolatedTrainer
    X, off, words, world = ctx["X"], ctx["off"], ctx["words"], ctx["world"]
    N, B = int(X.shape[0]), len(off) - 1
    pm, _ = engine.PackedModel.from_words(ctx["models"])       # private copy: training moves it
    trans_np = np.stack([m[3] for m in ctx["models"]])
    tr = IsolatedTrainer(X, off, words, pm, trans_np, S_, NCOMP, fused=None if not args.unfused else False)
    sampler = ClockSampler(ctx["local"])
    sampler.start()
    total_ms, per = timed_steps(tr.em_iteration, args.steps, args.warmup, torch, dist, world)
    clocks = sampler.stop()
    mean_cost = tr.mean_cost()
    pk = peaks()
    peak_fma = ctx["peak_fma"]
    if tr.fused:
        # phase breakdown (outside the timed region): CUDA events around the fused kernel call and the rest
        names = ["em_fused", "allreduce", "mstep"]
        acc = dict.fromkeys(names, 0.0)
        reps = 5
        for _ in range(reps):
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
            ev[0].record()
            tr.fused_pass()
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
        flops = 2.0 * fma_pf * N / (kern_ms * 1e-3)
        gbs = (D_ * 4) * N / (kern_ms * 1e-3) / 1e9
        roof = {"bound": "fp32_fma", "achieved": flops / 1e12, "peak": 2 * peak_fma / 1e12, "unit": "TFLOP/s",
                "frac": flops / (2 * peak_fma), **ncu_traffic("em_isolated_kernel<39,5,2>", N),
                "kernel": "em_isolated_kernel<39,5,2>",
                "kernel_ms": kern_ms, "peak_source": "srb_fma_peak (FFMA2 probe) measured in this run",
                "algorithmic": f"{fma_pf} FMA/frame = {S_ * M_} Gaussians x 2*39 scoring + {NCOMP} x 2*39 accumulation; "
                               f"{D_ * 4} B/frame of X read once",
                "hbm": {"achieved_gbs": gbs, "peak_gbs": pk.get("hbm_gbs"), "frac": gbs / pk.get("hbm_gbs")},
                "phase_ms": acc}
        launches = 6          # em_isolated, reduce_slices, cost sum (torch, 1-2), mstep, pack
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
        "value": world * N * args.steps / (total_ms * 1e-3), "unit": "frames/s", "ms_per_step": total_ms / args.steps,
        "gpu_launches": args.steps * launches, "roofline": roof,
        "mean_viterbi_cost": mean_cost, "clocks": clocks, "fused": bool(tr.fused),
    }
    # ---- end to end: the shard comes from pinned host memory every step, the updated model goes back
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
    out["e2e"] = {"value": world * N * args.steps / (e2e_ms * 1e-3), "unit": "frames/s",
                  "h2d_bytes_per_step": int(Xh.numel() * 4), "d2h_bytes_per_step": int(host_model.numel() * 8),
                  "api": "srb200.trainer.IsolatedTrainer.em_iteration_from_host (pinned host shard in, statistics/model out)"}
    return out


def _roof_fma(fma_pf, N, ms, peak_fma, kernel, note):
    flops = 2.0 * fma_pf * N / (ms * 1e-3)
    return {"bound": "fp32_fma", "achieved": flops / 1e12, "peak": 2 * peak_fma / 1e12, "unit": "TFLOP/s",
            "frac": flops / (2 * peak_fma), "traffic": None, "kernel": kernel, "kernel_ms": ms,
            "peak_source": "srb_fma_peak (FFMA2 probe) measured in this run", "algorithmic": note}


def bench_connected(args, ctx):
    """Config 4: connected-digit decoding, 7 layers x 11 words (393 states, 847 arcs), 500-frame utterances:
    K1 scoring of the 55 tied states + K2 graph Viterbi with back-pointers, traceback and paths."""
    torch, dist, engine = ctx["torch"], ctx["dist"], ctx["engine"]
    from srb200 import synth
    from srb200.graph import GraphBatch, word_loop_graph
    from srb200.trainer import ViterbiPlan
    B, T, L = args.utts if args.utts != 100_000 else 4096, 500, 7
    X, off, strings = synth.make_strings(300 + ctx["rank"], ctx["base"], B, L, T)
    g, word_of = word_loop_graph(np.stack([m[3] for m in ctx["models"]]), L)
    gb = GraphBatch([g], engine.device())
    Xd = engine.as_frames(X)
    N = Xd.shape[0]
    plan = ViterbiPlan(off, gb, want_path=True, want_label=True)
    pm = ctx["pm"]
    E = torch.empty((N, pm.R), dtype=torch.float32, device="cuda")
    L_ = ctx["lib"]

    use_tc = pm.scorer_tensor_core_ok() and not args.ffma

    def score():
        if use_tc:
            engine.gmm_score_tc(Xd, pm, E)
        else:
            L_.check(L_.load().srb_gmm_score(L_.ptr(Xd), N, pm.D, L_.ptr(pm.ab), L_.ptr(pm.c), L_.ptr(pm.shift), pm.R,
                                             pm.Mp, L_.ptr(E), pm.R, None, None, None, L_.stream_ptr()))

    def step():
        score()
        plan.run(E)

    sampler = ClockSampler(ctx["local"])
    sampler.start()
    total_ms, per = timed_steps(step, args.steps, args.warmup, torch, dist, ctx["world"])
    clocks = sampler.stop()
    # decoded strings from the frame labels (main.py:61-67 equivalent): word of the state at each word-slot centre
    lab = plan.label.cpu().numpy().reshape(B, T)
    centres = ((np.arange(L) + 0.5) * T / L).astype(int)
    acc = float((word_of[lab[:, centres]] == strings).mean())
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    ev[0].record()
    score()
    ev[1].record()
    plan.run(E)
    ev[2].record()
    torch.cuda.synchronize()
    ms = float(np.mean(per))
    roof = _roof_fma(FMA_PER_FRAME_DECODE, N, ms, ctx["peak_fma"],
                     ("classify_tc_kernel<39> (scorer mode, tcgen05 3xTF32)" if use_tc else "score_kernel<39>") +
                     " + viterbi_banded1_kernel (one-graph banded form of srb_viterbi_banded)",
                     "17160 FMA/frame scoring (55 tied states x 4 x 2*39); DP 847 arcs + 393 states per frame; "
                     "393 B/frame back-pointers")
    roof["phase_ms"] = {"score": ev[0].elapsed_time(ev[1]), "viterbi": ev[1].elapsed_time(ev[2])}
    return {"metric": "decoded frames/s (connected-digit grammar, GMM scoring + graph Viterbi with back-pointers)",
            "value": ctx["world"] * N * args.steps / (total_ms * 1e-3), "unit": "frames/s",
            "ms_per_step": total_ms / args.steps, "gpu_launches": 2 * args.steps, "roofline": roof,
            "digit_accuracy": acc, "clocks": clocks,
            "_workload": f"connected: 7 layers x 11 words (393 states), {B} utterances x {T} frames per GPU",
            "_frames": N}


def bench_stress(args, ctx):
    """Config 5: likelihood-only sweep, 2000 tied states x 32 Gaussians x 39 dims; per-frame best state streamed out
    (the 8 kB/frame cost matrix is never materialised)."""
    torch, dist, engine = ctx["torch"], ctx["dist"], ctx["engine"]
    R, M = 2000, 32
    rng = np.random.default_rng(7)
    mu = rng.normal(size=(R, 1, D_)) + 0.5 * rng.normal(size=(R, M, D_))
    var = rng.uniform(0.5, 1.5, size=(R, M, D_))
    w = rng.dirichlet(np.ones(M), size=R)
    pm = engine.PackedModel(mu, var, w)
    N = args.utts * args.frames
    Xd = torch.randn((N, D_), dtype=torch.float32, device="cuda", generator=torch.Generator("cuda").manual_seed(9))
    bc = torch.empty(N, dtype=torch.float32, device="cuda")
    bs = torch.empty(N, dtype=torch.int32, device="cuda")
    L_ = ctx["lib"]

    use_tc = not args.ffma
    launches = 1

    def step():
        if use_tc:
            engine.gmm_best_gemm(Xd, pm, bc, bs)
        else:
            L_.check(L_.load().srb_gmm_score(L_.ptr(Xd), N, D_, L_.ptr(pm.ab), L_.ptr(pm.c), L_.ptr(pm.shift), R, pm.Mp,
                                             None, R, L_.ptr(bc), L_.ptr(bs), None, L_.stream_ptr()))

    sampler = ClockSampler(ctx["local"])
    sampler.start()
    total_ms, per = timed_steps(step, args.steps, args.warmup, torch, dist, ctx["world"])
    clocks = sampler.stop()
    ms = float(np.mean(per))
    roof = _roof_fma(R * M * 2 * D_, N, ms, ctx["peak_fma"],
                     "gmm_best_gemm_kernel<39> (persistent tcgen05 3xTF32 GEMM, model streamed by TMA bulk copies)"
                     if use_tc else "score_kernel<39>", "4.992 M FMA/frame = 64000 Gaussians x 2*39")
    if use_tc:
        roof["note"] = ("tcgen05 3xTF32: algorithmic FP32-equivalent rate over the measured FFMA peak, > 1 = above the "
                        "CUDA-core roofline; --ffma runs the FP32-FMA scorer")
    return {"metric": "GMM-scored frames/s (2000 states x 32 Gaussians, best state per frame)",
            "value": ctx["world"] * N * args.steps / (total_ms * 1e-3), "unit": "frames/s",
            "ms_per_step": total_ms / args.steps, "gpu_launches": args.steps * launches, "roofline": roof, "clocks": clocks,
            "_workload": f"stress: 2000 states x 32 Gaussians, {N} frames per GPU", "_frames": N}


def bench_fb(args, ctx):
    """Config 3: one Baum-Welch iteration, 8 states x 16 Gaussians per word (int(log 16) = 2 splits -> 4 trained
    components, hmm.py:104), ragged 50-300 frame utterances: per-word K1 scoring + K4 forward-backward + soft-count
    statistics + allreduce + M-step + transition re-estimation.  No reference counterpart (parity unpinned)."""
    torch, dist, engine = ctx["torch"], ctx["dist"], ctx["engine"]
    from srb200 import synth
    from srb200.trainer import BaumWelchTrainer
    S, M = 8, 16
    ncomp = 2 ** int(np.log(M))
    models, base = synth.make_models(11, W=W_, S=S, M=M)
    B = args.utts if args.utts != 100_000 else 50_000
    X, off, words = synth.make_utterances(400 + ctx["rank"], base, B, ragged=(50, 300))
    pm, _ = engine.PackedModel.from_words(models)
    tr = BaumWelchTrainer(X, off, words, pm, np.stack([m[3] for m in models]), S, ncomp)
    N = tr.N
    sampler = ClockSampler(ctx["local"])
    sampler.start()
    total_ms, per = timed_steps(tr.em_iteration, args.steps, args.warmup, torch, dist, ctx["world"])
    clocks = sampler.stop()
    names = ["score+forward_backward", "soft_stats", "allreduce", "mstep+transitions"]
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
    ev[0].record(); tr.e_step(); ev[1].record(); tr.accumulate(); ev[2].record(); tr.reduce(); ev[3].record()
    tr.update(); ev[4].record()
    torch.cuda.synchronize()
    ms = float(np.mean(per))
    roof = _roof_fma(S * M * 2 * D_ + S * ncomp * 2 * D_, N, ms, ctx["peak_fma"],
                     "classify_tc_kernel<39> (scorer mode, 32 sub-states x 4) x11 + neg_lse_groups_kernel x11 + forward_backward_banded_kernel<8> + gmm_stats_soft_kernel<4> x11",
                     "9984 FMA/frame scoring (8 states x 16 x 2*39) + up to 8 x 4 x 2*39 soft accumulation; "
                     "22 log-add arcs per frame, twice")
    roof["phase_ms"] = {k: ev[i].elapsed_time(ev[i + 1]) for i, k in enumerate(names)}
    return {"metric": "Baum-Welch EM-iteration frames/s (8 states x 16 Gaussians, ragged 50-300 frames)",
            "value": ctx["world"] * N * args.steps / (total_ms * 1e-3), "unit": "frames/s",
            "ms_per_step": total_ms / args.steps, "gpu_launches": (2 * W_ + 10) * args.steps, "roofline": roof,
            "mean_neg_log_z_per_utt": tr.mean_neg_log_z(), "clocks": clocks,
            "_workload": f"fb: 11 words x 8 states x 16 Gaussians ({ncomp} trained), {B} utterances of 50-300 frames per GPU",
            "_frames": N}


def run_ours(args):
    import torch
    import torch.distributed as dist
    rank, local, world = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU path for the product arm")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from srb200 import engine, synth
    from srb200.trainer import sort_by_word

    from srb200 import lib as lib_
    models, base = synth.make_models(1, W=W_, S=S_, M=M_)
    pm, trans = engine.PackedModel.from_words(models)
    peak_fma = max(engine.fma_peak(True, 8192) for _ in range(3))      # FMA/s, measured now on this GPU
    ctx = {"torch": torch, "dist": dist, "engine": engine, "pm": pm, "trans": trans, "models": models, "base": base,
           "world": world, "local": local, "rank": rank, "peak_fma": peak_fma, "lib": lib_}
    extra = {"connected": bench_connected, "stress": bench_stress, "fb": bench_fb, "frontend": bench_frontend}
    if args.workload in extra:
        # parity-test configurations of BASELINE.json (configs[2..4]) as optional device-resident sub-benchmarks
        line = extra[args.workload](args, ctx)
        wl, N = line.pop("_workload"), line.pop("_frames")
        line.update({"n_gpus": world, "steps": args.steps, "warmup": args.warmup, "higher_is_better": True,
                     "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                     "config": {"workload": wl, "frames_per_gpu": N, "l2": "inputs larger than L2",
                                "parallelism": f"utterance-sharded x{world}, no collective"}})
        if rank == 0:
            emit(line)
        if world > 1:
            dist.destroy_process_group()
        return
    X, off, words = synth.make_utterances(100 + rank, base, args.utts, T=args.frames)
    X, off, words, _ = sort_by_word(X, off, words)
    N, B = int(X.shape[0]), args.utts
    ctx.update({"X": X, "off": off, "words": words})
    line = (bench_em if args.workload == "em" else bench_decode)(args, ctx)
    if args.workload == "em" and world == 1 and not args.no_decode:
        d = bench_decode(args, ctx)
        line["decode"] = {k: d[k] for k in ("metric", "value", "unit", "ms_per_step", "roofline", "e2e", "accuracy")}
    line.update({
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.workload}: {W_} words x {S_} states x {M_} diag Gaussians, D={D_}, "
                               f"{B} utterances x {args.frames} frames per GPU"
                               + (f", EM over the first {NCOMP} components (hmm.py:104)" if args.workload == "em" else ""),
                   "utterances_per_gpu": B, "frames_per_gpu": N, "l2": "shard (%.2f GB) larger than L2" % (X.nbytes / 1e9),
                   "parallelism": f"utterance-sharded x{world}" + (", one fp64 allreduce of the statistics per step"
                                                                   if args.workload == "em" else ", no collective")},
    })
    if rank == 0 and not args.no_cpu:
        n_cpu = args.cpu_utts or min(args.utts, 100_000 if args.workload == "em" else 50_000)
        fn = cpu_em if args.workload == "em" else cpu_decode
        v, cores, ms = fn(models, base, n_cpu, args.frames)
        line["cpu_baseline"] = {"value": v, "unit": "frames/s", "cores": cores, "kind": "port",
                                "sample": f"{n_cpu} utterances x {args.frames} frames, oracle/ float64 C restatement "
                                          f"(OpenMP), {ms:.0f} ms"}
    if rank == 0:
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def bench_frontend(args, ctx):
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
    sampler = ClockSampler(ctx["local"])
    sampler.start()
    total_ms, per = timed_steps(lambda: fe.run(st, out=out), args.steps, args.warmup, torch, dist, world)
    clocks = sampler.stop()
    F = st["F"]
    kern_ms = float(np.mean(per))
    # algorithmic bytes per frame: 160 new samples in, 13 cepstra out and back in, 39 features out, standardisation
    # re-reads them twice and rewrites once (L2-resident per utterance)
    bytes_pf = 160 * 4 + 13 * 4 * 2 + 39 * 4
    gbs = bytes_pf * F / (kern_ms * 1e-3) / 1e9
    pk = peaks()
    e2e_ms, _ = timed_steps(lambda: fe.features(sigs)["features"].cpu(), max(1, args.steps // 3), 1, torch, dist, world)
    from oracle import feature_oracle as fo
    t0 = time.perf_counter()
    n_cpu = 200
    for s in sigs[:n_cpu]:
        fo.load_features(s)
    cpu_s = time.perf_counter() - t0
    return {"metric": "MFCC feature frames/s (pre-emphasis, FFT-512, 40 mel filters, DCT, deltas, standardisation)",
            "value": world * F * args.steps / (total_ms * 1e-3), "unit": "frames/s", "ms_per_step": total_ms / args.steps,
            "gpu_launches": 2 * args.steps,
            "roofline": {"bound": "hbm", "achieved": gbs, "peak": pk.get("hbm_gbs"), "unit": "GB/s",
                         "frac": gbs / pk.get("hbm_gbs"), "traffic": None, "kernel": "mfcc_kernel", "kernel_ms": kern_ms,
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs" if "source" not in pk else "fallback",
                         "algorithmic": f"{bytes_pf} B/frame; the kernel is bound by the shared-memory FFT (one warp "
                                        f"per frame, 9 radix-2 stages), not by HBM"},
            "clocks": clocks,
            "e2e": {"value": world * F * max(1, args.steps // 3) / (e2e_ms * 1e-3), "unit": "frames/s",
                    "h2d_bytes_per_step": int(st["sig"].numel() * 2), "d2h_bytes_per_step": int(F * 39 * 4),
                    "api": "srb200.frontend.MfccFrontEnd.features (list of host int16 sample arrays in, features out; "
                           "includes the host-side concatenation of the arrays)"},
            "cpu_baseline": {"value": n_cpu * 100 / cpu_s, "unit": "frames/s", "cores": 1, "kind": "port",
                             "sample": f"{n_cpu} utterances of 1 s, oracle/feature_oracle.py (numpy float64), "
                                       f"{cpu_s * 1e3:.0f} ms"},
            "_workload": f"frontend: {B} utterances x 16000 samples (1 s at 16 kHz) per GPU -> 100 frames x 39 each",
            "_frames": F}


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {"hbm_gbs": 6650.0, "source": "fallback"}


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
        "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
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
