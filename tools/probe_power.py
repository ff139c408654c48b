"""Is classify power/clock limited?  Isolated launches after idle vs back-to-back, with NVML clocks + power."""
import os, sys, time, threading, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "speech-recognition_b200")]
import numpy as np, torch, pynvml
from srb200 import engine, synth
pynvml.nvmlInit(); h = pynvml.nvmlDeviceGetHandleByIndex(0)
models, base = synth.make_models(1)
pm, trans = engine.PackedModel.from_words(models)
def mk(nu):
    X, off, w = synth.make_utterances(2, base, nu, T=100)
    return engine.as_frames(X), engine.as_offsets(off), torch.empty((nu, 11), dtype=torch.float32, device="cuda"), torch.empty(nu, dtype=torch.int32, device="cuda")
small, big = mk(20000), mk(100000)
def run(a): engine.classify(a[0], a[1], pm, trans, 11, 5, a[2], a[3])
def ev_time(fn):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); fn(); e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1)
run(small); run(big); torch.cuda.synchronize()
out = {"isolated_2M_ms": [], "isolated_10M_ms": []}
for _ in range(5):
    time.sleep(0.5); out["isolated_2M_ms"].append(ev_time(lambda: run(small)))
for _ in range(3):
    time.sleep(0.5); out["isolated_10M_ms"].append(ev_time(lambda: run(big)))
samples = []; stop = False
def loop():
    while not stop:
        samples.append((time.perf_counter(), pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM), pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0, pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)))
        time.sleep(0.003)
th = threading.Thread(target=loop); th.start()
t0 = time.perf_counter()
per = []
evs = [torch.cuda.Event(enable_timing=True) for _ in range(61)]
evs[0].record()
for i in range(60):
    run(big); evs[i + 1].record()
torch.cuda.synchronize(); t1 = time.perf_counter()
stop = True; th.join()
out["b2b_10M_ms"] = [evs[i].elapsed_time(evs[i + 1]) for i in range(60)]
out["nvml"] = [(round(t - t0, 4), c, p, r) for t, c, p, r in samples if t0 <= t <= t1][::4]
for packed in (0, 1):
    out["fma_peak_%d" % packed] = [engine.fma_peak(bool(packed), it) for it in (1024, 8192, 32768)]
print(json.dumps(out))
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "probe_power.json"), "w"))
