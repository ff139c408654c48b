"""Round-1 GPU probe: FP32 FMA peak (plain / packed) and raw kernel timings. Output -> gpurun_out/probe.json"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "speech-recognition_b200")]
import numpy as np, torch
from srb200 import engine, synth

out = {}
for packed in (0, 1):
    r = [engine.fma_peak(bool(packed), 8192) for _ in range(3)]
    out["fma_peak_packed" if packed else "fma_peak_plain"] = max(r)
    print("fma peak packed=%d: %.2f TFMA/s" % (packed, max(r) / 1e12), flush=True)

def timeit(fn, n=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ts = []
    for _ in range(n):
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts), float(np.median(ts))

models, base = synth.make_models(1)
nu = int(os.environ.get("NU", 100000))
X, off, words = synth.make_utterances(2, base, nu, T=100)
Xd, offd = engine.as_frames(X), engine.as_offsets(off)
pm, trans = engine.PackedModel.from_words(models)
N = Xd.shape[0]
final = torch.empty((nu, 11), dtype=torch.float32, device="cuda"); label = torch.empty(nu, dtype=torch.int32, device="cuda")
tmin, tmed = timeit(lambda: engine.classify(Xd, offd, pm, trans, 11, 5, final, label))
out["classify_ms"] = tmed; out["classify_frames_per_s"] = N / (tmed * 1e-3)
out["classify_fma_per_s"] = N * 17160 / (tmed * 1e-3)
print("classify: %.3f ms  %.3e frames/s  %.2f TFMA/s  acc=%.4f" % (tmed, N / tmed * 1e3, N * 17160 / tmed * 1e3 / 1e12,
      float((label.cpu().numpy() == words).mean())), flush=True)
Ns = 2_000_000
E = torch.empty((Ns, 55), dtype=torch.float32, device="cuda")
from srb200 import lib as _l
def score():
    _l.check(_l.load().srb_gmm_score(_l.ptr(Xd), Ns, 39, _l.ptr(pm.ab), _l.ptr(pm.c), _l.ptr(pm.shift), 55, 4, _l.ptr(E), 55, None, None, None, _l.stream_ptr()))
tmin, tmed = timeit(score)
out["score_ms"] = tmed; out["score_fma_per_s"] = Ns * 17160 / (tmed * 1e-3)
print("score 2M frames: %.3f ms %.2f TFMA/s" % (tmed, Ns * 17160 / tmed * 1e3 / 1e12), flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "probe.json"), "w"), indent=1)
