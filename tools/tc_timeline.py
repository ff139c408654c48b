"""Pipeline timeline of the tensor-core classifier (srb_classify_tc_timeline): per tile, cycles of CTA 0's events relative to
the MMA issue of that tile.  Run on a B200:  python tools/tc_timeline.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "speech-recognition_b200"))
import torch  # noqa: E402
from srb200 import engine, synth, lib as _l  # noqa: E402

W, S, M = 11, 5, 4
models, base = synth.make_models(1, W=W, S=S, M=M)
X, off, words = synth.make_utterances(2, base, 20000, T=100)
pm, trans = engine.PackedModel.from_words(models)
Xd, offd = engine.as_frames(X), engine.as_offsets(off)
buf = torch.zeros((64, 10), dtype=torch.int64, device="cuda")
for _ in range(3):
    engine.classify(Xd, offd, pm, trans, W, S, tensor=True)
B = offd.numel() - 1
final = torch.empty((B, W), dtype=torch.float32, device="cuda")
label = torch.empty(B, dtype=torch.int32, device="cuda")
_l.check(_l.load().srb_classify_tc_timeline(_l.ptr(Xd), _l.ptr(offd), B, Xd.shape[0], pm.D, _l.ptr(pm.tcb), _l.ptr(pm.shift),
                                            W, S, pm.Mp, _l.ptr(trans), _l.ptr(final), _l.ptr(label), _l.ptr(buf),
                                            _l.stream_ptr()))
torch.cuda.synchronize()
t = buf.cpu().numpy()
names = ["mma:AR", "mma:DF", "mma:commit", "scoreA:HB", "scoreA:end", "vit:prep0", "vit:prep1", "vit:dp0", "vit:dp1", "scoreB:end"]
print("tile  period | " + "  ".join(f"{n:>10}" for n in names))
for j in range(8, 40):
    ref = t[j, 1]
    print(f"{j:4d} {t[j, 1] - t[j - 1, 1]:7d} | " + "  ".join(f"{int(v - ref):10d}" for v in t[j]))
