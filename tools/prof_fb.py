import sys, os, time
import numpy as np, torch
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/speech-recognition_b200')
import bench
from srb200 import engine, synth, lib as _l
from srb200.trainer import BaumWelchTrainer
S,M=8,16
models, base = synth.make_models(11, W=11, S=S, M=M)
X, off, words = bench.make_ragged_corpus_device(torch, base, 200000, 50, 300, 400)
pm,_=engine.PackedModel.from_words(models)
tr=BaumWelchTrainer(X, off, words, pm, np.stack([m[3] for m in models]), S, 4)
print('N',tr.N,'gemm',tr.gemm, file=sys.stderr)
L=_l.load(); p=_l.ptr
def t(fn,n=3):
    fn(); torch.cuda.synchronize()
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1)/n
def score():
    st=_l.stream_ptr()
    _l.check(L.srb_pack_model_blocks_sb(p(pm.ab), p(pm.c), pm.R, pm.Mp, pm.D, S, p(tr.blocks), st))
    _l.check(L.srb_gmm_score_gemm(p(tr.X), tr.N, pm.D, p(tr.blocks), p(pm.shift), pm.R, pm.Mp, S, p(tr.score_tiles), tr.word_tiles[-1], p(tr.E), S, st))
def fb():
    engine.forward_backward(tr.E, tr.off, tr.gb, utt_graph=tr.words, want_xi=True)
print('score ms', t(score), 'fb ms', t(fb), 'e_step', t(tr.e_step), 'acc', t(tr.accumulate), file=sys.stderr)
def fb_noxi():
    engine.forward_backward(tr.E, tr.off, tr.gb, utt_graph=tr.words, want_xi=False)
print('fb without xi ms', t(fb_noxi), file=sys.stderr)
import torch
offd = engine.as_offsets(tr.off)
def sort_only():
    torch.argsort(offd[1:] - offd[:-1], descending=True, stable=True).to(torch.int32)
print('argsort ms', t(sort_only), file=sys.stderr)
