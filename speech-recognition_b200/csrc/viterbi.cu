// viterbi.cu -- K2: batched Viterbi over arbitrary state graphs with non-emitting states (decode.py:80-146),
// back-pointers + traceback, and K4: batched log-domain forward-backward (no reference counterpart).
//
// One warp per utterance, states strided over the lanes, the two live cost columns in shared memory.
// The reference sweeps states in index order inside a column and lets a transition that touches a
// non-emitting state (NES) read the SAME column (decode.py:109-111); a same-column read of a state with a
// higher index sees +inf (not yet written).  That order is reproduced by cutting a column into the runs of
// emitting states between consecutive NES: run k is relaxed in parallel, then NES k is reduced over its
// predecessors, then run k+1 ...  Ties: the first minimum in origin order wins (np.argmin, decode.py:118).
#include "common.cuh"

namespace srb {

constexpr int VW = 4;   // warps (utterances in flight) per CTA

struct GraphBatch {
    const int32_t *utt_graph;     // [B] graph of each utterance, or null (all use graph 0)
    const int32_t *g_state_off;   // [ng+1] first global state index of each graph
    const int32_t *g_nes_off;     // [ng+1] range of nes_idx
    const int32_t *g_end_off;     // [ng+1] range of end_state
    const int32_t *emit;          // [sum R] column of E, or -1 for a NES (hmm_state.py:90-91: evaluate -> 0)
    const int32_t *pred_ptr;      // [sum R + 1] absolute CSR offsets into pred_idx / pred_cost
    const int32_t *pred_idx;      // origin state (graph local), ascending per target
    const float *pred_cost;       // transitions[target, origin]
    const int32_t *nes_idx;       // ascending local indices of the NES of each graph
    const int32_t *end_state;     // candidate end states (local), frame = last (main.py:60 passes -1)
    const int32_t *end_frame;     // optional: frame of each end point (negative counts from the end, decode.py:131-137)
};

__device__ __forceinline__ void lexmin(float &v, int &p, float ov, int op) {
    if (ov < v || (ov == v && op < p)) {
        v = ov;
        p = op;
    }
}

// Best end point ('>=': the last listed minimum wins, decode.py:131-134), traceback (decode.py:142-146) and frame
// labels for one utterance whose last cost column is `prev`.  Shared by the generic and the banded kernel.
__device__ __forceinline__ void finish_utterance(const GraphBatch &G, int g, int s0, int R, int T, int64_t f0, int u,
                                                 const float *prev, const int32_t *emit, const int32_t *pptr,
                                                 const uint8_t *bpu, int lane, float *best_cost, int32_t *best_end,
                                                 const int64_t *path_off, int32_t *path_state, int32_t *path_frame,
                                                 int32_t *path_len, int32_t *frame_label, const float *cu = nullptr) {
    const float INF = CUDART_INF_F;
    // ---- best end point: '>=' -> the last listed minimum wins (decode.py:131-134); traceback :142-146
    if (lane == 0) {
        float bc = INF;
        int bi = -1, bj = T - 1;
        for (int e = G.g_end_off[g]; e < G.g_end_off[g + 1]; ++e) {
            const int s = G.end_state[e];
            int c = T - 1;
            if (G.end_frame) {                                   // any [state, frame] end point (decode.py:126-134)
                c = G.end_frame[e];
                if (c < 0) c += T;                               // Python indexing: -1 = the last frame
                if (c < 0 || c >= T) continue;                   // the host layer rejects these (IndexError)
            }
            const float v = (c == T - 1 || !cu) ? prev[s] : cu[(size_t)c * R + s];
            if (bc >= v) {
                bc = v;
                bi = s;
                bj = c;
            }
        }
        best_cost[u] = bc;
        best_end[u] = bi;
        int len = 0;
        if (bc < INF && (path_state || frame_label)) {
            int32_t *ps = path_state ? path_state + path_off[u] : nullptr;
            int32_t *pf = path_state ? path_frame + path_off[u] : nullptr;
            const int cap = path_state ? (int)(path_off[u + 1] - path_off[u]) : 0x7fffffff;
            int i = bi, j = bj;
            if (frame_label) {
                for (int c = bj + 1; c < T; ++c) frame_label[f0 + c] = -1;   // frames past an early end point
                if (emit[i] >= 0) frame_label[f0 + j] = s0 + i;
            }
            while (j > 0 && len < cap) {
                const int code = bpu[(size_t)j * R + i];
                const int o = G.pred_idx[pptr[i] + code];
                if (!(emit[o] < 0 || emit[i] < 0)) --j;
                i = o;
                if (ps) {
                    ps[len] = i;
                    pf[len] = j;
                }
                if (frame_label && emit[i] >= 0) frame_label[f0 + j] = s0 + i;
                ++len;
            }
            if (j > 0) len = -1;   // capacity exhausted
        } else if (frame_label) {
            for (int c = 0; c < T; ++c) frame_label[f0 + c] = -1;   // no path: frames are skipped downstream
        }
        if (path_len) path_len[u] = len;
    }
}

__global__ void __launch_bounds__(VW * 32)
viterbi_graph_kernel(const float *__restrict__ E, int64_t ldE, const int32_t *__restrict__ off, int B, GraphBatch G,
                     int Rmax, const int64_t *__restrict__ bp_off, uint8_t *__restrict__ bp,
                     float *__restrict__ costs_out, float *__restrict__ best_cost, int32_t *__restrict__ best_end,
                     const int64_t *__restrict__ path_off, int32_t *__restrict__ path_state,
                     int32_t *__restrict__ path_frame, int32_t *__restrict__ path_len,
                     int32_t *__restrict__ frame_label) {
    extern __shared__ float vsm[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float *prev = vsm + (size_t)warp * 2 * Rmax, *cur = prev + Rmax;
    const float INF = CUDART_INF_F;

    for (int u = blockIdx.x * VW + warp; u < B; u += gridDim.x * VW) {
        const int g = G.utt_graph ? G.utt_graph[u] : 0;
        const int s0 = G.g_state_off[g], R = G.g_state_off[g + 1] - s0;
        const int n0 = G.g_nes_off[g], K = G.g_nes_off[g + 1] - n0;
        const int64_t f0 = off[u];
        const int T = (int)(off[u + 1] - f0);
        uint8_t *bpu = bp + bp_off[u];
        float *cu = costs_out ? costs_out + bp_off[u] : nullptr;
        const int32_t *emit = G.emit + s0, *pptr = G.pred_ptr + s0;
        if (T <= 0) {
            if (lane == 0) {
                best_cost[u] = INF;
                best_end[u] = -1;
                if (path_len) path_len[u] = 0;
            }
            continue;
        }
        for (int r = lane; r < R; r += 32) prev[r] = INF;   // column -1 == the still-untouched last column (A8)
        __syncwarp();
        for (int c = 0; c < T; ++c) {
            const float *Erow = E + (f0 + c) * ldE;
            uint8_t *bpc = bpu + (size_t)c * R;
            int lo = 0;
            for (int k = 0; k <= K; ++k) {
                const int hi = (k < K) ? G.nes_idx[n0 + k] : R;
                for (int r = lo + lane; r < hi; r += 32) {        // emitting states of run k
                    float out;
                    if (c == 0 && r == 0) {
                        out = Erow[emit[0]];                       // decode.py:99-100
                        bpc[0] = 0;
                    } else {
                        const int p0 = pptr[r], p1 = pptr[r + 1];
                        float best = INF;
                        int bj = 0;
                        for (int j = p0; j < p1; ++j) {
                            const int o = G.pred_idx[j];
                            const float v = (emit[o] < 0) ? (o < r ? cur[o] : INF) : prev[o];
                            const float cand = G.pred_cost[j] + v;
                            if (cand < best) {
                                best = cand;
                                bj = j - p0;
                            }
                        }
                        out = (p1 > p0) ? best + Erow[emit[r]] : INF;
                        bpc[r] = (uint8_t)bj;
                    }
                    cur[r] = out;
                }
                __syncwarp();
                if (k < K) {                                       // the NES closing run k: same-column reduce
                    const int n = hi;
                    const int p0 = pptr[n], p1 = pptr[n + 1];
                    if (c == 0 && n == 0) {
                        if (lane == 0) {
                            cur[0] = 0.f;
                            bpc[0] = 0;
                        }
                    } else {
                        float best = INF;
                        int bj = 0x7fffffff;
                        for (int j = p0 + lane; j < p1; j += 32) {
                            const int o = G.pred_idx[j];
                            const float cand = G.pred_cost[j] + (o < n ? cur[o] : INF);
                            lexmin(best, bj, cand, j - p0);
                        }
#pragma unroll
                        for (int d = 16; d; d >>= 1) {
                            const float ov = __shfl_xor_sync(0xffffffffu, best, d);
                            const int op = __shfl_xor_sync(0xffffffffu, bj, d);
                            lexmin(best, bj, ov, op);
                        }
                        if (lane == 0) {
                            cur[n] = (p1 > p0) ? best : INF;
                            bpc[n] = (uint8_t)(p1 > p0 ? bj : 0);
                        }
                    }
                    __syncwarp();
                    lo = hi + 1;
                }
            }
            if (cu)
                for (int r = lane; r < R; r += 32) cu[(size_t)c * R + r] = cur[r];
            float *t = prev;
            prev = cur;
            cur = t;
            __syncwarp();
        }
        __syncwarp();                                              // costs_out rows written by other lanes are visible
        finish_utterance(G, g, s0, R, T, f0, u, prev, emit, pptr, bpu, lane, best_cost, best_end, path_off, path_state,
                         path_frame, path_len, frame_label, cu);
        __syncwarp();
    }
}


// ------------------------------------------------------------------------------------------------ K2, banded form
// Fast path for the graphs build_state_sequences produces (continuous_speech.py:13-53) and for plain left-to-right
// models: every emitting state r has its emitting predecessors inside {r-2, r-1, r} plus at most one non-emitting
// predecessor n < r ("entry"), and the states feeding a NES are emitting and have no entry themselves.  Then a column
// needs three data-parallel sweeps instead of one sequential run per NES:
//   1. all emitting states from the previous column (band costs a2, a1, a0; no predecessor-list walk),
//   2. all NES from the finished emitting states of this column,
//   3. the entry states, re-relaxed with their NES of this column.
// Candidate order, first-minimum ties and the back-pointer codes (position in the ascending predecessor list) are
// those of the generic kernel, so costs, back-pointers and paths are identical.
struct BandForm {
    const float *band;            // [sum R][3] = a2, a1, a0 (+inf when the arc is absent)
    const int32_t *codes;         // [sum R] bytes: code(r-2) | code(r-1) << 8 | code(r) << 16 | code(entry NES) << 24
    const int32_t *entry_nes;     // [sum R] local index of the entry NES or -1
    const float *entry_cost;      // [sum R]
    const int32_t *g_entry_off;   // [ng+1] range of entry_state
    const int32_t *entry_state;   // local indices of the states with an entry
};

__global__ void __launch_bounds__(VW * 32)
viterbi_banded_kernel(const float *__restrict__ E, int64_t ldE, const int32_t *__restrict__ off, int B, GraphBatch G,
                      BandForm F, int Rmax, const int64_t *__restrict__ bp_off, uint8_t *__restrict__ bp,
                      float *__restrict__ costs_out, float *__restrict__ best_cost, int32_t *__restrict__ best_end,
                      const int64_t *__restrict__ path_off, int32_t *__restrict__ path_state,
                      int32_t *__restrict__ path_frame, int32_t *__restrict__ path_len,
                      int32_t *__restrict__ frame_label) {
    extern __shared__ float vsm[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float *prev = vsm + (size_t)warp * 2 * Rmax, *cur = prev + Rmax;
    const float INF = CUDART_INF_F;

    for (int u = blockIdx.x * VW + warp; u < B; u += gridDim.x * VW) {
        const int g = G.utt_graph ? G.utt_graph[u] : 0;
        const int s0 = G.g_state_off[g], R = G.g_state_off[g + 1] - s0;
        const int n0 = G.g_nes_off[g], K = G.g_nes_off[g + 1] - n0;
        const int q0 = F.g_entry_off[g], Q = F.g_entry_off[g + 1] - q0;
        const int64_t f0 = off[u];
        const int T = (int)(off[u + 1] - f0);
        uint8_t *bpu = bp + bp_off[u];
        float *cu = costs_out ? costs_out + bp_off[u] : nullptr;
        const int32_t *emit = G.emit + s0, *pptr = G.pred_ptr + s0;
        const float *band = F.band + (size_t)s0 * 3, *ecost = F.entry_cost + s0;
        const int32_t *codes = F.codes + s0, *enes = F.entry_nes + s0;
        if (T <= 0) {
            if (lane == 0) {
                best_cost[u] = INF;
                best_end[u] = -1;
                if (path_len) path_len[u] = 0;
            }
            continue;
        }
        for (int r = lane; r < R; r += 32) prev[r] = INF;
        __syncwarp();
        for (int c = 0; c < T; ++c) {
            const float *Erow = E + (f0 + c) * ldE;
            uint8_t *bpc = bpu + (size_t)c * R;
            // ---- 1. emitting states from the previous column
            for (int r = lane; r < R; r += 32) {
                const int em = emit[r];
                if (em < 0) continue;
                if (c == 0 && r == 0) {
                    cur[0] = Erow[em];                               // decode.py:99-100
                    bpc[0] = 0;
                    continue;
                }
                const int cd = codes[r];
                float best = INF;
                int code = 0;
                const float t2 = (r >= 2 ? prev[r - 2] : INF) + band[r * 3 + 0];
                const float t1 = (r >= 1 ? prev[r - 1] : INF) + band[r * 3 + 1];
                const float t0 = prev[r] + band[r * 3 + 2];
                if (t2 < best) { best = t2; code = cd & 0xff; }
                if (t1 < best) { best = t1; code = (cd >> 8) & 0xff; }
                if (t0 < best) { best = t0; code = (cd >> 16) & 0xff; }
                cur[r] = enes[r] < 0 ? best + Erow[em] : best;       // entry states get their emission in sweep 3
                bpc[r] = (uint8_t)code;
            }
            __syncwarp();
            // ---- 2. the non-emitting states (same column, first minimum in origin order)
            for (int k = lane; k < K; k += 32) {
                const int n = G.nes_idx[n0 + k];
                if (c == 0 && n == 0) {
                    cur[0] = 0.f;
                    bpc[0] = 0;
                    continue;
                }
                const int p0 = pptr[n], p1 = pptr[n + 1];
                float best = INF;
                int bj = 0;
                for (int j = p0; j < p1; ++j) {
                    const int o = G.pred_idx[j];
                    const float cand = G.pred_cost[j] + (o < n ? cur[o] : INF);
                    if (cand < best) {
                        best = cand;
                        bj = j - p0;
                    }
                }
                cur[n] = (p1 > p0) ? best : INF;
                bpc[n] = (uint8_t)bj;
            }
            __syncwarp();
            // ---- 3. entry states: the NES arc competes with the band candidates (lower list position wins ties)
            for (int i = lane; i < Q; i += 32) {
                const int r = F.entry_state[q0 + i];
                if (c == 0 && r == 0) continue;
                const float cand = ecost[r] + cur[enes[r]];
                const int cn = (codes[r] >> 24) & 0xff;
                float best = cur[r];
                int code = bpc[r];
                if (cand < best || (cand == best && cand < INF && cn < code)) {
                    best = cand;
                    code = cn;
                }
                cur[r] = best + Erow[emit[r]];
                bpc[r] = (uint8_t)code;
            }
            __syncwarp();
            if (cu)
                for (int r = lane; r < R; r += 32) cu[(size_t)c * R + r] = cur[r];
            float *t = prev;
            prev = cur;
            cur = t;
            __syncwarp();
        }
        finish_utterance(G, g, s0, R, T, f0, u, prev, emit, pptr, bpu, lane, best_cost, best_end, path_off, path_state,
                         path_frame, path_len, frame_label);
        __syncwarp();
    }
}


// ------------------------------------------------------------------------------------------------ K2, banded, one graph
// The same three sweeps for a batch whose utterances all use ONE graph (connected-digit decoding, main.py:40-60): the
// graph's tables are packed once per CTA into shared memory -- per state {a2, a1, a0, meta}, meta = the four 2-bit
// back-pointer codes | entry flag << 8 | emits flag << 9 | emit << 14 (bits 12.. = byte offset of the emission in its
// row; 0 for a NES, so meta >> 9 != 0 iff the state emits) -- next to the NES predecessor lists and the entry
// list, so a state update is one 16-byte and three 4-byte shared loads, the emission gather and two stores instead of
// seven table loads with 64-bit address arithmetic.  The NES sweep gives every NES a group of lanes (first minimum
// kept across the group by (cost, position) order).  Outputs are bit-identical to viterbi_banded_kernel.
constexpr int VB_NES_CAP = 64;    // NES whose descriptor sits in shared memory (more fall back to the tables in global)
constexpr int VB_ECOLS = 128;     // widest emission row staged in shared memory one frame ahead (else gathered from global)

struct BandState { float a2, a1, a0; uint32_t meta; };
struct BandEntry { uint32_t rn; float cost; };                    // rn = state | entry NES << 16
struct BandNes { int32_t n, p0, cnt, q0; };                       // p0: absolute CSR offset; q0: offset in the shared copy or -1

// What a lane does in sweeps 2 and 3 does not depend on the frame: when every NES fits one pass (K * L <= 32, lists in
// shared memory) and the entries fit VB_QH per lane, the descriptors are read once into registers (HOIST).
constexpr int VB_QH = 3;
struct LaneWork {
    int n, q0, cnt;                 // this lane group's NES (cnt = 0: none), list at pidx_s / pcost_s + q0
    uint32_t ern[VB_QH];            // entries lane, lane+32, ..: state | NES << 16, or 0xffffffff
    float ecost[VB_QH];
    uint32_t emeta[VB_QH];          // entry code << 16 | emit column
};

// Erow: the frame's emission row, staged in shared memory by the caller or in global
template <bool FIRST, bool HOIST>
__device__ __forceinline__ void banded_column(const BandState *tab, const BandEntry *ent, const BandNes *nes,
                                              const int32_t *pidx_s, const float *pcost_s, int pcap, const GraphBatch &G,
                                              int R, int K, int Q, int L, const float *prev, float *cur, uint8_t *code_s,
                                              const float *__restrict__ Erow, uint8_t *__restrict__ bpc, int lane,
                                              const LaneWork &lw) {
    const float INF = CUDART_INF_F;
    // ---- 1. emitting states from the previous column (prev[-2], prev[-1] are +inf pads).  Straight-line body (the
    //         tables are padded to a multiple of 32 states with non-emitting dummies) so that unrolled iterations overlap.
    const int R32 = (R + 31) & ~31;
#pragma unroll 4
    for (int r = lane; r < R32; r += 32) {
        const BandState t = tab[r];
        const uint32_t meta = t.meta;
        const bool emits = (meta >> 9) != 0;                         // else a NES (sweep 2) or padding
        const float t2 = prev[r - 2] + t.a2, t1 = prev[r - 1] + t.a1, t0 = prev[r] + t.a0;
        float best = t2;                                             // first minimum in list order r-2, r-1, r; the
        uint32_t sh = 0;                                             // code is cut out of meta once, after the choice
        if (t1 < best) { best = t1; sh = 2; }
        if (t0 < best) { best = t0; sh = 4; }
        uint32_t code = best < INF ? (meta >> sh) & 3 : 0;
        const bool entry = (meta & 0x100) != 0;
        const float e = *reinterpret_cast<const float *>(reinterpret_cast<const char *>(Erow) + (meta >> 12));
        float v = entry ? best : best + e;                           // entry states get their emission in sweep 3
        if (FIRST && r == 0) {                                       // decode.py:99-100
            v = e;
            code = 0;
        }
        if (emits) {
            cur[r] = v;
            bpc[r] = (uint8_t)code;
        }
        if (entry) code_s[r] = (uint8_t)code;
    }
    // ---- 2. the non-emitting states: L lanes per NES, each scanning every L-th predecessor in list order
    if (HOIST) {
        const int sub = lane & (L - 1);
        float best = INF;
        int bj = 0x7fffffff;
        for (int j = sub; j < lw.cnt; j += L) {
            const float cand = pcost_s[lw.q0 + j] + cur[pidx_s[lw.q0 + j]];
            if (cand < best) {
                best = cand;
                bj = j;
            }
        }
        for (int d = L >> 1; d; d >>= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, best, d);
            const int oj = __shfl_xor_sync(0xffffffffu, bj, d);
            if (ob < best || (ob == best && oj < bj)) {
                best = ob;
                bj = oj;
            }
        }
        if (lw.n >= 0 && sub == 0) {
            if (FIRST && lw.n == 0) {
                cur[0] = 0.f;
                bpc[0] = 0;
            } else {
                cur[lw.n] = best;
                bpc[lw.n] = (uint8_t)(best < INF ? bj : 0);
            }
        }
        __syncwarp();
        // ---- 3. entry states (descriptors in registers)
#pragma unroll
        for (int h = 0; h < VB_QH; ++h) {
            const uint32_t rn = lw.ern[h];
            const int r = (int)(rn & 0xffffu), en = (int)(rn >> 16);
            if (rn == 0xffffffffu || (FIRST && r == 0)) continue;
            const float cand = lw.ecost[h] + cur[en];
            const uint32_t cn = lw.emeta[h] >> 16;
            float best3 = cur[r];
            uint32_t code = code_s[r];
            if (cand < best3 || (cand == best3 && cand < INF && cn < code)) {
                best3 = cand;
                code = cn;
            }
            cur[r] = best3 + Erow[lw.emeta[h] & 0xffffu];
            bpc[r] = (uint8_t)code;
        }
        __syncwarp();
        return;
    }
    for (int k0 = 0; k0 < K; k0 += 32 / L) {
        const int k = k0 + lane / L, sub = lane & (L - 1);
        float best = INF;
        int bj = 0x7fffffff, n = 0, cnt = 0;
        if (k < K) {
            int p0, q0 = -1;
            if (k < VB_NES_CAP) {
                const BandNes d = nes[k];
                n = d.n; p0 = d.p0; cnt = d.cnt; q0 = d.q0;
            } else {
                n = G.nes_idx[k];
                p0 = G.pred_ptr[n];
                cnt = G.pred_ptr[n + 1] - p0;
            }
            if (q0 >= 0) {                                           // predecessor list in shared memory
                for (int j = sub; j < cnt; j += L) {
                    const float cand = pcost_s[q0 + j] + cur[pidx_s[q0 + j]];
                    if (cand < best) {
                        best = cand;
                        bj = j;
                    }
                }
            } else {
                for (int j = sub; j < cnt; j += L) {
                    const float cand = G.pred_cost[p0 + j] + cur[G.pred_idx[p0 + j]];
                    if (cand < best) {
                        best = cand;
                        bj = j;
                    }
                }
            }
        }
        for (int d = L >> 1; d; d >>= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, best, d);
            const int oj = __shfl_xor_sync(0xffffffffu, bj, d);
            if (ob < best || (ob == best && oj < bj)) {
                best = ob;
                bj = oj;
            }
        }
        if (k < K && sub == 0) {
            if (FIRST && n == 0) {
                cur[0] = 0.f;
                bpc[0] = 0;
            } else {
                cur[n] = best;
                bpc[n] = (uint8_t)(best < INF ? bj : 0);
            }
        }
    }
    __syncwarp();
    // ---- 3. entry states: the NES arc competes with the band candidates (lower list position wins ties)
    for (int i = lane; i < Q; i += 32) {
        const BandEntry e = ent[i];
        const int r = (int)(e.rn & 0xffffu), en = (int)(e.rn >> 16);
        if (FIRST && r == 0) continue;
        const uint32_t meta = tab[r].meta;
        const float cand = e.cost + cur[en];
        const uint32_t cn = (meta >> 6) & 3;
        float best = cur[r];
        uint32_t code = code_s[r];
        if (cand < best || (cand == best && cand < INF && cn < code)) {
            best = cand;
            code = cn;
        }
        cur[r] = best + Erow[meta >> 14];
        bpc[r] = (uint8_t)code;
    }
    __syncwarp();
}

// finish_utterance for the one-graph kernel: the walk back (decode.py:142-146) is done by the whole warp on a register
// window of back-pointers -- 16 frames x 16 states below the current (frame, state), 8 bytes per lane, fetched with one
// round trip and read with shuffles -- because the serial walk otherwise pays one DRAM latency per frame (an
// utterance's back-pointers have long left L2 when its last column is done).  A band step is decoded from the packed
// table (codes -> r-2 / r-1 / r); NES hops (7 per connected-digit string) go through the predecessor lists.
constexpr int VB_WROWS = 16, VB_WCOLS = 16;

__device__ __forceinline__ void finish_banded1(const GraphBatch &G, const BandState *tab, int R, int T, int64_t f0, int u,
                                               const float *prev, const uint8_t *bpu, int lane,
                                               float *best_cost, int32_t *best_end, const int64_t *path_off,
                                               int32_t *path_state, int32_t *path_frame, int32_t *path_len,
                                               int32_t *frame_label) {
    const float INF = CUDART_INF_F;
    float bc = INF;
    int bi = -1;
    for (int e = G.g_end_off[0]; e < G.g_end_off[1]; ++e) {          // '>=': the last listed minimum wins (decode.py:131-134)
        const int s = G.end_state[e];
        if (bc >= prev[s]) {
            bc = prev[s];
            bi = s;
        }
    }
    if (lane == 0) {
        best_cost[u] = bc;
        best_end[u] = bi;
    }
    int len = 0;
    if (bc < INF && (path_state || frame_label)) {
        int32_t *ps = path_state ? path_state + path_off[u] : nullptr;
        int32_t *pf = path_state ? path_frame + path_off[u] : nullptr;
        const int cap = path_state ? (int)(path_off[u + 1] - path_off[u]) : 0x7fffffff;
        int i = bi, j = T - 1;
        if (lane == 0 && frame_label && (tab[i].meta >> 9) != 0) frame_label[f0 + j] = i;
        int ja = -1, ia = 0;                                         // window: frames ja-15 .. ja, states ia-15 .. ia
        uint32_t w0 = 0, w1 = 0;                                     // lane: frames ja - (lane>>2) [- 8], states ia-15+4*(lane&3) ..+3
        while (j > 0 && len < cap) {
            if (j > ja || j <= ja - VB_WROWS || i > ia || i <= ia - VB_WCOLS) {
                ja = j;
                ia = i;
                const int sbase = ia - (VB_WCOLS - 1) + 4 * (lane & 3);
                uint32_t b[8];
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int row = ja - (lane >> 2) - 8 * h;
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int st = sbase + q;
                        b[h * 4 + q] = (row >= 0 && st >= 0) ? (uint32_t)bpu[(size_t)row * R + st] : 0u;
                    }
                }
                w0 = b[0] | (b[1] << 8) | (b[2] << 16) | (b[3] << 24);
                w1 = b[4] | (b[5] << 8) | (b[6] << 16) | (b[7] << 24);
                // pull the next window's sectors towards L2 while this one is walked
                const int prow = ja - VB_WROWS - (lane & 15), pst = ia - ((lane >> 4) ? 0 : 2 * VB_WCOLS - 1);
                if (prow >= 0 && pst >= 0)
                    asm volatile("prefetch.global.L2 [%0];" ::"l"(bpu + (size_t)prow * R + pst));
            }
            const int dr = ja - j, ds = i - (ia - (VB_WCOLS - 1));
            const uint32_t wsel = (dr & 8) ? w1 : w0;
            const uint32_t word = __shfl_sync(0xffffffffu, wsel, ((dr & 7) << 2) | (ds >> 2));
            const int code = (int)((word >> (8 * (ds & 3))) & 0xffu);
            const BandState t = tab[i];
            const bool i_emits = (t.meta >> 9) != 0;
            int o;
            if (i_emits && t.a2 < INF && code == (int)(t.meta & 3)) o = i - 2;
            else if (i_emits && t.a1 < INF && code == (int)((t.meta >> 2) & 3)) o = i - 1;
            else if (i_emits && t.a0 < INF && code == (int)((t.meta >> 4) & 3)) o = i;
            else o = G.pred_idx[G.pred_ptr[i] + code];               // entry arc of a word, or i is a NES
            const bool o_emits = (tab[o].meta >> 9) != 0;
            if (i_emits && o_emits) --j;
            i = o;
            if (lane == 0) {
                if (ps) {
                    ps[len] = i;
                    pf[len] = j;
                }
                if (frame_label && o_emits) frame_label[f0 + j] = i;
            }
            ++len;
        }
        if (j > 0) len = -1;   // capacity exhausted
    } else if (frame_label) {
        for (int c = lane; c < T; c += 32) frame_label[f0 + c] = -1;   // no path: frames are skipped downstream
    }
    if (lane == 0 && path_len) path_len[u] = len;
}

// All columns of one utterance; leaves the last column in prev (by reference).  FAST: emission rows staged and
// sweeps 2 / 3 on register descriptors.
template <bool FAST>
__device__ __forceinline__ void banded1_forward(const BandState *tab, const BandEntry *ent, const BandNes *nes,
                                                const int32_t *pidx_s, const float *pcost_s, int pcap, const GraphBatch &G,
                                                int R, int K, int Q, int L, float *&prev, float *&cur, uint8_t *code_s,
                                                float *erow_s, int ecols, bool staged, const float *__restrict__ Eg,
                                                int64_t ldE, int T, uint8_t *__restrict__ bpc, float *cu, int lane,
                                                const LaneWork &lw) {
    float en[VB_ECOLS / 32];                                         // row c+1 in flight while column c is relaxed
    if (staged) {
#pragma unroll
        for (int i = 0; i < VB_ECOLS / 32; ++i)
            if (32 * i < ecols) erow_s[lane + 32 * i] = lane + 32 * i < ecols ? Eg[lane + 32 * i] : 0.f;
        __syncwarp();
    }
    for (int c = 0; c < T; ++c) {
        if (staged && c + 1 < T) {
#pragma unroll
            for (int i = 0; i < VB_ECOLS / 32; ++i)
                if (32 * i < ecols) en[i] = lane + 32 * i < ecols ? Eg[ldE + lane + 32 * i] : 0.f;
        }
        if (c == 0)
            banded_column<true, false>(tab, ent, nes, pidx_s, pcost_s, pcap, G, R, K, Q, L, prev, cur, code_s,
                                       staged ? erow_s : Eg, bpc, lane, lw);
        else if (FAST)
            banded_column<false, true>(tab, ent, nes, pidx_s, pcost_s, pcap, G, R, K, Q, L, prev, cur, code_s, erow_s, bpc,
                                       lane, lw);
        else
            banded_column<false, false>(tab, ent, nes, pidx_s, pcost_s, pcap, G, R, K, Q, L, prev, cur, code_s,
                                        staged ? erow_s : Eg, bpc, lane, lw);
        if (cu)
            for (int r = lane; r < R; r += 32) cu[(size_t)c * R + r] = cur[r];
        if (staged && c + 1 < T) {                                   // every lane is past its reads of row c
#pragma unroll
            for (int i = 0; i < VB_ECOLS / 32; ++i)
                if (32 * i < ecols) erow_s[lane + 32 * i] = en[i];
            __syncwarp();
        }
        float *t = prev;
        prev = cur;
        cur = t;
        Eg += ldE;
        bpc += R;
    }
}

__global__ void __launch_bounds__(VW * 32, 7)
viterbi_banded1_kernel(const float *__restrict__ E, int64_t ldE, const int32_t *__restrict__ off, int B, GraphBatch G,
                       BandForm F, int Rmax, const int64_t *__restrict__ bp_off, uint8_t *__restrict__ bp,
                       float *__restrict__ costs_out, float *__restrict__ best_cost, int32_t *__restrict__ best_end,
                       const int64_t *__restrict__ path_off, int32_t *__restrict__ path_state,
                       int32_t *__restrict__ path_frame, int32_t *__restrict__ path_len,
                       int32_t *__restrict__ frame_label) {
    extern __shared__ __align__(16) unsigned char vb_sm[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, tid = threadIdx.x;
    const int R = G.g_state_off[1] - G.g_state_off[0], K = G.g_nes_off[1] - G.g_nes_off[0];
    const int Q = F.g_entry_off[1] - F.g_entry_off[0];
    const int RP = (Rmax + 31) & ~31, pcap = RP;              // tables padded to whole 32-state sweeps
    BandState *tab = reinterpret_cast<BandState *>(vb_sm);                       // RP
    BandNes *nes = reinterpret_cast<BandNes *>(tab + RP);                        // VB_NES_CAP
    BandEntry *ent = reinterpret_cast<BandEntry *>(nes + VB_NES_CAP);            // RP
    int32_t *pidx_s = reinterpret_cast<int32_t *>(ent + RP);                     // pcap
    float *pcost_s = reinterpret_cast<float *>(pidx_s + pcap);                   // pcap
    float *colbase = pcost_s + pcap;                                             // VW * 2 * (RP + 4)
    float *erowbase = colbase + (size_t)VW * 2 * (RP + 4);                      // VW * VB_ECOLS
    uint8_t *codebase = reinterpret_cast<uint8_t *>(erowbase + VW * VB_ECOLS);   // VW * RP
    const float INF = CUDART_INF_F;
    __shared__ int ecols_s;
    if (tid == 0) ecols_s = 0;
    __syncthreads();

    int emax = 0;
    for (int r = tid; r < R; r += VW * 32) {
        const uint32_t cd = (uint32_t)F.codes[r];
        const int em = G.emit[r];
        emax = max(emax, em + 1);
        BandState t;
        t.a2 = F.band[r * 3 + 0];
        t.a1 = F.band[r * 3 + 1];
        t.a0 = F.band[r * 3 + 2];
        t.meta = (cd & 3) | (((cd >> 8) & 3) << 2) | (((cd >> 16) & 3) << 4) | (((cd >> 24) & 3) << 6) |
                 (F.entry_nes[r] >= 0 ? 0x100u : 0u) | (em >= 0 ? 0x200u | ((uint32_t)em << 14) : 0u);
        tab[r] = t;
    }
    for (int i = tid; i < Q; i += VW * 32) {
        const int r = F.entry_state[i];
        BandEntry e;
        e.rn = (uint32_t)r | ((uint32_t)F.entry_nes[r] << 16);
        e.cost = F.entry_cost[r];
        ent[i] = e;
    }
    if (tid == 0) {                                                  // NES descriptors; their predecessor lists are packed
        int q = 0;                                                   // back to back into the shared copy while they fit
        for (int k = 0; k < min(K, VB_NES_CAP); ++k) {
            const int n = G.nes_idx[k];
            BandNes d;
            d.n = n;
            d.p0 = G.pred_ptr[n];
            d.cnt = G.pred_ptr[n + 1] - d.p0;
            d.q0 = q + d.cnt <= pcap ? q : -1;
            if (d.q0 >= 0) q += d.cnt;
            nes[k] = d;
        }
    }
    __syncthreads();
    for (int k = warp; k < min(K, VB_NES_CAP); k += VW) {
        const BandNes d = nes[k];
        if (d.q0 < 0) continue;
        for (int j = lane; j < d.cnt; j += 32) {
            pidx_s[d.q0 + j] = G.pred_idx[d.p0 + j];
            pcost_s[d.q0 + j] = G.pred_cost[d.p0 + j];
        }
    }
    for (int r = R + tid; r < RP; r += VW * 32) {                   // padding: never written, never emitting
        BandState t;
        t.a2 = t.a1 = t.a0 = INF;
        t.meta = 0;
        tab[r] = t;
    }
    atomicMax(&ecols_s, emax);
    __syncthreads();
    const int ecols = ecols_s;                                       // columns of E the graph reads
    const bool staged = ecols <= VB_ECOLS;
    float *erow_s = erowbase + warp * VB_ECOLS;
    int L = 1;                                                       // lanes per NES in sweep 2
    while (L < 32 && 2 * L * K <= 32) L <<= 1;

    // frame-invariant work of this lane in sweeps 2 and 3
    LaneWork lw;
    bool fast = staged && K * L <= 32 && K <= VB_NES_CAP && Q <= 32 * VB_QH;
    lw.n = -1;
    lw.q0 = lw.cnt = 0;
    if (fast && lane / L < K) {
        const BandNes d = nes[lane / L];
        lw.n = d.n;
        lw.q0 = d.q0;
        lw.cnt = d.cnt;
    }
    fast = fast && __all_sync(0xffffffffu, lw.q0 >= 0);
#pragma unroll
    for (int h = 0; h < VB_QH; ++h) {
        lw.ern[h] = 0xffffffffu;
        lw.ecost[h] = 0.f;
        lw.emeta[h] = 0;
        if (fast && lane + 32 * h < Q) {
            const BandEntry e = ent[lane + 32 * h];
            const uint32_t meta = tab[e.rn & 0xffffu].meta;
            lw.ern[h] = e.rn;
            lw.ecost[h] = e.cost;
            lw.emeta[h] = (((meta >> 6) & 3) << 16) | (meta >> 14);
        }
    }

    float *prev = colbase + (size_t)warp * 2 * (RP + 4) + 4, *cur = prev + RP + 4;
    uint8_t *code_s = codebase + (size_t)warp * RP;
    for (int u = blockIdx.x * VW + warp; u < B; u += gridDim.x * VW) {
        const int64_t f0 = off[u];
        const int T = (int)(off[u + 1] - f0);
        uint8_t *bpu = bp + bp_off[u];
        float *cu = costs_out ? costs_out + bp_off[u] : nullptr;
        if (T <= 0) {
            if (lane == 0) {
                best_cost[u] = INF;
                best_end[u] = -1;
                if (path_len) path_len[u] = 0;
            }
            continue;
        }
        for (int r = lane - 4; r < RP; r += 32) prev[r] = cur[r] = INF;   // incl. the pads in front of and behind both columns
        __syncwarp();
        if (fast)
            banded1_forward<true>(tab, ent, nes, pidx_s, pcost_s, pcap, G, R, K, Q, L, prev, cur, code_s, erow_s, ecols, true,
                                  E + f0 * ldE, ldE, T, bpu, cu, lane, lw);
        else
            banded1_forward<false>(tab, ent, nes, pidx_s, pcost_s, pcap, G, R, K, Q, L, prev, cur, code_s, erow_s, ecols,
                                   staged, E + f0 * ldE, ldE, T, bpu, cu, lane, lw);
        finish_banded1(G, tab, R, T, f0, u, prev, bpu, lane, best_cost, best_end, path_off, path_state, path_frame,
                       path_len, frame_label);
        __syncwarp();
    }
}

static size_t banded1_smem_bytes(int Rmax) {
    const size_t RP = (size_t)((Rmax + 31) & ~31);
    return RP * (sizeof(BandState) + sizeof(BandEntry) + 8) + VB_NES_CAP * sizeof(BandNes) + (size_t)VW * 2 * (RP + 4) * 4 +
           (size_t)VW * VB_ECOLS * 4 + (size_t)VW * RP;
}


// ------------------------------------------------------------------------------------------------ K4
// Scaled log-domain forward-backward for NES-free graphs.  Boundary conditions are those of the Viterbi
// above (start in state 0 at frame 0, end in one of the graph's end states at the last frame); replacing the
// log-sum-exp by min gives exactly the Viterbi costs (tests tie the two).  Every column is renormalised by
// its minimum (alpha_hat = alpha - min), the offsets summed in double, so posteriors keep FP32 relative
// accuracy for paths whose cumulative cost reaches 1e4:
//     gamma_c(r) = exp(-(alpha_hat_c(r) + beta_hat_c(r))) / sum_r(...)
//     xi_c(o->r) = exp(-(alpha_hat_{c-1}(o) + tr + E_c(r) + beta_hat_c(r))) / sum_edges(...)
// exp / log are the fast intrinsics (ex2.approx / lg2.approx): measured on the B200, the IEEE-accurate functions changed
// no digit of gamma / xi (the error that mattered came from adding un-renormalised emission costs, fixed above) and
// made the banded kernel three times slower.
struct OnlineNegLse {   // accumulates -log(sum exp(-x_i))
    float m, s;
    __device__ __forceinline__ void reset() { m = __int_as_float(0x7f800000); s = 0.f; }
    __device__ __forceinline__ void add(float x) {
        if (x == CUDART_INF_F) return;
        if (x < m) {
            s = s * __expf(x - m) + 1.f;   // m == inf: s == 0 and exp(-inf) == 0
            m = x;
        } else {
            s += __expf(m - x);
        }
    }
    __device__ __forceinline__ float value() const { return s > 0.f ? m - __logf(s) : CUDART_INF_F; }
};

__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
    for (int d = 16; d; d >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, d));
    return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int d = 16; d; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    return v;
}

// smallest emission cost of a column over the graph's states (0 when every state is impossible)
__device__ __forceinline__ float column_emin(const float *Erow, const int32_t *emit, int R, int lane) {
    float m = CUDART_INF_F;
    for (int r = lane; r < R; r += 32) m = fminf(m, Erow[emit[r]]);
    m = warp_min(m);
    return m < CUDART_INF_F ? m : 0.f;
}

__global__ void __launch_bounds__(VW * 32)
forward_backward_kernel(const float *__restrict__ E, int64_t ldE, const int32_t *__restrict__ off, int B,
                        GraphBatch G, const int32_t *__restrict__ succ_ptr, const int32_t *__restrict__ succ_idx,
                        const float *__restrict__ succ_cost, const int32_t *__restrict__ pred_tgt, int Rmax,
                        int NNZmax, const int64_t *__restrict__ ws_off, float *__restrict__ gamma,
                        double *__restrict__ neg_log_z, double *__restrict__ xi_part, int nnz_total) {
    extern __shared__ __align__(8) float vsm[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    double *xacc = reinterpret_cast<double *>(vsm) + (size_t)warp * NNZmax;   // this utterance's edge posteriors
    float *colA = vsm + (size_t)VW * NNZmax * 2 + (size_t)warp * (2 * Rmax + NNZmax), *colB = colA + Rmax,
          *xi = colB + Rmax;
    const float INF = CUDART_INF_F;
    // edge posteriors without atomics: every warp owns one row of xi_part (zeroed here, summed over the utterances the
    // warp walks, reduced over the rows in a fixed order by xi_reduce_kernel) -- the result does not depend on timing
    double *wrow = xi_part ? xi_part + (size_t)(blockIdx.x * VW + warp) * nnz_total : nullptr;
    if (wrow) {
        for (int j = lane; j < nnz_total; j += 32) wrow[j] = 0.0;
        __syncwarp();
    }

    for (int u = blockIdx.x * VW + warp; u < B; u += gridDim.x * VW) {
        const int g = G.utt_graph ? G.utt_graph[u] : 0;
        const int s0 = G.g_state_off[g], R = G.g_state_off[g + 1] - s0;
        const int64_t f0 = off[u];
        const int T = (int)(off[u + 1] - f0);
        const int32_t *emit = G.emit + s0, *pptr = G.pred_ptr + s0, *sptr = succ_ptr + s0;
        const int e0 = pptr[0], nnz = pptr[R] - e0;
        float *gu = gamma + ws_off[u];
        if (T <= 0) {
            if (lane == 0 && neg_log_z) neg_log_z[u] = (double)INF;
            continue;
        }
        // ---- forward
        float *prev = colA, *cur = colB;
        double scale = 0.0;
        for (int c = 0; c < T; ++c) {
            const float *Erow = E + (f0 + c) * ldE;
            // the column's smallest emission cost is taken out BEFORE the costs are added: emission costs are tens of
            // nats, adding one to a renormalised alpha (a few nats) and subtracting the minimum afterwards would round
            // at the emission's magnitude in every column (measured: 7e-5 relative on gamma after 300 frames)
            const float emin = column_emin(Erow, emit, R, lane);
            float mymin = INF;
            for (int r = lane; r < R; r += 32) {
                float a;
                if (c == 0) {
                    a = (r == 0) ? Erow[emit[0]] - emin : INF;
                } else {
                    OnlineNegLse acc; acc.reset();
                    for (int j = pptr[r]; j < pptr[r + 1]; ++j) acc.add(G.pred_cost[j] + prev[G.pred_idx[j]]);
                    a = acc.value() + (Erow[emit[r]] - emin);
                }
                cur[r] = a;
                mymin = fminf(mymin, a);
            }
            const float m = warp_min(mymin);
            const float sub = (m < INF) ? m : 0.f;
            scale += (double)sub + (double)emin;
            __syncwarp();
            for (int r = lane; r < R; r += 32) {
                const float a = cur[r] - sub;
                cur[r] = a;
                gu[(size_t)c * R + r] = a;           // alpha_hat, overwritten by gamma in the backward sweep
            }
            __syncwarp();
            float *t = prev;
            prev = cur;
            cur = t;
        }
        // log Z over the end states at the last column (prev = alpha_hat_{T-1})
        OnlineNegLse zacc; zacc.reset();
        for (int e = G.g_end_off[g] + lane; e < G.g_end_off[g + 1]; e += 32) zacc.add(prev[G.end_state[e]]);
        float zl = zacc.value();
        {   // merge the per-lane partial sums: -lse over lanes
            const float zm = warp_min(zl);
            const float zs = warp_sum(zl < INF ? __expf(zm - zl) : 0.f);
            zl = (zm < INF) ? zm - __logf(zs) : INF;
        }
        if (lane == 0 && neg_log_z) neg_log_z[u] = (zl < INF) ? scale + (double)zl : (double)INF;
        // ---- backward; colA/colB now hold beta_hat columns
        float *bnext = colA, *bcur = colB;
        for (int j = lane; j < nnz; j += 32) {
            xi[j] = 0.f;
            xacc[j] = 0.0;
        }
        for (int r = lane; r < R; r += 32) bnext[r] = INF;
        __syncwarp();
        for (int e = G.g_end_off[g] + lane; e < G.g_end_off[g + 1]; e += 32) bnext[G.end_state[e]] = 0.f;
        __syncwarp();
        for (int c = T - 1; c >= 0; --c) {
            // gamma of column c from alpha_hat_c (global) and beta_hat_c (bnext)
            float gs = 0.f;
            for (int r = lane; r < R; r += 32) {
                const float v = gu[(size_t)c * R + r] + bnext[r];
                const float p = (v < INF) ? __expf(-v) : 0.f;
                bcur[r] = p;                          // scratch
                gs += p;
            }
            gs = warp_sum(gs);
            const float inv = gs > 0.f ? 1.f / gs : 0.f;
            __syncwarp();
            if (c > 0) {
                // xi of the transitions into column c (needs alpha_hat_{c-1}, still in global)
                const float *Erow = E + (f0 + c) * ldE;
                const float emin = column_emin(Erow, emit, R, lane);
                float xs = 0.f, vmin = INF;
                for (int j = lane; j < nnz; j += 32) {
                    const int r = pred_tgt[e0 + j];
                    const int o = G.pred_idx[e0 + j];
                    const float v = gu[(size_t)(c - 1) * R + o] + G.pred_cost[e0 + j] + (Erow[emit[r]] - emin) + bnext[r];
                    xi[j] = v;
                    vmin = fminf(vmin, v);
                }
                vmin = warp_min(vmin);                 // E_c is not renormalised: exp(-v) alone would underflow
                for (int j = lane; j < nnz; j += 32) {
                    const float v = xi[j];
                    const float p = (v < INF) ? __expf(vmin - v) : 0.f;
                    xs += p;
                    xi[j] = p;                         // unnormalised, this column only (added below)
                }
                xs = warp_sum(xs);
                const float xinv = xs > 0.f ? 1.f / xs : 0.f;
                if (wrow)
                    for (int j = lane; j < nnz; j += 32) xacc[j] += (double)(xi[j] * xinv);   // lane-private entries
            }
            for (int r = lane; r < R; r += 32) gu[(size_t)c * R + r] = bcur[r] * inv;
            __syncwarp();
            if (c == 0) break;
            // beta_hat_{c-1}(o) = -lse over successors r of (tr + E_c(r) + beta_hat_c(r)), renormalised
            const float *Erow = E + (f0 + c) * ldE;
            const float emin = column_emin(Erow, emit, R, lane);
            float mymin = INF;
            for (int o = lane; o < R; o += 32) {
                OnlineNegLse acc; acc.reset();
                for (int j = sptr[o]; j < sptr[o + 1]; ++j) {
                    const int r = succ_idx[j];
                    acc.add(succ_cost[j] + (Erow[emit[r]] - emin) + bnext[r]);
                }
                const float b = acc.value();
                bcur[o] = b;
                mymin = fminf(mymin, b);
            }
            const float m = warp_min(mymin);
            const float sub = (m < INF) ? m : 0.f;
            __syncwarp();
            for (int o = lane; o < R; o += 32) bcur[o] -= sub;
            __syncwarp();
            float *t = bnext;
            bnext = bcur;
            bcur = t;
        }
        if (wrow)
            for (int j = lane; j < nnz; j += 32) wrow[e0 + j] += xacc[j];
        __syncwarp();
    }
}

// xi_sum[j] += sum over the per-warp rows of xi_part, in a FIXED order: thread t of block j adds rows t, t + 256, ...
// in ascending order, then the 256 partial sums are folded by a fixed-shape tree.
__global__ void __launch_bounds__(256)
xi_reduce_kernel(const double *__restrict__ part, int nparts, int nnz, double *__restrict__ xi_sum) {
    __shared__ double red[256];
    const int j = blockIdx.x, t = threadIdx.x;
    double s = 0.0;
    for (int p = t; p < nparts; p += 256) s += part[(size_t)p * nnz + j];
    red[t] = s;
    __syncthreads();
#pragma unroll
    for (int d = 128; d; d >>= 1) {
        if (t < d) red[t] += red[t + d];
        __syncthreads();
    }
    if (t == 0) xi_sum[j] += red[0];
}

// ------------------------------------------------------------------------------------------------ K4, banded
// The same recursion for left-to-right models (every predecessor of r inside {r-2, r-1, r}, no NES, at most 32 states:
// what HMM.fit / calc_transition_costs produce, kmeans.py:53-95).  A group of W = pow2 >= Rmax lanes owns one utterance
// (32 / W utterances per warp), a lane owns one state: alpha / beta live in a register, the band costs and the CSR
// positions of the three arcs are read once per utterance, predecessors / successors arrive by width-W shuffles, the
// emission and alpha_hat loads run one column ahead, and the edge posteriors are summed in registers and added once per
// UTTERANCE to the warp's private row of xi_part (no atomics: the groups of a warp take turns, rows are reduced in a
// fixed order by xi_reduce_kernel).  `order` (optional) lists the utterances by falling length so that the
// groups of a warp finish together.  Same boundary conditions, renormalisation and outputs as forward_backward_kernel.
template <int W>
__device__ __forceinline__ float group_min(float v) {
#pragma unroll
    for (int d = W >> 1; d; d >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, d));
    return v;
}
template <int W>
__device__ __forceinline__ float group_sum(float v) {
#pragma unroll
    for (int d = W >> 1; d; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    return v;
}
// -log(exp(-a) + exp(-b) + exp(-c)), +inf entries ignored
__device__ __forceinline__ float neg_lse3(float a, float b, float c) {
    const float m = fminf(a, fminf(b, c));
    if (!(m < CUDART_INF_F)) return CUDART_INF_F;
    return m - __logf(__expf(m - a) + __expf(m - b) + __expf(m - c));
}

template <int W>
__global__ void __launch_bounds__(VW * 32)
forward_backward_banded_kernel(const float *__restrict__ E, int64_t ldE, const int32_t *__restrict__ off, int B,
                               GraphBatch G, const float *__restrict__ band, const int32_t *__restrict__ codes,
                               const int32_t *__restrict__ order, const int64_t *__restrict__ ws_off,
                               float *__restrict__ gamma, double *__restrict__ neg_log_z, double *__restrict__ xi_part,
                               int nnz_total) {
    constexpr int GPW = 32 / W;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, sub = lane & (W - 1), grp = lane / W;
    const float INF = CUDART_INF_F;
    const int nslots = (B + GPW - 1) / GPW;
    double *wrow = xi_part ? xi_part + (size_t)(blockIdx.x * VW + warp) * nnz_total : nullptr;
    if (wrow) {
        for (int j = lane; j < nnz_total; j += 32) wrow[j] = 0.0;
        __syncwarp();
    }

    for (int slot = blockIdx.x * VW + warp; slot < nslots; slot += gridDim.x * VW) {
        const int idx = slot * GPW + grp;
        const bool valid = idx < B;
        const int u = valid ? (order ? order[idx] : idx) : 0;
        const int g = (valid && G.utt_graph) ? G.utt_graph[u] : 0;
        const int s0 = G.g_state_off[g], R = G.g_state_off[g + 1] - s0;
        const int64_t f0 = off[u];
        const int T = valid ? max((int)(off[u + 1] - f0), 0) : 0;
        const bool act = valid && sub < R;                           // this lane owns a state
        float a2 = INF, a1 = INF, a0 = INF;
        int em = 0, arc2 = 0, arc1 = 0, arc0 = 0;
        bool is_end = false;
        if (act) {
            a2 = band[(size_t)(s0 + sub) * 3 + 0];
            a1 = band[(size_t)(s0 + sub) * 3 + 1];
            a0 = band[(size_t)(s0 + sub) * 3 + 2];
            const uint32_t cd = (uint32_t)codes[s0 + sub];
            const int p0 = G.pred_ptr[s0 + sub];
            arc2 = p0 + (int)(cd & 0xffu);
            arc1 = p0 + (int)((cd >> 8) & 0xffu);
            arc0 = p0 + (int)((cd >> 16) & 0xffu);
            em = G.emit[s0 + sub];
            for (int e = G.g_end_off[g]; e < G.g_end_off[g + 1]; ++e) is_end |= (G.end_state[e] == sub);
        }
        if (sub < 2) a2 = INF;                                       // no state r-2 / r-1 in this group
        if (sub < 1) a1 = INF;
        int Tmax = T;
#pragma unroll
        for (int d = 16; d >= W; d >>= 1) Tmax = max(Tmax, __shfl_xor_sync(0xffffffffu, Tmax, d));
        if (valid && T == 0 && sub == 0 && neg_log_z) neg_log_z[u] = (double)INF;
        if (Tmax == 0) continue;
        const float *Eu = E + f0 * ldE + em;
        float *gu = gamma + (valid ? ws_off[u] : 0) + sub;

        // ---- forward: alpha_hat columns to gamma
        // Emission costs are tens of nats while the renormalised alpha / beta are a few: the column's smallest emission
        // cost is taken out BEFORE the costs are added, so no sum ever rounds at the emission's magnitude (measured:
        // gamma 7e-5 relative after 300 frames without this, 3e-6 with it; profiles/r02_fb_error.json).
        float alpha = INF;
        double scale = 0.0;
        float e_c = (act && 0 < T) ? Eu[0] : INF;
        for (int c = 0; c < Tmax; ++c) {
            const bool live = c < T;
            float e_n = INF;
            if (act && c + 1 < T) e_n = Eu[(int64_t)(c + 1) * ldE];
            float emin = group_min<W>(e_c);
            if (!(emin < INF)) emin = 0.f;
            const float eh = e_c - emin;
            const float p1 = __shfl_up_sync(0xffffffffu, alpha, 1, W), p2 = __shfl_up_sync(0xffffffffu, alpha, 2, W);
            float a;
            if (c == 0) a = (sub == 0) ? eh : INF;
            else a = neg_lse3(p2 + a2, p1 + a1, alpha + a0) + eh;
            if (!act) a = INF;
            const float m = group_min<W>(a);
            const float sh = (m < INF) ? m : 0.f;
            if (live) {
                scale += (double)sh + (double)emin;
                alpha = a - sh;
                if (act) gu[(size_t)c * R] = alpha;
            }
            e_c = e_n;
        }
        {   // log Z over the end states of the last column
            const float v = (is_end && T > 0) ? alpha : INF;
            const float zm = group_min<W>(v);
            const float zs = group_sum<W>(v < INF ? __expf(zm - v) : 0.f);
            const float zl = (zm < INF) ? zm - __logf(zs) : INF;
            if (valid && T > 0 && sub == 0 && neg_log_z) neg_log_z[u] = (zl < INF) ? scale + (double)zl : (double)INF;
        }
        // ---- backward: gamma over alpha_hat, edge posteriors into registers
        float beta = is_end ? 0.f : INF;
        double x2 = 0.0, x1 = 0.0, x0 = 0.0;
        int c = Tmax - 1;
        float al_c = (act && c < T) ? gu[(size_t)c * R] : INF;
        float al_p = (act && c >= 1 && c - 1 < T) ? gu[(size_t)(c - 1) * R] : INF;
        e_c = (act && c < T) ? Eu[(int64_t)c * ldE] : INF;
        for (; c >= 0; --c) {
            const bool live = c < T;
            float al_n = INF, e_n = INF;                             // alpha_hat_{c-2}, E_{c-1}
            if (act && c >= 2 && c - 2 < T) al_n = gu[(size_t)(c - 2) * R];
            if (act && c >= 1 && c - 1 < T) e_n = Eu[(int64_t)(c - 1) * ldE];
            // gamma_c
            const float v = al_c + beta;
            const float p = (v < INF) ? __expf(-v) : 0.f;
            const float gs = group_sum<W>(p);
            if (live && act) gu[(size_t)c * R] = p * (gs > 0.f ? 1.f / gs : 0.f);
            // xi of the arcs into column c, beta_hat_{c-1}
            float emin = group_min<W>(e_c);
            if (!(emin < INF)) emin = 0.f;
            const float w = (e_c - emin) + beta;                     // E_hat_c(r) + beta_hat_c(r)
            const float q1 = __shfl_up_sync(0xffffffffu, al_p, 1, W), q2 = __shfl_up_sync(0xffffffffu, al_p, 2, W);
            const float y2 = a2 + w, y1 = a1 + w, y0 = a0 + w;
            const float v2 = q2 + y2, v1 = q1 + y1, v0 = al_p + y0;
            const float vmin = group_min<W>(fminf(v2, fminf(v1, v0)));
            const float r2 = (v2 < INF) ? __expf(vmin - v2) : 0.f;
            const float r1 = (v1 < INF) ? __expf(vmin - v1) : 0.f;
            const float r0 = (v0 < INF) ? __expf(vmin - v0) : 0.f;
            const float xs = group_sum<W>(r2 + r1 + r0);
            const float xinv = xs > 0.f ? 1.f / xs : 0.f;
            float s1 = __shfl_down_sync(0xffffffffu, y1, 1, W), s2 = __shfl_down_sync(0xffffffffu, y2, 2, W);
            if (sub + 1 >= W) s1 = INF;
            if (sub + 2 >= W) s2 = INF;
            const float b = neg_lse3(y0, s1, s2);
            const float bm = group_min<W>(b);
            if (live && c > 0) {
                x2 += (double)(r2 * xinv);
                x1 += (double)(r1 * xinv);
                x0 += (double)(r0 * xinv);
                beta = b - ((bm < INF) ? bm : 0.f);
            }
            al_c = al_p;
            al_p = al_n;
            e_c = e_n;
        }
        if (wrow) {
            // the groups of a warp may sit on the same word (same arcs): they add in turn; inside a group every lane
            // owns its state's three arcs
#pragma unroll
            for (int gi = 0; gi < GPW; ++gi) {
                if (grp == gi && act && T > 0) {
                    if (a2 < INF) wrow[arc2] += x2;
                    if (a1 < INF) wrow[arc1] += x1;
                    if (a0 < INF) wrow[arc0] += x0;
                }
                __syncwarp();
            }
        }
    }
}


// ------------------------------------------------------------------------------------------------ path -> words
// main.py:59-67 after the decoder: the reversed path read forwards, consecutive duplicate states removed, split at the
// non-emitting states, the first state of every piece mapped to its word.  One warp per utterance: lane i looks at
// forward position 32c + i; a word starts where the state changes to an emitting one whose predecessor is a NES (or
// at position 0); ballot + popc give the output slot, so the words come out in order without a second pass.
__global__ void path_words_kernel(const int32_t *__restrict__ path_state, const int64_t *__restrict__ path_off,
                                  const int32_t *__restrict__ path_len, int B, const int32_t *__restrict__ utt_graph,
                                  const int32_t *__restrict__ g_state_off, const int32_t *__restrict__ state_word,
                                  int max_words, int32_t *__restrict__ words, int32_t *__restrict__ nwords) {
    const int lane = threadIdx.x & 31;
    const int u = (int)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    if (u >= B) return;
    const int len = max(path_len[u], 0);
    const int32_t *ps = path_state + path_off[u];
    const int32_t *sw = state_word + g_state_off[utt_graph ? utt_graph[u] : 0];
    int count = 0, carry = -1;
    for (int c0 = 0; c0 < len; c0 += 32) {
        const int p = c0 + lane;
        const int s = p < len ? ps[len - 1 - p] : -1;            // the path is stored reversed (decode.py:142-146)
        int prev = __shfl_up_sync(0xffffffffu, s, 1);
        if (lane == 0) prev = carry;
        const bool start = p < len && s != prev && sw[s] >= 0 && (p == 0 || sw[prev] < 0);
        const unsigned m = __ballot_sync(0xffffffffu, start);
        if (start) {
            const int k = count + __popc(m & ((1u << lane) - 1));
            if (k < max_words) words[(size_t)u * max_words + k] = sw[s];
        }
        count += __popc(m);
        carry = __shfl_sync(0xffffffffu, s, 31);
    }
    if (lane == 0) nwords[u] = count;
}

}  // namespace srb

using namespace srb;

template <int W>
static int launch_fb_banded(const float *E, int64_t ldE, const int32_t *utt_off, int32_t B, const GraphBatch &G,
                            const float *band, const int32_t *codes, const int32_t *order, const int64_t *ws_off,
                            float *gamma, double *neg_log_z, double *xi_sum, double *xi_part, int nnz_total,
                            void *stream) {
    const int64_t slots = ((int64_t)B * W + 31) / 32;
    int64_t grid = (slots + VW - 1) / VW;
    const int64_t cap = (int64_t)sm_count() * 16;
    if (grid > cap) grid = cap;
    forward_backward_banded_kernel<W><<<(int)grid, VW * 32, 0, (cudaStream_t)stream>>>(
        E, ldE, utt_off, B, G, band, codes, order, ws_off, gamma, neg_log_z, xi_sum ? xi_part : nullptr, nnz_total);
    SRB_CUDA(cudaGetLastError());
    if (xi_sum) {
        xi_reduce_kernel<<<nnz_total, 256, 0, (cudaStream_t)stream>>>(xi_part, (int)grid * VW, nnz_total, xi_sum);
        SRB_CUDA(cudaGetLastError());
    }
    return SRB_OK;
}

extern "C" {

static GraphBatch make_graph(const int32_t *utt_graph, const int32_t *g_state_off, const int32_t *g_nes_off,
                             const int32_t *g_end_off, const int32_t *emit, const int32_t *pred_ptr,
                             const int32_t *pred_idx, const float *pred_cost, const int32_t *nes_idx,
                             const int32_t *end_state, const int32_t *end_frame = nullptr) {
    GraphBatch G;
    G.utt_graph = utt_graph;
    G.g_state_off = g_state_off;
    G.g_nes_off = g_nes_off;
    G.g_end_off = g_end_off;
    G.emit = emit;
    G.pred_ptr = pred_ptr;
    G.pred_idx = pred_idx;
    G.pred_cost = pred_cost;
    G.nes_idx = nes_idx;
    G.end_state = end_state;
    G.end_frame = end_frame;
    return G;
}

int srb_viterbi_graph(const float *E, int64_t ldE, const int32_t *utt_off, int32_t B, const int32_t *utt_graph,
                      const int32_t *g_state_off, const int32_t *g_nes_off, const int32_t *g_end_off,
                      const int32_t *emit, const int32_t *pred_ptr, const int32_t *pred_idx, const float *pred_cost,
                      const int32_t *nes_idx, const int32_t *end_state, const int32_t *end_frame, int32_t Rmax,
                      const int64_t *bp_off, uint8_t *bp, float *costs_out, float *best_cost, int32_t *best_end,
                      const int64_t *path_off, int32_t *path_state, int32_t *path_frame, int32_t *path_len,
                      int32_t *frame_label, void *stream) {
    if (B == 0) return SRB_OK;
    SRB_REQUIRE(!end_frame || costs_out, SRB_EINVAL,
                "srb_viterbi_graph: end points off the last frame need costs_out (their costs are read from it)");
    SRB_REQUIRE(E && utt_off && g_state_off && g_nes_off && g_end_off && emit && pred_ptr && pred_idx && pred_cost &&
                    end_state && bp_off && bp && best_cost && best_end,
                SRB_EINVAL, "srb_viterbi_graph: null pointer");
    SRB_REQUIRE(!path_state || (path_off && path_frame && path_len), SRB_EINVAL,
                "srb_viterbi_graph: path_state needs path_off, path_frame and path_len");
    SRB_REQUIRE(B >= 0 && Rmax > 0 && ldE > 0, SRB_EINVAL, "srb_viterbi_graph: bad sizes");
    if (B == 0) return SRB_OK;
    const size_t bytes = (size_t)VW * 2 * Rmax * sizeof(float);
    SRB_REQUIRE(bytes <= (size_t)smem_optin(), SRB_EUNSUPPORTED, "srb_viterbi_graph: %d states do not fit shared memory",
                Rmax);
    SRB_CUDA(cudaFuncSetAttribute(viterbi_graph_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    const int sms = sm_count();
    int64_t grid = ((int64_t)B + VW - 1) / VW;
    const int64_t cap = (int64_t)sms * 16;
    if (grid > cap) grid = cap;
    GraphBatch G = make_graph(utt_graph, g_state_off, g_nes_off, g_end_off, emit, pred_ptr, pred_idx, pred_cost, nes_idx,
                              end_state, end_frame);
    viterbi_graph_kernel<<<(int)grid, VW * 32, bytes, (cudaStream_t)stream>>>(
        E, ldE, utt_off, B, G, Rmax, bp_off, bp, costs_out, best_cost, best_end, path_off, path_state, path_frame,
        path_len, frame_label);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int srb_viterbi_banded(const float *E, int64_t ldE, const int32_t *utt_off, int32_t B, const int32_t *utt_graph,
                       const int32_t *g_state_off, const int32_t *g_nes_off, const int32_t *g_end_off,
                       const int32_t *emit, const int32_t *pred_ptr, const int32_t *pred_idx, const float *pred_cost,
                       const int32_t *nes_idx, const int32_t *end_state, const float *band, const int32_t *codes,
                       const int32_t *entry_nes, const float *entry_cost, const int32_t *g_entry_off,
                       const int32_t *entry_state, int32_t Rmax, const int64_t *bp_off, uint8_t *bp, float *costs_out,
                       float *best_cost, int32_t *best_end, const int64_t *path_off, int32_t *path_state,
                       int32_t *path_frame, int32_t *path_len, int32_t *frame_label, void *stream) {
    if (B == 0) return SRB_OK;
    SRB_REQUIRE(E && utt_off && g_state_off && g_nes_off && g_end_off && emit && pred_ptr && pred_idx && pred_cost &&
                    end_state && bp_off && bp && best_cost && best_end && band && codes && entry_nes && entry_cost &&
                    g_entry_off && entry_state,
                SRB_EINVAL, "srb_viterbi_banded: null pointer");
    SRB_REQUIRE(!path_state || (path_off && path_frame && path_len), SRB_EINVAL,
                "srb_viterbi_banded: path_state needs path_off, path_frame and path_len");
    SRB_REQUIRE(B >= 0 && Rmax > 0 && ldE > 0, SRB_EINVAL, "srb_viterbi_banded: bad sizes");
    if (B == 0) return SRB_OK;
    const size_t bytes = (size_t)VW * 2 * Rmax * sizeof(float);
    SRB_REQUIRE(bytes <= (size_t)smem_optin(), SRB_EUNSUPPORTED, "srb_viterbi_banded: %d states do not fit shared memory",
                Rmax);
    SRB_CUDA(cudaFuncSetAttribute(viterbi_banded_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    int64_t grid = ((int64_t)B + VW - 1) / VW;
    const int64_t cap = (int64_t)sm_count() * 16;
    if (grid > cap) grid = cap;
    GraphBatch G = make_graph(utt_graph, g_state_off, g_nes_off, g_end_off, emit, pred_ptr, pred_idx, pred_cost, nes_idx,
                              end_state);
    BandForm F;
    F.band = band;
    F.codes = codes;
    F.entry_nes = entry_nes;
    F.entry_cost = entry_cost;
    F.g_entry_off = g_entry_off;
    F.entry_state = entry_state;
    const size_t bytes1 = banded1_smem_bytes(Rmax);
    if (!utt_graph && bytes1 <= (size_t)smem_optin() / 2 && ldE <= (1 << 18)) {   // emission column in 18 bits of meta
        // one graph for the whole batch: its tables live in shared memory (viterbi_banded1_kernel)
        SRB_CUDA(cudaFuncSetAttribute(viterbi_banded1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes1));
        viterbi_banded1_kernel<<<(int)grid, VW * 32, bytes1, (cudaStream_t)stream>>>(
            E, ldE, utt_off, B, G, F, Rmax, bp_off, bp, costs_out, best_cost, best_end, path_off, path_state, path_frame,
            path_len, frame_label);
        SRB_CUDA(cudaGetLastError());
        return SRB_OK;
    }
    viterbi_banded_kernel<<<(int)grid, VW * 32, bytes, (cudaStream_t)stream>>>(
        E, ldE, utt_off, B, G, F, Rmax, bp_off, bp, costs_out, best_cost, best_end, path_off, path_state, path_frame,
        path_len, frame_label);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int srb_forward_backward(const float *E, int64_t ldE, const int32_t *utt_off, int32_t B, const int32_t *utt_graph,
                         const int32_t *g_state_off, const int32_t *g_end_off, const int32_t *emit,
                         const int32_t *pred_ptr, const int32_t *pred_idx, const float *pred_cost,
                         const int32_t *pred_tgt, const int32_t *succ_ptr, const int32_t *succ_idx,
                         const float *succ_cost, const int32_t *end_state, int32_t Rmax, int32_t NNZmax,
                         const int64_t *ws_off, float *gamma, double *neg_log_z, double *xi_sum, double *xi_part,
                         int32_t nnz_total, void *stream) {
    SRB_REQUIRE(E && utt_off && g_state_off && g_end_off && emit && pred_ptr && pred_idx && pred_cost && pred_tgt &&
                    succ_ptr && succ_idx && succ_cost && end_state && ws_off && gamma,
                SRB_EINVAL, "srb_forward_backward: null pointer");
    SRB_REQUIRE(B >= 0 && Rmax > 0 && NNZmax >= 0 && ldE > 0, SRB_EINVAL, "srb_forward_backward: bad sizes");
    SRB_REQUIRE(!xi_sum || (xi_part && nnz_total > 0), SRB_EINVAL,
                "srb_forward_backward: xi_sum needs the xi_part workspace (srb_fb_xi_parts() * nnz_total doubles)");
    if (B == 0) return SRB_OK;
    const size_t bytes = (size_t)VW * ((2 * Rmax + NNZmax) * sizeof(float) + NNZmax * sizeof(double));
    SRB_REQUIRE(bytes <= (size_t)smem_optin(), SRB_EUNSUPPORTED, "srb_forward_backward: graph too large for shared memory");
    SRB_CUDA(cudaFuncSetAttribute(forward_backward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    const int sms = sm_count();
    int64_t grid = ((int64_t)B + VW - 1) / VW;
    const int64_t cap = (int64_t)sms * 16;
    if (grid > cap) grid = cap;
    GraphBatch G = make_graph(utt_graph, g_state_off, g_state_off, g_end_off, emit, pred_ptr, pred_idx, pred_cost, nullptr,
                              end_state);
    forward_backward_kernel<<<(int)grid, VW * 32, bytes, (cudaStream_t)stream>>>(
        E, ldE, utt_off, B, G, succ_ptr, succ_idx, succ_cost, pred_tgt, Rmax, NNZmax, ws_off, gamma, neg_log_z,
        xi_sum ? xi_part : nullptr, nnz_total);
    SRB_CUDA(cudaGetLastError());
    if (xi_sum) {
        xi_reduce_kernel<<<nnz_total, 256, 0, (cudaStream_t)stream>>>(xi_part, (int)grid * VW, nnz_total, xi_sum);
        SRB_CUDA(cudaGetLastError());
    }
    return SRB_OK;
}

int srb_fb_xi_parts(void) { return sm_count() * 16 * VW; }

int srb_forward_backward_banded(const float *E, int64_t ldE, const int32_t *utt_off, int32_t B, const int32_t *utt_graph,
                                const int32_t *g_state_off, const int32_t *g_end_off, const int32_t *emit,
                                const int32_t *pred_ptr, const int32_t *end_state, const float *band,
                                const int32_t *codes, const int32_t *order, int32_t Rmax, const int64_t *ws_off,
                                float *gamma, double *neg_log_z, double *xi_sum, double *xi_part, int32_t nnz_total,
                                void *stream) {
    SRB_REQUIRE(E && utt_off && g_state_off && g_end_off && emit && pred_ptr && end_state && band && codes && ws_off &&
                    gamma,
                SRB_EINVAL, "srb_forward_backward_banded: null pointer");
    SRB_REQUIRE(!xi_sum || (xi_part && nnz_total > 0), SRB_EINVAL,
                "srb_forward_backward_banded: xi_sum needs the xi_part workspace (srb_fb_xi_parts() * nnz_total doubles)");
    SRB_REQUIRE(B >= 0 && Rmax > 0 && ldE > 0, SRB_EINVAL, "srb_forward_backward_banded: bad sizes");
    SRB_REQUIRE(Rmax <= 32, SRB_EUNSUPPORTED, "srb_forward_backward_banded: %d states (at most 32: use srb_forward_backward)",
                Rmax);
    if (B == 0) return SRB_OK;
    GraphBatch G = make_graph(utt_graph, g_state_off, g_state_off, g_end_off, emit, pred_ptr, nullptr, nullptr, nullptr,
                              end_state);
    if (Rmax <= 4)
        return launch_fb_banded<4>(E, ldE, utt_off, B, G, band, codes, order, ws_off, gamma, neg_log_z, xi_sum, xi_part,
                                   nnz_total, stream);
    if (Rmax <= 8)
        return launch_fb_banded<8>(E, ldE, utt_off, B, G, band, codes, order, ws_off, gamma, neg_log_z, xi_sum, xi_part,
                                   nnz_total, stream);
    if (Rmax <= 16)
        return launch_fb_banded<16>(E, ldE, utt_off, B, G, band, codes, order, ws_off, gamma, neg_log_z, xi_sum, xi_part,
                                   nnz_total, stream);
    return launch_fb_banded<32>(E, ldE, utt_off, B, G, band, codes, order, ws_off, gamma, neg_log_z, xi_sum, xi_part,
                                nnz_total, stream);
}

int srb_path_words(const int32_t *path_state, const int64_t *path_off, const int32_t *path_len, int32_t B,
                   const int32_t *utt_graph, const int32_t *g_state_off, const int32_t *state_word, int32_t max_words,
                   int32_t *words, int32_t *nwords, void *stream) {
    SRB_REQUIRE(B >= 0 && max_words >= 0, SRB_EINVAL, "srb_path_words: bad sizes");
    if (B == 0) return SRB_OK;
    SRB_REQUIRE(path_state && path_off && path_len && g_state_off && state_word && nwords && (words || max_words == 0),
                SRB_EINVAL, "srb_path_words: null pointer");
    const int threads = 128;
    const int64_t grid = ((int64_t)B * 32 + threads - 1) / threads;
    path_words_kernel<<<(int)grid, threads, 0, (cudaStream_t)stream>>>(path_state, path_off, path_len, B, utt_graph,
                                                                        g_state_off, state_word, max_words, words, nwords);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // extern "C"
