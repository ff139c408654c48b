// em.cu -- fused segmental-K-means / EM pass over isolated-word utterances (BASELINE config 2):
//   K1 scoring of each utterance against ITS word's states  ->  left-to-right Viterbi with segment boundaries
//   ->  within-state posteriors  ->  sufficient statistics,   all in one kernel, X read from HBM exactly once.
// Replaces, per training iteration, align_gmm_states (kmeans.py:196-205: decode_hmm_states + get_segments_from_path
// + segment_data) followed by the E-step of GMM.em on every state's pool (hmm_state.py:122-136).
//
// Work unit = a TILE of whole utterances of ONE word with at most ETILE frames (planned on the host,
// srb_em_plan).  Per tile a CTA of 128 threads
//   1. stages the frames in shared memory and scores them (thread = 2 frames, register-resident, score_core.cuh)
//      against the word's S states: costs E, per-component log-likelihoods (kept for the posteriors);
//   2. runs Viterbi with one thread per utterance (threads spread over the warps).  The models are left-to-right
//      (origin <= target, skips of at most 2 states: kmeans.py:53-95), so the best path is monotone and the
//      thread carries, instead of back-pointers, the segment boundaries bd[r][k] = first frame at which the best
//      path into state r is in a state >= k.  No back-pointer array, no traceback: bd[S-1][1..S-1] IS
//      get_segments_from_path's result (kmeans.py:98-108);
//   3. labels every frame from the boundaries, normalises the posteriors of the first ncomp components;
//   4. accumulates [S0, S1, S2, count, loglik] rows with the lanes over the feature dimensions (moments about
//      each component's mean, like stats.cu) into per-warp shared partials, flushed to the CTA's private global
//      slice when the word changes.  No atomics; reduce_partials adds the slices in a fixed order in double.
// The scorer runs at ONE frame per thread here (~128 registers): 16 warps are resident per SM (4 CTAs of 128
// threads, or 2 of 256 when an utterance is longer than 128 frames), so one CTA's serial Viterbi phase overlaps the
// others' scoring and statistics.
#include <algorithm>
#include <type_traits>

#include "score_core.cuh"

namespace srb {

constexpr int EFR = 1;               // frames per thread (1: ~128 registers -> 16 warps per SM)
constexpr int ETILE_MAX = 256;       // threads per CTA = frames per tile: 128 (4 CTAs/SM) or 256 (2 CTAs/SM)
constexpr int EUTT = 32;             // utterances per tile
constexpr int EPARTS = 4;            // private statistics slices per SM (= CTAs per SM at the small tile)
constexpr int ESMAX = SRB_DP_MAX_STATES;
constexpr int EMAXC = 8;             // trained components per state handled in registers

struct EmSmem {                      // offsets in floats into the dynamic shared memory
    int tr, pc, pab, mean, xs, es, ps, ll, part, bnd, uoff, lab, seg, total;
};

__host__ __device__ inline EmSmem em_layout(int S, int Mp, int ncomp, int D, bool stats, int ETILE) {
    const int EWARPS = ETILE / 32;
    EmSmem L;
    int o = 64;                                           // shift_s
    L.tr = o;   o += ESMAX * ESMAX;
    L.pc = o;   o += (S * Mp + 3) & ~3;
    L.pab = o;  o += S * Mp * D * 2;
    L.mean = o; o += (S * ncomp * D + 3) & ~3;
    L.xs = o;   o += ETILE * D + 4;
    L.es = o;   o += ETILE * (S | 1);
    L.ps = o;   o += ETILE * ((S * ncomp) | 1);
    L.ll = o;   o += ETILE;
    L.part = o; o += stats ? EWARPS * S * ncomp * (2 * D + 3) : 0;
    L.bnd = o;  o += EUTT * ESMAX;
    L.uoff = o; o += EUTT + 4;
    L.lab = o;  o += ETILE / 4;
    L.seg = o;  o += 2 * 8 * 32;                          // segment transfer costs + boundaries (time-parallel Viterbi)
    L.total = o;
    return L;
}

template <int D, int ST, int NC, int ENT>
__global__ void __launch_bounds__(ENT, 512 / ENT)
em_isolated_kernel(const float *__restrict__ X, int64_t N, const int32_t *__restrict__ off,
                   const int4 *__restrict__ tile_desc, int ntiles,
                   const float *__restrict__ ab, const float *__restrict__ c, const float *__restrict__ shift, int S_rt,
                   int M, int Mp, const float *__restrict__ trans, int ncomp_rt, const double *__restrict__ mu,
                   float *__restrict__ partial, unsigned long long *__restrict__ slice_mask, int G,
                   float *__restrict__ best_cost, int32_t *__restrict__ seg_start, int32_t *__restrict__ frame_label) {
    extern __shared__ __align__(16) float smem[];
    const int S = ST ? ST : S_rt;
    const int ncomp = NC ? NC : ncomp_rt;
    const bool stats = partial != nullptr;
    constexpr int ETILE = ENT * EFR, EWARPS = ENT / 32;
    const EmSmem L = em_layout(S, Mp, ncomp, D, stats, ETILE);
    float *shift_s = smem, *tr_s = smem + L.tr, *pc = smem + L.pc, *pab = smem + L.pab, *mean_s = smem + L.mean;
    float *xs = smem + L.xs, *Es = smem + L.es, *Ps = smem + L.ps, *LLs = smem + L.ll, *part = smem + L.part;
    int *bnd = reinterpret_cast<int *>(smem + L.bnd), *uoff = reinterpret_cast<int *>(smem + L.uoff);
    unsigned char *lab = reinterpret_cast<unsigned char *>(smem + L.lab);
    const int ES = S | 1, PS = (S * ncomp) | 1, row = 2 * D + 3, GW = S * ncomp;
    const int tid = threadIdx.x, lane = tid & 31, wq = tid >> 5;
    const int nblk = Mp >> 2;
    const float INF = CUDART_INF_F;

    if (tid < D) shift_s[tid] = shift[tid];
    // contiguous range of tiles per CTA: a CTA sees one or two words, so parameters are (re)loaded rarely
    const int t_lo = (int)((int64_t)ntiles * blockIdx.x / gridDim.x);
    const int t_hi = (int)((int64_t)ntiles * (blockIdx.x + 1) / gridDim.x);
    // The CTA's private slice holds one block of rows per word.  Words arrive in ascending order, so a word is flushed
    // exactly once per launch: the flush WRITES its rows and sets the word's bit in the slice's mask; rows of other
    // words are never read by the reducers (no 42 MB memset and no 42 MB of zeros to add per iteration).
    float *slice = stats ? partial + (size_t)blockIdx.x * G * row : nullptr;
    unsigned long long wmask = 0ull;
    int cur_w = -1;

    int4 desc = t_lo < t_hi ? __ldg(tile_desc + t_lo) : make_int4(0, 0, 0, 0);
    for (int t = t_lo; t < t_hi; ++t) {
        // descriptor {first utterance, first frame, frames, word << 8 | utterances}; the next one is fetched now
        const int u0 = desc.x, nf = desc.z, nutt = desc.w & 0xff, w = desc.w >> 8;
        const int64_t f0 = desc.y;
        if (t + 1 < t_hi) desc = __ldg(tile_desc + t + 1);
        __syncthreads();                                   // previous tile fully consumed
        if (w != cur_w) {
            if (stats && cur_w >= 0) {                     // flush the finished word's statistics (fixed order)
                for (int i = tid; i < GW * row; i += ENT) {
                    float s = 0.f;
#pragma unroll
                    for (int q = 0; q < EWARPS; ++q) s += part[q * GW * row + i];
                    const int g = i / row, j = i - g * row;
                    slice[((size_t)(cur_w * S + g / ncomp) * M + g % ncomp) * row + j] = s;
                }
                wmask |= 1ull << (cur_w & 63);
            }
            __syncthreads();
            if (stats)
                for (int i = tid; i < EWARPS * GW * row; i += ENT) part[i] = 0.f;
            copy_to_smem(pab, ab + (size_t)w * S * Mp * D * 2, (int64_t)S * Mp * D * 2);
            for (int i = tid; i < S * Mp; i += ENT) pc[i] = c[(size_t)w * S * Mp + i];
            for (int i = tid; i < ESMAX * ESMAX; i += ENT) {
                const int r = i / ESMAX, o = i % ESMAX;
                tr_s[i] = (r < S && o < S) ? trans[((size_t)w * S + r) * S + o] : INF;
            }
            for (int i = tid; i < S * ncomp * D; i += ENT) {
                const int s = i / (ncomp * D), m = (i / D) % ncomp, d = i % D;
                mean_s[i] = (float)mu[((size_t)(w * S + s) * M + m) * D + d];
            }
            cur_w = w;
        }
        for (int i = tid; i <= nutt; i += ENT) uoff[i] = (int)(off[u0 + i] - f0);
        const int xoff = stage_frames_async<D, ENT>(xs, X, N, f0, nf);
        cp_async_wait_all();
        __syncthreads();

        // ---- 1. scoring: E[f][s], natural-log component scores Ps[f][s*Mp + m]
        {
            Frames<D, EFR> fr;
            load_frames<D, ENT, EFR>(fr, xs, xoff, nf, shift_s);
            for (int s = 0; s < S && (tid & ~31) < nf; ++s) {          // warps past the tile's last frame skip
                float cost[EFR];
                float *crow[EFR];
#pragma unroll
                for (int k = 0; k < EFR; ++k) crow[k] = Ps + (size_t)(tid + k * ENT) * PS + s * ncomp;
                score_state<D, EFR>(fr, pab + (size_t)s * Mp * D * 2, pc + s * Mp, nblk, cost, crow, true, ncomp);
#pragma unroll
                for (int k = 0; k < EFR; ++k) Es[(tid + k * ENT) * ES + s] = cost[k];
            }
        }
        __syncthreads();

        // ---- 2. Viterbi (decode.py:97-134 for a left-to-right model; ties: lowest origin wins like np.argmin).
        if constexpr (ST > 0 && ST * ST <= 32) {
            // TIME-PARALLEL form for S <= 5: the utterance is cut into one segment per warp; warp q computes the
            // (min,+) transfer matrix of its segment, lane (i, j) = best cost from state i at the segment start to
            // state j at its end (same banded recursion, 1/EWARPS of the steps), with the boundaries of that
            // partial path packed one byte each.  Warp 0 then chains the matrices from the start vector and splices
            // the boundaries.  The dependent chain is T/EWARPS + EWARPS steps instead of T.
            float *Mc = smem + L.seg;
            uint32_t *Mb = reinterpret_cast<uint32_t *>(smem + L.seg + 8 * 32);
            const int i = lane / S, j = lane % S;
            const bool live = lane < S * S;
            const float a0 = live ? tr_s[j * ESMAX + j] : INF;
            const float a1 = (live && j >= 1) ? tr_s[j * ESMAX + j - 1] : INF;
            const float a2 = (live && j >= 2) ? tr_s[j * ESMAX + j - 2] : INF;
            const uint32_t ones = 0x01010101u;
            const uint32_t keep1 = j >= 2 ? ((1u << (8 * (j - 1))) - 1) : 0u;
            const uint32_t keep2 = j >= 3 ? ((1u << (8 * (j - 2))) - 1) : 0u;
            for (int p = 0; p < nutt; ++p) {
                const int ts = uoff[p], T = uoff[p + 1] - ts;
                const int per = (T - 1 + EWARPS - 1) / EWARPS;
                const int t_lo = 1 + wq * per, t_hi = min(T, t_lo + per);
                float m = (live && i == j) ? 0.f : INF;
                uint32_t bd = 0, rep = ones * (uint32_t)(t_lo - 1);
                const float *Ep = Es + (size_t)ts * ES + j;
                float e_next = (t_lo < t_hi) ? Ep[(size_t)t_lo * ES] : INF;
                for (int tt = t_lo; tt < t_hi; ++tt) {
                    const float e = e_next;
                    if (tt + 1 < t_hi) e_next = Ep[(size_t)(tt + 1) * ES];
                    const float c1 = __shfl_up_sync(0xffffffffu, m, 1);
                    const float c2 = __shfl_up_sync(0xffffffffu, m, 2);
                    const uint32_t b1 = __shfl_up_sync(0xffffffffu, bd, 1);
                    const uint32_t b2 = __shfl_up_sync(0xffffffffu, bd, 2);
                    const float t2 = c2 + a2, t1 = c1 + a1, t0 = m + a0;       // a1/a2 = inf mask the row wrap-around
                    const bool p1 = t1 < t2;
                    const float b12 = p1 ? t1 : t2;
                    const bool p0 = t0 < b12;
                    rep += ones;
                    const uint32_t inh = p1 ? b1 : b2, keep = p1 ? keep1 : keep2;
                    bd = p0 ? bd : ((inh & keep) | (rep & ~keep));
                    m = (p0 ? t0 : b12) + e;
                }
                Mc[wq * 32 + lane] = m;
                Mb[wq * 32 + lane] = bd;
                __syncthreads();
                if (wq == 0) {
                    float v = (lane == 0 && T > 0) ? Es[(size_t)ts * ES] : INF;   // alpha at frame 0, lane = state
                    uint32_t vb = 0;
                    const int jj = lane < S ? lane : 0;
                    for (int q = 0; q < EWARPS; ++q) {
                        float best = INF;
                        int bi = 0;
#pragma unroll
                        for (int ii = 0; ii < ST; ++ii) {
                            const float cand = __shfl_sync(0xffffffffu, v, ii) + Mc[q * 32 + ii * S + jj];
                            if (cand < best) {
                                best = cand;
                                bi = ii;
                            }
                        }
                        const uint32_t vbi = __shfl_sync(0xffffffffu, vb, bi);
                        const uint32_t keep = bi > 0 ? ((bi >= 4) ? 0xffffffffu : ((1u << (8 * bi)) - 1)) : 0u;
                        vb = (vbi & keep) | (Mb[q * 32 + bi * S + jj] & ~keep);
                        v = lane < S ? best : INF;
                    }
                    if (lane == S - 1) {
                        bnd[p * ESMAX] = (v < INF) ? 0 : -1;       // -1: no path, the utterance's frames are skipped
                        for (int k = 1; k < S; ++k) bnd[p * ESMAX + k] = (int)((vb >> (8 * (k - 1))) & 0xff);
                        if (best_cost) best_cost[u0 + p] = v;
                        if (seg_start)
                            for (int k = 0; k < S; ++k)
                                seg_start[(size_t)(u0 + p) * S + k] = (v < INF && k) ? (int)((vb >> (8 * (k - 1))) & 0xff) : 0;
                    }
                }
                __syncthreads();
            }
        } else
        //         Sequential form: 8 lanes per utterance (lane = state), 4 utterances per warp.
        {
            using BT = typename std::conditional<(ST > 0 && ST <= 5), uint32_t, uint64_t>::type;
            const int grp = lane >> 3, r = lane & 7;
            const float a0 = tr_s[r * ESMAX + r];
            const float a1 = r >= 1 ? tr_s[r * ESMAX + r - 1] : INF;
            const float a2 = r >= 2 ? tr_s[r * ESMAX + r - 2] : INF;
            const BT ones = (BT)0x0101010101010101ull;
            // byte k-1 holds boundary k; an origin o keeps the bytes < o
            const BT keep1 = r >= 2 ? (((BT)1 << (8 * (r - 1))) - 1) : (BT)0;
            const BT keep2 = r >= 3 ? (((BT)1 << (8 * (r - 2))) - 1) : (BT)0;
            for (int pb = 0; pb < nutt; pb += EWARPS * 4) {
                if (pb + wq * 4 >= nutt) break;                       // warp uniform: nothing left for this warp
                const int p = pb + wq * 4 + grp;
                const bool act = p < nutt;
                const int ts = act ? uoff[p] : 0;
                const int T = act ? uoff[p + 1] - ts : 0;
                int Tmax = T;
#pragma unroll
                for (int d = 16; d >= 8; d >>= 1) Tmax = max(Tmax, __shfl_xor_sync(0xffffffffu, Tmax, d));
                const int rr = r < S ? r : 0;
                const float *Ep = Es + (size_t)ts * ES + rr;
                float cost = (T > 0 && r == 0) ? Ep[0] : INF;
                BT bd = 0, rep = 0;
                float e_next = (T > 1 && r < S) ? Ep[ES] : INF;
                for (int tt = 1; tt < Tmax; ++tt) {
                    const float e = e_next;
                    if (tt + 1 < T && r < S) e_next = Ep[(size_t)(tt + 1) * ES];
                    const float c1 = __shfl_up_sync(0xffffffffu, cost, 1, 8);
                    const float c2 = __shfl_up_sync(0xffffffffu, cost, 2, 8);
                    const BT b1 = __shfl_up_sync(0xffffffffu, bd, 1, 8);
                    const BT b2 = __shfl_up_sync(0xffffffffu, bd, 2, 8);
                    // candidates in ascending origin order, strict '<': the lowest origin wins ties (np.argmin)
                    const float t2 = c2 + a2, t1 = c1 + a1, t0 = cost + a0;
                    const bool p1 = t1 < t2;
                    const float b12 = p1 ? t1 : t2;
                    const bool p0 = t0 < b12;
                    const float best = p0 ? t0 : b12;
                    rep += ones;                                       // tt in every byte
                    const BT inh = p1 ? b1 : b2;                       // boundaries k <= origin are inherited,
                    const BT keep = p1 ? keep1 : keep2;                // the later ones become tt
                    const BT nb = p0 ? bd : ((inh & keep) | (rep & ~keep));
                    if (tt < T) {
                        cost = best + e;
                        bd = nb;
                    }
                }
                if (act && r == S - 1) {
                    const float fc = cost;
                    bnd[p * ESMAX] = (fc < INF) ? 0 : -1;      // -1: no path, the utterance's frames are skipped
                    for (int k = 1; k < S; ++k) bnd[p * ESMAX + k] = (int)((bd >> (8 * (k - 1))) & 0xff);
                    if (best_cost) best_cost[u0 + p] = fc;
                    if (seg_start)
                        for (int k = 0; k < S; ++k)
                            seg_start[(size_t)(u0 + p) * S + k] = (fc < INF && k) ? (int)((bd >> (8 * (k - 1))) & 0xff) : 0;
                }
            }
        }
        __syncthreads();

        // ---- 3. frame labels from the boundaries: state = number of boundaries <= t
        for (int f = tid; f < nf; f += ENT) {
            int lo = 0, hi = nutt;                         // last p with uoff[p] <= f
            while (hi - lo > 1) {
                const int mid = (lo + hi) >> 1;
                if (uoff[mid] <= f) lo = mid; else hi = mid;
            }
            const int tt = f - uoff[lo];
            int s = 255;
            if (bnd[lo * ESMAX] == 0) {
                s = 0;
                for (int k = 1; k < S; ++k) s += (bnd[lo * ESMAX + k] <= tt) ? 1 : 0;
            }
            lab[f] = (unsigned char)s;
            if (frame_label) frame_label[f0 + f] = (s == 255) ? -1 : w * S + s;
            if (stats && s != 255) {
                // posteriors over the first ncomp components of the ALIGNED state (hmm_state.py:126-136, log domain)
                float *prow = Ps + (size_t)f * PS + s * ncomp;
                float mx = -INF;
                for (int m = 0; m < ncomp; ++m) mx = fmaxf(mx, prow[m]);
                float den = 0.f;
                for (int m = 0; m < ncomp; ++m) {
                    const float e = (mx > -INF) ? __expf(prow[m] - mx) : 0.f;
                    prow[m] = e;
                    den += e;
                }
                const float inv = den > 0.f ? 1.f / den : 0.f;
                for (int m = 0; m < ncomp; ++m) prow[m] *= inv;
                LLs[f] = den > 0.f ? mx + __logf(den) : 0.f;
            }
        }
        if (!stats) continue;                              // uniform per CTA: alignment-only call
        __syncthreads();

        // ---- 4. statistics: each warp takes a quarter of the tile's frames and walks it RUN by RUN (a run = the
        //         frames of one utterance aligned to one state, known from the boundaries), lanes over the feature
        //         dimensions, the run's moments in registers, one flush per run into the warp's shared partials.
        {
            const int d0 = lane, d1 = lane + 32;
            const bool h0 = d0 < D, h1 = d1 < D;
            const int chunk = (nf + EWARPS - 1) / EWARPS;
            const int c0 = wq * chunk, c1 = min(nf, c0 + chunk);
            float *mine = part + (size_t)wq * GW * row;
            constexpr int NCC = NC ? NC : EMAXC;
            int f = c0;
            while (f < c1) {
                int lo = 0, hi = nutt;                         // utterance of frame f (warp uniform)
                while (hi - lo > 1) {
                    const int mid = (lo + hi) >> 1;
                    if (uoff[mid] <= f) lo = mid; else hi = mid;
                }
                const int us = uoff[lo], ue = uoff[lo + 1];
                const int st = lab[f];
                if (st == 255) {                               // utterance without a path: skipped
                    f = ue;
                    continue;
                }
                const int rend = (st + 1 < S) ? us + bnd[lo * ESMAX + st + 1] : ue;
                const int e = min(c1, rend);
                float m0[NCC], m1[NCC], s0[NCC], a0[NCC], a1[NCC], q0[NCC], q1[NCC];
#pragma unroll
                for (int m = 0; m < NCC; ++m) {
                    s0[m] = a0[m] = a1[m] = q0[m] = q1[m] = 0.f;
                    m0[m] = (m < ncomp && h0) ? mean_s[(st * ncomp + m) * D + d0] : 0.f;
                    m1[m] = (m < ncomp && h1) ? mean_s[(st * ncomp + m) * D + d1] : 0.f;
                }
                float llsum = 0.f;
                const float cnt = (float)(e - f);
                const float *xp = xs + xoff + f * D;
                const float *pp = Ps + (size_t)f * PS + st * ncomp;
                const float *lp = LLs + f;
                for (; f < e; ++f, xp += D, pp += PS, ++lp) {
                    const float xa = h0 ? xp[d0] : 0.f, xb = h1 ? xp[d1] : 0.f;
                    llsum += lp[0];
#pragma unroll
                    for (int m = 0; m < NCC; ++m)
                        if (m < ncomp) {
                            const float pw = pp[m];
                            const float e0 = xa - m0[m], e1 = xb - m1[m];
                            s0[m] += pw;
                            a0[m] = fmaf(pw, e0, a0[m]);
                            a1[m] = fmaf(pw, e1, a1[m]);
                            q0[m] = fmaf(pw * e0, e0, q0[m]);
                            q1[m] = fmaf(pw * e1, e1, q1[m]);
                        }
                }
#pragma unroll
                for (int m = 0; m < NCC; ++m) {
                    if (m < ncomp) {
                        float *r = mine + (size_t)(st * ncomp + m) * row;
                        if (lane == 0) {
                            r[0] += s0[m];
                            if (m == 0) {
                                r[2 * D + 1] += cnt;
                                r[2 * D + 2] += llsum;
                            }
                        }
                        if (h0) {
                            r[1 + d0] += a0[m];
                            r[1 + D + d0] += q0[m];
                        }
                        if (h1) {
                            r[1 + d1] += a1[m];
                            r[1 + D + d1] += q1[m];
                        }
                    }
                }
            }
        }
    }
    // ---- flush the last word's statistics
    __syncthreads();
    if (stats && cur_w >= 0) {
        for (int i = tid; i < GW * row; i += ENT) {
            float s = 0.f;
#pragma unroll
            for (int q = 0; q < EWARPS; ++q) s += part[q * GW * row + i];
            const int g = i / row, j = i - g * row;
            slice[((size_t)(cur_w * S + g / ncomp) * M + g % ncomp) * row + j] = s;
        }
        wmask |= 1ull << (cur_w & 63);
    }
    if (stats && tid == 0) slice_mask[blockIdx.x] = wmask;
}

// Fixed-order fp64 sum of the per-CTA slices.  Only the rows a slice's CTA has written are read: word w of slice p when
// bit w of mask[p] is set, and of those the rows of the first ncomp components (the others were never touched: 0).
__global__ void reduce_slices_kernel(const float *__restrict__ partial, const unsigned long long *__restrict__ mask,
                                     int64_t nparts, int64_t len, int word_len, int M, int ncomp, int row,
                                     double *__restrict__ out, int accumulate) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= len) return;
    double s = accumulate ? out[i] : 0.0;
    const int w = (int)(i / word_len);
    if ((int)((i / row) % M) < ncomp) {
        const unsigned long long bit = 1ull << (w & 63);
        for (int64_t p = 0; p < nparts; ++p)
            if (mask[p] & bit) s += (double)partial[p * len + i];
    }
    out[i] = s;
}

}  // namespace srb

using namespace srb;

extern "C" {

int srb_em_tile_frames(void) { return ETILE_MAX; }

int srb_em_parts(void) { return sm_count() * EPARTS; }

int64_t srb_em_partial_floats(int W, int S, int M, int D) {
    const int64_t nparts = (int64_t)sm_count() * EPARTS;
    return nparts * W * S * M * (2 * D + 3) + 2 * nparts;      // the slices + one 64-bit word mask per slice
}

int srb_em_plan(const int32_t *utt_off, const int32_t *utt_word, int32_t B, int32_t *tile_desc, int32_t *ntiles,
                int32_t *tile_frames) {
    SRB_REQUIRE(utt_off && utt_word && tile_desc && ntiles && tile_frames && B >= 0, SRB_EINVAL,
                "srb_em_plan: bad argument");
    int64_t longest = 0;
    for (int32_t u = 0; u < B; ++u) longest = std::max<int64_t>(longest, (int64_t)utt_off[u + 1] - utt_off[u]);
    const int ETILE = longest <= ETILE_MAX / 2 ? ETILE_MAX / 2 : ETILE_MAX;   // small tiles: 4 CTAs per SM
    *tile_frames = ETILE;
    int nt = 0;
    int32_t u = 0;
    while (u < B) {
        SRB_REQUIRE(u == 0 || utt_word[u] >= utt_word[u - 1], SRB_EINVAL,
                    "srb_em_plan: utterances must be sorted by word (utterance %d)", u);
        SRB_REQUIRE(utt_word[u] >= 0 && utt_word[u] < (1 << 22), SRB_EINVAL, "srb_em_plan: bad word id");
        const int64_t start = utt_off[u];
        int32_t v = u;
        while (v < B && utt_word[v] == utt_word[u] && (int64_t)utt_off[v + 1] - start <= ETILE && v - u < EUTT) {
            SRB_REQUIRE(utt_off[v + 1] > utt_off[v], SRB_EUNSUPPORTED, "srb_em_plan: empty utterance %d", v);
            ++v;
        }
        SRB_REQUIRE(v > u, SRB_EUNSUPPORTED, "srb_em_plan: utterance %d has more than %d frames", u, ETILE);
        int32_t *d = tile_desc + 4 * (size_t)nt++;
        d[0] = u;
        d[1] = utt_off[u];
        d[2] = utt_off[v] - utt_off[u];
        d[3] = (utt_word[u] << 8) | (v - u);
        u = v;
    }
    *ntiles = nt;
    return SRB_OK;
}

int srb_em_isolated(const float *X, int64_t N, int D, const int32_t *utt_off, int32_t B, const int32_t *tile_desc,
                    int32_t ntiles, int32_t tile_frames, const float *ab, const float *c, const float *shift,
                    int W, int S, int M, int Mp, const float *trans, int ncomp, const double *mu, float *partial,
                    double *stats_out, int accumulate, float *best_cost, int32_t *seg_start, int32_t *frame_label,
                    void *stream) {
    SRB_REQUIRE(X && utt_off && tile_desc && ab && c && shift && trans && ((uintptr_t)tile_desc & 15) == 0, SRB_EINVAL,
                "srb_em_isolated: null pointer");
    SRB_REQUIRE(!partial || ((stats_out || accumulate < 0) && mu), SRB_EINVAL,
                "srb_em_isolated: statistics need stats_out and mu");
    SRB_REQUIRE(W > 0 && S > 0 && S <= ESMAX && M > 0 && Mp >= M && (Mp & 3) == 0, SRB_EUNSUPPORTED,
                "srb_em_isolated: S=%d (max %d), M=%d, Mp=%d", S, ESMAX, M, Mp);
    SRB_REQUIRE(ncomp >= 1 && ncomp <= M && ncomp <= EMAXC, SRB_EUNSUPPORTED, "srb_em_isolated: ncomp=%d", ncomp);
    SRB_REQUIRE(((uintptr_t)X & 15) == 0 && ((uintptr_t)ab & 15) == 0, SRB_EINVAL, "srb_em_isolated: alignment");
    const int sms = sm_count();
    SRB_REQUIRE(sms > 0, SRB_ENODEVICE, "no CUDA device");
    const int G = W * S * M, row = 2 * D + 3;
    SRB_REQUIRE(tile_frames == ETILE_MAX || tile_frames == ETILE_MAX / 2, SRB_EINVAL,
                "srb_em_isolated: tile_frames must come from srb_em_plan");
    const int nparts = sms * EPARTS;
    cudaStream_t st = (cudaStream_t)stream;
    SRB_REQUIRE(!partial || W <= 64, SRB_EUNSUPPORTED, "srb_em_isolated: statistics for at most 64 word models per call");
    // the masks of ALL slices are cleared (a launch may use fewer CTAs than there are slices); the slices themselves
    // are not: a CTA writes the rows of the words it saw, the reducers read nothing else
    unsigned long long *slice_mask = partial ? reinterpret_cast<unsigned long long *>(partial + (size_t)nparts * G * row)
                                             : nullptr;
    if (partial) SRB_CUDA(cudaMemsetAsync(slice_mask, 0, sizeof(unsigned long long) * (size_t)nparts, st));
    if (B > 0 && ntiles > 0) {
        const EmSmem L = em_layout(S, Mp, ncomp, D, partial != nullptr, tile_frames);
        const size_t bytes = sizeof(float) * (size_t)L.total;
        SRB_REQUIRE(bytes <= (size_t)smem_optin(), SRB_EUNSUPPORTED, "srb_em_isolated: %zu B shared memory needed", bytes);
        const int ctas = sms * (512 / tile_frames);
        const int grid = ntiles < ctas ? ntiles : ctas;
#define SRB_EM_LAUNCH_T(kD, kS, kC, kT)                                                                            \
    {                                                                                                              \
        SRB_CUDA(cudaFuncSetAttribute(em_isolated_kernel<kD, kS, kC, kT>,                                          \
                                      cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));                   \
        em_isolated_kernel<kD, kS, kC, kT><<<grid, kT, bytes, st>>>(                                               \
            X, N, utt_off, reinterpret_cast<const int4 *>(tile_desc), ntiles, ab, c, shift, S, M, Mp, trans, ncomp, mu, \
            partial, slice_mask, G, best_cost, seg_start, frame_label);                                                        \
    }
#define SRB_EM_LAUNCH(kD, kS, kC)                                                                                  \
    {                                                                                                              \
        if (tile_frames == 128) SRB_EM_LAUNCH_T(kD, kS, kC, 128) else SRB_EM_LAUNCH_T(kD, kS, kC, 256)             \
    }
        // the 5-state / 4-Gaussian model of the reference's scripts (config.py, core.py:48; hmm.py:104 trains 2
        // components) gets a fully static instantiation
        if (D == 39 && S == 5 && ncomp == 2) SRB_EM_LAUNCH(39, 5, 2)
        else if (D == 39 && S == 5) SRB_EM_LAUNCH(39, 5, 0)
        else if (D == 39 && S == 8) SRB_EM_LAUNCH(39, 8, 0)
        else if (D == 39) SRB_EM_LAUNCH(39, 0, 0)
        else if (D == 26) SRB_EM_LAUNCH(26, 0, 0)
        else if (D == 13) SRB_EM_LAUNCH(13, 0, 0)
        else {
            set_error("feature dimension %d not built (13, 26, 39)", D);
            return SRB_EUNSUPPORTED;
        }
#undef SRB_EM_LAUNCH
#undef SRB_EM_LAUNCH_T
    }
    if (partial && accumulate >= 0) {                        // accumulate < 0: the slices are left for srb_em_push
        const int64_t len = (int64_t)G * row;
        reduce_slices_kernel<<<(int)((len + 255) / 256), 256, 0, st>>>(partial, slice_mask, nparts, len, S * M * row, M, ncomp,
                                                                       row, stats_out, accumulate);
    }
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // extern "C"
