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
// Two CTAs are resident per SM so that one CTA's serial Viterbi phase overlaps the other's FMA-bound scoring.
#include "score_core.cuh"

namespace srb {

constexpr int ENT = 128;             // threads per CTA
constexpr int ETILE = ENT * FR;      // frames per tile
constexpr int EWARPS = ENT / 32;
constexpr int ESMAX = SRB_DP_MAX_STATES;
constexpr int EMAXC = 8;             // trained components per state handled in registers

struct EmSmem {                      // offsets in floats into the dynamic shared memory
    int tr, pc, pab, mean, xs, es, ps, ll, part, bnd, uoff, lab, total;
};

__host__ __device__ inline EmSmem em_layout(int S, int Mp, int ncomp, int D, bool stats) {
    EmSmem L;
    int o = 64;                                           // shift_s
    L.tr = o;   o += ESMAX * ESMAX;
    L.pc = o;   o += (S * Mp + 3) & ~3;
    L.pab = o;  o += S * Mp * D * 2;
    L.mean = o; o += (S * ncomp * D + 3) & ~3;
    L.xs = o;   o += ETILE * D + 4;
    L.es = o;   o += ETILE * (S | 1);
    L.ps = o;   o += ETILE * ((S * Mp) | 1);
    L.ll = o;   o += ETILE * (S | 1);
    L.part = o; o += stats ? EWARPS * S * ncomp * (2 * D + 3) : 0;
    L.bnd = o;  o += ENT * ESMAX;
    L.uoff = o; o += ENT + 4;
    L.lab = o;  o += ETILE / 4;
    L.total = o;
    return L;
}

template <int D, int ST>
__global__ void __launch_bounds__(ENT, 2)
em_isolated_kernel(const float *__restrict__ X, int64_t N, const int32_t *__restrict__ off,
                   const int32_t *__restrict__ word, const int32_t *__restrict__ tile_u, int ntiles,
                   const float *__restrict__ ab, const float *__restrict__ c, const float *__restrict__ shift, int S_rt,
                   int M, int Mp, const float *__restrict__ trans, int ncomp, const double *__restrict__ mu,
                   float *__restrict__ partial, int G, float *__restrict__ best_cost, int32_t *__restrict__ seg_start,
                   int32_t *__restrict__ frame_label) {
    extern __shared__ __align__(16) float smem[];
    const int S = ST ? ST : S_rt;
    const bool stats = partial != nullptr;
    const EmSmem L = em_layout(S, Mp, ncomp, D, stats);
    float *shift_s = smem, *tr_s = smem + L.tr, *pc = smem + L.pc, *pab = smem + L.pab, *mean_s = smem + L.mean;
    float *xs = smem + L.xs, *Es = smem + L.es, *Ps = smem + L.ps, *LLs = smem + L.ll, *part = smem + L.part;
    int *bnd = reinterpret_cast<int *>(smem + L.bnd), *uoff = reinterpret_cast<int *>(smem + L.uoff);
    unsigned char *lab = reinterpret_cast<unsigned char *>(smem + L.lab);
    const int ES = S | 1, PS = (S * Mp) | 1, row = 2 * D + 3, GW = S * ncomp;
    const int tid = threadIdx.x, lane = tid & 31, wq = tid >> 5;
    const int nblk = Mp >> 2;
    const float INF = CUDART_INF_F;

    if (tid < D) shift_s[tid] = shift[tid];
    // contiguous range of tiles per CTA: a CTA sees one or two words, so parameters are (re)loaded rarely
    const int t_lo = (int)((int64_t)ntiles * blockIdx.x / gridDim.x);
    const int t_hi = (int)((int64_t)ntiles * (blockIdx.x + 1) / gridDim.x);
    float *slice = stats ? partial + (size_t)blockIdx.x * G * row : nullptr;
    int cur_w = -1;

    for (int t = t_lo; t < t_hi; ++t) {
        const int u0 = tile_u[t], u1 = tile_u[t + 1], nutt = u1 - u0;
        const int64_t f0 = off[u0];
        const int nf = (int)(off[u1] - f0);
        const int w = word[u0];
        __syncthreads();                                   // previous tile fully consumed
        if (w != cur_w) {
            if (stats && cur_w >= 0) {                     // flush the finished word's statistics (fixed order)
                for (int i = tid; i < GW * row; i += ENT) {
                    float s = 0.f;
#pragma unroll
                    for (int q = 0; q < EWARPS; ++q) s += part[q * GW * row + i];
                    const int g = i / row, j = i - g * row;
                    slice[((size_t)(cur_w * S + g / ncomp) * M + g % ncomp) * row + j] += s;
                }
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
        const int xoff = stage_frames<D, ENT>(xs, X, N, f0, nf);
        __syncthreads();

        // ---- 1. scoring: E[f][s], natural-log component scores Ps[f][s*Mp + m]
        {
            Frames<D> fr;
            load_frames<D, ENT>(fr, xs, xoff, nf, shift_s);
            for (int s = 0; s < S; ++s) {
                float cost[FR];
                float *crow[FR];
#pragma unroll
                for (int k = 0; k < FR; ++k) crow[k] = Ps + (size_t)(tid + k * ENT) * PS + s * Mp;
                score_state<D>(fr, pab + (size_t)s * Mp * D * 2, pc + s * Mp, nblk, cost, crow, true);
#pragma unroll
                for (int k = 0; k < FR; ++k) Es[(tid + k * ENT) * ES + s] = cost[k];
            }
            // posteriors over the first ncomp components (hmm_state.py:126-136, log domain); thread-private rows
#pragma unroll
            for (int k = 0; k < FR; ++k) {
                float *prow = Ps + (size_t)(tid + k * ENT) * PS;
                for (int s = 0; s < S; ++s) {
                    float mx = -INF;
                    for (int m = 0; m < ncomp; ++m) mx = fmaxf(mx, prow[s * Mp + m]);
                    float den = 0.f;
                    for (int m = 0; m < ncomp; ++m) {
                        const float e = (mx > -INF) ? __expf(prow[s * Mp + m] - mx) : 0.f;
                        prow[s * Mp + m] = e;
                        den += e;
                    }
                    const float inv = den > 0.f ? 1.f / den : 0.f;
                    for (int m = 0; m < ncomp; ++m) prow[s * Mp + m] *= inv;
                    LLs[(tid + k * ENT) * ES + s] = den > 0.f ? mx + __logf(den) : 0.f;
                }
            }
        }
        __syncthreads();

        // ---- 2. Viterbi, one thread per utterance, threads spread over the warps (decode.py:97-134 for a
        //         left-to-right model; ties: lowest origin wins like np.argmin)
        for (int p = (tid & 31) * EWARPS + wq; p < nutt; p += ENT) {
            const int ts = uoff[p], T = uoff[p + 1] - ts;
            float cost[ESMAX], a0[ESMAX], a1[ESMAX], a2[ESMAX];
            int bd[ESMAX][ESMAX];
#pragma unroll
            for (int r = 0; r < ESMAX; ++r) {
                cost[r] = INF;
                a0[r] = tr_s[r * ESMAX + r];
                a1[r] = r >= 1 ? tr_s[r * ESMAX + r - 1] : INF;
                a2[r] = r >= 2 ? tr_s[r * ESMAX + r - 2] : INF;
#pragma unroll
                for (int k = 0; k < ESMAX; ++k) bd[r][k] = 0;
            }
            const float *Erow = Es + (size_t)ts * ES;
            if (T > 0) cost[0] = Erow[0];
            for (int tt = 1; tt < T; ++tt) {
                Erow += ES;
#pragma unroll
                for (int r = ESMAX - 1; r >= 0; --r) {
                    if (r < S) {
                        float best;
                        int src;
                        if (r >= 2) {
                            best = cost[r - 2] + a2[r];
                            src = 2;
                            const float c1 = cost[r - 1] + a1[r];
                            if (c1 < best) { best = c1; src = 1; }
                        } else if (r == 1) {
                            best = cost[0] + a1[1];
                            src = 1;
                        } else {
                            best = INF;
                            src = 0;
                        }
                        const float c0 = cost[r] + a0[r];
                        if (r == 0 || c0 < best) { best = c0; src = 0; }
                        cost[r] = best + Erow[r];
#pragma unroll
                        for (int k = 1; k < ESMAX; ++k) {
                            if (k <= r) {
                                int v = bd[r][k];
                                if (src == 1) v = (k <= r - 1) ? bd[r >= 1 ? r - 1 : 0][k] : tt;
                                if (src == 2) v = (k <= r - 2) ? bd[r >= 2 ? r - 2 : 0][k] : tt;
                                bd[r][k] = v;
                            }
                        }
                    }
                }
            }
            float fc = INF;
#pragma unroll
            for (int r = 0; r < ESMAX; ++r)
                if (r == S - 1) {
                    fc = cost[r];
#pragma unroll
                    for (int k = 1; k < ESMAX; ++k)
                        if (k <= r) bnd[p * ESMAX + k] = bd[r][k];
                }
            bnd[p * ESMAX] = (fc < INF) ? 0 : -1;          // -1: no path, the utterance's frames are skipped
            if (best_cost) best_cost[u0 + p] = fc;
            if (seg_start) {
                for (int k = 0; k < S; ++k) seg_start[(size_t)(u0 + p) * S + k] = (fc < INF && k) ? bnd[p * ESMAX + k] : 0;
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
        }
        if (!stats) continue;                              // uniform per CTA: alignment-only call
        __syncthreads();

        // ---- 4. statistics: warp = 64-frame chunk, lanes over dims, runs of equal state in registers
        {
            const int d0 = lane, d1 = lane + 32;
            const bool h0 = d0 < D, h1 = d1 < D;
            const int c0 = wq * (ETILE / EWARPS), c1 = min(nf, c0 + ETILE / EWARPS);
            float *mine = part + (size_t)wq * GW * row;
            float m0[EMAXC], m1[EMAXC], s0[EMAXC], a0[EMAXC], a1[EMAXC], q0[EMAXC], q1[EMAXC];
            int curs = -1;
            float cnt = 0.f, llsum = 0.f;
            for (int f = c0; f <= c1 && c0 < c1; ++f) {
                int s = (f < c1) ? (int)lab[f] : -2;
                if (s == 255) s = -1;
                if (s != curs) {
                    if (curs >= 0) {
#pragma unroll
                        for (int m = 0; m < EMAXC; ++m) {
                            if (m < ncomp) {
                                float *r = mine + (size_t)(curs * ncomp + m) * row;
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
                    curs = s;
                    cnt = llsum = 0.f;
                    if (s >= 0) {
#pragma unroll
                        for (int m = 0; m < EMAXC; ++m) {
                            s0[m] = a0[m] = a1[m] = q0[m] = q1[m] = 0.f;
                            if (m < ncomp) {
                                m0[m] = h0 ? mean_s[(s * ncomp + m) * D + d0] : 0.f;
                                m1[m] = h1 ? mean_s[(s * ncomp + m) * D + d1] : 0.f;
                            }
                        }
                    }
                }
                if (s >= 0) {
                    const float *x = xs + xoff + f * D;
                    const float xa = h0 ? x[d0] : 0.f, xb = h1 ? x[d1] : 0.f;
                    const float *prow = Ps + (size_t)f * PS + s * Mp;
                    cnt += 1.f;
                    llsum += LLs[f * ES + s];
#pragma unroll
                    for (int m = 0; m < EMAXC; ++m)
                        if (m < ncomp) {
                            const float p = prow[m];
                            const float e0 = xa - m0[m], e1 = xb - m1[m];
                            s0[m] += p;
                            a0[m] = fmaf(p, e0, a0[m]);
                            a1[m] = fmaf(p, e1, a1[m]);
                            q0[m] = fmaf(p * e0, e0, q0[m]);
                            q1[m] = fmaf(p * e1, e1, q1[m]);
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
            slice[((size_t)(cur_w * S + g / ncomp) * M + g % ncomp) * row + j] += s;
        }
    }
}

__global__ void reduce_slices_kernel(const float *__restrict__ partial, int64_t nparts, int64_t len,
                                     double *__restrict__ out, int accumulate) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= len) return;
    double s = accumulate ? out[i] : 0.0;
    for (int64_t p = 0; p < nparts; ++p) s += (double)partial[p * len + i];
    out[i] = s;
}

}  // namespace srb

using namespace srb;

extern "C" {

int srb_em_tile_frames(void) { return ETILE; }

int srb_em_parts(void) { return sm_count() * 2; }

int srb_em_plan(const int32_t *utt_off, const int32_t *utt_word, int32_t B, int32_t *tile_first_utt, int32_t *ntiles) {
    SRB_REQUIRE(utt_off && utt_word && tile_first_utt && ntiles && B >= 0, SRB_EINVAL, "srb_em_plan: bad argument");
    int nt = 0;
    int32_t u = 0;
    while (u < B) {
        SRB_REQUIRE(u == 0 || utt_word[u] >= utt_word[u - 1], SRB_EINVAL,
                    "srb_em_plan: utterances must be sorted by word (utterance %d)", u);
        tile_first_utt[nt++] = u;
        const int64_t start = utt_off[u];
        int32_t v = u;
        while (v < B && utt_word[v] == utt_word[u] && (int64_t)utt_off[v + 1] - start <= ETILE && v - u < ENT) {
            SRB_REQUIRE(utt_off[v + 1] > utt_off[v], SRB_EUNSUPPORTED, "srb_em_plan: empty utterance %d", v);
            ++v;
        }
        SRB_REQUIRE(v > u, SRB_EUNSUPPORTED, "srb_em_plan: utterance %d has more than %d frames", u, ETILE);
        u = v;
    }
    tile_first_utt[nt] = B;
    *ntiles = nt;
    return SRB_OK;
}

int srb_em_isolated(const float *X, int64_t N, int D, const int32_t *utt_off, const int32_t *utt_word, int32_t B,
                    const int32_t *tile_first_utt, int32_t ntiles, const float *ab, const float *c, const float *shift,
                    int W, int S, int M, int Mp, const float *trans, int ncomp, const double *mu, float *partial,
                    double *stats_out, int accumulate, float *best_cost, int32_t *seg_start, int32_t *frame_label,
                    void *stream) {
    SRB_REQUIRE(X && utt_off && utt_word && tile_first_utt && ab && c && shift && trans, SRB_EINVAL,
                "srb_em_isolated: null pointer");
    SRB_REQUIRE(!partial || (stats_out && mu), SRB_EINVAL, "srb_em_isolated: statistics need stats_out and mu");
    SRB_REQUIRE(W > 0 && S > 0 && S <= ESMAX && M > 0 && Mp >= M && (Mp & 3) == 0, SRB_EUNSUPPORTED,
                "srb_em_isolated: S=%d (max %d), M=%d, Mp=%d", S, ESMAX, M, Mp);
    SRB_REQUIRE(ncomp >= 1 && ncomp <= M && ncomp <= EMAXC, SRB_EUNSUPPORTED, "srb_em_isolated: ncomp=%d", ncomp);
    SRB_REQUIRE(((uintptr_t)X & 15) == 0 && ((uintptr_t)ab & 15) == 0, SRB_EINVAL, "srb_em_isolated: alignment");
    const int sms = sm_count();
    SRB_REQUIRE(sms > 0, SRB_ENODEVICE, "no CUDA device");
    const int G = W * S * M, row = 2 * D + 3;
    const int nparts = sms * 2;
    cudaStream_t st = (cudaStream_t)stream;
    if (partial) SRB_CUDA(cudaMemsetAsync(partial, 0, sizeof(float) * (size_t)nparts * G * row, st));
    if (B > 0 && ntiles > 0) {
        const EmSmem L = em_layout(S, Mp, ncomp, D, partial != nullptr);
        const size_t bytes = sizeof(float) * (size_t)L.total;
        SRB_REQUIRE(bytes <= (size_t)smem_optin(), SRB_EUNSUPPORTED, "srb_em_isolated: %zu B shared memory needed", bytes);
        const int grid = ntiles < nparts ? ntiles : nparts;
#define SRB_EM_LAUNCH(kD, kS)                                                                                      \
    {                                                                                                              \
        SRB_CUDA(cudaFuncSetAttribute(em_isolated_kernel<kD, kS>, cudaFuncAttributeMaxDynamicSharedMemorySize,     \
                                      (int)bytes));                                                                \
        em_isolated_kernel<kD, kS><<<grid, ENT, bytes, st>>>(X, N, utt_off, utt_word, tile_first_utt, ntiles, ab,  \
                                                             c, shift, S, M, Mp, trans, ncomp, mu, partial, G,     \
                                                             best_cost, seg_start, frame_label);                   \
    }
        // the 5-state model of the reference's scripts (config.py, core.py:48) gets a fully static instantiation
        if (D == 39 && S == 5) SRB_EM_LAUNCH(39, 5)
        else if (D == 39 && S == 8) SRB_EM_LAUNCH(39, 8)
        else if (D == 39) SRB_EM_LAUNCH(39, 0)
        else if (D == 26) SRB_EM_LAUNCH(26, 0)
        else if (D == 13) SRB_EM_LAUNCH(13, 0)
        else {
            set_error("feature dimension %d not built (13, 26, 39)", D);
            return SRB_EUNSUPPORTED;
        }
#undef SRB_EM_LAUNCH
    }
    if (partial) {
        const int64_t len = (int64_t)G * row;
        reduce_slices_kernel<<<(int)((len + 255) / 256), 256, 0, st>>>(partial, nparts, len, stats_out, accumulate);
    }
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // extern "C"
