// classify_dp.cuh -- Viterbi phase of the fused isolated-word classifiers (score.cu: FFMA scorer, score_tc.cu:
// tcgen05 scorer).  Advances the per-(utterance, word) recursions of decode.py:97-124 (costs only) over one chunk of
// frames whose state costs E[frame - cf0][w*S + s] sit in shared memory; thread = (piece, word), states in registers.
// The costs of the utterance that straddles the chunk end are carried in shared memory (carry[par] in, carry[par^1]
// out).  Finished utterances write costs[-1,-1] per word (hmm.py:135) and, once all words are in, the arg-min label
// (core.py:82-87: strict '<', first model wins; fin_s, optional, is shared scratch of fin_cap x W floats that saves the
// round trip through L2 for the first fin_cap utterances of the chunk).  Must be called by the NTH threads (tid = 0 .. NTH-1) that share
// named barrier BAR_ID (the whole CTA with barrier 0 in score.cu) or, with BAR_ID < 0 and NTH = 32, by one warp (each
// Viterbi warp of score_tc.cu owns a sub-stream of utterances).
#pragma once
#include "common.cuh"

namespace srb {

__device__ __forceinline__ void named_sync(int id, int count) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory"); }
__device__ __forceinline__ void named_arrive(int id, int count) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(count) : "memory"); }

// SN live states of a left-to-right model with skips <= 2 (kmeans.py:53-95), nsteps frames; the next frame's costs
// are loaded while the current one is folded in.
template <int SN>
__device__ __forceinline__ void dp_banded_steps(float (&cost)[SRB_DP_MAX_STATES], const float *Erow, int ES, int nsteps,
                                                const float *trw) {
    constexpr int SMAX = SRB_DP_MAX_STATES;
    float a0[SN], a1[SN], a2[SN], e[SN];
#pragma unroll
    for (int r = 0; r < SN; ++r) {
        a0[r] = trw[r * SMAX + r];
        a1[r] = r >= 1 ? trw[r * SMAX + r - 1] : CUDART_INF_F;
        a2[r] = r >= 2 ? trw[r * SMAX + r - 2] : CUDART_INF_F;
        e[r] = nsteps > 0 ? Erow[r] : 0.f;
    }
    for (int n = 0; n < nsteps; ++n) {
        Erow += ES;
        float en[SN];
        const bool more = n + 1 < nsteps;
#pragma unroll
        for (int r = 0; r < SN; ++r) en[r] = more ? Erow[r] : 0.f;
#pragma unroll
        for (int r = SN - 1; r >= 0; --r) {
            float best = cost[r] + a0[r];
            if (r >= 1) best = fminf(best, cost[r - 1] + a1[r]);
            if (r >= 2) best = fminf(best, cost[r - 2] + a2[r]);
            cost[r] = best + e[r];
        }
#pragma unroll
        for (int r = 0; r < SN; ++r) e[r] = en[r];
    }
}

template <int NTH, int BAR_ID = 0>
__device__ __forceinline__ void classify_dp_chunk(const float *Es, int ES, const int32_t *__restrict__ off, int ue,
                                                  int &ufirst, int64_t cf0, int64_t cf1, int W, int S,
                                                  const float *tr_s, float *carry, int par, bool banded,
                                                  float *__restrict__ final_cost, int32_t *__restrict__ label,
                                                  int tid, float *fin_s = nullptr, int fin_cap = 0) {
    constexpr int SMAX = SRB_DP_MAX_STATES;
    while (ufirst + 1 < ue && (int64_t)off[ufirst + 1] <= cf0) ++ufirst;   // first utterance with frames >= cf0
    int ulast = ufirst;
    {
        // first u' with off >= cf1: utterances are not empty, so it is at most cf1 - cf0 + 1 entries ahead
        const int nfull = ue - ufirst + 1, nnear = (int)min((int64_t)nfull, cf1 - cf0 + 2);
        int nb = lower_bound_i32(off + ufirst, nnear, cf1);
        if (nb == nnear && nnear < nfull) nb = lower_bound_i32(off + ufirst, nfull, cf1);   // empty utterances in between
        ulast = ufirst + nb - 1;
        if (ulast >= ue) ulast = ue - 1;
    }
    const int npieces = ulast - ufirst + 1;
    for (int pw = tid; pw < npieces * W; pw += NTH) {
        const int p = pw / W, w = pw - p * W;
        const int u = ufirst + p;
        const int64_t us = off[u], uend = off[u + 1];
        const int64_t t0 = max(us, cf0), t1 = min(uend, cf1);
        if (t0 >= t1) continue;
        float cost[SMAX];
        const float *Erow = Es + (size_t)(t0 - cf0) * ES + w * S;
        int64_t t = t0;
        if (t0 == us) {   // decode.py:99-100: only (state 0, frame 0) is seeded
#pragma unroll
            for (int r = 0; r < SMAX; ++r) cost[r] = CUDART_INF_F;
            cost[0] = Erow[0];
            ++t;
            Erow += ES;
        } else {
#pragma unroll
            for (int r = 0; r < SMAX; ++r) cost[r] = carry[(par * W + w) * SMAX + r];
        }
        const float *trw = tr_s + w * SMAX * SMAX;
        if (banded && S == 5) {                   // the reference's 5-state models (config.py), fully static
            dp_banded_steps<5>(cost, Erow, ES, (int)(t1 - t), trw);
        } else if (banded) {
            float a0[SMAX], a1[SMAX], a2[SMAX];
#pragma unroll
            for (int r = 0; r < SMAX; ++r) {
                a0[r] = trw[r * SMAX + r];
                a1[r] = r >= 1 ? trw[r * SMAX + r - 1] : CUDART_INF_F;
                a2[r] = r >= 2 ? trw[r * SMAX + r - 2] : CUDART_INF_F;
            }
            for (; t < t1; ++t, Erow += ES) {
#pragma unroll
                for (int r = SMAX - 1; r >= 0; --r) {
                    if (r < S) {
                        float best = cost[r] + a0[r];
                        if (r >= 1) best = fminf(best, cost[r - 1] + a1[r]);
                        if (r >= 2) best = fminf(best, cost[r - 2] + a2[r]);
                        cost[r] = best + Erow[r];
                    }
                }
            }
        } else {
            for (; t < t1; ++t, Erow += ES) {
                float nc[SMAX];
#pragma unroll
                for (int r = 0; r < SMAX; ++r) {
                    float best = CUDART_INF_F;
#pragma unroll
                    for (int o = 0; o < SMAX; ++o) best = fminf(best, cost[o] + trw[r * SMAX + o]);
                    nc[r] = (r < S) ? best + Erow[r] : CUDART_INF_F;
                }
#pragma unroll
                for (int r = 0; r < SMAX; ++r) cost[r] = nc[r];
            }
        }
        if (t1 == uend) {
            float fc = CUDART_INF_F;
#pragma unroll
            for (int r = 0; r < SMAX; ++r)
                if (r == S - 1) fc = cost[r];
            final_cost[(size_t)u * W + w] = fc;           // hmm.py:135 costs[-1,-1]
            if (p < fin_cap) fin_s[p * W + w] = fc;       // shared copy for the arg-min below
        } else {
#pragma unroll
            for (int r = 0; r < SMAX; ++r) carry[((par ^ 1) * W + w) * SMAX + r] = cost[r];
        }
    }
    if (BAR_ID < 0) __syncwarp(); else named_sync(BAR_ID, NTH);   // BAR_ID < 0: the callers are one warp
    if (label) {   // core.py:82-87: strict '<', first model wins
        for (int p = tid; p < npieces; p += NTH) {
            const int u = ufirst + p;
            const int64_t uend = off[u + 1];
            if (uend > cf0 && uend <= cf1) {
                float bcst = CUDART_INF_F;
                int bl = 0;
                for (int w = 0; w < W; ++w) {
                    const float v = p < fin_cap ? fin_s[p * W + w] : __ldcg(final_cost + (size_t)u * W + w);
                    if (v < bcst) {
                        bcst = v;
                        bl = w;
                    }
                }
                label[u] = bl;
            }
        }
    }
    ufirst = ulast;
}

}  // namespace srb
