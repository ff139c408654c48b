// classify_dp.cuh -- Viterbi phase of the fused isolated-word classifiers (score.cu: FFMA scorer, score_tc.cu:
// tcgen05 scorer).  Advances the per-(utterance, word) recursions of decode.py:97-124 (costs only) over one chunk of
// frames whose state costs E[frame - cf0][w*S + s] sit in shared memory; thread = (piece, word), states in registers.
// The costs of the utterance that straddles the chunk end are carried in shared memory (carry[par] in, carry[par^1]
// out).  Finished utterances write costs[-1,-1] per word (hmm.py:135) and, once all words are in, the arg-min label
// (core.py:82-87: strict '<', first model wins).  Must be called by all NTH threads of the CTA.
#pragma once
#include "common.cuh"

namespace srb {

template <int NTH>
__device__ __forceinline__ void classify_dp_chunk(const float *Es, int ES, const int32_t *__restrict__ off, int ue,
                                                  int &ufirst, int64_t cf0, int64_t cf1, int W, int S,
                                                  const float *tr_s, float *carry, int par, bool banded,
                                                  float *__restrict__ final_cost, int32_t *__restrict__ label) {
    constexpr int SMAX = SRB_DP_MAX_STATES;
    while (ufirst + 1 < ue && (int64_t)off[ufirst + 1] <= cf0) ++ufirst;   // first utterance with frames >= cf0
    int ulast = ufirst;
    {
        const int nb = lower_bound_i32(off + ufirst, ue - ufirst + 1, cf1);   // first u' with off >= cf1
        ulast = ufirst + nb - 1;
        if (ulast >= ue) ulast = ue - 1;
    }
    const int npieces = ulast - ufirst + 1;
    for (int pw = threadIdx.x; pw < npieces * W; pw += NTH) {
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
        if (banded) {
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
        } else {
#pragma unroll
            for (int r = 0; r < SMAX; ++r) carry[((par ^ 1) * W + w) * SMAX + r] = cost[r];
        }
    }
    __syncthreads();
    if (label) {   // core.py:82-87: strict '<', first model wins
        for (int p = threadIdx.x; p < npieces; p += NTH) {
            const int u = ufirst + p;
            const int64_t uend = off[u + 1];
            if (uend > cf0 && uend <= cf1) {
                float bcst = CUDART_INF_F;
                int bl = 0;
                for (int w = 0; w < W; ++w) {
                    const float v = __ldcg(final_cost + (size_t)u * W + w);
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
