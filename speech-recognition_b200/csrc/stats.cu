// stats.cu -- K5: sufficient-statistic accumulation (hard alignment + within-state posteriors, GMM.em
// hmm_state.py:122-148; segment / cluster moments, kmeans.py:6-30,173-187), K6: K-means assignment,
// the deterministic partial-sum reduction and the M-step.
//
// Accumulation layout: one warp walks a contiguous chunk of frames with the LANES OVER THE FEATURE DIMENSIONS
// (d = lane and lane + 32, D <= 64).  Frames aligned to the same group (tied state / segment / cluster) form
// runs; the moments of the current run live in registers and are flushed to the warp's PRIVATE slice of a
// partial buffer when the group changes -- no atomics anywhere.  A second kernel adds the slices in a fixed
// order in double, which makes the result independent of scheduling (and ready for the NCCL allreduce).
//
// Moments are taken about a per-group shift vector (the Gaussian's current mean / the current centroid) so
// that  var = S2/S0 - (S1/S0)^2  does not cancel in FP32:  S0 = sum g, S1 = sum g (x - shift), S2 = sum g (x - shift)^2.
// Row layout of one group in every buffer:  [S0, S1[0..D), S2[0..D), count, loglik]  = 2*D + 3 values
// (count / loglik: frames and summed log-likelihood of the state, kept on its component 0 only).
#include <cstdlib>
#include <mutex>

#include "common.cuh"

namespace srb {

constexpr int SW = 8;          // warps per CTA
constexpr int SCH = 128;       // frames per warp chunk
constexpr int SCH_SOFT = 32;   // soft counts: no runs to keep, shorter chunks balance the ~3 chunks a warp gets per word
constexpr int MAXC = 8;        // components per state handled in registers by the GMM mode

__device__ __forceinline__ float wsum(float v) {
#pragma unroll
    for (int d = 16; d; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    return v;
}

// ---- group moments: label[n] in [0, K) (or < 0: skip), weight 1
__global__ void __launch_bounds__(SW * 32)
group_stats_kernel(const float *__restrict__ X, int64_t N, int D, const int32_t *__restrict__ label, int K,
                   const double *__restrict__ shift, float *__restrict__ partial) {
    const int lane = threadIdx.x & 31;
    const int64_t wg = (int64_t)blockIdx.x * SW + (threadIdx.x >> 5), nw = (int64_t)gridDim.x * SW;
    const int row = 2 * D + 3;
    float *mine = partial + (size_t)wg * K * row;
    const int d0 = lane, d1 = lane + 32;
    const bool h0 = d0 < D, h1 = d1 < D;
    for (int64_t c0 = wg * SCH; c0 < N; c0 += nw * SCH) {
        const int64_t c1 = min(N, c0 + SCH);
        int curg = -1;
        float s0 = 0.f, a0 = 0.f, a1 = 0.f, q0 = 0.f, q1 = 0.f, sh0 = 0.f, sh1 = 0.f;
        for (int64_t n = c0; n <= c1; ++n) {
            const int g = (n < c1) ? label[n] : -2;
            if (g != curg) {
                if (curg >= 0) {   // flush the run
                    float *r = mine + (size_t)curg * row;
                    if (lane == 0) r[0] += s0;
                    if (h0) {
                        r[1 + d0] += a0;
                        r[1 + D + d0] += q0;
                    }
                    if (h1) {
                        r[1 + d1] += a1;
                        r[1 + D + d1] += q1;
                    }
                }
                curg = g;
                s0 = a0 = a1 = q0 = q1 = 0.f;
                if (g >= 0) {
                    sh0 = h0 ? (float)shift[(size_t)g * D + d0] : 0.f;
                    sh1 = h1 ? (float)shift[(size_t)g * D + d1] : 0.f;
                }
            }
            if (g >= 0) {
                const float *x = X + n * D;
                const float x0 = h0 ? x[d0] - sh0 : 0.f, x1 = h1 ? x[d1] - sh1 : 0.f;
                s0 += 1.f;
                a0 += x0;
                a1 += x1;
                q0 = fmaf(x0, x0, q0);
                q1 = fmaf(x1, x1, q1);
            }
        }
    }
}

// ---- GMM mode: label[n] = tied state s; groups are the first ncomp components of s, weights are the
// within-state posteriors  p_m = w_m N_m(x) / sum_{m' < ncomp} w_m' N_m'(x)  (hmm_state.py:126-136, log domain).
__global__ void __launch_bounds__(SW * 32)
gmm_stats_kernel(const float *__restrict__ X, int64_t N, int D, const int32_t *__restrict__ label, int M, int ncomp,
                 const double *__restrict__ mu, const double *__restrict__ var, const double *__restrict__ lc,
                 float *__restrict__ partial, int G) {
    const int lane = threadIdx.x & 31;
    const int64_t wg = (int64_t)blockIdx.x * SW + (threadIdx.x >> 5), nw = (int64_t)gridDim.x * SW;
    const int row = 2 * D + 3;
    float *mine = partial + (size_t)wg * G * row;
    const int d0 = lane, d1 = lane + 32;
    const bool h0 = d0 < D, h1 = d1 < D;
    float m0[MAXC], m1[MAXC], v0[MAXC], v1[MAXC], cst[MAXC];
    float s0[MAXC], a0[MAXC], a1[MAXC], q0[MAXC], q1[MAXC];
    for (int64_t c0 = wg * SCH; c0 < N; c0 += nw * SCH) {
        const int64_t c1 = min(N, c0 + SCH);
        int curs = -1;
        float cnt = 0.f, llsum = 0.f;
        for (int64_t n = c0; n <= c1; ++n) {
            const int s = (n < c1) ? label[n] : -2;
            if (s != curs) {
                if (curs >= 0) {
#pragma unroll
                    for (int m = 0; m < MAXC; ++m) {
                        if (m < ncomp) {
                            float *r = mine + ((size_t)curs * M + m) * row;
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
                    for (int m = 0; m < MAXC; ++m) {
                        s0[m] = a0[m] = a1[m] = q0[m] = q1[m] = 0.f;
                        if (m < ncomp) {
                            const size_t g = (size_t)s * M + m;
                            m0[m] = h0 ? (float)mu[g * D + d0] : 0.f;
                            m1[m] = h1 ? (float)mu[g * D + d1] : 0.f;
                            v0[m] = h0 ? (float)(-0.5 / var[g * D + d0]) : 0.f;
                            v1[m] = h1 ? (float)(-0.5 / var[g * D + d1]) : 0.f;
                            cst[m] = (float)lc[g];
                        }
                    }
                }
            }
            if (s >= 0) {
                const float *x = X + n * D;
                const float xa = h0 ? x[d0] : 0.f, xb = h1 ? x[d1] : 0.f;
                float ll[MAXC], e0[MAXC], e1[MAXC];
                float mx = -CUDART_INF_F;
#pragma unroll
                for (int m = 0; m < MAXC; ++m) {
                    if (m < ncomp) {
                        e0[m] = xa - m0[m];
                        e1[m] = xb - m1[m];
                        const float part = e0[m] * e0[m] * v0[m] + e1[m] * e1[m] * v1[m];
                        ll[m] = wsum(part) + cst[m];
                        mx = fmaxf(mx, ll[m]);
                    }
                }
                float den = 0.f;
#pragma unroll
                for (int m = 0; m < MAXC; ++m)
                    if (m < ncomp) {
                        ll[m] = __expf(ll[m] - mx);
                        den += ll[m];
                    }
                const float inv = 1.f / den;
                cnt += 1.f;
                llsum += mx + __logf(den);
#pragma unroll
                for (int m = 0; m < MAXC; ++m)
                    if (m < ncomp) {
                        const float p = ll[m] * inv;
                        s0[m] += p;
                        a0[m] = fmaf(p, e0[m], a0[m]);
                        a1[m] = fmaf(p, e1[m], a1[m]);
                        q0[m] = fmaf(p * e0[m], e0[m], q0[m]);
                        q1[m] = fmaf(p * e1[m], e1[m], q1[m]);
                    }
            }
        }
    }
}

// ---- soft mode (Baum-Welch): frames [row0, row0+nrows) belong to ONE word whose S states carry the posteriors
// gamma[n, s] (K4).  Group (s, m) receives weight gamma[n,s] * p_m with p_m the within-state posterior over the
// first ncomp components (same log-domain arithmetic as gmm_stats_kernel).  States with gamma below `gmin` are
// skipped (a left-to-right model puts almost all mass on a few states per frame).
// A warp walks a chunk of frames.  Its lanes are NCP = pow2(ncomp) groups of LPC = 32 / NCP lanes: group m scores and
// accumulates component m, lane dl of a group owns the dimensions dl, dl + LPC, ...  (D = 39, ncomp = 4: 5 dimensions
// per lane, 39 of 40 lane slots used; the former lanes-over-dimensions layout ran 4 components x 2 half-empty passes
// with a full warp reduction each).  Row strides of the shared tables are = +-LPC mod 32, so the groups hit disjoint
// banks.  The frame's features and its posterior row are fetched one frame ahead.  Each warp owns a private
// [S*ncomp, rowp] block of shared memory (plain read-modify-write, no atomics, deterministic); the CTA adds its warps'
// blocks in a fixed order into its private global slice, laid out [S*M, row] like the final statistics.
__host__ __device__ inline int soft_stride(int n, int lpc) {       // smallest multiple of lpc >= n that is +-lpc mod 32
    int v = (n + lpc - 1) / lpc * lpc;
    while (v % 32 != lpc % 32 && v % 32 != (32 - lpc) % 32) v += lpc;
    return v;
}

template <int NCP>
__global__ void __launch_bounds__(SW * 32)
gmm_stats_soft_kernel(const float *__restrict__ X, int64_t row0, int64_t nrows, int D, const float *__restrict__ gamma,
                      int S, int M, int ncomp, const double *__restrict__ mu, const double *__restrict__ var,
                      const double *__restrict__ lc, float gmin, float *__restrict__ partial) {
    constexpr int LPC = 32 / NCP, KMAX = 64 / LPC;
    extern __shared__ float ssm[];
    const int row = 2 * D + 3, GW = S * ncomp, Dp = soft_stride(D, LPC), rowp = soft_stride(row, LPC);
    float *mean_s = ssm, *hiv_s = mean_s + GW * Dp, *lc_s = hiv_s + GW * Dp, *part = lc_s + ((GW + 3) & ~3);
    const int lane = threadIdx.x & 31, wq = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < GW * Dp; i += blockDim.x) {
        const int gi = i / Dp, d = i - gi * Dp;
        const size_t g = (size_t)(gi / ncomp) * M + gi % ncomp;
        mean_s[i] = d < D ? (float)mu[g * D + d] : 0.f;
        hiv_s[i] = d < D ? (float)(-0.5 / var[g * D + d]) : 0.f;       // zero beyond D: padded dimensions add nothing
    }
    for (int i = threadIdx.x; i < GW; i += blockDim.x) lc_s[i] = (float)lc[(size_t)(i / ncomp) * M + i % ncomp];
    for (int i = threadIdx.x; i < SW * GW * rowp; i += blockDim.x) part[i] = 0.f;
    __syncthreads();
    float *mine = part + (size_t)wq * GW * rowp;
    const int m = lane / LPC, dl = lane % LPC, mm = min(m, ncomp - 1);
    const bool has = m < ncomp;
    const int64_t wg = (int64_t)blockIdx.x * SW + wq, nw = (int64_t)gridDim.x * SW;
    for (int64_t c0 = wg * SCH_SOFT; c0 < nrows; c0 += nw * SCH_SOFT) {
        const int64_t nb = row0 + c0, ne = row0 + min(nrows, c0 + SCH_SOFT);
        float xn[KMAX], gn = 0.f;
#pragma unroll
        for (int k = 0; k < KMAX; ++k) {
            xn[k] = 0.f;
            if (k * LPC < D && k * LPC + dl < D) xn[k] = X[nb * D + k * LPC + dl];
        }
        if (lane < S) gn = gamma[nb * S + lane];
        for (int64_t n = nb; n < ne; ++n) {
            float x[KMAX];
            const float g0 = gn;
#pragma unroll
            for (int k = 0; k < KMAX; ++k) x[k] = xn[k];
            if (n + 1 < ne) {
#pragma unroll
                for (int k = 0; k < KMAX; ++k)
                    if (k * LPC < D && k * LPC + dl < D) xn[k] = X[(n + 1) * D + k * LPC + dl];
                if (lane < S) gn = gamma[(n + 1) * S + lane];
            }
            for (int sb = 0; sb < S; sb += 32) {
                const float gl = sb == 0 ? g0 : (sb + lane < S ? gamma[n * S + sb + lane] : 0.f);
                unsigned todo = __ballot_sync(0xffffffffu, sb + lane < S && gl > gmin);
                while (todo) {                                           // states in ascending order, warp uniform
                    const int bit = __ffs(todo) - 1;
                    todo &= todo - 1;
                    const int s = sb + bit;
                    const float g = __shfl_sync(0xffffffffu, gl, bit);
                    const float *mp = mean_s + (size_t)(s * ncomp + mm) * Dp + dl;
                    const float *hp = hiv_s + (size_t)(s * ncomp + mm) * Dp + dl;
                    float e[KMAX], acc = 0.f;
#pragma unroll
                    for (int k = 0; k < KMAX; ++k) {
                        e[k] = 0.f;
                        if (k * LPC < D) {
                            e[k] = x[k] - mp[k * LPC];                   // 0 - 0 beyond D
                            acc = fmaf(e[k] * e[k], hp[k * LPC], acc);
                        }
                    }
#pragma unroll
                    for (int d = LPC >> 1; d; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
                    const float ll = has ? acc + lc_s[s * ncomp + mm] : -CUDART_INF_F;
                    float mx = ll;
#pragma unroll
                    for (int d = LPC; d < 32; d <<= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, d));
                    const float pe = (mx > -CUDART_INF_F) ? __expf(ll - mx) : 0.f;
                    float den = pe;
#pragma unroll
                    for (int d = LPC; d < 32; d <<= 1) den += __shfl_xor_sync(0xffffffffu, den, d);
                    const float p = pe * (den > 0.f ? g / den : 0.f);
                    if (has) {
                        float *r = mine + (size_t)(s * ncomp + m) * rowp;
                        if (dl == 0) {
                            r[0] += p;
                            if (m == 0) r[2 * D + 1] += g;
                        }
#pragma unroll
                        for (int k = 0; k < KMAX; ++k)
                            if (k * LPC < D && k * LPC + dl < D) {
                                const int d = k * LPC + dl;
                                r[1 + d] = fmaf(p, e[k], r[1 + d]);
                                r[1 + D + d] = fmaf(p * e[k], e[k], r[1 + D + d]);
                            }
                    }
                }
            }
        }
    }
    __syncthreads();
    float *slice = partial + (size_t)blockIdx.x * S * M * row;
    for (int i = threadIdx.x; i < GW * row; i += blockDim.x) {
        const int gi = i / row, j = i - gi * row;
        float t = 0.f;
        for (int q = 0; q < SW; ++q) t += part[((size_t)q * GW + gi) * rowp + j];
        slice[(size_t)((gi / ncomp) * M + gi % ncomp) * row + j] = t;
    }
}

// Register form of the same kernel for models of at most SS states (everything HMM.fit builds): the lane that owns
// component m and dimensions dl, dl + LPC, ... keeps the moments of ALL states in registers (SS x (2 * KD + 2) of them),
// and the loop over the states with gamma > gmin is unrolled with warp-uniform predicates, so the per-frame read-modify-
// write of a private shared-memory block (two thirds of the old kernel's instructions, and the 11 KB per warp that capped
// it at 16 latency-bound warps per SM) is gone.  The registers are folded into ONE shared block per CTA, warp after warp
// in a fixed order, every SOFT_FLUSH chunks (bounds the FP32 accumulation length) and at the end; the block goes to the
// CTA's private global slice as before.  Same arithmetic per (frame, state, component) as gmm_stats_soft_kernel.
constexpr int SOFT_FLUSH = 128;    // chunks of SCH_SOFT frames a warp accumulates in registers between two folds

template <int NCP, int SS, int KD, int OCC>
__global__ void __launch_bounds__(SW * 32, OCC)
gmm_stats_soft_reg_kernel(const float *__restrict__ X, int64_t row0, int64_t nrows, int D, const float *__restrict__ gamma,
                          int S, int M, int ncomp, const double *__restrict__ mu, const double *__restrict__ var,
                          const double *__restrict__ lc, float gmin, float *__restrict__ partial) {
    constexpr int LPC = 32 / NCP;
    extern __shared__ float ssm[];
    const int row = 2 * D + 3, GW = S * ncomp, Dp = soft_stride(D, LPC);
    float *mean_s = ssm, *hiv_s = mean_s + GW * Dp, *lc_s = hiv_s + GW * Dp, *blk = lc_s + ((GW + 3) & ~3);   // blk[GW][row]
    const int lane = threadIdx.x & 31, wq = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < GW * Dp; i += blockDim.x) {
        const int gi = i / Dp, d = i - gi * Dp;
        const size_t g = (size_t)(gi / ncomp) * M + gi % ncomp;
        mean_s[i] = d < D ? (float)mu[g * D + d] : 0.f;
        hiv_s[i] = d < D ? (float)(-0.5 / var[g * D + d]) : 0.f;
    }
    for (int i = threadIdx.x; i < GW; i += blockDim.x) lc_s[i] = (float)lc[(size_t)(i / ncomp) * M + i % ncomp];
    for (int i = threadIdx.x; i < GW * row; i += blockDim.x) blk[i] = 0.f;
    __syncthreads();
    const int m = lane / LPC, dl = lane % LPC, mm = min(m, ncomp - 1);
    const bool has = m < ncomp;
    float a0[SS], ag[SS], a1[SS][KD], a2[SS][KD];
#pragma unroll
    for (int s = 0; s < SS; ++s) {
        a0[s] = ag[s] = 0.f;
#pragma unroll
        for (int k = 0; k < KD; ++k) a1[s][k] = a2[s][k] = 0.f;
    }
    const int64_t wg = (int64_t)blockIdx.x * SW + wq, nw = (int64_t)gridDim.x * SW;
    const int64_t nchunks = (nrows + SCH_SOFT - 1) / SCH_SOFT;
    const int64_t iters = (nchunks + nw - 1) / nw;                          // the same for every warp of the grid
    for (int64_t it = 0; it < iters; ++it) {
        const int64_t c0 = (wg + it * nw) * SCH_SOFT;
        if (c0 < nrows) {
            const int64_t nb = row0 + c0, ne = row0 + min(nrows, c0 + SCH_SOFT);
            float xn[KD], xm[KD], gn = 0.f, gm_ = 0.f;                       // the next pair of frames, fetched one step ahead
#pragma unroll
            for (int k = 0; k < KD; ++k) {
                xn[k] = (k * LPC + dl < D) ? X[nb * D + k * LPC + dl] : 0.f;
                xm[k] = (nb + 1 < ne && k * LPC + dl < D) ? X[(nb + 1) * D + k * LPC + dl] : 0.f;
            }
            if (lane < S) {
                gn = gamma[nb * S + lane];
                if (nb + 1 < ne) gm_ = gamma[(nb + 1) * S + lane];
            }
            // TWO frames per step: neighbouring frames put their mass on (nearly) the same states, so the unrolled state
            // loop runs over the union of their active states once, with two independent chains per lane (a state that
            // is inactive for one of the frames contributes exact zeros: its gamma is passed as 0)
            for (int64_t n = nb; n < ne; n += 2) {
                float xa[KD], xb[KD];
                const float gla = gn, glb = gm_;
#pragma unroll
                for (int k = 0; k < KD; ++k) {
                    xa[k] = xn[k];
                    xb[k] = xm[k];
                }
                gm_ = 0.f;
#pragma unroll
                for (int k = 0; k < KD; ++k) {
                    xm[k] = 0.f;
                    if (k * LPC + dl < D) {
                        if (n + 2 < ne) xn[k] = X[(n + 2) * D + k * LPC + dl];
                        if (n + 3 < ne) xm[k] = X[(n + 3) * D + k * LPC + dl];
                    }
                }
                if (lane < S) {
                    if (n + 2 < ne) gn = gamma[(n + 2) * S + lane];
                    if (n + 3 < ne) gm_ = gamma[(n + 3) * S + lane];
                }
                const float ga_ = gla > gmin ? gla : 0.f, gb_ = glb > gmin ? glb : 0.f;
                const unsigned todo = __ballot_sync(0xffffffffu, lane < S && (ga_ > 0.f || gb_ > 0.f));
#pragma unroll
                for (int s = 0; s < SS; ++s) {
                    if (!((todo >> s) & 1u)) continue;                      // warp uniform
                    const float g_a = __shfl_sync(0xffffffffu, ga_, s), g_b = __shfl_sync(0xffffffffu, gb_, s);
                    const float *mp = mean_s + (size_t)(s * ncomp + mm) * Dp + dl;
                    const float *hp = hiv_s + (size_t)(s * ncomp + mm) * Dp + dl;
                    float ea[KD], eb[KD], acca = 0.f, accb = 0.f;
#pragma unroll
                    for (int k = 0; k < KD; ++k) {
                        const float mk = mp[k * LPC], hk = hp[k * LPC];
                        ea[k] = xa[k] - mk;                                 // 0 - 0 beyond D
                        eb[k] = xb[k] - mk;
                        acca = fmaf(ea[k] * ea[k], hk, acca);
                        accb = fmaf(eb[k] * eb[k], hk, accb);
                    }
#pragma unroll
                    for (int d = LPC >> 1; d; d >>= 1) {
                        acca += __shfl_xor_sync(0xffffffffu, acca, d);
                        accb += __shfl_xor_sync(0xffffffffu, accb, d);
                    }
                    const float lcv = lc_s[s * ncomp + mm];
                    const float lla = has ? acca + lcv : -CUDART_INF_F, llb = has ? accb + lcv : -CUDART_INF_F;
                    float mxa = lla, mxb = llb;
#pragma unroll
                    for (int d = LPC; d < 32; d <<= 1) {
                        mxa = fmaxf(mxa, __shfl_xor_sync(0xffffffffu, mxa, d));
                        mxb = fmaxf(mxb, __shfl_xor_sync(0xffffffffu, mxb, d));
                    }
                    const float pea = (mxa > -CUDART_INF_F) ? __expf(lla - mxa) : 0.f;
                    const float peb = (mxb > -CUDART_INF_F) ? __expf(llb - mxb) : 0.f;
                    float dena = pea, denb = peb;
#pragma unroll
                    for (int d = LPC; d < 32; d <<= 1) {
                        dena += __shfl_xor_sync(0xffffffffu, dena, d);
                        denb += __shfl_xor_sync(0xffffffffu, denb, d);
                    }
                    const float pa = pea * (dena > 0.f ? g_a / dena : 0.f);
                    const float pb = peb * (denb > 0.f ? g_b / denb : 0.f);
                    a0[s] += pa;
                    a0[s] += pb;
                    ag[s] += g_a;
                    ag[s] += g_b;
#pragma unroll
                    for (int k = 0; k < KD; ++k) {
                        a1[s][k] = fmaf(pa, ea[k], a1[s][k]);
                        a2[s][k] = fmaf(pa * ea[k], ea[k], a2[s][k]);
                        a1[s][k] = fmaf(pb, eb[k], a1[s][k]);
                        a2[s][k] = fmaf(pb * eb[k], eb[k], a2[s][k]);
                    }
                }
            }
        }
        if ((it % SOFT_FLUSH) == SOFT_FLUSH - 1 || it + 1 == iters) {
            // fold the registers into the CTA's block, warp after warp (fixed order), and start again from zero
            for (int q = 0; q < SW; ++q) {
                if (wq == q && has) {
#pragma unroll
                    for (int s = 0; s < SS; ++s) {
                        if (s < S) {
                            float *r = blk + (size_t)(s * ncomp + m) * row;
                            if (dl == 0) {
                                r[0] += a0[s];
                                if (m == 0) r[2 * D + 1] += ag[s];
                            }
#pragma unroll
                            for (int k = 0; k < KD; ++k)
                                if (k * LPC + dl < D) {
                                    r[1 + k * LPC + dl] += a1[s][k];
                                    r[1 + D + k * LPC + dl] += a2[s][k];
                                }
                        }
                    }
                }
                __syncthreads();
            }
#pragma unroll
            for (int s = 0; s < SS; ++s) {
                a0[s] = ag[s] = 0.f;
#pragma unroll
                for (int k = 0; k < KD; ++k) a1[s][k] = a2[s][k] = 0.f;
            }
        }
    }
    float *slice = partial + (size_t)blockIdx.x * S * M * row;
    for (int i = threadIdx.x; i < GW * row; i += blockDim.x) {
        const int gi = i / row, j = i - gi * row;
        slice[(size_t)((gi / ncomp) * M + gi % ncomp) * row + j] = blk[i];
    }
}

template <int NCP>
static cudaError_t launch_stats_soft_reg(int grid, cudaStream_t st, const float *X, int64_t row0, int64_t nrows, int D,
                                         const float *gamma, int S, int M, int ncomp, const double *mu, const double *var,
                                         const double *lc, float gmin, float *partial) {
    constexpr int LPC = 32 / NCP, KD = (39 + LPC - 1) / LPC;                 // D <= 39 + what the padding of KD*LPC allows
    const int row = 2 * D + 3, GW = S * ncomp;
    const size_t bytes = sizeof(float) * ((size_t)2 * GW * soft_stride(D, LPC) + ((GW + 3) & ~3) + (size_t)GW * row);
    static const int occ = getenv("SRB_SOFT_OCC") ? atoi(getenv("SRB_SOFT_OCC")) : 1;   // CTAs per SM the registers allow
    if (occ >= 2) {
        cudaError_t rc = cudaFuncSetAttribute(gmm_stats_soft_reg_kernel<NCP, 8, KD, 2>,
                                              cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
        if (rc != cudaSuccess) return rc;
        gmm_stats_soft_reg_kernel<NCP, 8, KD, 2><<<grid, SW * 32, bytes, st>>>(X, row0, nrows, D, gamma, S, M, ncomp, mu, var,
                                                                               lc, gmin, partial);
        return cudaGetLastError();
    }
    cudaError_t rc = cudaFuncSetAttribute(gmm_stats_soft_reg_kernel<NCP, 8, KD, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)bytes);
    if (rc != cudaSuccess) return rc;
    gmm_stats_soft_reg_kernel<NCP, 8, KD, 1><<<grid, SW * 32, bytes, st>>>(X, row0, nrows, D, gamma, S, M, ncomp, mu, var, lc,
                                                                           gmin, partial);
    return cudaGetLastError();
}

template <int NCP>
static cudaError_t launch_stats_soft(int grid, size_t bytes, cudaStream_t st, const float *X, int64_t row0, int64_t nrows,
                                     int D, const float *gamma, int S, int M, int ncomp, const double *mu,
                                     const double *var, const double *lc, float gmin, float *partial) {
    cudaError_t rc = cudaFuncSetAttribute(gmm_stats_soft_kernel<NCP>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)bytes);
    if (rc != cudaSuccess) return rc;
    gmm_stats_soft_kernel<NCP><<<grid, SW * 32, bytes, st>>>(X, row0, nrows, D, gamma, S, M, ncomp, mu, var, lc, gmin,
                                                             partial);
    return cudaGetLastError();
}

// out[i] (+)= sum_p partial[p][i]  in double, fixed order
__global__ void reduce_partials_kernel(const float *__restrict__ partial, int64_t nparts, int64_t len,
                                       double *__restrict__ out, int accumulate) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= len) return;
    double s = accumulate ? out[i] : 0.0;
    for (int64_t p = 0; p < nparts; ++p) s += (double)partial[p * len + i];
    out[i] = s;
}


// ---- local distances of every frame to S templates (dtw's dist_fun, kmeans' mahalanobis):
// kind 0: ||x - y_s||_2 (kmeans.py:111 default np.linalg.norm)
// kind 1: mahalanobis(x, y_s, var_s) = 0.5*log((2 pi)^D prod var) + 0.5*sum (x-y)^2/var  (hmm_state.py:48-58)
// One warp per frame group, lanes over dims, templates in shared memory (double -> float once per CTA).
__global__ void __launch_bounds__(SW * 32)
local_dist_kernel(const float *__restrict__ X, int64_t N, int D, const double *__restrict__ Y,
                  const double *__restrict__ var, int S, int kind, int var_stride, float *__restrict__ out,
                  int64_t ldo, int32_t *__restrict__ argmin_out) {
    extern __shared__ float lsm[];
    float *ys = lsm, *iv = ys + (size_t)S * D, *cs = iv + (size_t)S * D;   // S*D, S*D, S
    for (int i = threadIdx.x; i < S * D; i += blockDim.x) {
        ys[i] = (float)Y[i];
        iv[i] = (kind == 1) ? (float)(0.5 / var[(size_t)(i / D) * var_stride + (i % D)]) : 1.f;
    }
    for (int s = threadIdx.x; s < S; s += blockDim.x) {
        double c = 0.0;
        if (kind == 1) {
            for (int d = 0; d < D; ++d) c += log(var[(size_t)s * var_stride + d]);
            c = 0.5 * (c + (double)D * 1.8378770664093454836);
        }
        cs[s] = (float)c;
    }
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int64_t wg = (int64_t)blockIdx.x * SW + (threadIdx.x >> 5), nw = (int64_t)gridDim.x * SW;
    const int d0 = lane, d1 = lane + 32;
    const bool h0 = d0 < D, h1 = d1 < D;
    for (int64_t n = wg; n < N; n += nw) {
        const float xa = h0 ? X[n * D + d0] : 0.f, xb = h1 ? X[n * D + d1] : 0.f;
        float best = CUDART_INF_F;
        int bi = 0;
        for (int s = 0; s < S; ++s) {
            float p = 0.f;
            if (h0) {
                const float e = xa - ys[s * D + d0];
                p = e * e * iv[s * D + d0];
            }
            if (h1) {
                const float e = xb - ys[s * D + d1];
                p = fmaf(e * e, iv[s * D + d1], p);
            }
            p = wsum(p);
            const float dist = (kind == 1) ? p + cs[s] : sqrtf(p);
            if (out && lane == 0) out[n * ldo + s] = dist;
            if (dist < best) {             // np.argmin: first minimum (kmeans.py:185)
                best = dist;
                bi = s;
            }
        }
        if (argmin_out && lane == 0) argmin_out[n] = bi;
    }
}

// ---- K-means assignment for MANY pools at once (kmeans.py:180-185 for every state of every word in one launch): frame n
// belongs to pool group[n] (< 0: skipped), whose k centroids are Y[group*k .. +k) and whose distance uses ONE variance
// row var[group] for all centroids (kmeans.py:183 passes cov[0]).  argmin_out[n] = first minimum of
// mahalanobis(centroid_c, x_n, var) over c (hmm_state.py:48-58); the constant term is the same for all c and dropped.
__global__ void __launch_bounds__(SW * 32)
kmeans_assign_kernel(const float *__restrict__ X, int64_t N, int D, const int32_t *__restrict__ group,
                     const double *__restrict__ Y, const double *__restrict__ var, int G, int k,
                     int32_t *__restrict__ argmin_out) {
    extern __shared__ float lsm[];
    float *ys = lsm, *iv = ys + (size_t)G * k * D;                         // G*k*D centroids, G*D half inverse variances
    for (int i = threadIdx.x; i < G * k * D; i += blockDim.x) ys[i] = (float)Y[i];
    for (int i = threadIdx.x; i < G * D; i += blockDim.x) iv[i] = (float)(0.5 / var[i]);
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int64_t wg = (int64_t)blockIdx.x * SW + (threadIdx.x >> 5), nw = (int64_t)gridDim.x * SW;
    const int d0 = lane, d1 = lane + 32;
    const bool h0 = d0 < D, h1 = d1 < D;
    for (int64_t n = wg; n < N; n += nw) {
        const int g = group[n];
        if (g < 0 || g >= G) {
            if (lane == 0) argmin_out[n] = -1;
            continue;
        }
        const float xa = h0 ? X[n * D + d0] : 0.f, xb = h1 ? X[n * D + d1] : 0.f;
        float best = CUDART_INF_F;
        int bi = 0;
        for (int c = 0; c < k; ++c) {
            const float *y = ys + (size_t)(g * k + c) * D;
            float p = 0.f;
            if (h0) {
                const float e = xa - y[d0];
                p = e * e * iv[g * D + d0];
            }
            if (h1) {
                const float e = xb - y[d1];
                p = fmaf(e * e, iv[g * D + d1], p);
            }
            p = wsum(p);
            if (p < best) {                // np.argmin: first minimum (kmeans.py:185); NaN centroids never win
                best = p;
                bi = c;
            }
        }
        if (lane == 0) argmin_out[n] = bi;
    }
}

// ---- M-step of GMM.em (hmm_state.py:138-148) from the reduced moments, for the first ncomp components of
// every state:  mu' = mu + S1/S0;  var' = S2/S0 - (S1/S0)^2  (== sum p (x-mu')^2 / sum p);  w' = S0 / count.
// Components with S0 == 0 are left untouched (the reference would divide by zero there).
__global__ void mstep_kernel(const double *__restrict__ stats, int R, int M, int ncomp, int D, double *__restrict__ mu,
                             double *__restrict__ var, double *__restrict__ w, double var_floor) {
    const int row = 2 * D + 3;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;   // (state, comp, d)
    if (idx >= R * ncomp * D) return;
    const int d = idx % D, m = (idx / D) % ncomp, s = idx / (D * ncomp);
    const size_t g = (size_t)s * M + m;
    const double *r = stats + g * row;
    const double S0 = r[0];
    if (!(S0 > 0.0)) return;
    const double m1 = r[1 + d] / S0;
    double v = r[1 + D + d] / S0 - m1 * m1;
    if (v < var_floor) v = var_floor;
    mu[g * D + d] += m1;
    var[g * D + d] = v;
    if (d == 0) {
        const double cnt = stats[((size_t)s * M) * row + 2 * D + 1];
        if (cnt > 0.0) w[g] = S0 / cnt;
    }
}

// natural-log constants  lc[g] = log w - 0.5*(D log 2pi + sum log var)  (hmm_state.py:39-40 in the log domain)
__global__ void gauss_consts_kernel(const double *__restrict__ var, const double *__restrict__ w, int G, int D,
                                    double *__restrict__ lc) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= G) return;
    double ld = 0.0;
    for (int d = 0; d < D; ++d) ld += log(var[(size_t)g * D + d]);
    lc[g] = (w[g] > 0.0 ? log(w[g]) : -CUDART_INF) - 0.5 * ((double)D * 1.8378770664093454836 + ld);
}

// Baum-Welch transition update on the device: a[o -> r] = xi[o -> r] / sum_r' xi[o -> r'], cost = -log a, written in
// place into every copy of the arc's cost the kernels read (predecessor CSR, successor CSR, band array).  One thread per
// origin state walks its successor list in a fixed order: deterministic, no host round trip.  Arcs never taken (xi == 0)
// and origins without mass keep their cost, like the host form this replaces.
__global__ void update_transitions_kernel(const double *__restrict__ xi, const int32_t *__restrict__ succ_ptr,
                                          const int32_t *__restrict__ succ_edge, const int32_t *__restrict__ band_slot,
                                          int n_states, float *__restrict__ pred_cost, float *__restrict__ succ_cost,
                                          float *__restrict__ band) {
    const int o = blockIdx.x * blockDim.x + threadIdx.x;
    if (o >= n_states) return;
    const int j0 = succ_ptr[o], j1 = succ_ptr[o + 1];
    double out = 0.0;
    for (int j = j0; j < j1; ++j) out += xi[succ_edge[j]];
    if (!(out > 0.0)) return;
    for (int j = j0; j < j1; ++j) {
        const int e = succ_edge[j];
        const double x = xi[e];
        if (!(x > 0.0)) continue;
        const float c = (float)(-log(x / out));
        pred_cost[e] = c;
        if (succ_cost) succ_cost[j] = c;
        if (band && band_slot && band_slot[e] >= 0) band[band_slot[e]] = c;
    }
}

}  // namespace srb

using namespace srb;

extern "C" {

int srb_stats_row(int D) { return 2 * D + 3; }

int srb_stats_parts(void) { return sm_count() * 2 * SW; }   // private slices (one per accumulating warp)

static int stats_grid() { return sm_count() * 2; }

int srb_group_stats(const float *X, int64_t N, int D, const int32_t *label, int K, const double *shift,
                    float *partial, double *out, int accumulate, void *stream) {
    SRB_REQUIRE((X || N == 0) && label && shift && partial && out, SRB_EINVAL, "srb_group_stats: null pointer");
    SRB_REQUIRE(N >= 0 && D > 0 && D <= 64 && K > 0, SRB_EINVAL, "srb_group_stats: bad N/D/K");
    const int grid = stats_grid();
    SRB_REQUIRE(grid > 0, SRB_ENODEVICE, "no CUDA device");
    const int64_t len = (int64_t)K * (2 * D + 3), parts = (int64_t)grid * SW;
    cudaStream_t st = (cudaStream_t)stream;
    SRB_CUDA(cudaMemsetAsync(partial, 0, sizeof(float) * len * parts, st));
    if (N > 0) group_stats_kernel<<<grid, SW * 32, 0, st>>>(X, N, D, label, K, shift, partial);
    reduce_partials_kernel<<<(int)((len + 255) / 256), 256, 0, st>>>(partial, parts, len, out, accumulate);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

// ---- srb_group_stats_static: srb_group_stats for callers that may not allocate (CUDA graph capture, serving loops).
// The private slices live in this file-scope device array instead of caller scratch.  Host code only ever uses the
// DEVICE address of the array, fetched once per device with cudaGetSymbolAddress (the bare symbol is a host shadow).
// Every call on a device shares the array: two calls that overlap on the GPU (different streams) race.
static __device__ __align__(256) float g_static_partial[SRB_STATIC_SCRATCH_FLOATS];

int64_t srb_static_scratch_floats(void) { return (int64_t)SRB_STATIC_SCRATCH_FLOATS; }

static int static_partial_ptr(float **out) {
    constexpr int MAXDEV = 64;
    static float *ptrs[MAXDEV] = {};
    static std::mutex mu;
    int dev = -1;
    SRB_CUDA(cudaGetDevice(&dev));
    SRB_REQUIRE(dev >= 0 && dev < MAXDEV, SRB_EUNSUPPORTED, "srb_group_stats_static: device ordinal %d", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[dev]) {
        float *p = nullptr;
        SRB_CUDA(cudaGetSymbolAddress((void **)&p, g_static_partial));       // with device `dev` current
        SRB_REQUIRE(p != nullptr, SRB_ECUDA, "srb_group_stats_static: no device address for the scratch array");
        ptrs[dev] = p;
    }
    *out = ptrs[dev];
    return SRB_OK;
}

int srb_group_stats_static(const float *X, int64_t N, int D, const int32_t *label, int K, const double *shift,
                           double *out, int accumulate, void *stream) {
    SRB_REQUIRE((X || N == 0) && label && shift && out, SRB_EINVAL, "srb_group_stats_static: null pointer");
    SRB_REQUIRE(N >= 0 && D > 0 && D <= 64 && K > 0, SRB_EINVAL, "srb_group_stats_static: bad N/D/K");
    const int grid = stats_grid();
    SRB_REQUIRE(grid > 0, SRB_ENODEVICE, "no CUDA device");
    const int64_t len = (int64_t)K * (2 * D + 3), parts = (int64_t)grid * SW;
    SRB_REQUIRE(len * parts <= (int64_t)SRB_STATIC_SCRATCH_FLOATS, SRB_EUNSUPPORTED,
                "srb_group_stats_static: K=%d groups x %d floats x %lld slices = %lld floats exceed the static scratch of "
                "%lld floats (at most %lld groups at this D); use srb_group_stats with caller scratch",
                K, 2 * D + 3, (long long)parts, (long long)(len * parts), (long long)SRB_STATIC_SCRATCH_FLOATS,
                (long long)(SRB_STATIC_SCRATCH_FLOATS / (parts * (2 * D + 3))));
    float *partial = nullptr;
    const int rc = static_partial_ptr(&partial);
    if (rc != SRB_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    SRB_CUDA(cudaMemsetAsync(partial, 0, sizeof(float) * len * parts, st));    // the array keeps the last call's sums
    if (N > 0) group_stats_kernel<<<grid, SW * 32, 0, st>>>(X, N, D, label, K, shift, partial);
    reduce_partials_kernel<<<(int)((len + 255) / 256), 256, 0, st>>>(partial, parts, len, out, accumulate);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int srb_gmm_stats(const float *X, int64_t N, int D, const int32_t *label, int R, int M, int ncomp, const double *mu,
                  const double *var, const double *lc, float *partial, double *out, int accumulate, void *stream) {
    SRB_REQUIRE((X || N == 0) && label && mu && var && lc && partial && out, SRB_EINVAL, "srb_gmm_stats: null pointer");
    SRB_REQUIRE(N >= 0 && D > 0 && D <= 64 && R > 0 && M > 0, SRB_EINVAL, "srb_gmm_stats: bad N/D/R/M");
    SRB_REQUIRE(ncomp >= 1 && ncomp <= M && ncomp <= MAXC, SRB_EUNSUPPORTED,
                "srb_gmm_stats: ncomp=%d outside 1..min(M,%d)", ncomp, MAXC);
    const int grid = stats_grid();
    SRB_REQUIRE(grid > 0, SRB_ENODEVICE, "no CUDA device");
    const int G = R * M;
    const int64_t len = (int64_t)G * (2 * D + 3), parts = (int64_t)grid * SW;
    cudaStream_t st = (cudaStream_t)stream;
    SRB_CUDA(cudaMemsetAsync(partial, 0, sizeof(float) * len * parts, st));
    if (N > 0) gmm_stats_kernel<<<grid, SW * 32, 0, st>>>(X, N, D, label, M, ncomp, mu, var, lc, partial, G);
    reduce_partials_kernel<<<(int)((len + 255) / 256), 256, 0, st>>>(partial, parts, len, out, accumulate);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int srb_gmm_stats_soft(const float *X, int64_t N, int64_t row0, int64_t nrows, int D, const float *gamma, int S, int M,
                       int ncomp, const double *mu, const double *var, const double *lc, double gmin, float *partial,
                       double *out, int accumulate, void *stream) {
    SRB_REQUIRE(X && gamma && mu && var && lc && partial && out, SRB_EINVAL, "srb_gmm_stats_soft: null pointer");
    SRB_REQUIRE(N >= 0 && row0 >= 0 && nrows >= 0 && row0 + nrows <= N && D > 0 && D <= 64 && S > 0 && M > 0, SRB_EINVAL,
                "srb_gmm_stats_soft: bad sizes");
    SRB_REQUIRE(ncomp >= 1 && ncomp <= M && ncomp <= MAXC, SRB_EUNSUPPORTED,
                "srb_gmm_stats_soft: ncomp=%d outside 1..min(M,%d)", ncomp, MAXC);
    const int grid = stats_grid();
    SRB_REQUIRE(grid > 0, SRB_ENODEVICE, "no CUDA device");
    const int row = 2 * D + 3, GW = S * ncomp;
    const int ncp = ncomp <= 1 ? 1 : ncomp <= 2 ? 2 : ncomp <= 4 ? 4 : 8, lpc = 32 / ncp;
    const size_t bytes = sizeof(float) * ((size_t)2 * GW * soft_stride(D, lpc) + ((GW + 3) & ~3) +
                                          (size_t)SW * GW * soft_stride(row, lpc));
    SRB_REQUIRE(bytes <= (size_t)smem_optin(), SRB_EUNSUPPORTED,
                "srb_gmm_stats_soft: %d states x %d components need %zu B shared memory", S, ncomp, bytes);
    const int64_t len = (int64_t)S * M * row;
    cudaStream_t st = (cudaStream_t)stream;
    SRB_CUDA(cudaMemsetAsync(partial, 0, sizeof(float) * len * grid, st));
    const float gm = (float)gmin;
    if (S <= 8 && D <= (39 + lpc - 1) / lpc * lpc && ncp >= 2 && !getenv("SRB_SOFT_SMEM")) {
        // register form: the moments of all states stay in the lanes' registers (see gmm_stats_soft_reg_kernel)
        if (ncp == 2)
            SRB_CUDA(launch_stats_soft_reg<2>(grid, st, X, row0, nrows, D, gamma, S, M, ncomp, mu, var, lc, gm, partial));
        else if (ncp == 4)
            SRB_CUDA(launch_stats_soft_reg<4>(grid, st, X, row0, nrows, D, gamma, S, M, ncomp, mu, var, lc, gm, partial));
        else
            SRB_CUDA(launch_stats_soft_reg<8>(grid, st, X, row0, nrows, D, gamma, S, M, ncomp, mu, var, lc, gm, partial));
    } else if (ncp == 1)
        SRB_CUDA(launch_stats_soft<1>(grid, bytes, st, X, row0, nrows, D, gamma, S, M, ncomp, mu, var, lc, gm, partial));
    else if (ncp == 2)
        SRB_CUDA(launch_stats_soft<2>(grid, bytes, st, X, row0, nrows, D, gamma, S, M, ncomp, mu, var, lc, gm, partial));
    else if (ncp == 4)
        SRB_CUDA(launch_stats_soft<4>(grid, bytes, st, X, row0, nrows, D, gamma, S, M, ncomp, mu, var, lc, gm, partial));
    else
        SRB_CUDA(launch_stats_soft<8>(grid, bytes, st, X, row0, nrows, D, gamma, S, M, ncomp, mu, var, lc, gm, partial));
    reduce_partials_kernel<<<(int)((len + 255) / 256), 256, 0, st>>>(partial, grid, len, out, accumulate);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int srb_local_dist(const float *X, int64_t N, int D, const double *Y, const double *var, int S, int kind,
                   int var_stride, float *out, int64_t ldo, int32_t *argmin_out, void *stream) {
    SRB_REQUIRE((X || N == 0) && Y && (out || argmin_out || N == 0), SRB_EINVAL, "srb_local_dist: null pointer");
    SRB_REQUIRE(kind == 0 || (kind == 1 && var), SRB_EINVAL, "srb_local_dist: kind 1 needs var");
    SRB_REQUIRE(N >= 0 && D > 0 && D <= 64 && S > 0 && (!out || ldo >= S), SRB_EINVAL, "srb_local_dist: bad sizes");
    if (N == 0) return SRB_OK;
    const size_t bytes = sizeof(float) * ((size_t)2 * S * D + S);
    SRB_REQUIRE(bytes <= (size_t)smem_optin(), SRB_EUNSUPPORTED, "srb_local_dist: %d templates do not fit shared memory", S);
    SRB_CUDA(cudaFuncSetAttribute(local_dist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    int64_t grid = (N + SW - 1) / SW;
    const int64_t cap = (int64_t)sm_count() * 8;
    if (grid > cap) grid = cap;
    local_dist_kernel<<<(int)grid, SW * 32, bytes, (cudaStream_t)stream>>>(X, N, D, Y, var, S, kind, var_stride, out,
                                                                          ldo, argmin_out);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int srb_kmeans_assign(const float *X, int64_t N, int D, const int32_t *group, const double *Y, const double *var, int G,
                      int k, int32_t *argmin_out, void *stream) {
    SRB_REQUIRE(N >= 0 && D > 0 && D <= 64 && G > 0 && k > 0, SRB_EINVAL, "srb_kmeans_assign: bad sizes");
    if (N == 0) return SRB_OK;
    SRB_REQUIRE(X && group && Y && var && argmin_out, SRB_EINVAL, "srb_kmeans_assign: null pointer");
    const size_t bytes = sizeof(float) * ((size_t)G * k * D + (size_t)G * D);
    SRB_REQUIRE(bytes <= (size_t)smem_optin(), SRB_EUNSUPPORTED,
                "srb_kmeans_assign: %d pools x %d centroids do not fit shared memory", G, k);
    SRB_CUDA(cudaFuncSetAttribute(kmeans_assign_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    int64_t grid = (N + SW - 1) / SW;
    const int64_t cap = (int64_t)sm_count() * 8;
    if (grid > cap) grid = cap;
    kmeans_assign_kernel<<<(int)grid, SW * 32, bytes, (cudaStream_t)stream>>>(X, N, D, group, Y, var, G, k, argmin_out);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int srb_mstep(const double *stats, int R, int M, int ncomp, int D, double *mu, double *var, double *w,
              double var_floor, void *stream) {
    SRB_REQUIRE(stats && mu && var && w, SRB_EINVAL, "srb_mstep: null pointer");
    SRB_REQUIRE(R > 0 && M > 0 && ncomp >= 1 && ncomp <= M && D > 0, SRB_EINVAL, "srb_mstep: bad sizes");
    const int n = R * ncomp * D;
    mstep_kernel<<<(n + 127) / 128, 128, 0, (cudaStream_t)stream>>>(stats, R, M, ncomp, D, mu, var, w, var_floor);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int srb_update_transitions(const double *xi, const int32_t *succ_ptr, const int32_t *succ_edge, const int32_t *band_slot,
                           int32_t n_states, float *pred_cost, float *succ_cost, float *band, void *stream) {
    SRB_REQUIRE(n_states >= 0, SRB_EINVAL, "srb_update_transitions: bad sizes");
    if (n_states == 0) return SRB_OK;
    SRB_REQUIRE(xi && succ_ptr && succ_edge && pred_cost, SRB_EINVAL, "srb_update_transitions: null pointer");
    update_transitions_kernel<<<(n_states + 127) / 128, 128, 0, (cudaStream_t)stream>>>(xi, succ_ptr, succ_edge, band_slot,
                                                                                          n_states, pred_cost, succ_cost, band);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int srb_gauss_consts(const double *var, const double *w, int G, int D, double *lc, void *stream) {
    SRB_REQUIRE(var && w && lc && G > 0 && D > 0, SRB_EINVAL, "srb_gauss_consts: bad argument");
    gauss_consts_kernel<<<(G + 127) / 128, 128, 0, (cudaStream_t)stream>>>(var, w, G, D, lc);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // extern "C"
