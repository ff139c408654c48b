// pool_stats.cu -- the statistics of one K-means or one EM iteration for MANY pools (HMM states) in one pass over X.
//
// The reference fits one state at a time (hmm.py:97-124): kmeans (kmeans.py:166-193) assigns every frame of the state's
// pool to the nearest of k centroids and re-estimates them; GMM.em (hmm_state.py:122-159) weighs every frame by its
// posterior over the state's first k components.  Both are "weights over the k components of the frame's own state,
// then weighted moments": here one kernel does either for all W*S pools at once, and the K-means assignment is fused
// with the moments (X is read once per iteration instead of twice).
//
// A warp owns chunks of 32 consecutive frames.  The chunk is staged in shared memory by coalesced loads (the next chunk
// is already in flight in registers).  Phase A: lane = frame, scores its k components against the parameter tables in
// shared memory (frames of one state broadcast).  Phase B: lanes = dimensions, a loop over the 32 frames adds the
// weighted first and second moments about the component's shift to register accumulators; when the state changes (the
// frames of an utterance come in runs of one state) the accumulators are added to the warp's PRIVATE slice of `partial`
// with fire-and-forget reductions -- one thread per address, program order, so the sums do not depend on scheduling.
// The slices are then summed in a fixed order in double.  HBM traffic: 4*D bytes of X + 4 bytes of label per frame.
#include <cstdint>
#include <cstdio>

#include "common.cuh"

namespace srb {
namespace {

constexpr int PW = 16;                 // warps per CTA, one CTA per SM
constexpr int MAX_SMEM = 227 * 1024;

template <int KC, int MODE, int DMAX>   // MODE 0: EM posteriors (hmm_state.py:122-136)  1: K-means one-hot (kmeans.py:180-187)
__global__ void __launch_bounds__(PW * 32, 1)
pool_stats_kernel(const float *__restrict__ X, int64_t N, int D, const int32_t *__restrict__ state, int G, int M,
                  const double *__restrict__ center, const double *__restrict__ var, const double *__restrict__ lc,
                  const double *__restrict__ shift, float *__restrict__ partial, int32_t *__restrict__ argmin_out) {
    extern __shared__ float sm[];
    const int Dp = D | 1;                                  // odd row stride: lane = frame reads are conflict-free
    const int GKD = G * KC * D;
    float *cen = sm;                                       // [G, KC, D] centres the frames are scored against
    float *vv = cen + GKD;                                 // EM: [G, KC, D] -0.5/var     K-means: [G, D] 0.5/var0
    float *aux = vv + (MODE == 0 ? GKD : G * D);           // EM: [G, KC] log w - 0.5 log|2 pi var|   K-means: [G, KC, D] shift
    float *xs_all = aux + (MODE == 0 ? G * KC : GKD);
    for (int i = threadIdx.x; i < GKD; i += blockDim.x) {
        const int g = i / (KC * D), r = i - g * (KC * D), m = r / D, d = r - m * D;
        const size_t src = ((size_t)g * M + m) * D + d;
        cen[i] = (float)center[src];
        if (MODE == 0) vv[i] = (float)(-0.5 / var[src]);
        else aux[i] = (float)shift[src];
    }
    if (MODE == 0) {
        for (int i = threadIdx.x; i < G * KC; i += blockDim.x) aux[i] = (float)lc[(size_t)(i / KC) * M + (i % KC)];
    } else {
        for (int i = threadIdx.x; i < G * D; i += blockDim.x) vv[i] = (float)(0.5 / var[i]);
    }
    __syncthreads();

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float *xs = xs_all + (size_t)warp * 32 * Dp;
    const int64_t wg = (int64_t)blockIdx.x * PW + warp, nw = (int64_t)gridDim.x * PW;
    const int row = 2 * D + 3;
    float *mine = partial + (size_t)wg * G * M * row;
    const int64_t nchunks = (N + 31) >> 5, total = N * D;
    const int d0 = lane, d1 = lane + 32;
    const bool h0 = d0 < D, h1 = d1 < D;

    float pre[DMAX];
    auto fetch = [&](int64_t c) {
        const int64_t base = c * 32 * D + lane;
#pragma unroll
        for (int i = 0; i < DMAX; ++i) {
            const int64_t idx = base + (int64_t)i * 32;
            pre[i] = (i < D && idx < total) ? X[idx] : 0.f;
        }
    };

    int cur = -1;
    float s0[KC], a0[KC], a1[KC], q0[KC], q1[KC], sh0[KC], sh1[KC];
    float cnt = 0.f, llsum = 0.f;
#pragma unroll
    for (int m = 0; m < KC; ++m) s0[m] = a0[m] = a1[m] = q0[m] = q1[m] = sh0[m] = sh1[m] = 0.f;

    auto flush = [&]() {
        if (cur < 0) return;
#pragma unroll
        for (int m = 0; m < KC; ++m) {
            float *r = mine + ((size_t)cur * M + m) * row;
            if (lane == 0) {
                atomicAdd(r, s0[m]);
                if (MODE == 0 && m == 0) {
                    atomicAdd(r + 2 * D + 1, cnt);
                    atomicAdd(r + 2 * D + 2, llsum);
                }
            }
            if (h0) {
                atomicAdd(r + 1 + d0, a0[m]);
                atomicAdd(r + 1 + D + d0, q0[m]);
            }
            if (h1) {
                atomicAdd(r + 1 + d1, a1[m]);
                atomicAdd(r + 1 + D + d1, q1[m]);
            }
        }
    };

    if (wg < nchunks) fetch(wg);
    for (int64_t c = wg; c < nchunks; c += nw) {
        // ---- stage the chunk: element e = i*32 + lane of the [32, D] block goes to row e / D, column e % D
        {
            int rr = lane / D, cc = lane - rr * D;
#pragma unroll
            for (int i = 0; i < DMAX; ++i) {
                if (i < D) xs[rr * Dp + cc] = pre[i];
                cc += 32;
                while (cc >= D) {
                    cc -= D;
                    ++rr;
                }
            }
        }
        __syncwarp();
        if (c + nw < nchunks) fetch(c + nw);               // in flight while this chunk is worked on

        // ---- phase A: lane = frame
        const int64_t n = c * 32 + lane;
        int s = (n < N) ? state[n] : -1;
        if (s >= G) s = -1;
        float w[KC], llf = 0.f;
#pragma unroll
        for (int m = 0; m < KC; ++m) w[m] = 0.f;
        if (s >= 0) {
            const float *xr = xs + lane * Dp;
            const float *cg = cen + (size_t)s * KC * D;
            const float *vg = (MODE == 0) ? vv + (size_t)s * KC * D : vv + (size_t)s * D;
            float ll[KC];
#pragma unroll
            for (int m = 0; m < KC; ++m) ll[m] = 0.f;
#pragma unroll 3
            for (int d = 0; d < D; ++d) {
                const float x = xr[d];
#pragma unroll
                for (int m = 0; m < KC; ++m) {
                    const float e = x - cg[m * D + d];
                    ll[m] = fmaf(e * e, (MODE == 0) ? vg[m * D + d] : vg[d], ll[m]);
                }
            }
            if (MODE == 0) {
                float mx = -CUDART_INF_F;
#pragma unroll
                for (int m = 0; m < KC; ++m) {
                    ll[m] += aux[s * KC + m];
                    mx = fmaxf(mx, ll[m]);
                }
                float den = 0.f;
#pragma unroll
                for (int m = 0; m < KC; ++m) {
                    w[m] = __expf(ll[m] - mx);
                    den += w[m];
                }
                const float inv = 1.f / den;
#pragma unroll
                for (int m = 0; m < KC; ++m) w[m] *= inv;
                llf = mx + __logf(den);
            } else {
                float best = CUDART_INF_F;
                int bi = 0;
#pragma unroll
                for (int m = 0; m < KC; ++m)
                    if (ll[m] < best) {                    // np.argmin: first minimum (kmeans.py:185); NaN never wins
                        best = ll[m];
                        bi = m;
                    }
#pragma unroll
                for (int m = 0; m < KC; ++m) w[m] = (m == bi) ? 1.f : 0.f;
                if (argmin_out) argmin_out[n] = bi;
            }
        } else if (MODE == 1 && argmin_out && n < N) {
            argmin_out[n] = -1;
        }

        // ---- phase B: lanes = dimensions, loop over the frames of the chunk
#pragma unroll 4
        for (int j = 0; j < 32; ++j) {
            const int sj = __shfl_sync(0xffffffffu, s, j);
            if (sj != cur) {
                flush();
                cur = sj;
                cnt = llsum = 0.f;
#pragma unroll
                for (int m = 0; m < KC; ++m) {
                    s0[m] = a0[m] = a1[m] = q0[m] = q1[m] = 0.f;
                    if (sj >= 0) {
                        const float *t = (MODE == 0 ? cen : aux) + ((size_t)sj * KC + m) * D;
                        sh0[m] = h0 ? t[d0] : 0.f;
                        sh1[m] = h1 ? t[d1] : 0.f;
                    }
                }
            }
            float wj[KC];
#pragma unroll
            for (int m = 0; m < KC; ++m) wj[m] = __shfl_sync(0xffffffffu, w[m], j);
            const float lj = (MODE == 0) ? __shfl_sync(0xffffffffu, llf, j) : 0.f;
            if (sj < 0) continue;
            const float x0 = h0 ? xs[j * Dp + d0] : 0.f, x1 = h1 ? xs[j * Dp + d1] : 0.f;
            cnt += 1.f;
            llsum += lj;
#pragma unroll
            for (int m = 0; m < KC; ++m) {
                const float e0 = x0 - sh0[m], e1 = x1 - sh1[m];
                s0[m] += wj[m];
                a0[m] = fmaf(wj[m], e0, a0[m]);
                a1[m] = fmaf(wj[m], e1, a1[m]);
                q0[m] = fmaf(wj[m] * e0, e0, q0[m]);
                q1[m] = fmaf(wj[m] * e1, e1, q1[m]);
            }
        }
        __syncwarp();
    }
    flush();
}

// out[i] (+)= sum_p partial[p][i] in double, fixed order: 8 lanes of slices per element, slice p goes to lane p % 8,
// the 8 partial sums are added in lane order.
__global__ void __launch_bounds__(256)
reduce_slices8_kernel(const float *__restrict__ partial, int64_t nparts, int64_t len, double *__restrict__ out,
                      int accumulate) {
    __shared__ double part[8][32];
    const int e = threadIdx.x & 31, j = threadIdx.x >> 5;
    const int64_t i = (int64_t)blockIdx.x * 32 + e;
    double s = 0.0;
    if (i < len) {
        int64_t p = j;
        for (; p + 24 < nparts; p += 32) {
            const float v0 = partial[p * len + i], v1 = partial[(p + 8) * len + i], v2 = partial[(p + 16) * len + i],
                        v3 = partial[(p + 24) * len + i];
            s += (double)v0;
            s += (double)v1;
            s += (double)v2;
            s += (double)v3;
        }
        for (; p < nparts; p += 8) s += (double)partial[p * len + i];
    }
    part[j][e] = s;
    __syncthreads();
    if (j == 0 && i < len) {
        double t = accumulate ? out[i] : 0.0;
#pragma unroll
        for (int q = 0; q < 8; ++q) t += part[q][e];
        out[i] = t;
    }
}

size_t smem_bytes(int G, int k, int D, int mode) {
    const size_t gkd = (size_t)G * k * D;
    const size_t tables = mode == 0 ? 2 * gkd + (size_t)G * k : 2 * gkd + (size_t)G * D;
    return (tables + (size_t)PW * 32 * (D | 1)) * sizeof(float);
}

int pow2_at_least(int k) { return k <= 1 ? 1 : k <= 2 ? 2 : k <= 4 ? 4 : 8; }

template <int KC, int MODE>
int launch(const float *X, int64_t N, int D, const int32_t *state, int G, int M, const double *center, const double *var,
           const double *lc, const double *shift, float *partial, int32_t *argmin_out, int grid, size_t smem,
           cudaStream_t st) {
    if (D <= 40) {
        auto kfn = pool_stats_kernel<KC, MODE, 40>;
        SRB_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kfn<<<grid, PW * 32, smem, st>>>(X, N, D, state, G, M, center, var, lc, shift, partial, argmin_out);
    } else {
        auto kfn = pool_stats_kernel<KC, MODE, 64>;
        SRB_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kfn<<<grid, PW * 32, smem, st>>>(X, N, D, state, G, M, center, var, lc, shift, partial, argmin_out);
    }
    return SRB_OK;
}

}  // namespace
}  // namespace srb

using namespace srb;

extern "C" {

int srb_pool_stats_parts(void) { return sm_count() * PW; }

int srb_pool_stats_supported(int G, int k, int D) {
    return G > 0 && k >= 1 && k <= 8 && D > 0 && D <= 64 && smem_bytes(G, pow2_at_least(k), D, 1) <= (size_t)MAX_SMEM &&
           smem_bytes(G, pow2_at_least(k), D, 0) <= (size_t)MAX_SMEM;
}

int srb_pool_stats(const float *X, int64_t N, int D, const int32_t *state, int G, int M, int k, int mode,
                   const double *center, const double *var, const double *lc, const double *shift, float *partial,
                   double *out, int32_t *argmin_out, int accumulate, void *stream) {
    SRB_REQUIRE((X || N == 0) && (state || N == 0) && center && var && partial && out, SRB_EINVAL,
                "srb_pool_stats: null pointer");
    SRB_REQUIRE(mode == 0 || mode == 1, SRB_EINVAL, "srb_pool_stats: mode %d", mode);
    SRB_REQUIRE(mode == 0 ? lc != nullptr : shift != nullptr, SRB_EINVAL, "srb_pool_stats: lc (EM) / shift (K-means) missing");
    SRB_REQUIRE(N >= 0 && D > 0 && D <= 64 && G > 0 && M > 0 && k >= 1 && k <= M, SRB_EINVAL, "srb_pool_stats: bad N/D/G/M/k");
    SRB_REQUIRE(k == pow2_at_least(k) && k <= 8, SRB_EUNSUPPORTED, "srb_pool_stats: k=%d is not 1, 2, 4 or 8", k);
    const size_t smem = smem_bytes(G, k, D, mode);
    SRB_REQUIRE(smem <= (size_t)MAX_SMEM, SRB_EUNSUPPORTED,
                "srb_pool_stats: the parameter tables of %d pools x %d components x %d dims need %zu bytes of shared "
                "memory (limit %d); use srb_kmeans_assign + srb_group_stats / srb_gmm_stats", G, k, D, smem, MAX_SMEM);
    const int grid = sm_count();
    SRB_REQUIRE(grid > 0, SRB_ENODEVICE, "no CUDA device");
    const int64_t len = (int64_t)G * M * (2 * D + 3), parts = (int64_t)grid * PW;
    cudaStream_t st = (cudaStream_t)stream;
    SRB_CUDA(cudaMemsetAsync(partial, 0, sizeof(float) * len * parts, st));
    if (N > 0) {
        int rc = SRB_OK;
#define SRB_POOL_CASE(KC)                                                                                              \
    case KC:                                                                                                           \
        rc = mode == 0 ? launch<KC, 0>(X, N, D, state, G, M, center, var, lc, shift, partial, argmin_out, grid, smem, st) \
                       : launch<KC, 1>(X, N, D, state, G, M, center, var, lc, shift, partial, argmin_out, grid, smem, st); \
        break;
        switch (k) {
            SRB_POOL_CASE(1)
            SRB_POOL_CASE(2)
            SRB_POOL_CASE(4)
            SRB_POOL_CASE(8)
        }
#undef SRB_POOL_CASE
        if (rc != SRB_OK) return rc;
    }
    reduce_slices8_kernel<<<(int)((len + 31) / 32), 256, 0, st>>>(partial, parts, len, out, accumulate);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // extern "C"
