// fit_update.cu -- the per-iteration update of the batched HMM.fit (srb200/fit.py): what the reference does in numpy
// float64 between two passes over the data, for every pool (HMM state) at once and without leaving the device.
//
//   srb_fit_em_update      GMM.em's M-step and stop test (hmm_state.py:138-159) for the first k components of every
//                          state: mu += S1/S0, var = S2/S0 - (S1/S0)^2, w = S0/count; a state whose new parameters
//                          satisfy np.allclose against the previous ones (atol 1e-8, rtol 1e-5) stops HERE and is frozen;
//                          the others copy the new parameters to *_old.  Also refreshes the constants lc the next
//                          E-step needs (srb_gauss_consts).
//   srb_fit_kmeans_update  kmeans.py:187-192: new centroid = shift + S1/S0; a pool whose new centroids equal the previous
//                          ones exactly (np.array_equal) stops and keeps them.
//
// One CTA per pool.  The arithmetic is the reference's, operation by operation, in IEEE double WITHOUT fused
// multiply-adds (numpy does not contract): the stop iteration of a state is the one the reference's loop would take on the
// same statistics.
#include <cstdint>

#include "common.cuh"

namespace srb {
namespace {

constexpr int UT = 128;

__device__ __forceinline__ bool np_isclose(double a, double b) {          // numpy.isclose(a, b): |a - b| <= atol + rtol |b|
    return fabs(__dsub_rn(a, b)) <= __dadd_rn(1e-8, __dmul_rn(1e-5, fabs(b)));
}

__global__ void __launch_bounds__(UT)
fit_em_update_kernel(const double *__restrict__ stats, int M, int k, int D, double *__restrict__ mu,
                     double *__restrict__ var, double *__restrict__ w, double *__restrict__ mu_old,
                     double *__restrict__ var_old, double *__restrict__ w_old, uint8_t *__restrict__ active,
                     int64_t *__restrict__ its, int64_t it, double *__restrict__ lc) {
    const int g = blockIdx.x, row = 2 * D + 3;
    const double *st = stats + (size_t)g * M * row;
    const bool act = active[g] != 0;
    const double cntg = st[2 * D + 1];
    bool conv = true;
    for (int i = threadIdx.x; i < k * D; i += UT) {
        const int m = i / D, d = i - m * D;
        const size_t p = ((size_t)g * M + m) * D + d;
        const double S0 = st[(size_t)m * row];
        double a = mu[p], v = var[p];
        if (act && S0 > 0.0) {                              // empty components stay as they are
            const double m1 = __ddiv_rn(st[(size_t)m * row + 1 + d], S0);
            a = __dadd_rn(a, m1);
            v = __dsub_rn(__ddiv_rn(st[(size_t)m * row + 1 + D + d], S0), __dmul_rn(m1, m1));
            mu[p] = a;
            var[p] = v;
        }
        conv = conv && np_isclose(a, mu_old[p]) && np_isclose(v, var_old[p]);
    }
    for (int m = threadIdx.x; m < k; m += UT) {
        const size_t p = (size_t)g * M + m;
        const double S0 = st[(size_t)m * row];
        double x = w[p];
        if (act && S0 > 0.0 && cntg > 0.0) {
            x = __ddiv_rn(S0, cntg);
            w[p] = x;
        }
        conv = conv && np_isclose(x, w_old[p]);
    }
    const bool all_conv = __syncthreads_and(conv) != 0;     // also orders the writes above before the reads below
    const bool still = act && !all_conv;
    if (still) {
        for (int i = threadIdx.x; i < k * D; i += UT) {
            const int m = i / D, d = i - m * D;
            const size_t p = ((size_t)g * M + m) * D + d;
            mu_old[p] = mu[p];
            var_old[p] = var[p];
        }
        for (int m = threadIdx.x; m < k; m += UT) w_old[(size_t)g * M + m] = w[(size_t)g * M + m];
    }
    if (threadIdx.x == 0) {
        if (act) its[g] = it;
        active[g] = still ? 1 : 0;
    }
    // lc[g, m] = log w - 0.5 (D log 2 pi + sum log var) for all M components (gauss_consts_kernel's formula); a warp per
    // component, lanes over dimensions, fixed-order tree
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int m = warp; m < M; m += UT / 32) {
        const size_t p = (size_t)g * M + m;
        double ld = 0.0;
        for (int d = lane; d < D; d += 32) ld += log(var[p * D + d]);
#pragma unroll
        for (int o = 16; o; o >>= 1) ld += __shfl_xor_sync(0xffffffffu, ld, o);
        if (lane == 0) lc[p] = (w[p] > 0.0 ? log(w[p]) : -CUDART_INF) - 0.5 * ((double)D * 1.8378770664093454836 + ld);
    }
}

__global__ void __launch_bounds__(UT)
fit_kmeans_update_kernel(const double *__restrict__ stats, int k, int D, const double *__restrict__ shift,
                         double *__restrict__ cen, uint8_t *__restrict__ active, int64_t *__restrict__ its, int64_t it,
                         double *__restrict__ cnt) {
    const int g = blockIdx.x, row = 2 * D + 3;
    const double *st = stats + (size_t)g * k * row;
    const bool act = active[g] != 0;
    bool same = true;
    for (int i = threadIdx.x; i < k * D; i += UT) {
        const int m = i / D, d = i - m * D;
        const size_t p = ((size_t)g * k + m) * D + d;
        const double nc = __dadd_rn(shift[p], __ddiv_rn(st[(size_t)m * row + 1 + d], st[(size_t)m * row]));
        same = same && (nc == cen[p]);                      // NaN never equal, like np.array_equal
    }
    const bool all_same = __syncthreads_and(same) != 0;
    const bool still = act && !all_same;
    if (still)
        for (int i = threadIdx.x; i < k * D; i += UT) {
            const int m = i / D, d = i - m * D;
            const size_t p = ((size_t)g * k + m) * D + d;
            cen[p] = __dadd_rn(shift[p], __ddiv_rn(st[(size_t)m * row + 1 + d], st[(size_t)m * row]));
        }
    if (act)
        for (int m = threadIdx.x; m < k; m += UT) cnt[(size_t)g * k + m] = st[(size_t)m * row];
    if (threadIdx.x == 0) {
        if (act) its[g] = it;
        active[g] = still ? 1 : 0;
    }
}

}  // namespace
}  // namespace srb

using namespace srb;

extern "C" {

int srb_fit_em_update(const double *stats, int G, int M, int k, int D, double *mu, double *var, double *w, double *mu_old,
                      double *var_old, double *w_old, uint8_t *active, int64_t *its, int64_t it, double *lc,
                      void *stream) {
    SRB_REQUIRE(stats && mu && var && w && mu_old && var_old && w_old && active && its && lc, SRB_EINVAL,
                "srb_fit_em_update: null pointer");
    SRB_REQUIRE(G > 0 && M > 0 && k >= 1 && k <= M && D > 0, SRB_EINVAL, "srb_fit_em_update: bad G/M/k/D");
    fit_em_update_kernel<<<G, UT, 0, (cudaStream_t)stream>>>(stats, M, k, D, mu, var, w, mu_old, var_old, w_old, active,
                                                             its, it, lc);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int srb_fit_kmeans_update(const double *stats, int G, int k, int D, const double *shift, double *cen, uint8_t *active,
                          int64_t *its, int64_t it, double *cnt, void *stream) {
    SRB_REQUIRE(stats && shift && cen && active && its && cnt, SRB_EINVAL, "srb_fit_kmeans_update: null pointer");
    SRB_REQUIRE(G > 0 && k >= 1 && D > 0, SRB_EINVAL, "srb_fit_kmeans_update: bad G/k/D");
    fit_kmeans_update_kernel<<<G, UT, 0, (cudaStream_t)stream>>>(stats, k, D, shift, cen, active, its, it, cnt);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // extern "C"
