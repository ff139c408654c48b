// pool_stats64.cu -- srb_pool_stats in float64: the same one-pass K-means / EM statistics for many pools, with the
// reference's own arithmetic precision.
//
// Why: kmeans.py:178-193 iterates a DISCRETE map (nearest-centroid partitions) to a fixed point and stops on bitwise
// equality; GMM.em iterates to np.allclose at 1e-8 / 1e-5.  On pools of 10^4..10^5 frames the K-means objective is
// nearly flat across boundary rotations: FP32 distances / FP32-accumulated centroids send the trajectory to a different
// nearby fixed point (cluster sizes off by ~1 %, tests/test_gpu_fullsize.py measured it), which no tolerance on a single
// iteration can express.  With float64 scoring, weights and moments (X itself is float32 and converts exactly) the
// trajectory follows numpy's to ~1e-14 per step, so the batched fit reaches the reference's fixed points and stop
// iterations.  B200 runs FP64 at half the FP32 rate, and the FP32 kernel is issue-bound on packed operations that also
// take two pipe cycles: the float64 pass costs about the same per frame.
//
// Structure = pool_stats.cu (chunks of 32 frames by TMA bulk copy into the warp's shared-memory buffer, phase A lane =
// frame, phase B lanes = dimensions over runs of one state, write-once private slices with a pool mask), scalar double
// arithmetic, double slices; one stage per warp (the double parameter table takes the room of the second).
#include <cstdint>

#include "common.cuh"

namespace srb {
namespace {

constexpr int PW = 16;
constexpr int MAX_SMEM = 227 * 1024;

__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bar_init(uint64_t *bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count));
}
__device__ __forceinline__ void bar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(smem_addr(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_load(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_addr(dst)), "l"(src), "r"(bytes), "r"(smem_addr(bar)) : "memory");
}

// shared memory, in bytes: tab double[G, D, 2*KC] = {-centre[KC], scale[KC]}; aux double: K-means [G, KC, D] negated shifts,
// EM [G, KC] constants; rec double[PW, 32, 2 + KC] = {state, loglik, w[KC]}; xs float[PW, 32*ld]; PW mbarriers.
struct Layout64 {
    size_t tab, aux, rec, xs, lab, bar, bytes;
    int ld;
};
__host__ __device__ inline Layout64 layout64(int G, int KC, int D, int mode) {
    Layout64 L;
    L.ld = D | 1;
    L.tab = 0;
    L.aux = (size_t)G * D * 2 * KC * 8;
    L.rec = L.aux + (size_t)(mode == 1 ? G * KC * D : G * KC) * 8;
    L.xs = L.rec + (size_t)PW * 32 * (2 + KC) * 8;
    L.xs = (L.xs + 15) & ~(size_t)15;
    L.lab = L.xs + (size_t)PW * 32 * L.ld * 4;
    L.lab = (L.lab + 15) & ~(size_t)15;
    L.bar = L.lab + (size_t)PW * 32 * 4;
    L.bytes = L.bar + (size_t)PW * 8;
    return L;
}

template <int KC, int MODE>
__global__ void __launch_bounds__(PW * 32, 1)
pool_stats64_kernel(const float *__restrict__ X, int64_t N, int D, const int32_t *__restrict__ state, int G, int M,
                    const double *__restrict__ center, const double *__restrict__ var, const double *__restrict__ lc,
                    const double *__restrict__ shift, double *__restrict__ partial, unsigned long long *__restrict__ masks,
                    int32_t *__restrict__ argmin_out, int bulk) {
    extern __shared__ __align__(128) unsigned char smraw[];
    constexpr int RW = 2 + KC;
    const Layout64 L = layout64(G, KC, D, MODE);
    double *tab = reinterpret_cast<double *>(smraw + L.tab), *aux = reinterpret_cast<double *>(smraw + L.aux);
    const int ld = L.ld;
    for (int i = threadIdx.x; i < G * D * KC; i += blockDim.x) {
        const int g = i / (D * KC), r = i - g * (D * KC), d = r / KC, m = r - d * KC;
        const size_t src = ((size_t)g * M + m) * D + d;
        double *t = tab + ((size_t)g * D + d) * 2 * KC;
        t[m] = -center[src];
        t[KC + m] = (MODE == 0) ? -0.5 / var[src] : 0.5 / var[(size_t)g * D + d];
        if (MODE == 1) aux[((size_t)g * KC + m) * D + d] = -shift[src];
    }
    if (MODE == 0)
        for (int i = threadIdx.x; i < G * KC; i += blockDim.x) aux[i] = lc[(size_t)(i / KC) * M + (i % KC)];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    double *rec = reinterpret_cast<double *>(smraw + L.rec) + warp * 32 * RW;
    float *xs = reinterpret_cast<float *>(smraw + L.xs) + (size_t)warp * 32 * ld;
    int *lab = reinterpret_cast<int *>(smraw + L.lab) + warp * 32;
    uint64_t *bar = reinterpret_cast<uint64_t *>(smraw + L.bar) + warp;
    if (lane == 0) bar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();

    const int64_t wg = (int64_t)blockIdx.x * PW + warp, nw = (int64_t)gridDim.x * PW;
    const int row = 2 * D + 3;
    double *mine = partial + (size_t)wg * G * KC * row;
    const int64_t nchunks = (N + 31) >> 5, nfull = N >> 5;
    const int64_t cpw = (nchunks + nw - 1) / nw, c_begin = wg * cpw, c_end = min(nchunks, c_begin + cpw);
    const int d0 = lane, d1 = lane + 32;
    const bool h0 = d0 < D, h1 = d1 < D;
    const int o0 = h0 ? d0 : 0, o1 = h1 ? d1 : 0;          // lanes without a dimension read column 0, never flushed
    const uint32_t chunk_bytes = 32u * (uint32_t)D * 4u;

    int cur = -1, run_frames = 0;
    // moments are accumulated about ONE shift per state (component 0's) and moved to each component's own shift when
    // the run is flushed (exact in float64 to ~1e-16 of the moment): y and y^2 are formed once per frame, not per
    // component.  Occupancies and log-likelihoods are summed by the lane that owns the frame and reduced at the flush.
    double ps0[KC], a0[KC], a1[KC], q0[KC], q1[KC], n0[KC], n1[KC], pll = 0.0;
#pragma unroll
    for (int m = 0; m < KC; ++m) ps0[m] = a0[m] = a1[m] = q0[m] = q1[m] = n0[m] = n1[m] = 0.0;

    unsigned long long touched = 0ull;
    auto flush = [&]() {
        if (cur < 0) return;
        const bool first = masks != nullptr && !((touched >> cur) & 1ull);
        touched |= 1ull << (cur & 63);
        double ll = pll;
        if (MODE == 0) {
#pragma unroll
            for (int o = 16; o; o >>= 1) ll += __shfl_xor_sync(0xffffffffu, ll, o);
        }
#pragma unroll
        for (int m = 0; m < KC; ++m) {
            double S = ps0[m];
#pragma unroll
            for (int o = 16; o; o >>= 1) S += __shfl_xor_sync(0xffffffffu, S, o);
            // from the state's shift (component 0's) to this component's: e = y - dlt
            const double dl0 = n0[0] - n0[m], dl1 = n1[0] - n1[m];
            const double A0 = a0[m] - dl0 * S, A1 = a1[m] - dl1 * S;
            const double Q0 = q0[m] - 2.0 * dl0 * a0[m] + dl0 * dl0 * S, Q1 = q1[m] - 2.0 * dl1 * a1[m] + dl1 * dl1 * S;
            double *r = mine + ((size_t)cur * KC + m) * row;
            const double c1 = (MODE == 0 && m == 0) ? (double)run_frames : 0.0, c2 = (MODE == 0 && m == 0) ? ll : 0.0;
            if (first) {
                if (lane == 0) {
                    r[0] = S;
                    r[2 * D + 1] = c1;
                    r[2 * D + 2] = c2;
                }
                if (h0) {
                    r[1 + d0] = A0;
                    r[1 + D + d0] = Q0;
                }
                if (h1) {
                    r[1 + d1] = A1;
                    r[1 + D + d1] = Q1;
                }
            } else {
                if (lane == 0) {
                    atomicAdd(r, S);
                    if (MODE == 0 && m == 0) {
                        atomicAdd(r + 2 * D + 1, c1);
                        atomicAdd(r + 2 * D + 2, c2);
                    }
                }
                if (h0) {
                    atomicAdd(r + 1 + d0, A0);
                    atomicAdd(r + 1 + D + d0, Q0);
                }
                if (h1) {
                    atomicAdd(r + 1 + d1, A1);
                    atomicAdd(r + 1 + D + d1, Q1);
                }
            }
            ps0[m] = a0[m] = a1[m] = q0[m] = q1[m] = 0.0;
        }
        pll = 0.0;
        run_frames = 0;
    };

    if (bulk && lane == 0 && c_begin < c_end && c_begin < nfull) {
        bar_expect_tx(bar, chunk_bytes + 128u);
        bulk_load(xs, X + c_begin * 32 * D, chunk_bytes, bar);
        bulk_load(lab, state + c_begin * 32, 128u, bar);
    }
    int it = 0;
    for (int64_t c = c_begin; c < c_end; ++c, ++it) {
        if (bulk && c < nfull) {
            bar_wait(bar, (uint32_t)(it & 1));
        } else {
            const int64_t base = c * 32 * D;
            const int nel = (int)min((int64_t)32 * D, N * D - base);
            for (int e = lane; e < nel; e += 32) xs[(e / D) * ld + (e % D)] = X[base + e];
            __syncwarp();
        }
        // ---- phase A: lane = frame
        const int64_t n = c * 32 + lane;
        int s = (bulk && c < nfull) ? lab[lane] : ((n < N) ? state[n] : -1);   // labels ride in the same bulk transaction
        if (s >= G) s = -1;
        double w[KC], llf = 0.0;
#pragma unroll
        for (int m = 0; m < KC; ++m) w[m] = 0.0;
        if (s >= 0) {
            const float *xr = xs + lane * ld;
            const double *tg = tab + (size_t)s * D * 2 * KC;
            double ll[KC];
#pragma unroll
            for (int m = 0; m < KC; ++m) ll[m] = 0.0;
#pragma unroll 39
            for (int d = 0; d < D; ++d) {
                const double x = (double)xr[d];
#pragma unroll
                for (int m = 0; m < KC; ++m) {
                    const double e = x + tg[d * 2 * KC + m];
                    ll[m] = fma(e * e, tg[d * 2 * KC + KC + m], ll[m]);
                }
            }
            if (MODE == 0) {
                double mx = -CUDART_INF;
#pragma unroll
                for (int m = 0; m < KC; ++m) {
                    ll[m] += aux[s * KC + m];
                    mx = fmax(mx, ll[m]);
                }
                double den = 0.0;
#pragma unroll
                for (int m = 0; m < KC; ++m) {
                    w[m] = exp(ll[m] - mx);
                    den += w[m];
                }
#pragma unroll
                for (int m = 0; m < KC; ++m) w[m] /= den;
                llf = mx + log(den);
            } else {
                double best = CUDART_INF;
                int bi = 0;
#pragma unroll
                for (int m = 0; m < KC; ++m)
                    if (ll[m] < best) {                    // np.argmin: first minimum (kmeans.py:185); NaN never wins
                        best = ll[m];
                        bi = m;
                    }
#pragma unroll
                for (int m = 0; m < KC; ++m) w[m] = (m == bi) ? 1.0 : 0.0;
                if (argmin_out) argmin_out[n] = bi;
            }
        } else if (MODE == 1 && argmin_out && n < N) {
            argmin_out[n] = -1;
        }
        {
            double *r = rec + lane * RW;
            r[0] = (double)s;
            r[1] = llf;
#pragma unroll
            for (int m = 0; m < KC; ++m) r[2 + m] = w[m];
        }
        __syncwarp();

        // ---- phase B: lanes = dimensions (d0, d1), runs of one state, accumulators carried across chunks
        {
            const int s_up = __shfl_up_sync(0xffffffffu, s, 1);
            const unsigned starts = __ballot_sync(0xffffffffu, lane == 0 || s != s_up);
            int j0 = 0;
            while (j0 < 32) {
                const unsigned rest = starts & ~((2u << j0) - 1u);
                const int j1 = rest ? __ffs(rest) - 1 : 32;
                const int sj = __shfl_sync(0xffffffffu, s, j0);
                if (sj != cur) {
                    flush();
                    cur = sj;
                    if (sj >= 0) {
#pragma unroll
                        for (int m = 0; m < KC; ++m) {
                            if (MODE == 0) {
                                const double *t = tab + (size_t)sj * D * 2 * KC + m;
                                n0[m] = t[o0 * 2 * KC];
                                n1[m] = t[o1 * 2 * KC];
                            } else {
                                const double *t = aux + ((size_t)sj * KC + m) * D;
                                n0[m] = t[o0];
                                n1[m] = t[o1];
                            }
                        }
                    }
                }
                if (sj >= 0) {
                    run_frames += j1 - j0;
                    const bool own = lane >= j0 && lane < j1;
#pragma unroll
                    for (int m = 0; m < KC; ++m) ps0[m] += own ? w[m] : 0.0;
                    if (MODE == 0) pll += own ? llf : 0.0;
                    const float *xp0 = xs + j0 * ld + o0, *xp1 = xs + j0 * ld + o1;
                    const double *rp = rec + j0 * RW + 2;
                    const double c0 = n0[0], c1 = n1[0];
#pragma unroll 8
                    for (int j = j0; j < j1; ++j, xp0 += ld, xp1 += ld, rp += RW) {
                        const double y0 = (double)*xp0 + c0, y1 = (double)*xp1 + c1;
                        const double z0 = y0 * y0, z1 = y1 * y1;
#pragma unroll
                        for (int m = 0; m < KC; ++m) {
                            const double wm = rp[m];
                            a0[m] = fma(wm, y0, a0[m]);
                            a1[m] = fma(wm, y1, a1[m]);
                            q0[m] = fma(wm, z0, q0[m]);
                            q1[m] = fma(wm, z1, q1[m]);
                        }
                    }
                }
                j0 = j1;
            }
        }
        __syncwarp();
        if (bulk && lane == 0 && c + 1 < c_end && c + 1 < nfull) {
            bar_expect_tx(bar, chunk_bytes + 128u);
            bulk_load(xs, X + (c + 1) * 32 * D, chunk_bytes, bar);
            bulk_load(lab, state + (c + 1) * 32, 128u, bar);
        }
    }
    flush();
    if (masks != nullptr && lane == 0) masks[wg] = touched;
}

// out[G*M rows] (+)= sum over the double slices [G*k rows], fixed order (32 lanes of slices per element, lane order)
__global__ void __launch_bounds__(1024)
reduce_slices64_kernel(const double *__restrict__ partial, const unsigned long long *__restrict__ masks, int64_t nparts,
                       int M, int k, int row, int64_t len_out, double *__restrict__ out, int accumulate) {
    __shared__ double part[32][33];
    const int e = threadIdx.x & 31, j = threadIdx.x >> 5;
    const int64_t i = (int64_t)blockIdx.x * 32 + e;
    double s = 0.0;
    if (i < len_out) {
        const int64_t g = i / ((int64_t)M * row);
        const int r = (int)(i - g * M * row), m = r / row, col = r - m * row;
        if (m < k) {
            const int64_t len = len_out / M * k, src = (g * k + m) * row + col;
            for (int64_t p = j; p < nparts; p += 32)
                if (masks == nullptr || ((masks[p] >> g) & 1ull)) s += partial[p * len + src];
        }
    }
    part[j][e] = s;
    __syncthreads();
    if (j == 0 && i < len_out) {
        double t = accumulate ? out[i] : 0.0;
#pragma unroll
        for (int q = 0; q < 32; ++q) t += part[q][e];
        out[i] = t;
    }
}

template <int KC, int MODE>
int launch64(const float *X, int64_t N, int D, const int32_t *state, int G, int M, const double *center, const double *var,
             const double *lc, const double *shift, double *partial, unsigned long long *masks, int32_t *argmin_out,
             int grid, cudaStream_t st) {
    const size_t smem = layout64(G, KC, D, MODE).bytes;
    const int bulk = (D & 1) && ((reinterpret_cast<uintptr_t>(X) | reinterpret_cast<uintptr_t>(state)) & 15) == 0;
    auto kfn = pool_stats64_kernel<KC, MODE>;
    SRB_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kfn<<<grid, PW * 32, smem, st>>>(X, N, D, state, G, M, center, var, lc, shift, partial, masks, argmin_out, bulk);
    return SRB_OK;
}

}  // namespace
}  // namespace srb

using namespace srb;

extern "C" {

int srb_pool_stats_f64_supported(int G, int k, int D) {
    return G > 0 && (k == 1 || k == 2 || k == 4) && D > 0 && D <= 64 && layout64(G, k, D, 1).bytes <= (size_t)MAX_SMEM &&
           layout64(G, k, D, 0).bytes <= (size_t)MAX_SMEM;
}

int64_t srb_pool_stats_f64_partial_doubles(int G, int k, int D) {
    const int64_t parts = (int64_t)sm_count() * PW;
    return parts * G * k * (2 * D + 3) + parts;
}

int srb_pool_stats_f64(const float *X, int64_t N, int D, const int32_t *state, int G, int M, int k, int mode,
                       const double *center, const double *var, const double *lc, const double *shift, double *partial,
                       double *out, int32_t *argmin_out, int accumulate, void *stream) {
    SRB_REQUIRE((X || N == 0) && (state || N == 0) && center && var && partial && out, SRB_EINVAL,
                "srb_pool_stats_f64: null pointer");
    SRB_REQUIRE(mode == 0 || mode == 1, SRB_EINVAL, "srb_pool_stats_f64: mode %d", mode);
    SRB_REQUIRE(mode == 0 ? lc != nullptr : shift != nullptr, SRB_EINVAL, "srb_pool_stats_f64: lc (EM) / shift (K-means) missing");
    SRB_REQUIRE(N >= 0 && D > 0 && D <= 64 && G > 0 && M > 0 && k >= 1 && k <= M, SRB_EINVAL, "srb_pool_stats_f64: bad N/D/G/M/k");
    SRB_REQUIRE(k == 1 || k == 2 || k == 4, SRB_EUNSUPPORTED, "srb_pool_stats_f64: k=%d is not 1, 2 or 4", k);
    const size_t smem = layout64(G, k, D, mode).bytes;
    SRB_REQUIRE(smem <= (size_t)MAX_SMEM, SRB_EUNSUPPORTED,
                "srb_pool_stats_f64: the double parameter tables of %d pools x %d components x %d dims need %zu bytes of "
                "shared memory (limit %d); use srb_pool_stats", G, k, D, smem, MAX_SMEM);
    const int grid = sm_count();
    SRB_REQUIRE(grid > 0, SRB_ENODEVICE, "no CUDA device");
    const int row = 2 * D + 3;
    const int64_t len_out = (int64_t)G * M * row, len = (int64_t)G * k * row, parts = (int64_t)grid * PW;
    cudaStream_t st = (cudaStream_t)stream;
    unsigned long long *masks = G <= 64 ? reinterpret_cast<unsigned long long *>(partial + len * parts) : nullptr;
    if (masks) SRB_CUDA(cudaMemsetAsync(masks, 0, sizeof(unsigned long long) * parts, st));
    else SRB_CUDA(cudaMemsetAsync(partial, 0, sizeof(double) * len * parts, st));
    if (N > 0) {
        int rc = SRB_OK;
#define SRB_POOL64_CASE(KC)                                                                                            \
    case KC:                                                                                                           \
        rc = mode == 0 ? launch64<KC, 0>(X, N, D, state, G, M, center, var, lc, shift, partial, masks, argmin_out, grid, st) \
                       : launch64<KC, 1>(X, N, D, state, G, M, center, var, lc, shift, partial, masks, argmin_out, grid, st); \
        break;
        switch (k) {
            SRB_POOL64_CASE(1)
            SRB_POOL64_CASE(2)
            SRB_POOL64_CASE(4)
        }
#undef SRB_POOL64_CASE
        if (rc != SRB_OK) return rc;
    }
    reduce_slices64_kernel<<<(int)((len_out + 31) / 32), 1024, 0, st>>>(partial, masks, parts, M, k, row, len_out, out, accumulate);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // extern "C"
