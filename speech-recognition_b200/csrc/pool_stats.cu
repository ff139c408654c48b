// pool_stats.cu -- the statistics of one K-means or one EM iteration for MANY pools (HMM states) in one pass over X.
//
// The reference fits one state at a time (hmm.py:97-124): kmeans (kmeans.py:166-193) assigns every frame of the state's
// pool to the nearest of k centroids and re-estimates them; GMM.em (hmm_state.py:122-159) weighs every frame by its
// posterior over the state's first k components.  Both are "weights over the k components of the frame's own state,
// then weighted moments": here one kernel does either for all W*S pools at once, and the K-means assignment is fused
// with the moments (X is read once per iteration instead of twice).
//
// A warp owns chunks of 32 consecutive frames = 128*D contiguous bytes, brought into the warp's own shared-memory ring
// (two stages when they fit) by ONE bulk copy of the TMA engine per chunk (cp.async.bulk + mbarrier complete_tx; plain
// loads for the ragged last chunk, an even D or a misaligned X).  Phase A: lane = frame, scores its k components against
// the parameter table in shared memory ({-centre, scale} interleaved over the components: one 128-bit broadcast load and
// three packed fma.rn.f32x2 per dimension and component pair).  Phase B: lanes = dimensions (d, d+32 packed in a float2),
// a loop over the 32 frames reads the frame's {state, weights} record (one broadcast load) and adds the weighted first
// and second moments about the component's shift to register accumulators; when the state changes (the
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
// TMA engine, bulk (1-D) form: global -> shared, completion counted in bytes on the mbarrier
__device__ __forceinline__ void bulk_load(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_addr(dst)), "l"(src), "r"(bytes), "r"(smem_addr(bar)) : "memory");
}

__host__ __device__ constexpr int rec_floats(int KC) { return KC <= 2 ? 4 : KC == 4 ? 8 : 12; }   // {state, loglik, w[KC]}
__host__ __device__ inline int up4(int n) { return (n + 3) & ~3; }

// shared memory, in floats: tab [G, D, 2*KC] = {-centre[KC], scale[KC]} per (pool, dim); shf: K-means [G, KC, D] negated
// shifts, EM [G, KC] constants log w - 0.5 log|2 pi var| (EM shifts by the centre); rec [PW, 32, rec_floats]; xs [PW, nst, 32*ld]; then PW*nst mbarriers.
struct PoolLayout {
    int tab, shf, rec, xs, lab, bar_bytes_off, ld, stage;
    size_t bytes;
};
__host__ __device__ inline PoolLayout pool_layout(int G, int KC, int D, int mode, int nst) {
    PoolLayout L;
    L.ld = D | 1;
    L.stage = 32 * L.ld;
    L.tab = 0;
    L.shf = up4(G * D * 2 * KC);
    L.rec = L.shf + (mode == 1 ? up4(G * KC * D) : up4(G * KC));
    L.xs = L.rec + PW * 32 * rec_floats(KC);
    L.lab = L.xs + PW * nst * L.stage;                     // the chunk's 32 labels per stage (same bulk-copy transaction)
    L.bar_bytes_off = (L.lab + PW * nst * 32) * 4;
    L.bytes = (size_t)L.bar_bytes_off + (size_t)PW * nst * 8;
    return L;
}

template <int KC, int MODE>   // MODE 0: EM posteriors (hmm_state.py:122-136)  1: K-means one-hot (kmeans.py:180-187)
__global__ void __launch_bounds__(PW * 32, 1)
pool_stats_kernel(const float *__restrict__ X, int64_t N, int D, const int32_t *__restrict__ state, int G, int M,
                  const double *__restrict__ center, const double *__restrict__ var, const double *__restrict__ lc,
                  const double *__restrict__ shift, float *__restrict__ partial, unsigned long long *__restrict__ masks,
                  int32_t *__restrict__ argmin_out, int nst, int bulk) {
    extern __shared__ __align__(128) float sm[];
    constexpr int RW = rec_floats(KC), K2 = (KC + 1) / 2;
    const PoolLayout L = pool_layout(G, KC, D, MODE, nst);
    float *tab = sm + L.tab, *shf = sm + L.shf;
    const int ld = L.ld;
    // scale: EM -0.5/var of the component, K-means 0.5/var0 of the pool (kmeans.py:183: one variance row for all centroids)
    for (int i = threadIdx.x; i < G * D * KC; i += blockDim.x) {
        const int g = i / (D * KC), r = i - g * (D * KC), d = r / KC, m = r - d * KC;
        const size_t src = ((size_t)g * M + m) * D + d;
        float *t = tab + ((size_t)g * D + d) * 2 * KC;
        t[m] = -(float)center[src];
        t[KC + m] = (MODE == 0) ? (float)(-0.5 / var[src]) : (float)(0.5 / var[(size_t)g * D + d]);
        if (MODE == 1) shf[((size_t)g * KC + m) * D + d] = -(float)shift[src];
    }
    if (MODE == 0)
        for (int i = threadIdx.x; i < G * KC; i += blockDim.x) shf[i] = (float)lc[(size_t)(i / KC) * M + (i % KC)];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float *rec = sm + L.rec + warp * 32 * RW;
    float *xs0 = sm + L.xs + (size_t)warp * nst * L.stage;
    int *lab0 = reinterpret_cast<int *>(sm + L.lab) + warp * nst * 32;
    uint64_t *bars = reinterpret_cast<uint64_t *>(reinterpret_cast<char *>(sm) + L.bar_bytes_off) + warp * nst;
    if (lane == 0)
        for (int i = 0; i < nst; ++i) bar_init(&bars[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();

    const int64_t wg = (int64_t)blockIdx.x * PW + warp, nw = (int64_t)gridDim.x * PW;
    const int row = 2 * D + 3;
    float *mine = partial + (size_t)wg * G * KC * row;          // private slice: the KC scored components of every pool
    const int64_t nchunks = (N + 31) >> 5, nfull = N >> 5;      // chunks [0, nfull) hold 32 frames
    // a warp works on a CONTIGUOUS range of chunks: runs of one state continue across chunks, and the few pools a warp
    // meets keep its slice rows in L2
    const int64_t cpw = (nchunks + nw - 1) / nw, c_begin = wg * cpw, c_end = min(nchunks, c_begin + cpw);
    const int d0 = lane, d1 = lane + 32;
    const bool h0 = d0 < D, h1 = d1 < D;
    const uint32_t chunk_bytes = 32u * (uint32_t)D * 4u;

    int cur = -1;
    float s0[KC];
    float2 a2[KC], q2[KC], nsh[KC];
    float llsum = 0.f;
    int run_frames = 0;                                     // frames of the current run (the count column)
#pragma unroll
    for (int m = 0; m < KC; ++m) {
        s0[m] = 0.f;
        a2[m] = q2[m] = nsh[m] = make_float2(0.f, 0.f);
    }

    // masks != nullptr (G <= 64): the slice is NOT zeroed beforehand; the first flush of a pool STORES its rows and sets
    // the pool's bit in the warp's mask, later ones add; the reduction reads a slice's rows of pool g only if bit g is set.
    unsigned long long touched = 0ull;
    auto flush = [&]() {
        if (cur < 0) return;
        const bool first = masks != nullptr && !((touched >> cur) & 1ull);
        touched |= 1ull << (cur & 63);
#pragma unroll
        for (int m = 0; m < KC; ++m) {
            float *r = mine + ((size_t)cur * KC + m) * row;
            const float c1 = (MODE == 0 && m == 0) ? (float)run_frames : 0.f, c2 = (MODE == 0 && m == 0) ? llsum : 0.f;
            if (first) {
                if (lane == 0) {
                    r[0] = s0[m];
                    r[2 * D + 1] = c1;
                    r[2 * D + 2] = c2;
                }
                if (h0) {
                    r[1 + d0] = a2[m].x;
                    r[1 + D + d0] = q2[m].x;
                }
                if (h1) {
                    r[1 + d1] = a2[m].y;
                    r[1 + D + d1] = q2[m].y;
                }
            } else {
                if (lane == 0) {
                    atomicAdd(r, s0[m]);
                    if (MODE == 0 && m == 0) {
                        atomicAdd(r + 2 * D + 1, c1);
                        atomicAdd(r + 2 * D + 2, c2);
                    }
                }
                if (h0) {
                    atomicAdd(r + 1 + d0, a2[m].x);
                    atomicAdd(r + 1 + D + d0, q2[m].x);
                }
                if (h1) {
                    atomicAdd(r + 1 + d1, a2[m].y);
                    atomicAdd(r + 1 + D + d1, q2[m].y);
                }
            }
        }
    };

    // ---- prologue: the first nst chunks of this warp
    if (bulk && lane == 0)
        for (int i = 0; i < nst; ++i) {
            const int64_t c = c_begin + i;
            if (c < c_end && c < nfull) {
                bar_expect_tx(&bars[i], chunk_bytes + 128u);
                bulk_load(xs0 + (size_t)i * L.stage, X + c * 32 * D, chunk_bytes, &bars[i]);
                bulk_load(lab0 + i * 32, state + c * 32, 128u, &bars[i]);
            }
        }
    // lanes without a dimension (d >= D) read column 0 and feed accumulators that are never flushed
    const int o0 = h0 ? d0 : 0, o1 = h1 ? d1 : 0;
    int it = 0;
    for (int64_t c = c_begin; c < c_end; ++c, ++it) {
        const int stg = (nst == 2) ? (it & 1) : 0;
        float *xs = xs0 + (size_t)stg * L.stage;
        if (bulk && c < nfull) {
            bar_wait(&bars[stg], (uint32_t)((nst == 2 ? (it >> 1) : it) & 1));
        } else {                                            // tail chunk, or X / D not fit for the bulk copy
            const int64_t base = c * 32 * D;
            const int nel = (int)min((int64_t)32 * D, N * D - base);
            for (int e = lane; e < nel; e += 32) xs[(e / D) * ld + (e % D)] = X[base + e];
            __syncwarp();
        }

        // ---- phase A: lane = frame, scores the k components of its own pool (its label was fetched a chunk ahead)
        const int64_t n = c * 32 + lane;
        // (labels: with the bulk copy they arrived in shared memory beside the features -- a global load here would share
        // a scoreboard with everything else in flight and stall the chunk's first instructions)
        int s = (bulk && c < nfull) ? lab0[stg * 32 + lane] : ((n < N) ? state[n] : -1);
        if (s >= G) s = -1;
        float w[KC], llf = 0.f;
#pragma unroll
        for (int m = 0; m < KC; ++m) w[m] = 0.f;
        if (s >= 0) {
            const float *xr = xs + lane * ld;
            const float *tg = tab + (size_t)s * D * 2 * KC;
            float ll[KC];
            if constexpr (KC == 1) {
                float acc = 0.f;
#pragma unroll 13
                for (int d = 0; d < D; ++d) {
                    const float2 t = *reinterpret_cast<const float2 *>(tg + 2 * d);
                    const float e = xr[d] + t.x;
                    acc = fmaf(e * e, t.y, acc);
                }
                ll[0] = acc;
            } else {
                float2 acc[K2];
#pragma unroll
                for (int h = 0; h < K2; ++h) acc[h] = make_float2(0.f, 0.f);
#pragma unroll 13
                for (int d = 0; d < D; ++d) {
                    const float x = xr[d];
                    const float2 x2 = make_float2(x, x);
                    const float4 *t4 = reinterpret_cast<const float4 *>(tg + (size_t)d * 2 * KC);
                    if constexpr (KC == 2) {
                        const float4 t = t4[0];                                   // {-c0, -c1, v0, v1}
                        const float2 e = __fadd2_rn(x2, make_float2(t.x, t.y));
                        acc[0] = __ffma2_rn(__fmul2_rn(e, e), make_float2(t.z, t.w), acc[0]);
                    } else {
#pragma unroll
                        for (int h = 0; h < KC / 4; ++h) {
                            const float4 tc = t4[h], tv = t4[KC / 4 + h];         // {-c[4h..4h+4)}, {v[4h..4h+4)}
                            const float2 ea = __fadd2_rn(x2, make_float2(tc.x, tc.y));
                            const float2 eb = __fadd2_rn(x2, make_float2(tc.z, tc.w));
                            acc[2 * h] = __ffma2_rn(__fmul2_rn(ea, ea), make_float2(tv.x, tv.y), acc[2 * h]);
                            acc[2 * h + 1] = __ffma2_rn(__fmul2_rn(eb, eb), make_float2(tv.z, tv.w), acc[2 * h + 1]);
                        }
                    }
                }
#pragma unroll
                for (int h = 0; h < K2; ++h) {
                    ll[2 * h] = acc[h].x;
                    ll[2 * h + 1] = acc[h].y;
                }
            }
            if (MODE == 0) {
                float mx = -CUDART_INF_F;
#pragma unroll
                for (int m = 0; m < KC; ++m) {
                    ll[m] += shf[s * KC + m];                                     // log w - 0.5 log|2 pi var|
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
        {   // the frame's record for phase B: {state, loglik, w[KC]}
            float *r = rec + lane * RW;
            if constexpr (KC == 1) {
                *reinterpret_cast<float4 *>(r) = make_float4(__int_as_float(s), llf, w[0], 0.f);
            } else {
                *reinterpret_cast<float4 *>(r) = make_float4(__int_as_float(s), llf, w[0], w[1]);
#pragma unroll
                for (int h = 1; h < RW / 4; ++h) {
                    const int b = 4 * h - 2;
                    *reinterpret_cast<float4 *>(r + 4 * h) = make_float4(b < KC ? w[b] : 0.f, b + 1 < KC ? w[b + 1] : 0.f,
                                                                         b + 2 < KC ? w[b + 2] : 0.f, b + 3 < KC ? w[b + 3] : 0.f);
                }
            }
        }
        __syncwarp();

        // ---- phase B: lanes = dimensions (d0, d1 packed in one float2).  The chunk is cut into runs of one state (a
        // ballot of the state changes); a run's frames go through a branch-free loop, the accumulators are flushed
        // between runs of different states and carried across chunks.  Runs of skipped frames (state < 0) cost nothing.
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
                    llsum = 0.f;
                    run_frames = 0;
#pragma unroll
                    for (int m = 0; m < KC; ++m) {
                        s0[m] = 0.f;
                        a2[m] = q2[m] = make_float2(0.f, 0.f);
                        if (sj >= 0) {
                            if (MODE == 0) {
                                const float *t = tab + (size_t)sj * D * 2 * KC + m;
                                nsh[m] = make_float2(t[o0 * 2 * KC], t[o1 * 2 * KC]);
                            } else {
                                const float *t = shf + ((size_t)sj * KC + m) * D;
                                nsh[m] = make_float2(t[o0], t[o1]);
                            }
                        }
                    }
                }
                if (sj >= 0) {
                    run_frames += j1 - j0;
                    const float *xp0 = xs + j0 * ld + o0, *xp1 = xs + j0 * ld + o1;
                    const float *rp = rec + j0 * RW;
#pragma unroll 4
                    for (int j = j0; j < j1; ++j, xp0 += ld, xp1 += ld, rp += RW) {
                        const float4 r0 = *reinterpret_cast<const float4 *>(rp);  // broadcast {state, loglik, w0, w1}
                        float wj[KC];
                        if constexpr (KC == 1) {
                            wj[0] = r0.z;
                        } else {
                            wj[0] = r0.z;
                            wj[1] = r0.w;
#pragma unroll
                            for (int h = 1; h < RW / 4; ++h) {
                                const float4 rh = *reinterpret_cast<const float4 *>(rp + 4 * h);
                                const int b = 4 * h - 2;
                                if (b < KC) wj[b] = rh.x;
                                if (b + 1 < KC) wj[b + 1] = rh.y;
                                if (b + 2 < KC) wj[b + 2] = rh.z;
                                if (b + 3 < KC) wj[b + 3] = rh.w;
                            }
                        }
                        const float2 x2 = make_float2(*xp0, *xp1);
                        if (MODE == 0) llsum += r0.y;
#pragma unroll
                        for (int m = 0; m < KC; ++m) {
                            const float2 w2 = make_float2(wj[m], wj[m]);
                            const float2 e = __fadd2_rn(x2, nsh[m]);
                            s0[m] += wj[m];
                            a2[m] = __ffma2_rn(w2, e, a2[m]);
                            q2[m] = __ffma2_rn(__fmul2_rn(w2, e), e, q2[m]);
                        }
                    }
                }
                j0 = j1;
            }
        }
        __syncwarp();                                       // every lane is done with xs[stg] and rec

        // ---- refill this stage with the chunk nst rounds ahead
        if (bulk && lane == 0) {
            const int64_t cn = c + nst;
            if (cn < c_end && cn < nfull) {
                bar_expect_tx(&bars[stg], chunk_bytes + 128u);
                bulk_load(xs, X + cn * 32 * D, chunk_bytes, &bars[stg]);
                bulk_load(lab0 + stg * 32, state + cn * 32, 128u, &bars[stg]);
            }
        }
    }
    flush();
    if (masks != nullptr && lane == 0) masks[wg] = touched;
}

// out[G*M rows] (+)= sum over the slices [G*k rows] in double, fixed order: 32 lanes of slices per element, slice p
// goes to lane p % 32, the 32 partial sums are added in lane order.  Rows of the untrained components (m >= k) get zero.
__global__ void __launch_bounds__(1024)
reduce_slices32_kernel(const float *__restrict__ partial, const unsigned long long *__restrict__ masks, int64_t nparts,
                       int M, int k, int row, int64_t len_out, double *__restrict__ out, int accumulate) {
    __shared__ double part[32][33];
    const int e = threadIdx.x & 31, j = threadIdx.x >> 5;
    const int64_t i = (int64_t)blockIdx.x * 32 + e;
    double s = 0.0;
    bool live = false;
    if (i < len_out) {
        const int64_t g = i / ((int64_t)M * row);
        const int r = (int)(i - g * M * row), m = r / row, col = r - m * row;
        live = m < k;
        if (live) {
            const int64_t len = len_out / M * k, src = (g * k + m) * row + col;
            int64_t p = j;
            if (masks != nullptr) {                         // only the slices whose warp met pool g; 8 loads in flight
                for (; p < nparts; p += 256) {
                    float v[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        const int64_t q = p + 32 * u;
                        v[u] = (q < nparts && ((masks[q] >> g) & 1ull)) ? partial[q * len + src] : 0.f;
                    }
#pragma unroll
                    for (int u = 0; u < 8; ++u) s += (double)v[u];
                }
                p = nparts;
            }
            for (; p + 96 < nparts; p += 128) {
                const float v0 = partial[p * len + src], v1 = partial[(p + 32) * len + src],
                            v2 = partial[(p + 64) * len + src], v3 = partial[(p + 96) * len + src];
                s += (double)v0;
                s += (double)v1;
                s += (double)v2;
                s += (double)v3;
            }
            for (; p < nparts; p += 32) s += (double)partial[p * len + src];
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

size_t smem_bytes(int G, int k, int D, int mode, int nst = 1) { return pool_layout(G, k, D, mode, nst).bytes; }

int pow2_at_least(int k) { return k <= 1 ? 1 : k <= 2 ? 2 : k <= 4 ? 4 : 8; }

template <int KC, int MODE>
int launch(const float *X, int64_t N, int D, const int32_t *state, int G, int M, const double *center, const double *var,
           const double *lc, const double *shift, float *partial, unsigned long long *masks, int32_t *argmin_out, int grid,
           cudaStream_t st) {
    const int nst = smem_bytes(G, KC, D, MODE, 2) <= (size_t)MAX_SMEM ? 2 : 1;
    const size_t smem = smem_bytes(G, KC, D, MODE, nst);
    const int bulk = (D & 1) && ((reinterpret_cast<uintptr_t>(X) | reinterpret_cast<uintptr_t>(state)) & 15) == 0;   // dense odd rows, 16-byte aligned sources
    auto kfn = pool_stats_kernel<KC, MODE>;
    SRB_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kfn<<<grid, PW * 32, smem, st>>>(X, N, D, state, G, M, center, var, lc, shift, partial, masks, argmin_out, nst, bulk);
    return SRB_OK;
}

}  // namespace
}  // namespace srb

using namespace srb;

extern "C" {

int srb_pool_stats_parts(void) { return sm_count() * PW; }   // private slices of G*k rows

int64_t srb_pool_stats_partial_floats(int G, int k, int D) {    // slices + one 64-bit pool mask per slice
    const int64_t parts = srb_pool_stats_parts();
    return ((parts * G * k * (2 * D + 3) + 1) & ~(int64_t)1) + 2 * parts;
}

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
    const int row = 2 * D + 3;
    const int64_t len_out = (int64_t)G * M * row, len = (int64_t)G * k * row, parts = (int64_t)grid * PW;
    cudaStream_t st = (cudaStream_t)stream;
    // G <= 64: write-once slices with one 64-bit pool mask per warp behind them (no memset of the slices, the reduction
    // skips what a warp never met); else zeroed slices
    unsigned long long *masks = G <= 64 ? reinterpret_cast<unsigned long long *>(partial + ((len * parts + 1) & ~(int64_t)1))
                                        : nullptr;
    if (masks) SRB_CUDA(cudaMemsetAsync(masks, 0, sizeof(unsigned long long) * parts, st));
    else SRB_CUDA(cudaMemsetAsync(partial, 0, sizeof(float) * len * parts, st));
    if (N > 0) {
        int rc = SRB_OK;
#define SRB_POOL_CASE(KC)                                                                                              \
    case KC:                                                                                                           \
        rc = mode == 0 ? launch<KC, 0>(X, N, D, state, G, M, center, var, lc, shift, partial, masks, argmin_out, grid, st) \
                       : launch<KC, 1>(X, N, D, state, G, M, center, var, lc, shift, partial, masks, argmin_out, grid, st); \
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
    reduce_slices32_kernel<<<(int)((len_out + 31) / 32), 1024, 0, st>>>(partial, masks, parts, M, k, row, len_out, out, accumulate);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // extern "C"
