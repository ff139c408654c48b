// score_core.cuh -- register-resident diagonal-GMM scorer shared by the score / classify / train kernels.
//
// Replaces GMM.evaluate -> MultivariateNormal.pdf (hmm_state.py:36-45,114-120): one thread owns FR frames,
// keeps the (x'^2, x') pair of every feature dimension in registers (x' = x - shift) and sweeps the packed
// Gaussians from shared memory with broadcast 128-bit loads.  One fma.rn.f32x2 per (frame, Gaussian, dim)
// computes x'^2*A + x'*B in its two halves, i.e. exactly the 2*D FMA per (frame, Gaussian) of SURVEY 8d.
#pragma once
#include "common.cuh"

namespace srb {

constexpr int NT = 256;        // threads per CTA of the scorer
constexpr int FR = 2;          // frames per thread
constexpr int TF = NT * FR;    // frames per tile

template <int D, int FRN = FR>
struct Frames {
    float2 xx[FRN][D];         // (x'^2, x')
};

// Stage rows [f0, f0+nf) of X[N,D] into shared memory with 128-bit loads.  Returns the offset (0..3 floats)
// of row f0 inside xs.  xs must hold TF*D + 4 floats.
template <int D, int NTH = NT>
__device__ __forceinline__ int stage_frames(float *xs, const float *__restrict__ X, int64_t N, int64_t f0, int nf) {
    const int64_t g0 = f0 * D, g1 = (f0 + nf) * D, gend = N * D;
    const int64_t ga = g0 & ~(int64_t)3;
    const int nvec = (int)((g1 - ga + 3) >> 2);
    const float4 *src = reinterpret_cast<const float4 *>(X + ga);
    float4 *dst = reinterpret_cast<float4 *>(xs);
    for (int i = threadIdx.x; i < nvec; i += NTH) {
        const int64_t g = ga + 4 * (int64_t)i;
        float4 v;
        if (g + 4 <= gend) {
            v = __ldg(src + i);
        } else {  // last vector of the whole array: do not read past X
            v.x = g + 0 < gend ? X[g + 0] : 0.f;
            v.y = g + 1 < gend ? X[g + 1] : 0.f;
            v.z = g + 2 < gend ? X[g + 2] : 0.f;
            v.w = 0.f;
        }
        dst[i] = v;
    }
    return (int)(g0 - ga);
}

// 16-byte async copies global -> shared (LDGSTS): all of a tile's loads are in flight at once
__device__ __forceinline__ void cp_async16(void *dst, const void *src) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(src));
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory"); }

template <int D, int NTH>
__device__ __forceinline__ int stage_frames_async(float *xs, const float *__restrict__ X, int64_t N, int64_t f0, int nf) {
    const int64_t g0 = f0 * D, g1 = (f0 + nf) * D, gend = N * D;
    const int64_t ga = g0 & ~(int64_t)3;
    const int nvec = (int)((g1 - ga + 3) >> 2);
    for (int i = threadIdx.x; i < nvec; i += NTH) {
        const int64_t g = ga + 4 * (int64_t)i;
        if (g + 4 <= gend) {
            cp_async16(xs + 4 * i, X + g);
        } else {   // last vector of the whole array: do not read past X
            xs[4 * i + 0] = g + 0 < gend ? X[g + 0] : 0.f;
            xs[4 * i + 1] = g + 1 < gend ? X[g + 1] : 0.f;
            xs[4 * i + 2] = g + 2 < gend ? X[g + 2] : 0.f;
            xs[4 * i + 3] = 0.f;
        }
    }
    return (int)(g0 - ga);
}

// shared tile -> registers.  Frame slot k of this thread is local frame threadIdx.x + k*NT; frames past nf
// read frame 0 (their results are discarded by the caller).
template <int D, int NTH = NT, int FRN = FR>
__device__ __forceinline__ void load_frames(Frames<D, FRN> &fr, const float *xs, int xoff, int nf, const float *shift_s) {
#pragma unroll
    for (int k = 0; k < FRN; ++k) {
        int lf = threadIdx.x + k * NTH;
        if (lf >= nf) lf = 0;
        const float *row = xs + xoff + lf * D;   // stride D words: odd D -> conflict free
#pragma unroll
        for (int d = 0; d < D; ++d) {
            const float v = row[d] - shift_s[d];
            fr.xx[k][d] = make_float2(v * v, v);
        }
    }
}

// log2-domain scores of the 4 Gaussians of one packed block for the thread's FR frames.
// pab: shared, float4[D][2] = (A0,B0,A1,B1),(A2,B2,A3,B3) per dim; pc: shared float[4].
template <int D, int FRN = FR>
__device__ __forceinline__ void score_block4(const Frames<D, FRN> &fr, const float4 *__restrict__ pab,
                                             const float *__restrict__ pc, float (&l)[FRN][4]) {
    float2 acc[FRN][4];
#pragma unroll
    for (int k = 0; k < FRN; ++k)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[k][j] = make_float2(0.f, 0.f);
#pragma unroll
    for (int d = 0; d < D; ++d) {
        const float4 p01 = pab[2 * d], p23 = pab[2 * d + 1];
#pragma unroll
        for (int k = 0; k < FRN; ++k) {
            acc[k][0] = ffma2(fr.xx[k][d], make_float2(p01.x, p01.y), acc[k][0]);
            acc[k][1] = ffma2(fr.xx[k][d], make_float2(p01.z, p01.w), acc[k][1]);
            acc[k][2] = ffma2(fr.xx[k][d], make_float2(p23.x, p23.y), acc[k][2]);
            acc[k][3] = ffma2(fr.xx[k][d], make_float2(p23.z, p23.w), acc[k][3]);
        }
    }
    const float4 cc = *reinterpret_cast<const float4 *>(pc);
#pragma unroll
    for (int k = 0; k < FRN; ++k) {
        l[k][0] = acc[k][0].x + acc[k][0].y + cc.x;
        l[k][1] = acc[k][1].x + acc[k][1].y + cc.y;
        l[k][2] = acc[k][2].x + acc[k][2].y + cc.z;
        l[k][3] = acc[k][3].x + acc[k][3].y + cc.w;
    }
}

// cost (natural-log units) of one state = -ln sum_m 2^l_m over its Mp/4 blocks (online log-sum-exp).
// comp (optional, global) receives ln(w_m N_m(x)) per Gaussian: comp_row[k] points at this state's Mp floats.
template <int D, int FRN = FR>
__device__ __forceinline__ void score_state(const Frames<D, FRN> &fr, const float *__restrict__ pab_state,
                                            const float *__restrict__ pc_state, int nblk, float (&cost)[FRN],
                                            float *const (&comp_row)[FRN], bool want_comp, int nstore = 1 << 30) {
    float mrun[FRN], srun[FRN];
#pragma unroll
    for (int k = 0; k < FRN; ++k) {
        mrun[k] = -CUDART_INF_F;
        srun[k] = 0.f;
    }
    for (int b = 0; b < nblk; ++b) {
        float l[FRN][4];
        score_block4<D, FRN>(fr, reinterpret_cast<const float4 *>(pab_state) + (size_t)b * D * 2, pc_state + 4 * b, l);
#pragma unroll
        for (int k = 0; k < FRN; ++k) {
            if (want_comp && comp_row[k]) {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (4 * b + j < nstore) comp_row[k][4 * b + j] = l[k][j] * LN2_F;
            }
            const float m4 = fmaxf(fmaxf(l[k][0], l[k][1]), fmaxf(l[k][2], l[k][3]));
            const float mn = fmaxf(mrun[k], m4);
            if (mn > -CUDART_INF_F) {
                srun[k] = srun[k] * ex2(mrun[k] - mn) + ex2(l[k][0] - mn) + ex2(l[k][1] - mn) + ex2(l[k][2] - mn) +
                          ex2(l[k][3] - mn);
                mrun[k] = mn;
            }
        }
    }
#pragma unroll
    for (int k = 0; k < FRN; ++k) cost[k] = -LN2_F * (mrun[k] + lg2(srun[k]));   // s == 0 -> +inf like -log(0)
}

// cooperative copy of `nfloats` (multiple of 4, 16-byte aligned both sides) global -> shared
__device__ __forceinline__ void copy_to_smem(float *dst, const float *__restrict__ src, int64_t nfloats) {
    const float4 *s = reinterpret_cast<const float4 *>(src);
    float4 *d = reinterpret_cast<float4 *>(dst);
    for (int64_t i = threadIdx.x; i < (nfloats >> 2); i += blockDim.x) d[i] = __ldg(s + i);
}

}  // namespace srb
