// score.cu -- K1 (batched GMM scoring) and the fused isolated-word classifier (score + Viterbi).
#include "score_core.cuh"
#include "classify_dp.cuh"

namespace srb {

constexpr int SMAX = SRB_DP_MAX_STATES;

// ------------------------------------------------------------------------------------------------ K1
// grid = persistent CTAs over 512-frame tiles; parameters are swept through shared memory in groups of SG
// states (all of them when the model fits, as in the 220-Gaussian digit model).
template <int D>
__global__ void __launch_bounds__(NT, 1)
score_kernel(const float *__restrict__ X, int64_t N, int64_t row0, int64_t nrows, const float *__restrict__ ab, const float *__restrict__ c,
             const float *__restrict__ shift, int R, int Mp, int SG, float *__restrict__ E, int64_t ldE,
             float *__restrict__ best_cost, int32_t *__restrict__ best_state, float *__restrict__ comp) {
    extern __shared__ __align__(16) float smem[];
    float *shift_s = smem;                               // 64
    float *pc = shift_s + 64;                            // SG*Mp
    float *pab = pc + (size_t)SG * Mp;                   // SG*Mp*D*2
    float *xs = pab + (size_t)SG * Mp * D * 2;           // union: x tile / E tile
    const int ES = SG | 1;
    const int nblk = Mp >> 2;
    const int ngroups = (R + SG - 1) / SG;
    const int64_t ntiles = (nrows + TF - 1) / TF;

    if (threadIdx.x < D) shift_s[threadIdx.x] = shift[threadIdx.x];
    bool params_loaded = false;

    for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int64_t f0 = row0 + tile * TF;
        const int nf = (int)min((int64_t)TF, row0 + nrows - f0);
        __syncthreads();
        const int xoff = stage_frames_async<D, NT>(xs, X, N, f0, nf);
        cp_async_wait_all();
        __syncthreads();
        Frames<D> fr;
        load_frames<D>(fr, xs, xoff, nf, shift_s);
        __syncthreads();

        float bc[FR];
        int bs[FR];
#pragma unroll
        for (int k = 0; k < FR; ++k) {
            bc[k] = CUDART_INF_F;
            bs[k] = 0;
        }
        for (int g = 0; g < ngroups; ++g) {
            const int r0 = g * SG, rn = min(SG, R - r0);
            if (ngroups > 1 || !params_loaded) {
                copy_to_smem(pab, ab + (size_t)r0 * Mp * D * 2, (int64_t)rn * Mp * D * 2);
                copy_to_smem(pc, c + (size_t)r0 * Mp, (int64_t)rn * Mp);
                params_loaded = true;
                __syncthreads();
            }
            for (int s = 0; s < rn; ++s) {
                float cost[FR];
                float *crow[FR];
#pragma unroll
                for (int k = 0; k < FR; ++k) {
                    const int lf = threadIdx.x + k * NT;
                    crow[k] = (comp && lf < nf) ? comp + ((size_t)(f0 + lf) * R + (r0 + s)) * Mp : nullptr;
                }
                score_state<D>(fr, pab + (size_t)s * Mp * D * 2, pc + s * Mp, nblk, cost, crow, comp != nullptr);
#pragma unroll
                for (int k = 0; k < FR; ++k) {
                    if (E) xs[(threadIdx.x + k * NT) * ES + s] = cost[k];
                    if (cost[k] < bc[k]) {   // first minimum wins
                        bc[k] = cost[k];
                        bs[k] = r0 + s;
                    }
                }
            }
            __syncthreads();
            if (E) {
                for (int i = threadIdx.x; i < nf * rn; i += NT) {
                    const int f = i / rn, s = i - f * rn;
                    E[(f0 + f) * ldE + r0 + s] = xs[f * ES + s];
                }
            }
            if (ngroups > 1) __syncthreads();
        }
#pragma unroll
        for (int k = 0; k < FR; ++k) {
            const int lf = threadIdx.x + k * NT;
            if (lf < nf) {
                if (best_cost) best_cost[f0 + lf] = bc[k];
                if (best_state) best_state[f0 + lf] = bs[k];
            }
        }
    }
}

static size_t score_smem_bytes(int SG, int Mp, int D) {
    const size_t un = max((size_t)TF * D + 4, (size_t)TF * (SG | 1));
    return 4 * (64 + (size_t)SG * Mp + (size_t)SG * Mp * D * 2 + un);
}

// ------------------------------------------------------------------------------------------------ fused classifier
// One persistent CTA owns a contiguous run of whole utterances (balanced by frames).  Per 512-frame chunk:
// phase 1 scores all W*S states into a shared E tile, phase 2 advances the per-(utterance, word) Viterbi
// recursions over the chunk (thread = (piece, word), states in registers); costs of an utterance that
// straddles a chunk boundary are carried in shared memory.  Only costs[-1,-1] leaves the chip.
template <int D>
__global__ void __launch_bounds__(NT, 1)
classify_kernel(const float *__restrict__ X, const int32_t *__restrict__ off, int B, int64_t N,
                const float *__restrict__ ab, const float *__restrict__ c, const float *__restrict__ shift, int W,
                int S, int Mp, const float *__restrict__ trans, float *__restrict__ final_cost,
                int32_t *__restrict__ label) {
    extern __shared__ __align__(16) float smem[];
    const int R = W * S;
    const int ES = R | 1;
    float *shift_s = smem;                               // 64
    float *tr_s = shift_s + 64;                          // W*SMAX*SMAX  tr_s[w][i][j] = cost j->i
    float *carry = tr_s + W * SMAX * SMAX;               // 2*W*SMAX
    float *pc = carry + 2 * W * SMAX;                    // R*Mp
    float *pab = pc + (((size_t)R * Mp + 3) & ~(size_t)3);   // R*Mp*D*2
    float *xs = pab + (size_t)R * Mp * D * 2;            // union: x tile / E tile
    __shared__ int band_flag;
    const int nblk = Mp >> 2;

    if (threadIdx.x == 0) band_flag = 1;
    if (threadIdx.x < D) shift_s[threadIdx.x] = shift[threadIdx.x];
    __syncthreads();
    for (int i = threadIdx.x; i < W * SMAX * SMAX; i += NT) {
        const int w = i / (SMAX * SMAX), r = (i / SMAX) % SMAX, o = i % SMAX;
        float v = CUDART_INF_F;
        if (r < S && o < S) v = trans[((size_t)w * S + r) * S + o];
        tr_s[i] = v;
        if (v < CUDART_INF_F && (o > r || r - o > 2)) band_flag = 0;   // benign race: all writers store 0
    }
    copy_to_smem(pab, ab, (int64_t)R * Mp * D * 2);
    copy_to_smem(pc, c, (int64_t)R * Mp);
    __syncthreads();
    const bool banded = band_flag != 0;

    // utterance range of this CTA: boundaries at the first utterance starting at or after b*N/grid
    const int G = gridDim.x, b = blockIdx.x;
    const int ub = (b == 0) ? 0 : lower_bound_i32(off, B + 1, (int64_t)b * N / G);
    const int ue = (b == G - 1) ? B : lower_bound_i32(off, B + 1, (int64_t)(b + 1) * N / G);
    if (ub >= ue) return;
    const int64_t fbeg = off[ub], fend = off[ue];
    int ufirst = ub;
    int par = 0;

    for (int64_t cf0 = fbeg; cf0 < fend; cf0 += TF, par ^= 1) {
        const int nf = (int)min((int64_t)TF, fend - cf0);
        const int64_t cf1 = cf0 + nf;
        __syncthreads();
        const int xoff = stage_frames_async<D, NT>(xs, X, N, cf0, nf);
        cp_async_wait_all();
        __syncthreads();
        Frames<D> fr;
        load_frames<D>(fr, xs, xoff, nf, shift_s);
        __syncthreads();
        // ---- phase 1: E tile
        for (int s = 0; s < R; ++s) {
            float cost[FR];
            float *crow[FR] = {nullptr, nullptr};
            score_state<D>(fr, pab + (size_t)s * Mp * D * 2, pc + s * Mp, nblk, cost, crow, false);
#pragma unroll
            for (int k = 0; k < FR; ++k) xs[(threadIdx.x + k * NT) * ES + s] = cost[k];
        }
        __syncthreads();
        // ---- phase 2: Viterbi over the pieces of this chunk (classify_dp.cuh)
        classify_dp_chunk<NT>(xs, ES, off, ue, ufirst, cf0, cf1, W, S, tr_s, carry, par, banded, final_cost, label,
                              threadIdx.x);
    }
}

static size_t classify_smem_bytes(int W, int S, int Mp, int D) {
    const int R = W * S;
    const size_t un = max((size_t)TF * D + 4, (size_t)TF * (R | 1));
    return 4 * (64 + (size_t)W * SMAX * SMAX + 2 * (size_t)W * SMAX + (((size_t)R * Mp + 3) & ~(size_t)3) +
                (size_t)R * Mp * D * 2 + un);
}

template <typename K>
static int set_smem(K kernel, size_t bytes) {
    return check_cuda(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes),
                      "cudaFuncSetAttribute(MaxDynamicSharedMemorySize)");
}

// E[n, r] = -log sum_j exp(-P[n, r*J + j]): the cost of a state whose Gaussians were scored as J sub-states of 4 (the
// tensor-core scorer handles 4 Gaussians per column group; a 16-Gaussian state is 4 of them).  One thread per (n, r).
__global__ void neg_lse_groups_kernel(const float *__restrict__ P, int64_t N, int R, int J, float *__restrict__ E,
                                      int64_t ldE) {
    const int64_t total = N * R;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t n = i / R;
        const int r = (int)(i - n * R);
        const float *p = P + i * J;
        float m = CUDART_INF_F;
        for (int j = 0; j < J; ++j) m = fminf(m, p[j]);
        float v = m;
        if (m < CUDART_INF_F) {
            float sum = 0.f;
            for (int j = 0; j < J; ++j) sum += __expf(m - p[j]);
            v = m - __logf(sum);
        }
        E[n * ldE + r] = v;
    }
}

// Running best state over blocks of states scored as sub-states: P[n, s*J + j] are the sub-state costs of states
// state0 .. state0+Rb-1; the state cost is merged as in neg_lse_groups_kernel and replaces best_cost[n] / best_state[n]
// when strictly smaller (blocks arrive in ascending state order: the first minimum wins, as in score_kernel).
template <bool VEC4>   // VEC4: J % 4 == 0 and P 16-byte aligned -- whole 16-byte loads, every fetched sector used by one thread
__global__ void merge_best_kernel(const float *__restrict__ P, int64_t N, int Rb, int J, int state0, int first,
                                  float *__restrict__ best_cost, int32_t *__restrict__ best_state) {
    for (int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; n < N; n += (int64_t)gridDim.x * blockDim.x) {
        const float *p = P + n * Rb * J;
        float bc = first ? CUDART_INF_F : best_cost[n];
        int bs = first ? 0 : best_state[n];
        for (int r = 0; r < Rb; ++r, p += J) {
            float m = CUDART_INF_F, sum = 0.f;
            if (VEC4) {
                const float4 *q = reinterpret_cast<const float4 *>(p);
                float4 v[16];                                            // J <= 64
#pragma unroll
                for (int j = 0; j < 16; ++j)
                    if (4 * j < J) {
                        v[j] = q[j];
                        m = fminf(m, fminf(fminf(v[j].x, v[j].y), fminf(v[j].z, v[j].w)));
                    }
                if (!(m < bc + 0.75f * (float)J)) continue;              // -log sum >= m - log J: cannot win
#pragma unroll
                for (int j = 0; j < 16; ++j)
                    if (4 * j < J) sum += __expf(m - v[j].x) + __expf(m - v[j].y) + __expf(m - v[j].z) + __expf(m - v[j].w);
            } else {
                for (int j = 0; j < J; ++j) m = fminf(m, p[j]);
                if (!(m < bc + 0.75f * (float)J)) continue;
                for (int j = 0; j < J; ++j) sum += __expf(m - p[j]);
            }
            const float c = m - __logf(sum);
            if (c < bc) {
                bc = c;
                bs = state0 + r;
            }
        }
        best_cost[n] = bc;
        best_state[n] = bs;
    }
}

}  // namespace srb

using namespace srb;

#define SRB_DISPATCH_D(D, ...)                                        \
    switch (D) {                                                      \
        case 13: { constexpr int kD = 13; __VA_ARGS__; } break;       \
        case 26: { constexpr int kD = 26; __VA_ARGS__; } break;       \
        case 39: { constexpr int kD = 39; __VA_ARGS__; } break;       \
        default:                                                      \
            set_error("feature dimension %d not built (13, 26, 39)", D); \
            return SRB_EUNSUPPORTED;                                  \
    }

extern "C" {

int srb_gmm_score_rows(const float *X, int64_t N, int64_t row0, int64_t nrows, int D, const float *ab, const float *c,
                       const float *shift, int R, int Mp, float *E, int64_t ldE, float *best_cost, int32_t *best_state,
                       float *comp, void *stream) {
    SRB_REQUIRE((X || nrows == 0) && ab && c && shift, SRB_EINVAL, "srb_gmm_score: null pointer");
    SRB_REQUIRE(N >= 0 && R > 0 && Mp > 0 && (Mp & 3) == 0, SRB_EINVAL, "srb_gmm_score: bad N/R/Mp");
    SRB_REQUIRE(row0 >= 0 && nrows >= 0 && row0 + nrows <= N, SRB_EINVAL, "srb_gmm_score: rows outside X");
    SRB_REQUIRE(!E || ldE >= R, SRB_EINVAL, "srb_gmm_score: ldE < R");
    SRB_REQUIRE(((uintptr_t)X & 15) == 0 && ((uintptr_t)ab & 15) == 0 && ((uintptr_t)c & 15) == 0, SRB_EINVAL,
                "srb_gmm_score: X, ab, c must be 16-byte aligned");
    if (nrows == 0) return SRB_OK;
    const int sms = sm_count(), cap = smem_optin();
    SRB_REQUIRE(sms > 0, SRB_ENODEVICE, "no CUDA device");
    int SG = R;
    while (SG > 1 && score_smem_bytes(SG, Mp, D) > (size_t)cap) --SG;
    SRB_REQUIRE(score_smem_bytes(SG, Mp, D) <= (size_t)cap, SRB_EUNSUPPORTED,
                "srb_gmm_score: one state of %d Gaussians x %d dims does not fit shared memory", Mp, D);
    const size_t bytes = score_smem_bytes(SG, Mp, D);
    const int64_t ntiles = (nrows + TF - 1) / TF;
    const int grid = (int)(ntiles < sms ? ntiles : sms);
    SRB_DISPATCH_D(D, {
        int rc = set_smem(score_kernel<kD>, bytes);
        if (rc) return rc;
        score_kernel<kD><<<grid, NT, bytes, (cudaStream_t)stream>>>(X, N, row0, nrows, ab, c, shift, R, Mp, SG, E, ldE,
                                                                    best_cost, best_state, comp);
    });
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int srb_gmm_score(const float *X, int64_t N, int D, const float *ab, const float *c, const float *shift, int R,
                  int Mp, float *E, int64_t ldE, float *best_cost, int32_t *best_state, float *comp, void *stream) {
    return srb_gmm_score_rows(X, N, 0, N, D, ab, c, shift, R, Mp, E, ldE, best_cost, best_state, comp, stream);
}

int srb_classify(const float *X, const int32_t *utt_off, int32_t B, int64_t N, int D, const float *ab,
                 const float *c, const float *shift, int W, int S, int Mp, const float *trans, float *final_cost,
                 int32_t *label, void *stream) {
    SRB_REQUIRE(((X && final_cost) || B == 0 || N == 0) && utt_off && ab && c && shift && trans, SRB_EINVAL,
                "srb_classify: null pointer");
    SRB_REQUIRE(B >= 0 && N >= 0 && W > 0 && S > 0 && Mp > 0 && (Mp & 3) == 0, SRB_EINVAL, "srb_classify: bad sizes");
    SRB_REQUIRE(S <= SMAX, SRB_EUNSUPPORTED, "srb_classify: S=%d > %d states per word", S, SMAX);
    SRB_REQUIRE(((uintptr_t)X & 15) == 0 && ((uintptr_t)ab & 15) == 0 && ((uintptr_t)c & 15) == 0, SRB_EINVAL,
                "srb_classify: X, ab, c must be 16-byte aligned");
    if (B == 0 || N == 0) return SRB_OK;
    const int sms = sm_count(), cap = smem_optin();
    SRB_REQUIRE(sms > 0, SRB_ENODEVICE, "no CUDA device");
    const size_t bytes = classify_smem_bytes(W, S, Mp, D);
    SRB_REQUIRE(bytes <= (size_t)cap, SRB_EUNSUPPORTED,
                "srb_classify: model (%d words x %d states x %d Gaussians) needs %zu B shared memory > %d", W, S, Mp,
                bytes, cap);
    int64_t want = (N + TF - 1) / TF;
    const int grid = (int)(want < sms ? want : sms);
    SRB_DISPATCH_D(D, {
        int rc = set_smem(classify_kernel<kD>, bytes);
        if (rc) return rc;
        classify_kernel<kD><<<grid, NT, bytes, (cudaStream_t)stream>>>(X, utt_off, B, N, ab, c, shift, W, S, Mp, trans,
                                                                       final_cost, label);
    });
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int srb_neg_lse_groups(const float *P, int64_t N, int R, int J, float *E, int64_t ldE, void *stream) {
    SRB_REQUIRE(N >= 0 && R > 0 && J > 0 && ldE >= R, SRB_EINVAL, "srb_neg_lse_groups: bad sizes");
    if (N == 0) return SRB_OK;
    SRB_REQUIRE(P && E, SRB_EINVAL, "srb_neg_lse_groups: null pointer");
    int64_t grid = (N * R + 255) / 256;
    const int64_t cap = (int64_t)sm_count() * 32;
    if (grid > cap) grid = cap;
    neg_lse_groups_kernel<<<(int)grid, 256, 0, (cudaStream_t)stream>>>(P, N, R, J, E, ldE);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int srb_merge_best(const float *P, int64_t N, int Rb, int J, int state0, int first, float *best_cost,
                   int32_t *best_state, void *stream) {
    SRB_REQUIRE(N >= 0 && Rb > 0 && J > 0 && J <= 64 && state0 >= 0, SRB_EINVAL, "srb_merge_best: bad sizes");
    if (N == 0) return SRB_OK;
    SRB_REQUIRE(P && best_cost && best_state, SRB_EINVAL, "srb_merge_best: null pointer");
    int64_t grid = (N + 127) / 128;
    const int64_t cap = (int64_t)sm_count() * 32;
    if (grid > cap) grid = cap;
    if (J % 4 == 0 && ((uintptr_t)P & 15) == 0)
        merge_best_kernel<true><<<(int)grid, 128, 0, (cudaStream_t)stream>>>(P, N, Rb, J, state0, first, best_cost,
                                                                             best_state);
    else
        merge_best_kernel<false><<<(int)grid, 128, 0, (cudaStream_t)stream>>>(P, N, Rb, J, state0, first, best_cost,
                                                                              best_state);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // extern "C"
