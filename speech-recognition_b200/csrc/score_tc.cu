// score_tc.cu -- fused isolated-word classifier with the GMM contraction on the 5th-generation tensor cores.
//
// Same contract as classify_kernel (score.cu; core.py:79-87 over hmm.py:133-135 / decode.py:97-124), different engine
// for the [x'^2, x'] . [A, B] contraction of GMM.evaluate (hmm_state.py:36-45,114-120):
//   * a tile is 128 frames = the 128 lanes of tensor memory.  Thread t builds frame t's 2*D features in registers,
//     splits each into a TF32 pair (hi = rna(x), lo = x - hi, exact) and writes both halves straight into TMEM
//     (tcgen05.st, thread = lane): the A operand never touches shared memory;
//   * the model's Gaussians are the N columns (220 -> 224 for the 11-word digit model), split the same way once per
//     model update (srb_pack_model_tc) and resident in shared memory for the whole kernel;
//   * one thread issues K/8 x 3 tcgen05.mma.kind::tf32 (lo*hi + hi*lo + hi*hi: the 3xTF32 scheme of the north star;
//     measured error 4.4e-7 of sum|a*b|, the FP32 FMA chain's is 2.3e-7 -- tools/probe_tc.cu) accumulating the log2
//     densities of all Gaussians of all words for the 128 frames in TMEM;
//   * all 8 warps read the accumulators back (tcgen05.ld, thread = frame), add the constants, take the per-state
//     log-sum-exp and write state costs into a shared 256-frame ring;
//   * the Viterbi recursions of the previous 256 frames (classify_dp.cuh) run while the tensor core works on the
//     current tile; X of the next tile streams in with cp.async at the same time.
// Shared memory: B operand 2 x 70 KB + cost ring 55 KB + frame staging 20 KB; TMEM: 224 accumulator + 160 operand columns.
#include "classify_dp.cuh"
#include "score_core.cuh"
#include "tc05.cuh"

namespace srb {

constexpr int TC_M = 128;              // frames per MMA tile (TMEM lanes)
constexpr int TC_MAIN = 256;           // threads of the scoring warps (features, MMA issue, epilogue)
constexpr int TC_DP = 128;             // threads of the Viterbi warps
constexpr int TC_MMA_WARP = (TC_MAIN + TC_DP) / 32;   // the warp whose lane 0 issues the MMAs
constexpr int TC_NT = TC_MAIN + TC_DP + 32;   // threads per CTA
constexpr int TC_CHUNK = 2 * TC_M;     // frames of the cost ring (two tiles)
constexpr int TC_SUB = 4;              // sub-streams of utterances per CTA (one per Viterbi warp)
constexpr int TC_SF = TC_M / TC_SUB;   // frames of one sub-stream in a tile
constexpr int TC_XSQ = (TC_SF * 39 + 4 + 3) & ~3;   // staging floats per sub-stream (D = 39)
constexpr int TC_FIN = 16;             // utterance ends per sub-stream and tile whose final costs stay in shared memory
constexpr int TC_BAR_FULL = 1, TC_BAR_FREE_A = 3, TC_BAR_FREE_B = 5;   // named barriers (+ ring slot)
constexpr int TC_NMAX = 256;           // Gaussian columns of one MMA
constexpr int TC_COL_A = 256;          // TMEM column of the A operand (accumulators sit in [0, NP))
constexpr int TC_MP = 4;               // Gaussians per state handled by the epilogue
constexpr int TC_TL_TILES = 64, TC_TL_EVENTS = 10;   // srb_classify_tc_timeline buffer: int64[TC_TL_TILES][TC_TL_EVENTS]
constexpr float TC_CMIN = -1e30f;      // stands in for log2(0) inside the MMA

__host__ __device__ constexpr int tc_kp(int D) { return (2 * D + 7) & ~7; }
static int tc_np(int R, int Mp) { return (R * Mp + 15) & ~15; }   // M = 128 needs N % 16 == 0

// Gaussian parameters -> TF32 (hi, lo) pair in the K-major chunk-plane layout of tc05.cuh: tcb[2][KP/4][NP][4].
// Feature k = 2*d is x'^2 (coefficient A), k = 2*d + 1 is x' (coefficient B), k = 2*D is the constant 1 whose
// coefficient is the Gaussian's log2 constant c (-inf of zero-weight padding is clamped to -1e30: 2^-1e30 == 0 all the
// same, and 0 * c stays finite in the low-half product); rows past R*Mp and the remaining k are zero.
__global__ void pack_tc_kernel(const float *__restrict__ ab, const float *__restrict__ c, int R, int Mp, int D, int NP,
                               int KP, uint32_t *__restrict__ tcb) {
    const int total = KP * NP;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int kc = i / (NP * 4), n = (i >> 2) % NP, k = kc * 4 + (i & 3);
        float v = 0.f;
        if (n < R * Mp && k < 2 * D) {
            const int r = n / Mp, m = n % Mp;
            v = ab[(((size_t)r * (Mp >> 2) + (m >> 2)) * D + (k >> 1)) * 8 + 2 * (m & 3) + (k & 1)];
        } else if (n < R * Mp && k == 2 * D) {
            v = fmaxf(c[n], TC_CMIN);
        }
        uint32_t h, l;
        tc::split_tf32(v, h, l);
        tcb[i] = h;
        tcb[total + i] = l;
    }
}

// state costs of 4 states (16 accumulator columns = log2 weighted densities incl. constants) -> erow[s0 .. s0+3]:
// cost = -ln sum_m 2^l_m (hmm_state.py:114-120 in the log domain), straight-line so that the four chains interleave.
// The largest term is 2^0 = 1: a 4-element min/max network separates it, so a state costs 3 EX2 + 1 LG2.
__device__ __forceinline__ void tc_state_costs(const uint32_t (&r)[16], float *erow, int s0, int s_hi) {
    float m[4], x0[4], x1[4], x2[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const float l0 = __uint_as_float(r[4 * q]), l1 = __uint_as_float(r[4 * q + 1]);
        const float l2 = __uint_as_float(r[4 * q + 2]), l3 = __uint_as_float(r[4 * q + 3]);
        const float a = fmaxf(l0, l1), b = fminf(l0, l1), c = fmaxf(l2, l3), d = fminf(l2, l3);
        m[q] = fmaxf(a, c);
        x0[q] = b - m[q];
        x1[q] = d - m[q];
        x2[q] = fminf(a, c) - m[q];
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const float sum = 1.f + ex2(x0[q]) + ex2(x1[q]) + ex2(x2[q]);
        float cost = -LN2_F * (m[q] + lg2(sum));
        cost = m[q] > 0.5f * TC_CMIN ? cost : CUDART_INF_F;   // every Gaussian of the state has zero weight
        if (s0 + q < s_hi) erow[s0 + q] = cost;
    }
}

template <int D>
__global__ void __launch_bounds__(TC_NT, 1)
classify_tc_kernel(const float *__restrict__ X, const int32_t *__restrict__ off, int B, int64_t N,
                   const uint32_t *__restrict__ tcb, const float *__restrict__ shift, int W,
                   int S, int NP, const float *__restrict__ trans, float *__restrict__ final_cost,
                   int32_t *__restrict__ label, float *__restrict__ E_out, int64_t ldE,
                   long long *__restrict__ timeline) {
    constexpr int KP = tc_kp(D), SMAX = SRB_DP_MAX_STATES;
    extern __shared__ __align__(128) float smem[];
    __shared__ uint64_t bar[4];
    __shared__ int sub_u[8];
    __shared__ uint32_t tslot;
    __shared__ int band_flag;
    const int R = W * S, ES = R | 1;
    const int Wc = E_out ? 0 : W;                            // scorer mode (E_out): no Viterbi tables, the "utterances" are chunks
    float *shift_s = smem;                                   // 64
    float *tr_s = shift_s + 64;                              // W*SMAX*SMAX   tr_s[w][i][j] = cost j->i
    float *carry = tr_s + Wc * SMAX * SMAX;                  // TC_SUB*2*W*SMAX
    uint32_t *Bh = reinterpret_cast<uint32_t *>(carry + TC_SUB * 2 * Wc * SMAX);   // KP*NP, 16-byte aligned
    uint32_t *Bl = Bh + KP * NP;
    float *xs = reinterpret_cast<float *>(Bl + KP * NP);     // TC_SUB * TC_XSQ
    float *fin_s = xs + TC_SUB * TC_XSQ;                     // TC_SUB * TC_FIN * W
    float *Es = fin_s + TC_SUB * TC_FIN * Wc;                // TC_CHUNK * ES
    const int tid = threadIdx.x, wq = tid >> 5;
    // diagnostics (srb_classify_tc_timeline): CTA 0 stamps clock64 at its pipeline events of the first TC_TL_TILES tiles
    const bool tl_on = timeline != nullptr && blockIdx.x == 0 && (tid & 31) == 0;
#define TC_STAMP(j, k) do { if (tl_on && (j) < TC_TL_TILES) timeline[(j) * TC_TL_EVENTS + (k)] = clock64(); } while (0)

    // utterance range of this CTA: boundaries at the first utterance starting at or after b*N/grid
    const int G = gridDim.x, b = blockIdx.x;
    const int ub = (b == 0) ? 0 : lower_bound_i32(off, B + 1, (int64_t)b * N / G);
    const int ue = (b == G - 1) ? B : lower_bound_i32(off, B + 1, (int64_t)(b + 1) * N / G);
    if (ub >= ue) return;                                    // uniform: before any TMEM allocation

    if (wq == 0) tc::tmem_alloc<512>(&tslot);
    if (tid == 32) {
        tc::mbar_init(&bar[0], TC_M);                        // XR: group A's threads
        tc::mbar_init(&bar[1], TC_DP);                       // AR: the Viterbi warps' threads
        tc::mbar_init(&bar[2], 1);                           // HB: tcgen05.commit
        tc::mbar_init(&bar[3], TC_MAIN);                     // DF: all scoring threads
        tc::mbar_fence_init();
    }
    if (tid == 0) band_flag = 1;
    if (tid < D) shift_s[tid] = shift[tid];
    __syncthreads();
    for (int i = tid; i < Wc * SMAX * SMAX; i += TC_NT) {
        const int w = i / (SMAX * SMAX), r = (i / SMAX) % SMAX, o = i % SMAX;
        float v = CUDART_INF_F;
        if (r < S && o < S) v = trans[((size_t)w * S + r) * S + o];
        tr_s[i] = v;
        if (v < CUDART_INF_F && (o > r || r - o > 2)) band_flag = 0;   // benign race: all writers store 0
    }
    copy_to_smem(reinterpret_cast<float *>(Bh), reinterpret_cast<const float *>(tcb), (int64_t)2 * KP * NP);
    tc::fence_async_smem();                                  // B was written through the generic proxy
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const bool banded = band_flag != 0;
    const uint32_t tb = tslot;
    const uint32_t lane_base = tb + ((uint32_t)((wq & 3) * 32) << 16);
    const uint32_t idesc = tc::idesc_tf32(TC_M, NP);
    const uint32_t bH = tc::smem_u32(Bh), bL = tc::smem_u32(Bl);

    // The CTA's utterance range is cut into TC_SUB sub-streams of whole utterances (equal frame counts up to one
    // utterance).  Tile j = frames [32j, 32j+32) of EVERY sub-stream (TMEM lanes 32q .. 32q+31 = sub-stream q), so a
    // tile holds TC_SUB independent pieces of utterances and the Viterbi chains are 32 steps long, not 128.
    const int64_t fbeg = off[ub], fend = off[ue];
    if (tid <= TC_SUB) {
        int u = ub;
        if (tid == TC_SUB) u = ue;
        else if (tid > 0) u = ub + lower_bound_i32(off + ub, ue - ub + 1, fbeg + (fend - fbeg) * tid / TC_SUB);
        sub_u[tid] = u;
    }
    __syncthreads();
    int ntile = 0;
#pragma unroll
    for (int q = 0; q < TC_SUB; ++q) ntile = max(ntile, (int)((off[sub_u[q + 1]] - off[sub_u[q]] + TC_SF - 1) / TC_SF));

    // mbarriers, one completion per tile each (parity of tile j = j & 1):
    //   XR   frames of the tile staged (scoring group A -> Viterbi warps)
    //   AR   A operand in TMEM, staging buffer consumed (Viterbi warps -> MMA, group A)
    //   HB   accumulators complete (MMA -> scoring warps; also frees the A operand for the Viterbi warps)
    //   DF   accumulators read back (scoring warps -> MMA)
    uint64_t *XR = &bar[0], *AR = &bar[1], *HB = &bar[2], *DF = &bar[3];

    if (wq == TC_MMA_WARP) {
        // ================= MMA issuer: one thread
        if ((tid & 31) == 0) {
            for (int j = 0; j < ntile; ++j) {
                tc::mbar_wait(AR, j & 1);
                tc::fence_after_sync();
                TC_STAMP(j, 0);
                if (j > 0) {
                    tc::mbar_wait(DF, (j - 1) & 1);          // tile j-1's accumulators have been read back
                    tc::fence_after_sync();
                }
                TC_STAMP(j, 1);
                // one N = NP MMA per K step and product: 122 cycles each at N = 224, while anything narrower than 160
                // columns costs a flat 93 (tools/probe_tc.cu), so the accumulator is not split
#pragma unroll
                for (int kk = 0; kk < KP / 8; ++kk) {
                    const uint64_t dbh = tc::smem_desc(bH + kk * 2 * NP * 16, NP * 16, 128);
                    const uint64_t dbl = tc::smem_desc(bL + kk * 2 * NP * 16, NP * 16, 128);
                    tc::mma_tf32_ts(tb, tb + TC_COL_A + KP + kk * 8, dbh, idesc, kk > 0);
                    tc::mma_tf32_ts(tb, tb + TC_COL_A + kk * 8, dbl, idesc, true);
                    tc::mma_tf32_ts(tb, tb + TC_COL_A + kk * 8, dbh, idesc, true);
                }
                tc::mma_commit(HB);
                TC_STAMP(j, 2);
            }
        }
    } else if (wq >= TC_MAIN / 32) {
        // ================= Viterbi warps: warp q owns sub-stream q = TMEM lanes 32q .. 32q+31.  Per iteration it turns
        //                   tile j's frames into the A operand (as soon as tile j-1's MMAs are done), then advances the
        //                   recursions over tile j-1, whose costs the scoring warps have finished meanwhile.
        const int q = wq - TC_MAIN / 32, lane = tid & 31;
        const int sue = sub_u[q + 1];
        const int64_t sbeg = off[sub_u[q]], send = off[sue];
        int ufirst = sub_u[q], par = 0;
        float *carry_q = carry + (size_t)q * 2 * Wc * SMAX;
        for (int j = 0; j <= ntile; ++j) {
            if (j < ntile) {
                // ---- A operand: thread = frame = TMEM lane
                tc::mbar_wait(XR, j & 1);                    // frames of tile j are staged
                if (j > 0) {
                    tc::mbar_wait(HB, (j - 1) & 1);          // every MMA of tile j-1 is done with the A operand
                    tc::fence_after_sync();
                }
                if (q == 0) TC_STAMP(j, 5);
                const int64_t f0 = sbeg + (int64_t)j * TC_SF;
                const int nf = (int)max((int64_t)0, min((int64_t)TC_SF, send - f0));
                const int lf = lane < nf ? lane : max(nf - 1, 0);   // lanes past the end repeat (discarded)
                const float *row = xs + q * TC_XSQ + (int)((f0 * D) & 3) + lf * D;
#pragma unroll
                for (int k0 = 0; k0 < KP; k0 += 8) {
                    uint32_t h[8], l[8];
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int d = (k0 >> 1) + i;
                        if (d < D) {
                            const float v = nf > 0 ? row[d] - shift_s[d] : 0.f;
                            tc::split_tf32(v * v, h[2 * i], l[2 * i]);
                            tc::split_tf32(v, h[2 * i + 1], l[2 * i + 1]);
                        } else {                             // k = 2*D: the constant feature, then zero padding
                            h[2 * i] = d == D ? 0x3f800000u : 0u;
                            h[2 * i + 1] = l[2 * i] = l[2 * i + 1] = 0u;
                        }
                    }
                    tc::tmem_st8(lane_base + TC_COL_A + k0, h);
                    tc::tmem_st8(lane_base + TC_COL_A + KP + k0, l);
                }
                tc::tmem_st_wait();
                tc::fence_before_sync();
                tc::mbar_arrive(AR);
                if (q == 0) TC_STAMP(j, 6);
            }
            if (j >= 1) {
                const int t = j - 1;
                const int64_t cf0 = sbeg + (int64_t)t * TC_SF;
                named_sync(TC_BAR_FULL + (t & 1), TC_MAIN + TC_DP);   // costs of tile t are in ring slot t & 1
                if (q == 0) TC_STAMP(t, 7);
                if (cf0 < send && E_out) {
                    // scorer mode: the tile's cost rows of this sub-stream go to E[frame, 0..R) (coalesced along a row)
                    const int rows = (int)min((int64_t)TC_SF, send - cf0);
                    const float *src = Es + (size_t)((t & 1) * TC_M + q * TC_SF) * ES;
                    for (int i = lane; i < rows * R; i += 32) {
                        const int r = i / R, c = i - r * R;
                        E_out[(cf0 + r) * ldE + c] = src[r * ES + c];
                    }
                } else if (cf0 < send) {
                    classify_dp_chunk<32, -1>(Es + (size_t)((t & 1) * TC_M + q * TC_SF) * ES, ES, off, sue, ufirst, cf0,
                                              min(send, cf0 + (int64_t)TC_SF), W, S, tr_s, carry_q, par, banded,
                                              final_cost, label, lane, fin_s + (size_t)q * TC_FIN * W, TC_FIN);
                    par ^= 1;
                }
                if (q == 0) TC_STAMP(t, 8);
                __threadfence_block();
                named_arrive(TC_BAR_FREE_A + (t & 1), TC_M + TC_DP);  // slot may be overwritten (tile t + 2), told to
                named_arrive(TC_BAR_FREE_B + (t & 1), TC_M + TC_DP);  // both scoring groups
            }
        }
    } else {
        // ================= scoring warps: epilogue of both column halves, the states of each half split between group A
        //                   (warps 0-3) and group B (warps 4-7) so that every scheduler has two warps on it.  Group A
        //                   also stages the next tile's frames.
        const bool grpA = wq < 4;
        const int sq = wq & (TC_SUB - 1);
        const int64_t my_beg = off[sub_u[sq]], my_end = off[sub_u[sq + 1]];
        // staging: lane i of warp sq copies vectors i + 32k of sub-stream sq.  A tile advances every sub-stream by
        // 32 frames = 312 float4, so the 16-byte phase of a sub-stream's first row never changes.
        const int sti = tid & 31;
        const int64_t st_g0 = my_beg * D, st_ga = st_g0 & ~(int64_t)3, st_end = my_end * D, gend = N * D;
        float *st_dst = xs + sq * TC_XSQ;
        auto stage_tile = [&](int j) {
            const int64_t base = st_ga + (int64_t)j * (TC_SF * D);
            const int64_t lim = min(st_end, st_g0 + (int64_t)(j + 1) * (TC_SF * D));
#pragma unroll
            for (int k = 0; k < (TC_SF * D / 4 + 1 + 31) / 32; ++k) {
                const int i = sti + 32 * k;
                const int64_t g = base + 4 * (int64_t)i;
                if (i <= TC_SF * D / 4 && g < lim) {
                    if (g + 4 <= gend) {
                        cp_async16(st_dst + 4 * i, X + g);
                    } else {   // last vector of the whole array: do not read past X
                        st_dst[4 * i + 0] = g + 0 < gend ? X[g + 0] : 0.f;
                        st_dst[4 * i + 1] = g + 1 < gend ? X[g + 1] : 0.f;
                        st_dst[4 * i + 2] = g + 2 < gend ? X[g + 2] : 0.f;
                        st_dst[4 * i + 3] = 0.f;
                    }
                }
            }
            cp_async_wait_all();
            tc::mbar_arrive(XR);
        };
        const int sp = (((R + 1) >> 1) + 3) & ~3;            // group A takes states [0, sp), group B the rest
        const int s_lo = grpA ? 0 : sp, s_hi = grpA ? min(R, sp) : R;
        float *erow0 = Es + (size_t)(tid & (TC_M - 1)) * ES;
        if (grpA) stage_tile(0);
        for (int j = 0; j < ntile; ++j) {
            if (grpA && j + 1 < ntile) {
                tc::mbar_wait(AR, j & 1);                    // staging buffer consumed
                stage_tile(j + 1);
            }
            float *erow = erow0 + (size_t)(j & 1) * TC_M * ES;
            // ---- state cost = -ln sum_m 2^D  (constants are inside D) for this group's states
            tc::mbar_wait(HB, j & 1);
            tc::fence_after_sync();
            if (wq == 0) TC_STAMP(j, 3);
            if (j >= 2) named_sync((grpA ? TC_BAR_FREE_A : TC_BAR_FREE_B) + (j & 1), TC_M + TC_DP);   // ring slot free
            {
                // eight states (two TMEM loads) per round, reduced in one straight-line block so that the chains
                // interleave; the next round's loads are in flight meanwhile
                uint32_t va[16], vb[16], vc[16], vd[16];
                auto issue2 = [&](int s0, uint32_t (&p)[16], uint32_t (&q)[16]) {
                    if (s0 < s_hi) tc::tmem_ld16_issue(lane_base + s0 * TC_MP, p);
                    if (s0 + 4 < s_hi) tc::tmem_ld16_issue(lane_base + (s0 + 4) * TC_MP, q);
                };
                issue2(s_lo, va, vb);
                for (int s0 = s_lo; s0 < s_hi; s0 += 16) {
                    tc::tmem_ld16_fence(va);
                    tc::tmem_ld16_fence(vb);
                    issue2(s0 + 8, vc, vd);
                    tc_state_costs(va, erow, s0, s_hi);
                    if (s0 + 4 < s_hi) tc_state_costs(vb, erow, s0 + 4, s_hi);
                    if (s0 + 8 < s_hi) {
                        tc::tmem_ld16_fence(vc);
                        tc::tmem_ld16_fence(vd);
                        issue2(s0 + 16, va, vb);
                        tc_state_costs(vc, erow, s0 + 8, s_hi);
                        if (s0 + 12 < s_hi) tc_state_costs(vd, erow, s0 + 12, s_hi);
                    }
                }
                if (wq == 0) TC_STAMP(j, 4);
                if (wq == 4) TC_STAMP(j, 9);
                tc::fence_before_sync();
                tc::mbar_arrive(DF);                         // the accumulators may be overwritten
            }
            __threadfence_block();
            named_arrive(TC_BAR_FULL + (j & 1), TC_MAIN + TC_DP);   // hand tile j's costs to the Viterbi warps
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (wq == 0) tc::tmem_free<512>(tb);
#undef TC_STAMP
}

static size_t classify_tc_smem_bytes(int W, int S, int NP, int D, bool scorer = false) {
    const int R = W * S, KP = tc_kp(D), Wc = scorer ? 0 : W;
    return 4 * (64 + (size_t)Wc * SRB_DP_MAX_STATES * SRB_DP_MAX_STATES + TC_SUB * 2 * (size_t)Wc * SRB_DP_MAX_STATES +
                2 * (size_t)KP * NP + (size_t)TC_SUB * TC_XSQ + (size_t)TC_SUB * TC_FIN * Wc + (size_t)TC_CHUNK * (R | 1));
}

}  // namespace srb

using namespace srb;

extern "C" {

int64_t srb_tc_model_floats(int R, int Mp, int D) {
    if (R <= 0 || Mp <= 0 || (Mp & 3) || D <= 0 || tc_np(R, Mp) > TC_NMAX) return 0;
    return 2 * (int64_t)tc_kp(D) * tc_np(R, Mp);
}

int srb_pack_model_tc(const float *ab, const float *c, int R, int Mp, int D, float *tcb, void *stream) {
    SRB_REQUIRE(ab && c && tcb, SRB_EINVAL, "srb_pack_model_tc: null pointer");
    SRB_REQUIRE(srb_tc_model_floats(R, Mp, D) > 0, SRB_EUNSUPPORTED,
                "srb_pack_model_tc: %d states x %d Gaussians exceed %d tensor-core columns", R, Mp, TC_NMAX);
    SRB_REQUIRE(((uintptr_t)tcb & 15) == 0, SRB_EINVAL, "srb_pack_model_tc: tcb must be 16-byte aligned");
    const int NP = tc_np(R, Mp), KP = tc_kp(D);
    pack_tc_kernel<<<(KP * NP + 255) / 256, 256, 0, (cudaStream_t)stream>>>(ab, c, R, Mp, D, NP, KP,
                                                                              reinterpret_cast<uint32_t *>(tcb));
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int srb_classify_tc_supported(int D, int W, int S, int Mp) {
    if (D != 39 || Mp != TC_MP || W <= 0 || S <= 0 || S > SRB_DP_MAX_STATES) return 0;
    const int NP = tc_np(W * S, Mp);
    if (NP > TC_NMAX) return 0;
    const int cap = smem_optin();
    return cap > 0 && classify_tc_smem_bytes(W, S, NP, D) + 64 <= (size_t)cap;
}

int srb_classify_tc(const float *X, const int32_t *utt_off, int32_t B, int64_t N, int D, const float *tcb,
                    const float *shift, int W, int S, int Mp, const float *trans, float *final_cost, int32_t *label,
                    void *stream) {
    return srb_classify_tc_timeline(X, utt_off, B, N, D, tcb, shift, W, S, Mp, trans, final_cost, label, nullptr, stream);
}

int srb_classify_tc_timeline(const float *X, const int32_t *utt_off, int32_t B, int64_t N, int D, const float *tcb,
                             const float *shift, int W, int S, int Mp, const float *trans, float *final_cost,
                             int32_t *label, int64_t *timeline, void *stream) {
    SRB_REQUIRE(((X && final_cost) || B == 0 || N == 0) && utt_off && tcb && shift && trans, SRB_EINVAL,
                "srb_classify_tc: null pointer");
    SRB_REQUIRE(B >= 0 && N >= 0, SRB_EINVAL, "srb_classify_tc: bad sizes");
    SRB_REQUIRE(srb_classify_tc_supported(D, W, S, Mp), SRB_EUNSUPPORTED,
                "srb_classify_tc: D=%d, %d words x %d states x %d Gaussians not covered by the tensor-core classifier "
                "(D = 39, Mp = 4, at most %d Gaussians); use srb_classify", D, W, S, Mp, TC_NMAX);
    SRB_REQUIRE(((uintptr_t)X & 15) == 0 && ((uintptr_t)tcb & 15) == 0, SRB_EINVAL,
                "srb_classify_tc: X and tcb must be 16-byte aligned");
    if (B == 0 || N == 0) return SRB_OK;
    const int sms = sm_count();
    SRB_REQUIRE(sms > 0, SRB_ENODEVICE, "no CUDA device");
    const int NP = tc_np(W * S, Mp);
    const size_t bytes = classify_tc_smem_bytes(W, S, NP, D);
    const int64_t want = (N + TC_CHUNK - 1) / TC_CHUNK;
    const int grid = (int)(want < sms ? want : sms);
    SRB_CUDA(cudaFuncSetAttribute(classify_tc_kernel<39>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    classify_tc_kernel<39><<<grid, TC_NT, bytes, (cudaStream_t)stream>>>(
        X, utt_off, B, N, reinterpret_cast<const uint32_t *>(tcb), shift, W, S, NP, trans, final_cost, label, nullptr, 0,
        reinterpret_cast<long long *>(timeline));
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int srb_gmm_score_tc_supported(int D, int R, int Mp) {
    if (D != 39 || Mp != TC_MP || R <= 0) return 0;
    const int NP = tc_np(R, Mp);
    if (NP > TC_NMAX) return 0;
    const int cap = smem_optin();
    return cap > 0 && classify_tc_smem_bytes(R, 1, NP, D, true) + 64 <= (size_t)cap;
}

int srb_gmm_score_tc(const float *X, int64_t N, int D, const float *tcb, const float *shift, int R, int Mp,
                     const int32_t *chunk_off, int32_t nchunks, float *E, int64_t ldE, void *stream) {
    SRB_REQUIRE(N >= 0 && nchunks >= 0, SRB_EINVAL, "srb_gmm_score_tc: bad sizes");
    if (N == 0 || nchunks == 0) return SRB_OK;
    SRB_REQUIRE(X && tcb && shift && chunk_off && E, SRB_EINVAL, "srb_gmm_score_tc: null pointer");
    SRB_REQUIRE(ldE >= R, SRB_EINVAL, "srb_gmm_score_tc: ldE < R");
    SRB_REQUIRE(srb_gmm_score_tc_supported(D, R, Mp), SRB_EUNSUPPORTED,
                "srb_gmm_score_tc: D=%d, %d states x %d Gaussians not covered by the tensor-core scorer (D = 39, Mp = 4, "
                "at most 57 states = 228 Gaussians); use srb_gmm_score", D, R, Mp);
    SRB_REQUIRE(((uintptr_t)X & 15) == 0 && ((uintptr_t)tcb & 15) == 0, SRB_EINVAL,
                "srb_gmm_score_tc: X and tcb must be 16-byte aligned");
    const int sms = sm_count();
    SRB_REQUIRE(sms > 0, SRB_ENODEVICE, "no CUDA device");
    const int NP = tc_np(R, Mp);
    const size_t bytes = classify_tc_smem_bytes(R, 1, NP, D, true);
    const int64_t want = (N + TC_CHUNK - 1) / TC_CHUNK;
    const int grid = (int)(want < sms ? want : sms);
    SRB_CUDA(cudaFuncSetAttribute(classify_tc_kernel<39>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    classify_tc_kernel<39><<<grid, TC_NT, bytes, (cudaStream_t)stream>>>(
        X, chunk_off, nchunks, N, reinterpret_cast<const uint32_t *>(tcb), shift, R, 1, NP, nullptr, nullptr, nullptr, E, ldE,
        nullptr);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // extern "C"
