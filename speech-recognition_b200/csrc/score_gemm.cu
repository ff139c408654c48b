// score_gemm.cu -- GMM.evaluate (hmm_state.py:114-120) for models of ANY number of states as ONE persistent GEMM on the
// 5th-generation tensor cores, reduced on the fly to the best state per frame (BASELINE config 5: 2 000 states x 32
// Gaussians x 39 dims on 10 M frames; the 8 kB/frame cost matrix is never materialised).
//
//   log2(w_g N_g(x)) = [x'^2, x', 1] . [A_g, B_g, c_g]        (north star; 3xTF32: lo*hi + hi*lo + hi*hi, FP32 accumulate)
//
// Loop order (the round-1 path ran one launch per block of states, re-read X 286 times and round-tripped 1.9 GB of
// sub-state costs per pass through HBM):
//   * a CTA owns a tile of 128 frames = the 128 lanes of tensor memory.  The tile's split A operand (2 x 80 columns) is
//     written to TMEM ONCE (tcgen05.st, thread = frame) and stays there while ALL Gaussians stream past it;
//   * the model is pre-split and pre-blocked (srb_pack_model_blocks): block b = NB = 160 Gaussian columns (SB whole
//     states) in the K-major chunk-plane layout of tc05.cuh, 100 KB per block.  One producer thread streams the blocks
//     through a 2-stage shared-memory ring with the TMA engine's bulk copy (cp.async.bulk + mbarrier complete_tx); the
//     41 MB model stays L2 resident, HBM sees X once (156 B/frame) and 8 B/frame of results;
//   * one thread issues 30 tcgen05.mma.kind::tf32 (M 128 x N 160 x K 8) per block into one of TWO accumulators in TMEM
//     (2 x 160 + 160 operand columns = 480 of 512), committing to mbarriers: the accumulator's "full" barrier and the
//     ring stage's "empty" barrier;
//   * 8 epilogue warps (2 per TMEM lane quarter) read the other accumulator (tcgen05.ld, thread = frame), fold each
//     state's Mp columns into its log-sum-exp and keep the running (best, state) in registers across all blocks.
//     Only best_cost / best_state leave the SM.
#include "score_core.cuh"
#include "tc05.cuh"

namespace srb {

constexpr int GB_M = 128;                 // frames per tile (TMEM lanes)
constexpr int GB_NB = 160;                // Gaussian columns per block (N of one MMA): 93 cycles per MMA for any N <= 160
constexpr int GB_EPI_WARPS = 8;
constexpr int GB_NT = 64 + GB_EPI_WARPS * 32;   // warp 0: TMA producer, warp 1: MMA issuer, warps 2..9: epilogue
constexpr int GB_STAGES = 2;
constexpr int GB_COL_A = 2 * GB_NB;       // TMEM: accumulators [0,160) and [160,320), A operand hi [320,400), lo [400,480)
constexpr float GB_CMIN = -1e30f;

__host__ __device__ constexpr int gb_kp(int D) { return (2 * D + 7) & ~7; }
__device__ __forceinline__ void gb_bar_sync(int id, int count) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory"); }

__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(tc::smem_u32(bar)), "r"(bytes) : "memory");
}
// TMA engine, bulk (1-D) form: global -> shared, completion counted in bytes on the mbarrier
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     tc::smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(tc::smem_u32(bar))
                 : "memory");
}

// all blocks of the model in one launch: out[b][2][KP/4][NB][4], block b = states [b*SB, b*SB+SB); columns past the
// model's last Gaussian are zero (the epilogue never reads them)
__global__ void pack_blocks_kernel(const float *__restrict__ ab, const float *__restrict__ c, int R, int Mp, int D, int SB,
                                   int KP, uint32_t *__restrict__ out) {
    const int b = blockIdx.y;
    const int r0 = b * SB, nst = min(SB, R - r0);
    const int total = KP * GB_NB;
    uint32_t *dst = out + (size_t)b * 2 * total;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int kc = i / (GB_NB * 4), n = (i >> 2) % GB_NB, k = kc * 4 + (i & 3);
        float v = 0.f;
        if (n < nst * Mp) {
            const int r = r0 + n / Mp, m = n % Mp;
            if (k < 2 * D) v = ab[(((size_t)r * (Mp >> 2) + (m >> 2)) * D + (k >> 1)) * 8 + 2 * (m & 3) + (k & 1)];
            else if (k == 2 * D) v = fmaxf(c[(size_t)r * Mp + m], GB_CMIN);
        }
        uint32_t h, l;
        tc::split_tf32(v, h, l);
        dst[i] = h;
        dst[total + i] = l;
    }
}

// 16 accumulator columns of one state -> running (max, sum of 2^(v - max)) of the state's log-sum-exp
__device__ __forceinline__ void gb_fold16(const uint32_t (&r)[16], float &M, float &S) {
    float m = __uint_as_float(r[0]);
#pragma unroll
    for (int i = 1; i < 16; ++i) m = fmaxf(m, __uint_as_float(r[i]));
    const float Mn = fmaxf(M, m);
    float s0 = 0.f, s1 = 0.f;
#pragma unroll
    for (int i = 0; i < 16; i += 2) {
        s0 += ex2(__uint_as_float(r[i]) - Mn);
        s1 += ex2(__uint_as_float(r[i + 1]) - Mn);
    }
    S = S * ex2(M - Mn) + (s0 + s1);
    M = Mn;
}

template <int D>
__global__ void __launch_bounds__(GB_NT, 1)
gmm_best_gemm_kernel(const float *__restrict__ X, int64_t N, const uint32_t *__restrict__ blocks,
                     const float *__restrict__ shift, int R, int Mp, int SB, int nblocks,
                     float *__restrict__ best_cost, int32_t *__restrict__ best_state,
                     const int4 *__restrict__ tile_desc, int64_t ntiles_rows, float *__restrict__ E_out, int64_t ldE) {
    // Two modes.  Sweep (tile_desc == NULL): tile t = frames [128 t, 128 t + 128), ALL blocks of the model stream past it,
    // the running best state is kept.  Rows (tile_desc != NULL): tile t = {first frame, frames <= 128, block, -} scores
    // its frames against ONE block (the states of the tile's word) and writes their costs to E_out[frame, state].
    const bool rows = tile_desc != nullptr;
    constexpr int KP = gb_kp(D);
    constexpr uint32_t BLOCK_BYTES = 2u * KP * GB_NB * 4u;
    extern __shared__ __align__(128) uint8_t gsm[];
    __shared__ uint64_t bar_full[GB_STAGES], bar_empty[GB_STAGES], bar_acc_full[2], bar_acc_free[2], bar_a_ready;
    __shared__ uint32_t tslot;
    __shared__ float shift_s[64];
    __shared__ float comb_v[GB_M];
    __shared__ int comb_s[GB_M];
    const int tid = threadIdx.x, wq = tid >> 5, lane = tid & 31;
    const int64_t ntiles = rows ? ntiles_rows : (N + GB_M - 1) / GB_M;
    if ((int64_t)blockIdx.x >= ntiles) return;                // uniform: before any allocation
    const int nb_t = rows ? 1 : nblocks;                      // blocks per tile

    if (wq == 0) tc::tmem_alloc<512>(&tslot);
    if (tid == 32) {
        for (int s = 0; s < GB_STAGES; ++s) {
            tc::mbar_init(&bar_full[s], 1);                   // the producer's arrive.expect_tx (+ the bytes)
            tc::mbar_init(&bar_empty[s], 1);                  // tcgen05.commit of the block that read the stage
        }
        for (int b = 0; b < 2; ++b) {
            tc::mbar_init(&bar_acc_full[b], 1);               // tcgen05.commit
            tc::mbar_init(&bar_acc_free[b], GB_EPI_WARPS);    // one elected lane per epilogue warp
        }
        tc::mbar_init(&bar_a_ready, GB_M);                    // the 128 threads that build the A operand
        tc::mbar_fence_init();
    }
    if (tid < D) shift_s[tid] = shift[tid];
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tb = tslot;

    if (wq == 0) {
        // ================= producer: streams the model's blocks through the ring, tile after tile
        if (lane == 0) {
            uint32_t g = 0;
            for (int64_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
                const int b0 = rows ? __ldg(tile_desc + t).z : 0;
                for (int jb = 0; jb < nb_t; ++jb, ++g) {
                    const int s = g % GB_STAGES;
                    tc::mbar_wait(&bar_empty[s], ((g / GB_STAGES) & 1) ^ 1);
                    mbar_expect_tx(&bar_full[s], BLOCK_BYTES);
                    uint8_t *dst = gsm + (size_t)s * BLOCK_BYTES;
                    const uint8_t *src = reinterpret_cast<const uint8_t *>(blocks) + (size_t)(b0 + jb) * BLOCK_BYTES;
                    bulk_g2s(dst, src, BLOCK_BYTES / 2, &bar_full[s]);
                    bulk_g2s(dst + BLOCK_BYTES / 2, src + BLOCK_BYTES / 2, BLOCK_BYTES / 2, &bar_full[s]);
                }
            }
        }
    } else if (wq == 1) {
        // ================= MMA issuer: one thread
        if (lane == 0) {
            const uint32_t idesc = tc::idesc_tf32(GB_M, GB_NB);
            uint32_t g = 0, lt = 0;
            for (int64_t t = blockIdx.x; t < ntiles; t += gridDim.x, ++lt) {
                tc::mbar_wait(&bar_a_ready, lt & 1);
                tc::fence_after_sync();
                for (int jb = 0; jb < nb_t; ++jb, ++g) {
                    const int s = g % GB_STAGES, b = g & 1;
                    tc::mbar_wait(&bar_acc_free[b], ((g >> 1) & 1) ^ 1);   // the epilogue has drained this accumulator
                    tc::mbar_wait(&bar_full[s], (g / GB_STAGES) & 1);      // the block's bytes have landed
                    tc::fence_after_sync();
                    const uint32_t bH = tc::smem_u32(gsm + (size_t)s * BLOCK_BYTES), bL = bH + BLOCK_BYTES / 2;
                    const uint32_t acc = tb + b * GB_NB;
#pragma unroll
                    for (int kk = 0; kk < KP / 8; ++kk) {
                        const uint64_t dbh = tc::smem_desc(bH + kk * 2 * GB_NB * 16, GB_NB * 16, 128);
                        const uint64_t dbl = tc::smem_desc(bL + kk * 2 * GB_NB * 16, GB_NB * 16, 128);
                        tc::mma_tf32_ts(acc, tb + GB_COL_A + KP + kk * 8, dbh, idesc, kk > 0);   // lo * hi
                        tc::mma_tf32_ts(acc, tb + GB_COL_A + kk * 8, dbl, idesc, true);          // hi * lo
                        tc::mma_tf32_ts(acc, tb + GB_COL_A + kk * 8, dbh, idesc, true);          // hi * hi
                    }
                    tc::mma_commit(&bar_empty[s]);            // the stage may be refilled once these MMAs are done
                    tc::mma_commit(&bar_acc_full[b]);
                }
            }
        }
    } else {
        // ================= epilogue warps.  Warp w may touch TMEM lanes 32*(w%4) .. +31; the two warps of a lane quarter
        //                   take alternate states of a block (and swap from block to block), each keeping its own
        //                   running best; the halves are merged once per tile.
        const int q = wq & 3, half = (wq - 2) >> 2;
        const uint32_t lane_base = tb + ((uint32_t)(q * 32) << 16);
        const int fl = q * 32 + lane;                         // frame of the tile = TMEM lane
        const int C = Mp >> 4;                                // 16-column loads per state
        uint32_t g = 0;
        // The frames of a tile are fetched into registers ONE TILE AHEAD (half 0: thread = frame), so that the A operand
        // of tile t+1 is written to TMEM as soon as the last MMA of tile t has completed -- before this warp starts on
        // tile t's last accumulator -- and the tensor core only idles for the split + tcgen05.st of one tile.
        float xr[D];
        auto tile_of = [&](int64_t tt, int64_t &f0, int &nf, int &blk) {
            if (rows) {
                const int4 d = __ldg(tile_desc + tt);
                f0 = d.x;
                nf = d.y;
                blk = d.z;
            } else {
                f0 = tt * GB_M;
                nf = (int)min((int64_t)GB_M, N - f0);
                blk = 0;
            }
        };
        auto load_x = [&](int64_t tt) {
            int64_t f0;
            int nf, blk;
            tile_of(tt, f0, nf, blk);
            const bool live = fl < nf;
            const float *row = X + (live ? f0 + fl : 0) * D;
#pragma unroll
            for (int d = 0; d < D; ++d) xr[d] = live ? __ldg(row + d) : 0.f;
        };
        auto build_a = [&]() {
#pragma unroll
            for (int k0 = 0; k0 < KP; k0 += 8) {
                uint32_t h[8], l[8];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int d = (k0 >> 1) + i;
                    if (d < D) {
                        const float v = xr[d] - shift_s[d];
                        tc::split_tf32(v * v, h[2 * i], l[2 * i]);
                        tc::split_tf32(v, h[2 * i + 1], l[2 * i + 1]);
                    } else {                                  // k = 2*D: the constant feature, then zero padding
                        h[2 * i] = d == D ? 0x3f800000u : 0u;
                        h[2 * i + 1] = l[2 * i] = l[2 * i + 1] = 0u;
                    }
                }
                tc::tmem_st8(lane_base + GB_COL_A + k0, h);
                tc::tmem_st8(lane_base + GB_COL_A + KP + k0, l);
            }
            tc::tmem_st_wait();
            tc::fence_before_sync();
            tc::mbar_arrive(&bar_a_ready);
        };
        if (half == 0) {
            load_x(blockIdx.x);
            build_a();
            if ((int64_t)blockIdx.x + gridDim.x < ntiles) load_x((int64_t)blockIdx.x + gridDim.x);
        }
        for (int64_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
            int64_t f0;
            int nf, blk0;
            tile_of(t, f0, nf, blk0);
            const int64_t f = f0 + fl;
            const bool live = fl < nf;
            float best = -CUDART_INF_F;                       // largest log2 sum_m w_m N_m(x) so far = smallest cost
            int bstate = 0;
            for (int jb = 0; jb < nb_t; ++jb, ++g) {
                const int b = g & 1;
                const int state0 = (blk0 + jb) * SB;          // first state of the block (global index)
                const int nst = min(SB, R - state0);
                tc::mbar_wait(&bar_acc_full[b], (g >> 1) & 1);
                tc::fence_after_sync();
                if (half == 0 && jb == nb_t - 1 && t + gridDim.x < ntiles) {
                    // every MMA of this tile is done: the operand columns are free for the next tile's frames
                    build_a();
                    if (t + 2 * (int64_t)gridDim.x < ntiles) load_x(t + 2 * (int64_t)gridDim.x);
                }
                const uint32_t acc = lane_base + b * GB_NB;
                // states s = s0, s0 + 2, ...: two 16-column loads in flight
                const int s0 = (half ^ jb) & 1;
                uint32_t va[16], vb[16];
                int s = s0, ch = 0;                           // (state, chunk) of the load in va
                if (s < nst) tc::tmem_ld16_issue(acc + s * Mp, va);
                float M = -3e38f, S = 0.f;
                auto finish_state = [&](int st) {
                    const float lse = M + lg2(S);
                    const bool ok = M > 0.5f * GB_CMIN;       // else: every Gaussian of the state has zero weight
                    if (rows) {
                        if (live) E_out[f * ldE + st] = ok ? -LN2_F * lse : CUDART_INF_F;
                    } else if (ok && lse > best) {            // strict: the first (lowest) state keeps a tie
                        best = lse;
                        bstate = state0 + st;
                    }
                    M = -3e38f;
                    S = 0.f;
                };
                while (s < nst) {
                    // next (state, chunk)
                    int sn = s, chn = ch + 1;
                    if (chn == C) {
                        chn = 0;
                        sn = s + 2;
                    }
                    tc::tmem_ld16_fence(va);
                    if (sn < nst) tc::tmem_ld16_issue(acc + sn * Mp + chn * 16, vb);
                    gb_fold16(va, M, S);
                    if (ch == C - 1) finish_state(s);
                    s = sn;
                    ch = chn;
                    if (s >= nst) break;
                    // the same step with the buffers swapped
                    sn = s;
                    chn = ch + 1;
                    if (chn == C) {
                        chn = 0;
                        sn = s + 2;
                    }
                    tc::tmem_ld16_fence(vb);
                    if (sn < nst) tc::tmem_ld16_issue(acc + sn * Mp + chn * 16, va);
                    gb_fold16(vb, M, S);
                    if (ch == C - 1) finish_state(s);
                    s = sn;
                    ch = chn;
                }
                tc::fence_before_sync();
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(&bar_acc_free[b]);
            }
            if (rows) continue;
            // ---- merge the two halves of the lane quarter (ties: the lower state index, i.e. the first minimum)
            if (half == 1) {
                comb_v[fl] = best;
                comb_s[fl] = bstate;
            }
            gb_bar_sync(1, GB_EPI_WARPS * 32);
            if (half == 0) {
                const float ov = comb_v[fl];
                const int os = comb_s[fl];
                if (ov > best || (ov == best && os < bstate)) {
                    best = ov;
                    bstate = os;
                }
                if (live) {
                    if (best_cost) best_cost[f] = best > -CUDART_INF_F ? -LN2_F * best : CUDART_INF_F;
                    if (best_state) best_state[f] = bstate;
                }
            }
            gb_bar_sync(1, GB_EPI_WARPS * 32);                 // comb_* may be overwritten by the next tile
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (wq == 0) tc::tmem_free<512>(tb);
}

static int gb_states_per_block(int Mp) { return GB_NB / Mp; }

}  // namespace srb

using namespace srb;

extern "C" {

int srb_gmm_best_gemm_supported(int D, int R, int Mp) {
    if (D != 39 || R <= 0 || Mp < 16 || (Mp & 15) || Mp > GB_NB) return 0;
    const int cap = smem_optin();
    return cap > 0 && (size_t)GB_STAGES * 2 * gb_kp(D) * GB_NB * 4 + 2048 <= (size_t)cap;
}

int64_t srb_model_blocks_floats(int R, int Mp, int D) {
    if (!srb_gmm_best_gemm_supported(D, R, Mp)) return 0;
    const int SB = gb_states_per_block(Mp);
    return (int64_t)((R + SB - 1) / SB) * 2 * gb_kp(D) * GB_NB;
}

int64_t srb_model_blocks_floats_sb(int R, int Mp, int D, int SB) {
    if (!srb_gmm_best_gemm_supported(D, R, Mp) || SB < 1 || SB * Mp > GB_NB) return 0;
    return (int64_t)((R + SB - 1) / SB) * 2 * gb_kp(D) * GB_NB;
}

static int pack_blocks(const float *ab, const float *c, int R, int Mp, int D, int SB, float *blocks, void *stream) {
    SRB_REQUIRE(ab && c && blocks, SRB_EINVAL, "srb_pack_model_blocks: null pointer");
    SRB_REQUIRE(srb_gmm_best_gemm_supported(D, R, Mp) && SB >= 1 && SB * Mp <= GB_NB, SRB_EUNSUPPORTED,
                "srb_pack_model_blocks: D=%d, %d Gaussians per state, %d states per block not covered (D = 39, Mp a multiple "
                "of 16, at most %d columns per block)", D, Mp, SB, GB_NB);
    SRB_REQUIRE(((uintptr_t)blocks & 15) == 0, SRB_EINVAL, "srb_pack_model_blocks: blocks must be 16-byte aligned");
    const int nb = (R + SB - 1) / SB, KP = gb_kp(D);
    dim3 grid((KP * GB_NB + 255) / 256, nb);
    pack_blocks_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(ab, c, R, Mp, D, SB, KP, reinterpret_cast<uint32_t *>(blocks));
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int srb_pack_model_blocks(const float *ab, const float *c, int R, int Mp, int D, float *blocks, void *stream) {
    return pack_blocks(ab, c, R, Mp, D, gb_states_per_block(Mp > 0 ? Mp : 16), blocks, stream);
}

int srb_pack_model_blocks_sb(const float *ab, const float *c, int R, int Mp, int D, int SB, float *blocks, void *stream) {
    return pack_blocks(ab, c, R, Mp, D, SB, blocks, stream);
}

static int gemm_launch(const float *X, int64_t N, int D, const float *blocks, const float *shift, int R, int Mp, int SB,
                       float *best_cost, int32_t *best_state, const int32_t *tile_desc, int64_t ntiles_rows, float *E,
                       int64_t ldE, void *stream) {
    const int sms = sm_count();
    SRB_REQUIRE(sms > 0, SRB_ENODEVICE, "no CUDA device");
    const int nb = (R + SB - 1) / SB;
    const size_t bytes = (size_t)GB_STAGES * 2 * gb_kp(D) * GB_NB * 4;
    const int64_t ntiles = tile_desc ? ntiles_rows : (N + GB_M - 1) / GB_M;
    const int grid = (int)(ntiles < sms ? ntiles : sms);
    SRB_CUDA(cudaFuncSetAttribute(gmm_best_gemm_kernel<39>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    gmm_best_gemm_kernel<39><<<grid, GB_NT, bytes, (cudaStream_t)stream>>>(
        X, N, reinterpret_cast<const uint32_t *>(blocks), shift, R, Mp, SB, nb, best_cost, best_state,
        reinterpret_cast<const int4 *>(tile_desc), ntiles_rows, E, ldE);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int srb_gmm_best_gemm(const float *X, int64_t N, int D, const float *blocks, const float *shift, int R, int Mp,
                      float *best_cost, int32_t *best_state, void *stream) {
    SRB_REQUIRE(N >= 0, SRB_EINVAL, "srb_gmm_best_gemm: bad sizes");
    if (N == 0) return SRB_OK;
    SRB_REQUIRE(X && blocks && shift && (best_cost || best_state), SRB_EINVAL, "srb_gmm_best_gemm: null pointer");
    SRB_REQUIRE(srb_gmm_best_gemm_supported(D, R, Mp), SRB_EUNSUPPORTED,
                "srb_gmm_best_gemm: D=%d, %d Gaussians per state not covered (D = 39, Mp a multiple of 16 up to %d); use "
                "srb_gmm_score", D, Mp, GB_NB);
    SRB_REQUIRE(((uintptr_t)blocks & 15) == 0, SRB_EINVAL, "srb_gmm_best_gemm: blocks must be 16-byte aligned");
    return gemm_launch(X, N, D, blocks, shift, R, Mp, gb_states_per_block(Mp), best_cost, best_state, nullptr, 0, nullptr, 0,
                       stream);
}

int srb_gmm_score_gemm(const float *X, int64_t N, int D, const float *blocks, const float *shift, int R, int Mp, int SB,
                       const int32_t *tile_desc, int64_t ntiles, float *E, int64_t ldE, void *stream) {
    SRB_REQUIRE(N >= 0 && ntiles >= 0, SRB_EINVAL, "srb_gmm_score_gemm: bad sizes");
    if (N == 0 || ntiles == 0) return SRB_OK;
    SRB_REQUIRE(X && blocks && shift && tile_desc && E, SRB_EINVAL, "srb_gmm_score_gemm: null pointer");
    SRB_REQUIRE(srb_gmm_best_gemm_supported(D, R, Mp) && SB >= 1 && SB * Mp <= GB_NB && ldE >= SB, SRB_EUNSUPPORTED,
                "srb_gmm_score_gemm: D=%d, %d Gaussians per state, %d states per block not covered", D, Mp, SB);
    SRB_REQUIRE(((uintptr_t)blocks & 15) == 0 && ((uintptr_t)tile_desc & 15) == 0, SRB_EINVAL,
                "srb_gmm_score_gemm: blocks and tile_desc must be 16-byte aligned");
    return gemm_launch(X, N, D, blocks, shift, R, Mp, SB, nullptr, nullptr, tile_desc, ntiles, E, ldE, stream);
}

}  // extern "C"
