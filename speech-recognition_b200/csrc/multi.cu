// multi.cu -- the per-iteration tail of the fused isolated-word training pass (BASELINE config 2), written for the
// STRONG-scaling case where a GPU's shard is small and the tail IS the step: statistics exchange over NVLink peer memory
// and the replicated M-step, two launches in all.
//
//   srb_em_push     per-CTA fp32 slices of srb_em_isolated -> fixed-order fp64 sums (the reduce_slices step), the summed
//                   Viterbi cost and the utterance count -> written straight into EVERY rank's exchange buffer (own slot
//                   of the epoch's parity half) with peer stores; the last CTA to finish releases one flag per peer.
//   srb_em_update   one CTA per Gaussian: waits for the flags of all ranks, adds the ranks' slots IN RANK ORDER (every rank
//                   forms bit-identical totals, so the replicated models stay identical), applies the M-step of GMM.em
//                   (hmm_state.py:138-148) to its Gaussian and re-packs it for both scorers (srb_pack_model's ab / c and
//                   srb_pack_model_tc's split tensor-core operand).
// This replaces reduce_slices + 2 torch reductions + ncclAllReduce (143 kB: pure latency, 0.22 ms at 8 GPUs in round 1) +
// mstep + pack + pack_tc.  With one rank it is the same two launches without peers.
//
// Exchange buffer of a rank (srb_em_sync_bytes): flags uint64[16] | CTA counter | slots double[2][world][len + 17].
// Flags carry the epoch (monotone), the slot halves alternate with its parity: a rank can only be one epoch ahead of a
// peer that still reads, and then it writes the other half.
#include "common.cuh"
#include "tc05.cuh"

namespace srb {

constexpr int MAX_RANKS = 8;
constexpr int SYNC_HEADER = 256;          // bytes: 16 flags + the CTA counter
constexpr int COST_BLOCKS = 16;           // CTAs of srb_em_push that sum the shard's Viterbi costs (one partial each)
constexpr int SLOT_TAIL = COST_BLOCKS + 1; // doubles after the statistics in a slot: the cost partials and the count
constexpr int PUSH_MAX_PARTS = 1024;      // per-CTA slices srb_em_push can list (srb_em_parts() = 4 per SM)

struct PeerBufs {
    uint8_t *buf[MAX_RANKS];
};

__device__ __forceinline__ double *slots_of(uint8_t *buf) { return reinterpret_cast<double *>(buf + SYNC_HEADER); }
__device__ __forceinline__ unsigned long long *flags_of(uint8_t *buf) { return reinterpret_cast<unsigned long long *>(buf); }

__device__ __forceinline__ void st_release_sys(unsigned long long *p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

__global__ void __launch_bounds__(256)
em_push_kernel(const float *__restrict__ partial, const unsigned long long *__restrict__ mask, int nparts, int64_t len,
               int word_len, int M, int ncomp, int row, const float *__restrict__ best_cost, int B, PeerBufs P, int world,
               int rank, unsigned long long epoch) {
    __shared__ double red[256];
    const int64_t slot_len = len + SLOT_TAIL;
    const int64_t base = ((int64_t)(epoch & 1) * world + rank) * slot_len;
    if (blockIdx.x + COST_BLOCKS < gridDim.x) {
        // A CTA = 32 elements of ONE word x 8 part groups.  Warp 0 lists, in ascending order, the slices whose CTA saw the
        // word (mask bit); group q then adds its eighth of the list in that order (4 loads in flight) and the 8 group sums
        // are added in group order -- a fixed order whatever the timing.  Rows of untrained components were never
        // written by anybody: 0.
        __shared__ int list[PUSH_MAX_PARTS];
        __shared__ int nlist;
        const int e = threadIdx.x & 31, q = threadIdx.x >> 5;
        const int bpw = (word_len + 31) >> 5;
        const int w = blockIdx.x / bpw, j = (blockIdx.x - w * bpw) * 32 + e;
        const int64_t i = (int64_t)w * word_len + j;
        if (q == 0) {
            int n = 0;
            for (int p0 = 0; p0 < nparts; p0 += 32) {
                const int p = p0 + e;
                const bool on = p < nparts && ((mask[p] >> (w & 63)) & 1ull);
                const unsigned b = __ballot_sync(0xffffffffu, on);
                if (on) list[n + __popc(b & ((1u << e) - 1))] = p;
                n += __popc(b);
            }
            if (e == 0) nlist = n;
        }
        __syncthreads();
        const int n = nlist;
        double s = 0.0;
        if (j < word_len && (int)((j / row) % M) < ncomp) {
            int k = (int)((int64_t)n * q / 8);
            const int k1 = (int)((int64_t)n * (q + 1) / 8);
            for (; k + 4 <= k1; k += 4) {
                const float a = partial[(int64_t)list[k] * len + i], b = partial[(int64_t)list[k + 1] * len + i];
                const float c = partial[(int64_t)list[k + 2] * len + i], d = partial[(int64_t)list[k + 3] * len + i];
                s += (double)a;
                s += (double)b;
                s += (double)c;
                s += (double)d;
            }
            for (; k < k1; ++k) s += (double)partial[(int64_t)list[k] * len + i];
        }
        red[threadIdx.x] = s;
        __syncthreads();
        if (q == 0 && j < word_len) {
            double t = red[e];
#pragma unroll
            for (int k = 1; k < 8; ++k) t += red[k * 32 + e];
            for (int r = 0; r < world; ++r) slots_of(P.buf[r])[base + i] = t;
        }
    } else {
        // the last COST_BLOCKS CTAs: Viterbi costs of a sixteenth of the shard's utterances each (fixed-shape reduction);
        // srb_em_update adds the partials in order.  The first of them also writes the utterance count.
        const int cb = blockIdx.x - (gridDim.x - COST_BLOCKS);
        const int u0 = (int)((int64_t)B * cb / COST_BLOCKS), u1 = (int)((int64_t)B * (cb + 1) / COST_BLOCKS);
        double s = 0.0;
        for (int u = u0 + threadIdx.x; u < u1; u += 256) s += (double)best_cost[u];
        red[threadIdx.x] = s;
        __syncthreads();
#pragma unroll
        for (int d = 128; d; d >>= 1) {
            if ((int)threadIdx.x < d) red[threadIdx.x] += red[threadIdx.x + d];
            __syncthreads();
        }
        if (threadIdx.x == 0)
            for (int r = 0; r < world; ++r) {
                slots_of(P.buf[r])[base + len + cb] = red[0];
                if (cb == 0) slots_of(P.buf[r])[base + len + COST_BLOCKS] = (double)B;
            }
    }
    // every store of this CTA is visible system-wide before the CTA counts itself done; the last one raises the flags
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long *counter = flags_of(P.buf[rank]) + 16;
        const unsigned long long done = atomicAdd(counter, 1ull);
        if (done == gridDim.x - 1) {
            *counter = 0;
            __threadfence_system();
            for (int r = 0; r < world; ++r) st_release_sys(flags_of(P.buf[r]) + rank, epoch);
        }
    }
}

template <int D>
__global__ void __launch_bounds__(128)
em_update_kernel(uint8_t *__restrict__ buf, int world, unsigned long long epoch, int64_t len, int R, int M, int Mp,
                 int ncomp, double *__restrict__ mu, double *__restrict__ var, double *__restrict__ w,
                 const double *__restrict__ shift, double var_floor, float *__restrict__ ab, float *__restrict__ c,
                 uint32_t *__restrict__ tcb, int NP, int KP, double *__restrict__ stats_out) {
    constexpr int ROW = 2 * D + 3;
    __shared__ double row_s[ROW + 1];
    __shared__ double red[4];
    __shared__ float col_s[2 * D + 8];
    const int tid = threadIdx.x;
    const int64_t slot_len = len + SLOT_TAIL;
    const double *slots = slots_of(buf) + (int64_t)(epoch & 1) * world * slot_len;
    if (tid < world) {
        const unsigned long long *f = flags_of(buf) + tid;
        while (ld_acquire_sys(f) < epoch) {
        }
    }
    __syncthreads();
    const int g = blockIdx.x;                                 // real Gaussian index, or R*M for the scalar block
    if (g == R * M) {
        if (tid == 0) {                                       // summed Viterbi cost: ranks in order, partials in order
            double s = 0.0;
            for (int r = 0; r < world; ++r)
                for (int k = 0; k < COST_BLOCKS; ++k) s += __ldcg(slots + r * slot_len + len + k);
            stats_out[len] = s;
        } else if (tid == 1) {                                // utterances
            double s = 0.0;
            for (int r = 0; r < world; ++r) s += __ldcg(slots + r * slot_len + len + COST_BLOCKS);
            stats_out[len + 1] = s;
        }
        return;
    }
    const int st = g / M, m = g % M;
    if (tid < ROW) {
        double s = 0.0;
        for (int r = 0; r < world; ++r) s += __ldcg(slots + r * slot_len + (int64_t)g * ROW + tid);   // rank order
        row_s[tid] = s;
        stats_out[(int64_t)g * ROW + tid] = s;
    } else if (tid == ROW) {
        double s = 0.0;
        for (int r = 0; r < world; ++r) s += __ldcg(slots + r * slot_len + (int64_t)st * M * ROW + 2 * D + 1);
        row_s[ROW] = s;                                       // frames aligned to the state (count of its first row)
    }
    __syncthreads();
    // ---- M-step of this Gaussian (hmm_state.py:138-148): mu += S1/S0, var = S2/S0 - (S1/S0)^2, w = S0 / count
    const double S0 = row_s[0];
    double v = 1.0, mm = 0.0;
    if (tid < D) {
        double mu_d = mu[(size_t)g * D + tid];
        v = var[(size_t)g * D + tid];
        if (m < ncomp && S0 > 0.0) {
            const double m1 = row_s[1 + tid] / S0;
            v = row_s[1 + D + tid] / S0 - m1 * m1;
            if (v < var_floor) v = var_floor;
            mu_d += m1;
            mu[(size_t)g * D + tid] = mu_d;
            var[(size_t)g * D + tid] = v;
        }
        mm = mu_d - shift[tid];
    }
    double wg = w[g];
    if (m < ncomp && S0 > 0.0 && row_s[ROW] > 0.0) {
        wg = S0 / row_s[ROW];
        if (tid == 0) w[g] = wg;
    }
    // ---- re-pack: A = -0.5/var*log2e, B = (mu-shift)/var*log2e, C = (log w - 0.5(D log 2pi + sum log var) - 0.5 q)*log2e
    const double L2E = 1.4426950408889634074;
    const int gp = st * Mp + m;
    double part = 0.0;
    if (tid < D) {
        float2 *abp = reinterpret_cast<float2 *>(ab) + ((size_t)(gp >> 2) * D) * 4 + (gp & 3);
        const float a_ = (float)(-0.5 / v * L2E), b_ = (float)(mm / v * L2E);
        abp[(size_t)tid * 4] = make_float2(a_, b_);
        col_s[2 * tid] = a_;
        col_s[2 * tid + 1] = b_;
        part = log(v) + mm * mm / v;
    }
#pragma unroll
    for (int d = 16; d; d >>= 1) part += __shfl_xor_sync(0xffffffffu, part, d);
    if ((tid & 31) == 0) red[tid >> 5] = part;
    __syncthreads();
    if (tid == 0) {
        const double tot = red[0] + red[1] + red[2] + red[3];               // sum log var + q
        const double lw = wg > 0.0 ? log(wg) : -CUDART_INF;
        const float cf = (float)((lw - 0.5 * ((double)D * 1.8378770664093454836 + tot)) * L2E);
        c[gp] = cf;
        col_s[2 * D] = fmaxf(cf, -1e30f);
    }
    for (int k = 2 * D + 1 + tid; k < 2 * D + 8; k += 128) col_s[k] = 0.f;
    __syncthreads();
    // ---- the tensor-core operand's column of this Gaussian (srb_pack_model_tc layout: tcb[2][KP/4][NP][4])
    if (tcb && tid < KP) {
        uint32_t h, l;
        tc::split_tf32(col_s[tid], h, l);
        const size_t i = (size_t)(tid >> 2) * NP * 4 + (size_t)gp * 4 + (tid & 3);
        tcb[i] = h;
        tcb[(size_t)KP * NP + i] = l;
    }
}

}  // namespace srb

using namespace srb;

extern "C" {

int srb_peer_alloc(int64_t bytes, void **ptr) {
    SRB_REQUIRE(ptr && bytes > 0, SRB_EINVAL, "srb_peer_alloc: bad argument");
    SRB_CUDA(cudaMalloc(ptr, (size_t)bytes));
    SRB_CUDA(cudaMemset(*ptr, 0, (size_t)bytes));
    SRB_CUDA(cudaDeviceSynchronize());
    return SRB_OK;
}

int srb_peer_free(void *ptr) {
    if (ptr) SRB_CUDA(cudaFree(ptr));
    return SRB_OK;
}

int srb_peer_export(void *ptr, uint8_t *handle64) {
    SRB_REQUIRE(ptr && handle64, SRB_EINVAL, "srb_peer_export: null pointer");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    cudaIpcMemHandle_t h;
    SRB_CUDA(cudaIpcGetMemHandle(&h, ptr));
    memcpy(handle64, &h, 64);
    return SRB_OK;
}

int srb_peer_open(const uint8_t *handle64, void **ptr) {
    SRB_REQUIRE(ptr && handle64, SRB_EINVAL, "srb_peer_open: null pointer");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    SRB_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return SRB_OK;
}

int srb_peer_close(void *ptr) {
    if (ptr) SRB_CUDA(cudaIpcCloseMemHandle(ptr));
    return SRB_OK;
}

int64_t srb_em_sync_bytes(int world, int64_t len) {
    if (world < 1 || world > MAX_RANKS || len < 0) return 0;
    return SYNC_HEADER + (int64_t)sizeof(double) * 2 * world * (len + SLOT_TAIL);
}

int srb_em_push(const float *partial, int32_t nparts, int W, int S, int M, int ncomp, int D, const float *best_cost,
                int32_t B, void *const *peer_bufs, int world, int rank, uint64_t epoch, void *stream) {
    const int row = 2 * D + 3;
    const int64_t len = (int64_t)W * S * M * row;
    SRB_REQUIRE(partial && peer_bufs && (best_cost || B == 0), SRB_EINVAL, "srb_em_push: null pointer");
    SRB_REQUIRE(world >= 1 && world <= MAX_RANKS && rank >= 0 && rank < world && len > 0 && nparts > 0 && epoch > 0,
                SRB_EINVAL, "srb_em_push: bad sizes (at most %d ranks)", MAX_RANKS);
    PeerBufs P;
    for (int r = 0; r < MAX_RANKS; ++r) P.buf[r] = r < world ? static_cast<uint8_t *>(peer_bufs[r]) : nullptr;
    for (int r = 0; r < world; ++r) SRB_REQUIRE(P.buf[r], SRB_EINVAL, "srb_em_push: peer buffer %d is null", r);
    SRB_REQUIRE(nparts <= PUSH_MAX_PARTS, SRB_EUNSUPPORTED, "srb_em_push: more than %d slices", PUSH_MAX_PARTS);
    const int grid = W * ((S * M * row + 31) / 32) + COST_BLOCKS;
    const unsigned long long *mask = reinterpret_cast<const unsigned long long *>(partial + (size_t)nparts * len);
    em_push_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(partial, mask, nparts, len, S * M * row, M, ncomp, row, best_cost,
                                                            B, P, world, rank, (unsigned long long)epoch);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

int srb_em_update(void *own_buf, int world, uint64_t epoch, int R, int M, int Mp, int ncomp, int D, double *mu, double *var,
                  double *w, const double *shift, double var_floor, float *ab, float *c, float *tcb, double *stats_out,
                  void *stream) {
    SRB_REQUIRE(own_buf && mu && var && w && shift && ab && c && stats_out, SRB_EINVAL, "srb_em_update: null pointer");
    SRB_REQUIRE(world >= 1 && world <= MAX_RANKS && epoch > 0 && R > 0 && M > 0 && Mp >= M && ncomp >= 0 && ncomp <= M,
                SRB_EINVAL, "srb_em_update: bad sizes");
    SRB_REQUIRE(D == 39, SRB_EUNSUPPORTED, "srb_em_update: feature dimension %d not built (39)", D);
    const int64_t len = (int64_t)R * M * (2 * D + 3);
    const int KP = (2 * D + 7) & ~7, NP = (R * Mp + 15) & ~15;
    em_update_kernel<39><<<R * M + 1, 128, 0, (cudaStream_t)stream>>>(
        static_cast<uint8_t *>(own_buf), world, (unsigned long long)epoch, len, R, M, Mp, ncomp, mu, var, w, shift,
        var_floor, ab, c, reinterpret_cast<uint32_t *>(tcb), NP, KP, stats_out);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}

}  // extern "C"
