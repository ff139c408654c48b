// dtw.cu -- dtw with beam pruning (decode.py:7-77), the part of the reference's template aligner the graph Viterbi
// kernel does not carry: after every column the reference ranks the column's costs, marks each FINITE cell of rank
// >= beam with the sentinel -1 (decode.py:62-68) and the next column treats marked origins specially (decode.py:44-51):
//   * row 0 -- the first row swept -- SKIPS them (they are not candidates at all, so np.argmin over what is left can
//     fall on a later origin even when everything left is +inf) and resets the sentinel to +inf while scanning;
//   * rows 1.. therefore see them as +inf candidates (np.argmin's first minimum: origin 0 when all are +inf).
// The sentinels of the last column are never reset and stay in the returned cost matrix.
//
// One warp per utterance, template rows strided over lanes, the previous column, the current column and the pruning
// marks in shared memory.  Transitions are a dense [S, S] matrix with +inf entries kept as candidates.  The rank of a
// cell is its position in the (cost, row) order (numpy's argsort on a short column is an insertion sort, i.e. stable).
#include "common.cuh"

namespace srb {

constexpr int DTW_WARPS = 4;
constexpr int DTW_MAX_S = 4096;

__global__ void __launch_bounds__(DTW_WARPS * 32)
dtw_beam_kernel(const float *__restrict__ Dl, int64_t ldD, const int32_t *__restrict__ off, int B,
                const float *__restrict__ trans, int S, int beam, const int64_t *__restrict__ ws_off,
                uint16_t *__restrict__ bp, float *__restrict__ costs_out, const int64_t *__restrict__ path_off,
                int32_t *__restrict__ path_state, int32_t *__restrict__ path_frame, int32_t *__restrict__ path_len) {
    extern __shared__ float dsm[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float *prev = dsm + (size_t)warp * 3 * S, *cur = prev + S;
    int *mark = reinterpret_cast<int *>(prev + 2 * S);        // 1 = pruned by the beam in the previous column
    const float INF = CUDART_INF_F;

    for (int u = blockIdx.x * DTW_WARPS + warp; u < B; u += gridDim.x * DTW_WARPS) {
        const int64_t f0 = off[u];
        const int T = (int)(off[u + 1] - f0);
        uint16_t *bpu = bp + ws_off[u];
        float *cu = costs_out + ws_off[u];
        if (T < 2) {                                           // decode.py:22 asserts; the host layer raises first
            if (lane == 0 && path_len) path_len[u] = -2;
            continue;
        }
        bool empty = false;                                    // row 0 ran out of candidates (np.argmin of nothing)
        for (int j = 0; j < T; ++j) {
            const float *drow = Dl + (f0 + j) * ldD;
            for (int i = lane; i < S; i += 32) {
                float out;
                int bo = 0;
                if (j == 0) {
                    // (0, 0) is the seed (decode.py:34-40); the other rows read column -1, still +inf (decode.py:50)
                    out = i == 0 ? drow[0] : INF;
                } else {
                    float best = INF;
                    bo = -1;
                    const float *tr = trans + (size_t)i * S;
                    for (int o = 0; o < S; ++o) {
                        float v = prev[o];
                        if (mark[o]) {
                            if (i == 0) continue;              // decode.py:46-48: skipped by the row that resets it
                            v = INF;
                        }
                        const float cand = tr[o] + v;
                        if (bo < 0 || cand < best) {           // first candidate, then strict '<': np.argmin
                            best = cand;
                            bo = o;
                        }
                    }
                    if (bo < 0) {
                        empty = true;
                        bo = 0;
                    }
                    out = best + drow[i];
                }
                cur[i] = out;
                bpu[(size_t)j * S + i] = (uint16_t)bo;
            }
            __syncwarp();
            // ---- beam: finite cells of rank >= beam are marked (decode.py:62-68)
            for (int i = lane; i < S; i += 32) {
                const float ci = cur[i];
                int m = 0;
                if (beam >= 0 && ci < INF) {
                    int rank = 0;
                    for (int k = 0; k < S; ++k) {
                        const float ck = cur[k];
                        rank += (ck < ci || (ck == ci && k < i)) ? 1 : 0;
                    }
                    m = rank >= beam;
                }
                mark[i] = m;
                cu[(size_t)j * S + i] = m ? (j == T - 1 ? -1.f : INF) : ci;
            }
            __syncwarp();
            float *t = prev;
            prev = cur;
            cur = t;
        }
        empty = __any_sync(0xffffffffu, empty);
        // ---- the walk back (decode.py:71-77): from (S-1, T-1) to (0, 0), whatever the costs are
        if (lane == 0 && path_len) {
            int len = 0;
            if (empty) {
                len = -2;
            } else {
                int32_t *ps = path_state + path_off[u], *pf = path_frame + path_off[u];
                const int cap = (int)(path_off[u + 1] - path_off[u]);
                int i = S - 1, j = T - 1;
                while (i != 0 || j != 0) {
                    if (j == 0 || len >= cap) {                // the reference would index column -1 from here
                        len = -3;
                        break;
                    }
                    i = bpu[(size_t)j * S + i];
                    --j;
                    ps[len] = i;
                    pf[len] = j;
                    ++len;
                }
            }
            path_len[u] = len;
        }
        __syncwarp();
    }
}

}  // namespace srb

using namespace srb;

extern "C" int srb_dtw_beam(const float *Dl, int64_t ldD, const int32_t *utt_off, int32_t B, const float *trans,
                            int32_t S, int32_t beam, const int64_t *ws_off, uint16_t *bp, float *costs_out,
                            const int64_t *path_off, int32_t *path_state, int32_t *path_frame, int32_t *path_len,
                            void *stream) {
    SRB_REQUIRE(B >= 0 && ldD > 0, SRB_EINVAL, "srb_dtw_beam: bad sizes");
    if (B == 0) return SRB_OK;
    SRB_REQUIRE(Dl && utt_off && trans && ws_off && bp && costs_out, SRB_EINVAL, "srb_dtw_beam: null pointer");
    SRB_REQUIRE(!path_len || (path_off && path_state && path_frame), SRB_EINVAL,
                "srb_dtw_beam: path_len needs path_off, path_state and path_frame");
    SRB_REQUIRE(S >= 2 && S <= DTW_MAX_S && ldD >= S, SRB_EUNSUPPORTED, "srb_dtw_beam: %d template rows (2 .. %d)", S,
                DTW_MAX_S);
    const size_t bytes = (size_t)DTW_WARPS * 3 * S * sizeof(float);
    SRB_REQUIRE(bytes <= (size_t)smem_optin(), SRB_EUNSUPPORTED, "srb_dtw_beam: %d rows do not fit shared memory", S);
    SRB_CUDA(cudaFuncSetAttribute(dtw_beam_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    int64_t grid = ((int64_t)B + DTW_WARPS - 1) / DTW_WARPS;
    const int64_t cap = (int64_t)sm_count() * 8;
    if (grid > cap) grid = cap;
    dtw_beam_kernel<<<(int)grid, DTW_WARPS * 32, bytes, (cudaStream_t)stream>>>(
        Dl, ldD, utt_off, B, trans, S, beam, ws_off, bp, costs_out, path_off, path_state, path_frame, path_len);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}
