// lextree.cu -- text_viterbi (sr/langmodel/spellchecker.py:15-135): the edit-distance dynamic programme of a string
// against a lexical tree with a loop back to the root, SURVEY 8 row f4.
//
// Rows are the tree's nodes in depth-first order (spellchecker.py:7-12,35-37) plus a blank node; columns the characters
// of '*' + text.  Cell (r, c), node_dist = dist_fun(x[c], node r's character) (spellchecker.py:72):
//   root r = 0      c > 0: node_dist + loop_cost + min over the word ends (blank node first) of column c-1   (:92-100)
//   blank node      node_dist + space_cost + min over the word ends but the blank node of column c-1          (:80-82)
//   child of root   deletion (c > 0): node_dist + 1 + (r, c-1); insertion: node_dist + 1 + (root, c)           (:83-91)
//   other nodes     deletion; match (c > 0): node_dist + 0 + (parent, c-1); insertion: node_dist + 1 + (parent, c)
// np.argmin keeps the first minimum of [deletion, match, insertion, loop] / of the word-end list.  The only same-column
// predecessor of a node is its parent, so a column is swept level by level (nodes of one depth in parallel).
//
// One CTA per string: the two live columns sit in a caller-owned global workspace (L2 resident), the back-pointers are
// one byte per cell, the loop / blank-node choices one int per column.  The walk back (spellchecker.py:120-135) is done
// by one thread and writes the matched characters.
#include "common.cuh"

namespace srb {

constexpr int LT_THREADS = 1024;

__device__ __forceinline__ void lt_lexmin(float &v, int &p, float ov, int op) {
    if (ov < v || (ov == v && op < p)) {
        v = ov;
        p = op;
    }
}

// first minimum of val over positions [lo, n) of the word-end list: (value, position) lexicographic block reduction
__device__ void lt_block_min(const float *col, const int32_t *word_ends, int lo, int n, float *red_v, int *red_p,
                             float &out_v, int &out_p) {
    float v = CUDART_INF_F;
    int p = 0x7fffffff;
    for (int k = lo + threadIdx.x; k < n; k += blockDim.x) lt_lexmin(v, p, col[word_ends[k]], k);
#pragma unroll
    for (int d = 16; d; d >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, v, d);
        const int op = __shfl_xor_sync(0xffffffffu, p, d);
        lt_lexmin(v, p, ov, op);
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) {
        red_v[warp] = v;
        red_p[warp] = p;
    }
    __syncthreads();
    if (warp == 0) {
        v = lane < (int)(blockDim.x >> 5) ? red_v[lane] : CUDART_INF_F;
        p = lane < (int)(blockDim.x >> 5) ? red_p[lane] : 0x7fffffff;
#pragma unroll
        for (int d = 16; d; d >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, v, d);
            const int op = __shfl_xor_sync(0xffffffffu, p, d);
            lt_lexmin(v, p, ov, op);
        }
        if (lane == 0) {
            red_v[0] = v;
            red_p[0] = p;
        }
    }
    __syncthreads();
    out_v = red_v[0];
    out_p = red_p[0];
    __syncthreads();
}

__global__ void __launch_bounds__(LT_THREADS)
text_viterbi_kernel(const int32_t *__restrict__ x, const int64_t *__restrict__ x_off, int B,
                    const int32_t *__restrict__ node_val, const int32_t *__restrict__ node_parent,
                    const int32_t *__restrict__ level_nodes, const int32_t *__restrict__ level_off, int n_levels,
                    int n_rows, const int32_t *__restrict__ word_ends, int n_ends, const float *__restrict__ dist,
                    float *__restrict__ col_ws, uint8_t *__restrict__ bp, const int64_t *__restrict__ bp_off,
                    int32_t *__restrict__ bp_jump, float *__restrict__ best_cost, int32_t *__restrict__ out_nodes,
                    const int64_t *__restrict__ out_off, int32_t *__restrict__ out_len) {
    __shared__ float red_v[32];
    __shared__ int red_p[32];
    const float INF = CUDART_INF_F;
    const int space = n_rows - 1;
    for (int u = blockIdx.x; u < B; u += gridDim.x) {
        const int64_t c0 = x_off[u];
        const int n_cols = (int)(x_off[u + 1] - c0);
        float *prev = col_ws + (size_t)u * 2 * n_rows, *cur = prev + n_rows;
        uint8_t *bpu = bp + bp_off[u];
        int32_t *jump = bp_jump + 2 * c0;                       // [column][0] = loop choice, [1] = blank-node choice
        if (n_cols < 2) {                                        // '*' alone: the reference's walk does not terminate
            if (threadIdx.x == 0) {
                best_cost[u] = INF;
                out_len[u] = -2;
            }
            continue;
        }
        for (int c = 0; c < n_cols; ++c) {
            const int xc = x[c0 + c];
            const float *drow = dist ? dist + (c0 + c) * (int64_t)n_rows : nullptr;
            // ---- root and blank node: both read the word ends of the previous column
            if (c == 0) {
                if (threadIdx.x == 0) {
                    cur[0] = 0.f;                                // costs[0, :] = 0 and (0, 0) is skipped (:43,68-69)
                    cur[space] = INF;                            // column -1 is the untouched last column
                    jump[0] = -1;
                    jump[1] = 1;
                }
            } else {
                float m1;
                int p1;
                lt_block_min(prev, word_ends, 1, n_ends, red_v, red_p, m1, p1);
                if (threadIdx.x == 0) {
                    const float nd_sp = drow ? drow[space] : (xc != node_val[space] ? 1.f : 0.f);
                    cur[space] = (nd_sp + 0.f) + m1;             // :81  (all +inf -> argmin 0 -> the first listed end)
                    jump[2 * c + 1] = p1 == 0x7fffffff ? 1 : p1;
                    float m = prev[space];
                    int p = 0;
                    lt_lexmin(m, p, m1, p1);
                    const float nd_rt = drow ? drow[0] : (xc != node_val[0] ? 1.f : 0.f);
                    const float loopc = (nd_rt + 1.f) + m;       // :95-98
                    cur[0] = loopc;
                    jump[2 * c] = loopc < INF ? p : -1;          // -1: every sub-cost inf -> from_points[0] = [0, 0]
                }
            }
            __syncthreads();
            // ---- the tree, level by level
            for (int d = 0; d < n_levels; ++d) {
                const int l0 = level_off[d], l1 = level_off[d + 1];
                for (int k = l0 + threadIdx.x; k < l1; k += blockDim.x) {
                    const int r = level_nodes[k];
                    const int p = node_parent[r];
                    const float nd = drow ? drow[r] : (xc != node_val[r] ? 1.f : 0.f);
                    float best = INF;
                    int code = 0;
                    if (c > 0) best = (nd + 1.f) + prev[r];                      // deletion   (:85-87,103-105)
                    if (c > 0 && p != 0) {                                      // match      (:106-109)
                        const float t = (nd + 0.f) + prev[p];
                        if (t < best) {
                            best = t;
                            code = 1;
                        }
                    }
                    const float t = (nd + 1.f) + cur[p];                        // insertion  (:88-91,110-113)
                    if (t < best) {
                        best = t;
                        code = 2;
                    }
                    cur[r] = best;
                    bpu[(size_t)c * n_rows + r] = (uint8_t)code;
                }
                __syncthreads();
            }
            float *t = prev;
            prev = cur;
            cur = t;
        }
        // ---- best word end of the last column (:120-123) and the walk back (:126-135)
        float bv;
        int bpos;
        lt_block_min(prev, word_ends, 0, n_ends, red_v, red_p, bv, bpos);
        if (threadIdx.x == 0) {
            if (bpos == 0x7fffffff) bpos = 0;
            best_cost[u] = bv;
            int32_t *out = out_nodes + out_off[u];
            const int cap = (int)(out_off[u + 1] - out_off[u]);
            int r = word_ends[bpos], c = n_cols - 1, len = 0;
            if (cap > 0) out[len++] = r;
            while (c != 1 && len >= 0) {
                int nr, nc;
                if (r == 0) {
                    const int j = jump[2 * c];
                    nr = j < 0 ? 0 : word_ends[j];
                    nc = j < 0 ? 0 : c - 1;
                } else if (r == space) {
                    nr = word_ends[jump[2 * c + 1]];
                    nc = c - 1;
                } else {
                    const int code = bpu[(size_t)c * n_rows + r];
                    nr = code == 0 ? r : node_parent[r];
                    nc = code == 2 ? c : c - 1;
                }
                r = nr;
                c = nc;
                if (c < 1) {                                                    // jumped past column 1: never terminates
                    len = -3;
                    break;
                }
                if (r != 0) {
                    if (len >= cap) {
                        len = -1;
                        break;
                    }
                    out[len++] = r;
                }
            }
            out_len[u] = len;                                    // node indices, end of the string first
        }
        __syncthreads();
    }
}

}  // namespace srb

using namespace srb;

extern "C" int srb_text_viterbi(const int32_t *x, const int64_t *x_off, int32_t B, const int32_t *node_val,
                                const int32_t *node_parent, const int32_t *level_nodes, const int32_t *level_off,
                                int32_t n_levels, int32_t n_rows, const int32_t *word_ends, int32_t n_ends,
                                const float *dist, float *col_ws, uint8_t *bp, const int64_t *bp_off, int32_t *bp_jump,
                                float *best_cost, int32_t *out_nodes, const int64_t *out_off, int32_t *out_len,
                                void *stream) {
    SRB_REQUIRE(B >= 0 && n_rows >= 2 && n_levels >= 0 && n_ends >= 1, SRB_EINVAL, "srb_text_viterbi: bad sizes");
    if (B == 0) return SRB_OK;
    SRB_REQUIRE(x && x_off && node_val && node_parent && level_nodes && level_off && word_ends && col_ws && bp && bp_off &&
                    bp_jump && best_cost && out_nodes && out_off && out_len,
                SRB_EINVAL, "srb_text_viterbi: null pointer");
    SRB_REQUIRE(n_ends >= 2, SRB_EUNSUPPORTED,
                "srb_text_viterbi: a tree without word ends (the reference's np.argmin gets an empty list)");
    int grid = B < 4 * sm_count() ? B : 4 * sm_count();
    text_viterbi_kernel<<<grid, LT_THREADS, 0, (cudaStream_t)stream>>>(
        x, x_off, B, node_val, node_parent, level_nodes, level_off, n_levels, n_rows, word_ends, n_ends, dist, col_ws, bp,
        bp_off, bp_jump, best_cost, out_nodes, out_off, out_len);
    SRB_CUDA(cudaGetLastError());
    return SRB_OK;
}
