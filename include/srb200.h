/*
 * srb200.h -- C ABI of libsrb200.so: the B200 (sm_100a) GMM-HMM scoring / alignment / EM hot path.
 *
 * Drop-in boundary for the path SURVEY.md section 8 names.  The reference (tjysdsg/speech-recognition) has
 * no FFI: its boundary is the Python module API of sr/recognition (sr/recognition/__init__.py:1-5).  The
 * host-side mirror of that API lives in speech-recognition_b200/sr/recognition and calls THIS library
 * through ctypes (speech-recognition_b200/srb200/lib.py); INTEGRATION.md shows the stub.
 *
 * Conventions (all entry points):
 *   - plain C types, DEVICE pointers unless the name says host; the caller owns every allocation;
 *   - `stream` is a cudaStream_t passed as void*; calls are asynchronous on it, re-entrant per stream;
 *   - return 0 on success, a negative SRB_E* code otherwise; srb_last_error_string() explains (thread local);
 *   - costs are NEGATIVE log-likelihoods, smaller is better (decode.py:86-87); +inf = impossible;
 *   - ragged batches are CSR: frames of utterance u are rows utt_off[u] .. utt_off[u+1]-1 of X[N,D] (fp32);
 *   - a packed model (see srb_pack_model) holds R tied states x Mp Gaussians (Mp = M rounded up to 4).
 */
#ifndef SRB200_H
#define SRB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SRB_OK 0
#define SRB_EINVAL (-1)      /* bad argument (null pointer, size out of range) */
#define SRB_EUNSUPPORTED (-2) /* shape outside what the kernels are built for */
#define SRB_ECUDA (-3)       /* a CUDA runtime call failed */
#define SRB_ENODEVICE (-4)   /* no sm_100 device */

#define SRB_MAX_DIM 64       /* feature dimension limit of the register-resident scorer */
#define SRB_DP_MAX_STATES 8  /* per-word state limit of the fused classify / train kernels */

int srb_version(void);
const char *srb_last_error_string(void);
/* multiprocessor count and max opt-in shared memory of the current device */
int srb_device_info(int *sm_count, int *smem_optin_bytes, int *cc_major, int *cc_minor);

/* ---- packed model ------------------------------------------------------------------------------
 * Replaces MultivariateNormal.__init__/cov setter/pdf and GMM.evaluate's per-call arithmetic
 * (hmm_state.py:10-45,114-120): for Gaussian g, dim d (x' = x - shift[d]):
 *     log2 N_g(x) w_g = sum_d ( x'^2 * A[g,d] + x' * B[g,d] ) + C[g]
 *     A = -0.5/var * log2(e),  B = (mu-shift)/var * log2(e),
 *     C = (log w - 0.5*(D log 2pi + sum log var) - 0.5*sum (mu-shift)^2/var) * log2(e)
 * Layout: ab = float2[G4/4][D][4] (A,B pairs of 4 consecutive Gaussians per dim), c = float[G4],
 * G4 = R*Mp.  Padding Gaussians carry C = -inf.  Weights are NOT normalised.
 * Inputs are float64 device arrays mu[R*M,D], var[R*M,D], w[R*M], shift[D]. */
int64_t srb_packed_ab_floats(int R, int M, int D);
int64_t srb_packed_c_floats(int R, int M);
int srb_pack_model(const double *mu, const double *var, const double *w, const double *shift, int R, int M, int D,
                   float *ab, float *c, float *shift_f32, void *stream);

/* ---- K1: GMM scoring (GMM.evaluate, hmm_state.py:114-120, batched over frames and states) -------
 * E[n, r] = -log sum_m w_rm N_rm(X[n])  for r in [0,R).  Any output may be NULL:
 *   E        float[N, ldE]   costs
 *   best_cost/best_state     per-frame min over r (first minimum) -- the streaming reduction used for
 *                            the large sweep where E must not be materialised
 *   comp     float[N, R*Mp]  per-Gaussian natural-log  log(w_m N_m(x))  (GMM.evaluate(..., False), log domain) */
int srb_gmm_score(const float *X, int64_t N, int D, const float *ab, const float *c, const float *shift, int R,
                  int Mp, float *E, int64_t ldE, float *best_cost, int32_t *best_state, float *comp, void *stream);

/* Same on rows [row0, row0+nrows) of X only (outputs stay indexed by absolute row): lets a caller score each
 * word's contiguous block of a word-sorted corpus against that word's states (HMM.fit works per word, hmm.py:57). */
int srb_gmm_score_rows(const float *X, int64_t N, int64_t row0, int64_t nrows, int D, const float *ab, const float *c,
                       const float *shift, int R, int Mp, float *E, int64_t ldE, float *best_cost, int32_t *best_state,
                       float *comp, void *stream);

/* E[n, r] = -log sum_j exp(-P[n, r*J + j]), P float[N, R*J] dense: merges the costs of J sub-states into their state.
 * With it srb_gmm_score_tc (4 Gaussians per column group) scores states of 4*J Gaussians: the packed model of R states
 * x 4*J Gaussians IS the packed model of R*J states x 4 (same ab / c memory), see BaumWelchTrainer.e_step. */
int srb_neg_lse_groups(const float *P, int64_t N, int R, int J, float *E, int64_t ldE, void *stream);
/* Running best state for models too large for one tensor-core launch (BASELINE config 5: 2 000 states x 32): the
 * states are scored block by block as sub-states (srb_gmm_score_tc on the block's slice of ab / c) and each block's
 * P[N, Rb*J] is folded into best_cost / best_state (first != 0 on the first block: starts from +inf / state 0;
 * strict '<' keeps the first minimum like srb_gmm_score).  The cost matrix is never materialised. */
int srb_merge_best(const float *P, int64_t N, int Rb, int J, int state0, int first, float *best_cost,
                   int32_t *best_state, void *stream);

/* ---- the same reduction as ONE persistent GEMM (score_gemm.cu): a 128-frame tile's split [x'^2, x', 1] operand stays in
 * tensor memory while ALL of the model's Gaussians stream past it in blocks of 160 columns (TMA bulk copies through a
 * 2-stage shared-memory ring, two accumulators in tensor memory, 3xTF32), the per-state log-sum-exp and the running
 * best state are kept in registers; X is read once, only best_cost / best_state are written.
 *   srb_gmm_best_gemm_supported(D, R, Mp)  1 for D = 39 and Mp a multiple of 16 up to 160 (any number of states R)
 *   srb_model_blocks_floats(R, Mp, D)      floats of the blocked split model
 *   srb_pack_model_blocks(ab, c, ...)      ab, c (srb_pack_model) -> blocks[ceil(R / SB)][2][K/4][160][4], SB = 160 / Mp
 *                                          states per block; once per model update
 *   srb_gmm_best_gemm(...)                 best_cost / best_state as srb_gmm_score (first minimum) */
int srb_gmm_best_gemm_supported(int D, int R, int Mp);
int64_t srb_model_blocks_floats(int R, int Mp, int D);
int srb_pack_model_blocks(const float *ab, const float *c, int R, int Mp, int D, float *blocks, void *stream);
int srb_gmm_best_gemm(const float *X, int64_t N, int D, const float *blocks, const float *shift, int R, int Mp,
                      float *best_cost, int32_t *best_state, void *stream);
/* The same kernel in ROWS mode, for word-sorted corpora (HMM.fit scores each word's frames against that word's states,
 * hmm.py:57): blocks packed with SB = the states of one word (srb_pack_model_blocks_sb; SB * Mp <= 160), tile_desc
 * int32[ntiles][4] (device, 16-byte aligned) = {first frame, frames <= 128, block = word, 0}; every tile is scored
 * against ITS block only and E[frame, s] = states[block * SB + s].evaluate(X[frame]) for s < SB (ldE >= SB).  One launch
 * for all words, no sub-state scratch (the round-1 path wrote and re-read 256 B/frame of it). */
int64_t srb_model_blocks_floats_sb(int R, int Mp, int D, int SB);
int srb_pack_model_blocks_sb(const float *ab, const float *c, int R, int Mp, int D, int SB, float *blocks, void *stream);
int srb_gmm_score_gemm(const float *X, int64_t N, int D, const float *blocks, const float *shift, int R, int Mp, int SB,
                       const int32_t *tile_desc, int64_t ntiles, float *E, int64_t ldE, void *stream);

/* ---- fused isolated-word classifier (core.py:79-87 over hmm.py:126-135 over decode.py:80-146) ----
 * W word models of S states each (packed model has R = W*S states, word-major); trans float[W,S,S] with
 * trans[w,i,j] = cost j -> i, +inf = none (decode.py:86-87).  Start in state 0 at frame 0
 * (decode.py:99-100), end in state S-1 at the last frame (decode.py:126-127).
 *   final float[B,W]  = costs[-1,-1] per word (hmm.py:135);   label int32[B] = first argmin (core.py:85)
 * Utterances need >= 2 frames (decode.py reads column -1 at c == 0). */
int srb_classify(const float *X, const int32_t *utt_off, int32_t B, int64_t N, int D, const float *ab,
                 const float *c, const float *shift, int W, int S, int Mp, const float *trans, float *final,
                 int32_t *label, void *stream);

/* Same classifier with the Gaussian contraction on the tcgen05 tensor cores (3xTF32 split, FP32 accumulation in
 * tensor memory; score_tc.cu).  The packed model is split once per model update:
 *   srb_tc_model_floats(R, Mp, D)   floats of the split operand, 0 when the model does not fit one MMA (R*Mp <= 256)
 *   srb_pack_model_tc(ab, c, ...)   ab, c (srb_pack_model) -> tcb[2][K/4][NP][4]: TF32 high and low halves, K-major;
 *                                   K = 2*D features + the constant 1 (coefficient c), padded to a multiple of 8
 *   srb_classify_tc_supported(...)  1 when srb_classify_tc covers the shape (D = 39, Mp = 4, shared memory fits)
 *   srb_classify_tc(...)            arguments and results as srb_classify, tcb in place of ab and c
 * Costs agree with srb_classify within FP32 accumulation noise (~1e-6 relative); labels are equal wherever the two
 * best words differ by more than that. */
int64_t srb_tc_model_floats(int R, int Mp, int D);
/* diagnostics: srb_classify_tc with a caller-owned device buffer timeline int64[64][10] (or NULL): CTA 0 stamps clock64
 * at its pipeline events (score_tc.cu, TC_STAMP) for its first 64 tiles.  An argument, not library state: the library
 * keeps no hidden global state and every entry point is re-entrant per stream. */
int srb_pack_model_tc(const float *ab, const float *c, int R, int Mp, int D, float *tcb, void *stream);
int srb_classify_tc_supported(int D, int W, int S, int Mp);
/* The scorer alone on the same kernel: E[n, r] = states[r].evaluate(X[n]) like srb_gmm_score (E only).  chunk_off
 * int32[nchunks+1] is any partition of the N frames into chunks of >= 1 frame (e.g. the utterance offsets, or steps of
 * 32 frames); it only drives the split of the work over CTAs and sub-streams. */
int srb_gmm_score_tc_supported(int D, int R, int Mp);
int srb_gmm_score_tc(const float *X, int64_t N, int D, const float *tcb, const float *shift, int R, int Mp,
                     const int32_t *chunk_off, int32_t nchunks, float *E, int64_t ldE, void *stream);
int srb_classify_tc(const float *X, const int32_t *utt_off, int32_t B, int64_t N, int D, const float *tcb,
                    const float *shift, int W, int S, int Mp, const float *trans, float *final,
                    int32_t *label, void *stream);
int srb_classify_tc_timeline(const float *X, const int32_t *utt_off, int32_t B, int64_t N, int D, const float *tcb,
                             const float *shift, int W, int S, int Mp, const float *trans, float *final,
                             int32_t *label, int64_t *timeline, void *stream);


/* ---- K2: batched Viterbi over state graphs with non-emitting states ------------------------------
 * Replaces decode_hmm_states (decode.py:80-146) and, with E = local distances, dtw (decode.py:7-77).
 * One or more graphs are packed back to back (graph g owns states g_state_off[g] .. g_state_off[g+1]-1);
 * utterance u runs on graph utt_graph[u] (NULL: graph 0).  Per graph-local state r:
 *   emit[r]      column of E holding states[r].evaluate(x), or -1 for a NES (hmm_state.py:90-91, cost 0)
 *   pred_ptr     CSR (absolute offsets) over the finite entries of transitions[r, :] in ascending origin order
 *                (decode.py:104-107), pred_idx = origin (graph local), pred_cost = transitions[r, origin]
 *   nes_idx      ascending local indices of the NES (they split a column into runs, see viterbi.cu)
 *   end_state    candidate end states in caller order; the end frame is the last frame (callers pass -1 or T-1:
 *                main.py:60, continuous_speech.py:89, decode.py:126-127) unless end_frame is given
 *   end_frame    (srb_viterbi_graph only, optional) frame of each end point, negative values count from the end like
 *                the reference's Python indexing (decode.py:131-137 accepts any [state, frame]); needs costs_out,
 *                from which the costs of end points off the last frame are read.  The path then starts at the best
 *                end point's frame and frame_label is -1 past it.
 * At most 255 predecessors per state.  Start: only (state 0, frame 0) is seeded (decode.py:99-100).
 * Outputs: best_cost[B], best_end[B] (last minimum wins, decode.py:131-134; +inf / -1 when no path),
 *   bp[bp_off[u] + c*R + r] = position of the winning origin in r's predecessor list (first minimum wins),
 *   costs_out (optional) same indexing, float: costs[r, c] (decode.py returns it),
 *   path_state/path_frame[path_off[u] ...] the reversed path of decode.py:142-146 (end cell excluded,
 *   stops at frame 0), path_len[u] its length (0 when no path, -1 when the capacity path_off[u+1]-path_off[u]
 *   was too small; T + 2 * number of NES + 1 always suffices),
 *   frame_label[N] (optional) = g_state_off[graph] + state of the emitting state aligned to each frame (the
 *   hard alignment kmeans.py:196-205 / continuous_speech.py:83-106 pool by; -1 for utterances without a path;
 *   a frame emitted twice around a NES keeps the EARLIER state). */
int srb_viterbi_graph(const float *E, int64_t ldE, const int32_t *utt_off, int32_t B, const int32_t *utt_graph,
                      const int32_t *g_state_off, const int32_t *g_nes_off, const int32_t *g_end_off,
                      const int32_t *emit, const int32_t *pred_ptr, const int32_t *pred_idx, const float *pred_cost,
                      const int32_t *nes_idx, const int32_t *end_state, const int32_t *end_frame, int32_t Rmax,
                      const int64_t *bp_off, uint8_t *bp, float *costs_out, float *best_cost, int32_t *best_end,
                      const int64_t *path_off, int32_t *path_state, int32_t *path_frame, int32_t *path_len,
                      int32_t *frame_label, void *stream);

/* ---- dtw with beam pruning (decode.py:7-77 incl. the beam of :42-48,62-68) -----------------------------------------
 * Dl float[N, ldD] local distances of every frame to the S template rows (srb_local_dist); trans float[S,S] dense,
 * trans[i*S + o] = cost o -> i, +inf entries stay candidates as in the reference (decode.py:44-51 does not skip them).
 * After each column the reference sorts its costs and marks every FINITE cell of rank >= beam with -1; the next
 * column's row 0 skips those origins, the other rows see them as +inf (row 0's scan has reset them), and the marks of
 * the last column stay in the returned matrix.  beam < 0 = no pruning (np.inf).  Ranks are by (cost, row).
 *   costs_out float[ws_off[u] + j*S + i] = costs[i, j] (pruned cells: +inf, -1 in the last column)
 *   bp uint16 same indexing = winning origin row (the first candidate when every candidate is +inf, like np.argmin)
 *   path_state/path_frame[path_off[u] ...] the reversed path of decode.py:71-77 from (S-1, T-1) down to (0, 0),
 *   end cell excluded; path_len[u] = T-1, or -2 when row 0 of some column had no candidate left (the reference's
 *   np.argmin raises ValueError), -3 when the walk reaches column 0 off row 0 (the reference then wraps to column -1
 *   and does not terminate).  Utterances need >= 2 frames and S >= 2 (decode.py:22); S <= 4096. */
int srb_dtw_beam(const float *Dl, int64_t ldD, const int32_t *utt_off, int32_t B, const float *trans, int32_t S,
                 int32_t beam, const int64_t *ws_off, uint16_t *bp, float *costs_out, const int64_t *path_off,
                 int32_t *path_state, int32_t *path_frame, int32_t *path_len, void *stream);

/* Same contract and bit-identical outputs as srb_viterbi_graph, for graphs in BANDED FORM (what
 * build_state_sequences, continuous_speech.py:13-53, and plain left-to-right models produce): the emitting
 * predecessors of an emitting state r lie in {r-2, r-1, r}; it may have one non-emitting predecessor n < r (its
 * "entry"); states feeding a NES are emitting and have no entry.  Extra per-state arrays (graph-local, packed like
 * emit): band float[R,3] = costs of the arcs from r-2, r-1, r (+inf = absent); codes int32[R] = back-pointer codes
 * (positions in the ascending predecessor list) of those three arcs and of the entry arc, one byte each;
 * entry_nes int32[R] (-1 = none), entry_cost float[R]; entry_state lists the states with an entry per graph
 * (g_entry_off).  srb200/graph.py derives them (StateGraph.banded_form) or reports that a graph is not banded. */
int srb_viterbi_banded(const float *E, int64_t ldE, const int32_t *utt_off, int32_t B, const int32_t *utt_graph,
                       const int32_t *g_state_off, const int32_t *g_nes_off, const int32_t *g_end_off,
                       const int32_t *emit, const int32_t *pred_ptr, const int32_t *pred_idx, const float *pred_cost,
                       const int32_t *nes_idx, const int32_t *end_state, const float *band, const int32_t *codes,
                       const int32_t *entry_nes, const float *entry_cost, const int32_t *g_entry_off,
                       const int32_t *entry_state, int32_t Rmax, const int64_t *bp_off, uint8_t *bp, float *costs_out,
                       float *best_cost, int32_t *best_end, const int64_t *path_off, int32_t *path_state,
                       int32_t *path_frame, int32_t *path_len, int32_t *frame_label, void *stream);

/* ---- decoded path -> word string (main.py:59-67, the step right after decode_hmm_states) ----
 * path_state/path_off/path_len as written by srb_viterbi_graph / srb_viterbi_banded (reversed paths);
 * state_word[g_state_off[g] + r] = word of graph g's local state r, or -1 for a non-emitting state.
 * The path is read forwards, consecutive duplicates removed, split at the NES; the first state of each piece gives
 * the word.  words int32[B, max_words] receives the first max_words words, nwords[B] the full count. */
int srb_path_words(const int32_t *path_state, const int64_t *path_off, const int32_t *path_len, int32_t B,
                   const int32_t *utt_graph, const int32_t *g_state_off, const int32_t *state_word, int32_t max_words,
                   int32_t *words, int32_t *nwords, void *stream);

/* ---- text_viterbi over a lexical tree (sr/langmodel/spellchecker.py:15-135) --------------------------
 * The edit-distance dynamic programme of '*' + text against the flattened lextree with a loop back to the root.
 * Tree (built by the host, sr/langmodel/lextree.py FlatTree): n_rows nodes in the depth-first order of
 * spellchecker.py:7-12, the LAST row being the blank node the spell-checker appends (:37); node_val int32[n_rows]
 * (character code), node_parent int32[n_rows] (spellchecker.py:52-63; -1 for the root and the blank node);
 * level_nodes = node indices of depth >= 1 sorted by depth, level_off int32[n_levels+1] their ranges (a node's only
 * same-column predecessor is its parent: a column is swept level by level); word_ends int32[n_ends] = blank node
 * first, then the end-of-word nodes in index order (:54-59), n_ends >= 2.
 * Strings: x int32[sum n_cols] character codes INCLUDING the leading '*' (:26), x_off int64[B+1]; n_cols >= 2.
 * dist (optional) float[sum n_cols, n_rows] = dist_fun(x[c], node r) for a caller-supplied distance; NULL = the
 * default int(a != b) (:15).  Costs: deletion 1, insertion 1, match 0, space 0, loop 1 (:28-32); first minimum of
 * [deletion, match, insertion, loop] and of the word-end lists (np.argmin).
 * Workspaces (caller-owned): col_ws float[B * 2 * n_rows]; bp uint8[bp_off[u] + c*n_rows + r] (bp_off[u] =
 * x_off[u] * n_rows); bp_jump int32[sum n_cols, 2].
 * Outputs: best_cost[B] (:123); out_nodes[out_off[u] ...] the node indices the walk back of :126-135 appends, END OF
 * THE STRING FIRST (the caller maps them to characters and reverses, :135); out_len[u] their count, -1 when the
 * capacity out_off[u+1]-out_off[u] was too small (n_cols * (n_levels + 2) always suffices), -2 for an empty text and
 * -3 when the walk jumps past column 1 (in both cases the reference's `while 1` does not terminate). */
int srb_text_viterbi(const int32_t *x, const int64_t *x_off, int32_t B, const int32_t *node_val,
                     const int32_t *node_parent, const int32_t *level_nodes, const int32_t *level_off, int32_t n_levels,
                     int32_t n_rows, const int32_t *word_ends, int32_t n_ends, const float *dist, float *col_ws,
                     uint8_t *bp, const int64_t *bp_off, int32_t *bp_jump, float *best_cost, int32_t *out_nodes,
                     const int64_t *out_off, int32_t *out_len, void *stream);

/* ---- K4: batched forward-backward (Baum-Welch E-step) for NES-free graphs -------------------------
 * The reference has NO forward-backward; semantics are defined to be consistent with
 * srb_viterbi_graph: same graph packing, start in state 0 at frame 0, end in any end_state at the last frame;
 * min instead of log-sum-exp gives the Viterbi costs.  succ_* is the successor CSR of the same edges
 * (succ_ptr absolute, per origin), pred_tgt[j] the target of predecessor edge j.
 *   gamma[ws_off[u] + c*R + r]  state posterior (sums to 1 over r);  neg_log_z[u] = -log P(x) (double)
 *   xi_sum[j] += sum_c posterior of edge j (double, indexed like pred_idx; caller zeroes it; optional).  It is formed
 *   WITHOUT atomics: every warp adds its utterances' edge posteriors to a private row of the caller-owned workspace
 *   xi_part (srb_fb_xi_parts() rows of nnz_total doubles, nnz_total = number of edges of all packed graphs) and the
 *   rows are summed in a fixed order, so xi_sum is bit-reproducible run to run (needed when xi_sum != NULL). */
int srb_fb_xi_parts(void);
int srb_forward_backward(const float *E, int64_t ldE, const int32_t *utt_off, int32_t B, const int32_t *utt_graph,
                         const int32_t *g_state_off, const int32_t *g_end_off, const int32_t *emit,
                         const int32_t *pred_ptr, const int32_t *pred_idx, const float *pred_cost,
                         const int32_t *pred_tgt, const int32_t *succ_ptr, const int32_t *succ_idx,
                         const float *succ_cost, const int32_t *end_state, int32_t Rmax, int32_t NNZmax,
                         const int64_t *ws_off, float *gamma, double *neg_log_z, double *xi_sum, double *xi_part,
                         int32_t nnz_total, void *stream);

/* The same E-step for left-to-right models: every graph NES-free with all predecessors of r inside {r-2, r-1, r}
 * (what HMM.fit / calc_transition_costs build, kmeans.py:53-95) and Rmax <= 32.  band / codes as in
 * srb_viterbi_banded (band[r] = costs from r-2, r-1, r; codes[r] = their positions in r's predecessor list, one
 * byte each); order: optional int32[B] permutation of the utterances (falling length balances the lane groups).
 * One lane per state, 32 / pow2(Rmax) utterances per warp; outputs and the xi_part workspace as srb_forward_backward
 * (the per-utterance edge sums are formed in registers, then added to the warp's row). */
int srb_forward_backward_banded(const float *E, int64_t ldE, const int32_t *utt_off, int32_t B, const int32_t *utt_graph,
                                const int32_t *g_state_off, const int32_t *g_end_off, const int32_t *emit,
                                const int32_t *pred_ptr, const int32_t *end_state, const float *band,
                                const int32_t *codes, const int32_t *order, int32_t Rmax, const int64_t *ws_off,
                                float *gamma, double *neg_log_z, double *xi_sum, double *xi_part, int32_t nnz_total,
                                void *stream);


/* ---- K5 / K6: sufficient statistics, K-means assignment, M-step ---------------------------------
 * Statistics row of one group (Gaussian / segment / cluster), srb_stats_row(D) = 2*D + 3 doubles:
 *   [S0, S1[0..D), S2[0..D), count, loglik]   S1, S2 = first / second moments about the group's shift vector.
 * `partial` is caller-owned scratch of srb_stats_parts() * groups * srb_stats_row(D) floats (one private slice
 * per accumulating warp: no atomics); the slices are summed in a fixed order in double into `out`
 * (accumulate != 0 adds to `out`), so the result does not depend on scheduling and can be allreduced as is. */
int srb_stats_row(int D);
int srb_stats_parts(void);

/* moments of the frames of each group: label[n] in [0,K) or < 0 to skip; shift double[K,D].
 * Replaces segment_data/combine_templates (kmeans.py:15-50), calc_variance (kmeans.py:6-12: unbiased variance
 * = (S2 - S1^2/S0)/(S0-1)), cluster_centroids and the centroid update (kmeans.py:158-164,187). */
int srb_group_stats(const float *X, int64_t N, int D, const int32_t *label, int K, const double *shift,
                    float *partial, double *out, int accumulate, void *stream);

/* srb_group_stats without caller scratch and without any allocation: the private slices live in a file-scope
 * `static __device__` array of SRB_STATIC_SCRATCH_FLOATS floats (48 MiB, held by every GPU that loads the library until
 * the process exits).  No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (CUDA graph capture, tight serving loops).  Same arguments and results as srb_group_stats minus
 * `partial`.  Limit: srb_stats_parts() * K * srb_stats_row(D) <= srb_static_scratch_floats() (K <= 65 groups at D = 39
 * on a 148-SM B200), else SRB_EUNSUPPORTED.  Calls cudaMemsetAsync + two kernels only: no cudaMalloc / cudaFree.
 * All calls on one device share the array (it is cleared at the start of each call): calls that overlap on the GPU,
 * e.g. on two streams, race -- serialise them. */
#define SRB_STATIC_SCRATCH_FLOATS (12 * 1024 * 1024)
int64_t srb_static_scratch_floats(void);
int srb_group_stats_static(const float *X, int64_t N, int D, const int32_t *label, int K, const double *shift,
                           double *out, int accumulate, void *stream);

/* E-step of GMM.em (hmm_state.py:122-136) for frames hard-aligned to tied states: label[n] = state in [0,R)
 * (< 0: skip); posteriors over the FIRST ncomp components of the state (hmm.py:104-124 trains 2^i of them),
 * computed in the log domain (no underflow rows); shift = the component's current mean.
 * mu/var double[R*M,D], lc double[R*M] from srb_gauss_consts.  ncomp <= 8.  out: double[R*M, row]. */
int srb_gmm_stats(const float *X, int64_t N, int D, const int32_t *label, int R, int M, int ncomp, const double *mu,
                  const double *var, const double *lc, float *partial, double *out, int accumulate, void *stream);

/* Baum-Welch (soft-count) variant for rows [row0, row0+nrows) of X that belong to ONE word model: gamma float[N, S]
 * state posteriors of that word's S states from srb_forward_backward (ws_off[u] = utt_off[u]*S), mu/var/lc/out
 * already offset to the word's first state (rows [S*M]).  Group (s, m) gets weight gamma[n,s] * p_m; `count` of a
 * state = sum of its gamma, so srb_mstep's w = S0/count is the soft update.  States with gamma <= gmin are skipped.
 * partial: srb_stats_parts()/8 * S*M * srb_stats_row(D) floats suffice (srb_stats_parts() * ... always does).
 * The reference has no soft counts: parity unpinned, defined to reduce to srb_gmm_stats when gamma
 * is the 0/1 indicator of the Viterbi alignment. */
int srb_gmm_stats_soft(const float *X, int64_t N, int64_t row0, int64_t nrows, int D, const float *gamma, int S, int M,
                       int ncomp, const double *mu, const double *var, const double *lc, double gmin, float *partial,
                       double *out, int accumulate, void *stream);

/* local distances to S templates Y double[S,D]: kind 0 = Euclidean norm (kmeans.py:111), kind 1 =
 * mahalanobis(x, y_s, var_s) (hmm_state.py:48-58) with var double[S, var_stride>=D or 0 for one shared row]
 * (kmeans.py:183 uses cov[0] for every centroid -> var_stride 0).  out float[N,ldo] and/or argmin_out int32[N]
 * (first minimum, kmeans.py:185). */
int srb_local_dist(const float *X, int64_t N, int D, const double *Y, const double *var, int S, int kind,
                   int var_stride, float *out, int64_t ldo, int32_t *argmin_out, void *stream);

/* K-means assignment for many pools in one launch (kmeans.py:180-185 for every state of every word at once): frame n
 * belongs to pool group[n] in [0, G) (< 0: skipped, argmin -1); Y double[G*k, D] the pools' centroids, var double[G, D] the
 * ONE variance row each pool uses for all its centroids (kmeans.py:183 passes cov[0]).  argmin_out[n] = first minimum of
 * mahalanobis(centroid, x, var) over the pool's k centroids. */
int srb_kmeans_assign(const float *X, int64_t N, int D, const int32_t *group, const double *Y, const double *var, int G,
                      int k, int32_t *argmin_out, void *stream);

/* The statistics of ONE K-means or ONE EM iteration for many pools (HMM states) at once, X read once: frame n belongs to
 * pool state[n] in [0, G) (< 0 or >= G: skipped) and is weighed over the pool's first k components, then the weighted
 * moments of every (pool, component) are accumulated (rows [G*M, srb_stats_row(D)] like srb_gmm_stats; k in {1,2,4,8}).
 *   mode 0  GMM.em E-step (hmm_state.py:122-136): weights = posteriors under center/var double[G*M, D] and
 *           lc double[G*M] (srb_gauss_consts), moments about the component's own mean; `shift` unused.
 *           Same results as srb_gmm_stats.
 *   mode 1  kmeans.py:180-187: weights = one-hot first minimum of mahalanobis(center_c, x, var[pool]) with the ONE
 *           variance row var double[G, D] per pool (kmeans.py:183), moments about shift double[G*M, D]; M must equal
 *           k.  Same results as srb_kmeans_assign followed by srb_group_stats; argmin_out int32[N] optional (-1 = skipped).
 * partial: caller scratch of srb_pool_stats_partial_floats(G, k, D) floats, 8-byte aligned (srb_pool_stats_parts() private
 * slices of G*k rows, one per warp, and one 64-bit pool mask per slice: with G <= 64 the slices are write-once and the
 * reduction skips the pools a warp never met).  The parameter tables live in shared memory:
 * srb_pool_stats_supported(G, k, D) says whether they fit (55 pools x 4 x 39 do), else SRB_EUNSUPPORTED.
 * Deterministic: private slices, fixed-order double reduction. */
int srb_pool_stats_parts(void);
int64_t srb_pool_stats_partial_floats(int G, int k, int D);
int srb_pool_stats_supported(int G, int k, int D);
int srb_pool_stats(const float *X, int64_t N, int D, const int32_t *state, int G, int M, int k, int mode,
                   const double *center, const double *var, const double *lc, const double *shift, float *partial,
                   double *out, int32_t *argmin_out, int accumulate, void *stream);

/* srb_pool_stats in float64 (scoring, weights, moments, slices): same arguments and rows, `partial` is caller scratch of
 * srb_pool_stats_f64_partial_doubles(G, k, D) doubles; k in {1, 2, 4}; the double parameter tables must fit shared memory
 * (srb_pool_stats_f64_supported; 55 x 2 x 39 and 55 x 4 x 39 do).  For the iterated loops of the batched HMM.fit:
 * kmeans.py:178-193 is a discrete map run to bitwise equality and GMM.em stops on np.allclose at 1e-8 / 1e-5; with the
 * reference's own precision the trajectories, fixed points and stop iterations are the reference's (X is float32 and
 * converts exactly), where FP32 rounding sends K-means on 10^4..10^5-frame pools to a different nearby fixed point. */
int srb_pool_stats_f64_supported(int G, int k, int D);
int64_t srb_pool_stats_f64_partial_doubles(int G, int k, int D);
int srb_pool_stats_f64(const float *X, int64_t N, int D, const int32_t *state, int G, int M, int k, int mode,
                       const double *center, const double *var, const double *lc, const double *shift, double *partial,
                       double *out, int32_t *argmin_out, int accumulate, void *stream);

/* Per-iteration updates of the batched HMM.fit, one CTA per pool, float64 without fused multiply-adds (the reference's
 * numpy arithmetic operation by operation, so a pool stops at the iteration the reference's loop would):
 *   srb_fit_em_update      GMM.em's M-step + stop test (hmm_state.py:138-159) on the first k components of every ACTIVE
 *                          state from stats double[G*M, row]: mu += S1/S0, var = S2/S0 - (S1/S0)^2, w = S0/count
 *                          (components with S0 == 0 stay); a state whose new parameters satisfy np.allclose (atol 1e-8,
 *                          rtol 1e-5) against *_old is frozen (active[g] = 0), else *_old <- new.  its[g] = it for the
 *                          states that were active.  lc double[G*M] <- srb_gauss_consts of the updated model.
 *   srb_fit_kmeans_update  kmeans.py:187-192 from stats double[G*k, row]: new centroid = shift + S1/S0; an active pool
 *                          whose new centroids equal cen exactly (np.array_equal) is frozen, else cen <- new;
 *                          cnt double[G*k] <- cluster sizes of the pools that were active. */
int srb_fit_em_update(const double *stats, int G, int M, int k, int D, double *mu, double *var, double *w, double *mu_old,
                      double *var_old, double *w_old, uint8_t *active, int64_t *its, int64_t it, double *lc,
                      void *stream);
int srb_fit_kmeans_update(const double *stats, int G, int k, int D, const double *shift, double *cen, uint8_t *active,
                          int64_t *its, int64_t it, double *cnt, void *stream);

/* M-step of GMM.em (hmm_state.py:138-148) in place on mu/var/w double (first ncomp components per state):
 * mu += S1/S0, var = S2/S0 - (S1/S0)^2 (floored at var_floor), w = S0/count. */
int srb_mstep(const double *stats, int R, int M, int ncomp, int D, double *mu, double *var, double *w,
              double var_floor, void *stream);

/* Baum-Welch transition re-estimation on the device (no reference counterpart): for every origin state o
 * of the packed graphs, a[o -> r] = xi[o -> r] / sum_r' xi[o -> r'] and cost = -log a is written IN PLACE into the arc's
 * cost in the predecessor CSR (pred_cost, indexed like xi), the successor CSR (succ_cost) and the band array of
 * srb_forward_backward_banded (band_slot[e] = index into band of edge e, -1 = none).  succ_ptr int32[n_states+1] is the
 * successor CSR over ALL packed states (absolute offsets), succ_edge[j] = predecessor-CSR index of successor entry j.
 * Arcs with xi == 0 and origins without mass keep their cost.  Deterministic (one thread per origin, fixed order). */
int srb_update_transitions(const double *xi, const int32_t *succ_ptr, const int32_t *succ_edge, const int32_t *band_slot,
                           int32_t n_states, float *pred_cost, float *succ_cost, float *band, void *stream);

/* lc[g] = log w_g - 0.5*(D log 2pi + sum_d log var[g,d])  (natural log) */
int srb_gauss_consts(const double *var, const double *w, int G, int D, double *lc, void *stream);

/* ---- fused isolated-word training pass (BASELINE config 2) ----------------------------------------
 * One kernel per iteration replaces align_gmm_states (kmeans.py:196-205: decode_hmm_states per utterance,
 * get_segments_from_path, segment_data) followed by the E-step of GMM.em on every state's pool
 * (hmm_state.py:122-136): each utterance is scored against the S states of ITS word (packed model, word-major,
 * R = W*S), Viterbi-aligned (left-to-right transitions with skips of at most 2 states, kmeans.py:53-95; start in
 * state 0, end in state S-1 at the last frame), and the posterior-weighted moments of the first ncomp components
 * are accumulated.  X is read from HBM once.
 *   srb_em_plan   (HOST arrays) groups consecutive utterances of one word into tiles of *tile_frames frames
 *                 (128 when no utterance is longer, else srb_em_tile_frames() = 256; pass it on to srb_em_isolated) and writes one 16-byte descriptor per tile, int32 {first utterance, first frame, frames,
 *                 word << 8 | utterances} (capacity 4*B ints; copy it to 16-byte aligned device memory).
 *                 utt_word must be non-decreasing; an utterance longer than a tile gives SRB_EUNSUPPORTED (use
 *                 srb_gmm_score_rows + srb_viterbi_graph + srb_gmm_stats instead).
 *   srb_em_isolated  partial: caller scratch of srb_em_partial_floats(W, S, M, D) floats (srb_em_parts() private slices of
 *                 W*S*M * srb_stats_row(D) floats, one per CTA, followed by one 64-bit word mask per slice: a CTA WRITES
 *                 the rows of the words it saw and sets their bits, nothing is zero-filled; W <= 64), or NULL for an
 *                 alignment-only call; stats_out double[W*S*M, row] as srb_gmm_stats (accumulate < 0: the per-CTA
 *                 slices stay in `partial` for srb_em_push, stats_out may be NULL); optional outputs
 *                 best_cost[B] (costs[-1,-1] of hmm.py:135, +inf when no path), seg_start int32[B,S] (segment start
 *                 frames = seg_starts of kmeans.py:200-204), frame_label int32[N] (w*S+s, -1 without a path). */
int srb_em_tile_frames(void);
int srb_em_parts(void);
int64_t srb_em_partial_floats(int W, int S, int M, int D);
int srb_em_plan(const int32_t *utt_off_host, const int32_t *utt_word_host, int32_t B, int32_t *tile_desc_host,
                int32_t *ntiles, int32_t *tile_frames);
int srb_em_isolated(const float *X, int64_t N, int D, const int32_t *utt_off, int32_t B, const int32_t *tile_desc,
                    int32_t ntiles, int32_t tile_frames, const float *ab, const float *c, const float *shift,
                    int W, int S, int M, int Mp, const float *trans, int ncomp, const double *mu, float *partial,
                    double *stats_out, int accumulate, float *best_cost, int32_t *seg_start, int32_t *frame_label,
                    void *stream);

/* ---- multi-GPU tail of the fused training pass: statistics exchange over NVLink peer memory + replicated M-step -------
 * One process per GPU; the only coupling of the ranks is the sum of the sufficient statistics (kmeans.py:41-50 pools the
 * frames of ALL utterances per state).  Instead of reduce + ncclAllReduce + M-step + two packing launches:
 *   srb_peer_alloc / _export / _open / _close / _free   a device buffer other processes of the box can map (CUDA IPC):
 *                   the owner exports a 64-byte handle, the peers open it (the host layer ships the handles)
 *   srb_em_sync_bytes(world, len)   size of a rank's exchange buffer for statistics of `len` doubles (W*S*M*(2D+3))
 *   srb_em_push     sums the per-CTA slices srb_em_isolated left in `partial` (call it with accumulate = -1) in a fixed
 *                   order, adds 16 partial sums of best_cost and the utterance count, and stores these doubles into slot
 *                   [epoch & 1][rank] of EVERY rank's buffer (peer_bufs: HOST array of `world` device pointers, entry
 *                   `rank` the own buffer), then releases flag[rank] = epoch on every peer.  epoch = 1, 2, 3, ...
 *   srb_em_update   waits until all `world` flags of the own buffer reach `epoch`, adds the slots IN RANK ORDER (all
 *                   ranks get bit-identical totals) into stats_out[len+2], applies the M-step of GMM.em
 *                   (hmm_state.py:138-148, as srb_mstep) to mu / var / w and re-packs ab / c (srb_pack_model) and, when
 *                   tcb != NULL, the tensor-core operand (srb_pack_model_tc) -- one CTA per Gaussian, D = 39.
 * world <= 8.  With world == 1 the pair is the single-GPU tail (no peers involved). */
int srb_peer_alloc(int64_t bytes, void **ptr);
int srb_peer_free(void *ptr);
int srb_peer_export(void *ptr, uint8_t *handle64);
int srb_peer_open(const uint8_t *handle64, void **ptr);
int srb_peer_close(void *ptr);
int64_t srb_em_sync_bytes(int world, int64_t len);
int srb_em_push(const float *partial, int32_t nparts, int W, int S, int M, int ncomp, int D, const float *best_cost,
                int32_t B, void *const *peer_bufs, int world, int rank, uint64_t epoch, void *stream);
int srb_em_update(void *own_buf, int world, uint64_t epoch, int R, int M, int Mp, int ncomp, int D, double *mu, double *var,
                  double *w, const double *shift, double var_floor, float *ab, float *c, float *tcb, double *stats_out,
                  void *stream);

/* ---- FP32 FMA peak probe (roofline denominator for K1; not part of the reference path) ----------
 * Runs `iters` dependent-chain-free FMA rounds per thread on every SM and returns elapsed ms through
 * *ms_out (host pointer) plus the FMA count (host pointer).  packed != 0 uses fma.rn.f32x2. */
int srb_fma_peak(int packed, int iters, float *ms_out, double *fma_count_out, void *stream);

/* ---- MFCC front-end: the step right before the scoring path ----------------------------------------
 * Replaces sr/feature/feature.py:7-88 and core.py:13-31 for a batch of 16 kHz utterances.
 *   srb_mfcc_num_frames(n, sr)   frames of an n-sample utterance: ceil(n / 160) (feature.py:7-16; 0 for sr != 16000)
 *   srb_mfcc                     signal float[sum n] (the int16 samples as floats), sig_off int64[B+1], frame_off
 *                                int64[B+1] (prefix sums of srb_mfcc_num_frames), F = frame_off[B];
 *                                fb_w float[40,257] + fb_range int32[40,2] = the mel filter bank of feature.py:57-76
 *                                built by the host (dense weights, [first, last) non-zero bin per filter);
 *                                fbank float[F,40] (optional, log10 energies), mfcc float[F,13] (DCT-II ortho, c 1..13)
 *   srb_delta_standardize        mfcc float[F,13] -> features float[F,39] = [c, delta, delta-delta] (core.py:13-22),
 *                                standardised per utterance and dimension when standardize != 0 (feature.py:85-88);
 *                                utterances need >= 2 frames (the reference indexes feat[1]) */
int64_t srb_mfcc_num_frames(int64_t nsamples, int sample_rate);
int srb_mfcc(const float *signal, const int64_t *sig_off, const int64_t *frame_off, int32_t B, int64_t F, int sample_rate,
             const float *fb_w, const int32_t *fb_range, float *fbank, float *mfcc, void *stream);
int srb_delta_standardize(const float *mfcc, const int64_t *frame_off, int32_t B, float *features, int standardize,
                          void *stream);

#ifdef __cplusplus
}
#endif
#endif /* SRB200_H */
