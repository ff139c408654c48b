/*
 * sr_oracle.c -- TEST INFRASTRUCTURE ONLY (CPU float64 restatement of the reference hot path).
 *
 * This file is the parity oracle for the CUDA kernels in speech-recognition_b200/csrc.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it.  The product
 * never links or calls it.
 *
 * Every function restates, loop for loop and in IEEE float64 like numpy, one function of the reference
 * (paths relative to /root/reference).  Parity status: PINNED -- tests/test_oracle_golden.py checks each
 * function against tests/golden/ref_*.npz, which tests/golden/make_golden.py produced by executing the
 * unmodified reference.  Exception: orc_forward_backward has no reference counterpart ("parity
 * unpinned"); it is tied to orc_decode by the max-product identity instead.
 *
 * Build: gcc -O2 -fPIC -shared -fopenmp -o oracle/_build/libsroracle.so oracle/sr_oracle.c -lm
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_INF (INFINITY)

/* ---- hmm_state.py:36-43  MultivariateNormal.pdf (linear domain, diagonal covariance).
 * inv_cov is inv(diag(cov)) (hmm_state.py:13-17), i.e. 1/cov on the diagonal; the quadratic form
 * x_mu.dot(inv_cov).dot(x_mu) therefore reduces to sum_d (x_mu_d * (1/cov_d)) * x_mu_d. */
static double pdf_linear(const double *x, const double *mean, const double *cov, int D) {
    double det = 1.0;
    for (int d = 0; d < D; ++d) det *= cov[d];                               /* np.prod(self.cov) :39 */
    double norm_const = 1.0 / (pow(2.0 * M_PI, (double)D / 2.0) * sqrt(det)); /* :40 */
    double q = 0.0;
    for (int d = 0; d < D; ++d) {
        double xm = x[d] - mean[d];                                          /* :41 */
        q += (xm * (1.0 / cov[d])) * xm;                                     /* :42 */
    }
    return norm_const * exp(-0.5 * q);                                       /* :42-43 */
}

/* ---- log-domain restatement of the same density (no underflow); equals log(pdf_linear) to ~1e-13
 * wherever pdf_linear is a normal number. */
static double pdf_log(const double *x, const double *mean, const double *cov, int D) {
    double logdet = 0.0, q = 0.0;
    for (int d = 0; d < D; ++d) {
        logdet += log(cov[d]);
        double xm = x[d] - mean[d];
        q += xm * xm / cov[d];
    }
    return -0.5 * ((double)D * log(2.0 * M_PI) + logdet) - 0.5 * q;
}

/* ---- hmm_state.py:114-120  GMM.evaluate for N frames.
 * comps_out[N,M] (optional) = w_m * pdf_m(x)  (return_neg_log_likelihood=False)
 * cost_out[N]   (optional) = -log(sum_m ...)  (inf when every component underflowed, A7)
 * log_domain != 0 switches to the log-sum-exp restatement (finite everywhere). */
void orc_gmm_evaluate(const double *X, int64_t N, int D, const double *mu, const double *cov, const double *w,
                      int M, int log_domain, double *cost_out, double *comps_out) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < N; ++i) {
        const double *x = X + i * D;
        if (!log_domain) {
            double s = 0.0;
            for (int m = 0; m < M; ++m) {
                double v = pdf_linear(x, mu + (size_t)m * D, cov + (size_t)m * D, D) * w[m]; /* :115 */
                if (comps_out) comps_out[i * M + m] = v;
                s += v;
            }
            if (cost_out) cost_out[i] = -log(s);                                               /* :118 */
        } else {
            double l[1024], mx = -ORC_INF;                       /* oracle.py checks M <= 1024 */
            for (int m = 0; m < M; ++m) {
                l[m] = log(w[m]) + pdf_log(x, mu + (size_t)m * D, cov + (size_t)m * D, D);
                if (l[m] > mx) mx = l[m];
                if (comps_out) comps_out[i * M + m] = l[m];      /* log of the :115 value */
            }
            double s = 0.0;
            for (int m = 0; m < M; ++m) s += exp(l[m] - mx);
            if (cost_out) cost_out[i] = -(mx + log(s));
        }
    }
}

/* ---- hmm_state.py:48-58  mahalanobis(v1, v2, variance) */
double orc_mahalanobis(const double *v1, const double *v2, const double *var, int D) {
    double prod = 1.0, q = 0.0;
    for (int d = 0; d < D; ++d) prod *= var[d];
    for (int d = 0; d < D; ++d) {
        double m = v1[d] - v2[d];
        q += m / var[d] * m;
    }
    return 0.5 * log(pow(2.0 * M_PI, (double)D) * prod) + 0.5 * q;
}

/* local distances for dtw: kind 0 = np.linalg.norm(x - y) (kmeans.py:111 default dist_fun),
 * kind 1 = mahalanobis(x, y, var[s]) (hmm.py:134).  out[T,S]. */
void orc_local_dist(const double *X, int64_t T, int D, const double *Y, const double *var, int S, int kind,
                    double *out) {
    for (int64_t t = 0; t < T; ++t)
        for (int s = 0; s < S; ++s) {
            if (kind == 0) {
                double q = 0.0;
                for (int d = 0; d < D; ++d) {
                    double m = X[t * D + d] - Y[s * D + d];
                    q += m * m;
                }
                out[t * S + s] = sqrt(q);
            } else {
                out[t * S + s] = orc_mahalanobis(X + t * D, Y + s * D, var + s * D, D);
            }
        }
}

/* ---- decode.py:80-146  decode_hmm_states.
 * E[T,R]   emission cost of state r at frame c (states[r].evaluate(x[c]); 0 for NES, hmm_state.py:90-91)
 * trans[R,R] trans[r,o] = cost of o -> r, inf = none (decode.py:86-87)
 * ends[n_ends,2] = (state, frame) candidates, frame may be -1 (main.py:60)
 * outputs: costs[R,T]; pm[R,T,2] back-pointers (INT64_MIN where untouched, like np.full(inf, dtype=int));
 *          path[(<=max_path),2] reversed (decode.py:142-146); returns path length, or -1 if the
 *          origin_point == [r,c] guard (decode.py:120-121) fired, -2 if path buffer too small. */
int64_t orc_decode(const double *E, int64_t T, int R, const double *trans, const uint8_t *is_nes,
                   const int64_t *ends, int n_ends, double *costs, int64_t *pm, int64_t *path, int64_t max_path,
                   int64_t *best_end_out) {
    for (int64_t i = 0; i < (int64_t)R * T; ++i) costs[i] = ORC_INF;
    for (int64_t i = 0; i < (int64_t)R * T * 2; ++i) pm[i] = INT64_MIN;
#define COST(r, c) costs[(int64_t)(r) * T + (((c) % T + T) % T)] /* python negative indexing: c-1 == -1 at c=0 */
    for (int64_t c = 0; c < T; ++c) {
        for (int r = 0; r < R; ++r) {
            if (r == 0 && c == 0) {
                COST(0, 0) = E[0];                                          /* :99-100 */
                continue;
            }
            double best = 0.0;
            int64_t bo = -1, bc = 0;
            int n = 0;
            for (int o = 0; o < R; ++o) {
                double tr = trans[(int64_t)r * R + o];
                if (isinf(tr)) continue;                                    /* :106 */
                double pc;
                int64_t fc;
                if (is_nes[o] || is_nes[r]) {                               /* :109-111 same column */
                    pc = tr + COST(o, c);
                    fc = c;
                } else {
                    pc = tr + COST(o, c - 1);                               /* :113 */
                    fc = c - 1;
                }
                if (n == 0 || pc < best) {                                  /* np.argmin: first minimum :118 */
                    best = pc;
                    bo = o;
                    bc = fc;
                }
                ++n;
            }
            if (n == 0) continue;                                           /* :116-117 */
            if (bo == r && bc == c) return -1;                              /* :120-121 */
            pm[((int64_t)r * T + c) * 2] = bo;
            pm[((int64_t)r * T + c) * 2 + 1] = bc;                          /* may be -1 at c == 0 */
            double cur = best + E[c * R + r];                               /* :123 */
            if (cur < COST(r, c)) COST(r, c) = cur;                         /* :124 (min; NaN-free inputs) */
        }
    }
    double best_cost = ORC_INF;
    int64_t bi = 0, bj = 0;
    int have = 0;
    for (int e = 0; e < n_ends; ++e) {                                      /* :131-134, '>=' -> last wins */
        int64_t er = ends[2 * e], ec = ends[2 * e + 1];
        if (best_cost >= COST(er, ec)) {
            best_cost = COST(er, ec);
            bi = er;
            bj = ec;
            have = 1;
        }
    }
    if (!have) return -3;
    if (best_end_out) {
        best_end_out[0] = bi;
        best_end_out[1] = bj;
    }
    int64_t len = 0, i = bi, j = bj;
    while (j != 0) {                                                        /* :143-145 */
        int64_t jj = ((j % T) + T) % T;
        int64_t ni = pm[((int64_t)i * T + jj) * 2], nj = pm[((int64_t)i * T + jj) * 2 + 1];
        if (ni == INT64_MIN) return -4;                                     /* walked into an untouched cell */
        i = ni;
        j = nj;
        if (len >= max_path) return -2;
        path[2 * len] = i;
        path[2 * len + 1] = j;
        ++len;
    }
#undef COST
    return len;
}

/* ---- decode.py:7-77  dtw (beam = inf, the only value any caller uses; A15).
 * Dl[T,S] local distance of frame j to template row i; trans[S,S].
 * outputs costs[S,T], path reversed [[i,j],...] down to (0,0) exclusive of the end cell. */
int64_t orc_dtw(const double *Dl, int64_t T, int S, const double *trans, double *costs, int64_t *path,
                int64_t max_path) {
    if (!(T > 1 && S > 1)) return -5;                                       /* assert :22 */
    int64_t *pm = (int64_t *)malloc(sizeof(int64_t) * S * T * 2);
    for (int64_t i = 0; i < (int64_t)S * T; ++i) costs[i] = ORC_INF;
    for (int64_t i = 0; i < (int64_t)S * T * 2; ++i) pm[i] = INT64_MIN;
#define COST(r, c) costs[(int64_t)(r) * T + (((c) % T + T) % T)]
    for (int64_t j = 0; j < T; ++j) {
        for (int i = 0; i < S; ++i) {
            if (i == 0 && j == 0) {
                COST(0, 0) = Dl[0];                                         /* :33-40 */
                continue;
            }
            double best = 0.0;
            int bo = 0;
            for (int o = 0; o < S; ++o) {                                   /* :44-51 (no inf skip here) */
                double pc = trans[i * S + o] + COST(o, j - 1);
                if (o == 0 || pc < best) {
                    best = pc;
                    bo = o;
                }
            }
            pm[((int64_t)i * T + j) * 2] = bo;
            pm[((int64_t)i * T + j) * 2 + 1] = j - 1;
            double cur = best + Dl[j * S + i];
            if (cur < COST(i, j)) COST(i, j) = cur;                         /* :60 */
        }
    }
    int64_t len = 0, i = S - 1, j = T - 1;
    while (i != 0 || j != 0) {                                              /* :74-76 */
        if (j < 0 || len >= max_path) {
            free(pm);
            return -2;                                                      /* no finite alignment */
        }
        int64_t ni = pm[((int64_t)i * T + j) * 2], nj = pm[((int64_t)i * T + j) * 2 + 1];
        i = ni;
        j = nj;
        path[2 * len] = i;
        path[2 * len + 1] = j;
        ++len;
    }
#undef COST
    free(pm);
    return len;
}

/* ---- hmm_state.py:124-148  one E+M pass of GMM.em over the first n components.
 * returns new mu[n,D], sigma[n,D], w[n].  Linear-domain posteriors incl. the 1e-5 zero guards (A7). */
void orc_em_iteration(const double *X, int64_t N, int D, const double *mu, const double *cov, const double *w, int M,
                      int n, double *mu_out, double *sigma_out, double *w_out) {
    double *p = (double *)malloc(sizeof(double) * N * n);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < N; ++i) {
        double s = 0.0;
        for (int c = 0; c < n; ++c) {
            p[i * n + c] = pdf_linear(X + i * D, mu + (size_t)c * D, cov + (size_t)c * D, D) * w[c]; /* :128 */
            s += p[i * n + c];
        }
        if (s == 0.0) s = 1e-5;                                             /* :132 */
        for (int c = 0; c < n; ++c) p[i * n + c] /= s;                      /* :133 */
    }
    (void)M;
    for (int c = 0; c < n; ++c) {
        double ps = 0.0;
        for (int64_t i = 0; i < N; ++i) ps += p[i * n + c];                 /* :134 */
        double psg = (ps == 0.0) ? 1e-5 : ps;                               /* :136 */
        for (int d = 0; d < D; ++d) {
            double a = 0.0;
            for (int64_t i = 0; i < N; ++i) a += X[i * D + d] * p[i * n + c];       /* :138 */
            mu_out[c * D + d] = a / psg;                                            /* :139 */
        }
        for (int d = 0; d < D; ++d) {
            double a = 0.0, m = mu_out[c * D + d];
            for (int64_t i = 0; i < N; ++i) {
                double t = X[i * D + d] - m;                                        /* :141 new mean */
                a += t * t * p[i * n + c];                                          /* :142 */
            }
            sigma_out[c * D + d] = a / psg;                                         /* :143 */
        }
        w_out[c] = ps / (double)N;                                                  /* :148 p.mean(axis=0) */
    }
    free(p);
}

/* ---- kmeans.py:180-187  one assignment + centroid pass.
 * dists[i,c] = mahalanobis(centroids[c], data[i], cov0) (cov[0] for every centroid, A3);
 * clusters = argmin (first minimum); new centroids = cluster means (NaN for an empty cluster). */
void orc_kmeans_iteration(const double *X, int64_t N, int D, const double *cent, int k, const double *cov0,
                          int64_t *clusters, double *new_cent) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < N; ++i) {
        double best = 0.0;
        int bc = 0;
        for (int c = 0; c < k; ++c) {
            double dd = orc_mahalanobis(cent + (size_t)c * D, X + i * D, cov0, D);
            if (c == 0 || dd < best) {
                best = dd;
                bc = c;
            }
        }
        clusters[i] = bc;
    }
    for (int c = 0; c < k; ++c) {
        int64_t cnt = 0;
        for (int d = 0; d < D; ++d) new_cent[c * D + d] = 0.0;
        for (int64_t i = 0; i < N; ++i)
            if (clusters[i] == c) {
                ++cnt;
                for (int d = 0; d < D; ++d) new_cent[c * D + d] += X[i * D + d];
            }
        for (int d = 0; d < D; ++d) new_cent[c * D + d] = cnt ? new_cent[c * D + d] / (double)cnt : NAN;
    }
}

/* ---- kmeans.py:6-12  calc_variance: np.cov(data).diagonal() = unbiased variance of each row of a
 * [D,N] matrix; here X is [N,D] (the callers pass seg.T). */
void orc_calc_variance(const double *X, int64_t N, int D, double *out) {
    for (int d = 0; d < D; ++d) {
        double m = 0.0;
        for (int64_t i = 0; i < N; ++i) m += X[i * D + d];
        m /= (double)N;
        double a = 0.0;
        for (int64_t i = 0; i < N; ++i) {
            double t = X[i * D + d] - m;
            a += t * t;
        }
        out[d] = a / (double)(N - 1);
    }
}

/* ---- isolated-word classifier: core.py:79-87 over hmm.py:126-135 (cost = costs[-1,-1]).
 * X ragged [total,D] with off[B+1]; model arrays [W,S,M,D] / [W,S,M] / [W,S,S].
 * final[B,W] and label[B] (strict '<' -> first wins).  Linear-domain emissions like the reference.
 * This is also the CPU baseline for the decode workload (OpenMP over utterances). */
void orc_classify(const double *X, const int64_t *off, int64_t B, int D, const double *mu, const double *cov,
                  const double *w, const double *trans, int W, int S, int M, int log_domain, double *final,
                  int64_t *label) {
#pragma omp parallel for schedule(dynamic, 1)
    for (int64_t u = 0; u < B; ++u) {
        int64_t T = off[u + 1] - off[u];
        double *E = (double *)malloc(sizeof(double) * T * S);
        double *costs = (double *)malloc(sizeof(double) * T * S);
        int64_t *pm = (int64_t *)malloc(sizeof(int64_t) * T * S * 2);
        int64_t *path = (int64_t *)malloc(sizeof(int64_t) * (T + 1) * 2 * 2);
        uint8_t nes[64] = {0};
        double bc = ORC_INF;
        int64_t bl = 0;
        for (int wd = 0; wd < W; ++wd) {
            for (int s = 0; s < S; ++s) {
                size_t g = ((size_t)wd * S + s) * M;
                double *col = (double *)malloc(sizeof(double) * T);
                /* single-threaded inner call: we are already inside a parallel region */
                for (int64_t t = 0; t < T; ++t) {
                    const double *x = X + (off[u] + t) * D;
                    if (!log_domain) {
                        double sm = 0.0;
                        for (int m = 0; m < M; ++m) sm += pdf_linear(x, mu + (g + m) * D, cov + (g + m) * D, D) * w[g + m];
                        col[t] = -log(sm);
                    } else {
                        double l[64], mx = -ORC_INF, sm = 0.0;
                        for (int m = 0; m < M; ++m) {
                            l[m] = log(w[g + m]) + pdf_log(x, mu + (g + m) * D, cov + (g + m) * D, D);
                            if (l[m] > mx) mx = l[m];
                        }
                        for (int m = 0; m < M; ++m) sm += exp(l[m] - mx);
                        col[t] = -(mx + log(sm));
                    }
                }
                for (int64_t t = 0; t < T; ++t) E[t * S + s] = col[t];
                free(col);
            }
            int64_t ends[2] = {S - 1, T - 1};
            orc_decode(E, T, S, trans + (size_t)wd * S * S, nes, ends, 1, costs, pm, path, (T + 1) * 2, NULL);
            double c = costs[(int64_t)(S - 1) * T + (T - 1)];                /* hmm.py:135 */
            final[u * W + wd] = c;
            if (c < bc) {                                                   /* core.py:85 */
                bc = c;
                bl = wd;
            }
        }
        label[u] = bl;
        free(E);
        free(costs);
        free(pm);
        free(path);
    }
}

/* ---- forward-backward for a NES-free model (NO reference counterpart: "parity unpinned").
 * Same conventions and boundary conditions as orc_decode: trans[r,o] = cost o->r, start only in state 0 at
 * frame 0 (decode.py:99-100), end in state R-1 at frame T-1 (decode.py:126-127).
 * alpha/beta are NEGATIVE log (costs), gamma[T,R] posterior, returns -log Z.  Replacing the
 * log-sum-exp with min reproduces orc_decode's costs exactly (tested). */
static double neg_lse2(double a, double b) { /* -log(exp(-a)+exp(-b)) */
    if (isinf(a) && a > 0) return b;
    if (isinf(b) && b > 0) return a;
    double m = a < b ? a : b;
    return m - log(exp(-(a - m)) + exp(-(b - m)));
}

static double neg_min2(double a, double b) { return a < b ? a : b; }

/* semiring != 0: (min, +) instead of (-log-sum-exp, +), i.e. the SAME recursion with every sum over paths replaced
 * by the best path -- alpha then equals orc_decode's cost matrix bit for bit (same additions in the same order),
 * which ties this restatement to decode_hmm_states (decode.py:97-124). */
double orc_forward_backward_sr(const double *E, int64_t T, int R, const double *trans, double *alpha, double *beta,
                               double *gamma, int semiring) {
    double (*comb)(double, double) = semiring ? neg_min2 : neg_lse2;
    for (int64_t i = 0; i < T * R; ++i) alpha[i] = beta[i] = ORC_INF;
    alpha[0] = E[0];
    for (int64_t c = 1; c < T; ++c)
        for (int r = 0; r < R; ++r) {
            double a = ORC_INF;
            for (int o = 0; o < R; ++o) {
                double tr = trans[(int64_t)r * R + o];
                if (isinf(tr)) continue;
                a = comb(a, tr + alpha[(c - 1) * R + o]);
            }
            alpha[c * R + r] = a + E[c * R + r];
        }
    beta[(T - 1) * R + (R - 1)] = 0.0;
    for (int64_t c = T - 2; c >= 0; --c)
        for (int o = 0; o < R; ++o) {
            double b = ORC_INF;
            for (int r = 0; r < R; ++r) {
                double tr = trans[(int64_t)r * R + o];
                if (isinf(tr)) continue;
                b = comb(b, tr + E[(c + 1) * R + r] + beta[(c + 1) * R + r]);
            }
            beta[c * R + o] = b;
        }
    double nlz = alpha[(T - 1) * R + (R - 1)];
    if (gamma)
        for (int64_t i = 0; i < T * R; ++i) gamma[i] = exp(-(alpha[i] + beta[i] - nlz));
    return nlz;
}

double orc_forward_backward(const double *E, int64_t T, int R, const double *trans, double *alpha, double *beta,
                            double *gamma) {
    return orc_forward_backward_sr(E, T, R, trans, alpha, beta, gamma, 0);
}


/* ---- align_gmm_states (kmeans.py:196-205) for a labelled batch: every utterance is Viterbi-aligned to the S
 * states of its own word (decode_hmm_states with the word's GMM states), OpenMP over utterances.
 * labels[N] = tied state (word*S + state) of every frame, cost[B] = costs[-1,-1].  This is the alignment half of
 * the CPU baseline of the EM workload; the E/M half is orc_em_iteration per tied state. */
void orc_align_labels(const double *X, const int64_t *off, int64_t B, int D, const int64_t *word, const double *mu,
                      const double *cov, const double *w, const double *trans, int S, int M, int log_domain,
                      int64_t *labels, double *cost) {
#pragma omp parallel for schedule(dynamic, 4)
    for (int64_t u = 0; u < B; ++u) {
        int64_t T = off[u + 1] - off[u];
        int wd = (int)word[u];
        double *E = (double *)malloc(sizeof(double) * T * S);
        double *costs = (double *)malloc(sizeof(double) * T * S);
        int64_t *pm = (int64_t *)malloc(sizeof(int64_t) * T * S * 2);
        int64_t *path = (int64_t *)malloc(sizeof(int64_t) * (T + 2) * 2);
        double *col = (double *)malloc(sizeof(double) * T);
        uint8_t nes[64] = {0};
        for (int s = 0; s < S; ++s) {
            size_t g = ((size_t)wd * S + s) * M;
            orc_gmm_evaluate(X + off[u] * D, T, D, mu + g * D, cov + g * D, w + g, M, log_domain, col, NULL);
            for (int64_t t = 0; t < T; ++t) E[t * S + s] = col[t];
        }
        int64_t ends[2] = {S - 1, T - 1};
        int64_t n = orc_decode(E, T, S, trans + (size_t)wd * S * S, nes, ends, 1, costs, pm, path, T + 2, NULL);
        cost[u] = costs[(int64_t)(S - 1) * T + (T - 1)];
        int64_t *lab = labels + off[u];
        for (int64_t t = 0; t < T; ++t) lab[t] = -1;
        if (n >= 0 && !isinf(cost[u])) {
            lab[T - 1] = (int64_t)wd * S + (S - 1);
            for (int64_t k = 0; k < n; ++k) lab[path[2 * k + 1]] = (int64_t)wd * S + path[2 * k];
        }
        free(E);
        free(costs);
        free(pm);
        free(path);
        free(col);
    }
}

int orc_num_threads(void) {
#ifdef _OPENMP
    extern int omp_get_max_threads(void);
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* thread count of the OpenMP loops (torchrun exports OMP_NUM_THREADS=1; the CPU baseline wants all host cores) */
void orc_set_num_threads(int n) {
#ifdef _OPENMP
    extern void omp_set_num_threads(int);
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}
