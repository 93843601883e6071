/*
 * oracle/ctc_oracle_impl.h -- TEST INFRASTRUCTURE ONLY (never linked or imported by the product).
 *
 * CPU restatement of the CTC loss forward+backward that the reference applies at
 *   /root/reference/train.py:600 (train), :352 and :457 (eval)  via  `from warpctc_pytorch import CTCLoss`
 *   (/root/reference/train.py:12, ctor :904).
 *
 * PARITY UNPINNED: the arithmetic of this path lives in the third-party package `warpctc_pytorch`
 * (SeanNaren/warp-ctc binding over baidu-research/warp-ctc), which the reference neither vendors nor pins
 * (no requirements.txt / setup.py / submodule) and which is absent from this image.  The reference holds no
 * tests, golden vectors or fixtures for the path.  This file therefore restates the
 * published algorithm (Graves et al. 2006, as written out in SURVEY.md section 8(c)) against the reference's
 * call contract, and is pinned against (i) closed-form enumeration over all C^T alignments on tiny cases and
 * (ii) torch 2.11 CPU fp64 F.ctc_loss(log_softmax(.)) fixtures under tests/golden/ (generator committed).
 *
 * Contract restated:
 *   acts    fp32 [T,B,C] unnormalised logits, row (t,b,:) contiguous      (train.py:593 "TxNxH")
 *   labels  int32 flat concatenation of the B targets, values 1..C-1      (data/data_loader_aug.py:546-547)
 *   act_lens, label_lens int32 [B]                                        (train.py:600, data_loader_aug.py:545)
 *   blank = 0 (labels.json "_" at index 0), reduction = per-utterance costs (caller sums; train.py:601 divides by B)
 *   grads = d(sum_b cost_b)/d acts, zero for t >= act_lens[b]
 *   infeasible utterance (L + repeats > T): cost = 0 under zero_infinity, +inf otherwise; its grads are 0 either way.
 *
 * The file is included twice (REAL = double -> golden oracle, REAL = float -> timed "reference-class" baseline,
 * fp32 log-space like warp-ctc's CPU path [external, from memory]).  OpenMP parallel-for over the minibatch.
 */

#ifndef REAL
#error "define REAL, SUFFIX, R_EXP, R_LOG, R_LOG1P before including"
#endif

#define CAT_(a, b) a##b
#define CAT(a, b) CAT_(a, b)
#define FN(name) CAT(name, SUFFIX)

static inline REAL FN(lse2_)(REAL a, REAL b) {
    if (a == -INFINITY) return b;
    if (b == -INFINITY) return a;
    REAL m = a > b ? a : b;
    REAL d = a > b ? b - a : a - b;
    return m + R_LOG1P(R_EXP(d));
}

/* one utterance; returns cost (= -log-likelihood); writes grad rows [0,Tb) if grads != NULL */
static double FN(ctc_one_)(const float *acts, size_t row_stride /* = B*C */, int Tb, int C, const int *lab, int L,
                           int blank, REAL *grads_out /* strided like acts, may be NULL */, int *feasible) {
    const int S = 2 * L + 1;
    *feasible = 1;
    if (Tb <= 0) {
        if (L == 0) return 0.0;
        *feasible = 0;
        return INFINITY;
    }
    REAL *lp = (REAL *)malloc(sizeof(REAL) * (size_t)Tb * C);
    REAL *alpha = (REAL *)malloc(sizeof(REAL) * (size_t)Tb * S);
    REAL *beta = (REAL *)malloc(sizeof(REAL) * 2 * (size_t)S);
    REAL *acc = (REAL *)malloc(sizeof(REAL) * (size_t)C);
    int *ext = (int *)malloc(sizeof(int) * (size_t)S);
    for (int s = 0; s < S; ++s) ext[s] = (s & 1) ? lab[s >> 1] : blank;

    /* log-softmax per frame */
    for (int t = 0; t < Tb; ++t) {
        const float *a = acts + (size_t)t * row_stride;
        REAL m = (REAL)a[0];
        for (int c = 1; c < C; ++c)
            if ((REAL)a[c] > m) m = (REAL)a[c];
        REAL z = 0;
        for (int c = 0; c < C; ++c) z += R_EXP((REAL)a[c] - m);
        REAL lz = m + R_LOG(z);
        for (int c = 0; c < C; ++c) lp[(size_t)t * C + c] = (REAL)a[c] - lz;
    }

    /* alpha */
    for (int s = 0; s < S; ++s) alpha[s] = -INFINITY;
    alpha[0] = lp[ext[0]];
    if (S > 1) alpha[1] = lp[ext[1]];
    for (int t = 1; t < Tb; ++t) {
        const REAL *ap = alpha + (size_t)(t - 1) * S;
        REAL *an = alpha + (size_t)t * S;
        const REAL *l = lp + (size_t)t * C;
        for (int s = 0; s < S; ++s) {
            REAL v = ap[s];
            if (s >= 1) v = FN(lse2_)(v, ap[s - 1]);
            if (s >= 2 && ext[s] != blank && ext[s] != ext[s - 2]) v = FN(lse2_)(v, ap[s - 2]);
            an[s] = (v == -INFINITY) ? -INFINITY : v + l[ext[s]];
        }
    }
    const REAL *al = alpha + (size_t)(Tb - 1) * S;
    REAL ll = al[S - 1];
    if (S > 1) ll = FN(lse2_)(ll, al[S - 2]);
    if (ll == -INFINITY) {
        *feasible = 0;
        free(lp); free(alpha); free(beta); free(acc); free(ext);
        return INFINITY;
    }

    if (grads_out) {
        /* beta (includes the emission at t, like alpha) fused with the gradient */
        REAL *bn = beta, *bp = beta + S; /* bn: row t, bp: row t+1 */
        for (int t = Tb - 1; t >= 0; --t) {
            const REAL *l = lp + (size_t)t * C;
            if (t == Tb - 1) {
                for (int s = 0; s < S; ++s) bn[s] = -INFINITY;
                bn[S - 1] = l[ext[S - 1]];
                if (S > 1) bn[S - 2] = l[ext[S - 2]];
            } else {
                for (int s = 0; s < S; ++s) {
                    REAL v = bp[s];
                    if (s + 1 < S) v = FN(lse2_)(v, bp[s + 1]);
                    if (s + 2 < S && ext[s + 2] != blank && ext[s + 2] != ext[s]) v = FN(lse2_)(v, bp[s + 2]);
                    bn[s] = (v == -INFINITY) ? -INFINITY : v + l[ext[s]];
                }
            }
            const REAL *at = alpha + (size_t)t * S;
            for (int c = 0; c < C; ++c) acc[c] = -INFINITY;
            for (int s = 0; s < S; ++s) acc[ext[s]] = FN(lse2_)(acc[ext[s]], at[s] + bn[s]);
            REAL *g = grads_out + (size_t)t * row_stride;
            for (int c = 0; c < C; ++c) {
                REAL y = R_EXP(l[c]);
                REAL occ = (acc[c] == -INFINITY) ? (REAL)0 : R_EXP(acc[c] - l[c] - ll);
                g[c] = y - occ;
            }
            REAL *tmp = bn; bn = bp; bp = tmp;
        }
    }
    free(lp); free(alpha); free(beta); free(acc); free(ext);
    return -(double)ll;
}

/*
 * Batched entry.  grads (REAL [T,B,C]) may be NULL for score-only.  costs is double [B].
 * Returns 0 on success, 1 on invalid arguments (label out of range / equal to blank / negative length).
 */
int FN(ctc_oracle_compute_)(const float *acts, int T, int B, int C, const int *labels, const int *label_lens,
                            const int *act_lens, int blank, int zero_infinity, double *costs, REAL *grads,
                            int num_threads) {
    if (T < 0 || B < 0 || C < 1 || blank < 0 || blank >= C) return 1;
    size_t *off = (size_t *)malloc(sizeof(size_t) * (size_t)(B + 1));
    off[0] = 0;
    for (int b = 0; b < B; ++b) {
        if (label_lens[b] < 0 || act_lens[b] < 0 || act_lens[b] > T) { free(off); return 1; }
        off[b + 1] = off[b] + (size_t)label_lens[b];
    }
    for (size_t i = 0; i < off[B]; ++i)
        if (labels[i] < 0 || labels[i] >= C || labels[i] == blank) { free(off); return 1; }
    if (grads) memset(grads, 0, sizeof(REAL) * (size_t)T * B * C);
    const size_t rs = (size_t)B * C;
#ifdef _OPENMP
    if (num_threads > 0) omp_set_num_threads(num_threads);
#endif
#pragma omp parallel for schedule(dynamic, 1)
    for (int b = 0; b < B; ++b) {
        int feasible = 1;
        REAL *g = grads ? grads + (size_t)b * C : NULL;
        double cost = FN(ctc_one_)(acts + (size_t)b * C, rs, act_lens[b], C, labels + off[b], label_lens[b], blank,
                                   g, &feasible);
        if (!feasible) {
            cost = zero_infinity ? 0.0 : INFINITY;
            if (g)
                for (int t = 0; t < act_lens[b]; ++t) memset(g + (size_t)t * rs, 0, sizeof(REAL) * (size_t)C);
        }
        costs[b] = cost;
    }
    free(off);
    return 0;
}

#undef FN
#undef CAT
#undef CAT_
