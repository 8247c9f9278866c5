/*
 * ORACLE / TEST INFRASTRUCTURE ONLY -- never linked into libavenir_tree.so and never called by the
 * product.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load it.
 *
 * CPU restatement, in plain C, of the reference's tree-building arithmetic in HISTOGRAM form
 * (SURVEY.md Appendix F): the counts the reference accumulates as text records keyed by
 * (parent path, "splitId:predicate") and class value, and the impurity / weighted average / arg-min
 * it derives from them.  The literal (string-keyed) restatement lives in oracle/literal.py; the two
 * were written independently and tests/ checks that they agree.
 *
 * PARITY UNPINNED: the reference ships no tests and cannot run here (no JVM).  Pinned only by the
 * golden vectors in tests/golden/ (four gini values and the JSON shape of
 * resource/dec_tree_rules.json).
 *
 * Reference lines followed (paths relative to /root/reference/src/main/java/org/avenir/):
 *   counting          tree/DecisionTreeBuilder.java:323-352 (mapper emits per true predicate),
 *                     :730-767 (reducer counts class values per (parent, child) key)
 *   impurity          util/InfoContentStat.java:71-101
 *   weighted average  tree/DecisionTreeBuilder.java:534-548
 *   arg-min           tree/DecisionTreeBuilder.java:549-559 (strict <, first wins)
 *   CPG statistics    util/AttributeSplitStat.java:153-170, 191-218, 240-282, 298-336, 452-471
 *   log               fdlibm e_log.c (StrictMath.log), restated below
 *
 * Build: see oracle/Makefile (gcc -O3 -ffp-contract=off -fPIC -shared -pthread).
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORA_MAX_C 16
#define ORA_MAX_SEG 8
#define ORA_NODE_NONE 0xFFFFu

typedef struct ora_split { int32_t attr, kind, n_seg, off; } ora_split; /* same layout as avt_split */

typedef struct ora_best {
    double min_av, node_stat;
    int64_t population;
    int32_t best_split, n_ties;
} ora_best;

typedef struct ora_eval_t {
    double av;
    double seg_stat[ORA_MAX_SEG];
    uint32_t seg_count[ORA_MAX_SEG][ORA_MAX_C];
} ora_eval_t;

typedef struct ora_cpg_t {
    double stat, split_info;
    uint32_t seg_count[ORA_MAX_SEG][ORA_MAX_C];
} ora_cpg_t;

/* ---- fdlibm __ieee754_log ----------------------------------------------------------------- */
static double ora_log(double x) {
    static const double ln2_hi = 6.93147180369123816490e-01, ln2_lo = 1.90821492927058770002e-10,
                        two54 = 1.80143985094819840000e+16, Lg1 = 6.666666666666735130e-01,
                        Lg2 = 3.999999999940941908e-01, Lg3 = 2.857142874366239149e-01,
                        Lg4 = 2.222219843214978396e-01, Lg5 = 1.818357216161805012e-01,
                        Lg6 = 1.531383769920937332e-01, Lg7 = 1.479819860511658591e-01;
    double hfsq, f, s, z, R, w, t1, t2, dk;
    int32_t k, hx, i, j;
    uint32_t lx;
    uint64_t u;
    memcpy(&u, &x, 8);
    hx = (int32_t)(u >> 32);
    lx = (uint32_t)u;
    k = 0;
    if (hx < 0x00100000) {
        if (((hx & 0x7fffffff) | lx) == 0) return -INFINITY;
        if (hx < 0) return NAN;
        k -= 54;
        x *= two54;
        memcpy(&u, &x, 8);
        hx = (int32_t)(u >> 32);
    }
    if (hx >= 0x7ff00000) return x + x;
    k += (hx >> 20) - 1023;
    hx &= 0x000fffff;
    i = (hx + 0x95f64) & 0x100000;
    memcpy(&u, &x, 8);
    u = (u & 0xffffffffULL) | ((uint64_t)(uint32_t)(hx | (i ^ 0x3ff00000)) << 32);
    memcpy(&x, &u, 8);
    k += (i >> 20);
    f = x - 1.0;
    if ((0x000fffff & (2 + hx)) < 3) {
        if (f == 0.0) {
            if (k == 0) return 0.0;
            dk = (double)k;
            return dk * ln2_hi + dk * ln2_lo;
        }
        R = f * f * (0.5 - 0.33333333333333333 * f);
        if (k == 0) return f - R;
        dk = (double)k;
        return dk * ln2_hi - ((R - dk * ln2_lo) - f);
    }
    s = f / (2.0 + f);
    dk = (double)k;
    z = s * s;
    i = hx - 0x6147a;
    w = z * z;
    j = 0x6b851 - hx;
    t1 = w * (Lg2 + w * (Lg4 + w * Lg6));
    t2 = z * (Lg1 + w * (Lg3 + w * (Lg5 + w * Lg7)));
    i |= j;
    R = t2 + t1;
    if (i > 0) {
        hfsq = 0.5 * f * f;
        if (k == 0) return f - (hfsq - s * (hfsq + R));
        return dk * ln2_hi - ((hfsq - (s * (hfsq + R) + dk * ln2_lo)) - f);
    }
    if (k == 0) return f - s * (f - R);
    return dk * ln2_hi - ((s * (f - R) - dk * ln2_lo) - f);
}

double ora_strict_log(double x) { return ora_log(x); }

/* ---- InfoContentStat.processStat (util/InfoContentStat.java:71-101) -------------------------- */
/* Classes with zero count have no map entry; terms accumulate in `order`. */
double ora_impurity(const uint32_t* cnt, int C, const uint8_t* order, int entropy) {
    uint32_t total = 0;
    double stat = 0.0;
    int i;
    for (i = 0; i < C; ++i) total += cnt[i];
    if (total == 0) return 0.0;
    if (entropy) {
        const double log2v = ora_log(2.0);
        for (i = 0; i < C; ++i) {
            uint32_t c = cnt[order ? order[i] : i];
            double pr;
            if (!c) continue;
            pr = (double)c / (double)total;
            stat -= pr * ora_log(pr) / log2v;
        }
        return stat;
    } else {
        double prsq = 0.0;
        for (i = 0; i < C; ++i) {
            uint32_t c = cnt[order ? order[i] : i];
            double pr;
            if (!c) continue;
            pr = (double)c / (double)total;
            prsq += pr * pr;
        }
        return 1.0 - prsq;
    }
}

/* ---- counting -------------------------------------------------------------------------------- */
typedef struct hist_job {
    int64_t r0, r1;
    int A, C, T, TB;
    const int32_t* n_bins;
    const int32_t* bin_off;
    const void* const* cols;
    const uint8_t* cls;
    const uint16_t* const* nid;
    const uint8_t* const* w;
    const int32_t* node_base;
    uint32_t* H;
    size_t hwords;
} hist_job;

static void* hist_worker(void* arg) {
    hist_job* j = (hist_job*)arg;
    int64_t r;
    int t, a;
    for (r = j->r0; r < j->r1; ++r) {
        const uint32_t cl = j->cls[r];
        for (t = 0; t < j->T; ++t) {
            const uint32_t node = j->nid && j->nid[t] ? j->nid[t][r] : 0;
            const uint32_t wt = j->w && j->w[t] ? j->w[t][r] : 1;
            uint32_t* Hn;
            if (node == ORA_NODE_NONE || wt == 0) continue;
            Hn = j->H + (size_t)(j->node_base[t] + node) * j->TB * j->C;
            for (a = 0; a < j->A; ++a) {
                const uint32_t code = j->n_bins[a] > 256 ? ((const uint16_t*)j->cols[a])[r] : ((const uint8_t*)j->cols[a])[r];
                Hn[(size_t)(j->bin_off[a] + code) * j->C + cl] += wt;
            }
        }
    }
    return NULL;
}

/* H[tree][node][attr][bin][class] += weight: one increment per (row, candidate attribute), which is
 * what one emitted record per (row, split, true predicate) sums to for any split of that attribute. */
int ora_histogram(int64_t n_rows, int A, const int32_t* n_bins, const void* const* cols, const uint8_t* cls,
                  int C, int T, const uint16_t* const* nid, const uint8_t* const* w,
                  const int32_t* nodes_per_tree, uint32_t* H, int n_threads) {
    int32_t* bin_off = (int32_t*)malloc(sizeof(int32_t) * (size_t)(A + 1));
    int32_t* node_base = (int32_t*)malloc(sizeof(int32_t) * (size_t)(T + 1));
    int a, t, i;
    size_t hwords;
    hist_job* jobs;
    pthread_t* th;
    if (n_threads < 1) n_threads = 1;
    bin_off[0] = 0;
    for (a = 0; a < A; ++a) bin_off[a + 1] = bin_off[a] + n_bins[a];
    node_base[0] = 0;
    for (t = 0; t < T; ++t) node_base[t + 1] = node_base[t] + nodes_per_tree[t];
    hwords = (size_t)node_base[T] * bin_off[A] * C;
    memset(H, 0, hwords * sizeof(uint32_t));
    jobs = (hist_job*)calloc((size_t)n_threads, sizeof(hist_job));
    th = (pthread_t*)calloc((size_t)n_threads, sizeof(pthread_t));
    for (i = 0; i < n_threads; ++i) {
        hist_job* j = &jobs[i];
        j->r0 = n_rows * i / n_threads;
        j->r1 = n_rows * (i + 1) / n_threads;
        j->A = A; j->C = C; j->T = T; j->TB = bin_off[A];
        j->n_bins = n_bins; j->bin_off = bin_off; j->cols = cols; j->cls = cls; j->nid = nid; j->w = w;
        j->node_base = node_base; j->hwords = hwords;
        j->H = i == 0 ? H : (uint32_t*)calloc(hwords, sizeof(uint32_t));
    }
    for (i = 1; i < n_threads; ++i) pthread_create(&th[i], NULL, hist_worker, &jobs[i]);
    hist_worker(&jobs[0]);
    for (i = 1; i < n_threads; ++i) {
        size_t k;
        pthread_join(th[i], NULL);
        for (k = 0; k < hwords; ++k) H[k] += jobs[i].H[k];
        free(jobs[i].H);
    }
    free(jobs); free(th); free(bin_off); free(node_base);
    return 0;
}

/* ---- segments of a split --------------------------------------------------------------------- */
static void seg_counts(const ora_split* sp, int j, int B, int C, const int32_t* bounds, const uint8_t* lut,
                       const uint32_t* Hn, uint32_t* cnt) {
    int b, c;
    for (c = 0; c < C; ++c) cnt[c] = 0;
    if (sp->kind == 0) {
        /* numeric: segment j = bins (bounds[j-1], bounds[j]]; value <= p_k  <=>  bin <= k
         * (tree/SplitManager.java:635-661 for two segments; util/AttributeSplitHandler.java:148-155) */
        int lo = j == 0 ? -1 : bounds[sp->off + j - 1];
        int hi = j == sp->n_seg - 1 ? B - 1 : bounds[sp->off + j];
        for (b = lo + 1; b <= hi; ++b)
            for (c = 0; c < C; ++c) cnt[c] += Hn[b * C + c];
    } else if (sp->kind == 2) {
        /* multi-way numeric predicates as inclusive bin ranges that may overlap: `le p_last` has no lower bound
         * (tree/SplitManager.java:644-658), the double twins never have one (:677-689, 840-843) */
        int lo = bounds[sp->off + 2 * j], hi = bounds[sp->off + 2 * j + 1];
        if (hi > B - 1) hi = B - 1;
        for (b = lo; b <= hi; ++b)
            for (c = 0; c < C; ++c) cnt[c] += Hn[b * C + c];
    } else {
        /* categorical: bins whose value is in group j (tree/SplitManager.java:928-937) */
        for (b = 0; b < B; ++b)
            if (lut[sp->off + b] == j)
                for (c = 0; c < C; ++c) cnt[c] += Hn[b * C + c];
    }
}

/* weighted average impurity of a split (tree/DecisionTreeBuilder.java:534-548) */
static double split_av(const ora_split* sp, int B, int C, const int32_t* bounds, const uint8_t* lut,
                       const uint32_t* Hn, const uint8_t* order, int entropy, ora_eval_t* ev) {
    double weighted = 0.0;
    uint32_t total = 0;
    int j, c;
    for (j = 0; j < sp->n_seg; ++j) {
        uint32_t cnt[ORA_MAX_C], n = 0;
        double stat = 0.0;
        seg_counts(sp, j, B, C, bounds, lut, Hn, cnt);
        for (c = 0; c < C; ++c) n += cnt[c];
        if (n) {
            stat = ora_impurity(cnt, C, order, entropy);
            weighted += stat * n;
            total += n;
        }
        if (ev) {
            ev->seg_stat[j] = stat;
            for (c = 0; c < C; ++c) ev->seg_count[j][c] = cnt[c];
        }
    }
    return weighted / total;
}

/* BuilderReducer.expandTree's split loop for every frontier node (tree/DecisionTreeBuilder.java:
 * 512-565): av per candidate split, minimum, lowest table index attaining it, number of equal minima. */
int ora_scan(const uint32_t* H, int NF, int A, const int32_t* n_bins, int C, int S, const ora_split* splits,
             const int32_t* bounds, const uint8_t* lut, const uint32_t* cand_mask, const uint8_t* order,
             int entropy, double* av /* [NF][S] */, ora_best* best /* [NF] */) {
    int32_t* bin_off = (int32_t*)malloc(sizeof(int32_t) * (size_t)(A + 1));
    int a, g, s, c, b;
    const int mw = (A + 31) / 32;
    bin_off[0] = 0;
    for (a = 0; a < A; ++a) bin_off[a + 1] = bin_off[a] + n_bins[a];
    for (g = 0; g < NF; ++g) {
        const uint32_t* Hnode = H + (size_t)g * bin_off[A] * C;
        ora_best r;
        uint32_t cnt[ORA_MAX_C];
        r.best_split = -1; r.n_ties = 0; r.min_av = NAN;
        for (s = 0; s < S; ++s) {
            const ora_split* sp = &splits[s];
            double v;
            if (cand_mask && !((cand_mask[(size_t)g * mw + (sp->attr >> 5)] >> (sp->attr & 31)) & 1u)) {
                av[(size_t)g * S + s] = NAN;
                continue;
            }
            v = split_av(sp, n_bins[sp->attr], C, bounds, lut, Hnode + (size_t)bin_off[sp->attr] * C, order, entropy, NULL);
            av[(size_t)g * S + s] = v;
            if (v != v) continue;
            if (r.best_split < 0 || v < r.min_av) { r.min_av = v; r.best_split = s; r.n_ties = 1; }
            else if (v == r.min_av) r.n_ties += 1;
        }
        for (c = 0; c < C; ++c) cnt[c] = 0;
        for (b = 0; b < n_bins[0]; ++b)
            for (c = 0; c < C; ++c) cnt[c] += Hnode[b * C + c];
        r.population = 0;
        for (c = 0; c < C; ++c) r.population += cnt[c];
        r.node_stat = ora_impurity(cnt, C, order, entropy);
        best[g] = r;
    }
    free(bin_off);
    return 0;
}

int ora_eval(const uint32_t* H, int A, const int32_t* n_bins, int C, const ora_split* splits, const int32_t* bounds,
             const uint8_t* lut, const uint8_t* order, int entropy, int n, const int32_t* nodes, const int32_t* sidx,
             ora_eval_t* out) {
    int32_t* bin_off = (int32_t*)malloc(sizeof(int32_t) * (size_t)(A + 1));
    int a, i;
    bin_off[0] = 0;
    for (a = 0; a < A; ++a) bin_off[a + 1] = bin_off[a] + n_bins[a];
    for (i = 0; i < n; ++i) {
        const ora_split* sp = &splits[sidx[i]];
        memset(&out[i], 0, sizeof(ora_eval_t));
        out[i].av = split_av(sp, n_bins[sp->attr], C, bounds, lut,
                             H + ((size_t)nodes[i] * bin_off[A] + bin_off[sp->attr]) * C, order, entropy, &out[i]);
    }
    free(bin_off);
    return 0;
}

/* row -> child (tree/DecisionTreeBuilder.java:333-349 re-keys the row with the predicate it satisfies;
 * :279-292 drops rows whose path is not in the new list) */
int ora_repartition(int64_t n_rows, const int32_t* n_bins, const void* const* cols, int T, uint16_t* const* nid,
                    const int32_t* nodes_per_tree, const int32_t* node_attr, const uint16_t* child_lut, int lut_width) {
    int t;
    int64_t r;
    int base = 0;
    for (t = 0; t < T; ++t) {
        for (r = 0; r < n_rows; ++r) {
            uint32_t old = nid[t][r];
            int g, a;
            uint32_t code;
            if (old == ORA_NODE_NONE) continue;
            g = base + (int)old;
            a = node_attr[g];
            if (a < 0) { nid[t][r] = ORA_NODE_NONE; continue; }
            code = n_bins[a] > 256 ? ((const uint16_t*)cols[a])[r] : ((const uint8_t*)cols[a])[r];
            nid[t][r] = child_lut[(size_t)g * lut_width + code];
        }
        base += nodes_per_tree[t];
    }
    return 0;
}

/* ClassPartitionGenerator reducer statistics for every split at one node
 * (explore/ClassPartitionGenerator.java:514-553; util/AttributeSplitStat.java as cited above).
 * criterion: 0 gini, 1 entropy, 2 Hellinger, 3 class-confidence ratio. */
int ora_cpg(const uint32_t* H, int node, int A, const int32_t* n_bins, int C, int S, const ora_split* splits,
            const int32_t* bounds, const uint8_t* lut, const uint8_t* order, int criterion, ora_cpg_t* out) {
    int32_t* bin_off = (int32_t*)malloc(sizeof(int32_t) * (size_t)(A + 1));
    int a, s, j, c, i;
    const double log2v = ora_log(2.0);
    bin_off[0] = 0;
    for (a = 0; a < A; ++a) bin_off[a + 1] = bin_off[a] + n_bins[a];
    for (s = 0; s < S; ++s) {
        const ora_split* sp = &splits[s];
        const uint32_t* Hn = H + ((size_t)node * bin_off[A] + bin_off[sp->attr]) * C;
        ora_cpg_t r;
        uint32_t segn[ORA_MAX_SEG], total = 0;
        double info = 0.0, stat = 0.0;
        memset(&r, 0, sizeof r);
        for (j = 0; j < sp->n_seg; ++j) {
            uint32_t cnt[ORA_MAX_C];
            seg_counts(sp, j, n_bins[sp->attr], C, bounds, lut, Hn, cnt);
            segn[j] = 0;
            for (c = 0; c < C; ++c) { r.seg_count[j][c] = cnt[c]; segn[j] += cnt[c]; }
            total += segn[j];
        }
        for (j = 0; j < sp->n_seg; ++j) {          /* SplitStat.getInfoContent :153-170 */
            double pr;
            if (!segn[j]) continue;
            pr = (double)segn[j] / (double)total;
            info -= pr * ora_log(pr) / log2v;
        }
        if (criterion == 0 || criterion == 1) {    /* SplitInfoContent.processStat :191-218 */
            double sum = 0.0;
            for (j = 0; j < sp->n_seg; ++j) {
                if (!segn[j]) continue;
                sum += ora_impurity(r.seg_count[j], C, order, criterion == 1) * segn[j];
            }
            stat = sum / total;
        } else if (criterion == 2) {               /* SplitHellingerDistance.processStat :240-282 */
            const int c0 = order[0], c1 = order[1];
            uint32_t n0 = 0, n1 = 0;
            double sum = 0.0;
            for (j = 0; j < sp->n_seg; ++j) { n0 += r.seg_count[j][c0]; n1 += r.seg_count[j][c1]; }
            for (j = 0; j < sp->n_seg; ++j) {
                double v0, v1;
                if (!segn[j]) continue;
                v0 = sqrt((double)r.seg_count[j][c0] / (double)n0);
                v1 = sqrt((double)r.seg_count[j][c1] / (double)n1);
                sum += (v0 - v1) * (v0 - v1);
            }
            stat = sqrt(sum);
        } else {                                   /* SplitClassCofidenceRatio.processStat :298-336, :452-471 */
            uint32_t ctot[ORA_MAX_C];
            double sum = 0.0;
            for (c = 0; c < C; ++c) {
                ctot[c] = 0;
                for (j = 0; j < sp->n_seg; ++j) ctot[c] += r.seg_count[j][c];
            }
            for (j = 0; j < sp->n_seg; ++j) {
                double conf[ORA_MAX_C], tc = 0.0, ent = 0.0;
                if (!segn[j]) continue;
                for (i = 0; i < C; ++i) {
                    conf[i] = (double)r.seg_count[j][order[i]] / (double)ctot[order[i]];
                    tc += conf[i];
                }
                for (i = 0; i < C; ++i) {
                    double ccr = conf[i] / tc;
                    ent -= ccr * ora_log(ccr) / log2v;
                }
                sum += ent * segn[j];
            }
            stat = sum / total;
        }
        r.stat = stat;
        r.split_info = info;
        out[s] = r;
    }
    free(bin_off);
    return 0;
}

/* Batch prediction on admission masks (tree/DecisionTreeModel.java:56-93: first decision path, in list order,
 * whose predicates all hold; model/EnsemblePredictiveModel.java:69-110: weighted vote, stable sort by count).
 * Deliberately the slow way round: path by path, attribute by attribute. */
int ora_predict(int64_t n_rows, int A, const int32_t* n_bins, const void* const* cols, int M, const int32_t* n_paths,
                const uint32_t* masks, const uint8_t* path_class, const double* weights, int n_classes,
                const uint8_t* class_rank, uint8_t* out) {
    int32_t* bin_off = (int32_t*)malloc(sizeof(int32_t) * (size_t)(A + 1));
    int a, m, p, c;
    int64_t r;
    bin_off[0] = 0;
    for (a = 0; a < A; ++a) bin_off[a + 1] = bin_off[a] + n_bins[a];
    for (r = 0; r < n_rows; ++r) {
        double votes[ORA_MAX_C];
        int voted[ORA_MAX_C], unmatched = 0, last = 0, path_base = 0;
        const uint32_t* mk = masks;
        for (c = 0; c < n_classes; ++c) { votes[c] = 0.0; voted[c] = 0; }
        for (m = 0; m < M; ++m) {
            const int P = n_paths[m], W = (P + 31) / 32;
            int found = -1;
            for (p = 0; p < P && found < 0; ++p) {
                int ok = 1;
                for (a = 0; a < A && ok; ++a) {
                    const uint32_t code = n_bins[a] > 256 ? ((const uint16_t*)cols[a])[r] : ((const uint8_t*)cols[a])[r];
                    ok = (mk[(size_t)(bin_off[a] + code) * W + (p >> 5)] >> (p & 31)) & 1u;
                }
                if (ok) found = p;
            }
            if (found < 0) { unmatched = 1; break; }
            c = path_class[path_base + found];
            votes[c] += weights ? weights[m] : 1.0;
            voted[c] = 1;
            last = c;
            mk += (size_t)bin_off[A] * W;
            path_base += P;
        }
        if (unmatched) { out[r] = 0xFF; continue; }
        if (M == 1) { out[r] = (uint8_t)last; continue; }
        {
            int best = -1;
            for (c = 0; c < n_classes; ++c) {
                if (!voted[c]) continue;
                if (best < 0 || votes[c] > votes[best] || (votes[c] == votes[best] && class_rank[c] < class_rank[best])) best = c;
            }
            out[r] = (uint8_t)best;
        }
    }
    free(bin_off);
    return 0;
}

/* ------------------------------------------------------------------------------------------------
 * CSV encode.  Plain-C restatement of the reference's per-row parsing
 * (tree/DecisionTreeBuilder.java:268, 309-320, 749; grid semantics tree/SplitManager.java:256-330) in
 * the histogram form: checker for avt_dataset_load_csv at sizes the Python restatement
 * (oracle/encode_literal.py) cannot reach, and the timed CPU baseline of tools/encode_bench.py.
 * Same arguments as the C-ABI call; out_cols[a] is int32[n_rows], out_cls uint8[n_rows].
 * Returns the number of rows, or -(row+1) of the first row the reference would throw on.
 * ------------------------------------------------------------------------------------------------ */
typedef struct { int32_t ordinal, kind, n, off; } ora_csv_attr;

typedef struct {
    const char* text; int64_t b0, b1, row0, rows, bad;
    char delim; int skip, n_attrs; const ora_csv_attr* attrs;
    const int32_t* ip; const double* dp; const char* cb; const int32_t* co;
    int class_ord, n_class; const char* kb; const int32_t* ko;
    int32_t* const* out_cols; uint8_t* out_cls;
} csv_job;

static int find_str(const char* bytes, const int32_t* off, int n, const char* p, int len) {
    int k;
    for (k = 0; k < n; ++k)
        if (off[k + 1] - off[k] == len && memcmp(bytes + off[k], p, (size_t)len) == 0) return k;
    return -1;
}

static int parse_int32(const char* p, int len, long long* out) {
    int i = 0, neg = 0;
    long long v = 0;
    if (len <= 0) return 0;
    if (p[0] == '-' || p[0] == '+') { neg = p[0] == '-'; i = 1; }
    if (i >= len) return 0;
    for (; i < len; ++i) {
        if (p[i] < '0' || p[i] > '9') return 0;
        v = v * 10 + (p[i] - '0');
        if (v > 2147483648LL) return 0;
    }
    v = neg ? -v : v;
    if (v > 2147483647LL || v < -2147483648LL) return 0;
    *out = v;
    return 1;
}

static int encode_one(const csv_job* j, const char* row, int len, int64_t r) {
    const char* fs[64]; int fl[64];
    int nf = 0, s = 0, i, a, need = j->class_ord;
    if (len > 0 && row[len - 1] == '\r') --len;
    for (a = 0; a < j->n_attrs; ++a) if (j->attrs[a].ordinal > need) need = j->attrs[a].ordinal;
    need += j->skip + 1;
    if (need > 64) return 0;
    for (i = 0; i <= len && nf < 64; ++i)
        if (i == len || row[i] == j->delim) { fs[nf] = row + s; fl[nf] = i - s; ++nf; s = i + 1; }
    if (i <= len) { /* more than 64 fields: only whether the tail is empty matters for the length */
        int k; for (k = i; k < len; ++k) if (row[k] != j->delim) { nf = 65; break; }
    }
    while (nf > 1 && nf <= 64 && fl[nf - 1] == 0) --nf;          /* String.split drops trailing empty strings */
    if (nf == 1 && fl[0] == 0 && len > 0) nf = 0;
    for (a = 0; a < j->n_attrs; ++a) {
        const ora_csv_attr* A = &j->attrs[a];
        const int f = A->ordinal + j->skip;
        int bin = 0, k;
        if (f >= nf) return 0;
        if (A->kind == 0) {
            long long v;
            if (!parse_int32(fs[f], fl[f], &v)) return 0;
            for (k = 0; k < A->n && (long long)j->ip[A->off + k] < v; ++k) {}
            bin = k;
        } else if (A->kind == 1) {
            char buf[128]; char* end; double v; int b0 = 0, b1 = fl[f];
            while (b0 < b1 && (unsigned char)fs[f][b0] <= ' ') ++b0;
            while (b1 > b0 && (unsigned char)fs[f][b1 - 1] <= ' ') --b1;
            if (b1 - b0 <= 0 || b1 - b0 >= (int)sizeof buf) return 0;
            memcpy(buf, fs[f] + b0, (size_t)(b1 - b0)); buf[b1 - b0] = 0;
            v = strtod(buf, &end);
            if (*end) return 0;
            for (k = 0; k < A->n && j->dp[A->off + k] < v; ++k) {}
            bin = k;
        } else {
            bin = find_str(j->cb, j->co + A->off, A->n, fs[f], fl[f]);
            if (bin < 0) return 0;
        }
        j->out_cols[a][r] = bin;
    }
    if (j->class_ord < 0) { /* rows without a class field (tree/DataPartitioner.java's input): one class, code 0 */
        j->out_cls[r] = 0;
    } else {
        const int f = j->class_ord + j->skip;
        int c;
        if (f >= nf) return 0;
        c = find_str(j->kb, j->ko, j->n_class, fs[f], fl[f]);
        if (c < 0) return 0;
        j->out_cls[r] = (uint8_t)c;
    }
    return 1;
}

static void* csv_count_worker(void* arg) {
    csv_job* j = (csv_job*)arg;
    const char* p = j->text + j->b0; const char* e = j->text + j->b1;
    int64_t c = 0;
    while (p < e && (p = (const char*)memchr(p, '\n', (size_t)(e - p))) != NULL) { ++c; ++p; }
    j->rows = c;
    return NULL;
}

static void* csv_parse_worker(void* arg) {
    csv_job* j = (csv_job*)arg;
    int64_t pos = j->b0, r = j->row0;
    j->bad = -1;
    while (pos < j->b1) {
        const char* nl = (const char*)memchr(j->text + pos, '\n', (size_t)(j->b1 - pos));
        const int64_t end = nl ? (int64_t)(nl - j->text) : j->b1;
        if (!encode_one(j, j->text + pos, (int)(end - pos), r) && j->bad < 0) j->bad = r;
        ++r;
        pos = end + 1;
    }
    return NULL;
}

int64_t ora_encode_csv(const char* text, int64_t n_bytes, char delim, int skip, int n_attrs, const ora_csv_attr* attrs,
                       const int32_t* ip, const double* dp, const char* cb, const int32_t* co, int class_ord, int n_class,
                       const char* kb, const int32_t* ko, int32_t* const* out_cols, uint8_t* out_cls, int n_threads) {
    csv_job* jobs; pthread_t* th; int i; int64_t rows = 0, bad = -1;
    if (n_threads < 1) n_threads = 1;
    jobs = (csv_job*)calloc((size_t)n_threads, sizeof(csv_job));
    th = (pthread_t*)calloc((size_t)n_threads, sizeof(pthread_t));
    for (i = 0; i < n_threads; ++i) {   /* byte ranges cut at row boundaries */
        int64_t b = n_bytes * i / n_threads;
        if (i > 0) { while (b < n_bytes && text[b - 1] != '\n') ++b; }
        jobs[i].b0 = b;
        jobs[i].text = text; jobs[i].delim = delim; jobs[i].skip = skip; jobs[i].n_attrs = n_attrs; jobs[i].attrs = attrs;
        jobs[i].ip = ip; jobs[i].dp = dp; jobs[i].cb = cb; jobs[i].co = co; jobs[i].class_ord = class_ord; jobs[i].n_class = n_class;
        jobs[i].kb = kb; jobs[i].ko = ko; jobs[i].out_cols = out_cols; jobs[i].out_cls = out_cls;
    }
    for (i = 0; i < n_threads; ++i) jobs[i].b1 = i + 1 < n_threads ? jobs[i + 1].b0 : n_bytes;
    for (i = 1; i < n_threads; ++i) pthread_create(&th[i], NULL, csv_count_worker, &jobs[i]);
    csv_count_worker(&jobs[0]);
    for (i = 1; i < n_threads; ++i) pthread_join(th[i], NULL);
    for (i = 0; i < n_threads; ++i) {
        jobs[i].row0 = rows;
        rows += jobs[i].rows;
        if (i == n_threads - 1 && n_bytes > 0 && text[n_bytes - 1] != '\n') ++rows;   /* unterminated last row */
    }
    if (out_cols) {
        for (i = 1; i < n_threads; ++i) pthread_create(&th[i], NULL, csv_parse_worker, &jobs[i]);
        csv_parse_worker(&jobs[0]);
        for (i = 1; i < n_threads; ++i) pthread_join(th[i], NULL);
        for (i = 0; i < n_threads; ++i) if (jobs[i].bad >= 0 && (bad < 0 || jobs[i].bad < bad)) bad = jobs[i].bad;
    }
    free(jobs); free(th);
    return bad >= 0 ? -(bad + 1) : rows;
}

/* ---- explore/BaggingSampler.java:101-122 (emitBatch), SURVEY App. B-3 / B-5 / E --------------------------------
 * Bootstrap multiplicities of rows [row0, row0 + n) of a file of `total_rows` rows: rows are taken in file
 * order in batches of `batch` (bas.batch.size, default 10 000; the tail batch is shorter); a batch of m rows
 * emits m draws `batch.get((int)(Math.random() * m))`; the multiplicity of a row is the number of draws that
 * hit it.  Math.random() is java.util.Random.nextDouble(): seed' = (seed * 0x5DEECE66D + 0xB) mod 2^48,
 * next(b) = seed' >> (48 - b), nextDouble = ((next(26) << 27) + next(27)) * 2^-53.  The stream is seeded
 * (our addition, App. B-5) with `seed` scrambled as `new Random(seed)` does.  Sequential on purpose: this is
 * the checker of the device generator's skip-ahead.  Returns -1 when a multiplicity exceeds 255. */
int ora_bagging_weights(int64_t seed, int32_t batch, int64_t total_rows, int64_t row0, int64_t n, uint8_t* out) {
    const uint64_t MULT = 0x5DEECE66DULL, MASK = (1ULL << 48) - 1;
    uint64_t s = ((uint64_t)seed ^ MULT) & MASK;
    int64_t b0, i;
    memset(out, 0, (size_t)n);
    if (batch < 1) batch = 1;
    for (b0 = 0; b0 < total_rows; b0 += batch) {
        const int64_t m = total_rows - b0 < batch ? total_rows - b0 : batch;
        for (i = 0; i < m; ++i) {
            uint64_t hi, lo;
            double d;
            int64_t sel;
            s = (s * MULT + 0xBULL) & MASK; hi = s >> (48 - 26);
            s = (s * MULT + 0xBULL) & MASK; lo = s >> (48 - 27);
            d = (double)((hi << 27) + lo) * (1.0 / 9007199254740992.0);
            sel = (int64_t)(int32_t)(d * (double)m);
            if (sel >= m) sel = m - 1;
            sel += b0 - row0;
            if (sel >= 0 && sel < n) {
                if (out[sel] == 255) return -1;
                ++out[sel];
            }
        }
    }
    return 0;
}
