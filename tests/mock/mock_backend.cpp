// TEST-ONLY link-time stand-in for the kernel-level C-ABI (include/avenir_tree.h).
//
// The product's C++ tool layer (avenir_b200/csrc/host/*.cpp) calls avt_* by symbol.  In
// libavenir_tree.so those symbols are the CUDA engine.  For CPU-only tests of the HOST logic
// (encoding, split ids, HashMap-order replay, JSON) this file is linked INSTEAD of the CUDA engine
// into tests/_build/libavt_hostmock.so and answers from the C oracle (oracle/hist_oracle.c).
// It is never part of libavenir_tree.so; the product has no CPU path.
#include <cstring>
#include <string>
#include <vector>

#include "../../include/avenir_tree.h"

extern "C" {
struct ora_split { int32_t attr, kind, n_seg, off; };
struct ora_best { double min_av, node_stat; int64_t population; int32_t best_split, n_ties; };
int ora_histogram(int64_t, int, const int32_t*, const void* const*, const uint8_t*, int, int, const uint16_t* const*,
                  const uint8_t* const*, const int32_t*, uint32_t*, int);
int ora_scan(const uint32_t*, int, int, const int32_t*, int, int, const ora_split*, const int32_t*, const uint8_t*,
             const uint32_t*, const uint8_t*, int, double*, ora_best*);
int ora_eval(const uint32_t*, int, const int32_t*, int, const ora_split*, const int32_t*, const uint8_t*, const uint8_t*,
             int, int, const int32_t*, const int32_t*, avt_split_eval*);
int ora_repartition(int64_t, const int32_t*, const void* const*, int, uint16_t* const*, const int32_t*, const int32_t*,
                    const uint16_t*, int);
int ora_predict(int64_t, int, const int32_t*, const void* const*, int, const int32_t*, const uint32_t*, const uint8_t*, const double*, int,
                const uint8_t*, uint8_t*);
int ora_cpg(const uint32_t*, int, int, const int32_t*, int, int, const ora_split*, const int32_t*, const uint8_t*,
            const uint8_t*, int, avt_cpg_stat*);
int64_t ora_encode_csv(const char*, int64_t, char, int, int, const avt_csv_attr*, const int32_t*, const double*, const char*, const int32_t*, int, int,
                       const char*, const int32_t*, int32_t* const*, uint8_t*, int);
}

struct avt_engine {
    std::string err;
    int64_t N = 0;
    int A = 0, C = 0, TB = 0, T = 0, NF = 0;
    std::vector<int32_t> n_bins;
    std::vector<std::vector<uint8_t>> col8;
    std::vector<std::vector<uint16_t>> col16;
    std::vector<const void*> cols;
    std::vector<uint8_t> cls;
    std::vector<std::vector<uint16_t>> nid;
    std::vector<std::vector<uint8_t>> w;
    std::vector<int32_t> npt;
    std::vector<uint32_t> H;
    std::vector<avt_split> splits;
    std::vector<int32_t> bounds;
    std::vector<uint8_t> lut;
    std::vector<double> av;
    std::string csv_text;   // text of the last avt_dataset_load_csv (empty: the data set did not come from text)
    avt_stats stats{};
};

static std::string g_err;
static std::vector<uint8_t> order_or_identity(avt_engine* e, const uint8_t* o) {
    std::vector<uint8_t> v(e->C);
    for (int i = 0; i < e->C; ++i) v[i] = o ? o[i] : (uint8_t)i;
    return v;
}

extern "C" {
const char* avt_last_error(const avt_engine* e) { return e ? e->err.c_str() : g_err.c_str(); }
int avt_create(int, avt_engine** out) { *out = new avt_engine(); return 0; }
void avt_destroy(avt_engine* e) { delete e; }
int avt_dataset_load(avt_engine* e, int64_t n, int32_t A, const int32_t* nb, const void* const* columns, const uint8_t* cc, int32_t C) {
    e->N = n; e->A = A; e->C = C; e->n_bins.assign(nb, nb + A); e->TB = 0;
    e->col8.assign(A, {}); e->col16.assign(A, {}); e->cols.assign(A, nullptr);
    for (int a = 0; a < A; ++a) {
        e->TB += nb[a];
        if (nb[a] > 256) { e->col16[a].assign((const uint16_t*)columns[a], (const uint16_t*)columns[a] + n); e->cols[a] = e->col16[a].data(); }
        else { e->col8[a].assign((const uint8_t*)columns[a], (const uint8_t*)columns[a] + n); e->cols[a] = e->col8[a].data(); }
    }
    e->cls.assign(cc, cc + n);
    e->csv_text.clear();
    return 0;
}
int avt_dataset_fetch_column(avt_engine* e, int32_t attr, void* out) {
    if (attr < 0) memcpy(out, e->cls.data(), (size_t)e->N);
    else if (e->n_bins[attr] > 256) memcpy(out, e->col16[attr].data(), (size_t)e->N * 2);
    else memcpy(out, e->col8[attr].data(), (size_t)e->N);
    return 0;
}
int avt_rows_partition(avt_engine* e, int32_t attr, const uint16_t* seg_of_bin, int32_t n_seg, char* out, int64_t cap, int64_t* seg_bytes, int64_t* seg_rows) {
    // plain restatement: rows in input order per segment, '\r' before the terminator dropped, every row ends in '\n'
    if (e->csv_text.empty()) { e->err = "avt_rows_partition (mock): no text"; return AVT_E_STATE; }
    std::vector<std::string> seg((size_t)n_seg);
    std::vector<int64_t> rows((size_t)n_seg, 0);
    size_t pos = 0;
    for (int64_t r = 0; r < e->N; ++r) {
        size_t nl = e->csv_text.find('\n', pos);
        if (nl == std::string::npos) nl = e->csv_text.size();
        size_t end = nl;
        if (end > pos && e->csv_text[end - 1] == '\r') --end;
        const unsigned code = e->n_bins[attr] > 256 ? e->col16[attr][(size_t)r] : e->col8[attr][(size_t)r];
        const uint16_t s = seg_of_bin[code];
        if (s != AVT_NODE_NONE) { seg[s].append(e->csv_text, pos, end - pos); seg[s] += '\n'; ++rows[s]; }
        pos = nl + 1;
    }
    int64_t total = 0;
    for (int s = 0; s < n_seg; ++s) { seg_bytes[s] = (int64_t)seg[(size_t)s].size(); if (seg_rows) seg_rows[s] = rows[(size_t)s]; total += seg_bytes[s]; }
    if (!out) return 0;
    if (cap < total) { e->err = "avt_rows_partition (mock): output buffer too small"; return AVT_E_INVALID; }
    for (int s = 0; s < n_seg; ++s) { memcpy(out, seg[(size_t)s].data(), seg[(size_t)s].size()); out += seg[(size_t)s].size(); }
    return 0;
}
int avt_dataset_load_csv(avt_engine* e, const char* text, int64_t n_bytes, char delim, int32_t skip, int32_t A, const avt_csv_attr* attrs, const int32_t* ip,
                         const double* dp, const char* cb, const int32_t* co, int32_t class_ord, int32_t n_class, const char* kb, const int32_t* ko,
                         int64_t* n_rows, int64_t* first_row, int64_t* bad_row, int32_t* bad_reason) {
    // the C restatement of the encode (avt_csv_attr and ora_csv_attr have the same layout)
    if (bad_row) *bad_row = -1;
    if (bad_reason) *bad_reason = 0;
    const int64_t n = ora_encode_csv(text, n_bytes, delim, skip, A, attrs, ip, dp, cb, co, class_ord, n_class, kb, ko, nullptr, nullptr, 1);
    std::vector<std::vector<int32_t>> tmp((size_t)A, std::vector<int32_t>((size_t)n));
    std::vector<int32_t*> ptrs;
    for (auto& v : tmp) ptrs.push_back(v.data());
    std::vector<uint8_t> cls((size_t)n);
    const int64_t rc = ora_encode_csv(text, n_bytes, delim, skip, A, attrs, ip, dp, cb, co, class_ord, n_class, kb, ko, ptrs.data(), cls.data(), 1);
    if (rc < 0) {
        if (bad_row) *bad_row = -rc - 1;
        e->err = "avt_dataset_load_csv (mock): row " + std::to_string(-rc - 1) + " rejected";
        return AVT_E_REFERENCE;
    }
    if (n_rows) *n_rows = n;
    std::vector<int32_t> nb((size_t)A);
    for (int a = 0; a < A; ++a) nb[(size_t)a] = attrs[a].kind == AVT_CSV_CATEGORICAL ? attrs[a].n : attrs[a].n + 1;
    std::vector<std::vector<uint8_t>> c8((size_t)A);
    std::vector<std::vector<uint16_t>> c16((size_t)A);
    std::vector<const void*> cols((size_t)A);
    for (int a = 0; a < A; ++a) {
        if (nb[(size_t)a] > 256) { c16[(size_t)a].assign(tmp[(size_t)a].begin(), tmp[(size_t)a].end()); cols[(size_t)a] = c16[(size_t)a].data(); }
        else { c8[(size_t)a].assign(tmp[(size_t)a].begin(), tmp[(size_t)a].end()); cols[(size_t)a] = c8[(size_t)a].data(); }
    }
    if (first_row) {
        for (int k = 0; k < n_class; ++k) first_row[k] = -1;
        for (int64_t r = 0; r < n; ++r) if (first_row[cls[(size_t)r]] < 0) first_row[cls[(size_t)r]] = r;
    }
    const int rc2 = avt_dataset_load(e, n, A, nb.data(), cols.data(), cls.data(), n_class);
    e->csv_text.assign(text, (size_t)n_bytes);
    return rc2;
}
int avt_forest_init(avt_engine* e, int32_t T, const uint8_t* const* weights, const uint16_t* const* node_ids) {
    e->T = T; e->nid.assign(T, {}); e->w.assign(T, {});
    for (int t = 0; t < T; ++t) {
        if (node_ids && node_ids[t]) e->nid[t].assign(node_ids[t], node_ids[t] + e->N); else e->nid[t].assign(e->N, 0);
        if (weights && weights[t]) e->w[t].assign(weights[t], weights[t] + e->N);
    }
    return 0;
}
int avt_split_table_set(avt_engine* e, int32_t S, const avt_split* s, const int32_t* b, int32_t nb, const uint8_t* l, int64_t nl) {
    e->splits.assign(s, s + S); e->bounds.assign(b, b + nb); e->lut.assign(l, l + nl); return 0;
}
int avt_level_histogram(avt_engine* e, const int32_t* npt, uint32_t* out) {
    e->npt.assign(npt, npt + e->T); e->NF = 0;
    for (int t = 0; t < e->T; ++t) e->NF += npt[t];
    e->H.assign((size_t)e->NF * e->TB * e->C, 0);
    std::vector<const uint16_t*> np; std::vector<const uint8_t*> wp;
    for (int t = 0; t < e->T; ++t) { np.push_back(e->nid[t].data()); wp.push_back(e->w[t].empty() ? nullptr : e->w[t].data()); }
    ora_histogram(e->N, e->A, e->n_bins.data(), e->cols.data(), e->cls.data(), e->C, e->T, np.data(), wp.data(), npt, e->H.data(), 1);
    if (out) memcpy(out, e->H.data(), e->H.size() * 4);
    return 0;
}
int avt_level_scan(avt_engine* e, int32_t crit, const uint32_t* mask, const uint8_t* order, avt_node_best* out) {
    auto o = order_or_identity(e, order);
    e->av.assign((size_t)e->NF * e->splits.size(), 0);
    std::vector<ora_best> b(e->NF);
    ora_scan(e->H.data(), e->NF, e->A, e->n_bins.data(), e->C, (int)e->splits.size(), (const ora_split*)e->splits.data(), e->bounds.data(),
             e->lut.data(), mask, o.data(), crit == AVT_CRIT_ENTROPY, e->av.data(), b.data());
    for (int g = 0; g < e->NF; ++g) { out[g].min_av = b[g].min_av; out[g].node_stat = b[g].node_stat; out[g].population = b[g].population; out[g].best_split = b[g].best_split; out[g].n_ties = b[g].n_ties; }
    return 0;
}
int avt_level_fetch_av(avt_engine* e, int32_t node, double* out) { memcpy(out, e->av.data() + (size_t)node * e->splits.size(), e->splits.size() * 8); return 0; }
int avt_level_fetch_hist(avt_engine* e, int32_t node, uint32_t* out) { size_t w = (size_t)e->TB * e->C; memcpy(out, e->H.data() + node * w, w * 4); return 0; }
int avt_level_eval(avt_engine* e, int32_t crit, const uint8_t* order, int32_t n, const int32_t* nodes, const int32_t* sidx, avt_split_eval* out) {
    auto o = order_or_identity(e, order);
    return ora_eval(e->H.data(), e->A, e->n_bins.data(), e->C, (const ora_split*)e->splits.data(), e->bounds.data(), e->lut.data(), o.data(),
                    crit == AVT_CRIT_ENTROPY, n, nodes, sidx, out);
}
int avt_level_repartition(avt_engine* e, const int32_t* node_attr, const uint16_t* lut, int32_t w) {
    std::vector<uint16_t*> np;
    for (int t = 0; t < e->T; ++t) np.push_back(e->nid[t].data());
    return ora_repartition(e->N, e->n_bins.data(), e->cols.data(), e->T, np.data(), e->npt.data(), node_attr, lut, w);
}
int avt_cpg_sweep(avt_engine* e, int32_t node, int32_t crit, const uint8_t* order, avt_cpg_stat* out) {
    auto o = order_or_identity(e, order);
    return ora_cpg(e->H.data(), node, e->A, e->n_bins.data(), e->C, (int)e->splits.size(), (const ora_split*)e->splits.data(), e->bounds.data(),
                   e->lut.data(), o.data(), crit, out);
}
int avt_predict(avt_engine* e, int32_t M, const int32_t* n_paths, const uint32_t* masks, const uint8_t* path_class, const double* weights,
                int32_t n_classes, const uint8_t* class_rank, uint8_t* out) {
    return ora_predict(e->N, e->A, e->n_bins.data(), e->cols.data(), M, n_paths, masks, path_class, weights, n_classes, class_rank, out);
}
int avt_forest_node_ids(avt_engine* e, int32_t t, uint16_t* out) { memcpy(out, e->nid[t].data(), e->N * 2); return 0; }
int avt_level_class_counts(avt_engine* e, uint32_t* out) {
    for (int g = 0; g < e->NF; ++g)
        for (int c = 0; c < e->C; ++c) {
            uint32_t s = 0;
            for (int b = 0; b < e->n_bins[0]; ++b) s += e->H[((size_t)g * e->TB + b) * e->C + c];
            out[g * e->C + c] = s;
        }
    return 0;
}
int avt_timer_begin(avt_engine*) { return 0; }
int avt_timer_end(avt_engine*, float* ms) { *ms = 0; return 0; }
int avt_synchronize(avt_engine*) { return 0; }
int avt_get_stats(avt_engine* e, avt_stats* s) { *s = e->stats; return 0; }
int avt_reset_stats(avt_engine*) { return 0; }
int avt_set_hist_variant(avt_engine*, int32_t) { return 0; }
int avt_comm_unique_id(void*) { return AVT_E_COMM; }
int avt_comm_init_rank(avt_engine*, int32_t, int32_t, const void*) { return AVT_E_COMM; }
}
