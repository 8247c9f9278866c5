// org.avenir.explore.CramerCorrelation re-hosted on the histogram kernel (the count consumers in
// explore/ share kernel (a)'s sufficient statistic).  The reference's mapper keeps one ContingencyMatrix per
// (source attribute, destination attribute) pair and increments cell (index of the source value, index of the
// destination value) per row (explore/CramerCorrelation.java:88-183); the reducer sums the mappers' matrices and
// prints `srcName,dstName,cramerIndex` (:189-235, util/ContingencyMatrix.java:87-124).  Here the text is encoded on
// the device (avt_dataset_load_csv) once per destination attribute with that attribute as the "class": the root
// histogram H[bin of source attribute][class] IS the pair's contingency matrix, for all source attributes at once.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <memory>

#include "javaemu.hpp"
#include "tool.hpp"

namespace avt_host {

namespace {

std::vector<int> int_list(const Props& conf, const std::string& key) {
    if (!conf.has(key)) throw ReferenceError("NullPointerException: " + key + " is not set");
    std::vector<int> out;
    for (auto& s : split_str(conf.get(key), ",", false)) {
        long long v;
        if (!parse_java_int(s, v)) throw ReferenceError("NumberFormatException: For input string: \"" + s + "\"");
        out.push_back((int)v);
    }
    return out;
}

// ContingencyMatrix.cramerIndex (util/ContingencyMatrix.java:87-124): empty rows / columns count as 1 in the
// denominators, Pearson's sum minus one over (smaller dimension - 1); plain double arithmetic, no fused operations.
double cramer_index(const std::vector<long long>& t, int rows, int cols) {
    std::vector<long long> row_sum((size_t)rows, 0), col_sum((size_t)cols, 0);
    for (int i = 0; i < rows; ++i) {
        for (int j = 0; j < cols; ++j) row_sum[(size_t)i] += t[(size_t)i * cols + j];
        if (row_sum[(size_t)i] == 0) row_sum[(size_t)i] = 1;
    }
    for (int j = 0; j < cols; ++j) {
        for (int i = 0; i < rows; ++i) col_sum[(size_t)j] += t[(size_t)i * cols + j];
        if (col_sum[(size_t)j] == 0) col_sum[(size_t)j] = 1;
    }
    double pearson = 0;
    for (int i = 0; i < rows; ++i)
        for (int j = 0; j < cols; ++j) {
            const double c = (double)t[(size_t)i * cols + j];
            const double num = c * c;
            const double den = (double)row_sum[(size_t)i] * (double)col_sum[(size_t)j];
            pearson += num / den;
        }
    pearson -= 1.0;
    const int smaller = rows < cols ? rows : cols;
    return pearson / (smaller - 1);
}

// ContingencyMatrix.getAggregates + rowSumAsDouble / colSumAsDouble (util/ContingencyMatrix.java:62-85, 126-140)
struct Aggregates {
    std::vector<double> row, col;   // sums (0 counted as 1) over the total count
    long long total = 0;
};
Aggregates aggregates(const std::vector<long long>& t, int rows, int cols) {
    Aggregates a;
    std::vector<long long> rs((size_t)rows, 0), cs((size_t)cols, 0);
    for (int i = 0; i < rows; ++i) {
        for (int j = 0; j < cols; ++j) { rs[(size_t)i] += t[(size_t)i * cols + j]; a.total += t[(size_t)i * cols + j]; }
        if (rs[(size_t)i] == 0) rs[(size_t)i] = 1;
    }
    for (int j = 0; j < cols; ++j) {
        for (int i = 0; i < rows; ++i) cs[(size_t)j] += t[(size_t)i * cols + j];
        if (cs[(size_t)j] == 0) cs[(size_t)j] = 1;
    }
    for (int i = 0; i < rows; ++i) a.row.push_back((double)rs[(size_t)i] / (double)a.total);
    for (int j = 0; j < cols; ++j) a.col.push_back((double)cs[(size_t)j] / (double)a.total);
    return a;
}

// ContingencyMatrix.concentrationCoeff (util/ContingencyMatrix.java:142-164)
double concentration_coeff(const std::vector<long long>& t, int rows, int cols) {
    const Aggregates a = aggregates(t, rows, cols);
    double sum_one = 0;
    for (int i = 0; i < rows; ++i) {
        double sq = 0;
        for (int j = 0; j < cols; ++j) {
            const double elem = (double)t[(size_t)i * cols + j] / (double)a.total;
            const double p = elem * elem;
            sq += p;
        }
        sum_one += sq / a.row[(size_t)i];
    }
    double sum_two = 0;
    for (int j = 0; j < cols; ++j) {
        const double p = a.col[(size_t)j] * a.col[(size_t)j];
        sum_two += p;
    }
    return (sum_one - sum_two) / (1.0 - sum_two);
}

// ContingencyMatrix.uncertaintyCoeff (util/ContingencyMatrix.java:166-187); an empty cell gives 0 * log10(0) = NaN, as there
double uncertainty_coeff(const std::vector<long long>& t, int rows, int cols) {
    const Aggregates a = aggregates(t, rows, cols);
    double sum_one = 0;
    for (int i = 0; i < rows; ++i)
        for (int j = 0; j < cols; ++j) {
            const double elem = (double)t[(size_t)i * cols + j] / (double)a.total;
            const double arg = elem * a.col[(size_t)j] / a.row[(size_t)i];
            const double term = elem * log10(arg);
            sum_one += term;
        }
    double sum_two = 0;
    for (int j = 0; j < cols; ++j) {
        const double term = a.col[(size_t)j] * log10(a.col[(size_t)j]);
        sum_two += term;
    }
    return sum_one / sum_two;
}

struct CorrelationKeys {
    const char *schema, *source, *dest;
    double (*stat)(const std::vector<long long>&, int, int);
};

int run_correlation(const Props& conf, const std::string& in_path, const std::string& out_path, const CorrelationKeys& K);

}  // namespace

int run_cramer_correlation(const Props& conf, const std::string& in_path, const std::string& out_path) {
    return run_correlation(conf, in_path, out_path, CorrelationKeys{"crc.feature.schema.file.path", "crc.source.attributes", "crc.dest.attributes", cramer_index});
}

// org.avenir.explore.HeterogeneityReductionCorrelation: CategoricalCorrelation's mapper (explore/CategoricalCorrelation.java:66-150,
// the same pair bookkeeping, `cac.*` keys -- `cac.second.set..attributes` with two dots is the reference's spelling) and a reducer
// that prints the concentration ("gini") or the uncertainty coefficient (explore/HeterogeneityReductionCorrelation.java:71-88).
int run_heterogeneity_reduction_correlation(const Props& conf, const std::string& in_path, const std::string& out_path) {
    const bool gini = conf.get("hrc.heterogeneity.algorithm", "gini") == "gini";
    return run_correlation(conf, in_path, out_path,
                           CorrelationKeys{"cac.feature.schema.file.path", "cac.first.set.attributes", "cac.second.set..attributes", gini ? concentration_coeff : uncertainty_coeff});
}

namespace {

int run_correlation(const Props& conf, const std::string& in_path, const std::string& out_path, const CorrelationKeys& K) {
    const std::string delim = conf.get("field.delim.regex", ",");
    const std::string delim_out = conf.get("field.delim.out", ",");
    if (!conf.has(K.schema)) throw ReferenceError(std::string("NullPointerException: ") + K.schema + " is not set");
    const Schema schema = Schema::load(conf.get(K.schema));
    const std::vector<int> src = int_list(conf, K.source), dst = int_list(conf, K.dest);
    for (int s : src)
        for (int d : dst)
            if (s == d)
                throw Unsupported("the source and destination attribute sets overlap: the reference's mapper then walks its pair list out of step "
                                  "(explore/CramerCorrelation.java:170-180) and mis-assigns or overruns the matrices; not reproduced");
    if (delim.size() != 1) throw Unsupported("field.delim.regex must be a single literal character for the device encoder");
    std::vector<const Field*> sf, df;
    for (int s : src) sf.push_back(&schema.by_ordinal(s));
    for (int d : dst) df.push_back(&schema.by_ordinal(d));
    for (auto* f : sf)
        if (f->cardinality.empty()) throw ReferenceError("NullPointerException: attribute " + std::to_string(f->ordinal) + " has no cardinality");
    for (auto* f : df) {
        if (f->cardinality.empty()) throw ReferenceError("NullPointerException: attribute " + std::to_string(f->ordinal) + " has no cardinality");
        if ((int)f->cardinality.size() > AVT_MAX_CLASSES)
            throw Unsupported("destination attribute " + std::to_string(f->ordinal) + " has more than " + std::to_string(AVT_MAX_CLASSES) + " values");
    }
    // source attributes in field order for the encoder; `pos` maps back to the configured order
    std::vector<size_t> by_ord(sf.size());
    for (size_t i = 0; i < by_ord.size(); ++i) by_ord[i] = i;
    std::sort(by_ord.begin(), by_ord.end(), [&](size_t a, size_t b) { return sf[a]->ordinal < sf[b]->ordinal; });
    for (size_t i = 1; i < by_ord.size(); ++i)
        if (sf[by_ord[i]]->ordinal == sf[by_ord[i - 1]]->ordinal) throw Unsupported("the source attribute set lists an attribute twice");
    std::vector<avt_csv_attr> attrs;
    std::string cat_bytes;
    std::vector<int32_t> cat_off, bin_off{0};
    for (size_t i : by_ord) {
        attrs.push_back(avt_csv_attr{sf[i]->ordinal, AVT_CSV_CATEGORICAL, (int32_t)sf[i]->cardinality.size(), (int32_t)cat_off.size()});
        for (auto& v : sf[i]->cardinality) {
            cat_off.push_back((int32_t)cat_bytes.size());
            cat_bytes += v;
        }
        bin_off.push_back(bin_off.back() + (int32_t)sf[i]->cardinality.size());
    }
    cat_off.push_back((int32_t)cat_bytes.size());
    const int TB = bin_off.back();

    std::vector<std::string> lines = read_input_lines(in_path);
    std::string text;
    for (auto& ln : lines) {
        text += ln;
        text += '\n';
    }
    struct Pair { int src, dst; std::string line; };
    std::vector<Pair> out;
    std::unique_ptr<Engine> eng;
    for (size_t di = 0; di < df.size(); ++di) {
        const Field& D = *df[di];
        const int C = (int)D.cardinality.size();
        std::vector<uint32_t> H((size_t)TB * C, 0);
        if (!text.empty()) {
            if (!eng) eng.reset(new Engine(conf.get_int("avt.device", 0)));
            std::string cls_bytes;
            std::vector<int32_t> cls_off;
            for (auto& v : D.cardinality) {
                cls_off.push_back((int32_t)cls_bytes.size());
                cls_bytes += v;
            }
            cls_off.push_back((int32_t)cls_bytes.size());
            int64_t n_rows = 0, bad_row = -1;
            int32_t bad_reason = 0;
            const int rc = avt_dataset_load_csv(eng->raw(), text.data(), (int64_t)text.size(), delim[0], 0, (int32_t)attrs.size(), attrs.data(), nullptr, nullptr,
                                                cat_bytes.data(), cat_off.data(), D.ordinal, C, cls_bytes.data(), cls_off.data(), &n_rows, nullptr, &bad_row, &bad_reason);
            if (rc == AVT_E_REFERENCE && bad_row >= 0)   // items[ordinal] beyond the row, or cardinalityIndex == -1 into the matrix
                throw ReferenceError("ArrayIndexOutOfBoundsException in CorrelationMapper.map at row " + std::to_string(bad_row) + ": " +
                                     (bad_row < (int64_t)lines.size() ? lines[(size_t)bad_row] : std::string()));
            eng->check(rc);
            eng->check(avt_forest_init(eng->raw(), 1, nullptr, nullptr));
            const int32_t one = 1;
            eng->check(avt_level_histogram(eng->raw(), &one, nullptr));
            eng->check(avt_level_fetch_hist(eng->raw(), 0, H.data()));
        }
        for (size_t k = 0; k < by_ord.size(); ++k) {
            const Field& S = *sf[by_ord[k]];
            const int R = (int)S.cardinality.size();
            std::vector<long long> t((size_t)R * C);
            for (int i = 0; i < R; ++i)
                for (int j = 0; j < C; ++j) t[(size_t)i * C + j] = H[((size_t)bin_off[k] + i) * C + j];
            out.push_back(Pair{S.ordinal, D.ordinal, S.name + delim_out + D.name + delim_out + java_double_to_string(K.stat(t, R, C))});
        }
    }
    // one reducer: keys arrive sorted, Tuple (src, dst) compares element by element
    std::stable_sort(out.begin(), out.end(), [](const Pair& a, const Pair& b) { return a.src != b.src ? a.src < b.src : a.dst < b.dst; });
    std::vector<std::string> out_lines;
    for (auto& p : out) out_lines.push_back(p.line);
    write_output_lines(out_path, out_lines);
    return 0;
}

}  // namespace

}  // namespace avt_host
