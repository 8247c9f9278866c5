// org.avenir.explore.CategoricalContinuousEncoding re-hosted on the histogram kernel (a count consumer of
// explore/).  The reference's mapper emits, per row and listed categorical attribute, the key (ordinal, value) with (1, 0) for a
// row of the positive class and (0, 1) otherwise (explore/CategoricalContinuousEncoding.java:115-148); combiner and reducer add
// them up (:155-175, 243-256) and the reducer's cleanup turns every (ordinal, value) into an integer code: supervised ratio
// pos * scale / total in int arithmetic, or weight of evidence (int)(log((pos / allPos) / (max(neg, 1) / allNeg)) * scale)
// (:198-236).  Those counts are kernel (a)'s root histogram H[value][class] with the class values as classes.
//
// The job has no schema: attribute and class values are whatever strings occur.  One host pass over the rows collects them in
// first-seen order (it looks at the listed fields only); the text is then encoded on the device against those lists
// (avt_dataset_load_csv) and counted by one histogram launch.
// Output order: the reference iterates a HashMap keyed by chombo Tuples, whose hashCode is not in the reference's tree -- the
// order is UNPINNED.  Lines are written in the reducer's key order (ordinal, then value, ascending); compare as a set.
// Math.log is the platform's log here (it is an intrinsic, not StrictMath, in the reference's JVMs too; the (int) cast hides last-bit differences
// except at exact integer boundaries).
#include <algorithm>
#include <cmath>
#include <cstring>
#include <map>
#include <unordered_map>

#include "javaemu.hpp"
#include "tool.hpp"

namespace avt_host {

namespace {

std::string require(const Props& conf, const char* key, const char* what) {
    if (!conf.nonblank(key)) throw ReferenceError(std::string("IllegalStateException: ") + what);   // chombo Utility.assert*ConfigParam
    return conf.get(key);
}

long long require_int(const Props& conf, const char* key, const char* what) {
    const std::string v = trim(require(conf, key, what));
    long long out;
    if (!parse_java_int(v, out)) throw ReferenceError("NumberFormatException: For input string: \"" + v + "\"");
    return out;
}

// the k-th field of a row split on a one-byte delimiter, trailing empty fields kept (String.split(regex, -1)); false: row too short
bool field_at(const std::string& row, char d, int k, size_t& b, size_t& e) {
    size_t pos = 0;
    for (int f = 0; f < k; ++f) {
        const size_t q = row.find(d, pos);
        if (q == std::string::npos) return false;
        pos = q + 1;
    }
    b = pos;
    e = row.find(d, pos);
    if (e == std::string::npos) e = row.size();
    return true;
}

struct ValueList {
    std::vector<std::string> names;
    std::unordered_map<std::string, int> index;
    void add(const std::string& s) {
        if (index.emplace(s, (int)names.size()).second) names.push_back(s);
    }
};

}  // namespace

int run_categorical_continuous_encoding(const Props& conf, const std::string& in_path, const std::string& out_path) {
    const std::string delim = conf.get("field.delim.regex", ",");
    const std::string delim_out = conf.get("field.delim", ",");
    std::vector<int> ordinals;
    for (auto& s : split_str(require(conf, "coe.cat.attribute.ordinals", "missing categorical attribute ordinals"), ",", false)) {
        long long v;
        if (!parse_java_int(trim(s), v)) throw ReferenceError("NumberFormatException: For input string: \"" + s + "\"");
        ordinals.push_back((int)v);
    }
    const bool woe = require(conf, "coe.encoding.strategy", "missing encoding strategy") == "weightOfEvidence";
    const int class_ord = (int)require_int(conf, "coe.class.attr.ordinal", "missing class atrribute ordinal");
    const std::string pos_value = require(conf, "coe.pos.class.attr.value", "missing positive class attribute value");
    const int scale = (int)require_int(conf, "coe.output.scale", "missing output scale");
    if (delim.size() != 1 || strchr(".*+?()[]{}|^$\\", delim[0])) throw Unsupported("field.delim.regex must be a single literal character for the device encoder");
    // attributes in field order for the encoder (each listed ordinal once); the output walks the sorted keys anyway
    std::vector<int> attrs = ordinals;
    std::sort(attrs.begin(), attrs.end());
    attrs.erase(std::unique(attrs.begin(), attrs.end()), attrs.end());
    if (attrs.empty()) throw ReferenceError("IllegalStateException: missing categorical attribute ordinals");
    for (int o : attrs)
        if (o < 0 || o >= 1024 || o == class_ord) throw Unsupported("attribute ordinal " + std::to_string(o) + ": outside 0..1023 or equal to the class ordinal");
    if (class_ord < 0 || class_ord >= 1024) throw Unsupported("class ordinal outside 0..1023");
    // a listed ordinal that occurs twice is counted twice by the reference's mapper
    std::map<int, int> times;
    for (int o : ordinals) ++times[o];

    std::vector<std::string> lines = read_input_lines(in_path);
    std::vector<ValueList> values(attrs.size());
    ValueList classes;
    const char d = delim[0];
    for (auto& ln : lines) {
        size_t b, e;
        if (!field_at(ln, d, class_ord, b, e)) throw ReferenceError("ArrayIndexOutOfBoundsException: " + std::to_string(class_ord));
        if (e == b) throw Unsupported("empty class value: String.split(regex, -1) keeps trailing empty fields, the device encoder's split does not");
        classes.add(ln.substr(b, e - b));
        for (size_t a = 0; a < attrs.size(); ++a) {
            if (!field_at(ln, d, attrs[a], b, e)) throw ReferenceError("ArrayIndexOutOfBoundsException: " + std::to_string(attrs[a]));
            if (e == b) throw Unsupported("empty attribute value: String.split(regex, -1) keeps trailing empty fields, the device encoder's split does not");
            values[a].add(ln.substr(b, e - b));
        }
    }
    std::vector<std::string> out;
    if (!lines.empty()) {
        if ((int)classes.names.size() > AVT_MAX_CLASSES) throw Unsupported("more than " + std::to_string(AVT_MAX_CLASSES) + " class values");
        std::vector<avt_csv_attr> spec;
        std::string cat_bytes, cls_bytes, text;
        std::vector<int32_t> cat_off, cls_off, bin_off{0};
        for (size_t a = 0; a < attrs.size(); ++a) {
            if (values[a].names.size() > 65535) throw Unsupported("attribute " + std::to_string(attrs[a]) + " has more than 65 535 values");
            spec.push_back(avt_csv_attr{attrs[a], AVT_CSV_CATEGORICAL, (int32_t)values[a].names.size(), (int32_t)cat_off.size()});
            for (auto& v : values[a].names) {
                cat_off.push_back((int32_t)cat_bytes.size());
                cat_bytes += v;
            }
            bin_off.push_back(bin_off.back() + (int32_t)values[a].names.size());
        }
        cat_off.push_back((int32_t)cat_bytes.size());
        for (auto& v : classes.names) {
            cls_off.push_back((int32_t)cls_bytes.size());
            cls_bytes += v;
        }
        cls_off.push_back((int32_t)cls_bytes.size());
        for (auto& ln : lines) {
            text += ln;
            text += '\n';
        }
        const int C = (int)classes.names.size();
        Engine eng(conf.get_int("avt.device", 0));
        int64_t n_rows = 0, bad_row = -1;
        int32_t bad_reason = 0;
        eng.check(avt_dataset_load_csv(eng.raw(), text.data(), (int64_t)text.size(), d, 0, (int32_t)spec.size(), spec.data(), nullptr, nullptr, cat_bytes.data(),
                                       cat_off.data(), class_ord, C, cls_bytes.data(), cls_off.data(), &n_rows, nullptr, &bad_row, &bad_reason));
        eng.check(avt_forest_init(eng.raw(), 1, nullptr, nullptr));
        const int32_t one = 1;
        std::vector<uint32_t> H((size_t)bin_off.back() * C);
        eng.check(avt_level_histogram(eng.raw(), &one, nullptr));
        eng.check(avt_level_fetch_hist(eng.raw(), 0, H.data()));
        const auto pc = classes.index.find(pos_value);
        const int pos_code = pc == classes.index.end() ? -1 : pc->second;
        // totals over all rows: the (-1, "*") key of the weight-of-evidence form (every row has exactly one value of the first attribute)
        long long all_pos = 0, all_neg = 0;
        for (int b = 0; b < (int)values[0].names.size(); ++b)
            for (int c = 0; c < C; ++c) (c == pos_code ? all_pos : all_neg) += H[(size_t)b * C + c];
        for (size_t a = 0; a < attrs.size(); ++a) {
            std::vector<int> idx(values[a].names.size());
            for (size_t i = 0; i < idx.size(); ++i) idx[i] = (int)i;
            std::sort(idx.begin(), idx.end(), [&](int x, int y) { return values[a].names[(size_t)x] < values[a].names[(size_t)y]; });
            for (int i : idx) {
                long long pos = 0, neg = 0;
                for (int c = 0; c < C; ++c) (c == pos_code ? pos : neg) += H[((size_t)bin_off[a] + i) * C + c];
                pos *= times[attrs[a]];
                neg *= times[attrs[a]];
                int32_t value;
                if (woe) {
                    double w = (double)(int32_t)pos / (double)(int32_t)all_pos;
                    const int32_t neg_count = (int32_t)neg == 0 ? 1 : (int32_t)neg;
                    w /= (double)neg_count / (double)(int32_t)all_neg;
                    w = std::log(w);
                    value = java_int_cast(w * (double)scale);
                } else {
                    // (posCount * scale) / totalCount in Java int arithmetic: the product wraps, the division truncates
                    const int32_t prod = (int32_t)((uint32_t)(int32_t)pos * (uint32_t)scale);
                    const int32_t total = (int32_t)(pos + neg);
                    value = total == -1 && prod == INT32_MIN ? INT32_MIN : prod / total;
                }
                out.push_back(std::to_string(attrs[a]) + delim_out + values[a].names[(size_t)i] + delim_out + std::to_string(value));
            }
        }
    }
    write_output_lines(out_path, out);
    return 0;
}

}  // namespace avt_host
