// Host-side model of the reference's configuration surface: .properties knobs, the JSON feature
// schema, candidate-split enumeration and the bin encoding that turns CSV rows into the columnar
// device layout.  Stand-ins for chombo's FeatureSchema/FeatureField/Utility follow SURVEY App. D.
#pragma once
#include <algorithm>
#include <functional>
#include <cstdint>
#include <fstream>
#include <map>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../../include/avenir_tree.h"
#include "javaemu.hpp"
#include "json.hpp"

namespace avt_host {

// An error the reference itself would raise on this input/configuration (job fails, exit 1).
struct ReferenceError : std::runtime_error {
    using std::runtime_error::runtime_error;
};
// A configuration outside what this engine implements yet.
struct Unsupported : std::runtime_error {
    using std::runtime_error::runtime_error;
};

inline std::string read_file(const std::string& path) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw std::runtime_error("cannot open " + path);
    std::stringstream ss;
    ss << f.rdbuf();
    return ss.str();
}

inline std::string trim(const std::string& s) {
    size_t a = s.find_first_not_of(" \t\r\n"), b = s.find_last_not_of(" \t\r\n");
    return a == std::string::npos ? "" : s.substr(a, b - a + 1);
}

inline std::vector<std::string> split_str(const std::string& s, const std::string& delim, bool keep_trailing) {
    std::vector<std::string> out;
    size_t pos = 0;
    for (;;) {
        size_t q = s.find(delim, pos);
        if (q == std::string::npos) { out.push_back(s.substr(pos)); break; }
        out.push_back(s.substr(pos, q - pos));
        pos = q + delim.size();
    }
    if (!keep_trailing) {  // Java String.split(regex) drops trailing empty strings
        while (out.size() > 1 && out.back().empty()) out.pop_back();
    }
    return out;
}

inline std::string join_str(const std::vector<std::string>& v, const std::string& d) {
    std::string o;
    for (size_t i = 0; i < v.size(); ++i) {
        if (i) o += d;
        o += v[i];
    }
    return o;
}

// Hadoop Configuration populated from a .properties file (-Dconf.path) plus -Dkey=value overrides
// (chombo Utility.setConfiguration, tree/DecisionTreeBuilder.java:75).
class Props {
    std::map<std::string, std::string> kv_;

public:
    void load(const std::string& path) {
        std::istringstream in(read_file(path));
        std::string line;
        while (std::getline(in, line)) {
            std::string s = trim(line);
            if (s.empty() || s[0] == '#' || s[0] == '!') continue;
            size_t p = s.find_first_of("=:");
            std::string k = trim(p == std::string::npos ? s : s.substr(0, p));
            std::string v = p == std::string::npos ? "" : trim(s.substr(p + 1));
            kv_[k] = v;
        }
    }
    void set(const std::string& k, const std::string& v) { kv_[k] = v; }
    bool has(const std::string& k) const { auto it = kv_.find(k); return it != kv_.end(); }
    std::string get(const std::string& k, const std::string& def = "") const {
        auto it = kv_.find(k);
        return it == kv_.end() ? def : it->second;
    }
    bool nonblank(const std::string& k) const { return has(k) && !get(k).empty(); }
    int get_int(const std::string& k, int def) const {
        if (!nonblank(k)) return def;
        try { return std::stoi(get(k)); } catch (...) { throw ReferenceError("NumberFormatException: " + k + "=" + get(k)); }
    }
    double get_double(const std::string& k, double def) const {
        if (!nonblank(k)) return def;
        try { return std::stod(get(k)); } catch (...) { throw ReferenceError("NumberFormatException: " + k + "=" + get(k)); }
    }
    bool get_bool(const std::string& k, bool def) const { return has(k) ? get(k) == "true" : def; }
};

struct Field {
    std::string name, data_type;
    int ordinal = 0;
    bool is_id = false, feature = false;
    double min = 0, max = 0, split_scan_interval = 0;
    int max_split = 0, bucket_width = 0;
    std::vector<std::string> cardinality;
    bool is_int() const { return data_type == "int"; }
    bool is_double() const { return data_type == "double"; }
    bool is_categorical() const { return data_type == "categorical"; }
};

struct Schema {
    std::vector<Field> fields;
    static Schema load(const std::string& path) {
        Json j = JsonParser(read_file(path)).parse();
        const Json* fs = j.get("fields");
        if (!fs || fs->kind != Json::Arr) throw std::runtime_error("schema " + path + ": no \"fields\" array");
        Schema s;
        for (auto& f : fs->arr) {
            Field x;
            if (auto* v = f.get("name")) x.name = v->str;
            if (auto* v = f.get("ordinal")) x.ordinal = (int)v->num;
            if (auto* v = f.get("id")) x.is_id = v->b;
            if (auto* v = f.get("feature")) x.feature = v->b;
            if (auto* v = f.get("dataType")) x.data_type = v->str;
            if (auto* v = f.get("min")) x.min = v->num;
            if (auto* v = f.get("max")) x.max = v->num;
            if (auto* v = f.get("splitScanInterval")) x.split_scan_interval = v->num;
            if (auto* v = f.get("maxSplit")) x.max_split = (int)v->num;
            if (auto* v = f.get("bucketWidth")) x.bucket_width = (int)v->num;
            if (auto* v = f.get("cardinality"))
                for (auto& c : v->arr) x.cardinality.push_back(c.str);
            s.fields.push_back(x);
        }
        return s;
    }
    const Field& by_ordinal(int o) const {
        for (auto& f : fields)
            if (f.ordinal == o) return f;
        throw ReferenceError("NullPointerException: no schema field with ordinal " + std::to_string(o));
    }
    std::vector<int> feature_ordinals() const {
        std::vector<int> o;
        for (auto& f : fields)
            if (f.feature) o.push_back(f.ordinal);
        return o;
    }
    const Field& class_field() const {
        for (auto& f : fields)
            if (!f.is_id && !f.feature) return f;
        throw ReferenceError("NullPointerException: schema has no class attribute field");
    }
};

// ---- candidate splits -------------------------------------------------------------------------

struct HostSplit {
    int n_seg = 2;
    std::vector<int> bounds;                       // numeric: inclusive upper bin of segments 0..n_seg-2
    std::vector<std::pair<int, int>> ranges;       // DecisionTreeBuilder multi-way numeric (maxSplit > 2): inclusive bin range of every
                                                   // predicate; ranges may overlap (tree/SplitManager.java:635-694)
    std::vector<uint8_t> lut;                      // categorical: segment of each bin
    std::vector<std::string> pred_str;             // DecisionTreeBuilder: predicate string per segment
    std::string key;                               // ClassPartitionGenerator: split key
    std::vector<std::vector<std::string>> groups;  // categorical groups
    std::vector<std::string> points;               // numeric split points as printed
};

struct AttrSplits {
    int ordinal = 0;
    int n_bins = 0;
    bool numeric = false, dbl = false;
    std::vector<long long> int_points;  // bin(v) = #{k : p_k < v}
    std::vector<double> dbl_points;
    std::vector<HostSplit> splits;
    const Field* field = nullptr;
};

inline int java_int_cast(double x) {
    if (x != x) return 0;
    if (x >= 2147483647.0) return 2147483647;
    if (x <= -2147483648.0) return -2147483647 - 1;
    return (int)x;
}

// createCategoricalPartitions (tree/SplitManager.java:452-515) and its ClassPartitionGenerator twin
// (explore/ClassPartitionGenerator.java:322-390, which lacks the size > numGroups guard).
using CatSplit = std::vector<std::vector<std::string>>;
void cat_partitions(std::vector<CatSplit>& list, const std::vector<std::string>& card, size_t idx, int groups, bool guard);

inline std::vector<CatSplit> cat_partial(const std::vector<std::string>& card, size_t idx, int groups, bool guard) {
    std::vector<CatSplit> out;
    if (groups == 2) {
        CatSplit sp(1);
        for (size_t i = 0; i <= idx; ++i) sp[0].push_back(card.at(i));
        out.push_back(sp);
    } else {
        std::vector<std::string> part(card.begin(), card.begin() + std::min(card.size(), idx + 1));
        cat_partitions(out, part, 0, groups - 1, guard);
    }
    return out;
}

inline void cat_partitions(std::vector<CatSplit>& list, const std::vector<std::string>& card, size_t idx, int groups, bool guard) {
    if (idx == 0) {
        if ((int)card.size() < groups) throw ReferenceError("IndexOutOfBoundsException: cardinality smaller than number of split groups");
        CatSplit full;
        for (int i = 0; i < groups; ++i) full.push_back({card[(size_t)i]});
        if (guard) {
            list.push_back(full);
            if ((int)card.size() > groups) {
                auto partial = cat_partial(card, (size_t)groups - 1, groups, guard);
                list.insert(list.end(), partial.begin(), partial.end());
                cat_partitions(list, card, idx + (size_t)groups, groups, guard);
            }
        } else {
            auto partial = cat_partial(card, (size_t)groups - 1, groups, guard);
            list.push_back(full);
            list.insert(list.end(), partial.begin(), partial.end());
            cat_partitions(list, card, idx + (size_t)groups, groups, guard);
        }
    } else if (idx < card.size()) {
        std::vector<CatSplit> nl;
        const std::string& el = card[idx];
        for (auto& sp : list) {
            if ((int)sp.size() == groups) {
                for (int i = 0; i < groups; ++i) {
                    CatSplit ns = sp;
                    ns[(size_t)i].push_back(el);
                    nl.push_back(std::move(ns));
                }
            } else {
                CatSplit ns = sp;
                ns.push_back({el});
                nl.push_back(std::move(ns));
            }
            if (nl.size() > 4000000) throw Unsupported("categorical split enumeration exceeds 4M candidates (cardinality too high for exhaustive set partitions)");
        }
        if (idx + 1 < card.size()) {
            auto partial = cat_partial(card, idx, groups, guard);
            nl.insert(nl.end(), partial.begin(), partial.end());
        }
        list.swap(nl);
        cat_partitions(list, card, idx + 1, groups, guard);
    }
}

inline std::vector<uint8_t> groups_to_lut(const Field& f, const CatSplit& groups) {
    std::vector<uint8_t> lut(f.cardinality.size(), 0xFF);
    for (size_t g = 0; g < groups.size(); ++g)
        for (auto& v : groups[g])
            for (size_t b = 0; b < f.cardinality.size(); ++b)
                if (f.cardinality[b] == v && lut[b] == 0xFF) lut[b] = (uint8_t)g;  // first containing group wins
    return lut;
}

// createIntPartitions / createDoublePartitions + createIntAttrPredicates / createDoubleAttrPredicates
// (tree/SplitManager.java:288-330, 366-399, 635-694): every increasing tuple of up to `max_points` grid points in DFS order
// ([p0], [p0,p1], ..., [p1], ...), and its predicates.  One point: `le p`, `gt p`.  Several points, int: `le p0`,
// `le p_i p_{i-1}` (lower bound p_{i-1}) for the middle points, then `le p_last` WITHOUT a lower bound, `gt p_last`; double: the
// four-argument DoublePredicate constructor never stores the other bound (:840-843), so every `le` is unbounded and prints
// without it.  Bins: value <= p_k  <=>  bin <= k.
inline void numeric_split_lists(AttrSplits& as, const std::string& ord, const std::vector<std::string>& pts, int max_points, bool dbl) {
    const int K = (int)pts.size(), B = K + 1;
    std::vector<int> cur;
    std::function<void(int)> rec = [&](int first) {
        for (int k = first; k < K; ++k) {
            cur.push_back(k);
            HostSplit h;
            for (int i : cur) h.points.push_back(pts[(size_t)i]);
            const int m = (int)cur.size();
            if (m == 1) {
                h.n_seg = 2;
                h.bounds = {k};
                h.pred_str = {ord + " le " + pts[(size_t)k], ord + " gt " + pts[(size_t)k]};
            } else {
                h.n_seg = m + 1;
                for (int i = 0; i < m; ++i) {
                    const int pi = cur[(size_t)i];
                    if (i == m - 1) {
                        h.pred_str.push_back(ord + " le " + pts[(size_t)pi]);
                        h.ranges.emplace_back(0, pi);
                        h.pred_str.push_back(ord + " gt " + pts[(size_t)pi]);
                        h.ranges.emplace_back(pi + 1, B - 1);
                    } else if (i == 0) {
                        h.pred_str.push_back(ord + " le " + pts[(size_t)pi]);
                        h.ranges.emplace_back(0, pi);
                    } else {
                        // int: `attr le p_i p_{i-1}`; double: otherBound is never stored, so neither evaluate() nor toString() sees it
                        h.pred_str.push_back(dbl ? ord + " le " + pts[(size_t)pi] : ord + " le " + pts[(size_t)pi] + " " + pts[(size_t)cur[(size_t)i - 1]]);
                        h.ranges.emplace_back(dbl ? 0 : cur[(size_t)i - 1] + 1, pi);
                    }
                }
            }
            as.splits.push_back(h);
            if (as.splits.size() > 4000000) throw Unsupported("numeric split enumeration exceeds 4M candidates");
            if (m < max_points) rec(k + 1);
            cur.pop_back();
        }
    };
    rec(0);
}

// DecisionTreeBuilder candidate splits of one attribute (tree/SplitManager.java:256-445, 635-694).
inline AttrSplits dtb_attr_splits(const Field& f) {
    AttrSplits as;
    as.ordinal = f.ordinal;
    as.field = &f;
    const std::string ord = std::to_string(f.ordinal);
    if (f.is_int() || f.is_double()) {
        as.numeric = true;
        as.dbl = f.is_double();
        if (f.max_split - 1 > AVT_MAX_SEG - 2) throw Unsupported("attribute " + ord + ": maxSplit above " + std::to_string(AVT_MAX_SEG - 1));
        double itv = f.split_scan_interval;
        if (java_int_cast((f.max - f.min) / itv) == 0) itv = (f.max - f.min) / 2;
        if (f.is_int()) {
            int sp = java_int_cast(f.min + itv);
            while (sp < f.max) {
                as.int_points.push_back(sp);
                int nx = java_int_cast(sp + itv);  // `int split; split += double` truncates each step (App. B-8)
                if (nx <= sp) throw ReferenceError("attribute " + ord + ": splitScanInterval too small, the reference's split loop never terminates");
                sp = nx;
                if (as.int_points.size() > 65534) throw Unsupported("attribute " + ord + ": more than 65534 split points");
            }
            as.n_bins = (int)as.int_points.size() + 1;
            std::vector<std::string> pts;
            for (long long v : as.int_points) pts.push_back(std::to_string(v));
            numeric_split_lists(as, ord, pts, std::max(f.max_split, 2) - 1, false);
        } else {
            double sp = f.min + itv;
            while (sp < f.max) {
                as.dbl_points.push_back(sp);
                double nx = sp + itv;
                if (!(nx > sp)) throw ReferenceError("attribute " + ord + ": splitScanInterval too small, the reference's split loop never terminates");
                sp = nx;
                if (as.dbl_points.size() > 65534) throw Unsupported("attribute " + ord + ": more than 65534 split points");
            }
            as.n_bins = (int)as.dbl_points.size() + 1;
            std::vector<std::string> pts;
            for (double v : as.dbl_points) pts.push_back(java_double_to_string(v));
            numeric_split_lists(as, ord, pts, std::max(f.max_split, 2) - 1, true);
        }
    } else if (f.is_categorical()) {
        if (f.max_split > AVT_MAX_SEG) throw Unsupported("attribute " + ord + ": maxSplit above " + std::to_string(AVT_MAX_SEG));
        if (f.cardinality.empty()) throw ReferenceError("attribute " + ord + ": categorical attribute without cardinality");
        as.n_bins = (int)f.cardinality.size();
        for (int gr = 2; gr <= f.max_split; ++gr) {
            std::vector<CatSplit> list;
            cat_partitions(list, f.cardinality, 0, gr, true);
            for (auto& groups : list) {
                HostSplit h;
                h.n_seg = (int)groups.size();
                h.groups = groups;
                h.lut = groups_to_lut(f, groups);
                for (auto& g : groups) h.pred_str.push_back(ord + " in " + join_str(g, ":"));
                as.splits.push_back(h);
            }
        }
    } else {
        throw ReferenceError("NullPointerException: attribute " + ord + " of type '" + f.data_type + "' has no split predicates");
    }
    return as;
}

// ClassPartitionGenerator candidate splits (explore/ClassPartitionGenerator.java:235-443;
// util/AttributeSplitHandler.java:44, 148-155, 192-215).
inline void cpg_num_rec(const Field& f, int mn, int mx, int bw, std::vector<int>* cur, std::vector<std::vector<int>>& out) {
    if (!cur) {
        for (int sp = mn + bw; sp < mx; sp += bw) {
            std::vector<int> ns{sp};
            out.push_back(ns);
            cpg_num_rec(f, mn, mx, bw, &ns, out);
        }
    } else if ((int)cur->size() < f.max_split - 1) {
        for (int sp = cur->back() + bw; sp < mx; sp += bw) {
            std::vector<int> ns = *cur;
            ns.push_back(sp);
            out.push_back(ns);
            cpg_num_rec(f, mn, mx, bw, &ns, out);
            if (out.size() > 4000000) throw Unsupported("numeric split enumeration exceeds 4M candidates");
        }
    }
}

inline AttrSplits cpg_attr_splits(const Field& f, int max_cat_groups) {
    AttrSplits as;
    as.ordinal = f.ordinal;
    as.field = &f;
    const std::string ord = std::to_string(f.ordinal);
    if (f.is_int()) {
        as.numeric = true;
        const int mn = java_int_cast(f.min + 0.01), mx = java_int_cast(f.max + 0.01), bw = f.bucket_width;
        if (bw <= 0 && mn < mx) throw ReferenceError("attribute " + ord + ": bucketWidth 0, the reference's split loop never terminates");
        if (f.max_split - 1 > AVT_MAX_SEG - 1) throw Unsupported("attribute " + ord + ": maxSplit above " + std::to_string(AVT_MAX_SEG));
        std::vector<std::vector<int>> lists;
        cpg_num_rec(f, mn, mx, bw, nullptr, lists);
        for (int sp = mn + bw; sp < mx; sp += bw) as.int_points.push_back(sp);  // the grid = all single points
        as.n_bins = (int)as.int_points.size() + 1;
        for (auto& pts : lists) {
            HostSplit h;
            h.n_seg = (int)pts.size() + 1;
            std::vector<std::string> ps;
            for (int p : pts) {
                // segment index = #points < value  <=>  bins <= (grid index of p) fall left of p
                auto it = std::lower_bound(as.int_points.begin(), as.int_points.end(), (long long)p);
                h.bounds.push_back((int)(it - as.int_points.begin()));
                ps.push_back(std::to_string(p));
            }
            h.points = ps;
            h.key = join_str(ps, ";");
            as.splits.push_back(h);
        }
    } else if (f.is_categorical()) {
        if (f.max_split > max_cat_groups)
            throw ReferenceError("more than " + std::to_string(max_cat_groups) + " split groups not allwed for categorical attr");
        if (f.max_split > AVT_MAX_SEG) throw Unsupported("attribute " + ord + ": maxSplit above " + std::to_string(AVT_MAX_SEG));
        as.n_bins = (int)f.cardinality.size();
        for (int gr = 2; gr <= f.max_split; ++gr) {
            std::vector<CatSplit> list;
            cat_partitions(list, f.cardinality, 0, gr, false);
            for (auto& groups : list) {
                HostSplit h;
                h.n_seg = std::max<int>(2, (int)groups.size());
                h.groups = groups;
                h.lut = groups_to_lut(f, groups);
                std::vector<std::string> gs;
                for (auto& g : groups) gs.push_back("[" + join_str(g, ", ") + "]");  // java.util.List.toString
                h.key = join_str(gs, ":");
                as.splits.push_back(h);
            }
        }
    } else {
        throw ReferenceError("NullPointerException: ClassPartitionGenerator has no split list for attribute " + ord + " of type '" + f.data_type + "'");
    }
    return as;
}

// ---- encoding ---------------------------------------------------------------------------------

inline bool parse_java_int(const std::string& s, long long& out) {
    if (s.empty()) return false;
    size_t i = 0;
    bool neg = false;
    if (s[0] == '-' || s[0] == '+') { neg = s[0] == '-'; i = 1; }
    if (i >= s.size()) return false;
    long long v = 0;
    for (; i < s.size(); ++i) {
        if (s[i] < '0' || s[i] > '9') return false;
        v = v * 10 + (s[i] - '0');
        if (v > 2147483648LL) return false;
    }
    v = neg ? -v : v;
    if (v > 2147483647LL || v < -2147483648LL) return false;
    out = v;
    return true;
}

// bin of a raw CSV value for one attribute; throws what the reference would throw.
inline int encode_value(const AttrSplits& as, const std::string& raw) {
    if (as.numeric) {
        if (!as.dbl) {
            long long v;
            if (!parse_java_int(raw, v)) throw ReferenceError("NumberFormatException: For input string: \"" + raw + "\"");
            return (int)(std::lower_bound(as.int_points.begin(), as.int_points.end(), v) - as.int_points.begin());
        }
        char* end = nullptr;
        std::string t = trim(raw);
        double v = strtod(t.c_str(), &end);
        if (t.empty() || *end) throw ReferenceError("NumberFormatException: For input string: \"" + raw + "\"");
        return (int)(std::lower_bound(as.dbl_points.begin(), as.dbl_points.end(), v) - as.dbl_points.begin());
    }
    const auto& card = as.field->cardinality;
    for (size_t b = 0; b < card.size(); ++b)
        if (card[b] == raw) return (int)b;
    throw ReferenceError("no matching predicate for attribute: " + std::to_string(as.ordinal));
}

// Flattened split table in the C-ABI's form.
struct DeviceSplitTable {
    std::vector<avt_split> splits;
    std::vector<int32_t> bounds;
    std::vector<uint8_t> lut;
    std::vector<int> attr_first;  // [A+1]
};

inline DeviceSplitTable flatten_splits(const std::vector<AttrSplits>& attrs) {
    DeviceSplitTable t;
    for (size_t a = 0; a < attrs.size(); ++a) {
        t.attr_first.push_back((int)t.splits.size());
        for (auto& h : attrs[a].splits) {
            avt_split s;
            s.attr = (int)a;
            s.n_seg = h.n_seg;
            if (!h.ranges.empty()) {
                s.kind = AVT_SPLIT_RANGES;
                s.off = (int)t.bounds.size();
                for (auto& r : h.ranges) { t.bounds.push_back(r.first); t.bounds.push_back(r.second); }
            } else if (attrs[a].numeric) {
                s.kind = AVT_SPLIT_NUMERIC;
                s.off = (int)t.bounds.size();
                for (int b : h.bounds) t.bounds.push_back(b);
            } else {
                s.kind = AVT_SPLIT_LUT;
                s.off = (int)t.lut.size();
                for (uint8_t v : h.lut) t.lut.push_back(v);
            }
            t.splits.push_back(s);
        }
    }
    t.attr_first.push_back((int)t.splits.size());
    return t;
}

// does the bin fall in segment `sg` of the split (segments of a multi-way numeric split may overlap)
inline bool split_bin_in_segment(const AttrSplits& as, const HostSplit& h, int bin, int sg);

// segment a bin falls in for a host split (mirror of the device rule); for overlapping ranges: the first one
inline int split_segment_of_bin(const AttrSplits& as, const HostSplit& h, int bin) {
    if (!h.ranges.empty()) {
        for (int sg = 0; sg < (int)h.ranges.size(); ++sg)
            if (bin >= h.ranges[(size_t)sg].first && bin <= h.ranges[(size_t)sg].second) return sg;
        return -1;
    }
    if (as.numeric) {
        int j = 0;
        while (j < (int)h.bounds.size() && bin > h.bounds[(size_t)j]) ++j;
        return j;
    }
    return h.lut[(size_t)bin] == 0xFF ? -1 : h.lut[(size_t)bin];
}

inline bool split_bin_in_segment(const AttrSplits& as, const HostSplit& h, int bin, int sg) {
    if (!h.ranges.empty()) return bin >= h.ranges[(size_t)sg].first && bin <= h.ranges[(size_t)sg].second;
    return split_segment_of_bin(as, h, bin) == sg;
}

}  // namespace avt_host
