// org.avenir.explore.CategoricalClassAffinity re-hosted (SURVEY 8f row 4 names it among the count consumers of explore/).
//
// Unlike the other jobs of that row it does no per-row counting: its input is another job's output, one line per
// (feature ordinal, class value, feature value) with that value's class-conditional probability
// (explore/CategoricalClassAffinity.java:119-137), and its reducer scores every feature value's affinity to the positive class
// (odds ratio, distribution difference, minimum risk, KL term; :236-261).  There is nothing for a kernel here -- a few hundred
// numbers per feature -- so this is host code only; it is in the library because the reference's pipeline runs it between jobs
// that are.  What is kept exactly:
//   * keys (ordinal, 0 = positive class / 1 = any other) grouped by ordinal with the positive lines first
//     (Utility.setTuplePairSecondarySorting); inside a group the lines keep their input order;
//   * the two distribution maps are reducer FIELDS: map 0 is cleared at the first line of every group, map 1 only when a line's
//     class value differs from the previous line's -- a feature without a second class value is scored against the previous
//     feature's map 1, and with several non-positive class values only the last run of lines survives (:191-210);
//   * HashMap<String, Double> iteration order of the positive map (JDK 8 String hashes; clear() keeps the table length) decides
//     the order of equal scores (Collections.sort is stable); a feature value missing from map 1 is a NullPointerException; a
//     strategy that is none of the four leaves the score list empty and the job dies in substring(0, -1);
//   * one output record per strategy and group: "algorithm: <name>", then the group's lines joined by '\n'; Double.toString.
// Math.log is the platform's log.
#include <algorithm>
#include <cmath>
#include <cstring>

#include "javaemu.hpp"
#include "tool.hpp"

namespace avt_host {

int run_categorical_class_affinity(const Props& conf, const std::string& in_path, const std::string& out_path) {
    const std::string delim = conf.get("field.delim.regex", ",");
    const std::string delim_out = conf.get("field.delim", ",");
    if (!conf.nonblank("cca.pos.class.attr.value")) throw ReferenceError("IllegalStateException: missing positive class attribute value");
    const std::string pos_value = conf.get("cca.pos.class.attr.value");
    if (!conf.nonblank("cca.affinity.strategy")) throw ReferenceError("IllegalStateException: missing affinity strategy list");
    const std::vector<std::string> strategies = split_str(conf.get("cca.affinity.strategy"), ",", false);
    for (char c : delim)
        if (strchr(".*+?()[]{}|^$\\", c)) throw Unsupported("delimiter '" + delim + "' contains regex metacharacters; only literal delimiters are implemented");

    struct Line { int ord, flag; std::string cls, val; double distr; };
    std::vector<Line> recs;
    for (auto& ln : read_input_lines(in_path)) {
        const std::vector<std::string> items = split_str(ln, delim, true);
        long long ord;
        if (!parse_java_int(items[0], ord)) throw ReferenceError("NumberFormatException: For input string: \"" + items[0] + "\"");
        if (items.size() < 4) throw ReferenceError("ArrayIndexOutOfBoundsException: " + std::to_string(items.size()));
        const std::string t = trim(items[3]);
        char* end = nullptr;
        const double v = t.empty() ? 0.0 : strtod(t.c_str(), &end);
        if (t.empty() || *end) throw ReferenceError("NumberFormatException: For input string: \"" + items[3] + "\"");
        recs.push_back(Line{(int)ord, items[1] == pos_value ? 0 : 1, items[1], items[2], v});
    }
    std::stable_sort(recs.begin(), recs.end(), [](const Line& a, const Line& b) { return a.ord != b.ord ? a.ord < b.ord : a.flag < b.flag; });

    JStrMap<double> distr[2];   // reducer fields: they live across groups
    std::vector<std::string> out;
    size_t i = 0;
    while (i < recs.size()) {
        size_t j = i;
        while (j < recs.size() && recs[j].ord == recs[i].ord) ++j;
        const std::string* class_val = nullptr;
        JStrMap<double>* cur = nullptr;
        for (size_t k = i; k < j; ++k) {
            if (!class_val) {
                class_val = &recs[k].cls;
                cur = &distr[0];
                cur->clear();
            } else if (recs[k].cls != *class_val) {
                class_val = &recs[k].cls;
                cur = &distr[1];
                cur->clear();
            }
            cur->put(recs[k].val, recs[k].distr);
        }
        for (auto& strategy : strategies) {
            out.push_back("algorithm: " + strategy);
            std::vector<std::pair<std::string, double>> scores;
            for (auto& fv : distr[0].keys()) {
                const double p = *distr[0].find(fv);
                const double* np = distr[1].find(fv);
                if (!np) throw ReferenceError("NullPointerException: feature value '" + fv + "' of attribute " + std::to_string(recs[i].ord) + " has no distribution under the other class");
                const double n = *np;
                if (strategy == "oddsRatio") scores.emplace_back(fv, (p / (1 - p)) / (n / (1 - n)));
                else if (strategy == "distrDiff") scores.emplace_back(fv, p - n);
                else if (strategy == "minRisk") scores.emplace_back(fv, p * (1 - n));
                else if (strategy == "klDiff") scores.emplace_back(fv, p * std::log(p / n));
            }
            if (scores.empty()) throw ReferenceError("StringIndexOutOfBoundsException: String index out of range: -1");
            std::stable_sort(scores.begin(), scores.end(), [](const std::pair<std::string, double>& a, const std::pair<std::string, double>& b) { return a.second > b.second; });
            for (auto& sc : scores) out.push_back(std::to_string(recs[i].ord) + delim_out + sc.first + delim_out + java_double_to_string(sc.second));
        }
        i = j;
    }
    write_output_lines(out_path, out);
    return 0;
}

}  // namespace avt_host
