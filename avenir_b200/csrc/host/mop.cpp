// ModelPredictor re-hosted above the kernel-level C-ABI (model/ModelPredictor.java; tree/DecisionTreeModel.java;
// model/ProbabilisticPredictiveModel.java; model/EnsemblePredictiveModel.java; model/PredictiveModel.java).
//
// Same mop.* knobs, schema JSON, CSV rows and DecisionPathList model files in; `record,predictedClass` (or
// id / actual-class variants) lines out, plus the error-rate file.  The rows are encoded ONCE against the
// models' own thresholds and category lists, every model becomes per-(attribute, bin) admission masks over its
// decision paths, and avt_predict finds each row's first matching path per model and the ensemble vote.
//
// Differences worth knowing: every attribute a model mentions is parsed for every row (the reference parses an
// attribute only when a predicate on it is reached, so a malformed value in a never-reached attribute passes
// there and fails here); BasicUtils.formatDouble (chombo, absent) is assumed to print three decimals.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <memory>
#include <set>

#include "tool.hpp"

namespace avt_host {

namespace {

struct MPred {
    int attr = 0;
    std::string op;
    long long vint = 0;
    double vdbl = 0;
    std::vector<std::string> cats;
    bool has_other_int = false;
    long long other_int = 0;
};

struct MPath {
    std::vector<MPred> preds;                              // without the root pseudo-predicate
    std::vector<std::pair<std::string, double>> class_pr;  // file order (Jackson -> LinkedHashMap)
};

struct MModel {
    std::vector<MPath> paths;
};

MModel load_model(const std::string& path) {
    Json j = JsonParser(read_file(path)).parse();
    const Json* arr = j.get("decisionPaths");
    if (!arr || arr->kind != Json::Arr) throw ReferenceError("NullPointerException: model " + path + " has no decisionPaths");
    MModel m;
    for (auto& pj : arr->arr) {
        MPath p;
        const Json* preds = pj.get("predicates");
        if (!preds || preds->kind != Json::Arr)
            throw ReferenceError("NullPointerException: a decision path without predicates (root-only model) cannot predict");
        for (size_t i = 1; i < preds->arr.size(); ++i) {   // predicate 0 is the $root pseudo-predicate (tree/DecisionTreeModel.java:64)
            const Json& q = preds->arr[i];
            MPred r;
            if (auto* x = q.get("attribute")) r.attr = (int)x->num;
            if (auto* x = q.get("operator"); x && x->kind == Json::Str) r.op = x->str;
            if (auto* x = q.get("valueInt")) r.vint = (long long)x->num;
            if (auto* x = q.get("valueDbl")) r.vdbl = x->num;
            if (auto* x = q.get("categoricalValues"); x && x->kind == Json::Arr)
                for (auto& c : x->arr) r.cats.push_back(c.str);
            if (auto* x = q.get("otherBoundInt"); x && x->kind == Json::Num) { r.has_other_int = true; r.other_int = (long long)x->num; }
            p.preds.push_back(r);
        }
        if (auto* v = pj.get("classValPr"))
            for (auto& kv : v->obj) p.class_pr.emplace_back(kv.first, kv.second.num);
        m.paths.push_back(p);
    }
    if (m.paths.empty()) throw ReferenceError("NullPointerException: model " + path + " has no decision path");
    return m;
}

// bin axis of one attribute, built from what the models test on it
struct Axis {
    int ordinal = 0;
    const Field* field = nullptr;
    std::vector<long long> ithr;    // int thresholds, sorted: bin = #{t < x}
    std::vector<double> dthr;       // double thresholds
    std::vector<std::string> cats;  // categorical values mentioned; bin = index, cats.size() = "any other value"
    int n_bins() const {
        if (field->is_int()) return (int)ithr.size() + 1;
        if (field->is_double()) return (int)dthr.size() + 1;
        return (int)cats.size() + 1;
    }
};

std::string format_double3(double v) {
    char b[64];
    snprintf(b, sizeof b, "%.3f", v);
    return b;
}

}  // namespace

int run_model_predictor(const Props& conf, const std::string& in_path, const std::string& out_path) {
    const std::string delim_regex = conf.get("field.delim.regex", ",");
    const std::string delim_out = conf.get("field.delim.out", ",");
    for (char c : delim_regex)
        if (strchr(".*+?()[]{}|^$\\", c)) throw Unsupported("delimiter '" + delim_regex + "' contains regex metacharacters; only literal delimiters are implemented");
    if (!conf.nonblank("mop.feature.schema.file.path")) throw ReferenceError("NullPointerException: mop.feature.schema.file.path not set");
    Schema schema = Schema::load(conf.get("mop.feature.schema.file.path"));
    if (!conf.nonblank("mop.model.dir.path")) throw ReferenceError("missing model directory path");
    if (!conf.nonblank("mop.model.file.names")) throw ReferenceError("missing mode file names");
    if (!conf.nonblank("mop.classifier.type")) throw ReferenceError("missing classifier type");
    if (conf.get("mop.classifier.type") != "decTreeClassifier") throw ReferenceError("invalid classifier type");
    std::vector<std::string> model_files;
    for (auto& s : split_str(conf.get("mop.model.file.names"), ",", false)) model_files.push_back(trim(s));
    const bool error_counting = conf.get_bool("mop.error.counting.enabled", false);
    const bool cost_based = conf.get_bool("mop.cost.based.prediction.enabled", false);
    int class_attr_ord = -1;
    if (error_counting) {
        if (!conf.nonblank("mop.class.attr.ord")) throw ReferenceError("missing mop.class.attr.ord");
        class_attr_ord = conf.get_int("mop.class.attr.ord", -1);
    }
    std::vector<std::string> class_values;
    double fp_cost = 0, fn_cost = 0;
    if (cost_based || error_counting) {
        if (!conf.nonblank("mop.class.attr.values")) throw ReferenceError("missing class atrribute values, need for for cost based prediction");
        for (auto& s : split_str(conf.get("mop.class.attr.values"), ",", false)) class_values.push_back(trim(s));
        if (class_values.size() > 2) throw ReferenceError("cost based classification possible only for binary classification");
        if (class_values.size() < 2) throw ReferenceError("ArrayIndexOutOfBoundsException: mop.class.attr.values needs two values");
        if (!conf.nonblank("map.error.rate.file.path")) throw ReferenceError("missing error file path");
        if (cost_based) {
            if (!conf.nonblank("mop.miss.class.costs")) throw ReferenceError("missing misclassification costs");
            auto cs = split_str(conf.get("mop.miss.class.costs"), ",", false);
            if (cs.size() < 2) throw ReferenceError("ArrayIndexOutOfBoundsException: mop.miss.class.costs needs two values");
            fp_cost = std::stod(cs[0]);
            fn_cost = std::stod(cs[1]);
        }
    }
    const bool ensemble = model_files.size() > 1;
    if (ensemble && model_files.size() % 2 == 0) throw ReferenceError("neem odd number of models in ensemble");
    if (ensemble && error_counting)
        throw ReferenceError("NullPointerException: EnsemblePredictiveModel.predict never sets `items`, so countError() fails (model/EnsemblePredictiveModel.java:69-110)");
    std::vector<double> weights(model_files.size(), 1.0);
    if (ensemble && conf.nonblank("mop.ensemble.memeber.weights")) {
        auto ws = split_str(conf.get("mop.ensemble.memeber.weights"), ",", false);
        if (ws.size() < model_files.size()) throw ReferenceError("ArrayIndexOutOfBoundsException: mop.ensemble.memeber.weights");
        for (size_t i = 0; i < model_files.size(); ++i) weights[i] = std::stod(ws[i]);
    }
    const std::string output_mode = conf.get("mop.output.mode", "withRecord");
    int id_ord = -1, out_class_ord = -1;
    if (output_mode == "withKId") {
        if (!conf.nonblank("mop.rec.id.ordinal")) throw ReferenceError("missing id ordinal");
        id_ord = conf.get_int("mop.rec.id.ordinal", -1);
    } else if (output_mode == "withActualClassAttr") {
        if (!conf.nonblank("mop.rec.class.attr.ordinal")) throw ReferenceError("missing class attribute ordinal");
        out_class_ord = conf.get_int("mop.rec.class.attr.ordinal", -1);
    } else if (output_mode != "withRecord") {
        throw ReferenceError("invalid output mode");
    }

    std::vector<MModel> models;
    for (auto& f : model_files) models.push_back(load_model(conf.get("mop.model.dir.path") + "/" + f));

    // ---- bin axes: one per attribute any model tests
    std::vector<Axis> axes;
    auto axis_of = [&](int ord) -> Axis& {
        for (auto& a : axes)
            if (a.ordinal == ord) return a;
        Axis a;
        a.ordinal = ord;
        a.field = &schema.by_ordinal(ord);
        if (!a.field->is_int() && !a.field->is_double() && !a.field->is_categorical())
            throw Unsupported("attribute " + std::to_string(ord) + " of type '" + a.field->data_type + "' in a model predicate");
        axes.push_back(a);
        return axes.back();
    };
    for (auto& m : models)
        for (auto& p : m.paths)
            for (auto& q : p.preds) {
                Axis& ax = axis_of(q.attr);
                if (ax.field->is_int()) {
                    if (q.op != "le" && q.op != "gt") throw ReferenceError("invalid operator type for int attribute");
                    ax.ithr.push_back(q.vint);
                    if (q.op == "gt" && q.has_other_int) ax.ithr.push_back(q.other_int);
                } else if (ax.field->is_double()) {
                    if (q.op != "le" && q.op != "gt") throw ReferenceError("invalid operator type for double attribute");
                    ax.dthr.push_back(q.vdbl);
                    if (q.op == "gt" && q.has_other_int) ax.dthr.push_back((double)q.other_int);   // the reference reads otherBoundInt here too
                } else {
                    if (q.op != "in") throw ReferenceError("invalid operator type for categorical attribute");
                    for (auto& c : q.cats)
                        if (std::find(ax.cats.begin(), ax.cats.end(), c) == ax.cats.end()) ax.cats.push_back(c);
                }
            }
    if (axes.empty()) {   // models whose paths carry no predicate beyond $root: every row matches the first path
        Axis a;
        a.ordinal = -1;
        static Field dummy;
        dummy.data_type = "categorical";
        a.field = &dummy;
        axes.push_back(a);
    }
    std::sort(axes.begin(), axes.end(), [](const Axis& x, const Axis& y) { return x.ordinal < y.ordinal; });
    std::vector<int32_t> n_bins, bin_off{0};
    for (auto& ax : axes) {
        std::sort(ax.ithr.begin(), ax.ithr.end());
        ax.ithr.erase(std::unique(ax.ithr.begin(), ax.ithr.end()), ax.ithr.end());
        std::sort(ax.dthr.begin(), ax.dthr.end());
        ax.dthr.erase(std::unique(ax.dthr.begin(), ax.dthr.end()), ax.dthr.end());
        if (ax.n_bins() > 65535) throw Unsupported("more than 65534 distinct thresholds on attribute " + std::to_string(ax.ordinal));
        n_bins.push_back(ax.n_bins());
        bin_off.push_back(bin_off.back() + ax.n_bins());
    }
    const int A = (int)axes.size(), TB = bin_off.back();

    // ---- output classes, the class each path predicts (DecisionPath.doPrediction + cost-based threshold)
    std::vector<std::string> out_classes;
    auto class_code = [&](const std::string& s) {
        for (size_t i = 0; i < out_classes.size(); ++i)
            if (out_classes[i] == s) return (int)i;
        out_classes.push_back(s);
        if (out_classes.size() > AVT_MAX_CLASSES) throw Unsupported("more than 16 distinct predicted classes");
        return (int)out_classes.size() - 1;
    };
    std::vector<int32_t> n_paths;
    std::vector<uint8_t> path_class;
    std::vector<uint32_t> masks;
    for (auto& m : models) {
        const int P = (int)m.paths.size(), W = (P + 31) / 32;
        n_paths.push_back(P);
        const size_t base = masks.size();
        masks.resize(base + (size_t)TB * W, 0u);
        for (int p = 0; p < P; ++p) {
            const MPath& path = m.paths[(size_t)p];
            // doPrediction (tree/DecisionPathList.java:356-366): strict > from 0 over classValPr in file order
            std::string pred_cl;
            bool has = false;
            double max_pr = 0;
            for (auto& kv : path.class_pr)
                if (kv.second > max_pr) { pred_cl = kv.first; max_pr = kv.second; has = true; }
            std::string final_cl = has ? pred_cl : "null";
            if (cost_based) {   // model/ProbabilisticPredictiveModel.java:45-52
                if (!has) throw ReferenceError("NullPointerException: a decision path predicts no class (cost based prediction)");
                const double thr = fp_cost / (fp_cost + fn_cost);
                const double pos_prob = pred_cl == class_values[0] ? max_pr : 1.0 - max_pr;
                final_cl = pos_prob > thr ? class_values[0] : class_values[1];
            }
            path_class.push_back((uint8_t)class_code(final_cl));
            // admission of every bin of every axis
            for (int a = 0; a < A; ++a) {
                const Axis& ax = axes[(size_t)a];
                for (int b = 0; b < ax.n_bins(); ++b) {
                    bool ok = true;
                    for (auto& q : path.preds) {
                        if (q.attr != ax.ordinal) continue;
                        if (ax.field->is_int()) {
                            const int k = (int)(std::lower_bound(ax.ithr.begin(), ax.ithr.end(), q.vint) - ax.ithr.begin());   // x <= t_k  <=>  bin <= k
                            if (q.op == "le") ok = ok && b <= k;
                            else {
                                ok = ok && b > k;
                                if (q.has_other_int) ok = ok && b <= (int)(std::lower_bound(ax.ithr.begin(), ax.ithr.end(), q.other_int) - ax.ithr.begin());
                            }
                        } else if (ax.field->is_double()) {
                            const int k = (int)(std::lower_bound(ax.dthr.begin(), ax.dthr.end(), q.vdbl) - ax.dthr.begin());
                            if (q.op == "le") ok = ok && b <= k;
                            else {
                                ok = ok && b > k;
                                if (q.has_other_int) ok = ok && b <= (int)(std::lower_bound(ax.dthr.begin(), ax.dthr.end(), (double)q.other_int) - ax.dthr.begin());
                            }
                        } else {
                            ok = ok && b < (int)ax.cats.size() && std::find(q.cats.begin(), q.cats.end(), ax.cats[(size_t)b]) != q.cats.end();
                        }
                    }
                    if (ok) masks[base + (size_t)(bin_off[(size_t)a] + b) * W + (size_t)(p >> 5)] |= 1u << (p & 31);
                }
            }
        }
    }
    if (error_counting || cost_based)
        for (auto& s : class_values) class_code(s);
    std::vector<uint8_t> class_rank(out_classes.size(), 0);
    {
        auto order = hashmap_key_order(out_classes);   // votes.keySet() iteration order (model/EnsemblePredictiveModel.java:88-92)
        for (size_t r = 0; r < order.size(); ++r)
            for (size_t c = 0; c < out_classes.size(); ++c)
                if (out_classes[c] == order[r]) class_rank[c] = (uint8_t)r;
    }

    // ---- encode the rows (model/ModelPredictor.java:231: split without limit; tree/DecisionTreeModel.java:101-141)
    std::vector<std::string> lines = read_input_lines(in_path);
    const size_t n = lines.size();
    if (n == 0) {
        write_output_lines(out_path, {});
        return 0;
    }
    std::vector<std::vector<uint8_t>> col8((size_t)A);
    std::vector<std::vector<uint16_t>> col16((size_t)A);
    for (int a = 0; a < A; ++a) {
        if (n_bins[(size_t)a] > 256) col16[(size_t)a].assign(n, 0);
        else col8[(size_t)a].assign(n, 0);
    }
    std::vector<std::vector<std::string>> kept(n);   // the fields the output needs
    for (size_t r = 0; r < n; ++r) {
        std::vector<std::string> items = split_str(lines[r], delim_regex, false);
        for (int a = 0; a < A; ++a) {
            const Axis& ax = axes[(size_t)a];
            if (ax.ordinal < 0) continue;
            if (ax.ordinal >= (int)items.size()) throw ReferenceError("ArrayIndexOutOfBoundsException: attribute " + std::to_string(ax.ordinal) + " missing in row: " + lines[r]);
            const std::string& raw = items[(size_t)ax.ordinal];
            int b;
            if (ax.field->is_int()) {
                long long v;
                if (!parse_java_int(raw, v)) throw ReferenceError("NumberFormatException: For input string: \"" + raw + "\"");
                b = (int)(std::lower_bound(ax.ithr.begin(), ax.ithr.end(), v) - ax.ithr.begin());
            } else if (ax.field->is_double()) {
                char* end = nullptr;
                std::string t = trim(raw);
                double v = strtod(t.c_str(), &end);
                if (t.empty() || *end) throw ReferenceError("NumberFormatException: For input string: \"" + raw + "\"");
                b = (int)(std::lower_bound(ax.dthr.begin(), ax.dthr.end(), v) - ax.dthr.begin());
            } else {
                b = (int)(std::find(ax.cats.begin(), ax.cats.end(), raw) - ax.cats.begin());   // cats.size() = not mentioned by any model
            }
            if (n_bins[(size_t)a] > 256) col16[(size_t)a][r] = (uint16_t)b;
            else col8[(size_t)a][r] = (uint8_t)b;
        }
        auto need = [&](int ord, const char* what) -> std::string {
            if (ord < 0 || ord >= (int)items.size()) throw ReferenceError(std::string("ArrayIndexOutOfBoundsException: ") + what + " ordinal " + std::to_string(ord));
            return items[(size_t)ord];
        };
        if (id_ord >= 0 || output_mode == "withKId") kept[r].push_back(need(id_ord, "id"));
        if (output_mode == "withActualClassAttr") kept[r].push_back(need(out_class_ord, "class attribute"));
        if (error_counting) kept[r].push_back(need(class_attr_ord, "class attribute"));
    }

    // ---- predict on the device
    Engine eng(conf.get_int("avt.device", 0));
    std::vector<const void*> cols;
    for (int a = 0; a < A; ++a) cols.push_back(n_bins[(size_t)a] > 256 ? (const void*)col16[(size_t)a].data() : (const void*)col8[(size_t)a].data());
    std::vector<uint8_t> dummy_cls(n, 0);
    eng.check(avt_dataset_load(eng.raw(), (int64_t)n, A, n_bins.data(), cols.data(), dummy_cls.data(), 1));
    std::vector<uint8_t> pred(n, 0xFF);
    eng.check(avt_predict(eng.raw(), (int32_t)models.size(), n_paths.data(), masks.data(), path_class.data(), weights.data(),
                          (int32_t)out_classes.size(), class_rank.data(), pred.data()));

    // ---- output lines + error counting (model/ModelPredictor.java:229-254, 178-191; model/PredictiveModel.java:84-96)
    std::vector<std::string> out;
    out.reserve(n);
    long long total = 0, errors = 0, false_pos = 0, false_neg = 0;
    for (size_t r = 0; r < n; ++r) {
        if (pred[r] == 0xFF) throw ReferenceError("NullPointerException: no decision path matches row: " + lines[r]);
        const std::string& pc = out_classes[pred[r]];
        if (error_counting) {
            ++total;
            if (kept[r].back() != pc) {
                if (pc == class_values[0]) ++false_pos;
                else ++false_neg;
                ++errors;
            }
        }
        if (output_mode == "withRecord") out.push_back(lines[r] + delim_out + pc);
        else out.push_back(kept[r][0] + delim_out + pc);
    }
    write_output_lines(out_path, out);
    if (error_counting) {
        std::string txt = "total error: " + format_double3((double)errors / (double)total) + "\n" +
                          "false positive error: " + format_double3((double)false_pos / (double)total) + "\n" +
                          "false negative error: " + format_double3((double)false_neg / (double)total);
        const std::string path = conf.get("map.error.rate.file.path");
        FILE* f = fopen(path.c_str(), "wb");
        if (!f) throw std::runtime_error("cannot write " + path);
        fwrite(txt.data(), 1, txt.size(), f);
        fclose(f);
    }
    return 0;
}

}  // namespace avt_host
