// DecisionTreeBuilder re-hosted above the kernel-level C-ABI.
//
// Keeps the reference job's surface (tree/DecisionTreeBuilder.java): same dtb.* knobs, the JSON
// feature schema, CSV rows (path-tagged after the first pass), the DecisionPathList JSON written to
// dtb.decision.file.path.out and the path-tagged data file <outPath>/part-r-00000.  What changes is
// who does the work: rows are binned once into columns, kernel (a) builds the per-(node, attribute,
// bin, class) counts that BuilderMapper.pathMapHelper + the shuffle + BuilderReducer.reduce produce
// as N x S text records, kernel (b) computes what BuilderReducer.expandTree computes, kernel (c)
// re-keys the rows.  The host keeps the string world: split ids, predicate strings, HashMap
// iteration orders, stopping flags, JSON.
//
// Extensions (absent from the reference, default = reference behaviour):
//   avt.dtb.num.levels   (1)  levels run inside this one invocation (root pass counts as one when
//                             dtb.decision.file.path.in is absent); the reference needs one job and
//                             a manual file shuffle (resource/detr.sh:34-54) per level
//   avt.forest.num.trees (1)  bagged trees built together, one launch per level for all of them
//   avt.random.seed      (0)  seed of the java.util.Random stream that replaces Math.random()
//   avt.device           (0)  CUDA device ordinal
//   avt.encode      (device)  device: the first pass hands the rows' text to avt_dataset_load_csv; host: the C++ encoder below
#include <dirent.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <memory>

#include "tool.hpp"

namespace avt_host {

static const char* kRoot = "$root";

// ---- shared helpers ---------------------------------------------------------------------------

Engine::Engine(int device) {
    int rc = avt_create(device, &e_);
    if (rc != AVT_OK) throw EngineError(rc, avt_last_error(nullptr));
}
Engine::~Engine() { avt_destroy(e_); }
void Engine::check(int rc) const {
    if (rc != AVT_OK) throw EngineError(rc, avt_last_error(e_));
}

int ClassTable::code_of(const std::string& s) {
    for (size_t i = 0; i < names.size(); ++i)
        if (names[i] == s) return (int)i;
    names.push_back(s);
    if (names.size() > 12)
        throw Unsupported("more than 12 class values: HashMap<String,Integer> resizes and its iteration order becomes node-dependent (SURVEY App. B-7); not implemented");
    return (int)names.size() - 1;
}
void ClassTable::finish() {
    order.clear();
    for (auto& k : hashmap_key_order(names))
        for (size_t i = 0; i < names.size(); ++i)
            if (names[i] == k) order.push_back((uint8_t)i);
}

std::vector<std::string> read_input_lines(const std::string& path) {
    struct stat st;
    if (stat(path.c_str(), &st) != 0) throw std::runtime_error("input path does not exist: " + path);
    std::vector<std::string> files;
    if (S_ISDIR(st.st_mode)) {
        DIR* d = opendir(path.c_str());
        if (!d) throw std::runtime_error("cannot list " + path);
        while (dirent* de = readdir(d)) {
            std::string n = de->d_name;
            if (n.empty() || n[0] == '.' || n[0] == '_') continue;
            std::string p = path + "/" + n;
            struct stat s2;
            if (stat(p.c_str(), &s2) == 0 && S_ISREG(s2.st_mode)) files.push_back(p);
        }
        closedir(d);
        std::sort(files.begin(), files.end());
    } else {
        files.push_back(path);
    }
    std::vector<std::string> lines;
    for (auto& f : files) {
        std::string data = read_file(f);
        size_t pos = 0;
        while (pos < data.size()) {
            size_t q = data.find('\n', pos);
            if (q == std::string::npos) q = data.size();
            size_t end = q;
            if (end > pos && data[end - 1] == '\r') --end;
            lines.emplace_back(data, pos, end - pos);
            pos = q + 1;
        }
    }
    return lines;
}

static void mkdirs(const std::string& dir) {
    std::string cur;
    for (size_t i = 0; i <= dir.size(); ++i) {
        if (i == dir.size() || dir[i] == '/') {
            if (!cur.empty()) mkdir(cur.c_str(), 0777);
        }
        if (i < dir.size()) cur.push_back(dir[i]);
    }
}

static void write_text(const std::string& path, const std::string& text) {
    size_t slash = path.rfind('/');
    if (slash != std::string::npos && slash > 0) mkdirs(path.substr(0, slash));
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) throw std::runtime_error("cannot write " + path);
    fwrite(text.data(), 1, text.size(), f);
    fclose(f);
}

void write_output_lines(const std::string& out_dir, const std::vector<std::string>& lines, int part) {
    mkdirs(out_dir);
    char name[32];
    snprintf(name, sizeof name, "/part-r-%05d", part);
    std::string path = out_dir + name;
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) throw std::runtime_error("cannot write " + path);
    for (auto& l : lines) {
        fwrite(l.data(), 1, l.size(), f);
        fputc('\n', f);
    }
    fclose(f);
}

// ---- DecisionPathList beans (tree/DecisionPathList.java) ----------------------------------------

struct PathPred {
    int attribute = 0;
    std::string predicate_str, op;
    bool has_op = false;
    int value_int = 0;
    double value_dbl = 0.0;
    bool has_cat = false;
    std::vector<std::string> cat;
    bool has_other_int = false, has_other_dbl = false;
    int other_int = 0;
    double other_dbl = 0;
};

struct DecPath {
    bool has_preds = false;
    std::vector<PathPred> preds;
    long long population = 0;
    double info_content = 0;
    bool stopped = false;
    std::vector<std::pair<std::string, double>> class_val_pr;  // HashMap iteration order
    std::string output_class_val;
    double confidence = 0, support = 0;
};

static PathPred root_pred() {
    PathPred p;
    p.predicate_str = kRoot;
    return p;
}

// DecisionPathPredicate.create{Int,Double,Categorical}Predicate (tree/DecisionPathList.java:402-447)
static PathPred make_pred(const std::string& s, const Field& f) {
    std::vector<std::string> it;
    {
        std::istringstream is(s);
        std::string w;
        while (is >> w) it.push_back(w);
    }
    if (it.size() < 3) throw ReferenceError("ArrayIndexOutOfBoundsException: predicate '" + s + "'");
    PathPred p;
    p.attribute = std::stoi(it[0]);
    p.op = it[1];
    p.has_op = true;
    p.predicate_str = s;
    if (f.is_int()) {
        p.value_int = std::stoi(it[2]);
        if (it.size() == 4) { p.has_other_int = true; p.other_int = std::stoi(it[3]); }
    } else if (f.is_double()) {
        p.value_dbl = std::stod(it[2]);
        if (it.size() == 4) { p.has_other_dbl = true; p.other_dbl = std::stod(it[3]); }
    } else if (f.is_categorical()) {
        p.has_cat = true;
        p.cat = split_str(it[2], ":", false);
    } else {
        throw ReferenceError("invalid data type for predicates");
    }
    return p;
}

static std::string dbl(double x) { return java_double_to_string(x); }

static std::string pred_json(const PathPred& p) {
    std::string o = "{";
    o += "\"attribute\":" + std::to_string(p.attribute);
    o += ",\"predicateStr\":" + json_escape(p.predicate_str);
    o += ",\"operator\":" + (p.has_op ? json_escape(p.op) : std::string("null"));
    o += ",\"valueInt\":" + std::to_string(p.value_int);
    o += ",\"valueDbl\":" + dbl(p.value_dbl);
    if (p.has_cat) {
        o += ",\"categoricalValues\":[";
        for (size_t i = 0; i < p.cat.size(); ++i) o += (i ? "," : "") + json_escape(p.cat[i]);
        o += "]";
    } else {
        o += ",\"categoricalValues\":null";
    }
    o += ",\"otherBoundInt\":" + (p.has_other_int ? std::to_string(p.other_int) : std::string("null"));
    o += ",\"otherBoundDbl\":" + (p.has_other_dbl ? dbl(p.other_dbl) : std::string("null"));
    return o + "}";
}

// Field order of the sample model resource/dec_tree_rules.json, then the three newer bean fields.
static std::string path_json(const DecPath& d) {
    std::string o = "{\"stopped\":";
    o += d.stopped ? "true" : "false";
    o += ",\"classValPr\":{";
    for (size_t i = 0; i < d.class_val_pr.size(); ++i)
        o += (i ? "," : "") + json_escape(d.class_val_pr[i].first) + ":" + dbl(d.class_val_pr[i].second);
    o += "},\"infoContent\":" + dbl(d.info_content);
    if (d.has_preds) {
        o += ",\"predicates\":[";
        for (size_t i = 0; i < d.preds.size(); ++i) o += (i ? "," : "") + pred_json(d.preds[i]);
        o += "]";
    } else {
        o += ",\"predicates\":null";
    }
    o += ",\"population\":" + std::to_string(d.population);
    o += ",\"outputClassVal\":" + json_escape(d.output_class_val);
    o += ",\"confidence\":" + dbl(d.confidence);
    o += ",\"support\":" + dbl(d.support);
    return o + "}";
}

static std::string path_list_json(const std::vector<DecPath>* paths) {
    if (!paths) return "{\"decisionPaths\":null}";
    std::string o = "{\"decisionPaths\":[";
    for (size_t i = 0; i < paths->size(); ++i) o += (i ? "," : "") + path_json((*paths)[i]);
    return o + "]}";
}

static DecPath path_from_json(const Json& j) {
    DecPath d;
    if (auto* v = j.get("population")) d.population = (long long)v->num;
    if (auto* v = j.get("infoContent")) d.info_content = v->num;
    if (auto* v = j.get("stopped")) d.stopped = v->b;
    if (auto* v = j.get("classValPr"))
        for (auto& kv : v->obj) d.class_val_pr.emplace_back(kv.first, kv.second.num);
    if (auto* v = j.get("outputClassVal")) d.output_class_val = v->str;
    if (auto* v = j.get("confidence")) d.confidence = v->num;
    if (auto* v = j.get("support")) d.support = v->num;
    if (auto* v = j.get("predicates"); v && v->kind == Json::Arr) {
        d.has_preds = true;
        for (auto& pj : v->arr) {
            PathPred p;
            if (auto* x = pj.get("attribute")) p.attribute = (int)x->num;
            if (auto* x = pj.get("predicateStr")) p.predicate_str = x->str;
            if (auto* x = pj.get("operator"); x && x->kind == Json::Str) { p.has_op = true; p.op = x->str; }
            if (auto* x = pj.get("valueInt")) p.value_int = (int)x->num;
            if (auto* x = pj.get("valueDbl")) p.value_dbl = x->num;
            if (auto* x = pj.get("categoricalValues"); x && x->kind == Json::Arr) {
                p.has_cat = true;
                for (auto& c : x->arr) p.cat.push_back(c.str);
            }
            if (auto* x = pj.get("otherBoundInt"); x && x->kind == Json::Num) { p.has_other_int = true; p.other_int = (int)x->num; }
            if (auto* x = pj.get("otherBoundDbl"); x && x->kind == Json::Num) { p.has_other_dbl = true; p.other_dbl = x->num; }
            d.preds.push_back(p);
        }
    }
    return d;
}

// DecisionPathList.stripSplitId (tree/DecisionPathList.java:97-107)
static std::string strip_id(const std::string& item) {
    if (item == kRoot) return item;
    size_t p = item.find(':');
    if (p == std::string::npos) throw ReferenceError("split id delimiter not found in path element '" + item + "'");
    return item.substr(p + 1);
}

// DecisionPath.findMajorityClass (tree/DecisionPathList.java:175-192)
static void majority(DecPath& d) {
    double maj = -1.0;
    for (auto& kv : d.class_val_pr)
        if (kv.second > maj) { maj = kv.second; d.output_class_val = kv.first; }
    double mn = 1.0 - maj;
    if (mn > 0) {
        d.confidence = maj / mn;
        if (d.confidence > 100) d.confidence = 100;
    } else {
        d.confidence = 100;
    }
}

// ---- the job ------------------------------------------------------------------------------------

namespace {

struct Node {
    std::vector<std::string> tagged;  // key items as the reducer sees them: "$root", "3:2 le 20", ...
    DecPath path;                     // its entry in the current decision-path list
    std::vector<int> cand;            // candidate attribute columns (indices into attrs_)
    // Where the node's rows live on the device.  An engine "tree" is a LAYER: one node id per row.  With binary splits a
    // reference tree is one layer.  The predicates of a multi-way numeric split overlap (`le p_last` has no lower bound,
    // SURVEY App. B-1), so a row can sit in several frontier nodes at once: children whose bins overlap a sibling's are
    // given a layer of their own (same rows and multiplicities, node ids of their own).
    int layer = 0, id = 0;
};

struct Tree {
    std::vector<Node> frontier;
    std::vector<uint8_t> weights;  // empty = all ones
    JavaRandom rng{0};             // the tree's Math.random() stand-in: sampling first, then randomAmongTop draws
};

class DtbJob {
    const Props& conf_;
    std::string in_path_, out_path_;
    // knobs (tree/DecisionTreeBuilder.java:137-193, 414-463)
    std::string delim_regex_, delim_out_, path_delim_, attr_strategy_, sampling_, algorithm_, split_sel_, stop_strategy_;
    int sampling_rate_ = 0, sampling_buffer_ = 10000, total_population_ = -1, top_split_count_ = 3;
    int max_depth_ = -1, min_population_ = -1;
    double min_info_gain_ = -1;
    int num_levels_ = 1, num_trees_ = 1, device_ = 0;
    long long seed_ = 0;
    bool debug_ = false;
    // model
    Schema schema_;
    std::vector<AttrSplits> attrs_;  // base attributes in getBaseAttributeOrdinals() order = device columns
    DeviceSplitTable table_;
    int class_ord_ = 0;
    ClassTable classes_;
    // data
    std::vector<std::string> rows_;  // row text without the path prefix
    std::vector<std::vector<uint8_t>> col8_;
    std::vector<std::vector<uint16_t>> col16_;
    std::vector<uint8_t> cls_;
    bool device_encode_ = false;   // the rows are encoded by avt_dataset_load_csv (first pass over untagged rows); the host keeps the text only
    std::vector<Tree> trees_;
    std::vector<int> layer_tree_;    // device layer (engine tree) -> reference tree
    // Row sharding over several GPUs (avt.num.ranks > 1; one process per GPU, launched like the ranks of any multi-GPU job):
    // every rank reads the whole input, so class codes, bootstrap samples and every host-side decision are identical, but
    // uploads and partitions only rows [lo_, hi_); the per-level histograms are summed over the ranks inside
    // avt_level_histogram (avt_comm_init_rank: NCCL over NVLink).  Rank 0 writes the decision-path lists, every rank its
    // own part-r-<rank> of the level data.  This is the reference's shuffle between BuilderMapper and BuilderReducer
    // (tree/DecisionTreeBuilder.java:80-93) with the mappers' rows living on different GPUs.
    int n_ranks_ = 1, rank_ = 0;
    size_t lo_ = 0, hi_ = 0;
    size_t n_local() const { return hi_ - lo_; }
    bool multiway_ = false;          // some candidate split has more than two predicates
    std::unique_ptr<Engine> eng_;
    JavaRandom rng_{0};
    int pass_ = 0;  // passes written so far (0 = root pass)

public:
    DtbJob(const Props& conf, std::string in, std::string out) : conf_(conf), in_path_(std::move(in)), out_path_(std::move(out)) {}

    int run() {
        read_conf();
        const bool tree_available = conf_.nonblank("dtb.decision.file.path.in") && file_exists(conf_.get("dtb.decision.file.path.in"));
        std::vector<std::string> lines = read_input_lines(in_path_);
        std::vector<std::vector<std::string>> row_tagged;  // per row: tagged path items (tree available)
        std::vector<DecPath> dec_in;
        if (tree_available) {
            Json j = JsonParser(read_file(conf_.get("dtb.decision.file.path.in"))).parse();
            const Json* arr = j.get("decisionPaths");
            if (!arr || arr->kind != Json::Arr) throw ReferenceError("NullPointerException: decision path list has no decisionPaths");
            for (auto& pj : arr->arr) dec_in.push_back(path_from_json(pj));
        }
        // Encoding (avt.encode = device | host, default device): the first pass over untagged rows hands the text to the device
        // encoder after one light host pass that only collects the class values in first-seen order -- the
        // schemas do not list them (call_hangup.json's class field has no cardinality) and the reference's HashMap order
        // depends on it.  Tagged rows (passes that resume from a tree file) need their path strings on the host anyway.
        device_encode_ = conf_.get("avt.encode", "device") == "device" && !tree_available && class_pass_for_device_encode(lines);
        if (!device_encode_) encode(lines, tree_available, row_tagged);
        lo_ = rows_.size() * (size_t)rank_ / (size_t)n_ranks_;
        hi_ = rows_.size() * (size_t)(rank_ + 1) / (size_t)n_ranks_;
        if (hi_ == lo_) throw Unsupported("rank " + std::to_string(rank_) + " of " + std::to_string(n_ranks_) + " has no rows");
        eng_.reset(new Engine(device_));
        if (n_ranks_ > 1) join_ranks();
        upload();
        int levels_left = num_levels_;
        if (!tree_available) {
            init_root_pass();
            std::vector<std::vector<DecPath>> lists = root_pass();
            --levels_left;
            if (levels_left == 0) {
                write_lists(lists);
                write_root_data();
                return 0;
            }
            write_lists(lists);
        } else {
            init_from_paths(dec_in, row_tagged);
        }
        std::vector<std::vector<DecPath>> lists;
        std::vector<bool> list_null;
        for (; levels_left > 0; --levels_left) {
            bool any = expand_level(lists, list_null);
            write_lists(lists, &list_null);
            if (!any) break;
        }
        write_level_data();
        return 0;
    }

private:
    static bool file_exists(const std::string& p) {
        struct stat st;
        return stat(p.c_str(), &st) == 0;
    }

    void read_conf() {
        debug_ = conf_.get_bool("debug.on", false);
        n_ranks_ = conf_.get_int("avt.num.ranks", 1);
        rank_ = conf_.get_int("avt.rank", 0);
        if (n_ranks_ < 1 || rank_ < 0 || rank_ >= n_ranks_) throw Unsupported("avt.rank " + std::to_string(rank_) + " outside 0.." + std::to_string(n_ranks_ - 1));
        delim_regex_ = conf_.get("field.delim.regex", ",");
        delim_out_ = conf_.get("field.delim.out", ",");
        path_delim_ = conf_.get("dtb.dec.path.delim", ";");
        for (const std::string* d : {&delim_regex_, &path_delim_})
            for (char c : *d)
                if (strchr(".*+?()[]{}|^$\\", c)) throw Unsupported("delimiter '" + *d + "' contains regex metacharacters; only literal delimiters are implemented");
        if (!conf_.nonblank("dtb.feature.schema.file.path")) throw ReferenceError("NullPointerException: dtb.feature.schema.file.path not set");
        schema_ = Schema::load(conf_.get("dtb.feature.schema.file.path"));
        attr_strategy_ = conf_.get("dtb.split.attribute.selection.strategy", "notUsedYet");
        if (attr_strategy_ == "randomAll" || attr_strategy_ == "randomNotUsedYet")
            throw ReferenceError("NullPointerException: dtb.split.attribute.selection.strategy=" + attr_strategy_ +
                                 " never assigns splitAttrs in the reference (tree/DecisionTreeBuilder.java:372-378)");
        if (attr_strategy_ != "all" && attr_strategy_ != "notUsedYet") throw ReferenceError("invalid splitting attribute selection strategy");
        sampling_ = conf_.get("dtb.sub.sampling.strategy", "withReplace");
        if (sampling_ == "withoutReplace") {
            if (!conf_.nonblank("dtb.sub.sampling.rate")) throw ReferenceError("samling rate should be provided for sampling without replacement");
            sampling_rate_ = conf_.get_int("dtb.sub.sampling.rate", 0);
        }
        sampling_buffer_ = conf_.get_int("dtb.sub.sampling.buffer.size", 10000);
        algorithm_ = conf_.get("dtb.split.algorithm", "giniIndex");
        total_population_ = conf_.get_int("dtb.total.population", -1);
        split_sel_ = conf_.get("dtb.split.select.strategy", "best");
        if (split_sel_ != "best" && split_sel_ != "randomAmongTop") throw ReferenceError("ivalid split slection strategy");
        top_split_count_ = conf_.get_int("dtb.top.split.count", 3);
        stop_strategy_ = conf_.get("dtb.path.stopping.strategy", "minInfoGain");
        if (stop_strategy_ == "maxDepth") {
            if (!conf_.nonblank("dtb.max.depth.limit")) throw ReferenceError("missing max depth limit for tree");
            max_depth_ = conf_.get_int("dtb.max.depth.limit", -1);
        } else if (stop_strategy_ == "minInfoGain") {
            if (!conf_.nonblank("dtb.min.info.gain.limit")) throw ReferenceError("missing min info gain limit");
            min_info_gain_ = conf_.get_double("dtb.min.info.gain.limit", -1);
        } else if (stop_strategy_ == "minPopulation") {
            if (!conf_.nonblank("dtb.min.population.limit")) throw ReferenceError("missing min population limit");
            min_population_ = conf_.get_int("dtb.min.population.limit", -1);
        } else {
            throw ReferenceError("invalid stopping strategy " + stop_strategy_);
        }
        if (!conf_.nonblank("dtb.decision.file.path.out")) throw ReferenceError("NullPointerException: dtb.decision.file.path.out not set");
        num_levels_ = std::max(1, conf_.get_int("avt.dtb.num.levels", 1));
        num_trees_ = std::max(1, conf_.get_int("avt.forest.num.trees", 1));
        seed_ = conf_.get_int("avt.random.seed", 0);
        device_ = conf_.get_int("avt.device", 0);
        rng_ = JavaRandom(seed_);
        // base attributes (tree/SplitManager.java:141-157)
        std::vector<int> base;
        if (conf_.nonblank("dtb.custom.base.attributes")) {
            for (auto& s : split_str(conf_.get("dtb.custom.base.attributes"), ",", false)) base.push_back(std::stoi(trim(s)));
        } else {
            base = schema_.feature_ordinals();
        }
        for (int o : base) attrs_.push_back(dtb_attr_splits(schema_.by_ordinal(o)));
        if (attrs_.empty()) throw ReferenceError("no candidate attributes in the schema");
        table_ = flatten_splits(attrs_);
        for (auto& as : attrs_)
            for (auto& h : as.splits) multiway_ = multiway_ || h.n_seg > 2;
        class_ord_ = schema_.class_field().ordinal;
        int rec_len = conf_.get_int("dtb.rec.len", -1);
        if (rec_len > 0)
            for (auto& f : schema_.fields)
                if (f.ordinal >= rec_len) throw ReferenceError("field ordinal " + std::to_string(f.ordinal) + " beyond record length");
    }

    // CSV rows -> columnar bin codes + class codes; with a tree available also the tagged path of
    // every row (tree/DecisionTreeBuilder.java:268-320).
    void encode(const std::vector<std::string>& lines, bool tagged, std::vector<std::vector<std::string>>& row_tagged) {
        const size_t n = lines.size();
        if (n == 0) throw ReferenceError("NullPointerException: no input rows (the reducer never sees a key)");
        col8_.assign(attrs_.size(), {});
        col16_.assign(attrs_.size(), {});
        for (size_t a = 0; a < attrs_.size(); ++a) {
            if (attrs_[a].n_bins > 256) col16_[a].resize(n);
            else col8_[a].resize(n);
        }
        cls_.resize(n);
        rows_.resize(n);
        if (tagged) row_tagged.resize(n);
        const int shift = tagged ? 1 : 0;
        for (size_t r = 0; r < n; ++r) {
            std::vector<std::string> items = split_str(lines[r], delim_regex_, true);
            if (tagged) {
                row_tagged[r] = split_str(items[0], path_delim_, false);
                size_t pos = lines[r].find(delim_regex_);
                if (pos == std::string::npos) throw ReferenceError("StringIndexOutOfBoundsException: row without a path field: " + lines[r]);
                rows_[r] = lines[r].substr(pos + delim_regex_.size());
            } else {
                rows_[r] = lines[r];
            }
            for (size_t a = 0; a < attrs_.size(); ++a) {
                const int ord = attrs_[a].ordinal;
                if (ord + shift >= (int)items.size() || (tagged && ord > (int)items.size() - 2))
                    throw ReferenceError("attrbite index out of bound attr:" + std::to_string(ord) + " rec length:" + std::to_string(items.size() - 1));
                int b = encode_value(attrs_[a], items[(size_t)(ord + shift)]);
                if (attrs_[a].n_bins > 256) col16_[a][r] = (uint16_t)b;
                else col8_[a][r] = (uint8_t)b;
            }
            if (class_ord_ + shift >= (int)items.size()) throw ReferenceError("ArrayIndexOutOfBoundsException: class attribute missing in row: " + lines[r]);
            cls_[r] = (uint8_t)classes_.code_of(items[(size_t)(class_ord_ + shift)]);
        }
        classes_.finish();
    }

    // Device encoding is possible when the delimiter is one literal byte and the attributes come in field order.  Collects the
    // class values (first-seen order) by walking every row up to its class field only; returns false -- the caller then runs the
    // host encoder, which also raises the reference's exception for the row -- on anything the two encoders could treat differently.
    bool class_pass_for_device_encode(const std::vector<std::string>& lines) {
        if (lines.empty() || delim_regex_.size() != 1) return false;
        int prev = -1;
        for (auto& as : attrs_) {
            if (as.ordinal <= prev || as.ordinal == class_ord_ || as.ordinal >= 1024) return false;
            prev = as.ordinal;
            for (long long p : as.int_points)
                if (p < -2147483647LL - 1 || p > 2147483647LL) return false;
        }
        if (class_ord_ < 0 || class_ord_ >= 1024) return false;
        const char d = delim_regex_[0];
        ClassTable found;
        for (auto& ln : lines) {
            size_t pos = 0;
            for (int f = 0; f < class_ord_; ++f) {
                const size_t q = ln.find(d, pos);
                if (q == std::string::npos) return false;   // short row: the host encoder decides which exception comes first
                pos = q + 1;
            }
            size_t end = ln.find(d, pos);
            if (end == std::string::npos) end = ln.size();
            if (end == pos || ln.find('\n') != std::string::npos) return false;   // empty class value: String.split(regex, -1) keeps it, the device encoder's split does not
            try {
                found.code_of(ln.substr(pos, end - pos));
            } catch (const Unsupported&) {
                return false;
            }
        }
        classes_ = found;
        classes_.finish();
        rows_ = lines;
        return true;
    }

    void upload_text() {
        std::string text;
        size_t bytes = 0;
        for (size_t r = lo_; r < hi_; ++r) bytes += rows_[r].size() + 1;
        text.reserve(bytes);
        for (size_t r = lo_; r < hi_; ++r) {
            text += rows_[r];
            text += '\n';
        }
        std::vector<avt_csv_attr> spec;
        std::vector<int32_t> ip, cat_off, cls_off;
        std::vector<double> dp;
        std::string cat_bytes, cls_bytes;
        for (auto& as : attrs_) {
            if (as.numeric && !as.dbl) {
                spec.push_back(avt_csv_attr{as.ordinal, AVT_CSV_INT, (int32_t)as.int_points.size(), (int32_t)ip.size()});
                for (long long p : as.int_points) ip.push_back((int32_t)p);
            } else if (as.numeric) {
                spec.push_back(avt_csv_attr{as.ordinal, AVT_CSV_DOUBLE, (int32_t)as.dbl_points.size(), (int32_t)dp.size()});
                dp.insert(dp.end(), as.dbl_points.begin(), as.dbl_points.end());
            } else {
                spec.push_back(avt_csv_attr{as.ordinal, AVT_CSV_CATEGORICAL, (int32_t)as.field->cardinality.size(), (int32_t)cat_off.size()});
                for (auto& v : as.field->cardinality) {
                    cat_off.push_back((int32_t)cat_bytes.size());
                    cat_bytes += v;
                }
            }
        }
        cat_off.push_back((int32_t)cat_bytes.size());
        for (auto& v : classes_.names) {
            cls_off.push_back((int32_t)cls_bytes.size());
            cls_bytes += v;
        }
        cls_off.push_back((int32_t)cls_bytes.size());
        int64_t n_rows = 0, bad_row = -1;
        int32_t bad_reason = 0;
        const int rc = avt_dataset_load_csv(eng_->raw(), text.data(), (int64_t)text.size(), delim_regex_[0], 0, (int32_t)spec.size(), spec.data(), ip.data(), dp.data(),
                                            cat_bytes.data(), cat_off.data(), class_ord_, (int32_t)classes_.names.size(), cls_bytes.data(), cls_off.data(), &n_rows,
                                            nullptr, &bad_row, &bad_reason);
        if (rc == AVT_E_REFERENCE && bad_row >= 0) {
            // a row the reference throws on: the host encoder walks the rows in order and raises that exception with its message
            std::vector<std::vector<std::string>> unused;
            std::vector<std::string> lines = rows_;
            encode(lines, false, unused);
            throw std::runtime_error("the device encoder rejected row " + std::to_string(lo_ + (size_t)bad_row) + " but the host encoder accepts it");
        }
        eng_->check(rc);
        if ((size_t)n_rows != n_local()) throw std::runtime_error("the device encoder saw " + std::to_string(n_rows) + " rows, the host " + std::to_string(n_local()));
    }

    // bin codes of one attribute for this rank's rows on the host (multi-way spawns): fetched from the device when it did the encoding
    void need_host_column(size_t a) {
        if (!device_encode_) return;
        if (col8_.size() != attrs_.size()) {
            col8_.assign(attrs_.size(), {});
            col16_.assign(attrs_.size(), {});
        }
        const bool wide = attrs_[a].n_bins > 256;
        if (wide ? !col16_[a].empty() : !col8_[a].empty()) return;
        if (wide) {
            col16_[a].assign(rows_.size(), 0);
            eng_->check(avt_dataset_fetch_column(eng_->raw(), (int32_t)a, col16_[a].data() + lo_));
        } else {
            col8_[a].assign(rows_.size(), 0);
            eng_->check(avt_dataset_fetch_column(eng_->raw(), (int32_t)a, col8_[a].data() + lo_));
        }
    }

    // NCCL rendezvous through a file both ranks can see (avt.comm.id.file, default <decision file out>.ncclid): rank 0 creates the
    // 128-byte id and publishes it with an atomic rename, the others poll for it.  Any launcher that starts one process per GPU
    // with avt.rank / avt.num.ranks set (torchrun, mpirun, a shell loop) works; nothing else is exchanged on the host side.
    void join_ranks() {
        std::string path = conf_.get("avt.comm.id.file", conf_.get("dtb.decision.file.path.out", "avt") + ".ncclid");
        unsigned char id[128];
        if (rank_ == 0) {
            eng_->check(avt_comm_unique_id(id));
            const std::string tmp = path + ".tmp";
            FILE* f = fopen(tmp.c_str(), "wb");
            if (!f || fwrite(id, 1, sizeof id, f) != sizeof id) throw std::runtime_error("cannot write " + tmp);
            fclose(f);
            if (rename(tmp.c_str(), path.c_str()) != 0) throw std::runtime_error("cannot publish " + path);
        } else {
            bool ok = false;
            for (int tries = 0; tries < 1200 && !ok; ++tries) {   // up to two minutes
                FILE* f = fopen(path.c_str(), "rb");
                if (f) {
                    ok = fread(id, 1, sizeof id, f) == sizeof id;
                    fclose(f);
                }
                if (!ok) usleep(100000);
            }
            if (!ok) throw std::runtime_error("rank " + std::to_string(rank_) + ": no communicator id at " + path + " after 120 s");
        }
        eng_->check(avt_comm_init_rank(eng_->raw(), n_ranks_, rank_, id));
        if (rank_ == 0 && conf_.get_bool("avt.comm.id.cleanup", false)) remove(path.c_str());
    }

    void upload() {
        if (device_encode_) {
            upload_text();
        } else {
            std::vector<int32_t> nb;
            std::vector<const void*> cols;
            for (size_t a = 0; a < attrs_.size(); ++a) {
                nb.push_back(attrs_[a].n_bins);
                cols.push_back(attrs_[a].n_bins > 256 ? (const void*)(col16_[a].data() + lo_) : (const void*)(col8_[a].data() + lo_));
            }
            eng_->check(avt_dataset_load(eng_->raw(), (int64_t)n_local(), (int32_t)attrs_.size(), nb.data(), cols.data(), cls_.data() + lo_, (int32_t)classes_.names.size()));
        }
        eng_->check(avt_split_table_set(eng_->raw(), (int32_t)table_.splits.size(), table_.splits.data(), table_.bounds.data(), (int32_t)table_.bounds.size(),
                                        table_.lut.data(), (int64_t)table_.lut.size()));
    }

    // First pass: sampling (tree/DecisionTreeBuilder.java:209-243; `withReplace` = the intended
    // buffer bootstrap, identical to explore/BaggingSampler.java:101-122).
    void init_root_pass() {
        const size_t n = rows_.size();
        trees_.assign((size_t)num_trees_, Tree());
        std::vector<const uint8_t*> wptr((size_t)num_trees_, nullptr);
        for (int t = 0; t < num_trees_; ++t) {
            Tree& tr = trees_[(size_t)t];
            tr.rng = JavaRandom(seed_ + t);
            JavaRandom& rng = tr.rng;
            if (sampling_ == "withReplace") {
                tr.weights.assign(n, 0);
                const size_t B = (size_t)std::max(1, sampling_buffer_);
                for (size_t s = 0; s < n; s += B) {
                    const size_t m = std::min(B, n - s);
                    for (size_t i = 0; i < m; ++i) {
                        size_t sel = (size_t)java_int_cast(rng.next_double() * (double)m);
                        if (sel == m) sel = m - 1;
                        if (tr.weights[s + sel] == 255) throw Unsupported("bootstrap multiplicity above 255");
                        ++tr.weights[s + sel];
                    }
                }
            } else if (sampling_ == "withoutReplace") {
                tr.weights.assign(n, 0);
                for (size_t r = 0; r < n; ++r)
                    if (java_int_cast(rng.next_double() * 100) < sampling_rate_) tr.weights[r] = 1;
            }
            if (!tr.weights.empty()) wptr[(size_t)t] = tr.weights.data() + lo_;
            Node root;
            root.tagged = {kRoot};
            root.layer = t;
            tr.frontier.push_back(root);
        }
        layer_tree_.resize((size_t)num_trees_);
        for (int t = 0; t < num_trees_; ++t) layer_tree_[(size_t)t] = t;
        eng_->check(avt_forest_init(eng_->raw(), num_trees_, wptr.data(), nullptr));
    }

    // Tree available: node of a row = first decision path matched by its id-stripped path
    // (tree/DecisionTreeBuilder.java:268-292; tree/DecisionPathList.java:71-81,198-214).
    void init_from_paths(const std::vector<DecPath>& dec_in, const std::vector<std::vector<std::string>>& row_tagged) {
        if (num_trees_ != 1) throw Unsupported("avt.forest.num.trees > 1 needs the in-process level loop (no dtb.decision.file.path.in)");
        trees_.assign(1, Tree());
        Tree& tr = trees_[0];
        tr.rng = JavaRandom(seed_);
        tr.frontier.resize(dec_in.size());
        for (size_t i = 0; i < dec_in.size(); ++i) {
            tr.frontier[i].path = dec_in[i];
            tr.frontier[i].id = (int)i;   // one layer: every input line carries one path, hence one node
        }
        layer_tree_.assign(1, 0);
        if (dec_in.size() >= AVT_NODE_NONE) throw Unsupported("more than 65534 frontier nodes");
        std::vector<uint16_t> nid(rows_.size(), (uint16_t)AVT_NODE_NONE);
        JStrMap<int> cache;  // validDecPaths
        for (size_t r = 0; r < rows_.size(); ++r) {
            std::vector<std::string> stripped;
            for (auto& it : row_tagged[r]) stripped.push_back(strip_id(it));
            const std::string key = join_str(stripped, path_delim_);
            int* hit = cache.find(key);
            int node;
            if (hit) {
                node = *hit;
            } else {
                node = -1;
                for (size_t i = 0; i < dec_in.size() && node < 0; ++i) {
                    const DecPath& d = dec_in[i];
                    if (!d.has_preds) {
                        if (stripped[0] == kRoot) node = (int)i;
                        continue;
                    }
                    bool ok = true;
                    for (size_t k = 0; k < d.preds.size(); ++k) {
                        if (k >= stripped.size()) throw ReferenceError("ArrayIndexOutOfBoundsException: row path shorter than decision path");
                        if (d.preds[k].predicate_str != stripped[k]) { ok = false; break; }
                    }
                    if (ok) node = (int)i;
                }
                cache.put(key, node);
            }
            if (node >= 0) {
                nid[r] = (uint16_t)node;
                if (tr.frontier[(size_t)node].tagged.empty()) tr.frontier[(size_t)node].tagged = row_tagged[r];
            }
        }
        const uint16_t* np = nid.data() + lo_;
        eng_->check(avt_forest_init(eng_->raw(), 1, nullptr, &np));
    }

    // frontier nodes of every device layer (node ids inside a layer are 0..n-1)
    std::vector<int32_t> nodes_per_tree() const {
        std::vector<int32_t> v(layer_tree_.size(), 0);
        for (auto& t : trees_)
            for (auto& nd : t.frontier) v[(size_t)nd.layer] = std::max(v[(size_t)nd.layer], (int32_t)nd.id + 1);
        return v;
    }

    std::vector<std::pair<std::string, double>> class_pr(const uint32_t* cnt, long long total) const {
        std::vector<std::pair<std::string, double>> pr;
        for (uint8_t c : classes_.order)
            if (cnt[c]) pr.emplace_back(classes_.names[c], (double)cnt[c] / (double)total);
        return pr;
    }

    int criterion() const { return algorithm_ == "entropy" ? AVT_CRIT_ENTROPY : AVT_CRIT_GINI; }

    // BuilderReducer.generateRoot (tree/DecisionTreeBuilder.java:478-493)
    std::vector<std::vector<DecPath>> root_pass() {
        auto npt = nodes_per_tree();
        eng_->check(avt_level_histogram(eng_->raw(), npt.data(), nullptr));
        std::vector<avt_node_best> best((size_t)num_trees_);
        eng_->check(avt_level_scan(eng_->raw(), criterion(), nullptr, classes_.order.data(), best.data()));
        std::vector<uint32_t> cc((size_t)num_trees_ * classes_.names.size());
        eng_->check(avt_level_class_counts(eng_->raw(), cc.data()));
        std::vector<std::vector<DecPath>> lists;
        for (int t = 0; t < num_trees_; ++t) {
            DecPath d;
            d.population = best[(size_t)t].population;
            if (d.population == 0) throw ReferenceError("NullPointerException: no rows reach the reducer in the root pass");
            d.info_content = best[(size_t)t].node_stat;
            d.class_val_pr = class_pr(&cc[(size_t)t * classes_.names.size()], d.population);
            if (total_population_ > 0) d.support = (double)d.population / (double)total_population_;
            trees_[(size_t)t].frontier[0].path = d;
            lists.push_back({d});
        }
        return lists;
    }

    // candidate attribute columns of a node (tree/DecisionTreeBuilder.java:365-381;
    // tree/SplitManager.java:195-232)
    std::vector<int> candidates(const Node& nd) const {
        std::vector<int> out;
        if (attr_strategy_ == "all") {
            for (size_t a = 0; a < attrs_.size(); ++a) out.push_back((int)a);
            return out;
        }
        std::vector<int> used;
        for (auto& it : nd.tagged) {
            if (it == kRoot) continue;
            std::string s = strip_id(it);
            used.push_back(std::stoi(s.substr(0, s.find(' '))));
        }
        for (size_t a = 0; a < attrs_.size(); ++a)
            if (std::find(used.begin(), used.end(), attrs_[a].ordinal) == used.end()) out.push_back((int)a);
        return out;
    }

    struct NodeOrder {
        std::vector<int> split_order;                       // table indices in splitInfoContent.keySet() order
        std::vector<std::vector<int>> seg_order_by_split;   // per table index: segments in predInfoContent.keySet() order
    };

    // Replays the insertion chain that fixes the reference's iteration orders for one parent:
    // reducer keys arrive sorted -> childStats (HashMap<"id:pred">) -> infoContentBySplit regroups
    // into splitInfoContent (HashMap<id>) and predInfoContent (HashMap<pred>)
    // (tree/DecisionTreeBuilder.java:730-767, 636-668).  `hist` = the node's [TB][C] counts.
    NodeOrder java_order(const Node& nd, const std::vector<uint32_t>& hist) const {
        const int C = (int)classes_.names.size();
        struct Key { std::string s; int table_idx; int seg; };
        std::vector<Key> keys;
        int split_id = 0;
        std::vector<int> bin_off(attrs_.size() + 1, 0);
        for (size_t a = 0; a < attrs_.size(); ++a) bin_off[a + 1] = bin_off[a] + attrs_[a].n_bins;
        for (int a : nd.cand) {
            const AttrSplits& as = attrs_[(size_t)a];
            for (size_t k = 0; k < as.splits.size(); ++k) {
                ++split_id;
                const HostSplit& h = as.splits[k];
                std::vector<char> present((size_t)h.n_seg, 0);
                for (int b = 0; b < as.n_bins; ++b) {
                    bool any = false;
                    for (int c = 0; c < C; ++c) any = any || hist[(size_t)(bin_off[(size_t)a] + b) * C + c] != 0;
                    if (!any) continue;
                    for (int sg = 0; sg < h.n_seg; ++sg)
                        if (split_bin_in_segment(as, h, b, sg)) present[(size_t)sg] = 1;
                }
                for (int sg = 0; sg < h.n_seg; ++sg)
                    if (present[(size_t)sg]) keys.push_back({std::to_string(split_id) + ":" + h.pred_str[(size_t)sg], table_.attr_first[(size_t)a] + (int)k, sg});
            }
        }
        std::sort(keys.begin(), keys.end(), [](const Key& x, const Key& y) { return x.s < y.s; });
        JStrMap<int> child_stats;
        for (size_t i = 0; i < keys.size(); ++i) child_stats.put(keys[i].s, (int)i);
        JStrMap<int> split_map;                    // split id string -> slot
        std::vector<JStrMap<int>> pred_maps;       // per slot: predicate -> key index
        std::vector<int> slot_table_idx;
        child_stats.for_each([&](const std::string& k, int& ki) {
            size_t p = k.find(':');
            std::string sid = k.substr(0, p), pred = k.substr(p + 1);
            int* slot = split_map.find(sid);
            if (!slot) {
                pred_maps.emplace_back();
                slot_table_idx.push_back(keys[(size_t)ki].table_idx);
                slot = &split_map.put(sid, (int)pred_maps.size() - 1);
            }
            pred_maps[(size_t)*slot].put(pred, ki);
        });
        NodeOrder o;
        o.seg_order_by_split.resize(table_.splits.size());
        split_map.for_each([&](const std::string&, int& slot) {
            const int ti = slot_table_idx[(size_t)slot];
            o.split_order.push_back(ti);
            pred_maps[(size_t)slot].for_each([&](const std::string&, int& ki) { o.seg_order_by_split[(size_t)ti].push_back(keys[(size_t)ki].seg); });
        });
        return o;
    }

    // Every tag chain under which the reference's reducer sees this node's rows.  With multi-way splits the same predicate
    // string occurs in several candidate splits of an attribute (`5 le 60` in [60], [60,120], [60,180], ...), the next
    // mapper matches rows by their id-STRIPPED path (tree/DecisionTreeBuilder.java:274-292), so rows written under rejected
    // splits survive with their own split ids: identical rows, different keys.  They change nothing in
    // the statistics (infoContentBySplit overwrites equal entries, :652-665) but they are keys of `decPaths`, whose size and
    // iteration order decide the order of the emitted list.  With binary splits every predicate string belongs to one split.
    std::vector<std::vector<std::string>> tag_variants(const Node& nd) const {
        std::vector<std::vector<std::string>> out{{}};
        Node prefix;
        for (size_t k = 0; k < nd.tagged.size(); ++k) {
            const std::string& item = nd.tagged[k];
            std::vector<std::string> alts;
            if (item == kRoot || !multiway_) {
                alts.push_back(item);
            } else {
                const std::string pred = strip_id(item);
                int split_id = 0;
                for (int a : candidates(prefix))
                    for (auto& h : attrs_[(size_t)a].splits) {
                        ++split_id;
                        for (auto& ps : h.pred_str)
                            if (ps == pred) { alts.push_back(std::to_string(split_id) + ":" + pred); break; }
                    }
                if (alts.empty()) alts.push_back(item);
            }
            std::vector<std::vector<std::string>> next;
            for (auto& pre : out)
                for (auto& alt : alts) {
                    next.push_back(pre);
                    next.back().push_back(alt);
                }
            if (next.size() > 200000) throw Unsupported("more than 200000 duplicate decision paths (multi-way splits, SURVEY App. B-2)");
            out.swap(next);
            prefix.tagged.push_back(item);
        }
        return out;
    }

    // Parent iteration order of decPathsInfoContentBySplit.keySet() (tree/DecisionTreeBuilder.java:
    // 512, 640-650): parents enter `decPaths` in reducer key order (Tuple order, element-wise),
    // are re-inserted id-stripped in decPaths.keySet() order, then iterated.
    std::vector<int> parent_order(const Tree& tr, const std::vector<int>& live) const {
        std::vector<std::pair<std::vector<std::string>, int>> keys;
        for (int i : live)
            for (auto& v : tag_variants(tr.frontier[(size_t)i])) keys.emplace_back(v, i);
        std::stable_sort(keys.begin(), keys.end(), [](const auto& x, const auto& y) { return x.first < y.first; });
        JStrMap<int> dec_paths;
        for (auto& kv : keys) dec_paths.put(join_str(kv.first, path_delim_), kv.second);
        JStrMap<int> by_split;
        dec_paths.for_each([&](const std::string&, int& i) {
            std::vector<std::string> st;
            for (auto& it : tr.frontier[(size_t)i].tagged) st.push_back(strip_id(it));
            std::string k = join_str(st, path_delim_);
            if (!by_split.find(k)) by_split.put(k, i);
        });
        std::vector<int> out;
        by_split.for_each([&](const std::string&, int& i) { out.push_back(i); });
        return out;
    }

    // DecisionPathStoppingStrategy.shouldStop (tree/DecisionPathStoppingStrategy.java:57-70)
    bool should_stop(long long population, double stat, double parent_stat, int depth) const {
        if (stop_strategy_ == "minPopulation") return population < min_population_;
        if (stop_strategy_ == "minInfoGain") return (parent_stat - stat) < min_info_gain_;
        return depth >= max_depth_;
    }

    // One level for every tree: BuilderReducer.expandTree (tree/DecisionTreeBuilder.java:499-616).
    bool expand_level(std::vector<std::vector<DecPath>>& lists, std::vector<bool>& list_null) {
        const int C = (int)classes_.names.size();
        const int A = (int)attrs_.size();
        const int mw = (A + 31) / 32;
        auto npl = nodes_per_tree();                       // per device layer
        const int n_layers = (int)npl.size();
        std::vector<int> base((size_t)n_layers + 1, 0);
        for (int l = 0; l < n_layers; ++l) base[(size_t)l + 1] = base[(size_t)l] + npl[(size_t)l];
        const int NF = base[(size_t)n_layers];
        auto gidx = [&](const Node& nd) { return base[(size_t)nd.layer] + nd.id; };
        lists.assign((size_t)num_trees_, {});
        list_null.assign((size_t)num_trees_, true);
        if (NF == 0) return false;
        eng_->check(avt_level_histogram(eng_->raw(), npl.data(), nullptr));
        std::vector<uint32_t> mask((size_t)NF * mw, 0);
        for (int t = 0; t < num_trees_; ++t)
            for (Node& nd : trees_[(size_t)t].frontier) {
                nd.cand = candidates(nd);
                for (int a : nd.cand) mask[(size_t)gidx(nd) * mw + (a >> 5)] |= 1u << (a & 31);
            }
        std::vector<avt_node_best> best((size_t)NF);
        eng_->check(avt_level_scan(eng_->raw(), criterion(), mask.data(), classes_.order.data(), best.data()));

        // choose the split of every live node; ties go through the HashMap-order replay
        std::vector<int> chosen((size_t)NF, -1);
        std::vector<NodeOrder> orders((size_t)NF);
        std::vector<char> have_order((size_t)NF, 0);
        std::vector<uint32_t> hist((size_t)table_bins() * C);
        std::vector<double> av(table_.splits.size());
        auto need_order = [&](int g, const Node& nd) {
            if (have_order[(size_t)g]) return;
            eng_->check(avt_level_fetch_hist(eng_->raw(), g, hist.data()));
            orders[(size_t)g] = java_order(nd, hist);
            have_order[(size_t)g] = 1;
        };
        // av of every candidate split of node g, exactly as the reference sums it.  Two-predicate splits: the device value
        // (two terms commute).  More predicates: the weighted sum runs in predInfoContent.keySet() order and the terms do not
        // commute in floating point, so the per-predicate statistics come from the device and the sum is redone here.
        auto node_av = [&](int g, const Node& nd) {
            eng_->check(avt_level_fetch_av(eng_->raw(), g, av.data()));
            if (!multiway_) return;
            need_order(g, nd);
            std::vector<int32_t> nodes, splits;
            for (int a : nd.cand)
                for (size_t k = 0; k < attrs_[(size_t)a].splits.size(); ++k)
                    if (attrs_[(size_t)a].splits[k].n_seg > 2) { nodes.push_back(g); splits.push_back(table_.attr_first[(size_t)a] + (int)k); }
            if (nodes.empty()) return;
            std::vector<avt_split_eval> ev(nodes.size());
            eng_->check(avt_level_eval(eng_->raw(), criterion(), classes_.order.data(), (int32_t)nodes.size(), nodes.data(), splits.data(), ev.data()));
            for (size_t k = 0; k < splits.size(); ++k) {
                const int ti = splits[k];
                double weighted = 0.0;
                long long total = 0;
                for (int sg : orders[(size_t)g].seg_order_by_split[(size_t)ti]) {
                    long long n = 0;
                    for (int c = 0; c < C; ++c) n += ev[k].seg_count[sg][c];
                    weighted += ev[k].seg_stat[sg] * (double)n;
                    total += n;
                }
                av[(size_t)ti] = weighted / (double)total;
            }
        };
        for (int t = 0; t < num_trees_; ++t)
            for (const Node& nd : trees_[(size_t)t].frontier) {
                const int g = gidx(nd);
                const avt_node_best& b = best[(size_t)g];
                if (b.population == 0 || b.best_split < 0) continue;  // no rows or no candidate split: nothing is emitted for this parent
                chosen[(size_t)g] = b.best_split;
                if (split_sel_ != "best") continue;
                if (multiway_) {
                    // strict `<` from 1000000 over splitInfoContent.keySet() (tree/DecisionTreeBuilder.java:549-559)
                    node_av(g, nd);
                    double mn = 1000000.0;
                    for (int ti : orders[(size_t)g].split_order)
                        if (av[(size_t)ti] < mn) { mn = av[(size_t)ti]; chosen[(size_t)g] = ti; }
                } else if (b.n_ties > 1) {
                    need_order(g, nd);
                    eng_->check(avt_level_fetch_av(eng_->raw(), g, av.data()));
                    for (int ti : orders[(size_t)g].split_order)
                        if (av[(size_t)ti] == b.min_av) { chosen[(size_t)g] = ti; break; }
                }
            }
        if (split_sel_ == "randomAmongTop") {
            // selectRandomSplitFromTop (tree/DecisionTreeBuilder.java:549-572, 622-631): a TreeMap keyed by av
            // (equal keys overwrite, in splitInfoContent.keySet() order), its first dtb.top.split.count entries,
            // one uniform draw per parent -- parents in decPathsInfoContentBySplit.keySet() order
            for (int t = 0; t < num_trees_; ++t) {
                Tree& tr = trees_[(size_t)t];
                std::vector<int> live;
                for (size_t i = 0; i < tr.frontier.size(); ++i)
                    if (chosen[(size_t)gidx(tr.frontier[i])] >= 0) live.push_back((int)i);
                for (int i : parent_order(tr, live)) {
                    const Node& nd = tr.frontier[(size_t)i];
                    const int g = gidx(nd);
                    need_order(g, nd);
                    node_av(g, nd);
                    std::map<double, int> ranked;
                    for (int ti : orders[(size_t)g].split_order) ranked[av[(size_t)ti]] = ti;
                    std::vector<int> top;
                    for (auto& kv : ranked) {
                        top.push_back(kv.second);
                        if ((int)top.size() == top_split_count_) break;
                    }
                    if (top.empty()) continue;
                    chosen[(size_t)g] = top[(size_t)java_int_cast(tr.rng.next_double() * (double)top.size())];
                }
            }
        }
        std::vector<int32_t> ev_nodes, ev_splits;
        for (int g = 0; g < NF; ++g)
            if (chosen[(size_t)g] >= 0) { ev_nodes.push_back(g); ev_splits.push_back(chosen[(size_t)g]); }
        std::vector<avt_split_eval> evals(ev_nodes.size());
        eng_->check(avt_level_eval(eng_->raw(), criterion(), classes_.order.data(), (int32_t)ev_nodes.size(), ev_nodes.data(), ev_splits.data(), evals.data()));
        std::vector<int> eval_of((size_t)NF, -1);
        for (size_t k = 0; k < ev_nodes.size(); ++k) eval_of[(size_t)ev_nodes[k]] = (int)k;

        int lut_w = 1;
        for (auto& as : attrs_) lut_w = std::max(lut_w, as.n_bins);
        std::vector<int32_t> node_attr((size_t)NF, -1);
        std::vector<uint16_t> child_lut((size_t)NF * lut_w, (uint16_t)AVT_NODE_NONE);
        std::vector<int> next_id((size_t)n_layers, 0);   // next free node id of every existing layer
        struct Spawn { int layer, parent_layer, parent_id, attr, lo, hi; };
        std::vector<Spawn> spawns;                        // children that overlap a sibling: a layer of their own
        std::vector<int> new_layer_tree = layer_tree_;
        bool any = false;
        for (int t = 0; t < num_trees_; ++t) {
            Tree& tr = trees_[(size_t)t];
            std::vector<int> live;
            for (size_t i = 0; i < tr.frontier.size(); ++i)
                if (chosen[(size_t)gidx(tr.frontier[i])] >= 0) live.push_back((int)i);
            std::vector<Node> next;
            std::vector<DecPath>& out_list = lists[(size_t)t];
            for (int i : parent_order(tr, live)) {
                const Node& nd = tr.frontier[(size_t)i];
                const int g = gidx(nd);
                const int ti = chosen[(size_t)g];
                const avt_split& ds = table_.splits[(size_t)ti];
                const AttrSplits& as = attrs_[(size_t)ds.attr];
                const HostSplit& hs = as.splits[(size_t)(ti - table_.attr_first[(size_t)ds.attr])];
                const avt_split_eval& ev = evals[(size_t)eval_of[(size_t)g]];
                // running split id of the chosen split among this node's candidates (:299,325)
                int split_id = 0;
                for (int a : nd.cand) {
                    if (a == ds.attr) { split_id += ti - table_.attr_first[(size_t)a] + 1; break; }
                    split_id += (int)attrs_[(size_t)a].splits.size();
                }
                // children in predInfoContent.keySet() order
                std::vector<int> segs;
                std::vector<long long> seg_n((size_t)hs.n_seg, 0);
                for (int sg = 0; sg < hs.n_seg; ++sg) {
                    for (int c = 0; c < C; ++c) seg_n[(size_t)sg] += ev.seg_count[sg][c];
                    if (seg_n[(size_t)sg]) segs.push_back(sg);
                }
                if (segs.size() == 2) {
                    uint32_t b0 = StrHasher()(hs.pred_str[(size_t)segs[0]]) & 15u, b1 = StrHasher()(hs.pred_str[(size_t)segs[1]]) & 15u;
                    if (b0 == b1) {
                        need_order(g, nd);
                        segs = orders[(size_t)g].seg_order_by_split[(size_t)ti];
                    } else if (b1 < b0) {
                        std::swap(segs[0], segs[1]);
                    }
                } else if (segs.size() > 2) {
                    need_order(g, nd);
                    segs = orders[(size_t)g].seg_order_by_split[(size_t)ti];
                }
                // which children stay in the parent's layer: greedily, in predicate order, those whose bins are still free
                std::vector<char> bin_taken((size_t)as.n_bins, 0);
                std::vector<char> in_parent_layer((size_t)hs.n_seg, 0);
                for (int sg = 0; sg < hs.n_seg; ++sg) {
                    if (!seg_n[(size_t)sg]) continue;
                    bool free = true;
                    for (int b = 0; b < as.n_bins && free; ++b) free = !(split_bin_in_segment(as, hs, b, sg) && bin_taken[(size_t)b]);
                    if (!free) continue;
                    in_parent_layer[(size_t)sg] = 1;
                    for (int b = 0; b < as.n_bins; ++b)
                        if (split_bin_in_segment(as, hs, b, sg)) bin_taken[(size_t)b] = 1;
                }
                // parent predicates (tree/DecisionPathList.java:454-472)
                std::vector<PathPred> parent_preds;
                for (auto& it : nd.tagged) {
                    if (it == kRoot) { parent_preds.push_back(root_pred()); continue; }
                    std::string s = strip_id(it);
                    parent_preds.push_back(make_pred(s, schema_.by_ordinal(std::stoi(s.substr(0, s.find(' '))))));
                }
                const double parent_stat = nd.path.info_content;
                node_attr[(size_t)g] = ds.attr;
                for (int sg : segs) {
                    DecPath d;
                    d.has_preds = true;
                    d.preds = parent_preds;
                    d.preds.push_back(make_pred(hs.pred_str[(size_t)sg], *as.field));
                    d.population = seg_n[(size_t)sg];
                    d.info_content = ev.seg_stat[sg];
                    d.stopped = should_stop(d.population, d.info_content, parent_stat, (int)parent_preds.size() + 1);
                    d.class_val_pr = class_pr(ev.seg_count[sg], d.population);
                    majority(d);
                    if (total_population_ > 0) d.support = (double)d.population / (double)total_population_;
                    Node ch;
                    ch.tagged = nd.tagged;
                    ch.tagged.push_back(std::to_string(split_id) + ":" + hs.pred_str[(size_t)sg]);
                    ch.path = d;
                    if (in_parent_layer[(size_t)sg]) {
                        ch.layer = nd.layer;
                        ch.id = next_id[(size_t)nd.layer]++;
                        if (ch.id >= (int)AVT_NODE_NONE - 1) throw Unsupported("more than 65534 frontier nodes in one tree");
                        for (int b = 0; b < as.n_bins; ++b)
                            if (split_bin_in_segment(as, hs, b, sg)) child_lut[(size_t)g * lut_w + b] = (uint16_t)ch.id;
                    } else {
                        ch.layer = (int)new_layer_tree.size();
                        ch.id = 0;
                        new_layer_tree.push_back(t);
                        int lo = as.n_bins, hi = -1;
                        for (int b = 0; b < as.n_bins; ++b)
                            if (split_bin_in_segment(as, hs, b, sg)) { lo = std::min(lo, b); hi = std::max(hi, b); }
                        spawns.push_back({ch.layer, nd.layer, nd.id, ds.attr, lo, hi});
                    }
                    next.push_back(ch);
                    out_list.push_back(d);
                }
            }
            list_null[(size_t)t] = live.empty();
            any = any || !next.empty();
            tr.frontier.swap(next);
        }
        // rows of the spawned children: members of the parent node (ids BEFORE the repartition) whose bin is in the child's range
        std::vector<std::vector<uint16_t>> spawn_nid;
        if (!spawns.empty()) {
            std::map<int, std::vector<uint16_t>> parent_nid;
            for (auto& sp : spawns)
                if (!parent_nid.count(sp.parent_layer)) {
                    auto& v = parent_nid[sp.parent_layer];
                    v.resize(n_local());
                    eng_->check(avt_forest_node_ids(eng_->raw(), sp.parent_layer, v.data()));
                }
            for (auto& sp : spawns) {
                const auto& pn = parent_nid[sp.parent_layer];
                std::vector<uint16_t> v(n_local(), (uint16_t)AVT_NODE_NONE);
                const bool wide = attrs_[(size_t)sp.attr].n_bins > 256;
                need_host_column((size_t)sp.attr);
                for (size_t r = 0; r < n_local(); ++r) {
                    if (pn[r] != (uint16_t)sp.parent_id) continue;
                    const int b = wide ? col16_[(size_t)sp.attr][lo_ + r] : col8_[(size_t)sp.attr][lo_ + r];
                    if (b >= sp.lo && b <= sp.hi) v[r] = 0;
                }
                spawn_nid.push_back(std::move(v));
            }
        }
        eng_->check(avt_level_repartition(eng_->raw(), node_attr.data(), child_lut.data(), lut_w));
        if (!spawns.empty()) {
            // the forest grows by the new layers: existing layers keep their (repartitioned) node ids
            std::vector<std::vector<uint16_t>> old_nid((size_t)n_layers, std::vector<uint16_t>(n_local()));
            for (int l = 0; l < n_layers; ++l) eng_->check(avt_forest_node_ids(eng_->raw(), l, old_nid[(size_t)l].data()));
            layer_tree_ = new_layer_tree;
            std::vector<const uint16_t*> np;
            std::vector<const uint8_t*> wp;
            for (int l = 0; l < (int)layer_tree_.size(); ++l) {
                np.push_back(l < n_layers ? old_nid[(size_t)l].data() : spawn_nid[(size_t)(l - n_layers)].data());
                const Tree& tr = trees_[(size_t)layer_tree_[(size_t)l]];
                wp.push_back(tr.weights.empty() ? nullptr : tr.weights.data() + lo_);
            }
            eng_->check(avt_forest_init(eng_->raw(), (int32_t)layer_tree_.size(), wp.data(), np.data()));
        }
        return any;
    }

    int table_bins() const {
        int tb = 0;
        for (auto& a : attrs_) tb += a.n_bins;
        return tb;
    }

    std::string list_path(int t) const {
        std::string p = conf_.get("dtb.decision.file.path.out");
        return num_trees_ == 1 ? p : p + ".tree" + std::to_string(t);
    }

    // writeDecisioList (tree/DecisionTreeBuilder.java:713-725)
    void write_lists(const std::vector<std::vector<DecPath>>& lists, const std::vector<bool>* is_null = nullptr) {
        for (int t = 0; t < num_trees_; ++t) {
            if (rank_ != 0) {   // every rank holds the same lists; one writes them
                const bool null_list = is_null && (*is_null)[(size_t)t];
                if (null_list && total_population_ > 0) throw ReferenceError("NullPointerException in DecisionPathList.withTotalPopulation: no decision path was generated");
                continue;
            }
            const bool null_list = is_null && (*is_null)[(size_t)t];
            if (null_list && total_population_ > 0) throw ReferenceError("NullPointerException in DecisionPathList.withTotalPopulation: no decision path was generated");
            const std::string text = path_list_json(null_list ? nullptr : &lists[(size_t)t]) + "\n";
            write_text(list_path(t), text);
            // with the in-process level loop every pass is also kept (the reference overwrites one file per job)
            if (num_levels_ > 1) write_text(list_path(t) + ".level" + std::to_string(pass_), text);
        }
        ++pass_;
    }

    std::string data_dir(int t) const { return num_trees_ == 1 ? out_path_ : out_path_ + "/tree" + std::to_string(t); }

    // "$root,<row>" for every sampled row (tree/DecisionTreeBuilder.java:248-258, 761-766)
    void write_root_data() {
        for (int t = 0; t < num_trees_; ++t) {
            std::vector<std::string> out;
            const Tree& tr = trees_[(size_t)t];
            for (size_t r = lo_; r < hi_; ++r) {
                int w = tr.weights.empty() ? 1 : tr.weights[r];
                for (int k = 0; k < w; ++k) out.push_back(std::string(kRoot) + delim_out_ + rows_[r]);
            }
            write_output_lines(data_dir(t), out, rank_);
        }
    }

    // After expansions: every row under its new path, selected split only (the records the next
    // pass's mapper would keep, tree/DecisionTreeBuilder.java:279-292), in
    // reducer key order.
    void write_level_data() {
        std::vector<std::vector<uint16_t>> nid(layer_tree_.size());
        for (int t = 0; t < num_trees_; ++t) {
            const Tree& tr = trees_[(size_t)t];
            std::vector<int> order(tr.frontier.size());
            for (size_t i = 0; i < order.size(); ++i) order[i] = (int)i;
            std::sort(order.begin(), order.end(), [&](int x, int y) { return tr.frontier[(size_t)x].tagged < tr.frontier[(size_t)y].tagged; });
            // rows of every frontier node: its layer's node ids
            std::map<std::pair<int, int>, int> node_of;   // (layer, id) -> index in the frontier
            for (size_t i = 0; i < tr.frontier.size(); ++i) node_of[{tr.frontier[i].layer, tr.frontier[i].id}] = (int)i;
            std::vector<std::vector<size_t>> rows_of(tr.frontier.size());
            for (size_t l = 0; l < layer_tree_.size(); ++l) {
                if (layer_tree_[l] != t) continue;
                nid[l].resize(n_local());
                eng_->check(avt_forest_node_ids(eng_->raw(), (int32_t)l, nid[l].data()));
                for (size_t r = 0; r < n_local(); ++r) {
                    if (nid[l][r] == AVT_NODE_NONE) continue;
                    auto it = node_of.find({(int)l, (int)nid[l][r]});
                    if (it != node_of.end()) rows_of[(size_t)it->second].push_back(r);
                }
                nid[l].clear();
                nid[l].shrink_to_fit();
            }
            std::vector<std::string> out;
            for (int i : order) {
                const std::string key = join_str(tr.frontier[(size_t)i].tagged, path_delim_);
                for (size_t r : rows_of[(size_t)i]) {   // r is local to the shard
                    int w = tr.weights.empty() ? 1 : tr.weights[lo_ + r];
                    for (int k = 0; k < w; ++k) out.push_back(key + delim_out_ + rows_[lo_ + r]);
                }
            }
            write_output_lines(data_dir(t), out, rank_);
        }
    }
};

}  // namespace

int run_decision_tree_builder(const Props& conf, const std::string& in_path, const std::string& out_path) {
    DtbJob job(conf, in_path, out_path);
    return job.run();
}

}  // namespace avt_host
