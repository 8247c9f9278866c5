// ClassPartitionGenerator re-hosted above the kernel-level C-ABI
// (explore/ClassPartitionGenerator.java; util/AttributeSplitHandler.java; util/AttributeSplitStat.java).
//
// Same cpg.* knobs, schema JSON, CSV rows in, `attr,splitKey,gainRatio[,seg,class,prob...]` (or
// `attr,splitKey,stat`) lines out.  Kernel (a) produces the (attr, bin, class) counts the mapper +
// combiner + reducer accumulate as text tuples; the CPG scan kernel evaluates every candidate split.
#include <algorithm>
#include <memory>

#include "tool.hpp"

namespace avt_host {

namespace {

int criterion_of(const std::string& alg) {
    if (alg == "entropy") return AVT_CRIT_ENTROPY;
    if (alg == "giniIndex") return AVT_CRIT_GINI;
    if (alg == "hellingerDistance") return AVT_CRIT_HELLINGER;
    return AVT_CRIT_CLASS_CONF;  // util/AttributeSplitStat.java:70-78: anything else
}

}  // namespace

int run_class_partition_generator(const Props& conf, const std::string& in_path, const std::string& out_path) {
    const std::string delim_regex = conf.get("field.delim.regex", ",");
    const std::string delim_out = conf.get("field.delim.out", ",");
    for (char c : delim_regex)
        if (strchr(".*+?()[]{}|^$\\", c)) throw Unsupported("delimiter '" + delim_regex + "' contains regex metacharacters; only literal delimiters are implemented");
    if (!conf.nonblank("cpg.feature.schema.file.path")) throw ReferenceError("NullPointerException: cpg.feature.schema.file.path not set");
    Schema schema = Schema::load(conf.get("cpg.feature.schema.file.path"));
    const std::string algorithm = conf.get("cpg.split.algorithm", "giniIndex");
    const int max_cat = conf.get_int("cpg.max.cat.attr.split.groups", 3);
    const bool mapper_at_root = conf.get_bool("cpg.at.root", false);
    const std::string strategy = conf.get("cpg.split.attribute.selection.strategy", "userSpecified");
    const bool reducer_at_root = !conf.has("cpg.split.attributes");
    const bool output_prob = conf.get_bool("cpg.output.split.prob", false);
    if (!conf.nonblank("cpg.parent.info")) throw ReferenceError("NullPointerException: cpg.parent.info not set (Double.parseDouble(null))");
    const double parent_info = conf.get_double("cpg.parent.info", 0);
    const int class_ord = schema.class_field().ordinal;

    // mapper-side attribute choice (explore/ClassPartitionGenerator.java:160-196)
    std::vector<int> split_attrs;
    if (!mapper_at_root) {
        if (strategy == "userSpecified") {
            if (reducer_at_root) throw ReferenceError("NullPointerException: cpg.split.attributes not set");
            for (auto& s : split_str(conf.get("cpg.split.attributes"), ",", false)) split_attrs.push_back(std::stoi(trim(s)));
        } else if (strategy == "all") {
            split_attrs = schema.feature_ordinals();
        } else if (strategy == "notUsedYet") {
            throw ReferenceError("NullPointerException: cpg.split.attribute.selection.strategy=notUsedYet is a TODO in the reference (explore/ClassPartitionGenerator.java:171-176)");
        } else if (strategy == "random") {
            throw Unsupported("cpg.split.attribute.selection.strategy=random draws from unseeded Math.random(); not implemented");
        } else {
            throw ReferenceError("invalid splitting attribute selection strategy");
        }
    }
    if (mapper_at_root != reducer_at_root)
        throw ReferenceError(mapper_at_root ? "NullPointerException: cpg.at.root=true but cpg.split.attributes is set (reducer looks up attribute -1)"
                                            : "NullPointerException: cpg.split.attributes missing while cpg.at.root=false");

    std::vector<AttrSplits> attrs;
    if (mapper_at_root) {
        // only the class column matters; a constant one-bin attribute carries the rows
        AttrSplits as;
        as.ordinal = -1;
        as.n_bins = 1;
        as.numeric = true;
        HostSplit h;
        h.bounds = {0};
        as.splits.push_back(h);
        attrs.push_back(as);
    } else {
        for (int o : split_attrs) attrs.push_back(cpg_attr_splits(schema.by_ordinal(o), max_cat));
    }

    std::vector<std::string> lines = read_input_lines(in_path);
    const size_t n = lines.size();
    if (n == 0) {
        if (reducer_at_root) {
            write_output_lines(out_path, {"0.0"});
            return 0;
        }
        write_output_lines(out_path, {});
        return 0;
    }
    std::vector<std::vector<uint8_t>> col8(attrs.size());
    std::vector<std::vector<uint16_t>> col16(attrs.size());
    for (size_t a = 0; a < attrs.size(); ++a) {
        if (attrs[a].n_bins > 256) col16[a].assign(n, 0);
        else col8[a].assign(n, 0);
    }
    std::vector<uint8_t> cls(n);
    ClassTable classes;
    for (size_t r = 0; r < n; ++r) {
        std::vector<std::string> items = split_str(lines[r], delim_regex, false);
        if (class_ord >= (int)items.size()) throw ReferenceError("ArrayIndexOutOfBoundsException: class attribute missing in row: " + lines[r]);
        cls[r] = (uint8_t)classes.code_of(items[(size_t)class_ord]);
        if (mapper_at_root) continue;
        for (size_t a = 0; a < attrs.size(); ++a) {
            const int ord = attrs[a].ordinal;
            if (ord >= (int)items.size()) throw ReferenceError("ArrayIndexOutOfBoundsException: attribute " + std::to_string(ord) + " missing in row: " + lines[r]);
            int b;
            try {
                b = encode_value(attrs[a], items[(size_t)ord]);
            } catch (const ReferenceError& e) {
                if (!attrs[a].numeric) throw ReferenceError("split segment not found for " + items[(size_t)ord]);
                throw;
            }
            if (attrs[a].n_bins > 256) col16[a][r] = (uint16_t)b;
            else col8[a][r] = (uint8_t)b;
        }
    }
    classes.finish();
    const int C = (int)classes.names.size();
    const int crit = criterion_of(algorithm);
    if (!reducer_at_root && crit == AVT_CRIT_HELLINGER && C != 2)
        throw ReferenceError("Hellinger distance algorithm is only valid for binary valued class attributes");

    Engine eng(conf.get_int("avt.device", 0));
    std::vector<int32_t> nb;
    std::vector<const void*> cols;
    for (size_t a = 0; a < attrs.size(); ++a) {
        nb.push_back(attrs[a].n_bins);
        cols.push_back(attrs[a].n_bins > 256 ? (const void*)col16[a].data() : (const void*)col8[a].data());
    }
    eng.check(avt_dataset_load(eng.raw(), (int64_t)n, (int32_t)attrs.size(), nb.data(), cols.data(), cls.data(), C));
    DeviceSplitTable table = flatten_splits(attrs);
    eng.check(avt_split_table_set(eng.raw(), (int32_t)table.splits.size(), table.splits.data(), table.bounds.data(), (int32_t)table.bounds.size(),
                                  table.lut.data(), (int64_t)table.lut.size()));
    eng.check(avt_forest_init(eng.raw(), 1, nullptr, nullptr));
    const int32_t one = 1;
    eng.check(avt_level_histogram(eng.raw(), &one, nullptr));

    std::vector<std::string> out;
    if (reducer_at_root) {
        // data-set impurity (explore/ClassPartitionGenerator.java:516-519)
        avt_node_best nb0;
        eng.check(avt_level_scan(eng.raw(), algorithm == "entropy" ? AVT_CRIT_ENTROPY : AVT_CRIT_GINI, nullptr, classes.order.data(), &nb0));
        out.push_back(java_double_to_string(nb0.node_stat));
        write_output_lines(out_path, out);
        return 0;
    }
    std::vector<avt_cpg_stat> stats(table.splits.size());
    eng.check(avt_cpg_sweep(eng.raw(), 0, crit, classes.order.data(), stats.data()));

    // reducer cleanup (explore/ClassPartitionGenerator.java:520-553): attributes in
    // cpg.split.attributes order; split keys in the iteration order of `stats`, a HashMap filled in
    // splitStats.keySet() order, itself filled in reducer key order (sorted split keys).
    std::vector<int> attr_ordinals;
    for (auto& s : split_str(conf.get("cpg.split.attributes"), ",", false)) attr_ordinals.push_back(std::stoi(trim(s)));
    for (int ord : attr_ordinals) {
        int a = -1;
        for (size_t i = 0; i < attrs.size(); ++i)
            if (attrs[i].ordinal == ord) { a = (int)i; break; }
        if (a < 0) continue;  // the mapper emitted nothing for it: AttributeSplitStat stays empty
        const AttrSplits& as = attrs[(size_t)a];
        std::vector<std::pair<std::string, int>> keys;
        for (size_t k = 0; k < as.splits.size(); ++k) keys.emplace_back(as.splits[k].key, (int)k);
        std::sort(keys.begin(), keys.end());
        JStrMap<int> split_stats;
        for (auto& kv : keys) split_stats.put(kv.first, kv.second);  // duplicate keys collapse like the reference's map
        JStrMap<int> stat_map;
        split_stats.for_each([&](const std::string& k, int& v) { stat_map.put(k, v); });
        stat_map.for_each([&](const std::string& key, int& k) {
            const avt_cpg_stat& st = stats[(size_t)(table.attr_first[(size_t)a] + k)];
            std::string line = std::to_string(ord) + delim_out + key + delim_out;
            if (crit == AVT_CRIT_GINI || crit == AVT_CRIT_ENTROPY) {
                const double gain = parent_info - st.stat;
                line += java_double_to_string(gain / st.split_info);
                if (output_prob) {
                    // serializeClassProbab (:557-566): segments in HashMap<Integer> order, classes in
                    // the segment's classValPr order
                    const HostSplit& h = as.splits[(size_t)k];
                    std::string frag;
                    for (int sg = 0; sg < h.n_seg; ++sg) {
                        long long tot = 0;
                        for (int c = 0; c < C; ++c) tot += st.seg_count[sg][c];
                        if (!tot) continue;
                        for (uint8_t c : classes.order) {
                            if (!st.seg_count[sg][c]) continue;
                            frag += delim_out + std::to_string(sg) + delim_out + classes.names[c] + delim_out +
                                    java_double_to_string((double)st.seg_count[sg][c] / (double)tot);
                        }
                    }
                    line += frag;
                }
            } else {
                line += java_double_to_string(st.stat);
            }
            out.push_back(line);
        });
    }
    write_output_lines(out_path, out);
    return 0;
}

}  // namespace avt_host
