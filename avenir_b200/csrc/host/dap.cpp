// org.avenir.tree.DataPartitioner re-hosted above the kernel-level C-ABI (the older node-wise pipeline,
// SplitGenerator -> DataPartitioner).
//
// The reference (tree/DataPartitioner.java) reads the candidate-split lines SplitGenerator left in the node's sibling
// directory `splits/part-r-00000`, sorts them by their third `;`-separated field in descending order (:157-226), takes the
// first one (or a random one of the top few), and runs a map/reduce job that keys every row of the node's data with the
// segment its value of the split attribute falls into (:325-337, util/AttributeSplitHandler.java:146-153, 190-204), one
// reducer per segment, and finally renames the part files into `split=<line index>/segment=<i>/data/partition.txt` (:114-129).
// Here the rows' text goes to the device once (avt_dataset_load_csv with the split attribute as the only encoded field; the
// bins are the split's own points or value groups), and avt_rows_partition scatters the rows into their segments there.
//
// Kept quirks: the lines are split on ';' whatever `field.delim.out` the generator used, and the segment count is the number
// of ':'-separated pieces of the key -- an integer split's key written by ClassPartitionGenerator has no
// ':' (its points are joined with ';'), so it has ONE reducer and both sides of the split land in segment=0, rows below the
// point first (reducer input is sorted by key).  Assumptions where the reference leans on code that is not in its tree
// (chombo): SecondarySort.RawIntKeyTextPartitioner sends key k to reducer k mod numReducers; Utility.getFileLines returns
// the file's lines; Utility.intArrayFromString is split + Integer.parseInt; rows of one key keep their input order.
// Extension: dap.split.selection.strategy=randomFromTop draws from java.util.Random(avt.random.seed) instead of the
// reference's unseeded Math.random().
#include <sys/stat.h>

#include <algorithm>
#include <cstring>
#include <memory>

#include "javaemu.hpp"
#include "tool.hpp"

namespace avt_host {

namespace {

void make_dirs(const std::string& dir) {
    std::string cur;
    for (size_t i = 0; i <= dir.size(); ++i) {
        if (i == dir.size() || dir[i] == '/') {
            if (!cur.empty()) mkdir(cur.c_str(), 0777);
        }
        if (i < dir.size()) cur.push_back(dir[i]);
    }
}

void write_file(const std::string& path, const char* data, size_t n) {
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) throw std::runtime_error("cannot write " + path);
    if (n) fwrite(data, 1, n, f);
    fclose(f);
}

bool blank(const std::string& s) { return trim(s).empty(); }

// Double.parseDouble as far as a statistics line can exercise it: surrounding white space ignored, decimal / exponent forms
double parse_double_or_throw(const std::string& raw) {
    const std::string t = trim(raw);
    char* end = nullptr;
    const double v = t.empty() ? 0.0 : strtod(t.c_str(), &end);
    if (t.empty() || *end) throw ReferenceError("NumberFormatException: For input string: \"" + raw + "\"");
    return v;
}

// DataPartitioner.Split (tree/DataPartitioner.java:208-275)
struct CandidateLine {
    int index;
    std::vector<std::string> items;
};

const std::string& item(const CandidateLine& c, size_t k) {
    if (k >= c.items.size()) throw ReferenceError("ArrayIndexOutOfBoundsException: " + std::to_string(k));
    return c.items[k];
}

}  // namespace

int run_data_partitioner(const Props& conf) {
    // getNodePath (:137-151)
    const std::string base = conf.get("dap.project.base.path");
    if (blank(base)) throw ReferenceError("IllegalStateException: base path not defined");
    const std::string split_path = conf.get("dap.split.path");
    const std::string in_path = blank(split_path) ? base + "/split=root/data" : base + "/split=root/data/" + split_path;
    // findBestSplitKey (:160-206)
    const std::string candidates = in_path.substr(0, in_path.find_last_of('/')) + "/splits/part-r-00000";   // Utility.getSiblingPath
    std::vector<CandidateLine> splits;
    {
        std::vector<std::string> lines = read_input_lines(candidates);
        for (size_t i = 0; i < lines.size(); ++i) splits.push_back(CandidateLine{(int)i, split_str(lines[i], ";", false)});
    }
    if (splits.size() > 1) {
        // Arrays.sort(Object[]) is a stable merge sort and compares every element at least once: a line without a third
        // field or with an unparsable one fails the job here
        std::vector<double> val;
        for (auto& c : splits) {
            val.push_back(parse_double_or_throw(item(c, 2)));
            if (val.back() != val.back()) throw Unsupported("a NaN statistic in " + candidates + ": the reference's comparator is then inconsistent");
        }
        std::vector<size_t> order(splits.size());
        for (size_t i = 0; i < order.size(); ++i) order[i] = i;
        std::stable_sort(order.begin(), order.end(), [&](size_t a, size_t b) { return val[a] > val[b]; });
        std::vector<CandidateLine> sorted;
        for (size_t i : order) sorted.push_back(splits[i]);
        splits.swap(sorted);
    }
    size_t pick = 0;
    if (conf.get("dap.split.selection.strategy", "best") == "randomFromTop") {
        if (!conf.nonblank("avt.random.seed"))
            throw Unsupported("dap.split.selection.strategy=randomFromTop draws from unseeded Math.random(); set avt.random.seed to draw from java.util.Random(seed)");
        JavaRandom rng(conf.get_int("avt.random.seed", 0));
        pick = (size_t)(int)(rng.next_double() * conf.get_int("dap.num.top.splits", 5));
    }
    if (pick >= splits.size()) throw ReferenceError("ArrayIndexOutOfBoundsException: " + std::to_string(pick));
    const CandidateLine& chosen = splits[pick];
    long long attr_ord;
    if (!parse_java_int(item(chosen, 0), attr_ord)) throw ReferenceError("NumberFormatException: For input string: \"" + item(chosen, 0) + "\"");
    const std::string key = item(chosen, 1);
    const std::vector<std::string> pieces = split_str(key, ":", false);
    const int n_reducers = (int)pieces.size();
    const std::string out_path = in_path + "/split=" + std::to_string(chosen.index);

    // PartitionerMapper.setup (:291-318)
    const std::string delim = conf.get("field.delim.regex", ",");
    if (!conf.nonblank("dap.feature.schema.file.path")) throw ReferenceError("NullPointerException: dap.feature.schema.file.path not set");
    const Schema schema = Schema::load(conf.get("dap.feature.schema.file.path"));
    const Field* field = nullptr;
    for (auto& f : schema.fields)
        if (f.ordinal == (int)attr_ord) field = &f;
    if (!field) throw ReferenceError("NullPointerException: no field with ordinal " + std::to_string(attr_ord) + " in the schema");
    if (!field->is_int() && !field->is_categorical())
        throw ReferenceError("NullPointerException: split attribute " + std::to_string(attr_ord) + " is neither integer nor categorical (split stays null)");
    std::vector<int32_t> points;               // IntegerSplit.fromString
    std::vector<std::vector<std::string>> groups;   // CategoricalSplit.fromString
    if (field->is_int()) {
        for (auto& p : pieces) {
            long long v;
            if (!parse_java_int(p, v)) throw ReferenceError("NumberFormatException: For input string: \"" + p + "\"");
            points.push_back((int32_t)v);
        }
        for (size_t i = 1; i < points.size(); ++i)
            if (points[i] <= points[i - 1]) throw Unsupported("split points of key '" + key + "' are not increasing");
    } else {
        for (auto& p : pieces) {
            if (p.size() < 2) throw ReferenceError("StringIndexOutOfBoundsException: String index out of range: " + std::to_string((int)p.size() - 1));
            std::vector<std::string> g;
            for (auto& v : split_str(p.substr(1, p.size() - 2), ",", false)) g.push_back(trim(v));
            groups.push_back(g);
        }
    }
    // number of segments a row can be keyed with: points + 1, or the groups
    const int n_keys = field->is_int() ? (int)points.size() + 1 : (int)groups.size();

    if (delim.size() != 1 || strchr(".*+?()[]{}|^$\\", delim[0])) throw Unsupported("field.delim.regex must be a single literal character for the device encoder");
    std::vector<std::string> lines = read_input_lines(in_path);
    std::string text;
    for (auto& ln : lines) {
        text += ln;
        text += '\n';
    }
    // the split attribute as the only encoded field: bins = the split's own grid, or one bin per listed value
    avt_csv_attr attr{(int32_t)attr_ord, field->is_int() ? AVT_CSV_INT : AVT_CSV_CATEGORICAL, 0, 0};
    std::string cat_bytes;
    std::vector<int32_t> cat_off;
    std::vector<uint16_t> seg_of_bin;   // bin -> reducer (key mod reducers); reducer input is sorted by key
    std::vector<int> key_of_bin;
    if (field->is_int()) {
        attr.n = (int32_t)points.size();
        for (int k = 0; k < n_keys; ++k) key_of_bin.push_back(k);
    } else {
        std::vector<std::string> seen;
        for (size_t g = 0; g < groups.size(); ++g)
            for (auto& v : groups[g]) {
                if (std::find(seen.begin(), seen.end(), v) != seen.end()) continue;   // List.contains: the first group that lists a value wins
                seen.push_back(v);
                cat_off.push_back((int32_t)cat_bytes.size());
                cat_bytes += v;
                key_of_bin.push_back((int)g);
            }
        cat_off.push_back((int32_t)cat_bytes.size());
        attr.n = (int32_t)seen.size();
        if (seen.empty()) throw ReferenceError("IllegalArgumentException: split segment not found for " + (lines.empty() ? std::string() : lines[0]));
    }
    // Rows keyed k go to reducer k mod n_reducers; inside a reducer they arrive sorted by key.  The device scatters by KEY
    // (n_keys segments, input order inside each); the reducers' files are those segments concatenated in key order.
    for (int k : key_of_bin) seg_of_bin.push_back((uint16_t)k);

    std::vector<std::string> reducer_out((size_t)n_reducers);
    if (!text.empty()) {
        Engine eng(conf.get_int("avt.device", 0));
        int64_t n_rows = 0, bad_row = -1;
        int32_t bad_reason = 0;
        const int rc = avt_dataset_load_csv(eng.raw(), text.data(), (int64_t)text.size(), delim[0], 0, 1, &attr, points.empty() ? nullptr : points.data(), nullptr,
                                            cat_bytes.data(), cat_off.empty() ? nullptr : cat_off.data(), -1, 1, nullptr, nullptr, &n_rows, nullptr, &bad_row, &bad_reason);
        if (rc == AVT_E_REFERENCE && bad_row >= 0 && bad_row < (int64_t)lines.size()) {
            // what PartitionerMapper.map throws on that row (:325-337)
            const std::vector<std::string> items = split_str(lines[(size_t)bad_row], delim, false);
            if ((size_t)attr_ord >= items.size()) throw ReferenceError("ArrayIndexOutOfBoundsException: " + std::to_string(attr_ord));
            const std::string& v = items[(size_t)attr_ord];
            if (field->is_int()) throw ReferenceError("NumberFormatException: For input string: \"" + v + "\"");
            throw ReferenceError("IllegalArgumentException: split segment not found for " + v);
        }
        eng.check(rc);
        std::vector<int64_t> seg_bytes((size_t)n_keys), seg_rows((size_t)n_keys);
        std::string out((size_t)text.size() + 1, '\0');
        eng.check(avt_rows_partition(eng.raw(), 0, seg_of_bin.data(), n_keys, &out[0], (int64_t)out.size(), seg_bytes.data(), seg_rows.data()));
        size_t at = 0;
        for (int k = 0; k < n_keys; ++k) {
            reducer_out[(size_t)(k % n_reducers)].append(out, at, (size_t)seg_bytes[(size_t)k]);
            at += (size_t)seg_bytes[(size_t)k];
        }
    }
    // moveOutputToSegmentDir (:114-129): part-r-0000<i> -> segment=<i>/data/partition.txt
    if (n_reducers > 10) throw Unsupported("more than 10 segments: the reference's rename looks for part-r-0000<i> with a two-digit i, which the job never writes");
    for (int i = 0; i < n_reducers; ++i) {
        const std::string dir = out_path + "/segment=" + std::to_string(i) + "/data";
        make_dirs(dir);
        write_file(dir + "/partition.txt", reducer_out[(size_t)i].data(), reducer_out[(size_t)i].size());
    }
    write_file(out_path + "/_SUCCESS", "", 0);
    return 0;
}

}  // namespace avt_host
