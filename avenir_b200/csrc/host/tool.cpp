// avt_tool_run: ToolRunner.run(new <Tool>(), args) for the two tree-path jobs
// (tree/DecisionTreeBuilder.java:773-776; explore/ClassPartitionGenerator.java:611-614).
#include "tool.hpp"

#include <cstdio>
#include <cstring>

using namespace avt_host;

namespace {

void set_err(char* errbuf, size_t errlen, const std::string& m) {
    if (errbuf && errlen) {
        snprintf(errbuf, errlen, "%s", m.c_str());
    }
}

bool ends_with(const std::string& s, const std::string& suf) {
    return s.size() >= suf.size() && s.compare(s.size() - suf.size(), suf.size(), suf) == 0;
}

}  // namespace

extern "C" int avt_tool_run(const char* tool, int argc, const char* const* argv, char* errbuf, size_t errlen) {
    if (errbuf && errlen) errbuf[0] = 0;
    try {
        if (!tool) throw std::runtime_error("tool name is NULL");
        const std::string name(tool);
        Props conf;
        std::vector<std::string> positional;
        std::vector<std::pair<std::string, std::string>> overrides;
        for (int i = 0; i < argc; ++i) {
            std::string a = argv[i] ? argv[i] : "";
            if (a.rfind("-D", 0) == 0) {
                std::string kv = a.size() > 2 ? a.substr(2) : (i + 1 < argc ? std::string(argv[++i]) : std::string());
                size_t p = kv.find('=');
                if (p == std::string::npos) throw std::runtime_error("malformed -D option: " + a);
                overrides.emplace_back(kv.substr(0, p), kv.substr(p + 1));
            } else {
                positional.push_back(a);
            }
        }
        // chombo Utility.setConfiguration(conf, "avenir"): the properties file named by -Dconf.path
        for (auto& kv : overrides)
            if (kv.first == "conf.path") conf.load(kv.second);
        for (auto& kv : overrides)
            if (kv.first != "conf.path") conf.set(kv.first, kv.second);
        if (ends_with(name, "SplitGenerator")) {
            // tree/SplitGenerator.java:39-54: ClassPartitionGenerator whose paths come from the project layout instead of the arguments
            const std::string base = conf.get("project.base.path");
            if (trim(base).empty()) throw ReferenceError("IllegalStateException: base path not defined");
            const std::string split_path = conf.get("split.path");
            const std::string in = trim(split_path).empty() ? base + "/split=root/data" : base + "/split=root/data/" + split_path;
            const size_t cut = in.find_last_of('/');
            return run_class_partition_generator(conf, in, in.substr(0, cut) + "/splits");   // Utility.getSiblingPath(inPath, "splits")
        }
        if (ends_with(name, "DataPartitioner")) return run_data_partitioner(conf);   // paths come from dap.project.base.path / dap.split.path
        if (positional.size() < 2) throw std::runtime_error("usage: <tool> -Dconf.path=<file.properties> [-Dkey=value ...] <inPath> <outPath>");
        if (ends_with(name, "DecisionTreeBuilder")) return run_decision_tree_builder(conf, positional[0], positional[1]);
        if (ends_with(name, "ClassPartitionGenerator")) return run_class_partition_generator(conf, positional[0], positional[1]);
        if (ends_with(name, "ModelPredictor")) return run_model_predictor(conf, positional[0], positional[1]);
        if (ends_with(name, "CramerCorrelation")) return run_cramer_correlation(conf, positional[0], positional[1]);
        if (ends_with(name, "HeterogeneityReductionCorrelation")) return run_heterogeneity_reduction_correlation(conf, positional[0], positional[1]);
        if (ends_with(name, "CategoricalClassAffinity")) return run_categorical_class_affinity(conf, positional[0], positional[1]);
        if (ends_with(name, "CategoricalContinuousEncoding")) return run_categorical_continuous_encoding(conf, positional[0], positional[1]);
        throw std::runtime_error("unknown tool '" + name + "' (DecisionTreeBuilder | ClassPartitionGenerator | ModelPredictor | SplitGenerator | DataPartitioner | CramerCorrelation | HeterogeneityReductionCorrelation | CategoricalContinuousEncoding | CategoricalClassAffinity)");
    } catch (const ReferenceError& e) {
        set_err(errbuf, errlen, std::string("reference failure reproduced: ") + e.what());
    } catch (const Unsupported& e) {
        set_err(errbuf, errlen, std::string("unsupported: ") + e.what());
    } catch (const EngineError& e) {
        set_err(errbuf, errlen, std::string("engine: ") + e.what());
    } catch (const std::exception& e) {
        set_err(errbuf, errlen, e.what());
    }
    return 1;
}
