// Tool-level host layer: the reference's job entry points, re-hosted above the kernel-level C-ABI.
#pragma once
#include <string>
#include <vector>

#include "model.hpp"

namespace avt_host {

// org.avenir.tree.DecisionTreeBuilder.run(args)  (tree/DecisionTreeBuilder.java:70-95)
int run_decision_tree_builder(const Props& conf, const std::string& in_path, const std::string& out_path);
// org.avenir.explore.ClassPartitionGenerator.run(args)  (explore/ClassPartitionGenerator.java:64-91)
int run_class_partition_generator(const Props& conf, const std::string& in_path, const std::string& out_path);

// org.avenir.model.ModelPredictor.run(args)  (model/ModelPredictor.java:52-68)
int run_model_predictor(const Props& conf, const std::string& in_path, const std::string& out_path);

// org.avenir.explore.CramerCorrelation.run(args)  (explore/CramerCorrelation.java:56-81)
int run_cramer_correlation(const Props& conf, const std::string& in_path, const std::string& out_path);

// org.avenir.explore.HeterogeneityReductionCorrelation.run(args)  (explore/HeterogeneityReductionCorrelation.java:40-65)
int run_heterogeneity_reduction_correlation(const Props& conf, const std::string& in_path, const std::string& out_path);

// org.avenir.explore.CategoricalContinuousEncoding.run(args)  (explore/CategoricalContinuousEncoding.java:54-82)
int run_categorical_continuous_encoding(const Props& conf, const std::string& in_path, const std::string& out_path);

// org.avenir.explore.CategoricalClassAffinity.run(args)  (explore/CategoricalClassAffinity.java:58-89); host code only
int run_categorical_class_affinity(const Props& conf, const std::string& in_path, const std::string& out_path);

// org.avenir.tree.DataPartitioner.run(args)  (tree/DataPartitioner.java:59-105); takes no path arguments
int run_data_partitioner(const Props& conf);

// Input "path" as Hadoop takes it: a file, or a directory whose regular files (not starting with
// '_' or '.') are read in name order.
std::vector<std::string> read_input_lines(const std::string& path);
void write_output_lines(const std::string& out_dir, const std::vector<std::string>& lines, int part = 0);

// RAII wrapper over the kernel-level C-ABI; every failure becomes an exception carrying
// avt_last_error().
struct EngineError : std::runtime_error {
    int code;
    EngineError(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};

class Engine {
    avt_engine* e_ = nullptr;

public:
    explicit Engine(int device);
    ~Engine();
    Engine(const Engine&) = delete;
    Engine& operator=(const Engine&) = delete;
    avt_engine* raw() { return e_; }
    void check(int rc) const;
};

// class strings of a data set: code = first-seen index; order = HashMap<String,...> iteration order
// of the class strings (util/InfoContentStat.java:31,74,82,91).
struct ClassTable {
    std::vector<std::string> names;
    std::vector<uint8_t> order;  // iteration order as class codes
    int code_of(const std::string& s);
    void finish();
};

}  // namespace avt_host
