"""avenir_b200: B200-native split-evaluation engine behind avenir's tree-building path.

Only what the hot path needs lives here: csrc/ (sm_100a kernels, the C-ABI of include/avenir_tree.h and
the C++ tool layer), a ctypes binding, the reference-named entry points, and synthetic-data helpers.
"""
from .tools import CategoricalClassAffinity, CategoricalContinuousEncoding, ClassPartitionGenerator, CramerCorrelation, DataPartitioner, DecisionTreeBuilder, HeterogeneityReductionCorrelation, ModelPredictor, SplitGenerator, ToolFailure  # noqa: F401

__all__ = ["DecisionTreeBuilder", "ClassPartitionGenerator", "ModelPredictor", "CramerCorrelation", "HeterogeneityReductionCorrelation", "SplitGenerator", "DataPartitioner", "CategoricalContinuousEncoding", "CategoricalClassAffinity", "ToolFailure"]
