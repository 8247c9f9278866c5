import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from avenir_b200 import synth
from avenir_b200._lib import CRIT_ENTROPY
from avenir_b200.engine import Engine, binary_split_table
n = int(os.environ.get("ROWS", 300000))
cols, cls, n_bins, _ = synth.call_hangup_encoded(n)
eng = Engine(0)
eng.load_dataset(cols, cls, n_bins, 2)
eng.set_splits(binary_split_table(n_bins, [True, True, True, False]))
eng.forest_init(1)
eng.set_profiling(False)
print(eng.grow(8, CRIT_ENTROPY, [1])[1])
