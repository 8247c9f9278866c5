"""GPU parity on randomised shapes of the level loop: 60 cases of tools/fuzz_gpu.py (narrow / one chunk / several chunks /
uint16 columns / many bins; 1..6 trees with and without multiplicities; `all` and `notUsedYet`; 1..9 levels) against the
C oracle -- records incl. av bits, ties, child populations, and every row's final node."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("seed", [11, 12, 13])
def test_random_level_loops_equal_oracle(seed):
    from tools import fuzz_gpu
    rng = np.random.default_rng(seed)
    seen = set()
    for k in range(20):
        line = fuzz_gpu.one_case(rng, k)
        seen.add(line.split(" variant ")[1].split()[0])
    assert seen      # (which kernels ran depends on the drawn shapes; profiles/fuzz_r02.txt has the long run)
