"""Bootstrap multiplicities (explore/BaggingSampler.java:101-122): the sequential C
restatement against the Python java.util.Random model, and the device generator (LCG skip-ahead, one block per
batch) against the C restatement -- bit-exact, incl. a tail batch, shards that cut batches, and withoutReplace."""
import ctypes as C

import numpy as np
import pytest

from oracle.javaemu import JavaRandom
from oracle.literal import java_int_cast
from tests import helpers


def python_bagging(n, batch, seed):
    rng = JavaRandom(seed)
    w = np.zeros(n, dtype=np.int64)
    for s in range(0, n, batch):
        m = min(batch, n - s)
        for _ in range(m):
            w[s + java_int_cast(rng.next_double() * m)] += 1
    return w


oracle_bagging = helpers.oracle_bagging


def test_c_restatement_equals_java_random_model():
    # 3 full batches + a tail batch
    n, batch = 3 * 1000 + 137, 1000
    for seed in (0, 100, -5, 2 ** 40 + 17):
        want = python_bagging(n, batch, seed)
        got = oracle_bagging(seed, batch, n, 0, n)
        assert want.sum() == n and np.array_equal(got, want.astype(np.uint8))


def test_c_restatement_shards_add_up():
    n, batch = 25_431, 10_000
    full = oracle_bagging(7, batch, n, 0, n)
    cuts = [0, 3_333, 10_000, 19_999, n]          # cuts inside batches and on a batch boundary
    parts = [oracle_bagging(7, batch, n, a, b - a) for a, b in zip(cuts, cuts[1:])]
    assert np.array_equal(np.concatenate(parts), full)


@pytest.mark.gpu
@pytest.mark.parametrize("n,batch,row0,total", [(35_017, 10_000, 0, 35_017), (20_000, 10_000, 13_333, 60_001), (5_003, 700, 2_100, 9_999),
                                                 (130_000, 60_000, 0, 130_000)])
def test_device_generator_equals_c_restatement(n, batch, row0, total):
    from avenir_b200 import _lib
    from avenir_b200.engine import Engine
    rng = np.random.default_rng(1)
    cols = [rng.integers(0, 4, n).astype(np.uint8)]
    cls = rng.integers(0, 2, n).astype(np.uint8)
    g = Engine(0)
    g.load_dataset(cols, cls, [4], 2)
    T, seed = 3, 100
    g.forest_init_sampled(T, _lib.SAMPLE_WITH_REPLACE, seed, batch, total, row0)
    for t in range(T):
        assert np.array_equal(g.weights(t), oracle_bagging(seed + t, batch, total, row0, n)), t
        assert np.array_equal(g.node_ids(t), np.zeros(n, dtype=np.uint16))
    # the sampled forest counts like a forest started from the same multiplicities on the host
    h1 = g.histogram([1] * T, fetch=True)
    g.forest_init(T, [oracle_bagging(seed + t, batch, total, row0, n) for t in range(T)], None)
    assert np.array_equal(g.histogram([1] * T, fetch=True), h1)
    g.close()


@pytest.mark.gpu
def test_device_bernoulli_equals_java_random_model():
    from avenir_b200 import _lib
    from avenir_b200.engine import Engine
    n, row0, total, rate = 10_007, 5_000, 20_000, 63
    rng = np.random.default_rng(2)
    g = Engine(0)
    g.load_dataset([rng.integers(0, 4, n).astype(np.uint8)], rng.integers(0, 2, n).astype(np.uint8), [4], 2)
    g.forest_init_sampled(2, _lib.SAMPLE_WITHOUT_REPLACE, 11, rate, total, row0)
    for t in range(2):
        jr = JavaRandom(11 + t)
        keep = np.array([1 if java_int_cast(jr.next_double() * 100) < rate else 0 for _ in range(row0 + n)], dtype=np.uint8)[row0:]
        assert np.array_equal(g.weights(t), keep)
    g.close()
