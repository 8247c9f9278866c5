"""Oracle platform emulation (oracle/javaemu.py) against the known answers of SURVEY App. C / E."""
import math
import random

from oracle.javaemu import (JavaHashMap, JavaRandom, double_to_string, fdlibm_log, hashmap_key_order, string_hash)


def test_string_hash_known_values():
    assert string_hash("") == 0
    assert string_hash("a") == 97
    assert string_hash("hello") == 99162322
    assert string_hash("T") == 84 and string_hash("F") == 70
    # int32 wrap-around
    assert string_hash("The quick brown fox jumps over the lazy dog") == -609428141


def test_hashmap_orders_appendix_c7():
    assert hashmap_key_order(["F", "T"]) == ["T", "F"]          # resource/dec_tree_rules.json classValPr order
    assert hashmap_key_order(["N", "Y"]) == ["Y", "N"]
    assert hashmap_key_order(["3", "1", "0", "2"]) == ["0", "1", "2", "3"]
    ids = [str(i) for i in range(1, 19)]
    assert hashmap_key_order(ids) == [str(i) for i in range(11, 19)] + [str(i) for i in range(1, 10)] + ["10"]


def test_hashmap_collision_keeps_insertion_order_and_resizes():
    # "100" and "221" collide modulo 1024 (hash difference 1024)
    assert (string_hash("221") - string_hash("100")) == 1024
    m = JavaHashMap()
    for i in range(1, 417):
        m.put(str(i), i)
    keys = list(m.keys())
    assert len(keys) == 416 and keys.index("100") + 1 == keys.index("221")
    m2 = JavaHashMap()
    for i in range(416, 0, -1):
        m2.put(str(i), i)
    k2 = list(m2.keys())
    assert k2.index("221") + 1 == k2.index("100")
    assert m.get("416") == 416 and m.get("417") is None and len(m) == 416


def test_java_random_stream():
    r = JavaRandom(42)
    assert r.next_double() == 0.7275636800328681     # new java.util.Random(42).nextDouble()
    r0 = JavaRandom(0)
    assert r0.next_double() == 0.730967787376657


def test_double_to_string():
    cases = {0.39594381212292684: "0.39594381212292684", 1.0: "1.0", 100.0: "100.0", 1e7: "1.0E7", 1e-4: "1.0E-4",
             0.001: "0.001", 123456.789: "123456.789", 9999999.0: "9999999.0", 1.9999999999999998: "1.9999999999999998",
             float("inf"): "Infinity", float("-inf"): "-Infinity", 0.0: "0.0", -2.5: "-2.5", 1.0e-5: "1.0E-5",
             12345678.9: "1.23456789E7"}
    for v, s in cases.items():
        assert double_to_string(v) == s
    assert double_to_string(float("nan")) == "NaN"
    assert double_to_string(-0.0) == "-0.0"


def test_fdlibm_log_within_one_ulp_of_libm():
    rnd = random.Random(7)
    assert fdlibm_log(1.0) == 0.0
    assert fdlibm_log(2.0) == 0.6931471805599453
    assert fdlibm_log(0.0) == float("-inf")
    assert math.isnan(fdlibm_log(-1.0))
    for _ in range(20000):
        x = rnd.random() * 10 ** rnd.randint(-8, 8)
        a, b = fdlibm_log(x), math.log(x)
        assert abs(a - b) <= abs(math.ulp(b))
    # probabilities count/total as the tree path produces them
    for n in (2, 3, 7, 331, 550, 881, 2119, 3000):
        for c in range(1, n):
            pr = c / n
            assert abs(fdlibm_log(pr) - math.log(pr)) <= abs(math.ulp(math.log(pr)))
