"""
ORACLE / TEST INFRASTRUCTURE ONLY -- never imported by the product (avenir_b200/).

Java platform behaviours the reference's results depend on, restated in
pure Python so that the literal restatement in oracle/literal.py can reproduce iteration orders,
random streams, number formatting and log() the way a JDK-8 JVM would.

The reference (kutieme/avenir) is Java 1.8 (pom.xml:23); no JVM exists in this environment, so
nothing here could be checked against a running JVM: PARITY UNPINNED for these platform pieces,
except where a golden vector exists (resource/dec_tree_rules.json key order "T" before "F",
SURVEY App. C item 7).

Restated behaviours
  * String.hashCode                (JLS: s[0]*31^(n-1) + ... + s[n-1], int32 wrap)
  * java.util.HashMap (JDK 8)      iteration order: table index = (h ^ h>>>16) & (cap-1),
                                   chains in insertion order, resize keeps relative order,
                                   resize at ++size > 0.75*cap, chain of 9 with cap < 64 resizes
  * java.util.Random               48-bit LCG, nextDouble() = ((next(26)<<27)+next(27)) * 2^-53
  * Double.toString                shortest round-trip digits, Java layout
  * StrictMath.log                 fdlibm __ieee754_log (the algorithm published in e_log.c)
"""
from __future__ import annotations

import struct

_M32 = 0xFFFFFFFF


def string_hash(s: str) -> int:
    """java.lang.String.hashCode() as a signed int32 (used by InfoContentStat's and
    DecisionTreeBuilder's HashMap<String,...>: util/InfoContentStat.java:31,
    tree/DecisionTreeBuilder.java:394-396)."""
    b = s.encode("utf-16-be")  # iterate UTF-16 code units
    h = 0
    for i in range(0, len(b), 2):
        cu = (b[i] << 8) | b[i + 1]
        h = (31 * h + cu) & _M32
    return h - (1 << 32) if h & 0x80000000 else h


def integer_hash(v: int) -> int:
    """java.lang.Integer.hashCode() == value."""
    v &= _M32
    return v - (1 << 32) if v & 0x80000000 else v


def _spread(h: int) -> int:
    """HashMap.hash(): h ^ (h >>> 16) on the int32 bit pattern."""
    h &= _M32
    return (h ^ (h >> 16)) & _M32


class TreeifiedBin(Exception):
    """A chain reached 9 entries with capacity >= 64: JDK 8 would turn the bin into a red-black
    tree whose iteration order this emulation does not restate."""


class JavaHashMap:
    """Insertion/iteration-order model of java.util.HashMap (JDK 8, default constructor).

    Keys are Python str (-> String.hashCode) or int (-> Integer.hashCode).  Values arbitrary.
    Only what the reference uses: get / put / containsKey / keySet() iteration / size.
    """

    __slots__ = ("_tab", "_cap", "_size", "_thr")

    def __init__(self):
        self._tab = None
        self._cap = 0
        self._size = 0
        self._thr = 0

    @staticmethod
    def _hash_of(key) -> int:
        if isinstance(key, str):
            return _spread(string_hash(key))
        if isinstance(key, bool):
            raise TypeError("bool key")
        if isinstance(key, int):
            return _spread(integer_hash(key))
        raise TypeError(f"unsupported key type {type(key)}")

    def _resize(self):
        if self._tab is None:
            self._cap = 16
            self._thr = 12
            self._tab = [None] * 16
            return
        old = self._tab
        ncap = self._cap * 2
        new = [None] * ncap
        for j, chain in enumerate(old):
            if not chain:
                continue
            lo = [e for e in chain if (e[0] & self._cap) == 0]
            hi = [e for e in chain if (e[0] & self._cap) != 0]
            if lo:
                new[j] = lo
            if hi:
                new[j + self._cap] = hi
        self._tab = new
        self._cap = ncap
        self._thr = ncap * 3 // 4

    def _find(self, key):
        if self._tab is None:
            return None
        h = self._hash_of(key)
        chain = self._tab[h & (self._cap - 1)]
        if chain:
            for e in chain:
                if e[0] == h and e[1] == key:
                    return e
        return None

    def get(self, key, default=None):
        e = self._find(key)
        return default if e is None else e[2]

    def __contains__(self, key):
        return self._find(key) is not None

    def __getitem__(self, key):
        e = self._find(key)
        if e is None:
            raise KeyError(key)
        return e[2]

    def put(self, key, value):
        if self._tab is None:
            self._resize()
        h = self._hash_of(key)
        i = h & (self._cap - 1)
        chain = self._tab[i]
        if chain is None:
            self._tab[i] = [[h, key, value]]
        else:
            for e in chain:
                if e[0] == h and e[1] == key:
                    e[2] = value
                    return
            chain.append([h, key, value])
            if len(chain) >= 9:  # binCount >= TREEIFY_THRESHOLD-1 -> treeifyBin()
                if self._cap < 64:
                    self._resize()
                else:
                    raise TreeifiedBin(f"bin {i} treeified at capacity {self._cap}")
        self._size += 1
        if self._size > self._thr:
            self._resize()

    __setitem__ = put

    def clear(self):
        """HashMap.clear() (JDK 8): every bucket emptied, the table keeps its length."""
        if self._tab is not None:
            self._tab = [None] * self._cap
        self._size = 0

    def __len__(self):
        return self._size

    def keys(self):
        """keySet() iteration order."""
        if self._tab is None:
            return
        for chain in self._tab:
            if chain:
                for e in chain:
                    yield e[1]

    __iter__ = keys

    def items(self):
        if self._tab is None:
            return
        for chain in self._tab:
            if chain:
                for e in chain:
                    yield e[1], e[2]

    def values(self):
        for _, v in self.items():
            yield v


def hashmap_key_order(keys):
    """Iteration order of a HashMap after put()-ing `keys` in the given order."""
    m = JavaHashMap()
    for k in keys:
        m.put(k, None)
    return list(m.keys())


class JavaRandom:
    """java.util.Random (the generator behind Math.random(), which the reference uses unseeded at
    tree/DecisionTreeBuilder.java:200,221,232 and explore/BaggingSampler.java:119).  A seed is our
    addition: the reference's stream is nanoTime-seeded and unrepeatable."""

    _MULT = 0x5DEECE66D
    _MASK = (1 << 48) - 1

    def __init__(self, seed: int):
        self.seed = (seed ^ self._MULT) & self._MASK

    def next(self, bits: int) -> int:
        self.seed = (self.seed * self._MULT + 0xB) & self._MASK
        r = self.seed >> (48 - bits)
        return r  # callers only use it unsigned (bits <= 27)

    def next_double(self) -> float:
        return ((self.next(26) << 27) + self.next(27)) * (1.0 / (1 << 53))


# ----------------------------------------------------------------------------------------------
# Double.toString
# ----------------------------------------------------------------------------------------------
def double_to_string(x: float) -> str:
    """java.lang.Double.toString(double) (JDK >= 19 shortest-digits behaviour).
    Used by Jackson for JSON numbers, by DoublePredicate.toString (tree/SplitManager.java:880-889)
    and by string concatenation in explore/ClassPartitionGenerator.java:538,544."""
    if x != x:
        return "NaN"
    if x == float("inf"):
        return "Infinity"
    if x == float("-inf"):
        return "-Infinity"
    if x == 0.0:
        return "-0.0" if struct.pack(">d", x)[0] & 0x80 else "0.0"
    sign = "-" if x < 0 else ""
    a = abs(x)
    r = repr(a)  # Python repr = shortest round-trip digits
    # split into digits and decimal exponent
    if "e" in r or "E" in r:
        mant, exp = r.lower().split("e")
        exp = int(exp)
    else:
        mant, exp = r, 0
    if "." in mant:
        ip, fp = mant.split(".")
    else:
        ip, fp = mant, ""
    digits = (ip + fp).lstrip("0")
    # value = 0.d1d2... * 10^pointpos  where pointpos is relative to the stripped digits
    lead_zeros = len(ip + fp) - len((ip + fp).lstrip("0"))
    pointpos = len(ip) - lead_zeros + exp
    digits = digits.rstrip("0") or "0"
    # scientific exponent e: a = d1.d2d3... * 10^e
    e = pointpos - 1
    if 1e-3 <= a < 1e7:
        if pointpos <= 0:
            s = "0." + "0" * (-pointpos) + digits
        elif pointpos >= len(digits):
            s = digits + "0" * (pointpos - len(digits)) + ".0"
        else:
            s = digits[:pointpos] + "." + digits[pointpos:]
    else:
        frac = digits[1:] or "0"
        s = digits[0] + "." + frac + "E" + str(e)
    return sign + s


# ----------------------------------------------------------------------------------------------
# fdlibm log  (StrictMath.log)
# ----------------------------------------------------------------------------------------------
_LN2_HI = 6.93147180369123816490e-01
_LN2_LO = 1.90821492927058770002e-10
_TWO54 = 1.80143985094819840000e+16
_LG1 = 6.666666666666735130e-01
_LG2 = 3.999999999940941908e-01
_LG3 = 2.857142874366239149e-01
_LG4 = 2.222219843214978396e-01
_LG5 = 1.818357216161805012e-01
_LG6 = 1.531383769920937332e-01
_LG7 = 1.479819860511658591e-01


def _hi_lo(x: float):
    b = struct.unpack(">II", struct.pack(">d", x))
    return b[0], b[1]


def _from_hi_lo(hi: int, lo: int) -> float:
    return struct.unpack(">d", struct.pack(">II", hi & _M32, lo & _M32))[0]


def _s32(v: int) -> int:
    v &= _M32
    return v - (1 << 32) if v & 0x80000000 else v


def fdlibm_log(x: float) -> float:
    """__ieee754_log of fdlibm 5.3 (the algorithm StrictMath.log is specified to follow),
    evaluated with Python floats = IEEE binary64, one rounding per operation, no FMA.
    The reference calls Math.log (util/InfoContentStat.java:80,83), which the JDK may intrinsify
    with <= 1 ulp error; StrictMath semantics are the reproducible choice."""
    hx, lx = _hi_lo(x)
    hx = _s32(hx)
    k = 0
    if hx < 0x00100000:  # x < 2**-1022
        if ((hx & 0x7FFFFFFF) | lx) == 0:
            return float("-inf")
        if hx < 0:
            return float("nan")
        k -= 54
        x *= _TWO54
        hx, lx = _hi_lo(x)
        hx = _s32(hx)
    if hx >= 0x7FF00000:
        return x + x
    k += (hx >> 20) - 1023
    hx &= 0x000FFFFF
    i = (hx + 0x95F64) & 0x100000
    x = _from_hi_lo(hx | (i ^ 0x3FF00000), lx)  # normalize x or x/2
    k += i >> 20
    f = x - 1.0
    if (0x000FFFFF & (2 + hx)) < 3:  # |f| < 2**-20
        if f == 0.0:
            if k == 0:
                return 0.0
            dk = float(k)
            return dk * _LN2_HI + dk * _LN2_LO
        R = f * f * (0.5 - 0.33333333333333333 * f)
        if k == 0:
            return f - R
        dk = float(k)
        return dk * _LN2_HI - ((R - dk * _LN2_LO) - f)
    s = f / (2.0 + f)
    dk = float(k)
    z = s * s
    i = hx - 0x6147A
    w = z * z
    j = 0x6B851 - hx
    t1 = w * (_LG2 + w * (_LG4 + w * _LG6))
    t2 = z * (_LG1 + w * (_LG3 + w * (_LG5 + w * _LG7)))
    i |= j
    R = t2 + t1
    if i > 0:
        hfsq = 0.5 * f * f
        if k == 0:
            return f - (hfsq - s * (hfsq + R))
        return dk * _LN2_HI - ((hfsq - (s * (hfsq + R) + dk * _LN2_LO)) - f)
    if k == 0:
        return f - s * (f - R)
    return dk * _LN2_HI - ((s * (f - R) - dk * _LN2_LO) - f)


LOG2 = fdlibm_log(2.0)
