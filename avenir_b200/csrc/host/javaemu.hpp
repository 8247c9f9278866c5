// Java platform behaviours that decide the reference's results, for the C++ host
// layer: String.hashCode, JDK-8 HashMap iteration order, java.util.Random, Double.toString.
// Written independently of oracle/javaemu.py (the oracle is test infrastructure; the product never
// links or imports it).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace avt_host {

// java.lang.String.hashCode over the UTF-16 code units of a UTF-8 std::string.
inline int32_t java_string_hash(const std::string& s) {
    uint32_t h = 0;
    size_t i = 0, n = s.size();
    while (i < n) {
        uint32_t c = (unsigned char)s[i], cp;
        if (c < 0x80) { cp = c; i += 1; }
        else if ((c >> 5) == 6 && i + 1 < n) { cp = ((c & 0x1F) << 6) | (s[i + 1] & 0x3F); i += 2; }
        else if ((c >> 4) == 14 && i + 2 < n) { cp = ((c & 0x0F) << 12) | ((s[i + 1] & 0x3F) << 6) | (s[i + 2] & 0x3F); i += 3; }
        else if ((c >> 3) == 30 && i + 3 < n) {
            cp = ((c & 0x07) << 18) | ((s[i + 1] & 0x3F) << 12) | ((s[i + 2] & 0x3F) << 6) | (s[i + 3] & 0x3F);
            i += 4;
        } else { cp = c; i += 1; }
        if (cp >= 0x10000) {
            cp -= 0x10000;
            h = 31u * h + (0xD800u + (cp >> 10));
            h = 31u * h + (0xDC00u + (cp & 0x3FF));
        } else {
            h = 31u * h + cp;
        }
    }
    return (int32_t)h;
}

inline uint32_t java_spread(int32_t h) {
    uint32_t u = (uint32_t)h;
    return u ^ (u >> 16);
}

struct TreeifiedBin : std::runtime_error {
    TreeifiedBin() : std::runtime_error("HashMap bin treeified (9 colliding keys at capacity >= 64): iteration order not restated") {}
};

// Iteration-order model of java.util.HashMap (JDK 8, default constructor): table index =
// spread(hash) & (cap-1); chains keep insertion order; resize (at ++size > 0.75 cap, or when a chain
// reaches 9 entries below capacity 64) preserves relative order inside the lo/hi halves.
template <class K, class V, class Hasher>
class JavaHashMap {
    struct Entry {
        uint32_t h;
        K k;
        V v;
    };
    std::vector<std::vector<Entry>> tab_;
    size_t size_ = 0, thr_ = 0;

    void resize() {
        if (tab_.empty()) {
            tab_.resize(16);
            thr_ = 12;
            return;
        }
        const size_t oc = tab_.size();
        std::vector<std::vector<Entry>> nt(oc * 2);
        for (size_t j = 0; j < oc; ++j)
            for (auto& e : tab_[j]) nt[(e.h & oc) ? j + oc : j].push_back(std::move(e));
        tab_.swap(nt);
        thr_ = tab_.size() * 3 / 4;
    }

public:
    V* find(const K& k) {
        if (tab_.empty()) return nullptr;
        const uint32_t h = Hasher()(k);
        for (auto& e : tab_[h & (tab_.size() - 1)])
            if (e.h == h && e.k == k) return &e.v;
        return nullptr;
    }
    // returns reference to the (possibly new) value
    V& put(const K& k, V v) {
        if (tab_.empty()) resize();
        const uint32_t h = Hasher()(k);
        size_t idx = h & (tab_.size() - 1);
        for (auto& e : tab_[idx])
            if (e.h == h && e.k == k) {
                e.v = std::move(v);
                return e.v;
            }
        tab_[idx].push_back(Entry{h, k, std::move(v)});
        if (tab_[idx].size() >= 9) {
            if (tab_.size() < 64) resize();
            else throw TreeifiedBin();
        }
        if (++size_ > thr_) resize();
        return *find(k);
    }
    V& get_or_create(const K& k) {
        if (V* p = find(k)) return *p;
        return put(k, V());
    }
    size_t size() const { return size_; }
    // HashMap.clear() (JDK 8): every bucket emptied, the table keeps its length (later iteration orders depend on it)
    void clear() {
        for (auto& chain : tab_) chain.clear();
        size_ = 0;
    }
    template <class F>
    void for_each(F&& f) {
        for (auto& chain : tab_)
            for (auto& e : chain) f(e.k, e.v);
    }
    std::vector<K> keys() const {
        std::vector<K> out;
        for (auto& chain : tab_)
            for (auto& e : chain) out.push_back(e.k);
        return out;
    }
};

struct StrHasher {
    uint32_t operator()(const std::string& s) const { return java_spread(java_string_hash(s)); }
};
struct IntHasher {
    uint32_t operator()(int v) const { return java_spread((int32_t)v); }
};
template <class V>
using JStrMap = JavaHashMap<std::string, V, StrHasher>;

inline std::vector<std::string> hashmap_key_order(const std::vector<std::string>& insertion) {
    JStrMap<char> m;
    for (auto& k : insertion) m.put(k, 0);
    return m.keys();
}

// java.util.Random
class JavaRandom {
    uint64_t seed_;
    static constexpr uint64_t kMult = 0x5DEECE66DULL, kMask = (1ULL << 48) - 1;

public:
    explicit JavaRandom(int64_t seed) : seed_(((uint64_t)seed ^ kMult) & kMask) {}
    uint32_t next(int bits) {
        seed_ = (seed_ * kMult + 0xBULL) & kMask;
        return (uint32_t)(seed_ >> (48 - bits));
    }
    double next_double() { return (double)(((uint64_t)next(26) << 27) + next(27)) * (1.0 / 9007199254740992.0); }
};

// java.lang.Double.toString (shortest round-trip digits, Java layout).
inline std::string java_double_to_string(double x) {
    if (x != x) return "NaN";
    if (std::isinf(x)) return x > 0 ? "Infinity" : "-Infinity";
    if (x == 0.0) return std::signbit(x) ? "-0.0" : "0.0";
    char buf[64];
    int prec = 1;
    for (; prec <= 17; ++prec) {
        snprintf(buf, sizeof buf, "%.*e", prec - 1, x);
        if (strtod(buf, nullptr) == x) break;
    }
    // buf = [-]d.ddddde[+-]xx
    std::string s(buf);
    bool neg = s[0] == '-';
    if (neg) s.erase(0, 1);
    size_t epos = s.find('e');
    int exp10 = atoi(s.c_str() + epos + 1);
    std::string digits;
    for (size_t i = 0; i < epos; ++i)
        if (s[i] != '.') digits.push_back(s[i]);
    while (digits.size() > 1 && digits.back() == '0') digits.pop_back();
    std::string out;
    const double a = std::fabs(x);
    if (a >= 1e-3 && a < 1e7) {
        int pointpos = exp10 + 1;  // digits before the decimal point
        if (pointpos <= 0) {
            out = "0." + std::string((size_t)(-pointpos), '0') + digits;
        } else if ((size_t)pointpos >= digits.size()) {
            out = digits + std::string((size_t)pointpos - digits.size(), '0') + ".0";
        } else {
            out = digits.substr(0, (size_t)pointpos) + "." + digits.substr((size_t)pointpos);
        }
    } else {
        out = digits.substr(0, 1) + "." + (digits.size() > 1 ? digits.substr(1) : std::string("0")) + "E" + std::to_string(exp10);
    }
    return neg ? "-" + out : out;
}

}  // namespace avt_host
