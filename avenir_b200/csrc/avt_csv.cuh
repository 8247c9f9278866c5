// On-device encode: CSV text -> columnar bin codes + class codes (the step before the path).
// Restates, per row, what BuilderMapper.pathMapHelper does before it can evaluate a predicate
// (tree/DecisionTreeBuilder.java:268 `value.toString().split(fieldDelimRegex)`, :309-320 Integer.parseInt /
// Double.parseDouble per attribute) and the predicate grid itself in the histogram form used by the engine
// (bin = number of grid points < value, tree/SplitManager.java:256-330; categorical bin = index in the value list),
// plus the class extraction of BuilderReducer.reduce (:749).  Three kernels:
//   csv_count_kernel   newlines per 128-byte chunk          (then an exclusive prefix sum over the chunks)
//   csv_lines_kernel   start offset of every row
//   csv_encode_kernel  one thread per row: fields -> codes, first offending row, values for the host's strtod
// Included by avt_kernels.cuh.
#pragma once

namespace avt {

constexpr int kCsvChunk = 128;          // bytes per thread of the line-indexing kernels
constexpr int kCsvMaxFields = 1024;     // highest field index + 1 that may be referenced
constexpr int kCsvNoClassField = -(1 << 20);   // CsvArgs::class_field of rows without a class field: no field index reaches it
constexpr int kCsvSlowCap = 65536;      // values left to the host's strtod per call
constexpr int kCsvStageBytes = 24 * 1024;   // text span of one block of rows staged in shared memory (longer spans: read in place)

struct CsvSlow {                        // one value the device parser could not convert exactly
    long long row;
    long long start;                    // byte offset of the field in the text
    int32_t len, attr;
};

struct CsvArgs {
    const char* text;
    long long n_bytes;
    const long long* line_start;        // [n_rows + 1]; row r = [line_start[r], line_start[r+1] - 1)
    long long n_rows;
    char delim;
    int skip_fields;
    int n_attrs;
    const avt_csv_attr* attrs;          // [n_attrs]
    const int16_t* field_attr;          // [n_fields_used] field index (after skip) -> attribute, -1 none
    int n_fields_used;
    const int32_t* int_points;
    const double* dbl_points;
    const char* cat_bytes;
    const int32_t* cat_offsets;
    const int32_t* cat_sorted;          // per categorical attribute: its value indices sorted by bytes
    int class_field;                    // field index (after skip) of the class
    int n_class_values;
    const char* class_bytes;
    const int32_t* class_offsets;
    const int32_t* class_sorted;
    void* const* cols;                  // [n_attrs] uint8 / uint16 columns
    const int32_t* wide;                // [n_attrs]
    uint8_t* cls;
    long long* class_first_row;         // [n_class_values], initialised to LLONG_MAX
    unsigned long long* bad;            // min over offending rows of (row << 8 | reason), initialised to ~0
    CsvSlow* slow;                      // [kCsvSlowCap]
    uint32_t* n_slow;
};

__global__ void __launch_bounds__(256) csv_count_kernel(const char* text, long long n_bytes, uint32_t* counts, long long n_chunks) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_chunks) return;
    const long long b0 = i * kCsvChunk;
    uint32_t c = 0;
    if (b0 + kCsvChunk <= n_bytes) {
        const uint4* p = reinterpret_cast<const uint4*>(text + b0);   // the text buffer is 256-byte aligned
#pragma unroll
        for (int k = 0; k < kCsvChunk / 16; ++k) {
            const uint4 v = p[k];
            const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                // bytes equal to '\n': zero bytes of w ^ 0x0A0A0A0A, counted exactly (no borrow tricks)
                const uint32_t x = w[j] ^ 0x0A0A0A0Au;
                const uint32_t z = ~(((x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | x | 0x7F7F7F7Fu);   // 0x80 where the byte is zero
                c += __popc(z);
            }
        }
    } else {
        for (long long b = b0; b < n_bytes; ++b) c += text[b] == '\n';
    }
    counts[i] = c;
}

__global__ void __launch_bounds__(256) csv_lines_kernel(const char* text, long long n_bytes, const unsigned long long* chunk_first,
                                                        long long* line_start, long long n_chunks) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_chunks) return;
    const long long b0 = i * kCsvChunk, b1 = min(n_bytes, b0 + kCsvChunk);
    unsigned long long k = chunk_first[i];   // newlines before this chunk = index of the row that ends next
    for (long long b = b0; b < b1; ++b)
        if (text[b] == '\n') line_start[++k] = b + 1;
}

// index of the value equal to [p, p+len) in a value list given as offsets + a byte-sorted permutation; -1 = absent
__device__ __forceinline__ int csv_find_value(const char* bytes, const int32_t* offsets, const int32_t* sorted, int n, const char* p, int len) {
    int lo = 0, hi = n - 1;
    while (lo <= hi) {
        const int mid = (lo + hi) >> 1;
        const int v = sorted[mid];
        const char* q = bytes + offsets[v];
        const int ql = offsets[v + 1] - offsets[v];
        int cmp = 0;
        const int m = len < ql ? len : ql;
        for (int i = 0; i < m && cmp == 0; ++i) cmp = (int)(unsigned char)p[i] - (int)(unsigned char)q[i];
        if (cmp == 0) cmp = len - ql;
        if (cmp == 0) return v;
        if (cmp < 0) hi = mid - 1;
        else lo = mid + 1;
    }
    return -1;
}

// Integer.parseInt: optional sign, decimal digits only, int32 range
__device__ __forceinline__ bool csv_parse_int(const char* p, int len, long long& out) {
    if (len <= 0) return false;
    int i = 0;
    bool neg = false;
    if (p[0] == '-' || p[0] == '+') { neg = p[0] == '-'; i = 1; }
    if (i >= len) return false;
    long long v = 0;
    for (; i < len; ++i) {
        const int d = p[i] - '0';
        if (d < 0 || d > 9) return false;
        v = v * 10 + d;
        if (v > 2147483648LL) return false;
    }
    v = neg ? -v : v;
    if (v > 2147483647LL || v < -2147483648LL) return false;
    out = v;
    return true;
}

// Decimal -> double, only where one IEEE operation gives the correctly rounded result (<= 15 significant digits,
// |decimal exponent| <= 22: integer mantissa < 2^53 times or divided by an exactly representable power of ten).
// Returns false for everything else, valid or not: the host decides with strtod.
__device__ __forceinline__ bool csv_parse_double_exact(const char* p, int len, double& out) {
    const double p10[23] = {1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11, 1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
    while (len > 0 && (unsigned char)p[0] <= ' ') { ++p; --len; }            // String.trim()
    while (len > 0 && (unsigned char)p[len - 1] <= ' ') --len;
    if (len <= 0) return false;
    int i = 0;
    bool neg = false;
    if (p[0] == '-' || p[0] == '+') { neg = p[0] == '-'; i = 1; }
    unsigned long long m = 0;
    int nd = 0, n_int = 0, n_frac = 0, frac_exp = 0;
    for (; i < len && p[i] >= '0' && p[i] <= '9'; ++i, ++n_int) {
        if (m || p[i] != '0') { if (++nd > 15) return false; m = m * 10 + (unsigned)(p[i] - '0'); }
    }
    if (i < len && p[i] == '.') {
        for (++i; i < len && p[i] >= '0' && p[i] <= '9'; ++i, ++n_frac) {
            if (m || p[i] != '0') { if (++nd > 15) return false; m = m * 10 + (unsigned)(p[i] - '0'); }
            ++frac_exp;
        }
    }
    if (n_int + n_frac == 0) return false;
    int e = 0;
    if (i < len && (p[i] == 'e' || p[i] == 'E')) {
        ++i;
        bool eneg = false;
        if (i < len && (p[i] == '-' || p[i] == '+')) { eneg = p[i] == '-'; ++i; }
        if (i >= len) return false;
        for (; i < len && p[i] >= '0' && p[i] <= '9'; ++i) {
            e = e * 10 + (p[i] - '0');
            if (e > 10000) return false;
        }
        e = eneg ? -e : e;
    }
    if (i != len) return false;   // trailing characters (type suffixes, hex, NaN, Infinity ...): host
    const int e10 = e - frac_exp;
    double v;
    if (m == 0) v = 0.0;
    else if (e10 >= 0 && e10 <= 22) v = __dmul_rn((double)m, p10[e10]);
    else if (e10 < 0 && e10 >= -22) v = __ddiv_rn((double)m, p10[-e10]);
    else return false;
    out = neg ? -v : v;
    return true;
}

// One thread per row.  Fields are visited left to right, which is also attribute order (ordinals increase), so the
// first failure met is the one the reference's loop over the attributes would throw on; Java's String.split drops
// trailing empty fields, so a row is as long as its last non-empty field.
__global__ void __launch_bounds__(128) csv_encode_kernel(const CsvArgs a) {
    // the block's rows are one contiguous span of the text: stage it in shared memory with coalesced 16-byte loads
    // (a thread walking its own row straight from global memory touches a new sector every few bytes)
    __shared__ __align__(16) char stage[kCsvStageBytes + 16];
    const long long r0 = (long long)blockIdx.x * blockDim.x;
    const long long r1 = min(a.n_rows, r0 + blockDim.x);
    const long long span0 = a.line_start[r0], span1 = min(a.n_bytes, a.line_start[r1]);
    const long long a0 = span0 & ~15LL;                    // the text buffer is 16-byte aligned and padded by 16 bytes
    const bool staged = span1 - a0 <= kCsvStageBytes;
    if (staged) {
        const int n16 = (int)((span1 - a0 + 15) >> 4);
        const uint4* src = reinterpret_cast<const uint4*>(a.text + a0);
        uint4* dst = reinterpret_cast<uint4*>(stage);
        for (int i = threadIdx.x; i < n16; i += blockDim.x) dst[i] = src[i];
    }
    __syncthreads();
    const long long r = r0 + threadIdx.x;
    if (r >= a.n_rows) return;
    const long long row_at = a.line_start[r];
    const char* row = staged ? stage + (row_at - a0) : a.text + row_at;
    long long len_ll = a.line_start[r + 1] - 1 - row_at;
    int len = (int)len_ll;
    if (len > 0 && row[len - 1] == '\r') --len;
    int field = -a.skip_fields;     // index after the skipped fields
    int n_eff = 0;                  // (index of the last non-empty field + 1), in the same numbering
    int fail_attr = -1, fail_reason = 0;
    int class_state = 0;            // 0 not seen, 1 ok, 2 unknown value
    if (a.class_field == kCsvNoClassField) {   // no class field: one class, code 0
        a.cls[r] = 0;
        class_state = 1;
    }
    int s = 0;
    for (int i = 0; i <= len; ++i) {
        if (i < len && row[i] != a.delim) continue;
        const char* p = row + s;
        const int fl = i - s;
        if (fl > 0) n_eff = field + 1;
        if (field >= 0 && field < a.n_fields_used) {
            const int at = a.field_attr[field];
            if (at >= 0) {
                const avt_csv_attr A = a.attrs[at];
                int bin = -1, reason = 0;
                if (A.kind == AVT_CSV_INT) {
                    long long v;
                    if (csv_parse_int(p, fl, v)) {
                        int lo = 0, hi = A.n;   // number of points < v
                        const int32_t* pts = a.int_points + A.off;
                        while (lo < hi) {
                            const int mid = (lo + hi) >> 1;
                            if ((long long)pts[mid] < v) lo = mid + 1;
                            else hi = mid;
                        }
                        bin = lo;
                    } else {
                        reason = AVT_CSV_BAD_INT;
                    }
                } else if (A.kind == AVT_CSV_DOUBLE) {
                    double v;
                    if (csv_parse_double_exact(p, fl, v)) {
                        int lo = 0, hi = A.n;
                        const double* pts = a.dbl_points + A.off;
                        while (lo < hi) {
                            const int mid = (lo + hi) >> 1;
                            if (pts[mid] < v) lo = mid + 1;
                            else hi = mid;
                        }
                        bin = lo;
                    } else {
                        // valid or not is the host's call (strtod); bin 0 until it patches the value in
                        const uint32_t k = atomicAdd(a.n_slow, 1u);
                        if (k < (uint32_t)kCsvSlowCap) a.slow[k] = CsvSlow{r, row_at + (long long)(p - row), fl, at};
                        bin = 0;
                    }
                } else {
                    bin = csv_find_value(a.cat_bytes, a.cat_offsets + A.off, a.cat_sorted + A.off, A.n, p, fl);
                    if (bin < 0) reason = AVT_CSV_BAD_CATEGORY;
                }
                if (reason) {
                    if (fail_attr < 0) { fail_attr = at; fail_reason = reason; }
                } else if (a.wide[at]) {
                    ((uint16_t*)a.cols[at])[r] = (uint16_t)bin;
                } else {
                    ((uint8_t*)a.cols[at])[r] = (uint8_t)bin;
                }
            }
        }
        if (field == a.class_field) {
            const int c = csv_find_value(a.class_bytes, a.class_offsets, a.class_sorted, a.n_class_values, p, fl);
            if (c >= 0) {
                a.cls[r] = (uint8_t)c;
                // first row of the class value: the stored minimum only ever decreases, so a plain look first spares almost
                // every row the atomic (millions of them on two or three addresses serialise in L2: 1 ms per 2 M rows)
                if (r < *reinterpret_cast<volatile long long*>(a.class_first_row + c))
                    atomicMin(reinterpret_cast<unsigned long long*>(a.class_first_row + c), (unsigned long long)r);
                class_state = 1;
            } else {
                class_state = 2;
            }
        }
        ++field;
        s = i + 1;
    }
    if (len == 0) n_eff = 1 - a.skip_fields;   // "".split(...) is [""]: one (empty) field
    // the reference's order of checks: attributes in order (length first, then the parse), then the class
    int reason = 0;
    int first_short = a.n_attrs;
    for (int at = 0; at < a.n_attrs; ++at)
        if (a.attrs[at].ordinal >= n_eff) { first_short = at; break; }
    if (fail_attr >= 0 && fail_attr < first_short) reason = fail_reason;
    else if (first_short < a.n_attrs) reason = AVT_CSV_BAD_SHORT_ROW;
    else if (a.class_field >= n_eff) reason = AVT_CSV_BAD_SHORT_ROW;
    else if (class_state != 1) reason = AVT_CSV_BAD_CLASS;
    if (reason) atomicMin(a.bad, ((unsigned long long)r << 8) | (unsigned long long)reason);
}

// codes computed on the host for the values the device parser left out
struct CsvPatch {
    long long row;
    int32_t attr, bin;
};
__global__ void csv_patch_kernel(const CsvPatch* patches, int n, void* const* cols, const int32_t* wide) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const CsvPatch p = patches[i];
    if (wide[p.attr]) ((uint16_t*)cols[p.attr])[p.row] = (uint16_t)p.bin;
    else ((uint8_t*)cols[p.attr])[p.row] = (uint8_t)p.bin;
}

// ---- rows of the encoded text, physically partitioned (avt_rows_partition; tree/DataPartitioner.java:325-352) ------------------
// A row's segment is a table look-up on one attribute's bin code; the segment's output is its rows in input order, each
// followed by '\n' (a '\r' in front of the source's line terminator is dropped, as LineRecordReader does).  Three small
// kernels per call + one exclusive prefix sum and one copy kernel per segment; all of it streams at HBM rate.
__global__ void rows_len_kernel(const char* text, const long long* line_start, long long n_rows, int32_t* len) {
    const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_rows) return;
    const long long s = line_start[r];
    long long l = line_start[r + 1] - 1 - s;
    if (l > 0 && text[s + l - 1] == '\r') --l;
    len[r] = (int32_t)l + 1;   // content + the '\n' written behind it
}

template <typename CodeT>
__global__ void rows_mask_kernel(const CodeT* codes, const uint16_t* seg_of_bin, const int32_t* len, long long n_rows, int seg, long long* masked,
                                 unsigned long long* seg_rows) {
    const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const bool mine = r < n_rows && seg_of_bin[codes[r]] == (uint16_t)seg;
    if (r < n_rows) masked[r] = mine ? (long long)len[r] : 0LL;
    const int c = __syncthreads_count(mine);
    if (threadIdx.x == 0 && c) atomicAdd(seg_rows + seg, (unsigned long long)c);
}

// base[seg + 1] = base[seg] + bytes of segment `seg`
template <typename CodeT>
__global__ void rows_base_kernel(const CodeT* codes, const uint16_t* seg_of_bin, const int32_t* len, long long n_rows, int seg, const long long* off, long long* base) {
    const long long last = n_rows - 1;
    base[seg + 1] = base[seg] + off[last] + (seg_of_bin[codes[last]] == (uint16_t)seg ? (long long)len[last] : 0LL);
}

// one warp per row: coalesced byte copy of the row to its place in the segment
template <typename CodeT>
__global__ void __launch_bounds__(256) rows_copy_kernel(const char* text, const long long* line_start, const CodeT* codes, const uint16_t* seg_of_bin,
                                                        const int32_t* len, long long n_rows, int seg, const long long* off, const long long* base, char* out) {
    const long long r = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (r >= n_rows || seg_of_bin[codes[r]] != (uint16_t)seg) return;
    const char* src = text + line_start[r];
    char* dst = out + base[seg] + off[r];
    const int l = len[r] - 1;
    for (int i = lane; i < l; i += 32) dst[i] = src[i];
    if (lane == 0) dst[l] = '\n';
}

}  // namespace avt
