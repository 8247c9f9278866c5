// libavenir_tree.so -- kernel-level C-ABI (include/avenir_tree.h) on top of the sm_100a kernels.
// Plain CUDA runtime + NCCL; no torch types anywhere.  There is no CPU path: every entry point needs
// a CUDA device and fails with AVT_E_CUDA when there is none.
#include <cuda_runtime.h>
#include <nccl.h>
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "avt_kernels.cuh"

using namespace avt;

namespace {

thread_local std::string g_create_error;

constexpr int kSmCountFallback = 148;
constexpr size_t kMaxDynSmem = 226 * 1024;        // B200: 227 KB of shared memory per CTA, 1 KB left for static shared variables
constexpr int kHistWordsBudget = 200 * 1024 / 4;  // shared-memory words of one block's histogram copies (the rest: repartition tables, skip marks)

struct TimerSlot {
    cudaEvent_t a, b;
    double* acc;
};

template <class T>
struct DevBuf {
    T* p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t n) {
        if (n <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        size_t want = n + n / 4 + 64;
        cudaError_t e = cudaMalloc((void**)&p, want * sizeof(T));
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};

}  // namespace

struct avt_engine {
    int device = 0;
    int sm_count = kSmCountFallback;
    cudaStream_t stream = nullptr;
    std::string err;
    // data set
    long long N = 0, Npad = 0;
    int A = 0, C = 0, TB = 0;
    std::vector<int> n_bins, bin_off, wide;
    std::vector<void*> cols;
    DevBuf<void*> d_colptrs;
    DevBuf<uint8_t> d_cls;
    DevBuf<int32_t> d_bin_off, d_wide;
    // forest
    int T = 0;
    bool any_w = false;
    DevBuf<uint16_t> d_nid, d_nid2;
    uint16_t* nid_cur = nullptr;  // buffer holding the current node ids (d_nid or d_nid2)
    DevBuf<uint8_t> d_w, d_has_w;
    std::vector<uint8_t> h_has_w;
    // frontier
    std::vector<int> node_base;
    int NF = 0;
    bool hist_valid = false, scan_valid = false;
    DevBuf<int32_t> d_node_base;
    DevBuf<uint32_t> d_H;
    // histogram group tables
    DevBuf<int32_t> d_grp;  // arena of packed group tables, one per distinct frontier bound
    std::map<int, size_t> grp_cache;
    size_t grp_used = 0;
    // splits
    int S = 0;
    std::vector<avt_split> splits;
    std::vector<int32_t> h_bounds;
    std::vector<uint8_t> h_seg_lut;
    DevBuf<avt_split> d_splits;
    DevBuf<int32_t> d_bounds, d_attr_first;
    DevBuf<uint8_t> d_seg_lut;
    // on-device CSV encode (avt_csv.cuh): text, line index, scan scratch, tables, results
    DevBuf<char> d_csv_text;
    DevBuf<uint32_t> d_csv_counts;
    DevBuf<unsigned long long> d_csv_first;
    DevBuf<long long> d_csv_lines, d_csv_ll;
    DevBuf<uint8_t> d_csv_tmp, d_csv_tables;
    int rot_pf = 4;   // hist_rot_kernel: L2 prefetch distance in trees (AVT_ROT_PF)
    // hist_smem_kernel: shared-memory words an attribute group may take (AVT_HIST_BUDGET_KB).  100 KB instead of the full 200 KB: two
    // blocks per SM instead of one -- C4's histogram launch (40 categorical columns x 1 000 levels) 0.855 -> 0.595 ms, 64 KB the same,
    // 32 KB 0.67, 16 KB 1.30 (more groups re-read the node ids and classes)
    long long hist_words_budget = 100 * 1024 / 4;
    // hist_sorted_kernel (avt_sorted.cuh): 0 off, 1 instead of the global-atomics fallback, 2 whenever hist_rot_kernel does not fit (AVT_HIST_SORTED)
    int hist_sorted = 2;
    DevBuf<uint16_t> d_skeys;
    DevBuf<uint32_t> d_siota, d_sperm;
    DevBuf<uint8_t> d_scls, d_sw, d_stmp;
    long long siota_rows = 0;
    long long csv_rows = 0, csv_bytes = 0;   // rows / bytes of the text the installed data set was encoded from (0: not from text)
    // avt_rows_partition scratch: row lengths, per-segment offsets, segment bases and row counts, bin -> segment table, output text
    DevBuf<int32_t> d_part_len;
    DevBuf<long long> d_part_off, d_part_base;
    DevBuf<uint16_t> d_part_lut;
    DevBuf<char> d_part_out;
    // all attributes' codes in one byte per row (narrow schemas; see HistArgs::packed)
    DevBuf<uint8_t> d_packed;
    bool packed_ok = false;
    // rotated row-major copy of the codes for hist_rot_kernel (forest shapes: >= 16 uint8 attributes), [chunks][rowrot_rows][32]
    DevBuf<uint8_t> d_rowrot;
    long long rowrot_rows = 0;
    bool rowrot_ok = false;
    uint32_t pk_shift[4] = {0, 0, 0, 0}, pk_mask4[4] = {0, 0, 0, 0};
    // item-parallel scan (avt_scan_fast.cuh): (split, segment) items, first item of every split, per-node child counts
    DevBuf<uint32_t> d_items;
    DevBuf<int32_t> d_seg_first, d_nkids;
    DevBuf<uint32_t> d_pub;        // look-back planning: published child counts (FastTable::pub)
    uint32_t scan_epoch = 0;       // tag of the last scan_fast_kernel launch that used d_pub
    int n_items = 0;
    bool fast_scan_ok = false;   // the split table is small enough for scan_fast_kernel
    // scan
    DevBuf<double> d_av;
    DevBuf<avt_node_best> d_best;
    DevBuf<uint32_t> d_mask;
    DevBuf<uint16_t> d_lut2;       // level loop: packed byte -> child rows (joint-code histogram kernel)
    DevBuf<uint32_t> d_cand[2];    // level loop: candidate attribute sets of the current / next frontier (ping-pong by level)
    DevBuf<uint8_t> d_order;
    DevBuf<int32_t> d_pairs;
    DevBuf<avt_split_eval> d_eval;
    DevBuf<avt_cpg_stat> d_cpg;
    // prediction
    DevBuf<uint32_t> d_pmask;
    DevBuf<int64_t> d_pmoff;
    DevBuf<int32_t> d_pint, d_ppoff;
    DevBuf<uint8_t> d_pcls, d_pout;
    DevBuf<double> d_pw;
    DevBuf<uint32_t> d_ccnt;
    DevBuf<int32_t> d_node_attr;
    DevBuf<uint16_t> d_lut;
    // device-driven level loop
    DevBuf<int32_t> d_prev_node_base, d_next_count, d_err, d_nb_all;
    DevBuf<uint32_t> d_ticket;
    DevBuf<avt_split_eval> d_best_eval;
    DevBuf<avt_grow_record> d_records;
    // timing
    std::vector<TimerSlot> pending;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> free_events;
    cudaEvent_t tev0 = nullptr, tev1 = nullptr;
    cudaEvent_t cev[4] = {nullptr, nullptr, nullptr, nullptr};   // avt_dataset_load_csv phases
    cudaEvent_t gev0 = nullptr, gev1 = nullptr;   // around one avt_grow_levels call (stats.last_grow_ms)
    bool gev1_recorded = false;
    avt_stats stats{};
    int force_variant = -1;
    bool skip_h_memset = false;
    DevBuf<unsigned char> d_levels;
    DevBuf<unsigned long long> d_phase;
    // level loop control block: [0] error flag, [1] last-block ticket, [16 ...] node_base of every level; initialised by one
    // copy from, and read back by one copy into, page-locked host memory (pageable copies cost ~10 us each and serialise)
    DevBuf<int32_t> d_ctrl;
    int32_t* h_ctrl = nullptr;      // page-locked, kCtrlHostInts ints
    avt_grow_record* h_rec = nullptr;   // page-locked staging for the records, kRecHostCount of them
    uint8_t order_cached[AVT_MAX_CLASSES];
    int order_cached_n = -1;        // classes covered by order_cached (-1: nothing uploaded yet)
    bool hist_pack = true;    // AVT_HIST_PACK=0: never use the bit-packed copy of the codes
    bool hist_joint = false;  // the joint-code kernel (avt_joint.cuh) is chosen only when forced (avt_set_hist_variant(3)) or with AVT_HIST_JOINT=1:
                              // its launches are shorter than hist_smem_kernel's (ncu: 193 vs 219 us over the 8 levels of C2) but the level
                              // loop as a whole measured slower with it on the same box (0.312 vs 0.290 ms per step), see profiles/README.md
    bool lookback = true;     // AVT_LOOKBACK=0: the last scan block plans the level (ticket) even for small frontiers
    bool peer_trailing_round = false;   // (round-1 experiment, kept off) end every level loop with a round of flags over all ranks
    bool scan_fast = true;    // AVT_SCAN_FAST=0: always the general scan kernel in the level loop
    bool pdl = true;          // AVT_PDL=0: plain stream launches in the level loop
    bool pdl_chain = false;   // the launch being issued directly follows a kernel of the level loop (no memset/copy between)
    bool profiling = true;  // CUDA-event timing of every launch (avt_set_profiling)
    long long hist_target_addrs = 4096;  // replication target of hist_smem_kernel: distinct words a warp's updates spread over
    int hist_max_bpsm = 2;               // histogram blocks per SM.  2 x 512 threads x 64 registers fill the register file, so more
                                         // blocks only queue behind them and add flushes (measured on C2: 0.383 ms/step with 2, 0.403 with 4, 0.52 with 1)
    // comm
    ncclComm_t comm = nullptr;
    int n_ranks = 1, rank = 0;
    // peer exchange over NVLink (avt_peer_export / avt_peer_import): the level loop's all-reduce fused into the scan
    struct Peer {
        bool active = false;
        int n = 0, rank = 0;
        uint32_t* arena = nullptr;     // local: [64 flag words | pad to 256 B | H buffer 0 | H buffer 1]
        size_t cap_words = 0;          // words per H buffer
        void* base[16] = {};           // every rank's arena mapped here (own arena for own rank)
        uint32_t epoch = 0;
        // Peer levels are numbered across calls; level k uses exchange buffer k & 1.  A rank zeroes or writes buffer k & 1
        // only after it has seen every rank's flag of level k-1, i.e. after every rank finished reading that buffer at
        // level k-2 -- so consecutive level-loop calls need no extra round of flags between them.
        size_t recv_off_words = 0;     // receive slots of the push exchange behind the histogram buffers (in uint32 words from the arena's start)
        unsigned long long seq = 0;
        bool lazy = false;             // the last level loop ended without a trailing flag round (see grow_device_loop)
    } peer;
    DevBuf<const uint32_t*> d_peer_H[2];
    DevBuf<unsigned long long*> d_push_dst[2];   // per buffer: this rank's slot array inside every rank's arena
    DevBuf<uint32_t*> d_remote_flags;
    DevBuf<uint32_t> d_hsum, d_hsum2, d_H2, d_ticket2;
    DevBuf<uint8_t> d_skip;
    DevBuf<int32_t> d_drv_parent, d_kids;
    int subtract_mode = -1;            // sibling subtraction: -1 automatic (on for the shapes the forest kernel takes), AVT_SUBTRACT=0/1 forces
    const uint8_t* skip_cur = nullptr; // skip marks for the pass being launched
    int32_t* hist_err = nullptr;       // error word of the level loop that is launching the pass (nullptr: d_err)
    int g_max_splits = 2;              // most segments of any installed split (bounds the counted nodes under sibling subtraction)
    bool ticket2_zeroed = false;
    uint32_t* H_cur = nullptr;         // histogram buffer of the pass being launched (nullptr: d_H)
    bool signal_pending = false;       // the pass being launched must signal peer.epoch to every rank
    bool peer_push = true;             // AVT_PEER_PUSH=0: always the pull form of the peer exchange
    bool peer_signal_scan = true;      // the scan kernel's first block signals the peers (false: the histogram pass's last block does -- fence + election in every block, 4.5 us per level slower)
    bool scan_will_signal = false;     // set by the stream-ordered level loop around its launch_histogram call
    bool signal_in_hist = false;       // ... and did so from inside the histogram kernel
};

namespace {

int fail(avt_engine* e, int code, const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (e) e->err = buf;
    else g_create_error = buf;
    return code;
}

#define CUDA_TRY(e, call)                                                                      \
    do {                                                                                       \
        cudaError_t _st = (call);                                                              \
        if (_st != cudaSuccess)                                                                \
            return fail((e), AVT_E_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(_st), \
                        __FILE__, __LINE__);                                                   \
    } while (0)

#define NCCL_TRY(e, call)                                                                      \
    do {                                                                                       \
        ncclResult_t _st = (call);                                                             \
        if (_st != ncclSuccess)                                                                \
            return fail((e), AVT_E_COMM, "%s failed: %s (%s:%d)", #call, ncclGetErrorString(_st), \
                        __FILE__, __LINE__);                                                   \
    } while (0)

int use_device(avt_engine* e) {
    CUDA_TRY(e, cudaSetDevice(e->device));
    return AVT_OK;
}

// device timing: events around a launch, resolved at the next synchronisation point
int timer_open(avt_engine* e, double* acc, TimerSlot* slot) {
    slot->acc = nullptr;
    if (!e->profiling) return AVT_OK;
    if (e->free_events.empty()) {
        cudaEvent_t a, b;
        CUDA_TRY(e, cudaEventCreate(&a));
        CUDA_TRY(e, cudaEventCreate(&b));
        e->free_events.push_back({a, b});
    }
    auto ev = e->free_events.back();
    e->free_events.pop_back();
    slot->a = ev.first;
    slot->b = ev.second;
    slot->acc = acc;
    CUDA_TRY(e, cudaEventRecord(slot->a, e->stream));
    return AVT_OK;
}
int timer_close(avt_engine* e, const TimerSlot& slot) {
    if (!slot.acc) return AVT_OK;
    CUDA_TRY(e, cudaEventRecord(slot.b, e->stream));
    e->pending.push_back(slot);
    return AVT_OK;
}
int sync_and_resolve(avt_engine* e) {
    CUDA_TRY(e, cudaStreamSynchronize(e->stream));
    for (auto& s : e->pending) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, s.a, s.b) == cudaSuccess) *s.acc += ms;
        e->free_events.push_back({s.a, s.b});
    }
    e->pending.clear();
    return AVT_OK;
}

template <class T>
int upload(avt_engine* e, DevBuf<T>& buf, const T* host, size_t n) {
    CUDA_TRY(e, buf.ensure(n ? n : 1));
    if (n) CUDA_TRY(e, cudaMemcpyAsync(buf.p, host, n * sizeof(T), cudaMemcpyHostToDevice, e->stream));
    return AVT_OK;
}

constexpr size_t kCtrlHostInts = 1 << 16;
constexpr size_t kRecHostCount = 4096;

SplitTable make_table(avt_engine* e) {
    SplitTable st;
    st.splits = e->d_splits.p;
    st.bounds = e->d_bounds.p;
    st.seg_lut = e->d_seg_lut.p;
    st.attr_first = e->d_attr_first.p;
    st.bin_off = e->d_bin_off.p;
    st.S = e->S;
    st.A = e->A;
    st.TB = e->TB;
    st.C = e->C;
    return st;
}

int upload_order(avt_engine* e, const uint8_t* class_order) {
    uint8_t ord[AVT_MAX_CLASSES];
    bool seen[AVT_MAX_CLASSES] = {false};
    for (int i = 0; i < e->C; ++i) {
        ord[i] = class_order ? class_order[i] : (uint8_t)i;
        if (ord[i] >= e->C || seen[ord[i]]) return fail(e, AVT_E_INVALID, "class_order is not a permutation of 0..%d", e->C - 1);
        seen[ord[i]] = true;
    }
    if (e->order_cached_n == e->C && memcmp(ord, e->order_cached, (size_t)e->C) == 0) return AVT_OK;   // already on the device
    if (int rc = upload(e, e->d_order, ord, (size_t)e->C)) return rc;
    memcpy(e->order_cached, ord, (size_t)e->C);
    e->order_cached_n = e->C;
    return AVT_OK;
}

}  // namespace

extern "C" {

const char* avt_last_error(const avt_engine* e) { return e ? e->err.c_str() : g_create_error.c_str(); }

int avt_create(int device, avt_engine** out) {
    if (!out) return fail(nullptr, AVT_E_INVALID, "avt_create: out is NULL");
    *out = nullptr;
    int count = 0;
    cudaError_t st = cudaGetDeviceCount(&count);
    if (st != cudaSuccess || count == 0)
        return fail(nullptr, AVT_E_CUDA, "avt_create: no CUDA device (%s); libavenir_tree has no CPU path",
                    st == cudaSuccess ? "device count is 0" : cudaGetErrorString(st));
    if (device < 0 || device >= count) return fail(nullptr, AVT_E_INVALID, "avt_create: device %d out of range [0,%d)", device, count);
    avt_engine* e = new avt_engine();
    e->device = device;
    st = cudaSetDevice(device);
    if (st == cudaSuccess) st = cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking);
    if (st == cudaSuccess) st = cudaEventCreate(&e->tev0);
    if (st == cudaSuccess) st = cudaEventCreate(&e->tev1);
    if (st == cudaSuccess) st = cudaEventCreate(&e->gev0);
    if (st == cudaSuccess) st = cudaEventCreate(&e->gev1);
    cudaDeviceProp prop;
    if (st == cudaSuccess) st = cudaGetDeviceProperties(&prop, device);
    if (st == cudaSuccess) {
        e->sm_count = prop.multiProcessorCount;
        if (prop.major < 10) {
            delete e;
            return fail(nullptr, AVT_E_CUDA, "avt_create: device %d is sm_%d%d; this library is built for sm_100a (B200) only", device, prop.major, prop.minor);
        }
    }
    {
        const void* kernels[] = {(const void*)hist_smem_kernel<false, false, false>, (const void*)hist_smem_kernel<false, false, true>,
                                 (const void*)hist_smem_kernel<false, true, false>,  (const void*)hist_smem_kernel<false, true, true>,
                                 (const void*)hist_smem_kernel<true, false, false>,  (const void*)hist_smem_kernel<true, false, true>,
                                 (const void*)hist_smem_kernel<true, true, false>,   (const void*)hist_smem_kernel<true, true, true>};
        for (const void* k : kernels)
            if (st == cudaSuccess) st = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem);
    }
    {
        const void* kernels[] = {(const void*)hist_rot_kernel<false, false>, (const void*)hist_rot_kernel<false, true>,
                                 (const void*)hist_rot_kernel<true, false>, (const void*)hist_rot_kernel<true, true>,
                                 (const void*)hist_joint_kernel<false, false>, (const void*)hist_joint_kernel<false, true>,
                                 (const void*)hist_joint_kernel<true, false>, (const void*)hist_joint_kernel<true, true>};
        for (const void* k : kernels)
            if (st == cudaSuccess) st = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem);
    }
    if (st != cudaSuccess) {
        int rc = fail(nullptr, AVT_E_CUDA, "avt_create: %s", cudaGetErrorString(st));
        delete e;
        return rc;
    }
    if (const char* v = getenv("AVT_PDL")) e->pdl = atoi(v) != 0;
    if (const char* v = getenv("AVT_SCAN_FAST")) e->scan_fast = atoi(v) != 0;
    if (const char* v = getenv("AVT_PEER_PUSH")) e->peer_push = atoi(v) != 0;
    if (const char* v = getenv("AVT_LOOKBACK")) e->lookback = atoi(v) != 0;
    if (const char* v = getenv("AVT_HIST_PACK")) e->hist_pack = atoi(v) != 0;
    if (const char* v = getenv("AVT_HIST_JOINT")) e->hist_joint = atoi(v) != 0;
    if (const char* v = getenv("AVT_SUBTRACT")) e->subtract_mode = atoi(v) != 0 ? 1 : 0;
    if (const char* v = getenv("AVT_ROT_PF")) e->rot_pf = std::max(0, std::min(64, atoi(v)));
    if (const char* v = getenv("AVT_HIST_BUDGET_KB")) e->hist_words_budget = std::max(8, std::min(200, atoi(v))) * 1024LL / 4;
    if (const char* v = getenv("AVT_HIST_SORTED")) e->hist_sorted = std::max(0, std::min(2, atoi(v)));
    *out = e;
    return AVT_OK;
}

void avt_destroy(avt_engine* e) {
    if (!e) return;
    cudaSetDevice(e->device);
    if (e->stream && e->peer.active && e->peer.lazy) {
        // the last level loop left out its trailing round of flags: another rank may still be reading this rank's exchange
        // buffers.  Tell every rank we are done and wait (bounded) until all of them are, before the arena is freed.
        ++e->peer.epoch;
        peer_signal_kernel<<<1, 32, 0, e->stream>>>(e->d_remote_flags.p, e->peer.n, e->peer.epoch);
        peer_wait_bounded_kernel<<<1, 32, 0, e->stream>>>(e->peer.arena, e->peer.n, e->peer.epoch, 1 << 24);
        e->peer.lazy = false;
    }
    if (e->stream) cudaStreamSynchronize(e->stream);
    if (e->comm) ncclCommDestroy(e->comm);
    for (int r = 0; r < e->peer.n; ++r)
        if (r != e->peer.rank && e->peer.base[r]) cudaIpcCloseMemHandle(e->peer.base[r]);
    if (e->peer.arena) cudaFree(e->peer.arena);
    e->d_peer_H[0].release(); e->d_peer_H[1].release(); e->d_push_dst[0].release(); e->d_push_dst[1].release(); e->d_remote_flags.release(); e->d_hsum.release(); e->d_ticket2.release(); e->d_hsum2.release(); e->d_H2.release(); e->d_skip.release(); e->d_drv_parent.release(); e->d_kids.release();
    for (void* c : e->cols)
        if (c) cudaFree(c);
    e->d_colptrs.release(); e->d_cls.release(); e->d_bin_off.release(); e->d_wide.release();
    e->d_nid.release(); e->d_levels.release(); e->d_nid2.release(); e->d_nb_all.release(); e->d_ticket.release(); e->d_prev_node_base.release(); e->d_next_count.release(); e->d_err.release(); e->d_best_eval.release(); e->d_records.release(); e->d_w.release(); e->d_has_w.release(); e->d_node_base.release();
    e->d_pmask.release(); e->d_pmoff.release(); e->d_pint.release(); e->d_ppoff.release(); e->d_pcls.release(); e->d_pout.release(); e->d_pw.release();
    e->d_H.release(); e->d_grp.release(); e->d_splits.release(); e->d_bounds.release();
    e->d_attr_first.release(); e->d_seg_lut.release(); e->d_items.release(); e->d_seg_first.release(); e->d_nkids.release(); e->d_csv_text.release(); e->d_csv_counts.release(); e->d_csv_first.release(); e->d_csv_lines.release(); e->d_csv_ll.release(); e->d_csv_tmp.release(); e->d_csv_tables.release(); e->d_part_len.release(); e->d_part_off.release(); e->d_part_base.release(); e->d_part_lut.release(); e->d_part_out.release(); e->d_skeys.release(); e->d_siota.release(); e->d_sperm.release(); e->d_scls.release(); e->d_sw.release(); e->d_stmp.release(); e->d_pub.release(); e->d_packed.release(); e->d_rowrot.release(); e->d_av.release(); e->d_best.release();
    e->d_mask.release(); e->d_lut2.release(); e->d_cand[0].release(); e->d_cand[1].release(); e->d_order.release(); e->d_pairs.release(); e->d_eval.release();
    e->d_cpg.release(); e->d_ccnt.release(); e->d_node_attr.release(); e->d_lut.release(); e->d_ctrl.release();
    if (e->h_ctrl) cudaFreeHost(e->h_ctrl);
    if (e->h_rec) cudaFreeHost(e->h_rec);
    for (auto& s : e->pending) { cudaEventDestroy(s.a); cudaEventDestroy(s.b); }
    for (auto& p : e->free_events) { cudaEventDestroy(p.first); cudaEventDestroy(p.second); }
    if (e->tev0) cudaEventDestroy(e->tev0);
    if (e->tev1) cudaEventDestroy(e->tev1);
    for (auto& ev : e->cev) if (ev) cudaEventDestroy(ev);
    if (e->gev0) cudaEventDestroy(e->gev0);
    if (e->gev1) cudaEventDestroy(e->gev1);
    if (e->stream) cudaStreamDestroy(e->stream);
    delete e;
}

// Shape of a new data set: metadata, column / class buffers (kept when the shape is unchanged: re-loading costs no
// cudaMalloc/cudaFree), zeroed padding rows.  The caller fills the first n_rows entries, then calls dataset_install.
static int dataset_shape(avt_engine* e, int64_t n_rows, int32_t n_attrs, const int32_t* n_bins, int32_t n_classes, const char* who) {
    if (n_classes < 1 || n_classes > AVT_MAX_CLASSES)
        return fail(e, AVT_E_INVALID, "%s: n_classes %d outside 1..%d", who, n_classes, AVT_MAX_CLASSES);
    const long long new_npad = (n_rows + 15) & ~15LL;
    bool same_shape = (int)e->cols.size() == n_attrs && e->Npad == new_npad && (int)e->wide.size() == n_attrs;
    for (int a = 0; same_shape && a < n_attrs; ++a) same_shape = e->cols[a] != nullptr && e->wide[a] == (n_bins[a] > 256 ? 1 : 0);
    if (!same_shape) {
        for (void* c : e->cols)
            if (c) cudaFree(c);
        e->cols.assign(n_attrs, nullptr);
    }
    e->N = n_rows;
    e->Npad = new_npad;
    e->A = n_attrs;
    e->C = n_classes;
    e->n_bins.assign(n_bins, n_bins + n_attrs);
    e->bin_off.assign(n_attrs + 1, 0);
    e->wide.assign(n_attrs, 0);
    for (int a = 0; a < n_attrs; ++a) {
        if (n_bins[a] < 1 || n_bins[a] > 65535) return fail(e, AVT_E_INVALID, "%s: n_bins[%d]=%d outside 1..65535", who, a, n_bins[a]);
        e->bin_off[a + 1] = e->bin_off[a] + n_bins[a];
        e->wide[a] = n_bins[a] > 256;
    }
    e->TB = e->bin_off[n_attrs];
    for (int a = 0; a < n_attrs; ++a) {
        const size_t w = e->wide[a] ? 2 : 1;
        if (!e->cols[a]) CUDA_TRY(e, cudaMalloc(&e->cols[a], ((size_t)e->Npad + 32) * w));   // + one tile of slack (hist_rot_kernel reads whole 32-row tiles)
        CUDA_TRY(e, cudaMemsetAsync((char*)e->cols[a] + (size_t)n_rows * w, 0, (size_t)(e->Npad - n_rows) * w, e->stream));
    }
    CUDA_TRY(e, e->d_cls.ensure((size_t)e->Npad + 32));
    CUDA_TRY(e, cudaMemsetAsync(e->d_cls.p + n_rows, 0, (size_t)(e->Npad - n_rows), e->stream));
    if (int rc = upload(e, e->d_colptrs, e->cols.data(), (size_t)n_attrs)) return rc;
    if (int rc = upload(e, e->d_bin_off, e->bin_off.data(), (size_t)n_attrs + 1)) return rc;
    if (int rc = upload(e, e->d_wide, e->wide.data(), (size_t)n_attrs)) return rc;
    return AVT_OK;
}

// The columns are filled (stream-ordered): derived copies, state reset.
static int dataset_install(avt_engine* e) {
    const int n_attrs = e->A;
    {
        // bit-packed copy of the codes for the histogram pass: one byte per row instead of one per attribute
        int bits = 0;
        bool ok = e->hist_pack && n_attrs <= 4;
        for (int a = 0; a < n_attrs && ok; ++a) {
            int w = 0;
            while ((1 << w) < e->n_bins[a]) ++w;
            e->pk_shift[a] = (uint32_t)bits;
            e->pk_mask4[a] = ((1u << w) - 1u) * 0x01010101u;
            bits += w;
            ok = !e->wide[a] && bits <= 8;
        }
        e->packed_ok = ok;
        if (ok) {
            CUDA_TRY(e, e->d_packed.ensure((size_t)e->Npad));
            const long long groups = e->Npad / 16;
            pack_codes_kernel<<<(unsigned)((groups + 255) / 256), 256, 0, e->stream>>>((const void* const*)e->d_colptrs.p, n_attrs,
                                                                                      make_uint4(e->pk_shift[0], e->pk_shift[1], e->pk_shift[2], e->pk_shift[3]), e->Npad, e->d_packed.p);
            CUDA_TRY(e, cudaGetLastError());
        }
    }
    {
        bool shape = n_attrs >= 16;
        for (int a = 0; a < n_attrs && shape; ++a) shape = !e->wide[a];
        e->rowrot_ok = false;
        if (shape) {
            const int chunks = (n_attrs + 31) / 32;
            e->rowrot_rows = e->Npad + 32;
            CUDA_TRY(e, e->d_rowrot.ensure((size_t)chunks * e->rowrot_rows * 32));
            rowrot_kernel<<<dim3((unsigned)((e->rowrot_rows + 255) / 256), (unsigned)chunks), 256, 0, e->stream>>>((const void* const*)e->d_colptrs.p, n_attrs, e->N,
                                                                                                               e->rowrot_rows, e->d_rowrot.p);
            CUDA_TRY(e, cudaGetLastError());
            e->rowrot_ok = true;
        }
    }
    e->T = 0;
    e->S = 0;
    e->grp_cache.clear();
    e->grp_used = 0;
    e->hist_valid = e->scan_valid = false;
    return AVT_OK;
}

int avt_dataset_load(avt_engine* e, int64_t n_rows, int32_t n_attrs, const int32_t* n_bins,
                     const void* const* columns, const uint8_t* class_codes, int32_t n_classes) {
    if (!e) return AVT_E_INVALID;
    if (n_rows <= 0 || n_attrs <= 0 || !n_bins || !columns || !class_codes)
        return fail(e, AVT_E_INVALID, "avt_dataset_load: empty data set or NULL argument");
    if (int rc = use_device(e)) return rc;
    TimerSlot ts;
    if (int rc = timer_open(e, &e->stats.h2d_ms, &ts)) return rc;
    e->csv_rows = 0;
    if (int rc = dataset_shape(e, n_rows, n_attrs, n_bins, n_classes, "avt_dataset_load")) return rc;
    for (int a = 0; a < n_attrs; ++a) {
        const size_t w = e->wide[a] ? 2 : 1;
        CUDA_TRY(e, cudaMemcpyAsync(e->cols[a], columns[a], (size_t)n_rows * w, cudaMemcpyDefault, e->stream));   // host or device source
    }
    CUDA_TRY(e, cudaMemcpyAsync(e->d_cls.p, class_codes, (size_t)n_rows, cudaMemcpyDefault, e->stream));
    if (int rc = dataset_install(e)) return rc;
    if (int rc = timer_close(e, ts)) return rc;
    return sync_and_resolve(e);
}

int avt_dataset_load_csv(avt_engine* e, const char* text, int64_t n_bytes, char delim, int32_t skip_fields,
                         int32_t n_attrs, const avt_csv_attr* attrs, const int32_t* int_points,
                         const double* dbl_points, const char* cat_bytes, const int32_t* cat_offsets,
                         int32_t class_ordinal, int32_t n_class_values, const char* class_bytes,
                         const int32_t* class_offsets, int64_t* n_rows_out, int64_t* class_first_row,
                         int64_t* bad_row, int32_t* bad_reason) {
    if (!e) return AVT_E_INVALID;
    if (bad_row) *bad_row = -1;
    if (bad_reason) *bad_reason = 0;
    const bool no_class = class_ordinal == -1;   // rows without a class field (DataPartitioner): one class, every row has code 0
    static const int32_t kNoClassOffsets[2] = {0, 0};
    if (no_class) {
        if (n_class_values != 1) return fail(e, AVT_E_INVALID, "avt_dataset_load_csv: class_ordinal -1 (no class field) needs n_class_values == 1");
        class_bytes = "";
        class_offsets = kNoClassOffsets;
    }
    if (!text || n_bytes <= 0 || n_attrs <= 0 || !attrs || n_class_values < 1 || !class_bytes || !class_offsets || skip_fields < 0)
        return fail(e, AVT_E_INVALID, "avt_dataset_load_csv: empty text or NULL argument");
    if (delim == '\n' || delim == '\r') return fail(e, AVT_E_INVALID, "avt_dataset_load_csv: the field delimiter cannot be a line terminator");
    if (class_ordinal < -1 || class_ordinal >= kCsvMaxFields) return fail(e, AVT_E_INVALID, "avt_dataset_load_csv: class ordinal %d outside -1..%d", class_ordinal, kCsvMaxFields - 1);
    e->csv_rows = 0;   // the text of an earlier call is gone from here on (avt_rows_partition)
    // ---- tables: attribute of every field, grids, value lists with their byte-sorted permutations
    std::vector<int32_t> n_bins((size_t)n_attrs);
    int n_int = 0, n_dbl = 0, n_cat = 0, max_ord = std::max(class_ordinal, 0);
    for (int a = 0; a < n_attrs; ++a) {
        const avt_csv_attr& A = attrs[a];
        if (A.ordinal < 0 || A.ordinal >= kCsvMaxFields || (a && A.ordinal <= attrs[a - 1].ordinal) || A.ordinal == class_ordinal)
            return fail(e, AVT_E_INVALID, "avt_dataset_load_csv: attribute %d: ordinals must increase, stay below %d and differ from the class ordinal", a, kCsvMaxFields);
        if (A.n < 0 || A.off < 0) return fail(e, AVT_E_INVALID, "avt_dataset_load_csv: attribute %d: negative size or offset", a);
        max_ord = std::max(max_ord, A.ordinal);
        if (A.kind == AVT_CSV_INT) {
            if (A.n && !int_points) return fail(e, AVT_E_INVALID, "avt_dataset_load_csv: int_points is NULL");
            n_int = std::max(n_int, A.off + A.n);
            n_bins[(size_t)a] = A.n + 1;
        } else if (A.kind == AVT_CSV_DOUBLE) {
            if (A.n && !dbl_points) return fail(e, AVT_E_INVALID, "avt_dataset_load_csv: dbl_points is NULL");
            n_dbl = std::max(n_dbl, A.off + A.n);
            n_bins[(size_t)a] = A.n + 1;
        } else if (A.kind == AVT_CSV_CATEGORICAL) {
            if (A.n < 1 || !cat_bytes || !cat_offsets) return fail(e, AVT_E_INVALID, "avt_dataset_load_csv: attribute %d: empty value list", a);
            n_cat = std::max(n_cat, A.off + A.n);
            n_bins[(size_t)a] = A.n;
        } else {
            return fail(e, AVT_E_INVALID, "avt_dataset_load_csv: attribute %d: unknown kind %d", a, A.kind);
        }
    }
    auto sorted_perm = [](const char* bytes, const int32_t* off, int n, int32_t* out) {
        for (int k = 0; k < n; ++k) out[k] = k;
        std::sort(out, out + n, [&](int32_t x, int32_t y) {
            const int lx = off[x + 1] - off[x], ly = off[y + 1] - off[y];
            const int c = memcmp(bytes + off[x], bytes + off[y], (size_t)std::min(lx, ly));
            return c != 0 ? c < 0 : lx < ly;
        });
    };
    std::vector<int16_t> field_attr((size_t)max_ord + 1, (int16_t)-1);
    for (int a = 0; a < n_attrs; ++a) field_attr[(size_t)attrs[a].ordinal] = (int16_t)a;
    std::vector<int32_t> cat_sorted((size_t)std::max(n_cat, 1)), class_sorted((size_t)n_class_values);
    for (int a = 0; a < n_attrs; ++a)
        if (attrs[a].kind == AVT_CSV_CATEGORICAL) sorted_perm(cat_bytes, cat_offsets + attrs[a].off, attrs[a].n, cat_sorted.data() + attrs[a].off);
    sorted_perm(class_bytes, class_offsets, n_class_values, class_sorted.data());
    const size_t cat_nbytes = n_cat ? (size_t)cat_offsets[n_cat] : 0, class_nbytes = (size_t)class_offsets[n_class_values];
    // one packed upload: every table (and the two byte pools) 16-byte aligned inside d_csv_tables
    std::vector<uint8_t> pack;
    auto put = [&](const void* src, size_t bytes) {
        const size_t at = (pack.size() + 15) & ~(size_t)15;
        pack.resize(at + bytes);
        if (bytes) memcpy(pack.data() + at, src, bytes);
        return at;
    };
    const size_t o_attrs = put(attrs, (size_t)n_attrs * sizeof(avt_csv_attr)), o_fa = put(field_attr.data(), field_attr.size() * sizeof(int16_t)),
                 o_ip = put(int_points, (size_t)n_int * 4), o_dp = put(dbl_points, (size_t)n_dbl * 8), o_co = put(cat_offsets, n_cat ? ((size_t)n_cat + 1) * 4 : 0),
                 o_cs = put(cat_sorted.data(), (size_t)n_cat * 4), o_klo = put(class_offsets, ((size_t)n_class_values + 1) * 4),
                 o_kls = put(class_sorted.data(), (size_t)n_class_values * 4), o_cb = put(cat_bytes, cat_nbytes), o_kb = put(class_bytes, class_nbytes);
    if (int rc = use_device(e)) return rc;
    TimerSlot ts;
    if (int rc = timer_open(e, &e->stats.h2d_ms, &ts)) return rc;
    for (auto& ev : e->cev) if (!ev) CUDA_TRY(e, cudaEventCreate(&ev));
    CUDA_TRY(e, cudaEventRecord(e->cev[0], e->stream));
    CUDA_TRY(e, e->d_csv_tables.ensure(pack.size()));
    CUDA_TRY(e, cudaMemcpyAsync(e->d_csv_tables.p, pack.data(), pack.size(), cudaMemcpyHostToDevice, e->stream));
    CUDA_TRY(e, e->d_csv_text.ensure((size_t)n_bytes + 16));
    CUDA_TRY(e, cudaMemcpyAsync(e->d_csv_text.p, text, (size_t)n_bytes, cudaMemcpyHostToDevice, e->stream));
    CUDA_TRY(e, cudaEventRecord(e->cev[1], e->stream));
    // ---- line index: newlines per chunk, exclusive prefix sum, row starts
    const long long n_chunks = (n_bytes + kCsvChunk - 1) / kCsvChunk;
    CUDA_TRY(e, e->d_csv_counts.ensure((size_t)n_chunks));
    CUDA_TRY(e, e->d_csv_first.ensure((size_t)n_chunks + 1));
    csv_count_kernel<<<(unsigned)((n_chunks + 255) / 256), 256, 0, e->stream>>>(e->d_csv_text.p, n_bytes, e->d_csv_counts.p, n_chunks);
    CUDA_TRY(e, cudaGetLastError());
    size_t tmp_bytes = 0;
    CUDA_TRY(e, cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, e->d_csv_counts.p, e->d_csv_first.p, (int)n_chunks, e->stream));
    CUDA_TRY(e, e->d_csv_tmp.ensure(tmp_bytes));
    CUDA_TRY(e, cub::DeviceScan::ExclusiveSum(e->d_csv_tmp.p, tmp_bytes, e->d_csv_counts.p, e->d_csv_first.p, (int)n_chunks, e->stream));
    unsigned long long last_first = 0;
    uint32_t last_count = 0;
    CUDA_TRY(e, cudaMemcpyAsync(&last_first, e->d_csv_first.p + (n_chunks - 1), sizeof last_first, cudaMemcpyDeviceToHost, e->stream));
    CUDA_TRY(e, cudaMemcpyAsync(&last_count, e->d_csv_counts.p + (n_chunks - 1), sizeof last_count, cudaMemcpyDeviceToHost, e->stream));
    if (int rc = sync_and_resolve(e)) return rc;
    const long long newlines = (long long)last_first + last_count;
    const bool open_tail = text[n_bytes - 1] != '\n';
    const long long n_rows = newlines + (open_tail ? 1 : 0);
    if (n_rows_out) *n_rows_out = n_rows;
    if (n_rows <= 0) return fail(e, AVT_E_INVALID, "avt_dataset_load_csv: no rows");
    CUDA_TRY(e, e->d_csv_lines.ensure((size_t)n_rows + 1));
    const long long first_start = 0, tail_end = n_bytes + 1;   // a virtual '\n' behind an unterminated last row
    CUDA_TRY(e, cudaMemcpyAsync(e->d_csv_lines.p, &first_start, sizeof first_start, cudaMemcpyHostToDevice, e->stream));
    if (open_tail) CUDA_TRY(e, cudaMemcpyAsync(e->d_csv_lines.p + n_rows, &tail_end, sizeof tail_end, cudaMemcpyHostToDevice, e->stream));
    csv_lines_kernel<<<(unsigned)((n_chunks + 255) / 256), 256, 0, e->stream>>>(e->d_csv_text.p, n_bytes, e->d_csv_first.p, e->d_csv_lines.p, n_chunks);
    CUDA_TRY(e, cudaGetLastError());
    CUDA_TRY(e, cudaEventRecord(e->cev[2], e->stream));
    // ---- the data set's buffers, then one thread per row
    if (int rc = dataset_shape(e, n_rows, n_attrs, n_bins.data(), n_class_values, "avt_dataset_load_csv")) return rc;
    CUDA_TRY(e, e->d_csv_ll.ensure((size_t)n_class_values + 2 + ((size_t)kCsvSlowCap * sizeof(CsvSlow) + 7) / 8));
    std::vector<long long> init((size_t)n_class_values + 2, 0x7fffffffffffffffLL);
    init[(size_t)n_class_values] = -1LL;   // bad: all ones
    init[(size_t)n_class_values + 1] = 0;  // n_slow
    CUDA_TRY(e, cudaMemcpyAsync(e->d_csv_ll.p, init.data(), init.size() * 8, cudaMemcpyHostToDevice, e->stream));
    CsvArgs ca{};
    const uint8_t* tb = e->d_csv_tables.p;
    ca.text = e->d_csv_text.p; ca.n_bytes = n_bytes; ca.line_start = e->d_csv_lines.p; ca.n_rows = n_rows;
    ca.delim = delim; ca.skip_fields = skip_fields; ca.n_attrs = n_attrs;
    ca.attrs = (const avt_csv_attr*)(tb + o_attrs); ca.field_attr = (const int16_t*)(tb + o_fa); ca.n_fields_used = (int)field_attr.size();
    ca.int_points = (const int32_t*)(tb + o_ip); ca.dbl_points = (const double*)(tb + o_dp);
    ca.cat_bytes = (const char*)(tb + o_cb); ca.cat_offsets = (const int32_t*)(tb + o_co); ca.cat_sorted = (const int32_t*)(tb + o_cs);
    ca.class_field = no_class ? kCsvNoClassField : class_ordinal; ca.n_class_values = n_class_values;
    ca.class_bytes = (const char*)(tb + o_kb); ca.class_offsets = (const int32_t*)(tb + o_klo); ca.class_sorted = (const int32_t*)(tb + o_kls);
    ca.cols = (void* const*)e->d_colptrs.p; ca.wide = e->d_wide.p; ca.cls = e->d_cls.p;
    ca.class_first_row = e->d_csv_ll.p;
    ca.bad = (unsigned long long*)(e->d_csv_ll.p + n_class_values);
    ca.n_slow = (uint32_t*)(e->d_csv_ll.p + n_class_values + 1);
    ca.slow = (CsvSlow*)(e->d_csv_ll.p + n_class_values + 2);
    csv_encode_kernel<<<(unsigned)((n_rows + 127) / 128), 128, 0, e->stream>>>(ca);
    CUDA_TRY(e, cudaGetLastError());
    CUDA_TRY(e, cudaEventRecord(e->cev[3], e->stream));
    std::vector<long long> back((size_t)n_class_values + 2);
    CUDA_TRY(e, cudaMemcpyAsync(back.data(), e->d_csv_ll.p, back.size() * 8, cudaMemcpyDeviceToHost, e->stream));
    if (int rc = sync_and_resolve(e)) return rc;
    {
        float ms[3] = {0.f, 0.f, 0.f};
        for (int k = 0; k < 3; ++k) CUDA_TRY(e, cudaEventElapsedTime(&ms[k], e->cev[k], e->cev[k + 1]));
        e->stats.last_csv_h2d_ms = ms[0];
        e->stats.last_csv_index_ms = ms[1];
        e->stats.last_csv_encode_ms = ms[2];   // includes the allocation of the columns on a first call of this shape
    }
    unsigned long long bad = (unsigned long long)back[(size_t)n_class_values];
    const uint32_t n_slow = (uint32_t)(back[(size_t)n_class_values + 1] & 0xffffffffLL);
    if (n_slow > (uint32_t)kCsvSlowCap)
        return fail(e, AVT_E_INVALID, "avt_dataset_load_csv: %u values need the host's strtod (limit %d): encode this input on the host", n_slow, kCsvSlowCap);
    if (n_slow) {
        // Double.parseDouble inputs outside the device parser's exact range: convert here, patch the codes in
        std::vector<CsvSlow> slow(n_slow);
        CUDA_TRY(e, cudaMemcpy(slow.data(), ca.slow, (size_t)n_slow * sizeof(CsvSlow), cudaMemcpyDeviceToHost));
        std::vector<CsvPatch> patches;
        for (const CsvSlow& sv : slow) {
            std::string t(text + sv.start, (size_t)sv.len);
            size_t b0 = 0, b1 = t.size();
            while (b0 < b1 && (unsigned char)t[b0] <= ' ') ++b0;
            while (b1 > b0 && (unsigned char)t[b1 - 1] <= ' ') --b1;
            t = t.substr(b0, b1 - b0);
            char* end = nullptr;
            const double v = t.empty() ? 0.0 : strtod(t.c_str(), &end);
            if (t.empty() || *end) {
                bad = std::min(bad, ((unsigned long long)sv.row << 8) | (unsigned long long)AVT_CSV_BAD_DOUBLE);
                continue;
            }
            const avt_csv_attr& A = attrs[sv.attr];
            const int bin = (int)(std::lower_bound(dbl_points + A.off, dbl_points + A.off + A.n, v) - (dbl_points + A.off));
            patches.push_back(CsvPatch{sv.row, sv.attr, bin});
        }
        if (!patches.empty()) {
            CsvPatch* d_p = (CsvPatch*)ca.slow;   // the list has been copied out: reuse its storage (24 -> 16 bytes per entry)
            CUDA_TRY(e, cudaMemcpyAsync(d_p, patches.data(), patches.size() * sizeof(CsvPatch), cudaMemcpyHostToDevice, e->stream));
            csv_patch_kernel<<<(unsigned)((patches.size() + 127) / 128), 128, 0, e->stream>>>(d_p, (int)patches.size(), (void* const*)e->d_colptrs.p, e->d_wide.p);
            CUDA_TRY(e, cudaGetLastError());
            if (int rc = sync_and_resolve(e)) return rc;   // `patches` goes out of scope
        }
    }
    if (bad != ~0ULL) {
        const long long row = (long long)(bad >> 8);
        const int reason = (int)(bad & 0xff);
        if (bad_row) *bad_row = row;
        if (bad_reason) *bad_reason = reason;
        e->N = 0;   // nothing is installed
        static const char* const what[] = {"", "row shorter than an attribute's or the class's ordinal", "NumberFormatException (Integer.parseInt)",
                                           "NumberFormatException (Double.parseDouble)", "no matching predicate for a categorical value", "class value not in the class list"};
        return fail(e, AVT_E_REFERENCE, "avt_dataset_load_csv: row %lld: %s", row, what[reason >= 1 && reason <= 5 ? reason : 0]);
    }
    if (class_first_row)
        for (int k = 0; k < n_class_values; ++k) class_first_row[k] = no_class ? 0 : back[(size_t)k] == 0x7fffffffffffffffLL ? -1 : back[(size_t)k];
    if (int rc = dataset_install(e)) return rc;
    e->csv_rows = n_rows;   // d_csv_text / d_csv_lines stay valid for avt_rows_partition until the next data set is loaded
    e->csv_bytes = n_bytes;
    if (int rc = timer_close(e, ts)) return rc;
    return sync_and_resolve(e);
}

int avt_dataset_fetch_column(avt_engine* e, int32_t attr, void* out) {
    if (!e) return AVT_E_INVALID;
    if (e->N == 0) return fail(e, AVT_E_STATE, "avt_dataset_fetch_column: no data set loaded");
    if (!out || attr < -1 || attr >= e->A) return fail(e, AVT_E_INVALID, "avt_dataset_fetch_column: attribute %d out of range", attr);
    if (int rc = use_device(e)) return rc;
    if (attr < 0) CUDA_TRY(e, cudaMemcpyAsync(out, e->d_cls.p, (size_t)e->N, cudaMemcpyDeviceToHost, e->stream));
    else CUDA_TRY(e, cudaMemcpyAsync(out, e->cols[(size_t)attr], (size_t)e->N * (e->wide[(size_t)attr] ? 2 : 1), cudaMemcpyDeviceToHost, e->stream));
    return sync_and_resolve(e);
}

}  // extern "C"

template <typename CodeT>
static int rows_partition_device(avt_engine* e, const CodeT* codes, int n_segments) {
    const long long N = e->N;
    const unsigned gb = (unsigned)((N + 255) / 256);
    long long* base = e->d_part_base.p;
    unsigned long long* seg_rows = reinterpret_cast<unsigned long long*>(base + n_segments + 1);
    rows_len_kernel<<<gb, 256, 0, e->stream>>>(e->d_csv_text.p, e->d_csv_lines.p, N, e->d_part_len.p);
    CUDA_TRY(e, cudaGetLastError());
    size_t tmp_bytes = 0;
    CUDA_TRY(e, cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, e->d_part_off.p, e->d_part_off.p, (int)N, e->stream));
    CUDA_TRY(e, e->d_csv_tmp.ensure(tmp_bytes));
    for (int s = 0; s < n_segments; ++s) {
        rows_mask_kernel<CodeT><<<gb, 256, 0, e->stream>>>(codes, e->d_part_lut.p, e->d_part_len.p, N, s, e->d_part_off.p, seg_rows);
        CUDA_TRY(e, cub::DeviceScan::ExclusiveSum(e->d_csv_tmp.p, tmp_bytes, e->d_part_off.p, e->d_part_off.p, (int)N, e->stream));
        rows_base_kernel<CodeT><<<1, 1, 0, e->stream>>>(codes, e->d_part_lut.p, e->d_part_len.p, N, s, e->d_part_off.p, base);
        rows_copy_kernel<CodeT><<<(unsigned)((N * 32 + 255) / 256), 256, 0, e->stream>>>(e->d_csv_text.p, e->d_csv_lines.p, codes, e->d_part_lut.p, e->d_part_len.p, N, s,
                                                                                          e->d_part_off.p, base, e->d_part_out.p);
        CUDA_TRY(e, cudaGetLastError());
    }
    return AVT_OK;
}

extern "C" {

int avt_rows_partition(avt_engine* e, int32_t attr, const uint16_t* segment_of_bin, int32_t n_segments, char* out_text, int64_t out_capacity,
                       int64_t* segment_bytes, int64_t* segment_rows) {
    if (!e) return AVT_E_INVALID;
    if (e->N == 0 || e->csv_rows != e->N)
        return fail(e, AVT_E_STATE, "avt_rows_partition: the installed data set was not encoded from text (avt_dataset_load_csv keeps the text on the device)");
    if (attr < 0 || attr >= e->A || !segment_of_bin || n_segments < 1 || n_segments > 4096 || !segment_bytes)
        return fail(e, AVT_E_INVALID, "avt_rows_partition: attribute %d / %d segments out of range, or NULL argument", attr, n_segments);
    if (e->N >= (1LL << 31)) return fail(e, AVT_E_INVALID, "avt_rows_partition: more than 2^31 - 1 rows");
    const int nb = e->n_bins[(size_t)attr];
    for (int b = 0; b < nb; ++b)
        if (segment_of_bin[b] != AVT_NODE_NONE && segment_of_bin[b] >= n_segments)
            return fail(e, AVT_E_INVALID, "avt_rows_partition: bin %d maps to segment %d of %d", b, (int)segment_of_bin[b], n_segments);
    if (int rc = use_device(e)) return rc;
    CUDA_TRY(e, e->d_part_len.ensure((size_t)e->N));
    CUDA_TRY(e, e->d_part_off.ensure((size_t)e->N));
    CUDA_TRY(e, e->d_part_base.ensure(2 * (size_t)n_segments + 1));
    CUDA_TRY(e, e->d_part_lut.ensure((size_t)nb));
    CUDA_TRY(e, e->d_part_out.ensure((size_t)e->csv_bytes + 1));   // no row grows: its terminator becomes one '\n'; an open last row gains one
    CUDA_TRY(e, cudaMemcpyAsync(e->d_part_lut.p, segment_of_bin, (size_t)nb * sizeof(uint16_t), cudaMemcpyHostToDevice, e->stream));
    CUDA_TRY(e, cudaMemsetAsync(e->d_part_base.p, 0, (2 * (size_t)n_segments + 1) * sizeof(long long), e->stream));
    if (int rc = e->wide[(size_t)attr] ? rows_partition_device<uint16_t>(e, (const uint16_t*)e->cols[(size_t)attr], n_segments)
                                       : rows_partition_device<uint8_t>(e, (const uint8_t*)e->cols[(size_t)attr], n_segments))
        return rc;
    std::vector<long long> back(2 * (size_t)n_segments + 1);
    CUDA_TRY(e, cudaMemcpyAsync(back.data(), e->d_part_base.p, back.size() * sizeof(long long), cudaMemcpyDeviceToHost, e->stream));
    if (int rc = sync_and_resolve(e)) return rc;
    for (int s = 0; s < n_segments; ++s) {
        segment_bytes[s] = back[(size_t)s + 1] - back[(size_t)s];
        if (segment_rows) segment_rows[s] = back[(size_t)n_segments + 1 + s];
    }
    const long long total = back[(size_t)n_segments];
    if (!out_text) return AVT_OK;   // sizes only
    if (out_capacity < total) return fail(e, AVT_E_INVALID, "avt_rows_partition: the segments hold %lld bytes, the output buffer %lld", total, (long long)out_capacity);
    if (total) CUDA_TRY(e, cudaMemcpyAsync(out_text, e->d_part_out.p, (size_t)total, cudaMemcpyDeviceToHost, e->stream));
    return sync_and_resolve(e);
}

int avt_forest_init(avt_engine* e, int32_t n_trees, const uint8_t* const* weights,
                    const uint16_t* const* node_ids) {
    if (!e) return AVT_E_INVALID;
    if (e->N == 0) return fail(e, AVT_E_STATE, "avt_forest_init: no data set loaded");
    if (n_trees < 1) return fail(e, AVT_E_INVALID, "avt_forest_init: n_trees %d < 1", n_trees);
    if (int rc = use_device(e)) return rc;
    e->T = n_trees;
    e->any_w = false;
    std::vector<uint8_t> has_w(n_trees, 0);
    if (weights)
        for (int t = 0; t < n_trees; ++t)
            if (weights[t]) { has_w[t] = 1; e->any_w = true; }
    TimerSlot ts;
    if (int rc = timer_open(e, &e->stats.h2d_ms, &ts)) return rc;
    CUDA_TRY(e, e->d_nid.ensure((size_t)n_trees * e->Npad + 32));
    // 0xFFFF everywhere (covers the padding), then node 0 / caller ids for the real rows
    CUDA_TRY(e, cudaMemsetAsync(e->d_nid.p, 0xFF, (size_t)n_trees * e->Npad * 2, e->stream));
    for (int t = 0; t < n_trees; ++t) {
        uint16_t* dst = e->d_nid.p + (size_t)t * e->Npad;
        e->nid_cur = e->d_nid.p;
        if (node_ids && node_ids[t]) CUDA_TRY(e, cudaMemcpyAsync(dst, node_ids[t], (size_t)e->N * 2, cudaMemcpyHostToDevice, e->stream));
        else CUDA_TRY(e, cudaMemsetAsync(dst, 0, (size_t)e->N * 2, e->stream));
    }
    if (e->any_w) {
        CUDA_TRY(e, e->d_w.ensure((size_t)n_trees * e->Npad + 32));
        CUDA_TRY(e, cudaMemsetAsync(e->d_w.p, 0, (size_t)n_trees * e->Npad, e->stream));
        for (int t = 0; t < n_trees; ++t)
            if (has_w[t]) CUDA_TRY(e, cudaMemcpyAsync(e->d_w.p + (size_t)t * e->Npad, weights[t], (size_t)e->N, cudaMemcpyHostToDevice, e->stream));
    }
    if (int rc = upload(e, e->d_has_w, has_w.data(), (size_t)n_trees)) return rc;
    e->h_has_w = has_w;
    if (int rc = timer_close(e, ts)) return rc;
    e->hist_valid = e->scan_valid = false;
    return sync_and_resolve(e);
}

int avt_forest_init_sampled(avt_engine* e, int32_t n_trees, int32_t mode, int64_t seed, int32_t param, int64_t total_rows, int64_t row0) {
    if (!e) return AVT_E_INVALID;
    if (e->N == 0) return fail(e, AVT_E_STATE, "avt_forest_init_sampled: no data set loaded");
    if (n_trees < 1) return fail(e, AVT_E_INVALID, "avt_forest_init_sampled: n_trees %d < 1", n_trees);
    if (mode != AVT_SAMPLE_WITH_REPLACE && mode != AVT_SAMPLE_WITHOUT_REPLACE) return fail(e, AVT_E_INVALID, "avt_forest_init_sampled: mode %d", mode);
    if (param < 1) return fail(e, AVT_E_INVALID, "avt_forest_init_sampled: %s %d < 1", mode == AVT_SAMPLE_WITH_REPLACE ? "batch size" : "sampling rate", param);
    if (row0 < 0 || total_rows < row0 + e->N) return fail(e, AVT_E_INVALID, "avt_forest_init_sampled: rows [%lld, %lld) are not inside a file of %lld rows", (long long)row0, (long long)(row0 + e->N), (long long)total_rows);
    if (total_rows >= (1LL << 46)) return fail(e, AVT_E_INVALID, "avt_forest_init_sampled: more rows than the 48-bit stream has draws");
    if (int rc = use_device(e)) return rc;
    e->T = n_trees;
    e->any_w = true;
    std::vector<uint8_t> has_w(n_trees, 1);
    TimerSlot ts;
    if (int rc = timer_open(e, &e->stats.sample_ms, &ts)) return rc;
    CUDA_TRY(e, e->d_nid.ensure((size_t)n_trees * e->Npad + 32));
    CUDA_TRY(e, e->d_w.ensure((size_t)n_trees * e->Npad + 32));
    CUDA_TRY(e, e->d_err.ensure(1));
    CUDA_TRY(e, cudaMemsetAsync(e->d_err.p, 0, sizeof(int32_t), e->stream));
    e->nid_cur = e->d_nid.p;
    if (e->Npad == e->N) {
        CUDA_TRY(e, cudaMemsetAsync(e->d_nid.p, 0, (size_t)n_trees * e->Npad * 2, e->stream));
    } else {
        CUDA_TRY(e, cudaMemsetAsync(e->d_nid.p, 0xFF, (size_t)n_trees * e->Npad * 2, e->stream));
        for (int t = 0; t < n_trees; ++t) CUDA_TRY(e, cudaMemsetAsync(e->d_nid.p + (size_t)t * e->Npad, 0, (size_t)e->N * 2, e->stream));
    }
    CUDA_TRY(e, cudaMemsetAsync(e->d_w.p, 0, (size_t)n_trees * e->Npad, e->stream));
    BagArgs ba{};
    ba.w = e->d_w.p;
    ba.Npad = e->Npad;
    ba.total_rows = total_rows;
    ba.row0 = row0;
    ba.n_local = e->N;
    ba.seed = seed;
    ba.batch = param;
    ba.err = e->d_err.p;
    if (mode == AVT_SAMPLE_WITH_REPLACE) {
        ba.first_batch = row0 / param;
        const long long last_batch = (row0 + e->N - 1) / param;
        const long long n_batches = last_batch - ba.first_batch + 1;
        if (n_batches > 0x7FFFFFFFLL) return fail(e, AVT_E_INVALID, "avt_forest_init_sampled: too many batches");
        const size_t smem = (size_t)param * sizeof(uint32_t);
        if (smem <= kMaxDynSmem) {
            CUDA_TRY(e, cudaFuncSetAttribute(bagging_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxDynSmem));
            // grid.y is limited to 65535 trees
            for (int t0 = 0; t0 < n_trees; t0 += 65535) {
                BagArgs b2 = ba;
                b2.w = ba.w + (size_t)t0 * e->Npad;
                b2.seed = seed + t0;
                bagging_kernel<<<dim3((unsigned)n_batches, (unsigned)std::min(65535, n_trees - t0)), 512, smem, e->stream>>>(b2);
            }
        } else {
            const unsigned slices = (unsigned)std::min<long long>(1024, ((long long)param + 256 * 64 - 1) / (256 * 64));
            for (int t0 = 0; t0 < n_trees; t0 += 65535) {
                BagArgs b2 = ba;
                b2.w = ba.w + (size_t)t0 * e->Npad;
                b2.seed = seed + t0;
                bagging_global_kernel<<<dim3((unsigned)n_batches, (unsigned)std::min(65535, n_trees - t0), slices), 256, 0, e->stream>>>(b2);
            }
        }
    } else {
        const int rpt = 64;
        const long long threads = (e->N + rpt - 1) / rpt;
        for (int t0 = 0; t0 < n_trees; t0 += 65535) {
            BagArgs b2 = ba;
            b2.w = ba.w + (size_t)t0 * e->Npad;
            b2.seed = seed + t0;
            bernoulli_kernel<<<dim3((unsigned)((threads + 255) / 256), (unsigned)std::min(65535, n_trees - t0)), 256, 0, e->stream>>>(b2, param, rpt);
        }
    }
    CUDA_TRY(e, cudaGetLastError());
    if (int rc = upload(e, e->d_has_w, has_w.data(), (size_t)n_trees)) return rc;
    e->h_has_w = has_w;
    int32_t err = 0;
    CUDA_TRY(e, cudaMemcpyAsync(&err, e->d_err.p, sizeof(int32_t), cudaMemcpyDeviceToHost, e->stream));
    if (int rc = timer_close(e, ts)) return rc;
    e->hist_valid = e->scan_valid = false;
    if (int rc = sync_and_resolve(e)) return rc;
    if (err) {
        e->T = 0;
        return fail(e, AVT_E_INVALID, "avt_forest_init_sampled: a bootstrap multiplicity exceeds 255");
    }
    return AVT_OK;
}

int avt_forest_reset(avt_engine* e) {
    if (!e) return AVT_E_INVALID;
    if (e->T == 0) return fail(e, AVT_E_STATE, "avt_forest_reset: no forest (avt_forest_init / avt_forest_init_sampled first)");
    if (int rc = use_device(e)) return rc;
    e->nid_cur = e->d_nid.p;
    if (e->Npad == e->N) {
        CUDA_TRY(e, cudaMemsetAsync(e->d_nid.p, 0, (size_t)e->T * e->Npad * 2, e->stream));
    } else {
        CUDA_TRY(e, cudaMemsetAsync(e->d_nid.p, 0xFF, (size_t)e->T * e->Npad * 2, e->stream));
        for (int t = 0; t < e->T; ++t) CUDA_TRY(e, cudaMemsetAsync(e->d_nid.p + (size_t)t * e->Npad, 0, (size_t)e->N * 2, e->stream));
    }
    e->hist_valid = e->scan_valid = false;
    e->NF = 0;
    return AVT_OK;
}

int avt_forest_weights(avt_engine* e, int32_t tree, uint8_t* out) {
    if (!e) return AVT_E_INVALID;
    if (tree < 0 || tree >= e->T || !out) return fail(e, AVT_E_INVALID, "avt_forest_weights: tree %d out of range", tree);
    if (int rc = use_device(e)) return rc;
    if (!e->any_w || !e->h_has_w[(size_t)tree]) {
        memset(out, 1, (size_t)e->N);
        return AVT_OK;
    }
    CUDA_TRY(e, cudaMemcpyAsync(out, e->d_w.p + (size_t)tree * e->Npad, (size_t)e->N, cudaMemcpyDeviceToHost, e->stream));
    return sync_and_resolve(e);
}

int avt_split_table_set(avt_engine* e, int32_t n_splits, const avt_split* splits, const int32_t* bounds,
                        int32_t n_bounds, const uint8_t* seg_lut, int64_t n_lut) {
    if (!e) return AVT_E_INVALID;
    if (e->N == 0) return fail(e, AVT_E_STATE, "avt_split_table_set: no data set loaded");
    if (n_splits < 0 || (n_splits && !splits)) return fail(e, AVT_E_INVALID, "avt_split_table_set: bad split table");
    if (int rc = use_device(e)) return rc;
    std::vector<int32_t> attr_first(e->A + 1, 0);
    int prev = 0;
    for (int s = 0; s < n_splits; ++s) {
        const avt_split& sp = splits[s];
        if (sp.attr < 0 || sp.attr >= e->A) return fail(e, AVT_E_INVALID, "split %d: attribute %d out of range", s, sp.attr);
        if (sp.attr < prev) return fail(e, AVT_E_INVALID, "split %d: splits must be grouped by non-decreasing attribute", s);
        if (sp.n_seg < 2 || sp.n_seg > AVT_MAX_SEG) return fail(e, AVT_E_INVALID, "split %d: n_seg %d outside 2..%d", s, sp.n_seg, AVT_MAX_SEG);
        if (sp.kind == AVT_SPLIT_NUMERIC) {
            if (sp.off < 0 || sp.off + sp.n_seg - 1 > n_bounds) return fail(e, AVT_E_INVALID, "split %d: bounds out of range", s);
            for (int j = 0; j < sp.n_seg - 1; ++j) {
                int b = bounds[sp.off + j];
                if (b < -1 || b >= e->n_bins[sp.attr]) return fail(e, AVT_E_INVALID, "split %d: bound %d outside the attribute's bins", s, b);
            }
        } else if (sp.kind == AVT_SPLIT_RANGES) {
            if (sp.off < 0 || sp.off + 2 * sp.n_seg > n_bounds) return fail(e, AVT_E_INVALID, "split %d: ranges out of range", s);
            for (int j = 0; j < sp.n_seg; ++j) {
                const int lo = bounds[sp.off + 2 * j], hi = bounds[sp.off + 2 * j + 1];
                if (lo < 0 || hi >= e->n_bins[sp.attr]) return fail(e, AVT_E_INVALID, "split %d: range [%d, %d] outside the attribute's bins", s, lo, hi);
            }
        } else if (sp.kind == AVT_SPLIT_LUT) {
            if (sp.off < 0 || (int64_t)sp.off + e->n_bins[sp.attr] > n_lut) return fail(e, AVT_E_INVALID, "split %d: segment table out of range", s);
        } else {
            return fail(e, AVT_E_INVALID, "split %d: unknown kind %d", s, sp.kind);
        }
        for (int a = prev + 1; a <= sp.attr; ++a) attr_first[a] = s;
        prev = sp.attr;
    }
    for (int a = prev + 1; a <= e->A; ++a) attr_first[a] = n_splits;
    e->S = n_splits;
    e->splits.assign(splits, splits + n_splits);
    e->g_max_splits = 2;
    for (int i = 0; i < n_splits; ++i) e->g_max_splits = std::max(e->g_max_splits, splits[i].n_seg);
    e->h_bounds.assign(bounds, bounds + std::max(n_bounds, 0));
    e->h_seg_lut.assign(seg_lut, seg_lut + std::max<int64_t>(n_lut, 0));
    if (int rc = upload(e, e->d_splits, splits, (size_t)n_splits)) return rc;
    if (int rc = upload(e, e->d_bounds, bounds, (size_t)std::max(n_bounds, 0))) return rc;
    if (int rc = upload(e, e->d_seg_lut, seg_lut, (size_t)std::max<int64_t>(n_lut, 0))) return rc;
    if (int rc = upload(e, e->d_attr_first, attr_first.data(), attr_first.size())) return rc;
    {
        // tables of the item-parallel scan, and whether this split table fits it (limits: avt_scan_fast.cuh)
        std::vector<uint32_t> items;
        std::vector<int32_t> seg_first((size_t)n_splits + 1, 0);
        bool ok = e->A <= kScanWarps && n_splits >= 1 && n_splits <= kScanThreads;
        for (int s = 0; s < n_splits; ++s) {
            seg_first[(size_t)s] = (int32_t)items.size();
            for (int j = 0; j < splits[s].n_seg; ++j) items.push_back(((uint32_t)s << 4) | (uint32_t)j);
            const int B = e->n_bins[splits[s].attr];
            ok = ok && (long long)B * e->C <= kFastWordsPerAttr && (splits[s].kind == AVT_SPLIT_NUMERIC || (splits[s].kind == AVT_SPLIT_LUT && B <= 64));
        }
        seg_first[(size_t)n_splits] = (int32_t)items.size();
        ok = ok && items.size() <= (size_t)kFastItems && items.size() * (size_t)e->C <= (size_t)kFastCntWords;
        e->n_items = (int)items.size();
        e->fast_scan_ok = ok;
        if (int rc = upload(e, e->d_items, items.data(), items.size())) return rc;
        if (int rc = upload(e, e->d_seg_first, seg_first.data(), seg_first.size())) return rc;
    }
    e->scan_valid = false;
    return sync_and_resolve(e);
}

// Repartition state folded into a histogram pass (device-driven level loop).
struct FusedRepart {
    const int32_t* prev_node_base;
    const int32_t* node_attr;
    const uint16_t* child_lut;
    int lut_w;
    int prev_n_max;  // bound on the previous frontier's nodes per tree (sizes the shared-memory copy of the tables)
    const uint16_t* lut2;  // packed byte -> child rows of the previous frontier written by its scan kernel (nullptr: none)
};

// Plans and launches kernel (a) for the frontier in e->node_base (already uploaded to d_node_base).
// With `fr` the previous level's repartition is applied on the fly (node ids ping-pong between the two
// buffers); without it the node ids are read in place.
static int launch_histogram(avt_engine* e, int n_max, const FusedRepart* fr, const int32_t* d_node_base = nullptr) {
    const size_t hwords = (size_t)e->NF * e->TB * e->C;
    if (!e->H_cur) CUDA_TRY(e, e->d_H.ensure(hwords));
    uint32_t* const Hbuf = e->H_cur ? e->H_cur : e->d_H.p;
    // ---- plan: attribute groups whose per-tree histogram fits the shared-memory budget
    const long long per_bin = (long long)(std::max(n_max, 1) + 1) * e->C;  // words per bin: all nodes of a tree + the phantom node
    int max_bins = 0;
    bool any_wide = false;
    for (int a = 0; a < e->A; ++a) {
        max_bins = std::max(max_bins, e->n_bins[a]);
        any_wide = any_wide || e->wide[a];
    }
    int variant = (per_bin * max_bins <= kHistWordsBudget) ? 0 : 1;
    // forest kernel (avt_forest.cuh: hist_rot_kernel): wants the rotated row-major copy of the codes (built for data sets with >= 16
    // uint8 attributes) and room for at least one tree's counted nodes in shared memory
    int f_slots = std::max(n_max, 1), f_G = 0;
    size_t f_smem = 0;
    {
        if (e->skip_cur) f_slots = std::max(1, n_max - n_max / std::max(2, e->g_max_splits));   // one child of every expanded parent is derived, not counted
        const size_t n_prev = fr ? (size_t)std::max(fr->prev_n_max, 1) : 0;
        const size_t per_tree = (size_t)f_slots * e->C * max_bins * 128 + n_prev * ((size_t)(fr ? fr->lut_w : 0) * 4 + 2) + ((size_t)n_max + 2 + f_slots) * 2 + 1;
        const size_t fixed = (size_t)e->C * max_bins * 128 /* phantom slot */ + (fr ? (size_t)kForestWarps * 1024 : 0) + 64;
        if (e->rowrot_ok && max_bins <= 256 && per_tree + fixed <= kMaxDynSmem) {
            f_G = (int)std::min<size_t>((size_t)e->T, (kMaxDynSmem - fixed) / per_tree);
            f_G = (int)std::min<long long>(f_G, std::max<long long>(1, 0xFFFFFFFFLL / std::max<long long>(e->Npad, 1)));   // 32-bit element offsets inside a tree group
            f_smem = fixed + per_tree * (size_t)f_G;
        }
        if (f_G >= 1 && e->force_variant < 0 && variant == 0) variant = 2;
    }
    // sorted-rows kernel (avt_sorted.cuh): one tree, the rotated copy, a frontier whose counters hist_rot_kernel cannot hold
    int s_S = 0;
    {
        const size_t per_node = (size_t)e->C * max_bins * 128;
        const bool usable = e->T == 1 && e->rowrot_ok && max_bins <= 256 && e->N < (1LL << 31) && per_node + 18 * 1024 <= kMaxDynSmem && !e->signal_pending;
        // two blocks per SM: each has 17 KB of static shared memory (the warps' row tiles) beside its window of S nodes
        if (usable) s_S = (int)std::min<size_t>(std::min<size_t>(std::max<size_t>((kMaxDynSmem / 2 - 18 * 1024) / per_node, 1), 8), (size_t)std::max(n_max, 1));
        if (s_S >= 1 && ((e->force_variant < 0 && f_G == 0 && (e->hist_sorted == 2 || (e->hist_sorted == 1 && variant == 1))) || e->force_variant == 4)) variant = 4;
        else s_S = 0;
    }
    // joint-code kernel (avt_joint.cuh): the data set has a packed copy (<= 4 narrow attributes in one byte) and one block can hold
    // the frontier's joint counters; it has no form for skip marks and does not signal peers itself
    int j_J = 0;
    size_t j_smem = 0;
    if (e->packed_ok && (e->hist_joint || e->force_variant == 3) && !e->skip_cur && (!e->signal_pending || (e->peer_signal_scan && e->scan_will_signal))) {
        int maxv = 0;
        for (int a = 0; a < e->A; ++a) maxv += (e->n_bins[a] - 1) << e->pk_shift[a];
        j_J = maxv + 1;
        const size_t n_prev = fr ? (size_t)std::max(fr->prev_n_max, 1) : 0;
        j_smem = ((size_t)std::max(n_max, 1) + 1) * j_J * e->C * 4 + (fr ? (n_prev + 1) * j_J * 2 : 0) + 16;
        if (j_smem > kMaxDynSmem) j_J = 0;
        else if (e->force_variant < 0 && variant == 0) variant = 3;
    }
    if (e->force_variant >= 0) {
        if (e->force_variant == 3 && j_J == 0)
            return fail(e, AVT_E_INVALID, "avt_level_histogram: joint-code variant forced but it needs a packed data set (<= 4 narrow attributes) and the frontier's joint counters in shared memory");
        if (e->force_variant == 0 && variant == 1)
            return fail(e, AVT_E_INVALID, "avt_level_histogram: shared-memory variant forced but the frontier needs %lld words", per_bin * max_bins);
        if (e->force_variant == 4 && s_S == 0)
            return fail(e, AVT_E_INVALID, "avt_level_histogram: sorted-rows variant forced but it needs one tree, >= 16 uint8 attributes and two nodes' counters in shared memory");
        if (e->force_variant == 2 && f_G == 0)
            return fail(e, AVT_E_INVALID, "avt_level_histogram: forest variant forced but it needs >= 16 attributes, uint8 codes and one tree's counters in shared memory");
        variant = e->force_variant;
    }
    uint16_t* nid_in = e->nid_cur;
    uint16_t* nid_out = nullptr;
    if (fr) {
        CUDA_TRY(e, e->d_nid2.ensure((size_t)e->T * e->Npad + 32));
        nid_out = (e->nid_cur == e->d_nid.p) ? e->d_nid2.p : e->d_nid.p;
    }
    if (fr && (variant == 1 || variant == 4)) {
        // the global-atomics fallback and the sorted-rows kernel have no fused form: repartition in place first
        TimerSlot tr;
        if (int rc = timer_open(e, &e->stats.repart_ms, &tr)) return rc;
        long long blocks = std::min<long long>((e->N + 256 * 8 - 1) / (256 * 8), (long long)e->sm_count * 8);
        dim3 grid((unsigned)blocks, (unsigned)e->T);
        repartition_kernel<<<grid, 256, 0, e->stream>>>(e->nid_cur, (const void* const*)e->d_colptrs.p, e->d_wide.p, fr->prev_node_base,
                                                        fr->node_attr, fr->child_lut, fr->lut_w, e->N, e->Npad);
        CUDA_TRY(e, cudaGetLastError());
        if (int rc = timer_close(e, tr)) return rc;
        e->stats.repart_launches += 1;
        e->stats.repart_alg_bytes += (long long)e->N * e->T * 5;
        fr = nullptr;
        nid_out = nullptr;
    }
    HistArgs ha{};
    ha.cols = (const void* const*)e->d_colptrs.p;
    ha.cls = e->d_cls.p;
    ha.nid = nid_in;
    ha.wts = e->any_w ? e->d_w.p : nullptr;
    ha.has_w = e->d_has_w.p;
    ha.H = Hbuf;
    ha.node_base = d_node_base ? d_node_base : e->d_node_base.p;
    ha.N = e->N;
    ha.Npad = e->Npad;
    ha.C = e->C;
    ha.TB = e->TB;
    ha.T = e->T;
    ha.wide = e->d_wide.p;
    ha.skip = e->skip_cur;
    if (fr) {
        ha.nid_out = nid_out;
        ha.prev_node_base = fr->prev_node_base;
        ha.node_attr = fr->node_attr;
        ha.child_lut = fr->child_lut;
        ha.lut_w = fr->lut_w;
    }

    TimerSlot ts;
    ts.acc = nullptr;
    if (int rc = timer_open(e, &e->stats.hist_ms, &ts)) return rc;
    if (!e->skip_h_memset) CUDA_TRY(e, cudaMemsetAsync(Hbuf, 0, hwords * sizeof(uint32_t), e->stream));
    if (variant == 0) {
        // Replication policy: enough copies that a warp's 32 updates spread over >= kTargetAddrs distinct
        // words, but no more -- smaller histograms mean more resident blocks per SM (latency hiding).
        const int bins_budget = (int)(e->hist_words_budget / per_bin);
        std::vector<int32_t> ent_attr, ent_loc, ent_glob, ent_wide, grp_start{0}, grp_bins, grp_rlog, l2g, l2g_off;
        int cur = 0;
        for (int a = 0; a < e->A; ++a) {
            if (cur && cur + e->n_bins[a] > bins_budget) {
                grp_bins.push_back(cur);
                grp_start.push_back((int)ent_attr.size());
                cur = 0;
            }
            if (cur == 0) l2g_off.push_back((int)l2g.size());
            ent_attr.push_back(a);
            ent_loc.push_back(cur);
            ent_glob.push_back(e->bin_off[a]);
            ent_wide.push_back(e->wide[a]);
            for (int b = 0; b < e->n_bins[a]; ++b) l2g.push_back(e->bin_off[a] + b);
            cur += e->n_bins[a];
        }
        grp_bins.push_back(cur);
        grp_start.push_back((int)ent_attr.size());
        const int G = (int)grp_bins.size();
        size_t smem_words = 0;
        int rlog_min = 5;
        for (int g = 0; g < G; ++g) {
            const long long keys = per_bin * grp_bins[g];
            int rlog = 0;
            while (rlog < 5 && (keys << (rlog + 1)) <= e->hist_words_budget && (keys << rlog) < e->hist_target_addrs) ++rlog;
            grp_rlog.push_back(rlog);
            rlog_min = std::min(rlog_min, rlog);
            smem_words = std::max(smem_words, (size_t)(keys << rlog));
        }
        std::vector<int32_t> pack;
        auto put = [&](const std::vector<int32_t>& v) {
            size_t o = pack.size();
            pack.insert(pack.end(), v.begin(), v.end());
            return o;
        };
        const size_t o_attr = put(ent_attr), o_loc = put(ent_loc), o_glob = put(ent_glob), o_wide = put(ent_wide),
                     o_gs = put(grp_start), o_gb = put(grp_bins), o_gr = put(grp_rlog), o_l2g = put(l2g), o_lo = put(l2g_off);
        // The tables depend only on (data set, frontier bound): keep every distinct one in a device arena so
        // that steady-state levels issue no host->device copy at all.
        constexpr size_t kGrpArenaInts = 1u << 20;
        if (e->d_grp.cap < kGrpArenaInts) {
            CUDA_TRY(e, e->d_grp.ensure(kGrpArenaInts));
            e->grp_cache.clear();
            e->grp_used = 0;
        }
        size_t at;
        auto hit = e->grp_cache.find(n_max);
        if (hit != e->grp_cache.end()) {
            at = hit->second;
        } else {
            if (e->grp_used + pack.size() > e->d_grp.cap) {  // arena full: start over (in-flight launches are stream-ordered before the overwrite)
                e->grp_cache.clear();
                e->grp_used = 0;
            }
            at = e->grp_used;
            CUDA_TRY(e, cudaMemcpyAsync(e->d_grp.p + at, pack.data(), pack.size() * sizeof(int32_t), cudaMemcpyHostToDevice, e->stream));
            e->grp_cache[n_max] = at;
            e->grp_used += pack.size();
        }
        const int32_t* base = e->d_grp.p + at;
        ha.ent_attr = base + o_attr; ha.ent_locoff = base + o_loc; ha.ent_globoff = base + o_glob; ha.ent_wide = base + o_wide;
        ha.grp_start = base + o_gs; ha.grp_bins = base + o_gb; ha.grp_rlog = base + o_gr; ha.grp_l2g = base + o_l2g; ha.grp_l2g_off = base + o_lo;
        ha.codes_in_regs = (G == 1 && !any_wide && e->A <= 4) ? 1 : 0;
        ha.zero_words = (int)smem_words;
        if (ha.codes_in_regs && e->packed_ok) {
            ha.packed = e->d_packed.p;
            for (int q = 0; q < 4; ++q) { ha.pk_shift[q] = e->pk_shift[q]; ha.pk_mask4[q] = e->pk_mask4[q]; }
        }
        ha.packed_bins = 1;
        for (int g = 0; g < G; ++g) ha.packed_bins = ha.packed_bins && grp_bins[(size_t)g] <= 256;
        size_t smem = smem_words * sizeof(uint32_t);
        if (fr && ha.codes_in_regs) {
            // previous frontier (bounded like this one) + phantom entry: selector word + one bin->child row each
            const size_t tbl = ((size_t)fr->prev_n_max + 2) * (4 + (size_t)fr->lut_w * 2) + 8;
            if (tbl <= 16 * 1024 && smem + tbl <= kMaxDynSmem) {
                ha.repart_smem = 1;
                smem += tbl;
            }
        }
        int bpsm = (int)std::min<size_t>((size_t)e->hist_max_bpsm, std::max<size_t>(1, (kMaxDynSmem + 1024) / (smem + 1024)));
        long long resident = (long long)e->sm_count * bpsm;
        long long chunks = std::max<long long>(1, resident / ((long long)e->T * G));
        const long long tiles = (e->N + (long long)kHistThreads * kRowsPerThread - 1) / ((long long)kHistThreads * kRowsPerThread);
        chunks = std::min(chunks, tiles);
        chunks = std::min<long long>(chunks, 65535);
        if (ha.skip) smem += (size_t)n_max + 16;   // skip marks of the frontier, staged behind the histogram (and the repartition tables)
        dim3 grid((unsigned)e->T, (unsigned)chunks, (unsigned)G);
        e->signal_in_hist = false;
        if (e->signal_pending && e->peer_signal_scan && e->scan_will_signal) {
            e->signal_in_hist = true;   // nothing to do here: the scan kernel behind this pass signals (PlanArgs::signal_flags)
        } else if (e->signal_pending) {
            CUDA_TRY(e, e->d_ticket2.ensure(1));
            if (!e->ticket2_zeroed) {
                CUDA_TRY(e, cudaMemsetAsync(e->d_ticket2.p, 0, sizeof(uint32_t), e->stream));
                e->ticket2_zeroed = true;
            }
            ha.signal_flags = e->d_remote_flags.p;
            ha.signal_ticket = e->d_ticket2.p;
            ha.signal_epoch = e->peer.epoch;
            ha.signal_ranks = e->peer.n;
            e->signal_in_hist = true;
        }
        const int sel = (e->any_w ? 4 : 0) | (any_wide ? 2 : 0) | (fr ? 1 : 0);
        void (*kern)(const HistArgs) = nullptr;
        switch (sel) {
            case 0: kern = hist_smem_kernel<false, false, false>; break;
            case 1: kern = hist_smem_kernel<false, false, true>; break;
            case 2: kern = hist_smem_kernel<false, true, false>; break;
            case 3: kern = hist_smem_kernel<false, true, true>; break;
            case 4: kern = hist_smem_kernel<true, false, false>; break;
            case 5: kern = hist_smem_kernel<true, false, true>; break;
            case 6: kern = hist_smem_kernel<true, true, false>; break;
            default: kern = hist_smem_kernel<true, true, true>; break;
        }
        {
            cudaLaunchConfig_t cfg{};
            cfg.gridDim = grid;
            cfg.blockDim = dim3(kHistThreads);
            cfg.dynamicSmemBytes = smem;
            cfg.stream = e->stream;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[0].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = attr;
            cfg.numAttrs = (e->pdl && e->pdl_chain) ? 1 : 0;   // only directly behind another kernel of the level loop
            CUDA_TRY(e, cudaLaunchKernelEx(&cfg, kern, ha));
        }
        e->stats.last_hist_replicas = 1 << rlog_min;
        e->stats.last_hist_grid = (int)(grid.x * grid.y * grid.z);
        e->stats.last_hist_block = kHistThreads;
        e->stats.last_hist_smem = (int)smem;
    } else if (variant == 3) {
        JointArgs ja{};
        ja.bin_off = e->d_bin_off.p;
        ja.A = e->A;
        ja.J = j_J;
        ja.n_prev_max = fr ? std::max(fr->prev_n_max, 1) : 1;
        ja.zero_words = (std::max(n_max, 1) + 1) * j_J * e->C;
        ja.lut2_g = fr ? fr->lut2 : nullptr;
        for (int q = 0; q < 4; ++q) {
            ja.shift[q] = e->pk_shift[q];
            ja.mask[q] = e->pk_mask4[q] & 0xFFu;
        }
        ha.packed = e->d_packed.p;
        // the per-block table build and flush grow with the frontier: deep levels run one block per SM
        const int bpsm = n_max > 32 ? 1 : (int)std::min<size_t>((size_t)e->hist_max_bpsm, std::max<size_t>(1, (kMaxDynSmem + 1024) / (j_smem + 1024)));
        long long chunks = std::max<long long>(1, (long long)e->sm_count * bpsm / e->T);
        const long long tiles = (e->N + (long long)kJointThreads * kRowsPerThread - 1) / ((long long)kJointThreads * kRowsPerThread);
        chunks = std::min(std::min(chunks, tiles), 65535LL);
        dim3 grid((unsigned)e->T, (unsigned)chunks);
        void (*kern)(const HistArgs, const JointArgs) = e->any_w ? (fr ? hist_joint_kernel<true, true> : hist_joint_kernel<true, false>)
                                                                 : (fr ? hist_joint_kernel<false, true> : hist_joint_kernel<false, false>);
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = grid;
        cfg.blockDim = dim3(kJointThreads);
        cfg.dynamicSmemBytes = j_smem;
        cfg.stream = e->stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        cfg.numAttrs = (e->pdl && e->pdl_chain) ? 1 : 0;
        CUDA_TRY(e, cudaLaunchKernelEx(&cfg, kern, ha, ja));
        e->signal_in_hist = e->signal_pending;   // (only reached when the scan kernel behind this pass signals the peers)
        e->stats.last_hist_replicas = 1;
        e->stats.last_hist_grid = (int)(grid.x * grid.y);
        e->stats.last_hist_block = kJointThreads;
        e->stats.last_hist_smem = (int)j_smem;
    } else if (variant == 2) {
        ForestArgs fa{};
        fa.bin_off = e->d_bin_off.p;
        fa.A = e->A;
        fa.G = f_G;
        fa.n_slots = f_slots;
        fa.Bmax = max_bins;
        fa.n_nodes_max = std::max(n_max, 1);
        fa.n_prev_max = fr ? std::max(fr->prev_n_max, 1) : 1;
        if (!e->hist_err) CUDA_TRY(e, e->d_err.ensure(1));
        fa.err = e->hist_err ? e->hist_err : e->d_err.p;
        const int groups = (e->T + f_G - 1) / f_G, achunks = (e->A + 31) / 32;
        const long long n_tiles = (e->N + 31) / 32;
        // grid.x = tree group (fastest): blocks that run together read the same rows, so the codes come out of L2 once per group
        long long parts = std::max<long long>(1, ((long long)e->sm_count * 4 + (long long)groups * achunks - 1) / ((long long)groups * achunks));
        long long tpp = (n_tiles + parts - 1) / parts;
        tpp = std::max<long long>((tpp + kForestWarps - 1) / kForestWarps * kForestWarps, (long long)kForestWarps * 8);
        parts = std::min<long long>((n_tiles + tpp - 1) / tpp, 65535);
        tpp = (n_tiles + parts - 1) / parts;
        fa.tiles_per_part = tpp;
        dim3 grid((unsigned)groups, (unsigned)parts, (unsigned)achunks);
        fa.rowrot = e->d_rowrot.p;
        fa.pf = e->rot_pf;
        fa.rowrot_rows = e->rowrot_rows;
        void (*kern)(const HistArgs, const ForestArgs) = e->any_w ? (fr ? hist_rot_kernel<true, true> : hist_rot_kernel<true, false>)
                                                                  : (fr ? hist_rot_kernel<false, true> : hist_rot_kernel<false, false>);
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = grid;
        cfg.blockDim = dim3(kForestThreads);
        cfg.dynamicSmemBytes = f_smem;
        cfg.stream = e->stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        cfg.numAttrs = (e->pdl && e->pdl_chain) ? 1 : 0;
        CUDA_TRY(e, cudaLaunchKernelEx(&cfg, kern, ha, fa));
        e->signal_in_hist = false;
        e->stats.last_hist_replicas = f_G;
        e->stats.last_hist_grid = (int)(grid.x * grid.y * grid.z);
        e->stats.last_hist_block = kForestThreads;
        e->stats.last_hist_smem = (int)f_smem;
    } else if (variant == 4) {
        // node-contiguous row order: radix sort of (node id, row) pairs, class / multiplicity copies in that order, then spans
        const long long N = e->N;
        CUDA_TRY(e, e->d_skeys.ensure((size_t)N));
        CUDA_TRY(e, e->d_sperm.ensure((size_t)N));
        CUDA_TRY(e, e->d_scls.ensure((size_t)N));
        if (e->any_w) CUDA_TRY(e, e->d_sw.ensure((size_t)N));
        const unsigned gb = (unsigned)((N + 255) / 256);
        if (e->siota_rows != N) {
            CUDA_TRY(e, e->d_siota.ensure((size_t)N));
            iota_kernel<<<gb, 256, 0, e->stream>>>(e->d_siota.p, N);
            e->siota_rows = N;
        }
        size_t tmp_bytes = 0;
        CUDA_TRY(e, cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, (const uint16_t*)nid_in, e->d_skeys.p, (const uint32_t*)e->d_siota.p, e->d_sperm.p, (int)N, 0, 16, e->stream));
        CUDA_TRY(e, e->d_stmp.ensure(tmp_bytes));
        CUDA_TRY(e, cub::DeviceRadixSort::SortPairs(e->d_stmp.p, tmp_bytes, (const uint16_t*)nid_in, e->d_skeys.p, (const uint32_t*)e->d_siota.p, e->d_sperm.p, (int)N, 0, 16, e->stream));
        sorted_gather_kernel<<<gb, 256, 0, e->stream>>>(e->d_sperm.p, e->d_cls.p, e->any_w ? e->d_w.p : nullptr, N, e->d_scls.p, e->d_sw.p);
        SortedArgs sa{};
        sa.keys = e->d_skeys.p; sa.perm = e->d_sperm.p; sa.cls_s = e->d_scls.p; sa.w_s = e->d_sw.p;
        sa.rowrot = e->d_rowrot.p; sa.rowrot_rows = e->rowrot_rows;
        sa.bin_off = e->d_bin_off.p; sa.skip = e->skip_cur; sa.H = Hbuf;
        sa.N = N; sa.A = e->A; sa.Bmax = max_bins; sa.S = s_S; sa.TB = e->TB; sa.C = e->C;
        const int achunks = (e->A + 31) / 32;
        const size_t smem = (size_t)s_S * e->C * max_bins * 128;
        const int bpsm = (int)std::max<size_t>(1, std::min<size_t>(4, (kMaxDynSmem + 1024) / (smem + 18 * 1024)));
        // spans: about four waves of blocks over all chunks, at least 2048 positions each
        long long spans = std::max<long long>(1, ((long long)e->sm_count * bpsm * 4 + achunks - 1) / achunks);
        sa.span = std::max<long long>(2048, (N + spans - 1) / spans);
        sa.span = (sa.span + 31) / 32 * 32;
        spans = (N + sa.span - 1) / sa.span;
        dim3 grid((unsigned)spans, (unsigned)achunks);
        if (e->any_w) {
            CUDA_TRY(e, cudaFuncSetAttribute(hist_sorted_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kMaxDynSmem - 18 * 1024)));
            hist_sorted_kernel<true><<<grid, kSortedThreads, smem, e->stream>>>(sa);
        } else {
            CUDA_TRY(e, cudaFuncSetAttribute(hist_sorted_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kMaxDynSmem - 18 * 1024)));
            hist_sorted_kernel<false><<<grid, kSortedThreads, smem, e->stream>>>(sa);
        }
        e->stats.last_hist_replicas = s_S;
        e->stats.last_hist_grid = (int)(grid.x * grid.y);
        e->stats.last_hist_block = kSortedThreads;
        e->stats.last_hist_smem = (int)smem;
    } else {
        long long blocks = std::min<long long>((e->N + 255) / 256, (long long)e->sm_count * 8);
        dim3 grid((unsigned)blocks, (unsigned)e->T);
        hist_global_kernel<<<grid, 256, 0, e->stream>>>(ha, e->A, e->d_bin_off.p, e->d_wide.p);
        e->stats.last_hist_replicas = 1;
        e->stats.last_hist_grid = (int)(grid.x * grid.y);
        e->stats.last_hist_block = 256;
        e->stats.last_hist_smem = 0;
    }
    CUDA_TRY(e, cudaGetLastError());
    if (int rc = timer_close(e, ts)) return rc;
    if (fr) e->nid_cur = nid_out;
    e->stats.last_hist_variant = variant;
    e->stats.hist_launches += 1;
    {
        long long code_bytes = 0;
        for (int a = 0; a < e->A; ++a) code_bytes += e->wide[a] ? 2 : 1;
        long long per_row = code_bytes + 1;
        for (int t = 0; t < e->T; ++t) per_row += 2 + ((e->any_w) ? 1 : 0) + (fr ? 2 : 0);  // fused: + the node-id write
        e->stats.hist_alg_bytes += per_row * e->N;
    }
    if (e->comm && !e->signal_pending) {
        TimerSlot tc;
        if (int rc = timer_open(e, &e->stats.comm_ms, &tc)) return rc;
        NCCL_TRY(e, ncclAllReduce(Hbuf, Hbuf, hwords, ncclUint32, ncclSum, e->comm, e->stream));
        if (int rc = timer_close(e, tc)) return rc;
        e->stats.comm_calls += 1;
    }
    e->hist_valid = true;
    return AVT_OK;
}

static int set_frontier(avt_engine* e, const int32_t* nodes_per_tree, int* n_max_out, const char* who) {
    e->node_base.assign(e->T + 1, 0);
    int n_max = 0;
    for (int t = 0; t < e->T; ++t) {
        if (nodes_per_tree[t] < 0 || nodes_per_tree[t] >= (int)AVT_NODE_NONE)
            return fail(e, AVT_E_INVALID, "%s: tree %d has %d frontier nodes (limit %d)", who, t, nodes_per_tree[t], AVT_NODE_NONE - 1);
        e->node_base[t + 1] = e->node_base[t] + nodes_per_tree[t];
        n_max = std::max(n_max, nodes_per_tree[t]);
    }
    e->NF = e->node_base[e->T];
    e->hist_valid = e->scan_valid = false;
    *n_max_out = n_max;
    return AVT_OK;
}

int avt_level_histogram(avt_engine* e, const int32_t* nodes_per_tree, uint32_t* hist_out) {
    if (!e) return AVT_E_INVALID;
    if (e->T == 0) return fail(e, AVT_E_STATE, "avt_level_histogram: avt_forest_init has not been called");
    if (!nodes_per_tree) return fail(e, AVT_E_INVALID, "avt_level_histogram: nodes_per_tree is NULL");
    if (int rc = use_device(e)) return rc;
    int n_max = 0;
    if (int rc = set_frontier(e, nodes_per_tree, &n_max, "avt_level_histogram")) return rc;
    if (e->NF == 0) return fail(e, AVT_E_INVALID, "avt_level_histogram: empty frontier");
    if (int rc = upload(e, e->d_node_base, e->node_base.data(), e->node_base.size())) return rc;
    if (int rc = launch_histogram(e, n_max, nullptr)) return rc;
    if (hist_out) {
        const size_t hwords = (size_t)e->NF * e->TB * e->C;
        CUDA_TRY(e, cudaMemcpyAsync(hist_out, e->d_H.p, hwords * sizeof(uint32_t), cudaMemcpyDeviceToHost, e->stream));
        return sync_and_resolve(e);
    }
    return AVT_OK;
}

int avt_level_scan(avt_engine* e, int32_t criterion, const uint32_t* cand_mask, const uint8_t* class_order,
                   avt_node_best* out) {
    if (!e) return AVT_E_INVALID;
    if (!e->hist_valid) return fail(e, AVT_E_STATE, "avt_level_scan: no histogram for the current frontier");
    if (e->S == 0) return fail(e, AVT_E_STATE, "avt_level_scan: no split table installed");
    if (criterion != AVT_CRIT_GINI && criterion != AVT_CRIT_ENTROPY) return fail(e, AVT_E_INVALID, "avt_level_scan: criterion %d", criterion);
    if (!out) return fail(e, AVT_E_INVALID, "avt_level_scan: out is NULL");
    if (int rc = use_device(e)) return rc;
    if (int rc = upload_order(e, class_order)) return rc;
    const int mw = (e->A + 31) / 32;
    const uint32_t* dmask = nullptr;
    if (cand_mask) {
        if (int rc = upload(e, e->d_mask, cand_mask, (size_t)e->NF * mw)) return rc;
        dmask = e->d_mask.p;
    }
    CUDA_TRY(e, e->d_av.ensure((size_t)e->NF * e->S));
    CUDA_TRY(e, e->d_best.ensure((size_t)e->NF));
    TimerSlot ts;
    if (int rc = timer_open(e, &e->stats.scan_ms, &ts)) return rc;
    scan_kernel<<<e->NF, kScanThreads, 0, e->stream>>>(e->d_H.p, make_table(e), dmask, e->d_order.p, criterion, e->d_av.p, e->d_best.p, nullptr, nullptr, PlanArgs{});
    CUDA_TRY(e, cudaGetLastError());
    if (int rc = timer_close(e, ts)) return rc;
    e->stats.scan_launches += 1;
    CUDA_TRY(e, cudaMemcpyAsync(out, e->d_best.p, (size_t)e->NF * sizeof(avt_node_best), cudaMemcpyDeviceToHost, e->stream));
    e->scan_valid = true;
    return sync_and_resolve(e);
}

int avt_level_fetch_av(avt_engine* e, int32_t node, double* av_out) {
    if (!e) return AVT_E_INVALID;
    if (!e->scan_valid) return fail(e, AVT_E_STATE, "avt_level_fetch_av: no scan for the current frontier");
    if (node < 0 || node >= e->NF || !av_out) return fail(e, AVT_E_INVALID, "avt_level_fetch_av: node %d out of range", node);
    if (int rc = use_device(e)) return rc;
    CUDA_TRY(e, cudaMemcpyAsync(av_out, e->d_av.p + (size_t)node * e->S, (size_t)e->S * sizeof(double), cudaMemcpyDeviceToHost, e->stream));
    return sync_and_resolve(e);
}

int avt_level_fetch_hist(avt_engine* e, int32_t node, uint32_t* hist_out) {
    if (!e) return AVT_E_INVALID;
    if (!e->hist_valid) return fail(e, AVT_E_STATE, "avt_level_fetch_hist: no histogram for the current frontier");
    if (node < 0 || node >= e->NF || !hist_out) return fail(e, AVT_E_INVALID, "avt_level_fetch_hist: node %d out of range", node);
    if (int rc = use_device(e)) return rc;
    const size_t words = (size_t)e->TB * e->C;
    CUDA_TRY(e, cudaMemcpyAsync(hist_out, e->d_H.p + (size_t)node * words, words * sizeof(uint32_t), cudaMemcpyDeviceToHost, e->stream));
    return sync_and_resolve(e);
}

int avt_level_eval(avt_engine* e, int32_t criterion, const uint8_t* class_order, int32_t n, const int32_t* nodes,
                   const int32_t* splits, avt_split_eval* out) {
    if (!e) return AVT_E_INVALID;
    if (!e->hist_valid) return fail(e, AVT_E_STATE, "avt_level_eval: no histogram for the current frontier");
    if (criterion != AVT_CRIT_GINI && criterion != AVT_CRIT_ENTROPY) return fail(e, AVT_E_INVALID, "avt_level_eval: criterion %d", criterion);
    if (n < 0 || (n && (!nodes || !splits || !out))) return fail(e, AVT_E_INVALID, "avt_level_eval: NULL argument");
    if (n == 0) return AVT_OK;
    for (int i = 0; i < n; ++i)
        if (nodes[i] < 0 || nodes[i] >= e->NF || splits[i] < 0 || splits[i] >= e->S)
            return fail(e, AVT_E_INVALID, "avt_level_eval: pair %d (node %d, split %d) out of range", i, nodes[i], splits[i]);
    if (int rc = use_device(e)) return rc;
    if (int rc = upload_order(e, class_order)) return rc;
    std::vector<int32_t> pairs(2 * (size_t)n);
    std::copy(nodes, nodes + n, pairs.begin());
    std::copy(splits, splits + n, pairs.begin() + n);
    if (int rc = upload(e, e->d_pairs, pairs.data(), pairs.size())) return rc;
    CUDA_TRY(e, e->d_eval.ensure((size_t)n));
    TimerSlot ts;
    if (int rc = timer_open(e, &e->stats.eval_ms, &ts)) return rc;
    eval_kernel<<<(n + 63) / 64, 64, 0, e->stream>>>(e->d_H.p, make_table(e), e->d_order.p, criterion, n, e->d_pairs.p, e->d_pairs.p + n, e->d_eval.p);
    CUDA_TRY(e, cudaGetLastError());
    if (int rc = timer_close(e, ts)) return rc;
    e->stats.eval_launches += 1;
    CUDA_TRY(e, cudaMemcpyAsync(out, e->d_eval.p, (size_t)n * sizeof(avt_split_eval), cudaMemcpyDeviceToHost, e->stream));
    return sync_and_resolve(e);
}

int avt_level_repartition(avt_engine* e, const int32_t* node_attr, const uint16_t* child_lut, int32_t lut_width) {
    if (!e) return AVT_E_INVALID;
    if (e->NF == 0) return fail(e, AVT_E_STATE, "avt_level_repartition: no frontier (call avt_level_histogram first)");
    if (!node_attr || !child_lut || lut_width < 1) return fail(e, AVT_E_INVALID, "avt_level_repartition: NULL argument");
    for (int g = 0; g < e->NF; ++g) {
        if (node_attr[g] >= e->A) return fail(e, AVT_E_INVALID, "avt_level_repartition: node %d attribute %d out of range", g, node_attr[g]);
        if (node_attr[g] >= 0 && e->n_bins[node_attr[g]] > lut_width)
            return fail(e, AVT_E_INVALID, "avt_level_repartition: lut_width %d < bins of attribute %d", lut_width, node_attr[g]);
    }
    if (int rc = use_device(e)) return rc;
    if (int rc = upload(e, e->d_node_attr, node_attr, (size_t)e->NF)) return rc;
    if (int rc = upload(e, e->d_lut, child_lut, (size_t)e->NF * lut_width)) return rc;
    TimerSlot ts;
    if (int rc = timer_open(e, &e->stats.repart_ms, &ts)) return rc;
    long long blocks = std::min<long long>((e->N + 256 * 8 - 1) / (256 * 8), (long long)e->sm_count * 8);
    dim3 grid((unsigned)blocks, (unsigned)e->T);
    repartition_kernel<<<grid, 256, 0, e->stream>>>(e->nid_cur, (const void* const*)e->d_colptrs.p, e->d_wide.p, e->d_node_base.p,
                                                    e->d_node_attr.p, e->d_lut.p, lut_width, e->N, e->Npad);
    CUDA_TRY(e, cudaGetLastError());
    if (int rc = timer_close(e, ts)) return rc;
    e->stats.repart_launches += 1;
    e->stats.repart_alg_bytes += (long long)e->N * e->T * 5;
    e->hist_valid = e->scan_valid = false;
    e->NF = 0;
    return AVT_OK;
}

// Asynchronous form of the device-driven loop: nothing is read back until the last level.  The host
// sizes every launch from an UPPER BOUND of the frontier (each node has at most max-n_seg children);
// the kernels read the true frontier (node_base of every level lives on the device, written by the
// planning kernel of the level before).  Levels whose bound exceeds kAsyncNodeLimit fall back to one
// read-back so that buffers follow the real frontier.
constexpr int kAsyncNodeLimit = 16384;

static int grow_device_loop(avt_engine* e, int32_t n_levels, int32_t criterion, const uint8_t* class_order, int32_t strategy,
                            const uint32_t* root_cand, int32_t* nodes_per_tree, avt_grow_record* records, int64_t max_records,
                            int64_t* n_records) {
    const int T = e->T;
    const int mw = (e->A + 31) / 32;
    const bool masks_on = strategy == AVT_ATTRS_NOT_USED_YET || root_cand != nullptr;   // per-node candidate attribute sets on the device
    int lut_w = 1, g_max = 2;
    for (int a = 0; a < e->A; ++a) lut_w = std::max(lut_w, e->n_bins[a]);
    for (auto& sp : e->splits) g_max = std::max(g_max, sp.n_seg);
    if (n_levels <= 0) {
        if (n_records) *n_records = 0;
        return AVT_OK;
    }
    if (int rc = upload_order(e, class_order)) return rc;
    const bool want_records = records != nullptr || n_records != nullptr;
    CUDA_TRY(e, e->d_next_count.ensure((size_t)T));
    // control block (see d_ctrl): error flag, ticket and level 0's node_base go down in one copy from page-locked memory
    const size_t ctrl_ints = 16 + (size_t)(n_levels + 1) * (T + 1);
    const bool ctrl_pinned = ctrl_ints <= kCtrlHostInts;
    if (!e->h_ctrl) CUDA_TRY(e, cudaHostAlloc((void**)&e->h_ctrl, kCtrlHostInts * sizeof(int32_t), cudaHostAllocDefault));
    if (!e->h_rec) CUDA_TRY(e, cudaHostAlloc((void**)&e->h_rec, kRecHostCount * sizeof(avt_grow_record), cudaHostAllocDefault));
    CUDA_TRY(e, e->d_ctrl.ensure(ctrl_ints));
    int32_t* const d_err_p = e->d_ctrl.p;
    uint32_t* const d_ticket_p = reinterpret_cast<uint32_t*>(e->d_ctrl.p + 1);
    int32_t* const d_nb_all_p = e->d_ctrl.p + 16;
    std::vector<avt_grow_record> all;
    std::vector<int32_t> ub(nodes_per_tree, nodes_per_tree + T), nb_host((size_t)T + 1);
    struct RecRegion { int level; size_t host_at, count; };
    std::vector<RecRegion> rec_regions;  // where each level's record slots end up in `all`
    size_t flushed = 0;                  // records already copied to `all` by a mid-loop flush
    auto nb_level = [&](int level) { return d_nb_all_p + (size_t)level * (T + 1); };
    // level 0: the caller's frontier is exact
    nb_host[0] = 0;
    for (int t = 0; t < T; ++t) nb_host[(size_t)t + 1] = nb_host[(size_t)t] + ub[(size_t)t];
    {
        std::vector<int32_t> init_pageable;
        int32_t* init = e->h_ctrl;
        if (!ctrl_pinned || 16 + (size_t)T + 1 > kCtrlHostInts) {
            init_pageable.resize(16 + (size_t)T + 1);
            init = init_pageable.data();
        }
        for (int i = 0; i < 16; ++i) init[i] = 0;
        for (int t = 0; t <= T; ++t) init[16 + t] = nb_host[(size_t)t];
        CUDA_TRY(e, cudaMemcpyAsync(e->d_ctrl.p, init, (16 + (size_t)T + 1) * sizeof(int32_t), cudaMemcpyHostToDevice, e->stream));
        if (init != e->h_ctrl) CUDA_TRY(e, cudaStreamSynchronize(e->stream));   // the pageable source goes out of scope
    }
    if (masks_on) {
        // both buffers sized for the largest frontier bound any level of this call can reach (nothing reallocates mid-loop)
        long long worst = 0, cur_bound = 0;
        std::vector<long long> b(ub.begin(), ub.end());
        for (int level = 0; level <= n_levels; ++level) {
            cur_bound = 0;
            for (int t = 0; t < T; ++t) {
                cur_bound += b[(size_t)t];
                b[(size_t)t] = std::min<long long>(b[(size_t)t] * g_max, (long long)AVT_NODE_NONE - 1);
            }
            worst = std::max(worst, cur_bound);
        }
        if ((size_t)worst * mw > ((size_t)1 << 28)) return fail(e, AVT_E_NOMEM, "avt_grow_levels_ex: candidate sets for up to %lld nodes do not fit", worst);
        CUDA_TRY(e, e->d_cand[0].ensure((size_t)worst * mw));
        CUDA_TRY(e, e->d_cand[1].ensure((size_t)worst * mw));
        const size_t nf0 = (size_t)nb_host[(size_t)T];
        std::vector<uint32_t> init(nf0 * mw);
        for (size_t g = 0; g < nf0; ++g)
            for (int w = 0; w < mw; ++w) {
                const int bits = std::min(32, e->A - 32 * w);
                const uint32_t all = bits >= 32 ? 0xFFFFFFFFu : ((1u << bits) - 1u);
                init[g * mw + w] = root_cand ? (root_cand[g * mw + w] & all) : all;
            }
        CUDA_TRY(e, cudaMemcpyAsync(e->d_cand[0].p, init.data(), init.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, e->stream));
        CUDA_TRY(e, cudaStreamSynchronize(e->stream));   // `init` goes out of scope
    }
    const bool peer_on = e->peer.active && e->peer.n > 1;
    // Sibling subtraction (count the smaller children, derive the largest as parent - siblings; exact in integers): pays off where
    // the pass is bound by counter updates, i.e. on the shapes the forest kernel takes; AVT_SUBTRACT=0/1 overrides.
    bool forest_shape = e->A >= 16;
    for (int a = 0; a < e->A; ++a) forest_shape = forest_shape && !e->wide[a];
    const bool subtract = e->subtract_mode < 0 ? forest_shape : e->subtract_mode != 0;
    const bool fast_scan = e->scan_fast && e->fast_scan_ok && !subtract;   // latency-optimised scan + planning (avt_scan_fast.cuh)
    // complete counts of a level: single GPU = the (ping-pong) histogram buffer itself; peer mode = the local sums
    auto H_level = [&](int level) { return (level & 1) ? e->d_H2.p : e->d_H.p; };
    auto hsum_level = [&](int level) { return (level & 1) ? e->d_hsum2.p : e->d_hsum.p; };
    bool prev_peer = false;
    auto arena_H = [&](int parity) { return e->peer.arena + 64 + (size_t)parity * e->peer.cap_words; };
    bool peer_used = false;
    bool all_peer = true;        // every level of this call exchanged through peer memory
    bool pending = false;
    bool planned_skip = false;   // the previous level's planning marked this level's derived nodes
    bool derive_ok = true;       // the previous level's complete counts survived (no reallocation since)
    bool h_prezeroed = false;  // the previous level's planning block already cleared this level's histogram region
    FusedRepart fr{};
    size_t rec_off = 0;
    int last_level = -1;
    for (int level = 0; level < n_levels; ++level) {
        long long nf_ub = 0;
        int n_max = 0;
        for (int t = 0; t < T; ++t) {
            nf_ub += ub[(size_t)t];
            n_max = std::max(n_max, ub[(size_t)t]);
        }
        if (level > 0 && nf_ub > kAsyncNodeLimit) {
            // bound too loose to size buffers from: read the true frontier back once
            CUDA_TRY(e, cudaMemcpyAsync(nb_host.data(), nb_level(level), ((size_t)T + 1) * sizeof(int32_t), cudaMemcpyDeviceToHost, e->stream));
            if (int rc = sync_and_resolve(e)) return rc;
            nf_ub = 0;
            n_max = 0;
            for (int t = 0; t < T; ++t) {
                ub[(size_t)t] = nb_host[(size_t)t + 1] - nb_host[(size_t)t];
                nf_ub += ub[(size_t)t];
                n_max = std::max(n_max, ub[(size_t)t]);
            }
        }
        if (nf_ub == 0) break;
        if (n_max >= (int)AVT_NODE_NONE) return fail(e, AVT_E_INVALID, "avt_grow_levels: a tree exceeded %d frontier nodes", AVT_NODE_NONE - 1);
        const int NF = (int)nf_ub;
        e->NF = NF;
        e->hist_valid = e->scan_valid = false;
        const bool peer_level = peer_on && (size_t)NF * e->TB * e->C <= e->peer.cap_words;   // same decision on every rank (same bounds)
        if (!peer_level) h_prezeroed = h_prezeroed && !peer_on;
        // every buffer of this level is sized BEFORE anything is enqueued (a reallocation frees memory
        // that earlier, still running launches may use)
        if (e->d_av.cap < (size_t)NF * e->S || e->d_best.cap < (size_t)NF || e->d_best_eval.cap < (size_t)NF || e->d_node_attr.cap < (size_t)NF ||
            e->d_lut.cap < (size_t)NF * lut_w || (!peer_level && (e->d_H.cap < (size_t)NF * e->TB * e->C || e->d_H2.cap < (size_t)NF * e->TB * e->C)) ||
            (peer_level && (e->d_hsum.cap < (size_t)NF * e->TB * e->C || e->d_hsum2.cap < (size_t)NF * e->TB * e->C)) ||
            (subtract && (e->d_skip.cap < (size_t)NF * g_max || e->d_drv_parent.cap < (size_t)NF * g_max || e->d_kids.cap < (size_t)NF * kMaxSeg)) || (want_records && e->d_records.cap < rec_off + (size_t)NF) ||
            e->d_nkids.cap < (size_t)NF + 1) {
            if (int rc = sync_and_resolve(e)) return rc;
            // flush records gathered so far: the record buffer may move
            if (want_records && rec_off) {
                const size_t at = all.size();
                all.resize(at + rec_off);
                CUDA_TRY(e, cudaMemcpy(all.data() + at, e->d_records.p, rec_off * sizeof(avt_grow_record), cudaMemcpyDeviceToHost));
                flushed += rec_off;
                rec_off = 0;
            }
            // tables of the previous level are still needed by this level's fused pass: preserve them
            std::vector<int32_t> keep_attr;
            std::vector<uint16_t> keep_lut;
            size_t prev_nf = 0;
            if (pending) {
                prev_nf = std::min(e->d_node_attr.cap, e->d_lut.cap / (size_t)lut_w);
                keep_attr.resize(prev_nf);
                keep_lut.resize(prev_nf * lut_w);
                CUDA_TRY(e, cudaMemcpy(keep_attr.data(), e->d_node_attr.p, prev_nf * sizeof(int32_t), cudaMemcpyDeviceToHost));
                CUDA_TRY(e, cudaMemcpy(keep_lut.data(), e->d_lut.p, prev_nf * lut_w * sizeof(uint16_t), cudaMemcpyDeviceToHost));
            }
            const size_t grow = (size_t)NF * (size_t)g_max * 2;  // room for the next levels as well
            CUDA_TRY(e, e->d_av.ensure(grow * e->S));
            CUDA_TRY(e, e->d_best.ensure(grow));
            CUDA_TRY(e, e->d_best_eval.ensure(grow));
            CUDA_TRY(e, e->d_node_attr.ensure(grow));
            CUDA_TRY(e, e->d_lut.ensure(grow * lut_w));
            CUDA_TRY(e, e->d_nkids.ensure(grow + 1));   // child counts of the look-back scan / prefix scratch of plan_level
            // (a reallocation here drops the previous level's complete counts: no derivation on this level)
            const bool had_prev = level > 0;
            if (!peer_level) {
                if (e->d_H.cap < grow * e->TB * e->C || e->d_H2.cap < grow * e->TB * e->C) derive_ok = !had_prev;
                CUDA_TRY(e, e->d_H.ensure(grow * e->TB * e->C));
                CUDA_TRY(e, e->d_H2.ensure(grow * e->TB * e->C));
            } else {
                const size_t hs = std::min(grow * e->TB * e->C, e->peer.cap_words);
                if (e->d_hsum.cap < hs || e->d_hsum2.cap < hs) derive_ok = !had_prev;
                CUDA_TRY(e, e->d_hsum.ensure(hs));
                CUDA_TRY(e, e->d_hsum2.ensure(hs));
            }
            if (subtract) {
                // plan tables of the previous level are consumed by this level: keep them across the reallocation
                std::vector<uint8_t> k_skip(e->d_skip.cap);
                std::vector<int32_t> k_par(e->d_drv_parent.cap), k_kids(e->d_kids.cap);
                if (had_prev && e->d_skip.cap) {
                    CUDA_TRY(e, cudaMemcpy(k_skip.data(), e->d_skip.p, k_skip.size(), cudaMemcpyDeviceToHost));
                    CUDA_TRY(e, cudaMemcpy(k_par.data(), e->d_drv_parent.p, k_par.size() * 4, cudaMemcpyDeviceToHost));
                    CUDA_TRY(e, cudaMemcpy(k_kids.data(), e->d_kids.p, k_kids.size() * 4, cudaMemcpyDeviceToHost));
                }
                CUDA_TRY(e, e->d_skip.ensure(grow * g_max));
                CUDA_TRY(e, e->d_drv_parent.ensure(grow * g_max));
                CUDA_TRY(e, e->d_kids.ensure(grow * kMaxSeg));
                if (had_prev && !k_skip.empty()) {
                    CUDA_TRY(e, cudaMemcpy(e->d_skip.p, k_skip.data(), k_skip.size(), cudaMemcpyHostToDevice));
                    CUDA_TRY(e, cudaMemcpy(e->d_drv_parent.p, k_par.data(), k_par.size() * 4, cudaMemcpyHostToDevice));
                    CUDA_TRY(e, cudaMemcpy(e->d_kids.p, k_kids.data(), k_kids.size() * 4, cudaMemcpyHostToDevice));
                }
            }
            if (want_records) CUDA_TRY(e, e->d_records.ensure(grow * 2));
            if (pending) {
                CUDA_TRY(e, cudaMemcpy(e->d_node_attr.p, keep_attr.data(), prev_nf * sizeof(int32_t), cudaMemcpyHostToDevice));
                CUDA_TRY(e, cudaMemcpy(e->d_lut.p, keep_lut.data(), prev_nf * lut_w * sizeof(uint16_t), cudaMemcpyHostToDevice));
                fr.node_attr = e->d_node_attr.p;
                fr.child_lut = e->d_lut.p;
            }
        }
        if (pending) fr.prev_node_base = nb_level(level - 1);
        // derivation needs this level's skip marks (from the previous planning) AND the previous level's complete counts
        // (reallocations lose them); the previous level may have been exchanged the other way (peer memory for small levels,
        // NCCL for large ones): its complete counts are read from wherever that mode left them
        const bool derive = subtract && level > 0 && planned_skip && derive_ok;
        if (subtract && level > 0 && planned_skip && !derive) {
            // the rows of "derived" nodes must be counted after all: clear the marks
            CUDA_TRY(e, cudaMemsetAsync(e->d_skip.p, 0, (size_t)NF, e->stream));
        }
        derive_ok = true;
        e->skip_cur = derive ? e->d_skip.p : nullptr;   // (without derivation the marks were just cleared: every node is counted)
        e->hist_err = d_err_p;
        e->skip_h_memset = h_prezeroed;
        // programmatic dependent launch: legal when nothing but the previous scan kernel precedes this pass
        e->pdl_chain = level > 0 && h_prezeroed && !e->profiling && !(subtract && planned_skip && !derive);
        const int par = (int)(e->peer.seq & 1);   // exchange buffer of this level (peer levels only)
        e->H_cur = peer_level ? arena_H(par) : H_level(level);
        if (peer_level) {
            ++e->peer.epoch;   // the pass signals this epoch when its last block finishes
            ++e->peer.seq;
        } else {
            all_peer = false;
        }
        const bool push_ok = peer_level && fast_scan && e->peer_push && e->peer.n > 1 && (long long)e->TB * e->C <= kPushWords &&
                             (long long)NF * e->TB * e->C <= kPushCap;
        e->signal_pending = peer_level;
        e->scan_will_signal = peer_level;
        if (int rc = launch_histogram(e, n_max, pending ? &fr : nullptr, nb_level(level))) return rc;
        e->scan_will_signal = false;
        // small levels: push form of the exchange (no flags at all); needs the item-parallel scan kernel behind a shared-memory pass
        const bool push_level = peer_level && push_ok;
        const bool scan_signals = peer_level && !push_level && e->peer_signal_scan && (e->stats.last_hist_variant == 0 || e->stats.last_hist_variant == 3);
        e->skip_h_memset = false;
        e->pdl_chain = false;
        e->H_cur = nullptr;
        e->skip_cur = nullptr;
        e->hist_err = nullptr;
        e->signal_pending = false;
        if (peer_level) {
            if (!e->signal_in_hist) {
                // global-atomics fallback pass: signal with a separate launch
                TimerSlot tc;
                if (int rc = timer_open(e, &e->stats.comm_ms, &tc)) return rc;
                peer_signal_kernel<<<1, 32, 0, e->stream>>>(e->d_remote_flags.p, e->peer.n, e->peer.epoch);
                CUDA_TRY(e, cudaGetLastError());
                if (int rc = timer_close(e, tc)) return rc;
            }
            e->stats.comm_calls += 1;
            peer_used = true;
        }
        pending = false;
        PlanArgs pa{};
        pa.best = e->d_best.p;
        pa.best_eval = e->d_best_eval.p;
        pa.node_base = nb_level(level);
        pa.node_attr = e->d_node_attr.p;
        pa.child_lut = e->d_lut.p;
        pa.next_count = e->d_next_count.p;
        pa.prefix = e->d_nkids.p;
        pa.records = want_records ? e->d_records.p + rec_off : nullptr;
        pa.err = d_err_p;
        pa.next_node_base = nb_level(level + 1);
        pa.ticket = d_ticket_p;
        pa.lut_w = lut_w;
        pa.level = level;
        pa.n_trees = T;
        // the planning block also clears the next level's histogram region when that is small
        {
            long long next_nodes = 0;
            for (int t = 0; t < T; ++t) next_nodes += std::min<long long>((long long)ub[(size_t)t] * g_max, (long long)AVT_NODE_NONE - 1);
            const long long next_words = next_nodes * e->TB * e->C;
            const bool next_peer = peer_on && (size_t)next_words <= e->peer.cap_words;
            h_prezeroed = level + 1 < n_levels && next_nodes <= kAsyncNodeLimit && next_words <= (1 << 16) &&
                          (next_peer || (!peer_on && (size_t)next_words <= std::min(e->d_H.cap, e->d_H2.cap)));
            pa.zero_ptr = h_prezeroed ? (next_peer ? arena_H((int)(e->peer.seq & 1)) : H_level(level + 1)) : nullptr;
            pa.zero_words = h_prezeroed ? next_words : 0;
        }
        uint32_t* const Hc = peer_level ? hsum_level(level) : H_level(level);   // this level's complete counts
        if (peer_level) {
            pa.peer_H = e->d_peer_H[par].p;
            pa.peer_flags = e->peer.arena;
            pa.h_sum = Hc;
            pa.peer_epoch = e->peer.epoch;
            pa.n_ranks = e->peer.n;
            if (scan_signals) pa.signal_flags = e->d_remote_flags.p;
            if (push_level) {
                pa.peer_H = nullptr;   // nothing is pulled
                pa.push_dst = e->d_push_dst[par].p;
                pa.push_recv = reinterpret_cast<const unsigned long long*>(e->peer.arena + e->peer.recv_off_words) + (size_t)par * 16 * kPushCap;
                pa.H_local = arena_H(par);
                pa.push_cap = kPushCap;
                pa.my_rank = e->peer.rank;
            }
        }
        if (subtract) {
            const bool next_plannable = level + 1 < n_levels;
            pa.kids = e->d_kids.p;
            pa.next_skip = next_plannable ? e->d_skip.p : nullptr;
            pa.next_parent = next_plannable ? e->d_drv_parent.p : nullptr;
            if (derive) {
                pa.drv_parent = e->d_drv_parent.p;
                pa.prev_kids = e->d_kids.p;
                pa.H_prev = prev_peer ? hsum_level(level - 1) : H_level(level - 1);
            }
            pa.H_mut = Hc;
            planned_skip = next_plannable;
        }
        prev_peer = peer_level;
        bool lut2_written = false;
        TimerSlot ts;
        if (int rc = timer_open(e, &e->stats.scan_ms, &ts)) return rc;
        {
            cudaLaunchConfig_t cfg{};
            cfg.gridDim = dim3((unsigned)NF);
            cfg.blockDim = dim3(kScanThreads);
            cfg.stream = e->stream;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[0].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = attr;
            // directly behind the histogram kernel unless an NCCL all-reduce or a signalling launch sits between
            const bool chain = e->pdl && !e->profiling && (e->stats.last_hist_variant == 0 || e->stats.last_hist_variant == 3) && (peer_level ? e->signal_in_hist : e->comm == nullptr);
            cfg.numAttrs = chain ? 1 : 0;
            if (fast_scan) {
                FastTable ft{e->d_items.p, e->d_seg_first.p, e->d_nkids.p, e->n_items, masks_on ? e->d_cand[level & 1].p : nullptr, nullptr, 0u};
                if (e->lookback && NF <= kLookbackNodes) {
                    if (e->d_pub.cap < (size_t)kLookbackNodes || e->scan_epoch >= (1u << 24) - 1) {
                        // (re)start the tags: no earlier launch may have left this value behind
                        CUDA_TRY(e, e->d_pub.ensure((size_t)kLookbackNodes));
                        CUDA_TRY(e, cudaMemsetAsync(e->d_pub.p, 0, (size_t)kLookbackNodes * sizeof(uint32_t), e->stream));
                        e->scan_epoch = 0;
                    }
                    ft.pub = e->d_pub.p;
                    ft.epoch = ++e->scan_epoch;
                    if (e->packed_ok && (e->hist_joint || e->force_variant == 3) && level + 1 < n_levels) {
                        // the next pass may be the joint-code kernel: have the node blocks expand their rows to packed bytes
                        int maxv = 0;
                        for (int a = 0; a < e->A; ++a) maxv += (e->n_bins[a] - 1) << e->pk_shift[a];
                        const size_t need = (size_t)kLookbackNodes * (maxv + 1);
                        if (e->d_lut2.cap < need) {
                            // (the table is written by this launch and read by the next pass: a reallocation here can only
                            // happen before the first scan that uses it)
                            if (int rc = sync_and_resolve(e)) return rc;
                            CUDA_TRY(e, e->d_lut2.ensure(need));
                        }
                        pa.lut2 = e->d_lut2.p;
                        pa.lut2_J = maxv + 1;
                        for (int q = 0; q < 4; ++q) { pa.pk_shift[q] = e->pk_shift[q]; pa.pk_mask[q] = e->pk_mask4[q] & 0xFFu; }
                        lut2_written = true;
                    }
                }
                CUDA_TRY(e, cudaLaunchKernelEx(&cfg, scan_fast_kernel, (const uint32_t*)Hc, make_table(e), ft, (const uint8_t*)e->d_order.p, (int)criterion,
                                               (const int32_t*)(nb_level(level) + T), pa));
            } else
            CUDA_TRY(e, cudaLaunchKernelEx(&cfg, scan_kernel, (const uint32_t*)Hc, make_table(e), (const uint32_t*)(masks_on ? e->d_cand[level & 1].p : nullptr), (const uint8_t*)e->d_order.p,
                                           (int)criterion, e->d_av.p, e->d_best.p, e->d_best_eval.p, (const int32_t*)(nb_level(level) + T), pa));
        }
        CUDA_TRY(e, cudaGetLastError());
        if (masks_on && level + 1 < n_levels) {
            inherit_masks_kernel<<<(NF + 127) / 128, 128, 0, e->stream>>>(nb_level(level), nb_level(level + 1), T, e->d_node_attr.p, e->d_lut.p, lut_w,
                                                                          e->d_cand[level & 1].p, e->d_cand[(level + 1) & 1].p, mw,
                                                                          strategy == AVT_ATTRS_NOT_USED_YET ? 1 : 0);
            CUDA_TRY(e, cudaGetLastError());
        }
        if (int rc = timer_close(e, ts)) return rc;
        e->stats.scan_launches += 1;
        if (want_records) {
            rec_regions.push_back({level, flushed + rec_off, (size_t)NF});
            rec_off += (size_t)NF;
        }
        fr.node_attr = e->d_node_attr.p;
        fr.child_lut = e->d_lut.p;
        fr.lut_w = lut_w;
        fr.prev_n_max = n_max;
        fr.lut2 = lut2_written ? e->d_lut2.p : nullptr;
        pending = true;
        last_level = level;
        for (int t = 0; t < T; ++t) ub[(size_t)t] = (int)std::min<long long>((long long)ub[(size_t)t] * g_max, (long long)AVT_NODE_NONE - 1);
    }
    if (pending) {
        // apply the last planned split so that the node ids match the returned frontier
        TimerSlot tr;
        if (int rc = timer_open(e, &e->stats.repart_ms, &tr)) return rc;
        bool narrow = e->A <= 4;
        for (int a = 0; a < e->A; ++a) narrow = narrow && !e->wide[a];
        const size_t tbl = ((size_t)fr.prev_n_max + 2) * (4 + (size_t)lut_w * 2) + 8;
        if (narrow && tbl <= 16 * 1024) {
            RepartFastArgs ra{};
            ra.nid = e->nid_cur;
            ra.cols = (const void* const*)e->d_colptrs.p;
            if (e->packed_ok) {
                ra.packed = e->d_packed.p;
                for (int q = 0; q < 4; ++q) { ra.pk_shift[q] = e->pk_shift[q]; ra.pk_mask4[q] = e->pk_mask4[q]; }
            }
            ra.node_base = nb_level(last_level);
            ra.node_attr = e->d_node_attr.p;
            ra.child_lut = e->d_lut.p;
            ra.lut_w = lut_w;
            ra.n_attrs = e->A;
            ra.N = e->N;
            ra.Npad = e->Npad;
            const long long tiles = (e->N + 256 * kRowsPerThread - 1) / (256 * kRowsPerThread);
            dim3 grid((unsigned)std::min<long long>(tiles, std::max<long long>(1, (long long)e->sm_count * 8 / T)), (unsigned)T);
            repartition_fast_kernel<<<grid, 256, tbl, e->stream>>>(ra);
        } else if (e->rowrot_ok && e->A <= 32 && T >= 4) {
            // forests over one 32-attribute chunk: codes once per row from the rotated copy, loop over all trees (avt_forest.cuh)
            const long long n_tiles = (e->N + 31) / 32;
            long long parts = std::min<long long>((long long)e->sm_count * 4, std::max<long long>(1, n_tiles / (kForestWarps * 4)));
            const long long tpp = ((n_tiles + parts - 1) / parts + kForestWarps - 1) / kForestWarps * kForestWarps;
            parts = (n_tiles + tpp - 1) / tpp;
            repartition_rot_kernel<<<(unsigned)parts, kForestThreads, 0, e->stream>>>(e->nid_cur, e->d_rowrot.p, nb_level(last_level), e->d_node_attr.p, e->d_lut.p, lut_w, T,
                                                                                       e->N, e->Npad, tpp);
        } else {
            long long blocks = std::min<long long>((e->N + 256 * 8 - 1) / (256 * 8), (long long)e->sm_count * 8);
            dim3 grid((unsigned)blocks, (unsigned)T);
            repartition_kernel<<<grid, 256, 0, e->stream>>>(e->nid_cur, (const void* const*)e->d_colptrs.p, e->d_wide.p, nb_level(last_level),
                                                            e->d_node_attr.p, e->d_lut.p, lut_w, e->N, e->Npad);
        }
        CUDA_TRY(e, cudaGetLastError());
        if (int rc = timer_close(e, tr)) return rc;
        e->stats.repart_launches += 1;
        e->stats.repart_alg_bytes += (long long)e->N * T * 5;
    }
    if (peer_used && all_peer && !e->peer_trailing_round) {
        e->peer.lazy = true;   // the buffer numbering continues in the next call (see Peer::seq): no trailing round
    } else if (peer_used) {
        // mixed call: finish with a round of flags so that any numbering is safe afterwards
        e->peer.lazy = false;
        ++e->peer.epoch;
        peer_signal_kernel<<<1, 32, 0, e->stream>>>(e->d_remote_flags.p, e->peer.n, e->peer.epoch);
        peer_wait_kernel<<<1, 32, 0, e->stream>>>(e->peer.arena, e->peer.n, e->peer.epoch, d_err_p);
        CUDA_TRY(e, cudaGetLastError());
    }
    e->hist_valid = e->scan_valid = false;
    e->NF = 0;
    // error flag + every level's node_base in one copy, the records in another, both into page-locked memory: one wait
    const size_t back_ints = 16 + (size_t)(last_level + 2) * (T + 1);
    std::vector<int32_t> back_pageable;
    int32_t* back = e->h_ctrl;
    if (!ctrl_pinned) {
        back_pageable.resize(back_ints);
        back = back_pageable.data();
    }
    CUDA_TRY(e, cudaMemcpyAsync(back, e->d_ctrl.p, back_ints * sizeof(int32_t), cudaMemcpyDeviceToHost, e->stream));
    const size_t tail_at = all.size();
    const bool rec_pinned = want_records && rec_off && rec_off <= kRecHostCount;
    if (want_records && rec_off) {
        all.resize(tail_at + rec_off);
        CUDA_TRY(e, cudaMemcpyAsync(rec_pinned ? e->h_rec : all.data() + tail_at, e->d_records.p, rec_off * sizeof(avt_grow_record), cudaMemcpyDeviceToHost, e->stream));
    }
    CUDA_TRY(e, cudaEventRecord(e->gev1, e->stream));   // end of the call's device work (stats.last_grow_ms)
    e->gev1_recorded = true;
    if (int rc = sync_and_resolve(e)) return rc;
    if (rec_pinned) memcpy(all.data() + tail_at, e->h_rec, rec_off * sizeof(avt_grow_record));
    const int32_t err = back[0];
    const int32_t* nb_all_host = back + 16;
    if (last_level >= 0)
        for (int t = 0; t <= T; ++t) nb_host[(size_t)t] = nb_all_host[(size_t)(last_level + 1) * (T + 1) + t];
    // slots of a level's region beyond its true frontier were never written: mark them unused
    for (auto& rg : rec_regions) {
        const size_t true_nf = (size_t)nb_all_host[(size_t)rg.level * (T + 1) + T];
        for (size_t i = true_nf; i < rg.count; ++i) all[rg.host_at + i].level = -1;
    }
    if (err == 3) return fail(e, AVT_E_COMM, "avt_grow_levels: another rank's histogram did not arrive within %d s (a rank died, failed or skipped this call); the results of this call are invalid", (int)(kPeerWaitNs / 1000000000ull));
    if (err == 4) return fail(e, AVT_E_STATE, "avt_grow_levels: a tree had more counted nodes than the histogram pass reserved slots for (internal bound violated)");
    if (err == 2) return fail(e, AVT_E_CUDA, "avt_grow_levels: a scan block waited too long for its predecessors (look-back planning)");
    if (err) return fail(e, AVT_E_INVALID, "avt_grow_levels: a tree exceeded %d frontier nodes", AVT_NODE_NONE - 1);
    if (last_level >= 0)
        for (int t = 0; t < T; ++t) nodes_per_tree[t] = nb_host[(size_t)t + 1] - nb_host[(size_t)t];
    int64_t nrec = 0;
    for (size_t i = 0; i < all.size(); ++i) {
        if (all[i].level < 0 || all[i].split < 0) continue;  // unused slot of a level's bound, or a node that was not expanded
        if (records && nrec < max_records) records[nrec] = all[i];
        ++nrec;
    }
    if (n_records) *n_records = nrec;
    return AVT_OK;
}

int avt_grow_levels(avt_engine* e, int32_t n_levels, int32_t criterion, const uint8_t* class_order, int32_t* nodes_per_tree,
                    avt_grow_record* records, int64_t max_records, int64_t* n_records) {
    return avt_grow_levels_ex(e, n_levels, criterion, class_order, AVT_ATTRS_ALL, nullptr, nodes_per_tree, records, max_records, n_records);
}

int avt_grow_levels_ex(avt_engine* e, int32_t n_levels, int32_t criterion, const uint8_t* class_order, int32_t strategy,
                       const uint32_t* root_cand_mask, int32_t* nodes_per_tree, avt_grow_record* records, int64_t max_records,
                       int64_t* n_records) {
    if (!e) return AVT_E_INVALID;
    if (strategy != AVT_ATTRS_ALL && strategy != AVT_ATTRS_NOT_USED_YET) return fail(e, AVT_E_INVALID, "avt_grow_levels_ex: strategy %d", strategy);
    for (auto& sp : e->splits)
        if (sp.kind == AVT_SPLIT_RANGES)
            return fail(e, AVT_E_INVALID, "avt_grow_levels: the split table has AVT_SPLIT_RANGES splits (segments may overlap, a row can have several children); "
                                          "drive such tables level by level (avt_level_histogram / scan / eval / repartition)");
    if (!nodes_per_tree) return fail(e, AVT_E_INVALID, "avt_grow_levels: nodes_per_tree is NULL");
    if (e->T == 0 || e->S == 0) return fail(e, AVT_E_STATE, "avt_grow_levels: needs avt_forest_init and avt_split_table_set first");
    if (criterion != AVT_CRIT_GINI && criterion != AVT_CRIT_ENTROPY) return fail(e, AVT_E_INVALID, "avt_grow_levels: criterion %d", criterion);
    if (int rc = use_device(e)) return rc;
    CUDA_TRY(e, cudaEventRecord(e->gev0, e->stream));
    e->gev1_recorded = false;
    if (int rc = grow_device_loop(e, n_levels, criterion, class_order, strategy, root_cand_mask, nodes_per_tree, records, max_records, n_records)) return rc;
    if (!e->gev1_recorded) {   // loops that read back between levels: the stream is idle here
        CUDA_TRY(e, cudaEventRecord(e->gev1, e->stream));
        CUDA_TRY(e, cudaEventSynchronize(e->gev1));
    }
    float ms = 0.f;
    CUDA_TRY(e, cudaEventElapsedTime(&ms, e->gev0, e->gev1));
    e->stats.last_grow_ms = ms;
    return AVT_OK;
}

int avt_cpg_sweep(avt_engine* e, int32_t node, int32_t criterion, const uint8_t* class_order, avt_cpg_stat* out) {
    if (!e) return AVT_E_INVALID;
    if (!e->hist_valid) return fail(e, AVT_E_STATE, "avt_cpg_sweep: no histogram for the current frontier");
    if (e->S == 0) return fail(e, AVT_E_STATE, "avt_cpg_sweep: no split table installed");
    if (node < 0 || node >= e->NF || !out) return fail(e, AVT_E_INVALID, "avt_cpg_sweep: node %d out of range", node);
    if (criterion < AVT_CRIT_GINI || criterion > AVT_CRIT_CLASS_CONF) return fail(e, AVT_E_INVALID, "avt_cpg_sweep: criterion %d", criterion);
    if (criterion == AVT_CRIT_HELLINGER && e->C != 2)
        return fail(e, AVT_E_REFERENCE, "Hellinger distance algorithm is only valid for binary valued class attributes");
    if (int rc = use_device(e)) return rc;
    if (int rc = upload_order(e, class_order)) return rc;
    CUDA_TRY(e, e->d_cpg.ensure((size_t)e->S));
    TimerSlot ts;
    if (int rc = timer_open(e, &e->stats.scan_ms, &ts)) return rc;
    cpg_kernel<<<(e->S + kCpgWarps - 1) / kCpgWarps, kCpgWarps * 32, 0, e->stream>>>(e->d_H.p, make_table(e), node, e->d_order.p, criterion, e->d_cpg.p);
    CUDA_TRY(e, cudaGetLastError());
    if (int rc = timer_close(e, ts)) return rc;
    e->stats.scan_launches += 1;
    CUDA_TRY(e, cudaMemcpyAsync(out, e->d_cpg.p, (size_t)e->S * sizeof(avt_cpg_stat), cudaMemcpyDeviceToHost, e->stream));
    return sync_and_resolve(e);
}

int avt_predict(avt_engine* e, int32_t n_models, const int32_t* n_paths, const uint32_t* masks, const uint8_t* path_class,
                const double* weights, int32_t n_classes, const uint8_t* class_rank, uint8_t* pred_out) {
    if (!e) return AVT_E_INVALID;
    if (e->N == 0) return fail(e, AVT_E_STATE, "avt_predict: no data set loaded");
    if (n_models < 1 || !n_paths || !masks || !path_class || !class_rank || !pred_out) return fail(e, AVT_E_INVALID, "avt_predict: NULL argument");
    if (n_classes < 1 || n_classes > AVT_MAX_CLASSES) return fail(e, AVT_E_INVALID, "avt_predict: n_classes %d outside 1..%d", n_classes, AVT_MAX_CLASSES);
    if (int rc = use_device(e)) return rc;
    std::vector<int64_t> mask_off((size_t)n_models);
    std::vector<int32_t> path_off((size_t)n_models);
    std::vector<double> w((size_t)n_models, 1.0);
    int64_t mo = 0;
    int32_t po = 0;
    for (int m = 0; m < n_models; ++m) {
        if (n_paths[m] < 1) return fail(e, AVT_E_INVALID, "avt_predict: model %d has no decision path", m);
        mask_off[(size_t)m] = mo;
        path_off[(size_t)m] = po;
        mo += (int64_t)e->TB * ((n_paths[m] + 31) / 32);
        po += n_paths[m];
        if (weights) w[(size_t)m] = weights[m];
    }
    for (int p = 0; p < po; ++p)
        if (path_class[p] >= n_classes) return fail(e, AVT_E_INVALID, "avt_predict: path %d predicts class code %d >= n_classes", p, path_class[p]);
    if (int rc = upload(e, e->d_pmask, masks, (size_t)mo)) return rc;
    if (int rc = upload(e, e->d_pmoff, mask_off.data(), mask_off.size())) return rc;
    if (int rc = upload(e, e->d_pint, n_paths, (size_t)n_models)) return rc;
    if (int rc = upload(e, e->d_ppoff, path_off.data(), path_off.size())) return rc;
    if (int rc = upload(e, e->d_pcls, path_class, (size_t)po)) return rc;
    if (int rc = upload(e, e->d_pw, w.data(), w.size())) return rc;
    if (int rc = upload(e, e->d_order, class_rank, (size_t)n_classes)) return rc;
    CUDA_TRY(e, e->d_pout.ensure((size_t)e->N));
    PredictArgs pa{};
    pa.cols = (const void* const*)e->d_colptrs.p;
    pa.wide = e->d_wide.p;
    pa.bin_off = e->d_bin_off.p;
    pa.masks = e->d_pmask.p;
    pa.mask_off = e->d_pmoff.p;
    pa.n_paths = e->d_pint.p;
    pa.path_off = e->d_ppoff.p;
    pa.path_class = e->d_pcls.p;
    pa.weights = e->d_pw.p;
    pa.class_rank = e->d_order.p;
    pa.out = e->d_pout.p;
    pa.N = e->N;
    pa.A = e->A;
    pa.M = n_models;
    pa.n_classes = n_classes;
    TimerSlot ts;
    if (int rc = timer_open(e, &e->stats.eval_ms, &ts)) return rc;
    const long long blocks = std::min<long long>((e->N + 255) / 256, (long long)e->sm_count * 8);
    predict_kernel<<<(unsigned)blocks, 256, 0, e->stream>>>(pa);
    CUDA_TRY(e, cudaGetLastError());
    if (int rc = timer_close(e, ts)) return rc;
    e->stats.eval_launches += 1;
    CUDA_TRY(e, cudaMemcpyAsync(pred_out, e->d_pout.p, (size_t)e->N, cudaMemcpyDeviceToHost, e->stream));
    return sync_and_resolve(e);
}

int avt_forest_node_ids(avt_engine* e, int32_t tree, uint16_t* out) {
    if (!e) return AVT_E_INVALID;
    if (tree < 0 || tree >= e->T || !out) return fail(e, AVT_E_INVALID, "avt_forest_node_ids: tree %d out of range", tree);
    if (int rc = use_device(e)) return rc;
    CUDA_TRY(e, cudaMemcpyAsync(out, e->nid_cur + (size_t)tree * e->Npad, (size_t)e->N * 2, cudaMemcpyDeviceToHost, e->stream));
    return sync_and_resolve(e);
}

int avt_level_class_counts(avt_engine* e, uint32_t* out) {
    if (!e) return AVT_E_INVALID;
    if (!e->hist_valid || !out) return fail(e, AVT_E_STATE, "avt_level_class_counts: no histogram for the current frontier");
    if (int rc = use_device(e)) return rc;
    const int n = e->NF * e->C;
    CUDA_TRY(e, e->d_ccnt.ensure((size_t)n));
    class_counts_kernel<<<(n + 127) / 128, 128, 0, e->stream>>>(e->d_H.p, e->NF, e->TB, e->C, e->n_bins[0], e->d_ccnt.p);
    CUDA_TRY(e, cudaGetLastError());
    CUDA_TRY(e, cudaMemcpyAsync(out, e->d_ccnt.p, (size_t)n * sizeof(uint32_t), cudaMemcpyDeviceToHost, e->stream));
    return sync_and_resolve(e);
}

int avt_timer_begin(avt_engine* e) {
    if (!e) return AVT_E_INVALID;
    if (int rc = use_device(e)) return rc;
    CUDA_TRY(e, cudaEventRecord(e->tev0, e->stream));
    return AVT_OK;
}

int avt_timer_end(avt_engine* e, float* ms) {
    if (!e || !ms) return AVT_E_INVALID;
    if (int rc = use_device(e)) return rc;
    CUDA_TRY(e, cudaEventRecord(e->tev1, e->stream));
    if (int rc = sync_and_resolve(e)) return rc;
    CUDA_TRY(e, cudaEventElapsedTime(ms, e->tev0, e->tev1));
    return AVT_OK;
}

int avt_synchronize(avt_engine* e) {
    if (!e) return AVT_E_INVALID;
    if (int rc = use_device(e)) return rc;
    return sync_and_resolve(e);
}

int avt_get_stats(avt_engine* e, avt_stats* out) {
    if (!e || !out) return AVT_E_INVALID;
    if (int rc = use_device(e)) return rc;
    if (int rc = sync_and_resolve(e)) return rc;
    *out = e->stats;
    return AVT_OK;
}

int avt_reset_stats(avt_engine* e) {
    if (!e) return AVT_E_INVALID;
    if (int rc = use_device(e)) return rc;
    if (int rc = sync_and_resolve(e)) return rc;
    avt_stats z{};
    z.last_hist_variant = e->stats.last_hist_variant;
    e->stats = z;
    return AVT_OK;
}

int avt_set_hist_variant(avt_engine* e, int32_t variant) {
    if (!e) return AVT_E_INVALID;
    if (variant < -1 || variant > 4) return fail(e, AVT_E_INVALID, "avt_set_hist_variant: %d", variant);
    e->force_variant = variant;
    return AVT_OK;
}

int avt_peer_export(avt_engine* e, void* handle64) {
    if (!e || !handle64) return fail(e, AVT_E_INVALID, "avt_peer_export: NULL argument");
    if (int rc = use_device(e)) return rc;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
    if (!e->peer.arena) {
        e->peer.cap_words = (size_t)2 << 20;  // 8 MiB per histogram buffer
        e->peer.recv_off_words = 64 + 2 * e->peer.cap_words;
        // + receive slots of the push exchange: 2 buffers x 16 source ranks x kPushCap 64-bit (tag, count) words
        const size_t bytes = e->peer.recv_off_words * sizeof(uint32_t) + (size_t)2 * 16 * kPushCap * sizeof(unsigned long long);
        CUDA_TRY(e, cudaMalloc((void**)&e->peer.arena, bytes));
        CUDA_TRY(e, cudaMemset(e->peer.arena, 0, bytes));
    }
    cudaIpcMemHandle_t h;
    CUDA_TRY(e, cudaIpcGetMemHandle(&h, e->peer.arena));
    memcpy(handle64, &h, 64);
    return AVT_OK;
}

int avt_peer_import(avt_engine* e, int32_t n_ranks, int32_t rank, const void* handles) {
    if (!e || !handles || n_ranks < 1 || n_ranks > 16 || rank < 0 || rank >= n_ranks) return fail(e, AVT_E_INVALID, "avt_peer_import: bad argument (1..16 ranks)");
    if (!e->peer.arena) return fail(e, AVT_E_STATE, "avt_peer_import: call avt_peer_export first");
    // one import per engine: the arena's flag words and tagged receive slots of an earlier session would pass for this one's
    // (the epoch restarts at 0), and the earlier IPC mappings would leak
    if (e->peer.n > 0) return fail(e, AVT_E_STATE, "avt_peer_import: this engine already imported its peers; create a new engine for a new set of ranks");
    if (int rc = use_device(e)) return rc;
    if (int rc = sync_and_resolve(e)) return rc;
    std::vector<const uint32_t*> h0((size_t)n_ranks), h1((size_t)n_ranks);
    std::vector<uint32_t*> rf((size_t)n_ranks);
    std::vector<void*> bases((size_t)n_ranks, nullptr);
    for (int r = 0; r < n_ranks; ++r) {
        void* base = e->peer.arena;
        if (r != rank) {
            cudaIpcMemHandle_t h;
            memcpy(&h, (const char*)handles + (size_t)r * 64, 64);
            const cudaError_t st = cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess);
            if (st != cudaSuccess) {
                for (int q = 0; q < r; ++q)
                    if (q != rank && bases[(size_t)q]) cudaIpcCloseMemHandle(bases[(size_t)q]);   // nothing of a failed import stays behind
                return fail(e, AVT_E_CUDA, "avt_peer_import: cudaIpcOpenMemHandle of rank %d: %s", r, cudaGetErrorString(st));
            }
        }
        bases[(size_t)r] = base;
        uint32_t* b = (uint32_t*)base;
        h0[(size_t)r] = b + 64;
        h1[(size_t)r] = b + 64 + e->peer.cap_words;
        rf[(size_t)r] = b + rank;  // this rank's flag word inside rank r's arena
    }
    for (int r = 0; r < n_ranks; ++r) e->peer.base[r] = bases[(size_t)r];
    e->peer.n = n_ranks;
    e->peer.rank = rank;
    if (int rc = upload(e, e->d_peer_H[0], h0.data(), h0.size())) return rc;
    if (int rc = upload(e, e->d_peer_H[1], h1.data(), h1.size())) return rc;
    if (int rc = upload(e, e->d_remote_flags, rf.data(), rf.size())) return rc;
    for (int par = 0; par < 2; ++par) {
        std::vector<unsigned long long*> dst((size_t)n_ranks);
        for (int r = 0; r < n_ranks; ++r)
            dst[(size_t)r] = reinterpret_cast<unsigned long long*>((uint32_t*)e->peer.base[r] + e->peer.recv_off_words) + ((size_t)par * 16 + (size_t)rank) * kPushCap;
        if (int rc = upload(e, e->d_push_dst[par], dst.data(), dst.size())) return rc;
    }
    e->peer.epoch = 0;
    e->peer.seq = 0;
    e->peer.lazy = false;
    e->peer.active = getenv("AVT_PEER_OFF") == nullptr;
    return sync_and_resolve(e);
}

int avt_set_profiling(avt_engine* e, int32_t on) {
    if (!e) return AVT_E_INVALID;
    if (int rc = use_device(e)) return rc;
    if (int rc = sync_and_resolve(e)) return rc;
    e->profiling = on != 0;
    return AVT_OK;
}

int avt_comm_unique_id(void* id128) {
    if (!id128) return AVT_E_INVALID;
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    ncclResult_t st = ncclGetUniqueId(&id);
    if (st != ncclSuccess) return fail(nullptr, AVT_E_COMM, "ncclGetUniqueId: %s", ncclGetErrorString(st));
    memcpy(id128, &id, 128);
    return AVT_OK;
}

int avt_comm_init_rank(avt_engine* e, int32_t n_ranks, int32_t rank, const void* id128) {
    if (!e || !id128 || n_ranks < 1 || rank < 0 || rank >= n_ranks) return fail(e, AVT_E_INVALID, "avt_comm_init_rank: bad argument");
    if (int rc = use_device(e)) return rc;
    if (e->comm) {
        ncclCommDestroy(e->comm);
        e->comm = nullptr;
    }
    ncclUniqueId id;
    memcpy(&id, id128, 128);
    NCCL_TRY(e, ncclCommInitRank(&e->comm, n_ranks, id, rank));
    e->n_ranks = n_ranks;
    e->rank = rank;
    return AVT_OK;
}

}  // extern "C"

// ================================================================================================
// avt_histogram_static -- histogram counts for callers that must not allocate device memory
// ================================================================================================
//
// Who this is for.  Some hosts of this library run where a device allocation is either forbidden or
// unaffordable at call time: a task thread inside a service whose GPU memory was carved up at start-up and
// handed to other components (a later cudaMalloc can fail or push a neighbour out of memory); a caller on
// a latency budget that cannot absorb cudaMalloc/cudaFree, which take driver-wide locks and may
// synchronise the whole device; or code that issues its work while another part of the process has a
// stream capture open, during which allocation calls are not permitted.  The regular entry points
// (avt_create, avt_dataset_load, avt_level_histogram, ...) allocate and grow device buffers on demand, so
// they are not usable from such places.
//
// What it does.  The same counts as avt_dataset_load + avt_forest_init(1 tree) + avt_level_histogram:
//     hist_out[node][bin over all attributes][class] = number of rows of that node with that bin and class
// computed by the warp-aggregated global-atomics histogram kernel (hist_global_kernel).  Everything the
// kernel needs on the device -- the staged columns, class codes, node ids, its small index tables and
// the count array it accumulates into -- lives in ONE fixed array that is part of the library image
// (g_static_scratch below, AVT_STATIC_SCRATCH_BYTES = 8 MiB per GPU that loads the library, held until
// the process exits).  The call performs no cudaMalloc / cudaFree, creates no stream or event, and uses
// no allocating library routine; copies and the launch go through the legacy default stream.
//
// Limits.  All staged data plus the count array must fit the 8 MiB array; a larger request is refused with
// AVT_E_INVALID and a message that states the bytes needed and the bytes available (avt_last_error(NULL)).
// At most AVT_STATIC_MAX_ATTRS attributes.
//
// Set-up.  The array's name is only a host-side handle; the device address is obtained once per GPU, with
// that GPU current, through cudaGetSymbolAddress, and cached.  This happens inside the FIRST call on each
// device -- make that first call before any stream capture begins in the process.
//
// Sharing.  Every call on a device uses the same array and finds whatever the previous call left there.
// The count region is cleared at the start of every call (it is accumulated into with atomics and must
// start at zero); all other regions are fully overwritten before they are read.  Two calls that overlap
// on the same GPU (from two threads, or two processes' worth of threads in one process) race on the array:
// serialise them.
#include <mutex>

#define AVT_STATIC_SCRATCH_BYTES (8u << 20)
#define AVT_STATIC_MAX_ATTRS 64
#define AVT_STATIC_MAX_DEVICES 64

static __device__ __align__(256) unsigned char g_static_scratch[AVT_STATIC_SCRATCH_BYTES];

namespace {

struct StaticScratchDev {
    bool ready = false;
    unsigned char* base = nullptr;  // device address of g_static_scratch on this GPU
};
StaticScratchDev g_static_dev[AVT_STATIC_MAX_DEVICES];
std::mutex g_static_mu;

inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

}  // namespace

extern "C" int avt_histogram_static(int device, int64_t n_rows, int32_t n_attrs, const int32_t* n_bins, const void* const* columns,
                                    const uint8_t* class_codes, int32_t n_classes, const uint16_t* node_ids, int32_t n_nodes,
                                    uint32_t* hist_out) {
    if (n_rows <= 0 || n_attrs <= 0 || !n_bins || !columns || !class_codes || !hist_out)
        return fail(nullptr, AVT_E_INVALID, "avt_histogram_static: empty data set or NULL argument");
    if (n_attrs > AVT_STATIC_MAX_ATTRS) return fail(nullptr, AVT_E_INVALID, "avt_histogram_static: %d attributes, limit %d", n_attrs, AVT_STATIC_MAX_ATTRS);
    if (n_classes < 1 || n_classes > AVT_MAX_CLASSES) return fail(nullptr, AVT_E_INVALID, "avt_histogram_static: n_classes %d outside 1..%d", n_classes, AVT_MAX_CLASSES);
    if (n_nodes < 1 || n_nodes >= (int)AVT_NODE_NONE) return fail(nullptr, AVT_E_INVALID, "avt_histogram_static: n_nodes %d outside 1..%d", n_nodes, AVT_NODE_NONE - 1);
    if (device < 0 || device >= AVT_STATIC_MAX_DEVICES) return fail(nullptr, AVT_E_INVALID, "avt_histogram_static: device %d outside 0..%d", device, AVT_STATIC_MAX_DEVICES - 1);
    if (n_rows > (int64_t)AVT_STATIC_SCRATCH_BYTES)   // before any size arithmetic: a huge row count must not wrap the carve below
        return fail(nullptr, AVT_E_INVALID, "avt_histogram_static: %lld rows cannot fit the built-in %u-byte array", (long long)n_rows, (unsigned)AVT_STATIC_SCRATCH_BYTES);

    // ---- carve the static array (fixed-size host bookkeeping only: nothing is allocated anywhere)
    int32_t bin_off[AVT_STATIC_MAX_ATTRS + 1], wide[AVT_STATIC_MAX_ATTRS];
    size_t col_off[AVT_STATIC_MAX_ATTRS];
    size_t cur = 0;
    bin_off[0] = 0;
    for (int a = 0; a < n_attrs; ++a) {
        if (n_bins[a] < 1 || n_bins[a] > 65535) return fail(nullptr, AVT_E_INVALID, "avt_histogram_static: n_bins[%d]=%d outside 1..65535", a, n_bins[a]);
        if (!columns[a]) return fail(nullptr, AVT_E_INVALID, "avt_histogram_static: columns[%d] is NULL", a);
        bin_off[a + 1] = bin_off[a] + n_bins[a];
        wide[a] = n_bins[a] > 256;
        col_off[a] = cur;
        cur += align256((size_t)n_rows * (wide[a] ? 2 : 1));
    }
    const int TB = bin_off[n_attrs];
    const size_t off_cls = cur;       cur += align256((size_t)n_rows);
    const size_t off_nid = cur;       cur += align256((size_t)n_rows * 2);
    const size_t off_ptrs = cur;      cur += align256((size_t)n_attrs * sizeof(void*));
    const size_t off_binoff = cur;    cur += align256((size_t)(n_attrs + 1) * sizeof(int32_t));
    const size_t off_wide = cur;      cur += align256((size_t)n_attrs * sizeof(int32_t));
    const size_t off_nb = cur;        cur += align256(2 * sizeof(int32_t));
    const size_t off_bad = cur;       cur += align256(sizeof(int32_t));
    const size_t off_hist = cur;
    const size_t hist_bytes = (size_t)n_nodes * TB * n_classes * sizeof(uint32_t);
    cur += align256(hist_bytes);
    if (cur > AVT_STATIC_SCRATCH_BYTES)
        return fail(nullptr, AVT_E_INVALID,
                    "avt_histogram_static: this request needs %zu bytes of device scratch (%lld rows x %d attributes staged + %d nodes x %d bins x %d classes of "
                    "counts) but the built-in array holds %u bytes; use avt_level_histogram (which allocates) or split the rows",
                    cur, (long long)n_rows, n_attrs, n_nodes, TB, n_classes, (unsigned)AVT_STATIC_SCRATCH_BYTES);

    // the caller's current device is put back on every return path
    struct DeviceRestore {
        int prev = -1;
        ~DeviceRestore() { if (prev >= 0) cudaSetDevice(prev); }
    } restore;
    if (cudaGetDevice(&restore.prev) != cudaSuccess) restore.prev = -1;
    cudaError_t st = cudaSetDevice(device);
    if (st != cudaSuccess) return fail(nullptr, AVT_E_CUDA, "avt_histogram_static: cudaSetDevice(%d): %s", device, cudaGetErrorString(st));
    unsigned char* base = nullptr;
    {
        // once per GPU, with that GPU current: turn the array's host-side handle into its device address
        std::lock_guard<std::mutex> lk(g_static_mu);
        StaticScratchDev& d = g_static_dev[device];
        if (!d.ready) {
            void* p = nullptr;
            st = cudaGetSymbolAddress(&p, g_static_scratch);
            if (st != cudaSuccess || !p)
                return fail(nullptr, AVT_E_CUDA, "avt_histogram_static: cudaGetSymbolAddress on device %d: %s", device, cudaGetErrorString(st));
            d.base = (unsigned char*)p;
            d.ready = true;
        }
        base = d.base;
    }
#define STATIC_TRY(call)                                                                                              \
    do {                                                                                                              \
        cudaError_t _e = (call);                                                                                      \
        if (_e != cudaSuccess) return fail(nullptr, AVT_E_CUDA, "avt_histogram_static: %s: %s", #call, cudaGetErrorString(_e)); \
    } while (0)
    // ---- stage the caller's host buffers (synchronous copies on the legacy default stream)
    const void* col_ptrs[AVT_STATIC_MAX_ATTRS];
    for (int a = 0; a < n_attrs; ++a) {
        STATIC_TRY(cudaMemcpy(base + col_off[a], columns[a], (size_t)n_rows * (wide[a] ? 2 : 1), cudaMemcpyHostToDevice));
        col_ptrs[a] = base + col_off[a];
    }
    STATIC_TRY(cudaMemcpy(base + off_cls, class_codes, (size_t)n_rows, cudaMemcpyHostToDevice));
    if (node_ids) STATIC_TRY(cudaMemcpy(base + off_nid, node_ids, (size_t)n_rows * 2, cudaMemcpyHostToDevice));
    else STATIC_TRY(cudaMemset(base + off_nid, 0, (size_t)n_rows * 2));  // every row in node 0
    const int32_t node_base[2] = {0, n_nodes};
    STATIC_TRY(cudaMemcpy(base + off_ptrs, col_ptrs, (size_t)n_attrs * sizeof(void*), cudaMemcpyHostToDevice));
    STATIC_TRY(cudaMemcpy(base + off_binoff, bin_off, (size_t)(n_attrs + 1) * sizeof(int32_t), cudaMemcpyHostToDevice));
    STATIC_TRY(cudaMemcpy(base + off_wide, wide, (size_t)n_attrs * sizeof(int32_t), cudaMemcpyHostToDevice));
    STATIC_TRY(cudaMemcpy(base + off_nb, node_base, sizeof node_base, cudaMemcpyHostToDevice));
    // the counts are accumulated with atomics: whatever an earlier call left in this region must go
    STATIC_TRY(cudaMemset(base + off_hist, 0, hist_bytes));
    STATIC_TRY(cudaMemset(base + off_bad, 0, sizeof(int32_t)));

    HistArgs ha{};
    ha.cols = (const void* const*)(base + off_ptrs);
    ha.cls = base + off_cls;
    ha.nid = (const uint16_t*)(base + off_nid);
    ha.wts = nullptr;
    ha.has_w = nullptr;
    ha.H = (uint32_t*)(base + off_hist);
    ha.node_base = (const int32_t*)(base + off_nb);
    ha.N = n_rows;
    ha.Npad = n_rows;
    ha.C = n_classes;
    ha.TB = TB;
    ha.T = 1;
    ha.bad = (int32_t*)(base + off_bad);   // the caller's ids and codes are unvalidated: out-of-shape rows are dropped and reported
    int sm_count = kSmCountFallback;
    cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, device);
    const long long blocks = std::min<long long>((n_rows + 255) / 256, (long long)sm_count * 8);
    hist_global_kernel<<<dim3((unsigned)blocks, 1), 256, 0, 0>>>(ha, n_attrs, (const int32_t*)(base + off_binoff), (const int32_t*)(base + off_wide));
    STATIC_TRY(cudaGetLastError());
    // synchronous copy on the same stream: waits for the kernel
    int32_t bad = 0;
    STATIC_TRY(cudaMemcpy(&bad, base + off_bad, sizeof bad, cudaMemcpyDeviceToHost));
    if (bad)
        return fail(nullptr, AVT_E_INVALID, "avt_histogram_static: the data does not fit the declared shape (a node id >= n_nodes other than AVT_NODE_NONE, "
                    "a bin code >= n_bins[attribute] or a class code >= n_classes); nothing was written to hist_out");
    STATIC_TRY(cudaMemcpy(hist_out, base + off_hist, hist_bytes, cudaMemcpyDeviceToHost));
#undef STATIC_TRY
    return AVT_OK;
}
