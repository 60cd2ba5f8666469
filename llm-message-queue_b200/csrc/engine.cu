// engine.cu — the C ABI of include/triage.h: engine object, device buffers, streams, the host-side
// level queues (the part of MultiLevelQueue that is bookkeeping, not arithmetic) and the NCCL merge.
//
// No CPU fallback lives here: every compute entry point launches the sm_100a kernels of kernels.cu
// and returns TRIAGE_E_CUDA if that is impossible.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cstdlib>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/triage.h"
#include "dfa_build.h"
#include "kernels.cuh"
#include "regex_vm.h"
#include "ingest.h"
#include <unordered_map>

using namespace triage;

namespace {

thread_local std::string g_err;
int32_t fail(int32_t code, const std::string &msg) { g_err = msg; return code; }

#define CU(call)                                                                                          \
    do {                                                                                                  \
        cudaError_t _e = (call);                                                                          \
        if (_e != cudaSuccess)                                                                            \
            return fail(TRIAGE_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(_e));               \
    } while (0)

// ---- lazily bound NCCL (so the library loads on hosts without it) ----------------------------
struct NcclApi {
    void *h = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitRankConfig)(ncclComm_t *, int, ncclUniqueId, int, ncclConfig_t *) = nullptr;   // optional (NCCL >= 2.14)
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    bool load(std::string *err) {
        if (h) return true;
        h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!h) { *err = std::string("dlopen libnccl.so.2: ") + dlerror(); return false; }
        GetUniqueId = reinterpret_cast<decltype(GetUniqueId)>(dlsym(h, "ncclGetUniqueId"));
        CommInitRank = reinterpret_cast<decltype(CommInitRank)>(dlsym(h, "ncclCommInitRank"));
        CommInitRankConfig = reinterpret_cast<decltype(CommInitRankConfig)>(dlsym(h, "ncclCommInitRankConfig"));
        AllGather = reinterpret_cast<decltype(AllGather)>(dlsym(h, "ncclAllGather"));
        CommDestroy = reinterpret_cast<decltype(CommDestroy)>(dlsym(h, "ncclCommDestroy"));
        GetErrorString = reinterpret_cast<decltype(GetErrorString)>(dlsym(h, "ncclGetErrorString"));
        if (!GetUniqueId || !CommInitRank || !AllGather || !CommDestroy || !GetErrorString) { *err = "libnccl: missing symbols"; return false; }
        return true;
    }
};
NcclApi g_nccl;

template <class T>
struct DevBuf {
    T *p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t n, bool keep = false) {
        if (n <= cap) return cudaSuccess;
        size_t want = n + n / 8 + 64;
        T *np = nullptr;
        cudaError_t e = cudaMalloc(&np, want * sizeof(T));
        if (e != cudaSuccess) return e;
        if (p && keep) e = cudaMemcpy(np, p, cap * sizeof(T), cudaMemcpyDeviceToDevice);
        if (p) cudaFree(p);
        p = np; cap = want;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

struct DfaBufs {
    DevBuf<uint8_t> table, eom, cls, ftable, ccls;
    DevBuf<uint16_t> ctable;
    DevBuf<EventPacked> events;
    DevBuf<uint32_t> wtable, eomw;
    void release() { table.release(); eom.release(); cls.release(); ftable.release(); ccls.release(); ctable.release(); events.release(); wtable.release(); eomw.release(); }
};

struct QItem { int32_t prio; uint64_t ts; uint64_t handle; };
inline bool qless(const QItem &a, const QItem &b) {   // queue.go:22-27 + handle (sequence) as the final tie-break
    if (a.prio != b.prio) return a.prio < b.prio;
    if (a.ts != b.ts) return a.ts < b.ts;
    return a.handle < b.handle;
}
struct LevelQueue {
    std::vector<QItem> items;   // sorted by qless from head
    size_t head = 0;
    int64_t max_size = 0;       // <= 0: unbounded (queue.go:101)
    bool alive = true;
    triage_qstats st{0, 0, 0, 0};
    size_t size() const { return items.size() - head; }
    void compact() { if (head > 4096 && head * 2 > items.size()) { items.erase(items.begin(), items.begin() + head); head = 0; } }
};

// A batch whose host->device copy was started ahead of its submit (triage_prefetch): own device buffers, filled on the
// copy stream, `ready` recorded behind the last copy.  Matched to the later submit by the caller's host pointers.
struct PrefetchSlot {
    DevBuf<uint8_t> bytes;
    DevBuf<uint64_t> offsets, ts;
    DevBuf<uint32_t> lengths;
    DevBuf<triage_meta> meta;
    std::vector<uint64_t> rebased;   // must outlive the asynchronous copy of the offsets
    const void *k_offsets = nullptr, *k_bytes = nullptr, *k_meta = nullptr, *k_lengths = nullptr, *k_ts = nullptr;
    uint32_t n = 0;
    uint64_t total = 0;
    cudaEvent_t ready = nullptr;     // behind the last copy
    cudaEvent_t small_ready = nullptr;   // behind offsets / lengths / meta / times (copied first)
    cudaEvent_t t_begin = nullptr, t_end = nullptr;   // timing events around the copies (h2d_ms)
    std::vector<cudaEvent_t> chunk_ev;   // behind the copy of arena chunk i (kUploadChunk bytes each)
    uint32_t n_chunks = 0;
    bool valid = false;
    uint64_t seq = 0;                // age: the older valid slot is the one a new prefetch may reuse last
};

// The arena travels in chunks so that the scan of chunk i runs while chunk i+1 is still on the PCIe link; the size is a
// multiple of every tile size of the row-lane scan (32 rows of 320 / 640 / 1280 bytes).
constexpr uint64_t kUploadChunk = 800ull * 40960ull;   // 31.25 MiB

// what run_device needs to know about an upload that is still in flight (nullptr: everything is resident)
struct UploadPlan {
    const cudaEvent_t *chunk_ev;
    uint32_t n_chunks;
    cudaEvent_t small_ready, all_ready;
};

double trace_now_ms() {
    static const auto t0 = std::chrono::steady_clock::now();
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}
bool trace_on() {
    static const bool on = getenv("TRIAGE_TRACE") != nullptr;
    return on;
}
#define TRACE(...) do { if (trace_on()) { fprintf(stderr, "[triage %9.3f ms] ", trace_now_ms()); fprintf(stderr, __VA_ARGS__); fputc('\n', stderr); } } while (0)

constexpr int kHeaderWords = 64;   // per-rank merge header (u64): n_adm, 16 counts, 16 first keys, 16 last keys

}  // namespace

struct triage_engine {
    triage_opts opts{};
    int sm_count = 0;
    int max_smem_optin = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[6]{};
    cudaEvent_t user_ev[4]{};
    int32_t default_prio = TRIAGE_PRIO_NORMAL;
    std::unordered_map<std::string, int> user_prio;   // Preprocessor.userPriorities for the JSON ingest
    std::vector<DictEntry> dict;
    RegexSet regex;                // AddKeywordPattern patterns that are not plain literals (regex_vm.h)
    DevBuf<RegexProgram> d_rx_prog;
    DevBuf<RegexInst> d_rx_inst;
    DevBuf<uint64_t> d_trace;
    DevBuf<RegexClass> d_rx_cls;
    DevBuf<uint32_t> d_rx_rng;
    RegexView rxview{};            // device view, n_programs == 0: none
    bool dict_dirty = true;
    Dfa dfa;
    ScanGeometry geom{0, 0, 0};
    uint32_t pad_warps = 0;
    uint32_t row_warps = 0;        // > 0: the padded arena is scanned by the row-lane kernel (2-D TMA), arenas of >= 1 row
    uint32_t pair_warps = 0;       // > 0: ... by its two-bytes-per-lookup variant (filter automaton, scan_pair.cuh)
    uint32_t wide_chains = 0, wide_warps = 0, wide_row = 0;   // > 0: wide dictionaries stream rows from global memory, several per lane
    DfaBufs dm, dx;   // device copies of the scan automaton and of the exact (sequential) one
    DeviceDfa ddfa{};
    // exact (sequential, non-overlapping) automaton, only when the dictionary has self-overlapping literals
    bool has_bordered = false;
    DeviceDfa dxdfa{};
    // batch buffers (triage_submit) and scratch
    DevBuf<uint8_t> d_bytes;
    DevBuf<uint64_t> d_offsets, d_ts, d_pending;
    DevBuf<int64_t> d_maxsize;
    DevBuf<triage_meta> d_meta;
    DevBuf<triage_result> d_results;
    DevBuf<uint32_t> d_lengths, d_counters, d_tile_first, d_order, d_perm_a, d_perm_b, d_blockhist, d_totals, d_small;
    DevBuf<uint8_t> d_bucket, d_bucket2;
    PrefetchSlot pre[2];           // batches uploading ahead of their submit
    PrefetchSlot own;              // the batch a submit uploads itself (same chunked copy, same stream)
    DevBuf<uint32_t> d_tilectr;    // one dynamic-tile counter per upload chunk
    cudaStream_t copy_stream = nullptr;
    uint64_t pre_seq = 0;
    uint32_t *h_small = nullptr;   // pinned: [0] flags, [1] n_admitted, [2..17] bucket counts, [32..] digit hist
    std::vector<uint32_t> h_order;
    std::vector<triage_result> h_results_tmp;
    // host queues
    std::mutex mu;
    std::vector<LevelQueue> q;   // q[0] unused; 1..4 = realtime/high/normal/low; further ids from triage_queue_create
    uint64_t next_handle = 1;
    uint64_t next_stamp = 0;
    // DelayedQueue (delayed_queue.go): items waiting for their ReadyAt, sorted by (ready_at, arrival) from `dhead`
    struct Delayed { uint64_t ready_at, seq, handle; int32_t queue, prio; };
    std::vector<Delayed> delayed;
    size_t dhead = 0;
    uint64_t delayed_seq = 0;
    DevBuf<uint64_t> d_keys;       // device sorter for big requeue / schedule batches
    DevBuf<uint8_t> d_kprio;
    DevBuf<uint32_t> d_korder;
    triage_timings tm{};
    // last device run (for the multi-GPU merge)
    const uint32_t *last_order = nullptr;
    const uint8_t *last_bucket = nullptr;
    const uint64_t *last_ts = nullptr;
    uint32_t last_n = 0, last_n_adm = 0;
    uint32_t last_counts[16]{};
    const uint32_t *last_counts_dev = nullptr, *last_n_adm_dev = nullptr;   // the same numbers where the kernels left them
    // one-exchange merge on its own stream (triage_merge_allgather_async)
    cudaStream_t merge_stream = nullptr;
    cudaEvent_t ev_packed = nullptr;       // the merge has read the last submit's order and counts
    bool packed_pending = false;
    DevBuf<uint32_t> d_gsrc, d_gall;
    DevBuf<uint64_t> d_ntotal;
    uint64_t *h_ntotal = nullptr;          // pinned
    uint64_t merge_capacity = 0;
    bool merge_pending = false;
    // comm
    ncclComm_t comm = nullptr;
    int comm_sms = 0;              // SMs the persistent scan leaves to the exchange's kernels
    int nranks = 1, rank = 0;
    DevBuf<uint64_t> d_run_keys, d_all_keys, d_header, d_all_header;
    DevBuf<uint32_t> d_run_gidx, d_all_gidx;
    uint64_t *h_header = nullptr;   // pinned, nranks * kHeaderWords
};

namespace {

// host automaton -> device buffers (either table format) and the DeviceDfa view the kernels take
int32_t upload_dfa(const Dfa &dfa, DfaBufs *b, DeviceDfa *out) {
    std::vector<EventPacked> ev(dfa.events.size());
    for (size_t i = 0; i < dfa.events.size(); i++) pack_event(dfa.events[i], &ev[i].x);
    DeviceDfa d{};
    d.n_states = dfa.n_states; d.n_events = (uint32_t)ev.size(); d.warm = dfa.warm_bytes;
    d.head_pad = dfa.head_pad; d.tail_pad = dfa.tail_pad; d.reset_pad = dfa.reset_pad;
    d.wide = dfa.wide ? 1 : 0;
    CU(b->events.ensure(ev.size()));
    CU(cudaMemcpy(b->events.p, ev.data(), ev.size() * sizeof(EventPacked), cudaMemcpyHostToDevice));
    d.events = b->events.p;
    if (!dfa.wide) {
        d.start_state = dfa.start_state; d.neutral_state = dfa.neutral_state;
        std::vector<uint8_t> eom(256, 0);
        std::copy(dfa.eom.begin(), dfa.eom.end(), eom.begin());
        // rows padded to kRowBytes so that row r starts at shared-memory bank r (mod 32)
        std::vector<uint8_t> padded((size_t)dfa.n_states * kRowBytes, 0);
        for (uint32_t st = 0; st < dfa.n_states; st++)
            std::memcpy(padded.data() + (size_t)st * kRowBytes, dfa.table.data() + (size_t)st * 256, 512);
        CU(b->table.ensure(padded.size()));
        CU(b->eom.ensure(256));
        CU(cudaMemcpy(b->table.p, padded.data(), padded.size(), cudaMemcpyHostToDevice));
        CU(cudaMemcpy(b->eom.p, eom.data(), 256, cudaMemcpyHostToDevice));
        d.table = b->table.p; d.eom = b->eom.p;
        if (dfa.filter.ok) {   // filter automaton of the pair lanes, same padded row layout
            const std::vector<uint8_t> &fp = dfa.filter.mtable;
            CU(b->ftable.ensure(fp.size()));
            CU(cudaMemcpy(b->ftable.p, fp.data(), fp.size(), cudaMemcpyHostToDevice));
            CU(b->ctable.ensure(dfa.ctable.size()));
            CU(b->ccls.ensure(256));
            CU(cudaMemcpy(b->ctable.p, dfa.ctable.data(), dfa.ctable.size() * 2, cudaMemcpyHostToDevice));
            CU(cudaMemcpy(b->ccls.p, dfa.ccls, 256, cudaMemcpyHostToDevice));
            d.has_filter = 1; d.ftable = b->ftable.p; d.f_states = dfa.filter.m_states; d.f_start = dfa.filter.m_start; d.f_fired = dfa.filter.m_fired_from;
            d.f_mul = dfa.filter.mul; d.f_shift = dfa.filter.shift;
            d.ctable = b->ctable.p; d.ccls = b->ccls.p; d.n_cmp = dfa.n_cmp;
        }
    } else {
        d.start_state = dfa.start_state_w; d.neutral_state = dfa.neutral_state_w;
        d.n_classes = dfa.n_classes;
        CU(b->wtable.ensure(dfa.wtable.size()));
        CU(b->cls.ensure(256));
        CU(b->eomw.ensure(dfa.eom_w.size()));
        CU(cudaMemcpy(b->wtable.p, dfa.wtable.data(), dfa.wtable.size() * 4, cudaMemcpyHostToDevice));
        CU(cudaMemcpy(b->cls.p, dfa.cls_of, 256, cudaMemcpyHostToDevice));
        CU(cudaMemcpy(b->eomw.p, dfa.eom_w.data(), dfa.eom_w.size() * 4, cudaMemcpyHostToDevice));
        d.wtable = b->wtable.p; d.cls_of = b->cls.p; d.eom_w = b->eomw.p;
    }
    *out = d;
    return TRIAGE_OK;
}

int32_t commit_dict(triage_engine *e) {
    if (!e->dict_dirty) return TRIAGE_OK;
    Dfa dfa;
    std::string err = build_dfa(e->dict, &dfa);
    if (!err.empty()) return fail(TRIAGE_E_PATTERN, err);
    if (dfa.warm_bytes > kMaxWarm) return fail(TRIAGE_E_PATTERN, "pattern too long for the scan's warm-up window");
    if (dfa.wide && dfa.reset_pad < 0)
        return fail(TRIAGE_E_PATTERN, "large dictionary without a usable pad byte (every ASCII whitespace occurs in a pattern)");
    // the previous automaton may still be in use by a kernel on the stream
    CU(cudaStreamSynchronize(e->stream));
    DeviceDfa d{};
    int32_t rc = upload_dfa(dfa, &e->dm, &d);
    if (rc) return rc;
    uint32_t pad_warps = d.reset_pad >= 0 ? pick_pad_warps(d, (size_t)e->max_smem_optin) : 0;
    if (const char *w = getenv("TRIAGE_PAD_WARPS")) {   // tuning knob: fewer warps than fit (8, 12, 16, 20, 24)
        uint32_t want = (uint32_t)atoi(w);
        if (pad_warps && want >= 8 && want <= pad_warps && want % 4 == 0 && !d.wide) pad_warps = want;
    }
    if (pad_warps == 0) d.reset_pad = -1;   // the padded kernel would not fit: packed entry points only
    ScanGeometry geom{0, 0, 0};
    if (!dfa.wide) {
        geom = pick_scan_geometry(d, (size_t)e->max_smem_optin);
        if (geom.nt == 0)
            return fail(TRIAGE_E_PATTERN, "dictionary automaton (" + std::to_string(dfa.n_states) + " states) does not fit in shared memory");
    } else if (pad_warps == 0) {
        return fail(TRIAGE_E_PATTERN, "large dictionary: the padded scan does not fit in shared memory");
    }
    e->has_bordered = dict_has_bordered(e->dict);
    if (e->has_bordered) {
        Dfa xd;
        err = build_dfa(e->dict, &xd, true);
        if (!err.empty()) return fail(TRIAGE_E_PATTERN, "exact automaton: " + err);
        rc = upload_dfa(xd, &e->dx, &e->dxdfa);
        if (rc) return rc;
    }
    e->rxview = RegexView{};
    if (!e->regex.empty()) {
        const RegexSet &r = e->regex;
        CU(e->d_rx_prog.ensure(r.programs.size()));
        CU(e->d_rx_inst.ensure(r.insts.size()));
        CU(e->d_rx_cls.ensure(r.classes.size()));
        CU(e->d_rx_rng.ensure(r.ranges.size() + 2));
        CU(cudaMemcpy(e->d_rx_prog.p, r.programs.data(), r.programs.size() * sizeof(RegexProgram), cudaMemcpyHostToDevice));
        CU(cudaMemcpy(e->d_rx_inst.p, r.insts.data(), r.insts.size() * sizeof(RegexInst), cudaMemcpyHostToDevice));
        CU(cudaMemcpy(e->d_rx_cls.p, r.classes.data(), r.classes.size() * sizeof(RegexClass), cudaMemcpyHostToDevice));
        if (!r.ranges.empty()) CU(cudaMemcpy(e->d_rx_rng.p, r.ranges.data(), r.ranges.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
        e->rxview = RegexView{e->d_rx_prog.p, e->d_rx_inst.p, e->d_rx_cls.p, e->d_rx_rng.p, (uint32_t)r.programs.size()};
    }
    e->ddfa = d;
    e->pad_warps = pad_warps;
    e->row_warps = 0;
    if (pad_warps && !dfa.wide) {
        const char *rv = getenv("TRIAGE_ROWS");   // tuning knob: 0 = keep to the mini-tile kernel
        if (!rv || atoi(rv) != 0) e->row_warps = pick_row_warps(d, (size_t)e->max_smem_optin);
    }
    e->pair_warps = e->row_warps ? pick_pair_warps(d, (size_t)e->max_smem_optin) : 0;
    e->wide_chains = e->wide_warps = e->wide_row = 0;
    if (pad_warps && dfa.wide && d.warm == 16) {
        // measurement knobs: TRIAGE_WIDE_STREAM=0 keeps to the mini-tile kernel; TRIAGE_WIDE_CHAINS=1|2|3|4 rows per lane
        // (32, 32|24, 20, 16|12 warps: registers), TRIAGE_WIDE_ROW = row length in bytes (multiple of 16)
        const char *sv = getenv("TRIAGE_WIDE_STREAM");
        if (!sv || atoi(sv) != 0) {
            uint32_t ch = 1, nw = 32, row = 320;   // measured best on the 10 000-keyword set (more chains per lane thrash L1: DESIGN.md 4.1b)
            if (const char *v = getenv("TRIAGE_WIDE_CHAINS")) { uint32_t x = (uint32_t)atoi(v); if (x >= 1 && x <= 4) ch = x; }
            nw = ch <= 2 ? 32 : ch == 3 ? 20 : 16;
            if (const char *v = getenv("TRIAGE_WIDE_WARPS")) { uint32_t x = (uint32_t)atoi(v); if ((ch == 2 && x == 24) || (ch == 4 && x == 12)) nw = x; }
            if (const char *v = getenv("TRIAGE_WIDE_ROW")) { uint32_t x = (uint32_t)atoi(v); if (x >= 48 && x <= 4096 && x % 16 == 0) row = x; }
            e->wide_chains = ch; e->wide_warps = nw; e->wide_row = row;
        }
    }
    e->geom = geom;
    e->dfa = dfa;
    e->dict_dirty = false;
    return TRIAGE_OK;
}

// The whole device pipeline on e->stream.  All pointers are device pointers.
enum RunMode { RUN_ROUTED = 0, RUN_SINGLE_QUEUE = 1, RUN_PROCESS_ONLY = 2 };

// lengths == nullptr: packed arena (message i = bytes[offsets[i], offsets[i+1])).  lengths != nullptr: padded arena
// (16-byte aligned starts, 1..16 reset-pad bytes after every message, offsets[n] = padded total).
int32_t run_device(triage_engine *e, RunMode mode, const uint64_t *offsets, const uint32_t *lengths, const uint8_t *bytes, uint64_t total,
                   const triage_meta *meta, const uint64_t *ts, uint32_t n, triage_result *results, uint32_t *order,
                   const uint64_t pending[4], const int64_t max_size[4], uint32_t *n_admitted, const UploadPlan *plan = nullptr) {
    int32_t rc = commit_dict(e);
    if (rc) return rc;
    if ((reinterpret_cast<uintptr_t>(bytes) & 15) != 0) return fail(TRIAGE_E_ARG, "device arena must be 16-byte aligned");
    const bool padded = lengths != nullptr;
    if (padded && e->ddfa.reset_pad < 0)
        return fail(TRIAGE_E_ARG, "this dictionary has no reset pad byte: use the packed (unpadded) entry points");
    if (!padded && e->ddfa.wide)
        return fail(TRIAGE_E_ARG, "a dictionary this large (more than 256 automaton states) needs the padded-arena entry points");
    if (padded && (total & 15)) return fail(TRIAGE_E_ARG, "padded arena size must be a multiple of 16");
    cudaStream_t s = e->stream;
    uint32_t launches = 0;
    e->tm = triage_timings{};
    *n_admitted = 0;
    if (n == 0) return TRIAGE_OK;
    bool use_rows = false;
    if (padded && e->row_warps > 0) {
        ScanArgs probe{bytes, offsets, total, n, nullptr, nullptr, 0, 0};
        use_rows = scan_rows_usable(probe);
    }
    // an upload in flight is scanned chunk by chunk: the row length is chosen for the chunk, not for the arena
    const bool chunked = plan && use_rows && plan->n_chunks > 1;
    const bool use_pair = use_rows && e->pair_warps > 0;
    const uint32_t row_warps = use_pair ? e->pair_warps : e->row_warps;
    const uint32_t row_bytes = use_rows ? scan_rows_lane_bytes(chunked ? kUploadChunk : total, row_warps, e->sm_count, use_pair) : 0;
    // With a communicator the exchange of step k runs on a second stream under the scan of step k+1, and the scan is one
    // persistent CTA per SM that fills the SM: the exchange's kernels would only get an SM when a scan ends.  A few SMs are
    // left to them (triage_comm_init: TRIAGE_COMM_SMS, default 4).
    const int scan_sms = e->sm_count - (e->comm ? e->comm_sms : 0);
    const uint32_t stream_tile = 32u * e->wide_chains * e->wide_row;
    const bool use_stream = padded && e->wide_chains > 0 && total >= 4ull * stream_tile;   // wide dictionaries: rows streamed from global memory
    const uint32_t tile_bytes = !padded ? e->geom.tile
                              : use_rows ? scan_rows_super_bytes(row_bytes) : use_stream ? stream_tile : kPadMiniTile;
    const uint32_t ntiles = (uint32_t)((total + tile_bytes - 1) / tile_bytes);
    const uint32_t nblocks = (n + kOrderBlock - 1) / kOrderBlock;
    CU(e->d_counters.ensure((size_t)n * CNT_STRIDE));
    CU(e->d_tile_first.ensure((size_t)ntiles + 2));
    CU(e->d_bucket.ensure(n));
    CU(e->d_bucket2.ensure(n));
    CU(e->d_blockhist.ensure((size_t)nblocks * 256));
    CU(e->d_totals.ensure(260));
    CU(e->d_small.ensure(32 + 8 * 256));
    CU(e->d_pending.ensure(4));
    CU(e->d_maxsize.ensure(4));
    OrderScratch sc{e->d_blockhist.p, e->d_totals.p, nblocks};

    if (plan) CU(cudaStreamWaitEvent(s, plan->small_ready, 0));
    if (e->packed_pending) { CU(cudaStreamWaitEvent(s, e->ev_packed, 0)); e->packed_pending = false; }   // an asynchronous merge still reads the last order
    CU(cudaEventRecord(e->ev[0], s));
    ScanArgs sa{bytes, offsets, total, n, e->d_counters.p, e->d_tile_first.p, ntiles, tile_bytes, 0, ntiles};
    CU(launch_prepare(sa, e->d_small.p, 32 + 8 * 256, e->sm_count, s)); launches++;   // tile index + zeroed counters and scratch
    e->tm.scan_variant = use_pair ? 5u : use_rows ? 3u : use_stream ? 6u : padded ? (e->ddfa.wide ? 4u : 2u) : 1u;
    e->tm.scan_row_bytes = use_stream ? e->wide_row : row_bytes;
    e->tm.scan_warps = use_rows ? row_warps : use_stream ? e->wide_warps : padded ? e->pad_warps : e->geom.nt / 32;
    if (ntiles && chunked) {
        // scan of chunk i behind the copy of chunk i: the kernel walks the tiles of the part that has arrived
        CU(e->d_tilectr.ensure(plan->n_chunks));
        CU(cudaMemsetAsync(e->d_tilectr.p, 0, plan->n_chunks * sizeof(uint32_t), s));
        CU(cudaEventRecord(e->ev[1], s));
        const uint32_t tiles_per_chunk = (uint32_t)(kUploadChunk / tile_bytes);
        for (uint32_t c = 0; c < plan->n_chunks; c++) {
            sa.tile_begin = c * tiles_per_chunk;
            sa.tile_end = std::min<uint64_t>(ntiles, (uint64_t)(c + 1) * tiles_per_chunk);
            if (sa.tile_begin >= sa.tile_end) break;
            CU(cudaStreamWaitEvent(s, plan->chunk_ev[c], 0));
            if (use_pair) CU(launch_scan_pair(sa, e->ddfa, row_warps, row_bytes, e->d_tilectr.p + c, scan_sms, s));
            else CU(launch_scan_rows(sa, e->ddfa, row_warps, row_bytes, e->d_tilectr.p + c, scan_sms, s));
            launches++;
        }
        CU(cudaStreamWaitEvent(s, plan->all_ready, 0));
    } else if (ntiles) {
        if (plan) CU(cudaStreamWaitEvent(s, plan->all_ready, 0));
        CU(cudaEventRecord(e->ev[1], s));
#ifdef TRIAGE_SCAN_TRACE_BUILD
        const char *trace_path = use_pair ? getenv("TRIAGE_SCAN_TRACE") : nullptr;   // measurement aid: when each warp of the scan ended
#else
        const char *trace_path = nullptr;
#endif
        const size_t trace_words = 4ull * (size_t)e->sm_count * 32;
        if (trace_path) { CU(e->d_trace.ensure(trace_words)); CU(cudaMemsetAsync(e->d_trace.p, 0, trace_words * 8, s)); sa.trace = e->d_trace.p; }
        if (use_pair) CU(launch_scan_pair(sa, e->ddfa, row_warps, row_bytes, e->d_small.p + 20, scan_sms, s));   // d_small[20]: spare word, zeroed by prepare
        else if (use_rows) CU(launch_scan_rows(sa, e->ddfa, row_warps, row_bytes, e->d_small.p + 20, scan_sms, s));
        else if (use_stream) CU(launch_scan_stream_wide(sa, e->ddfa, e->wide_warps, e->wide_chains, e->wide_row, e->d_small.p + 20, scan_sms, s));
        else if (padded) CU(launch_scan_padded(sa, e->ddfa, e->pad_warps, scan_sms, s));
        else CU(launch_scan(sa, e->ddfa, e->geom, e->sm_count, s));
        launches++;
        if (trace_path) {
            std::vector<uint64_t> h(trace_words);
            CU(cudaMemcpyAsync(h.data(), e->d_trace.p, trace_words * 8, cudaMemcpyDeviceToHost, s));
            CU(cudaStreamSynchronize(s));
            if (FILE *f = fopen(trace_path, "w")) {
                for (size_t w = 0; w < trace_words / 4; w++)
                    if (h[4 * w + 1]) fprintf(f, "%zu %llu %llu %llu %llu\n", w, (unsigned long long)h[4 * w], (unsigned long long)h[4 * w + 1],
                                              (unsigned long long)h[4 * w + 2], (unsigned long long)h[4 * w + 3]);
                fclose(f);
            }
            sa.trace = nullptr;
        }
    } else {
        if (plan) CU(cudaStreamWaitEvent(s, plan->all_ready, 0));
        CU(cudaEventRecord(e->ev[1], s));
    }
    if (e->has_bordered && ntiles) { CU(launch_fixup(bytes, offsets, lengths, n, e->dxdfa, e->d_counters.p, s)); launches++; }
    if (e->rxview.n_programs && ntiles) { CU(launch_regex_count(e->rxview, bytes, offsets, lengths, n, e->d_counters.p, e->sm_count, s)); launches++; }
    CU(cudaEventRecord(e->ev[2], s));
    const int32_t single = mode == RUN_SINGLE_QUEUE ? 1 : 0;
    FinalizeArgs fa{lengths, bytes, e->d_counters.p, meta, offsets, ts, n, e->default_prio, mode == RUN_PROCESS_ONLY ? 0 : e->opts.rules_enabled,
                    single, results, e->d_bucket.p, e->d_small.p};
    CU(launch_finalize(fa, s)); launches++;
    CU(cudaEventRecord(e->ev[3], s));
    if (mode == RUN_PROCESS_ONLY) {
        CU(cudaEventRecord(e->ev[4], s));
        CU(cudaStreamSynchronize(s));
        float ms0 = 0;
        cudaEventElapsedTime(&ms0, e->ev[1], e->ev[2]); e->tm.scan_ms = ms0;
        cudaEventElapsedTime(&ms0, e->ev[2], e->ev[3]); e->tm.finalize_ms = ms0;
        cudaEventElapsedTime(&ms0, e->ev[0], e->ev[4]); e->tm.total_ms = ms0;
        e->tm.launches = launches;
        e->tm.scan_bytes = total + 8ull * (n + 1) + 32ull * n + (padded ? 4ull * n : 0);
        return TRIAGE_OK;
    }
    const uint8_t *bucket = e->d_bucket.p;
    bool bounded = false;
    for (int i = 0; i < 4; i++) if (max_size[i] > 0) bounded = true;
    if (bounded) {
        CU(cudaMemcpyAsync(e->d_pending.p, pending, 4 * sizeof(uint64_t), cudaMemcpyHostToDevice, s));
        CU(cudaMemcpyAsync(e->d_maxsize.p, max_size, 4 * sizeof(int64_t), cudaMemcpyHostToDevice, s));
        CU(launch_admit(e->d_bucket.p, e->d_bucket2.p, results, n, e->d_pending.p, e->d_maxsize.p, single, sc, s)); launches += 3;
        bucket = e->d_bucket2.p;
    }
    // fast path: arrival order is time order, so the dequeue order is one stable 16-way split
    CU(launch_partition_u8(bucket, nullptr, n, 16, true, order, e->d_small.p + 2, sc, s)); launches += 4;
    CU(cudaEventRecord(e->ev[4], s));
    CU(cudaMemcpyAsync(e->h_small, e->d_small.p, 32 * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    CU(cudaMemcpyAsync(e->h_small + 1, e->d_totals.p + 256, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    uint32_t passes = 1;
    if (e->h_small[0] & 1u) {
        // enqueue times are not monotone in arrival order: LSD radix over the time key first (stable, so equal
        // times keep arrival order = sequence), skipping digit positions on which every key agrees
        CU(e->d_perm_a.ensure(n));
        CU(e->d_perm_b.ensure(n));
        CU(launch_key_digit_hist(ts, nullptr, n, e->d_small.p + 32, s)); launches++;
        CU(cudaMemcpyAsync(e->h_small + 32, e->d_small.p + 32, 8 * 256 * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
        CU(cudaStreamSynchronize(s));
        const uint32_t *perm = nullptr;
        uint32_t *bufs[2] = {e->d_perm_a.p, e->d_perm_b.p};
        int which = 0;
        for (uint32_t b = 0; b < 8; b++) {
            bool trivial = false;
            for (uint32_t d = 0; d < 256; d++) if (e->h_small[32 + b * 256 + d] == n) trivial = true;
            if (trivial) continue;
            CU(launch_partition_key(ts, 8 * b, perm, n, bufs[which], sc, s)); launches += 3;
            perm = bufs[which];
            which ^= 1;
            passes++;
        }
        CU(launch_partition_u8(bucket, perm, n, 16, true, order, e->d_small.p + 2, sc, s)); launches += 4;
        CU(cudaEventRecord(e->ev[4], s));
        CU(cudaMemcpyAsync(e->h_small, e->d_small.p, 32 * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
        CU(cudaMemcpyAsync(e->h_small + 1, e->d_totals.p + 256, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
        CU(cudaStreamSynchronize(s));
    }
    *n_admitted = e->h_small[1];
    for (int b = 0; b < 16; b++) e->last_counts[b] = e->h_small[2 + b];
    e->last_order = order; e->last_bucket = bucket; e->last_ts = ts; e->last_n = n; e->last_n_adm = *n_admitted;
    e->last_counts_dev = e->d_small.p + 2; e->last_n_adm_dev = e->d_totals.p + 256;
    float ms = 0;
    cudaEventElapsedTime(&ms, e->ev[1], e->ev[2]); e->tm.scan_ms = ms;
    cudaEventElapsedTime(&ms, e->ev[2], e->ev[3]); e->tm.finalize_ms = ms;
    cudaEventElapsedTime(&ms, e->ev[3], e->ev[4]); e->tm.order_ms = ms;
    cudaEventElapsedTime(&ms, e->ev[0], e->ev[4]); e->tm.total_ms = ms;
    e->tm.launches = launches;
    e->tm.order_passes = passes;
    e->tm.scan_bytes = total + 8ull * (n + 1) + 32ull * n + (padded ? 4ull * n : 0);
    return TRIAGE_OK;
}

uint64_t now_ns() {
    return (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::system_clock::now().time_since_epoch()).count();
}

bool queue_ok(const triage_engine *e, int32_t q) { return q >= 1 && (size_t)q < e->q.size() && e->q[q].alive; }

}  // namespace

extern "C" {

const char *triage_last_error(void) { return g_err.c_str(); }

int32_t triage_create(const triage_opts *opts, triage_engine **out) {
    if (!opts || !out) return fail(TRIAGE_E_ARG, "null argument");
    int ndev = 0;
    CU(cudaGetDeviceCount(&ndev));
    if (opts->device < 0 || opts->device >= ndev) return fail(TRIAGE_E_CUDA, "no such CUDA device");
    CU(cudaSetDevice(opts->device));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, opts->device));
    if (prop.major != 10) return fail(TRIAGE_E_CUDA, std::string("this library is built for sm_100a only, found ") + prop.name);
    triage_engine *e = new triage_engine();
    e->opts = *opts;
    e->sm_count = prop.multiProcessorCount;
    e->max_smem_optin = (int)prop.sharedMemPerBlockOptin;
    cudaError_t ce = cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking);
    for (int i = 0; i < 6 && ce == cudaSuccess; i++) ce = cudaEventCreate(&e->ev[i]);
    for (int i = 0; i < 4 && ce == cudaSuccess; i++) ce = cudaEventCreate(&e->user_ev[i]);
    if (ce == cudaSuccess) ce = cudaStreamCreateWithFlags(&e->copy_stream, cudaStreamNonBlocking);
    for (PrefetchSlot *ps : {&e->pre[0], &e->pre[1], &e->own}) {
        if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&ps->ready, cudaEventDisableTiming);
        if (ce == cudaSuccess) ce = cudaEventCreateWithFlags(&ps->small_ready, cudaEventDisableTiming);
        if (ce == cudaSuccess) ce = cudaEventCreate(&ps->t_begin);
        if (ce == cudaSuccess) ce = cudaEventCreate(&ps->t_end);
    }
    if (ce == cudaSuccess) ce = cudaHostAlloc(reinterpret_cast<void **>(&e->h_small), (32 + 8 * 256) * sizeof(uint32_t), cudaHostAllocDefault);
    if (ce != cudaSuccess) { delete e; return fail(TRIAGE_E_CUDA, cudaGetErrorString(ce)); }
    e->q.resize(5);
    for (int l = 1; l <= 4; l++) e->q[l].max_size = opts->max_queue_size;
    // preprocessor.go:28-43 default dictionary
    triage_dict_reset(e, 1);
    *out = e;
    return TRIAGE_OK;
}

void triage_destroy(triage_engine *e) {
    if (!e) return;
    cudaSetDevice(e->opts.device);
    cudaStreamSynchronize(e->stream);
    if (e->copy_stream) cudaStreamSynchronize(e->copy_stream);
    for (PrefetchSlot *ps : {&e->pre[0], &e->pre[1], &e->own}) {
        ps->bytes.release(); ps->offsets.release(); ps->ts.release(); ps->lengths.release(); ps->meta.release();
        for (cudaEvent_t ev : ps->chunk_ev) cudaEventDestroy(ev);
        for (cudaEvent_t ev : {ps->ready, ps->small_ready, ps->t_begin, ps->t_end}) if (ev) cudaEventDestroy(ev);
    }
    e->d_tilectr.release(); e->d_trace.release();
    if (e->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(e->comm);
    e->d_rx_prog.release(); e->d_rx_inst.release(); e->d_rx_cls.release(); e->d_rx_rng.release();
    e->d_keys.release(); e->d_kprio.release(); e->d_korder.release();
    e->dm.release(); e->dx.release(); e->d_bytes.release(); e->d_offsets.release();
    e->d_ts.release(); e->d_pending.release(); e->d_maxsize.release(); e->d_meta.release(); e->d_results.release(); e->d_counters.release();
    e->d_tile_first.release(); e->d_lengths.release(); e->d_order.release(); e->d_perm_a.release(); e->d_perm_b.release(); e->d_blockhist.release();
    e->d_totals.release(); e->d_small.release(); e->d_bucket.release(); e->d_bucket2.release(); e->d_run_keys.release();
    e->d_all_keys.release(); e->d_header.release(); e->d_all_header.release(); e->d_run_gidx.release(); e->d_all_gidx.release();
    if (e->h_small) cudaFreeHost(e->h_small);
    if (e->h_header) cudaFreeHost(e->h_header);
    if (e->h_ntotal) cudaFreeHost(e->h_ntotal);
    if (e->merge_stream) { cudaStreamSynchronize(e->merge_stream); cudaStreamDestroy(e->merge_stream); }
    if (e->ev_packed) cudaEventDestroy(e->ev_packed);
    e->d_gsrc.release(); e->d_gall.release(); e->d_ntotal.release();
    for (auto &ev : e->ev) if (ev) cudaEventDestroy(ev);
    for (auto &ev : e->user_ev) if (ev) cudaEventDestroy(ev);
    if (e->copy_stream) cudaStreamDestroy(e->copy_stream);
    if (e->stream) cudaStreamDestroy(e->stream);
    delete e;
}

int32_t triage_dict_reset(triage_engine *e, int32_t with_defaults) {
    if (!e) return fail(TRIAGE_E_ARG, "null engine");
    e->dict.clear();
    e->regex.clear();
    if (with_defaults) {
        for (const char *w : {"immediate", "emergency", "asap", "right now"}) e->dict.push_back({TRIAGE_PRIO_REALTIME, true, w});
        for (const char *w : {"urgent", "important", "priority", "critical", "soon"}) e->dict.push_back({TRIAGE_PRIO_HIGH, true, w});
    }
    e->dict_dirty = true;
    return TRIAGE_OK;
}

int32_t triage_dict_add_literal(triage_engine *e, int32_t prio, const uint8_t *lit, uint32_t len, int32_t ci) {
    if (!e || !lit) return fail(TRIAGE_E_ARG, "null argument");
    if (prio < 1 || prio > 4) return fail(TRIAGE_E_PATTERN, "priority must be 1..4");
    std::string s(reinterpret_cast<const char *>(lit), len);
    std::string err = validate_literal(s, ci != 0);
    if (!err.empty()) return fail(TRIAGE_E_PATTERN, err);
    e->dict.push_back({(uint8_t)prio, ci != 0, s});
    e->dict_dirty = true;
    return TRIAGE_OK;
}

int32_t triage_dict_add_pattern(triage_engine *e, int32_t prio, const char *regex, uint32_t len) {
    if (!e || !regex) return fail(TRIAGE_E_ARG, "null argument");
    if (prio < 1 || prio > 4) return fail(TRIAGE_E_PATTERN, "priority must be 1..4");
    bool ci = false;
    std::string lit, err;
    // plain (optionally "(?i)") literals go into the scan automaton; anything else is compiled as a regular expression
    // of the supported subset (regex_vm.h) and counted by the regex kernel
    if (parse_regex_literal(regex, len, &ci, &lit, &err))
        return triage_dict_add_literal(e, prio, reinterpret_cast<const uint8_t *>(lit.data()), (uint32_t)lit.size(), ci ? 1 : 0);
    err = regex_compile(regex, len, (uint32_t)prio, &e->regex);
    if (!err.empty()) return fail(TRIAGE_E_PATTERN, "regexp outside the supported subset: " + err);
    e->dict_dirty = true;
    return TRIAGE_OK;
}

int32_t triage_dict_commit(triage_engine *e) {
    if (!e) return fail(TRIAGE_E_ARG, "null engine");
    CU(cudaSetDevice(e->opts.device));
    return commit_dict(e);
}

int32_t triage_dict_count(triage_engine *e, int32_t prio, uint32_t *n) {
    if (!e || !n) return fail(TRIAGE_E_ARG, "null argument");
    uint32_t c = 0;
    for (auto &d : e->dict) if (d.prio == prio) c++;
    for (auto &r : e->regex.programs) if ((int32_t)r.prio == prio) c++;
    *n = c;
    return TRIAGE_OK;
}

int32_t triage_set_default_priority(triage_engine *e, int32_t prio) {
    if (!e) return fail(TRIAGE_E_ARG, "null engine");
    e->default_prio = prio;   // preprocessor.go:187 accepts any value
    return TRIAGE_OK;
}

int32_t triage_dict_info(triage_engine *e, uint32_t *n_states, uint32_t *warm_bytes, uint32_t *table_bytes) {
    if (!e) return fail(TRIAGE_E_ARG, "null engine");
    CU(cudaSetDevice(e->opts.device));
    int32_t rc = commit_dict(e);
    if (rc) return rc;
    if (n_states) *n_states = e->dfa.n_states;
    if (warm_bytes) *warm_bytes = e->dfa.warm_bytes;
    if (table_bytes) *table_bytes = e->dfa.wide ? e->dfa.n_states * e->dfa.n_classes * 4u : e->dfa.n_states * kRowBytes;
    return TRIAGE_OK;
}

int32_t triage_host_alloc(void **p, uint64_t bytes) {
    if (!p) return fail(TRIAGE_E_ARG, "null argument");
    CU(cudaHostAlloc(p, bytes ? bytes : 1, cudaHostAllocDefault));
    return TRIAGE_OK;
}
int32_t triage_host_free(void *p) {
    CU(cudaFreeHost(p));
    return TRIAGE_OK;
}

int32_t triage_submit_device(triage_engine *e, const uint64_t *offsets_dev, const uint8_t *bytes_dev, uint64_t total_bytes,
                             const triage_meta *meta_dev, const uint64_t *enqueue_ns_dev, uint32_t n,
                             triage_result *results_dev, uint32_t *order_dev, uint32_t *n_admitted) {
    if (!e || !n_admitted || (n && (!offsets_dev || !meta_dev || !results_dev || !order_dev))) return fail(TRIAGE_E_ARG, "null argument");
    if (total_bytes && !bytes_dev) return fail(TRIAGE_E_ARG, "null arena");
    CU(cudaSetDevice(e->opts.device));
    const uint64_t pending[4] = {0, 0, 0, 0};
    const int64_t ms[4] = {e->opts.max_queue_size, e->opts.max_queue_size, e->opts.max_queue_size, e->opts.max_queue_size};
    return run_device(e, RUN_ROUTED, offsets_dev, nullptr, bytes_dev, total_bytes, meta_dev, enqueue_ns_dev, n, results_dev, order_dev,
                      pending, ms, n_admitted);
}

namespace {

// H2D of one batch into the engine's own buffers.  offsets are rebased to start at 0.
int32_t upload_batch(triage_engine *e, const uint64_t *offsets, const uint32_t *lengths, const uint8_t *bytes, const triage_meta *meta,
                     const uint64_t *enqueue_ns, uint32_t n, uint64_t *total_out) {
    const uint64_t total = offsets[n] - offsets[0];
    if (total && !bytes) return fail(TRIAGE_E_ARG, "null arena");
    if (lengths) {
        if ((offsets[0] & 15) || (total & 15)) return fail(TRIAGE_E_ARG, "padded arena: offsets must be multiples of 16");
        CU(e->d_lengths.ensure(n));
    }
    cudaStream_t s = e->stream;
    CU(e->d_bytes.ensure(total + 32));
    CU(e->d_offsets.ensure((size_t)n + 1));
    CU(e->d_meta.ensure(n));
    CU(e->d_results.ensure(n));
    CU(e->d_order.ensure(n));
    if (enqueue_ns) CU(e->d_ts.ensure(n));
    std::vector<uint64_t> rebased;
    const uint64_t *off_src = offsets;
    if (offsets[0] != 0) {
        rebased.resize((size_t)n + 1);
        for (uint32_t i = 0; i <= n; i++) rebased[i] = offsets[i] - offsets[0];
        off_src = rebased.data();
    }
    if (total) CU(cudaMemcpyAsync(e->d_bytes.p, bytes + offsets[0], total, cudaMemcpyHostToDevice, s));
    CU(cudaMemsetAsync(e->d_bytes.p + total, 0, 32, s));   // the last tile's bulk copy rounds up to 16 bytes
    CU(cudaMemcpyAsync(e->d_offsets.p, off_src, ((size_t)n + 1) * 8, cudaMemcpyHostToDevice, s));
    CU(cudaMemcpyAsync(e->d_meta.p, meta, (size_t)n * sizeof(triage_meta), cudaMemcpyHostToDevice, s));
    if (lengths) CU(cudaMemcpyAsync(e->d_lengths.p, lengths, (size_t)n * 4, cudaMemcpyHostToDevice, s));
    if (enqueue_ns) CU(cudaMemcpyAsync(e->d_ts.p, enqueue_ns, (size_t)n * 8, cudaMemcpyHostToDevice, s));
    CU(cudaStreamSynchronize(s));   // `rebased` goes out of scope; also separates the H2D time
    *total_out = total;
    return TRIAGE_OK;
}

}  // namespace

// Starts the host->device copy of one batch into `ps` on the copy stream and returns at once: the small arrays first
// (the tile index needs the offsets), then the arena in kUploadChunk pieces with an event behind each, so that the scan
// can follow the copy chunk by chunk.  The host buffers must stay unchanged until the copy has finished.
static int32_t start_upload(triage_engine *e, PrefetchSlot *ps, const uint64_t *offsets, const uint32_t *lengths, const uint8_t *bytes,
                            const triage_meta *meta, const uint64_t *enqueue_ns, uint32_t n) {
    const uint64_t total = offsets[n] - offsets[0];
    if (total && !bytes) return fail(TRIAGE_E_ARG, "null arena");
    if (lengths && ((offsets[0] & 15) || (total & 15))) return fail(TRIAGE_E_ARG, "padded arena: offsets must be multiples of 16");
    cudaStream_t cs = e->copy_stream;
    ps->valid = false;
    CU(ps->bytes.ensure(total + 32));
    CU(ps->offsets.ensure((size_t)n + 1));
    CU(ps->meta.ensure(n));
    if (lengths) CU(ps->lengths.ensure(n));
    if (enqueue_ns) CU(ps->ts.ensure(n));
    const uint32_t n_chunks = (uint32_t)std::max<uint64_t>(1, (total + kUploadChunk - 1) / kUploadChunk);
    while (ps->chunk_ev.size() < n_chunks) {
        cudaEvent_t ev = nullptr;
        CU(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        ps->chunk_ev.push_back(ev);
    }
    const uint64_t *off_src = offsets;
    if (offsets[0] != 0) {
        CU(cudaStreamSynchronize(cs));   // an earlier copy may still read the previous rebased offsets
        ps->rebased.resize((size_t)n + 1);
        for (uint32_t i = 0; i <= n; i++) ps->rebased[i] = offsets[i] - offsets[0];
        off_src = ps->rebased.data();
    }
    TRACE("upload start: n=%u total=%llu chunks=%u", n, (unsigned long long)total, n_chunks);
    CU(cudaEventRecord(ps->t_begin, cs));
    CU(cudaMemcpyAsync(ps->offsets.p, off_src, ((size_t)n + 1) * 8, cudaMemcpyHostToDevice, cs));
    CU(cudaMemcpyAsync(ps->meta.p, meta, (size_t)n * sizeof(triage_meta), cudaMemcpyHostToDevice, cs));
    if (lengths) CU(cudaMemcpyAsync(ps->lengths.p, lengths, (size_t)n * 4, cudaMemcpyHostToDevice, cs));
    if (enqueue_ns) CU(cudaMemcpyAsync(ps->ts.p, enqueue_ns, (size_t)n * 8, cudaMemcpyHostToDevice, cs));
    CU(cudaEventRecord(ps->small_ready, cs));
    const uint8_t *src = bytes + offsets[0];
    for (uint32_t c = 0; c < n_chunks; c++) {
        const uint64_t lo = (uint64_t)c * kUploadChunk, hi = std::min(total, lo + kUploadChunk);
        if (hi > lo) CU(cudaMemcpyAsync(ps->bytes.p + lo, src + lo, hi - lo, cudaMemcpyHostToDevice, cs));
        if (c + 1 == n_chunks) CU(cudaMemsetAsync(ps->bytes.p + total, 0, 32, cs));   // the last tile's bulk copy rounds up to 16 bytes
        CU(cudaEventRecord(ps->chunk_ev[c], cs));
    }
    CU(cudaEventRecord(ps->t_end, cs));
    CU(cudaEventRecord(ps->ready, cs));
    ps->k_offsets = offsets; ps->k_bytes = bytes; ps->k_meta = meta; ps->k_lengths = lengths; ps->k_ts = enqueue_ns;
    ps->n = n; ps->total = total; ps->n_chunks = n_chunks;
    ps->seq = ++e->pre_seq;
    ps->valid = true;
    return TRIAGE_OK;
}

// triage_prefetch: start the host->device copy of a batch on the copy stream, into a slot of its own
static int32_t prefetch_impl(triage_engine *e, const uint64_t *offsets, const uint32_t *lengths, const uint8_t *bytes,
                             const triage_meta *meta, const uint64_t *enqueue_ns, uint32_t n) {
    if (!e || (n && (!offsets || !meta))) return fail(TRIAGE_E_ARG, "null argument");
    CU(cudaSetDevice(e->opts.device));
    std::lock_guard<std::mutex> lock(e->mu);
    if (n == 0) return TRIAGE_OK;
    PrefetchSlot *ps = &e->pre[0];   // a free slot, else the older one
    if (e->pre[0].valid && (!e->pre[1].valid || e->pre[1].seq < e->pre[0].seq)) ps = &e->pre[1];
    return start_upload(e, ps, offsets, lengths, bytes, meta, enqueue_ns, n);
}

// the slot that holds exactly this batch (same host pointers and sizes), oldest first; nullptr: not prefetched
static PrefetchSlot *find_prefetched(triage_engine *e, const uint64_t *offsets, const uint32_t *lengths, const uint8_t *bytes,
                                     const triage_meta *meta, const uint64_t *enqueue_ns, uint32_t n) {
    PrefetchSlot *best = nullptr;
    for (auto &ps : e->pre) {
        if (!ps.valid || ps.n != n || ps.k_offsets != offsets || ps.k_bytes != bytes || ps.k_meta != meta || ps.k_lengths != lengths ||
            ps.k_ts != enqueue_ns || ps.total != offsets[n] - offsets[0])
            continue;
        if (!best || ps.seq < best->seq) best = &ps;
    }
    return best;
}

static int32_t process_impl(triage_engine *e, const uint64_t *offsets, const uint32_t *lengths, const uint8_t *bytes,
                            const triage_meta *meta, uint32_t n, triage_result *out) {
    if (!e || (n && (!offsets || !meta || !out))) return fail(TRIAGE_E_ARG, "null argument");
    CU(cudaSetDevice(e->opts.device));
    std::lock_guard<std::mutex> lock(e->mu);
    if (n == 0) return TRIAGE_OK;
    auto t0 = std::chrono::steady_clock::now();
    uint64_t total = 0;
    int32_t rc = upload_batch(e, offsets, lengths, bytes, meta, nullptr, n, &total);
    if (rc) return rc;
    auto t1 = std::chrono::steady_clock::now();
    const uint64_t pending[4] = {0, 0, 0, 0};
    const int64_t ms[4] = {0, 0, 0, 0};
    uint32_t n_adm = 0;
    rc = run_device(e, RUN_PROCESS_ONLY, e->d_offsets.p, lengths ? e->d_lengths.p : nullptr, e->d_bytes.p, total, e->d_meta.p, nullptr, n,
                    e->d_results.p, e->d_order.p, pending, ms, &n_adm);
    if (rc) return rc;
    auto t2 = std::chrono::steady_clock::now();
    CU(cudaMemcpyAsync(out, e->d_results.p, (size_t)n * sizeof(triage_result), cudaMemcpyDeviceToHost, e->stream));
    CU(cudaStreamSynchronize(e->stream));
    auto t3 = std::chrono::steady_clock::now();
    e->tm.h2d_ms = std::chrono::duration<float, std::milli>(t1 - t0).count();
    e->tm.d2h_ms = std::chrono::duration<float, std::milli>(t3 - t2).count();
    return TRIAGE_OK;
}

int32_t triage_process(triage_engine *e, const uint64_t *offsets, const uint8_t *bytes, const triage_meta *meta, uint32_t n,
                       triage_result *out) {
    return process_impl(e, offsets, nullptr, bytes, meta, n, out);
}

int32_t triage_process_padded(triage_engine *e, const uint64_t *offsets, const uint32_t *lengths, const uint8_t *bytes,
                              const triage_meta *meta, uint32_t n, triage_result *out) {
    if (n && !lengths) return fail(TRIAGE_E_ARG, "null lengths");
    return process_impl(e, offsets, lengths, bytes, meta, n, out);
}

static int32_t submit_impl(triage_engine *e, int32_t target_queue, const uint64_t *offsets, const uint32_t *lengths, const uint8_t *bytes,
                           const triage_meta *meta, const uint64_t *enqueue_ns, uint32_t n, triage_result *out, uint64_t *handle_base) {
    if (!e || (n && (!offsets || !meta))) return fail(TRIAGE_E_ARG, "null argument");
    CU(cudaSetDevice(e->opts.device));
    std::lock_guard<std::mutex> lock(e->mu);
    if (target_queue != 0 && !queue_ok(e, target_queue)) return fail(TRIAGE_E_NOT_FOUND, "queue not found");
    if (handle_base) *handle_base = e->next_handle;
    if (n == 0) return TRIAGE_OK;
    cudaStream_t s = e->stream;
    auto t0 = std::chrono::steady_clock::now();
    uint64_t total = 0;
    int32_t rc = TRIAGE_OK;
    const uint64_t *d_off = nullptr, *d_tsp = nullptr;
    const uint32_t *d_len = nullptr;
    const uint8_t *d_byt = nullptr;
    const triage_meta *d_met = nullptr;
    PrefetchSlot *ps = find_prefetched(e, offsets, lengths, bytes, meta, enqueue_ns, n);
    const bool prefetched = ps != nullptr;
    TRACE("submit enter: n=%u prefetched=%d", n, prefetched ? 1 : 0);
    if (!ps) {   // not uploaded ahead: start the same chunked copy now, the scan below follows it chunk by chunk
        ps = &e->own;
        rc = start_upload(e, ps, offsets, lengths, bytes, meta, enqueue_ns, n);
        if (rc) return rc;
    }
    CU(e->d_results.ensure(n));
    CU(e->d_order.ensure(n));
    total = ps->total;
    d_off = ps->offsets.p; d_byt = ps->bytes.p; d_met = ps->meta.p;
    d_len = lengths ? ps->lengths.p : nullptr;
    d_tsp = enqueue_ns ? ps->ts.p : nullptr;
    ps->valid = false;   // consumed (the buffers stay untouched until the next upload into this slot, i.e. after this call returns)
    const UploadPlan plan{ps->chunk_ev.data(), ps->n_chunks, ps->small_ready, ps->ready};
    auto t1 = std::chrono::steady_clock::now();
    uint64_t pending[4] = {0, 0, 0, 0};
    int64_t ms[4] = {0, 0, 0, 0};
    if (target_queue == 0) {
        for (int l = 1; l <= 4; l++) { pending[l - 1] = e->q[l].size(); ms[l - 1] = e->q[l].max_size; }
    } else {
        pending[0] = e->q[target_queue].size();
        ms[0] = e->q[target_queue].max_size;
    }
    uint32_t n_adm = 0;
    rc = run_device(e, target_queue ? RUN_SINGLE_QUEUE : RUN_ROUTED, d_off, d_len, d_byt, total, d_met, d_tsp, n, e->d_results.p,
                    e->d_order.p, pending, ms, &n_adm, &plan);
    if (rc) return rc;
    auto t2 = std::chrono::steady_clock::now();
    TRACE("submit: kernels done (scan %.3f ms incl. waiting for chunks)", e->tm.scan_ms);
    triage_result *res = out;
    if (!res) { e->h_results_tmp.resize(n); res = e->h_results_tmp.data(); }
    e->h_order.resize(n_adm);
    CU(cudaMemcpyAsync(res, e->d_results.p, (size_t)n * sizeof(triage_result), cudaMemcpyDeviceToHost, s));
    if (n_adm) CU(cudaMemcpyAsync(e->h_order.data(), e->d_order.p, (size_t)n_adm * 4, cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    auto t3 = std::chrono::steady_clock::now();
    // h2d_ms: the copy engine's time for this batch when the call uploaded it itself (0: it was prefetched, the copy ran
    // under the previous batch); the scan of that case overlaps the copy, so scan_ms includes waiting for chunks
    e->tm.h2d_ms = 0;
    if (!prefetched) { float hm = 0; if (cudaEventElapsedTime(&hm, ps->t_begin, ps->t_end) == cudaSuccess) e->tm.h2d_ms = hm; }
    (void)t0; (void)t1;
    e->tm.d2h_ms = std::chrono::duration<float, std::milli>(t3 - t2).count();
    TRACE("submit: results on the host (d2h %.3f ms)", e->tm.d2h_ms);
    // queue.go:105-116: admitted messages enter their queue; the device order is already
    // (queue, item priority, time, sequence), so each queue receives one sorted run
    uint64_t stamp0 = 0;
    if (!enqueue_ns) { stamp0 = std::max(now_ns(), e->next_stamp); e->next_stamp = stamp0 + n; }
    const uint64_t base = e->next_handle;
    // The device order is bucket-major: bucket b = (level-1)*4 + (prio-1) (routed) or min(prio, 15) (single queue), so every
    // destination queue receives one contiguous slice of `order`, already sorted by (prio, time, sequence); the item
    // priority is a function of the bucket, so res[] need not be read back per item.  Filled in parallel for big batches.
    uint32_t bstart[17];
    bstart[0] = 0;
    for (int b = 0; b < 16; b++) bstart[b + 1] = bstart[b] + e->last_counts[b];
    struct Slice { int q; uint32_t b_lo, b_hi; };   // buckets [b_lo, b_hi) go to queue q
    std::vector<Slice> slices;
    if (target_queue) slices.push_back({target_queue, 0, 16});
    else for (int l = 1; l <= 4; l++) slices.push_back({l, (uint32_t)(l - 1) * 4, (uint32_t)l * 4});
    const uint32_t *ord = e->h_order.data();
    for (const Slice &sl : slices) {
        LevelQueue &lq = e->q[sl.q];
        const uint32_t lo = bstart[sl.b_lo], hi = bstart[sl.b_hi];
        const size_t added = hi - lo;
        if (!added) continue;
        const size_t run_begin = lq.items.size();
        lq.items.resize(run_begin + added);
        QItem *dst = lq.items.data() + run_begin;
        auto fill = [&](uint32_t k0, uint32_t k1) {
            for (uint32_t b = sl.b_lo; b < sl.b_hi; b++) {
                uint32_t s0 = std::max(bstart[b], k0), s1 = std::min(bstart[b + 1], k1);
                const int32_t prio = target_queue ? (int32_t)b : (int32_t)(b & 3) + 1;
                const bool lumped = target_queue && b == 15;   // single-queue mode lumps priorities >= 15 into the last bucket
                for (uint32_t k = s0; k < s1; k++) {
                    uint32_t i = ord[k];
                    dst[k - lo] = QItem{lumped ? (int32_t)res[i].label : prio, enqueue_ns ? enqueue_ns[i] : stamp0 + i, base + i};
                }
            }
        };
        unsigned nthr = added > 65536 ? std::min(8u, std::max(1u, std::thread::hardware_concurrency())) : 1;
        if (nthr <= 1) fill(lo, hi);
        else {
            std::vector<std::thread> th;
            for (unsigned t = 0; t < nthr; t++)
                th.emplace_back(fill, lo + (uint32_t)((uint64_t)added * t / nthr), lo + (uint32_t)((uint64_t)added * (t + 1) / nthr));
            for (auto &x : th) x.join();
        }
        if (target_queue && e->last_counts[15]) std::stable_sort(dst + (bstart[15] - lo), dst + added, qless);
        lq.st.pending += (int64_t)added;
        if (run_begin > lq.head && qless(lq.items[run_begin], lq.items[run_begin - 1]))
            std::inplace_merge(lq.items.begin() + lq.head, lq.items.begin() + run_begin, lq.items.end(), qless);
    }
    e->next_handle = base + n;
    TRACE("submit exit (queues filled)");
    return TRIAGE_OK;
}

int32_t triage_submit(triage_engine *e, int32_t target_queue, const uint64_t *offsets, const uint8_t *bytes, const triage_meta *meta,
                      const uint64_t *enqueue_ns, uint32_t n, triage_result *out, uint64_t *handle_base) {
    return submit_impl(e, target_queue, offsets, nullptr, bytes, meta, enqueue_ns, n, out, handle_base);
}

int32_t triage_submit_padded(triage_engine *e, int32_t target_queue, const uint64_t *offsets, const uint32_t *lengths, const uint8_t *bytes,
                             const triage_meta *meta, const uint64_t *enqueue_ns, uint32_t n, triage_result *out, uint64_t *handle_base) {
    if (n && !lengths) return fail(TRIAGE_E_ARG, "null lengths");
    return submit_impl(e, target_queue, offsets, lengths, bytes, meta, enqueue_ns, n, out, handle_base);
}

int32_t triage_prefetch(triage_engine *e, const uint64_t *offsets, const uint32_t *lengths, const uint8_t *bytes, const triage_meta *meta,
                        const uint64_t *enqueue_ns, uint32_t n) {
    return prefetch_impl(e, offsets, lengths, bytes, meta, enqueue_ns, n);
}

int32_t triage_submit_device_padded(triage_engine *e, const uint64_t *offsets_dev, const uint32_t *lengths_dev, const uint8_t *bytes_dev,
                                    uint64_t total_bytes, const triage_meta *meta_dev, const uint64_t *enqueue_ns_dev, uint32_t n,
                                    triage_result *results_dev, uint32_t *order_dev, uint32_t *n_admitted) {
    if (!e || !n_admitted || (n && (!offsets_dev || !lengths_dev || !meta_dev || !results_dev || !order_dev))) return fail(TRIAGE_E_ARG, "null argument");
    if (total_bytes && !bytes_dev) return fail(TRIAGE_E_ARG, "null arena");
    CU(cudaSetDevice(e->opts.device));
    const uint64_t pending[4] = {0, 0, 0, 0};
    const int64_t ms[4] = {e->opts.max_queue_size, e->opts.max_queue_size, e->opts.max_queue_size, e->opts.max_queue_size};
    return run_device(e, RUN_ROUTED, offsets_dev, lengths_dev, bytes_dev, total_bytes, meta_dev, enqueue_ns_dev, n, results_dev, order_dev,
                      pending, ms, n_admitted);
}

int32_t triage_pad_byte(triage_engine *e, int32_t *pad) {
    if (!e || !pad) return fail(TRIAGE_E_ARG, "null argument");
    CU(cudaSetDevice(e->opts.device));
    int32_t rc = commit_dict(e);
    if (rc) return rc;
    *pad = e->ddfa.reset_pad;
    return TRIAGE_OK;
}

// Host-side arena builder for shims: packed (offsets, bytes) -> padded layout.  dst_bytes == NULL only computes
// dst_offsets / dst_lengths / *dst_total so the caller can size its (pinned) arena.
int32_t triage_pack_padded(const uint64_t *src_offsets, const uint8_t *src_bytes, uint32_t n, int32_t pad, uint64_t *dst_offsets,
                           uint32_t *dst_lengths, uint8_t *dst_bytes, uint64_t *dst_total) {
    if ((n && !src_offsets) || !dst_offsets || !dst_lengths || !dst_total || pad < 0 || pad > 255) return fail(TRIAGE_E_ARG, "bad argument");
    uint64_t o = 0;
    for (uint32_t i = 0; i < n; i++) {
        uint64_t len = src_offsets[i + 1] - src_offsets[i];
        if (len > 0xFFFFFFFFull) return fail(TRIAGE_E_ARG, "message longer than 4 GiB");
        dst_offsets[i] = o;
        dst_lengths[i] = (uint32_t)len;
        o += (len / 16 + 1) * 16;   // 1..16 pad bytes after every message
    }
    dst_offsets[n] = o;
    *dst_total = o;
    if (!dst_bytes) return TRIAGE_OK;
    unsigned hw = std::thread::hardware_concurrency();
    unsigned nthr = hw ? (hw > 32 ? 32 : hw) : 4;
    if (nthr > n) nthr = n ? n : 1;
    std::vector<std::thread> th;
    for (unsigned t = 0; t < nthr; t++) {
        th.emplace_back([=]() {
            uint32_t lo = (uint32_t)((uint64_t)n * t / nthr), hi = (uint32_t)((uint64_t)n * (t + 1) / nthr);
            for (uint32_t i = lo; i < hi; i++) {
                uint64_t len = dst_lengths[i];
                if (len) std::memcpy(dst_bytes + dst_offsets[i], src_bytes + src_offsets[i], len);
                std::memset(dst_bytes + dst_offsets[i] + len, pad, dst_offsets[i + 1] - dst_offsets[i] - len);
            }
        });
    }
    for (auto &x : th) x.join();
    return TRIAGE_OK;
}

int32_t triage_queue_create(triage_engine *e, int64_t max_size, int32_t *queue) {
    if (!e || !queue) return fail(TRIAGE_E_ARG, "null argument");
    std::lock_guard<std::mutex> lock(e->mu);
    e->q.emplace_back();
    e->q.back().max_size = max_size;
    *queue = (int32_t)e->q.size() - 1;
    return TRIAGE_OK;
}

int32_t triage_queue_delete(triage_engine *e, int32_t queue) {
    if (!e) return fail(TRIAGE_E_ARG, "null engine");
    std::lock_guard<std::mutex> lock(e->mu);
    if (!queue_ok(e, queue) || queue <= 4) return fail(TRIAGE_E_NOT_FOUND, "queue not found");
    e->q[queue] = LevelQueue();
    e->q[queue].alive = false;
    return TRIAGE_OK;
}

int32_t triage_set_limits(triage_engine *e, int64_t max_queue_size, int32_t rules_enabled) {
    if (!e) return fail(TRIAGE_E_ARG, "null engine");
    std::lock_guard<std::mutex> lock(e->mu);
    e->opts.max_queue_size = max_queue_size;
    e->opts.rules_enabled = rules_enabled;
    for (int l = 1; l <= 4; l++) e->q[l].max_size = max_queue_size;
    return TRIAGE_OK;
}

int32_t triage_push(triage_engine *e, int32_t level, int32_t item_priority, uint64_t enqueue_ns, uint64_t *handle) {
    if (!e) return fail(TRIAGE_E_ARG, "null engine");
    std::lock_guard<std::mutex> lock(e->mu);
    if (!queue_ok(e, level)) return fail(TRIAGE_E_NOT_FOUND, "queue not found");
    LevelQueue &lq = e->q[level];
    if (lq.max_size > 0 && (int64_t)lq.size() >= lq.max_size) return fail(TRIAGE_E_FULL, "queue is full");
    if (enqueue_ns == 0) { enqueue_ns = std::max(now_ns(), e->next_stamp); e->next_stamp = enqueue_ns + 1; }
    QItem it{item_priority, enqueue_ns, e->next_handle++};
    auto pos = std::upper_bound(lq.items.begin() + lq.head, lq.items.end(), it, qless);
    lq.items.insert(pos, it);
    lq.st.pending++;
    if (handle) *handle = it.handle;
    return TRIAGE_OK;
}

int32_t triage_pop(triage_engine *e, int32_t level, uint64_t *handle) {
    if (!e || !handle) return fail(TRIAGE_E_ARG, "null argument");
    std::lock_guard<std::mutex> lock(e->mu);
    if (!queue_ok(e, level)) return fail(TRIAGE_E_EMPTY, "queue is empty");   // queue.go:126-128: missing queue -> ErrQueueEmpty
    LevelQueue &lq = e->q[level];
    if (lq.size() == 0) return fail(TRIAGE_E_EMPTY, "queue is empty");
    *handle = lq.items[lq.head++].handle;
    lq.st.pending--;
    lq.st.processing++;
    lq.compact();
    return TRIAGE_OK;
}

int32_t triage_peek(triage_engine *e, int32_t level, uint64_t *handle) {
    if (!e || !handle) return fail(TRIAGE_E_ARG, "null argument");
    std::lock_guard<std::mutex> lock(e->mu);
    if (!queue_ok(e, level)) return fail(TRIAGE_E_EMPTY, "queue is empty");
    LevelQueue &lq = e->q[level];
    if (lq.size() == 0) return fail(TRIAGE_E_EMPTY, "queue is empty");
    *handle = lq.items[lq.head].handle;
    return TRIAGE_OK;
}

int32_t triage_drain(triage_engine *e, uint32_t max, uint64_t *handles, uint8_t *levels, uint32_t *n_out) {
    if (!e || !n_out || (max && !handles)) return fail(TRIAGE_E_ARG, "null argument");
    std::lock_guard<std::mutex> lock(e->mu);
    uint32_t k = 0;
    for (int l = 1; l <= 4 && k < max; l++) {   // cmd/queue-manager/main.go:112
        LevelQueue &lq = e->q[l];
        while (k < max && lq.size()) {
            handles[k] = lq.items[lq.head++].handle;
            if (levels) levels[k] = (uint8_t)l;
            lq.st.pending--;
            lq.st.processing++;
            k++;
        }
        lq.compact();
    }
    *n_out = k;
    return TRIAGE_OK;
}

int32_t triage_queue_size(triage_engine *e, int32_t level, uint64_t *size) {
    if (!e || !size) return fail(TRIAGE_E_ARG, "null argument");
    std::lock_guard<std::mutex> lock(e->mu);
    if (!queue_ok(e, level)) return fail(TRIAGE_E_NOT_FOUND, "queue not found");
    *size = e->q[level].size();
    return TRIAGE_OK;
}

int32_t triage_stats(triage_engine *e, int32_t level, triage_qstats *st) {
    if (!e || !st) return fail(TRIAGE_E_ARG, "null argument");
    std::lock_guard<std::mutex> lock(e->mu);
    if (!queue_ok(e, level)) return fail(TRIAGE_E_NOT_FOUND, "queue not found");
    *st = e->q[level].st;
    return TRIAGE_OK;
}

int32_t triage_complete(triage_engine *e, int32_t level) {
    if (!e) return fail(TRIAGE_E_ARG, "null engine");
    std::lock_guard<std::mutex> lock(e->mu);
    if (!queue_ok(e, level)) return TRIAGE_OK;   // updateStats ignores unknown queues (queue.go:192)
    e->q[level].st.processing--;
    e->q[level].st.completed++;
    return TRIAGE_OK;
}

int32_t triage_fail(triage_engine *e, int32_t level) {
    if (!e) return fail(TRIAGE_E_ARG, "null engine");
    std::lock_guard<std::mutex> lock(e->mu);
    if (!queue_ok(e, level)) return TRIAGE_OK;
    e->q[level].st.processing--;
    e->q[level].st.failed++;
    return TRIAGE_OK;
}

}  // extern "C"

namespace {

// Stable order of n items by (bucket, 64-bit key, index) on the device: the LSD radix passes and the bucket split that
// order a submit whose enqueue times are not monotone (run_device), applied to keys that come from the host.  Used for
// large retry / schedule batches; order_out receives the permutation.  bucket == nullptr: one bucket.
int32_t device_sort_keys(triage_engine *e, const uint8_t *bucket, const uint64_t *keys, uint32_t n, uint32_t *order_out) {
    cudaStream_t s = e->stream;
    const uint32_t nblocks = (n + kOrderBlock - 1) / kOrderBlock;
    CU(e->d_keys.ensure(n));
    CU(e->d_kprio.ensure(n));
    CU(e->d_korder.ensure(n));
    CU(e->d_perm_a.ensure(n));
    CU(e->d_perm_b.ensure(n));
    CU(e->d_blockhist.ensure((size_t)nblocks * 256));
    CU(e->d_totals.ensure(260));
    CU(e->d_small.ensure(32 + 8 * 256));
    OrderScratch sc{e->d_blockhist.p, e->d_totals.p, nblocks};
    CU(cudaMemcpyAsync(e->d_keys.p, keys, (size_t)n * 8, cudaMemcpyHostToDevice, s));
    if (bucket) CU(cudaMemcpyAsync(e->d_kprio.p, bucket, n, cudaMemcpyHostToDevice, s));
    else CU(cudaMemsetAsync(e->d_kprio.p, 0, n, s));
    CU(cudaMemsetAsync(e->d_small.p, 0, (32 + 8 * 256) * sizeof(uint32_t), s));
    CU(launch_key_digit_hist(e->d_keys.p, nullptr, n, e->d_small.p + 32, s));
    CU(cudaMemcpyAsync(e->h_small + 32, e->d_small.p + 32, 8 * 256 * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    const uint32_t *perm = nullptr;
    uint32_t *bufs[2] = {e->d_perm_a.p, e->d_perm_b.p};
    int which = 0;
    for (uint32_t b = 0; b < 8; b++) {
        bool trivial = false;
        for (uint32_t d = 0; d < 256; d++) if (e->h_small[32 + b * 256 + d] == n) trivial = true;
        if (trivial) continue;   // every key has the same digit here
        CU(launch_partition_key(e->d_keys.p, 8 * b, perm, n, bufs[which], sc, s));
        perm = bufs[which];
        which ^= 1;
    }
    CU(launch_partition_u8(e->d_kprio.p, perm, n, 16, false, e->d_korder.p, nullptr, sc, s));
    CU(cudaMemcpyAsync(order_out, e->d_korder.p, (size_t)n * 4, cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    return TRIAGE_OK;
}

constexpr uint32_t kDeviceSortMin = 1u << 15;   // below this a host sort is faster than the round trip

// order of items by (bucket, key, index); buckets must be < 16 for the device path
int32_t sort_keys(triage_engine *e, const uint8_t *bucket, const uint64_t *keys, uint32_t n, std::vector<uint32_t> *order) {
    order->resize(n);
    if (n >= kDeviceSortMin) {
        bool fits = true;
        if (bucket) for (uint32_t i = 0; i < n && fits; i++) fits = bucket[i] < 16;
        if (fits && cudaSetDevice(e->opts.device) == cudaSuccess) return device_sort_keys(e, bucket, keys, n, order->data());
    }
    for (uint32_t i = 0; i < n; i++) (*order)[i] = i;
    std::stable_sort(order->begin(), order->end(), [&](uint32_t a, uint32_t b) {
        const uint32_t ba = bucket ? bucket[a] : 0, bb = bucket ? bucket[b] : 0;
        if (ba != bb) return ba < bb;
        return keys[a] < keys[b];
    });
    return TRIAGE_OK;
}

// PushMessage for items that already have handles: capacity in arrival order (queue.go:101-103), then one sorted run
// merged into the queue.  prio: item priorities; ts: enqueue times.  Caller holds e->mu.
int32_t push_run(triage_engine *e, int32_t queue, const uint64_t *handles, const int32_t *prio, const uint64_t *ts, uint32_t n,
                 uint8_t *accepted, uint32_t *n_accepted) {
    LevelQueue &lq = e->q[queue];
    std::vector<uint32_t> adm;
    adm.reserve(n);
    for (uint32_t i = 0; i < n; i++) {
        const bool ok = lq.max_size <= 0 || (int64_t)(lq.size() + adm.size()) < lq.max_size;
        if (accepted) accepted[i] = ok ? 1 : 0;
        if (ok) adm.push_back(i);
    }
    *n_accepted = (uint32_t)adm.size();
    if (adm.empty()) return TRIAGE_OK;
    std::vector<uint8_t> bucket(adm.size());
    std::vector<uint64_t> keys(adm.size());
    bool small_prio = true;
    for (size_t k = 0; k < adm.size(); k++) {
        const int32_t p = prio[adm[k]];
        if (p < 0 || p > 15) small_prio = false;
        bucket[k] = (uint8_t)(p < 0 ? 0 : (p > 15 ? 15 : p));
        keys[k] = ts[adm[k]];
    }
    std::vector<uint32_t> order;
    if (small_prio) {
        int32_t rc = sort_keys(e, bucket.data(), keys.data(), (uint32_t)adm.size(), &order);
        if (rc) return rc;
    } else {   // priorities outside 0..15: plain comparison sort on the real values
        order.resize(adm.size());
        for (uint32_t i = 0; i < order.size(); i++) order[i] = i;
        std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) {
            if (prio[adm[a]] != prio[adm[b]]) return prio[adm[a]] < prio[adm[b]];
            return keys[a] < keys[b];
        });
    }
    const size_t run_begin = lq.items.size();
    lq.items.resize(run_begin + adm.size());
    for (size_t k = 0; k < adm.size(); k++) {
        const uint32_t i = adm[order[k]];
        lq.items[run_begin + k] = QItem{prio[i], ts[i], handles[i]};
    }
    // equal (priority, time): the queue's total order falls back to the handle; the run above is stable by arrival
    std::stable_sort(lq.items.begin() + run_begin, lq.items.end(), qless);
    if (run_begin > lq.head && qless(lq.items[run_begin], lq.items[run_begin - 1]))
        std::inplace_merge(lq.items.begin() + lq.head, lq.items.begin() + run_begin, lq.items.end(), qless);
    lq.st.pending += (int64_t)adm.size();
    return TRIAGE_OK;
}

}  // namespace

extern "C" {

int32_t triage_set_user_priority(triage_engine *e, const char *user_id, uint32_t len, int32_t prio) {
    if (!e || (len && !user_id)) return fail(TRIAGE_E_ARG, "null argument");
    std::lock_guard<std::mutex> lock(e->mu);
    std::string k(user_id ? user_id : "", len);
    if (prio == 0) e->user_prio.erase(k); else e->user_prio[k] = prio;
    return TRIAGE_OK;
}

int32_t triage_ingest_json(triage_engine *e, const uint8_t *bodies, const uint64_t *body_offsets, uint32_t n, uint8_t *arena,
                           uint64_t arena_cap, uint64_t *offsets, uint32_t *lengths, triage_meta *meta, uint8_t *status,
                           uint64_t *arena_total) {
    if (!e) return fail(TRIAGE_E_ARG, "null engine");
    CU(cudaSetDevice(e->opts.device));
    int32_t rc = commit_dict(e);   // the pad byte belongs to the dictionary
    if (rc) return rc;
    std::lock_guard<std::mutex> lock(e->mu);
    auto lookup = [](const void *ctx, const char *id, size_t len) -> int {
        const auto *m = static_cast<const std::unordered_map<std::string, int> *>(ctx);
        auto it = m->find(std::string(id, len));
        return it == m->end() ? 0 : it->second;
    };
    IngestOut out{arena, arena_cap, offsets, lengths, meta, status};
    std::string err = ingest_json(bodies, body_offsets, n, e->ddfa.reset_pad, lookup, &e->user_prio, out, arena_total);
    if (!err.empty()) return fail(TRIAGE_E_ARG, "triage_ingest_json: " + err);
    for (uint32_t i = 0; i < n; i++) meta[i].flags |= TRIAGE_META_ANALYZE;
    return TRIAGE_OK;
}

int32_t triage_emit_analysis_json(const triage_result *res, uint32_t n, char *out, uint64_t cap, uint64_t *out_offsets, uint64_t *needed) {
    if ((n && !res) || !needed) return fail(TRIAGE_E_ARG, "null argument");
    *needed = emit_analysis_json(res, n, out, cap, out_offsets);
    if (out && *needed > cap) return fail(TRIAGE_E_ARG, "output buffer too small");
    return TRIAGE_OK;
}

// QueueManager.BatchPopMessages(queueName, count) (queue_manager.go:326-367): pops until the queue is empty
int32_t triage_drain_queue(triage_engine *e, int32_t queue, uint32_t max, uint64_t *handles, uint32_t *n_out) {
    if (!e || !n_out || (max && !handles)) return fail(TRIAGE_E_ARG, "null argument");
    std::lock_guard<std::mutex> lock(e->mu);
    *n_out = 0;
    if (!queue_ok(e, queue)) return fail(TRIAGE_E_NOT_FOUND, "queue not found");   // :330-333
    LevelQueue &lq = e->q[queue];
    uint32_t k = 0;
    while (k < max && lq.size()) handles[k++] = lq.items[lq.head++].handle;
    lq.st.pending -= k;
    lq.st.processing += k;
    lq.compact();
    *n_out = k;
    return TRIAGE_OK;
}

int32_t triage_complete_batch(triage_engine *e, int32_t queue, uint32_t k) {
    if (!e) return fail(TRIAGE_E_ARG, "null engine");
    std::lock_guard<std::mutex> lock(e->mu);
    if (!queue_ok(e, queue)) return TRIAGE_OK;   // updateStats ignores unknown queues (queue.go:192)
    e->q[queue].st.processing -= k;
    e->q[queue].st.completed += k;
    return TRIAGE_OK;
}

int32_t triage_fail_batch(triage_engine *e, int32_t queue, uint32_t k) {
    if (!e) return fail(TRIAGE_E_ARG, "null engine");
    std::lock_guard<std::mutex> lock(e->mu);
    if (!queue_ok(e, queue)) return TRIAGE_OK;
    e->q[queue].st.processing -= k;
    e->q[queue].st.failed += k;
    return TRIAGE_OK;
}

int32_t triage_requeue(triage_engine *e, int32_t queue, const uint64_t *handles, const int32_t *item_priority, const uint64_t *enqueue_ns,
                       uint32_t n, uint8_t *accepted, uint32_t *n_accepted) {
    if (!e || !n_accepted || (n && (!handles || !item_priority))) return fail(TRIAGE_E_ARG, "null argument");
    std::lock_guard<std::mutex> lock(e->mu);
    *n_accepted = 0;
    if (!queue_ok(e, queue)) return fail(TRIAGE_E_NOT_FOUND, "queue not found");
    if (n == 0) return TRIAGE_OK;
    std::vector<uint64_t> stamped;
    if (!enqueue_ns) {   // queue.go:108: time.Now() inside Push, strictly increasing in push order
        stamped.resize(n);
        const uint64_t t0 = std::max(now_ns(), e->next_stamp);
        for (uint32_t i = 0; i < n; i++) stamped[i] = t0 + i;
        e->next_stamp = t0 + n;
        enqueue_ns = stamped.data();
    }
    return push_run(e, queue, handles, item_priority, enqueue_ns, n, accepted, n_accepted);
}

int32_t triage_schedule(triage_engine *e, int32_t queue, const uint64_t *handles, const int32_t *item_priority, const uint64_t *ready_at_ns,
                        uint32_t n) {
    if (!e || (n && (!handles || !item_priority || !ready_at_ns))) return fail(TRIAGE_E_ARG, "null argument");
    std::lock_guard<std::mutex> lock(e->mu);
    if (!queue_ok(e, queue)) return fail(TRIAGE_E_NOT_FOUND, "queue not found");
    if (n == 0) return TRIAGE_OK;
    std::vector<uint32_t> order;
    int32_t rc = sort_keys(e, nullptr, ready_at_ns, n, &order);   // DelayedQueue.Less: ReadyAt only (delayed_queue.go:37-39)
    if (rc) return rc;
    if (e->dhead > 4096 && e->dhead * 2 > e->delayed.size()) { e->delayed.erase(e->delayed.begin(), e->delayed.begin() + e->dhead); e->dhead = 0; }
    const size_t run_begin = e->delayed.size();
    e->delayed.resize(run_begin + n);
    const uint64_t seq0 = e->delayed_seq;
    for (uint32_t k = 0; k < n; k++) {
        const uint32_t i = order[k];
        e->delayed[run_begin + k] = triage_engine::Delayed{ready_at_ns[i], seq0 + i, handles[i], queue, item_priority[i]};
    }
    e->delayed_seq += n;
    auto less = [](const triage_engine::Delayed &a, const triage_engine::Delayed &b) {
        if (a.ready_at != b.ready_at) return a.ready_at < b.ready_at;
        return a.seq < b.seq;
    };
    if (run_begin > e->dhead && less(e->delayed[run_begin], e->delayed[run_begin - 1]))
        std::inplace_merge(e->delayed.begin() + e->dhead, e->delayed.begin() + run_begin, e->delayed.end(), less);
    return TRIAGE_OK;
}

int32_t triage_delayed_size(triage_engine *e, uint64_t *n) {
    if (!e || !n) return fail(TRIAGE_E_ARG, "null argument");
    std::lock_guard<std::mutex> lock(e->mu);
    *n = e->delayed.size() - e->dhead;
    return TRIAGE_OK;
}

int32_t triage_delayed_peek(triage_engine *e, uint64_t *handle, uint64_t *ready_at_ns) {
    if (!e || !handle || !ready_at_ns) return fail(TRIAGE_E_ARG, "null argument");
    std::lock_guard<std::mutex> lock(e->mu);
    if (e->delayed.size() == e->dhead) return fail(TRIAGE_E_EMPTY, "nothing is scheduled");
    *handle = e->delayed[e->dhead].handle;
    *ready_at_ns = e->delayed[e->dhead].ready_at;
    return TRIAGE_OK;
}

int32_t triage_release_due(triage_engine *e, uint64_t now, uint64_t *handles, int32_t *queues, uint32_t max, uint32_t *n_released,
                           uint32_t *n_rejected) {
    if (!e || !n_released) return fail(TRIAGE_E_ARG, "null argument");
    std::lock_guard<std::mutex> lock(e->mu);
    *n_released = 0;
    if (n_rejected) *n_rejected = 0;
    if (now == 0) now = now_ns();
    // the run loop of delayed_queue.go:113-170 hands every item whose ReadyAt has passed to processFn, earliest first;
    // processFn here is PushMessage into the queue the item was scheduled for, stamped at the release
    size_t end = e->dhead;
    while (end < e->delayed.size() && e->delayed[end].ready_at <= now && (max == 0 || end - e->dhead < max)) end++;
    const uint32_t due = (uint32_t)(end - e->dhead);
    if (!due) return TRIAGE_OK;
    uint64_t t0 = std::max(now_ns(), e->next_stamp);
    e->next_stamp = t0 + due;
    uint32_t released = 0, rejected = 0;
    // group by destination queue, keeping the ReadyAt order inside each
    std::vector<int32_t> qs;
    for (size_t k = e->dhead; k < end; k++) if (std::find(qs.begin(), qs.end(), e->delayed[k].queue) == qs.end()) qs.push_back(e->delayed[k].queue);
    for (int32_t q : qs) {
        std::vector<uint64_t> h, ts;
        std::vector<int32_t> pr;
        for (size_t k = e->dhead; k < end; k++) if (e->delayed[k].queue == q) { h.push_back(e->delayed[k].handle); pr.push_back(e->delayed[k].prio); ts.push_back(t0 + (k - e->dhead)); }
        if (!queue_ok(e, q)) { rejected += (uint32_t)h.size(); continue; }   // the queue was deleted meanwhile: PushMessage fails
        std::vector<uint8_t> acc(h.size());
        uint32_t na = 0;
        int32_t rc = push_run(e, q, h.data(), pr.data(), ts.data(), (uint32_t)h.size(), acc.data(), &na);
        if (rc) return rc;
        for (size_t k = 0; k < h.size(); k++) {
            if (!acc[k]) { rejected++; continue; }
            if (handles && (max == 0 || released < max)) { handles[released] = h[k]; if (queues) queues[released] = q; }
            released++;
        }
    }
    e->dhead = end;
    *n_released = released;
    if (n_rejected) *n_rejected = rejected;
    return TRIAGE_OK;
}

// ExponentialBackoff.NextBackoff (worker.go:274-290): float64 arithmetic exactly as written there
int64_t triage_backoff_ns(int64_t initial_ns, int64_t max_ns, double factor, int32_t retry_count) {
    if (retry_count <= 0) return initial_ns;
    double backoff = (double)initial_ns;
    for (int32_t i = 1; i < retry_count; i++) backoff *= factor;
    if (backoff > (double)max_ns) return max_ns;
    return (int64_t)backoff;
}

int32_t triage_last_timings(triage_engine *e, triage_timings *t) {
    if (!e || !t) return fail(TRIAGE_E_ARG, "null argument");
    *t = e->tm;
    return TRIAGE_OK;
}

int32_t triage_event_record(triage_engine *e, int32_t slot) {
    if (!e || slot < 0 || slot > 3) return fail(TRIAGE_E_ARG, "bad argument");
    CU(cudaSetDevice(e->opts.device));
    CU(cudaEventRecord(e->user_ev[slot], e->stream));
    return TRIAGE_OK;
}

int32_t triage_event_elapsed_ms(triage_engine *e, int32_t from_slot, int32_t to_slot, float *ms) {
    if (!e || !ms || from_slot < 0 || from_slot > 3 || to_slot < 0 || to_slot > 3) return fail(TRIAGE_E_ARG, "bad argument");
    CU(cudaSetDevice(e->opts.device));
    CU(cudaEventSynchronize(e->user_ev[to_slot]));
    CU(cudaEventElapsedTime(ms, e->user_ev[from_slot], e->user_ev[to_slot]));
    return TRIAGE_OK;
}

// ------------------------------------------------------------------------------ multi-GPU ---
int32_t triage_comm_unique_id(uint8_t id[128]) {
    std::string err;
    if (!g_nccl.load(&err)) return fail(TRIAGE_E_CUDA, err);
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId size");
    ncclUniqueId u;
    ncclResult_t r = g_nccl.GetUniqueId(&u);
    if (r != ncclSuccess) return fail(TRIAGE_E_CUDA, std::string("ncclGetUniqueId: ") + g_nccl.GetErrorString(r));
    std::memcpy(id, &u, 128);
    return TRIAGE_OK;
}

int32_t triage_comm_init(triage_engine *e, int32_t nranks, int32_t rank, const uint8_t id[128]) {
    if (!e || !id || nranks < 1 || rank < 0 || rank >= nranks) return fail(TRIAGE_E_ARG, "bad argument");
    std::string err;
    if (!g_nccl.load(&err)) return fail(TRIAGE_E_CUDA, err);
    CU(cudaSetDevice(e->opts.device));
    ncclUniqueId u;
    std::memcpy(&u, id, 128);
    // The exchange runs on a second stream under the next step's scan, and the scan is one persistent CTA per SM: every SM
    // a collective's CTAs hold is an SM whose scan CTA starts late.  The order all-gather is small (4 bytes a message),
    // so the communicator is capped at a few CTAs (TRIAGE_NCCL_CTAS, 0 = NCCL's own choice).
    // (measured on 8 B200s, 1 Mi messages per rank: 0 / 4 / 8 / 12 SMs left free -> 1.46 / 1.31 / 1.30 / 1.20 ms per step:
    // at 8 ranks the exchange chain - 32 MB gathered and re-ordered on every rank - is what a step waits for)
    int ctas = nranks >= 8 ? 8 : 4;
    if (const char *v = getenv("TRIAGE_NCCL_CTAS")) ctas = atoi(v);
    e->comm_sms = nranks >= 8 ? 12 : 4;
    if (const char *v = getenv("TRIAGE_COMM_SMS")) { int x = atoi(v); if (x >= 0 && x <= 32) e->comm_sms = x; }
    ncclResult_t r;
    if (g_nccl.CommInitRankConfig && ctas > 0) {
        ncclConfig_t cfg = NCCL_CONFIG_INITIALIZER;
        cfg.minCTAs = 1;
        cfg.maxCTAs = ctas;
        r = g_nccl.CommInitRankConfig(&e->comm, nranks, u, rank, &cfg);
    } else {
        r = g_nccl.CommInitRank(&e->comm, nranks, u, rank);
    }
    if (r != ncclSuccess) return fail(TRIAGE_E_CUDA, std::string("ncclCommInitRank: ") + g_nccl.GetErrorString(r));
    e->nranks = nranks;
    e->rank = rank;
    CU(cudaHostAlloc(reinterpret_cast<void **>(&e->h_header), (size_t)nranks * kHeaderWords * 8, cudaHostAllocDefault));
    CU(cudaStreamCreateWithFlags(&e->merge_stream, cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&e->ev_packed, cudaEventDisableTiming));
    CU(cudaHostAlloc(reinterpret_cast<void **>(&e->h_ntotal), 8, cudaHostAllocDefault));
    CU(e->d_ntotal.ensure(1));
    return TRIAGE_OK;
}

}  // extern "C"

// Contiguous sequence shards with engine-stamped enqueue times (the normal multi-GPU submit): inside every bucket rank q's
// messages precede rank q+1's, so the merged order is offset arithmetic over the ranks' bucket counts.  One kernel packs
// [counts | local order], ONE all-gather moves it, one kernel concatenates with the geometry worked out on the device -
// all on the merge stream, no host round trip; the next submit only waits until its predecessor's order has been packed.
static int32_t merge_one_exchange(triage_engine *e, const uint64_t *rank_base, uint32_t *order_global_dev, uint64_t capacity) {
    cudaStream_t ms = e->merge_stream;
    const uint32_t R = (uint32_t)e->nranks, n = e->last_n;
    const uint32_t stride = kMergeHeaderWords + n;
    CU(e->d_gsrc.ensure(stride));
    CU(e->d_gall.ensure((size_t)R * stride));
    MergeBases mb{};
    mb.n_ranks = R; mb.stride = stride;
    for (uint32_t q = 0; q < R; q++) mb.base[q] = (uint32_t)rank_base[q];
    CU(launch_pack_run(e->last_order, e->last_counts_dev, e->last_n_adm_dev, n, e->d_gsrc.p, ms));
    CU(cudaEventRecord(e->ev_packed, ms));
    e->packed_pending = true;
    ncclResult_t r = g_nccl.AllGather(e->d_gsrc.p, e->d_gall.p, stride, ncclUint32, e->comm, ms);
    if (r != ncclSuccess) return fail(TRIAGE_E_CUDA, std::string("ncclAllGather(run): ") + g_nccl.GetErrorString(r));
    CU(launch_concat_gathered(e->d_gall.p, mb, n, order_global_dev, capacity, e->d_ntotal.p, ms));
    CU(cudaMemcpyAsync(e->h_ntotal, e->d_ntotal.p, 8, cudaMemcpyDeviceToHost, ms));
    e->merge_capacity = capacity;
    e->merge_pending = true;
    return TRIAGE_OK;
}

extern "C" int32_t triage_merge_wait(triage_engine *e, uint64_t *n_total) {
    if (!e) return fail(TRIAGE_E_ARG, "null engine");
    if (!e->merge_pending) { if (n_total) *n_total = 0; return TRIAGE_OK; }
    CU(cudaSetDevice(e->opts.device));
    CU(cudaStreamSynchronize(e->merge_stream));
    e->merge_pending = false;
    e->packed_pending = false;
    if (n_total) *n_total = *e->h_ntotal;
    if (*e->h_ntotal > e->merge_capacity) return fail(TRIAGE_E_ARG, "order_global_dev is too small");
    return TRIAGE_OK;
}

extern "C" int32_t triage_merge_allgather_async(triage_engine *e, const uint64_t *rank_base, uint32_t *order_global_dev, uint64_t capacity) {
    if (!e || !rank_base || !order_global_dev) return fail(TRIAGE_E_ARG, "null argument");
    if (!e->comm) return fail(TRIAGE_E_ARG, "triage_comm_init has not been called");
    if (e->nranks > 8) return fail(TRIAGE_E_ARG, "at most 8 ranks");
    if (e->last_ts != nullptr || !e->last_counts_dev)
        return fail(TRIAGE_E_ARG, "the asynchronous merge needs engine-stamped enqueue times (pass no enqueue_ns to the submit); use triage_merge_allgather");
    CU(cudaSetDevice(e->opts.device));
    return merge_one_exchange(e, rank_base, order_global_dev, capacity);
}

extern "C" int32_t triage_merge_allgather(triage_engine *e, const uint64_t *rank_base, uint32_t *order_global_dev, uint64_t capacity,
                                          uint64_t *n_total) {
    if (!e || !rank_base || !order_global_dev || !n_total) return fail(TRIAGE_E_ARG, "null argument");
    if (!e->comm) return fail(TRIAGE_E_ARG, "triage_comm_init has not been called");
    if (e->nranks > 8) return fail(TRIAGE_E_ARG, "at most 8 ranks");
    CU(cudaSetDevice(e->opts.device));
    if (e->last_ts == nullptr && e->last_counts_dev && !getenv("TRIAGE_MERGE_TWO_STEP")) {   // contiguous shards: one exchange
        if (e->merge_pending) { int32_t rc0 = triage_merge_wait(e, nullptr); if (rc0) return rc0; }
        int32_t rc = merge_one_exchange(e, rank_base, order_global_dev, capacity);
        if (rc) return rc;
        return triage_merge_wait(e, n_total);
    }
    cudaStream_t s = e->stream;
    const uint32_t R = (uint32_t)e->nranks;
    const uint32_t n_adm = e->last_n_adm;
    // 1. header of my run (counts and first/last time key per bucket), gathered from every rank
    CU(e->d_header.ensure(kHeaderWords));
    CU(e->d_all_header.ensure((size_t)R * kHeaderWords));
    CU(launch_header(e->last_order, n_adm, e->last_counts, e->last_ts, rank_base[e->rank], e->d_header.p, s));
    ncclResult_t r = g_nccl.AllGather(e->d_header.p, e->d_all_header.p, kHeaderWords, ncclUint64, e->comm, s);
    if (r != ncclSuccess) return fail(TRIAGE_E_CUDA, std::string("ncclAllGather(header): ") + g_nccl.GetErrorString(r));
    CU(cudaMemcpyAsync(e->h_header, e->d_all_header.p, (size_t)R * kHeaderWords * 8, cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    // 2. geometry
    ConcatTables t{};
    t.n_ranks = R;
    uint64_t total = 0;
    uint32_t cap = 1;
    for (uint32_t q = 0; q < R; q++) {
        t.cnt[q] = (uint32_t)e->h_header[q * kHeaderWords];
        t.base[q] = (uint32_t)rank_base[q];
        total += t.cnt[q];
        cap = std::max(cap, t.cnt[q]);
    }
    t.cap = cap;
    *n_total = total;
    if (total > capacity) return fail(TRIAGE_E_ARG, "order_global_dev is too small");
    if (cap > e->last_n) return fail(TRIAGE_E_ARG, "ranks must submit equally sized batches (order buffers are gathered at a common stride)");
    bool disjoint = true;   // per bucket, rank q's keys all <= rank q+1's keys?
    uint32_t dst = 0;
    for (uint32_t b = 0; b < 16; b++) {
        uint64_t prev_last = 0; bool have_prev = false;
        for (uint32_t q = 0; q < R; q++) {
            const uint64_t *h = e->h_header + (size_t)q * kHeaderWords;
            uint32_t c = (uint32_t)h[1 + b];
            t.dst[q][b] = dst;
            dst += c;
            if (c) {
                if (have_prev && h[17 + b] < prev_last) disjoint = false;
                prev_last = h[33 + b]; have_prev = true;
            }
        }
    }
    for (uint32_t q = 0; q < R; q++) {
        uint32_t acc = 0;
        for (uint32_t b = 0; b < 16; b++) { t.lstart[q][b] = acc; acc += (uint32_t)e->h_header[(size_t)q * kHeaderWords + 1 + b]; }
    }
    // 3. all-gather the local dequeue orders (4 bytes per message; the caller's order buffer holds >= cap entries)
    CU(e->d_all_gidx.ensure((size_t)R * cap));
    r = g_nccl.AllGather(e->last_order, e->d_all_gidx.p, cap, ncclUint32, e->comm, s);
    if (r != ncclSuccess) return fail(TRIAGE_E_CUDA, std::string("ncclAllGather(order): ") + g_nccl.GetErrorString(r));
    // 4. merge: when the per-bucket key ranges of the ranks do not interleave (contiguous sequence shards, the
    //    normal case) the merged order is pure offset arithmetic; otherwise gather the time keys too and rank every
    //    key by binary search in the other runs
    if (disjoint) {
        CU(launch_concat(e->d_all_gidx.p, t, order_global_dev, s));
    } else {
        CU(e->d_run_keys.ensure(cap));
        CU(e->d_run_gidx.ensure(cap));
        CU(e->d_all_keys.ensure((size_t)R * cap));
        CU(launch_build_run(e->last_order, n_adm, e->last_bucket, e->last_ts, rank_base[e->rank], rank_base[e->rank], e->d_run_keys.p,
                            e->d_run_gidx.p, s));
        r = g_nccl.AllGather(e->d_run_keys.p, e->d_all_keys.p, cap, ncclUint64, e->comm, s);
        if (r != ncclSuccess) return fail(TRIAGE_E_CUDA, std::string("ncclAllGather(keys): ") + g_nccl.GetErrorString(r));
        r = g_nccl.AllGather(e->d_run_gidx.p, e->d_all_gidx.p, cap, ncclUint32, e->comm, s);
        if (r != ncclSuccess) return fail(TRIAGE_E_CUDA, std::string("ncclAllGather(gidx): ") + g_nccl.GetErrorString(r));
        CU(launch_merge_strided(e->d_all_keys.p, e->d_all_gidx.p, t, order_global_dev, s));
    }
    CU(cudaStreamSynchronize(s));
    return TRIAGE_OK;
}

// =================================================================================================
// triage_submit_device_static — the device-resident submit for callers that must not allocate.
//
// Who needs it: code that runs where cudaMalloc / cudaFree are off limits — a host thread inside a
// latency-critical serving loop that may not take the allocator's implicit device synchronisation, a
// plugin whose embedding application owns all device memory and forbids growth after start-up, or a
// process that has already handed every byte of HBM to its own pools.  The regular entry points grow
// their scratch buffers on demand (cudaMalloc inside DevBuf::ensure); this one never does: all the
// scratch the kernels need (per-message counters, tile index, buckets, split histograms, permutation
// ping-pong buffers) is carved out of ONE file-scope __device__ array below.
//
// Cost and limits: every GPU that loads this library keeps kStaticScratchBytes (3 MiB) of device memory
// for the life of the process, and a batch may hold at most kStaticMaxMsgs messages / kStaticMaxBytes
// arena bytes; larger batches are refused with TRIAGE_E_ARG.  The dictionary must already be committed
// (triage_dict_commit allocates its table, so it is not done here).
//
// Set-up: the array's name is only a host-side handle.  The first call on each device resolves the real
// device address with cudaGetSymbolAddress (with that device current) and caches it; make that first
// call before starting any stream capture.
//
// Sharing: all calls, from every engine on that device, use the same array and it keeps whatever the last
// call left behind.  Each call clears the parts that must start at zero (counters, flags), so back-to-back
// calls are fine; calls that overlap in time on one GPU race on the scratch and must be serialised by the
// caller.
// =================================================================================================
namespace {

constexpr uint32_t kStaticMaxMsgs = 65536;
constexpr uint64_t kStaticMaxBytes = 256ull << 20;
constexpr size_t kStaticScratchBytes = 3u << 20;
constexpr size_t kStaticAlign = 256;

constexpr size_t up256(size_t x) { return (x + kStaticAlign - 1) / kStaticAlign * kStaticAlign; }
constexpr size_t kStMaxTiles = kStaticMaxBytes / 16384 + 2;   // holds for any scan tile of >= 16 KiB
constexpr size_t kStMaxBlocks = (kStaticMaxMsgs + kOrderBlock - 1) / kOrderBlock;
constexpr size_t kStOffCounters = 0;
constexpr size_t kStOffTileFirst = kStOffCounters + up256((size_t)kStaticMaxMsgs * CNT_STRIDE * 4);
constexpr size_t kStOffBucket = kStOffTileFirst + up256(kStMaxTiles * 4);
constexpr size_t kStOffBucket2 = kStOffBucket + up256(kStaticMaxMsgs);
constexpr size_t kStOffBlockhist = kStOffBucket2 + up256(kStaticMaxMsgs);
constexpr size_t kStOffTotals = kStOffBlockhist + up256(kStMaxBlocks * 256 * 4);
constexpr size_t kStOffSmall = kStOffTotals + up256(260 * 4);
constexpr size_t kStOffPending = kStOffSmall + up256((32 + 8 * 256) * 4);
constexpr size_t kStOffMaxsize = kStOffPending + up256(4 * 8);
constexpr size_t kStOffPermA = kStOffMaxsize + up256(4 * 8);
constexpr size_t kStOffPermB = kStOffPermA + up256((size_t)kStaticMaxMsgs * 4);
constexpr size_t kStEnd = kStOffPermB + up256((size_t)kStaticMaxMsgs * 4);
static_assert(kStEnd <= kStaticScratchBytes, "static scratch layout does not fit");
static_assert(kMinScanTile >= 16384, "tile index bound assumes tiles of at least 16 KiB");

static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

std::mutex g_static_mu;
unsigned char *g_static_ptr[64] = {};   // per device ordinal: device address of g_static_scratch

// once per device, with that device current
int32_t static_scratch_base(int device, unsigned char **out) {
    if (device < 0 || device >= 64) return fail(TRIAGE_E_ARG, "device ordinal out of range for the static scratch table");
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (!g_static_ptr[device]) {
        void *p = nullptr;
        CU(cudaGetSymbolAddress(&p, g_static_scratch));
        g_static_ptr[device] = static_cast<unsigned char *>(p);
    }
    *out = g_static_ptr[device];
    return TRIAGE_OK;
}

}  // namespace

extern "C" int32_t triage_submit_device_static(triage_engine *e, const uint64_t *offsets_dev, const uint8_t *bytes_dev,
                                               uint64_t total_bytes, const triage_meta *meta_dev, const uint64_t *enqueue_ns_dev,
                                               uint32_t n, triage_result *results_dev, uint32_t *order_dev, uint32_t *n_admitted) {
    if (!e || !n_admitted || (n && (!offsets_dev || !meta_dev || !results_dev || !order_dev))) return fail(TRIAGE_E_ARG, "null argument");
    if (total_bytes && !bytes_dev) return fail(TRIAGE_E_ARG, "null arena");
    if (n > kStaticMaxMsgs)
        return fail(TRIAGE_E_ARG, "triage_submit_device_static: batch of " + std::to_string(n) + " messages exceeds the static scratch limit of " +
                                      std::to_string(kStaticMaxMsgs));
    if (total_bytes > kStaticMaxBytes)
        return fail(TRIAGE_E_ARG, "triage_submit_device_static: arena of " + std::to_string(total_bytes) + " bytes exceeds the static scratch limit of " +
                                      std::to_string(kStaticMaxBytes));
    if (e->dict_dirty) return fail(TRIAGE_E_ARG, "triage_submit_device_static: call triage_dict_commit first (committing allocates, this entry point must not)");
    if (e->ddfa.wide) return fail(TRIAGE_E_ARG, "triage_submit_device_static: large dictionaries need the padded-arena entry points");
    if ((reinterpret_cast<uintptr_t>(bytes_dev) & 15) != 0) return fail(TRIAGE_E_ARG, "device arena must be 16-byte aligned");
    CU(cudaSetDevice(e->opts.device));
    unsigned char *base = nullptr;
    int32_t rc = static_scratch_base(e->opts.device, &base);
    if (rc) return rc;
    *n_admitted = 0;
    e->tm = triage_timings{};
    if (n == 0) return TRIAGE_OK;

    uint32_t *counters = reinterpret_cast<uint32_t *>(base + kStOffCounters);
    uint32_t *tile_first = reinterpret_cast<uint32_t *>(base + kStOffTileFirst);
    uint8_t *bucket = base + kStOffBucket;
    uint8_t *bucket2 = base + kStOffBucket2;
    uint32_t *blockhist = reinterpret_cast<uint32_t *>(base + kStOffBlockhist);
    uint32_t *totals = reinterpret_cast<uint32_t *>(base + kStOffTotals);
    uint32_t *small = reinterpret_cast<uint32_t *>(base + kStOffSmall);
    uint64_t *pending_dev = reinterpret_cast<uint64_t *>(base + kStOffPending);
    int64_t *maxsize_dev = reinterpret_cast<int64_t *>(base + kStOffMaxsize);
    uint32_t *perm_a = reinterpret_cast<uint32_t *>(base + kStOffPermA);
    uint32_t *perm_b = reinterpret_cast<uint32_t *>(base + kStOffPermB);

    cudaStream_t s = e->stream;
    uint32_t launches = 0;
    const uint32_t ntiles = (uint32_t)((total_bytes + e->geom.tile - 1) / e->geom.tile);
    const uint32_t nblocks = (n + kOrderBlock - 1) / kOrderBlock;
    OrderScratch sc{blockhist, totals, (uint32_t)kStMaxBlocks};
    (void)nblocks;

    CU(cudaEventRecord(e->ev[0], s));
    // the array keeps the previous call's contents: counters and flags must start at zero
    CU(cudaMemsetAsync(counters, 0, (size_t)n * CNT_STRIDE * sizeof(uint32_t), s));
    CU(cudaMemsetAsync(small, 0, (32 + 8 * 256) * sizeof(uint32_t), s));
    ScanArgs sa{bytes_dev, offsets_dev, total_bytes, n, counters, tile_first, ntiles, e->geom.tile};
    if (ntiles) {
        CU(launch_tile_index(sa, s)); launches++;
        CU(cudaEventRecord(e->ev[1], s));
        CU(launch_scan(sa, e->ddfa, e->geom, e->sm_count, s)); launches++;
    } else {
        CU(cudaEventRecord(e->ev[1], s));
    }
    if (e->has_bordered && ntiles) { CU(launch_fixup(bytes_dev, offsets_dev, nullptr, n, e->dxdfa, counters, s)); launches++; }
    if (e->rxview.n_programs && ntiles) { CU(launch_regex_count(e->rxview, bytes_dev, offsets_dev, nullptr, n, counters, e->sm_count, s)); launches++; }
    CU(cudaEventRecord(e->ev[2], s));
    FinalizeArgs fa{nullptr, nullptr, counters, meta_dev, offsets_dev, enqueue_ns_dev, n, e->default_prio, e->opts.rules_enabled, 0, results_dev,
                    bucket, small};
    CU(launch_finalize(fa, s)); launches++;
    CU(cudaEventRecord(e->ev[3], s));
    const uint8_t *bk = bucket;
    if (e->opts.max_queue_size > 0) {
        const uint64_t pending[4] = {0, 0, 0, 0};
        const int64_t ms[4] = {e->opts.max_queue_size, e->opts.max_queue_size, e->opts.max_queue_size, e->opts.max_queue_size};
        CU(cudaMemcpyAsync(pending_dev, pending, sizeof(pending), cudaMemcpyHostToDevice, s));
        CU(cudaMemcpyAsync(maxsize_dev, ms, sizeof(ms), cudaMemcpyHostToDevice, s));
        CU(cudaStreamSynchronize(s));   // the two host arrays above live on this stack frame
        CU(launch_admit(bucket, bucket2, results_dev, n, pending_dev, maxsize_dev, 0, sc, s)); launches += 3;
        bk = bucket2;
    }
    CU(launch_partition_u8(bk, nullptr, n, 16, true, order_dev, small + 2, sc, s)); launches += 4;
    CU(cudaEventRecord(e->ev[4], s));
    CU(cudaMemcpyAsync(e->h_small, small, 32 * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    CU(cudaMemcpyAsync(e->h_small + 1, totals + 256, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    uint32_t passes = 1;
    if (e->h_small[0] & 1u) {   // enqueue times not monotone in arrival order: LSD radix over the time key first
        CU(launch_key_digit_hist(enqueue_ns_dev, nullptr, n, small + 32, s)); launches++;
        CU(cudaMemcpyAsync(e->h_small + 32, small + 32, 8 * 256 * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
        CU(cudaStreamSynchronize(s));
        const uint32_t *perm = nullptr;
        uint32_t *bufs[2] = {perm_a, perm_b};
        int which = 0;
        for (uint32_t b = 0; b < 8; b++) {
            bool trivial = false;
            for (uint32_t d = 0; d < 256; d++) if (e->h_small[32 + b * 256 + d] == n) trivial = true;
            if (trivial) continue;
            CU(launch_partition_key(enqueue_ns_dev, 8 * b, perm, n, bufs[which], sc, s)); launches += 3;
            perm = bufs[which];
            which ^= 1;
            passes++;
        }
        CU(launch_partition_u8(bk, perm, n, 16, true, order_dev, small + 2, sc, s)); launches += 4;
        CU(cudaEventRecord(e->ev[4], s));
        CU(cudaMemcpyAsync(e->h_small, small, 32 * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
        CU(cudaMemcpyAsync(e->h_small + 1, totals + 256, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
        CU(cudaStreamSynchronize(s));
    }
    *n_admitted = e->h_small[1];
    // what triage_merge_allgather reads of "the last device submit" (the bucket bytes live in the shared array: merge
    // before the next static call on this GPU)
    for (int b = 0; b < 16; b++) e->last_counts[b] = e->h_small[2 + b];
    e->last_order = order_dev; e->last_bucket = bk; e->last_ts = enqueue_ns_dev; e->last_n = n; e->last_n_adm = *n_admitted;
    e->last_counts_dev = small + 2; e->last_n_adm_dev = totals + 256;
    float ms_ = 0;
    cudaEventElapsedTime(&ms_, e->ev[1], e->ev[2]); e->tm.scan_ms = ms_;
    cudaEventElapsedTime(&ms_, e->ev[2], e->ev[3]); e->tm.finalize_ms = ms_;
    cudaEventElapsedTime(&ms_, e->ev[3], e->ev[4]); e->tm.order_ms = ms_;
    cudaEventElapsedTime(&ms_, e->ev[0], e->ev[4]); e->tm.total_ms = ms_;
    e->tm.launches = launches;
    e->tm.order_passes = passes;
    e->tm.scan_bytes = total_bytes + 8ull * (n + 1) + 32ull * n;
    return TRIAGE_OK;
}
