// kernels.cu — hand-written sm_100a kernels of the triage path.
//
//   tile_index_kernel   offsets -> first message of every arena tile (scratch)
//   scan_kernel         the hot kernel: DFA byte scan of the packed arena, HBM-bound by design.
//                       Persistent CTAs, one per SM; arena tiles are brought into shared memory by
//                       1-D TMA bulk copies (cp.async.bulk + mbarrier, double buffered); the automaton
//                       table lives in shared memory; every lane walks its own contiguous chunk with
//                       conflict-free 128-bit shared loads; per-message word counts are combined across
//                       neighbouring lanes with shuffles before one atomic per (warp, message).
//   finalize_kernel     fused score -> label -> reason -> queue bucket (ProcessMessage decision tree)
//   admit/partition     stable multi-split = LSD radix passes that emit the dequeue order
//
// Reference semantics: internal/preprocessor/preprocessor.go:56-249, internal/priorityqueue/queue.go:22-27,
// 92-119, queue_manager.go:451-466, queue_factory.go:211-233, cmd/queue-manager/main.go:111-124.
#include "kernels.cuh"

#include <cuda.h>   // CUtensorMap + the cuTensorMapEncodeTiled prototype; the entry point is fetched at run time, libcuda is not linked
#include <cstdlib>
#include <type_traits>

namespace triage {

// ------------------------------------------------------------------------------ PTX helpers --
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// 1-D TMA bulk copy global -> shared, completion signalled on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void tma_load_1d(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// 2-D TMA tile copy global -> shared through a tensor map (SASS: UTMALDG); c0 = byte column, c1 = row
__device__ __forceinline__ void tma_load_2d(void *dst, const CUtensorMap *tm, int32_t c0, int32_t c1, uint64_t *bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                     smem_u32(dst)),
                 "l"(tm), "r"(c0), "r"(c1), "r"(smem_u32(bar))
                 : "memory");
}

// Raises a kernel's dynamic shared-memory limit once per (kernel, device): the attribute is per device, and one process
// may hold engines on several GPUs.
constexpr int kMaxDevices = 64;
struct SmemConfigured { size_t bytes[kMaxDevices] = {0}; };
template <class Kernel>
static cudaError_t ensure_dynamic_smem(Kernel kernel, size_t smem, SmemConfigured &done) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
    if (smem > done.bytes[dev]) {
        e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        done.bytes[dev] = smem;
    }
    return cudaSuccess;
}

// ------------------------------------------------------------------------------ tile index --
__global__ void tile_index_kernel(const uint64_t *__restrict__ offsets, uint32_t n, uint32_t ntiles, uint32_t tile_bytes,
                                  uint32_t *__restrict__ tile_first) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t > ntiles) return;
    if (t == ntiles) { tile_first[t] = n - 1; return; }
    uint64_t pos = (uint64_t)t * tile_bytes;
    uint32_t lo = 0, hi = n + 1;   // first index with offsets[idx] > pos
    while (lo < hi) {
        uint32_t mid = (lo + hi) >> 1;
        if (offsets[mid] > pos) hi = mid; else lo = mid + 1;
    }
    tile_first[t] = lo - 1;
}

cudaError_t launch_tile_index(const ScanArgs &a, cudaStream_t s) {
    uint32_t cnt = a.ntiles + 1;
    tile_index_kernel<<<(cnt + 255) / 256, 256, 0, s>>>(a.offsets, a.n, a.ntiles, a.tile_bytes, a.tile_first);
    return cudaGetLastError();
}

// One launch before the scan: the tile index, and the zeroing of the per-message counters and of the small scratch
// (two memsets and a kernel otherwise: ~45 us of a 1.2 ms step).  counters: n * CNT_STRIDE words, 16-byte aligned.
__global__ void __launch_bounds__(256) prepare_kernel(const uint64_t *__restrict__ offsets, uint32_t n, uint32_t ntiles, uint32_t tile_bytes,
                                                      uint32_t *__restrict__ tile_first, uint4 *__restrict__ counters4, uint64_t n_vec,
                                                      uint32_t *__restrict__ small, uint32_t small_words) {
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, nthreads = gridDim.x * blockDim.x;
    for (uint32_t t = tid; t <= ntiles; t += nthreads) {
        if (t == ntiles) { tile_first[t] = n - 1; continue; }
        uint64_t pos = (uint64_t)t * tile_bytes;
        uint32_t lo = 0, hi = n + 1;   // first index with offsets[idx] > pos
        while (lo < hi) {
            uint32_t mid = (lo + hi) >> 1;
            if (offsets[mid] > pos) hi = mid; else lo = mid + 1;
        }
        tile_first[t] = lo - 1;
    }
    const uint4 z = make_uint4(0, 0, 0, 0);
    for (uint64_t i = tid; i < n_vec; i += nthreads) counters4[i] = z;
    for (uint32_t i = tid; i < small_words; i += nthreads) small[i] = 0;
}

cudaError_t launch_prepare(const ScanArgs &a, uint32_t *small, uint32_t small_words, int sm_count, cudaStream_t s) {
    const uint64_t n_vec = (uint64_t)a.n * CNT_STRIDE / 4;   // CNT_STRIDE = 8 words per message
    prepare_kernel<<<sm_count * 8, 256, 0, s>>>(a.offsets, a.n, a.ntiles, a.tile_bytes, a.tile_first, reinterpret_cast<uint4 *>(a.counters),
                                                n_vec, small, small_words);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------------ scan --
// shared memory layout (bytes): [0,16) two mbarriers | table | events | eom(256) | stage0 | stage1 | list0 | list1
struct ScanSmem {
    uint32_t table, events, eom, stage0, list0, total, stage_bytes;
};
__host__ __device__ inline ScanSmem scan_layout(uint32_t n_states, uint32_t n_events, uint32_t warm, uint32_t tile) {
    ScanSmem L;
    uint32_t o = 128;
    L.table = o; o += ((n_states * kRowBytes + 127u) / 128u) * 128u;
    L.events = o; o += ((n_events * 16u + 127u) / 128u) * 128u;
    L.eom = o; o += 256u;
    L.stage_bytes = ((warm + tile + 16u + 127u) / 128u) * 128u;
    L.stage0 = o; o += 2 * L.stage_bytes;
    L.list0 = o; o += 2 * kListCap * 4u;
    L.total = o;
    return L;
}
size_t scan_smem_bytes(const DeviceDfa &d, const ScanGeometry &g) { return scan_layout(d.n_states, d.n_events, d.warm, g.tile).total; }

__device__ __forceinline__ int32_t rel_clamp(uint64_t off, uint64_t tile_start) {
    int64_t r = (int64_t)off - (int64_t)tile_start;
    const int64_t lim = 1 << 30;
    return (int32_t)(r < -lim ? -lim : (r > lim ? lim : r));
}

template <int NT, int C, bool FAST>
__global__ void __launch_bounds__(NT, 1) scan_kernel(ScanArgs a, DeviceDfa d) {
    extern __shared__ __align__(128) uint8_t smem[];
    constexpr uint32_t TILE = NT * C;
    constexpr int LPT = (kListCap + NT - 1) / NT;
    const ScanSmem L = scan_layout(d.n_states, d.n_events, d.warm, TILE);
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem);
    uint8_t *tbl = smem + L.table;
    EventPacked *events = reinterpret_cast<EventPacked *>(smem + L.events);
    uint8_t *eom = smem + L.eom;
    // buffers are addressed arithmetically (not through a pointer array) so the compiler keeps them in the
    // shared state space (LDS, not generic LD)
    auto stage_of = [&](uint32_t b) -> uint8_t * { return smem + L.stage0 + b * L.stage_bytes; };
    auto list_of = [&](uint32_t b) -> int32_t * { return reinterpret_cast<int32_t *>(smem + L.list0 + b * (kListCap * 4u)); };
    const uint32_t tid = threadIdx.x, lane = tid & 31;
    const uint32_t G = gridDim.x;
    const uint32_t warm = d.warm;

    if (blockIdx.x >= a.ntiles) return;
    const uint32_t nmine = (a.ntiles - blockIdx.x + G - 1) / G;

    if (tid == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // automaton -> shared memory (rows are 516 bytes: copy as 32-bit words)
    {
        const uint32_t *src = reinterpret_cast<const uint32_t *>(d.table);
        uint32_t *dst = reinterpret_cast<uint32_t *>(tbl);
        for (uint32_t i = tid; i < d.n_states * (kRowBytes / 4); i += NT) dst[i] = src[i];
        const uint4 *es = reinterpret_cast<const uint4 *>(d.events);
        uint4 *ed = reinterpret_cast<uint4 *>(events);
        for (uint32_t i = tid; i < d.n_events; i += NT) ed[i] = es[i];
        for (uint32_t i = tid; i < 256; i += NT) eom[i] = d.eom[i];
    }
    __syncthreads();

    auto issue_tile = [&](uint32_t tile, uint32_t buf) {   // one thread
        uint64_t ts = (uint64_t)tile * TILE;
        uint64_t te = ts + TILE < a.total ? ts + TILE : a.total;
        uint64_t cs = ts >= warm ? ts - warm : 0;
        uint32_t dst_off = (uint32_t)(cs + warm - ts);
        uint32_t bytes = (uint32_t)(((te - cs) + 15u) & ~(uint64_t)15u);
        mbar_expect_tx(&bars[buf], bytes);
        tma_load_1d(stage_of(buf) + dst_off, a.bytes + cs, bytes, &bars[buf]);
    };

    // prologue: tile 0 of this CTA
    uint32_t tile = blockIdx.x;
    uint32_t first_cur = a.tile_first[tile], last_cur = a.tile_first[tile + 1];
    if (tid == 0) issue_tile(tile, 0);
    {
        uint32_t cnt = last_cur - first_cur + 2;
        if (cnt <= kListCap)
            for (uint32_t i = tid; i < cnt; i += NT) list_of(0)[i] = rel_clamp(a.offsets[first_cur + i], (uint64_t)tile * TILE);
    }
    uint32_t first_nxt = 0, last_nxt = 0;
    if (nmine > 1) { first_nxt = a.tile_first[tile + G]; last_nxt = a.tile_first[tile + G + 1]; }
    __syncthreads();

    for (uint32_t k = 0; k < nmine; k++, tile += G) {
        const uint32_t buf = k & 1;
        const bool has_next = k + 1 < nmine;
        if (has_next && tid == 0) issue_tile(tile + G, buf ^ 1);
        // start the loads that build the NEXT tile's message list; they are consumed after the scan
        uint64_t nxt_off[LPT];
        uint32_t cnt_nxt = 0;
        uint32_t first_nn = 0, last_nn = 0;
        if (has_next) {
            cnt_nxt = last_nxt - first_nxt + 2;
            if (cnt_nxt <= kListCap) {
#pragma unroll
                for (int j = 0; j < LPT; j++) {
                    uint32_t i = tid + j * NT;
                    nxt_off[j] = i < cnt_nxt ? a.offsets[first_nxt + i] : 0;
                }
            }
            if (k + 2 < nmine) { first_nn = a.tile_first[tile + 2 * G]; last_nn = a.tile_first[tile + 2 * G + 1]; }
        }

        const uint64_t ts = (uint64_t)tile * TILE;
        const int32_t tile_len = (int32_t)((ts + TILE < a.total ? ts + TILE : a.total) - ts);
        const uint32_t cnt_cur = last_cur - first_cur + 2;

        LaneCtx c;
        c.tbl = tbl; c.events = events; c.eom = eom; c.stage = stage_of(buf);
        c.list = cnt_cur <= kListCap ? list_of(buf) : nullptr;
        c.list_abs = a.offsets; c.tile_start = (int64_t)ts;
        c.list_n = cnt_cur; c.msg_lo = first_cur; c.counters = a.counters; c.warm = (int32_t)warm;
        c.start_state = d.start_state; c.neutral_state = d.neutral_state;
        c.head_pad4 = 0x01010101u * (uint32_t)(d.head_pad & 0xFF); c.tail_pad4 = 0x01010101u * (uint32_t)(d.tail_pad & 0xFF);
        c.tbl_s = smem_u32(tbl);

        mbar_wait(&bars[buf], (k >> 1) & 1);

        uint32_t hm = 0xFFFFFFFFu, hw = 0;
        const int32_t chunk = (int32_t)(tid * C);
        // the warp's first lane finds its message by binary search, the others walk forward from there
        const int32_t warp_chunk = (int32_t)((tid & ~31u) * C);
        uint32_t khint = 0;
        if (warp_chunk < tile_len) {
            if (lane == 0) khint = find_msg(c, warp_chunk);
            khint = __shfl_sync(0xFFFFFFFFu, khint, 0);
        }
        if (chunk < tile_len) {
            int32_t ce = chunk + C < tile_len ? chunk + C : tile_len;
            uint32_t kk = find_msg_from(c, chunk, khint);
            if (FAST) lane_scan_fast(c, chunk, ce, kk, &hm, &hw);
            else lane_scan_slow(c, chunk, ce, kk, &hm, &hw);
        }
        // combine the head partials of neighbouring lanes that sit in the same message
        {
            const unsigned full = 0xFFFFFFFFu;
            uint32_t prev = __shfl_up_sync(full, hm, 1);
            bool head = lane == 0 || hm != prev;
            uint32_t heads = __ballot_sync(full, head);
            uint32_t v = hw;
#pragma unroll
            for (int dlt = 1; dlt < 32; dlt <<= 1) {
                uint32_t t = __shfl_up_sync(full, v, dlt);
                if ((int)lane >= dlt) v += t;
            }
            uint32_t above = heads & ~((2u << lane) - 1u);             // heads strictly above this lane
            uint32_t end = above ? (uint32_t)__ffs(above) - 2u : 31u;   // last lane of this lane's group
            uint32_t upto = __shfl_sync(full, v, end);
            uint32_t before = __shfl_up_sync(full, v, 1);
            if (lane == 0) before = 0;
            if (head && hm != 0xFFFFFFFFu) {
                uint32_t sum = upto - before;
                if (sum) atomicAdd(a.counters + (uint64_t)hm * CNT_STRIDE + CNT_WORDS, sum);
            }
        }
        // publish the next tile's list
        if (has_next && cnt_nxt <= kListCap) {
#pragma unroll
            for (int j = 0; j < LPT; j++) {
                uint32_t i = tid + j * NT;
                if (i < cnt_nxt) list_of(buf ^ 1)[i] = rel_clamp(nxt_off[j], ts + (uint64_t)G * TILE);
            }
        }
        first_cur = first_nxt; last_cur = last_nxt;
        first_nxt = first_nn; last_nxt = last_nn;
        __syncthreads();
    }
}

// geometries in order of preference (more bytes per lane first: less warm-up overhead)
static const ScanGeometry kGeometries[] = {{768, 80, 768 * 80}, {1024, 48, 1024 * 48}, {512, 48, 512 * 48}};

ScanGeometry pick_scan_geometry(const DeviceDfa &d, size_t max_smem) {
    for (const ScanGeometry &g : kGeometries)
        if (scan_smem_bytes(d, g) <= max_smem) return g;
    return ScanGeometry{0, 0, 0};
}

template <int NT, int C, bool FAST>
static cudaError_t launch_scan_variant(const ScanArgs &a, const DeviceDfa &d, size_t smem, uint32_t grid, cudaStream_t s) {
    auto kern = scan_kernel<NT, C, FAST>;
    static SmemConfigured configured;
    if (cudaError_t e = ensure_dynamic_smem(kern, smem, configured)) return e;
    kern<<<grid, NT, smem, s>>>(a, d);
    return cudaGetLastError();
}

cudaError_t launch_scan(const ScanArgs &a, const DeviceDfa &d, const ScanGeometry &g, int sm_count, cudaStream_t s) {
    static_assert(768 * 80 >= kMinScanTile && 1024 * 48 >= kMinScanTile && 512 * 48 >= kMinScanTile, "tile too small");
    if (a.ntiles == 0) return cudaSuccess;
    size_t smem = scan_smem_bytes(d, g);
    uint32_t grid = a.ntiles < (uint32_t)sm_count ? a.ntiles : (uint32_t)sm_count;
    const bool fast = d.head_pad >= 0 && d.tail_pad >= 0;
#define TRIAGE_SCAN_CASE(NT_, C_)                                                                        \
    if (g.nt == NT_ && g.chunk == C_)                                                                    \
        return fast ? launch_scan_variant<NT_, C_, true>(a, d, smem, grid, s) : launch_scan_variant<NT_, C_, false>(a, d, smem, grid, s);
    TRIAGE_SCAN_CASE(768, 80)
    TRIAGE_SCAN_CASE(1024, 48)
    TRIAGE_SCAN_CASE(512, 48)
#undef TRIAGE_SCAN_CASE
    return cudaErrorInvalidValue;
}

// ----------------------------------------------------------------------------- padded scan --
// Warp-autonomous variant for the padded arena: every warp owns a double-buffered staging area, its own pair of
// mbarriers and a small message list, takes 32*C-byte mini-tiles round-robin and never meets a CTA-wide barrier
// after the automaton has been loaded.
struct PadSmem { uint32_t table, events, stage0, list0, total, stage_bytes, list_bytes; };
constexpr uint32_t kPadListCap = 34;
__host__ __device__ inline PadSmem pad_layout(uint32_t n_states, uint32_t n_events, uint32_t warm, uint32_t nw, bool wide) {
    PadSmem L;
    uint32_t o = 1024;   // mbarriers
    // dense: the whole table and the event list; wide: only the 256-byte class map (table + events stay in global memory)
    L.table = o; o += wide ? 256u : ((n_states * kRowBytes + 127u) / 128u) * 128u;
    L.events = o; o += wide ? 0u : ((n_events * 16u + 127u) / 128u) * 128u;
    L.stage_bytes = ((warm + kPadMiniTile + 16u + 127u) / 128u) * 128u;
    L.stage0 = o; o += nw * 2 * L.stage_bytes;
    L.list_bytes = 160;
    L.list0 = o; o += nw * 2 * L.list_bytes;
    L.total = o;
    return L;
}
size_t scan_padded_smem_bytes(const DeviceDfa &d, uint32_t nw) { return pad_layout(d.n_states, d.n_events, d.warm, nw, d.wide != 0).total; }
// most warps per CTA whose staging buffers still fit next to the automaton (0: none)
uint32_t pick_pad_warps(const DeviceDfa &d, size_t max_smem) {
    if (d.wide && !getenv("TRIAGE_WIDE_24") && scan_padded_smem_bytes(d, 32) <= max_smem) return 32;
    for (uint32_t nw : {24u, 20u, 16u, 12u, 8u})
        if (scan_padded_smem_bytes(d, nw) <= max_smem) return nw;
    return 0;
}

template <int NW, bool WIDE>
__global__ void __launch_bounds__(NW * 32, 1) scan_padded_kernel(ScanArgs a, DeviceDfa d) {
    constexpr uint32_t kPadWarps = NW;
    extern __shared__ __align__(128) uint8_t smem[];
    constexpr int C = kPadChunk;
    constexpr uint32_t MT = kPadMiniTile;
    const PadSmem L = pad_layout(d.n_states, d.n_events, d.warm, kPadWarps, WIDE);
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem) + warp * 2;
    uint8_t *tbl = smem + L.table;
    EventPacked *events = reinterpret_cast<EventPacked *>(smem + L.events);
    uint8_t *stage_base = smem + L.stage0 + warp * 2 * L.stage_bytes;
    int32_t *list_base = reinterpret_cast<int32_t *>(smem + L.list0 + warp * 2 * L.list_bytes);
    const uint32_t warm = d.warm;

    if (lane == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if constexpr (WIDE) {
        for (uint32_t i = tid; i < 256; i += kPadWarps * 32) tbl[i] = d.cls_of[i];
    } else {
        const uint32_t *src = reinterpret_cast<const uint32_t *>(d.table);
        uint32_t *dst = reinterpret_cast<uint32_t *>(tbl);
        for (uint32_t i = tid; i < d.n_states * (kRowBytes / 4); i += kPadWarps * 32) dst[i] = src[i];
        const uint4 *es = reinterpret_cast<const uint4 *>(d.events);
        uint4 *ed = reinterpret_cast<uint4 *>(events);
        for (uint32_t i = tid; i < d.n_events; i += kPadWarps * 32) ed[i] = es[i];
    }
    __syncthreads();   // the only CTA-wide barrier

    const uint32_t stride = gridDim.x * kPadWarps;
    uint32_t t = blockIdx.x * kPadWarps + warp;
    if (t >= a.ntiles) return;
    const uint32_t nmine = (a.ntiles - t + stride - 1) / stride;
    const uint32_t last_tile = a.ntiles - 1;
    const uint32_t last_bytes = (uint32_t)(a.total - (uint64_t)last_tile * MT);   // multiple of 16 (padded arena)

    // one lane: bulk copy of mini-tile mt (plus the warm-up bytes before it) into buffer buf
    auto issue = [&](uint32_t mt, uint32_t buf) {
        const uint8_t *src = a.bytes + (uint64_t)mt * MT;
        uint32_t bytes = mt == last_tile ? last_bytes : MT;
        uint8_t *dst = stage_base + buf * L.stage_bytes + warm;
        if (mt != 0) { src -= warm; dst -= warm; bytes += warm; }
        mbar_expect_tx(&bars[buf], bytes);
        tma_load_1d(dst, src, bytes, &bars[buf]);
    };

    uint32_t first_cur = a.tile_first[t], last_cur = a.tile_first[t + 1];
    if (lane == 0) issue(t, 0);
    uint64_t ts = (uint64_t)t * MT;
    {
        uint32_t cnt = last_cur - first_cur + 2;
        if (cnt <= kPadListCap) {
            if (lane < cnt) list_base[lane] = rel_clamp(a.offsets[first_cur + lane], ts);
            if (lane + 32 < cnt) list_base[lane + 32] = rel_clamp(a.offsets[first_cur + lane + 32], ts);
        }
    }
    uint32_t first_nxt = 0, last_nxt = 0;
    if (nmine > 1) { first_nxt = a.tile_first[t + stride]; last_nxt = a.tile_first[t + stride + 1]; }
    __syncwarp();

    using Ctx = typename std::conditional<WIDE, WideCtx, LaneCtx>::type;
    Ctx c;
    if constexpr (WIDE) {
        c.wtbl = d.wtable; c.cls = tbl; c.ncls = d.n_classes; c.events = d.events;
    } else {
        c.tbl = tbl; c.events = events; c.eom = nullptr;
        c.head_pad4 = c.tail_pad4 = 0;
        c.tbl_s = smem_u32(tbl);
    }
    c.list_abs = a.offsets; c.counters = a.counters; c.warm = (int32_t)warm;
    c.start_state = d.start_state; c.neutral_state = d.neutral_state;
    const int32_t chunk = (int32_t)(lane * C);
    const uint64_t ts_step = (uint64_t)stride * MT;

    for (uint32_t k = 0; k < nmine; k++, t += stride, ts += ts_step) {
        const uint32_t buf = k & 1;
        const bool has_next = k + 1 < nmine;
        if (has_next && lane == 0) issue(t + stride, buf ^ 1);
        uint64_t n0 = 0, n1 = 0;
        uint32_t cnt_nxt = 0, first_nn = 0, last_nn = 0;
        if (has_next) {
            cnt_nxt = last_nxt - first_nxt + 2;
            if (cnt_nxt <= kPadListCap) {
                if (lane < cnt_nxt) n0 = a.offsets[first_nxt + lane];
                if (lane + 32 < cnt_nxt) n1 = a.offsets[first_nxt + lane + 32];
            }
            if (k + 2 < nmine) { first_nn = a.tile_first[t + 2 * stride]; last_nn = a.tile_first[t + 2 * stride + 1]; }
        }
        const int32_t len = t == last_tile ? (int32_t)last_bytes : (int32_t)MT;
        const uint32_t cnt_cur = last_cur - first_cur + 2;
        const int32_t *lst = list_base + buf * (L.list_bytes / 4);
        c.stage = stage_base + buf * L.stage_bytes;
        c.list = cnt_cur <= kPadListCap ? lst : nullptr;
        c.tile_start = (int64_t)ts;
        c.list_n = cnt_cur; c.msg_lo = first_cur;

        mbar_wait(&bars[buf], (k >> 1) & 1);

        uint32_t hm = 0xFFFFFFFFu, hw = 0;
        if (chunk < len) {
            int32_t ce = chunk + C < len ? chunk + C : len;
            uint32_t km;
            if (c.list) {   // few messages per mini-tile: count the starts at or before this lane's chunk
                km = 0;
                for (uint32_t j = 1; j + 1 < cnt_cur; j++) km += lst[j] <= chunk ? 1u : 0u;
            } else {
                km = find_msg(c, chunk);
            }
            lane_scan_padded(c, chunk, ce, km, t == 0 && lane == 0, &hm, &hw);
        }
        {   // combine the head partials of neighbouring lanes that sit in the same message
            const unsigned full = 0xFFFFFFFFu;
            uint32_t prev = __shfl_up_sync(full, hm, 1);
            bool head = lane == 0 || hm != prev;
            uint32_t heads = __ballot_sync(full, head);
            uint32_t v = hw;
#pragma unroll
            for (int dlt = 1; dlt < 32; dlt <<= 1) {
                uint32_t tt = __shfl_up_sync(full, v, dlt);
                if ((int)lane >= dlt) v += tt;
            }
            uint32_t above = heads & ~((2u << lane) - 1u);
            uint32_t end = above ? (uint32_t)__ffs(above) - 2u : 31u;
            uint32_t upto = __shfl_sync(full, v, end);
            uint32_t before = __shfl_up_sync(full, v, 1);
            if (lane == 0) before = 0;
            if (head && hm != 0xFFFFFFFFu) {
                uint32_t sum = upto - before;
                if (sum) atomicAdd(a.counters + (uint64_t)hm * CNT_STRIDE + CNT_WORDS, sum);
            }
        }
        if (has_next && cnt_nxt <= kPadListCap) {
            int32_t *ln = list_base + (buf ^ 1) * (L.list_bytes / 4);
            const uint64_t tsn = ts + ts_step;
            if (lane < cnt_nxt) ln[lane] = rel_clamp(n0, tsn);
            if (lane + 32 < cnt_nxt) ln[lane + 32] = rel_clamp(n1, tsn);
        }
        first_cur = first_nxt; last_cur = last_nxt;
        first_nxt = first_nn; last_nxt = last_nn;
        __syncwarp();
    }
}

template <int NW, bool WIDE = false>
static cudaError_t launch_scan_padded_variant(const ScanArgs &a, const DeviceDfa &d, int sm_count, cudaStream_t s) {
    size_t smem = scan_padded_smem_bytes(d, NW);
    static SmemConfigured configured;
    if (cudaError_t e = ensure_dynamic_smem(scan_padded_kernel<NW, WIDE>, smem, configured)) return e;
    uint32_t need = (a.ntiles + NW - 1) / NW;
    uint32_t grid = need < (uint32_t)sm_count ? need : (uint32_t)sm_count;
    scan_padded_kernel<NW, WIDE><<<grid, NW * 32, smem, s>>>(a, d);
    return cudaGetLastError();
}

cudaError_t launch_scan_padded(const ScanArgs &a, const DeviceDfa &d, uint32_t nw, int sm_count, cudaStream_t s) {
    if (a.ntiles == 0) return cudaSuccess;
    if (d.wide) {   // latency-bound on L1/L2 gathers: as many warps as the staging allows
        if (nw >= 32) return launch_scan_padded_variant<32, true>(a, d, sm_count, s);
        return launch_scan_padded_variant<24, true>(a, d, sm_count, s);
    }
    switch (nw) {
        case 24: return launch_scan_padded_variant<24>(a, d, sm_count, s);
        case 20: return launch_scan_padded_variant<20>(a, d, sm_count, s);
        case 16: return launch_scan_padded_variant<16>(a, d, sm_count, s);
        case 12: return launch_scan_padded_variant<12>(a, d, sm_count, s);
        case 8: return launch_scan_padded_variant<8>(a, d, sm_count, s);
        default: return cudaErrorInvalidValue;
    }
}

// ---------------------------------------------------------------- wide format, streamed rows --
// The wide format's walk is one dependent gather per input byte into a table that lives in L2/L1, so what it needs is
// gathers in flight and L1 capacity, not staging: here every lane walks N rows of `cl` bytes at once (chains
// interleaved transition by transition, scan_core.cuh long_stream_n) and reads its bytes straight from global memory,
// 16 bytes a row at a time.  Shared memory holds the 256-byte class map and one message list per warp, the rest of
// the SM's 256 KB stays L1 for the table.  Rows cost one 16-byte warm-up per `cl` bytes instead of one per 80.
// Tiles of 32 * N rows are handed out dynamically.
constexpr uint32_t kStreamListCap = 98;   // message starts of a tile kept in shared memory (lane, lane + 32, lane + 64)
template <int NW, int N>
__global__ void __launch_bounds__(NW * 32, 1) scan_stream_wide_kernel(ScanArgs a, DeviceDfa d, uint32_t cl, uint32_t *__restrict__ tile_counter) {
    __shared__ uint8_t cls[256];
    __shared__ int32_t lists[NW][kStreamListCap + 2];
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (uint32_t i = tid; i < 256; i += NW * 32) cls[i] = d.cls_of[i];
    __syncthreads();   // the only CTA-wide barrier

    const uint32_t nwarps = gridDim.x * NW, t_end = a.tile_end;
    uint32_t t = a.tile_begin + blockIdx.x * NW + warp;
    WideCtx c;
    c.wtbl = d.wtable; c.cls = cls; c.ncls = d.n_classes; c.events = d.events; c.stage = nullptr;
    c.list_abs = a.offsets; c.counters = a.counters; c.warm = (int32_t)d.warm;
    c.start_state = d.start_state; c.neutral_state = d.neutral_state;
    const uint32_t super = 32u * N * cl;
    int32_t *lst = lists[warp];
    const uint8_t *limit = a.bytes + a.total;

    while (t < t_end) {
        const uint64_t ts = (uint64_t)t * super;
        const uint32_t first = a.tile_first[t], last = a.tile_first[t + 1];
        const uint32_t cnt = last - first + 2;
        if (cnt <= kStreamListCap) {
            for (uint32_t j = lane; j < cnt; j += 32) lst[j] = rel_clamp(a.offsets[first + j], ts);
        }
        __syncwarp();
        c.list = cnt <= kStreamListCap ? lst : nullptr;
        c.tile_start = (int64_t)ts;
        c.list_n = cnt; c.msg_lo = first;

        LongLane L[N];
        const uint8_t *src[N];
        int32_t pos[N], nb[N];
        bool all_full = true;
#pragma unroll
        for (int ch = 0; ch < N; ch++) {
            const uint64_t cs = ts + (uint64_t)(ch * 32 + lane) * cl;
            nb[ch] = cs >= a.total ? 0 : (int32_t)(a.total - cs < cl ? a.total - cs : cl);
            all_full = all_full && nb[ch] == (int32_t)cl;
            src[ch] = a.bytes + cs;
            pos[ch] = (int32_t)(cs - ts);
            if (nb[ch]) long_begin(c, L[ch], find_msg(c, pos[ch]), cs == 0, cs ? a.bytes + cs - c.warm : a.bytes);
        }
        if (all_full) {
            long_stream_n<N>(c, L, src, pos, (int32_t)cl, limit);
        } else {   // the arena's last tile: rows cut short (or absent) are walked one at a time
#pragma unroll
            for (int ch = 0; ch < N; ch++) {
                if (!nb[ch]) continue;
                LongLane L1[1] = {L[ch]};
                const uint8_t *const s1[1] = {src[ch]};
                const int32_t p1[1] = {pos[ch]};
                long_stream_n<1>(c, L1, s1, p1, nb[ch], limit);
                L[ch] = L1[0];
            }
        }
#pragma unroll
        for (int ch = 0; ch < N; ch++) {
            if (!nb[ch]) continue;
            long_end(c, L[ch]);
            if (L[ch].head_words) atomicAdd(a.counters + (uint64_t)L[ch].head_msg * CNT_STRIDE + CNT_WORDS, L[ch].head_words);
        }
        __syncwarp();   // the list is rewritten for the next tile
        uint32_t v = 0;
        if (lane == 0) v = atomicAdd(tile_counter, 1u) + nwarps + a.tile_begin;
        t = __shfl_sync(0xFFFFFFFFu, v, 0);
    }
}

template <int NW, int N>
static cudaError_t launch_scan_stream_wide_variant(const ScanArgs &a, const DeviceDfa &d, uint32_t cl, uint32_t *tile_counter, int sm_count,
                                                   cudaStream_t s) {
    uint32_t need = (a.tile_end - a.tile_begin + NW - 1) / NW;
    uint32_t grid = need < (uint32_t)sm_count ? need : (uint32_t)sm_count;
    static bool carveout_set = false;
    if (!carveout_set) {   // as much L1 as the SM can give: the kernel's shared memory is a few KB
        cudaFuncSetAttribute(scan_stream_wide_kernel<NW, N>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxL1);
        carveout_set = true;
    }
    scan_stream_wide_kernel<NW, N><<<grid, NW * 32, 0, s>>>(a, d, cl, tile_counter);
    return cudaGetLastError();
}

// a.tile_bytes must be 32 * chains * row_bytes (scan_stream_wide_super_bytes)
cudaError_t launch_scan_stream_wide(const ScanArgs &a, const DeviceDfa &d, uint32_t nw, uint32_t chains, uint32_t row_bytes, uint32_t *tile_counter,
                                    int sm_count, cudaStream_t s) {
    if (a.tile_end <= a.tile_begin) return cudaSuccess;
    if (!d.wide || row_bytes == 0 || (row_bytes & 15u) || a.tile_bytes != 32u * chains * row_bytes) return cudaErrorInvalidValue;
    if (chains == 1 && nw == 32) return launch_scan_stream_wide_variant<32, 1>(a, d, row_bytes, tile_counter, sm_count, s);
    if (chains == 2 && nw == 32) return launch_scan_stream_wide_variant<32, 2>(a, d, row_bytes, tile_counter, sm_count, s);
    if (chains == 2 && nw == 24) return launch_scan_stream_wide_variant<24, 2>(a, d, row_bytes, tile_counter, sm_count, s);
    if (chains == 3 && nw == 20) return launch_scan_stream_wide_variant<20, 3>(a, d, row_bytes, tile_counter, sm_count, s);
    if (chains == 4 && nw == 16) return launch_scan_stream_wide_variant<16, 4>(a, d, row_bytes, tile_counter, sm_count, s);
    if (chains == 4 && nw == 12) return launch_scan_stream_wide_variant<12, 4>(a, d, row_bytes, tile_counter, sm_count, s);
    return cudaErrorInvalidValue;
}

// ---------------------------------------------------------------------------------- row lanes --
// Padded arena, dense table, long lanes fed by the TMA: the arena is viewed as a matrix of LANE-byte rows, lane l
// of a warp owns row 32 * tile + l, and ONE tensor copy per warp and window brings the next 80-byte window of all 32
// rows into shared memory (box 80 x 32), a second small one (box 16 x 32 at column LANE - 16 of the rows before) the
// warm-up bytes of a new tile.  The load path issues no load/store-unit instructions at all - on this part the scan is bound by
// the rate at which the SM accepts shared-memory loads (profiles/microbench/lds_chain.cu), so the staging copies a lane
// would otherwise issue itself (cp.async) cost as much as the warm-up lookups they save.  The automaton state, word
// count and current message stay in registers along the row.  The last, partial row of the arena is read straight from
// global memory by the lane that owns it.
// Window: 80 bytes (28 warps x 64-byte windows and 32 x 48 were measured slower than 24 x 80: the 16-byte reads of a
// 64-byte-pitch buffer collide on banks, and 64 registers per thread cost more than 8 more warps bring).  Row: 1280, 640
// or 320 bytes by arena size (longer rows: less warm-up, tile bookkeeping once per 40 / 20 / 10 KB).
constexpr uint32_t kRowWin = 80, kRowLaneLong = 1280, kRowLaneMid = 640, kRowLaneShort = 320;
__host__ __device__ constexpr uint32_t row_win(uint32_t) { return kRowWin; }
// Longest row that still leaves every warp several tiles: tiles are handed out dynamically, so the kernel's tail is
// half a tile per warp on average (1280-byte rows: 40 KB tiles, 0.955 ms on the 2.3 GB bench arena; 640: 0.98; 320: 1.03).
// pair: the two-bytes-per-lookup lanes, whose tiles vary more in duration (flagged groups): 640-byte rows at most
// (2.29 GB: 1280 -> 0.923 ms, 640 -> 0.880).
uint32_t scan_rows_lane_bytes(uint64_t total, uint32_t nw, int sm_count, bool pair) {
    if (const char *v = getenv("TRIAGE_ROW_BYTES")) {   // test / tuning knob: force a row length
        const uint32_t want = (uint32_t)atoi(v);
        if ((want == kRowLaneLong || want == kRowLaneMid || want == kRowLaneShort) && total >= want) return want;
    }
    const uint64_t warp_rows = (uint64_t)nw * (uint32_t)sm_count * 32;   // rows one round of tiles covers
    // measured on one box: 2.29 GB (16 tiles of 40 KB per warp) 1280 -> 0.955 ms, 640 -> 0.98, 320 -> 1.03; 1.69 GB mixed
    // lengths 0.783 / 0.778 / 0.810; 1.34 GB of 32-128 KB payloads 0.680 / 0.608 / 0.601; 575 MB 2.39 / 2.43 / 2.31 TB/s
    if (!pair && total >= 12 * warp_rows * kRowLaneLong) return kRowLaneLong;
    if (total >= 3 * warp_rows * kRowLaneMid) return kRowLaneMid;
    return kRowLaneShort;
}
constexpr uint32_t kRowListCap = 34;     // message starts of a tile kept in shared memory (lane, lane + 32)
constexpr uint32_t kRowQueueCap = 24;    // deferred event replays per warp (12 bytes each)
constexpr uint32_t kRowQueueFlush = 10;  // replayed by the whole warp once this many are waiting
struct RowSmem { uint32_t table, events, warp0, aux0, total, warp_bytes, aux_bytes, warm_off, list_off, queue_off; };

// Per-warp queue of "8 bytes whose walk met an event": (message | entry state << 24, bytes 0..3, bytes 4..7).  A lane
// that meets an event in the fast pass only records it; the warp replays the recorded halves together (one per lane)
// at the end of a window.  Full queue or message index >= 2^24: the lane replays on the spot.
struct RowEventQueue {
    uint32_t base_s;   // shared address: count word, then the slots
    __device__ __forceinline__ void push(const LaneCtx &c, uint32_t e0, uint32_t wa, uint32_t wb, uint32_t msg) {
        uint32_t slot = kRowQueueCap;
        if ((msg >> 24) == 0) asm volatile("atom.shared.add.u32 %0, [%1], 1;" : "=r"(slot) : "r"(base_s) : "memory");
        if (slot < kRowQueueCap) {
            const uint32_t at = base_s + 4 + slot * 12;
            asm volatile("st.shared.u32 [%0], %1;" ::"r"(at), "r"(msg | (e0 << 24)) : "memory");
            asm volatile("st.shared.u32 [%0], %1;" ::"r"(at + 4), "r"(wa) : "memory");
            asm volatile("st.shared.u32 [%0], %1;" ::"r"(at + 8), "r"(wb) : "memory");
        } else {
            replay_half(c, wa, wb, e0 & 0xFFu, msg);
        }
    }
    // whole warp, converged; force: also when only a few are waiting
    __device__ __forceinline__ void flush(const LaneCtx &c, uint32_t lane, bool force) {
        uint32_t n;
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(n) : "r"(base_s) : "memory");
        if (n == 0 || (!force && n < kRowQueueFlush)) return;
        if (n > kRowQueueCap) n = kRowQueueCap;
        if (lane < n) {
            uint32_t m, wa, wb;
            const uint32_t at = base_s + 4 + lane * 12;
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(m) : "r"(at) : "memory");
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(wa) : "r"(at + 4) : "memory");
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(wb) : "r"(at + 8) : "memory");
            uint32_t e = m >> 24, wdummy = 0;
            half8_events(c, wa, wb, e, wdummy, m & 0xFFFFFFu);
        }
        __syncwarp();
        if (lane == 0) asm volatile("st.shared.u32 [%0], %1;" ::"r"(base_s), "r"(0u) : "memory");
        __syncwarp();
    }
};
__host__ __device__ inline RowSmem row_layout(uint32_t n_states, uint32_t n_events, uint32_t nw) {
    RowSmem L;
    uint32_t o = 1024;   // mbarriers
    L.table = o; o += ((n_states * kRowBytes + 127u) / 128u) * 128u;
    L.events = o; o += ((n_events * 16u + 127u) / 128u) * 128u;
    // TMA destinations (two windows + the warm-up box, 128-byte aligned) of all warps first, the small per-warp areas after
    L.warm_off = 2 * 32 * row_win(nw);
    L.warp_bytes = L.warm_off + 512;
    L.warp0 = o; o += nw * L.warp_bytes;
    L.list_off = 0;
    L.queue_off = 2 * 4 * kRowListCap;
    L.aux_bytes = ((L.queue_off + 4 + 12 * kRowQueueCap + 15u) / 16u) * 16u;
    L.aux0 = o; o += nw * L.aux_bytes;
    L.total = o;
    return L;
}
size_t scan_rows_smem_bytes(const DeviceDfa &d, uint32_t nw) { return row_layout(d.n_states, d.n_events, nw).total; }
uint32_t scan_rows_super_bytes(uint32_t lane_bytes) { return 32 * lane_bytes; }
uint32_t pick_row_warps(const DeviceDfa &d, size_t max_smem) {
    if (d.wide || d.warm != 16 || d.reset_pad < 0) return 0;
    if (const char *w = getenv("TRIAGE_ROW_WARPS")) {   // tuning knob
        uint32_t nw = (uint32_t)atoi(w);
        if ((nw == 16 || nw == 20 || nw == 24) && scan_rows_smem_bytes(d, nw) <= max_smem) return nw;
    }
    for (uint32_t nw : {24u, 20u, 16u})
        if (scan_rows_smem_bytes(d, nw) <= max_smem) return nw;
    return 0;
}

template <int NW, int LANE>
__global__ void __launch_bounds__(NW * 32, 1) scan_rows_kernel(ScanArgs a, DeviceDfa d, const __grid_constant__ CUtensorMap tm_win,
                                                              const __grid_constant__ CUtensorMap tm_warm, uint32_t nrows_full,
                                                              uint32_t *__restrict__ tile_counter) {
    extern __shared__ __align__(128) uint8_t smem[];
    constexpr int32_t CL = LANE, WIN = (int32_t)row_win(NW), NWIN = CL / WIN;
    constexpr uint32_t SUPER = 32 * CL;
    const RowSmem L = row_layout(d.n_states, d.n_events, NW);
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem) + warp * 2;
    uint8_t *tbl = smem + L.table;
    EventPacked *events = reinterpret_cast<EventPacked *>(smem + L.events);
    uint8_t *wbase = smem + L.warp0 + warp * L.warp_bytes;
    uint8_t *warm_buf = wbase + L.warm_off;
    uint8_t *aux = smem + L.aux0 + warp * L.aux_bytes;
    int32_t *list_base = reinterpret_cast<int32_t *>(aux + L.list_off);

    RowEventQueue queue;
    queue.base_s = smem_u32(aux + L.queue_off);
    if (lane == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("st.shared.u32 [%0], %1;" ::"r"(queue.base_s), "r"(0u) : "memory");
    }
    {
        const uint32_t *src = reinterpret_cast<const uint32_t *>(d.table);
        uint32_t *dst = reinterpret_cast<uint32_t *>(tbl);
        for (uint32_t i = tid; i < d.n_states * (kRowBytes / 4); i += NW * 32) dst[i] = src[i];
        const uint4 *es = reinterpret_cast<const uint4 *>(d.events);
        uint4 *ed = reinterpret_cast<uint4 *>(events);
        for (uint32_t i = tid; i < d.n_events; i += NW * 32) ed[i] = es[i];
    }
    __syncthreads();   // the only CTA-wide barrier

    // Tiles are handed out dynamically: the first one per warp by position, every further one from a global counter
    // (zeroed by prepare_kernel), fetched one tile ahead so that the next tile's message list and first window can be
    // prefetched while the current one is walked.
    // One launch walks the tiles [a.tile_begin, t_end): the whole arena, or the part of it that has already arrived when
    // the upload is pipelined chunk by chunk (engine.cu, submit from host buffers).
    const uint32_t nwarps = gridDim.x * NW;
    const uint32_t t_end = a.tile_end;
    const uint32_t t0 = a.tile_begin + blockIdx.x * NW + warp;
    if (t0 >= t_end) return;
    auto grab = [&]() {
        uint32_t v = 0;
        if (lane == 0) v = atomicAdd(tile_counter, 1u) + nwarps + a.tile_begin;
        return __shfl_sync(0xFFFFFFFFu, v, 0);
    };

    // window w of tile `tile` -> buffer buf (one lane issues)
    auto issue = [&](uint32_t tile, uint32_t w, uint32_t buf) {
        const int32_t row0 = (int32_t)(tile * 32u);
        mbar_expect_tx(&bars[buf], 32 * WIN + (w == 0 ? 512u : 0u));
        tma_load_2d(wbase + buf * (32 * WIN), &tm_win, (int32_t)(w * WIN), row0, &bars[buf]);
        if (w == 0) tma_load_2d(warm_buf, &tm_warm, CL - 16, row0 - 1, &bars[buf]);
    };

    uint32_t t = t0, t_nxt = grab();
    uint32_t first_cur = a.tile_first[t0], last_cur = a.tile_first[t0 + 1];
    if (lane == 0) issue(t0, 0, 0);
    uint64_t ts = (uint64_t)t0 * SUPER;
    {
        uint32_t cnt = last_cur - first_cur + 2;
        if (cnt <= kRowListCap) {
            if (lane < cnt) list_base[lane] = rel_clamp(a.offsets[first_cur + lane], ts);
            if (lane + 32 < cnt) list_base[lane + 32] = rel_clamp(a.offsets[first_cur + lane + 32], ts);
        }
    }
    uint32_t first_nxt = 0, last_nxt = 0;
    __syncwarp();

    LaneCtx c;
    c.tbl = tbl; c.events = events; c.eom = nullptr; c.stage = nullptr;
    c.list_abs = a.offsets; c.counters = a.counters; c.warm = 16;
    c.start_state = d.start_state; c.neutral_state = d.neutral_state;
    c.head_pad4 = c.tail_pad4 = 0;
    c.tbl_s = smem_u32(tbl);
    const int32_t chunk = (int32_t)(lane * CL);
    uint32_t g = 0;

    for (uint32_t sk = 0; t < t_end; sk++) {
        const bool has_next = t_nxt < t_end;
        ts = (uint64_t)t * SUPER;
        uint64_t n0 = 0, n1 = 0;
        uint32_t cnt_nxt = 0;
        if (has_next) {
            first_nxt = a.tile_first[t_nxt]; last_nxt = a.tile_first[t_nxt + 1];
            cnt_nxt = last_nxt - first_nxt + 2;
            if (cnt_nxt <= kRowListCap) {
                if (lane < cnt_nxt) n0 = a.offsets[first_nxt + lane];
                if (lane + 32 < cnt_nxt) n1 = a.offsets[first_nxt + lane + 32];
            }
        }
        const uint32_t cnt_cur = last_cur - first_cur + 2;
        const int32_t *lst = list_base + (sk & 1) * kRowListCap;
        c.list = cnt_cur <= kRowListCap ? lst : nullptr;
        c.tile_start = (int64_t)ts;
        c.list_n = cnt_cur; c.msg_lo = first_cur;
        const uint32_t row = t * 32 + lane;
        const bool full_row = row < nrows_full;                                  // fed by the TMA
        const uint64_t row_start = (uint64_t)row * CL;
        const bool tail_row = row == nrows_full && row_start < a.total;          // the partial last row: read directly

        LongLane ll;
        if (full_row || tail_row) {
            uint32_t km;
            if (c.list) {
                km = 0;
                for (uint32_t j = 1; j + 1 < cnt_cur; j++) km += lst[j] <= chunk ? 1u : 0u;
            } else {
                km = find_msg(c, chunk);
            }
            ll.k = km;   // long_begin needs the warm-up bytes: after the first window has landed
        }
#pragma unroll 1
        for (int32_t w = 0; w < NWIN; w++, g++) {
            if (lane == 0) {
                if (w + 1 < NWIN) issue(t, (uint32_t)w + 1, (g + 1) & 1);
                else if (has_next) issue(t_nxt, 0, (g + 1) & 1);
            }
            mbar_wait(&bars[g & 1], (g >> 1) & 1);
            if (full_row) {
                if (w == 0) long_begin(c, ll, ll.k, row == 0, warm_buf + lane * 16);
                const uint8_t *wb = wbase + (g & 1) * (32 * WIN) + lane * WIN;
                const int32_t pos = chunk + w * WIN;
#pragma unroll 1
                for (int32_t q = 0; q < WIN; q += 16) {
                    long_boundary(c, ll, pos + q);
                    group16_sink(c, *reinterpret_cast<const Quad *>(wb + q), ll.e, ll.words, ll.msg, queue);
                }
            }
            __syncwarp();   // every lane is done with this buffer before lane 0 refills it (two windows on)
            if ((w & 3) == 3) queue.flush(c, lane, false);   // the check is a shared-memory load: every fourth window
        }
        if (tail_row) {
            const uint8_t *src = a.bytes + row_start;
            long_begin(c, ll, ll.k, row == 0, row == 0 ? src : src - 16);
            LongLane l1[1] = {ll};
            const uint8_t *s1[1] = {src};
            const int32_t p1[1] = {chunk};
            long_stream_n<1>(c, l1, s1, p1, (int32_t)(a.total - row_start), a.bytes + a.total);
            ll = l1[0];
        }
        uint32_t hm = 0xFFFFFFFFu, hw = 0;
        if (full_row || tail_row) { long_end(c, ll); hm = ll.head_msg; hw = ll.head_words; }
        {   // combine the head partials of neighbouring lanes that sit in the same message
            const unsigned full = 0xFFFFFFFFu;
            uint32_t prev = __shfl_up_sync(full, hm, 1);
            bool head = lane == 0 || hm != prev;
            uint32_t heads = __ballot_sync(full, head);
            uint32_t v = hw;
#pragma unroll
            for (int dlt = 1; dlt < 32; dlt <<= 1) {
                uint32_t tt = __shfl_up_sync(full, v, dlt);
                if ((int)lane >= dlt) v += tt;
            }
            uint32_t above = heads & ~((2u << lane) - 1u);
            uint32_t end = above ? (uint32_t)__ffs(above) - 2u : 31u;
            uint32_t upto = __shfl_sync(full, v, end);
            uint32_t before = __shfl_up_sync(full, v, 1);
            if (lane == 0) before = 0;
            if (head && hm != 0xFFFFFFFFu) {
                uint32_t sum = upto - before;
                if (sum) atomicAdd(a.counters + (uint64_t)hm * CNT_STRIDE + CNT_WORDS, sum);
            }
        }
        if (has_next && cnt_nxt <= kRowListCap) {
            int32_t *ln = list_base + ((sk + 1) & 1) * kRowListCap;
            const uint64_t tsn = (uint64_t)t_nxt * SUPER;
            if (lane < cnt_nxt) ln[lane] = rel_clamp(n0, tsn);
            if (lane + 32 < cnt_nxt) ln[lane + 32] = rel_clamp(n1, tsn);
        }
        first_cur = first_nxt; last_cur = last_nxt;
        t = t_nxt;
        t_nxt = t < t_end ? grab() : t;   // one tile ahead only: a tile a warp has claimed cannot be taken over by an idle one (warp-uniform: no more fetches past the end)
        __syncwarp();
    }
    queue.flush(c, lane, true);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

template <int NW, int LANE>
static cudaError_t launch_scan_rows_variant(const ScanArgs &a, const DeviceDfa &d, const CUtensorMap &tw, const CUtensorMap &tm,
                                            uint32_t nrows_full, uint32_t *tile_counter, int sm_count, cudaStream_t s) {
    size_t smem = scan_rows_smem_bytes(d, NW);
    static SmemConfigured configured;
    if (cudaError_t e = ensure_dynamic_smem(scan_rows_kernel<NW, LANE>, smem, configured)) return e;
    uint32_t need = (a.tile_end - a.tile_begin + NW - 1) / NW;
    uint32_t grid = need < (uint32_t)sm_count ? need : (uint32_t)sm_count;
    scan_rows_kernel<NW, LANE><<<grid, NW * 32, smem, s>>>(a, d, tw, tm, nrows_full, tile_counter);
    return cudaGetLastError();
}

// needs at least one full row and a 16-byte aligned arena; the caller falls back to the mini-tile kernel otherwise
bool scan_rows_usable(const ScanArgs &a) {
    return a.total >= kRowLaneShort && (reinterpret_cast<uintptr_t>(a.bytes) & 15u) == 0 && encode_tiled_fn() != nullptr;
}

cudaError_t launch_scan_rows(const ScanArgs &a, const DeviceDfa &d, uint32_t nw, uint32_t lane_bytes, uint32_t *tile_counter, int sm_count,
                             cudaStream_t s) {
    if (a.tile_end <= a.tile_begin) return cudaSuccess;
    EncodeTiledFn enc = encode_tiled_fn();
    if (!enc) return cudaErrorNotSupported;
    const uint32_t win = row_win(nw);
    if (lane_bytes != kRowLaneLong && lane_bytes != kRowLaneMid && lane_bytes != kRowLaneShort) return cudaErrorInvalidValue;
    const uint64_t nrows_full = a.total / lane_bytes;
    CUtensorMap tw, tm;
    const cuuint64_t dims[2] = {lane_bytes, nrows_full};
    const cuuint64_t strides[1] = {lane_bytes};
    const cuuint32_t box_w[2] = {win, 32}, box_m[2] = {16, 32}, es[2] = {1, 1};
    void *base = const_cast<uint8_t *>(a.bytes);
    if (enc(&tw, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, base, dims, strides, box_w, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
            CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS ||
        enc(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, base, dims, strides, box_m, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
            CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
        return cudaErrorInvalidValue;
    const uint32_t nrows = (uint32_t)nrows_full;
#define TRIAGE_ROWS_CASE(NW_)                                                                                                        \
    case NW_:                                                                                                                        \
        if (lane_bytes == kRowLaneLong) return launch_scan_rows_variant<NW_, kRowLaneLong>(a, d, tw, tm, nrows, tile_counter, sm_count, s); \
        if (lane_bytes == kRowLaneMid) return launch_scan_rows_variant<NW_, kRowLaneMid>(a, d, tw, tm, nrows, tile_counter, sm_count, s);   \
        return launch_scan_rows_variant<NW_, kRowLaneShort>(a, d, tw, tm, nrows, tile_counter, sm_count, s);
    switch (nw) {
        TRIAGE_ROWS_CASE(24)
        TRIAGE_ROWS_CASE(20)
        TRIAGE_ROWS_CASE(16)
        default: return cudaErrorInvalidValue;
    }
#undef TRIAGE_ROWS_CASE
}

// ---------------------------------------------------------------------------------- pair lanes --
// Row lanes with two bytes per lookup (scan_pair.cuh): same tiles, windows, 2-D TMA feed and dynamic tile hand-out as
// scan_rows_kernel, but the table in shared memory is the FILTER automaton over 4-bit class pairs (8 lookups per 16
// bytes), word starts are counted by SIMD-in-register compares, and the groups the filter flags (or that hold one of the
// special lead bytes, or follow such a group) are queued per warp - (message, position) - and walked through the exact
// automaton (its column-merged copy in shared memory) by the whole warp, one queued group per lane, which reads the
// group's bytes back from the arena.  The default scan of dictionaries that have a pair filter.
constexpr uint32_t kPairQueueCap = 64;     // flagged groups per warp: (message, arena position / 16 | "words too" << 31), 8 bytes each
constexpr uint32_t kPairQueueFlush = 32;   // walked by the whole warp, one group per lane, once this many wait (more can arrive in one window: those beyond the cap are walked by their own lane)
// Shared-memory layout.  The filter table sits at the shared address kPairTableShared (scan_pair.cuh: its entries are
// then addressed by a multiply-add and an immediate); the kernel's dynamic shared memory starts at kPairSmemBase (the
// first KB of the window is reserved by the system - the kernel checks), so the table is at offset 0xFC00 and the other
// pieces are laid out around it: the small ones and the first warps' staging buffers before, the other warps' after.
constexpr uint32_t kPairSmemBase = 0x400u;
struct PairSmem { uint32_t table, xtable, xcls, warp_a, warp_b, n_a, aux0, total, warp_bytes, aux_bytes, warm_off, list_off, queue_off; };
__host__ __device__ inline PairSmem pair_layout(uint32_t f_states, uint32_t x_states, uint32_t x_cls, uint32_t nw) {
    PairSmem L;
    L.warm_off = 2 * 32 * kRowWin;
    L.warp_bytes = L.warm_off + 512;
    L.list_off = 0;
    L.queue_off = 2 * 4 * kRowListCap;
    L.aux_bytes = ((L.queue_off + 8 + 8 * kPairQueueCap + 15u) / 16u) * 16u;
    uint32_t o = 1024;   // mbarriers
    L.xtable = o; o += ((x_states * x_cls * 2u + 127u) / 128u) * 128u;   // exact automaton, column-merged: flagged groups only
    L.xcls = o; o += 256u;
    L.aux0 = o; o += nw * L.aux_bytes;
    o = ((o + 127u) / 128u) * 128u;
    L.table = kPairTableShared - kPairSmemBase;                          // filter automaton, Moore form: one byte per entry
    L.warp_a = o;
    L.n_a = o <= L.table ? (L.table - o) / L.warp_bytes : 0u;
    if (L.n_a > nw) L.n_a = nw;
    L.warp_b = L.table + ((f_states * kPairRowBytes + 127u) / 128u) * 128u;
    L.total = o <= L.table ? L.warp_b + (nw - L.n_a) * L.warp_bytes : 0xFFFFFFFFu;   // (no room before the table: no such launch)
    return L;
}
size_t scan_pair_smem_bytes(const DeviceDfa &d, uint32_t nw) { return pair_layout(d.f_states, d.n_states, d.n_cmp, nw).total; }
uint32_t pick_pair_warps(const DeviceDfa &d, size_t max_smem) {
    if (!d.has_filter || d.wide || d.warm != 16 || d.reset_pad < 0) return 0;
    if (const char *v = getenv("TRIAGE_PAIR")) { if (atoi(v) == 0) return 0; }   // measurement knob: keep to the one-lookup-per-byte lanes
    if (const char *w = getenv("TRIAGE_PAIR_WARPS")) {
        uint32_t nw = (uint32_t)atoi(w);
        if ((nw == 16 || nw == 20 || nw == 24 || nw == 28) && scan_pair_smem_bytes(d, nw) <= max_smem) return nw;
    }
    for (uint32_t nw : {24u, 20u, 16u})
        if (scan_pair_smem_bytes(d, nw) <= max_smem) return nw;
    return 0;
}

struct PairEventQueue {
    uint32_t base_s;   // shared address: count word (+ 4 bytes of padding), then 8-byte slots
    int64_t tile_start;
    // the bytes are not copied: the flush reads the group and the 16 bytes before it back from the arena (L2)
    __device__ __forceinline__ void push(const PairCtx &c, const Quad &, const Quad &, uint32_t msg, int32_t pos, bool count_words) {
        const uint32_t where = (uint32_t)((uint64_t)(tile_start + pos) >> 4) | (count_words ? 0x80000000u : 0u);
        uint32_t slot;
        asm volatile("atom.shared.add.u32 %0, [%1], 1;" : "=r"(slot) : "r"(base_s) : "memory");
        if (slot < kPairQueueCap) asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(base_s + 8 + slot * 8), "r"(msg), "r"(where) : "memory");
        else walk(c, msg, where);   // queue full: walk it now
    }
    static __device__ __noinline__ void walk(const PairCtx &c, uint32_t msg, uint32_t where) {
        const uint64_t at = (uint64_t)(where & 0x7FFFFFFFu) << 4;
        uint4 w = make_uint4(c.pad4, c.pad4, c.pad4, c.pad4);
        if (at) w = __ldg(reinterpret_cast<const uint4 *>(c.arena + at - 16));
        const uint4 q = __ldg(reinterpret_cast<const uint4 *>(c.arena + at));
        const uint32_t v[8] = {w.x, w.y, w.z, w.w, q.x, q.y, q.z, q.w};
        pair_replay_words(c, [&](int i) { return v[i]; }, msg, (where >> 31) != 0);
    }
    __device__ __noinline__ void flush(const PairCtx &c, uint32_t lane) {
        uint32_t n;
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(n) : "r"(base_s) : "memory");
        if (n > kPairQueueCap) n = kPairQueueCap;
        for (uint32_t k = lane; k < n; k += 32) {
            uint32_t msg, where;
            asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(msg), "=r"(where) : "r"(base_s + 8 + k * 8) : "memory");
            walk(c, msg, where);
        }
        __syncwarp();
        if (lane == 0) asm volatile("st.shared.u32 [%0], %1;" ::"r"(base_s), "r"(0u) : "memory");
        __syncwarp();
    }
};

template <int NW, int LANE, bool SIMPLE>
__global__ void __launch_bounds__(NW * 32, 1) scan_pair_kernel(ScanArgs a, DeviceDfa d, const __grid_constant__ CUtensorMap tm_win,
                                                              const __grid_constant__ CUtensorMap tm_warm, uint32_t nrows_full,
                                                              uint32_t *__restrict__ tile_counter) {
    extern __shared__ __align__(128) uint8_t smem[];
    constexpr int32_t CL = LANE, WIN = (int32_t)kRowWin, NWIN = CL / WIN;
    constexpr uint32_t SUPER = 32 * CL;
    const PairSmem L = pair_layout(d.f_states, d.n_states, d.n_cmp, NW);
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem) + warp * 2;
    uint8_t *tbl = smem + L.table;
    uint8_t *wbase = smem + (warp < L.n_a ? L.warp_a + warp * L.warp_bytes : L.warp_b + (warp - L.n_a) * L.warp_bytes);
    uint8_t *warm_buf = wbase + L.warm_off;
    uint8_t *aux = smem + L.aux0 + warp * L.aux_bytes;
    int32_t *list_base = reinterpret_cast<int32_t *>(aux + L.list_off);
    if (smem_u32(smem) != kPairSmemBase) __trap();   // the table's address is compiled into the lookups

    PairEventQueue queue;
    queue.base_s = smem_u32(aux + L.queue_off);
    if (lane == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("st.shared.u32 [%0], %1;" ::"r"(queue.base_s), "r"(0u) : "memory");
    }
    {
        const uint32_t *src = reinterpret_cast<const uint32_t *>(d.ftable);
        uint32_t *dst = reinterpret_cast<uint32_t *>(tbl);
        for (uint32_t i = tid; i < d.f_states * 64u; i += NW * 32) dst[(i >> 6) * (kPairRowBytes / 4) + (i & 63u)] = src[i];   // rows of 260 bytes: row r starts at bank r mod 32
        uint16_t *xd = reinterpret_cast<uint16_t *>(smem + L.xtable);
        for (uint32_t i = tid; i < d.n_states * d.n_cmp; i += NW * 32) xd[i] = d.ctable[i];
        for (uint32_t i = tid; i < 256; i += NW * 32) smem[L.xcls + i] = d.ccls[i];
    }
    __syncthreads();   // the only CTA-wide barrier

    const uint32_t nwarps = gridDim.x * NW;
    const uint32_t t_end = a.tile_end;
    const uint32_t t0 = a.tile_begin + blockIdx.x * NW + warp;
    if (t0 >= t_end) return;
#ifdef TRIAGE_SCAN_TRACE_BUILD
    uint64_t trace_t0 = 0, trace_last = 0, trace_worst = 0;
    uint32_t trace_tiles = 0, trace_worst_tile = 0;
    if (a.trace) { asm volatile("mov.u64 %0, %globaltimer;" : "=l"(trace_t0)); trace_last = trace_t0; }
#endif
    auto grab = [&]() {
        uint32_t v = 0;
        if (lane == 0) v = atomicAdd(tile_counter, 1u) + nwarps + a.tile_begin;
        return __shfl_sync(0xFFFFFFFFu, v, 0);
    };
    auto issue = [&](uint32_t tile, uint32_t w, uint32_t buf) {
        const int32_t row0 = (int32_t)(tile * 32u);
        mbar_expect_tx(&bars[buf], 32 * WIN + (w == 0 ? 512u : 0u));
        tma_load_2d(wbase + buf * (32 * WIN), &tm_win, (int32_t)(w * WIN), row0, &bars[buf]);
        if (w == 0) tma_load_2d(warm_buf, &tm_warm, CL - 16, row0 - 1, &bars[buf]);
    };

    uint32_t t = t0, t_nxt = grab();
    uint32_t first_cur = a.tile_first[t0], last_cur = a.tile_first[t0 + 1];
    if (lane == 0) issue(t0, 0, 0);
    uint64_t ts = (uint64_t)t0 * SUPER;
    {
        uint32_t cnt = last_cur - first_cur + 2;
        if (cnt <= kRowListCap) {
            if (lane < cnt) list_base[lane] = rel_clamp(a.offsets[first_cur + lane], ts);
            if (lane + 32 < cnt) list_base[lane + 32] = rel_clamp(a.offsets[first_cur + lane + 32], ts);
        }
    }
    uint32_t first_nxt = 0, last_nxt = 0;
    __syncwarp();

    PairCtx c;
    c.ftbl = tbl; c.fstart = d.f_start; c.ffired = d.f_fired; c.mul = d.f_mul; c.shift = d.f_shift;
    c.xtbl = reinterpret_cast<const uint16_t *>(smem + L.xtable); c.xcls = smem + L.xcls; c.xn = d.n_cmp;
    c.xtbl_s = smem_u32(smem + L.xtable); c.xcls_s = smem_u32(smem + L.xcls);
    c.events = d.events; c.neutral_state = d.neutral_state;
    c.pad4 = 0x01010101u * (uint32_t)(d.reset_pad & 0xFF);
    c.list_abs = a.offsets; c.counters = a.counters; c.arena = a.bytes;
    Quad padq;
    padq.a = padq.b = padq.c = padq.d = c.pad4;
    const int32_t chunk = (int32_t)(lane * CL);
    uint32_t g = 0;
    uint32_t queued = 0;   // groups waiting in the warp's queue (warp-uniform, carried from tile to tile)

    for (uint32_t sk = 0; t < t_end; sk++) {
        const bool has_next = t_nxt < t_end;
        ts = (uint64_t)t * SUPER;
        uint64_t n0 = 0, n1 = 0;
        uint32_t cnt_nxt = 0;
        if (has_next) {
            first_nxt = a.tile_first[t_nxt]; last_nxt = a.tile_first[t_nxt + 1];
            cnt_nxt = last_nxt - first_nxt + 2;
            if (cnt_nxt <= kRowListCap) {
                if (lane < cnt_nxt) n0 = a.offsets[first_nxt + lane];
                if (lane + 32 < cnt_nxt) n1 = a.offsets[first_nxt + lane + 32];
            }
        }
        const uint32_t cnt_cur = last_cur - first_cur + 2;
        const int32_t *lst = list_base + (sk & 1) * kRowListCap;
        c.list = cnt_cur <= kRowListCap ? lst : nullptr;
        c.tile_start = (int64_t)ts;
        queue.tile_start = (int64_t)ts;
        c.list_n = cnt_cur; c.msg_lo = first_cur;
        const uint32_t row = t * 32 + lane;
        const bool full_row = row < nrows_full;
        const uint64_t row_start = (uint64_t)row * CL;
        const bool tail_row = row == nrows_full && row_start < a.total;

        uint32_t all_full = (uint64_t)t * 32 + 32 <= nrows_full ? 1u : 0u;   // warp-uniform: every lane of this tile owns a full row
        asm volatile("" : "+r"(all_full));                                    // (kept in a register: not recomputed every window)
        PairLane P;
        uint32_t km = 0;
        if (full_row || tail_row) {
            if (c.list) { for (uint32_t j = 1; j + 1 < cnt_cur; j++) km += lst[j] <= chunk ? 1u : 0u; }
            else km = find_msg(c, chunk);
        }
#pragma unroll 1
        for (int32_t w = 0; w < NWIN; w++, g++) {
            if (lane == 0) {
                if (w + 1 < NWIN) issue(t, (uint32_t)w + 1, (g + 1) & 1);
                else if (has_next) issue(t_nxt, 0, (g + 1) & 1);
            }
            mbar_wait(&bars[g & 1], (g >> 1) & 1);
            if (full_row) {
                if (w == 0) pair_begin<SIMPLE>(c, P, km, row == 0 ? padq : *reinterpret_cast<const Quad *>(warm_buf + lane * 16));
                const Quad *win = reinterpret_cast<const Quad *>(wbase + (g & 1) * (32 * WIN) + lane * WIN);
                // The queue is emptied as soon as enough groups are waiting (a warp-uniform count kept in a register): text
                // full of multi-byte whitespace would otherwise overflow it and fall back to one-lane walks.  All lanes of
                // the warp own full rows here, or none does, except in the arena's last tile.
                const uint32_t pushed = pair_window<SIMPLE, WIN / 16>(c, P, win, chunk + w * WIN, queue);
                if (all_full) {
                    queued += __reduce_add_sync(0xFFFFFFFFu, pushed);
                    if (queued >= kPairQueueFlush) { queue.flush(c, lane); queued = 0; }
                }
            }
            __syncwarp();
            if (!all_full) queue.flush(c, lane);   // last tile: lanes diverge above, empty it between windows
        }
        if (tail_row) {   // the arena's partial last row: straight from global memory, flagged groups walked on the spot
            const uint8_t *src = a.bytes + row_start;
            PairSinkNow now;
            Quad wq = padq;
            if (row != 0) { const uint4 v = __ldg(reinterpret_cast<const uint4 *>(src - 16)); wq.a = v.x; wq.b = v.y; wq.c = v.z; wq.d = v.w; }
            pair_begin<SIMPLE>(c, P, km, wq);
            const int32_t nb = (int32_t)(a.total - row_start);
#pragma unroll 1
            for (int32_t q = 0; q < nb; q += 16) pair_group<SIMPLE>(c, P, stream_quad(src + q), chunk + q, now);
        }
        uint32_t hm = 0xFFFFFFFFu, hw = 0;
        if (full_row || tail_row) { pair_end(c, P); hm = P.L.head_msg; hw = P.L.head_words; }
        {   // combine the head partials of neighbouring lanes that sit in the same message
            const unsigned full = 0xFFFFFFFFu;
            uint32_t prev = __shfl_up_sync(full, hm, 1);
            bool head = lane == 0 || hm != prev;
            uint32_t heads = __ballot_sync(full, head);
            uint32_t v = hw;
#pragma unroll
            for (int dlt = 1; dlt < 32; dlt <<= 1) {
                uint32_t tt = __shfl_up_sync(full, v, dlt);
                if ((int)lane >= dlt) v += tt;
            }
            uint32_t above = heads & ~((2u << lane) - 1u);
            uint32_t end = above ? (uint32_t)__ffs(above) - 2u : 31u;
            uint32_t upto = __shfl_sync(full, v, end);
            uint32_t before = __shfl_up_sync(full, v, 1);
            if (lane == 0) before = 0;
            if (head && hm != 0xFFFFFFFFu) {
                uint32_t sum = upto - before;
                if (sum) atomicAdd(a.counters + (uint64_t)hm * CNT_STRIDE + CNT_WORDS, sum);
            }
        }
        if (has_next && cnt_nxt <= kRowListCap) {
            int32_t *ln = list_base + ((sk + 1) & 1) * kRowListCap;
            const uint64_t tsn = (uint64_t)t_nxt * SUPER;
            if (lane < cnt_nxt) ln[lane] = rel_clamp(n0, tsn);
            if (lane + 32 < cnt_nxt) ln[lane + 32] = rel_clamp(n1, tsn);
        }
        first_cur = first_nxt; last_cur = last_nxt;
        t = t_nxt;
#ifdef TRIAGE_SCAN_TRACE_BUILD
        trace_tiles++;
        if (a.trace) {
            uint64_t now;
            asm volatile("mov.u64 %0, %globaltimer;" : "=l"(now));
            if (now - trace_last > trace_worst) { trace_worst = now - trace_last; trace_worst_tile = (uint32_t)(ts / SUPER); }
            trace_last = now;
        }
#endif
        // one tile ahead only: a tile a warp has claimed cannot be taken over by an idle one (warp-uniform: no more fetches
        // past the end).  Claiming the last rounds' tiles only when the current one is done was measured: slower (0.91 ms
        // against 0.88), the cold starts cost more than the shorter tail brings.
        t_nxt = t < t_end ? grab() : t;
        __syncwarp();
    }
    queue.flush(c, lane);
#ifdef TRIAGE_SCAN_TRACE_BUILD
    if (a.trace && lane == 0) {
        uint64_t t1;
        uint32_t smid;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t1));
        asm volatile("mov.u32 %0, %smid;" : "=r"(smid));
        uint64_t *o = a.trace + 4ull * (blockIdx.x * NW + warp);
        o[0] = trace_t0; o[1] = t1; o[2] = trace_tiles | (trace_worst << 32); o[3] = smid | ((uint64_t)trace_worst_tile << 32);
    }
#endif
}

template <int NW, int LANE, bool SIMPLE>
static cudaError_t launch_scan_pair_variant(const ScanArgs &a, const DeviceDfa &d, const CUtensorMap &tw, const CUtensorMap &tm,
                                            uint32_t nrows_full, uint32_t *tile_counter, int sm_count, cudaStream_t s) {
    size_t smem = scan_pair_smem_bytes(d, NW);
    static SmemConfigured configured;
    if (cudaError_t e = ensure_dynamic_smem(scan_pair_kernel<NW, LANE, SIMPLE>, smem, configured)) return e;
    uint32_t need = (a.tile_end - a.tile_begin + NW - 1) / NW;
    uint32_t grid = need < (uint32_t)sm_count ? need : (uint32_t)sm_count;
    scan_pair_kernel<NW, LANE, SIMPLE><<<grid, NW * 32, smem, s>>>(a, d, tw, tm, nrows_full, tile_counter);
    return cudaGetLastError();
}

cudaError_t launch_scan_pair(const ScanArgs &a, const DeviceDfa &d, uint32_t nw, uint32_t lane_bytes, uint32_t *tile_counter, int sm_count,
                             cudaStream_t s) {
    if (a.tile_end <= a.tile_begin) return cudaSuccess;
    EncodeTiledFn enc = encode_tiled_fn();
    if (!enc) return cudaErrorNotSupported;
    if (lane_bytes != kRowLaneLong && lane_bytes != kRowLaneMid && lane_bytes != kRowLaneShort) return cudaErrorInvalidValue;
    const uint64_t nrows_full = a.total / lane_bytes;
    CUtensorMap tw, tm;
    const cuuint64_t dims[2] = {lane_bytes, nrows_full};
    const cuuint64_t strides[1] = {lane_bytes};
    const cuuint32_t box_w[2] = {kRowWin, 32}, box_m[2] = {16, 32}, es[2] = {1, 1};
    void *base = const_cast<uint8_t *>(a.bytes);
    if (enc(&tw, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, base, dims, strides, box_w, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
            CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS ||
        enc(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, base, dims, strides, box_m, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
            CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
        return cudaErrorInvalidValue;
    const uint32_t nrows = (uint32_t)nrows_full;
    const bool simple = d.f_mul == 1 && d.f_shift == 0;
#define TRIAGE_PAIR_LANE(NW_, LANE_)                                                                                          \
    return simple ? launch_scan_pair_variant<NW_, LANE_, true>(a, d, tw, tm, nrows, tile_counter, sm_count, s)                 \
                  : launch_scan_pair_variant<NW_, LANE_, false>(a, d, tw, tm, nrows, tile_counter, sm_count, s);
#define TRIAGE_PAIR_CASE(NW_)                                                     \
    case NW_:                                                                     \
        if (lane_bytes == kRowLaneLong) { TRIAGE_PAIR_LANE(NW_, kRowLaneLong) }   \
        if (lane_bytes == kRowLaneMid) { TRIAGE_PAIR_LANE(NW_, kRowLaneMid) }     \
        TRIAGE_PAIR_LANE(NW_, kRowLaneShort)
    switch (nw) {
        TRIAGE_PAIR_CASE(28)
        TRIAGE_PAIR_CASE(24)
        TRIAGE_PAIR_CASE(20)
        TRIAGE_PAIR_CASE(16)
        default: return cudaErrorInvalidValue;
    }
#undef TRIAGE_PAIR_CASE
#undef TRIAGE_PAIR_LANE
}

// ---------------------------------------------------------------------------------- fix-up --
// Messages in which a self-overlapping keyword ("aa", "abab") may have overlapping occurrences are re-counted by one
// thread each, walking the whole message from its start through the exact automaton (matching such a pattern kills its
// other threads = FindAllString's leftmost non-overlapping count).  Rare by construction; HBM traffic negligible.
__global__ void __launch_bounds__(128) fixup_kernel(const uint8_t *__restrict__ bytes, const uint64_t *__restrict__ offsets,
                                                    const uint32_t *__restrict__ lengths, uint32_t n, DeviceDfa x,
                                                    uint32_t *__restrict__ counters) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        uint32_t *cnt = counters + (uint64_t)i * CNT_STRIDE;
        if (!(cnt[CNT_Q] & 2u)) continue;
        const uint64_t off = offsets[i];
        const uint64_t len = lengths ? (uint64_t)lengths[i] : offsets[i + 1] - off;
        uint32_t acc[CNT_STRIDE] = {0, 0, 0, 0, 0, 0, 0, 0};
        auto event = [&](uint32_t id) {
            const EventPacked e = x.events[id];
            acc[CNT_KW + 0] += e.x & 0xFFFFu; acc[CNT_KW + 1] += e.x >> 16;
            acc[CNT_KW + 2] += e.y & 0xFFFFu; acc[CNT_KW + 3] += e.y >> 16;
            acc[CNT_POS] += e.z & 0xFFu; acc[CNT_NEG] += (e.z >> 8) & 0xFFu;
            if ((e.z >> 16) & 0xFFu) acc[CNT_Q] |= 1u;
        };
        uint32_t st = x.start_state;
        if (!x.wide) {
            for (uint64_t p = 0; p < len; p++) {
                uint32_t e = *reinterpret_cast<const uint16_t *>(x.table + st * kRowBytes + 2u * bytes[off + p]);
                st = e & 0xFFu;
                acc[CNT_WORDS] += e >> 15;
                if (e & 0x7F00u) event((e >> 8) & 0x7Fu);
            }
            const uint32_t eo = x.eom[st];
            acc[CNT_WORDS] += eo >> 7;
            if (eo & 0x7Fu) event(eo & 0x7Fu);
        } else {
            for (uint64_t p = 0; p < len; p++) {
                uint32_t e = x.wtable[(uint64_t)st * x.n_classes + x.cls_of[bytes[off + p]]];
                st = e & 0xFFFFFu;
                acc[CNT_WORDS] += e >> 31;
                if (e & 0x7FF00000u) event((e >> 20) & 0x7FFu);
            }
            const uint32_t eo = x.eom_w[st];
            acc[CNT_WORDS] += eo >> 31;
            if (eo & 0x7FFFFFFFu) event(eo & 0x7FFFFFFFu);
        }
        if (len && bytes[off + len - 1] == '?') acc[CNT_Q] |= 1u;
#pragma unroll
        for (int k = 0; k < CNT_STRIDE; k++) cnt[k] = acc[k];
    }
}

cudaError_t launch_fixup(const uint8_t *bytes, const uint64_t *offsets, const uint32_t *lengths, uint32_t n, const DeviceDfa &exact,
                         uint32_t *counters, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    uint32_t blocks = (n + 127) / 128;
    if (blocks > 148 * 16) blocks = 148 * 16;
    fixup_kernel<<<blocks, 128, 0, s>>>(bytes, offsets, lengths, n, exact, counters);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------------ regex patterns --
// AddKeywordPattern with a real regular expression (regex_vm.h): one message per thread walks every compiled pattern over
// its own bytes (ordered-thread simulation, regex_walk.cuh) and adds len(FindAllString) to the pattern's priority score.
// Runs after the scan and the fix-up (which overwrites the counters of the messages it recounts) and before finalize.
// Not a fast path: a dictionary pays for it only if it holds such patterns.
__global__ void __launch_bounds__(128) regex_count_kernel(RegexView v, const uint8_t *__restrict__ bytes, const uint64_t *__restrict__ offsets,
                                                          const uint32_t *__restrict__ lengths, uint32_t n, uint32_t *__restrict__ counters) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint64_t off = offsets[i];
        const uint64_t len = lengths ? (uint64_t)lengths[i] : offsets[i + 1] - off;
        if (len == 0) continue;
        for (uint32_t p = 0; p < v.n_programs; p++) {
            const uint32_t c = rx_count(v, p, bytes + off, len);
            if (c) counters[(uint64_t)i * CNT_STRIDE + CNT_KW + (v.programs[p].prio - 1)] += c;
        }
    }
}

cudaError_t launch_regex_count(const RegexView &v, const uint8_t *bytes, const uint64_t *offsets, const uint32_t *lengths, uint32_t n,
                               uint32_t *counters, int sm_count, cudaStream_t s) {
    if (n == 0 || v.n_programs == 0) return cudaSuccess;
    uint32_t blocks = (n + 127) / 128;
    const uint32_t cap = (uint32_t)sm_count * 16;
    if (blocks > cap) blocks = cap;
    regex_count_kernel<<<blocks, 128, 0, s>>>(v, bytes, offsets, lengths, n, counters);
    return cudaGetLastError();
}

// -------------------------------------------------------------------------------- finalize --
__global__ void __launch_bounds__(256) finalize_kernel(FinalizeArgs a) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n) return;
    const uint4 c0 = reinterpret_cast<const uint4 *>(a.counters)[2 * (uint64_t)i];
    const uint4 c1 = reinterpret_cast<const uint4 *>(a.counters)[2 * (uint64_t)i + 1];
    const uint4 mraw = reinterpret_cast<const uint4 *>(a.meta)[i];
    const uint64_t off_i = a.offsets[i];
    const uint64_t len = a.lengths ? (uint64_t)a.lengths[i] : a.offsets[i + 1] - off_i;
    const uint32_t preset = mraw.x & 0xFFu, up = (mraw.x >> 8) & 0xFFu, ud = (mraw.x >> 16) & 0xFFu, mp = mraw.x >> 24;
    const uint32_t mflags = mraw.y & 0xFFu;
    const uint32_t sc[5] = {0, c0.y, c0.z, c0.w, c1.x};

    uint32_t label = preset, reason = TRIAGE_REASON_NONE, rflags = 0, words = 0, question = 0, srt = 0, shi = 0;
    int32_t sentiment = 0;
    if (preset == TRIAGE_PRIO_NORMAL || preset == 0) {           // preprocessor.go:63-65
        uint32_t tie = 0;
        if (up != TRIAGE_UP_ABSENT) {                            // :68-82
            if (up >= 1 && up <= 4) { label = up; reason = TRIAGE_REASON_USER_OVERRIDE; }
        } else if (ud != 0) {                                    // :83-86
            label = ud; reason = TRIAGE_REASON_USER_DEFAULT;
        } else {
            if (len == 0) {                                      // :120-138
                label = (mp >= 1 && mp <= 4) ? mp : (uint32_t)a.default_prio;
            } else {                                             // :141-165, ties: most urgent wins + TIE flag
                uint32_t best = 0, bp = (uint32_t)a.default_prio;
#pragma unroll
                for (uint32_t p = 1; p <= 4; p++) if (sc[p] > best) { best = sc[p]; bp = p; }
                if (best > 0) for (uint32_t p = bp + 1; p <= 4; p++) if (sc[p] == best) tie = 1;
                label = bp;
                srt = sc[1] > 0xFFFFu ? 0xFFFFu : sc[1];
                shi = sc[2] > 0xFFFFu ? 0xFFFFu : sc[2];
            }
            if (label != preset) reason = TRIAGE_REASON_CONTENT_KEYWORDS;
        }
        rflags = TRIAGE_RF_ANALYZED | (tie ? TRIAGE_RF_TIE : 0);
        if (len > 0) {                                           // :199-249
            words = c0.x;
            sentiment = c1.y > c1.z ? 1 : (c1.z > c1.y ? -1 : 0);
            question = (c1.w & 1u) ? 1 : 0;
            if (a.lengths && a.bytes[off_i + len - 1] == '?') question = 1;   // HasSuffix(content, "?") (the unpadded scan does it itself)
            rflags |= TRIAGE_RF_HAS_ANALYSIS;
        }
    } else if ((mflags & TRIAGE_META_ANALYZE) && len > 0) {
        // api/handlers.go:180-190 runs AnalyzeMessageContent on every message, also on those ProcessMessage leaves alone:
        // the analysis fields are filled in for the emit, Metadata (TRIAGE_RF_HAS_ANALYSIS) stays untouched
        words = c0.x;
        sentiment = c1.y > c1.z ? 1 : (c1.z > c1.y ? -1 : 0);
        question = (c1.w & 1u) ? 1 : 0;
        if (a.lengths && a.bytes[off_i + len - 1] == '?') question = 1;
    }
    // queue level = pre-rule label (api/handlers.go:202); rules mutate msg.Priority before the push
    uint32_t level = label, bucket = 0xFF, admit = TRIAGE_E_NOT_FOUND, accepted = 0;
    if (a.single_queue) {                                        // BatchPushMessages(queueName, ...): one heap, any int priority
        if (a.rules_on) {
            if (mflags & TRIAGE_META_VIP) label = TRIAGE_PRIO_HIGH;
            else if (len > 10000) label = TRIAGE_PRIO_LOW;
        }
        bucket = label < 15 ? label : 15;
        admit = TRIAGE_OK;
        accepted = 1;
    } else if (level >= 1 && level <= 4) {
        if (a.rules_on) {                                        // queue_manager.go:451-466 / queue_factory.go:211-233
            if (mflags & TRIAGE_META_VIP) label = TRIAGE_PRIO_HIGH;
            else if (len > 10000) label = TRIAGE_PRIO_LOW;
        }
        bucket = (level - 1) * 4 + (label - 1);
        admit = TRIAGE_OK;     // provisional: launch_admit applies the capacity rule
        accepted = 1;
    }
    uint4 r;
    r.x = label | (reason << 8) | ((uint32_t)(uint8_t)(int8_t)sentiment << 16) | (question << 24);
    r.y = accepted | (admit << 8) | (rflags << 16) | (level << 24);
    r.z = words;
    r.w = srt | (shi << 16);
    reinterpret_cast<uint4 *>(a.results)[i] = r;
    a.bucket[i] = (uint8_t)bucket;
    if (a.ts && i > 0 && a.ts[i] < a.ts[i - 1]) atomicOr(a.flags, 1u);
}

cudaError_t launch_finalize(const FinalizeArgs &a, cudaStream_t s) {
    if (a.n == 0) return cudaSuccess;
    finalize_kernel<<<(a.n + 255) / 256, 256, 0, s>>>(a);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------- order ---
// Stable multi-split with up to 256 buckets.  Position k (in perm_in order, identity when null)
// holds item perm_in[k]; digit(item) in [0,256) or 256 = leave out.
struct DigU8 {
    const uint8_t *d; bool drop_ff;
    __device__ uint32_t operator()(uint32_t item) const { uint32_t v = d[item]; return (drop_ff && v == 0xFFu) ? 256u : v; }
};
struct DigLevel {
    const uint8_t *b; uint32_t shift;   // shift 2: routed buckets (level-1)*4+(prio-1); shift 8: single queue
    __device__ uint32_t operator()(uint32_t item) const { uint32_t v = b[item]; return v == 0xFFu ? 256u : (v >> shift); }
};
struct DigKey {
    const uint64_t *k; uint32_t shift;
    __device__ uint32_t operator()(uint32_t item) const { return (uint32_t)(k[item] >> shift) & 0xFFu; }
};

template <class Dig>
__global__ void __launch_bounds__(kOrderBlock) split_hist_kernel(Dig dig, const uint32_t *__restrict__ perm_in, uint32_t n,
                                                                uint32_t *__restrict__ blockhist) {
    __shared__ uint32_t h[256];
    for (uint32_t i = threadIdx.x; i < 256; i += blockDim.x) h[i] = 0;
    __syncthreads();
    uint32_t k = blockIdx.x * kOrderBlock + threadIdx.x;
    if (k < n) {
        uint32_t item = perm_in ? perm_in[k] : k;
        uint32_t dg = dig(item);
        if (dg < 256) atomicAdd(&h[dg], 1u);
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < 256; i += blockDim.x) blockhist[(uint64_t)blockIdx.x * 256 + i] = h[i];
}

// exclusive scan in digit-major order over (digit, block); blockhist[b][d] <- global exclusive offset,
// totals[d] <- offset of (d, block 0), totals[256] <- total kept.  One CTA; thread t owns histogram rows t, t + 1024, ...
// and takes 16 digits of a row at a time (one 64-byte line, four 128-bit loads issued together): the 16 block-wide
// scans of a digit group run on registers with warp shuffles.
__global__ void __launch_bounds__(1024) split_scan_kernel(uint32_t *__restrict__ blockhist, uint32_t nblocks, uint32_t nb,
                                                          uint32_t *__restrict__ totals) {
    __shared__ uint32_t wt[32][17];     // per warp and digit of the group: warp total, then its exclusive prefix
    __shared__ uint32_t group_tot[16];  // per digit of the group: total over the current chunk of 1024 rows
    const uint32_t tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const unsigned full = 0xFFFFFFFFu;
    uint32_t run = 0;                   // offset of (first digit of the group, block 0)
    auto load_row = [&](uint32_t b, uint32_t d0, uint32_t (&v)[16]) {
        if (b < nblocks) {
            const uint4 *p = reinterpret_cast<const uint4 *>(blockhist + (size_t)b * 256 + d0);
#pragma unroll
            for (int q = 0; q < 4; q++) { uint4 x = p[q]; v[4 * q] = x.x; v[4 * q + 1] = x.y; v[4 * q + 2] = x.z; v[4 * q + 3] = x.w; }
#pragma unroll
            for (int j = 0; j < 16; j++) if (d0 + j >= nb) v[j] = 0;
        } else {
#pragma unroll
            for (int j = 0; j < 16; j++) v[j] = 0;
        }
    };
    // block-wide exclusive scan of v[j] over the 1024 threads, for all 16 j; group_tot[j] <- sum over the threads
    auto scan16 = [&](uint32_t (&v)[16], uint32_t (&ex)[16]) {
#pragma unroll
        for (int j = 0; j < 16; j++) {
            uint32_t inc = v[j];
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { uint32_t t = __shfl_up_sync(full, inc, d); if ((int)lane >= d) inc += t; }
            ex[j] = inc - v[j];
            if (lane == 31) wt[wid][j] = inc;
        }
        __syncthreads();
        if (wid < 16) {   // warp j scans the 32 warp totals of digit j
            uint32_t w = wt[lane][wid], inc = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { uint32_t t = __shfl_up_sync(full, inc, d); if ((int)lane >= d) inc += t; }
            wt[lane][wid] = inc - w;
            if (lane == 31) group_tot[wid] = inc;
        }
        __syncthreads();
#pragma unroll
        for (int j = 0; j < 16; j++) ex[j] += wt[wid][j];
    };
    for (uint32_t d0 = 0; d0 < nb; d0 += 16) {
        // pass 1 (only when the rows do not fit one chunk): totals of the group's digits over all rows
        uint32_t dig_tot[16];
        if (nblocks > 1024) {
            uint32_t acc[16], ex[16];
#pragma unroll
            for (int j = 0; j < 16; j++) acc[j] = 0;
            for (uint32_t b = tid; b < nblocks; b += 1024) {
                uint32_t v[16];
                load_row(b, d0, v);
#pragma unroll
                for (int j = 0; j < 16; j++) acc[j] += v[j];
            }
            scan16(acc, ex);
#pragma unroll
            for (int j = 0; j < 16; j++) dig_tot[j] = group_tot[j];
            __syncthreads();
        }
        // pass 2: chunk by chunk, digit-major inside the group
        uint32_t carry[16];
#pragma unroll
        for (int j = 0; j < 16; j++) carry[j] = 0;
        for (uint32_t base = 0; base < nblocks; base += 1024) {
            const uint32_t b = base + tid;
            uint32_t v[16], ex[16];
            load_row(b, d0, v);
            scan16(v, ex);
            if (nblocks <= 1024) {
#pragma unroll
                for (int j = 0; j < 16; j++) dig_tot[j] = group_tot[j];
            }
            uint32_t dbase = run;
#pragma unroll
            for (int j = 0; j < 16; j++) {
                if (b < nblocks && d0 + j < nb) blockhist[(size_t)b * 256 + d0 + j] = dbase + carry[j] + ex[j];
                if (b == 0 && d0 + j < nb) totals[d0 + j] = dbase;
                carry[j] += group_tot[j];
                dbase += dig_tot[j];
            }
            __syncthreads();   // group_tot / wt are reused by the next chunk
        }
#pragma unroll
        for (int j = 0; j < 16; j++) run += dig_tot[j];
    }
    for (uint32_t dgt = nb + tid; dgt <= 256; dgt += 1024) totals[dgt] = run;
}

struct ActPerm {
    uint32_t *out;
    __device__ void operator()(uint32_t item, uint32_t dg, uint32_t global_pos, uint32_t rank_in_bucket) const { out[global_pos] = item; }
    __device__ void dropped(uint32_t item) const {}
};
struct ActAdmit {   // capacity rule of MultiLevelQueue.Push (queue.go:101-103), evaluated in arrival order
    const uint64_t *pending; const int64_t *max_size; const uint8_t *bucket_in; uint8_t *bucket_out; triage_result *results;
    __device__ void operator()(uint32_t item, uint32_t dg, uint32_t global_pos, uint32_t rank_in_bucket) const {
        int64_t cap = max_size[dg];
        bool ok = cap <= 0 || (int64_t)(pending[dg] + rank_in_bucket) < cap;
        if (!ok) {
            uint8_t *r = reinterpret_cast<uint8_t *>(results + item);
            r[4] = 0;                 // accepted
            r[5] = TRIAGE_E_FULL;     // admit
            bucket_out[item] = 0xFF;
        } else bucket_out[item] = bucket_in[item];
    }
    __device__ void dropped(uint32_t item) const { bucket_out[item] = bucket_in[item]; }
};

template <class Dig, class Act>
__global__ void __launch_bounds__(kOrderBlock) split_scatter_kernel(Dig dig, Act act, const uint32_t *__restrict__ perm_in, uint32_t n,
                                                                   const uint32_t *__restrict__ blockhist, const uint32_t *__restrict__ totals) {
    __shared__ uint32_t wh[kOrderBlock / 32][256];   // per-warp digit counts -> exclusive prefix over warps
    const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (uint32_t i = tid; i < (kOrderBlock / 32) * 256; i += kOrderBlock) (&wh[0][0])[i] = 0;
    __syncthreads();
    uint32_t k = blockIdx.x * kOrderBlock + tid;
    uint32_t item = 0, dg = 256;
    if (k < n) { item = perm_in ? perm_in[k] : k; dg = dig(item); }
    uint32_t peers = __match_any_sync(0xFFFFFFFFu, dg);
    uint32_t rank = __popc(peers & ((1u << lane) - 1u));
    if (dg < 256 && rank == 0) wh[warp][dg] = __popc(peers);
    __syncthreads();
    if (tid < 256) {
        uint32_t run = 0;
#pragma unroll 4
        for (uint32_t w = 0; w < kOrderBlock / 32; w++) { uint32_t c = wh[w][tid]; wh[w][tid] = run; run += c; }
    }
    __syncthreads();
    if (dg < 256) {
        uint32_t base = blockhist[(uint64_t)blockIdx.x * 256 + dg];
        uint32_t local = wh[warp][dg] + rank;
        act(item, dg, base + local, base - totals[dg] + local);
    } else if (k < n) {
        act.dropped(item);
    }
}

template <class Dig, class Act>
static cudaError_t run_split(Dig dig, Act act, const uint32_t *perm_in, uint32_t n, uint32_t nb, OrderScratch sc, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    uint32_t nblocks = (n + kOrderBlock - 1) / kOrderBlock;
    if (nblocks > sc.nblocks_cap) return cudaErrorInvalidValue;
    split_hist_kernel<Dig><<<nblocks, kOrderBlock, 0, s>>>(dig, perm_in, n, sc.blockhist);
    split_scan_kernel<<<1, 1024, 0, s>>>(sc.blockhist, nblocks, nb, sc.totals);
    split_scatter_kernel<Dig, Act><<<nblocks, kOrderBlock, 0, s>>>(dig, act, perm_in, n, sc.blockhist, sc.totals);
    return cudaGetLastError();
}

cudaError_t launch_admit(const uint8_t *bucket_in, uint8_t *bucket_out, triage_result *results, uint32_t n,
                         const uint64_t *pending_dev, const int64_t *max_size_dev, int32_t single_queue, OrderScratch sc,
                         cudaStream_t s) {
    // pending / max_size are indexed by level-1 (digit of DigLevel); a single target queue uses index 0
    return run_split(DigLevel{bucket_in, single_queue ? 8u : 2u}, ActAdmit{pending_dev, max_size_dev, bucket_in, bucket_out, results},
                     nullptr, n, 4, sc, s);
}

__global__ void copy_totals_kernel(const uint32_t *totals, uint32_t *out) {
    uint32_t i = threadIdx.x;   // 256 threads; totals[d] are exclusive offsets, totals[256] the grand total
    uint32_t nxt = totals[i + 1];
    out[i] = nxt - totals[i];
}

cudaError_t launch_partition_u8(const uint8_t *digits, const uint32_t *perm_in, uint32_t n, uint32_t n_buckets, bool drop_ff,
                                uint32_t *perm_out, uint32_t *bucket_counts_dev, OrderScratch sc, cudaStream_t s) {
    if (n == 0) {
        if (bucket_counts_dev) return cudaMemsetAsync(bucket_counts_dev, 0, 256 * sizeof(uint32_t), s);
        return cudaSuccess;
    }
    cudaError_t e = run_split(DigU8{digits, drop_ff}, ActPerm{perm_out}, perm_in, n, n_buckets, sc, s);
    if (e != cudaSuccess) return e;
    if (bucket_counts_dev) copy_totals_kernel<<<1, 256, 0, s>>>(sc.totals, bucket_counts_dev);
    return cudaGetLastError();
}

cudaError_t launch_partition_key(const uint64_t *keys, uint32_t shift, const uint32_t *perm_in, uint32_t n, uint32_t *perm_out,
                                 OrderScratch sc, cudaStream_t s) {
    return run_split(DigKey{keys, shift}, ActPerm{perm_out}, perm_in, n, 256, sc, s);
}

__global__ void __launch_bounds__(256) key_digit_hist_kernel(const uint64_t *__restrict__ keys, const uint32_t *__restrict__ perm, uint32_t n,
                                                             uint32_t *__restrict__ hist) {
    __shared__ uint32_t h[8 * 256];
    for (uint32_t i = threadIdx.x; i < 8 * 256; i += blockDim.x) h[i] = 0;
    __syncthreads();
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
        uint64_t key = keys[perm ? perm[k] : k];
#pragma unroll
        for (int b = 0; b < 8; b++) atomicAdd(&h[b * 256 + ((key >> (8 * b)) & 0xFF)], 1u);
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < 8 * 256; i += blockDim.x) if (h[i]) atomicAdd(&hist[i], h[i]);
}

cudaError_t launch_key_digit_hist(const uint64_t *keys, const uint32_t *perm, uint32_t n, uint32_t *hist, cudaStream_t s) {
    if (n == 0) return cudaSuccess;
    uint32_t blocks = (n + 255) / 256;
    if (blocks > 1184) blocks = 1184;
    key_digit_hist_kernel<<<blocks, 256, 0, s>>>(keys, perm, n, hist);
    return cudaGetLastError();
}

// ------------------------------------------------------------------------ multi-GPU merge ---
__global__ void build_run_kernel(const uint32_t *__restrict__ order, uint32_t n_adm, const uint8_t *__restrict__ bucket,
                                 const uint64_t *__restrict__ ts, uint64_t ts_base, uint64_t gidx_base,
                                 uint64_t *__restrict__ run_keys, uint32_t *__restrict__ run_gidx) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_adm) return;
    uint32_t item = order[k];
    (void)bucket;
    run_keys[k] = ts ? ts[item] : ts_base + item;
    run_gidx[k] = (uint32_t)(gidx_base + item);
}

cudaError_t launch_build_run(const uint32_t *order, uint32_t n_adm, const uint8_t *bucket, const uint64_t *ts, uint64_t ts_base,
                             uint64_t gidx_base, uint64_t *run_keys, uint32_t *run_gidx, cudaStream_t s) {
    if (n_adm == 0) return cudaSuccess;
    build_run_kernel<<<(n_adm + 255) / 256, 256, 0, s>>>(order, n_adm, bucket, ts, ts_base, gidx_base, run_keys, run_gidx);
    return cudaGetLastError();
}

// per-rank header for the merge: [0] admitted, [1..16] bucket counts, [17..32] first time key, [33..48] last time key per
// bucket, read straight from the local order (key = enqueue time, or rank base + arrival index when times are implicit)
struct Counts16 { uint32_t c[16]; };
__global__ void header_kernel(const uint32_t *__restrict__ order, uint32_t n_adm, Counts16 cnt, const uint64_t *__restrict__ ts,
                              uint64_t ts_base, uint64_t *__restrict__ header) {
    uint32_t b = threadIdx.x;
    if (b == 0) header[0] = n_adm;
    if (b < 16) {
        uint32_t start = 0;
        for (uint32_t i = 0; i < b; i++) start += cnt.c[i];
        header[1 + b] = cnt.c[b];
        uint64_t first = 0, last = 0;
        if (cnt.c[b]) {
            uint32_t i0 = order[start], i1 = order[start + cnt.c[b] - 1];
            first = ts ? ts[i0] : ts_base + i0;
            last = ts ? ts[i1] : ts_base + i1;
        }
        header[17 + b] = first;
        header[33 + b] = last;
    }
}

cudaError_t launch_header(const uint32_t *order, uint32_t n_adm, const uint32_t *counts16_host, const uint64_t *ts, uint64_t ts_base,
                          uint64_t *header_dev, cudaStream_t s) {
    Counts16 c;
    for (int i = 0; i < 16; i++) c.c[i] = counts16_host[i];
    header_kernel<<<1, 32, 0, s>>>(order, n_adm, c, ts, ts_base, header_dev);
    return cudaGetLastError();
}

// fast merge: the ranks' key ranges do not interleave inside any bucket, so element j of (rank, bucket) goes to
// dst[rank][bucket] + (j - lstart[rank][bucket])
__global__ void __launch_bounds__(256) concat_kernel(const uint32_t *__restrict__ all_order, ConcatTables t,
                                                     uint32_t *__restrict__ order_out) {
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t q = blockIdx.y;
    if (j >= t.cnt[q]) return;
    uint64_t idx = (uint64_t)q * t.cap + j;
    uint32_t b = 0;
#pragma unroll
    for (uint32_t i = 1; i < 16; i++) if (j >= t.lstart[q][i]) b = i;   // lstart is non-decreasing; the last i wins
    order_out[t.dst[q][b] + (j - t.lstart[q][b])] = all_order[idx] + t.base[q];   // local index -> global index
}

cudaError_t launch_concat(const uint32_t *all_order, const ConcatTables &t, uint32_t *order_out, cudaStream_t s) {
    dim3 grid((t.cap + 255) / 256, t.n_ranks);
    concat_kernel<<<grid, 256, 0, s>>>(all_order, t, order_out);
    return cudaGetLastError();
}

// general merge: element j of run q goes to  j + sum over other runs of (#keys < mine, or <= mine for
// lower-numbered runs), which keeps equal keys in rank order (lower rank = earlier sequence numbers)
__global__ void __launch_bounds__(256) merge_strided_kernel(const uint64_t *__restrict__ all_keys, const uint32_t *__restrict__ all_gidx,
                                                            ConcatTables t, uint32_t *__restrict__ order_out) {
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t q = blockIdx.y;
    if (j >= t.cnt[q]) return;
    const uint64_t key = all_keys[(uint64_t)q * t.cap + j];
    uint32_t b = 0;
#pragma unroll
    for (uint32_t i = 1; i < 16; i++) if (j >= t.lstart[q][i]) b = i;
    // destination = start of bucket b in the merged order + my rank inside the bucket across all runs
    uint64_t pos = t.dst[0][b] + (j - t.lstart[q][b]);
    for (uint32_t o = 0; o < t.n_ranks; o++) {
        if (o == q) continue;
        const uint64_t *k = all_keys + (uint64_t)o * t.cap;
        const uint32_t ob = t.lstart[o][b];
        const uint32_t oe = b < 15 ? t.lstart[o][b + 1] : t.cnt[o];
        uint32_t lo = ob, hi = oe;
        const bool le = o < q;
        while (lo < hi) {
            uint32_t mid = (lo + hi) >> 1;
            uint64_t km = k[mid];
            if (le ? km <= key : km < key) lo = mid + 1; else hi = mid;
        }
        pos += lo - ob;
    }
    order_out[pos] = all_gidx[(uint64_t)q * t.cap + j];
}

// ---- one-exchange merge for contiguous sequence shards (engine-stamped enqueue times): every rank contributes
// [32-word header: admitted, 16 bucket counts | its local order] at a common stride, one all-gather moves all of it, and
// the concatenation works out its own geometry from the gathered headers - no host round trip between the two.
__global__ void __launch_bounds__(32) pack_header_kernel(const uint32_t *__restrict__ counts16, const uint32_t *__restrict__ n_adm,
                                                         uint32_t *__restrict__ gsrc) {
    const uint32_t i = threadIdx.x;
    if (i < kMergeHeaderWords) gsrc[i] = i == 0 ? *n_adm : (i <= 16 ? counts16[i - 1] : 0u);
}

// header by one warp, the order by the copy engine: no SM is taken from the scan that runs next to it
cudaError_t launch_pack_run(const uint32_t *order, const uint32_t *counts16_dev, const uint32_t *n_adm_dev, uint32_t n, uint32_t *gsrc,
                            cudaStream_t s) {
    static_assert(kMergeHeaderWords <= 32, "one warp writes the header");
    pack_header_kernel<<<1, 32, 0, s>>>(counts16_dev, n_adm_dev, gsrc);
    if (cudaError_t e = cudaGetLastError()) return e;
    if (n == 0) return cudaSuccess;
    return cudaMemcpyAsync(gsrc + kMergeHeaderWords, order, (size_t)n * sizeof(uint32_t), cudaMemcpyDeviceToDevice, s);
}

// The kernel runs on the few SMs the persistent scan of the next step leaves free (engine.cu: comm_sms), so it is a small
// grid of grid-stride copies, and the geometry is worked out once per CTA from a parallel read of the gathered headers.
__global__ void __launch_bounds__(256) concat_gathered_kernel(const uint32_t *__restrict__ all, MergeBases mb, uint32_t *__restrict__ order_out,
                                                              uint64_t capacity, uint64_t *__restrict__ n_total_out) {
    __shared__ uint32_t hdr[8][17], lstart[8][16], dst[8][16];
    __shared__ uint64_t total_s;
    for (uint32_t i = threadIdx.x; i < mb.n_ranks * 17u; i += blockDim.x) hdr[i / 17u][i % 17u] = all[(uint64_t)(i / 17u) * mb.stride + (i % 17u)];
    __syncthreads();
    if (threadIdx.x == 0) {   // geometry from the gathered headers: bucket-major, rank-minor
        uint64_t total = 0;
        for (uint32_t q = 0; q < mb.n_ranks; q++) total += hdr[q][0];
        uint32_t d = 0;
        for (uint32_t b = 0; b < 16; b++)
            for (uint32_t q = 0; q < mb.n_ranks; q++) { dst[q][b] = d; d += hdr[q][1 + b]; }
        for (uint32_t q = 0; q < mb.n_ranks; q++) {
            uint32_t acc = 0;
            for (uint32_t b = 0; b < 16; b++) { lstart[q][b] = acc; acc += hdr[q][1 + b]; }
        }
        total_s = total;
    }
    __syncthreads();
    const uint32_t q = blockIdx.y;
    if (q == 0 && blockIdx.x == 0 && threadIdx.x == 0 && n_total_out) *n_total_out = total_s;
    if (total_s > capacity) return;   // too small an output buffer: nothing is written, the host reports it
    const uint32_t cnt = hdr[q][0];
    const uint32_t *src = all + (uint64_t)q * mb.stride + kMergeHeaderWords;
    const uint32_t base = mb.base[q];
    for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < cnt; j += gridDim.x * blockDim.x) {
        uint32_t b = 0;
#pragma unroll
        for (uint32_t i = 1; i < 16; i++) if (j >= lstart[q][i]) b = i;
        order_out[dst[q][b] + (j - lstart[q][b])] = src[j] + base;
    }
}

cudaError_t launch_concat_gathered(const uint32_t *all, const MergeBases &mb, uint32_t n_per_rank, uint32_t *order_out, uint64_t capacity,
                                   uint64_t *n_total_dev, cudaStream_t s) {
    uint32_t gx = (n_per_rank + 4095) / 4096;   // 16 elements a thread at least
    if (gx > 48) gx = 48;
    if (gx == 0) gx = 1;
    dim3 grid(gx, mb.n_ranks);
    concat_gathered_kernel<<<grid, 256, 0, s>>>(all, mb, order_out, capacity, n_total_dev);
    return cudaGetLastError();
}

cudaError_t launch_merge_strided(const uint64_t *all_keys, const uint32_t *all_gidx, const ConcatTables &t, uint32_t *order_out, cudaStream_t s) {
    dim3 grid((t.cap + 255) / 256, t.n_ranks);
    merge_strided_kernel<<<grid, 256, 0, s>>>(all_keys, all_gidx, t, order_out);
    return cudaGetLastError();
}

}  // namespace triage
