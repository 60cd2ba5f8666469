// kernels.cuh — launch wrappers of the sm_100a kernels (kernels.cu).  Internal to the library.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/triage.h"
#include "scan_core.cuh"
#include "scan_wide.cuh"
#include "scan_pair.cuh"
#include "regex_walk.cuh"

namespace triage {

// Geometry of the scan: NT lanes per tile, each owning C contiguous bytes.  C is an odd multiple of 16 so that the
// lanes' 128-bit shared loads are bank-conflict free.  The variant is picked per dictionary by what fits in shared memory.
struct ScanGeometry { uint32_t nt, chunk, tile; };
constexpr uint32_t kMinScanTile = 16384;
constexpr uint32_t kListCap = 1024;   // message starts per tile kept in shared memory (else global offsets)
constexpr uint32_t kMaxWarm = 256;

struct DeviceDfa {
    const uint8_t *table;        // n_states * kRowBytes (padded rows)
    const EventPacked *events;   // n_events
    const uint8_t *eom;          // 256 (padded)
    uint32_t n_states, n_events, warm, start_state, neutral_state;
    int32_t head_pad, tail_pad;  // -1: no inert pad byte exists, use the byte-at-a-time boundary path
    int32_t reset_pad;           // fill byte of the padded arena layout (-1: this dictionary has none)
    // wide format (more than 256 states): class map + 32-bit table in global memory, padded arena only
    int32_t wide;
    const uint32_t *wtable;      // n_states * n_classes
    const uint8_t *cls_of;       // 256
    const uint32_t *eom_w;       // n_states (fix-up kernel)
    uint32_t n_classes;
    // filter automaton of the two-bytes-per-lookup lanes (scan_pair.cuh), padded rows like `table`
    int32_t has_filter;
    const uint8_t *ftable;
    uint32_t f_states, f_start, f_fired, f_mul, f_shift;   // ftable: Moore form, 256 one-byte entries per state; states >= f_fired were entered by a firing entry
    const uint16_t *ctable;      // column-merged copy of `table`: entry(state, ccls[byte]), n_cmp per state
    const uint8_t *ccls;         // 256
    uint32_t n_cmp;
};

struct ScanArgs {
    const uint8_t *bytes;        // 16-byte aligned, readable up to round_up(total, 16)
    const uint64_t *offsets;     // n + 1
    uint64_t total;
    uint32_t n;
    uint32_t *counters;          // n * CNT_STRIDE, zeroed
    uint32_t *tile_first;        // ntiles + 1 (scratch)
    uint32_t ntiles;
    uint32_t tile_bytes;         // = geometry.tile
    uint32_t tile_begin, tile_end;   // row-lane scan only: the tiles this launch walks (the whole arena: 0, ntiles)
    uint64_t *trace = nullptr;       // measurement aid (TRIAGE_SCAN_TRACE): per warp {start ns, end ns, tiles walked, SM id}
};

// padded-arena scan: warps are autonomous (own staging buffers, own mbarriers, no CTA barrier in the loop)
constexpr uint32_t kPadChunk = 80;
constexpr uint32_t kPadMiniTile = 32 * kPadChunk;   // bytes per warp iteration
size_t scan_padded_smem_bytes(const DeviceDfa &d, uint32_t nw);
uint32_t pick_pad_warps(const DeviceDfa &d, size_t max_smem);   // 24 for the default dictionary; fewer when the table is large
cudaError_t launch_scan_padded(const ScanArgs &a, const DeviceDfa &d, uint32_t nw, int sm_count, cudaStream_t s);

// row-lane variant of the padded scan (dense table, 16-byte warm-up): rows of 1280 / 640 / 320 bytes fed by 2-D TMA tile copies
size_t scan_rows_smem_bytes(const DeviceDfa &d, uint32_t nw);
uint32_t scan_rows_lane_bytes(uint64_t total, uint32_t nw, int sm_count, bool pair = false);   // row length for an arena of that size (1280, 640 or 320)
uint32_t scan_rows_super_bytes(uint32_t lane_bytes);
uint32_t pick_row_warps(const DeviceDfa &d, size_t max_smem);   // 0: this automaton cannot use it
bool scan_rows_usable(const ScanArgs &a);
// tile_counter: one zeroed device word (tiles beyond the first per warp are handed out through it)
cudaError_t launch_scan_rows(const ScanArgs &a, const DeviceDfa &d, uint32_t nw, uint32_t lane_bytes, uint32_t *tile_counter, int sm_count,
                             cudaStream_t s);

// pair-lane variant of the row scan (filter automaton over class pairs + SIMD word count + exact walk of flagged groups)
size_t scan_pair_smem_bytes(const DeviceDfa &d, uint32_t nw);
// wide format, rows streamed from global memory, `chains` rows per lane (tile = 32 * chains * row_bytes bytes)
cudaError_t launch_scan_stream_wide(const ScanArgs &a, const DeviceDfa &d, uint32_t nw, uint32_t chains, uint32_t row_bytes, uint32_t *tile_counter,
                                    int sm_count, cudaStream_t s);
uint32_t pick_pair_warps(const DeviceDfa &d, size_t max_smem);   // 0: this dictionary / automaton cannot use it
cudaError_t launch_scan_pair(const ScanArgs &a, const DeviceDfa &d, uint32_t nw, uint32_t lane_bytes, uint32_t *tile_counter, int sm_count,
                             cudaStream_t s);

// sequential exact recount of the messages the scan flagged (self-overlapping keywords); `exact` points at global memory
cudaError_t launch_fixup(const uint8_t *bytes, const uint64_t *offsets, const uint32_t *lengths, uint32_t n, const DeviceDfa &exact,
                         uint32_t *counters, cudaStream_t s);

// regex patterns of the dictionary (regex_vm.h): adds every pattern's FindAllString count to its priority's score counter
cudaError_t launch_regex_count(const RegexView &v, const uint8_t *bytes, const uint64_t *offsets, const uint32_t *lengths, uint32_t n,
                               uint32_t *counters, int sm_count, cudaStream_t s);

struct FinalizeArgs {
    const uint32_t *lengths;     // padded layout: true content lengths (else nullptr: len = offsets[i+1]-offsets[i])
    const uint8_t *bytes;        // padded layout: for the HasSuffix "?" test
    const uint32_t *counters;
    const triage_meta *meta;
    const uint64_t *offsets;
    const uint64_t *ts;          // nullable: arrival order is the time order
    uint32_t n;
    int32_t default_prio;
    int32_t rules_on;
    int32_t single_queue;        // 1: whole batch goes to one queue, bucket = min(int(Priority), 15)
    triage_result *results;
    uint8_t *bucket;             // per message: (level-1)*4 + (final-1), or 0xFF when the level queue does not exist
    uint32_t *flags;             // [0] |= 1 when ts is not non-decreasing in arrival order
};

// largest-throughput geometry whose shared memory fits; nt == 0 if none does
ScanGeometry pick_scan_geometry(const DeviceDfa &d, size_t max_smem);
size_t scan_smem_bytes(const DeviceDfa &d, const ScanGeometry &g);
cudaError_t launch_tile_index(const ScanArgs &a, cudaStream_t s);
// tile index + zeroing of a.counters (n * CNT_STRIDE words) and of `small` in one launch
cudaError_t launch_prepare(const ScanArgs &a, uint32_t *small, uint32_t small_words, int sm_count, cudaStream_t s);
cudaError_t launch_scan(const ScanArgs &a, const DeviceDfa &d, const ScanGeometry &g, int sm_count, cudaStream_t s);
cudaError_t launch_finalize(const FinalizeArgs &a, cudaStream_t s);

// ---- ordering --------------------------------------------------------------------------------
// Stable partition of items by an 8-bit digit.  perm_in == nullptr means the identity (arrival
// order).  Items whose digit is 0xFF when `drop_ff` are left out.
struct OrderScratch {
    uint32_t *blockhist;   // nblocks * 256
    uint32_t *totals;      // 256 + 1
    uint32_t nblocks_cap;
};
constexpr int kOrderBlock = 1024;

// level ranks -> admission (capacity), writes accepted/admit into results and turns the bucket of a
// rejected message into 0xFF.  pending[4] = current queue lengths, max_size <= 0: unbounded.
cudaError_t launch_admit(const uint8_t *bucket_in, uint8_t *bucket_out, triage_result *results, uint32_t n,
                         const uint64_t *pending_dev, const int64_t *max_size_dev, int32_t single_queue, OrderScratch sc,
                         cudaStream_t s);
// digits: u8 per ITEM INDEX (not per position); out_perm[k] = item index at rank k.
cudaError_t launch_partition_u8(const uint8_t *digits, const uint32_t *perm_in, uint32_t n, uint32_t n_buckets,
                                bool drop_ff, uint32_t *perm_out, uint32_t *bucket_counts_dev /* 256 */,
                                OrderScratch sc, cudaStream_t s);
// one LSD pass over 64-bit keys: digit = (key[item] >> shift) & 0xFF
cudaError_t launch_partition_key(const uint64_t *keys, uint32_t shift, const uint32_t *perm_in, uint32_t n,
                                 uint32_t *perm_out, OrderScratch sc, cudaStream_t s);
// per-digit-position histograms of keys[perm] over all 8 byte positions (decides which passes to run)
cudaError_t launch_key_digit_hist(const uint64_t *keys, const uint32_t *perm, uint32_t n, uint32_t *hist /* 8*256, zeroed */,
                                  cudaStream_t s);
// run (time key, gidx) for the multi-GPU merge, in dequeue order (bucket-major)
cudaError_t launch_build_run(const uint32_t *order, uint32_t n_adm, const uint8_t *bucket, const uint64_t *ts,
                             uint64_t ts_base_if_null, uint64_t gidx_base, uint64_t *run_keys, uint32_t *run_gidx,
                             cudaStream_t s);
// ---- multi-GPU merge of the all-gathered runs (rank q's run at [q*cap, q*cap + cnt[q])) ----
struct ConcatTables {
    uint32_t cap;                 // stride between ranks in the gathered buffers
    uint32_t n_ranks;
    uint32_t cnt[8];              // admitted per rank
    uint32_t base[8];             // first global index of every rank
    uint32_t dst[8][16];          // destination of the first element of (rank, bucket)
    uint32_t lstart[8][16];       // start of bucket inside the rank's run
};
// one-exchange merge (contiguous sequence shards): [header | order] per rank at stride kMergeHeaderWords + n
constexpr uint32_t kMergeHeaderWords = 32;
struct MergeBases { uint32_t base[8]; uint32_t n_ranks; uint32_t stride; };
cudaError_t launch_pack_run(const uint32_t *order, const uint32_t *counts16_dev, const uint32_t *n_adm_dev, uint32_t n, uint32_t *gsrc,
                            cudaStream_t s);
cudaError_t launch_concat_gathered(const uint32_t *all, const MergeBases &mb, uint32_t n_per_rank, uint32_t *order_out, uint64_t capacity,
                                   uint64_t *n_total_dev, cudaStream_t s);
cudaError_t launch_header(const uint32_t *order, uint32_t n_adm, const uint32_t *counts16_host, const uint64_t *ts, uint64_t ts_base,
                          uint64_t *header_dev, cudaStream_t s);
cudaError_t launch_concat(const uint32_t *all_order, const ConcatTables &t, uint32_t *order_out, cudaStream_t s);
cudaError_t launch_merge_strided(const uint64_t *all_keys, const uint32_t *all_gidx, const ConcatTables &t, uint32_t *order_out, cudaStream_t s);

}  // namespace triage
