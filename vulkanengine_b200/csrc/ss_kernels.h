// Internal launch interface between the context (context.cu) and the kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "ss_device.cuh"

namespace ss {

// World rows live at a 64-byte stride (rows 0..2 = TransformData, row 3 = 0,0,0,1): every row is one
// aligned 64 B DRAM burst, which halves the bytes the sorted-order gather in k_emit pulls from HBM
// (a 48 B stride makes half the rows straddle two bursts).
constexpr int kWorldStride = 4;  // float4 per instance

// ---- transform / cull / key (transform_cull.cu) ----------------------------------------------
struct NodeLevelArgs {
    const float* trs;          // [n_nodes][9]
    const int32_t* parent;     // [n_nodes] or nullptr
    float4* node_world;        // [n_nodes][3]
    uint32_t lo, hi;           // node range of this level
    int trig_mode;
};
void launch_node_level(const NodeLevelArgs& a, cudaStream_t st);
// TRS records of scattered nodes: trs[node_index[k]] = records[k] for k in order (a later record wins)
void launch_trs_scatter(float* trs, const uint32_t* node_index, const float* records, uint32_t count, cudaStream_t st);
// physics poses {position xyz, quaternion xyzw} of scattered nodes -> Translation and Rotation (Euler) of their TRS records
void launch_pose_scatter(float* trs, const uint32_t* node_index, const float* poses, uint32_t count, cudaStream_t st);

struct InstanceArgs {
    const float* trs;
    const int32_t* parent;
    const float4* node_world;
    const uint32_t* node;      // explicit instance -> node map or nullptr
    uint32_t node_base;
    const uint32_t* slot;
    const uint32_t* matword;
    const float* inst_aabb;    // [n][6] or nullptr
    const SlotDev* slots;
    float4* world;             // [n][world_stride] in/out
    uint32_t world_stride;     // float4 per instance: kWorldStride (4: one aligned 64 B block per row set, what the sorted-order
                               // gather wants) or 3 (packed TransformData, for the sequential reader of the resident layout)
    uint8_t* mask;
    uint64_t* keys;
    uint32_t* vals;            // iota payload for the sort
    uint32_t n;
    uint32_t tess_shift;       // mb + sb + hb
    uint32_t idx_bits;         // > 0: write [key | mask | i] into keys (packed sort), vals unused
    uint32_t mask_bits;        // 8 when the visibility byte rides in the packed word, else 0
    int world_given;           // SS_FRAME_WORLD_ROWS_GIVEN
    int trig_mode;
    // optional fusion of the sort's digit histograms (fast path only): hist = SortControl::hist, zeroed;
    // digit p of the KEY (before packing) is (key >> hist_shift[p]) & ((1 << hist_bits[p]) - 1)
    uint32_t* hist;
    int hist_passes;           // 0 = do not accumulate
    int hist_shift[3];
    int hist_bits[3];
};
// Returns true when the launched kernel also accumulated the sort histograms requested in `a`.
bool launch_instances(const InstanceArgs& a, const CullParams& cp, cudaStream_t st);
// persistent TMA + packed-FP32 fast path, two instances per thread (transform_cull_x2.cu); same results bit for bit
bool instances_x2_eligible(const InstanceArgs& a);
void launch_instances_x2(const InstanceArgs& a, const CullParams& cp, cudaStream_t st);

// SelectTransformData records {rows 48 B, entity id} of a view's compacted list (transform_cull.cu)
void launch_select_records(const float4* rows, const uint32_t* idx, const int32_t* entity, uint32_t visible, uint32_t* out13,
                           cudaStream_t st);
// light uniform blocks from the TRS stream (transform_cull.cu); out = [496 B UBPointLights | 656 B UBSpotLights | positions]
void launch_pack_lights(const float* trs, const ss_point_light* point, uint32_t n_point, const ss_spot_light* spot,
                        uint32_t n_spot, unsigned char* out, cudaStream_t st);

// ---- onesweep radix sort (radix_sort.cu) -------------------------------------------------------
constexpr int kMaxRadixBits = 9;
constexpr int kMaxRadix = 1 << kMaxRadixBits;
constexpr int kMaxPasses = 8;
constexpr int kSortThreads = 512;
constexpr int kSortItems = 16;
constexpr int kSortTile = kSortThreads * kSortItems;

// Device-resident control block of one sort (lives at the head of the scratch buffer).
struct SortControl {
    uint32_t hist[kMaxPasses][kMaxRadix];  // global digit histograms, then exclusive bin offsets
    uint32_t ticket[kMaxPasses];           // dynamic tile ids
    uint32_t skip[kMaxPasses];             // 1 = digit constant over all keys: pass moves nothing
    uint32_t src_sel[kMaxPasses + 1];      // buffer (0 = a, 1 = b) holding the input of pass p; [n_passes] = result
    uint32_t passes_run;
    uint32_t pad[7];
};

// digit positions of a sort over bits [begin_bit, end_bit)
struct SortPlan {
    int n_passes;
    int shift[kMaxPasses];
    int bits[kMaxPasses];
};
SortPlan sort_plan(int begin_bit, int end_bit);
inline int sort_num_passes(int begin_bit, int end_bit) { return sort_plan(begin_bit, end_bit).n_passes; }

size_t sort_scratch_bytes(uint64_t n);
// Enqueues clear + histogram + scan + passes.  vals_a == nullptr: keys only.
// Result location: control->src_sel[n_passes] (device).
// prepared_with_hist: sort_prepare() ran earlier on the stream AND a producer kernel has already filled
// control->hist with the digit histograms, so the clear and the histogram pass are skipped.
cudaError_t sort_prepare(void* scratch, uint64_t n, int begin_bit, int end_bit, cudaStream_t st);
cudaError_t sort_async(uint64_t* keys_a, uint32_t* vals_a, uint64_t* keys_b, uint32_t* vals_b, uint64_t n,
                       int begin_bit, int end_bit, void* scratch, cudaStream_t st, int* launches,
                       bool prepared_with_hist);
cudaError_t sort_pairs_async(uint64_t* keys_a, uint32_t* vals_a, uint64_t* keys_b, uint32_t* vals_b, uint64_t n,
                             int begin_bit, int end_bit, void* scratch, cudaStream_t st, int* launches);

// text returned by ss_last_error(NULL); defined in context.cu
void set_global_error(const char* msg);

// ---- compaction + emit (emit.cu) ---------------------------------------------------------------
constexpr int kEmitThreads = 256;
constexpr int kEmitItems = 4;
constexpr int kEmitTile = kEmitThreads * kEmitItems;
constexpr int kEmitWords = SS_MAX_VIEWS + 1;  // per-tile look-back words: V visible counts + segment heads

struct ViewOut {
    float4* rows;        // [cap][3]
    uint64_t* keys;      // [cap]
    uint32_t* idx;       // [cap]
    ss_draw_cmd* draws;  // [max_draws]
    uint64_t* draw_keys; // [max_draws]
    uint32_t* counts;    // {visible, draws}
};

struct EmitArgs {
    const uint64_t* keys_a; const uint32_t* vals_a;
    const uint64_t* keys_b; const uint32_t* vals_b;
    const SortControl* sort_ctl;   // result selector
    uint32_t sort_passes;
    uint32_t idx_bits;             // > 0: sorted words are [key | mask | instance]
    uint32_t mask_bits;            // 8: a visibility byte sits above the instance index (layout of the word)
    uint32_t gather_mask;          // 1: read visibility from mask[] (no byte in the word, or a reused, stale one)
    uint32_t world_stride;         // float4 per entry of `world`
    uint32_t seq;                  // 1: world rows and masks lie in WORD order: entry j belongs to sorted word j, so the
                                   //    "gather" is a sequential read (draw-order-resident layout, resident.cu)
    const uint8_t* mask;
    const float4* world;
    uint32_t n;
    uint32_t n_views;
    ViewOut out[SS_MAX_VIEWS];
    uint32_t cap_visible;
    uint32_t max_draws;
    // scratch
    uint32_t* lookback;      // [tiles][kEmitWords], zeroed per frame
    uint32_t* ticket;        // zeroed per frame
    uint32_t* seg_count;     // [n_views][max_draws], zero at rest
    uint64_t* seg_key;       // [max_draws]
    uint32_t* seg_inst;      // [max_draws]
    uint32_t* totals;        // [kEmitWords] final totals (visible per view, n_segs)
    uint32_t* overflow;      // set non-zero on capacity overflow
};
size_t emit_lookback_bytes(uint64_t n);
void launch_emit(const EmitArgs& a, cudaStream_t st);
void launch_merge_place_peers(const int32_t* const* rank_lists, const int32_t* seg_rank, const int64_t* seg_out,
                              const int64_t* seg_src, const int64_t* seg_add, const int64_t* seg_key, uint32_t n_segs,
                              int32_t* out_idx, int64_t* out_keys, cudaStream_t st);
// multi-GPU merge: out[e] = lists[seg_src[s] + (e - seg_out[s])] + seg_add[s], keys[e] = seg_key[s]
void launch_merge_place(const int32_t* lists, const int64_t* seg_out, const int64_t* seg_src, const int64_t* seg_add,
                        const int64_t* seg_key, uint32_t n_segs, int32_t* out_idx, int64_t* out_keys, cudaStream_t st);

struct DrawArgs {
    uint32_t n_views;
    uint32_t max_draws;
    const uint32_t* totals;
    const uint32_t* seg_count;     // [n_views][max_draws] per-view run counts of this frame (left as they are)
    uint32_t* seg_count_next;      // the other buffer: zeroed here for the next frame's k_emit
    uint32_t* runs_cur;            // out: number of runs this frame put into seg_count
    const uint32_t* runs_next;     // runs the frame before last left in seg_count_next (bounds the zeroing)
    const uint32_t* passes_run;    // SortControl::passes_run, forwarded to the host
    uint32_t* host_stats;          // mapped pinned host words: [stats_views][2] visible/draws, overflow, passes_run
    uint32_t stats_views;
    const uint64_t* seg_key;
    const uint32_t* seg_inst;
    const uint32_t* slot;
    const SlotDev* slots;
    ViewOut out[SS_MAX_VIEWS];
    uint32_t* overflow;
};
void launch_draws(const DrawArgs& a, cudaStream_t st);

// ---- draw-order-resident layout (resident.cu) -------------------------------------------------------
// sort words [key | instance] from the static id streams (+ the sort's digit histograms)
struct WordsArgs {
    const uint32_t* slot;
    const uint32_t* matword;
    const SlotDev* slots;
    uint64_t* words;
    uint32_t n;
    uint32_t tess_shift;
    uint32_t idx_bits;
    uint32_t* hist;            // SortControl::hist, zeroed; nullptr = no histograms
    int hist_passes;
    int hist_shift[3];         // of the word
    int hist_bits[3];
};
void launch_build_words(const WordsArgs& a, cudaStream_t st);

// streams in draw order: entry j of every *_out describes instance order[j] & idx_mask
struct PermuteArgs {
    const uint64_t* order;     // sorted words
    uint64_t idx_mask;
    uint32_t n;
    int trs_only;              // 1: only the transforms changed
    const float* trs;          // node stream, already offset to the first instance node
    const float* box;          // [n][6] or nullptr
    const uint32_t* slot;
    const uint32_t* matword;
    const int32_t* parent;     // node stream, offset like trs, or nullptr
    float* trs_out;
    float* box_out;
    uint32_t* slot_out;
    uint32_t* matword_out;
    int32_t* parent_out;
    uint32_t* pos_out;         // pos[instance] = j
};
void launch_permute(const PermuteArgs& a, cudaStream_t st);

// rows / bytes of instances [first, first + count) out of word-ordered arrays (downloads, parity tests)
void launch_unpermute(const uint64_t* words, uint64_t idx_mask, uint32_t n, uint32_t first, uint32_t count,
                      const float4* world_words, uint32_t stride4, float4* rows_out, const uint8_t* mask_words,
                      uint8_t* mask_out, cudaStream_t st);

// ---- depth + id rasterisation of a view's compacted list (raster.cu) -------------------------------
struct RasterArgs {
    const float4* rows;        // [visible][3] TransformData of the view, draw order
    const uint32_t* idx;       // [visible] instance index of each row
    const ss_draw_cmd* draws;  // the view's indirect draws (index range + instance range of each run)
    const uint32_t* counts;    // {visible, draws} of the view, device side
    uint32_t max_draws;
    // scratch of the depth order the pass is drawn in (front to back): sort words [depth key : 32 | row : 32] in two
    // buffers (the sort's control block says which one holds the result), the draw of every row, and the number of
    // triangles before every rank of that order ([visible + 1])
    unsigned long long* order_a;
    unsigned long long* order_b;
    void* sort_scratch;
    int sort_passes;
    uint32_t* draw_of;
    float* draw_aabb;          // [max_draws][6] bounds of the vertices each draw's index range names (min > max: none valid)
    uint4* bound;              // [visible] per row: cells {x0 | x1 << 16, y0 | y1 << 16} its samples can lie in (x0 > x1: none),
                               // depth floor bits, instance
    uint32_t* row_of;          // [n_instances], DepthClamp passes: the sort word carries the instance, this maps it back to its row
    unsigned long long* tri_prefix;
    unsigned long long* block_sums;
    const float* positions;    // [n_vertices][3]
    const uint32_t* indices;   // [n_indices]
    uint32_t n_vertices, n_indices;
    uint32_t visible;
    float vp[16];              // column-major view-projection of the view
    uint32_t width, height;
    uint32_t flags;            // SS_RASTER_*
    unsigned long long* target;  // [height][width] words [depth bits | instance], SS_RASTER_EMPTY where untouched
    // queue of huge triangles (nullptr: none, they are walked where they are found)
    uint32_t* queue_works;     // [queue_work_cap][20]
    uint32_t* queue_tiles;     // [queue_tile_cap][2] {work, which 64 x 64 tile of its bounds} of each entry
    uint32_t* queue_counts;    // {works, tile entries, next chunk of pairs, unused}
    uint32_t queue_work_cap, queue_tile_cap;
    // coarse depth: per 8 x 8-sample cell an upper bound of the words its samples hold (all ones: unknown), refreshed
    // between the phases the pairs are drawn in; nullptr: none.  phase_lo/hi: this launch's part of the pairs, 16.16
    unsigned long long* hiz;
    uint32_t hiz_w, hiz_h;
    unsigned long long* hiz2;  // the same per 64 x 64-sample tile (8 x 8 cells): one look for a large triangle or a queue entry
    uint32_t hiz2_w, hiz2_h;
    uint32_t phase_lo, phase_hi;
};
size_t raster_hiz_words(uint32_t width, uint32_t height);
size_t raster_order_bytes(uint32_t n_instances, uint32_t max_draws);  // one allocation, carved up by raster_order_carve
void raster_order_carve(void* blob, uint32_t n_instances, uint32_t max_draws, RasterArgs* a);
int device_sm_count();  // SMs of the current device, cached per device
void launch_raster_clear(unsigned long long* target, size_t n, cudaStream_t st);
void launch_raster(const RasterArgs& a, cudaStream_t st);

}  // namespace ss
