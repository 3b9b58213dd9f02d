// libscenesubmit C ABI: context, device streams, frame orchestration.  See include/scenesubmit.h for the
// contract and the reference code each entry point replaces.
#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <unordered_map>
#include <vector>

#include "ss_kernels.h"

using namespace ss;

namespace {

thread_local std::string g_create_error;

struct Slot {
    ViewOut out[SS_MAX_VIEWS];
    uint32_t* counts_dev = nullptr;   // [max_views][2]
    cudaEvent_t done = nullptr, start = nullptr;
    cudaEvent_t stage[4] = {nullptr, nullptr, nullptr, nullptr};  // after hierarchy, instances, sort, emit
    bool pending = false;
    uint32_t n_views = 0, n_instances = 0;
    int launches = 0;
    ss_view views[SS_MAX_VIEWS];      // as given to ss_frame (ss_raster_view projects with them)
    int pipeline = 0;                 // ss_pipeline that produced the slot
    const uint64_t* words_known = nullptr;  // resident layout: the sorted words
    uint32_t idx_bits = 0;

    // pinned host mirror
    uint32_t* h_counts = nullptr;     // [max_views][2] + overflow + passes_run: mapped pinned, written by k_draws
    uint32_t* h_counts_dev = nullptr; // the same words as the device sees them
};

}  // namespace

namespace ss {
void set_global_error(const char* msg) { g_create_error = msg ? msg : ""; }
}  // namespace ss

struct ss_ctx {
    ss_config cfg{};
    int device = 0;
    cudaStream_t stream = nullptr;
    std::string err;
    std::vector<void*> allocs;

    // static streams
    float* d_trs = nullptr;            // = d_trs_buf[trs_cur]
    // Whole-stream TRS uploads go to the other buffer on their own stream, so that the copy of frame f+1's
    // transforms overlaps frame f's kernels; ss_frame adopts the new buffer.  Partial (dirty-range) uploads
    // write into the current buffer on the frame stream.
    float* d_trs_buf[2] = {nullptr, nullptr};
    int trs_cur = 0;
    bool trs_swap_pending = false;
    cudaStream_t upload_stream = nullptr;
    cudaStream_t download_stream = nullptr;  // reads of a fenced slot must not queue behind later frames
    cudaEvent_t ev_upload_done = nullptr;
    cudaEvent_t ev_trs_read[2] = {nullptr, nullptr};  // last frame that read each buffer has passed its transform stage
    int32_t* d_parent = nullptr;
    float4* d_node_world = nullptr;
    uint32_t* d_node = nullptr;
    uint32_t* d_slot = nullptr;
    uint32_t* d_matword = nullptr;
    float* d_inst_aabb = nullptr;
    SlotDev* d_slots = nullptr;
    // per-frame products
    float4* d_world = nullptr;
    uint8_t* d_mask = nullptr;
    uint64_t *d_keys_a = nullptr, *d_keys_b = nullptr;
    uint32_t *d_vals_a = nullptr, *d_vals_b = nullptr;
    void* d_sort_scratch = nullptr;
    // emit scratch: [lookback | ticket | totals] is zeroed every frame with one memset
    unsigned char* d_emit_zero = nullptr;
    size_t emit_zero_bytes = 0;
    uint32_t* d_lookback = nullptr;
    uint32_t* d_ticket = nullptr;
    uint32_t* d_totals = nullptr;
    uint32_t* d_overflow = nullptr;
    uint32_t* d_seg_count = nullptr;   // [2][max_views][max_draws]: per-view run counters, double-buffered by frame
    uint32_t* d_seg_runs = nullptr;    // [2] runs left in each buffer
    // staging of the sparse uploads: [count] indices then [count][R] records; grows on demand.  One pinned host buffer
    // (pageable callers only) and two device buffers, so that the copy of one call (upload stream) runs under the frame
    // issued before it while the scatter of the call before still reads the other buffer.
    unsigned char* h_sparse = nullptr;
    unsigned char* d_sparse[2] = {nullptr, nullptr};
    size_t sparse_cap = 0;
    int sparse_cur = 0;
    cudaEvent_t ev_sparse = nullptr;             // the host staging buffer's last copy has left the host
    cudaEvent_t ev_sparse_copied = nullptr;      // upload stream: this call's records are in d_sparse[cur]
    cudaEvent_t ev_sparse_scattered[2] = {nullptr, nullptr};  // frame stream: the scatter that read d_sparse[b] is done
    int32_t* d_entity = nullptr;       // ss_set_instance_entities (allocated on first use)
    uint32_t* d_select = nullptr;      // ss_select_view: SelectTransformData records of one view (allocated on first use)
    unsigned char* d_lights = nullptr; // ss_pack_lights: light descriptions in, uniform blocks out
    // geometry + raster target of the consumer pass (ss_set_geometry / ss_raster_view); allocated on first use
    float* d_positions = nullptr;
    uint32_t* d_indices = nullptr;
    uint32_t n_vertices = 0, n_indices = 0;
    unsigned long long* d_raster[SS_MAX_VIEWS] = {};  // one target per view: [raster_h[v]][raster_w[v]]
    void* d_raster_order = nullptr;                // depth-order scratch of the raster pass (raster_order_bytes(n_instances))
    uint32_t* d_raster_queue = nullptr;            // works | tile entries | 2 counters (kRasterQueueWorks / kRasterQueueTiles)
    unsigned long long* d_raster_hiz = nullptr;              // coarse depth of the pass being drawn (raster_hiz_words of the largest target)
    size_t raster_hiz_cap = 0;
    uint32_t raster_w[SS_MAX_VIEWS] = {}, raster_h[SS_MAX_VIEWS] = {};
    uint32_t seg_parity = 0;
    uint64_t* d_seg_key = nullptr;
    uint32_t* d_seg_inst = nullptr;

    // draw-order-resident layout (resident.cu): the per-instance streams once more, in (key, instance) order
    float* d_trs_o = nullptr;
    float* d_box_o = nullptr;
    uint32_t* d_slot_o = nullptr;
    uint32_t* d_matword_o = nullptr;
    int32_t* d_parent_o = nullptr;
    uint32_t* d_pos = nullptr;             // pos[instance] = sorted position
    bool res_static_dirty = true;          // box / slot / material / parent / order of the permuted streams is stale
    bool res_trs_dirty = true;             // transforms changed since the permuted TRS stream was built
    bool resident_valid = false;           // the sort buffers hold the sorted words of the current static streams
    const uint64_t* d_words_resident = nullptr;
    uint32_t words_idx_bits = 0;
    int last_pipeline = 0;                 // ss_pipeline of the most recent frame
    bool nodes_valid = false;              // d_node_world holds the parent levels of the current transforms

    std::vector<Slot> slots;
    uint64_t frame_no = 0;

    // host state
    uint32_t n_nodes = 0, n_instances = 0, n_slots = 0, n_levels = 0;
    std::vector<uint32_t> level_offsets;
    bool has_parent = false, has_node_map = false, has_inst_aabb = false, world_uploaded = false;
    bool world_valid = false;  // d_world holds the rows of a finished transform stage (SS_FRAME_REUSE_WORLD)
    bool sort_valid = false;   // the sort buffers hold the order of the current static streams (SS_FRAME_REUSE_SORT)
    uint32_t node_base = 0;
    uint32_t max_material_rank = 0;  // over the instance stream (ss_set_key_layout re-checks it)
    uint32_t mesh_bits = 22, submesh_bits = 16, material_bits = 24;
    std::vector<ss_submesh> host_table;
    uint32_t max_draws = 0;
    uint32_t cap_visible = 0;

    ss_status fail(ss_status st, const char* fmt, ...)
    {
        char buf[512];
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(buf, sizeof buf, fmt, ap);
        va_end(ap);
        err = buf;
        return st;
    }
};

#define SS_CUDA(ctx, call)                                                                              \
    do {                                                                                                \
        cudaError_t e__ = (call);                                                                       \
        if (e__ != cudaSuccess)                                                                         \
            return (ctx)->fail(SS_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__),    \
                               __FILE__, __LINE__);                                                     \
    } while (0)

namespace {

template <class T>
cudaError_t dev_alloc(ss_ctx* c, T** p, size_t count)
{
    void* q = nullptr;
    cudaError_t e = cudaMalloc(&q, std::max<size_t>(count * sizeof(T), 256));
    if (e == cudaSuccess) {
        c->allocs.push_back(q);
        *p = static_cast<T*>(q);
    }
    return e;
}

ss_status upload_table(ss_ctx* c)
{
    std::vector<SlotDev> dev(c->n_slots);
    for (uint32_t i = 0; i < c->n_slots; ++i) {
        const ss_submesh& s = c->host_table[i];
        SlotDev& d = dev[i];
        const float* m = s.local;  // m[col*4 + row]
        d.r0 = make_float4(m[0], m[4], m[8], m[12]);
        d.r1 = make_float4(m[1], m[5], m[9], m[13]);
        d.r2 = make_float4(m[2], m[6], m[10], m[14]);
        for (int k = 0; k < 3; ++k) {
            d.mn[k] = s.aabb_min[k];
            d.mx[k] = s.aabb_max[k];
        }
        d.index_count = s.index_count;
        d.first_index = s.first_index;
        d.vertex_offset = s.vertex_offset;
        d.pad = 0;
        d.key_hi = ((uint64_t)s.mesh_rank << (c->material_bits + c->submesh_bits)) |
                   ((uint64_t)s.submesh_index << c->material_bits);
    }
    if (c->n_slots)
        SS_CUDA(c, cudaMemcpyAsync(c->d_slots, dev.data(), dev.size() * sizeof(SlotDev), cudaMemcpyHostToDevice,
                                   c->stream));
    SS_CUDA(c, cudaStreamSynchronize(c->stream));
    return SS_OK;
}

// N3: Gribb-Hartmann planes of a zero-to-one depth clip matrix, L R B T N F; plain adds, same as the oracle.
void planes_from_view_proj(const float* vp, ViewPlanes& out)
{
    float pl[6][4];
    for (int col = 0; col < 4; ++col) {
        const float r0 = vp[col * 4 + 0], r1 = vp[col * 4 + 1], r2 = vp[col * 4 + 2], r3 = vp[col * 4 + 3];
        pl[0][col] = r3 + r0;
        pl[1][col] = r3 - r0;
        pl[2][col] = r3 + r1;
        pl[3][col] = r3 - r1;
        pl[4][col] = r2;
        pl[5][col] = r3 - r2;
    }
    for (int p = 0; p < 6; ++p) out.pl[p] = make_float4(pl[p][0], pl[p][1], pl[p][2], pl[p][3]);
    out.flags = 0;
    for (int ax = 0; ax < 3; ++ax) {  // exactly opposite normals: the kernel shares the dot products
        const float* a = pl[2 * ax];
        const float* b = pl[2 * ax + 1];
        // compare as values: -0.0f == 0.0f is fine here, the negation of either zero gives an equal product
        if (b[0] == -a[0] && b[1] == -a[1] && b[2] == -a[2]) out.flags |= 1u << ax;
    }
}

// Implicit node map (instance i sits on node node_base + i) under a parent chain: ss_frame composes the levels ABOVE
// node_base with k_node_level and the instance kernel composes the instances' own level from world[parent], so every
// instance node must belong to the last level.  Without parents every node is a root and any range is fine.
// a static stream changed: kept world rows, sorted orders and the draw-order streams are all stale
void invalidate_static(ss_ctx* c)
{
    c->world_valid = c->sort_valid = c->nodes_valid = false;
    c->resident_valid = false;
    c->res_static_dirty = true;
}

bool implicit_nodes_in_last_level(const ss_ctx* c, uint32_t node_base, uint32_t n)
{
    if (!c->has_parent || c->n_levels == 0) return true;
    return node_base >= c->level_offsets[c->n_levels - 1] && (uint64_t)node_base + n <= c->level_offsets[c->n_levels];
}

}  // namespace

extern "C" {

SS_API uint32_t ss_version(void) { return (SS_VERSION_MAJOR << 16) | SS_VERSION_MINOR; }

SS_API const char* ss_last_error(const ss_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

SS_API void ss_destroy(ss_ctx* c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->upload_stream) cudaStreamSynchronize(c->upload_stream);
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->ev_upload_done) cudaEventDestroy(c->ev_upload_done);
    for (cudaEvent_t e : c->ev_trs_read)
        if (e) cudaEventDestroy(e);
    if (c->upload_stream) cudaStreamDestroy(c->upload_stream);
    if (c->download_stream) cudaStreamDestroy(c->download_stream);
    for (Slot& s : c->slots) {
        if (s.done) cudaEventDestroy(s.done);
        if (s.start) cudaEventDestroy(s.start);
        for (cudaEvent_t e : s.stage)
            if (e) cudaEventDestroy(e);
        if (s.h_counts) cudaFreeHost(s.h_counts);
    }
    for (void* p : c->allocs) cudaFree(p);
    if (c->h_sparse) cudaFreeHost(c->h_sparse);
    for (int b = 0; b < 2; ++b) {
        if (c->d_sparse[b]) cudaFree(c->d_sparse[b]);
        if (c->ev_sparse_scattered[b]) cudaEventDestroy(c->ev_sparse_scattered[b]);
    }
    if (c->ev_sparse) cudaEventDestroy(c->ev_sparse);
    if (c->ev_sparse_copied) cudaEventDestroy(c->ev_sparse_copied);
    if (c->d_lights) cudaFree(c->d_lights);
    if (c->d_entity) cudaFree(c->d_entity);
    if (c->d_select) cudaFree(c->d_select);
    if (c->d_positions) cudaFree(c->d_positions);
    if (c->d_indices) cudaFree(c->d_indices);
    for (unsigned long long* p : c->d_raster)
        if (p) cudaFree(p);
    if (c->d_raster_order) cudaFree(c->d_raster_order);
    if (c->d_raster_queue) cudaFree(c->d_raster_queue);
    if (c->d_raster_hiz) cudaFree(c->d_raster_hiz);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

SS_API ss_status ss_create(const ss_config* cfg, ss_ctx** out)
{
    if (!cfg || !out) {
        g_create_error = "ss_create: null argument";
        return SS_ERR_INVALID;
    }
    *out = nullptr;
    if (cfg->struct_size != sizeof(ss_config) || cfg->max_views < 1 || cfg->max_views > SS_MAX_VIEWS ||
        cfg->max_instances == 0 || cfg->max_nodes == 0 || cfg->max_submesh_slots == 0 ||
        cfg->max_instances >= (1u << 30) || cfg->trig_mode > SS_TRIG_LIBM ||
        (cfg->flags & ~(uint32_t)SS_CONFIG_DRAW_ORDER_RESIDENT)) {
        g_create_error = "ss_create: invalid ss_config (struct_size, capacities 1..2^30-1, max_views 1..8, trig_mode, flags)";
        return SS_ERR_INVALID;
    }
    int n_dev = 0;
    cudaError_t e = cudaGetDeviceCount(&n_dev);
    if (e != cudaSuccess || n_dev == 0 || cfg->device < 0 || cfg->device >= n_dev) {
        // No CPU fallback exists: without a CUDA device the library refuses to run.
        g_create_error = std::string("ss_create: no usable CUDA device (") +
                         (e != cudaSuccess ? cudaGetErrorString(e) : "bad device ordinal") + ")";
        return SS_ERR_CUDA;
    }
    ss_ctx* c = new (std::nothrow) ss_ctx();
    if (!c) {
        g_create_error = "ss_create: out of host memory";
        return SS_ERR_NOMEM;
    }
    c->cfg = *cfg;
    c->device = cfg->device;
    const uint32_t fif = cfg->frames_in_flight ? cfg->frames_in_flight : 3;
    c->max_draws = cfg->max_draws_per_view ? cfg->max_draws_per_view : std::min<uint32_t>(cfg->max_instances, 1u << 20);
    c->cap_visible = cfg->max_visible_per_view
                         ? (uint32_t)std::min<uint64_t>(cfg->max_visible_per_view, cfg->max_instances)
                         : cfg->max_instances;
    const size_t N = cfg->max_instances, NN = cfg->max_nodes, V = cfg->max_views;

    auto bail = [&](cudaError_t err, const char* what) {
        g_create_error = std::string("ss_create: ") + what + ": " + cudaGetErrorString(err);
        ss_destroy(c);
        return err == cudaErrorMemoryAllocation ? SS_ERR_NOMEM : SS_ERR_CUDA;
    };
#define TRY(call)                                  \
    do {                                           \
        cudaError_t e2 = (call);                   \
        if (e2 != cudaSuccess) return bail(e2, #call); \
    } while (0)

    TRY(cudaSetDevice(c->device));
    TRY(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    TRY(cudaStreamCreateWithFlags(&c->upload_stream, cudaStreamNonBlocking));
    TRY(cudaStreamCreateWithFlags(&c->download_stream, cudaStreamNonBlocking));
    TRY(cudaEventCreateWithFlags(&c->ev_upload_done, cudaEventDisableTiming));
    for (cudaEvent_t& e : c->ev_trs_read) TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    if (const char* g = getenv("SS_L2_FETCH_GRANULARITY")) {  // experiment knob: 32 / 64 / 128 bytes
        const int v = atoi(g);
        if (v == 32 || v == 64 || v == 128) cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)v);
    }
    TRY(dev_alloc(c, &c->d_trs_buf[0], NN * 9));
    TRY(dev_alloc(c, &c->d_trs_buf[1], NN * 9));
    c->d_trs = c->d_trs_buf[0];
    TRY(dev_alloc(c, &c->d_parent, NN));
    TRY(dev_alloc(c, &c->d_node_world, NN * 3));
    TRY(dev_alloc(c, &c->d_node, N));
    TRY(dev_alloc(c, &c->d_slot, N));
    TRY(dev_alloc(c, &c->d_matword, N));
    TRY(dev_alloc(c, &c->d_inst_aabb, N * 6));
    TRY(dev_alloc(c, &c->d_slots, (size_t)cfg->max_submesh_slots));
    TRY(dev_alloc(c, &c->d_world, N * kWorldStride));
    TRY(dev_alloc(c, &c->d_mask, N));
    TRY(dev_alloc(c, &c->d_keys_a, N));
    TRY(dev_alloc(c, &c->d_keys_b, N));
    if (cfg->flags & SS_CONFIG_DRAW_ORDER_RESIDENT) {
        TRY(dev_alloc(c, &c->d_trs_o, N * 9));
        TRY(dev_alloc(c, &c->d_box_o, N * 6));
        TRY(dev_alloc(c, &c->d_slot_o, N));
        TRY(dev_alloc(c, &c->d_matword_o, N));
        TRY(dev_alloc(c, &c->d_parent_o, N));
        TRY(dev_alloc(c, &c->d_pos, N));
    }
    TRY(dev_alloc(c, &c->d_vals_a, N));
    TRY(dev_alloc(c, &c->d_vals_b, N));
    TRY(dev_alloc(c, reinterpret_cast<unsigned char**>(&c->d_sort_scratch), sort_scratch_bytes(N)));

    const size_t lb_bytes = (emit_lookback_bytes(N) + 255) & ~size_t(255);
    c->emit_zero_bytes = lb_bytes + 256;
    TRY(dev_alloc(c, &c->d_emit_zero, c->emit_zero_bytes));
    c->d_lookback = reinterpret_cast<uint32_t*>(c->d_emit_zero);
    c->d_ticket = reinterpret_cast<uint32_t*>(c->d_emit_zero + lb_bytes);
    c->d_totals = c->d_ticket + 16;
    c->d_overflow = c->d_ticket + 32;
    TRY(dev_alloc(c, &c->d_seg_count, 2 * V * c->max_draws));
    TRY(cudaMemsetAsync(c->d_seg_count, 0, 2 * V * c->max_draws * sizeof(uint32_t), c->stream));
    TRY(dev_alloc(c, &c->d_seg_runs, (size_t)8));
    TRY(cudaMemsetAsync(c->d_seg_runs, 0, 8 * sizeof(uint32_t), c->stream));
    TRY(dev_alloc(c, &c->d_seg_key, (size_t)c->max_draws));
    TRY(dev_alloc(c, &c->d_seg_inst, (size_t)c->max_draws));

    c->slots.resize(fif);
    for (Slot& s : c->slots) {
        TRY(dev_alloc(c, &s.counts_dev, V * 2 + 8));
        TRY(cudaMemsetAsync(s.counts_dev, 0, (V * 2 + 8) * sizeof(uint32_t), c->stream));
        for (size_t v = 0; v < V; ++v) {
            TRY(dev_alloc(c, &s.out[v].rows, (size_t)c->cap_visible * 3));
            TRY(dev_alloc(c, &s.out[v].keys, (size_t)c->cap_visible));
            TRY(dev_alloc(c, &s.out[v].idx, (size_t)c->cap_visible));
            TRY(dev_alloc(c, &s.out[v].draws, (size_t)c->max_draws));
            TRY(dev_alloc(c, &s.out[v].draw_keys, (size_t)c->max_draws));
            s.out[v].counts = s.counts_dev + v * 2;
        }
        TRY(cudaEventCreate(&s.done));
        TRY(cudaEventCreate(&s.start));
        for (cudaEvent_t& e : s.stage) TRY(cudaEventCreate(&e));
        TRY(cudaHostAlloc(reinterpret_cast<void**>(&s.h_counts), (V * 2 + 8) * sizeof(uint32_t), cudaHostAllocMapped));
        memset(s.h_counts, 0, (V * 2 + 8) * sizeof(uint32_t));
        TRY(cudaHostGetDevicePointer(reinterpret_cast<void**>(&s.h_counts_dev), s.h_counts, 0));
    }
    TRY(cudaStreamSynchronize(c->stream));
#undef TRY
    // defaults: one level of roots, no parents
    c->level_offsets = {0, 0};
    *out = c;
    return SS_OK;
}

SS_API ss_status ss_set_key_layout(ss_ctx* c, uint32_t mesh_bits, uint32_t submesh_bits, uint32_t material_bits)
{
    if (!c) return SS_ERR_INVALID;
    if (mesh_bits == 0 || submesh_bits == 0 || material_bits == 0 || material_bits > 31 ||
        mesh_bits + submesh_bits + material_bits > 62)
        return c->fail(SS_ERR_INVALID, "ss_set_key_layout: widths must be >= 1, material <= 31, sum <= 62");
    // streams set under the old layout must still fit the new one, or key fields would overlap
    for (uint32_t i = 0; i < c->n_slots; ++i)
        if (c->host_table[i].mesh_rank >> mesh_bits || c->host_table[i].submesh_index >> submesh_bits)
            return c->fail(SS_ERR_INVALID, "ss_set_key_layout: submesh table row %u does not fit %u/%u bits", i, mesh_bits,
                           submesh_bits);
    if (c->n_instances && (c->max_material_rank >> material_bits))
        return c->fail(SS_ERR_INVALID, "ss_set_key_layout: material rank %u of the instance stream does not fit %u bits",
                       c->max_material_rank, material_bits);
    SS_CUDA(c, cudaSetDevice(c->device));
    invalidate_static(c);
    c->mesh_bits = mesh_bits;
    c->submesh_bits = submesh_bits;
    c->material_bits = material_bits;
    return c->n_slots ? upload_table(c) : SS_OK;
}

SS_API ss_status ss_set_submesh_table(ss_ctx* c, const ss_submesh* table, uint32_t count)
{
    if (!c) return SS_ERR_INVALID;
    if (!table || count == 0 || count > c->cfg.max_submesh_slots)
        return c->fail(SS_ERR_INVALID, "ss_set_submesh_table: %u rows, capacity %u", count, c->cfg.max_submesh_slots);
    for (uint32_t i = 0; i < count; ++i) {
        const float* m = table[i].local;
        if (m[3] != 0.f || m[7] != 0.f || m[11] != 0.f || m[15] != 1.f)
            return c->fail(SS_ERR_INVALID, "ss_set_submesh_table: row %u LocalTransform is not affine", i);
        if (table[i].mesh_rank >> c->mesh_bits || table[i].submesh_index >> c->submesh_bits)
            return c->fail(SS_ERR_INVALID, "ss_set_submesh_table: row %u rank/index exceeds the key layout", i);
    }
    SS_CUDA(c, cudaSetDevice(c->device));
    invalidate_static(c);
    c->host_table.assign(table, table + count);
    c->n_slots = count;
    return upload_table(c);
}

SS_API ss_status ss_set_nodes(ss_ctx* c, uint32_t n_nodes, const int32_t* parent, const uint32_t* level_offsets,
                              uint32_t n_levels)
{
    if (!c) return SS_ERR_INVALID;
    if (n_nodes == 0 || n_nodes > c->cfg.max_nodes)
        return c->fail(SS_ERR_INVALID, "ss_set_nodes: %u nodes, capacity %u", n_nodes, c->cfg.max_nodes);
    if (parent) {
        // nodes past the last level are "external": never composed from TRS, their world rows are given
        // through ss_upload_node_world_rows (ad-hoc submitters such as SubmitPhysicsMesh)
        if (!level_offsets || n_levels == 0 || n_levels > SS_MAX_LEVELS || level_offsets[0] != 0 ||
            level_offsets[n_levels] > n_nodes)
            return c->fail(SS_ERR_INVALID, "ss_set_nodes: level_offsets must start at 0 and end <= n_nodes in 1..%u levels",
                           SS_MAX_LEVELS);
        for (uint32_t l = 0; l < n_levels; ++l) {
            if (level_offsets[l] > level_offsets[l + 1])
                return c->fail(SS_ERR_INVALID, "ss_set_nodes: level_offsets not monotone at level %u", l);
            for (uint32_t i = level_offsets[l]; i < level_offsets[l + 1]; ++i) {
                const int32_t p = parent[i];
                const bool ok = l == 0 ? p < 0 : (p >= (int32_t)level_offsets[l - 1] && p < (int32_t)level_offsets[l]);
                if (!ok)
                    return c->fail(SS_ERR_INVALID, "ss_set_nodes: node %u (level %u) has parent %d outside level %d", i,
                                   l, p, (int)l - 1);
            }
        }
        c->level_offsets.assign(level_offsets, level_offsets + n_levels + 1);
        c->n_levels = n_levels;
    } else {
        c->level_offsets = {0, n_nodes};
        c->n_levels = 1;
    }
    SS_CUDA(c, cudaSetDevice(c->device));
    invalidate_static(c);
    c->n_nodes = n_nodes;
    c->has_parent = parent != nullptr;
    if (parent) {
        SS_CUDA(c, cudaMemcpyAsync(c->d_parent, parent, (size_t)n_nodes * 4, cudaMemcpyHostToDevice, c->stream));
        SS_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return SS_OK;
}

SS_API ss_status ss_set_instances(ss_ctx* c, uint32_t n, const uint32_t* node, uint32_t node_base,
                                  const uint32_t* submesh_slot, const uint32_t* material_word)
{
    if (!c) return SS_ERR_INVALID;
    if (!submesh_slot || !material_word || n > c->cfg.max_instances)
        return c->fail(SS_ERR_INVALID, "ss_set_instances: %u instances, capacity %u", n, c->cfg.max_instances);
    if (c->n_nodes == 0) return c->fail(SS_ERR_STATE, "ss_set_instances: call ss_set_nodes first");
    if (!node && (uint64_t)node_base + n > c->n_nodes)
        return c->fail(SS_ERR_INVALID, "ss_set_instances: node_base %u + %u instances exceeds %u nodes", node_base, n,
                       c->n_nodes);
    if (!node && n && !implicit_nodes_in_last_level(c, node_base, n))
        return c->fail(SS_ERR_INVALID,
                       "ss_set_instances: with node == NULL the nodes [%u, %u) must all lie in the last level [%u, %u) "
                       "(only the levels above it are composed before the instance kernel); pass an explicit node map",
                       node_base, node_base + n, c->level_offsets[c->n_levels - 1], c->level_offsets[c->n_levels]);
    uint32_t max_rank = 0;
    for (uint32_t i = 0; i < n; ++i) {
        max_rank = std::max(max_rank, material_word[i] & ~SS_MATERIAL_TESS_BIT);
        if (submesh_slot[i] != SS_SLOT_NULL && submesh_slot[i] >= c->n_slots)
            return c->fail(SS_ERR_INVALID, "ss_set_instances: instance %u names submesh slot %u of %u", i,
                           submesh_slot[i], c->n_slots);
        if ((material_word[i] & ~SS_MATERIAL_TESS_BIT) >> c->material_bits)
            return c->fail(SS_ERR_INVALID, "ss_set_instances: instance %u material rank exceeds the key layout", i);
        if (node && node[i] >= c->n_nodes)
            return c->fail(SS_ERR_INVALID, "ss_set_instances: instance %u names node %u of %u", i, node[i], c->n_nodes);
    }
    SS_CUDA(c, cudaSetDevice(c->device));
    c->n_instances = n;
    c->max_material_rank = max_rank;
    c->has_node_map = node != nullptr;
    c->node_base = node ? 0 : node_base;
    c->has_inst_aabb = false;
    c->world_uploaded = false;
    invalidate_static(c);
    if (n) {
        if (node) SS_CUDA(c, cudaMemcpyAsync(c->d_node, node, (size_t)n * 4, cudaMemcpyHostToDevice, c->stream));
        SS_CUDA(c, cudaMemcpyAsync(c->d_slot, submesh_slot, (size_t)n * 4, cudaMemcpyHostToDevice, c->stream));
        SS_CUDA(c, cudaMemcpyAsync(c->d_matword, material_word, (size_t)n * 4, cudaMemcpyHostToDevice, c->stream));
        SS_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return SS_OK;
}

SS_API ss_status ss_set_instance_aabbs(ss_ctx* c, uint32_t first, uint32_t count, const float* minmax6)
{
    if (!c) return SS_ERR_INVALID;
    c->res_static_dirty = true;  // the draw-order streams carry the boxes
    if (!minmax6) {
        c->has_inst_aabb = false;
        return SS_OK;
    }
    if ((uint64_t)first + count > c->n_instances)
        return c->fail(SS_ERR_INVALID, "ss_set_instance_aabbs: range [%u, +%u) exceeds %u instances", first, count,
                       c->n_instances);
    SS_CUDA(c, cudaSetDevice(c->device));
    if (count) {
        SS_CUDA(c, cudaMemcpyAsync(c->d_inst_aabb + (size_t)first * 6, minmax6, (size_t)count * 24,
                                   cudaMemcpyHostToDevice, c->stream));
        SS_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    c->has_inst_aabb = true;
    return SS_OK;
}

namespace {

constexpr uint32_t kAsyncUploadMinNodes = 1u << 16;  // below this a second stream buys nothing

// A whole-stream upload is waiting in the other TRS buffer: the frame stream takes it over.
cudaError_t adopt_pending_upload(ss_ctx* c)
{
    if (!c->trs_swap_pending) return cudaSuccess;
    cudaError_t e = cudaStreamWaitEvent(c->stream, c->ev_upload_done, 0);
    if (e != cudaSuccess) return e;
    c->trs_cur ^= 1;
    c->d_trs = c->d_trs_buf[c->trs_cur];
    c->trs_swap_pending = false;
    return cudaSuccess;
}

}  // namespace

SS_API ss_status ss_upload_trs(ss_ctx* c, uint32_t first, uint32_t count, const float* trs9)
{
    if (!c) return SS_ERR_INVALID;
    if (!trs9 || (uint64_t)first + count > c->n_nodes)
        return c->fail(SS_ERR_INVALID, "ss_upload_trs: range [%u, +%u) exceeds %u nodes", first, count, c->n_nodes);
    SS_CUDA(c, cudaSetDevice(c->device));
    if (!count) return SS_OK;
    c->world_valid = c->nodes_valid = false;  // transforms changed
    c->res_trs_dirty = true;
    if (first == 0 && count == c->n_nodes && count >= kAsyncUploadMinNodes) {
        const int tgt = c->trs_cur ^ 1;
        SS_CUDA(c, cudaStreamWaitEvent(c->upload_stream, c->ev_trs_read[tgt], 0));  // its last reader is past the transform stage
        SS_CUDA(c, cudaMemcpyAsync(c->d_trs_buf[tgt], trs9, (size_t)count * 36, cudaMemcpyHostToDevice, c->upload_stream));
        SS_CUDA(c, cudaEventRecord(c->ev_upload_done, c->upload_stream));
        c->trs_swap_pending = true;
        return SS_OK;
    }
    SS_CUDA(c, adopt_pending_upload(c));  // a range lands on top of whatever whole-stream upload came before it
    SS_CUDA(c, cudaMemcpyAsync(c->d_trs + (size_t)first * 9, trs9, (size_t)count * 36, cudaMemcpyHostToDevice, c->stream));
    return SS_OK;
}

// Shared path of the sparse uploads: (node, record of R floats) pairs, staged and copied packed (4 + 4R bytes per node),
// then scattered by a kernel.  R = 9: TransformComponent records; R = 7: physics poses (position + quaternion).
static ss_status upload_sparse_records(ss_ctx* c, uint32_t count, const uint32_t* node_index, const float* recs, uint32_t R,
                                       const char* who)
{
    if (!c) return SS_ERR_INVALID;
    if (count && (!node_index || !recs)) return c->fail(SS_ERR_INVALID, "%s: null array", who);
    if (!count) return SS_OK;
    // one pass over the list: bounds and strict ascent (what a producer walking its bodies in order sends)
    uint32_t over = node_index[0] >= c->n_nodes, not_asc = 0;
    for (uint32_t k = 1; k < count; ++k) {
        over |= node_index[k] >= c->n_nodes;
        not_asc |= node_index[k] <= node_index[k - 1];
    }
    if (over)
        for (uint32_t k = 0; k < count; ++k)
            if (node_index[k] >= c->n_nodes)
                return c->fail(SS_ERR_INVALID, "%s: record %u names node %u of %u", who, k, node_index[k], c->n_nodes);
    const bool ascending = !not_asc;
    SS_CUDA(c, cudaSetDevice(c->device));
    const size_t rec_bytes = (size_t)R * 4, bytes = (size_t)count * (4 + rec_bytes);
    if (!c->ev_sparse) {
        SS_CUDA(c, cudaEventCreateWithFlags(&c->ev_sparse, cudaEventDisableTiming));
        SS_CUDA(c, cudaEventCreateWithFlags(&c->ev_sparse_copied, cudaEventDisableTiming));
        for (int b = 0; b < 2; ++b) SS_CUDA(c, cudaEventCreateWithFlags(&c->ev_sparse_scattered[b], cudaEventDisableTiming));
    }
    if (bytes > c->sparse_cap) {
        SS_CUDA(c, cudaStreamSynchronize(c->upload_stream));
        SS_CUDA(c, cudaStreamSynchronize(c->stream));
        if (c->h_sparse) cudaFreeHost(c->h_sparse);
        for (int b = 0; b < 2; ++b) {
            if (c->d_sparse[b]) cudaFree(c->d_sparse[b]);
            c->d_sparse[b] = nullptr;
        }
        c->h_sparse = nullptr;
        c->sparse_cap = 0;
        const size_t cap = std::max<size_t>(bytes * 2, 1 << 16);
        SS_CUDA(c, cudaHostAlloc(reinterpret_cast<void**>(&c->h_sparse), cap, cudaHostAllocDefault));
        for (int b = 0; b < 2; ++b) SS_CUDA(c, cudaMalloc(reinterpret_cast<void**>(&c->d_sparse[b]), cap));
        c->sparse_cap = cap;
    }
    // Pinned caller memory with an ascending list (no duplicates possible) is read by the copy engine where it lies (same
    // rule as ss_upload_trs: unchanged until ss_flush_uploads or the next frame's ss_frame_sync).  Anything else is packed
    // into the pinned staging buffer, keeping only the LAST record of a node named more than once.
    bool in_place = false;
    if (ascending) {
        cudaPointerAttributes pa_i{}, pa_r{};
        in_place = cudaPointerGetAttributes(&pa_i, node_index) == cudaSuccess && pa_i.type == cudaMemoryTypeHost &&
                   cudaPointerGetAttributes(&pa_r, recs) == cudaSuccess && pa_r.type == cudaMemoryTypeHost;
        cudaGetLastError();
    }
    uint32_t kept = 0;
    if (in_place) {
        kept = count;
    } else {
        SS_CUDA(c, cudaEventSynchronize(c->ev_sparse));  // the previous staged copy must have left the staging buffer
        uint32_t* hi = reinterpret_cast<uint32_t*>(c->h_sparse);
        if (ascending) {
            kept = count;
            memcpy(hi, node_index, (size_t)count * 4);
            memcpy(c->h_sparse + (size_t)count * 4, recs, (size_t)count * rec_bytes);
        } else {
            std::unordered_map<uint32_t, uint32_t> last;
            last.reserve(count * 2);
            for (uint32_t k = 0; k < count; ++k) last[node_index[k]] = k;
            for (uint32_t k = 0; k < count; ++k)
                if (last[node_index[k]] == k) hi[kept++] = k;          // source positions for now
            float* hr = reinterpret_cast<float*>(c->h_sparse + (size_t)kept * 4);
            for (uint32_t j = 0; j < kept; ++j) {
                const uint32_t k = hi[j];
                memcpy(hr + (size_t)j * R, recs + (size_t)k * R, rec_bytes);
                hi[j] = node_index[k];
            }
        }
    }
    c->world_valid = c->nodes_valid = false;  // transforms changed
    c->res_trs_dirty = true;
    SS_CUDA(c, adopt_pending_upload(c));  // lands on top of whatever whole-stream upload came before it
    // copy on the upload stream (under the frames already issued), scatter on the frame stream once the copy is in
    const int b = c->sparse_cur;
    c->sparse_cur ^= 1;
    unsigned char* dst = c->d_sparse[b];
    SS_CUDA(c, cudaStreamWaitEvent(c->upload_stream, c->ev_sparse_scattered[b], 0));
    if (in_place) {
        SS_CUDA(c, cudaMemcpyAsync(dst, node_index, (size_t)kept * 4, cudaMemcpyHostToDevice, c->upload_stream));
        SS_CUDA(c, cudaMemcpyAsync(dst + (size_t)kept * 4, recs, (size_t)kept * rec_bytes, cudaMemcpyHostToDevice, c->upload_stream));
    } else {
        SS_CUDA(c, cudaMemcpyAsync(dst, c->h_sparse, (size_t)kept * (4 + rec_bytes), cudaMemcpyHostToDevice, c->upload_stream));
        SS_CUDA(c, cudaEventRecord(c->ev_sparse, c->upload_stream));
    }
    SS_CUDA(c, cudaEventRecord(c->ev_sparse_copied, c->upload_stream));
    SS_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_sparse_copied, 0));
    const uint32_t* d_idx = reinterpret_cast<const uint32_t*>(dst);
    const float* d_rec = reinterpret_cast<const float*>(dst + (size_t)kept * 4);
    if (R == 9)
        launch_trs_scatter(c->d_trs, d_idx, d_rec, kept, c->stream);
    else
        launch_pose_scatter(c->d_trs, d_idx, d_rec, kept, c->stream);
    SS_CUDA(c, cudaGetLastError());
    SS_CUDA(c, cudaEventRecord(c->ev_sparse_scattered[b], c->stream));
    return SS_OK;
}

SS_API ss_status ss_upload_trs_sparse(ss_ctx* c, uint32_t count, const uint32_t* node_index, const float* trs9)
{
    return upload_sparse_records(c, count, node_index, trs9, 9, "ss_upload_trs_sparse");
}

SS_API ss_status ss_upload_pose_sparse(ss_ctx* c, uint32_t count, const uint32_t* node_index, const float* pose7)
{
    return upload_sparse_records(c, count, node_index, pose7, 7, "ss_upload_pose_sparse");
}

SS_API ss_status ss_pack_lights(ss_ctx* c, const ss_point_light* point, uint32_t n_point, const ss_spot_light* spot, uint32_t n_spot,
                                void* ub_point_out, void* ub_spot_out, float* positions_out)
{
    if (!c) return SS_ERR_INVALID;
    if (n_point > SS_MAX_LIGHTS || n_spot > SS_MAX_LIGHTS || (n_point && !point) || (n_spot && !spot))
        return c->fail(SS_ERR_INVALID, "ss_pack_lights: at most %u lights of a kind (the reference's uniform arrays)", SS_MAX_LIGHTS);
    for (uint32_t k = 0; k < n_point; ++k)
        if (point[k].node >= c->n_nodes) return c->fail(SS_ERR_INVALID, "ss_pack_lights: point light %u names node %u of %u", k, point[k].node, c->n_nodes);
    for (uint32_t k = 0; k < n_spot; ++k)
        if (spot[k].node >= c->n_nodes) return c->fail(SS_ERR_INVALID, "ss_pack_lights: spot light %u names node %u of %u", k, spot[k].node, c->n_nodes);
    SS_CUDA(c, cudaSetDevice(c->device));
    constexpr size_t kIn = SS_MAX_LIGHTS * (sizeof(ss_point_light) + sizeof(ss_spot_light));
    constexpr size_t kOut = SS_UB_POINT_LIGHTS_BYTES + SS_UB_SPOT_LIGHTS_BYTES + 2 * SS_MAX_LIGHTS * 16;
    if (!c->d_lights) SS_CUDA(c, cudaMalloc(reinterpret_cast<void**>(&c->d_lights), kIn + kOut));
    SS_CUDA(c, adopt_pending_upload(c));  // the newest transforms
    unsigned char* d_out = c->d_lights + kIn;
    ss_point_light* d_point = reinterpret_cast<ss_point_light*>(c->d_lights);
    ss_spot_light* d_spot = reinterpret_cast<ss_spot_light*>(c->d_lights + SS_MAX_LIGHTS * sizeof(ss_point_light));
    SS_CUDA(c, cudaMemsetAsync(d_out, 0, kOut, c->stream));
    if (n_point) SS_CUDA(c, cudaMemcpyAsync(d_point, point, n_point * sizeof(ss_point_light), cudaMemcpyHostToDevice, c->stream));
    if (n_spot) SS_CUDA(c, cudaMemcpyAsync(d_spot, spot, n_spot * sizeof(ss_spot_light), cudaMemcpyHostToDevice, c->stream));
    launch_pack_lights(c->d_trs, d_point, n_point, d_spot, n_spot, d_out, c->stream);
    SS_CUDA(c, cudaGetLastError());
    unsigned char host[kOut];
    SS_CUDA(c, cudaMemcpyAsync(host, d_out, kOut, cudaMemcpyDeviceToHost, c->stream));
    SS_CUDA(c, cudaStreamSynchronize(c->stream));
    if (ub_point_out) memcpy(ub_point_out, host, SS_UB_POINT_LIGHTS_BYTES);
    if (ub_spot_out) memcpy(ub_spot_out, host + SS_UB_POINT_LIGHTS_BYTES, SS_UB_SPOT_LIGHTS_BYTES);
    if (positions_out) memcpy(positions_out, host + SS_UB_POINT_LIGHTS_BYTES + SS_UB_SPOT_LIGHTS_BYTES, (size_t)(n_point + n_spot) * 16);
    return SS_OK;
}

SS_API ss_status ss_upload_node_world_rows(ss_ctx* c, uint32_t first, uint32_t count, const float* rows12)
{
    if (!c) return SS_ERR_INVALID;
    const uint32_t computed = c->level_offsets.empty() ? 0 : c->level_offsets.back();
    if (!rows12 || first < computed || (uint64_t)first + count > c->n_nodes)
        return c->fail(SS_ERR_INVALID, "ss_upload_node_world_rows: range [%u, +%u) must lie in the external nodes [%u, %u)",
                       first, count, computed, c->n_nodes);
    SS_CUDA(c, cudaSetDevice(c->device));
    if (count) {
        c->world_valid = c->nodes_valid = false;  // instances on these nodes move: SS_FRAME_REUSE_WORLD must not keep their old rows
        SS_CUDA(c, cudaMemcpyAsync(c->d_node_world + (size_t)first * 3, rows12, (size_t)count * 48, cudaMemcpyHostToDevice,
                                   c->stream));
    }
    return SS_OK;
}

SS_API ss_status ss_flush_uploads(ss_ctx* c)
{
    if (!c) return SS_ERR_INVALID;
    SS_CUDA(c, cudaSetDevice(c->device));
    SS_CUDA(c, cudaStreamSynchronize(c->upload_stream));
    SS_CUDA(c, cudaStreamSynchronize(c->stream));
    return SS_OK;
}

SS_API ss_status ss_upload_world_rows(ss_ctx* c, uint32_t first, uint32_t count, const float* rows12)
{
    if (!c) return SS_ERR_INVALID;
    if (!rows12 || (uint64_t)first + count > c->n_instances)
        return c->fail(SS_ERR_INVALID, "ss_upload_world_rows: range [%u, +%u) exceeds %u instances", first, count,
                       c->n_instances);
    SS_CUDA(c, cudaSetDevice(c->device));
    if (count)
        SS_CUDA(c, cudaMemcpy2DAsync(c->d_world + (size_t)first * kWorldStride, kWorldStride * 16, rows12, 48, 48, count,
                                     cudaMemcpyHostToDevice, c->stream));
    c->world_uploaded = true;
    return SS_OK;
}

SS_API ss_status ss_frame(ss_ctx* c, const ss_view* views, uint32_t n_views, uint32_t flags, uint32_t* slot_out)
{
    if (!c) return SS_ERR_INVALID;
    if (!views || n_views < 1 || n_views > c->cfg.max_views)
        return c->fail(SS_ERR_INVALID, "ss_frame: %u views, capacity %u", n_views, c->cfg.max_views);
    if (c->n_slots == 0 || c->n_nodes == 0) return c->fail(SS_ERR_STATE, "ss_frame: submesh table / nodes not set");
    if ((flags & SS_FRAME_WORLD_ROWS_GIVEN) && !c->world_uploaded)
        return c->fail(SS_ERR_STATE, "ss_frame: SS_FRAME_WORLD_ROWS_GIVEN without ss_upload_world_rows");
    if (!c->has_node_map && c->n_instances && !implicit_nodes_in_last_level(c, c->node_base, c->n_instances))
        return c->fail(SS_ERR_STATE, "ss_frame: ss_set_nodes changed the levels under an implicit node map; call ss_set_instances again");
    SS_CUDA(c, cudaSetDevice(c->device));

    const uint32_t si = (uint32_t)(c->frame_no % c->slots.size());
    Slot& s = c->slots[si];
    if (s.pending) {  // the slot's previous frame must have been consumed (FramesInFlight back-pressure)
        SS_CUDA(c, cudaEventSynchronize(s.done));
        s.pending = false;
    }
    cudaStream_t st = c->stream;
    int launches = 0;
    const uint32_t n = c->n_instances;

    CullParams cp;
    memset(&cp, 0, sizeof cp);
    cp.n_views = n_views;
    for (uint32_t v = 0; v < n_views; ++v) {
        planes_from_view_proj(views[v].view_proj, cp.v[v]);
        if (views[v].flags & SS_VIEW_SKIP_NEAR) cp.v[v].flags |= kViewSkipNear;
    }

    // bits the key layout can set: material | submesh | mesh | tess | dropped.  When they fit next to the
    // instance index in one 64-bit word, key and index are sorted as that single word (8 B/element/pass).
    const uint32_t tess_shift = c->mesh_bits + c->submesh_bits + c->material_bits;
    const int key_bits = (int)tess_shift + 2;
    int idx_bits = 1;
    while (idx_bits < 32 && (1ull << idx_bits) < (uint64_t)n) ++idx_bits;
    const bool packed = idx_bits + key_bits <= 64;

    // Which pipeline.  The draw-order-resident layout serves TRS-driven instances with an implicit node map and
    // deterministic trig whose key and index share a word; everything else takes the general three-kernel pipeline.
    const bool resident = (c->cfg.flags & SS_CONFIG_DRAW_ORDER_RESIDENT) && !c->has_node_map &&
                          c->cfg.trig_mode == SS_TRIG_DETERMINISTIC && !(flags & SS_FRAME_WORLD_ROWS_GIVEN) && n > 0 && packed;
    // world rows already on the device: uploaded for stage-wise parity, or kept from the previous frame
    const bool world_given = !resident && ((flags & SS_FRAME_WORLD_ROWS_GIVEN) || ((flags & SS_FRAME_REUSE_WORLD) && c->world_valid));
    const bool nodes_kept = resident && (flags & SS_FRAME_REUSE_WORLD) && c->nodes_valid;  // node rows of the parent levels
    const bool reuse_sort = !resident && (flags & SS_FRAME_REUSE_SORT) && c->sort_valid;

    // NVTX ranges carry the names of the reference's Optick scopes around the loops this frame replaces:
    // "Submit-SceneRenderer" (VC/Scene/SceneRenderer.cpp:1062), "Submit-SubmitMeshes" (VC/Scene/Scene.cpp:137),
    // "SubmitMesh" (SceneRenderer.cpp:1126)
    struct NvtxScope {
        bool open = true;
        explicit NvtxScope(const char* name) { nvtxRangePushA(name); }
        void end() { if (open) nvtxRangePop(); open = false; }
        ~NvtxScope() { end(); }
    } nvtx_frame("Submit-SceneRenderer");
    SS_CUDA(c, adopt_pending_upload(c));
    SS_CUDA(c, cudaEventRecord(s.start, st));
    SS_CUDA(c, cudaMemsetAsync(c->d_emit_zero, 0, c->emit_zero_bytes, st));

    // N1: levels above the instances (all levels when an explicit instance->node map is in use)
    if (!world_given && !nodes_kept && (c->has_parent || c->has_node_map)) {
        const uint32_t node_hi = c->has_node_map ? c->n_nodes : c->node_base;
        for (uint32_t l = 0; l < c->n_levels; ++l) {
            NodeLevelArgs na;
            na.trs = c->d_trs;
            na.parent = c->has_parent ? c->d_parent : nullptr;
            na.node_world = c->d_node_world;
            na.lo = std::min(c->level_offsets[l], node_hi);
            na.hi = std::min(c->level_offsets[l + 1], node_hi);
            na.trig_mode = (int)c->cfg.trig_mode;
            if (na.hi > na.lo) {
                launch_node_level(na, st);
                ++launches;
            }
        }
    }

    const size_t seg_words = (size_t)c->cfg.max_views * c->max_draws;
    const uint64_t* words_known = nullptr;    // resident layout: the sorted words

    NvtxScope nvtx_meshes("Submit-SubmitMeshes");
    if (resident) {
        // Draw-order-resident layout: the sort runs when a static stream changes, not per frame; the streams are kept
        // in sorted order, so the instance kernel writes world rows and masks in draw order and k_emit streams them.
        const int begin_bit = idx_bits, end_bit = idx_bits + key_bits;
        const SortPlan plan = sort_plan(begin_bit, end_bit);
        const uint64_t imask = (1ull << idx_bits) - 1ull;
        SortControl* ctl = reinterpret_cast<SortControl*>(c->d_sort_scratch);
        if (!c->resident_valid) {
            const bool with_hist = plan.n_passes <= 3;
            WordsArgs wa;
            memset(&wa, 0, sizeof wa);
            wa.slot = c->d_slot;
            wa.matword = c->d_matword;
            wa.slots = c->d_slots;
            wa.words = c->d_keys_a;
            wa.n = n;
            wa.tess_shift = tess_shift;
            wa.idx_bits = (uint32_t)idx_bits;
            if (with_hist) {
                SS_CUDA(c, sort_prepare(c->d_sort_scratch, n, begin_bit, end_bit, st));
                wa.hist = ctl->hist[0];
                wa.hist_passes = plan.n_passes;
                for (int p = 0; p < plan.n_passes; ++p) {
                    wa.hist_shift[p] = plan.shift[p];
                    wa.hist_bits[p] = plan.bits[p];
                }
            }
            launch_build_words(wa, st);
            ++launches;
            SS_CUDA(c, sort_async(c->d_keys_a, nullptr, c->d_keys_b, nullptr, n, begin_bit, end_bit, c->d_sort_scratch, st,
                                  &launches, with_hist));
            uint32_t sel = 0;
            SS_CUDA(c, cudaMemcpyAsync(&sel, &ctl->src_sel[plan.n_passes], sizeof sel, cudaMemcpyDeviceToHost, st));
            SS_CUDA(c, cudaStreamSynchronize(st));
            c->d_words_resident = sel ? c->d_keys_b : c->d_keys_a;
            c->resident_valid = true;
            c->res_static_dirty = true;
        }
        if (c->res_static_dirty || c->res_trs_dirty) {
            PermuteArgs pa;
            memset(&pa, 0, sizeof pa);
            pa.order = c->d_words_resident;
            pa.idx_mask = imask;
            pa.n = n;
            pa.trs_only = c->res_static_dirty ? 0 : 1;
            pa.trs = c->d_trs + (size_t)c->node_base * 9;
            pa.box = c->has_inst_aabb ? c->d_inst_aabb : nullptr;
            pa.slot = c->d_slot;
            pa.matword = c->d_matword;
            pa.parent = c->has_parent ? c->d_parent + c->node_base : nullptr;
            pa.trs_out = c->d_trs_o;
            pa.box_out = c->d_box_o;
            pa.slot_out = c->d_slot_o;
            pa.matword_out = c->d_matword_o;
            pa.parent_out = c->d_parent_o;
            pa.pos_out = c->d_pos;
            launch_permute(pa, st);
            ++launches;
            c->res_static_dirty = c->res_trs_dirty = false;
        }
        SS_CUDA(c, cudaEventRecord(s.stage[0], st));
        SS_CUDA(c, cudaEventRecord(c->ev_trs_read[c->trs_cur], st));  // the permuted stream holds the transforms now

        InstanceArgs ia;
        memset(&ia, 0, sizeof ia);
        ia.trs = c->d_trs_o;
        ia.parent = c->has_parent ? c->d_parent_o : nullptr;
        ia.node_world = c->d_node_world;
        ia.node = nullptr;
        ia.node_base = 0;
        ia.slot = c->d_slot_o;
        ia.matword = c->d_matword_o;
        ia.inst_aabb = c->has_inst_aabb ? c->d_box_o : nullptr;
        ia.slots = c->d_slots;
        ia.world = c->d_world;   // rows of sorted position j, packed 48 B apart: their only reader streams them
        ia.world_stride = 3;
        ia.mask = c->d_mask;
        ia.keys = nullptr;       // the order is kept: masks only
        ia.vals = nullptr;
        ia.n = n;
        ia.tess_shift = tess_shift;
        ia.trig_mode = (int)c->cfg.trig_mode;
        launch_instances(ia, cp, st);
        ++launches;
        SS_CUDA(c, cudaEventRecord(s.stage[1], st));
        SS_CUDA(c, cudaEventRecord(s.stage[2], st));  // no sort

        EmitArgs ea;
        memset(&ea, 0, sizeof ea);
        ea.keys_a = c->d_words_resident;
        ea.idx_bits = (uint32_t)idx_bits;
        ea.seq = 1;
        ea.world_stride = 3;
        ea.mask = c->d_mask;
        ea.world = c->d_world;
        ea.n = n;
        ea.n_views = n_views;
        for (uint32_t v = 0; v < n_views; ++v) ea.out[v] = s.out[v];
        ea.cap_visible = c->cap_visible;
        ea.max_draws = c->max_draws;
        ea.lookback = c->d_lookback;
        ea.ticket = c->d_ticket;
        ea.seg_count = c->d_seg_count + c->seg_parity * seg_words;
        ea.seg_key = c->d_seg_key;
        ea.seg_inst = c->d_seg_inst;
        ea.totals = c->d_totals;
        ea.overflow = c->d_overflow;
        launch_emit(ea, st);
        ++launches;
        words_known = c->d_words_resident;
        c->nodes_valid = true;
        c->world_valid = c->sort_valid = false;  // the instance-order rows / packed words of the general pipeline are gone
        c->words_idx_bits = (uint32_t)idx_bits;
        c->last_pipeline = SS_PIPELINE_RESIDENT;
        SS_CUDA(c, cudaEventRecord(s.stage[3], st));
    } else {
        SS_CUDA(c, cudaEventRecord(s.stage[0], st));

        InstanceArgs ia;
        ia.trs = c->d_trs;
        ia.parent = c->has_parent ? c->d_parent : nullptr;
        ia.node_world = c->d_node_world;
        ia.node = c->has_node_map ? c->d_node : nullptr;
        ia.node_base = c->node_base;
        ia.slot = c->d_slot;
        ia.matword = c->d_matword;
        ia.inst_aabb = c->has_inst_aabb ? c->d_inst_aabb : nullptr;
        ia.slots = c->d_slots;
        ia.world = c->d_world;
        ia.world_stride = kWorldStride;
        ia.mask = c->d_mask;
        ia.keys = c->d_keys_a;
        ia.vals = c->d_vals_a;
        ia.n = n;
        ia.tess_shift = tess_shift;
        ia.world_given = world_given ? 1 : 0;
        ia.trig_mode = (int)c->cfg.trig_mode;
        ia.idx_bits = packed ? (uint32_t)idx_bits : 0u;
        ia.mask_bits = (packed && idx_bits + key_bits + 8 <= 64) ? 8u : 0u;
        // The sort's digit plan is known up front, so its scratch is cleared before the instance kernel runs and
        // the fast path accumulates the digit histograms on the way (saves the sort's own pass over the keys).
        const int begin_bit = packed ? idx_bits + (int)ia.mask_bits : 0, end_bit = begin_bit + key_bits;
        const SortPlan plan = sort_plan(begin_bit, end_bit);
        ia.hist = reinterpret_cast<SortControl*>(c->d_sort_scratch)->hist[0];
        ia.hist_passes = plan.n_passes <= 3 ? plan.n_passes : 0;
        for (int p = 0; p < 3; ++p) {
            ia.hist_shift[p] = p < plan.n_passes ? plan.shift[p] - begin_bit : 0;
            ia.hist_bits[p] = p < plan.n_passes ? plan.bits[p] : 0;
        }
        if (reuse_sort) {  // keep the previous frame's sorted order: no keys, no histograms, no sort
            ia.keys = nullptr;
            ia.vals = nullptr;
            ia.hist_passes = 0;
        } else {
            SS_CUDA(c, sort_prepare(c->d_sort_scratch, n, begin_bit, end_bit, st));
        }
        bool hist_done = false;
        if (n) {
            hist_done = launch_instances(ia, cp, st);
            ++launches;
        }
        c->world_valid = true;
        c->nodes_valid = true;
        c->resident_valid = false;  // the sort buffers now belong to this pipeline
        c->last_pipeline = SS_PIPELINE_GENERAL;
        SS_CUDA(c, cudaEventRecord(s.stage[1], st));
        SS_CUDA(c, cudaEventRecord(c->ev_trs_read[c->trs_cur], st));  // the TRS buffer may be overwritten from here on

        if (!reuse_sort) {
            if (!hist_done)  // the general kernel did not count digits: clear again is harmless, the sort counts itself
                SS_CUDA(c, sort_prepare(c->d_sort_scratch, n, begin_bit, end_bit, st));
            SS_CUDA(c, sort_async(c->d_keys_a, packed ? nullptr : c->d_vals_a, c->d_keys_b, packed ? nullptr : c->d_vals_b,
                                  n, begin_bit, end_bit, c->d_sort_scratch, st, &launches, hist_done));
            c->sort_valid = true;
        }
        SS_CUDA(c, cudaEventRecord(s.stage[2], st));

        EmitArgs ea;
        memset(&ea, 0, sizeof ea);
        ea.keys_a = c->d_keys_a;
        ea.vals_a = c->d_vals_a;
        ea.keys_b = c->d_keys_b;
        ea.vals_b = c->d_vals_b;
        ea.sort_ctl = reinterpret_cast<const SortControl*>(c->d_sort_scratch);
        ea.sort_passes = (uint32_t)sort_num_passes(begin_bit, end_bit);
        ea.idx_bits = ia.idx_bits;
        ea.mask_bits = ia.mask_bits;
        // the visibility byte inside a sorted word is this frame's only if this frame built the words
        ea.gather_mask = (!packed || ia.mask_bits == 0 || reuse_sort) ? 1u : 0u;
        ea.mask = c->d_mask;
        ea.world = c->d_world;
        ea.world_stride = kWorldStride;
        ea.n = n;
        ea.n_views = n_views;
        for (uint32_t v = 0; v < n_views; ++v) ea.out[v] = s.out[v];
        ea.cap_visible = c->cap_visible;
        ea.max_draws = c->max_draws;
        ea.lookback = c->d_lookback;
        ea.ticket = c->d_ticket;
        ea.seg_count = c->d_seg_count + c->seg_parity * seg_words;
        ea.seg_key = c->d_seg_key;
        ea.seg_inst = c->d_seg_inst;
        ea.totals = c->d_totals;
        ea.overflow = c->d_overflow;
        if (n) {
            launch_emit(ea, st);
            ++launches;
        }
        SS_CUDA(c, cudaEventRecord(s.stage[3], st));
    }

    nvtx_meshes.end();
    DrawArgs da;
    memset(&da, 0, sizeof da);
    da.n_views = n_views;
    da.max_draws = c->max_draws;
    da.totals = c->d_totals;
    da.seg_count = c->d_seg_count + c->seg_parity * seg_words;
    da.seg_count_next = c->d_seg_count + (c->seg_parity ^ 1u) * seg_words;
    da.runs_cur = c->d_seg_runs + c->seg_parity;
    da.runs_next = c->d_seg_runs + (c->seg_parity ^ 1u);
    da.passes_run = &reinterpret_cast<SortControl*>(c->d_sort_scratch)->passes_run;
    da.host_stats = s.h_counts_dev;
    da.stats_views = c->cfg.max_views;
    c->seg_parity ^= 1u;
    da.seg_key = c->d_seg_key;
    da.seg_inst = c->d_seg_inst;
    da.slot = c->d_slot;
    da.slots = c->d_slots;
    for (uint32_t v = 0; v < n_views; ++v) da.out[v] = s.out[v];
    da.overflow = c->d_overflow;
    launch_draws(da, st);
    ++launches;
    SS_CUDA(c, cudaGetLastError());

    // k_draws has written the counts, the overflow flag and passes_run into s.h_counts (mapped pinned memory)
    SS_CUDA(c, cudaEventRecord(s.done, st));
    s.pending = true;
    s.n_views = n_views;
    memcpy(s.views, views, n_views * sizeof(ss_view));
    s.n_instances = n;
    s.launches = launches;
    s.pipeline = c->last_pipeline;
    s.words_known = words_known;
    s.idx_bits = resident ? (uint32_t)idx_bits : 0u;
    ++c->frame_no;
    if (slot_out) *slot_out = si;
    return SS_OK;
}

SS_API ss_status ss_frame_sync(ss_ctx* c, uint32_t slot, ss_frame_stats* stats)
{
    if (!c) return SS_ERR_INVALID;
    if (slot >= c->slots.size()) return c->fail(SS_ERR_INVALID, "ss_frame_sync: slot %u of %zu", slot, c->slots.size());
    Slot& s = c->slots[slot];
    if (s.n_views == 0) return c->fail(SS_ERR_STATE, "ss_frame_sync: slot %u has no frame", slot);
    SS_CUDA(c, cudaSetDevice(c->device));
    SS_CUDA(c, cudaEventSynchronize(s.done));
    s.pending = false;
    const size_t V = c->cfg.max_views;
    if (stats) {
        memset(stats, 0, sizeof *stats);
        stats->n_instances = s.n_instances;
        stats->n_views = s.n_views;
        for (uint32_t v = 0; v < s.n_views; ++v) {
            stats->visible[v] = std::min(s.h_counts[v * 2 + 0], c->cap_visible);
            stats->draws[v] = s.h_counts[v * 2 + 1];
        }
        stats->overflow = s.h_counts[V * 2];
        stats->sort_passes_run = s.h_counts[V * 2 + 1];
        stats->kernel_launches = (uint32_t)s.launches;
        stats->pipeline = (uint32_t)s.pipeline;
        float ms = 0.f;
        SS_CUDA(c, cudaEventElapsedTime(&ms, s.start, s.done));
        stats->gpu_ms = ms;
    }
    if (s.h_counts[V * 2]) return c->fail(SS_ERR_CAPACITY, "frame overflowed max_visible_per_view / max_draws_per_view");
    return SS_OK;
}

SS_API ss_status ss_map_outputs(ss_ctx* c, uint32_t slot, uint32_t view, ss_outputs* out)
{
    if (!c || !out) return SS_ERR_INVALID;
    if (slot >= c->slots.size() || view >= c->cfg.max_views)
        return c->fail(SS_ERR_INVALID, "ss_map_outputs: slot %u view %u out of range", slot, view);
    Slot& s = c->slots[slot];
    if (view >= s.n_views) return c->fail(SS_ERR_STATE, "ss_map_outputs: slot %u ran %u views", slot, s.n_views);
    SS_CUDA(c, cudaSetDevice(c->device));
    SS_CUDA(c, cudaEventSynchronize(s.done));
    s.pending = false;
    out->instance_rows = reinterpret_cast<const float*>(s.out[view].rows);
    out->keys = s.out[view].keys;
    out->instance_idx = s.out[view].idx;
    out->draws = s.out[view].draws;
    out->draw_keys = s.out[view].draw_keys;
    out->counts = s.out[view].counts;
    out->world_rows = reinterpret_cast<const float*>(c->d_world);
    out->masks = c->d_mask;
    out->visible = std::min(s.h_counts[view * 2 + 0], c->cap_visible);
    out->n_draws = s.h_counts[view * 2 + 1];
    out->world_words = nullptr;
    out->world_idx_bits = 0;
    out->world_stride = kWorldStride * 4;
    if (s.pipeline == SS_PIPELINE_RESIDENT) {
        out->world_words = s.words_known;
        out->world_idx_bits = s.idx_bits;
        out->world_stride = 12;
    }
    return SS_OK;
}

SS_API ss_status ss_download_view(ss_ctx* c, uint32_t slot, uint32_t view, float* rows12, uint64_t* keys,
                                  uint32_t* instance_idx, ss_draw_cmd* draws, uint64_t* draw_keys)
{
    ss_outputs o;
    ss_status st = ss_map_outputs(c, slot, view, &o);
    if (st != SS_OK) return st;
    // the slot's frame is complete (ss_map_outputs fenced it) and its buffers stay untouched until the slot comes
    // round again, so the copies run on their own stream instead of waiting for frames issued since
    cudaStream_t s = c->download_stream;
    if (rows12 && o.visible)
        SS_CUDA(c, cudaMemcpyAsync(rows12, o.instance_rows, (size_t)o.visible * 48, cudaMemcpyDeviceToHost, s));
    if (keys && o.visible) SS_CUDA(c, cudaMemcpyAsync(keys, o.keys, (size_t)o.visible * 8, cudaMemcpyDeviceToHost, s));
    if (instance_idx && o.visible)
        SS_CUDA(c, cudaMemcpyAsync(instance_idx, o.instance_idx, (size_t)o.visible * 4, cudaMemcpyDeviceToHost, s));
    if (draws && o.n_draws)
        SS_CUDA(c, cudaMemcpyAsync(draws, o.draws, (size_t)o.n_draws * sizeof(ss_draw_cmd), cudaMemcpyDeviceToHost, s));
    if (draw_keys && o.n_draws)
        SS_CUDA(c, cudaMemcpyAsync(draw_keys, o.draw_keys, (size_t)o.n_draws * 8, cudaMemcpyDeviceToHost, s));
    SS_CUDA(c, cudaStreamSynchronize(s));
    return SS_OK;
}

static ss_status download_range(ss_ctx* c, const void* src, size_t elem, uint32_t first, uint32_t count, void* dst,
                                const char* what)
{
    if (!c) return SS_ERR_INVALID;
    if (!dst || (uint64_t)first + count > c->n_instances)
        return c->fail(SS_ERR_INVALID, "%s: range [%u, +%u) exceeds %u instances", what, first, count, c->n_instances);
    SS_CUDA(c, cudaSetDevice(c->device));
    if (count)
        SS_CUDA(c, cudaMemcpyAsync(dst, static_cast<const unsigned char*>(src) + (size_t)first * elem,
                                   (size_t)count * elem, cudaMemcpyDeviceToHost, c->stream));
    SS_CUDA(c, cudaStreamSynchronize(c->stream));
    return SS_OK;
}

// Draw-order-resident layout: world rows and masks lie in word order; bring the rows / bytes of instances [first, first + count)
// back to instance order through a scratch buffer (a download is not a per-frame operation).
static ss_status download_word_ordered(ss_ctx* c, uint32_t first, uint32_t count, float* rows12, uint8_t* masks)
{
    SS_CUDA(c, cudaStreamSynchronize(c->stream));
    const uint64_t* words = c->d_words_resident;
    void* tmp = nullptr;
    SS_CUDA(c, cudaMalloc(&tmp, std::max<size_t>((size_t)count * (rows12 ? 48 : 1), 256)));
    launch_unpermute(words, (1ull << c->words_idx_bits) - 1ull, c->n_instances, first, count,
                     reinterpret_cast<const float4*>(c->d_world), 3u, rows12 ? static_cast<float4*>(tmp) : nullptr, c->d_mask,
                     masks ? static_cast<uint8_t*>(tmp) : nullptr, c->stream);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(rows12 ? static_cast<void*>(rows12) : static_cast<void*>(masks), tmp, (size_t)count * (rows12 ? 48 : 1),
                            cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(tmp);
    if (e != cudaSuccess) return c->fail(SS_ERR_CUDA, "download (word order): %s", cudaGetErrorString(e));
    return SS_OK;
}

SS_API ss_status ss_download_world(ss_ctx* c, uint32_t first, uint32_t count, float* rows12)
{
    if (!c) return SS_ERR_INVALID;
    if (!rows12 || (uint64_t)first + count > c->n_instances)
        return c->fail(SS_ERR_INVALID, "ss_download_world: range [%u, +%u) exceeds %u instances", first, count,
                       c->n_instances);
    SS_CUDA(c, cudaSetDevice(c->device));
    if (count && c->last_pipeline == SS_PIPELINE_RESIDENT) return download_word_ordered(c, first, count, rows12, nullptr);
    if (count)  // device rows sit at a 64 B stride; the host gets them packed as TransformData
        SS_CUDA(c, cudaMemcpy2DAsync(rows12, 48, c->d_world + (size_t)first * kWorldStride, kWorldStride * 16, 48, count,
                                     cudaMemcpyDeviceToHost, c->stream));
    SS_CUDA(c, cudaStreamSynchronize(c->stream));
    return SS_OK;
}
SS_API ss_status ss_download_masks(ss_ctx* c, uint32_t first, uint32_t count, uint8_t* masks)
{
    if (c && masks && count && (uint64_t)first + count <= c->n_instances && c->last_pipeline == SS_PIPELINE_RESIDENT) {
        SS_CUDA(c, cudaSetDevice(c->device));
        return download_word_ordered(c, first, count, nullptr, masks);
    }
    return download_range(c, c ? c->d_mask : nullptr, 1, first, count, masks, "ss_download_masks");
}

SS_API ss_status ss_frame_stage_ms(ss_ctx* c, uint32_t slot, float ms_out[5])
{
    if (!c || !ms_out) return SS_ERR_INVALID;
    if (slot >= c->slots.size()) return c->fail(SS_ERR_INVALID, "ss_frame_stage_ms: slot %u of %zu", slot, c->slots.size());
    Slot& s = c->slots[slot];
    if (s.n_views == 0) return c->fail(SS_ERR_STATE, "ss_frame_stage_ms: slot %u has no frame", slot);
    SS_CUDA(c, cudaSetDevice(c->device));
    SS_CUDA(c, cudaEventSynchronize(s.done));
    cudaEvent_t ev[6] = {s.start, s.stage[0], s.stage[1], s.stage[2], s.stage[3], s.done};
    for (int i = 0; i < 5; ++i) SS_CUDA(c, cudaEventElapsedTime(&ms_out[i], ev[i], ev[i + 1]));
    return SS_OK;
}

// ---- consumer pass: depth + id rasterisation -------------------------------------------------------
SS_API ss_status ss_set_geometry(ss_ctx* c, const float* positions_xyz, uint32_t n_vertices, const uint32_t* indices,
                                 uint32_t n_indices)
{
    if (!c) return SS_ERR_INVALID;
    if ((n_vertices && !positions_xyz) || (n_indices && !indices)) return c->fail(SS_ERR_INVALID, "ss_set_geometry: null array");
    SS_CUDA(c, cudaSetDevice(c->device));
    SS_CUDA(c, cudaStreamSynchronize(c->stream));  // a raster pass may still be reading the old arrays
    if (c->d_positions) cudaFree(c->d_positions);
    if (c->d_indices) cudaFree(c->d_indices);
    c->d_positions = nullptr;
    c->d_indices = nullptr;
    c->n_vertices = c->n_indices = 0;
    if (n_vertices) {
        SS_CUDA(c, cudaMalloc(&c->d_positions, (size_t)n_vertices * 12));
        SS_CUDA(c, cudaMemcpy(c->d_positions, positions_xyz, (size_t)n_vertices * 12, cudaMemcpyHostToDevice));
    }
    if (n_indices) {
        SS_CUDA(c, cudaMalloc(&c->d_indices, (size_t)n_indices * 4));
        SS_CUDA(c, cudaMemcpy(c->d_indices, indices, (size_t)n_indices * 4, cudaMemcpyHostToDevice));
    }
    c->n_vertices = n_vertices;
    c->n_indices = n_indices;
    return SS_OK;
}

SS_API ss_status ss_raster_view(ss_ctx* c, uint32_t slot, uint32_t view, uint32_t width, uint32_t height, uint32_t flags)
{
    if (!c) return SS_ERR_INVALID;
    if (!width || !height || width > 16384 || height > 16384)
        return c->fail(SS_ERR_INVALID, "ss_raster_view: target %u x %u", width, height);
    ss_outputs o;
    ss_status st = ss_map_outputs(c, slot, view, &o);  // fences the frame: counts are on the host
    if (st != SS_OK) return st;
    if (width != c->raster_w[view] || height != c->raster_h[view] || !c->d_raster[view]) {
        SS_CUDA(c, cudaStreamSynchronize(c->stream));
        if (c->d_raster[view]) cudaFree(c->d_raster[view]);
        c->d_raster[view] = nullptr;
        c->raster_w[view] = c->raster_h[view] = 0;
        SS_CUDA(c, cudaMalloc(&c->d_raster[view], (size_t)width * height * sizeof(unsigned long long)));
        c->raster_w[view] = width;
        c->raster_h[view] = height;
    }
    const size_t layer = (size_t)width * height;
    RasterArgs ra;
    memset(&ra, 0, sizeof ra);
    ra.rows = reinterpret_cast<const float4*>(o.instance_rows);
    ra.idx = o.instance_idx;
    ra.draws = o.draws;
    ra.counts = o.counts;
    ra.max_draws = c->max_draws;
    if (!c->d_raster_order) SS_CUDA(c, cudaMalloc(&c->d_raster_order, raster_order_bytes(c->n_instances, c->max_draws)));
    raster_order_carve(c->d_raster_order, c->n_instances, c->max_draws, &ra);
    constexpr uint32_t kRasterQueueWorks = 1u << 19, kRasterQueueTiles = 1u << 23;  // 40 MB + 64 MB (two words per entry), on first use
    if (!c->d_raster_queue)
        SS_CUDA(c, cudaMalloc(&c->d_raster_queue, ((size_t)kRasterQueueWorks * 20 + 2 * (size_t)kRasterQueueTiles + 4) * sizeof(uint32_t)));
    ra.queue_works = c->d_raster_queue;
    ra.queue_tiles = c->d_raster_queue + (size_t)kRasterQueueWorks * 20;
    ra.queue_counts = ra.queue_tiles + 2 * (size_t)kRasterQueueTiles;
    ra.queue_work_cap = kRasterQueueWorks;
    ra.queue_tile_cap = kRasterQueueTiles;
    ra.positions = c->d_positions;
    ra.indices = c->d_indices;
    ra.n_vertices = c->n_vertices;
    ra.n_indices = c->n_indices;
    ra.visible = o.visible;
    memcpy(ra.vp, c->slots[slot].views[view].view_proj, 64);
    ra.width = width;
    ra.height = height;
    ra.flags = flags;
    ra.target = c->d_raster[view];
    const size_t hiz_words = raster_hiz_words(width, height);
    if (hiz_words > c->raster_hiz_cap) {
        SS_CUDA(c, cudaStreamSynchronize(c->stream));
        if (c->d_raster_hiz) cudaFree(c->d_raster_hiz);
        c->d_raster_hiz = nullptr;
        c->raster_hiz_cap = 0;
        SS_CUDA(c, cudaMalloc(&c->d_raster_hiz, hiz_words * sizeof(unsigned long long)));
        c->raster_hiz_cap = hiz_words;
    }
    ra.hiz = c->d_raster_hiz;
    launch_raster_clear(ra.target, layer, c->stream);
    launch_raster(ra, c->stream);
    SS_CUDA(c, cudaGetLastError());
    return SS_OK;
}

SS_API ss_status ss_map_raster(ss_ctx* c, uint32_t view, const uint64_t** words_out)
{
    if (!c || !words_out) return SS_ERR_INVALID;
    if (view >= c->cfg.max_views || !c->d_raster[view]) return c->fail(SS_ERR_STATE, "ss_map_raster: no target for view %u", view);
    SS_CUDA(c, cudaSetDevice(c->device));
    SS_CUDA(c, cudaStreamSynchronize(c->stream));
    *words_out = reinterpret_cast<const uint64_t*>(c->d_raster[view]);
    return SS_OK;
}

SS_API ss_status ss_download_raster(ss_ctx* c, uint32_t view, uint64_t* words_out)
{
    if (!c || !words_out) return SS_ERR_INVALID;
    if (view >= c->cfg.max_views || !c->d_raster[view]) return c->fail(SS_ERR_STATE, "ss_download_raster: no target for view %u", view);
    SS_CUDA(c, cudaSetDevice(c->device));
    const size_t layer = (size_t)c->raster_w[view] * c->raster_h[view];
    SS_CUDA(c, cudaMemcpyAsync(words_out, c->d_raster[view], layer * 8, cudaMemcpyDeviceToHost, c->stream));
    SS_CUDA(c, cudaStreamSynchronize(c->stream));
    return SS_OK;
}

SS_API ss_status ss_read_pixel(ss_ctx* c, uint32_t view, uint32_t x, uint32_t y, uint32_t* instance_out, float* depth_out)
{
    if (!c) return SS_ERR_INVALID;
    if (view >= c->cfg.max_views || !c->d_raster[view] || x >= c->raster_w[view] || y >= c->raster_h[view])
        return c->fail(SS_ERR_INVALID, "ss_read_pixel: view %u pixel (%u, %u) outside its target", view, x, y);
    SS_CUDA(c, cudaSetDevice(c->device));
    unsigned long long w = 0;
    const size_t at = (size_t)y * c->raster_w[view] + x;
    SS_CUDA(c, cudaMemcpyAsync(&w, c->d_raster[view] + at, 8, cudaMemcpyDeviceToHost, c->stream));
    SS_CUDA(c, cudaStreamSynchronize(c->stream));
    if (instance_out) *instance_out = (uint32_t)(w & 0xFFFFFFFFull);
    if (depth_out) {
        const uint32_t bits = (uint32_t)(w >> 32);
        memcpy(depth_out, &bits, 4);
    }
    return SS_OK;
}

SS_API ss_status ss_merge_place(void* cuda_stream, const int32_t* lists, const int64_t* seg_out, const int64_t* seg_src,
                                const int64_t* seg_add, const int64_t* seg_key, uint32_t n_segs, int32_t* out_idx,
                                int64_t* out_keys)
{
    if (n_segs == 0) return SS_OK;
    if (!lists || !seg_out || !seg_src || !seg_add || !seg_key || !out_idx || !out_keys) return SS_ERR_INVALID;
    launch_merge_place(lists, seg_out, seg_src, seg_add, seg_key, n_segs, out_idx, out_keys, static_cast<cudaStream_t>(cuda_stream));
    return cudaGetLastError() == cudaSuccess ? SS_OK : SS_ERR_CUDA;
}

SS_API ss_status ss_export_idx_lists(ss_ctx* c, ss_ipc_handle* handles_out, void** own_pointers_out, uint32_t capacity,
                                     uint32_t* count_out)
{
    if (!c || !handles_out || !own_pointers_out || !count_out) return SS_ERR_INVALID;
    const uint32_t V = c->cfg.max_views, need = (uint32_t)c->slots.size() * V;
    *count_out = need;
    if (capacity < need) return c->fail(SS_ERR_INVALID, "ss_export_idx_lists: %u handles, room for %u", need, capacity);
    static_assert(sizeof(cudaIpcMemHandle_t) == sizeof(ss_ipc_handle), "handle size");
    SS_CUDA(c, cudaSetDevice(c->device));
    for (size_t si = 0; si < c->slots.size(); ++si)
        for (uint32_t v = 0; v < V; ++v) {
            cudaIpcMemHandle_t h;
            SS_CUDA(c, cudaIpcGetMemHandle(&h, c->slots[si].out[v].idx));
            memcpy(&handles_out[si * V + v], &h, sizeof h);
            own_pointers_out[si * V + v] = c->slots[si].out[v].idx;
        }
    return SS_OK;
}

SS_API ss_status ss_export_slot_buffers(ss_ctx* c, ss_ipc_handle* handles_out, uint32_t capacity, uint32_t* count_out)
{
    if (!c || !handles_out || !count_out) return SS_ERR_INVALID;
    const uint32_t V = c->cfg.max_views, F = (uint32_t)c->slots.size(), need = F * V * 3 + F;
    *count_out = need;
    if (capacity < need) return c->fail(SS_ERR_INVALID, "ss_export_slot_buffers: %u handles, room for %u", need, capacity);
    SS_CUDA(c, cudaSetDevice(c->device));
    uint32_t k = 0;
    for (uint32_t si = 0; si < F; ++si)
        for (uint32_t v = 0; v < V; ++v) {
            const void* bufs[3] = {c->slots[si].out[v].idx, c->slots[si].out[v].draws, c->slots[si].out[v].draw_keys};
            for (const void* b : bufs) {
                cudaIpcMemHandle_t h;
                SS_CUDA(c, cudaIpcGetMemHandle(&h, const_cast<void*>(b)));
                memcpy(&handles_out[k++], &h, sizeof h);
            }
        }
    for (uint32_t si = 0; si < F; ++si) {
        cudaIpcMemHandle_t h;
        SS_CUDA(c, cudaIpcGetMemHandle(&h, c->slots[si].counts_dev));
        memcpy(&handles_out[k++], &h, sizeof h);
    }
    return SS_OK;
}

SS_API ss_status ss_ipc_open(int device, const ss_ipc_handle* handle, void** device_ptr_out)
{
    if (!handle || !device_ptr_out) return SS_ERR_INVALID;
    if (cudaSetDevice(device) != cudaSuccess) return SS_ERR_CUDA;
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof h);
    const cudaError_t e = cudaIpcOpenMemHandle(device_ptr_out, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
        ss::set_global_error(cudaGetErrorString(e));
        return SS_ERR_CUDA;
    }
    return SS_OK;
}

SS_API ss_status ss_ipc_close(int device, void* device_ptr)
{
    if (!device_ptr) return SS_OK;
    if (cudaSetDevice(device) != cudaSuccess) return SS_ERR_CUDA;
    return cudaIpcCloseMemHandle(device_ptr) == cudaSuccess ? SS_OK : SS_ERR_CUDA;
}

SS_API ss_status ss_merge_place_peers(void* cuda_stream, const int32_t* const* rank_lists, const int32_t* seg_rank,
                                      const int64_t* seg_out, const int64_t* seg_src, const int64_t* seg_add,
                                      const int64_t* seg_key, uint32_t n_segs, int32_t* out_idx, int64_t* out_keys)
{
    if (n_segs == 0) return SS_OK;
    if (!rank_lists || !seg_rank || !seg_out || !seg_src || !seg_add || !seg_key || !out_idx || !out_keys) return SS_ERR_INVALID;
    launch_merge_place_peers(rank_lists, seg_rank, seg_out, seg_src, seg_add, seg_key, n_segs, out_idx, out_keys,
                             static_cast<cudaStream_t>(cuda_stream));
    return cudaGetLastError() == cudaSuccess ? SS_OK : SS_ERR_CUDA;
}

SS_API ss_status ss_set_instance_entities(ss_ctx* c, uint32_t first, uint32_t count, const int32_t* entity_ids)
{
    if (!c) return SS_ERR_INVALID;
    if (!entity_ids || (uint64_t)first + count > c->n_instances)
        return c->fail(SS_ERR_INVALID, "ss_set_instance_entities: range [%u, +%u) exceeds %u instances", first, count, c->n_instances);
    SS_CUDA(c, cudaSetDevice(c->device));
    if (!c->d_entity) {
        SS_CUDA(c, cudaMalloc(reinterpret_cast<void**>(&c->d_entity), std::max<size_t>((size_t)c->cfg.max_instances * 4, 256)));
        SS_CUDA(c, cudaMemsetAsync(c->d_entity, 0xFF, (size_t)c->cfg.max_instances * 4, c->stream));  // -1: no entity
    }
    if (count) {
        SS_CUDA(c, cudaMemcpyAsync(c->d_entity + first, entity_ids, (size_t)count * 4, cudaMemcpyHostToDevice, c->stream));
        SS_CUDA(c, cudaStreamSynchronize(c->stream));
    }
    return SS_OK;
}

SS_API ss_status ss_select_view(ss_ctx* c, uint32_t slot, uint32_t view, const void** device_out, void* host_out)
{
    ss_outputs o;
    ss_status st = ss_map_outputs(c, slot, view, &o);  // fences the frame: counts are on the host
    if (st != SS_OK) return st;
    if (!c->d_select) SS_CUDA(c, cudaMalloc(reinterpret_cast<void**>(&c->d_select), std::max<size_t>((size_t)c->cap_visible * 52, 256)));
    launch_select_records(reinterpret_cast<const float4*>(o.instance_rows), o.instance_idx, c->d_entity, o.visible, c->d_select,
                          c->stream);
    SS_CUDA(c, cudaGetLastError());
    if (host_out && o.visible)
        SS_CUDA(c, cudaMemcpyAsync(host_out, c->d_select, (size_t)o.visible * 52, cudaMemcpyDeviceToHost, c->stream));
    SS_CUDA(c, cudaStreamSynchronize(c->stream));
    if (device_out) *device_out = c->d_select;
    return SS_OK;
}

SS_API void* ss_stream(ss_ctx* c) { return c ? c->stream : nullptr; }

// ---- standalone sort -------------------------------------------------------------------------

SS_API size_t ss_sort_scratch_bytes(uint64_t n) { return sort_scratch_bytes(n); }

SS_API ss_status ss_sort_pairs(void* cuda_stream, uint64_t* keys_a, uint32_t* vals_a, uint64_t* keys_b,
                               uint32_t* vals_b, uint64_t n, int begin_bit, int end_bit, void* scratch,
                               size_t scratch_bytes, int* sorted_in_b)
{
    if (!keys_a || !vals_a || !keys_b || !vals_b || !scratch || begin_bit < 0 || end_bit > 64 || begin_bit > end_bit ||
        n >= (1ull << 30) || scratch_bytes < sort_scratch_bytes(n)) {
        g_create_error = "ss_sort_pairs: invalid argument (pointers, bit range, n < 2^30, scratch size)";
        return SS_ERR_INVALID;
    }
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    cudaError_t e = sort_pairs_async(keys_a, vals_a, keys_b, vals_b, n, begin_bit, end_bit, scratch, st, nullptr);
    uint32_t sel = 0;
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(&sel, &static_cast<SortControl*>(scratch)->src_sel[sort_num_passes(begin_bit, end_bit)],
                            sizeof sel, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) {
        g_create_error = std::string("ss_sort_pairs: ") + cudaGetErrorString(e);
        return SS_ERR_CUDA;
    }
    if (sorted_in_b) *sorted_in_b = (int)sel;
    return SS_OK;
}

SS_API ss_status ss_sort_pairs_host(int device, const uint64_t* keys_in, const uint32_t* vals_in, uint64_t* keys_out,
                                    uint32_t* vals_out, uint64_t n, int begin_bit, int end_bit)
{
    if ((n && (!keys_in || !vals_in || !keys_out || !vals_out)) || n >= (1ull << 30)) {
        g_create_error = "ss_sort_pairs_host: invalid argument";
        return SS_ERR_INVALID;
    }
    cudaError_t e = cudaSetDevice(device);
    uint64_t *ka = nullptr, *kb = nullptr;
    uint32_t *va = nullptr, *vb = nullptr;
    void* scratch = nullptr;
    const size_t nn = n ? n : 1, sb = sort_scratch_bytes(n);
    if (e == cudaSuccess) e = cudaMalloc(&ka, nn * 8);
    if (e == cudaSuccess) e = cudaMalloc(&kb, nn * 8);
    if (e == cudaSuccess) e = cudaMalloc(&va, nn * 4);
    if (e == cudaSuccess) e = cudaMalloc(&vb, nn * 4);
    if (e == cudaSuccess) e = cudaMalloc(&scratch, sb);
    if (e == cudaSuccess && n) e = cudaMemcpy(ka, keys_in, n * 8, cudaMemcpyHostToDevice);
    if (e == cudaSuccess && n) e = cudaMemcpy(va, vals_in, n * 4, cudaMemcpyHostToDevice);
    ss_status st = SS_OK;
    int in_b = 0;
    if (e == cudaSuccess) st = ss_sort_pairs(nullptr, ka, va, kb, vb, n, begin_bit, end_bit, scratch, sb, &in_b);
    if (e == cudaSuccess && st == SS_OK && n) e = cudaMemcpy(keys_out, in_b ? kb : ka, n * 8, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess && st == SS_OK && n) e = cudaMemcpy(vals_out, in_b ? vb : va, n * 4, cudaMemcpyDeviceToHost);
    cudaFree(ka);
    cudaFree(kb);
    cudaFree(va);
    cudaFree(vb);
    cudaFree(scratch);
    if (e != cudaSuccess) {
        g_create_error = std::string("ss_sort_pairs_host: ") + cudaGetErrorString(e);
        return SS_ERR_CUDA;
    }
    return st;
}

}  // extern "C"
