// merge_cli: the merged draw list of several entity-wise shards from a C++ host, through the C ABI only (no Python, no
// CUDA headers).  G contexts on one GPU stand in for G ranks (with one process per GPU the other ranks' buffers are
// opened with ss_export_slot_buffers / ss_ipc_open instead, the ss_merge_views call is the same); the merged lists are
// compared with ONE context that holds the whole scene.
//   merge_cli.bin <instances> <shards> [resident]
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../include/scenesubmit.h"

namespace {

struct Rng {  // splitmix64
    uint64_t s;
    uint64_t next() { uint64_t z = (s += 0x9E3779B97F4A7C15ull); z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; return z ^ (z >> 31); }
    float uni(float lo, float hi) { return lo + (hi - lo) * (float)((next() >> 40) * (1.0 / 16777216.0)); }
};

struct SceneData {
    std::vector<float> trs, box;
    std::vector<uint32_t> slot, mat;
};

#define CK(call)                                                                         \
    do {                                                                                 \
        ss_status st__ = (call);                                                         \
        if (st__ != SS_OK) {                                                             \
            std::fprintf(stderr, "%s -> %d: %s\n", #call, (int)st__, ss_last_error(ctx)); \
            return 2;                                                                    \
        }                                                                                \
    } while (0)

int make_ctx(const SceneData& sc, size_t lo, size_t hi, const std::vector<ss_submesh>& table, uint32_t flags, ss_ctx** out)
{
    const uint32_t n = (uint32_t)(hi - lo);
    ss_config cfg{};
    cfg.struct_size = sizeof cfg;
    cfg.device = 0;
    cfg.max_nodes = n;
    cfg.max_instances = n;
    cfg.max_submesh_slots = (uint32_t)table.size();
    cfg.max_views = 5;
    cfg.max_draws_per_view = 16384;
    cfg.flags = flags;
    ss_ctx* ctx = nullptr;
    CK(ss_create(&cfg, &ctx));
    CK(ss_set_key_layout(ctx, 8, 4, 8));
    CK(ss_set_submesh_table(ctx, table.data(), (uint32_t)table.size()));
    CK(ss_set_nodes(ctx, n, nullptr, nullptr, 0));
    CK(ss_set_instances(ctx, n, nullptr, 0, sc.slot.data() + lo, sc.mat.data() + lo));
    CK(ss_set_instance_aabbs(ctx, 0, n, sc.box.data() + lo * 6));
    CK(ss_upload_trs(ctx, 0, n, sc.trs.data() + lo * 9));
    CK(ss_flush_uploads(ctx));
    *out = ctx;
    return 0;
}

}  // namespace

int main(int argc, char** argv)
{
    const size_t n = argc > 1 ? (size_t)std::atoll(argv[1]) : 200000;
    const uint32_t shards = argc > 2 ? (uint32_t)std::atoi(argv[2]) : 3;
    const uint32_t flags = (argc > 3 && !std::strcmp(argv[3], "resident")) ? SS_CONFIG_DRAW_ORDER_RESIDENT : 0;
    if (n == 0 || shards == 0 || shards > 16) return 1;
    // a small procedural scene: 64 submesh rows over 16 meshes, 50 materials
    std::vector<ss_submesh> table(64);
    for (uint32_t i = 0; i < table.size(); ++i) {
        ss_submesh& s = table[i];
        std::memset(&s, 0, sizeof s);
        s.local[0] = s.local[5] = s.local[10] = s.local[15] = 1.0f;
        s.local[12] = 0.25f * (float)(i % 5);
        for (int k = 0; k < 3; ++k) { s.aabb_min[k] = -1.0f; s.aabb_max[k] = 1.0f; }
        s.index_count = 36 + 6 * i;
        s.first_index = 100 * i;
        s.vertex_offset = (int32_t)(10 * i);
        s.mesh_rank = i / 4;
        s.submesh_index = i % 4;
    }
    Rng rng{12345};
    SceneData sc;
    sc.trs.resize(n * 9); sc.box.resize(n * 6); sc.slot.resize(n); sc.mat.resize(n);
    for (size_t i = 0; i < n; ++i) {
        float* t = &sc.trs[i * 9];
        for (int k = 0; k < 3; ++k) t[k] = rng.uni(-500.f, 500.f);
        for (int k = 3; k < 6; ++k) t[k] = rng.uni(-3.14f, 3.14f);
        for (int k = 6; k < 9; ++k) t[k] = rng.uni(0.5f, 2.0f);
        float* b = &sc.box[i * 6];
        for (int k = 0; k < 3; ++k) { const float c = rng.uni(-0.5f, 0.5f), e = rng.uni(0.25f, 2.0f); b[k] = c - e; b[3 + k] = c + e; }
        sc.slot[i] = (rng.next() % 100 == 0) ? SS_SLOT_NULL : (uint32_t)(rng.next() % table.size());
        sc.mat[i] = (uint32_t)(rng.next() % 50) | ((rng.next() % 20 == 0) ? SS_MATERIAL_TESS_BIT : 0u);
    }
    // camera + 4 cascades, the reference's defaults (EditorLayer.cpp:75, SceneRenderer.h:173-181)
    ss_view views[5];
    std::memset(views, 0, sizeof views);
    {
        const float pos[3] = {0.f, 0.f, 600.f}, light[3] = {1.f, -1.f, 1.f};
        float v[16], p[16], splits[4], lm[4][16];
        if (ss_camera_view_proj(0.7853982f, 1.635005f, 0.1f, 1000.f, pos, 0.f, 0.f, v, p, views[0].view_proj) != SS_OK) return 2;
        if (ss_cascade_matrices(views[0].view_proj, 0.1f, 1000.f, 0.92f, -75.f, 75.f, light, splits, lm) != SS_OK) return 2;
        for (int c = 0; c < 4; ++c) { std::memcpy(views[1 + c].view_proj, lm[c], 64); views[1 + c].flags = SS_VIEW_SKIP_NEAR; }
    }
    ss_ctx* ctx = nullptr;
    if (int rc = make_ctx(sc, 0, n, table, flags, &ctx)) return rc;
    ss_ctx* whole = ctx;
    uint32_t slot_w = 0;
    ss_frame_stats st_w;
    CK(ss_frame(whole, views, 5, 0, &slot_w));
    CK(ss_frame_sync(whole, slot_w, &st_w));

    std::vector<ss_ctx*> rank(shards);
    std::vector<uint32_t> slot(shards);
    std::vector<size_t> lo(shards + 1);
    for (uint32_t r = 0; r <= shards; ++r) lo[r] = n * r / shards;
    for (uint32_t r = 0; r < shards; ++r) {
        if (int rc = make_ctx(sc, lo[r], lo[r + 1], table, flags, &rank[r])) return rc;
        ctx = rank[r];
        ss_frame_stats st;
        CK(ss_frame(ctx, views, 5, 0, &slot[r]));
        CK(ss_frame_sync(ctx, slot[r], &st));
    }
    std::vector<ss_merge_peer> peers(5 * shards);
    for (uint32_t v = 0; v < 5; ++v)
        for (uint32_t r = 0; r < shards; ++r) {
            ctx = rank[r];
            ss_outputs o;
            CK(ss_map_outputs(ctx, slot[r], v, &o));
            peers[v * shards + r] = {o.instance_idx, o.draws, o.draw_keys, o.counts, (uint64_t)lo[r]};
        }
    ss_merge* mg = nullptr;
    if (ss_merge_create(0, shards, 5, 16384, n, &mg) != SS_OK) { std::fprintf(stderr, "ss_merge_create: %s\n", ss_last_error(nullptr)); return 2; }
    if (ss_merge_views(mg, nullptr, peers.data()) != SS_OK) { std::fprintf(stderr, "ss_merge_views: %s\n", ss_merge_last_error(mg)); return 2; }
    bool same = true;
    for (uint32_t v = 0; v < 5; ++v) {
        const uint32_t m = st_w.visible[v], d = st_w.draws[v];
        std::vector<uint32_t> idx_w(m), idx_m(m);
        std::vector<uint64_t> key_w(m), key_m(m), dk_w(d), dk_m(d);
        std::vector<ss_draw_cmd> dr_w(d), dr_m(d);
        ctx = whole;
        CK(ss_download_view(whole, slot_w, v, nullptr, key_w.data(), idx_w.data(), dr_w.data(), dk_w.data()));
        uint32_t vis = 0, nd = 0;
        if (ss_merge_download(mg, nullptr, v, nullptr, nullptr, nullptr, nullptr, &vis, &nd) != SS_OK) return 2;
        if (vis != m || nd != d) { same = false; std::printf("view %u: merged %u instances / %u draws, one context %u / %u\n", v, vis, nd, m, d); continue; }
        if (ss_merge_download(mg, nullptr, v, idx_m.data(), key_m.data(), dr_m.data(), dk_m.data(), nullptr, nullptr) != SS_OK) return 2;
        const bool ok = idx_w == idx_m && key_w == key_m && dk_w == dk_m && !std::memcmp(dr_w.data(), dr_m.data(), d * sizeof(ss_draw_cmd));
        same = same && ok;
        std::printf("view %u: %u instances, %u draws: %s\n", v, m, d, ok ? "merged == one context" : "DIFFERENT");
    }
    ss_merge_destroy(mg);
    for (ss_ctx* r : rank) ss_destroy(r);
    ss_destroy(whole);
    std::printf("%s\n", same ? "merge identical" : "merge MISMATCH");
    return same ? 0 : 1;
}
