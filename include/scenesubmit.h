/*
 * libscenesubmit — C ABI of the B200-native scene-submission path.
 *
 * This header is the drop-in boundary.  The reference (PrakarshPanwar/VulkanEngine) has no FFI for
 * this path; the seam is a handful of C++ member functions.  Every entry point below names the
 * reference code it stands in for (paths relative to VulkanCore/src/, "VC/" = VulkanCore/,
 * "PV/" = Platform/Vulkan/).  INTEGRATION.md shows the C++ binding a maintainer would add behind the
 * unchanged Scene / SceneRenderer signatures.
 *
 * Conventions
 *   - plain pointers and sizes only; no C++ or torch types cross the boundary;
 *   - every call returns an ss_status; the library never throws and never aborts the host
 *     (reference convention on this path is a silent skip, VC/Scene/SceneRenderer.cpp:1128-1136);
 *   - host pointers passed to ss_set_* / ss_upload_* are copied (or queued from pinned memory and
 *     fenced) before the call returns: the caller owns them again immediately;
 *   - device buffers returned by ss_map_outputs stay valid until the same frame slot is reused,
 *     i.e. for frames_in_flight-1 further ss_frame calls (VC/Renderer/Renderer.h:28, FramesInFlight=3);
 *   - one context per GPU, single producer thread (VC/Core/Application.cpp:68-69); no global state.
 *
 * All floating point is IEEE binary32.  Matrices are column-major float[16] exactly as glm::mat4
 * lies in memory (m[col*4 + row]); instance rows are TransformData (3 x vec4, row r =
 * (M[0][r], M[1][r], M[2][r], M[3][r]); VC/Core/Components.h:95-98).
 */
#ifndef SCENESUBMIT_H
#define SCENESUBMIT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define SS_API __declspec(dllexport)
#else
#define SS_API __attribute__((visibility("default")))
#endif

#define SS_VERSION_MAJOR 0
#define SS_VERSION_MINOR 1

#define SS_MAX_VIEWS 8u          /* one visibility byte per instance */
#define SS_MAX_LEVELS 64u
#define SS_SLOT_NULL 0xFFFFFFFFu /* instance is dropped (null mesh / null table / zero vertices) */
#define SS_MATERIAL_TESS_BIT 0x80000000u /* material word: bit 31 = HasDisplacementTexture() */

typedef enum ss_status {
    SS_OK = 0,
    SS_ERR_INVALID = 1,   /* bad argument (null pointer, size beyond capacity, malformed table) */
    SS_ERR_CUDA = 2,      /* a CUDA runtime call failed; see ss_last_error */
    SS_ERR_CAPACITY = 3,  /* an output list overflowed its configured capacity */
    SS_ERR_STATE = 4,     /* call order violated (frame before streams were set, bad slot ...) */
    SS_ERR_NOMEM = 5
} ss_status;

typedef enum ss_trig_mode {
    /* sin/cos of the half Euler angles through the explicit polynomial in scenesubmit_dettrig.h:
       the whole chain (matrices, masks, draw order) is bit-identical to the CPU oracle. */
    SS_TRIG_DETERMINISTIC = 0,
    /* CUDA sinf/cosf, as glm::sin/glm::cos call libm (GLM/detail/type_quat.inl:208-217):
       matrices agree with the reference to 1e-6 relative, masks can differ on borderline boxes. */
    SS_TRIG_LIBM = 1
} ss_trig_mode;

typedef enum ss_view_flags {
    SS_VIEW_DEFAULT = 0,
    /* shadow cascades: the shadow pipeline clamps depth (VC/Scene/SceneRenderer.cpp:188), so casters
       in front of the near plane still land in the map: skip the near plane. */
    SS_VIEW_SKIP_NEAR = 1
} ss_view_flags;

typedef enum ss_frame_flags {
    SS_FRAME_DEFAULT = 0,
    /* stage-wise parity: do not run the transform stage, cull/sort/emit from the world rows given
       to ss_upload_world_rows */
    SS_FRAME_WORLD_ROWS_GIVEN = 1,
    /* Batched multi-view work (many cameras over one scene, BASELINE configs[4]): views are processed
       SS_MAX_VIEWS at a time, one ss_frame per batch.  After a first, full frame the transforms have not
       changed, so later batches may keep the world rows of the previous frame (cull + emit only) ... */
    SS_FRAME_REUSE_WORLD = 4,
    /* ... and, the sort keys being static too, its sorted (key, instance) order.  Both flags are ignored
       (the stage runs) when no valid result of a previous frame exists. */
    SS_FRAME_REUSE_SORT = 8
} ss_frame_flags;

/* ss_config.flags */
typedef enum ss_config_flags {
    SS_CONFIG_DEFAULT = 0,
    /* Draw-order-resident layout.  The sort keys are static between ss_set_* calls (in the reference the std::map nodes
       persist across frames too, VC/Scene/SceneRenderer.cpp:1559-1591), so the library sorts ONCE when a static stream
       changes and keeps its per-instance streams in (key, instance) order; a frame is then two streaming passes with no
       sort and no gather.  Transform uploads are re-ordered on the device.  Results are the same lists, bit for bit.
       Applies to TRS-driven instances on an implicit node map with deterministic trig and a key that shares a word
       with the instance index; other scenes silently take the general pipeline. */
    SS_CONFIG_DRAW_ORDER_RESIDENT = 2
} ss_config_flags;

/* which kernels produced a frame (ss_frame_stats.pipeline) */
typedef enum ss_pipeline {
    SS_PIPELINE_GENERAL = 0,  /* transform+cull+key -> sort -> compact + gather, every frame */
    SS_PIPELINE_RESIDENT = 2  /* transform+cull over draw-order streams -> compact (sequential); sorted at ss_set_* time */
} ss_pipeline;

typedef struct ss_ctx ss_ctx;

typedef struct ss_config {
    uint32_t struct_size;          /* sizeof(ss_config), for ABI growth */
    int32_t device;                /* CUDA device ordinal */
    uint32_t max_nodes;            /* transform (entity) stream capacity */
    uint32_t max_instances;        /* instance (entity x submesh) stream capacity */
    uint32_t max_submesh_slots;    /* rows of the global submesh table */
    uint32_t max_views;            /* 1..SS_MAX_VIEWS */
    uint32_t max_draws_per_view;   /* distinct keys per view; 0 = min(max_instances, 1<<20) */
    uint32_t frames_in_flight;     /* output slots; 0 = 3 (VC/Renderer/Renderer.h:28) */
    uint64_t max_visible_per_view; /* compacted list capacity per view; 0 = max_instances */
    uint32_t trig_mode;            /* ss_trig_mode */
    uint32_t flags;                /* ss_config_flags */
} ss_config;

/* One row of the global submesh table: the per-frame facts SubmitMesh reads from Submesh
   (VC/Mesh/Mesh.h:39-46) plus the bounds the reference lacks. */
typedef struct ss_submesh {
    float local[16];        /* Submesh::LocalTransform, glm::mat4 memory; must be affine (row 3 = 0,0,0,1) */
    float aabb_min[3];      /* over Vertex.Position of [BaseVertex, BaseVertex+VertexCount) */
    float aabb_max[3];
    uint32_t index_count;   /* Submesh::IndexCount  -> vkCmdDrawIndexed arg 1 (PV/VulkanRenderer.cpp:628) */
    uint32_t first_index;   /* Submesh::BaseIndex   -> arg 3 */
    int32_t vertex_offset;  /* Submesh::BaseVertex  -> arg 4 */
    uint32_t mesh_rank;     /* rank of Mesh::Handle among the sorted unique mesh handles */
    uint32_t submesh_index; /* index into MeshSource::GetSubmeshes(), MeshKey::SubmeshIndex */
    uint32_t reserved;
} ss_submesh;

typedef struct ss_view {
    float view_proj[16]; /* P*V (VC/Renderer/EditorCamera.h:45) or CascadeLightMatrices[i]
                            (VC/Scene/SceneRenderer.cpp:683), glm::mat4 memory */
    uint32_t flags;      /* ss_view_flags */
} ss_view;

/* Field for field VkDrawIndexedIndirectCommand; meanings from PV/VulkanRenderer.cpp:628. */
typedef struct ss_draw_cmd {
    uint32_t index_count;
    uint32_t instance_count;
    uint32_t first_index;
    int32_t vertex_offset;
    uint32_t first_instance;
} ss_draw_cmd;

/* Device pointers of one (slot, view) result; library-owned. */
typedef struct ss_outputs {
    const float* instance_rows;   /* [visible][12]  TransformData in draw order */
    const uint64_t* keys;         /* [visible]      sort key of each compacted instance */
    const uint32_t* instance_idx; /* [visible]      index into the instance stream (EntityID lookup) */
    const ss_draw_cmd* draws;     /* [draws]        one per run of equal keys, in MeshKey order */
    const uint64_t* draw_keys;    /* [draws]        key of each draw (to bind mesh + material) */
    const uint32_t* counts;       /* device: {visible, draws} of this view */
    /* Every instance's world rows (world_stride floats apart: 16 = row-major 4x4 with row 3 = 0,0,0,1 in the general
       pipeline, 12 = packed TransformData in the resident layout) and visibility byte.  NOT per
       slot: they belong to the most recent ss_frame of the context and are overwritten by the next one (the per-view
       lists above are what stays valid per slot).  General pipeline: entry i = instance i, world_words = NULL.
       Draw-order-resident layout: entry j = the instance named by world_words[j] & ((1 << world_idx_bits) - 1), i.e.
       the arrays lie in (key, instance) order.  ss_download_world / ss_download_masks return instance order either way. */
    const float* world_rows;
    const uint8_t* masks;         /* bit v = visible in view v */
    uint32_t visible;             /* host copies, valid after ss_frame_sync */
    uint32_t n_draws;
    const uint64_t* world_words;  /* resident layout: [instances] sorted words [key | instance]; NULL otherwise */
    uint32_t world_idx_bits;
    uint32_t world_stride;        /* floats per entry of world_rows */
} ss_outputs;

typedef struct ss_frame_stats {
    uint32_t n_instances;
    uint32_t n_views;
    uint32_t visible[SS_MAX_VIEWS]; /* RendererStats::InstanceCount per pass (PV/VulkanRenderer.h:13-17) */
    uint32_t draws[SS_MAX_VIEWS];   /* RendererStats::DrawCalls per pass */
    uint32_t sort_passes_run;       /* radix passes that moved data (others were skipped) */
    uint32_t overflow;              /* non-zero: a list hit its capacity (results truncated) */
    float gpu_ms;                   /* CUDA-event time of the frame on the context stream */
    uint32_t kernel_launches;       /* kernels of this library launched for the frame */
    uint32_t pipeline;              /* ss_pipeline */
} ss_frame_stats;

/* ---- lifecycle ------------------------------------------------------------------------------ */

/* Replaces SceneRenderer construction of the draw-list maps (VC/Scene/SceneRenderer.h:252-261):
   allocates every device stream once, sized from cfg. */
SS_API ss_status ss_create(const ss_config* cfg, ss_ctx** out);
SS_API void ss_destroy(ss_ctx* ctx);
/* Last error text of this context (or of ss_create when ctx is NULL). Never NULL. */
SS_API const char* ss_last_error(const ss_ctx* ctx);
SS_API uint32_t ss_version(void);

/* ---- static streams (scene load, SetActiveScene, UpdateMeshInstanceData) --------------------- */

/* Key layout [tess:1][meshRank:mesh_bits][submeshIndex:submesh_bits][materialRank:material_bits],
   packed right-aligned in a uint64; order-isomorphic to MeshKey::operator<
   (VC/Scene/SceneRenderer.h:64-85: Mesh, then Submesh, then Material) with the tessellated list
   after the static one (VC/Scene/SceneRenderer.cpp:1368-1382).  Sum of widths <= 62: the bit above
   tess marks dropped instances inside the sort so they collect behind every real key (they are
   never visible, so that bit never reaches an output).  Default 22/16/24.  Narrow widths let the
   radix sort skip the constant digits: only ceil((sum+2)/8) passes are even considered. */
SS_API ss_status ss_set_key_layout(ss_ctx* ctx, uint32_t mesh_bits, uint32_t submesh_bits,
                                   uint32_t material_bits);

/* Replaces the per-frame reads of meshSource->GetSubmeshes()[i] (VC/Scene/SceneRenderer.cpp:1132,1150). */
SS_API ss_status ss_set_submesh_table(ss_ctx* ctx, const ss_submesh* table, uint32_t count);

/* Transform hierarchy (new step N1; the reference has no parent component).
   parent[i] < 0 = root, otherwise parent[i] < i and nodes are level-ordered:
   level l occupies [level_offsets[l], level_offsets[l+1]).  parent may be NULL (all roots, 1 level). */
SS_API ss_status ss_set_nodes(ss_ctx* ctx, uint32_t n_nodes, const int32_t* parent,
                              const uint32_t* level_offsets, uint32_t n_levels);

/* Instance stream = what the loops of Scene::OnUpdateGeometry x SubmitMesh enumerate
   (VC/Scene/Scene.cpp:138-146, VC/Scene/SceneRenderer.cpp:1138), flattened in that order.
   node[i]          entity whose transform places the instance; NULL = node_base + i
   submesh_slot[i]  row of the submesh table, or SS_SLOT_NULL to drop (SceneRenderer.cpp:1128-1136)
   material_word[i] materialRank | SS_MATERIAL_TESS_BIT (fallback to base material already resolved
                    host-side, SceneRenderer.cpp:1141) */
SS_API ss_status ss_set_instances(ss_ctx* ctx, uint32_t n_instances, const uint32_t* node,
                                  uint32_t node_base, const uint32_t* submesh_slot,
                                  const uint32_t* material_word);

/* Optional per-instance local AABB stream {min.xyz, max.xyz} (synthetic scenes); NULL = use the
   submesh table's box. */
SS_API ss_status ss_set_instance_aabbs(ss_ctx* ctx, uint32_t first, uint32_t count,
                                       const float* minmax6);

/* ---- per-frame inputs ------------------------------------------------------------------------ */

/* Dirty range of TransformComponent {Translation, Rotation(euler xyz, radians), Scale}, 36 B each,
   exactly the entt pool's memory (VC/Core/Components.h:13-15).  Replaces the GetTransform() call per
   entity (VC/Scene/Scene.cpp:145).  Pageable memory is staged before the call returns; PINNED memory
   is read asynchronously by the copy engine and must stay unchanged until ss_flush_uploads or the
   ss_frame_sync of the next frame. */
SS_API ss_status ss_upload_trs(ss_ctx* ctx, uint32_t first, uint32_t count, const float* trs9);

/* The same for a scattered set of nodes: what a transform producer that touches few entities
   leaves behind — the physics write-back loop (Platform/JoltPhysics/JoltPhysicsWorld.cpp:112-127, translation and
   rotation of every rigid body) or a gizmo edit (VulkanEditor/src/EditorLayer.cpp:593-603).  node_index[k] names the
   node trs9[k] belongs to; a node named twice keeps the later record.  The records cross the bus packed
   (40 B per node) and a kernel scatters them into the TRS stream.  A strictly ascending node list in PINNED memory is
   read by the copy engine where it lies (keep both arrays unchanged until ss_flush_uploads or the next frame's
   ss_frame_sync); anything else is staged before the call returns. */
SS_API ss_status ss_upload_trs_sparse(ss_ctx* ctx, uint32_t count, const uint32_t* node_index, const float* trs9);

/* The physics write-back itself (Platform/JoltPhysics/JoltPhysicsWorld.cpp:112-127): pose7[k] = body position xyz and
   rotation quaternion xyzw of node node_index[k].  A kernel turns the quaternion into Euler angles the way
   JPH::Quat::GetEulerAngles does (vendor/JoltPhysics/Jolt/Math/Quat.inl:185-203) and writes Translation and Rotation of
   the node's TransformComponent; Scale stays.  32 B per body over the bus instead of the 40 B of a finished record, and
   no trigonometry on the host.  Angles agree with Jolt's to 1e-6 (Jolt's ATan2/ASin are its own polynomials). */
SS_API ss_status ss_upload_pose_sparse(ss_ctx* ctx, uint32_t count, const uint32_t* node_index, const float* pose7);

/* World rows (TransformData, 48 B each) of EXTERNAL nodes: nodes numbered past the last level given to
   ss_set_nodes are never composed from TRS.  This is how the ad-hoc submitters that hand the renderer
   a finished glm::mat4 keep working: SceneRenderer::SubmitMesh / SubmitSelectedMesh
   (VC/Scene/SceneRenderer.h:32-33) when called outside Scene::OnUpdateGeometry, and SubmitPhysicsMesh
   (VC/Scene/SceneRenderer.cpp:1211-1245, caller Platform/JoltPhysics/JoltDebugRenderer.cpp:63).
   Instances name such a node through the explicit node map of ss_set_instances. */
SS_API ss_status ss_upload_node_world_rows(ss_ctx* ctx, uint32_t first, uint32_t count,
                                           const float* rows12);

/* Wait until every ss_upload_* issued so far has left host memory. */
SS_API ss_status ss_flush_uploads(ss_ctx* ctx);

/* Stage-wise parity input: world rows per instance (see SS_FRAME_WORLD_ROWS_GIVEN). */
SS_API ss_status ss_upload_world_rows(ss_ctx* ctx, uint32_t first, uint32_t count,
                                      const float* rows12);

/* One frame: transform -> cull -> key -> sort -> compact -> emit, asynchronous on the context
   stream.  Replaces the body of ShadowPass/GeometryPass submission: Scene::OnUpdateGeometry
   (VC/Scene/Scene.cpp:135-147) + SubmitMesh (VC/Scene/SceneRenderer.cpp:1124-1167) for every view.
   *slot_out receives the output slot (frame number modulo frames_in_flight). */
SS_API ss_status ss_frame(ss_ctx* ctx, const ss_view* views, uint32_t n_views, uint32_t flags,
                          uint32_t* slot_out);

/* Wait for the slot's frame and fetch its counts (the fence the render thread waits on,
   VC/Renderer/RenderThread.cpp:71-80). */
SS_API ss_status ss_frame_sync(ss_ctx* ctx, uint32_t slot, ss_frame_stats* stats_out);

/* Device time of the slot's frame by stage, CUDA events on the context stream:
   {clear + hierarchy levels (+ re-ordering of changed streams in the resident layout), transform+cull+key, sort,
   compact+emit, draws + count read-back}.
   The counterpart of the reference's per-pass timestamp queries
   (PV/VulkanRenderCommandBuffer.cpp:150-172). */
SS_API ss_status ss_frame_stage_ms(ss_ctx* ctx, uint32_t slot, float ms_out[5]);

/* ---- outputs --------------------------------------------------------------------------------- */

/* Replaces `for (auto& [mk, dc] : m_MeshDrawList)` (VC/Scene/SceneRenderer.cpp:1336,1368,1381,1429):
   device pointers for one view of a finished slot. */
SS_API ss_status ss_map_outputs(ss_ctx* ctx, uint32_t slot, uint32_t view, ss_outputs* out);

/* Host copies for headless consumers and parity tests; any pointer may be NULL. Sizes in elements
   must be at least the counts reported by ss_frame_sync. */
SS_API ss_status ss_download_view(ss_ctx* ctx, uint32_t slot, uint32_t view, float* rows12,
                                  uint64_t* keys, uint32_t* instance_idx, ss_draw_cmd* draws,
                                  uint64_t* draw_keys);
SS_API ss_status ss_download_world(ss_ctx* ctx, uint32_t first, uint32_t count, float* rows12);
SS_API ss_status ss_download_masks(ss_ctx* ctx, uint32_t first, uint32_t count, uint8_t* masks);

/* ---- selection pass ------------------------------------------------------------ */

/* Entity id of every instance: what Scene::OnSelectGeometry hands to SubmitSelectedMesh per entity
   (VC/Scene/Scene.cpp:149-160, (int)(entt::entity)ent).  Optional; without it ss_select_view reports -1. */
SS_API ss_status ss_set_instance_entities(ss_ctx* ctx, uint32_t first, uint32_t count, const int32_t* entity_ids);
/* The instance stream of the selection pass for one view of a finished slot: SelectTransformData
   {vec4 MRow[3]; int EntityID}, 52 B per visible instance, in draw order (VC/Core/Components.h:100-104,
   VC/Scene/SceneRenderer.cpp:1169-1204), so the pass's vertex stage needs no table on the host.  device_out receives a
   library-owned device buffer (valid until the next ss_select_view); host_out, when not NULL, gets a copy of
   `visible` records (capacity in records is the caller's business: ss_frame_sync reported the count). */
SS_API ss_status ss_select_view(ss_ctx* ctx, uint32_t slot, uint32_t view, const void** device_out, void* host_out);

/* The CUDA stream (cudaStream_t) frames run on, so a consumer can order its own work after them. */
SS_API void* ss_stream(ss_ctx* ctx);

/* ---- consuming the lists: depth + id rasterisation ------------------------------ */

/* What the reference does with the instance buffer and the draws next: ShadowPass renders depth per cascade
   (assets/shaders/ShadowDepth.vert/.geom: clip = LightSpaceMatrices[c] * transform * position, DepthClamp,
   VC/Scene/SceneRenderer.cpp:188, :1266-1310) and SelectionPass renders the entity id per pixel and reads ONE
   pixel back (assets/shaders/CoreEditor.vert/.frag, VC/Scene/SceneRenderer.cpp:1421-1457).  Here both are one
   CUDA pass over (compacted instance, triangle): every covered sample does a 64-bit atomic min of
   [depth float bits : 32 | instance index : 32], so the result is independent of submission order.
   Vulkan conventions: y down, depth z/w in [0, 1], samples at pixel centres, an edge belongs to one side.
   The reference's own rasteriser is the Vulkan device (unavailable on this box); the checker is the CPU
   restatement sso_raster_view in oracle/. */
enum {
    SS_RASTER_DEFAULT = 0,
    SS_RASTER_DEPTH_CLAMP = 1 /* clamp z/w to [0, 1] instead of discarding (the shadow pipeline's DepthClamp) */
};
#define SS_RASTER_EMPTY 0xFFFFFFFFFFFFFFFFull

/* Vertex positions (xyz floats) and 32-bit indices the submesh table's first_index / index_count /
   vertex_offset refer to (MeshSource::m_Vertices / m_Indices, VC/Asset/MeshImporter.cpp:97-158). Host pointers. */
SS_API ss_status ss_set_geometry(ss_ctx* ctx, const float* positions_xyz, uint32_t n_vertices,
                                 const uint32_t* indices, uint32_t n_indices);
/* Rasterises view `view` of the frame in `slot` with that view's view_proj into the library's target of that
   view (width x height words, cleared to SS_RASTER_EMPTY first; every view keeps its own size).  Asynchronous on
   the frame stream.  The pass draws front to back against a coarse depth buffer it rebuilds as it goes (DESIGN.md
   section 7); the image is the same whatever the order.  The first call allocates the pass's scratch once per
   context: 48 bytes per instance of the context (sort words, triangle prefix, per-row bounds) + the sort's own
   scratch + 24 bytes per draw; the coarse depth (16 bytes per 64 samples) grows with the largest target. */
SS_API ss_status ss_raster_view(ss_ctx* ctx, uint32_t slot, uint32_t view, uint32_t width, uint32_t height,
                                uint32_t flags);
/* Device pointer of view `view`'s target, at the size of its last ss_raster_view (fences the stream), or a copy
   to the host. */
SS_API ss_status ss_map_raster(ss_ctx* ctx, uint32_t view, const uint64_t** words_out);
SS_API ss_status ss_download_raster(ss_ctx* ctx, uint32_t view, uint64_t* words_out);
/* One sample: the hovered-entity read-back of SelectionPass.  instance_out = 0xFFFFFFFF where nothing was drawn. */
SS_API ss_status ss_read_pixel(ss_ctx* ctx, uint32_t view, uint32_t x, uint32_t y, uint32_t* instance_out,
                               float* depth_out);

/* ---- merged draw list of several shards (multi-GPU) ------------------------------------ */

/* Placement step of the merge: after the ranks' compacted instance lists have been all-gathered, the merged list
   is "run k: rank 0's elements, rank 1's, ..." (rank order = input order, so this IS the single-GPU stable order).
   The caller describes that as segments in output order; this writes, for every output position e in segment s,
       out_idx[e]  = lists[seg_src[s] + (e - seg_out[s])] + seg_add[s]      (global instance number)
       out_keys[e] = seg_key[s]
   All pointers are device pointers; seg_out has n_segs + 1 entries (exclusive prefix of the segment lengths, last =
   total).  Asynchronous on `cuda_stream`.  No counterpart in the reference (it has one list on one CPU). */
SS_API ss_status ss_merge_place(void* cuda_stream, const int32_t* lists, const int64_t* seg_out, const int64_t* seg_src,
                                const int64_t* seg_add, const int64_t* seg_key, uint32_t n_segs, int32_t* out_idx,
                                int64_t* out_keys);

/* The same placement reading the ranks' lists where they lie — in each rank's own output buffers, over NVLink peer
   access — instead of from an all-gathered copy: the exchange and the placement are one kernel.
   ss_export_idx_lists: CUDA IPC handles of this context's instance-index output buffers, [frames_in_flight][max_views]
   in that order (each 64 bytes), and the buffers' addresses in this process; another process opens one with ss_ipc_open (once; the mapping stays valid until
   ss_ipc_close or the exporter's ss_destroy) and may read it after the exporter's ss_frame_sync of that slot and
   before the exporter reuses the slot.  rank_lists: device array of one pointer per rank (this rank's own buffer
   as it is, the others' as opened); seg_rank[s] picks the list, seg_src[s] is the offset inside it. */
typedef struct ss_ipc_handle {
    unsigned char bytes[64];
} ss_ipc_handle;
SS_API ss_status ss_export_idx_lists(ss_ctx* ctx, ss_ipc_handle* handles_out, void** own_pointers_out, uint32_t capacity,
                                     uint32_t* count_out);
SS_API ss_status ss_ipc_open(int device, const ss_ipc_handle* handle, void** device_ptr_out);
SS_API ss_status ss_ipc_close(int device, void* device_ptr);
SS_API ss_status ss_merge_place_peers(void* cuda_stream, const int32_t* const* rank_lists, const int32_t* seg_rank,
                                      const int64_t* seg_out, const int64_t* seg_src, const int64_t* seg_add,
                                      const int64_t* seg_key, uint32_t n_segs, int32_t* out_idx, int64_t* out_keys);

/* The whole merge as one library call (no Python, no framework tensors): every rank runs it and gets the merged list of
   every view, equal to what ONE context holding all shards would have produced (the single-GPU stable order).
   Peers: for every (view, rank) the device pointers of that rank's outputs of one finished frame slot as THIS process
   sees them: its own from ss_map_outputs, the others' opened once from ss_export_slot_buffers handles (ss_ipc_open),
   or, for several contexts in one process, every rank's ss_map_outputs pointers directly.  instance_base is the global
   number of the rank's instance 0.  The caller makes sure (a barrier of its own) that every rank's frame of the slot
   has completed and that no rank reuses the slot while another still merges from it.  Four kernels on cuda_stream for
   all views together: fetch of the ranks' draws (KBs), a K-way merge of the sorted draw keys by bisection, segment
   prefix + merged indirect draws, and a placement kernel that reads every rank's instance list in place over NVLink
   (plain per-thread loads).  No host synchronisation until ss_merge_map reads the sizes.
   No counterpart in the reference (one list on one CPU, VC/Scene/SceneRenderer.cpp:1336,1368,1381). */
typedef struct ss_merge ss_merge;
typedef struct ss_merge_peer {
    const uint32_t* idx;          /* ss_outputs.instance_idx of the rank's view */
    const ss_draw_cmd* draws;     /* ss_outputs.draws */
    const uint64_t* draw_keys;    /* ss_outputs.draw_keys */
    const uint32_t* counts;       /* ss_outputs.counts: {visible, draws} on the device */
    uint64_t instance_base;
} ss_merge_peer;
typedef struct ss_merged_view {
    const uint32_t* instance_idx; /* [visible] global instance numbers, draw order */
    const uint64_t* keys;         /* [visible] */
    const ss_draw_cmd* draws;     /* [n_draws] one per distinct key over all ranks */
    const uint64_t* draw_keys;    /* [n_draws] */
    uint32_t visible, n_draws;    /* host copies */
} ss_merged_view;
SS_API ss_status ss_merge_create(int device, uint32_t n_ranks, uint32_t n_views, uint32_t max_draws_per_rank,
                                 uint64_t max_visible_per_view, ss_merge** out);
SS_API void ss_merge_destroy(ss_merge* m);
/* peers: host array [n_views][n_ranks]; asynchronous on cuda_stream */
SS_API ss_status ss_merge_views(ss_merge* m, void* cuda_stream, const ss_merge_peer* peers);
/* CUDA-event times of the last ss_merge_views' four kernels: fetch, rank, segments, placement (waits for them) */
SS_API ss_status ss_merge_stage_ms(ss_merge* m, float ms_out[4]);
/* waits for cuda_stream, returns the view's merged buffers (library-owned, valid until the next ss_merge_views) */
SS_API ss_status ss_merge_map(ss_merge* m, void* cuda_stream, uint32_t view, ss_merged_view* out);
/* host copies of a merged view (any pointer may be NULL; sizes from a first call with NULL pointers) */
SS_API ss_status ss_merge_download(ss_merge* m, void* cuda_stream, uint32_t view, uint32_t* instance_idx, uint64_t* keys,
                                   ss_draw_cmd* draws, uint64_t* draw_keys, uint32_t* visible_out, uint32_t* draws_out);
SS_API const char* ss_merge_last_error(const ss_merge* m);
/* CUDA IPC handles of every buffer a peer needs of this context, for ss_ipc_open in another process:
   [frames_in_flight][max_views][3] in the order idx, draws, draw_keys, then [frames_in_flight] counts blocks
   (counts of view v of a slot = block + 2 * v words).  count_out = handles written. */
SS_API ss_status ss_export_slot_buffers(ss_ctx* ctx, ss_ipc_handle* handles_out, uint32_t capacity, uint32_t* count_out);

/* ---- the sort on its own ---------------------------------------------------------------------- */

/* Stable LSD radix sort (onesweep) of (uint64 key, uint32 value) pairs on bits [begin_bit, end_bit).
   Device pointers; a/b are ping-pong buffers of n elements; scratch from ss_sort_scratch_bytes(n).
   The grouping std::map<MeshKey,...> does on the CPU (VC/Scene/SceneRenderer.h:252-253) is this sort.
   On return *sorted_in_b (host) tells which buffer pair holds the result.  Synchronous w.r.t. the
   host only in that it reads one flag word back after the last pass. */
SS_API size_t ss_sort_scratch_bytes(uint64_t n);
SS_API ss_status ss_sort_pairs(void* cuda_stream, uint64_t* keys_a, uint32_t* vals_a,
                               uint64_t* keys_b, uint32_t* vals_b, uint64_t n, int begin_bit,
                               int end_bit, void* scratch, size_t scratch_bytes, int* sorted_in_b);
/* Convenience for tests: host arrays in, host arrays out; allocates and frees device memory. */
SS_API ss_status ss_sort_pairs_host(int device, const uint64_t* keys_in, const uint32_t* vals_in,
                                    uint64_t* keys_out, uint32_t* vals_out, uint64_t n,
                                    int begin_bit, int end_bit);

/* The same sort for callers that must not allocate.  No counterpart in the reference: it exists for
   hosts that cannot call cudaMalloc at the point where they need a sort, such as a render thread
   working inside a memory budget fixed at start-up, a plug-in under an application that owns the
   device allocator, or an engine that has already handed all of HBM to its own pools.  The scratch
   that ss_sort_pairs takes from the caller is replaced by a fixed array inside the library image
   (about 1 MiB on every GPU that loads the library, held until the process exits), so n is limited
   to SS_SORT_STATIC_MAX_N elements; a larger n returns SS_ERR_CAPACITY and ss_last_error(NULL) says
   why.  The call itself performs no cudaMalloc/cudaFree and uses no allocating library routine; the
   caller still provides both ping-pong buffer pairs.  The first call on each device looks up the
   array's device address (once per device, with that device current): make that first call before
   any stream capture begins.  Every call ends by copying one word (which buffer holds the result)
   back to the host and waits for the stream, so it returns with the sort finished.  cuda_stream must
   belong to the device that is current when the call is made (checked; SS_ERR_INVALID otherwise), and
   the looked-up address is remembered per device ordinal for the life of the process: it does not
   survive a cudaDeviceReset.  All calls on one device share the one array, which still holds the
   previous call's state; each call clears the words it needs before it sorts, so two calls that
   overlap on the GPU (different streams, different threads) would race: keep them ordered. */
#define SS_SORT_STATIC_MAX_N (1u << 20)
SS_API ss_status ss_sort_pairs_static(void* cuda_stream, uint64_t* keys_a, uint32_t* vals_a,
                                      uint64_t* keys_b, uint32_t* vals_b, uint64_t n, int begin_bit,
                                      int end_bit, int* sorted_in_b);

/* ---- light submission ----------------------------------------------------------------- */

/* Scene::UpdateLightsBuffer and Scene::OnUpdateLights (VC/Scene/Scene.cpp:167-264) for lights whose TransformComponent
   lives in the library's TRS stream (after ss_upload_pose_sparse the host no longer holds the Euler angles): a kernel
   reads Translation / Rotation of each light's node and fills the uniform blocks the shaders read, byte for byte:
     ub_point_out  UBPointLights  (VC/Core/Components.h:229-234): int LightCount, vec3 pad, PointLightComponent[10]
                   {vec4(Translation, 1), Color, Radius, Falloff}, 48 B each (alignas 16)                  = 496 B
     ub_spot_out   UBSpotLights   (:236-241): int LightCount, vec3 pad, SpotLightComponent[10]
                   {vec4(Translation, 1), Color, normalize(Rotation), InnerCutoff, OuterCutoff, Radius, Falloff}, 64 B = 656 B
     positions_out [n_point + n_spot][4]: vec4(Translation, Radius), point lights first (OnUpdateLights, :167-198)
   Lights are given in view order (the host mirror's Registry::view); at most 10 of each kind, as in the reference's
   arrays.  Any output may be NULL.  Synchronous (a few hundred bytes). */
typedef struct ss_point_light {
    uint32_t node;
    float color[4];
    float radius, falloff;
} ss_point_light;
typedef struct ss_spot_light {
    uint32_t node;
    float color[4];
    float inner_cutoff, outer_cutoff, radius, falloff;
} ss_spot_light;
#define SS_MAX_LIGHTS 10u
#define SS_UB_POINT_LIGHTS_BYTES 496u
#define SS_UB_SPOT_LIGHTS_BYTES 656u
SS_API ss_status ss_pack_lights(ss_ctx* ctx, const ss_point_light* point, uint32_t n_point, const ss_spot_light* spot,
                                uint32_t n_spot, void* ub_point_out, void* ub_spot_out, float* positions_out);

/* ---- host-side view matrices (stay on the host) --------------------------------- */

/* EditorCamera::GetViewProjection for a camera given by position + pitch/yaw
   (VC/Renderer/EditorCamera.cpp:15-33, EditorCamera.h:45): P = perspectiveRH_ZO, V = inverse(T*R). */
SS_API ss_status ss_camera_view_proj(float fov_y, float aspect, float z_near, float z_far,
                                     const float position[3], float pitch, float yaw,
                                     float view_out[16], float proj_out[16], float view_proj_out[16]);

/* SceneRenderer::UpdateCascadeMap (VC/Scene/SceneRenderer.cpp:579-687) with CascadeOffset = 0:
   4 split depths (stored negated, :682) and 4 light matrices. */
SS_API ss_status ss_cascade_matrices(const float view_proj[16], float z_near, float z_far,
                                     float split_lambda, float near_offset, float far_offset,
                                     const float light_dir[3], float splits_out[4],
                                     float light_matrices_out[4][16]);

/* The same with CascadeOffset set (VC/Scene/SceneRenderer.cpp:662-679, an ImGui toggle at :987-1000): each cascade's
   projection is shifted so that CascadeOrigin projects onto the texel grid of a map_size x map_size shadow map
   (CSMSettings::MapSize, SceneRenderer.h:179).  cascade_origin == NULL gives ss_cascade_matrices. */
SS_API ss_status ss_cascade_matrices_snapped(const float view_proj[16], float z_near, float z_far,
                                             float split_lambda, float near_offset, float far_offset,
                                             const float light_dir[3], const float cascade_origin[3], float map_size,
                                             float splits_out[4], float light_matrices_out[4][16]);

/* The same for many cameras at once, on the GPU (batched multi-camera work needs a cascade set per
   camera).  view_proj: [n_cameras][16], light_matrices_out: [n_cameras][4][16], host pointers; the split distances
   depend on the clip range only and come back once.  The per-camera arithmetic is the text ss_cascade_matrices
   compiles for the host, so the matrices are the same bits. */
SS_API ss_status ss_cascade_matrices_batch(int device, const float* view_proj, uint32_t n_cameras, float z_near,
                                           float z_far, float split_lambda, float near_offset, float far_offset,
                                           const float light_dir[3], float splits_out[4],
                                           float* light_matrices_out);

#ifdef __cplusplus
}
#endif
#endif /* SCENESUBMIT_H */
