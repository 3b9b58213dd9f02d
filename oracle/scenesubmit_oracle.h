/*
 * CPU oracle for libscenesubmit — TEST INFRASTRUCTURE, NOT PRODUCT.
 *
 * A plain-C restatement of the reference's scene-submission arithmetic (paths below are relative to
 * /root/reference/VulkanCore/src/VulkanCore unless prefixed GLM/ = VulkanCore/vendor/glm/glm) plus
 * the specified extensions of SURVEY.md §8a N1-N6.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load it; the product (libscenesubmit.so) never
 * links or calls it.
 *
 * Parity status: the reference holds NO tests, golden vectors or fixtures for this path
 * (SURVEY.md §4).  The restatement is pinned instead against the reference's own arithmetic run
 * here: oracle/ref_glm_entt.cpp compiles the vendored GLM 0.9.9.9 + entt 3.15.0 headers where they
 * lie under /root/reference into oracle/_ref/libss_ref.so; tests/test_oracle_vs_ref.py checks the
 * two bit-for-bit, and tests/golden/ holds vectors generated from that build
 * (tests/golden/make_golden.py).  Steps N1-N6 have no reference counterpart: for those the oracle
 * is the definition ("parity unpinned" in the strict sense).
 */
#ifndef SCENESUBMIT_ORACLE_H
#define SCENESUBMIT_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SSO_SLOT_NULL 0xFFFFFFFFu
#define SSO_TESS_BIT 0x80000000u
#define SSO_TRIG_DETERMINISTIC 0
#define SSO_TRIG_LIBM 1
#define SSO_VIEW_SKIP_NEAR 1u

/* same memory layout as ss_submesh (include/scenesubmit.h); restated, not included, so that the
   oracle does not depend on the product's header */
typedef struct sso_submesh {
    float local[16];
    float aabb_min[3];
    float aabb_max[3];
    uint32_t index_count;
    uint32_t first_index;
    int32_t vertex_offset;
    uint32_t mesh_rank;
    uint32_t submesh_index;
    uint32_t reserved;
} sso_submesh;

typedef struct sso_view {
    float view_proj[16];
    uint32_t flags;
} sso_view;

typedef struct sso_draw_cmd {
    uint32_t index_count, instance_count, first_index;
    int32_t vertex_offset;
    uint32_t first_instance;
} sso_draw_cmd;

typedef struct sso_scene {
    uint32_t n_nodes;
    const float* trs;          /* [n_nodes][9] T, R(euler), S */
    const int32_t* parent;     /* [n_nodes] or NULL; parent[i] < i */
    uint32_t n_instances;
    const uint32_t* node;      /* [n_instances] or NULL -> node_base + i */
    uint32_t node_base;
    const uint32_t* slot;      /* [n_instances] */
    const uint32_t* matword;   /* [n_instances] */
    const float* inst_aabb;    /* [n_instances][6] or NULL -> submesh table box */
    uint32_t n_slots;
    const sso_submesh* slots;
    uint32_t mesh_bits, submesh_bits, material_bits;
    uint32_t trig_mode;
} sso_scene;

/* ---- reference-faithful core ------------------------------------------------------------- */
void sso_mat4_mul(const float a[16], const float b[16], float out[16]);
void sso_trs_to_mat4(const float trs[9], int trig_mode, float out[16]);
void sso_pack_rows(const float m[16], float rows[12]);

/* node world matrices: world[i] = world[parent[i]] * local[i] (N1) */
void sso_node_worlds(const sso_scene* sc, float* node_world16 /* [n_nodes][16] */);
/* instance world rows: rows(node_world[node] * slots[slot].local) (SubmitMesh :1150-1153) */
void sso_instance_rows(const sso_scene* sc, const float* node_world16, float* rows12);

/* ---- specified extensions ---------------------------------------------------------------- */
void sso_planes_from_view_proj(const float vp[16], float planes[6][4]);
void sso_cull(const sso_scene* sc, const float* rows12, const sso_view* views, uint32_t n_views,
              uint8_t* masks);
void sso_keys(const sso_scene* sc, uint64_t* keys);

/* per view: visible (key, idx) sorted by key then idx; rows gathered; draws = runs of equal keys.
   Output arrays are caller-allocated with n_instances entries per view (draws too).
   Returns 0. */
int sso_view_lists(const sso_scene* sc, const float* rows12, const uint8_t* masks,
                   const uint64_t* keys, uint32_t n_views, uint32_t* n_visible /* [n_views] */,
                   uint32_t* n_draws /* [n_views] */, uint32_t* out_idx /* [n_views][n_inst] */,
                   uint64_t* out_keys, float* out_rows12 /* may be NULL */, sso_draw_cmd* out_draws,
                   uint64_t* out_draw_keys);

/* whole frame on `threads` host threads (range-partitioned; 1 = how the reference runs).
   rows12/masks/keys are outputs of n_instances entries; the per-view outputs as above.
   Returns elapsed seconds of the computation proper (allocation excluded). */
double sso_frame(const sso_scene* sc, const sso_view* views, uint32_t n_views, int threads,
                 float* rows12, uint8_t* masks, uint64_t* keys, uint32_t* n_visible,
                 uint32_t* n_draws, uint32_t* out_idx, uint64_t* out_keys, float* out_rows12,
                 sso_draw_cmd* out_draws, uint64_t* out_draw_keys);

/* ---- view matrices (host-side inputs, A9) ------------------------------------------------- */
void sso_camera_view_proj(float fov_y, float aspect, float z_near, float z_far,
                          const float position[3], float pitch, float yaw, float view[16],
                          float proj[16], float view_proj[16]);
void sso_pack_lights(const float* trs, const uint32_t* point_node, const float* point_par, uint32_t n_point,
                     const uint32_t* spot_node, const float* spot_par, uint32_t n_spot, void* ub_point, void* ub_spot,
                     float* positions);
void sso_quat_to_euler(const float* quat_xyzw, uint32_t n, float* euler_out);
void sso_cascade_matrices_snapped(const float view_proj[16], float z_near, float z_far, float lambda,
                                  float near_offset, float far_offset, const float light_dir[3],
                                  const float* cascade_origin, float map_size, float splits[4], float light_matrices[64]);
void sso_cascade_matrices(const float view_proj[16], float z_near, float z_far, float lambda,
                          float near_offset, float far_offset, const float light_dir[3],
                          float splits[4], float light_matrices[64]);
void sso_mat4_inverse(const float m[16], float out[16]);

/* ---- the consumer of the lists: depth + id rasterisation -------------------------
   CPU restatement of what the reference's ShadowPass / SelectionPass make the Vulkan device do with the instance
   buffer (assets/shaders/ShadowDepth.vert/.geom, CoreEditor.vert/.frag: clip = view_proj * (transform * position);
   SceneRenderer.cpp:1266-1310, :1421-1457).  Vulkan conventions: y down, depth z/w in [0, 1], samples at pixel
   centres.  Target word = [depth float bits : 32 | instance : 32], nearest wins, 0xFFFF...F where nothing was drawn.
   flags bit 0: clamp depth to [0, 1] instead of discarding (DepthClamp).  Parity with the reference's own rasteriser
   is unpinned (it is the Vulkan device: unavailable on this box); this function is the checker of raster.cu. */
void sso_raster_view(const float* rows12, const uint32_t* instance_idx, uint32_t visible, const uint32_t* slot_of_instance,
                     const sso_submesh* slots, const float* positions_xyz, uint32_t n_vertices, const uint32_t* indices,
                     uint32_t n_indices, const float view_proj[16], uint32_t width, uint32_t height, uint32_t flags,
                     uint64_t* target);

/* det sincos exposed for the accuracy test */
void sso_det_sincos(float x, float* s, float* c);

#ifdef __cplusplus
}
#endif
#endif
