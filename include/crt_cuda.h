/*
 * crt_cuda.h -- the C-ABI drop-in boundary of the B200 render path.
 *
 * The reference (bvpav/chaos-ray-tracing-course-2025) has no FFI layer: its one "operator" is
 *     crt::Image crt::render_image(const crt::Scene&, const crt::RendererSettings&)   src/core/crt_renderer.h:27
 * called from src/standalone/main.cpp:38 and src/python/py_crt_module.cpp:100.  The functions below are what a
 * replacement of that call binds to: plain C, plain pointers and sizes, int status codes, no STL / torch types.
 *
 * Two groups:
 *   crt_scene_*  host side: parse a .crtscene (same format as src/core/crt_json.cpp:541-647), build the
 *                reference's acceleration tree (src/core/crt_acceleration_tree.cpp:31-106) and flatten it.
 *   crt_cuda_*   device side: upload the flattened scene, run the sm_100a kernels, read results back.
 *
 * Status codes: 0 = ok, negative = error (text via crt_cuda_last_error / crt_scene_last_error).
 * Threading: a context is not thread-safe; distinct contexts are independent.  The caller owns every host
 * buffer it passes in; a context owns all device memory it allocates.
 */
#ifndef CRT_CUDA_H
#define CRT_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CRT_OK 0
#define CRT_ERR_INVALID (-1)   /* bad argument / malformed scene          */
#define CRT_ERR_CUDA (-2)      /* CUDA runtime error (see last_error)      */
#define CRT_ERR_IO (-3)        /* file could not be opened                 */
#define CRT_ERR_PARSE (-4)     /* scene failed the loader's validation     */
#define CRT_ERR_NO_DEVICE (-5) /* no CUDA device: there is NO CPU fallback */

/* ------------------------------------------------------------------------------------------------
 * Flattened scene (16-byte aligned POD arrays; host pointers, owned by whoever built the struct).
 * ---------------------------------------------------------------------------------------------- */

typedef struct crt_float4 {
    float x, y, z, w;
} crt_float4;

/* 32 B node of the reference-order ("literal") traversal: the node's AABB (src/core/crt_aabb.h:11-12)
 * plus two links.  inner: left/right = child indices (>= 1) or -1 when absent
 * (crt_acceleration_tree.cpp:63,74).  leaf: left = -(first_ref) - 2 (<= -2), right = triangle count (> 0). */
typedef struct crt_box_node {
    float min[3];
    int32_t left;
    float max[3];
    int32_t right;
} crt_box_node;

/* 16 B node of the k-d walk.  Because every child box is its parent's box cut at the midpoint of
 * axis depth%3 (crt_aabb.h:24-35), the tree is a k-d tree: an inner node is fully described by the cut.
 * inner: {split plane, child0 (or -1), child1 (or -1), axis 0..2};  leaf: {0, first_ref, count, 3}.
 * Two arrays use it: kd_nodes mirrors the reference tree node for node (same indices as box_nodes, leaf lists in
 * refs); fast_nodes is what CRT_TRAVERSAL_KD walks: OUR OWN surface-area-heuristic k-d tree over the root box
 * (core/crt_kdbuild.cpp; leaf lists in fast_refs), or -- $CRT_WALK_TREE=reference -- the reference tree after the
 * flattener merged the subtrees whose splits do not pay and stored identical leaf lists once. */
typedef struct crt_kd_node {
    float split;
    int32_t a;
    int32_t b;
    int32_t axis;
} crt_kd_node;

#define CRT_REF_CULL_BIT 0x80000000u /* leaf reference: triangle index | back_face_culling << 31 */
#define CRT_META_SMOOTH_BIT (1 << 24)
#define CRT_META_CULL_BIT (1 << 25)
#define CRT_META_MATERIAL_MASK 0x00FFFFFF

/* src/core/crt_material.h:5-16 */
typedef struct crt_material {
    int32_t type; /* 0 diffuse, 1 reflective, 2 refractive, 3 constant */
    int32_t texture_index; /* -1 for refractive (crt_json.cpp:493) */
    float ior;
    int32_t pad;
} crt_material;

/* src/core/crt_texture.h:8-45, one 48 B record per texture */
typedef struct crt_texture {
    int32_t type; /* 0 albedo, 1 edges, 2 checker, 3 bitmap */
    int32_t width, height; /* bitmap only */
    int32_t texel_offset;  /* bitmap only: first texel in crt_flat_scene.texels */
    float color0[3];       /* albedo | edge_color | color_a */
    float param;           /* edge_width | square_size */
    float color1[3];       /* inner_color | color_b */
    float pad;
} crt_texture;

/* src/core/crt_camera.h:66-69 with the per-frame constants of crt_camera.cpp:21-26 folded on the host */
typedef struct crt_camera {
    float position[3];
    float rotation[9];    /* row-major; rays are row-vector * matrix (crt_matrix.h:66-74) */
    float tan_half_fov;   /* tanf(fov_radians * 0.5f), host libm */
    float aspect;         /* float(width) / height */
    int32_t width, height;
} crt_camera;

typedef struct crt_flat_scene {
    crt_camera camera;
    float background[3];
    int32_t gi_on, reflections_on, refractions_on, bucket_size;
    float root_min[3], root_max[3];
    int32_t n_triangles, n_vertices, n_nodes, n_refs, n_lights, n_materials, n_textures, n_texels;
    int32_t max_depth; /* depth of the deepest node (root = 0) */
    int32_t fast_max_depth;
    int32_t n_fast_nodes, n_fast_refs;
    const crt_float4 *tri_verts;   /* 3 per triangle: {v0.xyz, N.x} {v1.xyz, N.y} {v2.xyz, N.z}; parse order */
    const int32_t *tri_meta;       /* 4 per triangle: vertex indices i0,i1,i2, material | flag bits         */
    const crt_float4 *vert_normal; /* 1 per vertex: smooth normal xyz (crt_mesh.cpp:19-29)                  */
    const float *vert_uv;          /* 2 per vertex                                                          */
    const crt_box_node *box_nodes; /* n_nodes, DFS pre-order exactly as the reference builds them           */
    const crt_kd_node *kd_nodes;   /* n_nodes, same indexing                                                */
    const uint32_t *refs;          /* n_refs leaf references, leaf order = parse order within a leaf        */
    const crt_kd_node *fast_nodes; /* n_fast_nodes, DFS pre-order                                           */
    const uint32_t *fast_refs;     /* n_fast_refs, ascending triangle index within a leaf                   */
    const crt_float4 *lights;      /* {position.xyz, intensity}, file order                                 */
    const crt_material *materials;
    const crt_texture *textures;
    const crt_float4 *texels;      /* all bitmap texels, rgb in xyz                                         */
} crt_flat_scene;

/* ------------------------------------------------------------------------------------------------
 * Host side: scene handle (parse + tree build + flatten).  Mirrors
 *   std::optional<Scene> crt::json::read_scene_from_istream(std::istream&, const path& asset_root)
 *   (src/core/crt_json.h:11); a validation failure is CRT_ERR_PARSE (the reference returns nullopt).
 * ---------------------------------------------------------------------------------------------- */
typedef struct crt_scene crt_scene;

int crt_scene_load_file(const char *path, crt_scene **out);
int crt_scene_load_json(const char *json, size_t length, const char *asset_root, crt_scene **out);
/* pointers inside *out stay valid until crt_scene_free.  crt_scene_flatten includes the reference's acceleration tree
 * (box_nodes / kd_nodes / refs, src/core/crt_acceleration_tree.cpp:31-106), which CRT_TRAVERSAL_LITERAL walks;
 * crt_scene_flatten_ex without CRT_FLATTEN_REFERENCE_TREE leaves it out (n_nodes == 0): the scene then renders with
 * CRT_TRAVERSAL_KD only and skips the build and the upload of that tree (0.47 GB of the 0.53 GB of the 1 M-triangle scene) */
#define CRT_FLATTEN_REFERENCE_TREE 1
int crt_scene_flatten(const crt_scene *scene, crt_flat_scene *out);
int crt_scene_flatten_ex(const crt_scene *scene, int flags, crt_flat_scene *out);
/* change the image size / camera after loading (orbit frames, 4K variants); rebuilds nothing else */
int crt_scene_set_camera(crt_scene *scene, const float position[3], const float rotation[9], float fov_degrees);
int crt_scene_set_resolution(crt_scene *scene, int width, int height);
void crt_scene_free(crt_scene *scene);
const char *crt_scene_last_error(void);

/* ------------------------------------------------------------------------------------------------
 * Device side
 * ---------------------------------------------------------------------------------------------- */
typedef struct crt_cuda_ctx crt_cuda_ctx;

#define CRT_TRAVERSAL_LITERAL 0 /* reference order: child1 first, 6-plane box test, no pruning (crt_intersection.cpp:109-136) */
#define CRT_TRAVERSAL_KD 1      /* front-to-back k-d interval traversal with strict pruning; same winners  */
#define CRT_TRAVERSAL_BRUTE 2   /* every triangle in parse order (crt_intersection.cpp:95-107), checker only */

#define CRT_PIPELINE_AUTO 0      /* wavefront (hybrid for depth >= 5 or band-sharded frames); persistent with GI or shadow rays; megakernel for those under the literal traversal */
#define CRT_PIPELINE_MEGAKERNEL 1 /* one thread per pixel evaluates the whole recursion depth-first     */
#define CRT_PIPELINE_WAVEFRONT 2  /* ray queues + compaction between bounce stages (no GI)              */
#define CRT_PIPELINE_HYBRID 4     /* primary level as wavefront kernels, all deeper levels in one persistent kernel (no GI) */
#define CRT_PIPELINE_PERSISTENT 3 /* one persistent kernel: a lane keeps its pixel through the recursion, warp-cooperative walk (GI too) */

/* crt::RendererSettings (src/core/crt_renderer.h:18-25) + the knobs that are new on this side */
#define CRT_MAX_RAY_DEPTH_LIMIT 15 /* the reference recurses without limit; the device's recursion state is sized for 0 .. 15 */
typedef struct crt_render_params {
    uint32_t max_ray_depth;                /* default 3; values above CRT_MAX_RAY_DEPTH_LIMIT are refused (CRT_ERR_INVALID) */
    uint32_t diffuse_reflection_ray_count; /* default 4  */
    float shadow_bias, reflection_bias, diffuse_reflection_bias, refraction_bias; /* default 1e-2 each */
    int32_t traversal;   /* CRT_TRAVERSAL_*  */
    int32_t pipeline;    /* CRT_PIPELINE_*   */
    int32_t shadows;     /* 0 = HEAD semantics (shadow query is dead code, crt_renderer.cpp:32-33); 1 = trace opaque shadow rays */
    int32_t collect_stats; /* 1 = count rays / box tests / triangle tests (slower; never use for timing) */
    /* row-band ownership for multi-GPU stills: this rank renders bands b with b % world == rank,
     * a band being band_rows consecutive image rows; output buffers are then COMPACT (owned rows only,
     * in increasing y).  world <= 1 renders everything. */
    int32_t rank, world, band_rows;
    int32_t pad;
} crt_render_params;

typedef struct crt_render_stats {
    float gpu_ms;             /* CUDA-event time of all kernels of the frame on the context's stream */
    int32_t kernel_launches;  /* kernels of OURS launched for the frame                               */
    int32_t rows_rendered;    /* rows this rank owns                                                  */
    int32_t pad;
    uint64_t rays;            /* closest-hit traversals issued (all depths)  [collect_stats]          */
    uint64_t rays_by_depth[16];
    uint64_t shadow_rays;     /* only in shadows mode                                                 */
    uint64_t node_tests;      /* boxes tested / k-d nodes visited                                     */
    uint64_t tri_tests;
    uint64_t entered_rays;    /* rays that enter the root box, i.e. that a walk is started for (<= rays)   */
} crt_render_stats;

void crt_render_params_default(crt_render_params *p);

int crt_cuda_device_count(void);
int crt_cuda_create(int device, crt_cuda_ctx **out);
void crt_cuda_destroy(crt_cuda_ctx *ctx);
const char *crt_cuda_last_error(const crt_cuda_ctx *ctx);
/* run on a caller-provided CUDA stream (a cudaStream_t passed as void*); NULL = the context's own stream.
 * The legacy default stream is cudaStreamLegacy ((void*)1), not NULL. */
int crt_cuda_set_stream(crt_cuda_ctx *ctx, void *stream);

/* cudaMemcpyAsync of every array from (internally pinned) staging; replaces the implicit "scene is in RAM" */
int crt_cuda_upload_scene(crt_cuda_ctx *ctx, const crt_flat_scene *scene);
/* Host-only helper (no device needed), exposed for tests and tools: the walk tree as crt_cuda_upload_scene lays it out
 * in device memory -- 8 bytes per node instead of crt_kd_node's 16, nodes re-numbered so that the two children of a
 * node are neighbours.  out: n_nodes records.  Inner node: low word = split (float bits), high word = axis (bits 0-1)
 * | has_a << 2 | has_b << 3 | index of the first present child << 4 (child b follows child a).  Leaf: low word = first
 * reference, high word = 3 | count << 2.  The root stays node 0.  CRT_ERR_INVALID for a broken tree (shared nodes
 * that do not unfold into n_nodes records, a bad axis, an index out of range, more than 2^28 nodes). */
int crt_cuda_pack_walk_tree(const crt_kd_node *nodes, int32_t n_nodes, uint64_t *out);
/* bytes of one walk-tree node as the kernels of this build read it (16: crt_kd_node as it is; 8 when built with
 * -DCRT_NODE8=1, the packed record of crt_cuda_pack_walk_tree): what the algorithmic bytes per ray of a measurement are
 * counted with */
int crt_cuda_walk_node_bytes(void);
/* What shading reads and a caller may change between two frames WITHOUT touching the geometry: lights, materials,
 * texture records (same counts as uploaded; bitmap texels stay), background, the gi / reflections / refractions flags,
 * bucket_size.  crt::render_image calls it on every render, because the reference's render_image reads the Scene on
 * every call (crt_renderer.cpp:157-199).  Nothing is copied when nothing changed.  CRT_ERR_INVALID when a count
 * changed: upload the scene again. */
int crt_cuda_update_shading(crt_cuda_ctx *ctx, const crt_flat_scene *scene);
/* cheap per-frame update (animation): only the 72 B camera block */
int crt_cuda_set_camera(crt_cuda_ctx *ctx, const crt_camera *camera);

/* Replaces crt::render_image (crt_renderer.cpp:157-199).  rgb_out: width*rows*3 floats (row 0 = top,
 * crt_image.h:11-27) or NULL; rgb8_out: width*rows*3 bytes quantised exactly like write_ppm
 * (crt_image_ppm.cpp:16-19) or NULL.  Host pointers; the call returns after the copies complete. */
int crt_cuda_render(crt_cuda_ctx *ctx, const crt_render_params *params, float *rgb_out, uint8_t *rgb8_out, crt_render_stats *stats);
/* The same render for ONE FRAME SHARED BY SEVERAL RANKS (params->world > 1): rgb_frame / rgb8_frame point at the FULL
 * frame in host memory (width*height*3 floats / bytes, row 0 = top) -- e.g. a shared-memory segment that every rank
 * has mapped and page-locked -- and this rank's row bands are copied to their place in it, over this GPU's own PCIe
 * link: the frame the reference's bucket threads assemble in `buffer[y*W + x]` (crt_renderer.cpp:171-196), assembled by
 * N copy engines.  With world <= 1 it is crt_cuda_render. */
int crt_cuda_render_frame(crt_cuda_ctx *ctx, const crt_render_params *params, float *rgb_frame, uint8_t *rgb8_frame, crt_render_stats *stats);
/* same, results stay in device memory (see crt_cuda_device_framebuffer); asynchronous on the stream */
int crt_cuda_render_device(crt_cuda_ctx *ctx, const crt_render_params *params, crt_render_stats *stats);
/* device pointers of the last frame: float rgb (12 B/px) and u8 rgb (3 B/px); rows = owned rows */
int crt_cuda_device_framebuffer(crt_cuda_ctx *ctx, void **rgb_dev, void **rgb8_dev, int *width, int *rows);
/* completes the asynchronous work and fills gpu_ms of the last crt_cuda_render_device */
int crt_cuda_synchronize(crt_cuda_ctx *ctx, crt_render_stats *stats);

/* Parity hook: primary rays only.  tri_id = parse-order triangle index or -1; t = hit distance. Host buffers of
 * width*height entries; any may be NULL.  node_tests/tri_tests are per-pixel counters (int32). */
int crt_cuda_primary_hits(crt_cuda_ctx *ctx, int traversal, int32_t *tri_id, float *t, int32_t *node_tests, int32_t *tri_tests, crt_render_stats *stats);
/* Parity hook for ray generation (crt_camera.cpp:7-35): 6 floats per pixel (origin, direction) */
int crt_cuda_primary_rays(crt_cuda_ctx *ctx, float *rays_out);

/* ------------------------------------------------------------------------------------------------
 * Multi-GPU stills (replaces the shared `buffer[y*W + x]` the reference's bucket threads write,
 * crt_renderer.cpp:171-196): rank 0 owns the full quantised frame, every rank pushes its row bands into it with
 * direct peer stores (NVLink), un-interleaving on the way.  One process per GPU: rank 0 calls frame_create and ships
 * the 64-byte handle to the other ranks (any transport), they call frame_open.  One process driving several GPUs, or
 * several contexts on one GPU: frame_attach with the raw device pointer.  After crt_cuda_render_device, every rank
 * calls crt_cuda_push_bands (asynchronous on the stream); the caller then synchronises the ranks (a barrier) and
 * rank 0 reads the frame (crt_cuda_frame_pointer, crt_cuda_encode_ppm_device). */
#define CRT_IPC_HANDLE_BYTES 64
int crt_cuda_frame_create(crt_cuda_ctx *ctx, int width, int height, unsigned char handle_out[CRT_IPC_HANDLE_BYTES]);
int crt_cuda_frame_open(crt_cuda_ctx *ctx, const unsigned char handle[CRT_IPC_HANDLE_BYTES], int width, int height);
int crt_cuda_frame_attach(crt_cuda_ctx *ctx, void *frame_dev, int width, int height);
int crt_cuda_frame_pointer(crt_cuda_ctx *ctx, void **frame_dev, int *width, int *height);
int crt_cuda_push_bands(crt_cuda_ctx *ctx);

/* One host thread, several GPUs of one box: a context per device, scene replicated, one frame cut into interleaved
 * row bands (band b -> devices[b % n]), every device pushing its bands into the frame on devices[0] (peer access).
 * group_render fills HOST buffers like crt_cuda_render: rgb_out = width*height*3 floats, rgb8_out = width*height*3
 * bytes (either may be NULL); params->rank/world are set by the group.  stats: gpu_ms = slowest device, counters summed.
 * The same device may be listed more than once (contexts share it; used by the tests). */
typedef struct crt_cuda_group crt_cuda_group;
int crt_cuda_group_create(const int *devices, int n, crt_cuda_group **out);
void crt_cuda_group_destroy(crt_cuda_group *group);
const char *crt_cuda_group_last_error(const crt_cuda_group *group);
int crt_cuda_group_size(const crt_cuda_group *group);
int crt_cuda_group_upload_scene(crt_cuda_group *group, const crt_flat_scene *scene);
int crt_cuda_group_set_camera(crt_cuda_group *group, const crt_camera *camera);
int crt_cuda_group_update_shading(crt_cuda_group *group, const crt_flat_scene *scene);
int crt_cuda_group_render(crt_cuda_group *group, const crt_render_params *params, float *rgb_out, uint8_t *rgb8_out, crt_render_stats *stats);
/* the assembled quantised frame on devices[0] (for crt_cuda_encode_ppm_device) */
int crt_cuda_group_frame(crt_cuda_group *group, void **frame_dev, int *width, int *height);
/* P3 text of that frame (needs a group_render with rgb8_out != NULL before) */
int crt_cuda_group_encode_ppm(crt_cuda_group *group, char *text_out, size_t capacity, size_t *length);

/* P3 text encoding of the last frame on the device (crt_image_ppm.cpp:9-23), byte-exact.
 * Returns the byte count in *length; text_out may be NULL to query the size. */
int crt_cuda_encode_ppm(crt_cuda_ctx *ctx, char *text_out, size_t capacity, size_t *length);
/* same for any quantised frame in device memory (width*height*3 bytes, row 0 = top), e.g. the frame rank 0 holds
 * after the band gather of a multi-GPU render */
int crt_cuda_encode_ppm_device(crt_cuda_ctx *ctx, const void *rgb8_dev, int width, int height, char *text_out, size_t capacity, size_t *length);
/* The same encoder for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Calls no cudaMalloc / cudaFree or any allocating routine: the encoder's scratch (per-block text lengths and their
 * prefix sums) is a file-scope `static __device__` array of CRT_PPM_NOALLOC_MAX_PIXELS / 1024 + 8 64-bit words (64 KiB,
 * held by every GPU that loads the library until the process exits), and the text is written straight into the
 * CALLER's device buffer text_dev[0, *length) (any alignment; capacity in bytes).  text_dev may be NULL to query the
 * length.  width * height beyond CRT_PPM_NOALLOC_MAX_PIXELS is refused with CRT_ERR_INVALID (crt_cuda_last_error says
 * why).  All calls on one device share the scratch array, which keeps the last call's contents (nothing in it has to
 * start at zero: every word is written before it is read); calls that OVERLAP on the GPU race on it -- serialise them.
 * Runs on the context's stream and returns after the text is complete. */
#define CRT_PPM_NOALLOC_MAX_PIXELS 8388608 /* a 3840x2160 frame fits */
int crt_cuda_encode_ppm_noalloc(crt_cuda_ctx *ctx, const void *rgb8_dev, int width, int height, void *text_dev, size_t capacity, size_t *length);

#ifdef __cplusplus
}
#endif
#endif /* CRT_CUDA_H */
