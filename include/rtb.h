/* rtb.h -- C ABI of the B200-native hot path of the glTF path tracer (librtb.so).
 *
 * The reference (Korinin38/raytracing-hw) has no plugin / FFI surface: it is one process,
 * `run.sh input.gltf width height samples [output.ppm]` (run.sh:2, src/main.cpp:22-58).  The seam this
 * library replaces is the pair of const member functions the reference's main() reaches after the Scene
 * constructor has built the BVH and the light distribution:
 *
 *     BVH::intersect(const std::vector<Primitive>&, Ray)      src/core/bvh.h:24,  src/core/bvh.cpp:239-243
 *     Scene::render(ProgressFunc)                             src/core/scene.h:19, src/core/scene.cpp:17-65
 *
 * Everything is `extern "C"`, plain pointers and sizes, int status codes (0 = ok); the message of the last
 * failure on the calling thread is returned by rtb_last_error().  No exception crosses this boundary; the C++
 * host (raytracing-hw_b200/host/main.cpp) re-throws std::runtime_error to keep the reference's
 * abort-on-error behaviour.  There is NO CPU fallback: every compute entry point fails with RTB_ERR_NO_DEVICE
 * when no CUDA device is usable.
 *
 * Threading: a scene may be used from one host thread at a time.  Device selection is per scene.
 */
#ifndef RTB_H_
#define RTB_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RTB_OK 0
#define RTB_ERR_INVALID 1      /* bad argument / malformed scene description            */
#define RTB_ERR_NO_DEVICE 2    /* no usable CUDA device (never falls back to the CPU)    */
#define RTB_ERR_CUDA 3         /* a CUDA runtime call or kernel failed                   */
#define RTB_ERR_IO 4           /* file could not be read / written / parsed              */

/* Closest-hit culling policy (see DESIGN.md "culling modes").
 *  REFERENCE: visits exactly the nodes BVH::intersectHelper visits (src/core/bvh.cpp:177-237): the near child is
 *             always entered, only the far sibling is culled, against the best hit of the near sibling's subtree.
 *             Bit-exact ids by construction.
 *  GLOBAL   : same visit order, but every box is culled against the best hit found so far.  Visits a subset; ids are
 *             identical unless a triangle's t rounds below the entry distance of a box that contains it while another
 *             hit lies between the two values: a documented NEAR-tie deviation, seen only on rays aimed at shared
 *             vertices / edges (up to 0.45 % of such rays, t up to 3.4e-4 relative apart, always the nearer hit;
 *             counts per scene in DESIGN.md section 3 and tests/test_gpu_closest_hit.py: NEAR_TIES).                 */
#define RTB_CULL_REFERENCE 0
#define RTB_CULL_GLOBAL 1

typedef struct rtb_scene rtb_scene;            /* device-resident scene (one GPU)            */
typedef struct rtb_host_scene rtb_host_scene;  /* host-side scene: glTF load + BVH build     */

typedef struct rtb_texture_desc {
    int32_t width, height;
    const uint8_t *rgba;        /* width*height*4 bytes, row-major, as tinygltf/stb_image deliver (always 4 channels) */
} rtb_texture_desc;

/* Flat scene description, all arrays borrowed for the duration of rtb_scene_create only.
 * Mirrors what Scene holds after its constructor (src/core/scene.cpp:197-249):                                     */
typedef struct rtb_scene_desc {
    /* scene BVH in REFERENCE node order (struct Node, src/core/bvh.h:10-17) */
    uint32_t n_nodes;
    const float *node_aabb;       /* n_nodes x 6 : min xyz, max xyz                                                  */
    const int32_t *node_links;    /* n_nodes x 5 : left, right, split_dim, first_primitive_id, primitive_count; -1 = invalidKey */
    /* triangles in POST-BUILD order (BVH::buildBVH reorders Scene::objects, src/core/bvh.cpp:166) */
    uint32_t n_tris;
    const float *tri_pos;         /* n x 9  : p0, U = p1-p0, V = p2-p0   (Primitive::position, src/geometry/primitive.h:101) */
    const float *tri_nrm;         /* n x 9  : per-vertex normals          (Primitive::normal)                          */
    const float *tri_tan;         /* n x 12 : per-vertex tangents xyzw    (Primitive::tangent)                         */
    const float *tri_uv;          /* n x 6  : per-vertex texcoords        (Primitive::texcoord)                        */
    const int32_t *tri_mesh;      /* n      : Primitive::mesh_id                                                       */
    /* meshes (struct Mesh / Material, src/geometry/primitive.h:76-94) */
    uint32_t n_meshes;
    const float *mesh_material;   /* n x 16 : ior, alpha, base rgb, emission rgb, metallic, roughness2,
                                              base_color_i, normal_i, metallic_roughness_i, emission_i (as float, -1 = none), 0, 0 */
    const double *mesh_normal_transform; /* n x 16, column-major matrix4d                                              */
    uint32_t n_textures;
    const rtb_texture_desc *textures;
    /* light list: emissive triangles re-ordered by their own BVH (rng::ManyLightsDistribution, src/utils/random.cpp:156-167) */
    uint32_t n_lights;
    const float *light_pos;       /* n_lights x 9 : p0, U, V                                                          */
    uint32_t n_light_nodes;
    const float *light_node_aabb; /* n_light_nodes x 6                                                                */
    const int32_t *light_node_links; /* n_light_nodes x 5                                                             */
    /* camera (class Camera, src/core/camera.h:13-45) */
    float cam_pos[3];
    float cam_axes[9];            /* right, up, forward                                                               */
    float cam_tan_half_fov[2];    /* tanf(fov_x/2), tanf(fov_y/2) evaluated by the HOST libm (camera.cpp:51-52)       */
    /* integrator constants */
    int32_t ray_depth;            /* 6  (src/io/scene_parser.cpp:17)                                                  */
    float max_distance;           /* zfar or 1e9 (src/io/scene_parser.cpp:22,119-120)                                 */
    float bg_color[3];            /* 0,0,0 (src/core/scene.cpp:203)                                                   */
} rtb_scene_desc;

typedef struct rtb_render_params {
    int32_t width, height;        /* canvas size (CLI arguments 2 and 3)                                              */
    int32_t samples;              /* samples per pixel (CLI argument 4)                                               */
    int32_t row_begin, row_end;   /* rows [row_begin,row_end) rendered by this call (tile split); 0,height = all     */
    uint32_t pixel0_seed;         /* engine seed of pixel 0 (the reference uses std::random_device there,
                                     src/utils/random.cpp:12-18); 0 is mapped to 1 as minstd_rand::seed does          */
    int32_t sample_begin;         /* 0, or -- rtb_render_resume only -- the samples per pixel an earlier call has
                                     already taken: all samples of a pixel share ONE sequential engine stream
                                     (src/core/scene.cpp:34-35), so a frame can only be continued, not cut freely      */
    int32_t flags;                /* RTB_RENDER_* bits                                                                */
    /* multi-GPU tile split: rows [row_begin,row_end) are cut into blocks of `interleave_rows` rows (a multiple of 4);
     * this call renders the blocks b with b % interleave_count == interleave_index and writes them COMPACTLY, in block
     * order, into accum / rgb8 (rtb_render_local_rows() rows).  interleave_count <= 1: every row, as one range.        */
    int32_t interleave_count, interleave_index, interleave_rows;
    /* sample-range split (SURVEY section 8e "optional sample-range split"; statistically equivalent to, NOT bit-identical
     * with, the reference's stream).  substreams = K > 1: every pixel owns K independent engine streams; stream k is
     * seeded with k*width*height + (j*width + i) -- the seed the reference would give a pixel of a K times taller canvas
     * (src/core/scene.cpp:34) -- and takes samples [k*samples/K, (k+1)*samples/K) of the pixel; the K sums are folded in
     * k order into the pixel's sum (src/core/scene.cpp:42).  Stream 0 of a pixel is the reference's own stream.  A
     * call renders streams [substream_begin, substream_end) (0,0 = all K); a call that does not hold all K streams
     * returns PARTIAL sums in accum (to be added to the other calls' -- across GPUs: rtb_render_dist) and must pass
     * rgb8 == NULL.  Wavefront integrator only; not combinable with rtb_render_resume.  0 or 1: one stream per pixel,
     * the reference's (bit-exact mode, the default).                                                                   */
    int32_t substreams, substream_begin, substream_end;
} rtb_render_params;

/* number of rows a call with these parameters renders (and its output buffers hold); -1 on invalid parameters */
int rtb_render_local_rows(const rtb_render_params *params);
/* global row of local row `local_row` of such a call */
int rtb_render_global_row(const rtb_render_params *params, int local_row);

#define RTB_RENDER_COUNTERS 1        /* also count box / triangle tests (slower instantiation of the kernel); rays and
                                        light queries are always counted                                                */
#define RTB_RENDER_REFERENCE_CULL 2  /* traverse with RTB_CULL_REFERENCE instead of the default RTB_CULL_GLOBAL
                                        (implies the megakernel integrator)                                              */
#define RTB_RENDER_STAGE_TIMES 8     /* record CUDA events around every closest-hit queue launch (rtb_scene_stage_ms)   */
#define RTB_RENDER_MEGAKERNEL 4      /* one thread per pixel runs the whole bounce loop (the first integrator; kept for
                                        A/B measurements).  Default: the wavefront / ray-queue integrator, whose sums
                                        are bit-identical to the megakernel's                                            */

/* set rtb_stats.box_tests to this value before a call to ask for box/triangle test counts (a slower, counting
 * instantiation of the kernel is launched); otherwise box_tests / tri_tests come back 0.                          */
#define RTB_REQUEST_COUNTERS 0xFFFFFFFFFFFFFFFFull

typedef struct rtb_stats {
    uint64_t closest_hit_rays;    /* Scene::intersect entries with power > 0 (primary + secondary)                    */
    uint64_t light_pdf_queries;   /* BVH::intersectAll queries (src/utils/random.cpp:183)                             */
    uint64_t box_tests;           /* ray/AABB tests (closest-hit + light queries), 0 unless requested (see above)     */
    uint64_t tri_tests;           /* ray/triangle tests,                           0 unless requested                 */
    double kernel_ms;             /* device time of the render / closest-hit kernels of this call (CUDA events)       */
    double h2d_ms, d2h_ms;        /* copies inside this call                                                          */
    uint64_t launches;            /* kernels launched by this call                                                    */
    uint64_t light_box_tests;     /* the part of box_tests / tri_tests done by the light-pdf queries (light BVH)      */
    uint64_t light_tri_tests;
} rtb_stats;

const char *rtb_last_error(void);
int rtb_device_count(void);

/* ------------------------------------------------------------------------------------------------ host side
 * glTF loading and BVH construction stay on the C++ host (north_star); these produce a flat rtb_scene_desc whose
 * node and triangle layout equals the reference's (verified node-for-node in tests/test_layout.py).               */

/* parse_scene_gltf + Scene::Scene (src/io/scene_parser.cpp:25-350, src/core/scene.cpp:197-249) */
int rtb_host_scene_load_gltf(const char *path, int width, int height, int samples, rtb_host_scene **out);
/* a triangle soup given as 3 world-space vertices per triangle (n x 9 floats); normals may be NULL (flat);
 * triangles [emissive_from, n) are emissive (1,1,1).  Camera as in rtb_scene_desc.                               */
int rtb_host_scene_from_triangles(const float *verts9, const float *nrm9, size_t n, size_t emissive_from,
                                  const float cam_pos[3], const float cam_axes[9], float fov_x, float fov_y,
                                  float max_distance, rtb_host_scene **out);
/* fills *desc with pointers into the host scene (valid until rtb_host_scene_destroy) */
int rtb_host_scene_desc(const rtb_host_scene *hs, rtb_scene_desc *desc);
/* BVH statistics the reference prints (src/core/scene.cpp:233-246): counts[0..4] = nodes with 0..4 primitives */
int rtb_host_scene_bvh_stats(const rtb_host_scene *hs, uint64_t *n_nodes, uint64_t counts5[5], uint64_t *max_leaf);
double rtb_host_scene_build_seconds(const rtb_host_scene *hs);
/* Where BVH::buildBVH (src/core/bvh.cpp:5-175) ran for the scene's triangles.  out8: {1 if the GPU builder cut the tree's
 * large nodes (csrc/rtb_build.cu) else 0, levels, partition rounds, heap-sorted ranges, nodes handed back to the host
 * uncut, GPU seconds incl. copies, kernel seconds, 0}.  The layout is the reference's either way (tests/test_layout.py,
 * tests/test_gpu_build.py).  RTB_BUILD_GPU=0|1 forces the choice (default: scenes of >= 500 k triangles when a device exists),
 * RTB_BUILD_GPU_BELOW=<n> (default 5: the whole tree on the GPU) sets the node size below which the host carries on.                              */
int rtb_host_scene_build_info(const rtb_host_scene *hs, double out8[8]);

/* BVH::buildBVH alone (src/core/bvh.cpp:166-175; called from Scene::Scene, src/core/scene.cpp:213) for a caller that keeps
 * its own scene: tri_pos9 = Primitive::position[0..2] of every triangle (p0, U = p1 - p0, V = p2 - p0; primitive.h:104,
 * scene_parser.cpp:306-312), n of them, in the order of the caller's `primitives` vector.
 *   order[i]            = index (into the caller's vector) of the triangle buildBVH leaves at position i
 *   node_aabb/node_links = the reference's `nodes` in the reference's order: min xyz, max xyz / {left, right, split_dim,
 *                         first_primitive_id, primitive_count}, -1 = invalidKey (src/core/bvh.h:8-17)
 * node_capacity: nodes the two arrays hold (2 n - 1 always suffices); *n_nodes receives the count (RTB_ERR_INVALID and the
 * needed count when the arrays are too small).  out8 (may be NULL): rtb_host_scene_build_info's record.  Runs on the GPU
 * under the same policy as the scene loaders (RTB_BUILD_GPU), on the host otherwise; the layout is the reference's either way. */
int rtb_build_bvh(const float *tri_pos9, size_t n, uint32_t *order, float *node_aabb, int32_t *node_links,
                  size_t node_capacity, size_t *n_nodes, double out8[8]);
void rtb_host_scene_destroy(rtb_host_scene *hs);

/* ------------------------------------------------------------------------------------------------ device side */

/* Uploads the scene to CUDA device `device` (converts nodes to the 64-byte two-child layout). */
int rtb_scene_create(const rtb_scene_desc *desc, int device, rtb_scene **out);

/* ---- device-side scene assembly ------------------------------------------------------------
 * The reference assembles every triangle on the host while it parses the file (src/io/scene_parser.cpp:280-345: index
 * fetch, world transform, p0 / p1-p0 / p2-p0, normal transform + normalise, texcoords, tangents), computes the geometric
 * normal lazily (src/geometry/primitive.cpp:77-84), and its 64-byte Node array is what BVH::intersect walks.
 * rtb_scene_create_raw takes the geometry AS THE FILE DELIVERS IT -- one rtb_raw_mesh per glTF primitive, pointing at the
 * raw accessor views -- plus the host builder's permutation, uploads those bytes unchanged and runs the assembly on the
 * GPU: one kernel writes the traversal, shading and tangent records of every triangle with the host's operation order
 * (every buffer is byte-equal to what rtb_scene_create builds from the host-assembled arrays:
 * tests/test_gpu_assembly.py), and a level-by-level pass turns the reference-order nodes into the 64-byte two-child
 * nodes.  desc->tri_pos / tri_nrm / tri_tan / tri_uv / tri_mesh are NOT read (they may be NULL); everything else in
 * desc is as for rtb_scene_create (the light list is small and stays host-assembled).                                 */
#define RTB_NORMALS_TRANSFORM 0   /* glTF: normalize(normal_transform * n)  (scene_parser.cpp:316-325); no view: (0,0,1) */
#define RTB_NORMALS_VERBATIM 1    /* the view holds finished world-space normals                                         */
#define RTB_NORMALS_FLAT 2        /* normalize(cross(U, V)) for all three vertices (no view)                             */
#define RTB_RAW_WORLD_POSITIONS 1 /* rtb_raw_mesh::flags                                                                    */
typedef struct rtb_raw_mesh {
    uint32_t n_tris;              /* indices.count / 3                                                                  */
    uint32_t n_vertices;          /* elements in each vertex view; every index must be below it                         */
    const void *indices;          /* u16 / u32 index view, or NULL: triangle i uses vertices 3i, 3i+1, 3i+2            */
    int32_t index_bytes;          /* 2 or 4 (0 with indices == NULL)                                                    */
    int32_t index_stride;         /* bytes between indices (= index_bytes when tightly packed)                          */
    const void *position;         /* float3 view; strides in bytes, multiples of 4; views 4-byte aligned                */
    const void *normal;           /* float3 view or NULL                                                                */
    const void *tangent;          /* float4 view or NULL (zeros)                                                        */
    const void *texcoord;         /* float2 view or NULL (zeros)                                                        */
    int32_t position_stride, normal_stride, tangent_stride, texcoord_stride;
    int32_t normal_mode;          /* RTB_NORMALS_*                                                                      */
    int32_t mesh_id;              /* Primitive::mesh_id of its triangles (index into desc->mesh_material / _normal_transform) */
    int32_t flags;                /* RTB_RAW_WORLD_POSITIONS: positions are used as they are (no transform)             */
    double transform[16];         /* node matrix, matrix4d data order (src/utils/matrix.h); positions go through
                                     multiply(transform, vector4f): double products, FLOAT accumulator (matrix.h:66-73) */
} rtb_raw_mesh;
typedef struct rtb_raw_geometry {
    uint32_t n_meshes;
    const rtb_raw_mesh *meshes;   /* in loader order: mesh m's triangles are numbered after those of meshes 0..m-1      */
    const uint32_t *tri_source;   /* desc->n_tris entries: post-build triangle i is loader-order triangle tri_source[i]
                                     (BVH::buildBVH reorders Scene::objects, src/core/bvh.cpp:50,110)                   */
} rtb_raw_geometry;
/* the raw views of a host scene (valid until rtb_host_scene_destroy), and -- desc != NULL -- the scene description that
 * goes with them: as rtb_host_scene_desc fills it, but tri_nrm / tri_tan / tri_uv stay NULL (the host never assembles them) */
int rtb_host_scene_raw_geometry(const rtb_host_scene *hs, rtb_scene_desc *desc, rtb_raw_geometry *raw);
int rtb_scene_create_raw(const rtb_scene_desc *desc, const rtb_raw_geometry *raw, int device, rtb_scene **out);
/* Test / inspection aid: copies device buffer `which` of a scene to the host (dst may be NULL: size query).  Returns the
 * buffer's size in bytes through *bytes; copies min(*bytes on entry, size).                                            */
#define RTB_BUF_NODES 0
#define RTB_BUF_TRIS 1
#define RTB_BUF_SHADE 2
#define RTB_BUF_TANGENTS 3
#define RTB_BUF_BIG_LEAVES 4
#define RTB_BUF_BVH_SCALARS 5     /* root_box[6] (float), root_link, n_internal, n_tris, max_depth (int32): 40 bytes     */
int rtb_scene_read_buffer(rtb_scene *scene, int which, void *dst, size_t *bytes);
void rtb_scene_destroy(rtb_scene *scene);
int rtb_scene_device(const rtb_scene *scene);
/* bytes of device memory holding nodes+triangles (the traversal working set) */
size_t rtb_scene_traversal_bytes(const rtb_scene *scene);
/* host->device bytes rtb_scene_create copied for this scene (nodes, triangles, shading records, textures, ...) */
size_t rtb_scene_upload_bytes(const rtb_scene *scene);

/* Device blocks of destroyed scenes (scene arena, path state, staging buffers) are kept for the next scene -- at most
 * 8 blocks per device and RTB_BLOCK_CACHE_MB megabytes (default 8192; 0 = none) per process -- so that re-creating a scene
 * per frame does not pay cudaMalloc / cudaFree of ~1 GB each time.  This call hands all of them back to the driver.
 * (The threads of rtb_render_multi never free: what they drop is freed when the call returns.)                          */
void rtb_release_cached_memory(void);
/* kernels launched so far by rtb_render / rtb_render_device calls on this scene */
unsigned long long rtb_scene_launches(const rtb_scene *scene);
/* After a render with RTB_RENDER_STAGE_TIMES (wavefront integrator): waits for the frame and returns the summed
 * device time (CUDA events on the launching stream) of the closest-hit queue kernel launches, of everything else
 * in the frame (shade + compaction), and the number of queue launches.                                           */
int rtb_scene_stage_ms(rtb_scene *scene, double *trace_ms, double *other_ms, unsigned long long *trace_launches);

/* Replaces BVH::intersect (src/core/bvh.cpp:239-243) for a batch.  rays6: n x 6 floats = origin xyz, direction xyz
 * (un-normalised; normalised on device exactly as Ray::Ray does, src/geometry/primitive.cpp:12-15).
 * Outputs (any may be NULL except id): id = index into the post-build triangle array, -1 on a miss; t,u,v as
 * Intersection::distance/local_coords; inside as Intersection::inside.  HOST pointers; copies are inside the call. */
int rtb_closest_hit(rtb_scene *scene, const float *rays6, size_t n, int cull_mode,
                    int32_t *id, float *t, float *u, float *v, uint8_t *inside, rtb_stats *stats);
/* same, DEVICE pointers (caller-owned, e.g. torch tensors), launched on `stream` (a cudaStream_t, may be 0);
 * asynchronous; counters (may be NULL) = device pointer to 2 uint64 {box tests, triangle tests}.                  */
int rtb_closest_hit_device(rtb_scene *scene, const float *d_rays6, size_t n, int cull_mode,
                           int32_t *d_id, float *d_t, float *d_u, float *d_v, uint8_t *d_inside,
                           unsigned long long *d_counters, void *stream);

/* rtb_closest_hit_device for callers that are not allowed to allocate device memory.  (No counterpart in the reference:
 * its BVH::intersect runs on the host and has no such concern.)
 *
 * Who this is for.  Every other compute entry point may reach cudaMalloc / cudaFree behind the scenes: scenes keep
 * cudaMalloc'ed cursors, staging and path-state blocks grow on demand and go through the block cache.  Some callers cannot
 * have that happen on their thread:
 *   - a rank thread that shares the process with an NCCL collective in flight: cudaMalloc / cudaFree wait for the whole
 *     device, the collective waits for this rank -- the 8-GPU hang recorded in DESIGN.md section 7 was exactly this;
 *   - a plugin inside a host that has handed all device memory to its own pool (a caching allocator, a renderer's arena):
 *     the driver has nothing left to give, or the host forbids calls that synchronise the device;
 *   - a frame loop with a latency budget, where an implicit device-wide synchronisation stalls every other stream.
 * This call allocates nothing, frees nothing and calls no library routine that does.  The scene, the rays and the outputs
 * are the caller's, as for rtb_closest_hit_device (same arguments, same results bit for bit, asynchronous on `stream`).
 * The one piece of working memory the queue kernel needs -- the cursor its warps advance to claim rays -- lives in a
 * 16 KB table compiled into librtb.so (64 cursors, one per 256-byte line, used in turn).  Consequences:
 *   - every GPU on which the library's kernels are loaded holds those 16 KB until the process exits;
 *   - a cursor is 32 bits wide: at most RTB_CLOSEST_HIT_STATIC_MAX_RAYS rays per call, more is refused with
 *     RTB_ERR_INVALID (the other entry point switches kernels instead; this one asks the caller to split the batch);
 *   - the table is shared by all scenes and threads of the process on that GPU and keeps what the last call left in it.
 *     The call clears its cursor itself (a memset ordered on `stream` in front of the kernel), so no call depends on
 *     earlier contents -- but calls that overlap on the GPU (different streams or threads, 64 or more in flight) can be
 *     handed the same cursor and then race; keep such calls ordered;
 *   - the first call on a device looks up where the table lives on that device (cudaGetSymbolAddress, once per device,
 *     with that device current).  Make that first call ahead of time -- n == 0 is enough -- and outside any stream
 *     capture.
 * RTB_CULL_REFERENCE is accepted too; its kernel keeps no cursor and touches the table not at all.                    */
#define RTB_CLOSEST_HIT_STATIC_MAX_RAYS 0xFFFFFEFFu
int rtb_closest_hit_static(rtb_scene *scene, const float *d_rays6, size_t n, int cull_mode,
                           int32_t *d_id, float *d_t, float *d_u, float *d_v, uint8_t *d_inside,
                           unsigned long long *d_counters, void *stream);

/* Measurement aid (no reference counterpart): L2 bandwidth of `device` on a `bytes`-sized buffer (choose one that fits the
 * L2) -- coalesced streaming reads, and 64-byte records gathered at random indices (an incoherent ray's node / triangle
 * fetch without the arithmetic).  bench.py reports the traversal's algorithmic bytes against both.                    */
int rtb_measure_l2(int device, size_t bytes, double *stream_gbs, double *gather64_gbs);

/* Replaces ManyLightsDistribution::pdf (src/utils/random.cpp:179-188) for a batch (test/diagnostic entry). */
int rtb_light_pdf(rtb_scene *scene, const float *rays6, size_t n, float *pdf);

/* Replaces Scene::render (src/core/scene.cpp:17-65).  Renders rows [row_begin,row_end) at `samples` spp.
 * accum (may be NULL): (row_end-row_begin) x width x 3 floats, the per-pixel radiance SUMS before the resolve;
 * rgb8  (may be NULL): (row_end-row_begin) x width x 3 bytes after x(1/samples), ACES, gamma 1/2.2, round.
 * HOST pointers; the device->host copy is inside the call.                                                         */
int rtb_render(rtb_scene *scene, const rtb_render_params *params, float *accum, uint8_t *rgb8, rtb_stats *stats);
/* same, DEVICE output pointers, enqueued on `stream`.  The ray-queue integrator reads its live-slot counts back before it
 * returns (one short stream synchronisation per frame); RTB_TRACE_TAIL=0 or the megakernel flag
 * keep the call fully asynchronous.  d_counters: 6 uint64 {rays, light queries, box, tri,
 * light-query box, light-query tri} or NULL (box / tri are totals over both BVHs and need RTB_RENDER_COUNTERS)    */
int rtb_render_device(rtb_scene *scene, const rtb_render_params *params, float *d_accum, uint8_t *d_rgb8,
                      unsigned long long *d_counters, void *stream);

/* ---- progressive rendering: checkpoint / resume and the progress line -------------------
 * The state of a pixel BETWEEN two samples is its engine (minstd_rand value + the cached half of the Marsaglia pair,
 * src/utils/random.cpp) and its radiance sum (src/core/scene.cpp:42): RTB_PIXEL_STATE_BYTES per pixel, in the order
 * of the call's local rows.  rtb_render_resume renders samples [params->sample_begin, params->samples) of every pixel:
 *   state in : read when sample_begin > 0 (what an earlier call with samples == sample_begin wrote);
 *   state out: the state after params->samples samples (always written);
 *   accum / rgb8: the sums / the resolved image after params->samples samples (divided by params->samples), i.e.
 *   exactly what ONE call with these `samples` and sample_begin == 0 returns -- bit for bit
 *   (tests/test_gpu_render.py::test_resume_is_bit_identical).  A caller can therefore write a preview image after
 *   every chunk, stop, store `state` in a file and carry on later -- on another GPU if it likes.
 * Wavefront integrator only (RTB_ERR_INVALID with RTB_RENDER_MEGAKERNEL / RTB_RENDER_REFERENCE_CULL).
 * rtb_render(scene, p, accum, rgb8, stats) == rtb_render_resume(scene, p, NULL, accum, rgb8, stats).                  */
#define RTB_PIXEL_STATE_BYTES 32
size_t rtb_render_state_bytes(const rtb_render_params *params);
int rtb_render_resume(rtb_scene *scene, const rtb_render_params *params, void *state, float *accum, uint8_t *rgb8,
                      rtb_stats *stats);
/* the same with device buffers, stream-ordered (d_pixel_state may be NULL when sample_begin == 0) */
int rtb_render_resume_device(rtb_scene *scene, const rtb_render_params *params, void *d_pixel_state, float *d_accum,
                             uint8_t *d_rgb8, unsigned long long *d_counters, void *stream);
/* Progress line: the reference's USE_CALLBACK build calls ProgressCallback(percent, timer) once per pixel
 * (src/main.cpp:8-20, src/core/scene.cpp:45-50).  Here fn(percent, user) is called on the rendering thread, with 0 when
 * the frame starts, with the share of the frame's queue iterations that have completed (<= 99) while it runs, and
 * with 100 when it is complete; with a callback set a render call returns only when the frame is done.  fn == NULL:
 * off (default).  The return value is ignored, as the reference ignores it.                                          */
typedef int (*rtb_progress_fn)(int percent, void *user);
int rtb_scene_set_progress(rtb_scene *scene, rtb_progress_fn fn, void *user);
/* resolve only (src/core/scene.cpp:54-63): d_accum (n_pixels x 3 sums) -> d_rgb8 */
int rtb_resolve_device(const float *d_accum, size_t n_pixels, int samples, uint8_t *d_rgb8, void *stream);

/* ---- multi-GPU: the partition of Scene::render over the GPUs of one box and the combine on rank 0 ------
 * The reference renders with one OpenMP loop (src/core/scene.cpp:31) and has no exchange step; the seam is main()'s
 * `render()` followed by `draw_into()` (src/main.cpp:53-57).  One RANK per GPU -- a process (bench.py under torchrun)
 * or a thread (rtb_render_multi, the CLI) -- and one communicator per rank over NCCL (loaded with dlopen at the first
 * call; librtb.so does not link it):
 *   rank 0: rtb_comm_unique_id(id); hand the RTB_UNIQUE_ID_BYTES bytes to every rank by any means;
 *   all   : rtb_comm_create(id, world, rank, device, &comm);   (collective, like ncclCommInitRank)
 *   all   : rtb_render_dist[_device](scene, comm, frame, split, ...)  per frame (collective).
 * `frame` describes the WHOLE frame (width, height, samples, pixel0_seed, flags; row range and interleave index are
 * set per rank by the call; interleave_rows, when > 0, is the block height of the tile split; substreams, when larger
 * than the number of ranks, is the K of the sample split).
 *   RTB_SPLIT_TILE  : interleaved row blocks, block b on rank b % world; every pixel keeps its own engine stream, the
 *                     frame is bit-identical to the one-GPU frame.  Combine: ncclSend / ncclRecv of the finished RGB8 rows
 *                     to rank 0 + one kernel that puts them in place.
 *   RTB_SPLIT_SAMPLE: every rank renders the whole canvas with its share of K = max(substreams, world) engine streams per
 *                     pixel (see rtb_render_params::substreams): statistically equivalent to the reference's stream, not
 *                     bit-identical.  Combine: ncclReduce(sum) of the fp32 partial sums to rank 0, resolve (src/core/scene.cpp:54-63)
 *                     behind it.  A frame is spp/world dependent samples long on every GPU whatever the canvas size.
 * comm == NULL: one rank.  Results (d_rgb8_root / d_accum_root: the whole frame) are written on rank 0 only; d_accum_root
 * is available with the sample split only.  d_counters as in rtb_render_device, per rank.                                   */
#define RTB_UNIQUE_ID_BYTES 128
#define RTB_SPLIT_TILE 0
#define RTB_SPLIT_SAMPLE 1
typedef struct rtb_comm rtb_comm;
int rtb_comm_unique_id(void *id);
int rtb_comm_create(const void *id, int world, int rank, int device, rtb_comm **out);
void rtb_comm_destroy(rtb_comm *comm);
int rtb_comm_world(const rtb_comm *comm);
int rtb_comm_rank(const rtb_comm *comm);
/* the rtb_render_params rank `rank` of `world` renders for `frame` under `split` (what rtb_render_dist passes to rtb_render_device) */
int rtb_dist_rank_params(const rtb_render_params *frame, int split, int world, int rank, rtb_render_params *out);
int rtb_render_dist_device(rtb_scene *scene, rtb_comm *comm, const rtb_render_params *frame, int split, float *d_accum_root,
                           uint8_t *d_rgb8_root, unsigned long long *d_counters, void *stream);
/* same with a HOST image on rank 0 (one device->host copy of the finished frame); stats per rank */
int rtb_render_dist(rtb_scene *scene, rtb_comm *comm, const rtb_render_params *frame, int split, uint8_t *rgb8_root, rtb_stats *stats);
/* one process, one thread per device: scenes[r] (on any n distinct devices) renders rank r's share; the frame arrives in
 * rgb8 (host).  Communicators are created on the first call and kept.  stats: sums over the ranks (kernel_ms: the longest). */
int rtb_render_multi(rtb_scene *const *scenes, int n, const rtb_render_params *frame, int split, uint8_t *rgb8, rtb_stats *stats);

/* Canvas::write_to (src/render/canvas.h:76-89): "P6\n<W> <H>\n255\n" + W*H*3 bytes */
int rtb_write_ppm(const char *path, int width, int height, const uint8_t *rgb8);

#ifdef __cplusplus
}
#endif
#endif /* RTB_H_ */
