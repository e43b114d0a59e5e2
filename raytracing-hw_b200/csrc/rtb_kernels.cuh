// Kernel launchers (defined in rtb_kernels.cu / rtb_integrator.cu), called by the C ABI in rtb_api.cu.
#pragma once
#include <vector>
#include "rtb_device.cuh"

namespace rtb {

void launch_closest_hit(const DScene &sc, const float *d_rays6, size_t n, int cull_mode, int *d_id, float *d_t, float *d_u,
                        float *d_v, uint8_t *d_inside, unsigned long long *d_counters, unsigned *d_queue_head, cudaStream_t stream);
// the same launch with the ray-queue cursor taken from a table inside the library image (no allocation anywhere on the
// path); `device` = the current device.  rtb_kernels.cu
cudaError_t launch_closest_hit_static(const DScene &sc, int device, const float *d_rays6, size_t n, int cull_mode, int *d_id, float *d_t,
                                      float *d_u, float *d_v, uint8_t *d_inside, unsigned long long *d_counters, cudaStream_t stream);
int persistent_grid(int threads_per_block, int blocks_per_sm);
constexpr int kDefaultTraceMode = 4;       // rtb_trace.cuh: 3 / 4 / 5 = one / two / three steps per refill vote (RTB_TRACE_MODE)
int trace_mode();
constexpr int kDefaultTraceBlocks = 10;   // sweeps: profiles/r01d_closest_hit_blocks_sweep.jsonl (first form), r01i_mode3_sweep.txt, r01l_leaf_link_sweep.txt (form 3)
int persistent_blocks_per_sm();   // CTAs per SM of the persistent queue kernels (RTB_TRACE_BLOCKS overrides)

int measure_l2(size_t bytes, double *stream_gbs, double *gather64_gbs);   // rtb_kernels.cu: L2 bandwidth probes
void launch_light_pdf(const DScene &sc, const float *d_rays6, size_t n, float *d_pdf, cudaStream_t stream);

struct RenderLaunch {
    int width, height, samples;
    int row_begin, row_end;
    int il_count, il_index, il_rows;    // row-block interleave (il_count <= 1: contiguous); see rtb_render_params
    int local_rows;                     // rows this launch renders
    unsigned pixel0_seed;
    float *d_accum;                     // (row_end-row_begin) x width x 3, may be null
    uint8_t *d_rgb8;                    // same extent, may be null
    unsigned long long *d_counters;     // 6 x u64 {rays, light queries, box, tri, light box, light tri} or null
    bool count_tests;                   // also count box / triangle tests (slower instantiation)
    int cull_mode;                      // RTB_CULL_*
    // progressive rendering (wavefront integrator only, rtb_render_resume): the call renders samples
    // [sample_begin, samples) of every pixel.  d_pixel_state: 2 x uint4 per local pixel {engine value, cached Marsaglia
    // half (bits), has-cached flag, samples done}, {sum r, g, b (bits), 0}; read when sample_begin > 0, written (if not
    // null) when a pixel has had its last sample.  A pixel's samples are one sequential engine stream
    // (src/core/scene.cpp:34-42), so this is everything a later call needs to carry on bit for bit.
    int sample_begin;
    uint4 *d_pixel_state;
    // sample-range split (rtb_render_params::substreams): this launch runs streams [sub_begin, sub_end) of K = substreams
    // per pixel; a stream is a slot, slot = (stream - sub_begin) * plane_slots + the pixel's slot in its 8x4 tile
    int substreams, sub_begin, sub_end;
    unsigned plane_slots;               // slots of one stream plane (tiles x 32); set by wf_slots()
};
void launch_render(const DScene &sc, const RenderLaunch &p, cudaStream_t stream);
// wavefront (ray-queue) integrator, rtb_wavefront.cu; `state` = wf_alloc_bytes() bytes of device memory
unsigned wf_plane_slots(const RenderLaunch &p);   // slots of one stream plane
unsigned wf_slots(const RenderLaunch &p);         // all planes of the launch
size_t wf_alloc_bytes(unsigned n_slots, int depth_cap, int bvh_depth);
// stage_events (may be null): filled with events bracketing every k_wf_trace launch, see rtb_scene_stage_ms()
constexpr int kWfMaxGroups = 16;
struct WfStreams {   // owned by the scene: one stream per slot group + fork / join events (null: one group, caller's stream)
    cudaStream_t stream[kWfMaxGroups];
    cudaEvent_t fork, join[kWfMaxGroups];
    // the captured block of iterations (see launch_render_wavefront) and the launch parameters it was captured for
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t graph_exec = nullptr;
    std::vector<unsigned char> graph_key;
    unsigned long long graph_launches = 0;   // kernels per replay
    // progress line (rtb_scene_set_progress): called on the rendering thread with 0..100 while the frame's iterations
    // complete (the reference's USE_CALLBACK build: src/main.cpp:8-20, scene.cpp:45-50)
    int (*progress)(int, void *) = nullptr;
    void *progress_user = nullptr;
};
// returns 0, 1 (a stream synchronisation failed) or 2 (slots still live after the bounded number of extra rounds)
int launch_render_wavefront(const DScene &sc, const RenderLaunch &p, void *state, cudaStream_t stream, unsigned long long *launches,
                            std::vector<cudaEvent_t> *stage_events, WfStreams *aux);
void launch_resolve(const float *d_accum, size_t n_pixels, int samples, uint8_t *d_rgb8, cudaStream_t stream);

// device-side scene assembly (rtb_assemble.cu / rtb_assemble.cuh)
struct DRawMesh;
struct NodeItem;
struct BigLeafIndex;
void launch_texture_quads(const DTexture &t, cudaStream_t st);
void launch_fuse_textures(const DFused &f, const uchar4 *base, const uchar4 *mr, const uchar4 *nrm, cudaStream_t st);   // fills f.data   // fills t.quads from t.texels (rtb_assemble.cu)
void launch_assemble_triangles(const DRawMesh *d_meshes, const uint32_t *d_mesh_first, int n_meshes, const uint32_t *d_tri_source, uint32_t n_tris,
                               float4 *d_tris, float4 *d_shade, float4 *d_tan, int *d_error, cudaStream_t stream);
int launch_convert_nodes(uint32_t n_nodes, uint32_t n_internal, uint32_t n_tris, const float *d_aabb, const int32_t *d_links, const BigLeafIndex &big,
                         DNode *d_nodes, NodeItem *d_items, unsigned *d_counts, int *d_error, cudaStream_t stream);

}  // namespace rtb
