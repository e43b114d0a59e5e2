// Host-side flat scene: what Scene holds after its constructor (src/core/scene.cpp:197-249), as SoA arrays that
// map 1:1 onto rtb_scene_desc (include/rtb.h).  glTF loading and BVH construction stay on the host (north_star).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "../csrc/rtb_build.h"

namespace rtb {

struct HostTexture {
    int width = 0, height = 0;
    std::vector<uint8_t> rgba;   // always 4 channels (tinygltf expands every image, tiny_gltf.h:2609,2681-2683)
};

// struct Node (src/core/bvh.h:10-17) flattened: aabb min/max + {left,right,split_dim,first,count}, -1 = invalidKey
struct HostBvh {
    std::vector<float> aabb;      // n x 6
    std::vector<int32_t> links;   // n x 5
    bool built_on_gpu = false;    // the large nodes were cut by csrc/rtb_build.cu (same layout)
    GpuBuildStats gpu;
    size_t size() const { return links.size() / 5; }
};

// struct Primitive (src/geometry/primitive.h:96-134) as SoA.  The loaders fill pos / mesh / source (what the BVH build
// and the light list need); nrm / tan / uv -- the part of scene_parser.cpp:316-343 that only shading reads -- are
// assembled on demand (HostScene::assemble_attributes) for callers of the flat rtb_scene_desc, and on the GPU from
// the raw views otherwise (rtb_scene_create_raw).
struct HostTriangles {
    std::vector<float> pos;   // n x 9 : p0, U, V
    std::vector<float> nrm;   // n x 9   (empty until assembled)
    std::vector<float> tan;   // n x 12  (empty until assembled)
    std::vector<float> uv;    // n x 6   (empty until assembled)
    std::vector<int32_t> mesh;
    std::vector<uint32_t> source;   // loader-order number of the triangle (the build's permutation once build() ran)
    size_t size() const { return mesh.size(); }
    void resize_core(size_t n) { pos.resize(n * 9); mesh.resize(n); source.resize(n); }
    void permute(const std::vector<uint32_t> &order);   // new[i] = old[order[i]] for every array that is filled
};

// One glTF primitive as the file delivers it: raw accessor views (pointers into HostScene::raw_buffers).
struct RawMesh {
    size_t n_tris = 0, n_vertices = 0, first_tri = 0;
    const uint8_t *idx = nullptr;    // null: non-indexed (vertex 3i + v)
    int index_bytes = 0;             // 2 / 4
    size_t idx_stride = 0;
    const uint8_t *pos = nullptr, *nrm = nullptr, *tan = nullptr, *uv = nullptr;
    size_t pos_stride = 0, nrm_stride = 0, tan_stride = 0, uv_stride = 0;
    int normal_mode = 0;             // RTB_NORMALS_* (include/rtb.h)
    bool transform_positions = true; // false: positions are world-space already (triangle soups)
    int mesh_id = 0;
    double transform[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    double normal_transform[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    size_t vertex(size_t tri, int v) const;   // the index fetch of scene_parser.cpp:285-304
};
// scene_parser.cpp:306-343 for ONE triangle of a raw mesh (any output may be null).  pos9 = p0, U, V.
void assemble_triangle(const RawMesh &m, size_t tri, float *pos9, float *nrm9, float *tan12, float *uv6);

struct HostScene {
    HostTriangles tris;          // post-build order once build() ran
    HostBvh bvh;
    std::vector<float> mesh_material;              // n x 16 (layout in rtb.h)
    std::vector<double> mesh_normal_transform;     // n x 16
    std::vector<HostTexture> textures;
    std::vector<RawMesh> raw_meshes;               // loader order; views point into raw_buffers
    std::vector<std::vector<uint8_t>> raw_buffers; // the file's buffers, kept for the device-side assembly
    HostTriangles lights;        // emissive triangles, re-ordered by the light BVH (random.cpp:156-167)
    HostBvh light_bvh;
    float cam_pos[3] = {0, 0, 0};
    float cam_axes[9] = {1, 0, 0, 0, 1, 0, 0, 0, -1};
    float cam_fov[2] = {0.7f, 0.7f};
    float cam_tan_half_fov[2] = {0, 0};
    int width = 0, height = 0, samples = 0;
    int ray_depth = 6;           // src/io/scene_parser.cpp:17
    float max_distance = 1e9f;   // src/io/scene_parser.cpp:22
    double build_seconds = 0.0;
    std::vector<std::string> warnings;

    size_t n_meshes() const { return mesh_material.size() / 16; }
    // Scene::Scene (src/core/scene.cpp:197-249): BVH over all triangles, then the emissive copy + its BVH.
    void finalize();
    // fills tris.nrm / tan / uv (post-build order) from the raw views; idempotent
    void assemble_attributes();
};

// BVH::buildBVH (src/core/bvh.cpp:5-175): widest-axis full-sweep SAH, leaf <= 4, libstdc++ sort/partition order.
// Reorders `tris` in place exactly as the reference reorders Scene::objects.  threads<=1: sequential.
void build_bvh(HostTriangles &tris, HostBvh &out, int threads = 0);

// parse_scene_gltf (src/io/scene_parser.cpp:25-350); throws std::runtime_error like the reference.
void load_gltf(const std::string &path, int width, int height, int samples, HostScene &scene);

// PNG (image_decode.cpp, zlib) or baseline JPEG (jpeg_decode.cpp) -> RGBA8, the bytes tinygltf / stb_image give.
bool decode_image(const std::vector<uint8_t> &bytes, HostTexture &out, std::string &err);
bool decode_jpeg(const std::vector<uint8_t> &bytes, HostTexture &out, std::string &err);

}  // namespace rtb
