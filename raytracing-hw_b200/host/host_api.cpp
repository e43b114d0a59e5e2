// Host half of the C ABI (include/rtb.h): glTF load + BVH build -> rtb_scene_desc, PPM writer.
#include "../../include/rtb.h"
#include "host_scene.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <fstream>
#include <string>

struct rtb_host_scene {
    rtb::HostScene s;
    std::vector<rtb_texture_desc> tex_desc;
    std::vector<rtb_raw_mesh> raw_desc;
};

namespace rtb {
thread_local std::string g_last_error;
void set_error(const std::string &e) { g_last_error = e; }
}  // namespace rtb

extern "C" {

const char *rtb_last_error(void) { return rtb::g_last_error.c_str(); }

int rtb_host_scene_load_gltf(const char *path, int width, int height, int samples, rtb_host_scene **out) {
    if (!path || !out || width <= 0 || height <= 0 || samples <= 0) {
        rtb::set_error("rtb_host_scene_load_gltf: invalid argument");
        return RTB_ERR_INVALID;
    }
    rtb_host_scene *hs = new rtb_host_scene();
    try {
        rtb::load_gltf(path, width, height, samples, hs->s);
        hs->s.finalize();
    } catch (const std::exception &e) {
        rtb::set_error(e.what());
        delete hs;
        return RTB_ERR_IO;
    }
    *out = hs;
    return RTB_OK;
}

int rtb_host_scene_from_triangles(const float *verts9, const float *nrm9, size_t n, size_t emissive_from,
                                  const float cam_pos[3], const float cam_axes[9], float fov_x, float fov_y,
                                  float max_distance, rtb_host_scene **out) {
    if (!verts9 || !out || n == 0 || !cam_pos || !cam_axes) {
        rtb::set_error("rtb_host_scene_from_triangles: invalid argument");
        return RTB_ERR_INVALID;
    }
    rtb_host_scene *hs = new rtb_host_scene();
    rtb::HostScene &s = hs->s;
    try {
        // the caller's arrays are the "file": kept as raw views (two meshes: plain / emissive), positions assembled here
        // for the BVH build, normals where they are consumed (HostScene::assemble_attributes / the GPU)
        s.raw_buffers.resize(nrm9 ? 2 : 1);
        s.raw_buffers[0].assign((const uint8_t *) verts9, (const uint8_t *) verts9 + n * 36);
        if (nrm9) s.raw_buffers[1].assign((const uint8_t *) nrm9, (const uint8_t *) nrm9 + n * 36);
        const size_t cut = std::min(emissive_from, n);
        for (int part = 0; part < 2; ++part) {
            rtb::RawMesh rm;
            rm.first_tri = part == 0 ? 0 : cut;
            rm.n_tris = part == 0 ? cut : n - cut;
            rm.n_vertices = rm.n_tris * 3;
            rm.pos = s.raw_buffers[0].data() + rm.first_tri * 36;
            rm.pos_stride = 12;
            if (nrm9) { rm.nrm = s.raw_buffers[1].data() + rm.first_tri * 36; rm.nrm_stride = 12; }
            rm.normal_mode = nrm9 ? RTB_NORMALS_VERBATIM : RTB_NORMALS_FLAT;
            rm.transform_positions = false;
            rm.mesh_id = part;
            s.raw_meshes.push_back(rm);
        }
        s.tris.resize_core(n);
#pragma omp parallel for schedule(static)
        for (long long i = 0; i < (long long) n; ++i) {
            const int part = (size_t) i >= cut ? 1 : 0;
            rtb::assemble_triangle(s.raw_meshes[(size_t) part], (size_t) i - s.raw_meshes[(size_t) part].first_tri, &s.tris.pos[(size_t) i * 9], nullptr, nullptr, nullptr);
            s.tris.mesh[(size_t) i] = part;
            s.tris.source[(size_t) i] = (uint32_t) i;
        }
        const float m0[16] = {1.f, 1.f, 1.f, 1.f, 1.f, 0.f, 0.f, 0.f, 0.f, 1.f, -1.f, -1.f, -1.f, -1.f, 0.f, 0.f};
        float m1[16];
        std::memcpy(m1, m0, sizeof(m0));
        m1[5] = m1[6] = m1[7] = 1.f;
        s.mesh_material.insert(s.mesh_material.end(), m0, m0 + 16);
        s.mesh_material.insert(s.mesh_material.end(), m1, m1 + 16);
        const double I[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
        for (int r = 0; r < 2; ++r) s.mesh_normal_transform.insert(s.mesh_normal_transform.end(), I, I + 16);
        std::memcpy(s.cam_pos, cam_pos, 12);
        std::memcpy(s.cam_axes, cam_axes, 36);
        s.cam_fov[0] = fov_x;
        s.cam_fov[1] = fov_y;
        s.cam_tan_half_fov[0] = std::tan(fov_x / 2);
        s.cam_tan_half_fov[1] = std::tan(fov_y / 2);
        s.max_distance = max_distance;
        s.finalize();
    } catch (const std::exception &e) {
        rtb::set_error(e.what());
        delete hs;
        return RTB_ERR_INVALID;
    }
    *out = hs;
    return RTB_OK;
}

namespace {
void fill_desc(rtb_host_scene *hs, rtb_scene_desc *d, bool attributes);
}

int rtb_host_scene_desc(const rtb_host_scene *hs_c, rtb_scene_desc *d) {
    if (!hs_c || !d) {
        rtb::set_error("rtb_host_scene_desc: invalid argument");
        return RTB_ERR_INVALID;
    }
    rtb_host_scene *hs = const_cast<rtb_host_scene *>(hs_c);
    try {
        hs->s.assemble_attributes();   // the flat description carries host-assembled normals / tangents / texcoords
    } catch (const std::exception &e) {
        rtb::set_error(e.what());
        return RTB_ERR_INVALID;
    }
    fill_desc(hs, d, true);
    return RTB_OK;
}

namespace {
void fill_desc(rtb_host_scene *hs, rtb_scene_desc *d, bool attributes) {
    const rtb::HostScene &s = hs->s;
    std::memset(d, 0, sizeof(*d));
    d->n_nodes = (uint32_t) s.bvh.size();
    d->node_aabb = s.bvh.aabb.data();
    d->node_links = s.bvh.links.data();
    d->n_tris = (uint32_t) s.tris.size();
    d->tri_pos = s.tris.pos.data();
    d->tri_nrm = attributes ? s.tris.nrm.data() : nullptr;
    d->tri_tan = attributes ? s.tris.tan.data() : nullptr;
    d->tri_uv = attributes ? s.tris.uv.data() : nullptr;
    d->tri_mesh = s.tris.mesh.data();
    d->n_meshes = (uint32_t) s.n_meshes();
    d->mesh_material = s.mesh_material.data();
    d->mesh_normal_transform = s.mesh_normal_transform.data();
    hs->tex_desc.resize(s.textures.size());
    for (size_t i = 0; i < s.textures.size(); ++i)
        hs->tex_desc[i] = {s.textures[i].width, s.textures[i].height, s.textures[i].rgba.data()};
    d->n_textures = (uint32_t) s.textures.size();
    d->textures = hs->tex_desc.data();
    d->n_lights = (uint32_t) s.lights.size();
    d->light_pos = s.lights.pos.data();
    d->n_light_nodes = (uint32_t) s.light_bvh.size();
    d->light_node_aabb = s.light_bvh.aabb.data();
    d->light_node_links = s.light_bvh.links.data();
    std::memcpy(d->cam_pos, s.cam_pos, 12);
    std::memcpy(d->cam_axes, s.cam_axes, 36);
    d->cam_tan_half_fov[0] = s.cam_tan_half_fov[0];
    d->cam_tan_half_fov[1] = s.cam_tan_half_fov[1];
    d->ray_depth = s.ray_depth;
    d->max_distance = s.max_distance;
    d->bg_color[0] = d->bg_color[1] = d->bg_color[2] = 0.f;
}
}  // namespace

int rtb_host_scene_raw_geometry(const rtb_host_scene *hs_c, rtb_scene_desc *desc, rtb_raw_geometry *raw) {
    if (!hs_c || !raw) {
        rtb::set_error("rtb_host_scene_raw_geometry: invalid argument");
        return RTB_ERR_INVALID;
    }
    rtb_host_scene *hs = const_cast<rtb_host_scene *>(hs_c);
    const rtb::HostScene &s = hs->s;
    if (desc) fill_desc(hs, desc, false);
    hs->raw_desc.clear();
    for (const rtb::RawMesh &m : s.raw_meshes) {
        rtb_raw_mesh r;
        std::memset(&r, 0, sizeof(r));
        r.n_tris = (uint32_t) m.n_tris;
        r.n_vertices = (uint32_t) m.n_vertices;
        r.indices = m.idx;
        r.index_bytes = m.idx ? m.index_bytes : 0;
        r.index_stride = (int32_t) m.idx_stride;
        r.position = m.pos; r.position_stride = (int32_t) m.pos_stride;
        r.normal = m.nrm; r.normal_stride = (int32_t) m.nrm_stride;
        r.tangent = m.tan; r.tangent_stride = (int32_t) m.tan_stride;
        r.texcoord = m.uv; r.texcoord_stride = (int32_t) m.uv_stride;
        r.normal_mode = m.normal_mode;
        r.flags = m.transform_positions ? 0 : RTB_RAW_WORLD_POSITIONS;
        r.mesh_id = m.mesh_id;
        std::memcpy(r.transform, m.transform, sizeof(r.transform));
        hs->raw_desc.push_back(r);
    }
    raw->n_meshes = (uint32_t) hs->raw_desc.size();
    raw->meshes = hs->raw_desc.data();
    raw->tri_source = s.tris.source.data();
    return RTB_OK;
}

int rtb_host_scene_bvh_stats(const rtb_host_scene *hs, uint64_t *n_nodes, uint64_t counts5[5], uint64_t *max_leaf) {
    if (!hs) return RTB_ERR_INVALID;
    const rtb::HostBvh &b = hs->s.bvh;
    uint64_t c[5] = {0, 0, 0, 0, 0}, mx = 0;
    for (size_t i = 0; i < b.size(); ++i) {
        const uint64_t k = (uint64_t) b.links[i * 5 + 4];
        if (k < 5) ++c[k];
        if (k > mx) mx = k;
    }
    if (n_nodes) *n_nodes = b.size();
    if (counts5) std::memcpy(counts5, c, sizeof(c));
    if (max_leaf) *max_leaf = mx;
    return RTB_OK;
}

double rtb_host_scene_build_seconds(const rtb_host_scene *hs) { return hs ? hs->s.build_seconds : 0.0; }

int rtb_host_scene_build_info(const rtb_host_scene *hs, double out8[8]) {
    if (!hs || !out8) return RTB_ERR_INVALID;
    const rtb::HostBvh &b = hs->s.bvh;
    out8[0] = b.built_on_gpu ? 1.0 : 0.0;
    out8[1] = b.gpu.levels;
    out8[2] = (double) b.gpu.rounds;
    out8[3] = (double) b.gpu.heap_ranges;
    out8[4] = (double) b.gpu.host_fallback_nodes;
    out8[5] = b.gpu.seconds;
    out8[6] = b.gpu.kernel_seconds;
    out8[7] = 0.0;
    return RTB_OK;
}

int rtb_build_bvh(const float *tri_pos9, size_t n, uint32_t *order, float *node_aabb, int32_t *node_links, size_t node_capacity,
                  size_t *n_nodes, double out8[8]) {
    if (!tri_pos9 || n == 0 || !order || !node_aabb || !node_links || !n_nodes) {
        rtb::set_error("rtb_build_bvh: invalid argument");
        return RTB_ERR_INVALID;
    }
    try {
        rtb::HostTriangles t;
        t.resize_core(n);
        std::memcpy(t.pos.data(), tri_pos9, n * 9 * sizeof(float));
        for (size_t i = 0; i < n; ++i) { t.mesh[i] = 0; t.source[i] = (uint32_t) i; }
        rtb::HostBvh b;
        rtb::build_bvh(t, b);
        *n_nodes = b.size();
        if (out8) {
            const double o[8] = {b.built_on_gpu ? 1.0 : 0.0, (double) b.gpu.levels, (double) b.gpu.rounds, (double) b.gpu.heap_ranges,
                                 (double) b.gpu.host_fallback_nodes, b.gpu.seconds, b.gpu.kernel_seconds, 0.0};
            std::memcpy(out8, o, sizeof o);
        }
        if (b.size() > node_capacity) {
            rtb::set_error("rtb_build_bvh: node arrays too small (" + std::to_string(b.size()) + " nodes)");
            return RTB_ERR_INVALID;
        }
        std::memcpy(order, t.source.data(), n * sizeof(uint32_t));
        std::memcpy(node_aabb, b.aabb.data(), b.aabb.size() * sizeof(float));
        std::memcpy(node_links, b.links.data(), b.links.size() * sizeof(int32_t));
        return RTB_OK;
    } catch (const std::exception &e) {
        rtb::set_error(std::string("rtb_build_bvh: ") + e.what());
        return RTB_ERR_INVALID;
    }
}

void rtb_host_scene_destroy(rtb_host_scene *hs) { delete hs; }

int rtb_write_ppm(const char *path, int width, int height, const uint8_t *rgb8) {
    if (!path || !rgb8 || width <= 0 || height <= 0) return RTB_ERR_INVALID;
    std::ofstream out(path, std::ios_base::binary);
    if (!out) {
        rtb::set_error("File open error");   // canvas.h:79-80
        return RTB_ERR_IO;
    }
    out << "P6\n" << width << " " << height << "\n" << "255" << "\n";
    out.write((const char *) rgb8, (std::streamsize) width * height * 3);
    out.close();
    return out ? RTB_OK : RTB_ERR_IO;
}

}  // extern "C"
