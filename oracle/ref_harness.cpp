// TEST INFRASTRUCTURE ONLY -- never linked into, imported by, or executed from the product path.
//
// Harness around the UNMODIFIED reference (Korinin38/raytracing-hw).  It is compiled by
// oracle/Makefile against the sources where they lie under $REF (/root/reference); nothing from the
// reference is copied into this repository.  The output (oracle/_ref/libref_oracle.so) is git-ignored
// and travels to the GPU box as a prebuilt binary.
//
// What it gives the tests / bench.py cpu_baseline leg:
//   * the reference's own BVH layout (nodes + post-build triangle order)    -> layout parity
//   * the reference's own BVH::intersect answers for a ray batch            -> closest-hit parity
//   * the reference's own Scene::render (OpenMP, as shipped)                -> CPU baseline timing
//   * a deterministic render that drives the reference's own Camera::cast_in_pixel and
//     Scene::intersect pixel by pixel, resetting the file-static normal distribution
//     (src/utils/random.cpp:22) per pixel and seeding pixel 0 with a fixed constant
//      -> image parity target for the CPU twin and the GPU integrator.
//
// random.cpp is #included (not linked) so that the harness can reach its file-static `normDist`.
// Build flags: -fno-access-control so private members (scene_distribution_, light BVH) can be dumped.

#include <cstdint>
#include <functional>

#include "utils/random.cpp"   // from $REF/src (see Makefile -I); gives access to rng::normDist

#include <core/scene.h>
#include <io/scene_parser.h>

#include <sys/mman.h>
#include <sys/wait.h>
#include <unistd.h>
#include <omp.h>

#include <chrono>
#include <cstring>
#include <iostream>
#include <sstream>

namespace {

struct RefScene {
    Scene *scene = nullptr;
    std::string error;
};

thread_local std::string g_err;

struct CoutSilencer {
    std::streambuf *old;
    std::ostringstream sink;
    explicit CoutSilencer(bool on) : old(nullptr) { if (on) old = std::cout.rdbuf(sink.rdbuf()); }
    ~CoutSilencer() { if (old) std::cout.rdbuf(old); }
};

const rng::ManyLightsDistribution &lights_of(const Scene *s) {
    return s->scene_distribution_->light;
}

void fill_nodes(const BVH &bvh, float *aabb6, int64_t *links5) {
    for (size_t i = 0; i < bvh.nodes.size(); ++i) {
        const Node &n = bvh.nodes[i];
        for (int k = 0; k < 3; ++k) {
            aabb6[i * 6 + k] = n.aabb.min[k];
            aabb6[i * 6 + 3 + k] = n.aabb.max[k];
        }
        links5[i * 5 + 0] = (int64_t) n.left;
        links5[i * 5 + 1] = (int64_t) n.right;
        links5[i * 5 + 2] = (int64_t) n.split_dim;
        links5[i * 5 + 3] = (int64_t) n.first_primitive_id;
        links5[i * 5 + 4] = (int64_t) n.primitive_count;
    }
}

void fill_tris(const std::vector<Primitive> &objs, float *pos9, float *nrm9, float *tan12, float *uv6, int32_t *mesh_id) {
    for (size_t i = 0; i < objs.size(); ++i) {
        const Primitive &p = objs[i];
        for (int v = 0; v < 3; ++v) {
            for (int k = 0; k < 3; ++k) {
                if (pos9) pos9[i * 9 + v * 3 + k] = p.position[v][k];
                if (nrm9) nrm9[i * 9 + v * 3 + k] = p.normal[v][k];
            }
            for (int k = 0; k < 4; ++k)
                if (tan12) tan12[i * 12 + v * 4 + k] = p.tangent[v][k];
            for (int k = 0; k < 2; ++k)
                if (uv6) uv6[i * 6 + v * 2 + k] = p.texcoord[v][k];
        }
        if (mesh_id) mesh_id[i] = p.mesh_id;
    }
}

}  // namespace

extern "C" {

const char *ref_last_error() { return g_err.c_str(); }

// parse_scene_gltf (src/io/scene_parser.cpp:25) + Scene ctor (src/core/scene.cpp:197): loads, builds both BVHs.
void *ref_scene_from_gltf(const char *path, int width, int height, int samples, int quiet) {
    try {
        CoutSilencer hush(quiet != 0);
        Scene *s = new Scene(parse_scene_gltf(path, width, height, samples));
        return s;
    } catch (const std::exception &e) {
        g_err = e.what();
        return nullptr;
    }
}

// Build a Scene directly from a triangle soup (3 world-space vertices per triangle, 9 floats each), one
// mesh with the default Material unless `emissive_from` < n: triangles [emissive_from, n) get a second mesh
// whose emission is (1,1,1).  Camera: given position / axes / fov.  Mirrors what scene_parser.cpp:306-313
// stores (p0, p1-p0, p2-p0) and the flat normal it would get from a NORMAL-less primitive is replaced by the
// supplied per-vertex normals (nrm9 may be null -> geometric normal of the triangle, normalised).
void *ref_scene_from_triangles(const float *verts9, const float *nrm9, size_t n, size_t emissive_from,
                               int width, int height, int samples,
                               const float *cam_pos3, const float *cam_axes9, float fov_x, float fov_y,
                               float max_distance, int quiet) {
    try {
        CoutSilencer hush(quiet != 0);
        std::vector<Mesh> meshes(2);
        meshes[0].normal_transform = matrix4d::eye();
        meshes[1].normal_transform = matrix4d::eye();
        meshes[0].material.metallic = 0.f;
        meshes[1].material.metallic = 0.f;
        meshes[1].material.emission = {1.f, 1.f, 1.f};
        std::vector<Primitive> objs(n);
        for (size_t i = 0; i < n; ++i) {
            Primitive &p = objs[i];
            vector3f v[3];
            for (int k = 0; k < 3; ++k)
                v[k] = {verts9[i * 9 + k * 3], verts9[i * 9 + k * 3 + 1], verts9[i * 9 + k * 3 + 2]};
            p.position[0] = v[0];
            p.position[1] = v[1] - v[0];
            p.position[2] = v[2] - v[0];
            for (int k = 0; k < 3; ++k) {
                if (nrm9)
                    p.normal[k] = {nrm9[i * 9 + k * 3], nrm9[i * 9 + k * 3 + 1], nrm9[i * 9 + k * 3 + 2]};
                else
                    p.normal[k] = ::normal(cross(p.position[1], p.position[2]));
            }
            p.mesh_id = (i >= emissive_from) ? 1 : 0;
        }
        vector3f axes[3];
        for (int k = 0; k < 3; ++k) axes[k] = {cam_axes9[k * 3], cam_axes9[k * 3 + 1], cam_axes9[k * 3 + 2]};
        camera_uniq_ptr cam = std::make_unique<Camera>(vector2i{width, height},
                                                       vector3f{cam_pos3[0], cam_pos3[1], cam_pos3[2]}, axes, fov_x, fov_y);
        std::vector<Texture> textures;
        Scene *s = new Scene(cam, meshes, objs, textures, 6, samples, max_distance);
        return s;
    } catch (const std::exception &e) {
        g_err = e.what();
        return nullptr;
    }
}

void ref_scene_destroy(void *h) { delete (Scene *) h; }

size_t ref_num_nodes(void *h) { return ((Scene *) h)->bvh.nodes.size(); }
size_t ref_num_tris(void *h) { return ((Scene *) h)->objects.size(); }
size_t ref_num_meshes(void *h) { return ((Scene *) h)->meshes.size(); }
size_t ref_num_textures(void *h) { return ((Scene *) h)->textures.size(); }
size_t ref_num_lights(void *h) { return lights_of((Scene *) h).objects.size(); }
size_t ref_num_light_nodes(void *h) { return lights_of((Scene *) h).bvh.nodes.size(); }

// nodes in reference order (src/core/bvh.h:10-17): aabb6 = min xyz, max xyz; links5 = left,right,split_dim,first,count
// (size_t invalidKey is reported as -1).
void ref_get_nodes(void *h, float *aabb6, int64_t *links5) { fill_nodes(((Scene *) h)->bvh, aabb6, links5); }
void ref_get_light_nodes(void *h, float *aabb6, int64_t *links5) { fill_nodes(lights_of((Scene *) h).bvh, aabb6, links5); }

// triangles in POST-BUILD order (BVH::buildBVH reorders Scene::objects in place, src/core/bvh.cpp:166).
void ref_get_tris(void *h, float *pos9, float *nrm9, float *tan12, float *uv6, int32_t *mesh_id) {
    fill_tris(((Scene *) h)->objects, pos9, nrm9, tan12, uv6, mesh_id);
}
void ref_get_light_tris(void *h, float *pos9, float *nrm9, float *tan12, float *uv6, int32_t *mesh_id) {
    fill_tris(lights_of((Scene *) h).objects, pos9, nrm9, tan12, uv6, mesh_id);
}

// material as 16 floats: ior, alpha, base rgb, emission rgb, metallic, roughness2, then 4 texture ids, pad 2
void ref_get_meshes(void *h, float *mat16, double *normal_transform16) {
    Scene *s = (Scene *) h;
    for (size_t i = 0; i < s->meshes.size(); ++i) {
        const Material &m = s->meshes[i].material;
        float *o = mat16 + i * 16;
        o[0] = m.ior; o[1] = m.alpha;
        o[2] = m.base_color.x; o[3] = m.base_color.y; o[4] = m.base_color.z;
        o[5] = m.emission.x; o[6] = m.emission.y; o[7] = m.emission.z;
        o[8] = m.metallic; o[9] = m.roughness2;
        o[10] = (float) m.base_color_i; o[11] = (float) m.normal_i;
        o[12] = (float) m.metallic_roughness_i; o[13] = (float) m.emission_i;
        o[14] = o[15] = 0.f;
        for (int k = 0; k < 16; ++k) normal_transform16[i * 16 + k] = s->meshes[i].normal_transform.data[k];
    }
}

void ref_get_texture_info(void *h, size_t i, int *w, int *hgt, int *channels) {
    const Texture &t = ((Scene *) h)->textures[i];
    *w = t.width; *hgt = t.height; *channels = t.channels;
}
void ref_get_texture_data(void *h, size_t i, uint8_t *out) {
    const Texture &t = ((Scene *) h)->textures[i];
    std::memcpy(out, t.data.data(), t.data.size());
}

// camera: pos3, axes9 (right, up, forward), fov2 (x, y), size2, then max_distance / ray_depth / samples
void ref_get_camera(void *h, float *pos3, float *axes9, float *fov2, int *size2, float *max_distance, int *ray_depth, int *samples) {
    Scene *s = (Scene *) h;
    vector3f p = s->camera->get_position();
    for (int k = 0; k < 3; ++k) pos3[k] = p[k];
    for (int a = 0; a < 3; ++a) {
        vector3f ax = s->camera->get_axis(a);
        for (int k = 0; k < 3; ++k) axes9[a * 3 + k] = ax[k];
    }
    fov2[0] = s->camera->get_fov().x; fov2[1] = s->camera->get_fov().y;
    size2[0] = s->camera->canvas.width(); size2[1] = s->camera->canvas.height();
    *max_distance = s->max_distance;
    *ray_depth = s->ray_depth;
    *samples = s->samples;
}

// The reference's own closest hit (BVH::intersect, src/core/bvh.cpp:239-243) for a batch of rays given as
// origin xyz + un-normalised direction xyz; Ray::Ray (src/geometry/primitive.cpp:12-15) normalises.
// id = -1 on a miss.  Returns seconds spent inside the loop.
double ref_closest_hit(void *h, const float *rays6, size_t n, int64_t *id, float *t, float *u, float *v,
                       uint8_t *inside, int threads) {
    Scene *s = (Scene *) h;
    if (threads <= 0) threads = omp_get_max_threads();
    auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for schedule(dynamic, 1024) num_threads(threads)
    for (size_t i = 0; i < n; ++i) {
        Ray r({rays6[i * 6], rays6[i * 6 + 1], rays6[i * 6 + 2]}, {rays6[i * 6 + 3], rays6[i * 6 + 4], rays6[i * 6 + 5]});
        Intersection a = s->bvh.intersect(s->objects, r);
        if (a) {
            id[i] = (int64_t) a.object_id;
            t[i] = a.distance;
            if (u) u[i] = a.local_coords.x;
            if (v) v[i] = a.local_coords.y;
            if (inside) inside[i] = a.inside ? 1 : 0;
        } else {
            id[i] = -1;
            t[i] = 0.f;
            if (u) u[i] = 0.f;
            if (v) v[i] = 0.f;
            if (inside) inside[i] = 0;
        }
    }
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

// The reference's light pdf (ManyLightsDistribution::pdf, src/utils/random.cpp:179-188) for a batch of
// (point, direction) pairs; direction is passed through Ray::Ray inside, as in the reference.
void ref_light_pdf(void *h, const float *rays6, size_t n, float *pdf) {
    Scene *s = (Scene *) h;
    const auto &l = lights_of(s);
    for (size_t i = 0; i < n; ++i) {
        vector3f p{rays6[i * 6], rays6[i * 6 + 1], rays6[i * 6 + 2]};
        vector3f d{rays6[i * 6 + 3], rays6[i * 6 + 4], rays6[i * 6 + 5]};
        pdf[i] = l.pdf(p, d);
    }
}

// Scene::render exactly as shipped (OpenMP, nondeterministic: SURVEY.md section 0.5).  Times render() only.
double ref_render(void *h, uint8_t *rgb, int threads, int quiet) {
    Scene *s = (Scene *) h;
    if (threads > 0) omp_set_num_threads(threads);
    CoutSilencer hush(quiet != 0);
    auto t0 = std::chrono::steady_clock::now();
    s->render();
    double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    const int W = s->camera->canvas.width(), H = s->camera->canvas.height();
    if (rgb)
        for (int j = 0; j < H; ++j)
            for (int i = 0; i < W; ++i) {
                vector3si c = s->camera->canvas.get({i, j});
                rgb[(j * W + i) * 3 + 0] = c.x; rgb[(j * W + i) * 3 + 1] = c.y; rgb[(j * W + i) * 3 + 2] = c.z;
            }
    return sec;
}

// Deterministic drive of the reference's own per-pixel code (the body of the loop at src/core/scene.cpp:32-43
// calling the reference's Camera::cast_in_pixel and Scene::intersect), with
//   * rng::normDist.reset() at the start of every pixel (per-pixel normal-distribution state),
//   * pixel 0 seeded with `pixel0_seed` instead of std::random_device,
// and the resolve of src/core/scene.cpp:54-63.  Work is split over `nproc` forked processes (pixel-interleaved)
// because normDist is a single file-static object.  Rows [row_begin, row_end) only (a band of a BASELINE-sized frame is
// what the parity tests can afford); accum ((row_end-row_begin)*W*3 floats, pre-resolve sums) may be null.
double ref_render_deterministic_rows(void *h, unsigned pixel0_seed, int nproc, int row_begin, int row_end, float *accum, uint8_t *rgb) {
    Scene *s = (Scene *) h;
    const int W = s->camera->canvas.width(), H = s->camera->canvas.height();
    if (row_begin < 0 || row_end > H || row_begin > row_end) return -1.0;
    const size_t npix = (size_t) W * (size_t) (row_end - row_begin);
    if (npix == 0) return 0.0;
    const size_t bytes = npix * 3 * sizeof(float);
    float *shared = (float *) mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0);
    if (shared == MAP_FAILED) return -1.0;
    if (nproc <= 0) nproc = (int) sysconf(_SC_NPROCESSORS_ONLN);
    auto t0 = std::chrono::steady_clock::now();
    std::vector<pid_t> kids;
    for (int p = 0; p < nproc; ++p) {
        pid_t pid = fork();
        if (pid == 0) {
            uniform_float_d offset(-0.5f, 0.5f);
            // pixels are independent (own engine, own normal-distribution state): deal them out in runs of 8 so that a
            // band of a few rows still keeps every process busy
            for (size_t q = (size_t) p * 8; q < npix; q += (size_t) nproc * 8) {
                for (size_t idx = q; idx < q + 8 && idx < npix; ++idx) {
                    const int j = row_begin + (int) (idx / (size_t) W), i = (int) (idx % (size_t) W);
                    rng::normDist.reset();
                    size_t seed = (size_t) j * W + i;
                    Engine e(seed != 0 ? (unsigned) seed : pixel0_seed);
                    vector3f sum{0.f, 0.f, 0.f};
                    for (int k = 0; k < s->samples; ++k) {
                        vector2f pix_offset{offset(e), offset(e)};
                        Ray r = s->camera->cast_in_pixel({i, j}, pix_offset);
                        r.power = s->ray_depth;
                        auto inter = s->intersect(r, e);
                        sum += inter.color;
                    }
                    float *o = shared + idx * 3;
                    o[0] = sum.x; o[1] = sum.y; o[2] = sum.z;
                }
            }
            _exit(0);
        }
        kids.push_back(pid);
    }
    int bad = 0;
    for (pid_t k : kids) {
        int st = 0;
        waitpid(k, &st, 0);
        if (!WIFEXITED(st) || WEXITSTATUS(st) != 0) bad = 1;
    }
    double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (accum) std::memcpy(accum, shared, bytes);
    if (rgb) {
        const float normalizer = 1.f / (float) s->samples;
        for (size_t i = 0; i < npix; ++i) {
            vector3f color{shared[i * 3], shared[i * 3 + 1], shared[i * 3 + 2]};
            color *= normalizer;
            color = aces_tonemap(color);
            color = pow(color, s->gamma_);
            vector3si c = normal_to_ch8bit(color);
            rgb[i * 3] = c.x; rgb[i * 3 + 1] = c.y; rgb[i * 3 + 2] = c.z;
        }
    }
    munmap(shared, bytes);
    return bad ? -1.0 : sec;
}
double ref_render_deterministic(void *h, unsigned pixel0_seed, int nproc, float *accum, uint8_t *rgb) {
    Scene *s = (Scene *) h;
    return ref_render_deterministic_rows(h, pixel0_seed, nproc, 0, s->camera->canvas.height(), accum, rgb);
}

// One path through the reference's Scene::intersect for an explicit ray and engine state, with a fresh
// normal-distribution state.  Returns the radiance and the engine state afterwards (stream-consumption check).
void ref_trace_path(void *h, const float *o3, const float *d3, unsigned engine_seed, float *rgb3, unsigned *engine_after) {
    Scene *s = (Scene *) h;
    rng::normDist.reset();
    Engine e(engine_seed);
    Ray r({o3[0], o3[1], o3[2]}, {d3[0], d3[1], d3[2]});
    r.power = s->ray_depth;
    auto inter = s->intersect(r, e);
    rgb3[0] = inter.color.x; rgb3[1] = inter.color.y; rgb3[2] = inter.color.z;
    std::ostringstream os;
    os << e;
    *engine_after = (unsigned) std::stoul(os.str());
}

// libstdc++ <random> known-answer helpers
unsigned ref_minstd_first(unsigned seed) { Engine e(seed); return (unsigned) e(); }
float ref_uniform_first(unsigned seed, float a, float b) { Engine e(seed); uniform_float_d d(a, b); return d(e); }
void ref_normal_seq(unsigned seed, int n, float *out) {
    Engine e(seed);
    normal_d d(0, 1.f);
    for (int i = 0; i < n; ++i) out[i] = d(e);
}
void ref_libm(int which, const float *in, size_t n, float *out, float arg) {
    for (size_t i = 0; i < n; ++i) {
        switch (which) {
            case 0: out[i] = std::log(in[i]); break;
            case 1: out[i] = std::pow(in[i], arg); break;
            case 2: out[i] = (float) std::pow(1.f - in[i], 5); break;
            case 3: out[i] = std::tan(in[i]); break;
            default: out[i] = 0.f;
        }
    }
}

}  // extern "C"
