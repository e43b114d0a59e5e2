// glTF 2.0 front end: restates parse_scene_gltf (src/io/scene_parser.cpp:25-350) -- node-matrix propagation in
// array order, first-camera-wins, material mapping, accessor views, triangle assembly -- with the reference's float
// / double op order, producing the flat HostScene.  The JSON/base64/PNG readers are our own (the reference uses
// tinygltf v2.8.10; only its observable behaviour on well-formed files is reproduced: external and data: URIs,
// every image expanded to RGBA8).
//
// Known reference quirks kept on purpose: parents must precede children in `nodes`;
// camera axes are divided by the node scale; zfar>0 becomes the hit-distance limit of every ray; a non-zero file
// aspectRatio overrides W/H; alphaMode / doubleSided / KHR_materials_specular are ignored.
// Beyond the reference: bufferView.byteStride (interleaved vertex buffers) and the binary
// .glb container are accepted; tightly packed .gltf files load exactly as before.
// One deliberate deviation: a primitive WITHOUT a NORMAL attribute reads uninitialised `indices_count` in the
// reference (scene_parser.cpp:204-234, UB); here it gets the accessor's count and normal (0,0,1).
#include "host_scene.h"
#include "json.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <fstream>
#include <iostream>
#include <sstream>

namespace rtb {
namespace {

struct Mat4d {
    double d[16] = {0};
    bool empty = true;
};

Mat4d eye() {
    Mat4d m;
    m.empty = false;
    m.d[0] = m.d[5] = m.d[10] = m.d[15] = 1.0;   // eye() is written with float literals; exact in double
    return m;
}

// matrix4<T>::TRS (src/utils/matrix.h:28-35): evaluated in FLOAT, stored as double
Mat4d trs(const float t[3], const float r[4], const float s[3]) {
    const float rx = r[0], ry = r[1], rz = r[2], rw = r[3];
    const float sx = s[0], sy = s[1], sz = s[2];
    Mat4d m;
    m.empty = false;
    const float v[16] = {
        (1.0f - 2.0f * (ry * ry + rz * rz)) * sx, (rx * ry + rz * rw) * sx * 2.0f, (rx * rz - ry * rw) * sx * 2.0f, 0.f,
        (rx * ry - rz * rw) * sy * 2.0f, (1.0f - 2.0f * (rx * rx + rz * rz)) * sy, (ry * rz + rx * rw) * sy * 2.0f, 0.f,
        (rx * rz + ry * rw) * sz * 2.0f, (ry * rz - rx * rw) * sz * 2.0f, (1.0f - 2.0f * (rx * rx + ry * ry)) * sz, 0.f,
        t[0], t[1], t[2], 1.f};
    for (int i = 0; i < 16; ++i) m.d[i] = (double) v[i];
    return m;
}

// multiply(matrix4, matrix4) (src/utils/matrix.h:44-63): dest[i][j] = sum_k a[i][k] * b[k][j], a[i] = &data[4*i]
Mat4d mul(const Mat4d &a, const Mat4d &b) {
    Mat4d r;
    r.empty = false;
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j)
            r.d[4 * i + j] = a.d[4 * i + 0] * b.d[0 + j] + a.d[4 * i + 1] * b.d[4 + j] + a.d[4 * i + 2] * b.d[8 + j] + a.d[4 * i + 3] * b.d[12 + j];
    return r;
}

// multiply(matrix4<double>, vector4f) (src/utils/matrix.h:66-73): FLOAT accumulators, double products
void mul_vec4(const Mat4d &m, const float t[4], float out[4]) {
    for (int i = 0; i < 4; ++i) {
        float acc = 0.f;
        for (int j = 0; j < 4; ++j) acc = (float) ((double) acc + m.d[4 * j + i] * (double) t[j]);
        out[i] = acc;
    }
}

Mat4d transpose(const Mat4d &m) {
    Mat4d r;
    r.empty = false;
    for (int i = 0; i < 4; ++i)
        for (int j = 0; j < 4; ++j) r.d[4 * i + j] = m.d[4 * j + i];
    return r;
}

// inverse(matrix4) (src/utils/matrix.h:88-216), cofactor expansion, same term order
Mat4d inverse(const Mat4d &mat) {
    const double *m = mat.d;
    Mat4d res;
    res.empty = false;
    double *inv = res.d;
    inv[0] = m[5] * m[10] * m[15] - m[5] * m[11] * m[14] - m[9] * m[6] * m[15] + m[9] * m[7] * m[14] + m[13] * m[6] * m[11] - m[13] * m[7] * m[10];
    inv[4] = -m[4] * m[10] * m[15] + m[4] * m[11] * m[14] + m[8] * m[6] * m[15] - m[8] * m[7] * m[14] - m[12] * m[6] * m[11] + m[12] * m[7] * m[10];
    inv[8] = m[4] * m[9] * m[15] - m[4] * m[11] * m[13] - m[8] * m[5] * m[15] + m[8] * m[7] * m[13] + m[12] * m[5] * m[11] - m[12] * m[7] * m[9];
    inv[12] = -m[4] * m[9] * m[14] + m[4] * m[10] * m[13] + m[8] * m[5] * m[14] - m[8] * m[6] * m[13] - m[12] * m[5] * m[10] + m[12] * m[6] * m[9];
    inv[1] = -m[1] * m[10] * m[15] + m[1] * m[11] * m[14] + m[9] * m[2] * m[15] - m[9] * m[3] * m[14] - m[13] * m[2] * m[11] + m[13] * m[3] * m[10];
    inv[5] = m[0] * m[10] * m[15] - m[0] * m[11] * m[14] - m[8] * m[2] * m[15] + m[8] * m[3] * m[14] + m[12] * m[2] * m[11] - m[12] * m[3] * m[10];
    inv[9] = -m[0] * m[9] * m[15] + m[0] * m[11] * m[13] + m[8] * m[1] * m[15] - m[8] * m[3] * m[13] - m[12] * m[1] * m[11] + m[12] * m[3] * m[9];
    inv[13] = m[0] * m[9] * m[14] - m[0] * m[10] * m[13] - m[8] * m[1] * m[14] + m[8] * m[2] * m[13] + m[12] * m[1] * m[10] - m[12] * m[2] * m[9];
    inv[2] = m[1] * m[6] * m[15] - m[1] * m[7] * m[14] - m[5] * m[2] * m[15] + m[5] * m[3] * m[14] + m[13] * m[2] * m[7] - m[13] * m[3] * m[6];
    inv[6] = -m[0] * m[6] * m[15] + m[0] * m[7] * m[14] + m[4] * m[2] * m[15] - m[4] * m[3] * m[14] - m[12] * m[2] * m[7] + m[12] * m[3] * m[6];
    inv[10] = m[0] * m[5] * m[15] - m[0] * m[7] * m[13] - m[4] * m[1] * m[15] + m[4] * m[3] * m[13] + m[12] * m[1] * m[7] - m[12] * m[3] * m[5];
    inv[14] = -m[0] * m[5] * m[14] + m[0] * m[6] * m[13] + m[4] * m[1] * m[14] - m[4] * m[2] * m[13] - m[12] * m[1] * m[6] + m[12] * m[2] * m[5];
    inv[3] = -m[1] * m[6] * m[11] + m[1] * m[7] * m[10] + m[5] * m[2] * m[11] - m[5] * m[3] * m[10] - m[9] * m[2] * m[7] + m[9] * m[3] * m[6];
    inv[7] = m[0] * m[6] * m[11] - m[0] * m[7] * m[10] - m[4] * m[2] * m[11] + m[4] * m[3] * m[10] + m[8] * m[2] * m[7] - m[8] * m[3] * m[6];
    inv[11] = -m[0] * m[5] * m[11] + m[0] * m[7] * m[9] + m[4] * m[1] * m[11] - m[4] * m[3] * m[9] - m[8] * m[1] * m[7] + m[8] * m[3] * m[5];
    inv[15] = m[0] * m[5] * m[10] - m[0] * m[6] * m[9] - m[4] * m[1] * m[10] + m[4] * m[2] * m[9] + m[8] * m[1] * m[6] - m[8] * m[2] * m[5];
    double det = m[0] * inv[0] + m[1] * inv[4] + m[2] * inv[8] + m[3] * inv[12];
    if (det == 0) throw std::runtime_error("Zero determinant");
    det = 1.0 / det;
    for (int i = 0; i < 16; i++) inv[i] *= det;
    return res;
}

// normal(vector3f) (src/utils/vector.h:169-174)
void normalize3(float v[3]) {
    float mod = 0.f;
    for (int i = 0; i < 3; ++i) mod += v[i] * v[i];
    mod = std::sqrt(mod);
    mod = 1.f / mod;
    for (int i = 0; i < 3; ++i) v[i] = v[i] * mod;
}

std::vector<uint8_t> read_file(const std::string &path) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw std::runtime_error("[gltf] File read error : " + path);
    f.seekg(0, std::ios::end);
    std::streamoff sz = f.tellg();
    f.seekg(0, std::ios::beg);
    std::vector<uint8_t> data((size_t) sz);
    if (sz > 0) f.read((char *) data.data(), sz);
    return data;
}

std::vector<uint8_t> base64_decode(const std::string &s, size_t start) {
    static int8_t T[256];
    static bool init = false;
    if (!init) {
        std::memset(T, -1, sizeof(T));
        const char *abc = "ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789+/";
        for (int i = 0; i < 64; ++i) T[(unsigned char) abc[i]] = (int8_t) i;
        init = true;
    }
    std::vector<uint8_t> out;
    out.reserve((s.size() - start) * 3 / 4);
    uint32_t acc = 0;
    int bits = 0;
    for (size_t i = start; i < s.size(); ++i) {
        const int8_t v = T[(unsigned char) s[i]];
        if (v < 0) continue;   // '=' padding / whitespace
        acc = (acc << 6) | (uint32_t) v;
        bits += 6;
        if (bits >= 8) {
            bits -= 8;
            out.push_back((uint8_t) ((acc >> bits) & 0xFF));
        }
    }
    return out;
}

std::vector<uint8_t> load_uri(const std::string &uri, const std::string &base_dir) {
    if (uri.compare(0, 5, "data:") == 0) {
        const size_t comma = uri.find(',');
        if (comma == std::string::npos || uri.find(";base64") == std::string::npos)
            throw std::runtime_error("[gltf] unsupported data: URI");
        return base64_decode(uri, comma + 1);
    }
    return read_file(base_dir + uri);
}

struct View {
    const uint8_t *ptr = nullptr;
    size_t count = 0;
    int component_type = 0;
    std::string type;
    size_t stride = 0;      // bytes between elements: bufferView.byteStride, or the element size when tightly packed
    const uint8_t *at(size_t i) const { return ptr + i * stride; }
};

}  // namespace

void load_gltf(const std::string &filename, int width, int height, int samples, HostScene &scene) {
    std::string text;
    std::vector<uint8_t> glb_bin;     // the BIN chunk of a binary glTF container (buffer 0 when it has no uri)
    bool is_glb = false;
    {
        std::vector<uint8_t> raw = read_file(filename);
        auto u32 = [&](size_t off) { uint32_t x; std::memcpy(&x, raw.data() + off, 4); return x; };
        if (raw.size() >= 20 && std::memcmp(raw.data(), "glTF", 4) == 0) {
            // .glb (beyond the reference, which only calls LoadASCIIFromFile: scene_parser.cpp:33): 12-byte header,
            // then chunks {length, type, data}: JSON first, an optional BIN after it
            is_glb = true;
            const size_t total = std::min<size_t>(u32(8), raw.size());
            size_t off = 12;
            bool have_json = false;
            while (off + 8 <= total) {
                const size_t len = u32(off);
                const uint32_t type = u32(off + 4);
                if (off + 8 + len > total) throw std::runtime_error("[tinygltf] Failed to parse \"" + filename + "\"");
                if (type == 0x4E4F534Au && !have_json) {          // "JSON"
                    text.assign(raw.begin() + (ptrdiff_t) (off + 8), raw.begin() + (ptrdiff_t) (off + 8 + len));
                    have_json = true;
                } else if (type == 0x004E4942u && glb_bin.empty()) {   // "BIN\0"
                    glb_bin.assign(raw.begin() + (ptrdiff_t) (off + 8), raw.begin() + (ptrdiff_t) (off + 8 + len));
                }
                off += 8 + ((len + 3) & ~(size_t) 3);
            }
            if (!have_json) throw std::runtime_error("[tinygltf] Failed to parse \"" + filename + "\"");
        } else {
            text.assign(raw.begin(), raw.end());
        }
    }
    Json model;
    try {
        model = JsonParser(text).parse();
    } catch (const std::exception &e) {
        std::cout << "[tinygltf] Error: " << e.what() << std::endl;
        throw std::runtime_error("[tinygltf] Failed to parse \"" + filename + "\"");
    }
    std::string base_dir;
    {
        const size_t slash = filename.find_last_of("/\\");
        if (slash != std::string::npos) base_dir = filename.substr(0, slash + 1);
    }
    static const Json kEmptyArray = [] { Json j; j.type = Json::Array; return j; }();
    auto arr = [&](const char *key) -> const Json & {
        const Json *j = model.find(key);
        return (j && j->is_array()) ? *j : kEmptyArray;
    };
    const Json &jnodes = arr("nodes");
    const size_t n_nodes = jnodes.size();

    // ---- node transforms, in array order, parents assumed before children (scene_parser.cpp:49-97)
    std::vector<Mat4d> node_matrix(n_nodes);
    int node_with_camera = -1;
    for (size_t i = 0; i < n_nodes; ++i) {
        const Json &jn = jnodes.at(i);
        if (node_matrix[i].empty) {
            std::vector<double> m = jn.numbers("matrix");
            if (m.size() == 16) {
                node_matrix[i].empty = false;
                for (int k = 0; k < 16; ++k) node_matrix[i].d[k] = m[k];
            }
        }
    }
    for (size_t i = 0; i < n_nodes; ++i) {
        const Json &jn = jnodes.at(i);
        if (jn.int_or("camera", -1) != -1 && node_with_camera == -1) node_with_camera = (int) i;
        const std::vector<double> T = jn.numbers("translation"), R = jn.numbers("rotation"), S = jn.numbers("scale");
        Mat4d transform;
        if (T.empty() && R.empty() && S.empty()) {
            transform = node_matrix[i].empty ? eye() : node_matrix[i];
        } else {
            float t[3] = {0.f, 0.f, 0.f}, r[4] = {0.f, 0.f, 0.f, 1.f}, s[3] = {1.f, 1.f, 1.f};
            if (!T.empty()) for (int j = 0; j < 3; ++j) t[j] = (float) T.at(j);
            if (!R.empty()) for (int j = 0; j < 4; ++j) r[j] = (float) R.at(j);
            if (!S.empty()) for (int j = 0; j < 3; ++j) s[j] = (float) S.at(j);
            transform = trs(t, r, s);
            if (!node_matrix[i].empty) transform = mul(node_matrix[i], transform);
        }
        node_matrix[i] = transform;
        const Json *children = jn.find("children");
        if (children && children->is_array())
            for (const Json &c : children->arr) {
                const size_t child = (size_t) c.num;
                if (child >= n_nodes) throw std::runtime_error("[gltf check] child index out of range");
                node_matrix[child] = node_matrix[child].empty ? transform : mul(transform, node_matrix[child]);
            }
    }
    if (node_with_camera == -1) throw std::runtime_error("[gltf check] No camera found");

    // ---- camera (scene_parser.cpp:102-127)
    {
        const Json &jn = jnodes.at((size_t) node_with_camera);
        const Mat4d &M = node_matrix[(size_t) node_with_camera];
        const float origin[4] = {0.f, 0.f, 0.f, 1.f};
        float p[4];
        mul_vec4(M, origin, p);
        for (int k = 0; k < 3; ++k) scene.cam_pos[k] = p[k];
        const std::vector<double> S = jn.numbers("scale");
        for (int i = 0; i < 3; ++i) {
            float dir[4] = {0.f, 0.f, 0.f, 0.f};
            dir[i] = 1.f;
            dir[2] = -dir[2];
            float res[4];
            mul_vec4(M, dir, res);
            if (!S.empty())
                for (int j = 0; j < 3; ++j) res[j] /= (float) S.at(j);
            normalize3(res);
            for (int k = 0; k < 3; ++k) scene.cam_axes[i * 3 + k] = res[k];
        }
        const Json &jcam = arr("cameras").at((size_t) jn.int_or("camera", 0));
        static const Json kEmptyObj = [] { Json j; j.type = Json::Object; return j; }();
        const Json *persp = jcam.find("perspective");
        const Json &P = persp ? *persp : kEmptyObj;
        const double zfar = P.number_or("zfar", 0.0);
        if (zfar > 0) scene.max_distance = (float) zfar;
        const float fov_y = (float) P.number_or("yfov", 0.0);
        const double ar = P.number_or("aspectRatio", 0.0);
        const float aspect = (ar != 0) ? (float) ar : (float) width / (float) height;
        const float fov_x = 2.f * std::atan(std::tan(fov_y * 0.5f) * aspect);
        scene.cam_fov[0] = fov_x;
        scene.cam_fov[1] = fov_y;
        scene.cam_tan_half_fov[0] = std::tan(fov_x / 2);   // Camera::cast_in_pixel (camera.cpp:51-52)
        scene.cam_tan_half_fov[1] = std::tan(fov_y / 2);
    }
    scene.width = width;
    scene.height = height;
    scene.samples = samples;

    // ---- buffers
    const Json &jbuffers = arr("buffers");
    std::vector<std::vector<uint8_t>> buffers(jbuffers.size());
    for (size_t i = 0; i < jbuffers.size(); ++i) {
        const std::string uri = jbuffers.at(i).string_or("uri", "");
        if (uri.empty() && is_glb && i == 0) {
            buffers[i] = std::move(glb_bin);
            continue;
        }
        if (uri.empty()) throw std::runtime_error("[tinygltf] Failed to parse \"" + filename + "\"");
        try {
            buffers[i] = load_uri(uri, base_dir);
        } catch (const std::exception &e) {
            std::cout << "[tinygltf] Error: " << e.what() << std::endl;
            throw std::runtime_error("[tinygltf] Failed to parse \"" + filename + "\"");
        }
    }

    // ---- textures = images, always RGBA8 (scene_parser.cpp:131-142)
    const Json &jimages = arr("images");
    std::vector<std::vector<uint8_t>> image_bytes(jimages.size());
    std::vector<std::string> image_uri(jimages.size());
    for (size_t i = 0; i < jimages.size(); ++i) {
        const Json &ji = jimages.at(i);
        std::vector<uint8_t> &bytes = image_bytes[i];
        const std::string uri = ji.string_or("uri", "");
        image_uri[i] = uri;
        if (!uri.empty()) {
            bytes = load_uri(uri, base_dir);
        } else {
            const int bv = ji.int_or("bufferView", -1);
            if (bv < 0) throw std::runtime_error("[tinygltf] image without uri / bufferView");
            const Json &jv = arr("bufferViews").at((size_t) bv);
            const std::vector<uint8_t> &buf = buffers.at((size_t) jv.int_or("buffer", 0));
            const size_t off = (size_t) jv.number_or("byteOffset", 0), len = (size_t) jv.number_or("byteLength", 0);
            if (off + len > buf.size()) throw std::runtime_error("[tinygltf] image bufferView out of range");
            bytes.assign(buf.begin() + (ptrdiff_t) off, buf.begin() + (ptrdiff_t) (off + len));
        }
    }
    // decoding is the largest host cost of a textured scene (Sponza: 69 PNGs, 17 s on one thread): one image per thread
    scene.textures.resize(jimages.size());
    std::vector<std::string> image_err(jimages.size());
#pragma omp parallel for schedule(dynamic, 1)
    for (long long i = 0; i < (long long) jimages.size(); ++i) {
        std::string err;
        if (!decode_image(image_bytes[(size_t) i], scene.textures[(size_t) i], err)) image_err[(size_t) i] = err.empty() ? "decode failed" : err;
        std::vector<uint8_t>().swap(image_bytes[(size_t) i]);
    }
    for (size_t i = 0; i < jimages.size(); ++i)   // the first failing image is reported, as a sequential loader would
        if (!image_err[i].empty())
            throw std::runtime_error("[tinygltf] Failed to load image " + std::to_string(i) + " (" + image_uri[i] + "): " + image_err[i]);

    // ---- meshes: one Mesh per glTF primitive, in node order (scene_parser.cpp:144-346)
    const Json &jmeshes = arr("meshes"), &jmaterials = arr("materials"), &jtextures = arr("textures");
    const Json &jaccessors = arr("accessors"), &jviews = arr("bufferViews");
    auto view_of = [&](int accessor_index) -> View {
        if (accessor_index < 0) throw std::runtime_error("[gltf check] missing accessor");
        const Json &a = jaccessors.at((size_t) accessor_index);
        const Json &bv = jviews.at((size_t) a.int_or("bufferView", 0));
        const std::vector<uint8_t> &buf = buffers.at((size_t) bv.int_or("buffer", 0));
        const size_t start = (size_t) bv.number_or("byteOffset", 0) + (size_t) a.number_or("byteOffset", 0);
        if (start > buf.size()) throw std::runtime_error("[gltf check] accessor out of range");
        View v;
        v.ptr = buf.data() + start;
        v.count = (size_t) a.number_or("count", 0);
        v.component_type = a.int_or("componentType", 0);
        v.type = a.string_or("type", "");
        // (beyond the reference, which assumes tightly packed views: scene_parser.cpp:209-262 index the raw pointers)
        const size_t comps = v.type == "SCALAR" ? 1 : v.type == "VEC2" ? 2 : v.type == "VEC3" ? 3 : v.type == "VEC4" ? 4 : 0;
        const size_t csize = (v.component_type == 5120 || v.component_type == 5121) ? 1 : (v.component_type == 5122 || v.component_type == 5123) ? 2 : 4;
        const size_t packed = comps * csize;
        const size_t bs = (size_t) bv.number_or("byteStride", 0);
        v.stride = bs ? bs : packed;
        // count elements of `packed` bytes, `stride` apart, must lie inside the buffer (no wrap-around for huge counts)
        if (v.count && packed) {
            const size_t room = buf.size() - start;
            if (packed > room || v.stride == 0 || v.count - 1 > (room - packed) / v.stride) throw std::runtime_error("[gltf check] accessor out of range");
        }
        return v;
    };
    auto image_of = [&](const Json *tex_info) -> int {
        if (!tex_info) return -1;
        const int ti = tex_info->int_or("index", -1);
        return (ti == -1) ? -1 : jtextures.at((size_t) ti).int_or("source", -1);
    };
    const int kFloat = 5126, kU16 = 5123, kU32 = 5125;

    size_t n_tris_total = 0;
    for (size_t ni = 0; ni < n_nodes; ++ni) {
        const Json &jn = jnodes.at(ni);
        const int mesh_index = jn.int_or("mesh", -1);
        if (mesh_index == -1) continue;
        const Json &jprims = jmeshes.at((size_t) mesh_index).at("primitives");
        for (size_t pi = 0; pi < jprims.size(); ++pi) {
            const Json &p = jprims.at(pi);
            const Mat4d &transform = node_matrix[ni];
            const Mat4d normal_transform = inverse(transpose(transform));
            // material (scene_parser.cpp:161-196); defaults of struct Material (primitive.h:76-89)
            float mat[16] = {1.f, 1.f, 1.f, 1.f, 1.f, 0.f, 0.f, 0.f, 1.f, 1.f, -1.f, -1.f, -1.f, -1.f, 0.f, 0.f};
            const int material_index = p.int_or("material", -1);
            if (material_index != -1) {
                const Json &gm = jmaterials.at((size_t) material_index);
                static const Json kObj = [] { Json j; j.type = Json::Object; return j; }();
                const Json *pbrp = gm.find("pbrMetallicRoughness");
                const Json &pbr = pbrp ? *pbrp : kObj;
                std::vector<double> bcf = pbr.numbers("baseColorFactor");
                if (bcf.size() != 4) bcf = {1.0, 1.0, 1.0, 1.0};
                for (int j = 0; j < 3; ++j) mat[2 + j] = (float) bcf[j];
                const Json *ext = gm.find("extensions");
                if (bcf[3] < 1.) {
                    mat[1] = (float) bcf[3];
                    const Json *ior = ext ? ext->find("KHR_materials_ior") : nullptr;
                    mat[0] = ior ? (float) ior->number_or("ior", 0.0) : 1.5f;
                }
                mat[8] = (float) pbr.number_or("metallicFactor", 1.0);
                float r2 = (float) pbr.number_or("roughnessFactor", 1.0);
                r2 *= r2;
                mat[9] = r2;
                std::vector<double> ef = gm.numbers("emissiveFactor");
                if (ef.size() != 3) ef = {0.0, 0.0, 0.0};
                for (int j = 0; j < 3; ++j) mat[5 + j] = (float) ef[j];
                const Json *es = ext ? ext->find("KHR_materials_emissive_strength") : nullptr;
                if (es) {
                    const float strength = (float) es->number_or("emissiveStrength", 0.0);
                    for (int j = 0; j < 3; ++j) mat[5 + j] *= strength;
                }
                mat[10] = (float) image_of(pbr.find("baseColorTexture"));
                mat[11] = (float) image_of(gm.find("normalTexture"));
                mat[12] = (float) image_of(pbr.find("metallicRoughnessTexture"));
                mat[13] = (float) image_of(gm.find("emissiveTexture"));
            }
            const int mesh_id = (int) scene.n_meshes();
            scene.mesh_material.insert(scene.mesh_material.end(), mat, mat + 16);
            scene.mesh_normal_transform.insert(scene.mesh_normal_transform.end(), normal_transform.d, normal_transform.d + 16);

            const Json &attrs = p.at("attributes");
            const View idx = view_of(p.int_or("indices", -1));
            if (idx.type != "SCALAR") throw std::runtime_error("Index type not supported");
            const View pos = view_of(attrs.int_or("POSITION", -1));
            if (pos.type != "VEC3") throw std::runtime_error("Position type not supported");
            if (pos.component_type != kFloat) throw std::runtime_error("Position component type not supported");
            View nrm, uv, tan;
            if (attrs.has("NORMAL")) {
                nrm = view_of(attrs.int_or("NORMAL", -1));
                if (nrm.type != "VEC3") throw std::runtime_error("Normal type not supported");
                if (nrm.component_type != kFloat) throw std::runtime_error("Normal component type not supported");
            }
            if (attrs.has("TEXCOORD_0")) {
                uv = view_of(attrs.int_or("TEXCOORD_0", -1));
                if (uv.type != "VEC2") throw std::runtime_error("Texcoord type not supported");
                if (uv.component_type != kFloat) throw std::runtime_error("Texcoord component type not supported");
            }
            if (attrs.has("TANGENT")) {
                tan = view_of(attrs.int_or("TANGENT", -1));
                if (tan.type != "VEC4") throw std::runtime_error("Tangent type not supported");
                if (tan.component_type != kFloat) throw std::runtime_error("Tangent component type not supported");
            }
            if (idx.component_type != kU16 && idx.component_type != kU32)
                throw std::runtime_error("Index component type not supported");

            const size_t ntri = idx.count / 3;
            const size_t base = n_tris_total;
            n_tris_total += ntri;
            const bool nv = nrm.ptr != nullptr, tv = uv.ptr != nullptr, gv = tan.ptr != nullptr;
            RawMesh rm;
            rm.n_tris = ntri;
            rm.first_tri = base;
            rm.idx = idx.ptr;
            rm.index_bytes = idx.component_type == kU16 ? 2 : 4;
            rm.idx_stride = idx.stride;
            rm.pos = pos.ptr; rm.pos_stride = pos.stride;
            rm.nrm = nrm.ptr; rm.nrm_stride = nrm.stride;
            rm.tan = tan.ptr; rm.tan_stride = tan.stride;
            rm.uv = uv.ptr; rm.uv_stride = uv.stride;
            rm.n_vertices = pos.count;
            if (nv) rm.n_vertices = std::min(rm.n_vertices, nrm.count);
            if (tv) rm.n_vertices = std::min(rm.n_vertices, uv.count);
            if (gv) rm.n_vertices = std::min(rm.n_vertices, tan.count);
            rm.normal_mode = 0;   // RTB_NORMALS_TRANSFORM
            rm.mesh_id = mesh_id;
            std::memcpy(rm.transform, transform.d, sizeof(rm.transform));
            std::memcpy(rm.normal_transform, normal_transform.d, sizeof(rm.normal_transform));
            scene.tris.resize_core(n_tris_total);
            // positions only: what the BVH build needs.  Normals, tangents and texcoords are assembled from the same views
            // where they are consumed: on the GPU (rtb_scene_create_raw), or by HostScene::assemble_attributes.
            bool tangent_warned = false;
            for (size_t i = 0; i < ntri; ++i) {
                for (int v = 0; v < 3; ++v)   // (the reference indexes the raw buffers unchecked: scene_parser.cpp:306-343)
                    if (rm.vertex(i, v) >= rm.n_vertices) throw std::runtime_error("[gltf check] vertex index out of range");
                const size_t t = base + i;
                assemble_triangle(rm, i, &scene.tris.pos[t * 9], nullptr, nullptr, nullptr);
                if (gv && !tangent_warned) {
                    float w[3];
                    for (int v = 0; v < 3; ++v) std::memcpy(&w[v], tan.at(rm.vertex(i, v)) + 12, 4);
                    if (w[1] != w[0] || w[2] != w[0]) {
                        std::cout << "Warning: tangent.w values differ within triangle" << std::endl;
                        tangent_warned = true;   // the reference prints this per triangle; once per primitive here
                    }
                }
                scene.tris.mesh[t] = mesh_id;
                scene.tris.source[t] = (uint32_t) t;
            }
            scene.raw_meshes.push_back(rm);
        }
    }
    scene.raw_buffers = std::move(buffers);   // (moving a vector keeps its storage: the views stay valid)
}

size_t RawMesh::vertex(size_t tri, int v) const {
    const size_t k = tri * 3 + (size_t) v;
    if (!idx) return k;
    if (index_bytes == 2) {
        uint16_t x;
        std::memcpy(&x, idx + k * idx_stride, 2);
        return x;
    }
    uint32_t x;
    std::memcpy(&x, idx + k * idx_stride, 4);
    return x;
}

void assemble_triangle(const RawMesh &m, size_t tri, float *pos9, float *nrm9, float *tan12, float *uv6) {
    size_t index[3];
    for (int v = 0; v < 3; ++v) index[v] = m.vertex(tri, v);
    Mat4d T, N;
    T.empty = N.empty = false;
    std::memcpy(T.d, m.transform, sizeof(T.d));
    std::memcpy(N.d, m.normal_transform, sizeof(N.d));
    float U[3] = {0.f, 0.f, 0.f}, V[3] = {0.f, 0.f, 0.f};
    if (pos9 || (nrm9 && m.normal_mode == 2)) {
        float P[3][4];
        for (int v = 0; v < 3; ++v) {
            float in[4];
            std::memcpy(in, m.pos + index[v] * m.pos_stride, 12);
            in[3] = 1.f;
            if (m.transform_positions) mul_vec4(T, in, P[v]);
            else std::memcpy(P[v], in, 16);
        }
        for (int k = 0; k < 3; ++k) {
            U[k] = P[1][k] + (-P[0][k]);   // operator-(a,b) = a + (-b) (vector.h:326-328)
            V[k] = P[2][k] + (-P[0][k]);
        }
        if (pos9)
            for (int k = 0; k < 3; ++k) {
                pos9[k] = P[0][k];
                pos9[3 + k] = U[k];
                pos9[6 + k] = V[k];
            }
    }
    if (nrm9) {
        if (m.normal_mode == 2) {   // RTB_NORMALS_FLAT: normal(cross(U, V)) (vector.h:169-174, 372-378)
            float c[3] = {U[1] * V[2] - V[1] * U[2], U[2] * V[0] - V[2] * U[0], U[0] * V[1] - V[0] * U[1]};
            normalize3(c);
            for (int v = 0; v < 3; ++v)
                for (int k = 0; k < 3; ++k) nrm9[v * 3 + k] = c[k];
        } else if (m.normal_mode == 1) {   // RTB_NORMALS_VERBATIM
            for (int v = 0; v < 3; ++v) std::memcpy(nrm9 + v * 3, m.nrm + index[v] * m.nrm_stride, 12);
        } else {
            for (int v = 0; v < 3; ++v) {
                float in[4] = {0.f, 0.f, 1.f, 0.f};
                if (m.nrm) std::memcpy(in, m.nrm + index[v] * m.nrm_stride, 12);
                in[3] = 0.f;
                float res[4];
                mul_vec4(N, in, res);
                normalize3(res);
                for (int k = 0; k < 3; ++k) nrm9[v * 3 + k] = res[k];
            }
        }
    }
    if (uv6) {
        if (m.uv) for (int v = 0; v < 3; ++v) std::memcpy(uv6 + v * 2, m.uv + index[v] * m.uv_stride, 8);
        else std::memset(uv6, 0, 24);
    }
    if (tan12) {
        if (m.tan) for (int v = 0; v < 3; ++v) std::memcpy(tan12 + v * 4, m.tan + index[v] * m.tan_stride, 16);
        else std::memset(tan12, 0, 48);
    }
}

void HostScene::assemble_attributes() {
    const size_t n = tris.size();
    if (tris.nrm.size() == n * 9 && tris.uv.size() == n * 6 && tris.tan.size() == n * 12) return;
    tris.nrm.assign(n * 9, 0.f);
    tris.tan.assign(n * 12, 0.f);
    tris.uv.assign(n * 6, 0.f);
    std::vector<size_t> first(raw_meshes.size() + 1, 0);
    for (size_t m = 0; m < raw_meshes.size(); ++m) first[m + 1] = first[m] + raw_meshes[m].n_tris;
#pragma omp parallel for schedule(static)
    for (long long i = 0; i < (long long) n; ++i) {
        const size_t src = tris.source[(size_t) i];
        const size_t m = (size_t) (std::upper_bound(first.begin(), first.end(), src) - first.begin()) - 1;
        assemble_triangle(raw_meshes[m], src - first[m], nullptr, &tris.nrm[(size_t) i * 9], &tris.tan[(size_t) i * 12], &tris.uv[(size_t) i * 6]);
    }
}

}  // namespace rtb
