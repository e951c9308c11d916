// loaders.cpp — settings / scene / OBJ / image readers of the host application.
// File formats and behaviour follow the reference's utils (path_tracer/src/utils/
// parser.cpp, mesh_loader.cpp, image_buffer.cpp) and camera.cpp.
#include "loaders.hpp"

#include <zlib.h>

#include <cmath>
#include <cstdarg>
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <filesystem>
#include <fstream>
#include <algorithm>
#include <atomic>
#include <map>
#include <mutex>
#include <thread>
#include <sstream>
#include <stdexcept>

#include "json.hpp"

namespace host {

void say(const char* fmt, ...) {
    static const bool quiet = getenv("B200PT_QUIET") != nullptr;
    if (quiet) return;
    va_list ap;
    va_start(ap, fmt);
    vprintf(fmt, ap);
    va_end(ap);
}

material_data default_material() {
    material_data m{};
    m.base_color[0] = m.base_color[1] = m.base_color[2] = 0.8f;
    m.subsurface = 0.0f;
    m.metallic = 0.0f;
    m.specular = 0.5f;
    m.specular_tint = 1.0f;
    m.roughness = 0.5f;
    m.anisotropic = 0.0f;
    m.sheen = 0.0f;
    m.sheen_tint = 1.0f;
    m.clearcoat = 0.0f;
    m.clearcoat_gloss = 0.03f;
    m.ior = 1.45f;
    m.specular_transmission = 0.0f;
    m.specular_transmission_roughness = 0.0f;
    m.emission = 0.0f;
    return m;
}

// ------------------------------------------------------------------ camera.cpp:3-22
static vec3 sub(vec3 a, vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
static vec3 mul(float s, vec3 a) { return {s * a.x, s * a.y, s * a.z}; }
static vec3 divs(vec3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
static vec3 cross(vec3 a, vec3 b) { return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y}; }
static vec3 normalize(vec3 v) {
    float r = 1.f / sqrtf(v.x * v.x + v.y * v.y + v.z * v.z);
    return {v.x * r, v.y * r, v.z * r};
}

camera_data to_camera_data(const camera& c, int width, int height) {
    const float pi = 3.14159265358979323f;
    const float aspect = static_cast<float>(width) / static_cast<float>(height);
    const float theta = c.vertical_fov * pi / 180.0f;
    const float h = tanf(theta / 2);
    const float viewport_height = 2.0f * h;
    const float viewport_width = aspect * viewport_height;
    const vec3 origin = c.look_from;
    const vec3 w = normalize(sub(c.look_from, c.look_at));
    const vec3 u = normalize(cross(c.look_up, w));
    const vec3 v = normalize(cross(w, u));
    const vec3 horizontal = mul(viewport_width, u);
    const vec3 vertical = mul(viewport_height, v);
    const vec3 llc = sub(sub(sub(origin, divs(horizontal, 2.0f)), divs(vertical, 2.0f)), w);
    return camera_data{origin, llc, horizontal, vertical};
}

// ------------------------------------------------------------------ parser.cpp
static vec3 vec3_of(const Json& a) { return {a[0].as_float(), a[1].as_float(), a[2].as_float()}; }

settings_data parse_settings(const std::string& path) {  // parser.cpp:81-117
    say("Parsing settings\n");
    const Json config = Json::load(path);
    settings_data data;
    const Json& test = config["test"];
    data.test.name = test["name"].as_string();
    data.test.material_name = test["material_name"].as_string();
    data.test.attribute_name = test["attribute_name"].as_string();
    data.test.material_type = test["material_type"].as_int();
    data.test.step_size = test["step_size"].as_float();
    const Json& values = test["values"];
    for (size_t i = 0; i < values.size(); ++i) {
        if (values[i].is_array())
            data.test.vec_values.push_back(vec3_of(values[i]));
        else
            data.test.flt_values.push_back(values[i].as_float());
    }
    data.scene = config["scene"].as_string();
    data.buffer_size[0] = config["buffer_size"][0].as_int();
    data.buffer_size[1] = config["buffer_size"][1].as_int();
    data.max_path_depth = config["max_path_depth"].as_int();
    data.max_samples = config["max_samples"].as_int();
    data.environment_use = config["environment_use"].as_bool();
    data.environment_auto = config["environment_auto"].as_bool();
    data.environment_color = vec3_of(config["environment_color"]);
    data.environment_intensity = config["environment_intensity"].as_float();
    return data;
}

camera parse_camera(const std::string& path) {  // parser.cpp:65-79 (to_camera_data is applied by the caller)
    say("Parsing camera\n");
    const Json config = Json::load(path);
    const Json& cam = config["camera"];
    camera c;
    c.look_from = vec3_of(cam["look_from"]);
    c.look_at = vec3_of(cam["look_at"]);
    c.look_up = vec3_of(cam["look_up"]);
    c.vertical_fov = cam["vertical_fov"].as_float();
    return c;
}

std::vector<material_entry> parse_materials(const std::string& path) {  // parser.cpp:19-63
    say("Parsing materials\n");
    const Json config = Json::load(path);
    std::vector<material_entry> out;
    const Json& mats = config["materials"];
    for (size_t i = 0; i < mats.size(); ++i) {
        const Json& m = mats[i];
        material_entry e;
        e.name = m["name"].as_string();
        say(" - \"%s\"\n", e.name.c_str());
        e.data = default_material();
        // superset of the reference: a missing "use_texture" counts as false
        // (assets/sphere.json has none and would make the reference throw)
        const bool use_texture = m.contains("use_texture") ? m["use_texture"].as_bool() : false;
        if (use_texture) {
            e.texture_file = e.name + "-textures/" + m["filename"].as_string();
        } else {
            e.data.base_color[0] = m["base_color"][0].as_float();
            e.data.base_color[1] = m["base_color"][1].as_float();
            e.data.base_color[2] = m["base_color"][2].as_float();
        }
        e.data.subsurface = m["subsurface"].as_float();
        e.data.metallic = m["metallic"].as_float();
        e.data.specular = m["specular"].as_float();
        e.data.specular_tint = m["specular_tint"].as_float();
        e.data.roughness = m["roughness"].as_float();
        e.data.anisotropic = m["anisotropic"].as_float();
        e.data.sheen = m["sheen"].as_float();
        e.data.sheen_tint = m["sheen_tint"].as_float();
        e.data.clearcoat = m["clearcoat"].as_float();
        e.data.clearcoat_gloss = m["clearcoat_gloss"].as_float();
        e.data.ior = m["ior"].as_float();
        e.data.specular_transmission = m["specular_transmission"].as_float();
        e.data.specular_transmission_roughness = m["specular_transmission_roughness"].as_float();
        e.data.emission = m["emission"].as_float();
        out.push_back(std::move(e));
    }
    return out;
}

// ------------------------------------------------------------------ OBJ reader
// The subset of tinyobjloader's ObjReader the reference relies on (triangulate=false):
// v / vn / vt / f / o / g; 1-based and negative indices; a shape is emitted for every
// o/g block that has faces.  Everything else (mtllib, usemtl, s, ...) is ignored.
namespace {
struct obj_index {
    int v = -1, vt = -1, vn = -1;
};
struct obj_shape {
    std::string name;
    std::vector<obj_index> indices;
    std::vector<int> face_sizes;
};
struct obj_attrib {
    std::vector<float> vertices, normals, texcoords;
};

inline int fix_index(int idx, int n) {
    if (idx > 0) return idx - 1;
    if (idx == 0) return -1;
    return n + idx;
}

// The file is cut into chunks at line ends and the chunks are parsed by concurrent threads (the
// reference parses serially through tinyobjloader).  A chunk cannot resolve its own indices — a negative index
// counts back from the attributes seen so far in the WHOLE file — so it keeps the raw integers together with its
// local attribute counts at each face; a serial merge adds the chunk bases, fixes the indices and cuts the shapes.
// One chunk is the serial parser: both paths are the same code, tests/test_host.py compares their output.
struct obj_chunk {
    struct raw_index {
        int v = 0, vt = 0, vn = 0;
        bool has_vt = false, has_vn = false;
    };
    struct shape_break {
        size_t face;       // the o/g line stands before this face of the chunk
        std::string name;
    };
    std::vector<float> vertices, normals, texcoords;
    std::vector<raw_index> indices;
    std::vector<int> face_sizes;
    std::vector<int> face_counts;   // per face: vertices, normals, texcoords of the chunk seen before it
    std::vector<shape_break> breaks;
};

void parse_obj_chunk(const char* p, const char* end, obj_chunk& c) {
    auto is_sp = [](char ch) { return ch == ' ' || ch == '\t'; };
    while (p < end) {
        const char* eol = (const char*)memchr(p, '\n', (size_t)(end - p));
        if (!eol) eol = end;
        const char* q = p;
        while (q < eol && is_sp(*q)) ++q;
        if (q + 1 < eol && q[0] == 'v' && is_sp(q[1])) {
            char* e;
            q += 2;
            for (int k = 0; k < 3; ++k) {
                c.vertices.push_back((float)strtod(q, &e));
                q = e;
            }
        } else if (q + 2 < eol && q[0] == 'v' && q[1] == 'n' && is_sp(q[2])) {
            char* e;
            q += 3;
            for (int k = 0; k < 3; ++k) {
                c.normals.push_back((float)strtod(q, &e));
                q = e;
            }
        } else if (q + 2 < eol && q[0] == 'v' && q[1] == 't' && is_sp(q[2])) {
            char* e;
            q += 3;
            for (int k = 0; k < 2; ++k) {
                c.texcoords.push_back((float)strtod(q, &e));
                q = e;
            }
        } else if (q + 1 < eol && q[0] == 'f' && is_sp(q[1])) {
            q += 2;
            int n = 0;
            c.face_counts.push_back((int)c.vertices.size() / 3);
            c.face_counts.push_back((int)c.normals.size() / 3);
            c.face_counts.push_back((int)c.texcoords.size() / 2);
            while (q < eol) {
                while (q < eol && (is_sp(*q) || *q == '\r')) ++q;
                if (q >= eol) break;
                obj_chunk::raw_index ix;
                char* e;
                ix.v = (int)strtol(q, &e, 10);
                q = e;
                if (q < eol && *q == '/') {
                    ++q;
                    if (q < eol && *q == '/') {
                        ++q;
                        ix.vn = (int)strtol(q, &e, 10);
                        ix.has_vn = true;
                        q = e;
                    } else {
                        ix.vt = (int)strtol(q, &e, 10);
                        ix.has_vt = true;
                        q = e;
                        if (q < eol && *q == '/') {
                            ++q;
                            ix.vn = (int)strtol(q, &e, 10);
                            ix.has_vn = true;
                            q = e;
                        }
                    }
                }
                c.indices.push_back(ix);
                ++n;
            }
            c.face_sizes.push_back(n);
        } else if (q + 1 < eol && (q[0] == 'o' || q[0] == 'g') && is_sp(q[1])) {
            q += 2;
            while (q < eol && is_sp(*q)) ++q;
            const char* e = eol;
            while (e > q && (e[-1] == '\r' || is_sp(e[-1]))) --e;
            c.breaks.push_back({c.face_sizes.size(), std::string(q, e)});
        }
        p = eol + 1;
    }
}

int obj_threads(size_t bytes) {
    if (const char* e = getenv("B200PT_OBJ_THREADS")) return std::max(1, atoi(e));
    const size_t by_size = bytes / (256u << 10);   // a chunk below 256 KB is not worth a thread
    return (int)std::max<size_t>(1, std::min<size_t>({by_size, (size_t)std::thread::hardware_concurrency(), (size_t)16}));
}

bool parse_obj(const std::string& path, obj_attrib& at, std::vector<obj_shape>& shapes, std::string& err) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) {
        err = "Cannot open file [" + path + "]";
        return false;
    }
    std::string data;
    {
        fseek(f, 0, SEEK_END);
        long sz = ftell(f);
        fseek(f, 0, SEEK_SET);
        data.resize((size_t)sz);
        if (sz > 0 && fread(&data[0], 1, (size_t)sz, f) != (size_t)sz) {
            fclose(f);
            err = "read error";
            return false;
        }
        fclose(f);
    }
    // chunks: cut at the first line end at or after i * size / T
    const int T = obj_threads(data.size());
    const char* base = data.c_str();
    const char* end = base + data.size();
    std::vector<const char*> cut((size_t)T + 1, end);
    cut[0] = base;
    for (int i = 1; i < T; ++i) {
        const char* p = base + data.size() * (size_t)i / (size_t)T;
        if (p < cut[(size_t)i - 1]) p = cut[(size_t)i - 1];
        const char* nl = (const char*)memchr(p, '\n', (size_t)(end - p));
        cut[(size_t)i] = nl ? nl + 1 : end;
    }
    std::vector<obj_chunk> chunks((size_t)T);
    {
        std::vector<std::thread> th;
        for (int i = 1; i < T; ++i) th.emplace_back(parse_obj_chunk, cut[(size_t)i], cut[(size_t)i + 1], std::ref(chunks[(size_t)i]));
        parse_obj_chunk(cut[0], cut[1], chunks[0]);
        for (auto& t : th) t.join();
    }
    // merge in file order
    size_t tv = 0, tn = 0, tt = 0;
    for (const obj_chunk& c : chunks) {
        tv += c.vertices.size();
        tn += c.normals.size();
        tt += c.texcoords.size();
    }
    at.vertices.reserve(tv);
    at.normals.reserve(tn);
    at.texcoords.reserve(tt);
    obj_shape cur;
    auto flush = [&]() {
        if (!cur.face_sizes.empty()) shapes.push_back(cur);
        cur.indices.clear();
        cur.face_sizes.clear();
    };
    for (const obj_chunk& c : chunks) {
        const int bv = (int)at.vertices.size() / 3, bn = (int)at.normals.size() / 3, bt = (int)at.texcoords.size() / 2;
        size_t next_break = 0, corner = 0;
        for (size_t face = 0; face <= c.face_sizes.size(); ++face) {
            while (next_break < c.breaks.size() && c.breaks[next_break].face == face) {
                flush();
                cur.name = c.breaks[next_break].name;
                ++next_break;
            }
            if (face == c.face_sizes.size()) break;
            const int nv = bv + c.face_counts[3 * face], nn = bn + c.face_counts[3 * face + 1], nt = bt + c.face_counts[3 * face + 2];
            for (int k = 0; k < c.face_sizes[face]; ++k) {
                const obj_chunk::raw_index& r = c.indices[corner++];
                obj_index ix;
                ix.v = fix_index(r.v, nv);
                if (r.has_vt) ix.vt = fix_index(r.vt, nt);
                if (r.has_vn) ix.vn = fix_index(r.vn, nn);
                cur.indices.push_back(ix);
            }
            cur.face_sizes.push_back(c.face_sizes[face]);
        }
        at.vertices.insert(at.vertices.end(), c.vertices.begin(), c.vertices.end());
        at.normals.insert(at.normals.end(), c.normals.begin(), c.normals.end());
        at.texcoords.insert(at.texcoords.end(), c.texcoords.begin(), c.texcoords.end());
    }
    flush();
    return true;
}

// create_mesh, mesh_loader.cpp:9-83: vertices are shared by POSITION index only; the
// normal / uv stored for a local vertex is the one of the corner that first grew the
// arrays past it.  Faces with more than 3 corners are fan-split
// (the reference assumes triangles and would mis-index them).
std::shared_ptr<mesh> create_mesh(const obj_shape& shape, const obj_attrib& at) {
    auto m = std::make_shared<mesh>();
    // position index -> local vertex id.  The reference keeps a std::map (mesh_loader.cpp:12); a flat table over the
    // index range the shape uses gives the same mapping without the tree (0.4 s of a 1 M-triangle load)
    int32_t lo = INT32_MAX, hi = -1;
    for (const obj_index& ix : shape.indices) {
        if (ix.v < 0) continue;
        lo = std::min(lo, (int32_t)ix.v);
        hi = std::max(hi, (int32_t)ix.v);
    }
    std::vector<int32_t> vertex_mapping(hi >= lo ? (size_t)(hi - lo) + 1 : 0, -1);
    size_t offset = 0;
    auto corner = [&](const obj_index& ix) -> int32_t {
        int32_t* it = &vertex_mapping[(size_t)(ix.v - lo)];
        if (*it < 0) {
            *it = (int32_t)(m->vertices.size() / 3);
            for (int k = 0; k < 3; ++k) m->vertices.push_back(at.vertices[3 * (size_t)ix.v + k]);
        }
        if (ix.vn >= 0)
            while (m->normals.size() < m->vertices.size())
                for (int k = 0; k < 3; ++k) m->normals.push_back(at.normals[3 * (size_t)ix.vn + k]);
        if (ix.vt >= 0)
            while (m->texcoords.size() / 2 < m->vertices.size() / 3)
                for (int k = 0; k < 2; ++k) m->texcoords.push_back(at.texcoords[2 * (size_t)ix.vt + k]);
        return *it;
    };
    for (int fs : shape.face_sizes) {
        for (int t = 1; t + 1 < fs; ++t) {
            const obj_index c[3] = {shape.indices[offset], shape.indices[offset + t], shape.indices[offset + t + 1]};
            for (int k = 0; k < 3; ++k) {
                if (c[k].v < 0 || 3 * (size_t)c[k].v + 2 >= at.vertices.size())
                    throw std::runtime_error("obj: vertex index out of range");
                m->indices.push_back(corner(c[k]));
            }
        }
        offset += (size_t)fs;
    }
    return m;
}
}  // namespace

std::vector<std::pair<std::string, std::shared_ptr<mesh>>> load_obj(const std::string& obj_file) {  // mesh_loader.cpp:85-121
    obj_attrib at;
    std::vector<obj_shape> shapes;
    std::string err;
    if (!parse_obj(obj_file, at, shapes, err)) throw std::runtime_error(err);
    // the shapes are independent: their vertex maps are built by concurrent threads, results keep the file order
    std::vector<std::pair<std::string, std::shared_ptr<mesh>>> out(shapes.size());
    std::atomic<size_t> next{0};
    std::mutex mu;
    std::string first_error;
    auto work = [&]() {
        for (;;) {
            const size_t i = next.fetch_add(1);
            if (i >= shapes.size()) break;
            try {
                out[i] = {shapes[i].name, create_mesh(shapes[i], at)};
            } catch (const std::exception& e) {
                std::lock_guard<std::mutex> g(mu);
                if (first_error.empty()) first_error = e.what();
            }
        }
    };
    size_t bytes = 0;
    for (const auto& sh : shapes) bytes += sh.indices.size() * sizeof(obj_index);
    const int T = (int)std::min<size_t>((size_t)obj_threads(bytes * 4), std::max<size_t>(shapes.size(), 1));
    std::vector<std::thread> th;
    for (int i = 1; i < T; ++i) th.emplace_back(work);
    work();
    for (auto& t : th) t.join();
    if (!first_error.empty()) throw std::runtime_error(first_error);
    return out;
}

// ------------------------------------------------------------------ PNG
static uint32_t be32(const unsigned char* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

void write_pfm(const std::string& file_path, int w, int h, const float* rgba_top_first) {
    FILE* f = fopen(file_path.c_str(), "wb");
    if (!f) throw std::runtime_error("cannot write " + file_path);
    fprintf(f, "PF\n%d %d\n-1.0\n", w, h);  // negative scale: little endian
    std::vector<float> row((size_t)w * 3);
    for (int y = h - 1; y >= 0; --y) {  // PFM stores the bottom row first
        const float* src = rgba_top_first + (size_t)y * (size_t)w * 4;
        for (int x = 0; x < w; ++x) {
            row[3 * (size_t)x + 0] = src[4 * (size_t)x + 0];
            row[3 * (size_t)x + 1] = src[4 * (size_t)x + 1];
            row[3 * (size_t)x + 2] = src[4 * (size_t)x + 2];
        }
        fwrite(row.data(), sizeof(float), row.size(), f);
    }
    fclose(f);
}

void write_png(const std::string& file_path, int w, int h, const uint32_t* rgba) {
    std::vector<unsigned char> raw((size_t)h * (1 + (size_t)w * 4));
    for (int y = 0; y < h; ++y) {
        unsigned char* row = &raw[(size_t)y * (1 + (size_t)w * 4)];
        row[0] = 0;
        memcpy(row + 1, rgba + (size_t)y * w, (size_t)w * 4);
    }
    uLongf clen = compressBound((uLong)raw.size());
    std::vector<unsigned char> comp(clen);
    if (compress2(comp.data(), &clen, raw.data(), (uLong)raw.size(), 6) != Z_OK) throw std::runtime_error("png: deflate failed");
    FILE* f = fopen(file_path.c_str(), "wb");
    if (!f) throw std::runtime_error("png: cannot write " + file_path);
    auto chunk = [&](const char* type, const unsigned char* d, uint32_t n) {
        unsigned char len[4] = {(unsigned char)(n >> 24), (unsigned char)(n >> 16), (unsigned char)(n >> 8), (unsigned char)n};
        fwrite(len, 1, 4, f);
        fwrite(type, 1, 4, f);
        if (n) fwrite(d, 1, n, f);
        uLong c = crc32(0L, (const Bytef*)type, 4);
        if (n) c = crc32(c, d, n);
        unsigned char cb[4] = {(unsigned char)(c >> 24), (unsigned char)(c >> 16), (unsigned char)(c >> 8), (unsigned char)c};
        fwrite(cb, 1, 4, f);
    };
    const unsigned char sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    fwrite(sig, 1, 8, f);
    unsigned char ihdr[13] = {(unsigned char)(w >> 24), (unsigned char)(w >> 16), (unsigned char)(w >> 8), (unsigned char)w,
                              (unsigned char)(h >> 24), (unsigned char)(h >> 16), (unsigned char)(h >> 8), (unsigned char)h,
                              8, 6, 0, 0, 0};
    chunk("IHDR", ihdr, 13);
    chunk("IDAT", comp.data(), (uint32_t)clen);
    chunk("IEND", nullptr, 0);
    fclose(f);
}

static bool read_png(const std::string& path, image_buffer& img) {
    std::ifstream f(path, std::ios::binary);
    if (!f) return false;
    std::vector<unsigned char> d((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
    if (d.size() < 8 || memcmp(d.data(), "\x89PNG\r\n\x1a\n", 8) != 0) return false;
    size_t p = 8;
    int w = 0, h = 0, depth = 0, ctype = 0, interlace = 0;
    std::vector<unsigned char> idat, plte, trns;
    while (p + 8 <= d.size()) {
        uint32_t n = be32(&d[p]);
        std::string type((const char*)&d[p + 4], 4);
        const unsigned char* body = &d[p + 8];
        if (p + 12 + n > d.size()) return false;
        if (type == "IHDR") {
            w = (int)be32(body);
            h = (int)be32(body + 4);
            depth = body[8];
            ctype = body[9];
            interlace = body[12];
        } else if (type == "IDAT")
            idat.insert(idat.end(), body, body + n);
        else if (type == "PLTE")
            plte.assign(body, body + n);
        else if (type == "tRNS")
            trns.assign(body, body + n);
        else if (type == "IEND")
            break;
        p += 12 + n;
    }
    if (w <= 0 || h <= 0 || interlace != 0 || (depth != 8 && depth != 16)) return false;
    const int channels = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : ctype == 6 ? 4 : 0;
    if (!channels || (ctype == 3 && depth != 8)) return false;
    const size_t bpp = (size_t)channels * depth / 8, stride = (size_t)w * bpp;
    std::vector<unsigned char> raw((size_t)h * (stride + 1));
    uLongf rl = (uLongf)raw.size();
    if (uncompress(raw.data(), &rl, idat.data(), (uLong)idat.size()) != Z_OK || rl != raw.size()) return false;
    std::vector<unsigned char> px((size_t)h * stride);
    for (int y = 0; y < h; ++y) {
        const unsigned char* in = &raw[(size_t)y * (stride + 1)];
        unsigned char* out = &px[(size_t)y * stride];
        const unsigned char* up = y ? out - stride : nullptr;
        const int ft = in[0];
        for (size_t i = 0; i < stride; ++i) {
            const int a = i >= bpp ? out[i - bpp] : 0, b = up ? up[i] : 0, c = (up && i >= bpp) ? up[i - bpp] : 0;
            int v = in[1 + i];
            switch (ft) {
                case 1: v += a; break;
                case 2: v += b; break;
                case 3: v += (a + b) >> 1; break;
                case 4: {
                    const int pa = abs(b - c), pb = abs(a - c), pc = abs(a + b - 2 * c);
                    v += (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
                    break;
                }
                default: break;
            }
            out[i] = (unsigned char)v;
        }
    }
    img.width = w;
    img.height = h;
    img.pixels.resize((size_t)w * h);
    const size_t step = depth / 8;
    for (size_t i = 0; i < (size_t)w * h; ++i) {
        const unsigned char* s = &px[i * bpp];
        unsigned r, g, b, a = 255;
        if (ctype == 0) { r = g = b = s[0]; }
        else if (ctype == 2) { r = s[0]; g = s[step]; b = s[2 * step]; }
        else if (ctype == 3) {
            const unsigned k = s[0];
            r = 3 * k + 2 < plte.size() ? plte[3 * k] : 0;
            g = 3 * k + 2 < plte.size() ? plte[3 * k + 1] : 0;
            b = 3 * k + 2 < plte.size() ? plte[3 * k + 2] : 0;
            a = k < trns.size() ? trns[k] : 255;
        } else if (ctype == 4) { r = g = b = s[0]; a = s[step]; }
        else { r = s[0]; g = s[step]; b = s[2 * step]; a = s[3 * step]; }
        img.pixels[i] = r | (g << 8) | (b << 16) | (a << 24);
    }
    return true;
}

// ------------------------------------------------------------------ Radiance .hdr -> RGBA8
// stbi_load(..., STBI_rgb_alpha) on an .hdr: decode RGBE to float (stb_image.h:7053-7078),
// then stbi__hdr_to_ldr (stb_image.h:1864-1889): pow(x, 1/2.2)*255 + 0.5, truncated.
static unsigned char hdr_channel_to_ldr(float v) {
    float z = (float)pow((double)(v * 1.0f), (double)(1.0f / 2.2f)) * 255 + 0.5f;
    if (z < 0) z = 0;
    if (z > 255) z = 255;
    return (unsigned char)(int)z;
}

image_buffer ldr_from_rgbe(int w, int h, const unsigned char* rgbe) {
    image_buffer img;
    img.width = w;
    img.height = h;
    img.pixels.resize((size_t)w * h);
    for (size_t i = 0; i < (size_t)w * h; ++i) {
        const unsigned char* e = &rgbe[i * 4];
        float c[3] = {0, 0, 0};
        if (e[3] != 0) {
            const float f1 = (float)ldexp(1.0f, (int)e[3] - (int)(128 + 8));
            c[0] = e[0] * f1;
            c[1] = e[1] * f1;
            c[2] = e[2] * f1;
        }
        img.pixels[i] = (uint32_t)hdr_channel_to_ldr(c[0]) | ((uint32_t)hdr_channel_to_ldr(c[1]) << 8) |
                        ((uint32_t)hdr_channel_to_ldr(c[2]) << 16) | (255u << 24);
    }
    return img;
}

void rgbe_from_float(const float* c, unsigned char* o) {
    float m = fmaxf(c[0], fmaxf(c[1], c[2]));
    if (m < 1e-32f) {
        o[0] = o[1] = o[2] = o[3] = 0;
    } else {
        int e;
        float s = frexpf(m, &e) * 256.0f / m;
        o[0] = (unsigned char)(c[0] * s);
        o[1] = (unsigned char)(c[1] * s);
        o[2] = (unsigned char)(c[2] * s);
        o[3] = (unsigned char)(e + 128);
    }
}

// Radiance file -> RGBE bytes (w*h*4), shared by the LDR route of the reference and the float route
static bool read_hdr_rgbe(const std::string& path, int& w_out, int& h_out, std::vector<unsigned char>& rgbe_out) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) return false;
    char line[1024];
    bool ok = false;
    if (fgets(line, sizeof line, f) && (strncmp(line, "#?RADIANCE", 10) == 0 || strncmp(line, "#?RGBE", 6) == 0)) {
        bool format = false;
        while (fgets(line, sizeof line, f)) {
            if (line[0] == '\n' || line[0] == '\r') break;
            if (strncmp(line, "FORMAT=32-bit_rle_rgbe", 22) == 0) format = true;
        }
        int w = 0, h = 0;
        if (format && fgets(line, sizeof line, f) && sscanf(line, "-Y %d +X %d", &h, &w) == 2 && w > 0 && h > 0) {
            std::vector<unsigned char> rgbe((size_t)w * h * 4);
            ok = true;
            std::vector<unsigned char> scan((size_t)w * 4);
            for (int y = 0; y < h && ok; ++y) {
                unsigned char hd[4];
                if (fread(hd, 1, 4, f) != 4) { ok = false; break; }
                unsigned char* row = &rgbe[(size_t)y * w * 4];
                if (w < 8 || w >= 32768 || hd[0] != 2 || hd[1] != 2 || (hd[2] & 0x80)) {
                    // flat data: this scanline (and, as in stb, the rest of the file) is raw RGBE
                    memcpy(row, hd, 4);
                    size_t rest = ((size_t)w * (h - y) - 1) * 4;
                    if (fread(row + 4, 1, rest, f) != rest) ok = false;
                    break;
                }
                if (((hd[2] << 8) | hd[3]) != w) { ok = false; break; }
                for (int k = 0; k < 4 && ok; ++k) {
                    int i = 0;
                    while (i < w) {
                        int count = fgetc(f);
                        if (count == EOF) { ok = false; break; }
                        if (count > 128) {
                            int value = fgetc(f);
                            count -= 128;
                            if (count > w - i) { ok = false; break; }
                            for (int z = 0; z < count; ++z) scan[(size_t)(i++) * 4 + k] = (unsigned char)value;
                        } else {
                            if (count == 0 || count > w - i) { ok = false; break; }
                            for (int z = 0; z < count; ++z) scan[(size_t)(i++) * 4 + k] = (unsigned char)fgetc(f);
                        }
                    }
                }
                if (ok) memcpy(row, scan.data(), (size_t)w * 4);
            }
            if (ok) {
                w_out = w;
                h_out = h;
                rgbe_out.swap(rgbe);
            }
        }
    }
    fclose(f);
    return ok;
}
static bool read_hdr(const std::string& path, image_buffer& img) {
    int w = 0, h = 0;
    std::vector<unsigned char> rgbe;
    if (!read_hdr_rgbe(path, w, h, rgbe)) return false;
    img = ldr_from_rgbe(w, h, rgbe.data());
    return true;
}
// stbi__hdr_convert (stb_image.h:7053-7078): mantissa * 2^(e - 136); rows flipped like flip_rows
void float_from_rgbe(int w, int h, const unsigned char* rgbe, std::vector<float>& rgb_flipped) {
    rgb_flipped.assign((size_t)w * h * 3, 0.0f);
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            const unsigned char* e = &rgbe[((size_t)y * w + x) * 4];
            float* o = &rgb_flipped[((size_t)(h - 1 - y) * w + x) * 3];
            if (e[3] != 0) {
                const float f1 = (float)ldexp(1.0f, (int)e[3] - (int)(128 + 8));
                o[0] = e[0] * f1;
                o[1] = e[1] * f1;
                o[2] = e[2] * f1;
            }
        }
}
bool load_hdr_float(const std::string& name, const std::string& path, int& w, int& h, std::vector<float>& rgb_flipped) {
    const std::string file_path = path + "/" + name;
    std::vector<unsigned char> rgbe;
    if (!std::filesystem::exists(file_path) || !read_hdr_rgbe(file_path, w, h, rgbe)) return false;
    float_from_rgbe(w, h, rgbe.data(), rgb_flipped);
    return true;
}

void write_hdr(const std::string& path, int w, int h, const float* rgb) {
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) throw std::runtime_error("hdr: cannot write " + path);
    fprintf(f, "#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n-Y %d +X %d\n", h, w);
    if (w < 8 || w >= 32768) {
        fclose(f);
        throw std::runtime_error("hdr: width must be in [8, 32768)");
    }
    std::vector<unsigned char> row((size_t)w * 4);
    for (int y = 0; y < h; ++y) {
        for (int x = 0; x < w; ++x) {
            rgbe_from_float(rgb + 3 * ((size_t)y * w + x), &row[(size_t)x * 4]);
        }
        // new-style RLE scanline made of literal runs only (valid, just not compressed)
        const unsigned char hd[4] = {2, 2, (unsigned char)(w >> 8), (unsigned char)(w & 255)};
        fwrite(hd, 1, 4, f);
        for (int k = 0; k < 4; ++k) {
            int i = 0;
            while (i < w) {
                const int n = (w - i) < 128 ? (w - i) : 128;
                fputc(n, f);
                for (int z = 0; z < n; ++z) fputc(row[(size_t)(i + z) * 4 + k], f);
                i += n;
            }
        }
    }
    fclose(f);
}

void flip_rows(image_buffer& img) {  // image_buffer.cpp:50-55
    for (int y = 0; y < img.height / 2; ++y) {
        uint32_t* a = &img.pixels[(size_t)y * img.width];
        uint32_t* b = &img.pixels[(size_t)(img.height - 1 - y) * img.width];
        for (int x = 0; x < img.width; ++x) std::swap(a[x], b[x]);
    }
}

image_buffer load_image(const std::string& name, const std::string& path) {  // image_buffer.cpp:36-59
    const std::string file_path = path + "/" + name;
    image_buffer img;
    if (!std::filesystem::exists(file_path)) {
        say("Image file %s does not exist. Continue with empty.\n", file_path.c_str());
        return img;
    }
    bool ok = false;
    if (file_path.size() > 4 && file_path.substr(file_path.size() - 4) == ".hdr")
        ok = read_hdr(file_path, img);
    else
        ok = read_png(file_path, img);
    if (!ok) {
        say("Image file %s could not be decoded. Continue with empty.\n", file_path.c_str());
        return image_buffer{};
    }
    flip_rows(img);
    return img;
}

void write_image(const image_buffer& i, const std::string& name, const std::string& path) {  // image_buffer.cpp:25-34
    if (i.pixels.empty()) throw std::runtime_error{"image buffer is null"};
    const std::string file_path = path + "/" + name;
    write_png(file_path, i.width, i.height, i.pixels.data());
    say("Image written to %s\n", std::filesystem::absolute(file_path).string().c_str());
}

}  // namespace host
