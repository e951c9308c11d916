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
// The subset of tinyobjloader's LoadObj (extern/tinyobjloader/tiny_obj_loader.h:2432-2994, the version the reference
// vendors) that mesh_loader.cpp relies on with triangulate=false: v / vn / vt / f / o / g; 1-based and negative indices;
// everything else (mtllib, usemtl, s, l, p, t, vw, vertex colours) is ignored.  The behaviours below are tinyobj's and
// are held to it byte for byte by tests/test_ref_host.py (oracle/_ref/libref_host.so = the reference's own loader):
//   * numbers are read by tinyobj's own decimal routine (tryParseDouble, :866-1010), which is NOT correctly rounded —
//     digits are accumulated in binary64 (integer part: m = 10 m + d; fraction: m += d * 10^-k; exponent: ldexp(m * 5^e, e))
//     and the result is cast to float; a token that does not parse yields 0;
//   * a line ends at "\n", "\r\n" or a lone "\r" (safeGetline);
//   * `f`: each corner is i, i/j, i//k or i/j/k read with atoi; index 0 or an unparsable corner fails the whole file
//     ("Failed parse `f' line"); faces with fewer than 3 corners are dropped;
//   * `o <rest of line>` names the following faces (verbatim, inner and trailing blanks kept); `g a b` joins its words with
//     one blank; a shape is emitted whenever an o/g line (or the end of file) follows at least one kept face.
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

inline int fix_index(int idx, int n) {   // tiny_obj_loader.h:800-822 (0 is rejected by the parser before this point)
    return idx > 0 ? idx - 1 : n + idx;
}

// tinyobj's decimal parser restated (tryParseDouble): same operations in the same order, so the same binary64 value.
bool tinyobj_parse_double(const char* s, const char* s_end, double* result) {
    if (s >= s_end) return false;
    static const double neg_pow10[] = {1.0, 0.1, 0.01, 0.001, 0.0001, 0.00001, 0.000001, 0.0000001};
    auto digit = [](char c) { return (unsigned)(c - '0') < 10u; };
    double mantissa = 0.0;
    int exponent = 0;
    bool negative = false, exp_negative = false, leading_dot = false;
    const char* c = s;
    if (*c == '+' || *c == '-') {
        negative = *c == '-';
        ++c;
        if (c != s_end && *c == '.') leading_dot = true;
    } else if (*c == '.') {
        leading_dot = true;
    } else if (!digit(*c)) {
        return false;
    }
    int read = 0;
    if (!leading_dot) {
        for (; c != s_end && digit(*c); ++c, ++read) {
            mantissa *= 10;
            mantissa += (int)(*c - '0');
        }
        if (read == 0) return false;
    }
    if (c != s_end) {
        bool more = true;
        if (*c == '.') {
            ++c;
            for (read = 1; c != s_end && digit(*c); ++c, ++read)
                mantissa += (int)(*c - '0') * (read < 8 ? neg_pow10[read] : std::pow(10.0, -read));
        } else if (*c != 'e' && *c != 'E') {
            more = false;
        }
        if (more && c != s_end && (*c == 'e' || *c == 'E')) {
            ++c;
            if (c != s_end && (*c == '+' || *c == '-')) {
                exp_negative = *c == '-';
                ++c;
            } else if (c == s_end || !digit(*c)) {
                return false;
            }
            for (read = 0; c != s_end && digit(*c); ++c, ++read) {
                if (exponent > 2147483647 / 10) return false;
                exponent *= 10;
                exponent += (int)(*c - '0');
            }
            if (exp_negative) exponent = -exponent;
            if (read == 0) return false;
        }
    }
    *result = (negative ? -1 : 1) * (exponent ? std::ldexp(mantissa * std::pow(5.0, exponent), exponent) : mantissa);
    return true;
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
    size_t lines = 0;               // line ends seen (for the line number of an error in a later chunk)
    size_t bad_face_line = 0;       // 1-based line of the chunk whose `f` failed to parse, 0 = none
};

void parse_obj_chunk(const char* p, const char* end, obj_chunk& c) {
    auto is_sp = [](char ch) { return ch == ' ' || ch == '\t'; };
    auto is_delim = [](char ch) { return ch == ' ' || ch == '\t' || ch == '\r'; };
    // next blank-delimited token of the line as a float (parseReal, tiny_obj_loader.h:1012-1020)
    auto real = [&](const char*& q, const char* eol) {
        while (q < eol && is_sp(*q)) ++q;
        const char* e = q;
        while (e < eol && !is_delim(*e)) ++e;
        double val = 0.0;
        tinyobj_parse_double(q, e, &val);
        q = e;
        return (float)val;
    };
    // atoi on a bounded line: optional blanks, sign, digits; anything else reads as 0
    auto integer = [&](const char* q, const char* eol) {
        while (q < eol && (is_sp(*q) || *q == '\r' || *q == '\v' || *q == '\f')) ++q;
        bool neg = false;
        if (q < eol && (*q == '+' || *q == '-')) neg = *q++ == '-';
        long v = 0;
        for (; q < eol && (unsigned)(*q - '0') < 10u; ++q) v = v * 10 + (*q - '0');
        return (int)(neg ? -v : v);
    };
    auto skip_index = [&](const char*& q, const char* eol) {   // token += strcspn(token, "/ \t\r")
        while (q < eol && *q != '/' && !is_delim(*q)) ++q;
    };
    while (p < end) {
        const char* eol = p;
        while (eol < end && *eol != '\n' && *eol != '\r') ++eol;
        const char* next = eol < end ? eol + ((*eol == '\r' && eol + 1 < end && eol[1] == '\n') ? 2 : 1) : end;
        ++c.lines;
        const char* q = p;
        while (q < eol && is_sp(*q)) ++q;
        if (q + 1 < eol && q[0] == 'v' && is_sp(q[1])) {
            q += 2;
            for (int k = 0; k < 3; ++k) c.vertices.push_back(real(q, eol));
        } else if (q + 2 < eol && q[0] == 'v' && q[1] == 'n' && is_sp(q[2])) {
            q += 3;
            for (int k = 0; k < 3; ++k) c.normals.push_back(real(q, eol));
        } else if (q + 2 < eol && q[0] == 'v' && q[1] == 't' && is_sp(q[2])) {
            q += 3;
            for (int k = 0; k < 2; ++k) c.texcoords.push_back(real(q, eol));
        } else if (q + 1 < eol && q[0] == 'f' && is_sp(q[1])) {
            q += 2;
            while (q < eol && is_sp(*q)) ++q;
            const size_t first = c.indices.size();
            int n = 0;
            bool bad = false;
            while (q < eol) {   // parseTriple, tiny_obj_loader.h:1136-1186
                obj_chunk::raw_index ix;
                ix.v = integer(q, eol);
                bad |= ix.v == 0;
                skip_index(q, eol);
                if (!bad && q < eol && *q == '/') {
                    ++q;
                    if (q < eol && *q == '/') {
                        ++q;
                        ix.vn = integer(q, eol);
                        ix.has_vn = true;
                        bad |= ix.vn == 0;
                        skip_index(q, eol);
                    } else {
                        ix.vt = integer(q, eol);
                        ix.has_vt = true;
                        bad |= ix.vt == 0;
                        skip_index(q, eol);
                        if (!bad && q < eol && *q == '/') {
                            ++q;
                            ix.vn = integer(q, eol);
                            ix.has_vn = true;
                            bad |= ix.vn == 0;
                            skip_index(q, eol);
                        }
                    }
                }
                if (bad) break;
                c.indices.push_back(ix);
                ++n;
                while (q < eol && is_delim(*q)) ++q;
            }
            if (bad) {
                if (!c.bad_face_line) c.bad_face_line = c.lines;
                c.indices.resize(first);
            } else if (n >= 3) {   // "Degenerated face found": dropped by exportGroupsToShape (:1416-1424)
                c.face_counts.push_back((int)c.vertices.size() / 3);
                c.face_counts.push_back((int)c.normals.size() / 3);
                c.face_counts.push_back((int)c.texcoords.size() / 2);
                c.face_sizes.push_back(n);
            } else {
                c.indices.resize(first);
            }
        } else if (q + 1 < eol && q[0] == 'g' && is_sp(q[1])) {   // :2778-2827: words after the `g`, joined by one blank
            std::string name;
            const char* w = q + 1;
            for (int words = 0;; ++words) {
                while (w < eol && is_sp(*w)) ++w;
                if (w >= eol) break;
                const char* e = w;
                while (e < eol && !is_delim(*e)) ++e;
                if (words) name += ' ';
                name.append(w, e);
                w = e;
                while (w < eol && is_delim(*w)) ++w;
            }
            c.breaks.push_back({c.face_sizes.size(), name});
        } else if (q + 1 < eol && q[0] == 'o' && is_sp(q[1])) {   // :2830-2849: the rest of the line, verbatim
            c.breaks.push_back({c.face_sizes.size(), std::string(q + 2, eol)});
        }
        p = next;
    }
}

int obj_threads(size_t bytes) {
    if (const char* e = getenv("B200PT_OBJ_THREADS")) return std::max(1, atoi(e));
    const size_t by_size = bytes / (256u << 10);   // a chunk below 256 KB is not worth a thread
    return (int)std::max<size_t>(1, std::min<size_t>({by_size, (size_t)std::thread::hardware_concurrency(), (size_t)16}));
}

bool parse_obj(const std::string& path, int threads, obj_attrib& at, std::vector<obj_shape>& shapes, std::string& err) {
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
    const int T = threads > 0 ? threads : obj_threads(data.size());
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
    size_t tv = 0, tn = 0, tt = 0, lines_before = 0;
    for (const obj_chunk& c : chunks) {
        if (c.bad_face_line) {   // tiny_obj_loader.h:2664-2671; mesh_loader.cpp:99-101 turns it into a runtime_error
            err = "Failed parse `f' line(e.g. zero value for face index. line " + std::to_string(lines_before + c.bad_face_line) + ".)\n";
            return false;
        }
        lines_before += c.lines;
    }
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
        // tinyobj only warns about attribute indices past the end and the reference then reads out of bounds
        // (mesh_loader.cpp:56-60, :69-72); here such a file is refused
        if (ix.vn >= 0) {
            if (3 * (size_t)ix.vn + 2 >= at.normals.size()) throw std::runtime_error("obj: normal index out of range");
            while (m->normals.size() < m->vertices.size())
                for (int k = 0; k < 3; ++k) m->normals.push_back(at.normals[3 * (size_t)ix.vn + k]);
        }
        if (ix.vt >= 0) {
            if (2 * (size_t)ix.vt + 1 >= at.texcoords.size()) throw std::runtime_error("obj: texcoord index out of range");
            while (m->texcoords.size() / 2 < m->vertices.size() / 3)
                for (int k = 0; k < 2; ++k) m->texcoords.push_back(at.texcoords[2 * (size_t)ix.vt + k]);
        }
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

std::vector<std::pair<std::string, std::shared_ptr<mesh>>> load_obj(const std::string& obj_file, int threads,
                                                                     const create_mesh_fn& create) {  // mesh_loader.cpp:85-121
    obj_attrib at;
    std::vector<obj_shape> shapes;
    std::string err;
    if (!parse_obj(obj_file, threads, at, shapes, err)) throw std::runtime_error(err);
    if (create) {   // a pluggable create_mesh (the GPU de-duplication): shapes one after the other, corners fan-split here
        std::vector<std::pair<std::string, std::shared_ptr<mesh>>> out;
        std::vector<int32_t> cv, cn, ct;
        for (const obj_shape& sh : shapes) {
            cv.clear(); cn.clear(); ct.clear();
            size_t offset = 0;
            for (int fs : sh.face_sizes) {
                for (int t = 1; t + 1 < fs; ++t)
                    for (size_t k : {offset, offset + (size_t)t, offset + (size_t)t + 1}) {
                        const obj_index& ix = sh.indices[k];
                        if (ix.v < 0 || 3 * (size_t)ix.v + 2 >= at.vertices.size()) throw std::runtime_error("obj: vertex index out of range");
                        if (ix.vn >= 0 && 3 * (size_t)ix.vn + 2 >= at.normals.size()) throw std::runtime_error("obj: normal index out of range");
                        if (ix.vt >= 0 && 2 * (size_t)ix.vt + 1 >= at.texcoords.size()) throw std::runtime_error("obj: texcoord index out of range");
                        cv.push_back(ix.v); cn.push_back(ix.vn); ct.push_back(ix.vt);
                    }
                offset += (size_t)fs;
            }
            const raw_shape raw{at.vertices.data(), at.vertices.size() / 3, at.normals.data(), at.normals.size() / 3,
                                at.texcoords.data(), at.texcoords.size() / 2, cv.data(), cn.data(), ct.data(), cv.size()};
            std::shared_ptr<mesh> m = create(raw);
            if (!m) m = create_mesh(sh, at);
            out.emplace_back(sh.name, std::move(m));
        }
        return out;
    }
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
    const int T = (int)std::min<size_t>((size_t)(threads > 0 ? threads : obj_threads(bytes * 4)), std::max<size_t>(shapes.size(), 1));
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

// OpenEXR, single part, scan lines, no compression, three FLOAT channels (stored in the file format's
// alphabetical channel order B, G, R; each scan line holds its channels one after the other)
void write_exr(const std::string& file_path, int w, int h, const float* rgba_top_first) {
    if (w <= 0 || h <= 0) throw std::runtime_error("write_exr: empty image");
    std::vector<unsigned char> head;
    auto put = [&](const void* p, size_t n) { head.insert(head.end(), (const unsigned char*)p, (const unsigned char*)p + n); };
    auto put_str = [&](const char* z) { put(z, strlen(z) + 1); };
    auto put_i32 = [&](int32_t v) { put(&v, 4); };
    auto put_f32 = [&](float v) { put(&v, 4); };
    auto attr = [&](const char* name, const char* type, int32_t size) {
        put_str(name);
        put_str(type);
        put_i32(size);
    };
    const unsigned char magic[8] = {0x76, 0x2f, 0x31, 0x01, 2, 0, 0, 0};  // magic, version 2, no flags
    put(magic, 8);
    attr("channels", "chlist", 3 * (2 + 16) + 1);
    for (const char* c : {"B", "G", "R"}) {
        put_str(c);
        put_i32(2);  // FLOAT
        put_i32(0);  // pLinear + 3 reserved bytes
        put_i32(1);  // x sampling
        put_i32(1);  // y sampling
    }
    head.push_back(0);
    attr("compression", "compression", 1);
    head.push_back(0);  // NO_COMPRESSION
    for (const char* win : {"dataWindow", "displayWindow"}) {
        attr(win, "box2i", 16);
        put_i32(0);
        put_i32(0);
        put_i32(w - 1);
        put_i32(h - 1);
    }
    attr("lineOrder", "lineOrder", 1);
    head.push_back(0);  // INCREASING_Y: the top row first, as the frame is stored
    attr("pixelAspectRatio", "float", 4);
    put_f32(1.0f);
    attr("screenWindowCenter", "v2f", 8);
    put_f32(0.0f);
    put_f32(0.0f);
    attr("screenWindowWidth", "float", 4);
    put_f32(1.0f);
    head.push_back(0);  // end of header
    FILE* f = fopen(file_path.c_str(), "wb");
    if (!f) throw std::runtime_error("cannot write " + file_path);
    fwrite(head.data(), 1, head.size(), f);
    const uint64_t line_bytes = (uint64_t)w * 3 * sizeof(float), chunk = 8 + line_bytes;
    for (int y = 0; y < h; ++y) {  // offset table
        const uint64_t off = head.size() + (uint64_t)h * 8 + (uint64_t)y * chunk;
        fwrite(&off, 8, 1, f);
    }
    std::vector<float> row((size_t)w * 3);
    for (int y = 0; y < h; ++y) {
        const float* src = rgba_top_first + (size_t)y * (size_t)w * 4;
        for (int x = 0; x < w; ++x) {
            row[(size_t)x] = src[4 * (size_t)x + 2];
            row[(size_t)w + x] = src[4 * (size_t)x + 1];
            row[2 * (size_t)w + x] = src[4 * (size_t)x + 0];
        }
        const int32_t yy = y, size = (int32_t)line_bytes;
        fwrite(&yy, 4, 1, f);
        fwrite(&size, 4, 1, f);
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

// PNG decoder with the coverage and the conversions of stb_image (extern/stb/stb_image.h, the decoder behind the
// reference's load_image): grey / RGB / palette / grey+alpha / RGBA, 1-16 bits, Adam7 interlace, tRNS (palette alpha,
// colour key for grey and RGB).  Output is RGBA8 the way stbi_load(..., STBI_rgb_alpha) delivers it: grey below 8 bits
// scaled by 255 / 85 / 17, 16-bit samples reduced to their high byte AFTER the colour key was applied on 16 bits.
// Held to stb byte for byte by tests/test_ref_host.py.
static bool png_unfilter(const unsigned char* in, size_t in_size, int w, int h, int channels, int depth,
                         std::vector<unsigned char>& out, size_t& consumed) {
    const size_t bits = (size_t)w * channels * depth, stride = (bits + 7) / 8;
    const size_t bpp = std::max<size_t>(1, (size_t)channels * depth / 8);
    if (in_size < (stride + 1) * (size_t)h) return false;
    out.assign(stride * (size_t)h, 0);
    for (int y = 0; y < h; ++y) {
        const unsigned char* src = in + (size_t)y * (stride + 1);
        unsigned char* row = &out[(size_t)y * stride];
        const unsigned char* up = y ? row - stride : nullptr;
        const int ft = src[0];
        if (ft > 4) return false;
        for (size_t i = 0; i < stride; ++i) {
            const int a = i >= bpp ? row[i - bpp] : 0, b = up ? up[i] : 0, c = (up && i >= bpp) ? up[i - bpp] : 0;
            int v = src[1 + i];
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
            row[i] = (unsigned char)v;
        }
    }
    consumed = (stride + 1) * (size_t)h;
    return true;
}

// sample k of a packed row (depth 1, 2, 4, 8 or 16 bits, big endian)
static inline unsigned png_sample(const unsigned char* row, size_t k, int depth) {
    if (depth == 8) return row[k];
    if (depth == 16) return ((unsigned)row[2 * k] << 8) | row[2 * k + 1];
    const size_t bit = k * (size_t)depth;
    return (row[bit >> 3] >> (8 - depth - (int)(bit & 7))) & ((1u << depth) - 1u);
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
        const uint32_t n = be32(&d[p]);
        const std::string type((const char*)&d[p + 4], 4);
        if (p + 12 + (size_t)n > d.size()) return false;
        const unsigned char* body = &d[p + 8];
        if (type == "IHDR") {
            if (n != 13) return false;
            w = (int)be32(body);
            h = (int)be32(body + 4);
            depth = body[8];
            ctype = body[9];
            interlace = body[12];
            if (body[10] != 0 || body[11] != 0) return false;
        } else if (type == "IDAT")
            idat.insert(idat.end(), body, body + n);
        else if (type == "PLTE")
            plte.assign(body, body + n);
        else if (type == "tRNS")
            trns.assign(body, body + n);
        else if (type == "IEND")
            break;
        p += 12 + (size_t)n;
    }
    if (w <= 0 || h <= 0 || interlace > 1) return false;
    if (depth != 1 && depth != 2 && depth != 4 && depth != 8 && depth != 16) return false;
    const int channels = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 3 ? 1 : ctype == 4 ? 2 : ctype == 6 ? 4 : 0;
    if (!channels) return false;
    if (ctype == 3 && (depth == 16 || plte.empty())) return false;
    if ((ctype == 2 || ctype == 4 || ctype == 6) && depth < 8) return false;

    // inflate: the size is known from the header (per pass when interlaced)
    static const int px0[7] = {0, 4, 0, 2, 0, 1, 0}, py0[7] = {0, 0, 4, 0, 2, 0, 1};
    static const int pdx[7] = {8, 8, 4, 4, 2, 2, 1}, pdy[7] = {8, 8, 8, 4, 4, 2, 2};
    auto row_bytes = [&](int pw) { return ((size_t)pw * channels * depth + 7) / 8 + 1; };
    size_t raw_size = 0;
    if (interlace)
        for (int k = 0; k < 7; ++k) {
            const int pw = (w - px0[k] + pdx[k] - 1) / pdx[k], ph = (h - py0[k] + pdy[k] - 1) / pdy[k];
            if (pw > 0 && ph > 0) raw_size += row_bytes(pw) * (size_t)ph;
        }
    else
        raw_size = row_bytes(w) * (size_t)h;
    std::vector<unsigned char> raw(raw_size);
    uLongf rl = (uLongf)raw.size();
    const int zr = uncompress(raw.data(), &rl, idat.data(), (uLong)idat.size());
    if ((zr != Z_OK && zr != Z_BUF_ERROR) || rl != raw.size()) return false;

    // samples of the whole image, 16 bits each
    std::vector<uint16_t> smp((size_t)w * h * channels);
    std::vector<unsigned char> rows;
    if (!interlace) {
        size_t used = 0;
        if (!png_unfilter(raw.data(), raw.size(), w, h, channels, depth, rows, used)) return false;
        const size_t stride = ((size_t)w * channels * depth + 7) / 8;
        for (int y = 0; y < h; ++y)
            for (size_t k = 0; k < (size_t)w * channels; ++k)
                smp[(size_t)y * w * channels + k] = (uint16_t)png_sample(&rows[(size_t)y * stride], k, depth);
    } else {
        size_t off = 0;
        for (int k = 0; k < 7; ++k) {
            const int pw = (w - px0[k] + pdx[k] - 1) / pdx[k], ph = (h - py0[k] + pdy[k] - 1) / pdy[k];
            if (pw <= 0 || ph <= 0) continue;
            size_t used = 0;
            if (!png_unfilter(raw.data() + off, raw.size() - off, pw, ph, channels, depth, rows, used)) return false;
            off += used;
            const size_t stride = ((size_t)pw * channels * depth + 7) / 8;
            for (int j = 0; j < ph; ++j)
                for (int i = 0; i < pw; ++i) {
                    const size_t dst = ((size_t)(py0[k] + j * pdy[k]) * w + (size_t)(px0[k] + i * pdx[k])) * channels;
                    for (int c = 0; c < channels; ++c)
                        smp[dst + c] = (uint16_t)png_sample(&rows[(size_t)j * stride], (size_t)i * channels + c, depth);
                }
        }
    }

    img.width = w;
    img.height = h;
    img.pixels.resize((size_t)w * h);
    // 16 -> 8 bits: high byte; below 8 bits grey is scaled to the full range (palette indices are not)
    const unsigned scale = depth == 1 ? 255u : depth == 2 ? 85u : depth == 4 ? 17u : 1u;
    auto to8 = [&](unsigned v) { return depth == 16 ? (v >> 8) & 255u : v; };
    // colour key (tRNS of grey / RGB): compared on the stored sample width
    unsigned key[3] = {0, 0, 0};
    const bool has_key = (ctype == 0 && trns.size() >= 2) || (ctype == 2 && trns.size() >= 6);
    if (has_key)
        for (int c = 0; c < channels; ++c) {
            const unsigned v = ((unsigned)trns[2 * (size_t)c] << 8) | trns[2 * (size_t)c + 1];
            key[c] = depth == 16 ? v : (v & 255u) * scale;
        }
    for (size_t i = 0; i < (size_t)w * h; ++i) {
        const uint16_t* s = &smp[i * channels];
        unsigned r, g, b, a = 255;
        if (ctype == 0) {
            const unsigned v = depth < 8 ? s[0] * scale : s[0];
            if (has_key && v == key[0]) a = 0;
            r = g = b = to8(v);
        } else if (ctype == 2) {
            if (has_key && s[0] == key[0] && s[1] == key[1] && s[2] == key[2]) a = 0;
            r = to8(s[0]);
            g = to8(s[1]);
            b = to8(s[2]);
        } else if (ctype == 3) {
            const size_t k = s[0];
            r = 3 * k + 2 < plte.size() ? plte[3 * k] : 0;
            g = 3 * k + 2 < plte.size() ? plte[3 * k + 1] : 0;
            b = 3 * k + 2 < plte.size() ? plte[3 * k + 2] : 0;
            a = k < trns.size() ? trns[k] : 255;
        } else if (ctype == 4) {
            r = g = b = to8(s[0]);
            a = to8(s[1]);
        } else {
            r = to8(s[0]);
            g = to8(s[1]);
            b = to8(s[2]);
            a = to8(s[3]);
        }
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
                    // not run-length encoded: stb (stbi__hdr_load, "yes, this makes no sense") takes these four bytes as
                    // pixel 0 OF THE IMAGE and reads everything after them as flat RGBE from pixel 1 on, whatever the
                    // scanline it met them in; bytes past the end of the file read as 0
                    memcpy(&rgbe[0], hd, 4);
                    const size_t rest = ((size_t)w * h - 1) * 4;
                    const size_t got = fread(&rgbe[4], 1, rest, f);
                    if (got < rest) memset(&rgbe[4 + got], 0, rest - got);
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
    // stbi_load decides by content, not by extension: Radiance files start with "#?RADIANCE" / "#?RGBE", PNG with its
    // signature.  stb also reads JPEG / BMP / TGA / GIF / PSD / PIC / PNM, which this loader does not: such a file is an
    // error here, never a silent stand-in.
    unsigned char magic[10] = {0};
    {
        FILE* f = fopen(file_path.c_str(), "rb");
        if (f) {
            const size_t got = fread(magic, 1, sizeof magic, f);
            (void)got;
            fclose(f);
        }
    }
    bool ok = false;
    if (memcmp(magic, "#?RADIANCE", 10) == 0 || memcmp(magic, "#?RGBE", 6) == 0)
        ok = read_hdr(file_path, img);
    else if (memcmp(magic, "\x89PNG\r\n\x1a\n", 8) == 0)
        ok = read_png(file_path, img);
    else
        throw std::runtime_error("unsupported image format (PNG and Radiance .hdr are read): " + file_path);
    if (!ok) throw std::runtime_error("image file could not be decoded: " + file_path);
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
