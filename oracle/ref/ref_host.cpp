// ref_host.cpp — C ABI around the REFERENCE'S OWN host loaders, compiled as they are.
// TEST INFRASTRUCTURE: built into oracle/_ref/libref_host.so (git-ignored); loaded only by tests/.
//
// Included translation units of /root/reference/path_tracer/src (copies under oracle/_ref/build/src):
//   utils/mesh_loader.cpp   load_obj / create_mesh on extern/tinyobjloader/tiny_obj_loader.h
//   utils/image_buffer.cpp  load_image / write_image on extern/stb/stb_image.h, stb_image_write.h
//   utils/parser.cpp        parse_materials / parse_camera / parse_settings on extern/json
//   camera.cpp              to_camera_data
// The product's own loaders (owl_path_tracer_b200/host/loaders.cpp) are diffed against these in
// tests/test_ref_host.py; nothing here is restated.
#include "utils/mesh_loader.cpp"
#include "utils/image_buffer.cpp"
#include "utils/parser.cpp"
#include "camera.cpp"

namespace {
thread_local std::string g_err;
using shape_list = std::vector<std::tuple<std::string, std::shared_ptr<mesh>>>;
using material_list = std::vector<std::tuple<std::string, material_data, std::string>>;
struct RefImage {
    int32_t w = 0, h = 0;
    uint32_t* pixels = nullptr;   // stbi allocation (image_buffer::tag::manual is never freed by the reference)
};
}  // namespace

extern "C" {

const char* ref_host_last_error() { return g_err.c_str(); }

// ---- mesh_loader.cpp:85-121
void* ref_obj_load(const char* obj_file) {
    try {
        return new shape_list(load_obj(obj_file));
    } catch (std::exception const& e) {
        g_err = e.what();
        return nullptr;
    }
}
void ref_obj_free(void* h) { delete (shape_list*)h; }
int ref_obj_count(void* h) { return (int)((shape_list*)h)->size(); }
const char* ref_obj_name(void* h, int i) { return std::get<0>((*(shape_list*)h)[i]).c_str(); }
// sizes: triangles, vertices, normals, texcoords
void ref_obj_sizes(void* h, int i, int64_t* out) {
    mesh const& m = *std::get<1>((*(shape_list*)h)[i]);
    out[0] = (int64_t)m.indices.size();
    out[1] = (int64_t)m.vertices.size();
    out[2] = (int64_t)m.normals.size();
    out[3] = (int64_t)m.texcoords.size();
}
void ref_obj_copy(void* h, int i, int32_t* idx, float* verts, float* normals, float* texcoords) {
    mesh const& m = *std::get<1>((*(shape_list*)h)[i]);
    static_assert(sizeof(ivec3) == 12 && sizeof(vec3) == 12 && sizeof(vec2) == 8, "owl vec layouts");
    if (idx && !m.indices.empty()) memcpy(idx, m.indices.data(), sizeof(ivec3) * m.indices.size());
    if (verts && !m.vertices.empty()) memcpy(verts, m.vertices.data(), sizeof(vec3) * m.vertices.size());
    if (normals && !m.normals.empty()) memcpy(normals, m.normals.data(), sizeof(vec3) * m.normals.size());
    if (texcoords && !m.texcoords.empty()) memcpy(texcoords, m.texcoords.data(), sizeof(vec2) * m.texcoords.size());
}

// ---- image_buffer.cpp:36-59 (v-flipped RGBA8; *.hdr goes through stb's LDR conversion)
void* ref_image_load(const char* name, const char* path, int32_t* w, int32_t* h) {
    image_buffer img = load_image(name, path);
    auto* r = new RefImage{img.width, img.height, const_cast<uint32_t*>(img.buffer)};
    *w = r->w;
    *h = r->h;
    return r;
}
const uint32_t* ref_image_pixels(void* h) { return ((RefImage*)h)->pixels; }
void ref_image_free(void* h) {
    auto* r = (RefImage*)h;
    if (r->pixels) stbi_image_free(r->pixels);
    delete r;
}
// image_buffer.cpp:25-34
int ref_image_write(int32_t w, int32_t h, const uint32_t* pixels, const char* name, const char* path) {
    try {
        image_buffer img{w, h, pixels, image_buffer::tag::referenced};
        write_image(img, name, path);
        return 0;
    } catch (std::exception const& e) {
        g_err = e.what();
        return -1;
    }
}

// ---- parser.cpp:19-63
void* ref_parse_materials(const char* config_path) {
    try {
        return new material_list(parse_materials(config_path));
    } catch (std::exception const& e) {
        g_err = e.what();
        return nullptr;
    }
}
void ref_materials_free(void* h) { delete (material_list*)h; }
int ref_materials_count(void* h) { return (int)((material_list*)h)->size(); }
const char* ref_material_name(void* h, int i) { return std::get<0>((*(material_list*)h)[i]).c_str(); }
const char* ref_material_file(void* h, int i) { return std::get<2>((*(material_list*)h)[i]).c_str(); }
void ref_material_data(void* h, int i, float* out17) {
    static_assert(sizeof(material_data) == 68, "17 floats");
    memcpy(out17, &std::get<1>((*(material_list*)h)[i]), sizeof(material_data));
}

// ---- parser.cpp:65-79 + camera.cpp:3-22
int ref_parse_camera(const char* config_path, int w, int h, float* out12) {
    try {
        camera_data d = parse_camera(config_path, ivec2{w, h});
        const vec3* v[4] = {&d.origin, &d.llc, &d.horizontal, &d.vertical};
        for (int i = 0; i < 4; i++) { out12[3 * i] = v[i]->x; out12[3 * i + 1] = v[i]->y; out12[3 * i + 2] = v[i]->z; }
        return 0;
    } catch (std::exception const& e) {
        g_err = e.what();
        return -1;
    }
}

// ---- parser.cpp:81-117
struct ref_settings {
    char scene[256], test_name[256], test_material[256], test_attribute[256];
    int32_t buffer_size[2], max_samples, max_path_depth, environment_use, environment_auto, test_material_type;
    float environment_color[3], environment_intensity, test_step_size;
    int32_t n_vec_values, n_flt_values;
    float vec_values[64][3], flt_values[64];
};
int ref_parse_settings(const char* config_path, ref_settings* o) {
    try {
        settings_data s = parse_settings(config_path);
        memset(o, 0, sizeof *o);
        snprintf(o->scene, 256, "%s", s.scene.c_str());
        snprintf(o->test_name, 256, "%s", s.test.name.c_str());
        snprintf(o->test_material, 256, "%s", s.test.material_name.c_str());
        snprintf(o->test_attribute, 256, "%s", s.test.attribute_name.c_str());
        o->buffer_size[0] = s.buffer_size.x;
        o->buffer_size[1] = s.buffer_size.y;
        o->max_samples = s.max_samples;
        o->max_path_depth = s.max_path_depth;
        o->environment_use = s.environment_use;
        o->environment_auto = s.environment_auto;
        o->test_material_type = (int32_t)s.test.material_type;
        o->environment_color[0] = s.environment_color.x;
        o->environment_color[1] = s.environment_color.y;
        o->environment_color[2] = s.environment_color.z;
        o->environment_intensity = s.environment_intensity;
        o->test_step_size = s.test.step_size;
        o->n_vec_values = (int32_t)std::min<size_t>(64, s.test.vec_values.size());
        o->n_flt_values = (int32_t)std::min<size_t>(64, s.test.flt_values.size());
        for (int i = 0; i < o->n_vec_values; i++) {
            o->vec_values[i][0] = s.test.vec_values[i].x;
            o->vec_values[i][1] = s.test.vec_values[i].y;
            o->vec_values[i][2] = s.test.vec_values[i].z;
        }
        for (int i = 0; i < o->n_flt_values; i++) o->flt_values[i] = s.test.flt_values[i];
        return 0;
    } catch (std::exception const& e) {
        g_err = e.what();
        return -1;
    }
}

}  // extern "C"
