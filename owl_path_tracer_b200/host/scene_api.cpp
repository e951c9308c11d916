// scene_api.cpp — C entry points of libb200pt_scene.so (declared in include/b200pt_host.h): scene loading, accessors,
// the loaders one stage at a time, image / OBJ writers.  No dependency on the CUDA library.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <memory>

#include "loaders.hpp"
#include "procedural.hpp"
#include "scene_handle.hpp"

namespace host {
std::string& api_error() {
    static thread_local std::string e;
    return e;
}
}  // namespace host

extern "C" {

const char* b200pt_host_last_error(void) { return host::api_error().c_str(); }

int b200pt_scene_load(const char* assets_path, const b200pt_load_options* o, b200pt_scene** out) {
    return host::api_guard([&] {
        host::load_options opt;
        if (o) {
            if (o->scene) opt.scene_override = o->scene;
            opt.width = o->width;
            opt.height = o->height;
            opt.max_samples = o->max_samples;
            opt.max_path_depth = o->max_path_depth;
            opt.env_mode = o->env_mode;
            opt.env_intensity = o->env_intensity;
            opt.detail = o->detail;
            opt.allow_procedural = o->no_procedural == 0;
            opt.env_float = o->env_float != 0;
        }
        auto* s = new b200pt_scene();
        try {
            host::init_program_data(s->pdata, s->tdata, assets_path, opt);
        } catch (...) {
            delete s;
            throw;
        }
        *out = s;
    });
}

void b200pt_scene_free(b200pt_scene* s) { delete s; }

int b200pt_scene_info(const b200pt_scene* s, b200pt_scene_info_t* info) {
    const host::program_data& p = s->pdata;
    memset(info, 0, sizeof *info);
    info->width = p.buffer_size[0];
    info->height = p.buffer_size[1];
    info->max_samples = p.max_samples;
    info->max_path_depth = p.max_path_depth;
    info->environment_use = p.environment_use;
    info->environment_auto = p.environment_auto;
    info->environment_color[0] = p.environment_color.x;
    info->environment_color[1] = p.environment_color.y;
    info->environment_color[2] = p.environment_color.z;
    info->environment_intensity = p.environment_intensity;
    info->env_w = p.environment_map.width;
    info->env_h = p.environment_map.height;
    info->env_rgba = p.environment_map.empty() ? nullptr : (const uint8_t*)p.environment_map.pixels.data();
    info->env_f32 = p.environment_float.empty() ? nullptr : p.environment_float.data();
    memcpy(info->cam_origin, &p.cam.origin.x, 12);
    memcpy(info->cam_llc, &p.cam.llc.x, 12);
    memcpy(info->cam_horizontal, &p.cam.horizontal.x, 12);
    memcpy(info->cam_vertical, &p.cam.vertical.x, 12);
    info->n_entities = (int32_t)p.entities.size();
    info->n_materials = (int32_t)p.materials.size();
    info->n_notes = (int32_t)p.notes.size();
    int64_t tris = 0;
    for (const host::entity& e : p.entities) tris += (int64_t)p.meshes[e.mesh_index].second->indices.size() / 3;
    info->n_triangles = tris;
    return 0;
}

int b200pt_scene_entity(const b200pt_scene* s, int32_t i, b200pt_entity_view* v) {
    const host::program_data& p = s->pdata;
    if (i < 0 || i >= (int32_t)p.entities.size()) return 1;
    const host::entity& e = p.entities[i];
    const host::mesh& m = *p.meshes[e.mesh_index].second;
    memset(v, 0, sizeof *v);
    v->name = p.meshes[e.mesh_index].first.c_str();
    v->verts = m.vertices.data();
    v->normals = m.normals.size() == m.vertices.size() ? m.normals.data() : nullptr;
    v->nv = (int32_t)(m.vertices.size() / 3);
    v->uvs = (m.texcoords.size() / 2 == (size_t)v->nv && v->nv > 0) ? m.texcoords.data() : nullptr;
    v->idx = m.indices.data();
    v->nt = (int32_t)(m.indices.size() / 3);
    v->material_index = e.material_id;
    v->tex_rgba = e.texture.empty() ? nullptr : (const uint8_t*)e.texture.pixels.data();
    v->tex_w = e.texture.width;
    v->tex_h = e.texture.height;
    return 0;
}

const b200pt_material* b200pt_scene_materials(b200pt_scene* s, int32_t* n) {
    s->mats = host::material_array(s->pdata);
    if (n) *n = (int32_t)s->mats.size();
    return s->mats.data();
}
const char* b200pt_scene_material_name(const b200pt_scene* s, int32_t i) {
    if (i < 0 || i >= (int32_t)s->pdata.materials.size()) return nullptr;
    return s->pdata.materials[i].name.c_str();
}
const char* b200pt_scene_note(const b200pt_scene* s, int32_t i) {
    if (i < 0 || i >= (int32_t)s->pdata.notes.size()) return nullptr;
    return s->pdata.notes[i].c_str();
}

int b200pt_scene_set_material_attr(b200pt_scene* s, const char* material, const char* attr, float value) {
    host::test_data t;
    t.material_name = material;
    t.attribute_name = attr;
    return host::modify_material(s->pdata, t, value) ? 0 : 1;
}
int b200pt_scene_set_material_color(b200pt_scene* s, const char* material, const float rgb[3]) {
    host::test_data t;
    t.material_name = material;
    return host::modify_material(s->pdata, t, host::vec3{rgb[0], rgb[1], rgb[2]}) ? 0 : 1;
}

int b200pt_write_png(const char* path, int32_t w, int32_t h, const uint32_t* rgba) {
    return host::api_guard([&] { host::write_png(path, w, h, rgba); });
}
int b200pt_write_exr(const char* path, int32_t w, int32_t h, const float* rgba_top_first) {
    return host::api_guard([&] { host::write_exr(path, w, h, rgba_top_first); });
}
int b200pt_write_pfm(const char* path, int32_t w, int32_t h, const float* rgba_top_first) {
    return host::api_guard([&] { host::write_pfm(path, w, h, rgba_top_first); });
}
int b200pt_write_obj(const b200pt_scene* s, const char* path) {
    return host::api_guard([&] {
        FILE* f = fopen(path, "w");
        if (!f) throw std::runtime_error(std::string("cannot write ") + path);
        size_t voff = 1;
        for (const host::entity& e : s->pdata.entities) {
            const host::mesh& m = *s->pdata.meshes[e.mesh_index].second;
            fprintf(f, "o %s\n", s->pdata.meshes[e.mesh_index].first.c_str());
            const size_t nv = m.vertices.size() / 3;
            for (size_t i = 0; i < nv; ++i) fprintf(f, "v %.9g %.9g %.9g\n", m.vertices[3 * i], m.vertices[3 * i + 1], m.vertices[3 * i + 2]);
            for (size_t i = 0; i < nv; ++i) fprintf(f, "vn %.9g %.9g %.9g\n", m.normals[3 * i], m.normals[3 * i + 1], m.normals[3 * i + 2]);
            for (size_t t = 0; t < m.indices.size() / 3; ++t) {
                const size_t a = voff + m.indices[3 * t], b = voff + m.indices[3 * t + 1], c = voff + m.indices[3 * t + 2];
                fprintf(f, "f %zu//%zu %zu//%zu %zu//%zu\n", a, a, b, b, c, c);
            }
            voff += nv;
        }
        fclose(f);
    });
}

// ---- single stages (differential tests against the reference's own loaders)
}  // extern "C"
struct b200pt_materials {
    std::vector<host::material_entry> entries;
};
extern "C" {

int b200pt_obj_load(const char* obj_file, int32_t threads, b200pt_shapes** out) {
    return host::api_guard([&] {
        auto* s = new b200pt_shapes();
        try {
            s->shapes = host::load_obj(obj_file, threads);
        } catch (...) {
            delete s;
            throw;
        }
        *out = s;
    });
}
int32_t b200pt_obj_count(const b200pt_shapes* s) { return (int32_t)s->shapes.size(); }
int b200pt_obj_shape(const b200pt_shapes* s, int32_t i, b200pt_entity_view* v) {
    return host::api_guard([&] {
        if (i < 0 || (size_t)i >= s->shapes.size()) throw std::out_of_range("shape index");
        const host::mesh& m = *s->shapes[(size_t)i].second;
        memset(v, 0, sizeof *v);
        v->name = s->shapes[(size_t)i].first.c_str();
        v->verts = m.vertices.data();
        v->normals = m.normals.empty() ? nullptr : m.normals.data();
        v->uvs = m.texcoords.empty() ? nullptr : m.texcoords.data();
        v->nv = (int32_t)(m.vertices.size() / 3);
        v->idx = m.indices.data();
        v->nt = (int32_t)(m.indices.size() / 3);
        v->material_index = -1;
        // normals / uvs may be shorter than nv (first-occurrence rule, mesh_loader.cpp:52-75): their own counts
        v->tex_w = (int32_t)(m.normals.size() / 3);
        v->tex_h = (int32_t)(m.texcoords.size() / 2);
    });
}
void b200pt_obj_free(b200pt_shapes* s) { delete s; }

int b200pt_image_load(const char* name, const char* path, int32_t* w, int32_t* h, uint32_t** pixels) {
    return host::api_guard([&] {
        host::image_buffer img = host::load_image(name, path);
        *w = img.width;
        *h = img.height;
        *pixels = nullptr;
        if (!img.pixels.empty()) {
            *pixels = (uint32_t*)malloc(img.pixels.size() * sizeof(uint32_t));
            memcpy(*pixels, img.pixels.data(), img.pixels.size() * sizeof(uint32_t));
        }
    });
}
void b200pt_image_free(uint32_t* pixels) { free(pixels); }

int b200pt_parse_camera(const char* config_path, int32_t width, int32_t height, float out12[12]) {
    return host::api_guard([&] {
        const host::camera_data d = host::to_camera_data(host::parse_camera(config_path), width, height);
        const host::vec3* v[4] = {&d.origin, &d.llc, &d.horizontal, &d.vertical};
        for (int i = 0; i < 4; ++i) {
            out12[3 * i] = v[i]->x;
            out12[3 * i + 1] = v[i]->y;
            out12[3 * i + 2] = v[i]->z;
        }
    });
}

int b200pt_parse_settings(const char* config_path, b200pt_settings_view* o) {
    return host::api_guard([&] {
        const host::settings_data s = host::parse_settings(config_path);
        memset(o, 0, sizeof *o);
        snprintf(o->scene, sizeof o->scene, "%s", s.scene.c_str());
        snprintf(o->test_name, sizeof o->test_name, "%s", s.test.name.c_str());
        snprintf(o->test_material, sizeof o->test_material, "%s", s.test.material_name.c_str());
        snprintf(o->test_attribute, sizeof o->test_attribute, "%s", s.test.attribute_name.c_str());
        o->buffer_size[0] = s.buffer_size[0];
        o->buffer_size[1] = s.buffer_size[1];
        o->max_samples = s.max_samples;
        o->max_path_depth = s.max_path_depth;
        o->environment_use = s.environment_use;
        o->environment_auto = s.environment_auto;
        o->test_material_type = s.test.material_type;
        o->environment_color[0] = s.environment_color.x;
        o->environment_color[1] = s.environment_color.y;
        o->environment_color[2] = s.environment_color.z;
        o->environment_intensity = s.environment_intensity;
        o->test_step_size = s.test.step_size;
        o->n_vec_values = (int32_t)std::min<size_t>(64, s.test.vec_values.size());
        o->n_flt_values = (int32_t)std::min<size_t>(64, s.test.flt_values.size());
        for (int i = 0; i < o->n_vec_values; ++i) {
            o->vec_values[i][0] = s.test.vec_values[(size_t)i].x;
            o->vec_values[i][1] = s.test.vec_values[(size_t)i].y;
            o->vec_values[i][2] = s.test.vec_values[(size_t)i].z;
        }
        for (int i = 0; i < o->n_flt_values; ++i) o->flt_values[i] = s.test.flt_values[(size_t)i];
    });
}

int b200pt_parse_materials(const char* config_path, b200pt_materials** out) {
    return host::api_guard([&] {
        auto* m = new b200pt_materials();
        try {
            m->entries = host::parse_materials(config_path);
        } catch (...) {
            delete m;
            throw;
        }
        *out = m;
    });
}
int32_t b200pt_materials_count(const b200pt_materials* m) { return (int32_t)m->entries.size(); }
const char* b200pt_materials_name(const b200pt_materials* m, int32_t i) { return m->entries[(size_t)i].name.c_str(); }
const char* b200pt_materials_file(const b200pt_materials* m, int32_t i) { return m->entries[(size_t)i].texture_file.c_str(); }
const b200pt_material* b200pt_materials_data(const b200pt_materials* m, int32_t i) { return &m->entries[(size_t)i].data; }
void b200pt_materials_free(b200pt_materials* m) { delete m; }

}  // extern "C"
