// application.cpp — host application stages (see application.hpp).
#include "application.hpp"

#include <algorithm>
#include <cstdio>
#include <filesystem>
#include <stdexcept>

#include "loaders.hpp"
#include "procedural.hpp"

namespace host {

static void check(int status, const char* what) {
    if (status != B200PT_OK) throw std::runtime_error(std::string(what) + ": " + b200pt_last_error());
}

// init_program_data, application.cpp:143-181
void init_program_data(program_data& pdata, test_data& tdata, const std::string& assets_path, const load_options& opt) {
    const settings_data settings = parse_settings(assets_path + "/settings.json");
    const std::string scene = opt.scene_override.empty() ? settings.scene : opt.scene_override;
    const std::string config_file = assets_path + "/" + scene + ".json";

    tdata = settings.test;
    pdata.assets_path = assets_path;
    pdata.scene = scene;
    pdata.buffer_size[0] = opt.width > 0 ? opt.width : settings.buffer_size[0];
    pdata.buffer_size[1] = opt.height > 0 ? opt.height : settings.buffer_size[1];
    pdata.max_path_depth = opt.max_path_depth > 0 ? opt.max_path_depth : settings.max_path_depth;
    pdata.max_samples = opt.max_samples > 0 ? opt.max_samples : settings.max_samples;
    pdata.environment_use = settings.environment_use;
    pdata.environment_auto = settings.environment_auto;
    pdata.environment_color = settings.environment_color;
    pdata.environment_intensity = settings.environment_intensity;
    if (opt.env_mode >= 0) {
        pdata.environment_use = opt.env_mode == 2;
        pdata.environment_auto = opt.env_mode == 1;
    }
    if (opt.env_intensity >= 0.0f) pdata.environment_intensity = opt.env_intensity;

    pdata.environment_map = load_image("environment.hdr", assets_path);
    if (pdata.environment_map.empty() && pdata.environment_use && opt.allow_procedural) {
        // stand-in sky, pushed through the same RGBE -> 8 bit route a file would take
        const int w = 2048, h = 1024;
        std::vector<float> rgb;
        make_procedural_environment(w, h, rgb);
        std::vector<unsigned char> rgbe((size_t)w * h * 4);
        for (size_t i = 0; i < (size_t)w * h; ++i) rgbe_from_float(&rgb[3 * i], &rgbe[4 * i]);
        pdata.environment_map = ldr_from_rgbe(w, h, rgbe.data());
        flip_rows(pdata.environment_map);
        if (opt.env_float) float_from_rgbe(w, h, rgbe.data(), pdata.environment_float);
        pdata.notes.push_back("environment.hdr: procedural stand-in (2048x1024)");
    } else if (opt.env_float && !pdata.environment_map.empty()) {
        int w = 0, h = 0;
        if (!load_hdr_float("environment.hdr", assets_path, w, h, pdata.environment_float) || w != pdata.environment_map.width ||
            h != pdata.environment_map.height)
            pdata.environment_float.clear();
    }

    pdata.cam = to_camera_data(parse_camera(config_file), pdata.buffer_size[0], pdata.buffer_size[1]);
    pdata.materials = parse_materials(config_file);

    const std::string obj_file = assets_path + "/" + scene + ".obj.scene";
    if (std::filesystem::exists(obj_file)) {
        pdata.meshes = load_obj(obj_file);
    } else if (opt.allow_procedural) {
        pdata.meshes = make_procedural_scene(scene, opt.detail);
        if (pdata.meshes.empty()) throw std::runtime_error("Cannot open file [" + obj_file + "]");
        pdata.notes.push_back(scene + ".obj.scene: procedural stand-in");
    } else {
        throw std::runtime_error("Cannot open file [" + obj_file + "]");
    }

    // application.cpp:166-179: a shape is bound to the first material of the same name;
    // shapes without one are dropped
    pdata.entities.clear();
    for (size_t mi = 0; mi < pdata.meshes.size(); ++mi) {
        int position = 0;
        for (const material_entry& me : pdata.materials) {
            if (me.name == pdata.meshes[mi].first) {
                entity e;
                e.mesh_index = (int)mi;
                e.material_id = position;
                pdata.entities.push_back(std::move(e));
                break;
            }
            ++position;
        }
    }

    // textures (loaded in bind_sbt_data by the reference, application.cpp:214-246)
    for (entity& e : pdata.entities) {
        const std::string& file = pdata.materials[e.material_id].texture_file;
        if (file.empty()) continue;
        const mesh& m = *pdata.meshes[e.mesh_index].second;
        if (m.texcoords.size() / 2 != m.vertices.size() / 3) {
            host::say("Mesh %s has no texture coordinates; texture %s ignored.\n", pdata.meshes[e.mesh_index].first.c_str(),
                   file.c_str());
            continue;
        }
        e.texture = load_image(file, assets_path);  // v-flipped, like application.cpp:233-238
        if (e.texture.empty() && opt.allow_procedural) {
            e.texture = make_procedural_texture(file, 512);
            flip_rows(e.texture);
            pdata.notes.push_back(file + ": procedural stand-in (512x512)");
        }
    }
}

std::vector<material_data> material_array(const program_data& pdata) {
    std::vector<material_data> v;
    v.reserve(pdata.materials.size());
    for (const material_entry& m : pdata.materials) v.push_back(m.data);
    return v;
}

// bind_sbt_data, application.cpp:184-294
void bind_scene(const program_data& pdata, b200pt_ctx* ctx) {
    for (const entity& e : pdata.entities) {
        const mesh& m = *pdata.meshes[e.mesh_index].second;
        const int nv = (int)(m.vertices.size() / 3);
        if (m.normals.size() != m.vertices.size())
            throw std::runtime_error("mesh '" + pdata.meshes[e.mesh_index].first + "' has no per-vertex normals");
        const bool has_uv = m.texcoords.size() / 2 == (size_t)nv && nv > 0;
        int mesh_id = -1;
        check(b200pt_add_mesh(ctx, m.vertices.data(), m.normals.data(), has_uv ? m.texcoords.data() : nullptr, nv,
                              m.indices.data(), (int)(m.indices.size() / 3), e.material_id,
                              e.texture.empty() ? nullptr : (const uint8_t*)e.texture.pixels.data(), e.texture.width,
                              e.texture.height, &mesh_id),
              "add_mesh");
    }
    check(b200pt_build_accel(ctx), "build_accel");  // init_owl_world
    const float color[3] = {pdata.environment_color.x, pdata.environment_color.y, pdata.environment_color.z};
    if (!pdata.environment_float.empty())  // float radiance instead of the 8-bit map
        check(b200pt_set_environment_f32(ctx, pdata.environment_float.data(), pdata.environment_map.width,
                                         pdata.environment_map.height, pdata.environment_use ? 1 : 0,
                                         pdata.environment_auto ? 1 : 0, color, pdata.environment_intensity),
              "set_environment_f32");
    else
        check(b200pt_set_environment(ctx,
                                     pdata.environment_map.empty() ? nullptr : (const uint8_t*)pdata.environment_map.pixels.data(),
                                     pdata.environment_map.width, pdata.environment_map.height, pdata.environment_use ? 1 : 0,
                                     pdata.environment_auto ? 1 : 0, color, pdata.environment_intensity),
              "set_environment");
    const std::vector<material_data> mats = material_array(pdata);
    check(b200pt_set_materials(ctx, mats.data(), (int)mats.size()), "set_materials");
    check(b200pt_set_camera(ctx, &pdata.cam.origin.x, &pdata.cam.llc.x, &pdata.cam.horizontal.x, &pdata.cam.vertical.x),
          "set_camera");
    check(b200pt_set_render(ctx, pdata.buffer_size[0], pdata.buffer_size[1], pdata.max_samples, pdata.max_path_depth),
          "set_render");
}

static material_data* find_material(program_data& pdata, const test_data& test) {  // application.cpp:307-317
    for (material_entry& m : pdata.materials)
        if (m.name == test.material_name) return &m.data;
    return nullptr;
}

bool modify_material(program_data& pdata, const test_data& test, vec3 value) {  // application.cpp:320-326
    material_data* m = find_material(pdata, test);
    if (!m) return false;
    m->base_color[0] = value.x;
    m->base_color[1] = value.y;
    m->base_color[2] = value.z;
    return true;
}

bool modify_material(program_data& pdata, const test_data& test, float value) {  // application.cpp:329-360
    material_data* m = find_material(pdata, test);
    if (!m) return false;
    const std::string& a = test.attribute_name;
    if (a == "subsurface") m->subsurface = value;
    else if (a == "metallic") m->metallic = value;
    else if (a == "specular") m->specular = value;
    else if (a == "specular_tint") m->specular_tint = value;
    else if (a == "roughness") m->roughness = value;
    else if (a == "anisotropic") m->anisotropic = value;
    else if (a == "sheen") m->sheen = value;
    else if (a == "sheen_tint") m->sheen_tint = value;
    else if (a == "clearcoat") m->clearcoat = value;
    else if (a == "clearcoat_gloss") m->clearcoat_gloss = value;
    else if (a == "ior") m->ior = value;
    else if (a == "specular_transmission") m->specular_transmission = value;
    else if (a == "specular_transmission_roughness") m->specular_transmission_roughness = value;
    return true;
}

// render_frame, application.cpp:363-371
void render_frame(b200pt_ctx* ctx, const program_data& pdata, const test_data& tdata, const std::string& values,
                  const std::string& out_dir) {
    host::say("TRACING\n");
    const std::string stem = pdata.scene + "_" + tdata.name + "_" + tdata.attribute_name + "(" + values + ")";
    if (pdata.progressive_pass > 0) {
        // the same frame in passes: every pass continues the pixels' RNG streams, so the last pass leaves the
        // image of the blocking launch (or, with a target, the reference's image at the sample count reached)
        struct pass_ctx {
            const program_data* pdata;
            const std::string* stem;
            const std::string* out_dir;
        } pc{&pdata, &stem, &out_dir};
        auto on_pass = [](void* user, const b200pt_pass_info* info, const uint32_t* rgba, const float*) -> int {
            const pass_ctx* pc = static_cast<const pass_ctx*>(user);
            char line[160];
            snprintf(line, sizeof line, "pass: %d spp, rel. variance %.3e, %.1f ms, %.0f Mrays/s\n", info->samples_done,
                     info->rel_variance, info->render_ms, info->render_ms > 0 ? info->rays / info->render_ms * 1e-3 : 0.0);
            host::say("%s", line);
            if (pc->pdata->save_passes) {
                image_buffer img;
                img.width = pc->pdata->buffer_size[0];
                img.height = pc->pdata->buffer_size[1];
                img.pixels.assign(rgba, rgba + (size_t)img.width * img.height);
                write_image(img, *pc->stem + "_pass" + std::to_string(info->samples_done) + ".png", *pc->out_dir);
            }
            return 0;
        };
        b200pt_pass_info last{};
        check(b200pt_render_progressive(ctx, pdata.progressive_pass, pdata.target_error, on_pass, &pc, &last), "render_progressive");
        if (last.stopped_early) host::say("target reached after %d samples per pixel\n", last.samples_done);
    } else {
        check(b200pt_render(ctx), "render");
    }
    const uint32_t* fb = nullptr;
    check(b200pt_framebuffer_rgba8(ctx, &fb), "framebuffer");
    image_buffer result;
    result.width = pdata.buffer_size[0];
    result.height = pdata.buffer_size[1];
    result.pixels.assign(fb, fb + (size_t)result.width * result.height);
    write_image(result, stem + ".png", out_dir);
    if (pdata.write_pfm) {  // float output next to the PNG (the reference only has the 8-bit image)
        const float* mean = nullptr;
        check(b200pt_framebuffer_f32(ctx, &mean), "framebuffer_f32");
        write_pfm(out_dir + "/" + stem + ".pfm", result.width, result.height, mean);
    }
}

// test_loop, application.hpp:89-108
void test_loop(b200pt_ctx* ctx, program_data& pdata, const test_data& tdata, const std::string& out_dir) {
    int vstep = static_cast<int32_t>(tdata.step_size * 100);
    if (vstep <= 0) vstep = 101;  // the reference would never terminate; render one frame
    const bool is_vec = !tdata.vec_values.empty();
    if ((is_vec && tdata.vec_values.size() < 2) || (!is_vec && tdata.flt_values.size() < 2))
        throw std::runtime_error("settings.test.values needs two entries");
    for (int32_t i = 0; i <= 100; i += vstep) {
        const float c = i / 100.0f;
        char buf[96];
        if (is_vec) {
            const vec3 a = tdata.vec_values[0], b = tdata.vec_values[1];
            const vec3 v{a.x + (b.x - a.x) * c, a.y + (b.y - a.y) * c, a.z + (b.z - a.z) * c};
            modify_material(pdata, tdata, v);
            snprintf(buf, sizeof buf, "%.1f,%.1f,%.1f", v.x, v.y, v.z);
        } else {
            const float a = tdata.flt_values[0], b = tdata.flt_values[1];
            const float v = a + (b - a) * c;
            modify_material(pdata, tdata, v);
            snprintf(buf, sizeof buf, "%.1f", v);
        }
        const std::vector<material_data> mats = material_array(pdata);  // reset_field, application.cpp:297-304
        check(b200pt_set_materials(ctx, mats.data(), (int)mats.size()), "set_materials");
        render_frame(ctx, pdata, tdata, buf, out_dir);
    }
}

}  // namespace host
