// scene_load.cpp — the stages of the host application that never touch the device: init_program_data (settings, scene
// file, OBJ, textures, environment; application.cpp:143-181 of the reference), to_vector and the host half of modify_sbt.
// Built into libb200pt_scene.so, which does NOT link the CUDA library: a CPU-only process (bench.py --impl reference,
// the oracle tests) can load scenes without mapping libb200pt.so.
#include <algorithm>
#include <cstdio>
#include <filesystem>
#include <stdexcept>

#include "application.hpp"
#include "loaders.hpp"
#include "procedural.hpp"

namespace host {

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
        pdata.meshes = load_obj(obj_file, 0, opt.create_mesh);
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

}  // namespace host
