// application.hpp — the host application: same stages as the reference's
// path_tracer/src/application.{hpp,cpp} + Main.cpp, talking to the device only through
// the C ABI of include/b200pt.h.
#pragma once
#include <functional>

#include "loaders.hpp"
#include "scene.hpp"

namespace host {

struct load_options {
    std::string scene_override;   // empty: use settings.json "scene"
    int width = 0, height = 0;    // 0: settings.json
    int max_samples = 0;          // 0: settings.json
    int max_path_depth = 0;       // 0: settings.json
    int env_mode = -1;            // -1: settings.json; 0 colour, 1 auto sky, 2 environment.hdr
    float env_intensity = -1.0f;  // <0: settings.json
    int detail = 0;               // procedural stand-in detail knob
    bool allow_procedural = true; // generate stand-ins for assets missing from the checkout
    bool env_float = false;       // keep environment.hdr as float radiance instead of the reference's 8-bit conversion
    create_mesh_fn create_mesh;   // optional replacement of the loader's create_mesh stage (gpu_create_mesh below)
};

// init_program_data, application.cpp:143-181
void init_program_data(program_data& pdata, test_data& tdata, const std::string& assets_path, const load_options& opt);

// create_mesh on the device (b200pt_dedup_mesh): plug into load_options::create_mesh.  Irregular shapes (some
// corners with a normal / texcoord index, some without) are handed back to the host's create_mesh.  (application.cpp, links CUDA)
create_mesh_fn gpu_create_mesh(b200pt_ctx* ctx);

// bind_sbt_data, application.cpp:184-294: meshes, accel, env, materials, camera, frame size
void bind_scene(const program_data& pdata, b200pt_ctx* ctx);

// modify_sbt, application.cpp:306-360 (host side only; returns false if the material is unknown)
bool modify_material(program_data& pdata, const test_data& test, float value);
bool modify_material(program_data& pdata, const test_data& test, vec3 value);
std::vector<material_data> material_array(const program_data& pdata);  // to_vector, application.cpp:14-27

// render_frame, application.cpp:363-371
void render_frame(b200pt_ctx* ctx, const program_data& pdata, const test_data& tdata, const std::string& values,
                  const std::string& out_dir);

// test_loop, application.hpp:89-108
void test_loop(b200pt_ctx* ctx, program_data& pdata, const test_data& tdata, const std::string& out_dir);

}  // namespace host
