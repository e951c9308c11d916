// scene.hpp — host-side data of the application, mirroring the reference's structs:
// camera / camera_data (path_tracer/src/camera.hpp), test_data / settings_data
// (utils/parser.hpp), mesh (utils/mesh_loader.hpp), image_buffer (utils/image_buffer.hpp),
// entity / program_data (application.hpp).  Containers are plain std::vector<float> so
// that they can be handed to the C ABI (include/b200pt.h) without conversion.
#pragma once
#include <cstdint>
#include <memory>
#include <string>
#include <tuple>
#include <vector>

#include "../../include/b200pt.h"

namespace host {

struct vec3 {
    float x = 0, y = 0, z = 0;
};

struct camera {
    vec3 look_from, look_at, look_up;
    float vertical_fov = 0;
};
struct camera_data {
    vec3 origin, llc, horizontal, vertical;
};

using material_data = b200pt_material;
material_data default_material();  // member initialisers of device_global.hpp:19-36

struct test_data {
    std::string name, material_name, attribute_name;
    int material_type = 0;
    std::vector<vec3> vec_values;
    std::vector<float> flt_values;
    float step_size = 0;
};

struct settings_data {
    std::string scene;
    int buffer_size[2] = {0, 0};
    int max_samples = 0, max_path_depth = 0;
    bool environment_use = false, environment_auto = false;
    vec3 environment_color;
    float environment_intensity = 0;
    test_data test;
};

struct mesh {
    std::vector<int32_t> indices;   // 3 per triangle, local to this mesh
    std::vector<float> vertices;    // 3 per vertex
    std::vector<float> normals;     // 3 per vertex
    std::vector<float> texcoords;   // 2 per vertex (may be empty)
};

struct image_buffer {
    int width = 0, height = 0;
    std::vector<uint32_t> pixels;   // RGBA8, little endian R,G,B,A
    bool empty() const { return pixels.empty(); }
};

struct material_entry {
    std::string name;
    material_data data;
    std::string texture_file;  // "<name>-textures/<filename>" or empty (parser.cpp:32-35)
};

struct entity {
    int mesh_index = -1;     // into program_data::meshes
    int material_id = -1;
    image_buffer texture;    // loaded (and v-flipped) texture or empty
};

struct program_data {
    std::string scene, assets_path;
    int buffer_size[2] = {0, 0};
    int max_samples = 0, max_path_depth = 0;
    bool environment_use = false, environment_auto = false;
    vec3 environment_color;
    float environment_intensity = 0;
    image_buffer environment_map;
    std::vector<float> environment_float;  // linear RGB of the same map (load_options.env_float), same size and row order
    camera_data cam;
    std::vector<material_entry> materials;
    std::vector<std::pair<std::string, std::shared_ptr<mesh>>> meshes;
    std::vector<entity> entities;
    std::vector<std::string> notes;  // which inputs were procedural stand-ins
    bool write_pfm = false;          // also write the float frame (mean radiance before quantisation) as .pfm
    bool write_exr = false;          // ... and / or as an uncompressed OpenEXR
    // progressive frames (b200pt_render_progressive): passes of this many samples (0: one blocking
    // launch, the reference's render_frame), stop early once the relative-variance figure reaches the target
    int progressive_pass = 0;
    double target_error = 0.0;
    bool save_passes = false;        // write "<stem>_pass<samples>.png" after every pass
    // several GPUs from one process (path_tracer --gpus N): the contexts of ranks 0 .. N-1, samples sharded over them,
    // one NCCL reduce per frame inside b200pt_render_frame_all.  Empty: the single context handed to test_loop.
    std::vector<b200pt_ctx*> ranks;
    // frames of the sweep back to back (path_tracer --pipeline N): further contexts of this rank bound to the same scene;
    // test_loop then queues the frames on a b200pt_pipeline (frame k+1 starts while frame k drains) and writes the images
    // when all are done.  Empty: one blocking frame after the other, like the reference.
    std::vector<b200pt_ctx*> pipeline_slots;
};

}  // namespace host
