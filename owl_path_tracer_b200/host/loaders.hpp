// loaders.hpp — see loaders.cpp
#pragma once
#include <functional>

#include "scene.hpp"

namespace host {

// console log of the stage banners (types.hpp:20-28 colours dropped); silent when B200PT_QUIET is set
void say(const char* fmt, ...);

camera_data to_camera_data(const camera& c, int width, int height);            // camera.cpp:3-22
settings_data parse_settings(const std::string& path);                         // parser.cpp:81-117
camera parse_camera(const std::string& path);                                  // parser.cpp:65-79
std::vector<material_entry> parse_materials(const std::string& path);          // parser.cpp:19-63
// One shape as create_mesh (mesh_loader.cpp:9-83) receives it: the file's attribute arrays and, per triangle corner (faces
// with more corners are fan-split first), the indices into them (-1 = the corner has none).  Vertex indices are in range.
struct raw_shape {
    const float* attrib_vertices; size_t n_vertices;     // 3 floats each
    const float* attrib_normals; size_t n_normals;       // 3 floats each
    const float* attrib_texcoords; size_t n_texcoords;   // 2 floats each
    const int32_t* corner_v;
    const int32_t* corner_vn;
    const int32_t* corner_vt;
    size_t n_corners;                                    // 3 per triangle
};
// create_mesh replacement (the vertex de-duplication on the GPU, b200pt_dedup_mesh): returns the mesh, or
// nullptr to say "not for me" (the loader then runs its own create_mesh).
using create_mesh_fn = std::function<std::shared_ptr<mesh>(const raw_shape&)>;
// threads: 0 = by file size (B200PT_OBJ_THREADS overrides), n = n chunks
std::vector<std::pair<std::string, std::shared_ptr<mesh>>> load_obj(const std::string& obj_file, int threads = 0,
                                                                     const create_mesh_fn& create = nullptr);  // mesh_loader.cpp:85-121
image_buffer load_image(const std::string& name, const std::string& path);     // image_buffer.cpp:36-59 (v-flipped)
void write_image(const image_buffer& i, const std::string& name, const std::string& path);  // image_buffer.cpp:25-34
void write_png(const std::string& file_path, int w, int h, const uint32_t* rgba);
// float route of the environment (b200pt_set_environment_f32): linear RGB radiance, rows flipped like load_image
void float_from_rgbe(int w, int h, const unsigned char* rgbe, std::vector<float>& rgb_flipped);
bool load_hdr_float(const std::string& name, const std::string& path, int& w, int& h, std::vector<float>& rgb_flipped);
// Portable Float Map ("PF", little endian, rows bottom to top) of a top-row-first float RGBA frame
void write_pfm(const std::string& file_path, int w, int h, const float* rgba_top_first);
void write_exr(const std::string& file_path, int w, int h, const float* rgba_top_first);
image_buffer ldr_from_rgbe(int w, int h, const unsigned char* rgbe);          // stb hdr -> 8 bit route
void rgbe_from_float(const float* rgb, unsigned char* rgbe);
void flip_rows(image_buffer& img);                                             // image_buffer.cpp:50-55
void write_hdr(const std::string& path, int w, int h, const float* rgb);

}  // namespace host
