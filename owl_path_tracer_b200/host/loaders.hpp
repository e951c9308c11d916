// loaders.hpp — see loaders.cpp
#pragma once
#include "scene.hpp"

namespace host {

// console log of the stage banners (types.hpp:20-28 colours dropped); silent when B200PT_QUIET is set
void say(const char* fmt, ...);

camera_data to_camera_data(const camera& c, int width, int height);            // camera.cpp:3-22
settings_data parse_settings(const std::string& path);                         // parser.cpp:81-117
camera parse_camera(const std::string& path);                                  // parser.cpp:65-79
std::vector<material_entry> parse_materials(const std::string& path);          // parser.cpp:19-63
// threads: 0 = by file size (B200PT_OBJ_THREADS overrides), n = n chunks
std::vector<std::pair<std::string, std::shared_ptr<mesh>>> load_obj(const std::string& obj_file, int threads = 0);  // mesh_loader.cpp:85-121
image_buffer load_image(const std::string& name, const std::string& path);     // image_buffer.cpp:36-59 (v-flipped)
void write_image(const image_buffer& i, const std::string& name, const std::string& path);  // image_buffer.cpp:25-34
void write_png(const std::string& file_path, int w, int h, const uint32_t* rgba);
// float route of the environment (b200pt_set_environment_f32): linear RGB radiance, rows flipped like load_image
void float_from_rgbe(int w, int h, const unsigned char* rgbe, std::vector<float>& rgb_flipped);
bool load_hdr_float(const std::string& name, const std::string& path, int& w, int& h, std::vector<float>& rgb_flipped);
// Portable Float Map ("PF", little endian, rows bottom to top) of a top-row-first float RGBA frame
void write_pfm(const std::string& file_path, int w, int h, const float* rgba_top_first);
image_buffer ldr_from_rgbe(int w, int h, const unsigned char* rgbe);          // stb hdr -> 8 bit route
void rgbe_from_float(const float* rgb, unsigned char* rgbe);
void flip_rows(image_buffer& img);                                             // image_buffer.cpp:50-55
void write_hdr(const std::string& path, int w, int h, const float* rgb);

}  // namespace host
