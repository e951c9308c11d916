/* b200pt_host.h — C entry points of the host application layer (libb200pt_host.so).
 *
 * The host side of the reference is compiled C++ (path_tracer/Main.cpp,
 * path_tracer/src/application.cpp, path_tracer/src/utils/...).  Ours is too
 * (owl_path_tracer_b200/host/...); this header exposes its stages with plain C types so
 * that scripts and tests can drive exactly the code the path_tracer executable runs:
 *   b200pt_scene_load      = init_program_data   (application.cpp:143-181)
 *   b200pt_scene_bind      = bind_sbt_data       (application.cpp:184-294)
 *   b200pt_scene_test_loop = test_loop           (application.hpp:89-108)
 * The accessors hand out the loaded arrays (borrowed pointers, valid until
 * b200pt_scene_free) in the layout b200pt_add_mesh takes.
 */
#ifndef B200PT_HOST_H
#define B200PT_HOST_H

#include "b200pt.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200pt_scene b200pt_scene;

/* all zero / NULL = behave like the reference (everything from assets/settings.json) */
typedef struct b200pt_load_options {
    const char* scene;        /* override settings.json "scene"                         */
    int32_t width, height;    /* override buffer_size                                   */
    int32_t max_samples;      /* override max_samples                                   */
    int32_t max_path_depth;   /* override max_path_depth                                */
    int32_t env_mode;         /* -1 settings.json; 0 environment_color; 1 auto sky;
                                 2 environment.hdr                                      */
    float env_intensity;      /* < 0: settings.json                                     */
    int32_t detail;           /* procedural stand-in detail (0 = default)               */
    int32_t no_procedural;    /* 1: fail like the reference when an asset is missing    */
    int32_t env_float;        /* 1: also keep environment.hdr as float radiance and bind it
                                 through b200pt_set_environment_f32                     */
} b200pt_load_options;

typedef struct b200pt_scene_info_t {
    int32_t width, height, max_samples, max_path_depth;
    int32_t environment_use, environment_auto;
    float environment_color[3];
    float environment_intensity;
    int32_t env_w, env_h;
    const uint8_t* env_rgba;  /* RGBA8, v-flipped like image_buffer.cpp:50-55, or NULL  */
    float cam_origin[3], cam_llc[3], cam_horizontal[3], cam_vertical[3];
    int32_t n_entities, n_materials, n_notes;
    int64_t n_triangles;
    const float* env_f32;     /* env_w*env_h*3 linear RGB (load option env_float) or NULL */
} b200pt_scene_info_t;

typedef struct b200pt_entity_view {
    const char* name;
    const float* verts;       /* nv*3 */
    const float* normals;     /* nv*3 or NULL (the reference would trap on the device)  */
    const float* uvs;         /* nv*2 or NULL                                            */
    int32_t nv;
    const int32_t* idx;       /* nt*3 */
    int32_t nt;
    int32_t material_index;
    const uint8_t* tex_rgba;  /* RGBA8 v-flipped, or NULL                                */
    int32_t tex_w, tex_h;
} b200pt_entity_view;

const char* b200pt_host_last_error(void);

int b200pt_scene_load(const char* assets_path, const b200pt_load_options* opt, b200pt_scene** out);
void b200pt_scene_free(b200pt_scene* s);
int b200pt_scene_info(const b200pt_scene* s, b200pt_scene_info_t* info);
int b200pt_scene_entity(const b200pt_scene* s, int32_t i, b200pt_entity_view* view);
const b200pt_material* b200pt_scene_materials(b200pt_scene* s, int32_t* n);
const char* b200pt_scene_material_name(const b200pt_scene* s, int32_t i);
const char* b200pt_scene_note(const b200pt_scene* s, int32_t i);

/* modify_sbt (application.cpp:320-360), host side; follow with b200pt_set_materials */
int b200pt_scene_set_material_attr(b200pt_scene* s, const char* material, const char* attr, float value);
int b200pt_scene_set_material_color(b200pt_scene* s, const char* material, const float rgb[3]);

int b200pt_scene_bind(const b200pt_scene* s, b200pt_ctx* ctx);
int b200pt_scene_test_loop(b200pt_scene* s, b200pt_ctx* ctx, const char* out_dir);

int b200pt_write_png(const char* path, int32_t w, int32_t h, const uint32_t* rgba);
/* float frame (b200pt_framebuffer_f32: RGBA, top row first) as a Portable Float Map, RGB, little endian */
int b200pt_write_pfm(const char* path, int32_t w, int32_t h, const float* rgba_top_first);
/* the same frame as OpenEXR: one part, scan lines top row first, uncompressed, FLOAT channels B G R */
int b200pt_write_exr(const char* path, int32_t w, int32_t h, const float* rgba_top_first);
int b200pt_write_obj(const b200pt_scene* s, const char* path);

/* ---- the loaders one stage at a time.  The application above is built from exactly these functions; they are exported
 * so that tests/test_ref_host.py can hold each of them to the reference's own loader (oracle/_ref/libref_host.so =
 * utils/mesh_loader.cpp on tinyobjloader, utils/image_buffer.cpp on stb, utils/parser.cpp on nlohmann json). */
typedef struct b200pt_shapes b200pt_shapes;
/* load_obj (mesh_loader.cpp:85-121): every shape of the file, before the material match of application.cpp:166-179.
 * threads: 0 = automatic, n = cut the file into n chunks. */
int b200pt_obj_load(const char* obj_file, int32_t threads, b200pt_shapes** out);
/* the same with create_mesh's vertex de-duplication run on the device (b200pt_dedup_mesh; libb200pt_host.so):
 * identical arrays, byte for byte */
int b200pt_obj_load_gpu(b200pt_ctx* ctx, const char* obj_file, int32_t threads, b200pt_shapes** out);
int32_t b200pt_obj_count(const b200pt_shapes* s);
/* material_index = -1, no texture */
int b200pt_obj_shape(const b200pt_shapes* s, int32_t i, b200pt_entity_view* view);
void b200pt_obj_free(b200pt_shapes* s);
/* load_image (image_buffer.cpp:36-59): RGBA8, rows flipped; *.hdr through the 8-bit conversion of stb.  A missing file
 * gives w = h = 0 and *pixels = NULL with status 0 (the reference continues with an empty image); a file that exists but
 * cannot be decoded is an error.  Release with b200pt_image_free. */
int b200pt_image_load(const char* name, const char* path, int32_t* w, int32_t* h, uint32_t** pixels);
void b200pt_image_free(uint32_t* pixels);
/* parse_camera (parser.cpp:65-79) + to_camera_data (camera.cpp:3-22): origin, llc, horizontal, vertical */
int b200pt_parse_camera(const char* config_path, int32_t width, int32_t height, float out12[12]);
typedef struct b200pt_settings_view {   /* settings_data / test_data of utils/parser.hpp */
    char scene[256], test_name[256], test_material[256], test_attribute[256];
    int32_t buffer_size[2], max_samples, max_path_depth, environment_use, environment_auto, test_material_type;
    float environment_color[3], environment_intensity, test_step_size;
    int32_t n_vec_values, n_flt_values;
    float vec_values[64][3], flt_values[64];
} b200pt_settings_view;
int b200pt_parse_settings(const char* config_path, b200pt_settings_view* out);
typedef struct b200pt_materials b200pt_materials;
/* parse_materials (parser.cpp:19-63): name, 17 floats, "<name>-textures/<filename>" or "" */
int b200pt_parse_materials(const char* config_path, b200pt_materials** out);
int32_t b200pt_materials_count(const b200pt_materials* m);
const char* b200pt_materials_name(const b200pt_materials* m, int32_t i);
const char* b200pt_materials_file(const b200pt_materials* m, int32_t i);
const b200pt_material* b200pt_materials_data(const b200pt_materials* m, int32_t i);
void b200pt_materials_free(b200pt_materials* m);

#ifdef __cplusplus
}
#endif
#endif
