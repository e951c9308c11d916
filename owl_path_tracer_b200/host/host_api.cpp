// host_api.cpp — the two C entry points of libb200pt_host.so that drive the device (declared in include/b200pt_host.h);
// everything else of that header lives in libb200pt_scene.so (scene_api.cpp), which this library links.
#include "loaders.hpp"
#include "scene_handle.hpp"

extern "C" {

int b200pt_scene_bind(const b200pt_scene* s, b200pt_ctx* ctx) {
    return host::api_guard([&] { host::bind_scene(s->pdata, ctx); });
}
int b200pt_scene_test_loop(b200pt_scene* s, b200pt_ctx* ctx, const char* out_dir) {
    return host::api_guard([&] { host::test_loop(ctx, s->pdata, s->tdata, out_dir); });
}

int b200pt_obj_load_gpu(b200pt_ctx* ctx, const char* obj_file, int32_t threads, b200pt_shapes** out) {
    return host::api_guard([&] {
        auto* s = new b200pt_shapes();
        try {
            s->shapes = host::load_obj(obj_file, threads, host::gpu_create_mesh(ctx));
        } catch (...) {
            delete s;
            throw;
        }
        *out = s;
    });
}

}  // extern "C"
