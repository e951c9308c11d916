// application.cpp — the device-facing stages of the host application (see application.hpp): bind_sbt_data, render_frame,
// test_loop.  The loading stages live in scene_load.cpp (libb200pt_scene.so, no CUDA dependency).
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

// create_mesh on the device (b200pt_dedup_mesh); irregular shapes go back to the host routine (nullptr)
create_mesh_fn gpu_create_mesh(b200pt_ctx* ctx) {
    return [ctx](const raw_shape& r) -> std::shared_ptr<mesh> {
        bool any_n = false, all_n = true, any_t = false, all_t = true;
        for (size_t i = 0; i < r.n_corners; ++i) {
            (r.corner_vn[i] >= 0 ? any_n : all_n) = r.corner_vn[i] >= 0;
            (r.corner_vt[i] >= 0 ? any_t : all_t) = r.corner_vt[i] >= 0;
        }
        if (r.n_corners == 0 || (any_n && !all_n) || (any_t && !all_t)) return nullptr;
        auto m = std::make_shared<mesh>();
        m->indices.resize(r.n_corners);
        m->vertices.resize(3 * r.n_corners);
        if (any_n) m->normals.resize(3 * r.n_corners);
        if (any_t) m->texcoords.resize(2 * r.n_corners);
        int32_t nv = 0;
        check(b200pt_dedup_mesh(ctx, r.attrib_vertices, (int32_t)r.n_vertices, r.attrib_normals, (int32_t)r.n_normals, r.attrib_texcoords,
                                (int32_t)r.n_texcoords, r.corner_v, any_n ? r.corner_vn : nullptr, any_t ? r.corner_vt : nullptr,
                                (int32_t)r.n_corners, m->vertices.data(), any_n ? m->normals.data() : nullptr,
                                any_t ? m->texcoords.data() : nullptr, m->indices.data(), &nv),
              "dedup_mesh");
        m->vertices.resize(3 * (size_t)nv);
        if (any_n) m->normals.resize(3 * (size_t)nv);
        if (any_t) m->texcoords.resize(2 * (size_t)nv);
        return m;
    };
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

// "<scene>_<test>_<attribute>(<values>)", the file stem of application.cpp:364
static std::string frame_stem(const program_data& pdata, const test_data& tdata, const std::string& values) {
    return pdata.scene + "_" + tdata.name + "_" + tdata.attribute_name + "(" + values + ")";
}

// render_frame, application.cpp:363-371
void render_frame(b200pt_ctx* ctx, const program_data& pdata, const test_data& tdata, const std::string& values,
                  const std::string& out_dir) {
    host::say("TRACING\n");
    const std::string stem = frame_stem(pdata, tdata, values);
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
    } else if (pdata.ranks.size() > 1) {
        check(b200pt_render_frame_all(pdata.ranks.data(), (int32_t)pdata.ranks.size()), "render_frame_all");
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
    if (pdata.write_pfm || pdata.write_exr) {  // float output next to the PNG (the reference only has the 8-bit image)
        const float* mean = nullptr;
        check(b200pt_framebuffer_f32(ctx, &mean), "framebuffer_f32");
        if (pdata.write_pfm) write_pfm(out_dir + "/" + stem + ".pfm", result.width, result.height, mean);
        if (pdata.write_exr) write_exr(out_dir + "/" + stem + ".exr", result.width, result.height, mean);
    }
}

// test_loop, application.hpp:89-108
void test_loop(b200pt_ctx* ctx, program_data& pdata, const test_data& tdata, const std::string& out_dir) {
    int vstep = static_cast<int32_t>(tdata.step_size * 100);
    if (vstep <= 0) vstep = 101;  // the reference would never terminate; render one frame
    const bool is_vec = !tdata.vec_values.empty();
    if ((is_vec && tdata.vec_values.size() < 2) || (!is_vec && tdata.flt_values.size() < 2))
        throw std::runtime_error("settings.test.values needs two entries");
    // --pipeline: the frames are queued, rendered overlapped, and written afterwards (same files, same bytes)
    const bool pipelined = !pdata.pipeline_slots.empty() && pdata.ranks.size() <= 1 && pdata.progressive_pass == 0;
    b200pt_pipeline* pipe = nullptr;
    std::vector<std::vector<uint32_t>> frames;
    std::vector<std::string> labels;
    if (pipelined) {
        std::vector<b200pt_ctx*> slots{ctx};
        slots.insert(slots.end(), pdata.pipeline_slots.begin(), pdata.pipeline_slots.end());
        check(b200pt_pipeline_create(slots.data(), (int32_t)slots.size(), 0.0f, &pipe), "pipeline_create");
        frames.reserve(128);   // submit() keeps pointers into the vectors: no reallocation of the outer one below
    }
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
        for (size_t r = 1; r < pdata.ranks.size(); ++r)
            check(b200pt_set_materials(pdata.ranks[r], mats.data(), (int)mats.size()), "set_materials");
        if (pipelined) {
            frames.emplace_back((size_t)pdata.buffer_size[0] * pdata.buffer_size[1]);
            labels.emplace_back(buf);
            check(b200pt_pipeline_submit(pipe, mats.data(), (int32_t)mats.size(), frames.back().data()), "pipeline_submit");
            continue;
        }
        render_frame(ctx, pdata, tdata, buf, out_dir);
    }
    if (pipelined) {
        const int st = b200pt_pipeline_wait(pipe);
        const std::string err = st != B200PT_OK ? b200pt_last_error() : "";
        b200pt_pipeline_destroy(pipe);
        if (st != B200PT_OK) throw std::runtime_error("pipeline: " + err);
        for (size_t k = 0; k < frames.size(); ++k) {
            image_buffer result;
            result.width = pdata.buffer_size[0];
            result.height = pdata.buffer_size[1];
            result.pixels = std::move(frames[k]);
            write_image(result, frame_stem(pdata, tdata, labels[k]) + ".png", out_dir);
        }
    }
}

}  // namespace host
