// main.cpp — the path_tracer executable (reference: path_tracer/Main.cpp:13-31).
//
// Without arguments it behaves like the reference: reads ./assets/settings.json, then
// ./assets/<scene>.json + <scene>.obj.scene + environment.hdr, runs the settings' test
// sweep and writes "<scene>_<test>_<attr>(<value>).png" files into the current
// directory.  Optional flags (all default to the reference behaviour):
//   --assets DIR  --scene NAME  --size W H  --spp N  --depth N  --env color|auto|map
//   --env-intensity F  --detail N  --device K  --out DIR  --stats
//   --pipeline N (N contexts on the device: the frames of the sweep are queued and frame k+1 starts while frame k drains;
//                 same files, same bytes)
//   --gpu-dedup (create_mesh's vertex de-duplication on the device, b200pt_dedup_mesh; same meshes byte for byte)
//   --gpus N (devices K .. K+N-1, samples sharded over them, one NCCL reduce per frame; the reference's OWL context can
//             span devices the same way, extern/owl/owl/RayGen.cpp:140-146, and asks for one, application.cpp:62)
//   --nee (light sampling + MIS mode)  --streams V (RNG streams per pixel)  --pfm / --exr (float frame next to the PNG)
//   --env-float (environment.hdr as float radiance instead of the reference's 8-bit conversion)
//   --progressive N (frame in passes of N samples)  --target-error E (stop once the relative-variance figure <= E)
//   --save-passes (PNG after every pass)  --gamma (RGBA8 through saturate(pow(c, 1/2.2)), the reference's device.cu:248)
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <filesystem>
#include <string>
#include <vector>

#include "application.hpp"

int main(int argc, char** argv) {
    std::string assets_path = std::filesystem::current_path().string() + "/assets";
    std::string out_dir = std::filesystem::current_path().string();
    host::load_options opt;
    int device = 0, flags = 0, streams = 1, gpus = 1, pipeline = 1;
    bool stats = false, pfm = false, exr = false, gamma = false, save_passes = false, gpu_dedup = false;
    int progressive = 0;
    double target_error = 0.0;
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        auto need = [&](int n) {
            if (i + n >= argc) {
                fprintf(stderr, "missing value for %s\n", a.c_str());
                exit(2);
            }
        };
        if (a == "--assets") { need(1); assets_path = argv[++i]; }
        else if (a == "--scene") { need(1); opt.scene_override = argv[++i]; }
        else if (a == "--size") { need(2); opt.width = atoi(argv[++i]); opt.height = atoi(argv[++i]); }
        else if (a == "--spp") { need(1); opt.max_samples = atoi(argv[++i]); }
        else if (a == "--depth") { need(1); opt.max_path_depth = atoi(argv[++i]); }
        else if (a == "--env") {
            need(1);
            const std::string m = argv[++i];
            opt.env_mode = m == "map" ? 2 : (m == "auto" ? 1 : 0);
        }
        else if (a == "--env-intensity") { need(1); opt.env_intensity = (float)atof(argv[++i]); }
        else if (a == "--detail") { need(1); opt.detail = atoi(argv[++i]); }
        else if (a == "--device") { need(1); device = atoi(argv[++i]); }
        else if (a == "--gpus") { need(1); gpus = atoi(argv[++i]); }
        else if (a == "--pipeline") { need(1); pipeline = atoi(argv[++i]); }
        else if (a == "--gpu-dedup") gpu_dedup = true;
        else if (a == "--out") { need(1); out_dir = argv[++i]; }
        else if (a == "--stats") stats = true;
        else if (a == "--nee") flags |= B200PT_FLAG_NEE;
        else if (a == "--streams") { need(1); streams = atoi(argv[++i]); }
        else if (a == "--pfm") pfm = true;
        else if (a == "--exr") exr = true;
        else if (a == "--env-float") opt.env_float = true;
        else if (a == "--progressive") { need(1); progressive = atoi(argv[++i]); }
        else if (a == "--target-error") { need(1); target_error = atof(argv[++i]); }
        else if (a == "--save-passes") save_passes = true;
        else if (a == "--gamma") gamma = true;
        else if (a == "--help" || a == "-h") {
            printf("path_tracer [--assets DIR] [--scene NAME] [--size W H] [--spp N] [--depth N] [--env color|auto|map]\n"
                   "            [--env-intensity F] [--detail N] [--device K] [--gpus N] [--pipeline N] [--gpu-dedup] [--out DIR] [--stats] [--nee] [--streams V] [--pfm]\n"
                   "            [--env-float] [--progressive N] [--target-error E] [--save-passes] [--gamma]\n"
                   "without arguments: ./assets/settings.json and its scene, as the reference's executable\n");
            return 0;
        }
        else {
            fprintf(stderr, "unknown argument %s\n", a.c_str());
            return 2;
        }
    }
    try {
        host::program_data pdata;
        host::test_data tdata;
        b200pt_ctx* loader_ctx = nullptr;   // --gpu-dedup: the loader's create_mesh stage runs on the device
        if (gpu_dedup) {
            b200pt_config lc{};
            lc.device_ordinal = device;
            lc.world_size = 1;
            if (b200pt_create(&lc, &loader_ctx) != B200PT_OK) {
                fprintf(stderr, "b200pt_create failed: %s\n", b200pt_last_error());
                return 1;
            }
            opt.create_mesh = host::gpu_create_mesh(loader_ctx);
        }
        host::init_program_data(pdata, tdata, assets_path, opt);
        if (loader_ctx) b200pt_destroy(loader_ctx);
        for (const std::string& n : pdata.notes) printf("NOTE: %s\n", n.c_str());

        pdata.write_pfm = pfm;
        pdata.write_exr = exr;
        pdata.progressive_pass = progressive;
        pdata.target_error = target_error;
        pdata.save_passes = save_passes;
        if (target_error > 0.0 && progressive <= 0) pdata.progressive_pass = 16;
        if (gpus < 1 || (gpus > 1 && pdata.progressive_pass > 0)) {
            fprintf(stderr, "--gpus needs N >= 1 and renders whole frames (no --progressive / --target-error)\n");
            return 2;
        }
        std::vector<b200pt_ctx*> ranks;
        for (int r = 0; r < gpus; ++r) {
            b200pt_config cfg{};
            cfg.device_ordinal = device + r;
            cfg.rank = r;
            cfg.world_size = gpus;
            cfg.shard_mode = B200PT_SHARD_SPP;
            cfg.flags = flags;
            cfg.streams_per_pixel = streams;
            b200pt_ctx* c = nullptr;
            if (b200pt_create(&cfg, &c) != B200PT_OK) {
                fprintf(stderr, "b200pt_create (rank %d) failed: %s\n", r, b200pt_last_error());
                return 1;
            }
            host::bind_scene(pdata, c);
            if (gamma && b200pt_set_output(c, B200PT_TRANSFER_GAMMA22) != B200PT_OK) {
                fprintf(stderr, "b200pt_set_output failed: %s\n", b200pt_last_error());
                return 1;
            }
            ranks.push_back(c);
        }
        b200pt_ctx* ctx = ranks[0];
        for (int k = 1; k < pipeline && gpus == 1; ++k) {   // further slots of the frame pipeline (one GPU)
            b200pt_config cfg{};
            cfg.device_ordinal = device;
            cfg.world_size = 1;
            cfg.flags = flags;
            cfg.streams_per_pixel = streams;
            b200pt_ctx* c = nullptr;
            if (b200pt_create(&cfg, &c) != B200PT_OK) {
                fprintf(stderr, "b200pt_create (pipeline slot %d) failed: %s\n", k, b200pt_last_error());
                return 1;
            }
            host::bind_scene(pdata, c);
            if (gamma) b200pt_set_output(c, B200PT_TRANSFER_GAMMA22);
            pdata.pipeline_slots.push_back(c);
        }
        if (gpus > 1) {
            if (b200pt_comm_init_all(ranks.data(), gpus) != B200PT_OK) {
                fprintf(stderr, "b200pt_comm_init_all failed: %s\n", b200pt_last_error());
                return 1;
            }
            pdata.ranks = ranks;
        }
        host::test_loop(ctx, pdata, tdata, out_dir);
        if (stats) {
            b200pt_stats st{};
            b200pt_get_stats(ctx, &st);
            printf("{\"render_ms\": %.3f, \"extend_ms\": %.3f, \"shade_ms\": %.3f, \"build_ms\": %.3f, \"rays\": %llu, "
                   "\"paths\": %llu, \"mrays_per_s\": %.1f, \"bvh_nodes\": %u, \"bvh_tris\": %u}\n",
                   st.render_ms, st.extend_ms, st.shade_ms, st.build_ms, (unsigned long long)st.rays,
                   (unsigned long long)st.paths, st.render_ms > 0 ? st.rays / st.render_ms * 1e-3 : 0.0, st.bvh_nodes,
                   st.bvh_tris);
        }
        for (b200pt_ctx* c : pdata.pipeline_slots) b200pt_destroy(c);
        for (b200pt_ctx* c : ranks) b200pt_destroy(c);
    } catch (const std::exception& e) {
        fprintf(stderr, "error: %s\n", e.what());
        return 1;
    }
    return 0;
}
