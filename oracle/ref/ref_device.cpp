// ref_device.cpp — C ABI around the REFERENCE'S OWN render-path sources, compiled for the host.
// TEST INFRASTRUCTURE: built into oracle/_ref/libref_device.so (git-ignored), loaded only by tests/,
// tools/make_golden.py and bench.py's parity check; the product never links or loads it.
//
// This translation unit #includes /root/reference/path_tracer/src/device/device.cu (the copy that
// patch.py made under oracle/_ref/build/src) — i.e. __raygen__ray_gen, trace_path, the closest-hit
// and miss programs, sample_disney and every lobe, math.hpp, sample_methods.hpp, random.hpp — and
// runs those functions on CPU threads.  Nothing of the estimator is restated here.  The only
// stand-ins are optixTrace (closed source; a closest-hit callback, normally oracle_trace) and
// tex2D, both in shim/owl/owl_device.h.
//
// What the entry points replace on the reference side (all application.cpp):
//   ref_add_mesh         :197-246  per-entity buffers, entity_data{mesh_index, material_index, has_texture, texture}
//   ref_set_materials    :257-259  material_buffer
//   ref_set_environment  :253-254, 277-289  environment texture + flags
//   ref_set_camera       :270-275  ray_gen_data.camera
//   ref_set_render       :271, 283-284   fb_size, max_samples, max_path_depth
//   ref_render           :363-371  owlLaunch2D(ray_gen, W, H): one call of __raygen__ray_gen per pixel
#include "device/device.cu"
#include "device/lights.hpp"

#include <atomic>
#include <memory>
#include <mutex>
#include <thread>

namespace {

struct RefMesh {
    std::vector<vec3> verts, normals;
    std::vector<vec2> uvs;
    std::vector<ivec3> idx;
    std::vector<unsigned char> tex_pixels;
    refshim::Texture tex{};
    entity_data record{};
};

struct RefScene {
    std::vector<std::unique_ptr<RefMesh>> meshes;
    std::vector<material_data> materials;
    std::vector<unsigned char> env_pixels;
    refshim::Texture env{};
    bool env_use = false, env_auto = false;
    vec3 env_color{0.f};
    float env_intensity = 0.f;
    camera_data cam{};
    int width = 0, height = 0, spp = 1, depth = 8;
    refshim::World world{};
    // device-side tables (Buffer of Buffers, device_global.hpp:46-66)
    std::vector<Buffer> vb, ib, nb, tb;
    std::vector<const void*> records;
};

std::mutex g_launch_mutex;   // optixLaunchParams is one global, like the __constant__ it stands for

Buffer make_buffer(size_t count, const void* data) { return Buffer{0, count, const_cast<void*>(data)}; }

}  // namespace

extern "C" {

RefScene* ref_create() { return new RefScene(); }
void ref_destroy(RefScene* s) { delete s; }

int ref_add_mesh(RefScene* s, const float* verts, const float* normals, const float* uvs, int nv, const int32_t* idx, int nt,
                 int material, const uint8_t* tex_rgba, int tw, int th) {
    auto m = std::make_unique<RefMesh>();
    m->verts.resize(nv);
    memcpy((void*)m->verts.data(), verts, sizeof(float) * 3 * nv);
    if (normals) {
        m->normals.resize(nv);
        memcpy((void*)m->normals.data(), normals, sizeof(float) * 3 * nv);
    }
    if (uvs) {
        m->uvs.resize(nv);
        memcpy((void*)m->uvs.data(), uvs, sizeof(float) * 2 * nv);
    }
    m->idx.resize(nt);
    memcpy((void*)m->idx.data(), idx, sizeof(int32_t) * 3 * nt);
    m->record.mesh_index = (int32_t)s->meshes.size();
    m->record.material_index = material;
    m->record.has_texture = false;
    m->record.texture = 0;
    if (tex_rgba && tw > 0 && th > 0) {
        m->tex_pixels.assign(tex_rgba, tex_rgba + (size_t)4 * tw * th);
        m->tex = refshim::Texture{tw, th, m->tex_pixels.data()};
        m->record.has_texture = true;
        m->record.texture = (cudaTextureObject_t)(uintptr_t)&m->tex;
    }
    s->meshes.push_back(std::move(m));
    return (int)s->meshes.size() - 1;
}

void ref_set_materials(RefScene* s, const float* m, int nm) {
    static_assert(sizeof(material_data) == 17 * sizeof(float), "material_data is 17 floats (device_global.hpp:19-36)");
    s->materials.resize(nm);
    memcpy((void*)s->materials.data(), m, sizeof(material_data) * nm);
}

void ref_set_environment(RefScene* s, const uint8_t* rgba, int w, int h, int use, int auto_, const float* color,
                         float intensity) {
    s->env_pixels.clear();
    s->env = refshim::Texture{};
    if (rgba && w > 0 && h > 0) {
        s->env_pixels.assign(rgba, rgba + (size_t)4 * w * h);
        s->env = refshim::Texture{w, h, s->env_pixels.data()};
    }
    s->env_use = use != 0;
    s->env_auto = auto_ != 0;
    s->env_color = vec3{color[0], color[1], color[2]};
    s->env_intensity = intensity;
}

void ref_set_camera(RefScene* s, const float* origin, const float* llc, const float* hor, const float* ver) {
    s->cam.origin = vec3{origin[0], origin[1], origin[2]};
    s->cam.llc = vec3{llc[0], llc[1], llc[2]};
    s->cam.horizontal = vec3{hor[0], hor[1], hor[2]};
    s->cam.vertical = vec3{ver[0], ver[1], ver[2]};
}

void ref_set_render(RefScene* s, int w, int h, int spp, int depth) {
    s->width = w;
    s->height = h;
    s->spp = spp;
    s->depth = depth;
}

// closest-hit provider standing in for the OptiX traversal (signature of oracle_trace)
void ref_set_tracer(RefScene* s, refshim::trace_fn fn, void* scene) {
    s->world.trace = fn;
    s->world.scene = scene;
}

// One launch (application.cpp:363-371) over the window [x0,x1) x [y0,y1) of launch indices.
// mean: (H, W, 3) float, the argument of make_rgba, stored like the framebuffer (row flip of
// device.cu:251 applied); rgba: (H, W) uint32 framebuffer.  Pixels outside the window stay untouched.
// Returns the number of traceRay calls.
uint64_t ref_render(RefScene* s, int x0, int y0, int x1, int y1, int nthreads, float* mean, uint32_t* rgba) {
    std::lock_guard<std::mutex> lock(g_launch_mutex);
    const int W = s->width, H = s->height;
    size_t nm = s->meshes.size();
    s->vb.resize(nm); s->ib.resize(nm); s->nb.resize(nm); s->tb.resize(nm); s->records.resize(nm);
    for (size_t i = 0; i < nm; i++) {
        RefMesh& m = *s->meshes[i];
        s->vb[i] = make_buffer(m.verts.size(), m.verts.data());
        s->ib[i] = make_buffer(m.idx.size(), m.idx.data());
        s->nb[i] = make_buffer(m.normals.size(), m.normals.empty() ? nullptr : m.normals.data());
        s->tb[i] = make_buffer(m.uvs.size(), m.uvs.empty() ? nullptr : m.uvs.data());
        s->records[i] = &m.record;
    }
    s->world.entity_records = s->records.data();
    // the SBT of application.cpp:79 ("triangle_hit", ray type 0), :98-101 ("miss" -> 0, "miss_shadow" -> 1)
    s->world.prog.closest_hit = &__closesthit__triangle_hit;
    s->world.prog.miss[0] = &__miss__miss;
    s->world.prog.miss[1] = &__miss__miss_shadow;

    launch_params_data& lp = optixLaunchParams;
    lp.max_path_depth = s->depth;
    lp.max_samples = s->spp;
    lp.material_buffer = make_buffer(s->materials.size(), s->materials.data());
    lp.vertices_buffer = make_buffer(nm, s->vb.data());
    lp.indices_buffer = make_buffer(nm, s->ib.data());
    lp.normals_buffer = make_buffer(nm, s->nb.data());
    lp.texcoords_buffer = make_buffer(nm, s->tb.data());
    lp.world = (OptixTraversableHandle)(uintptr_t)&s->world;
    lp.environment_map = s->env.rgba ? (cudaTextureObject_t)(uintptr_t)&s->env : 0;
    lp.environment_use = s->env_use;
    lp.environment_auto = s->env_auto;
    lp.environment_color = s->env_color;
    lp.environment_intensity = s->env_intensity;

    ray_gen_data rg{};
    rg.fb_ptr = rgba;
    rg.fb_size = ivec2{W, H};
    rg.camera = s->cam;

    std::atomic<int> next_row{y0};
    std::atomic<uint64_t> rays{0};
    auto worker = [&]() {
        refshim::Thread& t = refshim::tls();
        t = refshim::Thread{};
        t.raygen_data = &rg;
        for (;;) {
            int y = next_row.fetch_add(1);
            if (y >= y1) break;
            for (int x = x0; x < x1; x++) {
                t.launch_index = ivec2{x, y};
                t.hit_data = nullptr;
                __raygen__ray_gen();
                if (mean) {
                    size_t o = (size_t)x + (size_t)W * (size_t)(H - 1 - y);
                    mean[3 * o + 0] = t.last_color.x;
                    mean[3 * o + 1] = t.last_color.y;
                    mean[3 * o + 2] = t.last_color.z;
                }
            }
        }
        rays += t.rays;
    };
    nthreads = std::max(1, nthreads);
    std::vector<std::thread> th;
    for (int i = 1; i < nthreads; i++) th.emplace_back(worker);
    worker();
    for (auto& x : th) x.join();
    return rays.load();
}

// ---- unit entry points (same signatures as the oracle's, tests/checkers.py)

// sample_disney (disney.cuh:31-66) on n cases: material index, local wo, lobe of the previous bounce, RNG state in.
void ref_sample_disney(const float* mats, int n, const int32_t* material, const float* wo, const int32_t* lobe_in,
                       const uint32_t* rng_in, float* f_out, float* wi_out, float* pdf_out, int32_t* lobe_out,
                       uint32_t* rng_out) {
    for (int i = 0; i < n; i++) {
        material_data m;
        memcpy((void*)&m, mats + 17 * (size_t)material[i], sizeof m);
        struct random r;
        r.state = rng_in[i];
        vec3 wi{0.f};   // trace_path declares `local_wi{}` (device.cu:171)
        float pdf{};
        int lobe = lobe_in[i];
        vec3 f = sample_disney(m, vec3{wo[3 * i], wo[3 * i + 1], wo[3 * i + 2]}, r, wi, pdf, lobe);
        f_out[3 * i] = f.x; f_out[3 * i + 1] = f.y; f_out[3 * i + 2] = f.z;
        wi_out[3 * i] = wi.x; wi_out[3 * i + 1] = wi.y; wi_out[3 * i + 2] = wi.z;
        pdf_out[i] = pdf;
        lobe_out[i] = lobe;
        rng_out[i] = r.state;
    }
}

// random.hpp:34-85: seed (x, y), n float draws, final state
void ref_rng(uint32_t x, uint32_t y, int n, float* out, uint32_t* state_out) {
    struct random r{x, y};
    for (int i = 0; i < n; i++) out[i] = r();
    *state_out = r.state;
}

// camera.cpp:3-22
void ref_camera(const float* from, const float* at, const float* up, float vfov, int w, int h, float* out) {
    camera c{vec3{from[0], from[1], from[2]}, vec3{at[0], at[1], at[2]}, vec3{up[0], up[1], up[2]}, vfov};
    camera_data d = to_camera_data(c, ivec2{w, h});
    const vec3* v[4] = {&d.origin, &d.llc, &d.horizontal, &d.vertical};
    for (int i = 0; i < 4; i++) { out[3 * i] = v[i]->x; out[3 * i + 1] = v[i]->y; out[3 * i + 2] = v[i]->z; }
}

// math.hpp:86-107: onb of n, then to_local(w) and to_world(w).  out: t(3) b(3) local(3) world(3) per case
void ref_onb(const float* n_in, const float* w_in, int n, float* out) {
    for (int i = 0; i < n; i++) {
        vec3 nn{n_in[3 * i], n_in[3 * i + 1], n_in[3 * i + 2]}, w{w_in[3 * i], w_in[3 * i + 1], w_in[3 * i + 2]};
        vec3 t{}, b{};
        onb(nn, t, b);
        vec3 l = to_local(t, b, nn, w), g = to_world(t, b, nn, w);
        const vec3* v[4] = {&t, &b, &l, &g};
        for (int k = 0; k < 4; k++) { out[12 * i + 3 * k] = v[k]->x; out[12 * i + 3 * k + 1] = v[k]->y; out[12 * i + 3 * k + 2] = v[k]->z; }
    }
}

// device.cu:23-28 uv_on_sphere
void ref_uv_on_sphere(const float* d, int n, float* out) {
    for (int i = 0; i < n; i++) {
        vec2 uv = uv_on_sphere(vec3{d[3 * i], d[3 * i + 1], d[3 * i + 2]});
        out[2 * i] = uv.x; out[2 * i + 1] = uv.y;
    }
}

// owl_device.h:79-91
uint32_t ref_quantise(float r, float g, float b) { return owl::real::make_rgba(vec3{r, g, b}); }

// lights.hpp:15-61 (dead code in the reference; the opt-in NEE mode builds on these three)
float ref_triangle_area(const float* a, const float* b, const float* c) {
    return triangle_area(vec3{a[0], a[1], a[2]}, vec3{b[0], b[1], b[2]}, vec3{c[0], c[1], c[2]});
}
float ref_pdf_area_to_solid_angle(float pdf_area, float dist_sqr, float cos_theta) {
    return pdf_area_to_solid_angle(pdf_area, dist_sqr, cos_theta);
}
float ref_power_heuristic(float nf, float pf, float ng, float pg) { return power_heuristic(nf, pf, ng, pg); }
void ref_uniform_sample_triangle(float u0, float u1, float* out) {
    vec2 b = uniform_sample_triangle(vec2{u0, u1});
    out[0] = b.x; out[1] = b.y;
}

}  // extern "C"

// camera.cpp is plain host C++ of the reference: compiled into this library as is
#include "camera.cpp"
