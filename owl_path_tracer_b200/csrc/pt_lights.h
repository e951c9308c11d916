// pt_lights.h — host-side light table of the next-event estimation mode (B200PT_FLAG_NEE).
//
// The reference samples no lights (its lights.hpp is dead code); this mode is assembled
// from those helpers.  Every triangle of every mesh whose material has emission > 0 becomes one entry,
// in mesh order; an entry is chosen with probability proportional to triangle_area * emission
// (lights.hpp:8-20: power = L_emit * area * pi).  The cdf is accumulated in binary64 and rounded once, so
// the scalar oracle (oracle/oracle.cpp build_lights) arrives at the same bits independently.
// Used by api.cu (upload) and tests/emu (host emulation of the kernels).
#pragma once
#include <vector>

#include "pt_integrator.h"

namespace pt {

// meshes[].light_first is rewritten; lights/cdf are empty when nothing emits.
inline void build_light_table(const std::vector<f4>& pos, const std::vector<int32_t>& idx4, std::vector<MeshInfo>& meshes,
                              const std::vector<Material>& mats, std::vector<LightTri>& lights, std::vector<float>& cdf) {
    lights.clear();
    cdf.clear();
    std::vector<double> acc;
    double total = 0.0;
    const size_t n_tris = idx4.size() / 4;
    for (size_t mi = 0; mi < meshes.size(); ++mi) {
        meshes[mi].light_first = -1;
        const int32_t mat = meshes[mi].material;
        if (mat < 0 || (size_t)mat >= mats.size()) continue;
        const float emission = mats[(size_t)mat].emission;
        if (!(emission > 0.0f)) continue;
        const size_t first = (size_t)meshes[mi].first_tri;
        const size_t last = mi + 1 < meshes.size() ? (size_t)meshes[mi + 1].first_tri : n_tris;
        meshes[mi].light_first = (int32_t)lights.size();
        for (size_t t = first; t < last; ++t) {
            const f4 a = pos[(size_t)idx4[4 * t]], b = pos[(size_t)idx4[4 * t + 1]], c = pos[(size_t)idx4[4 * t + 2]];
            const f3 n = cross(f3{b.x, b.y, b.z} - f3{a.x, a.y, a.z}, f3{c.x, c.y, c.z} - f3{a.x, a.y, a.z});
            const float area = 0.5f * sqrtf(dot(n, n));
            const float w = area * emission;
            total += (double)w;
            acc.push_back(total);
            lights.push_back(LightTri{(uint32_t)t, area, 0.0f, emission});
        }
    }
    if (!(total > 0.0)) {
        lights.clear();
        for (MeshInfo& m : meshes) m.light_first = -1;
        return;
    }
    cdf.resize(lights.size());
    for (size_t i = 0; i < lights.size(); ++i) cdf[i] = (float)(acc[i] / total);
    cdf.back() = 1.0f;
    for (size_t i = 0; i < lights.size(); ++i) lights[i].prob = cdf[i] - (i ? cdf[i - 1] : 0.0f);
}

// Environment map of the NEE mode: P(texel) proportional to the Rec.709 luminance of the texel.  rows: marginal
// cdf over the h rows; cols: cdf of every row over its w texels (a black row gets a uniform cdf; it is never
// selected).  Sums in binary64, rounded once per entry.  Returns false (tables empty) when the whole map is black.
template <class LumFn>
inline bool build_env_table_fn(int w, int h, LumFn lum_of, std::vector<float>& rows, std::vector<float>& cols) {
    rows.clear();
    cols.clear();
    if (w <= 0 || h <= 0) return false;
    rows.resize((size_t)h);
    cols.resize((size_t)w * (size_t)h);
    std::vector<double> row_acc((size_t)h), col_acc((size_t)w);
    double total = 0.0;
    for (int j = 0; j < h; ++j) {
        double acc = 0.0;
        for (int i = 0; i < w; ++i) {
            acc += (double)lum_of((size_t)j * (size_t)w + (size_t)i);
            col_acc[(size_t)i] = acc;
        }
        float* row = cols.data() + (size_t)j * (size_t)w;
        for (int i = 0; i < w; ++i) row[i] = acc > 0.0 ? (float)(col_acc[(size_t)i] / acc) : (float)(i + 1) / (float)w;
        row[w - 1] = 1.0f;
        total += acc;
        row_acc[(size_t)j] = total;
    }
    if (!(total > 0.0)) {
        rows.clear();
        cols.clear();
        return false;
    }
    for (int j = 0; j < h; ++j) rows[(size_t)j] = (float)(row_acc[(size_t)j] / total);
    rows[(size_t)h - 1] = 1.0f;
    return true;
}
// RGBA8 map (LDR by the reference's loading route)
inline bool build_env_table(const uint32_t* rgba, int w, int h, std::vector<float>& rows, std::vector<float>& cols) {
    if (!rgba) return false;
    return build_env_table_fn(w, h, [rgba](size_t k) {
        const uint32_t p = rgba[k];
        const float r = (float)(p & 0xFFu) / 255.0f, g = (float)((p >> 8) & 0xFFu) / 255.0f, b = (float)((p >> 16) & 0xFFu) / 255.0f;
        return 0.2126f * r + 0.7152f * g + 0.0722f * b;
    }, rows, cols);
}
// float map, texels as f4 (rgb, -); negative or NaN luminance counts as 0
inline bool build_env_table_f32(const f4* texels, int w, int h, std::vector<float>& rows, std::vector<float>& cols) {
    if (!texels) return false;
    return build_env_table_fn(w, h, [texels](size_t k) {
        const float lum = 0.2126f * texels[k].x + 0.7152f * texels[k].y + 0.0722f * texels[k].z;
        return lum > 0.0f ? lum : 0.0f;
    }, rows, cols);
}

}  // namespace pt
