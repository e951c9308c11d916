// procedural.hpp — see procedural.cpp
#pragma once
#include "scene.hpp"

namespace host {

// shapes of scene "sphere" | "mitsuba" | "dragon" | "car" (empty vector for anything else).
// detail: 0 = default budget; dragon: target triangle count; car: resolution multiplier.
std::vector<std::pair<std::string, std::shared_ptr<mesh>>> make_procedural_scene(const std::string& scene, int detail);
void make_procedural_environment(int w, int h, std::vector<float>& rgb);  // float RGB, row 0 = top
image_buffer make_procedural_texture(const std::string& file, int size);  // row 0 = top (as decoded from a file)

}  // namespace host
