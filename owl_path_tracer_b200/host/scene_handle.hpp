// scene_handle.hpp — the object behind the C handle b200pt_scene (include/b200pt_host.h), shared by scene_api.cpp
// (libb200pt_scene.so) and host_api.cpp (libb200pt_host.so).
#pragma once
#include <stdexcept>
#include <string>
#include <memory>
#include <utility>
#include <vector>

#include "../../include/b200pt_host.h"
#include "application.hpp"

struct b200pt_scene {
    host::program_data pdata;
    host::test_data tdata;
    std::vector<host::material_data> mats;  // scratch for b200pt_scene_materials
};

struct b200pt_shapes {
    std::vector<std::pair<std::string, std::shared_ptr<host::mesh>>> shapes;
};

namespace host {
std::string& api_error();   // thread-local message behind b200pt_host_last_error
template <typename F>
int api_guard(F&& f) {
    try {
        f();
        return 0;
    } catch (const std::exception& e) {
        api_error() = e.what();
        return 1;
    }
}
}  // namespace host
