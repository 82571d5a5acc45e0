#pragma once
#include "../../../include/b200pt.h"
#include "json.h"

#include <string>

namespace b200pt {

struct ParsedConfig {
    b200pt_config config;
    std::string ppm_output_path, png_output_path, traced_paths_output_path, traced_paths_2d_output_path, scene_path;
    std::string render_type, integrator_type;
    bool should_trace_paths = false;
    float avg_number_of_paths = 100.f;
    int32_t device_count = 1;
};

ParsedConfig parse_config(const Json &j);

} // namespace b200pt
