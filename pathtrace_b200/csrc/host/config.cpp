// config.cpp -- config.json -> b200pt_config, key for key as Config(json) does (config.h:98-132).
#include "config.h"

#include <stdexcept>

namespace b200pt {

ParsedConfig parse_config(const Json &j)
{
    ParsedConfig out;
    b200pt_config &c = out.config;
    c = b200pt_config{};
    if (!j.contains("film"))
        throw std::runtime_error("config: missing \"film\" (config.h:100)");
    const Json &film = j["film"];
    c.width = (int32_t)film.value("width", 400);
    c.height = (int32_t)film.value("height", 300);
    // config.h:24-25: the two keys are read into each other's field; only `exposure` is ever used
    // (renderer.h:37: col *= 16 + exposure), so the display multiplier is 16 + JSON "gamma".
    c.exposure = (float)film.value("gamma", 2.2);
    c.gamma = (float)film.value("exposure", 0.0);

    out.ppm_output_path = j.value("ppm_output_path", "out.ppm");
    out.png_output_path = j.value("png_output_path", "out.png");
    // config.h:104-105: these two are read with operator[] + get<std::string>() and are therefore REQUIRED
    if (!j.contains("traced_paths_output_path") || !j["traced_paths_output_path"].is_string())
        throw std::runtime_error("config: \"traced_paths_output_path\" is required (config.h:104)");
    if (!j.contains("traced_paths_2d_output_path") || !j["traced_paths_2d_output_path"].is_string())
        throw std::runtime_error("config: \"traced_paths_2d_output_path\" is required (config.h:105)");
    out.traced_paths_output_path = j["traced_paths_output_path"].as_string();
    out.traced_paths_2d_output_path = j["traced_paths_2d_output_path"].as_string();
    out.scene_path = j.value("scene", "scenes/scene.json");
    out.should_trace_paths = j.value("should_trace_paths", false);
    out.avg_number_of_paths = (float)j.value("avg_number_of_paths", 100.0);
    c.only_direct_illumination = j.value("only_direct_illumination", false) ? 1 : 0;
    c.block_width = (int32_t)j.value("block_width", 64);
    c.block_height = (int32_t)j.value("block_height", 64);

    out.render_type = j.value("render_type", "progressive");
    out.integrator_type = j.value("integrator_type", "recursive path tracing");
    // get_integrator_type_for (config.h:59-72) + integrator_from_config (main.cpp:30-55): INEEPT and RNEEPT have
    // their own classes, every other string (known or not) ends in RecursivePT.
    if (out.integrator_type == "iterative nee path tracing")
        c.integrator = B200PT_INTEGRATOR_INEEPT;
    else if (out.integrator_type == "recursive nee path tracing")
        throw std::runtime_error("config: integrator \"recursive nee path tracing\" (NEERecursive, integrator.h:78-169) "
                                 "branches twice per vertex and is not provided by the wavefront backend");
    else
        c.integrator = B200PT_INTEGRATOR_RPT;
    c.max_bounces = (int32_t)j.value("max_bounces", 10);
    c.samples = (int32_t)j.value("samples", 20);
    c.threads = (int32_t)(uint16_t)j.value("threads", 1);
    c.normal_offset = (float)j.value("normal_offset", 0.0001);
    c.light_samples = (int32_t)j.value("light_samples", 1);
    c.russian_roulette = j.value("russian_roulette", true) ? 1 : 0;

    // extensions of this backend (absent from the reference's file)
    c.seed = (uint64_t)j.value("seed", 0L);
    c.device = (int32_t)j.value("device", 0);
    c.batch_samples = (int32_t)j.value("batch_samples", 0);
    // "devices": N renders on GPUs 0 .. N-1, "device_ids": [a, b, ...] on the listed ones (the film is combined on the
    // first); like the reference's "threads", a scheduling key that does not change the image
    out.device_count = (int32_t)j.value("devices", (long)j.value("device_count", 1));
    if (j.contains("device_ids") && j["device_ids"].is_array()) {
        const Json &ids = j["device_ids"];
        if (ids.size() < 1 || ids.size() > 8)
            throw std::runtime_error("config: \"device_ids\" must list 1 to 8 devices");
        c.device_count = (int32_t)ids.size();
        for (size_t k = 0; k < ids.size(); k++)
            c.device_ids[k] = (int32_t)ids.at(k).as_double();
        c.device = c.device_ids[0];
    } else if (out.device_count > 1) {
        if (out.device_count > 8)
            throw std::runtime_error("config: \"devices\" must be at most 8");
        c.device_count = out.device_count;
        for (int k = 0; k < out.device_count; k++)
            c.device_ids[k] = c.device + k;
    }
    out.device_count = c.device_count > 1 ? c.device_count : 1;
    c.async_render = j.value("async_render", false) ? 1 : 0;
    // config.h:122-131: `long min_camera_rays = samples * film.total_pixels`
    c.trace_probability = 0.f;
    if (out.should_trace_paths)
        c.trace_probability = (float)(j.value("avg_number_of_paths", 100.0) / (double)((long)c.samples * ((long)c.width * c.height)));
    c.trace_max_paths = (int32_t)j.value("trace_max_paths", 0);
    c.physically_corrected = j.value("physically_corrected", false) ? 1 : 0;

    if (c.width <= 0 || c.height <= 0)
        throw std::runtime_error("config: film size must be positive");
    if (c.samples <= 0)
        throw std::runtime_error("config: samples must be positive (helpers.h:149 asserts n_samples > 0)");
    if (c.max_bounces <= 0 && c.integrator == B200PT_INTEGRATOR_RPT)
        throw std::runtime_error("config: max_bounces must be positive (integrator.h:25 asserts)");
    return out;
}

} // namespace b200pt
