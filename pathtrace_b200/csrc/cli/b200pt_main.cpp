// b200pt_main.cpp -- drop-in for the reference's main() (main.cpp:108-168): reads ./config.json, the scene it
// names, renders through the C ABI with the Renderer lifecycle (start_render / sync_progress loop / finalize)
// and writes the same PPM and the same stdout rate lines (renderer.h:98-112, 700-724).
#include "b200pt.h"

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <sstream>
#include <string>
#include <algorithm>
#include <thread>

static std::string slurp(const std::string &path)
{
    std::ifstream f(path, std::ios::binary);
    if (!f.good()) {
        std::cerr << "cannot open " << path << std::endl;
        exit(2);
    }
    std::stringstream ss;
    ss << f.rdbuf();
    return ss.str();
}

static void die(const char *what)
{
    std::cerr << what << ": " << b200pt_last_error() << std::endl;
    exit(1);
}

int main(int argc, char *argv[])
{
    const char *config_path = argc > 1 ? argv[1] : "config.json"; // the reference hard-codes config.json
    std::string config_text = slurp(config_path);
    b200pt_config config;
    char scene_path[1024], ppm_path[1024];
    if (b200pt_config_from_json(config_text.c_str(), &config, scene_path, sizeof(scene_path), ppm_path, sizeof(ppm_path)))
        die("config");
    const long total_pixels = (long)config.width * config.height;
    const long min_camera_rays = (long)config.samples * total_pixels;
    std::cout << "trace probability is " << config.trace_probability << std::endl;
    std::cout << std::endl;
    std::cout << "config read complete, rendering image at " << config.width << "x" << config.height << "==" << total_pixels
              << "px^2" << std::endl;
    std::cout << "with " << config.samples << " samples per pixel, that sets a minimum number of camera rays at "
              << min_camera_rays << "\n\n";
    std::cout << "using " << config.threads << " threads (ignored: " << b200pt_version() << ", "
              << (config.device_count > 1 ? config.device_count : 1) << " GPU(s))\n";
    std::cout << "reading scene data" << std::endl;
    std::string scene_text = slurp(scene_path);

    auto t1 = std::chrono::high_resolution_clock::now();
    b200pt_scene *scene = nullptr;
    if (b200pt_scene_create_from_json(scene_text.c_str(), config.width, config.height, config.device, &scene))
        die("scene");
    auto t2 = std::chrono::high_resolution_clock::now();
    std::cout << "time taken to build bvh " << std::chrono::duration<double>(t2 - t1).count() << std::endl;

    // start_render (renderer.h:553-603) returns at once: one worker thread per GPU renders in the background
    // ("devices": N in config.json puts N GPUs on the frame); this thread is main()'s polling loop (main.cpp:158-163)
    config.async_render = 1;
    b200pt_render *render = nullptr;
    if (b200pt_render_start(scene, &config, &render))
        die("render_start");
    auto start = std::chrono::high_resolution_clock::now();
    b200pt_progress pr{};
    // tests shorten the 0.5 s of main.cpp:162 and keep every preview: B200PT_CLI_POLL_MS, B200PT_CLI_KEEP_PREVIEWS=<prefix>
    const char *poll_env = std::getenv("B200PT_CLI_POLL_MS");
    const int poll_ms = poll_env ? std::max(1, atoi(poll_env)) : 500;
    const char *keep = std::getenv("B200PT_CLI_KEEP_PREVIEWS");
    int previews = 0;
    while (true) {
        // sync_progress (renderer.h:605-620): progress line, then the film so far tonemapped with 1 + done / pixels
        // samples and written over the PPM
        if (b200pt_render_poll(render, 0, &pr))
            die("render_poll");
        if (pr.done)
            break;
        auto now = std::chrono::high_resolution_clock::now();
        double elapsed = std::chrono::duration<double>(now - start).count();
        float rate = (float)(pr.samples_done / (elapsed > 0 ? elapsed : 1e-9));
        std::cout << "samples left" << std::setw(20) << (pr.samples_total - pr.samples_done) << " rate " << std::setw(10)
                  << rate << " time left " << std::setw(5) << (pr.samples_total - pr.samples_done) / rate
                  << "                                           " << '\r' << std::flush;
        const long n_preview = 1 + pr.samples_done / total_pixels;
        if (b200pt_render_write_ppm(render, n_preview, ppm_path))
            die("write_ppm");
        if (keep) {
            const std::string copy = std::string(keep) + std::to_string(previews) + ".ppm";
            if (b200pt_render_write_ppm(render, n_preview, copy.c_str()))
                die("write_ppm");
            std::cout << "\npreview " << previews << " samples " << n_preview << std::endl;
        }
        previews++;
        std::this_thread::sleep_for(std::chrono::milliseconds(poll_ms));
    }
    if (b200pt_render_wait(render, &pr))
        die("render_wait");
    std::cout << " done\n";

    // finalize (renderer.h:693-727)
    double secs = std::chrono::duration<double>(std::chrono::high_resolution_clock::now() - start).count();
    long total_bounces = pr.rays_extension + pr.rays_shadow;
    std::cout << "time taken to compute " << secs << std::endl;
    std::cout << "computed " << min_camera_rays << " camera rays in " << secs << "s, at " << (float)(min_camera_rays / secs)
              << " rays per second" << std::endl;
    std::cout << "computed " << total_bounces << " rays, at " << (float)(total_bounces / secs) << " rays per second"
              << " (reference accounting: " << pr.rays_reference << " rays)" << std::endl;
    std::cout << "device time " << pr.seconds << " s: " << min_camera_rays / pr.seconds * 1e-6 << " Mpaths/s, "
              << total_bounces / pr.seconds * 1e-6 << " Mrays/s" << std::endl;
    float max_lum = 0, avg_lum = 0;
    if (b200pt_render_tonemap(render, config.samples, nullptr, &max_lum, &avg_lum))
        die("tonemap");
    std::cout << "avg lum " << avg_lum << std::endl;
    std::cout << "max lum " << max_lum << std::endl;
    if (b200pt_render_write_ppm(render, config.samples, ppm_path))
        die("write_ppm");
    if (config.trace_probability > 0.f) { // handle_traced_paths (renderer.h:57-96)
        char p3[1024], p2[1024];
        int32_t added = 0;
        if (b200pt_config_string(config_text.c_str(), "traced_paths_output_path", p3, sizeof(p3)) ||
            b200pt_config_string(config_text.c_str(), "traced_paths_2d_output_path", p2, sizeof(p2)) ||
            b200pt_render_write_paths(render, p3, p2, &added))
            die("write_paths");
        std::cout << "wrote " << added << " traced paths" << std::endl;
    }
    b200pt_render_destroy(render);
    b200pt_scene_destroy(scene);
    return 0;
}
