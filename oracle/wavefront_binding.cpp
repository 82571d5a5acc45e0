// oracle/wavefront_binding.cpp -- TEST INFRASTRUCTURE: the reference-side binding of INTEGRATION.md section B, compiled.
//
// `Wavefront : public Renderer` (renderer.h:114-150) over the C ABI of libb200pt.so, built against the UNMODIFIED reference
// headers where they lie (-I<reference root>; `#include "main.cpp"` with its main() renamed, as oracle/ref_driver.cpp does)
// and driven by the reference's own lifecycle, statement for statement main.cpp:108-168:
//
//     Renderer *renderer = ...;  renderer->start_render(t2);
//     while (!renderer->is_done()) { renderer->sync_progress(); sleep 0.5 s; }   renderer->finalize();
//
// The only line that differs from main() is the factory: `"render_type": "wavefront"` in config.json selects this class
// (read from the raw JSON, because config.h maps unknown strings to NAIVE: config.h:37-44) -- every other value goes to the
// reference's own renderer_from_config, so the same binary still runs the CPU renderers.
//
// What the class hands to the reference's code: `framebuffer` (vec3**, row 0 = image bottom) filled from the device film, so
// that sync_progress() and finalize() can call the reference's OWN calculate_luminance + output_to_file on it
// (helpers.h:146-168, renderer.h:24-55) -- the PPM this binary writes is the reference's film path applied to the GPU's
// film, which tests compare with the PPM the library's own tonemap writes.
//
// usage (cwd holds config.json, like main.exe):  wavefront_main [poll_ms]
#define main ref_main
#include "main.cpp"
#undef main

#include "b200pt.h"

class Wavefront : public Renderer
{
public:
    Wavefront(const std::string &config_json, const std::string &scene_json, camera cam, Config config)
        : Renderer(nullptr, cam, config)
    {
        char scene_path[1024], ppm_path[1024];
        check(b200pt_config_from_json(config_json.c_str(), &cfg, scene_path, sizeof(scene_path), ppm_path, sizeof(ppm_path)));
        cfg.async_render = 1; // start_render returns at once, like the reference's (renderer.h:553-603)
        check(b200pt_scene_create_from_json(scene_json.c_str(), cfg.width, cfg.height, cfg.device, &scene));
        sums.resize(3ull * cfg.width * cfg.height);
        completed = false;
    }
    void preprocess() override {}
    void start_render(std::chrono::high_resolution_clock::time_point program_start_time) override
    {
        render_start_time = std::chrono::high_resolution_clock::now();
        check(b200pt_render_start(scene, &cfg, &render));
        std::chrono::duration<double> setup = render_start_time - program_start_time;
        std::cout << "time taken to setup the rest and spawn threads " << setup.count() << std::endl;
    }
    void next_pixel_and_ray(int, ray &, int, int) override {}
    void compute(int) override {} // the work happens on the GPU(s)
    void sync_progress() override
    {
        // renderer.h:605-620: progress line, then the film so far through calculate_luminance + output_to_file
        check(b200pt_render_poll(render, 0, &progress));
        long num_samples_done = progress.samples_done;
        long num_samples_left = progress.samples_total - num_samples_done;
        print_out_progress(num_samples_done, num_samples_left, render_start_time);
        fetch_film();
        // `1 + num_samples_done / pixels` of the reference, with the samples the snapshot really holds (it is taken after
        // the poll, so it may be a batch ahead of `progress`)
        long n = 1 + (long)spp_in_film;
        float avg_luminance, max_luminance, total_luminance;
        calculate_luminance(framebuffer, film.width, film.height, n, film.width * film.height, max_luminance,
                            total_luminance, avg_luminance);
        output_to_file(output, framebuffer, film.width, film.height, n, max_luminance, film.exposure, film.gamma);
        previews++;
        completed = progress.done != 0;
    }
    bool is_done() override { return completed; }
    void finalize() override
    {
        // renderer.h:693-727
        check(b200pt_render_wait(render, &progress));
        auto t4 = std::chrono::high_resolution_clock::now();
        std::chrono::duration<double> elapsed = t4 - render_start_time;
        long total_bounces = progress.rays_reference; // what bounce_counts[] would hold (integrator.h:192,247)
        std::cout << "time taken to compute " << elapsed.count() << std::endl;
        float rate1 = film.total_pixels * config.samples / elapsed.count();
        float rate2 = total_bounces / elapsed.count();
        std::cout << "computed " << film.total_pixels * config.samples << " camera rays in " << elapsed.count() << "s, at "
                  << rate1 << " rays per second" << std::endl;
        std::cout << "computed " << total_bounces << " rays, at " << rate2 << " rays per second" << std::endl;
        int added_paths = 0;
        if (config.should_trace_paths)
            check(b200pt_render_write_paths(render, config.traced_paths_output_path.c_str(),
                                            config.traced_paths_2d_output_path.c_str(), &added_paths));
        std::cout << "added " << added_paths << " paths" << std::endl;
        fetch_film();
        float max_luminance, avg_luminance, total_luminance;
        calculate_luminance(framebuffer, film.width, film.height, config.samples, film.width * film.height, max_luminance,
                            total_luminance, avg_luminance);
        std::cout << "avg lum " << avg_luminance << std::endl;
        std::cout << "max lum " << max_luminance << std::endl;
        output_to_file(output, framebuffer, film.width, film.height, config.samples, max_luminance, film.exposure, film.gamma);
        std::cout << "wavefront previews " << previews << std::endl;
    }
    ~Wavefront()
    {
        b200pt_render_destroy(render);
        b200pt_scene_destroy(scene);
    }

private:
    static void check(int rc)
    {
        if (rc)
        {
            std::cerr << "b200pt: " << b200pt_last_error() << std::endl;
            exit(3);
        }
    }
    void fetch_film() // Renderer::framebuffer (renderer.h:141) <- the device film (all GPUs combined)
    {
        int64_t spp = 0;
        check(b200pt_render_read(render, sums.data(), &spp));
        spp_in_film = spp;
        for (int j = 0; j < cfg.height; j++)
            for (int i = 0; i < cfg.width; i++)
            {
                const float *q = &sums[3ull * ((size_t)j * cfg.width + i)];
                framebuffer[j][i] = vec3(q[0], q[1], q[2]);
            }
    }
    b200pt_config cfg{};
    b200pt_scene *scene = nullptr;
    b200pt_render *render = nullptr;
    b200pt_progress progress{};
    std::vector<float> sums;
    int64_t spp_in_film = 0;
    int previews = 0;
};

static std::string slurp_text(const std::string &path)
{
    std::ifstream f(path, std::ios::binary);
    std::stringstream ss;
    ss << f.rdbuf();
    return ss.str();
}

// main.cpp:108-168 with one changed line (the factory)
int main(int argc, char *argv[])
{
    const int poll_ms = argc > 1 ? atoi(argv[1]) : 500;
    std::string config_text = slurp_text("config.json");
    json jconfig = json::parse(config_text);
    Config config = Config(jconfig);
    s_film film = config.film;
    long min_camera_rays = config.samples * film.total_pixels;
    std::cout << "trace probability is " << config.trace_probability << std::endl;
    std::cout << "config read complete, rendering image at " << film.width << "x" << film.height << "==" << film.total_pixels
              << "px^2" << std::endl;
    std::cout << "with " << config.samples << " samples per pixel, that sets a minimum number of camera rays at "
              << min_camera_rays << "\n\n";
    std::cout << "reading scene data" << std::endl;
    std::string scene_text = slurp_text(config.scene_path);
    json scene = json::parse(scene_text);
    auto t1 = std::chrono::high_resolution_clock::now();
    const bool wavefront = jconfig.value("render_type", std::string()) == "wavefront";
    camera cam = setup_camera(scene["camera"], float(film.width) / float(film.height));
    Renderer *renderer;
    if (wavefront)
    {
        renderer = new Wavefront(config_text, scene_text, cam, config); // build_scene + BVH happen in the library
        std::cout << "selected and constructed Wavefront renderer" << std::endl;
    }
    else
    {
        World *world = build_scene(scene);
        world->config = config;
        renderer = renderer_from_config(world, cam, config);
    }
    auto t2 = std::chrono::high_resolution_clock::now();
    std::chrono::duration<double> elapsed_seconds = t2 - t1;
    std::cout << "time taken to build bvh " << elapsed_seconds.count() << std::endl;

    renderer->start_render(t2);
    while (!renderer->is_done())
    {
        renderer->sync_progress();
        std::this_thread::sleep_for(std::chrono::milliseconds(poll_ms));
    }
    std::cout << " done\n";
    renderer->finalize();
    return 0;
}
