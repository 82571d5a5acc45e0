// capi.cu -- the extern "C" layer (include/b200pt.h): scene upload, the render lifecycle that stands in for the
// reference's Renderer (renderer.h:114-150: start_render / sync_progress / is_done / finalize) and the
// closest-hit parity entry point.  Everything below the ABI runs on the GPU; there is no CPU fallback.
#include "../../include/b200pt.h"
#include "device/wavefront.cuh"
#include "host/bvh_build.h"
#include "host/config.h"
#include "host/flat_scene.h"
#include "host/scene_pack.h"

#include <cuda_runtime.h>

#include <cstdio>
#include <cstring>
#include <fstream>
#include <algorithm>
#include <cstdlib>
#include <map>
#include <memory>
#include <mutex>
#include <thread>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

using namespace b200pt;

namespace {

thread_local std::string g_last_error;

int fail(int code, const std::string &msg)
{
    g_last_error = msg;
    return code;
}

struct CudaError : std::runtime_error {
    explicit CudaError(const std::string &m) : std::runtime_error(m) {}
};
struct NoDevice : std::runtime_error {
    explicit NoDevice(const std::string &m) : std::runtime_error(m) {}
};

#define CK(expr)                                                                                                       \
    do {                                                                                                               \
        cudaError_t e_ = (expr);                                                                                       \
        if (e_ != cudaSuccess)                                                                                         \
            throw CudaError(std::string(#expr) + ": " + cudaGetErrorString(e_));                                       \
    } while (0)

template <typename F> int guarded(F &&f)
{
    try {
        f();
        return B200PT_OK;
    } catch (const NoDevice &e) {
        return fail(B200PT_ERR_NO_DEVICE, e.what());
    } catch (const CudaError &e) {
        return fail(B200PT_ERR_CUDA, e.what());
    } catch (const std::exception &e) {
        return fail(B200PT_ERR_INVALID, e.what());
    }
}

// Device and pinned-host blocks of destroyed handles are kept for the next handle that asks for the same size: a
// render allocates ~1.3 GB of wavefront state in 20 blocks, and cudaMalloc/cudaFree of those cost more than
// rendering a few samples per pixel.  Blocks enter the cache only after the owning stream has been synchronised.
// B200PT_CACHE_MB caps the cached device bytes (default 8192; 0 disables); b200pt_trim_cache() empties it.
class BlockCache {
  public:
    void *dev_get(int device, size_t bytes)
    {
        {
            std::lock_guard<std::mutex> g(mu_);
            auto it = dev_.find({device, bytes});
            if (it != dev_.end()) {
                void *p = it->second;
                dev_.erase(it);
                dev_bytes_ -= bytes;
                return p;
            }
        }
        void *p = nullptr;
        cudaError_t e = cudaMalloc(&p, bytes);
        if (e == cudaErrorMemoryAllocation) {
            cudaGetLastError();
            trim();
            e = cudaMalloc(&p, bytes);
        }
        if (e != cudaSuccess)
            throw CudaError(std::string("cudaMalloc: ") + cudaGetErrorString(e));
        return p;
    }
    void dev_put(int device, void *p, size_t bytes)
    {
        {
            std::lock_guard<std::mutex> g(mu_);
            if (dev_bytes_ + bytes <= cap()) {
                dev_.insert({{device, bytes}, p});
                dev_bytes_ += bytes;
                return;
            }
        }
        cudaFree(p);
    }
    void *pinned_get(size_t bytes)
    {
        {
            std::lock_guard<std::mutex> g(mu_);
            auto it = pinned_.find(bytes);
            if (it != pinned_.end()) {
                void *p = it->second;
                pinned_.erase(it);
                pinned_bytes_ -= bytes;
                return p;
            }
        }
        void *p = nullptr;
        CK(cudaHostAlloc(&p, bytes, cudaHostAllocDefault));
        return p;
    }
    void pinned_put(void *p, size_t bytes)
    {
        {
            std::lock_guard<std::mutex> g(mu_);
            if (pinned_bytes_ + bytes <= (cap() ? (size_t)1 << 30 : 0)) {
                pinned_.insert({bytes, p});
                pinned_bytes_ += bytes;
                return;
            }
        }
        cudaFreeHost(p);
    }
    void trim()
    {
        std::lock_guard<std::mutex> g(mu_);
        int cur = 0;
        cudaGetDevice(&cur);
        for (auto &kv : dev_) {
            cudaSetDevice(kv.first.first);
            cudaFree(kv.second);
        }
        cudaSetDevice(cur);
        for (auto &kv : pinned_)
            cudaFreeHost(kv.second);
        dev_.clear();
        pinned_.clear();
        dev_bytes_ = pinned_bytes_ = 0;
    }

  private:
    static size_t cap()
    {
        static const size_t c = [] {
            const char *e = std::getenv("B200PT_CACHE_MB");
            return (size_t)(e ? std::max(0ll, atoll(e)) : 8192ll) << 20;
        }();
        return c;
    }
    std::mutex mu_;
    std::multimap<std::pair<int, size_t>, void *> dev_;
    std::multimap<size_t, void *> pinned_;
    size_t dev_bytes_ = 0, pinned_bytes_ = 0;
};

// Device -> pinned staging -> caller memory in 4 MB pieces: the host threads copy piece i out of the staging buffer
// while the DMA engine is still filling the pieces behind it.  Caller memory has often never been touched (a fresh
// numpy array): page faults dominate the host copy, and they scale with threads.
cudaError_t copy_out_pipelined(void *dst, void *stage, const void *d_src, size_t bytes, cudaStream_t stream)
{
    const size_t piece = (size_t)4 << 20;
    const size_t n = (bytes + piece - 1) / piece;
    if (n <= 1) {
        cudaError_t e = cudaMemcpyAsync(stage, d_src, bytes, cudaMemcpyDeviceToHost, stream);
        if (e == cudaSuccess)
            e = cudaStreamSynchronize(stream);
        if (e == cudaSuccess)
            memcpy(dst, stage, bytes);
        return e;
    }
    std::vector<cudaEvent_t> ev(n, nullptr);
    cudaError_t err = cudaSuccess;
    size_t issued = 0;
    for (; issued < n && err == cudaSuccess; issued++) {
        const size_t off = issued * piece, len = std::min(piece, bytes - off);
        err = cudaEventCreateWithFlags(&ev[issued], cudaEventDisableTiming);
        if (err == cudaSuccess)
            err = cudaMemcpyAsync((char *)stage + off, (const char *)d_src + off, len, cudaMemcpyDeviceToHost, stream);
        if (err == cudaSuccess)
            err = cudaEventRecord(ev[issued], stream);
    }
    if (err == cudaSuccess) {
        const unsigned nt = (unsigned)std::min<size_t>(n, std::min(8u, std::max(1u, std::thread::hardware_concurrency())));
        std::vector<cudaError_t> errs(nt, cudaSuccess);
        std::vector<std::thread> pool;
        for (unsigned t = 0; t < nt; t++)
            pool.emplace_back([&, t] {
                for (size_t i = t; i < n; i += nt) {
                    const cudaError_t e = cudaEventSynchronize(ev[i]);
                    if (e != cudaSuccess) {
                        errs[t] = e;
                        return;
                    }
                    const size_t off = i * piece, len = std::min(piece, bytes - off);
                    memcpy((char *)dst + off, (const char *)stage + off, len);
                }
            });
        for (auto &th : pool)
            th.join();
        for (cudaError_t e : errs)
            if (e != cudaSuccess)
                err = e;
    } else {
        cudaStreamSynchronize(stream); // let the copies that were issued finish before the staging block is reused
    }
    for (cudaEvent_t e : ev)
        if (e)
            cudaEventDestroy(e);
    return err;
}

int sm_count(int device)
{
    static std::mutex mu;
    static std::map<int, int> known;
    std::lock_guard<std::mutex> g(mu);
    auto it = known.find(device);
    if (it != known.end())
        return it->second;
    int n = 0;
    CK(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device));
    return known[device] = n;
}

BlockCache &block_cache()
{
    static BlockCache *c = new BlockCache; // leaked on purpose: no CUDA calls from static destructors
    return *c;
}

struct DevBlock {
    void *p;
    size_t bytes;
};

template <typename T> T *dev_upload(const std::vector<T> &v, int device, std::vector<DevBlock> &owned)
{
    T *d = nullptr;
    size_t bytes = std::max<size_t>(v.size(), 1) * sizeof(T);
    d = (T *)block_cache().dev_get(device, bytes);
    owned.push_back({d, bytes});
    if (!v.empty())
        CK(cudaMemcpy(d, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
    return d;
}

void select_device(int device)
{
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        throw NoDevice(std::string("no CUDA device: ") + cudaGetErrorString(e) +
                       " (this backend has no CPU fallback)");
    if (device < 0 || device >= count)
        throw std::runtime_error("device ordinal out of range");
    CK(cudaSetDevice(device));
}

} // namespace

struct b200pt_scene {
    FlatScene flat;
    PackedScene packed;
    int device = 0;
    cudaStream_t stream = nullptr;
    pt::PtSceneView view{};
    std::vector<DevBlock> allocations;
    int n_media = 0;
    uint32_t shade_kinds = 1u << pt::SK_MISS; // shade queues that can ever be non-empty for this scene

    ~b200pt_scene()
    {
        if (device >= 0)
            cudaSetDevice(device);
        if (stream) {
            cudaStreamSynchronize(stream);
            cudaStreamDestroy(stream);
        }
        for (const DevBlock &b : allocations)
            block_cache().dev_put(device, b.p, b.bytes);
    }
};

namespace {

void upload_scene(b200pt_scene &sc)
{
    sc.packed = pack_scene(sc.flat);
    sc.n_media = sc.packed.n_media;
    for (const b200pt_instance &i : sc.flat.instances) {
        switch (sc.flat.materials[i.material].kind) {
        case B200PT_MAT_DIELECTRIC: sc.shade_kinds |= 1u << pt::SK_DIELECTRIC; break;
        case B200PT_MAT_DIFFUSE_LIGHT: sc.shade_kinds |= 1u << pt::SK_LIGHT; break;
        case B200PT_MAT_ISOTROPIC: sc.shade_kinds |= 1u << pt::SK_ISOTROPIC; break;
        default: sc.shade_kinds |= 1u << pt::SK_DIFFUSE; break;
        }
    }
    sc.view = sc.packed.host_view(sc.flat); // host pointers, replaced below when a device is attached
    if (sc.device < 0)
        return; // host-only handle: description and BVH can be inspected, nothing can be traced
    select_device(sc.device);
    CK(cudaStreamCreateWithFlags(&sc.stream, cudaStreamNonBlocking));
    auto up = [&](const auto &vec) { return dev_upload(vec, sc.device, sc.allocations); };
    pt::PtSceneView &v = sc.view;
    v.nodes = (up(sc.packed.bvh.nodes4));
    v.inst = (up(sc.packed.inst));
    v.fwd = (up(sc.packed.fwd));
    v.mats = (up(sc.flat.materials));
    v.tex = (up(sc.flat.textures));
    v.lights = (up(sc.flat.lights));
    v.emitters = (up(sc.packed.emitters));
    v.light_boxes = (up(sc.packed.light_boxes));
    v.images = (up(sc.packed.images));
    v.texels = (up(sc.packed.texels));
    v.small = (up(sc.packed.small));
    v.perlin_ranvec = (up(sc.packed.perlin_ranvec));
    v.perlin_perm = (up(sc.packed.perlin_perm));
}

void need_device(const b200pt_scene *sc)
{
    if (sc->device < 0)
        throw NoDevice("scene handle was created host-only (device = -1); this backend has no CPU fallback");
    select_device(sc->device);
}

} // namespace

struct b200pt_render {
    b200pt_scene *scene = nullptr;
    b200pt_config cfg{};
    pt::RenderParams rp{};
    pt::Wavefront wf{};
    pt::Film film{};
    pt::LaunchCfg lc{};
    std::vector<DevBlock> allocations;
    float *d_lum = nullptr;   // [2]
    uint8_t *d_rgb = nullptr; // H*W*3
    float *d_rgbf = nullptr;  // H*W*3 floats: the film repacked for b200pt_render_read (allocated on first use)
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    int batch_samples = 1;
    int batches_total = 0, batches_done = 0;
    int n_rows = 0;
    double seconds = 0.0;
    int64_t spp_done = 0;
    // config.profile: per-class launch counts, event pairs of the launches enqueued since the last sync
    b200pt_profile prof{};
    struct Timed {
        int cls;
        cudaEvent_t a, b;
    };
    std::vector<Timed> pending;
    std::vector<cudaEvent_t> event_pool;

    cudaEvent_t get_event()
    {
        if (!event_pool.empty()) {
            cudaEvent_t e = event_pool.back();
            event_pool.pop_back();
            return e;
        }
        cudaEvent_t e = nullptr;
        cudaEventCreate(&e);
        return e;
    }

    ~b200pt_render()
    {
        if (scene)
            cudaSetDevice(scene->device);
        for (auto &t : pending) {
            cudaEventDestroy(t.a);
            cudaEventDestroy(t.b);
        }
        for (cudaEvent_t e : event_pool)
            cudaEventDestroy(e);
        if (scene && scene->stream)
            cudaStreamSynchronize(scene->stream); // nothing in flight may still touch a block that enters the cache
        for (const DevBlock &b : allocations)
            block_cache().dev_put(scene ? scene->device : 0, b.p, b.bytes);
        if (ev0)
            cudaEventDestroy(ev0);
        if (ev1)
            cudaEventDestroy(ev1);
    }
};

namespace {

template <typename T> T *render_alloc(b200pt_render &r, size_t count, bool zero = false)
{
    const size_t bytes = std::max<size_t>(count, 1) * sizeof(T);
    T *p = (T *)block_cache().dev_get(r.scene->device, bytes);
    r.allocations.push_back({p, bytes});
    if (zero)
        CK(cudaMemsetAsync(p, 0, bytes, r.scene->stream));
    return p;
}

// one kernel launch (or a short group of one class), counted and -- with config.profile == 1 -- bracketed by events
template <typename F> void launch_class(b200pt_render &r, int cls, int n_launches, F &&f)
{
    cudaStream_t st = r.scene->stream;
    r.prof.launches[cls] += n_launches;
    if (r.cfg.profile != 1) {
        f();
        return;
    }
    b200pt_render::Timed t{cls, r.get_event(), r.get_event()};
    CK(cudaEventRecord(t.a, st));
    f();
    CK(cudaEventRecord(t.b, st));
    r.pending.push_back(t);
}

void collect_kernel_times(b200pt_render &r)
{
    for (auto &t : r.pending) {
        float ms = 0.f;
        CK(cudaEventSynchronize(t.b));
        CK(cudaEventElapsedTime(&ms, t.a, t.b));
        r.prof.ms[t.cls] += ms;
        r.event_pool.push_back(t.a);
        r.event_pool.push_back(t.b);
    }
    r.pending.clear();
}

void enqueue_batch(b200pt_render &r)
{
    cudaStream_t st = r.scene->stream;
    const pt::PtSceneView &s = r.scene->view;
    CK(cudaMemsetAsync(r.wf.counters, 0, sizeof(int32_t) * pt::CNT_ROW * r.n_rows, st));
    launch_class(r, B200PT_K_RAYGEN, 1, [&] { pt::launch_raygen(s, r.rp, r.wf, r.lc, st); });
    const bool nee = r.cfg.integrator == B200PT_INTEGRATOR_INEEPT;
    const int n_bounces = nee ? r.cfg.max_bounces : r.cfg.max_bounces + 1;
    for (int b = 0; b < n_bounces; b++) {
        launch_class(r, B200PT_K_EXTEND, 1, [&] { pt::launch_extend(s, r.rp, r.wf, b, r.lc, st); });
        // one launch per material kind the scene has (a queue of an absent kind stays empty)
        const uint32_t kinds = r.scene->shade_kinds;
        launch_class(r, B200PT_K_SHADE, __builtin_popcount(kinds), [&] {
            for (int k = 0; k < pt::SK_COUNT; k++)
                if (kinds & (1u << k))
                    pt::launch_shade(s, r.rp, r.wf, b, k, r.lc, st);
        });
        if (r.wf.sort_on) {
            const bool more = b + 1 < n_bounces;
            launch_class(r, B200PT_K_SORT, (nee ? 3 : 0) + (more ? 3 : 0), [&] {
                if (nee)
                    pt::launch_sort_queue(r.wf, r.wf.key_nee, r.wf.q_raw_nee, r.wf.q_nee, b * pt::CNT_ROW + pt::CNT_NEE, r.lc, st);
                if (more)
                    pt::launch_sort_queue(r.wf, r.wf.key_ext, r.wf.q_raw_ext, r.wf.q_ext[(b + 1) & 1],
                                          (b + 1) * pt::CNT_ROW + pt::CNT_EXT, r.lc, st);
            });
        }
        if (nee)
            launch_class(r, B200PT_K_NEE, 1, [&] { pt::launch_nee(s, r.rp, r.wf, b, r.lc, st); });
    }
    launch_class(r, B200PT_K_ACCUMULATE, 2, [&] { pt::launch_accumulate(r.rp, r.wf, r.film, n_bounces, st); });
    CK(cudaGetLastError());
}

} // namespace

// ---------------------------------------------------------------------------------------------------------------

extern "C" {

const char *b200pt_last_error(void) { return g_last_error.c_str(); }

const char *b200pt_version(void) { return "b200pt 0.1 (sm_100a, wavefront)"; }

int b200pt_config_from_json(const char *config_json, b200pt_config *out, char *scene_path, size_t scene_cap,
                            char *ppm_path, size_t ppm_cap)
{
    if (!config_json || !out)
        return fail(B200PT_ERR_INVALID, "null argument");
    return guarded([&] {
        ParsedConfig pc = parse_config(Json::parse(config_json));
        *out = pc.config;
        if (scene_path && scene_cap) {
            strncpy(scene_path, pc.scene_path.c_str(), scene_cap - 1);
            scene_path[scene_cap - 1] = 0;
        }
        if (ppm_path && ppm_cap) {
            strncpy(ppm_path, pc.ppm_output_path.c_str(), ppm_cap - 1);
            ppm_path[ppm_cap - 1] = 0;
        }
    });
}

// The parsed document of a large scene is ~100 000 small heap blocks; they are freed on a helper thread while the
// caller's thread packs and uploads (joined before the entry point returns).
#ifndef PT_DOM_ASYNC_BYTES
#define PT_DOM_ASYNC_BYTES (64 << 10)
#endif
struct DomRelease {
    std::thread worker;
    DomRelease(Json &&doc, size_t text_bytes)
    {
        if (text_bytes >= (size_t)PT_DOM_ASYNC_BYTES) // a small document is freed faster than a thread starts
            worker = std::thread([d = std::move(doc)]() mutable { Json gone(std::move(d)); });
    }
    ~DomRelease()
    {
        if (worker.joinable())
            worker.join();
    }
};

int b200pt_scene_create_from_json(const char *scene_json, int32_t film_width, int32_t film_height, int32_t device,
                                  b200pt_scene **out)
{
    if (!scene_json || !out || film_width <= 0 || film_height <= 0)
        return fail(B200PT_ERR_INVALID, "bad argument");
    *out = nullptr;
    return guarded([&] {
        std::unique_ptr<b200pt_scene> sc(new b200pt_scene);
        sc->device = device;
        Json doc = Json::parse(scene_json);
        sc->flat = compile_scene(doc, film_width, film_height);
        DomRelease release(std::move(doc), strlen(scene_json));
        upload_scene(*sc);
        *out = sc.release();
    });
}

int b200pt_scene_create_from_file(const char *scene_path, int32_t film_width, int32_t film_height, int32_t device,
                                  b200pt_scene **out)
{
    if (!scene_path || !out || film_width <= 0 || film_height <= 0)
        return fail(B200PT_ERR_INVALID, "bad argument");
    *out = nullptr;
    return guarded([&] {
        std::ifstream f(scene_path, std::ios::binary);
        if (!f.good())
            throw std::runtime_error(std::string("cannot open scene file '") + scene_path + "'");
        std::stringstream ss;
        ss << f.rdbuf();
        const std::string path(scene_path);
        const size_t slash = path.find_last_of('/');
        const std::string dir = slash == std::string::npos ? std::string(".") : path.substr(0, slash);
        std::unique_ptr<b200pt_scene> sc(new b200pt_scene);
        sc->device = device;
        const std::string text = ss.str();
        Json doc = Json::parse(text);
        sc->flat = compile_scene(doc, film_width, film_height, dir);
        DomRelease release(std::move(doc), text.size());
        upload_scene(*sc);
        *out = sc.release();
    });
}

int b200pt_scene_create(const b200pt_scene_desc *desc, int32_t device, b200pt_scene **out)
{
    if (!desc || !out)
        return fail(B200PT_ERR_INVALID, "null argument");
    *out = nullptr;
    return guarded([&] {
        std::unique_ptr<b200pt_scene> sc(new b200pt_scene);
        sc->device = device;
        sc->flat = FlatScene::from_desc(*desc);
        upload_scene(*sc);
        *out = sc.release();
    });
}

int b200pt_scene_get_desc(const b200pt_scene *scene, b200pt_scene_desc *out)
{
    if (!scene || !out)
        return fail(B200PT_ERR_INVALID, "null argument");
    *out = scene->flat.desc();
    return B200PT_OK;
}

int b200pt_scene_bvh_info(const b200pt_scene *scene, int32_t *n_nodes, int32_t *max_depth, float bounds[6])
{
    if (!scene)
        return fail(B200PT_ERR_INVALID, "null argument");
    if (n_nodes)
        *n_nodes = (int32_t)scene->packed.bvh.nodes4.size();
    if (max_depth)
        *max_depth = scene->packed.bvh.max_depth4;
    if (bounds)
        memcpy(bounds, scene->packed.bvh.bounds, sizeof(float) * 6);
    return B200PT_OK;
}

void b200pt_scene_destroy(b200pt_scene *scene) { delete scene; }

int b200pt_trace_batch_device(b200pt_scene *scene, const void *d_rays, size_t n, const void *d_medium_xi, void *d_hits,
                              int32_t repeat, float *ms)
{
    if (!scene || (!d_rays && n))
        return fail(B200PT_ERR_INVALID, "null argument");
    return guarded([&] {
        need_device(scene);
        cudaEvent_t e0 = nullptr, e1 = nullptr;
        if (ms) {
            CK(cudaEventCreate(&e0));
            CK(cudaEventCreate(&e1));
            CK(cudaEventRecord(e0, scene->stream));
        }
        for (int i = 0; i < (repeat < 1 ? 1 : repeat); i++)
            pt::launch_trace_batch(scene->view, (const float *)d_rays, n, (const float *)d_medium_xi,
                                   (b200pt_hit *)d_hits, nullptr, scene->stream);
        CK(cudaGetLastError());
        if (ms) {
            CK(cudaEventRecord(e1, scene->stream));
            CK(cudaEventSynchronize(e1));
            CK(cudaEventElapsedTime(ms, e0, e1));
            cudaEventDestroy(e0);
            cudaEventDestroy(e1);
        }
    });
}

int b200pt_trace_batch(b200pt_scene *scene, const float *rays, size_t n, const float *medium_xi, b200pt_hit *hits)
{
    if (!scene || ((!rays || !hits) && n))
        return fail(B200PT_ERR_INVALID, "null argument");
    return guarded([&] {
        need_device(scene);
        if (n == 0)
            return;
        float *d_rays = nullptr, *d_xi = nullptr;
        b200pt_hit *d_hits = nullptr;
        struct Guard {
            float *&a, *&b;
            b200pt_hit *&c;
            ~Guard()
            {
                cudaFree(a);
                cudaFree(b);
                cudaFree(c);
            }
        } guard{d_rays, d_xi, d_hits};
        CK(cudaMalloc(&d_rays, n * 8 * sizeof(float)));
        CK(cudaMalloc(&d_hits, n * sizeof(b200pt_hit)));
        CK(cudaMemcpyAsync(d_rays, rays, n * 8 * sizeof(float), cudaMemcpyHostToDevice, scene->stream));
        if (medium_xi) {
            CK(cudaMalloc(&d_xi, n * sizeof(float)));
            CK(cudaMemcpyAsync(d_xi, medium_xi, n * sizeof(float), cudaMemcpyHostToDevice, scene->stream));
        }
        pt::launch_trace_batch(scene->view, d_rays, n, d_xi, d_hits, nullptr, scene->stream);
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(hits, d_hits, n * sizeof(b200pt_hit), cudaMemcpyDeviceToHost, scene->stream));
        CK(cudaStreamSynchronize(scene->stream));
    });
}

int b200pt_trace_count(b200pt_scene *scene, const float *rays, size_t n, uint64_t counts[3])
{
    if (!scene || (!rays && n) || !counts)
        return fail(B200PT_ERR_INVALID, "null argument");
    return guarded([&] {
        need_device(scene);
        counts[0] = counts[1] = counts[2] = 0;
        if (n == 0)
            return;
        float *d_rays = nullptr;
        unsigned long long *d_counts = nullptr;
        struct Guard {
            float *&a;
            unsigned long long *&b;
            ~Guard()
            {
                cudaFree(a);
                cudaFree(b);
            }
        } guard{d_rays, d_counts};
        CK(cudaMalloc(&d_rays, n * 8 * sizeof(float)));
        CK(cudaMalloc(&d_counts, 3 * sizeof(unsigned long long)));
        CK(cudaMemsetAsync(d_counts, 0, 3 * sizeof(unsigned long long), scene->stream));
        CK(cudaMemcpyAsync(d_rays, rays, n * 8 * sizeof(float), cudaMemcpyHostToDevice, scene->stream));
        pt::launch_trace_batch(scene->view, d_rays, n, nullptr, nullptr, d_counts, scene->stream);
        CK(cudaGetLastError());
        unsigned long long h[3];
        CK(cudaMemcpyAsync(h, d_counts, sizeof(h), cudaMemcpyDeviceToHost, scene->stream));
        CK(cudaStreamSynchronize(scene->stream));
        counts[0] = h[0];
        counts[1] = h[1];
        counts[2] = h[2];
    });
}

int b200pt_philox4x32_10(int32_t device, const uint32_t *counters_keys, size_t n, uint32_t *out)
{
    if ((!counters_keys || !out) && n)
        return fail(B200PT_ERR_INVALID, "null argument");
    return guarded([&] {
        select_device(device);
        if (n == 0)
            return;
        uint32_t *d_in = nullptr, *d_out = nullptr;
        struct Guard {
            uint32_t *&a, *&b;
            ~Guard()
            {
                cudaFree(a);
                cudaFree(b);
            }
        } guard{d_in, d_out};
        CK(cudaMalloc(&d_in, n * 6 * sizeof(uint32_t)));
        CK(cudaMalloc(&d_out, n * 4 * sizeof(uint32_t)));
        CK(cudaMemcpy(d_in, counters_keys, n * 6 * sizeof(uint32_t), cudaMemcpyHostToDevice));
        pt::launch_philox(d_in, n, d_out, nullptr);
        CK(cudaGetLastError());
        CK(cudaMemcpy(out, d_out, n * 4 * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    });
}

int b200pt_render_start(b200pt_scene *scene, const b200pt_config *config, b200pt_render **out)
{
    if (!scene || !config || !out)
        return fail(B200PT_ERR_INVALID, "null argument");
    *out = nullptr;
    return guarded([&] {
        const b200pt_config &c = *config;
        if (c.width <= 0 || c.height <= 0 || c.samples <= 0)
            throw std::runtime_error("config: film size and samples must be positive");
        if (c.integrator != B200PT_INTEGRATOR_INEEPT && c.integrator != B200PT_INTEGRATOR_RPT)
            throw std::runtime_error("config: unsupported integrator");
        if (c.max_bounces < 0 || c.max_bounces > 4096)
            throw std::runtime_error("config: max_bounces out of range");
        // integrator.h:265 divides the light sum by light_samples: with 0 every sample that reaches a surface is NaN
        if (c.integrator == B200PT_INTEGRATOR_INEEPT && (c.light_samples < 1 || c.light_samples > 1024))
            throw std::runtime_error("config: light_samples must be in [1, 1024] for \"iterative nee path tracing\"");
        // global sample indices are keyed into Philox as 56 bits (pt_rng.h)
        if (c.sample_offset < 0 || c.sample_offset > (1ll << PT_RNG_MAX_SAMPLE_BITS) - (long long)c.samples)
            throw std::runtime_error("config: sample_offset must lie in [0, 2^56 - samples]");
        if (c.device != scene->device)
            throw std::runtime_error("config.device differs from the device the scene was uploaded to");
        need_device(scene);
        std::unique_ptr<b200pt_render> r(new b200pt_render);
        r->scene = scene;
        r->cfg = c;
        const long long npix = (long long)c.width * c.height;
        if (npix > (1ll << 30))
            throw std::runtime_error("film too large");
        // batch sizing: enough paths in flight to fill 148 SMs many times over, small enough to stay a few GB
        long long target_slots = 32ll << 20;
        if (const char *e = std::getenv("B200PT_BATCH_SLOTS")) // tuning aid: paths in flight per wavefront batch
            if (std::atoll(e) > 0)
                target_slots = std::atoll(e);
        int bs = c.batch_samples > 0 ? c.batch_samples : (int)std::max<long long>(1, target_slots / npix);
        bs = std::min(bs, c.samples);
        // slots (and slots x light_samples, the NEE index space) are 32-bit indices
        // (k_nee pads the samples of a vertex to a power-of-two lane group when light_samples <= 32)
        long long per_slot = 1;
        if (c.integrator == B200PT_INTEGRATOR_INEEPT && c.light_samples > 1) {
            per_slot = c.light_samples;
            if (c.light_samples <= 32)
                for (per_slot = 1; per_slot < c.light_samples;)
                    per_slot *= 2;
        }
        while ((long long)bs * npix * per_slot > (1ll << 30) && bs > 1)
            bs--;
        if ((long long)bs * npix * per_slot >= (1ll << 31))
            throw std::runtime_error("config: film x light_samples exceeds the 32-bit index space of one wavefront batch");
        r->batch_samples = bs;
        r->batches_total = (c.samples + bs - 1) / bs;
        const size_t cap = (size_t)bs * (size_t)npix;

        pt::RenderParams &rp = r->rp;
        rp.width = c.width;
        rp.height = c.height;
        rp.integrator = c.integrator;
        rp.max_bounces = c.max_bounces;
        rp.light_samples = c.integrator == B200PT_INTEGRATOR_INEEPT ? c.light_samples : 0;
        rp.light_group = rp.light_samples >= 1 && rp.light_samples <= 32 ? (int32_t)(rp.light_samples > 1 ? per_slot : 1) : 0;
        rp.russian_roulette = c.russian_roulette;
        rp.only_direct = c.only_direct_illumination;
        rp.corrected = c.physically_corrected ? 1 : 0;
        if (rp.corrected && c.profile == 2)
            throw std::runtime_error("config: profile = 2 (work counters) is not built for the physically_corrected mode");
        rp.normal_offset = c.normal_offset;
        rp.seed_lo = (uint32_t)(c.seed & 0xffffffffull);
        rp.seed_hi = (uint32_t)(c.seed >> 32);

        pt::Wavefront &wf = r->wf;
        wf.npix = (int32_t)npix;
        wf.capacity = (int32_t)cap;
        {
            // divide-by-npix constants (Granlund & Montgomery): m = ceil(2^(31+l) / d), l = ceil(log2 d)
            const uint32_t d = (uint32_t)npix;
            if (d == 1) {
                wf.npix_mul = 0;
                wf.npix_shr = 0;
            } else {
                uint32_t l = 0;
                while ((1ull << l) < d)
                    l++;
                const unsigned long long pw = 1ull << (31 + l);
                wf.npix_mul = (uint32_t)((pw + d - 1) / d);
                wf.npix_shr = l - 1;
            }
        }
        wf.ray = render_alloc<float4>(*r, 2 * cap);
        wf.thr = render_alloc<float4>(*r, 2 * cap);
        wf.hit = render_alloc<float4>(*r, 2 * cap);
        wf.acc = render_alloc<float4>(*r, 2 * cap);
        for (int k = 0; k < 2; k++)
            wf.q_ext[k] = render_alloc<int32_t>(*r, cap);
        for (int k = 0; k < pt::SK_COUNT; k++)
            wf.q_kind[k] = render_alloc<int32_t>(*r, cap);
        wf.q_nee = render_alloc<int32_t>(*r, cap);
        // coherence sort between bounces: scenes that walk the BVH (B200PT_SORT=0 switches it off, for A/B runs)
        wf.sort_on = scene->view.n_small == 0 ? 1 : 0;
        if (const char *e = std::getenv("B200PT_SORT"))
            wf.sort_on = wf.sort_on && atoi(e) != 0;
        if (wf.sort_on) {
            wf.q_raw_ext = render_alloc<int32_t>(*r, cap);
            wf.q_raw_nee = render_alloc<int32_t>(*r, cap);
            wf.key_ext = render_alloc<uint32_t>(*r, cap);
            wf.key_nee = render_alloc<uint32_t>(*r, cap);
            wf.sort_hist = render_alloc<uint32_t>(*r, (size_t)PT_SORT_BUCKETS + PT_SORT_BUCKETS / PT_SORT_CHUNK);
        }
        r->n_rows = c.max_bounces + 3;
        wf.counters = render_alloc<int32_t>(*r, (size_t)pt::CNT_ROW * r->n_rows, true);
        wf.ctl = render_alloc<pt::BatchCtl>(*r, 1, true);
        wf.trace_counts = render_alloc<unsigned long long>(*r, 8, true);
        pt::BatchCtl ctl{};
        ctl.sample_base = c.sample_offset;
        ctl.nsamp = bs;
        ctl.remaining = c.samples - bs;
        ctl.batch_samples = bs;
        CK(cudaMemcpyAsync(wf.ctl, &ctl, sizeof(ctl), cudaMemcpyHostToDevice, scene->stream));
        CK(cudaStreamSynchronize(scene->stream)); // `ctl` is a stack variable

        if (c.trace_probability > 0.f) {
            // should_trace_paths (config.h:122-131): the expected number of recorded paths is avg_number_of_paths
            const double expect = (double)c.trace_probability * (double)c.samples * (double)npix;
            wf.path_max = c.trace_max_paths > 0 ? c.trace_max_paths : (int32_t)std::max(1024.0, std::min(4.0 * expect, 4.0e6));
            wf.path_stride = c.max_bounces + 1;
            wf.path_prob = c.trace_probability;
            wf.path_id = render_alloc<int32_t>(*r, cap);
            wf.path_hdr = render_alloc<int4>(*r, (size_t)wf.path_max, true);
            wf.path_pts = render_alloc<float4>(*r, (size_t)wf.path_max * wf.path_stride, true);
            wf.path_count = render_alloc<int32_t>(*r, 1, true);
        }
        r->film.accum = render_alloc<float4>(*r, (size_t)npix, true);
        r->film.clamp = c.stats_clamp;
        if (c.stats_clamp > 0.f) {
            r->film.csum = render_alloc<float4>(*r, (size_t)npix, true);
            r->film.csq = render_alloc<float4>(*r, (size_t)npix, true);
        }
        r->d_lum = render_alloc<float>(*r, 2, true);
        r->d_rgb = render_alloc<uint8_t>(*r, (size_t)npix * 3);
        CK(cudaEventCreate(&r->ev0));
        CK(cudaEventCreate(&r->ev1));

        r->lc.threads = 256;
        r->lc.count_work = c.profile == 2;
        r->lc.blocks = sm_count(scene->device) * 8;
        *out = r.release();
    });
}

int b200pt_render_poll(b200pt_render *r, int32_t max_batches, b200pt_progress *progress)
{
    if (!r)
        return fail(B200PT_ERR_INVALID, "null argument");
    return guarded([&] {
        need_device(r->scene);
        cudaStream_t st = r->scene->stream;
        int todo = r->batches_total - r->batches_done;
        if (max_batches > 0 && todo > max_batches)
            todo = max_batches;
        if (todo > 0) {
            CK(cudaEventRecord(r->ev0, st));
            for (int b = 0; b < todo; b++)
                enqueue_batch(*r);
            CK(cudaEventRecord(r->ev1, st));
            CK(cudaEventSynchronize(r->ev1));
            float ms = 0.f;
            CK(cudaEventElapsedTime(&ms, r->ev0, r->ev1));
            r->seconds += ms * 1e-3;
            r->batches_done += todo;
            collect_kernel_times(*r);
        }
        pt::BatchCtl ctl;
        CK(cudaMemcpyAsync(&ctl, r->wf.ctl, sizeof(ctl), cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        r->spp_done = (int64_t)(ctl.samples_done / (unsigned long long)r->wf.npix);
        if (progress) {
            progress->samples_done = (int64_t)ctl.samples_done;
            progress->samples_total = (int64_t)r->cfg.samples * r->wf.npix;
            progress->rays_extension = (int64_t)ctl.rays_extension;
            progress->rays_shadow = (int64_t)ctl.rays_shadow;
            progress->rays_reference = (int64_t)ctl.rays_reference;
            progress->seconds = r->seconds;
            progress->done = r->batches_done >= r->batches_total;
        }
    });
}

int b200pt_render_read(b200pt_render *r, float *rgb_sums, int64_t *spp_done)
{
    if (!r || !rgb_sums)
        return fail(B200PT_ERR_INVALID, "null argument");
    return guarded([&] {
        need_device(r->scene);
        // float4 sums -> packed rgb on the device, one D2H copy into pinned staging, one memcpy to the caller
        const size_t npix = (size_t)r->wf.npix, bytes = npix * 3 * sizeof(float);
        if (!r->d_rgbf)
            r->d_rgbf = render_alloc<float>(*r, npix * 3);
        pt::launch_pack_rgb(r->film.accum, (int)npix, r->d_rgbf, r->scene->stream);
        void *stage = block_cache().pinned_get(bytes);
        cudaError_t e = copy_out_pipelined(rgb_sums, stage, r->d_rgbf, bytes, r->scene->stream);
        block_cache().pinned_put(stage, bytes);
        CK(e);
        if (spp_done)
            *spp_done = r->spp_done;
    });
}

int b200pt_render_read_stats(b200pt_render *r, float *clamped_sum, float *clamped_sumsq)
{
    if (!r || !clamped_sum || !clamped_sumsq)
        return fail(B200PT_ERR_INVALID, "null argument");
    if (!r->film.csum)
        return fail(B200PT_ERR_STATE, "render was started without config.stats_clamp");
    return guarded([&] {
        need_device(r->scene);
        const size_t npix = (size_t)r->wf.npix;
        std::vector<float4> a(npix), b(npix);
        CK(cudaMemcpyAsync(a.data(), r->film.csum, npix * sizeof(float4), cudaMemcpyDeviceToHost, r->scene->stream));
        CK(cudaMemcpyAsync(b.data(), r->film.csq, npix * sizeof(float4), cudaMemcpyDeviceToHost, r->scene->stream));
        CK(cudaStreamSynchronize(r->scene->stream));
        for (size_t i = 0; i < npix; i++) {
            clamped_sum[3 * i + 0] = a[i].x;
            clamped_sum[3 * i + 1] = a[i].y;
            clamped_sum[3 * i + 2] = a[i].z;
            clamped_sumsq[3 * i + 0] = b[i].x;
            clamped_sumsq[3 * i + 1] = b[i].y;
            clamped_sumsq[3 * i + 2] = b[i].z;
        }
    });
}

int b200pt_render_device_buffer(b200pt_render *r, void **d_accum, size_t *bytes)
{
    if (!r || !d_accum)
        return fail(B200PT_ERR_INVALID, "null argument");
    *d_accum = r->film.accum;
    if (bytes)
        *bytes = (size_t)r->wf.npix * sizeof(float4);
    return B200PT_OK;
}

int b200pt_render_write(b200pt_render *r, const float *rgba_sums, int64_t spp)
{
    if (!r || !rgba_sums)
        return fail(B200PT_ERR_INVALID, "null argument");
    return guarded([&] {
        need_device(r->scene);
        CK(cudaMemcpyAsync(r->film.accum, rgba_sums, (size_t)r->wf.npix * sizeof(float4), cudaMemcpyHostToDevice,
                           r->scene->stream));
        CK(cudaStreamSynchronize(r->scene->stream));
        r->spp_done = spp;
    });
}

int b200pt_render_tonemap(b200pt_render *r, int64_t n_samples, uint8_t *rgb, float *max_luminance, float *avg_luminance)
{
    if (!r || n_samples <= 0)
        return fail(B200PT_ERR_INVALID, "bad argument");
    return guarded([&] {
        need_device(r->scene);
        cudaStream_t st = r->scene->stream;
        const int npix = r->wf.npix;
        // col /= float(samples) is a multiplication by float(1.0 / t) (vec3.h:181-189)
        const float inv_n = (float)(1.0 / (float)n_samples);
        pt::launch_luminance(r->film.accum, npix, inv_n, r->d_lum, st);
        float lum[2];
        CK(cudaMemcpyAsync(lum, r->d_lum, sizeof(lum), cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        // renderer.h:37: col *= 16 + exposure
        pt::launch_tonemap(r->film.accum, r->cfg.width, r->cfg.height, inv_n, 16 + r->cfg.exposure, lum[0], r->d_rgb, st);
        CK(cudaGetLastError());
        if (rgb)
            CK(cudaMemcpyAsync(rgb, r->d_rgb, (size_t)npix * 3, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        if (max_luminance)
            *max_luminance = lum[0];
        if (avg_luminance) // helpers.h:167: total / (total_pixels * n_samples) (sic: divided by n twice)
            *avg_luminance = lum[1] / ((float)npix * (float)n_samples);
    });
}

int b200pt_render_write_ppm(b200pt_render *r, int64_t n_samples, const char *path)
{
    if (!r || !path)
        return fail(B200PT_ERR_INVALID, "null argument");
    std::vector<uint8_t> rgb((size_t)r->wf.npix * 3);
    int rc = b200pt_render_tonemap(r, n_samples, rgb.data(), nullptr, nullptr);
    if (rc)
        return rc;
    std::ofstream f(path, std::ios::binary);
    if (!f.good())
        return fail(B200PT_ERR_INVALID, std::string("cannot open ") + path);
    f << "P6\n" << r->cfg.width << " " << r->cfg.height << "\n255\n";
    f.write((const char *)rgb.data(), (std::streamsize)rgb.size());
    return f.good() ? B200PT_OK : fail(B200PT_ERR_INVALID, std::string("write failed: ") + path);
}

namespace {

struct TracedPath {
    int32_t pixel;
    int64_t sample;
    std::vector<float> pts; // 3 per vertex
};

// the device-side path store, ordered by (sample, pixel): the order does not depend on scheduling
std::vector<TracedPath> fetch_paths(b200pt_render &r)
{
    std::vector<TracedPath> out;
    if (!r.wf.path_id)
        return out;
    cudaStream_t st = r.scene->stream;
    int32_t n = 0;
    CK(cudaMemcpyAsync(&n, r.wf.path_count, sizeof(n), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    n = std::min(n, r.wf.path_max);
    if (n <= 0)
        return out;
    const int stride = r.wf.path_stride;
    std::vector<int4> hdr((size_t)n);
    std::vector<float4> pts((size_t)n * stride);
    CK(cudaMemcpyAsync(hdr.data(), r.wf.path_hdr, hdr.size() * sizeof(int4), cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(pts.data(), r.wf.path_pts, pts.size() * sizeof(float4), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    out.resize((size_t)n);
    for (int i = 0; i < n; i++) {
        out[i].pixel = hdr[i].x;
        out[i].sample = (int64_t)(uint32_t)hdr[i].y | ((int64_t)hdr[i].z << 32);
        const int len = std::min(hdr[i].w, stride);
        for (int k = 0; k < len; k++) {
            const float4 &p = pts[(size_t)i * stride + k];
            out[i].pts.insert(out[i].pts.end(), {p.x, p.y, p.z});
        }
    }
    std::sort(out.begin(), out.end(), [](const TracedPath &a, const TracedPath &b) {
        return a.sample != b.sample ? a.sample < b.sample : a.pixel < b.pixel;
    });
    return out;
}

// camera::project (camera.h:49-109), float arithmetic as written there: dot() is x*x + y*y + z*z with no zero to
// start from, which decides the SIGN of a zero denominator for a point in the camera's plane (t = +-inf)
float dot3(const float a[3], const float b[3]) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }

bool camera_project(const b200pt_camera &c, const float pt[3], float &x, float &y)
{
    float d[3], l[3];
    for (int k = 0; k < 3; k++) {
        d[k] = c.origin[k] - pt[k];     // ray(point, origin - point)
        l[k] = c.lower_left[k] - pt[k]; // lower_left_corner - r.origin()
    }
    const float t = dot3(l, c.w) / dot3(d, c.w);
    if (t < 0.0) {
        const float lim = 1.0e+10f;
        return !(fabsf(pt[0]) > lim || fabsf(pt[1]) > lim || fabsf(pt[2]) > lim);
    }
    float p[3];
    for (int k = 0; k < 3; k++)
        p[k] = pt[k] + t * d[k] - c.lower_left[k];
    x = dot3(c.horizontal, p) / dot3(c.horizontal, c.horizontal);
    y = dot3(c.vertical, p) / dot3(c.vertical, c.vertical);
    return true;
}

} // namespace

int b200pt_camera_project(const b200pt_scene *scene, const float *points, size_t n, float *xy, int32_t *ok)
{
    if (!scene || ((!points || !xy || !ok) && n))
        return fail(B200PT_ERR_INVALID, "null argument");
    for (size_t i = 0; i < n; i++) {
        float x = 0.f, y = 0.f; // the caller's initial values survive a behind-the-film point (renderer.h:76-78)
        ok[i] = camera_project(scene->flat.camera, points + 3 * i, x, y) ? 1 : 0;
        xy[2 * i] = x;
        xy[2 * i + 1] = y;
    }
    return B200PT_OK;
}

int b200pt_render_paths(b200pt_render *r, int32_t capacity, int32_t *n_paths, int32_t *vertices_per_path, int32_t *pixel,
                        int64_t *sample, int32_t *length, float *points)
{
    if (!r || !n_paths)
        return fail(B200PT_ERR_INVALID, "null argument");
    return guarded([&] {
        need_device(r->scene);
        std::vector<TracedPath> paths = fetch_paths(*r);
        *n_paths = (int32_t)paths.size();
        const int stride = r->wf.path_id ? r->wf.path_stride : 0;
        if (vertices_per_path)
            *vertices_per_path = stride;
        for (int i = 0; i < (int)paths.size() && i < capacity; i++) {
            if (pixel)
                pixel[i] = paths[i].pixel;
            if (sample)
                sample[i] = paths[i].sample;
            if (length)
                length[i] = (int32_t)(paths[i].pts.size() / 3);
            if (points)
                std::copy(paths[i].pts.begin(), paths[i].pts.end(), points + (size_t)i * stride * 3);
        }
    });
}

int b200pt_render_write_paths(b200pt_render *r, const char *path_3d, const char *path_2d, int32_t *n_paths)
{
    if (!r || !path_3d || !path_2d)
        return fail(B200PT_ERR_INVALID, "null argument");
    return guarded([&] {
        need_device(r->scene);
        std::vector<TracedPath> paths = fetch_paths(*r);
        std::ofstream o3(path_3d), o2(path_2d);
        if (!o3.good() || !o2.good())
            throw std::runtime_error(std::string("cannot open ") + (o3.good() ? path_2d : path_3d));
        const b200pt_camera &cam = r->scene->flat.camera;
        int added = 0;
        for (const TracedPath &p : paths) { // handle_traced_paths (renderer.h:57-96)
            if (p.pts.empty())
                continue;
            added++;
            for (size_t k = 0; k + 2 < p.pts.size(); k += 3) {
                o3 << p.pts[k] << ',' << p.pts[k + 1] << ',' << p.pts[k + 2] << '\n';
                float x = 0, y = 0;
                const bool hit_scene = camera_project(cam, &p.pts[k], x, y);
                if (0.0 < x && 0.0 < y && x <= 1.0 && y <= 1.0 && hit_scene)
                    o2 << x << ',' << y << std::endl;
                if (!hit_scene)
                    o2 << x << ',' << y << '!' << std::endl;
            }
            o3 << std::endl;
            o2 << std::endl;
        }
        if (n_paths)
            *n_paths = added;
    });
}

int b200pt_config_string(const char *config_json, const char *key, char *out, size_t cap)
{
    if (!config_json || !key || !out || !cap)
        return fail(B200PT_ERR_INVALID, "null argument");
    return guarded([&] {
        Json j = Json::parse(config_json);
        std::string v = j.contains(key) && j[key].is_string() ? j[key].as_string() : std::string();
        strncpy(out, v.c_str(), cap - 1);
        out[cap - 1] = 0;
    });
}

int b200pt_render_profile(b200pt_render *r, b200pt_profile *out)
{
    if (!r || !out)
        return fail(B200PT_ERR_INVALID, "null argument");
    return guarded([&] {
        need_device(r->scene);
        unsigned long long h[8];
        CK(cudaMemcpyAsync(h, r->wf.trace_counts, sizeof(h), cudaMemcpyDeviceToHost, r->scene->stream));
        CK(cudaStreamSynchronize(r->scene->stream));
        *out = r->prof;
        out->ext_rays = h[0], out->ext_boxes = h[1], out->ext_prims = h[2];
        out->nee_rays = h[3], out->nee_boxes = h[4], out->nee_prims = h[5];
        out->nee_traced = h[6];
        out->nee_queued = h[7];
    });
}

void b200pt_render_destroy(b200pt_render *r) { delete r; }

void b200pt_trim_cache(void) { block_cache().trim(); }

} // extern "C"
