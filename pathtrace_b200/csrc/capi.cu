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
#include <atomic>
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
// B200PT_CACHE_MB caps the cached device bytes (default 49152 = two full-size wavefront batches; 0 disables);
// b200pt_trim_cache() empties it.
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
            return (size_t)(e ? std::max(0ll, atoll(e)) : 49152ll) << 20;
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

// The host side of a scene (compiled description + packed arrays + BVH) is built once and shared by the replicas a
// multi-GPU render uploads to its other devices.
struct SceneHost {
    FlatScene flat;
    PackedScene packed;
};

struct b200pt_scene {
    std::shared_ptr<SceneHost> host;
    FlatScene &flat;
    PackedScene &packed;
    int device = 0;
    cudaStream_t stream = nullptr;
    pt::PtSceneView view{};
    std::vector<DevBlock> allocations;
    int n_media = 0;
    uint32_t shade_kinds = 1u << pt::SK_MISS; // shade queues that can ever be non-empty for this scene
    std::mutex replica_mu;
    std::map<int, std::unique_ptr<b200pt_scene>> replicas; // device -> upload of the same host scene (multi-GPU renders)
    // One reference for the caller's handle and one per live render: b200pt_scene_destroy on a scene that still has renders
    // only drops the caller's reference, the memory goes when the last render is destroyed (a garbage collector may well
    // finalize the scene object first).
    std::atomic<int> refs{1};

    b200pt_scene() : host(std::make_shared<SceneHost>()), flat(host->flat), packed(host->packed) {}
    explicit b200pt_scene(std::shared_ptr<SceneHost> h) : host(std::move(h)), flat(host->flat), packed(host->packed) {}

    ~b200pt_scene()
    {
        replicas.clear();
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

void scene_unref(b200pt_scene *sc)
{
    if (sc && sc->refs.fetch_sub(1) == 1)
        delete sc;
}

void upload_scene(b200pt_scene &sc, bool pack = true)
{
    if (pack)
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

// the upload of `sc`'s host scene on `device`: sc itself, or a replica made on first use
b200pt_scene *scene_on_device(b200pt_scene *sc, int device)
{
    if (device == sc->device)
        return sc;
    std::lock_guard<std::mutex> g(sc->replica_mu);
    auto it = sc->replicas.find(device);
    if (it != sc->replicas.end())
        return it->second.get();
    std::unique_ptr<b200pt_scene> rep(new b200pt_scene(sc->host));
    rep->device = device;
    upload_scene(*rep, false);
    b200pt_scene *out = rep.get();
    sc->replicas[device] = std::move(rep);
    return out;
}

void need_device(const b200pt_scene *sc)
{
    if (sc->device < 0)
        throw NoDevice("scene handle was created host-only (device = -1); this backend has no CPU fallback");
    select_device(sc->device);
}

} // namespace

struct Shard { // one device's share of a render: its scene upload, stream, wavefront state and film
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
    // Whatever is enqueued on the stream as a unit -- the launches of one batch, the commands of a film snapshot --
    // is enqueued under this mutex: the worker thread of an asynchronous render and the caller's thread share the stream.
    std::mutex mu;
    // progress as last published (by poll in the paced mode, by the worker thread in the asynchronous one)
    struct Progress {
        unsigned long long samples_done = 0, rays_extension = 0, rays_shadow = 0, rays_reference = 0;
        double seconds = 0.0;
        int batches_done = 0;
    } prog;
    std::mutex pmu;
    // asynchronous mode: two batches in flight; ctl is copied to pinned memory behind each
    pt::BatchCtl *h_ctl = nullptr; // [2], pinned
    cudaEvent_t ev_batch[2] = {nullptr, nullptr};
    cudaEvent_t ev_first = nullptr;

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

    ~Shard()
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
        for (cudaEvent_t e : ev_batch)
            if (e)
                cudaEventDestroy(e);
        if (ev_first)
            cudaEventDestroy(ev_first);
        if (h_ctl)
            block_cache().pinned_put(h_ctl, 2 * sizeof(pt::BatchCtl));
    }

    void publish(const pt::BatchCtl &ctl, double secs, int done)
    {
        std::lock_guard<std::mutex> g(pmu);
        prog.samples_done = ctl.samples_done;
        prog.rays_extension = ctl.rays_extension;
        prog.rays_shadow = ctl.rays_shadow;
        prog.rays_reference = ctl.rays_reference;
        prog.seconds = secs;
        prog.batches_done = done;
    }
    Progress progress()
    {
        std::lock_guard<std::mutex> g(pmu);
        return prog;
    }
};

namespace {

template <typename T> T *render_alloc(Shard &r, size_t count, bool zero = false)
{
    const size_t bytes = std::max<size_t>(count, 1) * sizeof(T);
    T *p = (T *)block_cache().dev_get(r.scene->device, bytes);
    r.allocations.push_back({p, bytes});
    if (zero)
        CK(cudaMemsetAsync(p, 0, bytes, r.scene->stream));
    return p;
}

// one kernel launch (or a short group of one class), counted and -- with config.profile == 1 -- bracketed by events
template <typename F> void launch_class(Shard &r, int cls, int n_launches, F &&f)
{
    cudaStream_t st = r.scene->stream;
    r.prof.launches[cls] += n_launches;
    if (r.cfg.profile != 1) {
        f();
        return;
    }
    Shard::Timed t{cls, r.get_event(), r.get_event()};
    CK(cudaEventRecord(t.a, st));
    f();
    CK(cudaEventRecord(t.b, st));
    r.pending.push_back(t);
}

void collect_kernel_times(Shard &r)
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

void enqueue_batch(Shard &r)
{
    cudaStream_t st = r.scene->stream;
    const pt::PtSceneView &s = r.scene->view;
    CK(cudaMemsetAsync(r.wf.counters, 0, sizeof(int32_t) * pt::CNT_ROW * r.n_rows, st));
    launch_class(r, B200PT_K_RAYGEN, 1, [&] { pt::launch_raygen(s, r.rp, r.wf, r.lc, st); });
    const bool nee = r.cfg.integrator == B200PT_INTEGRATOR_INEEPT;
    const int n_bounces = nee ? r.cfg.max_bounces : r.cfg.max_bounces + 1;
    if (n_bounces == 0) // the radiance sums are written by the bounce-0 shade kernels: with no bounce at all they are zero
        CK(cudaMemsetAsync(r.wf.acc, 0, sizeof(float4) * 2 * (size_t)r.wf.capacity, st));
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
            // (the camera vertices' NEE queue is in pixel order already, but sorting it too is worth it: 901 vs 894 Mpaths/s)
            const bool nee_sorted = nee;
            launch_class(r, B200PT_K_SORT, (nee_sorted ? 3 : 0) + (more ? 3 : 0), [&] {
                if (nee_sorted)
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

namespace {

// One device's share: validates the configuration, sizes the wavefront batch and allocates its state.  `c.samples`
// and `c.sample_offset` are THIS shard's; `scene` is the upload on the shard's device.
std::unique_ptr<Shard> make_shard(b200pt_scene *scene, const b200pt_config &c)
{
    {
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
        std::unique_ptr<Shard> r(new Shard);
        r->scene = scene;
        r->cfg = c;
        const long long npix = (long long)c.width * c.height;
        if (npix > (1ll << 30))
            throw std::runtime_error("film too large");
        // batch sizing: enough paths in flight to fill 148 SMs many times over, small enough to stay a few GB
        // 128 Mi slots (~24 GB of path state and queues): measured on book2_final 4K, 32 / 64 / 128 Mi slots give
        // 865 / 888 / 900 Mpaths/s -- the launches of the late bounces are short, and a larger batch makes their ramp-up and
        // tail weigh less; the memory is there (180 GB)
        long long target_slots = 128ll << 20;
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
        return r;
    }
}


} // namespace

// The handle of include/b200pt.h: one Shard per device.  shards[0] lives on the device the films are combined on.
struct b200pt_render {
    b200pt_scene *scene = nullptr;
    b200pt_config cfg{}; // the caller's configuration (ALL samples)
    std::vector<std::unique_ptr<Shard>> shards;
    int npix = 0;
    bool async = false;
    std::vector<std::thread> workers;
    std::atomic<bool> stop{false};
    std::atomic<int> workers_running{0};
    std::mutex err_mu;
    std::string worker_error;
    // snapshots / totals on shards[0]'s device: [0] film, [1] clamped sums, [2] clamped squares (allocated on first use)
    std::mutex snap_mu;
    float4 *d_total[3] = {nullptr, nullptr, nullptr};
    std::vector<float4 *> staged;      // per shard: landing buffer on device 0 when its memory cannot be peer-loaded
    std::vector<char> peer_ok;         // per shard: device 0 kernels may dereference its pointers
    std::vector<cudaEvent_t> ev_sync;  // per shard: "film is at a batch boundary" marker on its stream
    cudaEvent_t ev_combined = nullptr; // on stream 0: the combine kernel has read every film
    pt::BatchCtl *h_snap = nullptr;    // pinned, one per shard: batch control at the snapshot point
    int64_t spp_done = 0;

    Shard &s0() { return *shards[0]; }
    cudaStream_t stream0() { return shards[0]->scene->stream; }
    int device0() { return shards[0]->scene->device; }

    void join_workers()
    {
        stop.store(true);
        for (std::thread &t : workers)
            if (t.joinable())
                t.join();
        workers.clear();
    }

    ~b200pt_render()
    {
        join_workers();
        if (!shards.empty()) {
            cudaSetDevice(device0());
            cudaStreamSynchronize(stream0());
        }
        for (cudaEvent_t e : ev_sync)
            if (e)
                cudaEventDestroy(e);
        if (ev_combined)
            cudaEventDestroy(ev_combined);
        if (h_snap)
            block_cache().pinned_put(h_snap, sizeof(pt::BatchCtl) * 8);
        // shards[0] owns d_total / staged (render_alloc on it); destroy the others first: their streams are idle after
        // the joins and the synchronize of their own destructors
        while (shards.size() > 1)
            shards.pop_back();
        shards.clear();
        if (scene_ref_held)
            scene_unref(scene); // may be the last reference (the caller destroyed the scene handle first)
    }
    bool scene_ref_held = false;
};

namespace {

void set_worker_error(b200pt_render &r, const std::string &msg)
{
    std::lock_guard<std::mutex> g(r.err_mu);
    if (r.worker_error.empty())
        r.worker_error = msg;
}

// Asynchronous mode: the thread that drives one device (Renderer::start_render's workers, renderer.h:555-597).  Two
// batches are kept in flight; behind each the batch control block is copied to pinned memory, and progress is published
// when that copy has landed.
void worker_main(b200pt_render *r, Shard *s)
{
    try {
        CK(cudaSetDevice(s->scene->device));
        cudaStream_t st = s->scene->stream;
        int published = 0;
        auto retire = [&](int b) { // batch b has finished: publish its totals
            CK(cudaEventSynchronize(s->ev_batch[b & 1]));
            float ms = 0.f;
            CK(cudaEventElapsedTime(&ms, s->ev_first, s->ev_batch[b & 1]));
            s->publish(s->h_ctl[b & 1], ms * 1e-3, b + 1);
            published = b + 1;
        };
        for (int b = 0; b < s->batches_total && !r->stop.load(); b++) {
            if (b >= 2)
                retire(b - 2);
            std::lock_guard<std::mutex> g(s->mu);
            if (b == 0)
                CK(cudaEventRecord(s->ev_first, st));
            enqueue_batch(*s);
            CK(cudaMemcpyAsync(&s->h_ctl[b & 1], s->wf.ctl, sizeof(pt::BatchCtl), cudaMemcpyDeviceToHost, st));
            CK(cudaEventRecord(s->ev_batch[b & 1], st));
            s->batches_done = b + 1; // enqueued
        }
        for (int b = published; b < s->batches_done; b++)
            retire(b);
        collect_kernel_times(*s);
    } catch (const std::exception &e) {
        set_worker_error(*r, e.what());
    }
    r->workers_running.fetch_sub(1);
}

// Paced mode: the caller's thread advances every shard by up to `max_batches` (0 = all that is left), one batch per
// shard at a time so that all devices have work queued, then waits for all of them.
void run_paced(b200pt_render &r, int max_batches)
{
    std::vector<int> todo(r.shards.size());
    int most = 0;
    for (size_t i = 0; i < r.shards.size(); i++) {
        Shard &s = *r.shards[i];
        todo[i] = s.batches_total - s.batches_done;
        if (max_batches > 0 && todo[i] > max_batches)
            todo[i] = max_batches;
        most = std::max(most, todo[i]);
    }
    if (most == 0)
        return;
    for (size_t i = 0; i < r.shards.size(); i++)
        if (todo[i] > 0) {
            CK(cudaSetDevice(r.shards[i]->scene->device));
            CK(cudaEventRecord(r.shards[i]->ev0, r.shards[i]->scene->stream));
        }
    for (int b = 0; b < most; b++)
        for (size_t i = 0; i < r.shards.size(); i++)
            if (b < todo[i]) {
                CK(cudaSetDevice(r.shards[i]->scene->device));
                enqueue_batch(*r.shards[i]);
            }
    for (size_t i = 0; i < r.shards.size(); i++)
        if (todo[i] > 0) {
            CK(cudaSetDevice(r.shards[i]->scene->device));
            CK(cudaEventRecord(r.shards[i]->ev1, r.shards[i]->scene->stream));
        }
    for (size_t i = 0; i < r.shards.size(); i++) {
        Shard &s = *r.shards[i];
        if (todo[i] == 0)
            continue;
        CK(cudaSetDevice(s.scene->device));
        CK(cudaEventSynchronize(s.ev1));
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, s.ev0, s.ev1));
        s.seconds += ms * 1e-3;
        s.batches_done += todo[i];
        collect_kernel_times(s);
        pt::BatchCtl ctl;
        CK(cudaMemcpyAsync(&ctl, s.wf.ctl, sizeof(ctl), cudaMemcpyDeviceToHost, s.scene->stream));
        CK(cudaStreamSynchronize(s.scene->stream));
        s.publish(ctl, s.seconds, s.batches_done);
    }
}

void fill_progress(b200pt_render &r, b200pt_progress *progress)
{
    Shard::Progress tot;
    bool done = true;
    for (auto &sp : r.shards) {
        const Shard::Progress p = sp->progress();
        tot.samples_done += p.samples_done;
        tot.rays_extension += p.rays_extension;
        tot.rays_shadow += p.rays_shadow;
        tot.rays_reference += p.rays_reference;
        tot.seconds = std::max(tot.seconds, p.seconds); // the devices render side by side
        done = done && p.batches_done >= sp->batches_total;
    }
    r.spp_done = (int64_t)(tot.samples_done / (unsigned long long)r.npix);
    if (progress) {
        progress->samples_done = (int64_t)tot.samples_done;
        progress->samples_total = (int64_t)r.cfg.samples * r.npix;
        progress->rays_extension = (int64_t)tot.rays_extension;
        progress->rays_shadow = (int64_t)tot.rays_shadow;
        progress->rays_reference = (int64_t)tot.rays_reference;
        progress->seconds = tot.seconds;
        progress->done = done ? 1 : 0;
    }
}

void check_worker_error(b200pt_render &r)
{
    std::lock_guard<std::mutex> g(r.err_mu);
    if (!r.worker_error.empty())
        throw CudaError("render worker: " + r.worker_error);
}

} // namespace

namespace {

// A stable view of plane `plane` (0 film, 1 clamped sums, 2 clamped squares) of the WHOLE render on device 0, valid for
// work enqueued on stream 0 after this call.  One shard rendering in the paced mode: its own buffer, no copy.  Otherwise
// the per-device buffers are summed into d_total[plane] by k_combine -- each device's stream is marked at a batch
// boundary (under the shard mutexes: a batch is enqueued as a unit), stream 0 waits for the marks, the kernel reads
// the other devices' films through NVLink peer loads (a peer copy into a landing buffer where peer access is not
// available), and the other streams wait for it before their next batch may touch their film.  With lum != null the
// luminance reduce of the sum (helpers.h:146-168) is fused into the same pass (single shard: its own reduce kernel).
// *samples receives the camera samples the view holds.
const float4 *snapshot_plane(b200pt_render &r, int plane, float inv_n, float *lum, unsigned long long *samples)
{
    Shard &a = r.s0();
    const size_t npix = (size_t)r.npix;
    auto plane_of = [&](Shard &s) -> float4 * { return plane == 0 ? s.film.accum : (plane == 1 ? s.film.csum : s.film.csq); };
    CK(cudaSetDevice(r.device0()));
    cudaStream_t st0 = r.stream0();
    if (r.shards.size() == 1 && !r.async) {
        if (lum)
            pt::launch_luminance(plane_of(a), (int)npix, inv_n, lum, st0);
        if (samples)
            *samples = a.progress().samples_done;
        return plane_of(a);
    }
    if (!r.d_total[plane])
        r.d_total[plane] = render_alloc<float4>(a, npix);
    const size_t n = r.shards.size();
    if (!r.h_snap)
        r.h_snap = (pt::BatchCtl *)block_cache().pinned_get(sizeof(pt::BatchCtl) * 8);
    std::vector<std::unique_lock<std::mutex>> locks;
    for (auto &sp : r.shards)
        locks.emplace_back(sp->mu);
    pt::FilmSources fs{};
    fs.n = (int)n;
    for (size_t i = 0; i < n; i++) {
        Shard &s = *r.shards[i];
        CK(cudaSetDevice(s.scene->device));
        CK(cudaMemcpyAsync(&r.h_snap[i], s.wf.ctl, sizeof(pt::BatchCtl), cudaMemcpyDeviceToHost, s.scene->stream));
        if (i > 0) {
            CK(cudaEventRecord(r.ev_sync[i], s.scene->stream));
            CK(cudaSetDevice(r.device0()));
            CK(cudaStreamWaitEvent(st0, r.ev_sync[i], 0));
            if (!r.peer_ok[i]) {
                if (!r.staged[i])
                    r.staged[i] = render_alloc<float4>(a, npix);
                CK(cudaMemcpyPeerAsync(r.staged[i], r.device0(), plane_of(s), s.scene->device, npix * sizeof(float4), st0));
            }
        }
        fs.src[i] = i > 0 && !r.peer_ok[i] ? r.staged[i] : plane_of(s);
    }
    CK(cudaSetDevice(r.device0()));
    pt::launch_combine(fs, (int)npix, r.d_total[plane], inv_n, lum, st0);
    CK(cudaGetLastError());
    CK(cudaEventRecord(r.ev_combined, st0));
    for (size_t i = 1; i < n; i++) {
        CK(cudaSetDevice(r.shards[i]->scene->device));
        CK(cudaStreamWaitEvent(r.shards[i]->scene->stream, r.ev_combined, 0));
    }
    locks.clear();
    if (samples) {
        unsigned long long tot = 0;
        for (size_t i = 0; i < n; i++) {
            // the ctl copy sits in front of the mark (shard 0: in front of the combine) on its stream
            CK(cudaEventSynchronize(i == 0 ? r.ev_combined : r.ev_sync[i]));
            tot += r.h_snap[i].samples_done;
        }
        *samples = tot;
    }
    CK(cudaSetDevice(r.device0()));
    return r.d_total[plane];
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

void b200pt_scene_destroy(b200pt_scene *scene) { scene_unref(scene); }

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
        if (c.samples <= 0)
            throw std::runtime_error("config: film size and samples must be positive");
        // the devices: `device` alone, or device_ids[0 .. device_count) with the films combined on the first
        std::vector<int> devs;
        if (c.device_count > 1) {
            if (c.device_count > 8)
                throw std::runtime_error("config: device_count must be at most 8");
            int count = 0;
            cudaError_t e = cudaGetDeviceCount(&count);
            if (e != cudaSuccess || count == 0)
                throw NoDevice(std::string("no CUDA device: ") + cudaGetErrorString(e) + " (this backend has no CPU fallback)");
            for (int k = 0; k < c.device_count; k++) {
                const int d = c.device_ids[k];
                if (d < 0 || d >= count)
                    throw std::runtime_error("config: device_ids[" + std::to_string(k) + "] = " + std::to_string(d) +
                                             " is not a CUDA device of this machine (" + std::to_string(count) + " visible)");
                // B200PT_ALLOW_DUPLICATE_DEVICES (tests on a one-GPU box): several shards may share a device
                if (std::find(devs.begin(), devs.end(), d) != devs.end() && !std::getenv("B200PT_ALLOW_DUPLICATE_DEVICES"))
                    throw std::runtime_error("config: device_ids lists a device twice");
                devs.push_back(d);
            }
            if (c.samples < c.device_count)
                throw std::runtime_error("config: fewer samples than devices");
            if (scene->device < 0)
                throw NoDevice("scene handle was created host-only (device = -1); this backend has no CPU fallback");
        } else {
            devs.push_back(c.device);
        }
        std::unique_ptr<b200pt_render> r(new b200pt_render);
        r->scene = scene;
        scene->refs.fetch_add(1);
        r->scene_ref_held = true;
        r->cfg = c;
        r->async = c.async_render != 0;
        const int n = (int)devs.size();
        // spp sharding on GLOBAL sample indices (the reference's own split, renderer.h:352-360): contiguous shares
        // whose sizes differ by at most one -- the film is the same whatever the number of devices
        const int base = c.samples / n, extra = c.samples % n;
        long long first = c.sample_offset;
        for (int i = 0; i < n; i++) {
            b200pt_config sc = c;
            sc.device = devs[i];
            sc.device_count = 0;
            sc.samples = base + (i < extra ? 1 : 0);
            sc.sample_offset = first;
            first += sc.samples;
            if (n > 1 && c.trace_max_paths > 0)
                sc.trace_max_paths = (c.trace_max_paths + n - 1) / n;
            b200pt_scene *on_dev = n > 1 ? scene_on_device(scene, devs[i]) : scene;
            r->shards.push_back(make_shard(on_dev, sc));
        }
        r->npix = r->shards[0]->wf.npix;
        r->staged.assign(n, nullptr);
        r->peer_ok.assign(n, 1);
        r->ev_sync.assign(n, nullptr);
        CK(cudaSetDevice(r->device0()));
        CK(cudaEventCreateWithFlags(&r->ev_combined, cudaEventDisableTiming));
        for (int i = 1; i < n; i++) {
            int can = 0;
            CK(cudaSetDevice(r->device0()));
            if (devs[i] == r->device0())
                can = 2; // the same device (test hook): plain loads
            else
                CK(cudaDeviceCanAccessPeer(&can, r->device0(), devs[i]));
            if (std::getenv("B200PT_NO_PEER_LOADS"))
                can = 0;
            if (can == 1) {
                cudaError_t e = cudaDeviceEnablePeerAccess(devs[i], 0);
                if (e == cudaErrorPeerAccessAlreadyEnabled)
                    cudaGetLastError();
                else if (e != cudaSuccess) {
                    cudaGetLastError();
                    can = 0;
                }
            }
            r->peer_ok[i] = (char)(can != 0);
            CK(cudaSetDevice(devs[i]));
            CK(cudaEventCreateWithFlags(&r->ev_sync[i], cudaEventDisableTiming));
        }
        if (r->async) {
            for (auto &sp : r->shards) {
                CK(cudaSetDevice(sp->scene->device));
                sp->h_ctl = (pt::BatchCtl *)block_cache().pinned_get(2 * sizeof(pt::BatchCtl));
                CK(cudaEventCreate(&sp->ev_first));
                for (cudaEvent_t &e : sp->ev_batch)
                    CK(cudaEventCreate(&e));
            }
            r->workers_running.store(n);
            for (auto &sp : r->shards)
                r->workers.emplace_back(worker_main, r.get(), sp.get());
        }
        CK(cudaSetDevice(r->device0()));
        *out = r.release();
    });
}

int b200pt_render_poll(b200pt_render *r, int32_t max_batches, b200pt_progress *progress)
{
    if (!r)
        return fail(B200PT_ERR_INVALID, "null argument");
    return guarded([&] {
        need_device(r->shards[0]->scene);
        if (r->async)
            check_worker_error(*r); // the workers render; this only reports
        else
            run_paced(*r, max_batches);
        fill_progress(*r, progress);
    });
}

int b200pt_render_wait(b200pt_render *r, b200pt_progress *progress)
{
    if (!r)
        return fail(B200PT_ERR_INVALID, "null argument");
    return guarded([&] {
        need_device(r->shards[0]->scene);
        if (r->async) {
            for (std::thread &t : r->workers)
                if (t.joinable())
                    t.join();
            r->workers.clear();
            check_worker_error(*r);
        } else {
            run_paced(*r, 0);
        }
        fill_progress(*r, progress);
    });
}

int b200pt_render_read(b200pt_render *r, float *rgb_sums, int64_t *spp_done)
{
    if (!r || !rgb_sums)
        return fail(B200PT_ERR_INVALID, "null argument");
    return guarded([&] {
        need_device(r->shards[0]->scene);
        std::lock_guard<std::mutex> sg(r->snap_mu);
        Shard &a = r->s0();
        cudaStream_t st = r->stream0();
        // the film of ALL devices (k_combine) -> packed rgb on device 0, one D2H copy into pinned staging, one memcpy
        const size_t npix = (size_t)r->npix, bytes = npix * 3 * sizeof(float);
        if (!a.d_rgbf)
            a.d_rgbf = render_alloc<float>(a, npix * 3);
        unsigned long long samples = 0;
        const float4 *film = snapshot_plane(*r, 0, 0.f, nullptr, &samples);
        pt::launch_pack_rgb(film, (int)npix, a.d_rgbf, st);
        void *stage = block_cache().pinned_get(bytes);
        cudaError_t e = copy_out_pipelined(rgb_sums, stage, a.d_rgbf, bytes, st);
        block_cache().pinned_put(stage, bytes);
        CK(e);
        if (!r->async && r->shards.size() == 1)
            samples = (unsigned long long)r->spp_done * npix; // paced: as of the last poll (or b200pt_render_write)
        if (spp_done)
            *spp_done = (int64_t)(samples / npix);
    });
}

int b200pt_render_read_stats(b200pt_render *r, float *clamped_sum, float *clamped_sumsq)
{
    if (!r || !clamped_sum || !clamped_sumsq)
        return fail(B200PT_ERR_INVALID, "null argument");
    if (!r->shards[0]->film.csum)
        return fail(B200PT_ERR_STATE, "render was started without config.stats_clamp");
    return guarded([&] {
        need_device(r->shards[0]->scene);
        std::lock_guard<std::mutex> sg(r->snap_mu);
        const size_t npix = (size_t)r->npix;
        std::vector<float4> a(npix), b(npix);
        const float4 *cs = snapshot_plane(*r, 1, 0.f, nullptr, nullptr);
        CK(cudaMemcpyAsync(a.data(), cs, npix * sizeof(float4), cudaMemcpyDeviceToHost, r->stream0()));
        const float4 *cq = snapshot_plane(*r, 2, 0.f, nullptr, nullptr);
        CK(cudaMemcpyAsync(b.data(), cq, npix * sizeof(float4), cudaMemcpyDeviceToHost, r->stream0()));
        CK(cudaStreamSynchronize(r->stream0()));
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
    if (r->shards.size() != 1)
        return fail(B200PT_ERR_STATE, "b200pt_render_device_buffer: a multi-device render combines its films itself");
    *d_accum = r->shards[0]->film.accum;
    if (bytes)
        *bytes = (size_t)r->npix * sizeof(float4);
    return B200PT_OK;
}

int b200pt_render_write(b200pt_render *r, const float *rgba_sums, int64_t spp)
{
    if (!r || !rgba_sums)
        return fail(B200PT_ERR_INVALID, "null argument");
    if (r->async || r->shards.size() != 1)
        return fail(B200PT_ERR_STATE, "b200pt_render_write needs a single-device render in the paced mode");
    return guarded([&] {
        need_device(r->shards[0]->scene);
        CK(cudaMemcpyAsync(r->shards[0]->film.accum, rgba_sums, (size_t)r->npix * sizeof(float4), cudaMemcpyHostToDevice,
                           r->stream0()));
        CK(cudaStreamSynchronize(r->stream0()));
        r->spp_done = spp;
    });
}

int b200pt_render_tonemap(b200pt_render *r, int64_t n_samples, uint8_t *rgb, float *max_luminance, float *avg_luminance)
{
    if (!r || n_samples <= 0)
        return fail(B200PT_ERR_INVALID, "bad argument");
    return guarded([&] {
        need_device(r->shards[0]->scene);
        std::lock_guard<std::mutex> sg(r->snap_mu);
        Shard &a = r->s0();
        cudaStream_t st = r->stream0();
        const int npix = r->npix;
        // col /= float(samples) is a multiplication by float(1.0 / t) (vec3.h:181-189)
        const float inv_n = (float)(1.0 / (float)n_samples);
        // calculate_luminance on the film of all devices: fused into the pass that sums them (k_combine)
        const float4 *film = snapshot_plane(*r, 0, inv_n, a.d_lum, nullptr);
        float lum[2];
        CK(cudaMemcpyAsync(lum, a.d_lum, sizeof(lum), cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        // renderer.h:37: col *= 16 + exposure
        pt::launch_tonemap(film, r->cfg.width, r->cfg.height, inv_n, 16 + r->cfg.exposure, lum[0], a.d_rgb, st);
        CK(cudaGetLastError());
        if (rgb)
            CK(cudaMemcpyAsync(rgb, a.d_rgb, (size_t)npix * 3, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        if (max_luminance)
            *max_luminance = lum[0];
        if (avg_luminance) // helpers.h:167: total / (total_pixels * n_samples) (sic: divided by n twice)
            *avg_luminance = lum[1] / ((float)npix * (float)n_samples);
    });
}

int b200pt_tonemap_static(int32_t device, const void *d_rgba_sums, int32_t width, int32_t height, int64_t n_samples,
                          float exposure, void *d_rgb, void *stream, float *max_luminance, float *avg_luminance)
{
    // no cudaMalloc / cudaFree / allocating routine anywhere on this path: the scratch is kernels.cu's static array
    if (!d_rgba_sums || !d_rgb || width <= 0 || height <= 0 || n_samples <= 0)
        return fail(B200PT_ERR_INVALID, "b200pt_tonemap_static: bad argument");
    const long long npix = (long long)width * (long long)height;
    if (npix > (long long)PT_STATIC_MAX_PIXELS)
        return fail(B200PT_ERR_INVALID, "b200pt_tonemap_static: film of " + std::to_string(width) + " x " +
                                            std::to_string(height) + " = " + std::to_string(npix) +
                                            " pixels exceeds the static scratch of " +
                                            std::to_string((long long)PT_STATIC_MAX_PIXELS) +
                                            " pixels (use b200pt_render_tonemap, which allocates)");
    return guarded([&] {
        select_device(device);
        // col /= float(samples) is a multiplication by float(1.0 / t) (vec3.h:181-189); col *= 16 + exposure (renderer.h:37)
        const float inv_n = (float)(1.0 / (float)n_samples);
        float lum[2] = {0.f, 0.f};
        const bool want = max_luminance || avg_luminance;
        CK(pt::launch_tonemap_static(device, (const float4 *)d_rgba_sums, width, height, inv_n, 16 + exposure,
                                     (uint8_t *)d_rgb, want ? lum : nullptr, (cudaStream_t)stream));
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
    std::vector<uint8_t> rgb((size_t)r->npix * 3);
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
std::vector<TracedPath> fetch_paths(Shard &r)
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

// the traced paths of all devices, ordered by (sample, pixel): the shards hold disjoint global sample ranges
std::vector<TracedPath> fetch_all_paths(b200pt_render &r)
{
    std::vector<TracedPath> all;
    for (auto &sp : r.shards) {
        CK(cudaSetDevice(sp->scene->device));
        std::lock_guard<std::mutex> g(sp->mu);
        std::vector<TracedPath> part = fetch_paths(*sp);
        all.insert(all.end(), std::make_move_iterator(part.begin()), std::make_move_iterator(part.end()));
    }
    if (r.shards.size() > 1)
        std::sort(all.begin(), all.end(), [](const TracedPath &a, const TracedPath &b) {
            return a.sample != b.sample ? a.sample < b.sample : a.pixel < b.pixel;
        });
    CK(cudaSetDevice(r.device0()));
    return all;
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
        need_device(r->shards[0]->scene);
        std::vector<TracedPath> paths = fetch_all_paths(*r);
        *n_paths = (int32_t)paths.size();
        const int stride = r->s0().wf.path_id ? r->s0().wf.path_stride : 0;
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
        need_device(r->shards[0]->scene);
        std::vector<TracedPath> paths = fetch_all_paths(*r);
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
        need_device(r->shards[0]->scene);
        *out = b200pt_profile{};
        for (auto &sp : r->shards) { // launches and work counters add up over the devices; so do the per-class times
            CK(cudaSetDevice(sp->scene->device));
            unsigned long long h[8];
            {
                std::lock_guard<std::mutex> g(sp->mu);
                CK(cudaMemcpyAsync(h, sp->wf.trace_counts, sizeof(h), cudaMemcpyDeviceToHost, sp->scene->stream));
            }
            CK(cudaStreamSynchronize(sp->scene->stream));
            for (int k = 0; k < B200PT_K_COUNT; k++) {
                out->ms[k] += sp->prof.ms[k];
                out->launches[k] += sp->prof.launches[k];
            }
            out->ext_rays += h[0], out->ext_boxes += h[1], out->ext_prims += h[2];
            out->nee_rays += h[3], out->nee_boxes += h[4], out->nee_prims += h[5];
            out->nee_traced += h[6];
            out->nee_queued += h[7];
        }
        CK(cudaSetDevice(r->device0()));
    });
}

void b200pt_render_destroy(b200pt_render *r) { delete r; }

void b200pt_trim_cache(void) { block_cache().trim(); }

} // extern "C"
