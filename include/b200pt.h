/* b200pt.h -- C ABI of the B200-native wavefront path-tracing backend.
 *
 * Plain C: opaque handles, plain pointers and sizes, int status codes (0 = ok).  No torch types.
 *
 * The reference (gillett-hernandez/pathtrace) has no FFI layer; its seam for this path is the
 * C++ Renderer lifecycle driven by main():
 *     renderer_from_config(world, cam, config)            main.cpp:57-84,153
 *     renderer->start_render(t)                           renderer.h:553-603   (Tiled)
 *     while(!renderer->is_done()) renderer->sync_progress()   main.cpp:158-163, renderer.h:605-620
 *     renderer->finalize()                                renderer.h:693-727
 * on top of  build_scene(json) -> World*                  scene_parser.h:241-595
 *            setup_camera(json, aspect)                   main.cpp:86-104
 *            Config(json)                                 config.h:98-132
 *            World::hit(ray, tmin, tmax, rec)             world.h:17-20
 * Each entry point below names the reference interface it replaces.  INTEGRATION.md shows the
 * binding a maintainer of the reference would add (a `Wavefront : public Renderer`).
 *
 * Threading: a handle may be used from one host thread at a time.  All host buffers passed in or
 * out are owned by the caller.  A scene is immutable after creation (uploaded once).
 */
#ifndef B200PT_H
#define B200PT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200PT_OK 0
#define B200PT_ERR_INVALID 1 /* bad argument / malformed JSON / unsupported option */
#define B200PT_ERR_CUDA 2    /* CUDA runtime error (message in b200pt_last_error) */
#define B200PT_ERR_NO_DEVICE 3
#define B200PT_ERR_STATE 4 /* call made in the wrong lifecycle state */

/* ---- flat scene description -------------------------------------------------------------- */

/* primitive kinds (primitive.h:27,123,227; volume.h:7) */
#define B200PT_PRIM_SPHERE 0
#define B200PT_PRIM_RECT 1
#define B200PT_PRIM_BOX 2
/* rect plane (primitive.h:11-16) */
#define B200PT_PLANE_XY 0
#define B200PT_PLANE_XZ 1
#define B200PT_PLANE_YZ 2
/* instance flags */
#define B200PT_INST_MEDIUM 1u    /* wrapped in a constant_medium (volume.h:7-93) */
#define B200PT_INST_BVH_DUP 2u   /* sits alone in an n==1 leaf of the reference BVH, i.e. is
                                    intersected twice there (bvh.h:150-153) */
#define B200PT_INST_RECT_FLIP 4u /* rect built with flip=true (primitive.h:127) */

/* material kinds (material.h:27,76,114,177,247) */
#define B200PT_MAT_LAMBERTIAN 0
#define B200PT_MAT_METAL 1
#define B200PT_MAT_DIELECTRIC 2
#define B200PT_MAT_DIFFUSE_LIGHT 3
#define B200PT_MAT_ISOTROPIC 4

/* texture kinds (texture.h:14,33,185) */
#define B200PT_TEX_CONSTANT 0
#define B200PT_TEX_CHECKER 1
#define B200PT_TEX_PERLIN 2
#define B200PT_TEX_IMAGE 3 /* image_texture (image.h:7-51): texel lookup by (u, v), with an alpha mask */

typedef struct b200pt_instance {
    float fwd[12];      /* local->world 3x4 affine, row major (transform3.h:17-25) */
    int32_t prim_kind;  /* B200PT_PRIM_* */
    int32_t plane;      /* B200PT_PLANE_* (rect only) */
    uint32_t flags;     /* B200PT_INST_* */
    int32_t material;   /* index into materials[]; for a medium: its isotropic phase function */
    float params[6];    /* sphere: cx,cy,cz,r   rect: x0,z0,x1,z1,y   box: p0.xyz,p1.xyz */
    float density;      /* medium only */
    int32_t light_index; /* position in lights[] or -1 */
} b200pt_instance;

typedef struct b200pt_material {
    int32_t kind;    /* B200PT_MAT_* */
    int32_t texture; /* albedo / emission texture */
    float color[3];  /* metal albedo (material.h:109) */
    float ior;       /* dielectric */
    float power;     /* diffuse_light */
    int32_t two_sided;
} b200pt_material;

typedef struct b200pt_texture {
    int32_t kind; /* B200PT_TEX_* */
    float color[3];
    float alpha;
    float scale;  /* checker / perlin */
    int32_t odd;  /* checker: texture ids; image: index into images[] */
    int32_t even;
} b200pt_texture;

/* image_texture's storage (image.h:52-70): row y, column x at rgba[4 * (y * width + x)], components = byte / 255.0
 * as from_4byte_vector computes them; row 0 is the FIRST row of the decoded file. */
typedef struct b200pt_image {
    int32_t width, height;
    const float *rgba;
} b200pt_image;

typedef struct b200pt_camera { /* camera.h:9-36 */
    float origin[3], lower_left[3], horizontal[3], vertical[3], u[3], v[3], w[3];
    float lens_radius;
} b200pt_camera;

typedef struct b200pt_scene_desc {
    const b200pt_instance *instances;
    int32_t n_instances;
    const b200pt_material *materials;
    int32_t n_materials;
    const b200pt_texture *textures;
    int32_t n_textures;
    const int32_t *lights; /* instance ids, in the order of World::lights (scene_parser.h:541-549) */
    int32_t n_lights;
    int32_t background; /* texture id (scene_parser.h:552-577) */
    b200pt_camera camera;
    /* Perlin tables (texture.h:180-183), 256 entries each; may be NULL when no perlin texture */
    const float *perlin_ranvec; /* 256 x 3 */
    const int32_t *perlin_perm; /* 3 x 256: x, y, z */
    const b200pt_image *images; /* B200PT_TEX_IMAGE textures point here (texture.odd); may be NULL */
    int32_t n_images;
} b200pt_scene_desc;

/* ---- render configuration: same keys as config.json (config.h:98-132) ------------------- */

#define B200PT_INTEGRATOR_RPT 0    /* "recursive path tracing"      integrator.h:20-76   */
#define B200PT_INTEGRATOR_INEEPT 4 /* "iterative nee path tracing"  integrator.h:171-343 */

typedef struct b200pt_config {
    int32_t width, height; /* film.width / film.height */
    float exposure;        /* NB config.h:24-25: this is JSON "gamma" (default 2.2) */
    float gamma;           /* ... and this is JSON "exposure"; unused by the reference too */
    int32_t integrator;    /* B200PT_INTEGRATOR_* */
    int32_t max_bounces;
    int32_t samples;
    int32_t light_samples;
    int32_t russian_roulette;
    int32_t only_direct_illumination;
    float normal_offset;
    int32_t block_width, block_height, threads; /* parsed, reported, not used for scheduling */
    /* extensions (not in the reference's config.json; zero = default) */
    uint64_t seed;          /* Philox key */
    int64_t sample_offset;  /* first global sample index rendered by this handle (spp sharding) */
    int32_t device;         /* CUDA device ordinal */
    int32_t batch_samples;  /* samples per pixel per wavefront batch; 0 = auto (up to 128 Mi paths in flight, ~24 GB of state;
                               the environment variable B200PT_BATCH_SLOTS overrides that target) */
    float stats_clamp;      /* >0: also accumulate per-pixel sum / sum of squares of the per-sample
                               radiance clamped to [-stats_clamp, stats_clamp] (parity statistics; a Perlin albedo can be
                               negative, texture.h:191) */
    int32_t profile;        /* measurement aid: 1 = bracket every kernel launch with CUDA events on the render stream
                               (per-kernel device time), 2 = run the traversal kernels with work counters on
                               (BVH boxes / primitives tested); see b200pt_render_profile.  0 in production. */
    float trace_probability; /* config.h:122-131: should_trace_paths ? avg_number_of_paths / (samples * pixels) : 0.
                                Each camera sample is recorded with this probability (renderer.h:652-668) */
    int32_t trace_max_paths; /* capacity of the device-side path store; 0 = max(1024, 4 x the expected number) */
    int32_t physically_corrected; /* config.json "physically_corrected" (default false).  0 = the reference's estimator
                               as it is at HEAD, quirks included -- the only mode compared with the reference.  1 = a
                               standard path tracer over the same scene (device/pt_corrected.h): specular metal with
                               fuzz, refracting glass, emission counted once, outgoing-cosine throughput, 1/n light
                               pick.  metal.fuzz travels in b200pt_material.power. */
    /* Multi-GPU inside the library (config.json "devices": N or "device_ids": [...]); the reference's own analogue is
     * the spp split of its worker threads (renderer.h:352-360: config.samples / N_THREADS each).  device_count <= 1:
     * one GPU, `device`.  Otherwise the scene is replicated on device_ids[0 .. device_count), `samples` are split over
     * them by global sample index (same film for any device count), one host thread and stream per device drive the
     * wavefront batches, and the per-device films are summed on device_ids[0] over NVLink peer loads (falling back to
     * peer copies) in the kernel that also reduces the luminance for the tonemap. */
    int32_t device_count;
    int32_t device_ids[8];
    /* 0: b200pt_render_poll advances the render itself, `max_batches` at a time, and returns when they are done (the
     * caller paces the work; what tests and bench.py use).  1: b200pt_render_start behaves like Renderer::start_render
     * (renderer.h:553-603): worker threads start rendering at once and the call returns; b200pt_render_poll then only
     * REPORTS progress (it never blocks on the GPU beyond a batch boundary), read / tonemap / write_ppm return a
     * consistent snapshot of the film while the render goes on, b200pt_render_wait blocks until it is done. */
    int32_t async_render;
} b200pt_config;

typedef struct b200pt_scene b200pt_scene;
typedef struct b200pt_render b200pt_render;

/* Thread-local message for the last non-zero status returned on this thread. */
const char *b200pt_last_error(void);
/* Library build info, e.g. "b200pt 0.1 sm_100a". */
const char *b200pt_version(void);

/* ---- config.json -------------------------------------------------------------------------- */
/* Config(json) config.h:98-132.  `scene_path` receives the "scene" value, `ppm_path` the
 * "ppm_output_path" value (both NUL-terminated, truncated to cap). */
int b200pt_config_from_json(const char *config_json, b200pt_config *out, char *scene_path, size_t scene_cap,
                            char *ppm_path, size_t ppm_cap);

/* ---- scene --------------------------------------------------------------------------------- */
/* build_scene + setup_camera (scene_parser.h:241-595, main.cpp:86-104): parse the reference's scene JSON,
 * compile it to the flat description, build the BVH and upload everything to `device`.  device = -1 makes a
 * host-only handle (no CUDA call): the compiled description and the BVH can be inspected, every call that would
 * trace returns B200PT_ERR_NO_DEVICE -- there is no CPU fallback. */
int b200pt_scene_create_from_json(const char *scene_json, int32_t film_width, int32_t film_height, int32_t device,
                                  b200pt_scene **out);
/* Same, reading the scene file itself (main.cpp:119-121).  Relative "path" entries of png textures
 * (scene_parser.h:318-323) are looked up in the working directory first, like the reference, then next to the scene
 * file and in its parent directory. */
int b200pt_scene_create_from_file(const char *scene_path, int32_t film_width, int32_t film_height, int32_t device,
                                  b200pt_scene **out);
/* Same, from a caller-built flat description (copied). */
int b200pt_scene_create(const b200pt_scene_desc *desc, int32_t device, b200pt_scene **out);
/* Borrow the compiled flat description (valid until the scene is destroyed). */
int b200pt_scene_get_desc(const b200pt_scene *scene, b200pt_scene_desc *out);
/* Statistics of the tree the kernels walk (4 children per node): n_nodes, depth, and the scene bounds. */
int b200pt_scene_bvh_info(const b200pt_scene *scene, int32_t *n_nodes, int32_t *max_depth, float bounds[6]);
/* Drops the caller's reference.  A scene that still has renders stays alive until the last of them is destroyed (the
 * reference news everything and never frees, main.cpp:108-168; a binding whose garbage collector finalizes the scene object
 * before the renderer object must not crash). */
void b200pt_scene_destroy(b200pt_scene *scene);

/* ---- closest-hit queries (parity entry point) ------------------------------------------- */
typedef struct b200pt_hit {
    int32_t instance; /* index into instances[] (== rec.primitive, primitive.h:305); -1 = miss */
    float t;
    float p[3];
    float n[3];
    float u, v; /* rect only (primitive.h:206-207); 0 otherwise */
} b200pt_hit;

/* World::hit (world.h:17-20) for n rays.  rays: n x 8 float32 {o.xyz, d.xyz, tmin, tmax} in HOST memory;
 * medium_xi: optional n floats in (0,1] -- the uniform variate constant_medium::hit draws (volume.h:70);
 * every medium tested by ray i uses medium_xi[i].  NULL: a Philox stream keyed by the ray index.
 * hits: n records in HOST memory. */
int b200pt_trace_batch(b200pt_scene *scene, const float *rays, size_t n, const float *medium_xi, b200pt_hit *hits);
/* Same on DEVICE pointers, asynchronous on the scene's stream; `repeat` launches back to back (benchmarking).
 * Returns the elapsed device time of all launches in *ms when ms != NULL (synchronises). */
int b200pt_trace_batch_device(b200pt_scene *scene, const void *d_rays, size_t n, const void *d_medium_xi, void *d_hits,
                              int32_t repeat, float *ms);
/* Traversal work counters for the same batch: boxes tested and leaf primitives tested, summed
 * over rays (counts[0] = boxes, counts[1] = primitives, counts[2] = hits). HOST rays. */
int b200pt_trace_count(b200pt_scene *scene, const float *rays, size_t n, uint64_t counts[3]);

/* random_double (random.h:9-15) is replaced by Philox4x32-10 keyed (seed; pixel, global sample, bounce/stage/k, index)
 * (device/pt_rng.h).  This runs the bare generator on the device: counters_keys = n x 6 words {counter[4], key[2]},
 * out = n x 4 words, both in HOST memory -- for known-answer tests against the published Random123 vectors. */
int b200pt_philox4x32_10(int32_t device, const uint32_t *counters_keys, size_t n, uint32_t *out);

/* ---- render lifecycle --------------------------------------------------------------------- */
typedef struct b200pt_progress {
    int64_t samples_done;   /* camera samples finished (pixels x spp), as Tiled::samples_done */
    int64_t samples_total;
    int64_t rays_extension; /* closest-hit queries traced for path segments */
    int64_t rays_shadow;    /* closest-hit queries traced for light samples */
    int64_t rays_reference; /* what the reference's bounce counter would read (integrator.h:192,247) */
    double seconds;         /* device time spent rendering so far */
    int32_t done;
} b200pt_progress;

/* Renderer ctor + start_render (renderer.h:545-603): allocates the wavefront state, returns at once. */
int b200pt_render_start(b200pt_scene *scene, const b200pt_config *config, b200pt_render **out);
/* sync_progress/is_done (renderer.h:605-625): advances the render by at most `max_batches` wavefront batches
 * (0 = run to completion) and reports progress. */
int b200pt_render_poll(b200pt_render *render, int32_t max_batches, b200pt_progress *progress);
/* Blocks until every batch has been rendered (the join the reference's finalize() implies, renderer.h:693-699) and
 * reports the final progress.  Works in both modes (async_render = 0: renders whatever is left). */
int b200pt_render_wait(b200pt_render *render, b200pt_progress *progress);
/* Renderer::framebuffer (renderer.h:141): per-pixel radiance SUMS, H x W x 3 float32, row 0 = image bottom.
 * spp_done receives the number of samples per pixel the sums hold. */
int b200pt_render_read(b200pt_render *render, float *rgb_sums, int64_t *spp_done);
/* Parity statistics (config.stats_clamp > 0): per pixel sum and sum of squares of clamped samples, H x W x 3 each. */
int b200pt_render_read_stats(b200pt_render *render, float *clamped_sum, float *clamped_sumsq);
/* Device pointer of the accumulation buffer (H x W float4, .w unused) and its stream, for an external
 * collective (NCCL reduce across ranks); b200pt_render_add_device adds a peer buffer into it. */
int b200pt_render_device_buffer(b200pt_render *render, void **d_accum, size_t *bytes);
/* Replace the film content by `rgba_sums` (HOST, H x W x 4 float32) -- used after a cross-rank reduce. */
int b200pt_render_write(b200pt_render *render, const float *rgba_sums, int64_t spp);
/* calculate_luminance + output_to_file (helpers.h:146-168, renderer.h:24-55): device max-luminance reduce and
 * tonemap; rgb receives H x W x 3 bytes, TOP row first (the PPM payload). n_samples: the divisor (finalize uses
 * config.samples; sync_progress 1 + done/pixels). */
int b200pt_render_tonemap(b200pt_render *render, int64_t n_samples, uint8_t *rgb, float *max_luminance,
                          float *avg_luminance);
/* Allocation-free tonemap.  NO COUNTERPART IN THE REFERENCE: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The same result
 * as b200pt_render_tonemap -- calculate_luminance + output_to_file on a film of sums -- on buffers the caller owns:
 *   d_rgba_sums  DEVICE, H x W float4 radiance sums, row 0 = image bottom (the layout of b200pt_render_device_buffer)
 *   d_rgb        DEVICE, H x W x 3 bytes out, TOP row first
 *   stream       a cudaStream_t on `device` (NULL = the default stream); the three kernels are enqueued on it
 *   max_luminance / avg_luminance  HOST, optional.  Both NULL: nothing is copied back and the call returns without
 *                synchronising (the white point stays on the device); otherwise the call waits for the stream.
 * It calls no cudaMalloc, cudaFree or allocating library routine: the scratch of the luminance reduce is a file-scope
 * `static __device__` array in the library (128.25 KiB, held by every GPU that loads the library until the process exits),
 * one partial per 1024 pixels.  LIMIT: width x height <= 16 777 216 pixels (16384 partials; twice a 3840 x 2160 film);
 * a larger film is refused with B200PT_ERR_INVALID and a message naming the limit.  Every call on a GPU shares that one
 * array: calls whose kernels overlap on the same GPU (two streams, two threads) race -- serialise them.  The array keeps
 * the previous call's contents; no cell needs clearing, every cell a call reads it has written first.  The luminance
 * total is summed in a fixed order (bit-reproducible; b200pt_render_tonemap's uses float atomics). */
int b200pt_tonemap_static(int32_t device, const void *d_rgba_sums, int32_t width, int32_t height, int64_t n_samples,
                          float exposure, void *d_rgb, void *stream, float *max_luminance, float *avg_luminance);
/* Writes "P6\nW H\n255\n" + payload (renderer.h:26-53). */
int b200pt_render_write_ppm(b200pt_render *render, int64_t n_samples, const char *path);

/* Traced paths (should_trace_paths; renderer.h:57-96,652-678, integrator.h:195-197): the hit point of every vertex of
 * the recorded camera samples, kept on the device and ordered here by (sample, pixel).
 * b200pt_render_paths: n_paths receives the number recorded; when the arrays are non-NULL up to `capacity` paths are
 * copied: pixel[i], sample[i] (global sample index), length[i] (vertices), points[(i * stride + k) * 3 + c] with
 * stride = *vertices_per_path.  b200pt_render_write_paths writes the reference's two text files
 * (handle_traced_paths, renderer.h:57-96): "x,y,z" per vertex and a blank line per path; the same through
 * camera::project (camera.h:49-109) as "x,y" -- with a trailing '!' where project() reports failure. */
int b200pt_render_paths(b200pt_render *render, int32_t capacity, int32_t *n_paths, int32_t *vertices_per_path,
                        int32_t *pixel, int64_t *sample, int32_t *length, float *points);
int b200pt_render_write_paths(b200pt_render *render, const char *path_3d, const char *path_2d, int32_t *n_paths);
/* camera::project (camera.h:49-109) for n points (HOST, n x 3): xy receives n x 2 film coordinates, ok[i] its return
 * value.  For a point behind the film plane (t < 0) x = y = 0 and ok = 0 only when a coordinate exceeds 1e10 -- the
 * reference's own behaviour, which handle_traced_paths prints as "0,0!". */
int b200pt_camera_project(const b200pt_scene *scene, const float *points, size_t n, float *xy, int32_t *ok);
/* The string value of a top-level key of config.json (e.g. "traced_paths_output_path"); empty when absent. */
int b200pt_config_string(const char *config_json, const char *key, char *out, size_t cap);

/* Per-kernel measurements of a render started with config.profile != 0. */
#define B200PT_K_RAYGEN 0
#define B200PT_K_EXTEND 1
#define B200PT_K_SHADE 2
#define B200PT_K_NEE 3
#define B200PT_K_ACCUMULATE 4
#define B200PT_K_SORT 5 /* coherence sort of the ray queues between bounces (BVH scenes) */
#define B200PT_K_COUNT 6
typedef struct b200pt_profile {
    double ms[B200PT_K_COUNT];        /* profile == 1: device time per kernel class, summed over launches */
    int64_t launches[B200PT_K_COUNT]; /* kernel launches per class so far (always counted) */
    /* profile == 2: traversal work of the extension rays (k_extend) and of the shadow rays (k_nee) */
    uint64_t ext_rays, ext_boxes, ext_prims;
    uint64_t nee_rays, nee_boxes, nee_prims;
    /* profile == 2: shadow queries that were really traced.  nee_rays follows the reference's accounting
     * (integrator.h:247: light_samples per vertex) and includes the samples whose coefficient is zero -- light below
     * the horizon, pdf 0 -- for which nothing is intersected. */
    uint64_t nee_traced;
    uint64_t nee_queued; /* light samples that occupied a lane of k_nee: nee_rays minus those of the vertices whose lights all
                          * lie below their horizon (never queued) */
} b200pt_profile;
int b200pt_render_profile(b200pt_render *render, b200pt_profile *out);

void b200pt_render_destroy(b200pt_render *render);

/* Device and pinned-host blocks of destroyed handles are kept for reuse by later handles (the reference `new`s
 * everything once per process and never frees, main.cpp:108-168; a library that is called once per frame must not pay
 * cudaMalloc for 1.3 GB of wavefront state every time).  This releases them.  Env B200PT_CACHE_MB caps the cache
 * (default 49152, 0 = off). */
void b200pt_trim_cache(void);

#ifdef __cplusplus
}
#endif
#endif /* B200PT_H */
