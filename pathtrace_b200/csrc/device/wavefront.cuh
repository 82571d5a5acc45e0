// wavefront.cuh -- device-side layout of one wavefront batch and the launch interface of the kernels.
//
// A batch renders `nsamp` samples of every pixel: slot = s_local * npix + pixel.  All per-path state lives in
// float4 pairs indexed by slot (one 32-byte sector per record); the kernels of one bounce are connected by
// queues of slot indices, compacted with warp ballots and a block prefix sum in shared memory:
//
//   raygen -> [ extend -> shade{miss,diffuse,dielectric,light,isotropic} -> nee ] x bounces -> accumulate
//
// Queue sizes live in `counters` on the device (row = bounce), so the host never waits between kernels.
#pragma once
#include "pt_corrected.h"

#include <cuda_runtime.h>

namespace pt {

enum ShadeKind { SK_MISS = 0, SK_DIFFUSE = 1, SK_DIELECTRIC = 2, SK_LIGHT = 3, SK_ISOTROPIC = 4, SK_COUNT = 5 };

// counters row layout (int32 each)
enum { CNT_EXT = 0, CNT_KIND0 = 1 /* ..5 */, CNT_NEE = 6, CNT_NEE_SKIP = 7 /* vertices whose samples are all void */, CNT_ROW = 8 };

struct BatchCtl {           // device-resident, advanced by the tally kernel at the end of every batch
    int64_t sample_base;    // first global sample index of the current batch
    int32_t nsamp;          // samples per pixel in the current batch
    int32_t remaining;      // samples per pixel still to render after the current batch
    int32_t batch_samples;  // nominal samples per batch
    int32_t pad;
    // running totals
    unsigned long long rays_extension, rays_shadow, rays_reference, samples_done;
};

struct Wavefront {
    // Per-slot state in PAIRS of float4 = one 32-byte DRAM sector per record: once the queues are sorted for ray
    // coherence a kernel visits slots in no particular order, and a 16-byte field on its own would cost a whole sector.
    // k_raygen stores `ray` only: `thr` and `acc` of a slot are first written by the bounce-0 shade kernel the slot passes through
    // (k_shade FIRST), which knows what a camera vertex starts from.
    float4 *ray; // [2 slot] = o.xyz, last_bsdf_pdf      [2 slot + 1] = d.xyz, -
    float4 *thr; // [2 slot] = beta.xyz, -               [2 slot + 1] = attenuation.xyz of the previous vertex, -
    float4 *hit; // [2 slot] = p.xyz, instance | face    [2 slot + 1] = n.xyz, t
    float4 *acc; // [2 slot] = W.xyz = attenuation*beta, material kind    [2 slot + 1] = radiance of the sample so far
    int32_t *q_ext[2];
    int32_t *q_kind[SK_COUNT];
    int32_t *q_nee;
    // Coherence sort between bounces (sort_cells > 0; scenes walked through the BVH): k_shade appends to the RAW queues
    // together with a key per entry -- direction octant | Morton code of the ray origin for the next extension
    // rays, Morton code of the vertex for the light samples -- and a counting sort (k_sort_*) writes q_ext / q_nee in
    // key order, so that the 32 rays of a warp start close together and head the same way.
    int32_t *q_raw_ext, *q_raw_nee;
    uint32_t *key_ext, *key_nee;
    uint32_t *sort_hist;   // PT_SORT_BUCKETS counters + PT_SORT_BUCKETS / PT_SORT_CHUNK chunk totals
    int32_t sort_on;
    int32_t *counters; // (max_bounces + 2) rows of CNT_ROW
    BatchCtl *ctl;
    // traced paths (config.trace_probability > 0, else null): slot -> path id, per path {pixel, sample lo, sample hi,
    // vertices} and path_stride hit points
    int32_t *path_id;
    int4 *path_hdr;
    float4 *path_pts;
    int32_t *path_count;
    int32_t path_max, path_stride;
    float path_prob;
    unsigned long long *trace_counts; // profile == 2: {ext rays, boxes, prims, nee rays, boxes, prims, nee rays traced, -}
    int32_t npix;
    int32_t capacity; // slots
    uint32_t npix_mul, npix_shr; // slot / npix == umulhi(slot, npix_mul) >> npix_shr (npix_mul == 0: npix == 1)
};

#ifndef PT_SORT_DIR_BITS
#define PT_SORT_DIR_BITS 3                                     /* direction bin: 3 = octant, 2n = n x n-bit octahedral grid */
#endif
#define PT_SORT_BITS (3 * PT_SORT_CELL_BITS + PT_SORT_DIR_BITS)
#define PT_SORT_BUCKETS (1 << PT_SORT_BITS)
#define PT_SORT_CHUNK 2048                                    /* counters scanned by one block */

struct Film {
    float4 *accum; // per pixel radiance sums (row 0 = image bottom), .w unused
    float4 *csum;  // parity statistics: sum of clamped samples (or null)
    float4 *csq;   //                    sum of squares of clamped samples
    float clamp;
};

struct LaunchCfg {
    int blocks;  // grid size for the queue-driven kernels (persistent, grid-stride)
    int threads; // 256
    bool count_work; // run the traversal kernels with work counters on (config.profile == 2)
};

void launch_raygen(const PtSceneView &s, const RenderParams &rp, const Wavefront &wf, const LaunchCfg &lc, cudaStream_t st);
void launch_extend(const PtSceneView &s, const RenderParams &rp, const Wavefront &wf, int bounce, const LaunchCfg &lc,
                   cudaStream_t st);
void launch_shade(const PtSceneView &s, const RenderParams &rp, const Wavefront &wf, int bounce, int kind,
                  const LaunchCfg &lc, cudaStream_t st);
void launch_nee(const PtSceneView &s, const RenderParams &rp, const Wavefront &wf, int bounce, const LaunchCfg &lc,
                cudaStream_t st);
// counting sort of a raw queue by key into `dst`; the entry count is read from counters[count_index] on the device
void launch_sort_queue(const Wavefront &wf, const uint32_t *keys, const int32_t *raw, int32_t *dst, int count_index,
                       const LaunchCfg &lc, cudaStream_t st);
void launch_accumulate(const RenderParams &rp, const Wavefront &wf, const Film &film, int n_rows, cudaStream_t st);

// closest-hit batch (b200pt_trace_batch): rays n x 8 floats, hits n x b200pt_hit
void launch_trace_batch(const PtSceneView &s, const float *rays, size_t n, const float *medium_xi, b200pt_hit *hits,
                        unsigned long long *counts, cudaStream_t st);

// Philox4x32-10 on n x {counter[4], key[2]} words -> n x 4 words (known-answer tests)
void launch_philox(const uint32_t *in, size_t n, uint32_t *out, cudaStream_t st);

// film: luminance reduce (max, total of |sum/n|) and tonemap to bytes (top row first)
void launch_luminance(const float4 *accum, int npix, float inv_n, float *max_total /* [2] */, cudaStream_t st);
// multi-GPU: dst = sum of up to 8 films (device order), optionally with the luminance reduce of the sum fused in
struct FilmSources {
    const float4 *src[8];
    int n;
};
void launch_combine(const FilmSources &fs, int npix, float4 *dst, float inv_n, float *max_total /* [2] or null */,
                    cudaStream_t st);
void launch_pack_rgb(const float4 *accum, int npix, float *rgb, cudaStream_t st);
void launch_tonemap(const float4 *accum, int width, int height, float inv_n, float exposure_mul, float max_lum,
                    uint8_t *rgb, cudaStream_t st);

// Allocation-free tonemap (b200pt_tonemap_static): luminance reduce + tonemap with the reduce's scratch in a file-scope
// __device__ array of kernels.cu.  One tile of PT_STATIC_TILE pixels per partial, at most PT_STATIC_MAX_TILES tiles.
#define PT_STATIC_TILE 1024
#define PT_STATIC_MAX_TILES 16384
#define PT_STATIC_MAX_PIXELS (PT_STATIC_TILE * PT_STATIC_MAX_TILES) /* 16 777 216: twice a 3840 x 2160 film */
#define PT_STATIC_SCRATCH_FLOATS (2 * PT_STATIC_MAX_TILES + 64)
cudaError_t launch_tonemap_static(int device, const float4 *accum, int width, int height, float inv_n, float exposure_mul,
                                  uint8_t *rgb, float *h_max_total /* [2] host, or null */, cudaStream_t st);

} // namespace pt
