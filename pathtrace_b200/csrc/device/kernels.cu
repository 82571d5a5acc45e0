// kernels.cu -- the sm_100a kernels of the wavefront path tracer.
//
//   k_raygen      Tiled::compute's jitter + camera::get_ray (renderer.h:648-650, camera.h:38-47)
//   k_extend      World::hit for every live path (world.h:17-20) + sort by material kind into shade queues
//   k_shade       the vertex logic of NEEIterative::color / RecursivePT::color, one launch per material kind
//   k_nee         light_samples shadow rays per vertex with light-pdf / BSDF-pdf MIS (integrator.h:221-265)
//   k_accumulate  framebuffer[j][i] += de_nan(col) (renderer.h:670-682)
//   k_luminance / k_tonemap   calculate_luminance + output_to_file (helpers.h:146-168, renderer.h:24-55)
//   k_trace_batch the parity entry point (b200pt_trace_batch)
//
// Nothing here is a dense contraction, so tensor cores are not involved; the kernels are bound by fp32
// intersection math, divergence and the L2/HBM traffic of the path state.
#include "wavefront.cuh"

#include <algorithm>
#include <mutex>

namespace pt {

namespace {

constexpr int kThreads = 256;
constexpr int kFaceShift = 28; // hit_p.w = instance | box side << 28 (b200pt_scene_create refuses >= 2^28 instances)
constexpr int kWarps = kThreads / 32;

__device__ __forceinline__ V3 xyz(float4 a) { return v3(a.x, a.y, a.z); }
__device__ __forceinline__ float4 f4(V3 a, float w) { return make_float4(a.x, a.y, a.z, w); }

// Path state is streamed: every record is read once and written once per bounce and the whole state (24 GB per batch) is far
// beyond the L2, while the scene (1.6 MB for 8192 instances) is re-read by every ray.  The TRAVERSAL kernels therefore mark
// their state accesses evict-first (ld.global.cs / st.global.cs), so that these do not push node and instance records out
// of L1 / L2: k_extend -1.5 %, k_nee -1.1 % on book2_final.  The shade kernels do not: with the same hints k_shade is 8 %
// slower (its records come back in the next kernel; measured, profiles/r2r_*).  PT_STATE_HINT is a bit mask for A/B builds:
// 1 traversal loads, 2 traversal stores, 4 shade / raygen loads, 8 shade / raygen stores.
#ifndef PT_STATE_HINT
#define PT_STATE_HINT 3
#endif
template <bool TRAVERSAL = false> __device__ __forceinline__ float4 ld_state(const float4 *p)
{
    if (PT_STATE_HINT & (TRAVERSAL ? 1 : 4))
        return __ldcs(p);
    return *p;
}
template <bool TRAVERSAL = false> __device__ __forceinline__ void st_state(float4 *p, float4 v)
{
    if (PT_STATE_HINT & (TRAVERSAL ? 2 : 8))
        __stcs(p, v);
    else
        *p = v;
}

// Block-wide queue reservation: every thread passes a bit mask of the NQ queues it appends to.  Ranks come from
// warp ballots, warp totals are prefix-summed in shared memory and ONE atomicAdd per (block, queue) reserves the
// range.  Must be called by all threads of the block.  Two barriers, not three: the scratch is double-buffered by
// `phase` (the caller's loop iteration & 1) -- a thread that runs ahead into iteration i + 1 writes the other buffer,
// and it cannot reach iteration i + 2 (this buffer again) before every thread has passed the first barrier of i + 1,
// i.e. has finished reading iteration i's values.  (Reserving per warp instead -- no barrier at all, one atomic per warp and
// queue -- was measured in round 2 because barriers are a third of k_shade's stall samples: eight times the same-address
// atomics saturate the L2 atomic unit, k_shade 100 -> 141 ms per book2_final step, 51 -> 89 ms on cornell_box.)
template <int NQ, int NW = kWarps>
__device__ __forceinline__ void block_reserve(unsigned mask, int32_t *const (&counter)[NQ], int (&pos)[NQ], int phase)
{
    __shared__ int s_cnt[2][NQ][NW];
    __shared__ int s_base[2][NQ];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned lt = (1u << lane) - 1u;
    int rank[NQ];
#pragma unroll
    for (int q = 0; q < NQ; q++) {
        unsigned b = __ballot_sync(0xffffffffu, (mask >> q) & 1u);
        rank[q] = __popc(b & lt);
        if (lane == 0)
            s_cnt[phase][q][warp] = __popc(b);
    }
    __syncthreads();
    if (threadIdx.x < NQ) {
        int total = 0;
#pragma unroll
        for (int w = 0; w < NW; w++) {
            int c = s_cnt[phase][threadIdx.x][w];
            s_cnt[phase][threadIdx.x][w] = total;
            total += c;
        }
        s_base[phase][threadIdx.x] = total ? atomicAdd(counter[threadIdx.x], total) : 0;
    }
    __syncthreads();
#pragma unroll
    for (int q = 0; q < NQ; q++)
        pos[q] = s_base[phase][q] + s_cnt[phase][q][warp] + rank[q];
}

// A small scene in shared memory: stage() copies boxes, instances and forward transforms and points the view at
// the copies, so that nothing on the path of a ray reads the scene from global memory (materials aside).
struct SmallScene {
    PtBox box[PT_SMALL_MAX];
    PtInstance inst[PT_SMALL_MAX];
    PtFwd fwd[PT_SMALL_MAX];

    __device__ __forceinline__ void stage(PtSceneView &s) // all threads of the block; ends with a barrier
    {
        const int n = s.n_small;
        copy16(box, s.small, n * (int)(sizeof(PtBox) / 16));
        copy16(inst, s.inst, n * (int)(sizeof(PtInstance) / 16));
        copy16(fwd, s.fwd, n * (int)(sizeof(PtFwd) / 16));
        __syncthreads();
        s.small = box;
        s.inst = inst;
        s.fwd = fwd;
    }
    static __device__ __forceinline__ void copy16(void *dst, const void *src, int n16)
    {
        for (int i = threadIdx.x; i < n16; i += blockDim.x)
            reinterpret_cast<float4 *>(dst)[i] = __ldg(reinterpret_cast<const float4 *>(src) + i);
    }
};

// work counters of one thread -> global totals (profile == 2 only)
__device__ __forceinline__ void flush_counts(const TraceCounters &tc, unsigned long long rays, unsigned long long *dst,
                                             unsigned long long traced = 0, unsigned long long *dst_traced = nullptr)
{
    unsigned long long b = tc.boxes, p = tc.prims, r = rays, t = traced;
    for (int off = 16; off > 0; off >>= 1) {
        b += __shfl_down_sync(0xffffffffu, b, off);
        p += __shfl_down_sync(0xffffffffu, p, off);
        r += __shfl_down_sync(0xffffffffu, r, off);
        t += __shfl_down_sync(0xffffffffu, t, off);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(dst + 0, r);
        atomicAdd(dst + 1, b);
        atomicAdd(dst + 2, p);
        if (dst_traced) {
            atomicAdd(dst_traced, t);
            atomicAdd(dst_traced + 1, r); // queries that occupied a lane (queued), as opposed to merely counted
        }
    }
}

__device__ __forceinline__ RngKey make_key(const RenderParams &rp, const Wavefront &wf, int slot, long long sample_base)
{
    RngKey k;
    k.seed_lo = rp.seed_lo;
    k.seed_hi = rp.seed_hi;
    // slot / npix by multiply-high with a host-computed reciprocal (exact for slot < 2^31): an integer division is
    // ~20 instructions and every kernel keys its random numbers through here
    int s_local = wf.npix_mul ? (int)(__umulhi((uint32_t)slot, wf.npix_mul) >> wf.npix_shr) : slot;
    k.pixel = (uint32_t)(slot - s_local * wf.npix);
    key_set_sample(k, sample_base + s_local);
    return k;
}

__device__ __forceinline__ int shade_kind_of(int mat_kind)
{
    switch (mat_kind) {
    case B200PT_MAT_DIELECTRIC: return SK_DIELECTRIC;
    case B200PT_MAT_DIFFUSE_LIGHT: return SK_LIGHT;
    case B200PT_MAT_ISOTROPIC: return SK_ISOTROPIC;
    default: return SK_DIFFUSE; // lambertian and metal share one lobe (material.h:27-112)
    }
}

// ---- coherence sort keys ---------------------------------------------------------------------------------------------
// 10 bits -> every third bit
__device__ __forceinline__ uint32_t spread3(uint32_t x)
{
    x = (x | (x << 16)) & 0x030000FFu;
    x = (x | (x << 8)) & 0x0300F00Fu;
    x = (x | (x << 4)) & 0x030C30C3u;
    x = (x | (x << 2)) & 0x09249249u;
    return x;
}

// Morton code of the cell of p on a 2^bits grid over the scene's sort bounds (positions outside are clamped)
template <int BITS> __device__ __forceinline__ uint32_t morton_cell(const PtSceneView &s, V3 p)
{
    const float top = (float)((1 << BITS) - 1), up = (float)(1 << BITS);
    // fminf / fmaxf drop NaNs: a NaN coordinate lands in cell 0
    const uint32_t cx = (uint32_t)fminf(fmaxf((p.x - s.sort_lo[0]) * (s.sort_scale[0] * up), 0.f), top);
    const uint32_t cy = (uint32_t)fminf(fmaxf((p.y - s.sort_lo[1]) * (s.sort_scale[1] * up), 0.f), top);
    const uint32_t cz = (uint32_t)fminf(fmaxf((p.z - s.sort_lo[2]) * (s.sort_scale[2] * up), 0.f), top);
    return spread3(cx) | (spread3(cy) << 1) | (spread3(cz) << 2);
}

// next extension ray: direction bin (major) | cell of the origin -- rays of a warp start together and head the same way.
// PT_SORT_DIR_BITS = 3: the octant of d; 2n > 3: an n x n-bit grid over the octahedral map of d
__device__ __forceinline__ uint32_t sort_key_ext(const PtSceneView &s, V3 o, V3 d)
{
#if PT_SORT_DIR_BITS == 3
    const uint32_t dir = (d.x < 0.f ? 1u : 0u) | (d.y < 0.f ? 2u : 0u) | (d.z < 0.f ? 4u : 0u);
#else
    constexpr int H = PT_SORT_DIR_BITS / 2;
    const float inv = 1.0f / (fabsf(d.x) + fabsf(d.y) + fabsf(d.z));
    float u = d.x * inv, v = d.y * inv;
    if (d.z < 0.f) { // fold the lower half outwards
        const float fu = (1.f - fabsf(v)) * (u < 0.f ? -1.f : 1.f), fv = (1.f - fabsf(u)) * (v < 0.f ? -1.f : 1.f);
        u = fu, v = fv;
    }
    const float top = (float)((1 << H) - 1), half = (float)(1 << (H - 1));
    const uint32_t bu = (uint32_t)fminf(fmaxf((u + 1.f) * half, 0.f), top), bv = (uint32_t)fminf(fmaxf((v + 1.f) * half, 0.f), top);
    const uint32_t dir = (bv << H) | bu;
#endif
    return (dir << (3 * PT_SORT_CELL_BITS)) | morton_cell<PT_SORT_CELL_BITS>(s, o);
}

// light samples: all of them aim at the few lights, so the vertex position alone orders them (one more bit per axis)
__device__ __forceinline__ uint32_t sort_key_nee(const PtSceneView &s, V3 p) { return morton_cell<PT_SORT_CELL_BITS + 1>(s, p); }

// ---------------------------------------------------------------------------------------------------------------

__global__ void __launch_bounds__(kThreads) k_raygen(PtSceneView s, RenderParams rp, Wavefront wf)
{
    const long long sample_base = wf.ctl->sample_base;
    const int n = wf.ctl->nsamp * wf.npix;
    if (blockIdx.x == 0 && threadIdx.x == 0)
        wf.counters[CNT_EXT] = n; // bounce 0 extends every slot (identity queue)
    for (int slot = blockIdx.x * kThreads + threadIdx.x; slot < n; slot += gridDim.x * kThreads) {
        RngKey key = make_key(rp, wf, slot, sample_base);
        F4 r = rng4(key, stage_word(0, RNG_CAMERA, 0), 0);
        int j = key.pixel / rp.width, i = key.pixel - j * rp.width; // row 0 = image bottom
        float u = (float)(i + (double)r.x) / (float)rp.width;      // renderer.h:648-649
        float v = (float)(j + (double)r.y) / (float)rp.height;
        V3 o, d;
        camera_ray(s.cam, u, v, r.z, r.w, o, d);
        st_state(&wf.ray[2 * slot], f4(o, -1.f)); // last_bsdf_pdf = -1 (integrator.h:183)
        st_state(&wf.ray[2 * slot + 1], f4(d, 0.f));
        // throughput (1, 1, 1), no attenuation yet and an empty sum are what a camera vertex starts from: the bounce-0
        // instantiations of k_shade (FIRST) know that and WRITE the sum, so none of the three is stored here
        if (wf.path_id) { // `random_double() < trace_probability` (renderer.h:652)
            int id = -1;
            if (rng4(key, stage_word(0, RNG_PATH_TRACE, 0), 0).x < wf.path_prob) {
                id = atomicAdd(wf.path_count, 1);
                if (id < wf.path_max) {
                    wf.path_hdr[id] = make_int4((int)key.pixel, (int)key.sample, (int)key.sample_hi, 0);
                } else {
                    id = -1;
                }
            }
            wf.path_id[slot] = id;
        }
    }
}

// Small blocks: the queue reservation is a block-wide barrier, and a block waits for its slowest ray -- with 256
// threads a quarter of all stall samples on the 8192-instance scene were that barrier.
// the node phase of the walk ends when fewer than 1 / DIV of the warp's unfinished lanes want an inner node (0: when none does).
// Measured on book2_final (profiles/r2n_ab_experiments.txt): DIV = 3 gains 0.3 %, 4 / 6 / 8 lose 0.4 - 1.5 % -- the policy is not
// what limits the lanes of the node loop (14.7 of 32 in k_nee): the rays of a warp finish at different times.
#ifndef PT_NEE_NODE_DIV
#define PT_NEE_NODE_DIV 0
#endif
#ifndef PT_EXT_NODE_DIV
#define PT_EXT_NODE_DIV 0
#endif
#ifndef PT_EXT_THREADS
#define PT_EXT_THREADS 64
#endif
constexpr int kExtThreads = PT_EXT_THREADS;
template <bool COUNT, bool MEDIA, bool SMALL>
#ifndef PT_EXT_MINB
#define PT_EXT_MINB 16
#endif
__global__ void __launch_bounds__(kExtThreads, PT_EXT_MINB) k_extend(PtSceneView s, RenderParams rp, Wavefront wf, int bounce)
{
    __shared__ SmallScene s_small; // dropped by the compiler when SMALL is false
    if (SMALL)
        s_small.stage(s);
    TraceCounters tc;
    tc.boxes = tc.prims = 0;
    unsigned long long nrays = 0;
    const long long sample_base = wf.ctl->sample_base;
    int32_t *row = wf.counters + bounce * CNT_ROW;
    const int n = row[CNT_EXT];
    const int32_t *q_in = bounce == 0 ? nullptr : wf.q_ext[bounce & 1];
    int32_t *const kind_counter[SK_COUNT] = {row + CNT_KIND0, row + CNT_KIND0 + 1, row + CNT_KIND0 + 2,
                                             row + CNT_KIND0 + 3, row + CNT_KIND0 + 4};
    int phase = 0;
    for (int base = blockIdx.x * kExtThreads; base < n; base += gridDim.x * kExtThreads, phase ^= 1) {
        const int i = base + threadIdx.x;
        unsigned mask = 0;
        int slot = 0;
        const bool live = i < n;
        Traversal T;
        TraversalStack T_stack;
        T.stk = SMALL ? nullptr : &T_stack; // the table walk of a small scene has no pending nodes
        MediumRng mr;
        mr.xi = -1.f;
        if (live) {
            slot = q_in ? q_in[i] : i;
            PT_CHECK(slot >= 0 && slot < wf.capacity);
            float4 ro = ld_state<true>(&wf.ray[2 * slot]), rd = ld_state<true>(&wf.ray[2 * slot + 1]);
            mr.key = make_key(rp, wf, slot, sample_base);
            mr.stage = stage_word(bounce, RNG_EXT_MEDIUM, 0);
            if (MEDIA) {
                // the variates of media 0 and 1 now, with all lanes busy, instead of at the first medium leaf (k_extend -0.7 %
                // on book2_final, whose mist every ray meets; the same in k_nee costs it 2 %: a sixth of its lanes are void)
                mr.cached = rng4(mr.key, mr.stage, 0);
                mr.cached_block = 0;
            }
            T.init(xyz(ro), xyz(rd), 0.001f, PT_FLT_MAX);
            nrays++;
        }
        if (SMALL) {
            if (live)
                T.template run_small<COUNT, MEDIA>(s.small, s.inst, s.n_small, mr, &tc);
        } else {
#ifdef PT_EXT_VOTE // A/B: the warp votes on the kind of the next step (the driver of rounds 1-2: 158.0 vs 156.6 ms per step)
            T.template run_warp<COUNT, MEDIA>(s, live, mr, &tc);
#else
            T.template run_warp_nodes_first<COUNT, MEDIA, false, PT_EXT_NODE_DIV>(s, live, mr, &tc); // all 32 lanes take part
#endif
        }
        if (live) {
            Hit h;
            int kind = SK_MISS;
            if (T.template result<MEDIA>(s, h)) {
                st_state<true>(&wf.hit[2 * slot], f4(h.p, __int_as_float(h.inst | (h.face << kFaceShift))));
                st_state<true>(&wf.hit[2 * slot + 1], f4(h.n, h.t));
                kind = shade_kind_of(s.mats[meta_material(s.inst[h.inst].meta)].kind);
            }
            mask = 1u << kind;
        }
        int pos[SK_COUNT];
        block_reserve<SK_COUNT, kExtThreads / 32>(mask, kind_counter, pos, phase);
        if (mask) {
            int kind = __ffs(mask) - 1;
            PT_CHECK(pos[kind] >= 0 && pos[kind] < wf.capacity);
            wf.q_kind[kind][pos[kind]] = slot;
        }
    }
    if (COUNT)
        flush_counts(tc, nrays, wf.trace_counts);
}

// One instantiation per (shade queue, integrator): each holds the code of its own materials only.
#ifndef PT_SHADE_THREADS
#define PT_SHADE_THREADS 256
#endif
constexpr int kShadeThreads = PT_SHADE_THREADS;
#ifndef PT_SHADE_MINB
#define PT_SHADE_MINB (1024 / PT_SHADE_THREADS) /* blocks per SM the register allocation aims at */
#endif
// FIRST: the camera vertices (bounce 0).  Every slot of the batch passes through exactly one of the FIRST kernels, with
// throughput 1, no previous attenuation and an empty radiance sum: nothing of that is read, and the sum is stored (a whole
// 32-byte sector with W) instead of read, added to and written back.
template <int kind, bool NEE, bool CORR, bool FIRST>
__global__ void __launch_bounds__(kShadeThreads, PT_SHADE_MINB) k_shade(PtSceneView s, RenderParams rp, Wavefront wf, int bounce)
{
    const long long sample_base = wf.ctl->sample_base;
    int32_t *row = wf.counters + bounce * CNT_ROW;
    const int n = row[CNT_KIND0 + kind];
    const int32_t *q_in = wf.q_kind[kind];
    const bool sorted = wf.sort_on != 0;
    int32_t *q_next = sorted ? wf.q_raw_ext : wf.q_ext[(bounce + 1) & 1];
    int32_t *q_nee_out = sorted ? wf.q_raw_nee : wf.q_nee;
    int32_t *const counter[2] = {row + CNT_NEE, row + CNT_ROW + CNT_EXT};
    int phase = 0;
    for (int base = blockIdx.x * kShadeThreads; base < n; base += gridDim.x * kShadeThreads, phase ^= 1) {
        const int i = base + threadIdx.x;
        unsigned mask = 0;
        int slot = 0;
        bool nee_skipped = false;
        uint32_t key_nee = 0, key_ext = 0;
        if (i < n) {
            slot = q_in[i];
            PT_CHECK(slot >= 0 && slot < wf.capacity);
            if (kind == SK_MISS) {
                // integrator.h:325-335: sum += beta * background.  A constant background (every shipped scene) needs
                // no ray direction, and a zero product -- a black sky -- leaves the sum alone: one 32-byte sector per miss
                // instead of five.  (x + 0 == x for every x, NaN included; a NaN or infinite beta still gives a NaN product
                // and is added, as in the reference.)
                const float4 bt = FIRST ? make_float4(1.f, 1.f, 1.f, 0.f) : ld_state(&wf.thr[2 * slot]);
                const b200pt_texture &bg = s.tex[s.background];
                V3 add;
                if (bg.kind == B200PT_TEX_CONSTANT)
                    add = xyz(bt) * v3(bg.color[0], bg.color[1], bg.color[2]);
                else
                    add = shade_miss(s, xyz(ld_state(&wf.ray[2 * slot + 1])), xyz(bt));
                if (FIRST) {
                    st_state(&wf.acc[2 * slot], make_float4(0.f, 0.f, 0.f, 0.f));
                    st_state(&wf.acc[2 * slot + 1], make_float4(add.x, add.y, add.z, 0.f));
                } else if (add.x != 0.f || add.y != 0.f || add.z != 0.f) {
                    const float4 acc = ld_state(&wf.acc[2 * slot + 1]);
                    st_state(&wf.acc[2 * slot + 1], make_float4(acc.x + add.x, acc.y + add.y, acc.z + add.z, 0.f));
                }
                continue; // a miss never appends (every thread of the block takes the `continue` below)
            }
            float4 ro = ld_state(&wf.ray[2 * slot]), rd = ld_state(&wf.ray[2 * slot + 1]);
            const float4 bt = FIRST ? make_float4(1.f, 1.f, 1.f, 0.f) : ld_state(&wf.thr[2 * slot]);
            {
                float4 hp = ld_state(&wf.hit[2 * slot]), hn = ld_state(&wf.hit[2 * slot + 1]);
                PathVertexIn in;
                in.o = xyz(ro);
                in.d = xyz(rd);
                in.beta = xyz(bt);
                in.last_bsdf_pdf = ro.w;
                // only a light leaves `attenuation` as the previous vertex set it (material.h:187-191)
                in.attenuation = kind == SK_LIGHT && !FIRST ? xyz(ld_state(&wf.thr[2 * slot + 1])) : v3(0.f, 0.f, 0.f);
                Hit rec;
                rec.inst = __float_as_int(hp.w) & ((1 << kFaceShift) - 1);
                rec.face = __float_as_int(hp.w) >> kFaceShift;
                rec.p = xyz(hp);
                rec.n = xyz(hn);
                rec.t = hn.w;
                rec.u = rec.v = 0.f;
                if (s.n_images > 0) // only image textures look at u, v
                    hit_uv(s, rec.inst, rec.face, xyz(ro), xyz(rd), rec.t, rec.u, rec.v);
                if (wf.path_id) { // `_path->push_back(rec.p)` (integrator.h:34-37,195-198)
                    const int id = wf.path_id[slot];
                    if (id >= 0 && bounce < wf.path_stride) {
                        wf.path_pts[(size_t)id * wf.path_stride + bounce] = f4(rec.p, 0.f);
                        wf.path_hdr[id].w = bounce + 1;
                    }
                }
                RngKey key = make_key(rp, wf, slot, sample_base);
                F4 rnd = rng4(key, stage_word(bounce, RNG_SHADE, 0), 0);
                PathVertexOut out;
                if (CORR)
                    shade_hit_corrected<kind>(s, rp, in, rec, rnd, out);
                else if (NEE)
                    shade_hit_ineept<kind>(s, rp, in, rec, rnd, out);
                else
                    shade_hit_rpt<kind>(s, rp, in, rec, bounce, rnd, out);
                if (FIRST) {
                    st_state(&wf.acc[2 * slot + 1], make_float4(out.add.x, out.add.y, out.add.z, 0.f));
                } else if (out.add.x != 0.f || out.add.y != 0.f || out.add.z != 0.f) { // rare: the sum is read only then
                    const float4 acc = ld_state(&wf.acc[2 * slot + 1]);
                    st_state(&wf.acc[2 * slot + 1], make_float4(acc.x + out.add.x, acc.y + out.add.y, acc.z + out.add.z, 0.f));
                }
                bool w_stored = false;
                if (out.want_nee) {
                    // all lights below the horizon of this vertex: its light samples are all void (pt_shade.h), they
                    // are counted (k_tally) and never queued
                    if (NEE && !CORR && lights_below_horizon(s, rec.p, rec.n)) {
                        nee_skipped = true;
                    } else {
                        st_state(&wf.acc[2 * slot], f4(out.W, __int_as_float(out.mat_kind)));
                        w_stored = true;
                        mask |= 1u;
                        if (sorted)
                            key_nee = sort_key_nee(s, rec.p);
                    }
                }
                if (FIRST && !w_stored)
                    st_state(&wf.acc[2 * slot], make_float4(0.f, 0.f, 0.f, 0.f)); // the other half of the sum's sector
                if (out.cont) {
                    if (sorted)
                        key_ext = sort_key_ext(s, out.o, out.d);
                    st_state(&wf.ray[2 * slot], f4(out.o, out.last_bsdf_pdf));
                    st_state(&wf.ray[2 * slot + 1], f4(out.d, 0.f));
                    st_state(&wf.thr[2 * slot], f4(out.beta, 0.f));
                    st_state(&wf.thr[2 * slot + 1], f4(out.attenuation, 0.f));
                    mask |= 2u;
                }
            }
        }
        if (kind == SK_MISS)
            continue; // uniform: a miss never appends
        if (NEE && !CORR) {
            const unsigned b = __ballot_sync(0xffffffffu, nee_skipped);
            if (b && (threadIdx.x & 31) == 0)
                atomicAdd(row + CNT_NEE_SKIP, __popc(b));
        }
        int pos[2];
        block_reserve<2, kShadeThreads / 32>(mask, counter, pos, phase);
        PT_CHECK(!(mask & 1u) || (pos[0] >= 0 && pos[0] < wf.capacity));
        PT_CHECK(!(mask & 2u) || (pos[1] >= 0 && pos[1] < wf.capacity));
        if (mask & 1u) {
            q_nee_out[pos[0]] = slot;
            if (sorted)
                wf.key_nee[pos[0]] = key_nee;
        }
        if (mask & 2u) {
            q_next[pos[1]] = slot;
            if (sorted)
                wf.key_ext[pos[1]] = key_ext;
        }
    }
}

// One thread per (vertex, light sample).  With light_samples a power of two <= 32 the samples of one vertex sit in
// adjacent lanes: same origin, nearly the same target -> coherent traversal, and the sum is a fixed-order shuffle
// reduction; otherwise float atomics.
#ifndef PT_NEE_MEDIA_MINB
#define PT_NEE_MEDIA_MINB 4
#endif
// ANY: the scene has at most PT_MAX_DIRECT_EMITTERS emitters -> emitters first, then an any-hit walk (trace_shadow)
// CORR: the estimator of pt_corrected.h instead of the reference's
#ifndef PT_NEE_THREADS
#define PT_NEE_THREADS 256
#endif
constexpr int kNeeThreads = PT_NEE_THREADS; // block size of k_nee: 128 / 64 threads at the same occupancy measured +1.5 % / +1.7 % time
template <bool COUNT, bool MEDIA, bool ANY, bool SMALL, bool CORR>
__global__ void __launch_bounds__(kNeeThreads, (MEDIA ? PT_NEE_MEDIA_MINB : 4) * (kThreads / kNeeThreads)) k_nee(PtSceneView s, RenderParams rp, Wavefront wf, int bounce)
{
    static_assert(ANY || !SMALL, "the small-scene walk is instantiated for the any-hit shadow query only");
    __shared__ SmallScene s_small; // dropped by the compiler when SMALL is false
    if (SMALL)
        s_small.stage(s);
    TraceCounters tc;
    tc.boxes = tc.prims = 0;
    unsigned long long nrays = 0, ntraced = 0;
    const long long sample_base = wf.ctl->sample_base;
    const int L = rp.light_samples;
    // The samples of one vertex occupy a group of P adjacent lanes, P = light_samples rounded up to a power of two
    // (rp.light_group; lanes k >= L of a group idle), so that their sum is a fixed-order butterfly whatever L is.
    // P == 0 (light_samples > 32): one thread per sample, float atomics -- the one case whose film is not
    // bit-reproducible (include/b200pt.h).  capacity * P < 2^31 is guaranteed by b200pt_render_start.
    const int P = rp.light_group;
    const bool pow2 = P != 0;
    const int shift = pow2 ? __ffs(P) - 1 : 0;
    const unsigned n = (unsigned)wf.counters[bounce * CNT_ROW + CNT_NEE] * (unsigned)(pow2 ? P : L);
    const float inv_L_div = (float)L;
    for (unsigned base = blockIdx.x * kNeeThreads; base < n; base += gridDim.x * kNeeThreads) {
        const unsigned e = base + threadIdx.x;
        V3 c = v3(0.f, 0.f, 0.f);
        int slot = -1;
        Traversal T;
        TraversalStack T_stack;
        T.stk = SMALL ? nullptr : &T_stack; // the table walk of a small scene has no pending nodes
        MediumRng mr;
        mr.xi = -1.f;
        NeeSample ns;
        bool trace = false;
        int lit = -1;
        int q = 0, k = 0;
        if (pow2) {
            q = (int)(e >> shift);
            k = (int)(e & (unsigned)(P - 1));
        } else {
            q = (int)(e / (unsigned)L);
            k = (int)(e - (unsigned)q * (unsigned)L);
        }
        if (e < n && k < L) {
            PT_CHECK(q >= 0 && q < wf.counters[bounce * CNT_ROW + CNT_NEE] && q < wf.capacity);
            slot = wf.q_nee[q];
            PT_CHECK(slot >= 0 && slot < wf.capacity);
            float4 hp = ld_state<true>(&wf.hit[2 * slot]), hn = ld_state<true>(&wf.hit[2 * slot + 1]), w = ld_state<true>(&wf.acc[2 * slot]);
            RngKey key = make_key(rp, wf, slot, sample_base);
            nrays++;
            if (CORR)
                nee_prepare_corrected(s, key, bounce, k, xyz(hp), xyz(hn), __float_as_int(w.w), xyz(w), ns);
            else
                nee_prepare(s, key, bounce, k, xyz(hp), xyz(hn), __float_as_int(w.w), xyz(w), ns);
            mr = nee_medium_rng(key, bounce, k);
            T.init(xyz(hp), ns.dir, 0.001f, PT_FLT_MAX);
            trace = !nee_is_void(ns);
            if (COUNT && trace)
                ntraced++; // a query that really is traced (void samples are counted as rays, never walked)
            if (ANY && trace) {
                trace = T.template emitters_first<COUNT>(s, &tc, ns.known_inst, &ns.known);
                lit = T.best_inst;
            }
        }
        // BVH: inner nodes until no lane wants one, then the reached leaves together (measured against the warp vote
        // of k_extend and against every lane looping on its own: -7 % / -10 % on the 8192-instance scene)
        if (SMALL) {
            if (trace)
                T.template run_small<COUNT, MEDIA, ANY>(s.small, s.inst, s.n_small, mr, &tc);
        } else {
            T.template run_warp_nodes_first<COUNT, MEDIA, ANY, PT_NEE_NODE_DIV>(s, trace, mr, &tc);
        }
        if (trace && (!ANY || T.best_inst == lit)) {
            Hit lrec;
            const bool hit = T.template result<MEDIA>(s, lrec, !ANY || s.emit_needs_p != 0, !ANY || s.emit_needs_n != 0);
            c = CORR ? nee_finish_corrected(s, ns, hit, lrec) : nee_finish(s, ns, hit, lrec);
        }
        if (pow2) {
            // lanes [g*P, (g+1)*P) hold one vertex (kThreads and 32 are multiples of P): fixed-order butterfly over
            // the group.  Five unconditional stages (the stages >= P are discarded) keep the shuffles out of
            // divergent control flow.
            __syncwarp();
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const float x = __shfl_xor_sync(0xffffffffu, c.x, off);
                const float y = __shfl_xor_sync(0xffffffffu, c.y, off);
                const float z = __shfl_xor_sync(0xffffffffu, c.z, off);
                if (off < P) {
                    c.x += x;
                    c.y += y;
                    c.z += z;
                }
            }
            if (slot >= 0 && (threadIdx.x & (P - 1)) == 0 && (c.x != 0.f || c.y != 0.f || c.z != 0.f)) {
                // one lane per vertex and kernel touches sum[slot]: a reduction without a return value is as
                // deterministic as load-add-store and does not wait for the load
                float *dst = reinterpret_cast<float *>(wf.acc + 2 * slot + 1);
                atomicAdd(dst + 0, c.x / inv_L_div);
                atomicAdd(dst + 1, c.y / inv_L_div);
                atomicAdd(dst + 2, c.z / inv_L_div);
            }
        } else if (slot >= 0) {
            float *dst = reinterpret_cast<float *>(wf.acc + 2 * slot + 1);
            if (c.x != 0.f)
                atomicAdd(dst + 0, c.x / inv_L_div);
            if (c.y != 0.f)
                atomicAdd(dst + 1, c.y / inv_L_div);
            if (c.z != 0.f)
                atomicAdd(dst + 2, c.z / inv_L_div);
        }
    }
    if (COUNT)
        flush_counts(tc, nrays, wf.trace_counts + 3, ntraced, wf.trace_counts + 6);
}

// ---- counting sort of a queue by key (PT_SORT_BUCKETS buckets) ------------------------------------------------------
// Three launches: histogram (one atomic per distinct key of a warp), exclusive scan of the counters in chunks of
// PT_SORT_CHUNK (chunk totals aside), scatter (the chunk totals are scanned again by every block: <= 1024 values).
// Entries of one bucket keep no particular order; the film does not depend on queue order (per-slot sums, Philox
// keyed by pixel and sample).
__global__ void __launch_bounds__(kThreads) k_sort_hist(const uint32_t *__restrict__ keys, const int32_t *__restrict__ count,
                                                        uint32_t *__restrict__ hist)
{
    const int n = *count;
    const int lane = threadIdx.x & 31;
    for (int base = blockIdx.x * kThreads; base < n; base += gridDim.x * kThreads) {
        const int i = base + threadIdx.x;
        const bool live = i < n;
        const uint32_t key = live ? keys[i] : 0xffffffffu;
        const unsigned peers = __match_any_sync(0xffffffffu, key);
        PT_CHECK(!live || key < (uint32_t)PT_SORT_BUCKETS);
        if (live && lane == __ffs(peers) - 1)
            atomicAdd(hist + key, (uint32_t)__popc(peers));
    }
}

__global__ void __launch_bounds__(1024) k_sort_scan(uint32_t *__restrict__ hist)
{
    // one block per chunk of PT_SORT_CHUNK counters, two per thread
    __shared__ uint32_t s_warp[32];
    uint32_t *h = hist + (size_t)blockIdx.x * PT_SORT_CHUNK;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const uint32_t a = h[2 * t], b = h[2 * t + 1];
    uint32_t v = a + b, incl = v;
    for (int off = 1; off < 32; off <<= 1) {
        const uint32_t u = __shfl_up_sync(0xffffffffu, incl, off);
        if (lane >= off)
            incl += u;
    }
    if (lane == 31)
        s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = s_warp[lane], wi = w;
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t u = __shfl_up_sync(0xffffffffu, wi, off);
            if (lane >= off)
                wi += u;
        }
        s_warp[lane] = wi - w; // exclusive
        if (lane == 31)
            hist[PT_SORT_BUCKETS + blockIdx.x] = wi; // chunk total
    }
    __syncthreads();
    const uint32_t excl = s_warp[warp] + incl - v;
    h[2 * t] = excl;
    h[2 * t + 1] = excl + a;
}

__global__ void __launch_bounds__(kThreads) k_sort_scatter(const uint32_t *__restrict__ keys, const int32_t *__restrict__ raw,
                                                           const int32_t *__restrict__ count, uint32_t *__restrict__ hist,
                                                           int32_t *__restrict__ dst)
{
    constexpr int kChunks = PT_SORT_BUCKETS / PT_SORT_CHUNK;
    static_assert(kChunks <= 4096, "chunk totals are scanned by one block");
    __shared__ uint32_t s_chunk[kChunks];
    __shared__ uint32_t s_part[kThreads / 32];
    const int n = *count;
    if ((int)(blockIdx.x * kThreads) >= n)
        return;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    {
        // exclusive scan of the chunk totals: kChunks / kThreads consecutive values per thread
        constexpr int per = (kChunks + kThreads - 1) / kThreads;
        uint32_t v[per], sum = 0;
#pragma unroll
        for (int k = 0; k < per; k++) {
            const int c = threadIdx.x * per + k;
            v[k] = c < kChunks ? hist[PT_SORT_BUCKETS + c] : 0u;
            sum += v[k];
        }
        uint32_t incl = sum;
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t u = __shfl_up_sync(0xffffffffu, incl, off);
            if (lane >= off)
                incl += u;
        }
        if (lane == 31)
            s_part[warp] = incl;
        __syncthreads();
        uint32_t before = 0;
        for (int w2 = 0; w2 < warp; w2++)
            before += s_part[w2];
        uint32_t run = before + incl - sum;
#pragma unroll
        for (int k = 0; k < per; k++) {
            const int c = threadIdx.x * per + k;
            if (c < kChunks)
                s_chunk[c] = run;
            run += v[k];
        }
        __syncthreads();
    }
    for (int base = blockIdx.x * kThreads; base < n; base += gridDim.x * kThreads) {
        const int i = base + threadIdx.x;
        const bool live = i < n;
        const uint32_t key = live ? keys[i] : 0xffffffffu;
        const unsigned peers = __match_any_sync(0xffffffffu, key);
        const int leader = __ffs(peers) - 1;
        uint32_t first = 0;
        if (live && lane == leader)
            first = atomicAdd(hist + key, (uint32_t)__popc(peers)) + s_chunk[key / PT_SORT_CHUNK];
        first = __shfl_sync(0xffffffffu, first, leader);
        PT_CHECK(!live || first + (uint32_t)__popc(peers & ((1u << lane) - 1u)) < (uint32_t)n);
        if (live)
            dst[first + __popc(peers & ((1u << lane) - 1u))] = raw[i];
    }
}

// framebuffer[j][i] += de_nan(col): one thread per pixel walks the samples of the batch in order, so the film is
// bit-reproducible.  Thread 0 of the LAST block to finish tallies the counters and advances the batch control.
__global__ void __launch_bounds__(kThreads) k_accumulate(RenderParams rp, Wavefront wf, Film film, int n_rows)
{
    const int nsamp = wf.ctl->nsamp;
    const int pix = blockIdx.x * kThreads + threadIdx.x;
    if (pix < wf.npix) {
        float4 acc = film.accum[pix];
        float4 cs = make_float4(0.f, 0.f, 0.f, 0.f), cq = make_float4(0.f, 0.f, 0.f, 0.f);
        const bool stats = film.csum != nullptr;
        if (stats) {
            cs = film.csum[pix];
            cq = film.csq[pix];
        }
        for (int k = 0; k < nsamp; k++) {
            float4 v = ld_state(&wf.acc[2 * ((size_t)k * wf.npix + pix) + 1]);
            float x = is_nan(v.x) ? 0.f : v.x, y = is_nan(v.y) ? 0.f : v.y, z = is_nan(v.z) ? 0.f : v.z; // de_nan
            acc.x += x;
            acc.y += y;
            acc.z += z;
            if (stats) {
                // two-sided: a Perlin albedo can be negative (texture.h:191) and the estimator has no finite variance (Q4)
                const float lo = -film.clamp;
                float cx = fmaxf(fminf(x, film.clamp), lo), cy = fmaxf(fminf(y, film.clamp), lo), cz = fmaxf(fminf(z, film.clamp), lo);
                cs.x += cx;
                cs.y += cy;
                cs.z += cz;
                cq.x += cx * cx;
                cq.y += cy * cy;
                cq.z += cz * cz;
            }
        }
        film.accum[pix] = acc;
        if (stats) {
            film.csum[pix] = cs;
            film.csq[pix] = cq;
        }
    }
}

__global__ void k_tally(RenderParams rp, Wavefront wf, int n_rows)
{
    if (threadIdx.x != 0 || blockIdx.x != 0)
        return;
    BatchCtl *c = wf.ctl;
    unsigned long long ext = 0, shadow = 0, ref = 0;
    for (int b = 0; b < n_rows; b++) {
        const int32_t *row = wf.counters + b * CNT_ROW;
        unsigned long long n_ext = (unsigned long long)row[CNT_EXT];
        unsigned long long hits = n_ext - (unsigned long long)row[CNT_KIND0 + SK_MISS];
        ext += n_ext;
        if (rp.integrator == B200PT_INTEGRATOR_INEEPT) {
            const unsigned long long skipped = (unsigned long long)row[CNT_NEE_SKIP] * (unsigned long long)rp.light_samples;
            shadow += (unsigned long long)row[CNT_NEE] * (unsigned long long)rp.light_samples + skipped;
            if (wf.trace_counts)
                wf.trace_counts[3] += skipped; // work counters: these queries cost no box and no primitive test
            ref += n_ext + hits * (unsigned long long)rp.light_samples; // integrator.h:192,247
        } else if (b > 0) {
            ref += n_ext; // one count per scatter (integrator.h:45)
        }
    }
    c->rays_extension += ext;
    c->rays_shadow += shadow;
    c->rays_reference += ref;
    c->samples_done += (unsigned long long)c->nsamp * (unsigned long long)wf.npix;
    c->sample_base += c->nsamp;
    int next = c->remaining < c->batch_samples ? c->remaining : c->batch_samples;
    c->remaining -= next;
    c->nsamp = next;
}

// ---------------------------------------------------------------------------------------------------------------

template <bool COUNT, bool MEDIA, bool SMALL>
__global__ void __launch_bounds__(kThreads) k_trace_batch(PtSceneView s, const float *__restrict__ rays, size_t n,
                                                          const float *__restrict__ medium_xi,
                                                          b200pt_hit *__restrict__ hits, unsigned long long *counts)
{
    __shared__ SmallScene s_small; // dropped by the compiler when SMALL is false
    if (SMALL)
        s_small.stage(s);
    unsigned long long boxes = 0, prims = 0, nhit = 0;
    for (size_t base = (size_t)blockIdx.x * kThreads; base < n; base += (size_t)gridDim.x * kThreads) {
        const size_t i = base + threadIdx.x;
        const bool live = i < n;
        Traversal T;
        TraversalStack T_stack;
        T.stk = SMALL ? nullptr : &T_stack; // the table walk of a small scene has no pending nodes
        MediumRng mr;
        mr.xi = -1.f;
        TraceCounters tc;
        tc.boxes = tc.prims = 0;
        if (live) {
            const float4 *r4 = reinterpret_cast<const float4 *>(rays + i * 8);
            float4 a = __ldg(r4), b = __ldg(r4 + 1);
            mr.key.seed_lo = 0;
            mr.key.seed_hi = 0;
            mr.key.pixel = (uint32_t)(i & 0xffffffffull);
            mr.key.sample = (uint32_t)(i >> 32);
            mr.key.sample_hi = 0;
            mr.stage = stage_word(0, RNG_BATCH_MEDIUM, 0);
            mr.xi = medium_xi ? medium_xi[i] : -1.f;
            T.init(v3(a.x, a.y, a.z), v3(a.w, b.x, b.y), b.z, b.w);
        }
        if (SMALL) {
            if (live)
                T.template run_small<COUNT, MEDIA>(s.small, s.inst, s.n_small, mr, &tc);
        } else {
            T.template run_warp<COUNT, MEDIA>(s, live, mr, &tc); // all 32 lanes take part in the votes
        }
        if (!live)
            continue;
        Hit h;
        bool ok = T.template result<MEDIA>(s, h);
        if (COUNT) {
            boxes += tc.boxes;
            prims += tc.prims;
            nhit += ok ? 1 : 0;
        }
        if (hits) {
            b200pt_hit o;
            o.instance = ok ? h.inst : -1;
            o.t = ok ? h.t : 0.f;
            o.p[0] = ok ? h.p.x : 0.f, o.p[1] = ok ? h.p.y : 0.f, o.p[2] = ok ? h.p.z : 0.f;
            o.n[0] = ok ? h.n.x : 0.f, o.n[1] = ok ? h.n.y : 0.f, o.n[2] = ok ? h.n.z : 0.f;
            o.u = ok ? h.u : 0.f;
            o.v = ok ? h.v : 0.f;
            hits[i] = o;
        }
    }
    if (COUNT) {
        for (int off = 16; off > 0; off >>= 1) {
            boxes += __shfl_down_sync(0xffffffffu, boxes, off);
            prims += __shfl_down_sync(0xffffffffu, prims, off);
            nhit += __shfl_down_sync(0xffffffffu, nhit, off);
        }
        if ((threadIdx.x & 31) == 0) {
            atomicAdd(counts + 0, boxes);
            atomicAdd(counts + 1, prims);
            atomicAdd(counts + 2, nhit);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------

// the generator on its own: n x {counter[4], key[2]} -> n x 4 words (known-answer tests of pt_rng.h on the device)
__global__ void k_philox(const uint32_t *__restrict__ in, size_t n, uint32_t *__restrict__ out)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        U4 c;
        c.x = in[6 * i], c.y = in[6 * i + 1], c.z = in[6 * i + 2], c.w = in[6 * i + 3];
        const U4 r = philox4x32_10(c, in[6 * i + 4], in[6 * i + 5]);
        out[4 * i] = r.x, out[4 * i + 1] = r.y, out[4 * i + 2] = r.z, out[4 * i + 3] = r.w;
    }
}

// calculate_luminance (helpers.h:146-168): max and total over pixels of |de_nan(sum) / n|
__global__ void __launch_bounds__(kThreads) k_luminance(const float4 *__restrict__ accum, int npix, float inv_n,
                                                        float *max_total)
{
    float mx = 0.f, tot = 0.f;
    for (int i = blockIdx.x * kThreads + threadIdx.x; i < npix; i += gridDim.x * kThreads) {
        float4 v = accum[i];
        V3 c = v3(is_nan(v.x) ? 0.f : v.x, is_nan(v.y) ? 0.f : v.y, is_nan(v.z) ? 0.f : v.z) * inv_n;
        float f = length(c);
        tot += f;
        mx = fmaxf(mx, f);
    }
    for (int off = 16; off > 0; off >>= 1) {
        mx = fmaxf(mx, __shfl_down_sync(0xffffffffu, mx, off));
        tot += __shfl_down_sync(0xffffffffu, tot, off);
    }
    __shared__ float s_mx[kWarps], s_tot[kWarps];
    if ((threadIdx.x & 31) == 0) {
        s_mx[threadIdx.x >> 5] = mx;
        s_tot[threadIdx.x >> 5] = tot;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < kWarps; w++) {
            mx = fmaxf(mx, s_mx[w]);
            tot += s_tot[w];
        }
        // lengths are >= 0: their float order equals the order of their bit patterns
        atomicMax(reinterpret_cast<int *>(max_total), __float_as_int(mx));
        atomicAdd(max_total + 1, tot);
    }
}

// Multi-GPU: the films of all devices summed into `dst` on this device, in device order (a fixed order: the total is
// bit-reproducible), reading the other devices' buffers in place over NVLink peer loads (or their staged copies).  With
// max_total != null the same pass reduces calculate_luminance's max / total (helpers.h:146-168) of the summed film, so the
// tonemap that follows needs no second pass over it.
__global__ void __launch_bounds__(kThreads) k_combine(FilmSources fs, int npix, float4 *__restrict__ dst, float inv_n,
                                                      float *max_total)
{
    float mx = 0.f, tot = 0.f;
    for (int i = blockIdx.x * kThreads + threadIdx.x; i < npix; i += gridDim.x * kThreads) {
        float4 a = fs.src[0][i];
        for (int d = 1; d < fs.n; d++) {
            const float4 b = fs.src[d][i];
            a.x += b.x, a.y += b.y, a.z += b.z, a.w += b.w;
        }
        dst[i] = a;
        if (max_total) {
            V3 c = v3(is_nan(a.x) ? 0.f : a.x, is_nan(a.y) ? 0.f : a.y, is_nan(a.z) ? 0.f : a.z) * inv_n;
            const float f = length(c);
            tot += f;
            mx = fmaxf(mx, f);
        }
    }
    if (!max_total)
        return;
    for (int off = 16; off > 0; off >>= 1) {
        mx = fmaxf(mx, __shfl_down_sync(0xffffffffu, mx, off));
        tot += __shfl_down_sync(0xffffffffu, tot, off);
    }
    __shared__ float s_mx[kWarps], s_tot[kWarps];
    if ((threadIdx.x & 31) == 0) {
        s_mx[threadIdx.x >> 5] = mx;
        s_tot[threadIdx.x >> 5] = tot;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < kWarps; w++) {
            mx = fmaxf(mx, s_mx[w]);
            tot += s_tot[w];
        }
        atomicMax(reinterpret_cast<int *>(max_total), __float_as_int(mx));
        atomicAdd(max_total + 1, tot);
    }
}

// film sums float4 -> packed rgb floats (what b200pt_render_read hands to the caller)
__global__ void __launch_bounds__(kThreads) k_pack_rgb(const float4 *__restrict__ accum, int npix, float *__restrict__ rgb)
{
    for (int i = blockIdx.x * kThreads + threadIdx.x; i < npix; i += gridDim.x * kThreads) {
        float4 v = accum[i];
        rgb[3 * (size_t)i + 0] = v.x;
        rgb[3 * (size_t)i + 1] = v.y;
        rgb[3 * (size_t)i + 2] = v.z;
    }
}

__device__ __forceinline__ float uncharted2(float x)
{
    // tonemap.h:4-15
    const float A = 0.15f, B = 0.50f, C = 0.10f, D = 0.20f, E = 0.02f, F = 0.30f;
    return ((x * (A * x + C * B) + D * E) / (x * (A * x + B) + D * F)) - E / F;
}

__device__ __forceinline__ float to_srgb(float c)
{
    // helpers.h:78-88
    if (c < 0.0031308f)
        return 323 * c / 25;
    return (211 * powf(c, 5.0f / 12) - 11) / 200;
}

__device__ __forceinline__ float clamp01(float x) { return x > 1.f ? 1.f : (x < 0.f ? 0.f : x); }

// output_to_file (renderer.h:24-55): rows are written top (j = H-1) first
__global__ void __launch_bounds__(kThreads) k_tonemap(const float4 *__restrict__ accum, int width, int height,
                                                      float inv_n, float exposure_mul, float max_lum,
                                                      uint8_t *__restrict__ rgb)
{
    const int npix = width * height;
    const float white = uncharted2(max_lum);
    for (int o = blockIdx.x * kThreads + threadIdx.x; o < npix; o += gridDim.x * kThreads) {
        int row = o / width, i = o - row * width;
        int j = height - 1 - row;
        float4 v = accum[j * width + i];
        float c[3] = {v.x, v.y, v.z};
#pragma unroll
        for (int k = 0; k < 3; k++) {
            float x = c[k] * inv_n;
            x *= exposure_mul;
            x = clamp01(uncharted2(x) / white);
            x = 255 * to_srgb(x);
            rgb[(size_t)o * 3 + k] = (uint8_t)(int)x;
        }
    }
}

// ---- allocation-free tonemap (b200pt_tonemap_static) -----------------------------------------------------------------
// calculate_luminance + output_to_file for callers that may not allocate (CUDA graph capture, serving loops): the scratch
// of the luminance reduce lives in a file-scope __device__ array instead of a cudaMalloc'ed block.  Layout:
//   [2 b], [2 b + 1]              max / total of |sum / n| over the pixels of tile b (PT_STATIC_TILE pixels each)
//   [2 PT_STATIC_MAX_TILES + 0/1] max / total over the film (written by k_static_lum_final, read by k_static_tonemap)
// Every cell a call reads was written earlier in the same call, so nothing has to start at zero and the contents the
// previous call left behind are never seen; the totals are summed in a fixed order (bit-reproducible, unlike the atomics
// of k_luminance).  Every GPU that loads the library holds the array (PT_STATIC_SCRATCH_FLOATS * 4 bytes = 128.25 KiB)
// until the process exits.  All calls share it: two calls that overlap on one GPU race.
constexpr int kStaticTile = PT_STATIC_TILE;
static __device__ __align__(256) float g_static_scratch[PT_STATIC_SCRATCH_FLOATS];

__global__ void __launch_bounds__(kThreads) k_static_lum_partials(const float4 *__restrict__ accum, int npix, float inv_n,
                                                                  float *__restrict__ scratch)
{
    float mx = 0.f, tot = 0.f;
    const int base = blockIdx.x * kStaticTile;
#pragma unroll
    for (int k = 0; k < kStaticTile / kThreads; k++) { // fixed order within a thread
        const int i = base + k * kThreads + threadIdx.x;
        if (i < npix) {
            const float4 v = accum[i];
            V3 c = v3(is_nan(v.x) ? 0.f : v.x, is_nan(v.y) ? 0.f : v.y, is_nan(v.z) ? 0.f : v.z) * inv_n;
            const float f = length(c);
            tot += f;
            mx = fmaxf(mx, f);
        }
    }
    for (int off = 16; off > 0; off >>= 1) {
        mx = fmaxf(mx, __shfl_down_sync(0xffffffffu, mx, off));
        tot += __shfl_down_sync(0xffffffffu, tot, off);
    }
    __shared__ float s_mx[kWarps], s_tot[kWarps];
    if ((threadIdx.x & 31) == 0) {
        s_mx[threadIdx.x >> 5] = mx;
        s_tot[threadIdx.x >> 5] = tot;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < kWarps; w++) {
            mx = fmaxf(mx, s_mx[w]);
            tot += s_tot[w];
        }
        scratch[2 * blockIdx.x] = mx;
        scratch[2 * blockIdx.x + 1] = tot;
    }
}

// one block: the tile partials in a fixed order -> the two result cells
__global__ void __launch_bounds__(kThreads) k_static_lum_final(int n_tiles, float *__restrict__ scratch)
{
    float mx = 0.f, tot = 0.f;
    for (int b = threadIdx.x; b < n_tiles; b += kThreads) {
        mx = fmaxf(mx, scratch[2 * b]);
        tot += scratch[2 * b + 1];
    }
    for (int off = 16; off > 0; off >>= 1) {
        mx = fmaxf(mx, __shfl_down_sync(0xffffffffu, mx, off));
        tot += __shfl_down_sync(0xffffffffu, tot, off);
    }
    __shared__ float s_mx[kWarps], s_tot[kWarps];
    if ((threadIdx.x & 31) == 0) {
        s_mx[threadIdx.x >> 5] = mx;
        s_tot[threadIdx.x >> 5] = tot;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < kWarps; w++) {
            mx = fmaxf(mx, s_mx[w]);
            tot += s_tot[w];
        }
        scratch[2 * PT_STATIC_MAX_TILES] = mx;
        scratch[2 * PT_STATIC_MAX_TILES + 1] = tot;
    }
}

// output_to_file (renderer.h:24-55) with the white point read from the scratch on the device: no host round trip
__global__ void __launch_bounds__(kThreads) k_static_tonemap(const float4 *__restrict__ accum, int width, int height,
                                                             float inv_n, float exposure_mul,
                                                             const float *__restrict__ scratch, uint8_t *__restrict__ rgb)
{
    const int npix = width * height;
    const float white = uncharted2(scratch[2 * PT_STATIC_MAX_TILES]);
    for (int o = blockIdx.x * kThreads + threadIdx.x; o < npix; o += gridDim.x * kThreads) {
        int row = o / width, i = o - row * width;
        int j = height - 1 - row;
        float4 v = accum[j * width + i];
        float c[3] = {v.x, v.y, v.z};
#pragma unroll
        for (int k = 0; k < 3; k++) {
            float x = c[k] * inv_n;
            x *= exposure_mul;
            x = clamp01(uncharted2(x) / white);
            x = 255 * to_srgb(x);
            rgb[(size_t)o * 3 + k] = (uint8_t)(int)x;
        }
    }
}

} // namespace

// ---------------------------------------------------------------------------------------------------------------

void launch_raygen(const PtSceneView &s, const RenderParams &rp, const Wavefront &wf, const LaunchCfg &lc, cudaStream_t st)
{
    k_raygen<<<lc.blocks, kThreads, 0, st>>>(s, rp, wf);
}

void launch_extend(const PtSceneView &s, const RenderParams &rp, const Wavefront &wf, int bounce, const LaunchCfg &lc,
                   cudaStream_t st)
{
    const int blocks = lc.blocks * (kThreads / kExtThreads);
    const bool small = s.n_small > 0;
#define PT_LAUNCH_EXT(C, M)                                                        \
    do {                                                                           \
        if (small)                                                                 \
            k_extend<C, M, true><<<blocks, kExtThreads, 0, st>>>(s, rp, wf, bounce);  \
        else                                                                       \
            k_extend<C, M, false><<<blocks, kExtThreads, 0, st>>>(s, rp, wf, bounce); \
    } while (0)
    if (lc.count_work)
        PT_LAUNCH_EXT(true, true);
    else if (s.n_media > 0)
        PT_LAUNCH_EXT(false, true);
    else
        PT_LAUNCH_EXT(false, false);
#undef PT_LAUNCH_EXT
}

void launch_shade(const PtSceneView &s, const RenderParams &rp, const Wavefront &wf, int bounce, int kind,
                  const LaunchCfg &lc, cudaStream_t st)
{
    const bool nee = rp.integrator == B200PT_INTEGRATOR_INEEPT;
#define PT_LAUNCH_SHADE_F(K, F)                                                     \
    do {                                                                           \
        if (rp.corrected)                                                          \
            k_shade<K, false, true, F><<<lc.blocks * kThreads / kShadeThreads, kShadeThreads, 0, st>>>(s, rp, wf, bounce);  \
        else if (nee)                                                              \
            k_shade<K, true, false, F><<<lc.blocks * kThreads / kShadeThreads, kShadeThreads, 0, st>>>(s, rp, wf, bounce);  \
        else                                                                       \
            k_shade<K, false, false, F><<<lc.blocks * kThreads / kShadeThreads, kShadeThreads, 0, st>>>(s, rp, wf, bounce); \
    } while (0)
#define PT_LAUNCH_SHADE(K)                                                  \
    case K:                                                                 \
        if (bounce == 0)                                                           \
            PT_LAUNCH_SHADE_F(K, true);                                            \
        else                                                                       \
            PT_LAUNCH_SHADE_F(K, false);                                           \
        break;
    switch (kind) {
        PT_LAUNCH_SHADE(SK_MISS)
        PT_LAUNCH_SHADE(SK_DIFFUSE)
        PT_LAUNCH_SHADE(SK_DIELECTRIC)
        PT_LAUNCH_SHADE(SK_LIGHT)
        PT_LAUNCH_SHADE(SK_ISOTROPIC)
    }
#undef PT_LAUNCH_SHADE
#undef PT_LAUNCH_SHADE_F
}

void launch_nee(const PtSceneView &s, const RenderParams &rp, const Wavefront &wf, int bounce, const LaunchCfg &lc,
                cudaStream_t st)
{
    const bool any = s.n_emitters <= PT_MAX_DIRECT_EMITTERS;
    const bool small = any && s.n_small > 0;
#define PT_LAUNCH_NEE(C, M, X)                                                          \
    do {                                                                               \
        if (small)                                                                            \
            k_nee<C, M, true, true, X><<<lc.blocks * (kThreads / kNeeThreads), kNeeThreads, 0, st>>>(s, rp, wf, bounce);   \
        else if (any)                                                                         \
            k_nee<C, M, true, false, X><<<lc.blocks * (kThreads / kNeeThreads), kNeeThreads, 0, st>>>(s, rp, wf, bounce);  \
        else                                                                                  \
            k_nee<C, M, false, false, X><<<lc.blocks * (kThreads / kNeeThreads), kNeeThreads, 0, st>>>(s, rp, wf, bounce); \
    } while (0)
    if (rp.corrected && s.n_media > 0)
        PT_LAUNCH_NEE(false, true, true);
    else if (rp.corrected)
        PT_LAUNCH_NEE(false, false, true);
    else if (lc.count_work)
        PT_LAUNCH_NEE(true, true, false);
    else if (s.n_media > 0)
        PT_LAUNCH_NEE(false, true, false);
    else
        PT_LAUNCH_NEE(false, false, false);
#undef PT_LAUNCH_NEE
}

void launch_sort_queue(const Wavefront &wf, const uint32_t *keys, const int32_t *raw, int32_t *dst, int count_index,
                       const LaunchCfg &lc, cudaStream_t st)
{
    const int32_t *count = wf.counters + count_index;
    cudaMemsetAsync(wf.sort_hist, 0, sizeof(uint32_t) * PT_SORT_BUCKETS, st);
    k_sort_hist<<<lc.blocks, kThreads, 0, st>>>(keys, count, wf.sort_hist);
    k_sort_scan<<<PT_SORT_BUCKETS / PT_SORT_CHUNK, PT_SORT_CHUNK / 2, 0, st>>>(wf.sort_hist);
    k_sort_scatter<<<lc.blocks, kThreads, 0, st>>>(keys, raw, count, wf.sort_hist, dst);
}

void launch_accumulate(const RenderParams &rp, const Wavefront &wf, const Film &film, int n_rows, cudaStream_t st)
{
    int blocks = (wf.npix + kThreads - 1) / kThreads;
    k_accumulate<<<blocks, kThreads, 0, st>>>(rp, wf, film, n_rows);
    k_tally<<<1, 32, 0, st>>>(rp, wf, n_rows);
}

void launch_trace_batch(const PtSceneView &s, const float *rays, size_t n, const float *medium_xi, b200pt_hit *hits,
                        unsigned long long *counts, cudaStream_t st)
{
    if (n == 0)
        return;
    size_t want = (n + kThreads - 1) / kThreads;
    int blocks = (int)(want < (size_t)(148 * 16) ? want : (size_t)(148 * 16));
    const bool small = s.n_small > 0;
#define PT_LAUNCH_TB(C, M, cnt)                                                                      \
    do {                                                                                             \
        if (small)                                                                                   \
            k_trace_batch<C, M, true><<<blocks, kThreads, 0, st>>>(s, rays, n, medium_xi, hits, cnt);  \
        else                                                                                         \
            k_trace_batch<C, M, false><<<blocks, kThreads, 0, st>>>(s, rays, n, medium_xi, hits, cnt); \
    } while (0)
    if (counts)
        PT_LAUNCH_TB(true, true, counts);
    else if (s.n_media > 0)
        PT_LAUNCH_TB(false, true, nullptr);
    else
        PT_LAUNCH_TB(false, false, nullptr);
#undef PT_LAUNCH_TB
}

void launch_philox(const uint32_t *in, size_t n, uint32_t *out, cudaStream_t st)
{
    if (n)
        k_philox<<<(unsigned)std::min<size_t>((n + 255) / 256, 148 * 8), 256, 0, st>>>(in, n, out);
}

void launch_luminance(const float4 *accum, int npix, float inv_n, float *max_total, cudaStream_t st)
{
    cudaMemsetAsync(max_total, 0, 2 * sizeof(float), st);
    int blocks = (npix + kThreads - 1) / kThreads;
    if (blocks > 148 * 8)
        blocks = 148 * 8;
    k_luminance<<<blocks, kThreads, 0, st>>>(accum, npix, inv_n, max_total);
}

void launch_combine(const FilmSources &fs, int npix, float4 *dst, float inv_n, float *max_total, cudaStream_t st)
{
    if (max_total)
        cudaMemsetAsync(max_total, 0, 2 * sizeof(float), st);
    int blocks = (npix + kThreads - 1) / kThreads;
    if (blocks > 148 * 8)
        blocks = 148 * 8;
    k_combine<<<blocks, kThreads, 0, st>>>(fs, npix, dst, inv_n, max_total);
}

void launch_pack_rgb(const float4 *accum, int npix, float *rgb, cudaStream_t st)
{
    int blocks = (npix + kThreads - 1) / kThreads;
    if (blocks > 148 * 8)
        blocks = 148 * 8;
    k_pack_rgb<<<blocks, kThreads, 0, st>>>(accum, npix, rgb);
}

void launch_tonemap(const float4 *accum, int width, int height, float inv_n, float exposure_mul, float max_lum,
                    uint8_t *rgb, cudaStream_t st)
{
    int npix = width * height;
    int blocks = (npix + kThreads - 1) / kThreads;
    if (blocks > 148 * 8)
        blocks = 148 * 8;
    k_tonemap<<<blocks, kThreads, 0, st>>>(accum, width, height, inv_n, exposure_mul, max_lum, rgb);
}

// The device address of g_static_scratch on `device` (which must be current).  In host code the array's name is only a
// host-side stand-in for the symbol: the address a kernel can use comes from cudaGetSymbolAddress, once per device.
static cudaError_t static_scratch_address(int device, float **out)
{
    static std::mutex mu;
    static float *known[64] = {nullptr};
    if (device < 0 || device >= 64)
        return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> g(mu);
    if (!known[device]) {
        float *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void **)&p, g_static_scratch);
        if (e != cudaSuccess)
            return e;
        known[device] = p;
    }
    *out = known[device];
    return cudaSuccess;
}

cudaError_t launch_tonemap_static(int device, const float4 *accum, int width, int height, float inv_n, float exposure_mul,
                                  uint8_t *rgb, float *h_max_total /* [2] host, or null */, cudaStream_t st)
{
    float *scratch = nullptr;
    cudaError_t e = static_scratch_address(device, &scratch);
    if (e != cudaSuccess)
        return e;
    const int npix = width * height;
    const int n_tiles = (npix + kStaticTile - 1) / kStaticTile;
    k_static_lum_partials<<<n_tiles, kThreads, 0, st>>>(accum, npix, inv_n, scratch);
    k_static_lum_final<<<1, kThreads, 0, st>>>(n_tiles, scratch);
    int blocks = (npix + kThreads - 1) / kThreads;
    if (blocks > 148 * 8)
        blocks = 148 * 8;
    k_static_tonemap<<<blocks, kThreads, 0, st>>>(accum, width, height, inv_n, exposure_mul, scratch, rgb);
    e = cudaGetLastError();
    if (e != cudaSuccess || !h_max_total)
        return e;
    e = cudaMemcpyAsync(h_max_total, scratch + 2 * PT_STATIC_MAX_TILES, 2 * sizeof(float), cudaMemcpyDeviceToHost, st);
    if (e != cudaSuccess)
        return e;
    return cudaStreamSynchronize(st);
}

} // namespace pt
