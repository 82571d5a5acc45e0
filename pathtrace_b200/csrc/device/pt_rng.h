// pt_rng.h -- counter-based random numbers: Philox4x32-10 (Salmon et al., SC'11), written out from the paper.
// Replaces the reference's single shared std::mt19937 (random.h:9-15): every draw is a pure function of
// (seed; pixel, sample, stage word, index), so results do not depend on scheduling, batch size or GPU count.
#pragma once
#include "pt_math.h"

namespace pt {

struct U4 {
    uint32_t x, y, z, w;
};
struct F4 {
    float x, y, z, w;
};

PT_HD uint32_t mulhi32(uint32_t a, uint32_t b)
{
#if defined(__CUDA_ARCH__)
    return __umulhi(a, b);
#else
    return (uint32_t)(((uint64_t)a * (uint64_t)b) >> 32);
#endif
}

PT_HD U4 philox4x32_10(U4 ctr, uint32_t k0, uint32_t k1)
{
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int r = 0; r < 10; r++) {
        uint32_t hi0 = mulhi32(M0, ctr.x), lo0 = M0 * ctr.x;
        uint32_t hi1 = mulhi32(M1, ctr.z), lo1 = M1 * ctr.z;
        U4 n;
        n.x = hi1 ^ ctr.y ^ k0;
        n.y = lo1;
        n.z = hi0 ^ ctr.w ^ k1;
        n.w = lo0;
        ctr = n;
        k0 += W0;
        k1 += W1;
    }
    return ctr;
}

// 24-bit uniform in [0,1)
PT_HD float u01(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }

// stage words: (bounce << 16) | (stage << 12) | k   with bounce < 2^16, stage < 16, k < 2^12 (k = light sample index;
// b200pt_render_start caps light_samples at 1024 and max_bounces at 4096, so the three fields never overlap)
enum {
    RNG_CAMERA = 0,      // idx 0: {pixel jitter u, v, lens r1, lens r2}
    RNG_EXT_MEDIUM = 1,  // idx = medium slot: {xi, xi2 (BVH-duplicated medium)}
    RNG_SHADE = 2,       // idx 0: {r1, r2, r3, russian roulette}
    RNG_NEE = 3,         // k = light sample: {light pick, s1, s2}
    RNG_NEE_MEDIUM = 4,  // k = light sample, idx = medium slot
    RNG_PATH_TRACE = 6,  // idx 0: {record this camera sample? (renderer.h:652)}
    RNG_BATCH_MEDIUM = 5 // b200pt_trace_batch without explicit variates: (ray lo, ray hi) as (pixel, sample)
};

// The GLOBAL sample index is 64 bits wide (config.sample_offset is an int64): its low word is one counter word, its
// high bits (at most 24: b200pt_render_start refuses offsets beyond 2^56) share the fourth counter word with `idx`
// (< 256: half a medium slot index, or 0).  Two samples 2^32 apart draw different numbers.
struct RngKey {
    uint32_t seed_lo, seed_hi;
    uint32_t pixel, sample;
    uint32_t sample_hi;
};

#define PT_RNG_MAX_K 4096          /* exclusive bound of the k field of a stage word */
#define PT_RNG_MAX_SAMPLE_BITS 56  /* global sample indices are < 2^56 */

PT_HD uint32_t stage_word(int bounce, int stage, int k) { return ((uint32_t)bounce << 16) | ((uint32_t)stage << 12) | (uint32_t)k; }

PT_HD void key_set_sample(RngKey &key, long long global_sample)
{
    key.sample = (uint32_t)((unsigned long long)global_sample & 0xffffffffull);
    key.sample_hi = (uint32_t)((unsigned long long)global_sample >> 32);
}

PT_HD F4 rng4(const RngKey &key, uint32_t stage, uint32_t idx)
{
    U4 c;
    c.x = key.pixel;
    c.y = key.sample;
    c.z = stage;
    c.w = (key.sample_hi << 8) | idx;
    U4 r = philox4x32_10(c, key.seed_lo, key.seed_hi);
    F4 f;
    f.x = u01(r.x);
    f.y = u01(r.y);
    f.z = u01(r.z);
    f.w = u01(r.w);
    return f;
}

} // namespace pt
