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

// stage words: (bounce << 16) | (stage << 8) | k
enum {
    RNG_CAMERA = 0,      // idx 0: {pixel jitter u, v, lens r1, lens r2}
    RNG_EXT_MEDIUM = 1,  // idx = medium slot: {xi, xi2 (BVH-duplicated medium)}
    RNG_SHADE = 2,       // idx 0: {r1, r2, r3, russian roulette}
    RNG_NEE = 3,         // k = light sample: {light pick, s1, s2}
    RNG_NEE_MEDIUM = 4,  // k = light sample, idx = medium slot
    RNG_PATH_TRACE = 6,  // idx 0: {record this camera sample? (renderer.h:652)}
    RNG_BATCH_MEDIUM = 5 // b200pt_trace_batch without explicit variates: (ray lo, ray hi) as (pixel, sample)
};

struct RngKey {
    uint32_t seed_lo, seed_hi;
    uint32_t pixel, sample;
};

PT_HD uint32_t stage_word(int bounce, int stage, int k) { return ((uint32_t)bounce << 16) | ((uint32_t)stage << 8) | (uint32_t)k; }

PT_HD F4 rng4(const RngKey &key, uint32_t stage, uint32_t idx)
{
    U4 c;
    c.x = key.pixel;
    c.y = key.sample;
    c.z = stage;
    c.w = idx;
    U4 r = philox4x32_10(c, key.seed_lo, key.seed_hi);
    F4 f;
    f.x = u01(r.x);
    f.y = u01(r.y);
    f.z = u01(r.z);
    f.w = u01(r.w);
    return f;
}

} // namespace pt
