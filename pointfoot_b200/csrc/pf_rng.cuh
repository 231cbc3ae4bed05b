// Counter-based RNG contract, device statement.  The numpy statement used by the checker is oracle/rng.py;
// tests/test_rng.py compares the two bit for bit.  Every draw is a pure function of
// (seed, step, purpose, global_env_id, lane), so results do not depend on how envs are sharded over GPUs.
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define PF_RNG_HD __host__ __device__ __forceinline__
#else
#define PF_RNG_HD static inline
#endif

enum PfRngPurpose : uint32_t {
    PF_RNG_OBS_NOISE = 0,   // PF:332, 336
    PF_RNG_RESET_DOF = 1,   // PF:557
    PF_RNG_RESET_ROOT = 2,  // PF:577 (lanes 0..1), PF:583 (lanes 2..7)
    PF_RNG_CURRICULUM = 3,  // PF:632
    PF_RNG_PUSH = 4,        // PF:598
    PF_RNG_RESAMPLE = 5,    // LR:337-351 from the callback (LR:326): lanes 0, 1, 2 (third draw: heading OR yaw)
    PF_RNG_POLICY = 6,      // rsl_rl ActorCritic.act (Normal.sample)
    PF_RNG_TERRAIN = 9,     // terrain generation: 'env' = sub-terrain index, lanes as in csrc/pf_terrain.cu; step 0
    PF_RNG_RESAMPLE_RESET = 8   // LR:337-351 from reset_idx (LR:166): same lanes, own purpose (7 is the minibatch shuffle)
};

PF_RNG_HD uint32_t pf_fmix32(uint32_t h) {
    h ^= h >> 16; h *= 0x85EBCA6Bu; h ^= h >> 13; h *= 0xC2B2AE35u; h ^= h >> 16;
    return h;
}

// key for (seed, step, purpose): warp-uniform, computed once per kernel
PF_RNG_HD uint32_t pf_rng_key(uint32_t seed, uint32_t step, uint32_t purpose) {
    uint32_t k = pf_fmix32(seed * 0x9E3779B1u + step);
    return pf_fmix32(k ^ (purpose * 0x85EBCA77u));
}
PF_RNG_HD uint32_t pf_rng_env_key(uint32_t key, uint32_t env) {
    return pf_fmix32(key + env * 0xC2B2AE3Du);
}
PF_RNG_HD uint32_t pf_rng_u32(uint32_t env_key, uint32_t lane) {
    return pf_fmix32(env_key ^ (lane * 0x27D4EB2Fu));
}
// fp32 in [0,1): top 24 bits * 2^-24
PF_RNG_HD float pf_rng_uniform(uint32_t env_key, uint32_t lane) {
    return (float)(pf_rng_u32(env_key, lane) >> 8) * (1.0f / 16777216.0f);
}
#ifdef __CUDACC__
// Box-Muller on lanes (2*lane, 2*lane+1)
__device__ __forceinline__ float pf_rng_normal(uint32_t env_key, uint32_t lane) {
    float u1 = ((float)(pf_rng_u32(env_key, 2u * lane) >> 8) + 1.0f) * (1.0f / 16777216.0f);
    float u2 = (float)(pf_rng_u32(env_key, 2u * lane + 1u) >> 8) * (1.0f / 16777216.0f);
    return sqrtf(-2.0f * logf(u1)) * cosf(6.283185307179586f * u2);
}
#endif
