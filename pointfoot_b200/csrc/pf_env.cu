// Env-side kernels of the hot path (sm_100a): PD torques, the fused post_physics_step, reset compaction and
// episode statistics, push forces, reset-all.  The per-env arithmetic lives in pf_env_core.cuh (shared with
// the host emulation in tests/); this file is the staging, the thread mapping and the block reductions.
//
// Mapping: one WARP owns envs_per_warp (8/16/32) consecutive envs and never synchronises
// with another warp; phases are described at pf_post_physics_kernel.  A second, tiny kernel
// (pf_finalize_kernel) turns the per-warp partials into the ascending reset id list (PF:221 nonzero()), the
// device count and the episode statistics, deterministically (no float atomics).
#include <cuda_runtime.h>
#include <stdlib.h>
#include <stdio.h>
#include <string.h>
#include "pf_env_core.cuh"
#include "pf_ctx.h"
#include "pf_tc_ptx.cuh"      // mbarrier + bulk-copy wrappers
#include "pf_launch.h"

#define PF_THREADS 256

// ------------------------------------------------------------------------------------------------ K1
// _compute_torques, PF:523-547 (+ the action clip of PF:176-177 when actions_in != NULL).
// One thread per (env, dof): every stream is contiguous -> fully coalesced, dof_state read as float2.
__global__ void __launch_bounds__(256) pf_pd_torques_kernel(const pf_params P, const pf_env_buffers B,
                                                            const float* __restrict__ actions_in) {
    const int n = P.num_envs * P.num_dof;
    pf_grid_dep_launch();
    pf_grid_dep_wait();
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int d = i % P.num_dof;
        float a;
        if (actions_in) {
            a = pf_clipf(actions_in[i], -P.clip_actions, P.clip_actions);
            B.actions[i] = a;
        } else {
            a = B.actions[i];
        }
        const float2 s = reinterpret_cast<const float2*>(B.dof_state)[i];
        const float last_qd = P.control_type == PF_CONTROL_V ? B.last_dof_vel[i] : 0.0f;
        B.torques[i] = pf_pd_torque(P, d, a, s.x, s.y, last_qd);
    }
}

// The same law, one thread per PAIR of DOFs (even DOF counts, 16-byte aligned dof_state / 8-byte aligned [N,D] tensors):
// one 128-bit load of (q, qd, q', qd'), one 64-bit load of the two actions, one 64-bit store of the two torques --
// a quarter of the memory instructions of the scalar kernel, and no grid-stride loop (the grid covers every pair).
__global__ void __launch_bounds__(256) pf_pd_torques2_kernel(const pf_params P, const pf_env_buffers B,
                                                             const float* __restrict__ actions_in) {
    const int half = P.num_dof >> 1;
    const int n = P.num_envs * half;
    pf_grid_dep_launch();
    pf_timeline(0x10);
    pf_grid_dep_wait();
    pf_timeline(0x11);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int d = 2 * (i % half);
    float2 a;
    if (actions_in) {
        a = reinterpret_cast<const float2*>(actions_in)[i];
        a.x = pf_clipf(a.x, -P.clip_actions, P.clip_actions);
        a.y = pf_clipf(a.y, -P.clip_actions, P.clip_actions);
        reinterpret_cast<float2*>(B.actions)[i] = a;
    } else {
        a = reinterpret_cast<const float2*>(B.actions)[i];
    }
    const float4 s = reinterpret_cast<const float4*>(B.dof_state)[i];
    float2 lq = make_float2(0.0f, 0.0f);
    if (P.control_type == PF_CONTROL_V) lq = reinterpret_cast<const float2*>(B.last_dof_vel)[i];
    float2 t;
    t.x = pf_pd_torque(P, d, a.x, s.x, s.y, lq.x);
    t.y = pf_pd_torque(P, d + 1, a.y, s.z, s.w, lq.y);
    reinterpret_cast<float2*>(B.torques)[i] = t;
    pf_timeline(0x12);
}

// ------------------------------------------------------------------------------------------------ helpers
// Per-WARP shared memory slab of the step kernel (epw = envs per warp: 8, 16 or 32).  Warps never talk to each
// other: there is no __syncthreads in the step kernel, only __syncwarp, so every warp runs ahead on its own
// loads.  Row strides are odd (13, 27) or small (12) so that one-thread-per-env row reads are (nearly)
// conflict free.
struct PfWarpSmem {
    float* root;      // [epw][13]   pre-reset root rows
    float* dof;       // [epw][2D]
    float* cf;        // [epw][cfw]  contact forces; reused for the observation columns before noise
    float* misc;      // [epw][4]    yaw quaternion (z, w), post-reset root z, noise key (bits)
    float* bar;       // 8-byte mbarrier of the warp's bulk copies (16 bytes reserved)
};

__host__ __device__ inline int pf_cf_width(const pf_params& P) {
    const int a = 3 * P.num_bodies, b = 6 + 3 * P.num_dof + 3;
    return (a > b ? a : b) | 1;
}
__host__ __device__ inline int pf_warp_smem_floats(const pf_params& P, int epw) {
    const int n = epw * (13 + 2 * P.num_dof + pf_cf_width(P) + 4);
    return ((n + 3) & ~3) + 4;                            // + the mbarrier; every warp slab stays 16-byte aligned
}

__device__ __forceinline__ PfWarpSmem pf_carve_warp(const pf_params& P, int epw, float* f) {
    PfWarpSmem s;
    s.root = f; f += epw * 13;
    s.dof = f; f += epw * 2 * P.num_dof;
    s.cf = f; f += epw * pf_cf_width(P);
    s.misc = f; f += epw * 4;
    s.bar = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(f) + 15) & ~(uintptr_t)15);
    return s;
}

// coalesced copy of `n` floats global -> shared by one warp; 128-bit when both sides allow it
__device__ __forceinline__ void pf_warp_copy(float* __restrict__ dst, const float* __restrict__ src, int n, int lane) {
    if (((reinterpret_cast<uintptr_t>(src) | (uintptr_t)__cvta_generic_to_shared(dst)) & 15) == 0 && (n & 3) == 0) {
        const float4* s4 = reinterpret_cast<const float4*>(src);
        float4* d4 = reinterpret_cast<float4*>(dst);
        for (int i = lane; i < (n >> 2); i += 32) d4[i] = __ldg(s4 + i);
    } else {
        for (int i = lane; i < n; i += 32) dst[i] = __ldg(src + i);
    }
}

__device__ __forceinline__ double pf_warp_sum(double v) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);      // fixed tree: deterministic
    return v;
}

// Height scan of ONE env by the whole warp (PF:1048-1075, 1100-1112): lane l takes points l, l+32, ...
// Writes measured_heights (if materialised) and the height columns of obs / priv obs (PF:367-369, 328-338,
// 189-192) with `root_z`; returns (to every lane) sum_p (root_z_pre - h_p) in double for the base-height mean.
struct PfScanLane {               // hoisted per-lane constants of the scan loop (same for every env of the warp)
    float gx[PF_MAX_HEIGHT_POINTS / 32], gy[PF_MAX_HEIGHT_POINTS / 32], ns[PF_MAX_HEIGHT_POINTS / 32];
};

template <bool kWriteObs, bool kFast>
__device__ __forceinline__ double pf_scan_env(const pf_params& P, const pf_env_buffers& B, const PfScanLane& L,
                                              const float* __restrict__ root, float zn, float wn, float root_z_obs,
                                              uint32_t nkey, int e, int lane, bool want_sum, const PfRewTab* T) {
    const int Pn = P.num_height_points, base_w = pf_obs_base_width(P);
    const bool rough = P.terrain_kind == PF_TERRAIN_HEIGHTFIELD;
    const float bx = root[0], by = root[1], rz = root[2];
    float* mh = B.measured_heights ? B.measured_heights + (size_t)e * Pn : nullptr;
    float* oh = (kWriteObs && P.measure_heights_actor) ? B.obs_buf + (size_t)e * P.num_obs + base_w : nullptr;
    float* ph = (kWriteObs && P.measure_heights_critic && B.priv_obs_buf)
                    ? B.priv_obs_buf + (size_t)e * P.num_priv_obs + base_w : nullptr;
    double acc = 0.0;
#pragma unroll
    for (int k = 0; k < PF_MAX_HEIGHT_POINTS / 32; ++k) {
        const int p = lane + 32 * k;
        if (p < Pn) {
            int16_t r = 0;                                       // plane: zeros, PF:1061-1062
            if (rough) r = pf_scan_point(P, B.height_samples, zn, wn, L.gx[k], L.gy[k], bx, by, kFast ? T : nullptr);
            const float h = (float)r * P.vertical_scale;
            if (mh) mh[p] = h;
            if (want_sum) acc += (double)(rz - h);
            if (kWriteObs && (oh || ph)) {
                float v = pf_clipf((root_z_obs - 0.5f) - h, -1.0f, 1.0f) * P.obs_scale_height;
                if (P.add_noise && L.ns[k] != 0.0f)
                    v = v + (2.0f * pf_rng_uniform(nkey, (uint32_t)(base_w + p)) - 1.0f) * L.ns[k];
                v = pf_clipf(v, -P.clip_obs, P.clip_obs);
                if (oh) oh[p] = v;
                if (ph) ph[p] = v;
            }
        }
    }
    return want_sum ? pf_warp_sum(acc) : 0.0;
}

// The same scan for the rare caller (phase 4 of the step kernel: envs that reset), ROLLED: one copy of the point's code
// instead of eight, the grid point and noise scale read inside the loop.  The few warps that take this path are the
// last of the launch to retire and execute it exactly once, so what counts is how many cold instruction lines they
// have to fetch, not instructions per point (same arithmetic, same results).
template <bool kFast>
__device__ __forceinline__ void pf_scan_env_rolled(const pf_params& P, const pf_env_buffers& B, const PfCtxScratch& S,
                                                const float* __restrict__ root, float zn, float wn, float root_z_obs,
                                                uint32_t nkey, int e, int lane) {
    const int Pn = P.num_height_points, base_w = pf_obs_base_width(P);
    const bool rough = P.terrain_kind == PF_TERRAIN_HEIGHTFIELD;
    const float bx = root[0], by = root[1];
    float* mh = B.measured_heights ? B.measured_heights + (size_t)e * Pn : nullptr;
    float* oh = P.measure_heights_actor ? B.obs_buf + (size_t)e * P.num_obs + base_w : nullptr;
    float* ph = (P.measure_heights_critic && B.priv_obs_buf) ? B.priv_obs_buf + (size_t)e * P.num_priv_obs + base_w : nullptr;
#pragma unroll 1
    for (int p = lane; p < Pn; p += 32) {
        const float2 g = reinterpret_cast<const float2*>(S.height_points)[p];
        const float ns = S.noise_vec[base_w + p];
        int16_t r = 0;
        if (rough) r = pf_scan_point(P, B.height_samples, zn, wn, g.x, g.y, bx, by, kFast ? &S.rew : nullptr);
        const float h = (float)r * P.vertical_scale;
        if (mh) mh[p] = h;
        if (oh || ph) {
            float v = pf_clipf((root_z_obs - 0.5f) - h, -1.0f, 1.0f) * P.obs_scale_height;
            if (P.add_noise && ns != 0.0f) v = v + (2.0f * pf_rng_uniform(nkey, (uint32_t)(base_w + p)) - 1.0f) * ns;
            v = pf_clipf(v, -P.clip_obs, P.clip_obs);
            if (oh) oh[p] = v;
            if (ph) ph[p] = v;
        }
    }
}

__device__ __forceinline__ void pf_scan_lane_load(const pf_params& P, const PfCtxScratch& S, int lane, PfScanLane& L) {
    const int Pn = P.num_height_points, base_w = pf_obs_base_width(P);
#pragma unroll
    for (int k = 0; k < PF_MAX_HEIGHT_POINTS / 32; ++k) {
        const int p = lane + 32 * k;
        const bool ok = p < Pn;
        const float2 g = ok ? reinterpret_cast<const float2*>(S.height_points)[p] : make_float2(0.f, 0.f);
        L.gx[k] = g.x; L.gy[k] = g.y;
        L.ns[k] = ok ? S.noise_vec[base_w + p] : 0.0f;
    }
}

// ------------------------------------------------------------------------------------------------ K2a
// _get_heights (PF:1048-1075, 1100-1112) + the height columns of both observation buffers (PF:367-369, 328-338,
// 189-192) + the base-height mean of _reward_base_height (PF:1144-1147), for every env: one WARP per env, lanes =
// scan points.  A kernel of its own because it wants the opposite of the step kernel: few registers, every warp
// slot of the SM filled, tens of thousands of independent gathers in flight against the L2-resident heightfield.
// The step kernel (launched behind it) only reads height_mean[e].
template <bool kFast>
__global__ void __launch_bounds__(256, 4) pf_height_scan_kernel(const pf_params P, const pf_env_buffers B,
                                                                const PfCtxScratch S, const int epw,
                                                                const uint32_t step_arg) {
    const int lane = threadIdx.x & 31;
    const int gw = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5), nw = (int)((gridDim.x * blockDim.x) >> 5);
    pf_grid_dep_launch();
    PfScanLane L;
    pf_scan_lane_load(P, S, lane, L);                   // constants of the context: not written by any kernel
    pf_timeline(0x40);
    pf_grid_dep_wait();
    pf_timeline(0x41);
    const uint32_t step = B.step_counter ? (uint32_t)(*B.step_counter + 1) : step_arg;
    const uint32_t key0 = pf_rng_key(P.seed, step, PF_RNG_OBS_NOISE);
    const double pn = (double)P.num_height_points;
    // A warp takes `epw` (1..32) consecutive envs at a time: what is per env rather than per point -- the root row,
    // the yaw quaternion (a square root and two divisions), the noise key, the final division of the mean -- is done
    // once per batch with lane = env and handed to the scan by shuffles.  epw = 1 for small batches (one env per
    // warp: maximum parallelism), 32 for large ones (minimum instructions).
    for (int base = gw * epw; base < P.num_envs; base += nw * epw) {
        const int nb = min(epw, P.num_envs - base);
        float x = 0.0f, y = 0.0f, z = 0.0f, zn = 0.0f, wn = 1.0f;
        uint32_t nkey = 0;
        if (lane < nb) {
            const float* g = B.root_states + (size_t)(base + lane) * 13;
            x = g[0]; y = g[1]; z = g[2];
            pf_yaw_quat(g + 3, zn, wn);
            nkey = pf_rng_env_key(key0, (uint32_t)(P.env_id_offset + base + lane));
        }
        double hs = 0.0;
        for (int j = 0; j < nb; ++j) {
            float r3[3];
            r3[0] = __shfl_sync(0xffffffffu, x, j); r3[1] = __shfl_sync(0xffffffffu, y, j); r3[2] = __shfl_sync(0xffffffffu, z, j);
            const float zj = __shfl_sync(0xffffffffu, zn, j), wj = __shfl_sync(0xffffffffu, wn, j);
            const uint32_t kj = __shfl_sync(0xffffffffu, nkey, j);
            const double sum = pf_scan_env<true, kFast>(P, B, L, r3, zj, wj, r3[2], kj, base + j, lane, true, &S.rew);
            if (lane == j) hs = sum;
        }
        if (lane < nb) S.height_mean[base + lane] = (float)(hs / pn);
    }
    pf_timeline(0x42);
}

// The same kernel for THE rough configuration of both pointfoot cfgs, everything warp-uniform decided at compile time or on
// the host (pf_post_physics_impl): heightfield terrain, the critic's observations carry the height columns (both pointfoot cfgs:
// 27 + 121 privileged columns) and with kActor the actor's too, noise on,
// measured_heights not materialised, the index division of kind kDiv (1: x * RN(1/scale), 2: the checked Markstein step),
// 32 * (kPasses - 1) < num_height_points <= 32 * kPasses so that only the last pass tests its lane.  Same arithmetic in
// the same order as pf_scan_env (bit-identical; tests/test_env_gpu.py runs both): what goes is ~30 of ~100 issued
// instructions per point -- pointer null tests, the division-kind dispatch, the lane's share of the noise hash.
template <int kDiv>
__device__ __forceinline__ int16_t pf_terrain_min3_k(const pf_params& P, const PfRewTab& T, const int16_t* __restrict__ hs, float x, float y) {
    x = x + P.border_size;
    y = y + P.border_size;
    if (kDiv == 2) { x = pf_div_by(x, P.horizontal_scale, T.r_hscale); y = pf_div_by(y, P.horizontal_scale, T.r_hscale); }
    else { x = x * T.r_hscale; y = y * T.r_hscale; }
    const int px = (int)pf_minf(pf_maxf(x, 0.0f), T.rows_m2);
    const int py = (int)pf_minf(pf_maxf(y, 0.0f), T.cols_m2);
    const int16_t* r = hs + (px * P.hf_cols + py);
    const int16_t h1 = __ldg(r), h2 = __ldg(r + P.hf_cols), h3 = __ldg(r + 1);
    const int16_t m = h1 < h2 ? h1 : h2;
    return m < h3 ? m : h3;
}

template <int kDiv, int kPasses, bool kActor>
__global__ void __launch_bounds__(256, 4) pf_height_scan_std_kernel(const pf_params P, const pf_env_buffers B,
                                                                    const PfCtxScratch S, const int epw,
                                                                    const uint32_t step_arg) {
    const int lane = threadIdx.x & 31;
    const int gw = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5), nw = (int)((gridDim.x * blockDim.x) >> 5);
    pf_grid_dep_launch();
    const int Pn = P.num_height_points, base_w = pf_obs_base_width(P);
    float gx[kPasses], gy[kPasses], ns[kPasses];
    uint32_t hk[kPasses];                               // the point's share of pf_rng_u32: (column) * 0x27D4EB2F
#pragma unroll
    for (int k = 0; k < kPasses; ++k) {
        const int p = lane + 32 * k;
        const bool ok = p < Pn;
        const float2 g = ok ? reinterpret_cast<const float2*>(S.height_points)[p] : make_float2(0.f, 0.f);
        gx[k] = g.x; gy[k] = g.y;
        ns[k] = ok ? S.noise_vec[base_w + p] : 0.0f;
        hk[k] = (uint32_t)(base_w + p) * 0x27D4EB2Fu;
    }
    const bool last_ok = lane + 32 * (kPasses - 1) < Pn;
    pf_timeline(0x40);
    pf_grid_dep_wait();
    pf_timeline(0x41);
    const uint32_t step = B.step_counter ? (uint32_t)(*B.step_counter + 1) : step_arg;
    const uint32_t key0 = pf_rng_key(P.seed, step, PF_RNG_OBS_NOISE);
    const double pn = (double)Pn;
    for (int base = gw * epw; base < P.num_envs; base += nw * epw) {
        const int nb = min(epw, P.num_envs - base);
        float x = 0.0f, y = 0.0f, z = 0.0f, zn = 0.0f, wn = 1.0f;
        uint32_t nkey = 0;
        if (lane < nb) {
            const float* g = B.root_states + (size_t)(base + lane) * 13;
            x = g[0]; y = g[1]; z = g[2];
            pf_yaw_quat(g + 3, zn, wn);
            nkey = pf_rng_env_key(key0, (uint32_t)(P.env_id_offset + base + lane));
        }
        double hs = 0.0;
        for (int j = 0; j < nb; ++j) {
            const float bx = __shfl_sync(0xffffffffu, x, j), by = __shfl_sync(0xffffffffu, y, j), rz = __shfl_sync(0xffffffffu, z, j);
            const float zj = __shfl_sync(0xffffffffu, zn, j), wj = __shfl_sync(0xffffffffu, wn, j);
            const uint32_t kj = __shfl_sync(0xffffffffu, nkey, j);
            float* oh = kActor ? B.obs_buf + (size_t)(base + j) * P.num_obs + base_w + lane : nullptr;
            float* ph = B.priv_obs_buf + (size_t)(base + j) * P.num_priv_obs + base_w + lane;
            const float zo = rz - 0.5f;
            double acc = 0.0;
#pragma unroll
            for (int k = 0; k < kPasses; ++k) {
                if (k < kPasses - 1 || last_ok) {
                    const float t0 = -(zj * gy[k]) * 2.0f;                       // pf_scan_point
                    const float t1 = (zj * gx[k]) * 2.0f;
                    const float ox = (gx[k] + wj * t0) + (-(zj * t1));
                    const float oy = (gy[k] + wj * t1) + (zj * t0);
                    const int16_t r = pf_terrain_min3_k<kDiv>(P, S.rew, B.height_samples, ox + bx, oy + by);
                    const float h = (float)r * P.vertical_scale;
                    acc += (double)(rz - h);
                    float v = pf_clipf(zo - h, -1.0f, 1.0f) * P.obs_scale_height;
                    if (ns[k] != 0.0f)
                        v = v + (2.0f * ((float)(pf_fmix32(kj ^ hk[k]) >> 8) * (1.0f / 16777216.0f)) - 1.0f) * ns[k];
                    v = pf_clipf(v, -P.clip_obs, P.clip_obs);
                    if (kActor) oh[32 * k] = v;
                    ph[32 * k] = v;
                }
            }
            const double sum = pf_warp_sum(acc);
            if (lane == j) hs = sum;
        }
        if (lane < nb) S.height_mean[base + lane] = (float)(hs / pn);
    }
    pf_timeline(0x42);
}

// Warp partials for extras["episode"] (PF:276-288): count, terrain-level sum, per-term episode sums of the
// warp's reset envs (read back BEFORE pf_env_reset zeroes them).  Entry index = global warp-slab index.
__device__ __forceinline__ void pf_warp_partials(const pf_params& P, const pf_env_buffers& B, const PfCtxScratch& S,
                                                 int slab, int e, bool active, bool is_reset, int R, int lane) {
    const unsigned m = __ballot_sync(0xffffffffu, is_reset);
    if (m) {
        const int N = P.num_envs;
        if ((m & (m - 1)) == 0) {
            // one reset env in the slab (the usual case): its lane writes the sums itself -- the same values the
            // tree below would produce (x + 0 + ... + 0), without 19 double-precision warp reductions on the
            // critical path of the slowest warps of the launch
            if (is_reset)
                for (int k = 0; k < R; ++k)
                    S.blk_sums[(size_t)slab * PF_MAX_REWARD_TERMS + k] = (double)B.episode_sums[(size_t)k * N + e];
        } else {
            for (int k = 0; k < R; ++k) {
                const double v = is_reset ? (double)B.episode_sums[(size_t)k * N + e] : 0.0;
                const double t = pf_warp_sum(v);
                if (lane == 0) S.blk_sums[(size_t)slab * PF_MAX_REWARD_TERMS + k] = t;
            }
        }
    }
    if (lane == 0) S.blk_count[slab] = __popc(m);
    (void)active;
}

// ------------------------------------------------------------------------------------------------ K2
// Fused post_physics_step (PF:195-227) + obs clip (PF:189-192).  One warp owns `epw` consecutive envs.
//   phase 0  coalesced 128-bit copies of the warp's PhysX rows (root, dof, contact) into its smem slab
//   phase 1  height scan, warp-cooperative per env (lanes = scan points): int16 min-of-3 gathers from the L2
//            resident field, height observation columns written straight from registers, base-height sum
//            by warp shuffle
//   phase 2  one lane per env: pf_env_step (+ pf_env_reset for the few that reset)
//   phase 3  the 27 leading observation columns, flattened over the warp's envs -> coalesced row writes
//   phase 4  (rare) envs that reset: height columns rewritten with the post-reset root z (PF:223 runs
//            after reset_idx) -- the scan itself stays on the pre-reset pose (PF:212-213)
// kMinB: the register allocation aims at 2 * kMinB resident CTAs of 128 threads per SM.  2 = 128 registers, nothing
// spilled; 80- and 64-register builds were measured slower at every batch size, a 96-register build (5 CTAs, 20 bytes
// spilled) no faster (profiles/r1h_step_kernel_sweeps.txt): ncu shows no unit saturated (DRAM 48 %, L1 39 %, LSU 21 %,
// issue 37 % at 262,144 envs) -- the kernel is bound by how well 16 warps per SM overlap their memory round trips.  A larger shared-memory carve-out is slower as well: the
// per-env loads of the [N,6] / [N,2] state arrays (24- and 8-byte strides per lane) live off L1 sector reuse.
// kScan: the task measures terrain heights (num_height_points > 0): pf_height_scan_kernel ran in front of this
// kernel and left height_mean[e]; only the envs that reset redo their scan here (phase 4).
template <uint32_t kMask, int kMinB, bool kScan>
__global__ void __launch_bounds__(128, 2 * kMinB) pf_post_physics_kernel(const pf_params P, const pf_env_buffers B,
                                                                        const PfCtxScratch S, const int epw,
                                                                        const uint32_t step_arg) {
    extern __shared__ __align__(16) float smem_f[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int slab = blockIdx.x * (blockDim.x >> 5) + warp;
    const int e0 = slab * epw;
    pf_grid_dep_launch();
    if (e0 >= P.num_envs) return;
    const PfWarpSmem s = pf_carve_warp(P, epw, smem_f + (size_t)warp * pf_warp_smem_floats(P, epw));
    const int nenv = min(epw, P.num_envs - e0);
    pf_timeline(0x20);
    pf_grid_dep_wait();
    pf_timeline(0x21);
    const uint32_t step = B.step_counter ? (uint32_t)(*B.step_counter + 1) : step_arg;
    constexpr bool kFast = kMask != PF_MASK_GENERIC;
    constexpr int kD = kFast ? PF_MAX_DOF : 0;
    const int D = kFast ? PF_MAX_DOF : P.num_dof, NB = P.num_bodies, Pn = kScan ? P.num_height_points : 0;
    const int R = P.num_reward_terms + (P.has_termination_term ? 1 : 0);
    const int base_w = 6 + 3 * D + 3, cfw = pf_cf_width(P);
    const bool active = lane < nenv;
    const int e = e0 + lane;

    // ---- phase 0: AoS rows -> shared memory ----
    // Full, 16-byte aligned slabs: one lane hands the three row blocks to the TMA engine (cp.async.bulk completing
    // on the warp's own mbarrier), so the staging costs a dozen instructions instead of three copy loops.
    const float* g_root = B.root_states + (size_t)e0 * 13;
    const float* g_dof = B.dof_state + (size_t)e0 * 2 * D;
    const float* g_cf = B.contact_forces + (size_t)e0 * 3 * NB;
    const bool bulk = kFast && nenv == epw && (epw & 3) == 0 && cfw == 3 * NB &&
                      ((reinterpret_cast<uintptr_t>(g_root) | reinterpret_cast<uintptr_t>(g_dof) |
                        reinterpret_cast<uintptr_t>(g_cf)) & 15) == 0;
    if (bulk) {
        const uint32_t bar = smem_u32(s.bar);
        if (lane == 0) {
            mbar_init(bar, 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            fence_async_smem();
            const uint32_t b_root = epw * 13 * 4, b_dof = epw * 2 * D * 4, b_cf = epw * 3 * NB * 4;
            mbar_expect_tx(bar, b_root + b_dof + b_cf);
            bulk_g2s(smem_u32(s.root), g_root, b_root, bar);
            bulk_g2s(smem_u32(s.dof), g_dof, b_dof, bar);
            bulk_g2s(smem_u32(s.cf), g_cf, b_cf, bar);
        }
    } else {
        pf_warp_copy(s.root, g_root, nenv * 13, lane);
        pf_warp_copy(s.dof, g_dof, nenv * 2 * D, lane);
        if (cfw == 3 * NB) {
            pf_warp_copy(s.cf, g_cf, nenv * 3 * NB, lane);
        } else {
            for (int i = lane; i < nenv * 3 * NB; i += 32) s.cf[(i / (3 * NB)) * cfw + (i % (3 * NB))] = __ldg(g_cf + i);
        }
    }
    float feet[6];
    PfEnvPre pre;
    if (active) {
        pf_env_preload<kMask>(P, B, S.rew, e, pre);       // in flight together with the staging above
        for (int f = 0; f < PF_NUM_FEET; ++f) {
            const float* g = B.rigid_body_states + ((size_t)e * NB + P.feet_indices[f]) * 13;
            feet[3 * f] = __ldg(g); feet[3 * f + 1] = __ldg(g + 1); feet[3 * f + 2] = __ldg(g + 2);
        }
    }
    const uint32_t nkey = pf_rng_env_key(pf_rng_key(P.seed, step, PF_RNG_OBS_NOISE), (uint32_t)(P.env_id_offset + e));
    __syncwarp();
    if (bulk) mbar_wait(smem_u32(s.bar), 0);

    // ---- phase 1 (height scan) is pf_height_scan_kernel, launched in front of this kernel ----
    const float height_mean = (kScan && active) ? S.height_mean[e] : 0.0f;

    // ---- phase 2: one lane per env ----
    bool is_reset = false;
    PfEnvOut out;
    if (active) {
        PfEnvIn in;
        in.root = s.root + lane * 13;
        in.dof = s.dof + lane * 2 * D;
        in.cf = s.cf + lane * cfw;
        in.feet_pos[0] = feet;
        in.feet_pos[1] = feet + 3;
        in.have_heights = kScan;
        in.height_mean = height_mean;
        pf_env_step_t<kMask>(P, B, S.rew, e, in, pre, step, nullptr, out);
        is_reset = out.reset;
    }
    pf_warp_partials(P, B, S, slab, e, active, is_reset, R, lane);
    long long lvl = 0;
    if (active) {
        if (is_reset) pf_env_reset_t<kD>(P, B, e, s.root + lane * 13, out.commands, step, true, nullptr, out);
        if (B.terrain_levels) lvl = B.terrain_levels[e];
        pf_obs_base_row<kD>(P, out, s.cf + lane * cfw);    // own row: the contact forces are no longer needed
        s.misc[4 * lane + 2] = out.root_z;
        s.misc[4 * lane + 3] = __uint_as_float(nkey);
        // PF:225-227
        if (kFast) {
#pragma unroll
            for (int d = 0; d < 6; d += 2) {
                pf_st2(B.last_actions + (size_t)e * 6 + d, out.actions[d], out.actions[d + 1]);
                pf_st2(B.last_dof_vel + (size_t)e * 6 + d, out.dof_vel[d], out.dof_vel[d + 1]);
                pf_st2(B.last_root_vel + (size_t)e * 6 + d, out.root_vel[d], out.root_vel[d + 1]);
            }
        } else {
#pragma unroll
            for (int d = 0; d < PF_MAX_DOF; ++d)
                if (d < D) {
                    B.last_actions[(size_t)e * D + d] = out.actions[d];
                    B.last_dof_vel[(size_t)e * D + d] = out.dof_vel[d];
                }
#pragma unroll
            for (int i = 0; i < 6; ++i) B.last_root_vel[(size_t)e * 6 + i] = out.root_vel[i];
        }
    }
    {
        const unsigned lv = __reduce_add_sync(0xffffffffu, (unsigned)lvl);   // levels are small and >= 0: exact
        if (lane == 0) S.blk_level[slab] = (long long)lv;
    }
    __syncwarp();

    // ---- phase 3: leading observation columns, coalesced (PF:376-382, 328-338, 189-192) ----
    const int no = P.num_obs, np = P.num_priv_obs;
    float* go = B.obs_buf + (size_t)e0 * no;
    float* gp = B.priv_obs_buf ? B.priv_obs_buf + (size_t)e0 * np : nullptr;
    if (kFast) {
        // 27 columns: lane = column, one env per iteration.  What depends on the column alone (noise scale, the
        // column's share of the RNG hash) is hoisted; a row write is one 108-byte coalesced store.
        constexpr int W = 6 + 3 * PF_MAX_DOF + 3;
        const bool col = lane < W;
        const float ns = (col && P.add_noise) ? S.noise_vec[lane] : 0.0f;
        const uint32_t lane_mix = (uint32_t)lane * 0x27D4EB2Fu;
        const float lim = P.clip_obs;
        const float* src = s.cf + lane;
        const float* keys = s.misc + 3;
        float* o = go + lane;
        if (col) {
            if (gp) {
                float* p = gp + lane;
#pragma unroll 4
                for (int le = 0; le < nenv; ++le) {
                    float v = src[le * cfw];
                    if (ns != 0.0f) {
                        const float u = (float)(pf_fmix32(__float_as_uint(keys[4 * le]) ^ lane_mix) >> 8) * (1.0f / 16777216.0f);
                        v = v + (2.0f * u - 1.0f) * ns;
                    }
                    v = pf_clipf(v, -lim, lim);
                    *o = v; o += no;
                    *p = v; p += np;
                }
            } else {
#pragma unroll 4
                for (int le = 0; le < nenv; ++le) {
                    float v = src[le * cfw];
                    if (ns != 0.0f) {
                        const float u = (float)(pf_fmix32(__float_as_uint(keys[4 * le]) ^ lane_mix) >> 8) * (1.0f / 16777216.0f);
                        v = v + (2.0f * u - 1.0f) * ns;
                    }
                    v = pf_clipf(v, -lim, lim);
                    *o = v; o += no;
                }
            }
        }
    } else {
        const uint32_t magic = ((1u << 20) + (uint32_t)base_w - 1u) / (uint32_t)base_w;    // i / base_w for i < 2^20 / base_w
        for (int i = lane; i < nenv * base_w; i += 32) {
            const int le = (int)(((uint32_t)i * magic) >> 20), j = i - le * base_w;
            const float v = pf_obs_finish(P, s.cf[le * cfw + j], __float_as_uint(s.misc[4 * le + 3]), j, S.noise_vec);
            go[le * no + j] = v;
            if (gp) gp[le * np + j] = v;
        }
    }

    // ---- phase 4: height columns of the envs that reset ----
    if (kScan && (P.measure_heights_actor || P.measure_heights_critic)) {
        unsigned m = __ballot_sync(0xffffffffu, is_reset);
        while (m) {
            const int i = __ffs(m) - 1;
            m &= m - 1;
            const float* r = s.root + i * 13;
            float zn, wn;
            pf_yaw_quat(r + 3, zn, wn);
            pf_scan_env_rolled<kFast>(P, B, S, r, zn, wn, s.misc[4 * i + 2], __float_as_uint(s.misc[4 * i + 3]), e0 + i, lane);
        }
    }
    pf_timeline(0x22);
#ifdef PF_TIMELINE
    {   // latest end time over all warps, kept apart for warps with / without a reset env: log[4000], log[4001]
        unsigned long long* p = g_pf_timeline;
        const bool had_reset = __any_sync(0xffffffffu, is_reset);
        if (p && lane == 0) {
            unsigned long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            atomicMax(p + 4000 + (had_reset ? 1 : 0), t);
        }
    }
#endif
}

// the instantiation of the step kernel for a context: statement of the step (pf_step_mode) x with / without height scan
static pf_step_kernel_t pf_pick_step_kernel(int mode, bool scan) {
    if (mode == 2) return scan ? pf_post_physics_kernel<PF_MASK_POINTFOOT, 2, true> : pf_post_physics_kernel<PF_MASK_POINTFOOT, 2, false>;
    if (mode == 1) return scan ? pf_post_physics_kernel<PF_MASK_TABLE, 2, true> : pf_post_physics_kernel<PF_MASK_TABLE, 2, false>;
    return scan ? pf_post_physics_kernel<PF_MASK_GENERIC, 2, true> : pf_post_physics_kernel<PF_MASK_GENERIC, 2, false>;
}

// reset_idx over ALL envs (PF:137-139 reset(), before the zero-action step).
__global__ void __launch_bounds__(PF_THREADS) pf_reset_all_kernel(const pf_params P, const pf_env_buffers B,
                                                                  const PfCtxScratch S, const int epw,
                                                                  const uint32_t step_arg) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int slab = blockIdx.x * (blockDim.x >> 5) + warp;
    const int e0 = slab * epw;
    if (e0 >= P.num_envs) return;
    const int nenv = min(epw, P.num_envs - e0);
    const uint32_t step = B.step_counter ? (uint32_t)(*B.step_counter) : step_arg;   // reset() does not count a step
    const int R = P.num_reward_terms + (P.has_termination_term ? 1 : 0);
    const bool active = lane < nenv;
    const int e = e0 + lane;
    pf_warp_partials(P, B, S, slab, e, active, active, R, lane);
    long long lvl = 0;
    if (active) {
        PfEnvOut out;
        float root_pre[13], cmd[2];
        for (int i = 0; i < 13; ++i) root_pre[i] = B.root_states[(size_t)e * 13 + i];
        cmd[0] = B.commands[(size_t)e * P.num_commands];
        cmd[1] = B.commands[(size_t)e * P.num_commands + 1];
        pf_env_reset(P, B, e, root_pre, cmd, step, true, nullptr, out);
        if (B.terrain_levels) lvl = B.terrain_levels[e];
    }
    const unsigned lv = __reduce_add_sync(0xffffffffu, (unsigned)lvl);
    if (lane == 0) S.blk_level[slab] = (long long)lv;
}

// ------------------------------------------------------------------------------------------------ K3
// Partials -> ascending reset ids (PF:221), device count, extras["episode"] statistics (PF:276-288).
// Each CTA of 1024 threads compacts 1024 envs; its base offset is the sum of the step kernel's warp counts
// before it.  CTA 0 additionally reduces the per-term partials (one warp per statistic, fixed order).
__global__ void __launch_bounds__(1024) pf_finalize_kernel(const pf_params P, const pf_env_buffers B,
                                                           const PfCtxScratch S, const int epw,
                                                           const int num_slabs, const int count_step) {
    __shared__ int s_before[32], s_total[32], s_flags[32];
    __shared__ double s_stat[32][PF_MAX_REWARD_TERMS];
    __shared__ long long s_lvl[32];
    __shared__ int s_last;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int R = P.num_reward_terms + (P.has_termination_term ? 1 : 0);
    pf_grid_dep_launch();
    pf_timeline(0x30);
    pf_grid_dep_wait();
    pf_timeline(0x31);

    // One pass over the step kernel's per-slab counts (128-bit loads: `before` is a multiple of 32) gives both the
    // number of resets in front of this CTA's envs and the total; one barrier; every warp then reduces the 32
    // per-warp partials itself (no serial loops).
    const int spc = 1024 / epw;                             // slabs per CTA
    const int before = blockIdx.x * spc;
    int part_b = 0, part_t = 0;
    const int n4 = num_slabs >> 2;
    const int4* c4 = reinterpret_cast<const int4*>(S.blk_count);
#pragma unroll 2
    for (int i = tid; i < n4; i += 1024) {
        const int4 c = c4[i];
        const int sum = (c.x + c.y) + (c.z + c.w);
        part_t += sum;
        part_b += 4 * i < before ? sum : 0;
    }
    for (int i = 4 * n4 + tid; i < num_slabs; i += 1024) {
        const int c = S.blk_count[i];
        part_t += c;
        part_b += i < before ? c : 0;
    }
    const int e = blockIdx.x * 1024 + tid;
    const bool flag = e < P.num_envs && B.reset_buf[e] != 0;
    const bool tmo = e < P.num_envs && B.time_out_buf[e] != 0;
    // this CTA's own slabs, for the episode statistics: thread t < spc owns slab before + t
    // (solo: few enough slabs for CTA 0 to own them all -- small batches skip the ticket round trip)
    const bool solo = num_slabs <= 1024;
    const int my_slab = solo ? tid : before + tid;
    const bool own = solo ? (blockIdx.x == 0 && tid < num_slabs) : (tid < spc && my_slab < num_slabs);
    const int my_count = own ? S.blk_count[my_slab] : 0;
    long long lvl = own ? S.blk_level[my_slab] : 0;
    const unsigned m = __ballot_sync(0xffffffffu, flag);
    part_b = __reduce_add_sync(0xffffffffu, part_b);
    part_t = __reduce_add_sync(0xffffffffu, part_t);
    if (lane == 0) { s_before[warp] = part_b; s_total[warp] = part_t; s_flags[warp] = __popc(m); }
    __syncthreads();
    const int total = __reduce_add_sync(0xffffffffu, s_total[lane]);
    const int off = __reduce_add_sync(0xffffffffu, s_before[lane] + (lane < warp ? s_flags[lane] : 0));
    if (B.extras_time_outs && total > 0 && e < P.num_envs) B.extras_time_outs[e] = tmo;
    if (flag) B.reset_ids[off + __popc(m & ((1u << lane) - 1u))] = e;
    pf_timeline(0x32);
    if (blockIdx.x == 0 && tid == 0) {
        *B.reset_count = total;
        if (B.step_counter && count_step) *B.step_counter += 1;
    }
    if (total == 0) return;                       // PF:259: extras rebuilt only when >= 1 env resets (uniform over the grid)
    if (solo && blockIdx.x != 0) return;

    // extras["episode"] (PF:276-288): every CTA reduces the partials of its own slabs in a fixed order and publishes
    // them; the CTA that draws the last ticket adds the CTA partials up in CTA order.  Deterministic for a geometry,
    // no float atomics, no CTA that walks every slab.
    const int nw = solo ? 32 : (spc + 31) >> 5;   // warps that own slabs
    if (warp < nw) {
        const bool any = __any_sync(0xffffffffu, my_count > 0);
        const double* row = S.blk_sums + (size_t)my_slab * PF_MAX_REWARD_TERMS;
        if (any) {
            for (int k0 = 0; k0 < R; k0 += 8) {   // 8 independent loads / reduction chains at a time (ILP without a register blow-up)
                double v[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] = (my_count > 0 && k0 + j < R) ? row[k0 + j] : 0.0;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) v[j] += __shfl_xor_sync(0xffffffffu, v[j], o);
                }
                if (lane == 0) {
#pragma unroll
                    for (int j = 0; j < 8; ++j) if (k0 + j < PF_MAX_REWARD_TERMS) s_stat[warp][k0 + j] = v[j];
                }
            }
        } else if (lane < R) {
            s_stat[warp][lane] = 0.0;             // no reset in this warp's slabs: nothing to load or reduce
        }
        for (int o = 16; o > 0; o >>= 1) lvl += __shfl_xor_sync(0xffffffffu, lvl, o);
        if (lane == 0) s_lvl[warp] = lvl;
    }
    __syncthreads();
    if (warp < R) {
        const double t = pf_warp_sum(lane < nw ? s_stat[lane][warp] : 0.0);
        if (lane == 0) {
            if (solo) B.episode_stats[warp] = (float)(t / (double)total) / P.max_episode_length_s;
            else S.cta_sums[(size_t)blockIdx.x * PF_MAX_REWARD_TERMS + warp] = t;
        }
    } else if (warp == R) {
        long long t = lane < nw ? s_lvl[lane] : 0;
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        if (lane == 0) {
            if (solo) {
                B.episode_stats[R] = (float)((double)t / (double)P.num_envs);
                B.episode_stats[R + 1] = (float)total;
            } else {
                S.cta_lvl[blockIdx.x] = t;
            }
        }
    }
    if (solo) return;
    __threadfence();
    __syncthreads();
    if (tid == 0) s_last = atomicAdd(S.ticket, 1) == (int)gridDim.x - 1;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    const int nc = (int)gridDim.x;
    if (warp < R) {
        double acc = 0.0;
#pragma unroll 8
        for (int c = lane; c < nc; c += 32) acc += __ldcg(S.cta_sums + (size_t)c * PF_MAX_REWARD_TERMS + warp);
        acc = pf_warp_sum(acc);
        if (lane == 0) B.episode_stats[warp] = (float)(acc / (double)total) / P.max_episode_length_s;
    } else if (warp == R) {
        long long acc = 0;
        for (int c = lane; c < nc; c += 32) acc += __ldcg(S.cta_lvl + c);
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (lane == 0) {
            B.episode_stats[R] = (float)((double)acc / (double)P.num_envs);
            B.episode_stats[R + 1] = (float)total;
            *S.ticket = 0;
        }
    }
}

// ------------------------------------------------------------------------------------------------ push
// _push_robots, PF:590-611: forces[N,B,3], body 0 row = rotated random force, everything else zero.
__global__ void __launch_bounds__(256) pf_push_kernel(const pf_params P, const pf_env_buffers B, float* forces,
                                                      const float max_push_force, const uint32_t step) {
    const int row = 3 * P.num_bodies;
    const int n = P.num_envs * row;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int e = i / row, c = i - e * row;
        float v = 0.0f;
        if (c < 3) {
            float q[4], f[3];
            for (int k = 0; k < 4; ++k) q[k] = B.root_states[(size_t)e * 13 + 3 + k];
            pf_push_force(P, e, q, max_push_force, step, f);
            v = f[c];
        }
        forces[i] = v;
    }
}

// ================================================================================================ C ABI
static thread_local char g_err[512] = "";

static int pf_fail(int code, const char* fmt, const char* detail) {
    snprintf(g_err, sizeof(g_err), fmt, detail ? detail : "");
    return code;
}
#define PF_CUDA(call)                                                                      \
    do {                                                                                   \
        cudaError_t err__ = (call);                                                        \
        if (err__ != cudaSuccess) return pf_fail(2, #call ": %s", cudaGetErrorString(err__)); \
    } while (0)

extern "C" const char* pf_last_error(void) { return g_err; }
void pf_set_error(const char* msg) { snprintf(g_err, sizeof(g_err), "%s", msg); }   // used by pf_rl.cu / pf_mlp_tc.cu
// diagnostic: install (or with NULL remove) a device log of 4096 uint64 that the env kernels append timeline records to
extern "C" int pf_debug_timeline(unsigned long long* device_log) {
    return pf_timeline_install(device_log) == cudaSuccess ? 0 : pf_fail(7, "%s", "pf_debug_timeline: not compiled in (build with PF_BUILD_TIMELINE=1) or cudaMemcpyToSymbol failed");
}
unsigned long long g_pf_launch_count = 0;
extern "C" unsigned long long pf_debug_launch_count(void) { return g_pf_launch_count; }
extern "C" int pf_abi_version(void) { return PF_ABI_VERSION; }
extern "C" int pf_sizeof_params(void) { return (int)sizeof(pf_params); }
extern "C" int pf_sizeof_env_buffers(void) { return (int)sizeof(pf_env_buffers); }

static int pf_pick_epw(int num_envs, int num_sms) {
    // envs per warp: small batches are latency bound, so spread them over more warps (idle lanes in the
    // one-lane-per-env phase are cheaper than idle SMs); from ~16 warps per SM on, full warps
    if (const char* e = getenv("PF_EPW")) { const int v = atoi(e); if (v == 1 || v == 2 || v == 4 || v == 8 || v == 16 || v == 32) return v; }
    if (num_envs <= num_sms * 16 * 8) return 8;
    if (num_envs <= num_sms * 16 * 16) return 16;
    return 32;
}

static int pf_ctx_init(pf_ctx* c, const float* noise_scale_vec_host, const float* height_points_xy_host);

extern "C" int pf_ctx_create(const pf_params* params, const float* noise_scale_vec_host,
                             const float* height_points_xy_host, pf_ctx** out) {
    if (!params || !out) return pf_fail(1, "pf_ctx_create: %s", "null argument");
    if (params->abi_version != PF_ABI_VERSION) return pf_fail(1, "pf_ctx_create: %s", "ABI version mismatch");
    const pf_params& P = *params;
    if (P.num_envs <= 0 || P.num_dof <= 0 || P.num_dof > PF_MAX_DOF || P.num_bodies <= 0 ||
        P.num_bodies > PF_MAX_BODIES || P.num_reward_terms < 0 ||
        P.num_reward_terms + (P.has_termination_term ? 1 : 0) > PF_MAX_REWARD_TERMS ||
        P.num_height_points < 0 || P.num_height_points > PF_MAX_HEIGHT_POINTS || P.num_obs <= 0 ||
        P.num_obs > PF_MAX_OBS || P.num_priv_obs > PF_MAX_OBS || P.num_commands < 3 || P.num_commands > 4)
        return pf_fail(1, "pf_ctx_create: %s", "parameter out of range");
    // observation widths must match what compute_observations builds (RuntimeError at PF:363-365, 346-348)
    const int base_w = 6 + 3 * P.num_dof + 3;
    if (P.num_obs != base_w + (P.measure_heights_actor ? P.num_height_points : 0))
        return pf_fail(3, "pf_ctx_create: %s", "computed observation width does not match num_obs");
    if (P.num_priv_obs && P.num_priv_obs != base_w + (P.measure_heights_critic ? P.num_height_points : 0))
        return pf_fail(3, "pf_ctx_create: %s", "computed privileged observation width does not match num_privileged_obs");
    if (P.control_type < PF_CONTROL_P || P.control_type > PF_CONTROL_T)
        return pf_fail(4, "pf_ctx_create: %s", "unknown controller type");

    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return pf_fail(2, "pf_ctx_create: no CUDA device (%s); there is no CPU fallback", cudaGetErrorString(e));

    pf_ctx* c = new pf_ctx();
    memset(&c->S, 0, sizeof(c->S));
    c->d_actions_in = nullptr; c->io_stream = nullptr;
    c->ev_fork = c->ev_in = c->ev_post = c->ev_out = c->ev_fin = nullptr;
    c->P = P;
    if (int rc = pf_ctx_init(c, noise_scale_vec_host, height_points_xy_host)) { pf_ctx_destroy(c); return rc; }
    *out = c;
    return 0;
}

// everything of pf_ctx_create that can fail after the context object exists (the caller destroys it on a non-zero return)
static int pf_ctx_init(pf_ctx* c, const float* noise_scale_vec_host, const float* height_points_xy_host) {
    const pf_params& P = c->P;
    PF_CUDA(cudaGetDevice(&c->device));
    PF_CUDA(cudaDeviceGetAttribute(&c->num_sms, cudaDevAttrMultiProcessorCount, c->device));
    c->envs_per_warp = pf_pick_epw(P.num_envs, c->num_sms);
    c->num_slabs = (P.num_envs + c->envs_per_warp - 1) / c->envs_per_warp;
    // CTA size of the step kernel: warps are independent, so a small batch is cut into small CTAs that reach every SM
    c->threads = 128;          // measured: 128-thread CTAs beat 256 at large batches too (finer-grained SM fill)
    while (c->threads > 32 && (c->num_slabs + c->threads / 32 - 1) / (c->threads / 32) < 2 * c->num_sms) c->threads >>= 1;
    if (const char* e = getenv("PF_BLOCK")) { const int v = atoi(e); if (v == 32 || v == 64 || v == 128) c->threads = v; }
    c->num_blocks = (c->num_slabs + c->threads / 32 - 1) / (c->threads / 32);
    c->max_blocks = c->num_slabs;
    c->smem_bytes = (size_t)(c->threads / 32) * pf_warp_smem_floats(P, c->envs_per_warp) * sizeof(float);
    c->canonical = pf_step_mode(P, c->S.rew, getenv("PF_GENERIC_REWARDS") != nullptr, getenv("PF_TABLE_REWARDS") != nullptr);
    if (c->smem_bytes > 200 * 1024) return pf_fail(1, "pf_ctx_create: %s", "shared memory budget exceeded");
    c->step_kernel = pf_pick_step_kernel(c->canonical, P.num_height_points > 0);
    c->generic_kernel = pf_pick_step_kernel(0, P.num_height_points > 0);
    PF_CUDA(cudaFuncSetAttribute(c->step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c->smem_bytes));
    PF_CUDA(cudaFuncSetAttribute(c->generic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c->smem_bytes));

    const int nv = P.num_priv_obs > P.num_obs ? P.num_priv_obs : P.num_obs;
    float* d_noise = nullptr; float* d_pts = nullptr;
    PF_CUDA(cudaMalloc(&d_noise, sizeof(float) * nv));
    c->S.noise_vec = d_noise;
    if (noise_scale_vec_host) PF_CUDA(cudaMemcpy(d_noise, noise_scale_vec_host, sizeof(float) * nv, cudaMemcpyHostToDevice));
    else PF_CUDA(cudaMemset(d_noise, 0, sizeof(float) * nv));
    PF_CUDA(cudaMalloc(&d_pts, sizeof(float) * 2 * (P.num_height_points > 0 ? P.num_height_points : 1)));
    c->S.height_points = d_pts;
    if (P.num_height_points > 0) {
        if (!height_points_xy_host) return pf_fail(1, "pf_ctx_create: %s", "height points missing");
        PF_CUDA(cudaMemcpy(d_pts, height_points_xy_host, sizeof(float) * 2 * P.num_height_points, cudaMemcpyHostToDevice));
    }
    PF_CUDA(cudaMalloc(&c->S.height_mean, sizeof(float) * (size_t)P.num_envs));
    PF_CUDA(cudaMalloc(&c->S.blk_count, sizeof(int) * c->max_blocks));
    PF_CUDA(cudaMalloc(&c->S.blk_level, sizeof(long long) * c->max_blocks));
    PF_CUDA(cudaMalloc(&c->S.blk_sums, sizeof(double) * PF_MAX_REWARD_TERMS * c->max_blocks));
    PF_CUDA(cudaMemset(c->S.blk_count, 0, sizeof(int) * c->max_blocks));
    const int fin_ctas = (P.num_envs + 1023) / 1024;
    PF_CUDA(cudaMalloc(&c->S.cta_sums, sizeof(double) * PF_MAX_REWARD_TERMS * fin_ctas));
    PF_CUDA(cudaMalloc(&c->S.cta_lvl, sizeof(long long) * fin_ctas));
    PF_CUDA(cudaMalloc(&c->S.ticket, sizeof(int)));
    PF_CUDA(cudaMemset(c->S.ticket, 0, sizeof(int)));
    PF_CUDA(cudaMalloc(&c->d_actions_in, sizeof(float) * (size_t)P.num_envs * P.num_dof));
    PF_CUDA(cudaStreamCreateWithFlags(&c->io_stream, cudaStreamNonBlocking));
    for (cudaEvent_t* e : {&c->ev_fork, &c->ev_in, &c->ev_post, &c->ev_out, &c->ev_fin}) PF_CUDA(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
    c->out_pending = 0;
    c->fin_pending = 0;
    return 0;
}

extern "C" int pf_set_command_ranges(pf_ctx* c, const float* ranges_host) {
    if (!c || !ranges_host) return pf_fail(1, "%s", "pf_set_command_ranges: null argument");
    for (int i = 0; i < 4; ++i) { c->P.cmd_range[i][0] = ranges_host[2 * i]; c->P.cmd_range[i][1] = ranges_host[2 * i + 1]; }
    return 0;
}

extern "C" int pf_ctx_destroy(pf_ctx* c) {
    if (!c) return 0;
    cudaFree((void*)c->S.noise_vec); cudaFree((void*)c->S.height_points);
    cudaFree(c->S.height_mean); cudaFree(c->S.cta_sums); cudaFree(c->S.cta_lvl); cudaFree(c->S.ticket); cudaFree(c->S.blk_count); cudaFree(c->S.blk_level); cudaFree(c->S.blk_sums); cudaFree(c->d_actions_in);
    if (c->io_stream) cudaStreamDestroy(c->io_stream);
    for (cudaEvent_t e : {c->ev_fork, c->ev_in, c->ev_post, c->ev_out, c->ev_fin}) if (e) cudaEventDestroy(e);
    delete c;
    return 0;
}

static int pf_check_bufs(const pf_ctx* c, const pf_env_buffers* b) {
    if (!c || !b) return pf_fail(1, "%s", "null ctx / buffers");
    const pf_params& P = c->P;
    if (!b->root_states || !b->dof_state || !b->contact_forces || !b->rigid_body_states || !b->actions ||
        !b->torques || !b->commands || !b->episode_length_buf || !b->base_lin_vel || !b->base_ang_vel ||
        !b->projected_gravity || !b->last_actions || !b->last_dof_vel || !b->last_root_vel || !b->feet_air_time ||
        !b->last_feet_air_time || !b->current_max_feet_height || !b->last_max_feet_height || !b->feet_height ||
        !b->contact_filt || !b->first_contact || !b->last_contacts || !b->episode_sums || !b->rew_buf ||
        !b->reset_buf || !b->time_out_buf || !b->terminate || !b->obs_buf || !b->env_origins || !b->reset_ids ||
        !b->reset_count || !b->episode_stats)
        return pf_fail(1, "%s", "pf_env_buffers: a required pointer is NULL");
    if (P.num_priv_obs && !b->priv_obs_buf) return pf_fail(1, "%s", "pf_env_buffers: priv_obs_buf is NULL");
    if (P.terrain_kind == PF_TERRAIN_HEIGHTFIELD &&
        (!b->height_samples || !b->terrain_origins || !b->terrain_levels || !b->terrain_types))
        return pf_fail(1, "%s", "pf_env_buffers: terrain pointers are NULL for a heightfield task");
    return 0;
}

extern "C" int pf_pd_torques(pf_ctx* c, const pf_env_buffers* bufs, const float* actions_in, void* stream) {
    if (int rc = pf_check_bufs(c, bufs)) return rc;
    const int n = c->P.num_envs * c->P.num_dof;
    const uintptr_t al8 = reinterpret_cast<uintptr_t>(bufs->actions) | reinterpret_cast<uintptr_t>(bufs->torques) |
                          reinterpret_cast<uintptr_t>(bufs->last_dof_vel) | reinterpret_cast<uintptr_t>(actions_in);
    if ((c->P.num_dof & 1) == 0 && (al8 & 7) == 0 && (reinterpret_cast<uintptr_t>(bufs->dof_state) & 15) == 0) {
        const int pairs = n / 2;
        PF_CUDA(pf_launch_pdl(pf_pd_torques2_kernel, dim3((pairs + 255) / 256), dim3(256), 0, (cudaStream_t)stream, c->P, *bufs, actions_in));
        return 0;
    }
    int blocks = (n + 255) / 256;
    if (blocks > c->num_sms * 8) blocks = c->num_sms * 8;
    PF_CUDA(pf_launch_pdl(pf_pd_torques_kernel, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, c->P, *bufs, actions_in));
    return 0;
}

static int pf_launch_finalize(pf_ctx* c, const pf_env_buffers* bufs, cudaStream_t st, int count_step) {
    const int blocks = (c->P.num_envs + 1023) / 1024;
    PF_CUDA(pf_launch_pdl(pf_finalize_kernel, dim3(blocks), dim3(1024), 0, st, c->P, *bufs, c->S, c->envs_per_warp, c->num_slabs, count_step));
    return 0;
}

// post_physics kernel on `st`; the finalize kernel behind it on `st`, or (side) forked onto the context's io stream so
// that what the caller queues next on `st` runs beside it.
static int pf_post_physics_impl(pf_ctx* c, const pf_env_buffers* bufs, int64_t common_step_counter, cudaStream_t st, bool side) {
    if (int rc = pf_check_bufs(c, bufs)) return rc;
    const uint32_t step32 = (uint32_t)common_step_counter;
    const dim3 g(c->num_blocks), b(c->threads);
    if (c->P.num_height_points > 0) {
        // envs per warp of the scan: as many as keep >= 8 warps per SM sub-partition busy, 1..16 (32 measured no better)
        int sepw = 1;
        while (sepw < 16 && c->P.num_envs / (sepw * 2) >= c->num_sms * 4 * 8) sepw *= 2;
        if (const char* e = getenv("PF_SCAN_EPW")) { const int v = atoi(e); if (v >= 1 && v <= 32) sepw = v; }
        const int warps = (c->P.num_envs + sepw - 1) / sepw;
        int blocks = (warps + 7) / 8;
        if (blocks > c->num_sms * 8) blocks = c->num_sms * 8;
        // the specialised scan (pf_height_scan_std_kernel) where every condition it compiles in holds; PF_SCAN_NO_STD=1: never
        const char* nse = getenv("PF_SCAN_NO_STD");
        const int no_std = nse ? atoi(nse) : 0;
        const int Pn = c->P.num_height_points, kind = c->S.rew.index_div_kind;
        const bool stdcfg = !no_std && c->canonical && (kind == 1 || kind == 2) && c->P.terrain_kind == PF_TERRAIN_HEIGHTFIELD &&
                            !bufs->measured_heights && c->P.measure_heights_critic && bufs->priv_obs_buf &&
                            c->P.add_noise;
        if (stdcfg && ((Pn > 96 && Pn <= 128) || (Pn > 160 && Pn <= 192))) {
            using ScanFn = void (*)(const pf_params, const pf_env_buffers, const PfCtxScratch, const int, const uint32_t);
            static const ScanFn fns[2][2][2] = {       // [kind - 1][passes 4 / 6][actor]
                {{pf_height_scan_std_kernel<1, 4, false>, pf_height_scan_std_kernel<1, 4, true>},
                 {pf_height_scan_std_kernel<1, 6, false>, pf_height_scan_std_kernel<1, 6, true>}},
                {{pf_height_scan_std_kernel<2, 4, false>, pf_height_scan_std_kernel<2, 4, true>},
                 {pf_height_scan_std_kernel<2, 6, false>, pf_height_scan_std_kernel<2, 6, true>}}};
            PF_CUDA(pf_launch_pdl(fns[kind - 1][Pn > 128 ? 1 : 0][c->P.measure_heights_actor ? 1 : 0], dim3(blocks), dim3(256), 0, st,
                                  c->P, *bufs, c->S, sepw, step32));
        } else if (c->canonical) PF_CUDA(pf_launch_pdl(pf_height_scan_kernel<true>, dim3(blocks), dim3(256), 0, st, c->P, *bufs, c->S, sepw, step32));
        else PF_CUDA(pf_launch_pdl(pf_height_scan_kernel<false>, dim3(blocks), dim3(256), 0, st, c->P, *bufs, c->S, sepw, step32));
    }
    // the specialised kernels read / write pairs of the per-env state as 64-bit words: 8-byte aligned tensors only
    const uintptr_t al = reinterpret_cast<uintptr_t>(bufs->torques) | reinterpret_cast<uintptr_t>(bufs->last_dof_vel) |
                         reinterpret_cast<uintptr_t>(bufs->last_actions) | reinterpret_cast<uintptr_t>(bufs->actions) |
                         reinterpret_cast<uintptr_t>(bufs->last_root_vel) | reinterpret_cast<uintptr_t>(bufs->feet_air_time) |
                         reinterpret_cast<uintptr_t>(bufs->last_feet_air_time) | reinterpret_cast<uintptr_t>(bufs->feet_height) |
                         reinterpret_cast<uintptr_t>(bufs->current_max_feet_height) | reinterpret_cast<uintptr_t>(bufs->last_max_feet_height) |
                         ((reinterpret_cast<uintptr_t>(bufs->contact_filt) | reinterpret_cast<uintptr_t>(bufs->first_contact) |
                           reinterpret_cast<uintptr_t>(bufs->last_contacts)) << 2);
    const pf_step_kernel_t kern = (al & 7) == 0 ? c->step_kernel : c->generic_kernel;
    PF_CUDA(pf_launch_pdl(kern, g, b, c->smem_bytes, st, c->P, *bufs, c->S, c->envs_per_warp, step32));
    if (!side) return pf_launch_finalize(c, bufs, st, 1);
    PF_CUDA(cudaEventRecord(c->ev_post, st));
    PF_CUDA(cudaStreamWaitEvent(c->io_stream, c->ev_post, 0));
    if (int rc = pf_launch_finalize(c, bufs, c->io_stream, 1)) return rc;
    PF_CUDA(cudaEventRecord(c->ev_fin, c->io_stream));
    c->fin_pending = 1;
    return 0;
}

extern "C" int pf_post_physics_step(pf_ctx* c, const pf_env_buffers* bufs, int64_t common_step_counter, void* stream) {
    return pf_post_physics_impl(c, bufs, common_step_counter, (cudaStream_t)stream, false);
}

// The same step with its tail (reset id list, device count, extras["episode"] statistics, the step counter) left
// running on a context-owned stream: the observations, rewards and reset flags are complete in `stream` order, so
// the policy's act kernel queued next overlaps the finalize kernel.  pf_post_physics_step_join makes `stream` wait for
// the tail; it must come before anything reads reset_ids / reset_count / episode_stats / extras_time_outs, before
// the next step of this context, and before a stream capture ends.
extern "C" int pf_post_physics_step_begin(pf_ctx* c, const pf_env_buffers* bufs, int64_t common_step_counter, void* stream) {
    return pf_post_physics_impl(c, bufs, common_step_counter, (cudaStream_t)stream, true);
}

extern "C" int pf_post_physics_step_join(pf_ctx* c, void* stream) {
    if (!c) return pf_fail(1, "%s", "pf_post_physics_step_join: null context");
    if (c->fin_pending) {
        PF_CUDA(cudaStreamWaitEvent((cudaStream_t)stream, c->ev_fin, 0));
        c->fin_pending = 0;
    }
    return 0;
}

extern "C" int pf_reset_all(pf_ctx* c, const pf_env_buffers* bufs, int64_t common_step_counter, void* stream) {
    if (int rc = pf_check_bufs(c, bufs)) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    PF_CUDA(cudaMemsetAsync(bufs->reset_buf, 1, (size_t)c->P.num_envs, st));
    PF_LAUNCHED(), pf_reset_all_kernel<<<c->num_blocks, c->threads, 0, st>>>(c->P, *bufs, c->S, c->envs_per_warp,
                                                                          (uint32_t)common_step_counter);
    PF_CUDA(cudaGetLastError());
    return pf_launch_finalize(c, bufs, st, 0);
}

extern "C" int pf_push_forces(pf_ctx* c, const pf_env_buffers* bufs, float* forces, float max_push_force,
                              int64_t common_step_counter, void* stream) {
    if (int rc = pf_check_bufs(c, bufs)) return rc;
    if (!forces) return pf_fail(1, "%s", "pf_push_forces: forces is NULL");
    const int n = c->P.num_envs * c->P.num_bodies * 3;
    int blocks = (n + 255) / 256;
    if (blocks > c->num_sms * 8) blocks = c->num_sms * 8;
    PF_LAUNCHED(), pf_push_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(c->P, *bufs, forces, max_push_force,
                                                             (uint32_t)common_step_counter);
    PF_CUDA(cudaGetLastError());
    return 0;
}

// ---- pf_step_host data movement -----------------------------------------------------------------------------------
// With PINNED (mapped) host buffers the device reads / writes them directly over the bus: ONE gather kernel pulls the
// PhysX state + actions in, ONE scatter kernel pushes the results out, instead of a dozen cudaMemcpyAsync calls whose
// fixed costs add up (measured at 4,096 envs: 113 us of copies in for 1 MB, 59 us out).  Pageable buffers fall
// back to the copies.
#define PF_IO_SEGS 8
struct PfIoSeg { const void* src; void* dst; unsigned long long bytes; };
struct PfIoArgs {
    PfIoSeg seg[PF_IO_SEGS]; int nseg;
    // feet rows of rigid_body_states: xyz of body feet_idx[f] of every env (the only columns the path reads)
    const float* rbs_src; float* rbs_dst; int num_envs, num_bodies, feet_idx[PF_NUM_FEET];
};

__global__ void __launch_bounds__(256) pf_host_io_kernel(const PfIoArgs a) {
    const int s = blockIdx.y;
    const unsigned long long tid = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x, nth = (unsigned long long)gridDim.x * blockDim.x;
    if (s < a.nseg) {
        const PfIoSeg g = a.seg[s];
        if (((reinterpret_cast<uintptr_t>(g.src) | reinterpret_cast<uintptr_t>(g.dst)) & 15) == 0) {
            const unsigned long long n16 = g.bytes >> 4;
            const uint4* src = static_cast<const uint4*>(g.src);
            uint4* dst = static_cast<uint4*>(g.dst);
            for (unsigned long long i = tid; i < n16; i += nth) dst[i] = src[i];
            const unsigned char* sb = static_cast<const unsigned char*>(g.src);
            unsigned char* db = static_cast<unsigned char*>(g.dst);
            for (unsigned long long i = (n16 << 4) + tid; i < g.bytes; i += nth) db[i] = sb[i];
        } else {
            const unsigned char* sb = static_cast<const unsigned char*>(g.src);
            unsigned char* db = static_cast<unsigned char*>(g.dst);
            for (unsigned long long i = tid; i < g.bytes; i += nth) db[i] = sb[i];
        }
    } else if (a.rbs_src) {
        const unsigned long long n = (unsigned long long)a.num_envs * PF_NUM_FEET * 3;
        for (unsigned long long i = tid; i < n; i += nth) {
            const int e = (int)(i / (PF_NUM_FEET * 3)), r = (int)(i % (PF_NUM_FEET * 3)), f = r / 3, c = r % 3;
            const size_t off = ((size_t)e * a.num_bodies + a.feet_idx[f]) * 13 + c;
            a.rbs_dst[off] = a.rbs_src[off];
        }
    }
}

// device-visible alias of a pinned host pointer, or NULL if the memory is pageable
static const void* pf_mapped(const void* host) {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, host) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return (at.type == cudaMemoryTypeHost) ? at.devicePointer : nullptr;
}

static int pf_step_host_impl(pf_ctx* c, const pf_env_buffers* bufs, const pf_host_state* hs, const float* actions_host,
                             float* obs_host, float* priv_obs_host, float* rew_host, uint8_t* reset_host,
                             float* torques_host, int64_t common_step_counter, void* stream, bool sync, bool join) {
    if (int rc = pf_check_bufs(c, bufs)) return rc;
    if (!actions_host || !obs_host || !rew_host || !reset_host) return pf_fail(1, "%s", "pf_step_host: null host buffer");
    const pf_params& P = c->P;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t N = (size_t)P.num_envs;
    const cudaMemcpyKind H2D = cudaMemcpyHostToDevice, D2H = cudaMemcpyDeviceToHost;
    static const int dbg = [] { const char* e = getenv("PF_HOST_DEBUG"); return e ? atoi(e) : 0; }();   // experiments: 1 no copies in, 2 none out, 4 always cudaMemcpy, 8 no side stream
    const size_t rbs_pitch = sizeof(float) * 13 * P.num_bodies;
    const cudaStream_t io = (dbg & 8) ? st : c->io_stream;          // 8: no side stream (everything in order on `stream`)

    // ---- in ----
    // `in` = what the first torque kernel reads (dof_state, actions); `in2` = the rest (root, contacts, feet rows), which
    // only post_physics_step needs: on io_stream it crosses the bus while the torque sub-steps run.
    PfIoArgs in = {}, in2 = {};
    bool mapped = !(dbg & 4);
    bool forked = false;
    auto add = [&](PfIoArgs& a, const void* src, void* dst, size_t bytes, bool src_is_host) {
        if (!src || !dst) return;
        const void* h = src_is_host ? src : dst;
        const void* m = mapped ? pf_mapped(h) : nullptr;
        if (!m) { mapped = false; return; }
        PfIoSeg& g = a.seg[a.nseg++];
        g.src = src_is_host ? m : src; g.dst = src_is_host ? dst : const_cast<void*>(m); g.bytes = bytes;
    };
    auto io_grid = [&](size_t bytes, int ny) {
        const size_t gx = (bytes / 16 + 255) / 256;
        return dim3((unsigned)(gx < 1 ? 1 : gx > 592 ? 592 : gx), (unsigned)ny);
    };
    if (!(dbg & 1)) {
        if (hs) {
            add(in, hs->dof_state, bufs->dof_state, 4 * N * P.num_dof * 2, true);
            add(in2, hs->root_states, bufs->root_states, 4 * N * 13, true);
            add(in2, hs->contact_forces, (void*)bufs->contact_forces, 4 * N * P.num_bodies * 3, true);
            if (hs->rigid_body_states) {
                const void* m = mapped ? pf_mapped(hs->rigid_body_states) : nullptr;
                if (m) {
                    in2.rbs_src = static_cast<const float*>(m); in2.rbs_dst = const_cast<float*>(bufs->rigid_body_states);
                    in2.num_envs = P.num_envs; in2.num_bodies = P.num_bodies;
                    for (int f = 0; f < PF_NUM_FEET; ++f) in2.feet_idx[f] = P.feet_indices[f];
                } else mapped = false;
            }
        }
        add(in, actions_host, c->d_actions_in, sizeof(float) * N * P.num_dof, true);
        if (mapped) {
            PF_LAUNCHED(), pf_host_io_kernel<<<io_grid(4 * N * P.num_dof * 2, in.nseg), 256, 0, st>>>(in);
            PF_CUDA(cudaGetLastError());
            if (in2.nseg || in2.rbs_src) {           // forked behind the first gather: the bus is not shared with it
                PF_CUDA(cudaEventRecord(c->ev_fork, st));
                PF_CUDA(cudaStreamWaitEvent(io, c->ev_fork, 0));
                PF_LAUNCHED(), pf_host_io_kernel<<<io_grid(4 * N * P.num_bodies * 3, in2.nseg + (in2.rbs_src ? 1 : 0)), 256, 0, io>>>(in2);
                PF_CUDA(cudaGetLastError());
                PF_CUDA(cudaEventRecord(c->ev_in, io));
                forked = true;
            }
        } else {
            if (hs) {
                if (hs->root_states) PF_CUDA(cudaMemcpyAsync(bufs->root_states, hs->root_states, 4 * N * 13, H2D, st));
                if (hs->dof_state) PF_CUDA(cudaMemcpyAsync(bufs->dof_state, hs->dof_state, 4 * N * P.num_dof * 2, H2D, st));
                if (hs->contact_forces)
                    PF_CUDA(cudaMemcpyAsync((void*)bufs->contact_forces, hs->contact_forces, 4 * N * P.num_bodies * 3, H2D, st));
                if (hs->rigid_body_states)      // only the feet position columns (PF:1083-1086, 1114-1129): one pitched copy per foot
                    for (int f = 0; f < PF_NUM_FEET; ++f) {
                        const size_t off = (size_t)P.feet_indices[f] * 13;
                        PF_CUDA(cudaMemcpy2DAsync((void*)(bufs->rigid_body_states + off), rbs_pitch, hs->rigid_body_states + off,
                                                  rbs_pitch, 3 * sizeof(float), N, H2D, st));
                    }
            }
            PF_CUDA(cudaMemcpyAsync(c->d_actions_in, actions_host, sizeof(float) * N * P.num_dof, H2D, st));
        }
    }
    for (int i = 0; i < P.decimation; ++i)
        if (int rc = pf_pd_torques(c, bufs, i == 0 ? (const float*)c->d_actions_in : nullptr, stream)) return rc;
    if (forked) PF_CUDA(cudaStreamWaitEvent(st, c->ev_in, 0));
    // pinned buffers + side stream: the finalize kernel goes to the side stream too (in front of the scatter), off the
    // path of what the caller queues next on `stream`
    const bool side_tail = mapped && !(dbg & 8) && !(dbg & 2);
    if (int rc = pf_post_physics_impl(c, bufs, common_step_counter, st, side_tail)) return rc;

    // ---- out ----  (pinned: on io_stream, so that work the caller queues next -- the policy -- overlaps the transfer)
    if (!(dbg & 2)) {
        PfIoArgs out = {};
        mapped = !(dbg & 4);
        add(out, bufs->obs_buf, obs_host, sizeof(float) * N * P.num_obs, false);
        if (priv_obs_host && bufs->priv_obs_buf) add(out, bufs->priv_obs_buf, priv_obs_host, sizeof(float) * N * P.num_priv_obs, false);
        add(out, bufs->rew_buf, rew_host, sizeof(float) * N, false);
        add(out, bufs->reset_buf, reset_host, N, false);
        if (torques_host) add(out, bufs->torques, torques_host, sizeof(float) * N * P.num_dof, false);
        if (mapped) {
            if (!side_tail) {                                    // (side_tail: io already waits for the step kernel)
                PF_CUDA(cudaEventRecord(c->ev_post, st));
                PF_CUDA(cudaStreamWaitEvent(io, c->ev_post, 0));
            }
            PF_LAUNCHED(), pf_host_io_kernel<<<io_grid(sizeof(float) * N * (P.num_priv_obs > P.num_obs ? P.num_priv_obs : P.num_obs), out.nseg), 256, 0, io>>>(out);
            PF_CUDA(cudaGetLastError());
            PF_CUDA(cudaEventRecord(c->ev_out, io));
            c->out_pending = 1;
            if (join) { PF_CUDA(cudaStreamWaitEvent(st, c->ev_out, 0)); c->out_pending = 0; c->fin_pending = 0; }
        } else {
            if (c->fin_pending) { PF_CUDA(cudaStreamWaitEvent(st, c->ev_fin, 0)); c->fin_pending = 0; }
            PF_CUDA(cudaMemcpyAsync(obs_host, bufs->obs_buf, sizeof(float) * N * P.num_obs, D2H, st));
            if (priv_obs_host && bufs->priv_obs_buf)
                PF_CUDA(cudaMemcpyAsync(priv_obs_host, bufs->priv_obs_buf, sizeof(float) * N * P.num_priv_obs, D2H, st));
            PF_CUDA(cudaMemcpyAsync(rew_host, bufs->rew_buf, sizeof(float) * N, D2H, st));
            PF_CUDA(cudaMemcpyAsync(reset_host, bufs->reset_buf, N, D2H, st));
            if (torques_host) PF_CUDA(cudaMemcpyAsync(torques_host, bufs->torques, sizeof(float) * N * P.num_dof, D2H, st));
        }
    }
    if (sync) PF_CUDA(cudaStreamSynchronize(st));
    return 0;
}

extern "C" int pf_step_host(pf_ctx* c, const pf_env_buffers* bufs, const pf_host_state* hs, const float* actions_host,
                            float* obs_host, float* priv_obs_host, float* rew_host, uint8_t* reset_host,
                            float* torques_host, int64_t common_step_counter, void* stream) {
    return pf_step_host_impl(c, bufs, hs, actions_host, obs_host, priv_obs_host, rew_host, reset_host, torques_host,
                             common_step_counter, stream, true, true);
}

extern "C" int pf_step_host_async(pf_ctx* c, const pf_env_buffers* bufs, const pf_host_state* hs, const float* actions_host,
                                  float* obs_host, float* priv_obs_host, float* rew_host, uint8_t* reset_host,
                                  float* torques_host, int64_t common_step_counter, void* stream) {
    return pf_step_host_impl(c, bufs, hs, actions_host, obs_host, priv_obs_host, rew_host, reset_host, torques_host,
                             common_step_counter, stream, false, true);
}

// pf_step_host_async without the final join: the results' transfer to the pinned host buffers is left running on a
// context-owned stream, so that what the caller queues next on `stream` (the policy's act kernel on the device
// observations) overlaps it.  pf_step_host_join(ctx, stream) makes `stream` wait for that transfer; it must be called
// before the host reads the outputs, before the next step, and before a stream capture ends.
extern "C" int pf_step_host_begin(pf_ctx* c, const pf_env_buffers* bufs, const pf_host_state* hs, const float* actions_host,
                                  float* obs_host, float* priv_obs_host, float* rew_host, uint8_t* reset_host,
                                  float* torques_host, int64_t common_step_counter, void* stream) {
    return pf_step_host_impl(c, bufs, hs, actions_host, obs_host, priv_obs_host, rew_host, reset_host, torques_host,
                             common_step_counter, stream, false, false);
}

extern "C" int pf_step_host_join(pf_ctx* c, void* stream) {
    if (!c) return pf_fail(1, "%s", "pf_step_host_join: null context");
    if (c->out_pending) {
        PF_CUDA(cudaStreamWaitEvent((cudaStream_t)stream, c->ev_out, 0));
        c->out_pending = 0;
    }
    if (c->fin_pending) {
        PF_CUDA(cudaStreamWaitEvent((cudaStream_t)stream, c->ev_fin, 0));
        c->fin_pending = 0;
    }
    return 0;
}
