// Env-side kernels of the hot path (sm_100a): PD torques, the fused post_physics_step, reset compaction and
// episode statistics, push forces, reset-all.  The per-env arithmetic lives in pf_env_core.cuh (shared with
// the host emulation in tests/); this file is the staging, the thread mapping and the block reductions.
//
// Mapping: a CTA of PF_THREADS threads owns EPB consecutive envs.
//   phase 0  cooperative, coalesced copy of the CTA's PhysX rows (root [EPB,13], dof [EPB,2D], contact forces
//            [EPB,3B]) into shared memory -- the AoS -> per-thread transpose happens there, not in HBM;
//   phase A  height scan: the EPB*P scan points are spread over all threads (min of 3 int16 samples each, the
//            5.46 MB field stays L2 resident), raw samples parked in shared memory;
//   phase B  one thread per env: pf_env_step (+ pf_env_reset for the few that reset);
//   phase C  cooperative, coalesced write of the observation rows (noise + clip fused);
//   phase D  CTA partials for extras["episode"] (count, per-term sums over reset envs, terrain-level sum).
// A second, tiny kernel (pf_finalize_kernel) turns the per-CTA partials into the ascending reset id list
// (PF:221 nonzero()), the device count and the episode statistics, deterministically (no float atomics).
#include <cuda_runtime.h>
#include <stdio.h>
#include "pf_env_core.cuh"
#include "pf_ctx.h"

#define PF_THREADS 128

// ------------------------------------------------------------------------------------------------ K1
// _compute_torques, PF:523-547 (+ the action clip of PF:176-177 when actions_in != NULL).
// One thread per (env, dof): every stream is contiguous -> fully coalesced, dof_state read as float2.
__global__ void __launch_bounds__(256) pf_pd_torques_kernel(const pf_params P, const pf_env_buffers B,
                                                            const float* __restrict__ actions_in) {
    const int n = P.num_envs * P.num_dof;
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

// ------------------------------------------------------------------------------------------------ helpers
// dynamic shared memory carve-up for the step kernel
struct PfSmem {
    float* root;      // [EPB][13]
    float* dof;       // [EPB][2D+1]   (+1: odd stride, conflict-free per-thread rows)
    float* cf;        // [EPB][3B]
    float* feet;      // [EPB][6]
    float* base;      // [EPB][base_w] observation columns before noise
    float* rootz;     // [EPB] post-reset root z
    float* zw;        // [EPB][2] yaw quaternion (z, w)
    float* sums;      // [EPB][R] episode sums of reset envs
    uint32_t* nkey;   // [EPB] noise key
    int16_t* raw;     // [EPB][P]
};

__host__ __device__ inline size_t pf_smem_floats(const pf_params& P, int epb) {
    const int R = P.num_reward_terms + (P.has_termination_term ? 1 : 0);
    return (size_t)epb * (13 + (2 * P.num_dof + 1) + 3 * P.num_bodies + 6 + (6 + 3 * P.num_dof + 3) + 1 + 2 + R + 1);
}
__host__ __device__ inline size_t pf_smem_bytes(const pf_params& P, int epb) {
    size_t b = pf_smem_floats(P, epb) * 4 + (size_t)epb * P.num_height_points * 2;
    return (b + 15) & ~(size_t)15;
}

__device__ __forceinline__ PfSmem pf_carve(const pf_params& P, int epb, unsigned char* smem) {
    PfSmem s;
    const int R = P.num_reward_terms + (P.has_termination_term ? 1 : 0);
    float* f = reinterpret_cast<float*>(smem);
    s.root = f; f += epb * 13;
    s.dof = f; f += epb * (2 * P.num_dof + 1);
    s.cf = f; f += epb * 3 * P.num_bodies;
    s.feet = f; f += epb * 6;
    s.base = f; f += epb * (6 + 3 * P.num_dof + 3);
    s.rootz = f; f += epb;
    s.zw = f; f += epb * 2;
    s.sums = f; f += epb * R;
    s.nkey = reinterpret_cast<uint32_t*>(f); f += epb;
    s.raw = reinterpret_cast<int16_t*>(f);
    return s;
}

// CTA partials for extras["episode"] (PF:276-288).  `is_reset` per thread (threads >= nenv pass false),
// `lvl` = this env's terrain level after the update (0 if none).
__device__ __forceinline__ void pf_block_partials(const pf_params& P, const PfCtxScratch& S, const PfSmem& s,
                                                  bool is_reset, long long lvl, int R) {
    __shared__ unsigned s_mask[PF_THREADS / 32];
    __shared__ unsigned long long s_lvl;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) s_lvl = 0ull;
    const unsigned m = __ballot_sync(0xffffffffu, is_reset);
    if (lane == 0) s_mask[warp] = m;
    // levels are >= 0 and < max_terrain_level: exact integer sum
    unsigned lv = __reduce_add_sync(0xffffffffu, (unsigned)lvl);
    __syncthreads();
    if (lane == 0 && lv) atomicAdd(&s_lvl, (unsigned long long)lv);
    __syncthreads();
    int cnt = 0;
    for (int w = 0; w < PF_THREADS / 32; ++w) cnt += __popc(s_mask[w]);
    if (tid == 0) {
        S.blk_count[blockIdx.x] = cnt;
        S.blk_level[blockIdx.x] = (long long)s_lvl;
    }
    // per-term sums over this CTA's reset envs, in env order, double accumulation (deterministic)
    if (tid < R) {
        double acc = 0.0;
        if (cnt) {
            for (int w = 0; w < PF_THREADS / 32; ++w) {
                unsigned mm = s_mask[w];
                while (mm) {
                    const int b = __ffs(mm) - 1;
                    mm &= mm - 1;
                    acc += (double)s.sums[(w * 32 + b) * R + tid];
                }
            }
        }
        S.blk_sums[(size_t)blockIdx.x * PF_MAX_REWARD_TERMS + tid] = acc;
    }
}

// ------------------------------------------------------------------------------------------------ K2
// Fused post_physics_step (PF:195-227) + obs clip (PF:189-192).  `epb` envs per CTA (32, 64 or 128).
__global__ void __launch_bounds__(PF_THREADS) pf_post_physics_kernel(const pf_params P, const pf_env_buffers B,
                                                                     const PfCtxScratch S, const int epb,
                                                                     const uint32_t step_arg) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const PfSmem s = pf_carve(P, epb, smem_raw);
    const int tid = threadIdx.x;
    const int e0 = blockIdx.x * epb;
    const int nenv = min(epb, P.num_envs - e0);
    const uint32_t step = B.step_counter ? (uint32_t)(*B.step_counter + 1) : step_arg;
    const int D = P.num_dof, NB = P.num_bodies, Pn = P.num_height_points;
    const int R = P.num_reward_terms + (P.has_termination_term ? 1 : 0);
    const int base_w = pf_obs_base_width(P);

    // ---- phase 0: coalesced AoS rows -> shared memory ----
    {
        const float* g = B.root_states + (size_t)e0 * 13;
        for (int i = tid; i < nenv * 13; i += PF_THREADS) s.root[i] = g[i];
        g = B.dof_state + (size_t)e0 * 2 * D;
        const int w = 2 * D;
        for (int i = tid; i < nenv * w; i += PF_THREADS) s.dof[(i / w) * (w + 1) + (i % w)] = g[i];
        g = B.contact_forces + (size_t)e0 * 3 * NB;
        for (int i = tid; i < nenv * 3 * NB; i += PF_THREADS) s.cf[i] = g[i];
        for (int i = tid; i < nenv * 6; i += PF_THREADS) {
            const int e = i / 6, c = i % 6;
            s.feet[i] = B.rigid_body_states[((size_t)(e0 + e) * NB + P.feet_indices[c / 3]) * 13 + (c % 3)];
        }
    }
    __syncthreads();

    // ---- phase A: height scan (PF:1048-1075, 1100-1112) ----
    if (Pn > 0) {
        if (tid < nenv) {
            float zn, wn;
            pf_yaw_quat(s.root + tid * 13 + 3, zn, wn);
            s.zw[2 * tid] = zn; s.zw[2 * tid + 1] = wn;
        }
        __syncthreads();
        const bool rough = P.terrain_kind == PF_TERRAIN_HEIGHTFIELD;
        for (int i = tid; i < nenv * Pn; i += PF_THREADS) {
            const int e = i / Pn, p = i - e * Pn;
            int16_t r = 0;                                       // plane: zeros, PF:1061-1062
            if (rough) {
                const float2 g = reinterpret_cast<const float2*>(S.height_points)[p];
                r = pf_scan_point(P, B.height_samples, s.zw[2 * e], s.zw[2 * e + 1], g.x, g.y,
                                  s.root[e * 13], s.root[e * 13 + 1]);
            }
            s.raw[i] = r;
            if (B.measured_heights) B.measured_heights[(size_t)e0 * Pn + i] = (float)r * P.vertical_scale;
        }
        __syncthreads();
    }

    // ---- phase B: one thread per env ----
    bool is_reset = false;
    long long lvl = 0;
    if (tid < nenv) {
        const int e = e0 + tid;
        PfEnvIn in;
        PfEnvOut out;
        in.root = s.root + tid * 13;
        in.dof = s.dof + tid * (2 * D + 1);
        in.cf = s.cf + tid * 3 * NB;
        in.feet_pos[0] = s.feet + tid * 6;
        in.feet_pos[1] = s.feet + tid * 6 + 3;
        in.have_heights = Pn > 0;
        {
            const float rz = in.root[2];
            const int16_t* raw = s.raw + tid * Pn;
            double hs = 0.0;          // terms in fp32 like the reference, summed in double (see pf_env_core.cuh)
            for (int p = 0; p < Pn; ++p) hs += (double)(rz - (float)raw[p] * P.vertical_scale);
            in.height_mean = Pn > 0 ? (float)(hs / (double)Pn) : 0.0f;
        }
        pf_env_step(P, B, e, in, step, nullptr, out);
        is_reset = out.reset;
        if (is_reset) pf_env_reset(P, B, e, in.root, out.commands, step, true, s.sums + tid * R, out);
        if (B.terrain_levels) lvl = B.terrain_levels[e];
        const float* act = B.actions + (size_t)e * D;
        for (int j = 0; j < base_w; ++j) s.base[tid * base_w + j] = pf_obs_base_col(P, out, act, j);
        s.rootz[tid] = out.root_z;
        s.nkey[tid] = pf_rng_env_key(pf_rng_key(P.seed, step, PF_RNG_OBS_NOISE), (uint32_t)(P.env_id_offset + e));
        // PF:225-227
        for (int d = 0; d < D; ++d) {
            B.last_actions[(size_t)e * D + d] = act[d];
            B.last_dof_vel[(size_t)e * D + d] = out.dof_vel[d];
        }
        const float* rv = is_reset ? B.root_states + (size_t)e * 13 + 7 : in.root + 7;
        for (int i = 0; i < 6; ++i) B.last_root_vel[(size_t)e * 6 + i] = rv[i];
    }
    __syncthreads();

    // ---- phase C: observation rows, coalesced (PF:320-382, 189-192) ----
    {
        const int no = P.num_obs;
        float* g = B.obs_buf + (size_t)e0 * no;
        for (int i = tid; i < nenv * no; i += PF_THREADS) {
            const int e = i / no, j = i - e * no;
            const float v = j < base_w ? s.base[e * base_w + j]
                                       : pf_obs_height_col(P, s.rootz[e], s.raw[e * Pn + (j - base_w)]);
            g[i] = pf_obs_finish(P, v, s.nkey[e], j, S.noise_vec);
        }
        if (B.priv_obs_buf) {
            const int np = P.num_priv_obs;
            g = B.priv_obs_buf + (size_t)e0 * np;
            for (int i = tid; i < nenv * np; i += PF_THREADS) {
                const int e = i / np, j = i - e * np;
                const float v = j < base_w ? s.base[e * base_w + j]
                                           : pf_obs_height_col(P, s.rootz[e], s.raw[e * Pn + (j - base_w)]);
                g[i] = pf_obs_finish(P, v, s.nkey[e], j, S.noise_vec);
            }
        }
    }

    // ---- phase D: CTA partials ----
    pf_block_partials(P, S, s, is_reset, lvl, R);
}

// reset_idx over ALL envs (PF:137-139 reset(), before the zero-action step).
__global__ void __launch_bounds__(PF_THREADS) pf_reset_all_kernel(const pf_params P, const pf_env_buffers B,
                                                                  const PfCtxScratch S, const int epb,
                                                                  const uint32_t step_arg) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const PfSmem s = pf_carve(P, epb, smem_raw);
    const int tid = threadIdx.x;
    const int e0 = blockIdx.x * epb;
    const int nenv = min(epb, P.num_envs - e0);
    const uint32_t step = B.step_counter ? (uint32_t)(*B.step_counter) : step_arg;   // reset() does not count a step
    const int R = P.num_reward_terms + (P.has_termination_term ? 1 : 0);
    long long lvl = 0;
    const bool active = tid < nenv;
    if (active) {
        const int e = e0 + tid;
        PfEnvOut out;
        float root_pre[13], cmd[2];
        for (int i = 0; i < 13; ++i) root_pre[i] = B.root_states[(size_t)e * 13 + i];
        cmd[0] = B.commands[(size_t)e * P.num_commands];
        cmd[1] = B.commands[(size_t)e * P.num_commands + 1];
        pf_env_reset(P, B, e, root_pre, cmd, step, true, s.sums + tid * R, out);
        if (B.terrain_levels) lvl = B.terrain_levels[e];
    }
    __syncthreads();
    pf_block_partials(P, S, s, active, lvl, R);
}

// ------------------------------------------------------------------------------------------------ K3
// Partials -> ascending reset ids (PF:221), device count, extras["episode"] statistics (PF:276-288).
// Each CTA of 1024 threads compacts 1024 envs; its base offset is the sum of the step kernel's CTA counts
// before it.  CTA 0 additionally reduces the per-term partials (one warp per statistic, fixed order).
__global__ void __launch_bounds__(1024) pf_finalize_kernel(const pf_params P, const pf_env_buffers B,
                                                           const PfCtxScratch S, const int epb,
                                                           const int num_step_blocks, const int count_step) {
    __shared__ int s_warp[32];
    __shared__ int s_base;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int R = P.num_reward_terms + (P.has_termination_term ? 1 : 0);

    // offset of this CTA = sum of blk_count over the step CTAs in front of it
    const int before = (int)(((long long)blockIdx.x * 1024) / epb);
    int part = 0;
    for (int i = tid; i < before; i += 1024) part += S.blk_count[i];
    part = __reduce_add_sync(0xffffffffu, part);
    if (lane == 0) s_warp[warp] = part;
    __syncthreads();
    if (tid == 0) {
        int t = 0;
        for (int w = 0; w < 32; ++w) t += s_warp[w];
        s_base = t;
    }
    __syncthreads();
    const int base = s_base;
    __syncthreads();

    // total number of resets (every CTA needs it for extras["time_outs"])
    __shared__ int s_total;
    {
        int tot = 0;
        for (int i = tid; i < num_step_blocks; i += 1024) tot += S.blk_count[i];
        tot = __reduce_add_sync(0xffffffffu, tot);
        __syncthreads();
        if (lane == 0) s_warp[warp] = tot;
        __syncthreads();
        if (tid == 0) {
            int t = 0;
            for (int w = 0; w < 32; ++w) t += s_warp[w];
            s_total = t;
        }
        __syncthreads();
    }
    const int total = s_total;
    __syncthreads();

    const int e = blockIdx.x * 1024 + tid;
    if (B.extras_time_outs && total > 0 && e < P.num_envs) B.extras_time_outs[e] = B.time_out_buf[e];
    const bool flag = e < P.num_envs && B.reset_buf[e] != 0;
    const unsigned m = __ballot_sync(0xffffffffu, flag);
    if (lane == 0) s_warp[warp] = __popc(m);
    __syncthreads();
    int off = base;
    for (int w = 0; w < warp; ++w) off += s_warp[w];
    if (flag) B.reset_ids[off + __popc(m & ((1u << lane) - 1u))] = e;

    if (blockIdx.x == 0) {
        if (tid == 0) {
            *B.reset_count = total;
            if (B.step_counter && count_step) *B.step_counter += 1;
        }
        if (total > 0 && warp < R + 1) {          // PF:259: extras rebuilt only when >= 1 env resets
            if (warp < R) {
                double acc = 0.0;
                for (int i = lane; i < num_step_blocks; i += 32) acc += S.blk_sums[(size_t)i * PF_MAX_REWARD_TERMS + warp];
                for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
                if (lane == 0) B.episode_stats[warp] = (float)(acc / (double)total) / P.max_episode_length_s;
            } else {
                long long acc = 0;
                for (int i = lane; i < num_step_blocks; i += 32) acc += S.blk_level[i];
                for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
                if (lane == 0) {
                    B.episode_stats[R] = (float)((double)acc / (double)P.num_envs);
                    B.episode_stats[R + 1] = (float)total;
                }
            }
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
extern "C" int pf_abi_version(void) { return PF_ABI_VERSION; }
extern "C" int pf_sizeof_params(void) { return (int)sizeof(pf_params); }
extern "C" int pf_sizeof_env_buffers(void) { return (int)sizeof(pf_env_buffers); }

static int pf_pick_epb(int num_envs, int num_sms) {
    // enough CTAs to cover the SMs a couple of times before growing the per-CTA env count
    if (num_envs <= num_sms * 2 * 32) return 32;
    if (num_envs <= num_sms * 4 * 64) return 64;
    return 128;
}

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
    c->P = P;
    PF_CUDA(cudaGetDevice(&c->device));
    PF_CUDA(cudaDeviceGetAttribute(&c->num_sms, cudaDevAttrMultiProcessorCount, c->device));
    c->envs_per_block = pf_pick_epb(P.num_envs, c->num_sms);
    c->num_blocks = (P.num_envs + c->envs_per_block - 1) / c->envs_per_block;
    c->max_blocks = c->num_blocks;
    c->smem_bytes = pf_smem_bytes(P, c->envs_per_block);
    if (c->smem_bytes > 200 * 1024) { delete c; return pf_fail(1, "pf_ctx_create: %s", "shared memory budget exceeded"); }
    PF_CUDA(cudaFuncSetAttribute(pf_post_physics_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c->smem_bytes));
    PF_CUDA(cudaFuncSetAttribute(pf_reset_all_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c->smem_bytes));

    const int nv = P.num_priv_obs > P.num_obs ? P.num_priv_obs : P.num_obs;
    float* d_noise = nullptr; float* d_pts = nullptr;
    PF_CUDA(cudaMalloc(&d_noise, sizeof(float) * nv));
    if (noise_scale_vec_host) PF_CUDA(cudaMemcpy(d_noise, noise_scale_vec_host, sizeof(float) * nv, cudaMemcpyHostToDevice));
    else PF_CUDA(cudaMemset(d_noise, 0, sizeof(float) * nv));
    PF_CUDA(cudaMalloc(&d_pts, sizeof(float) * 2 * (P.num_height_points > 0 ? P.num_height_points : 1)));
    if (P.num_height_points > 0) {
        if (!height_points_xy_host) { delete c; return pf_fail(1, "pf_ctx_create: %s", "height points missing"); }
        PF_CUDA(cudaMemcpy(d_pts, height_points_xy_host, sizeof(float) * 2 * P.num_height_points, cudaMemcpyHostToDevice));
    }
    c->S.noise_vec = d_noise;
    c->S.height_points = d_pts;
    PF_CUDA(cudaMalloc(&c->S.blk_count, sizeof(int) * c->max_blocks));
    PF_CUDA(cudaMalloc(&c->S.blk_level, sizeof(long long) * c->max_blocks));
    PF_CUDA(cudaMalloc(&c->S.blk_sums, sizeof(double) * PF_MAX_REWARD_TERMS * c->max_blocks));
    PF_CUDA(cudaMemset(c->S.blk_count, 0, sizeof(int) * c->max_blocks));
    PF_CUDA(cudaMalloc(&c->d_actions_in, sizeof(float) * (size_t)P.num_envs * P.num_dof));
    *out = c;
    return 0;
}

extern "C" int pf_ctx_destroy(pf_ctx* c) {
    if (!c) return 0;
    cudaFree((void*)c->S.noise_vec); cudaFree((void*)c->S.height_points);
    cudaFree(c->S.blk_count); cudaFree(c->S.blk_level); cudaFree(c->S.blk_sums); cudaFree(c->d_actions_in);
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
    int blocks = (n + 255) / 256;
    if (blocks > c->num_sms * 8) blocks = c->num_sms * 8;
    pf_pd_torques_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(c->P, *bufs, actions_in);
    PF_CUDA(cudaGetLastError());
    return 0;
}

static int pf_launch_finalize(pf_ctx* c, const pf_env_buffers* bufs, cudaStream_t st, int count_step) {
    const int blocks = (c->P.num_envs + 1023) / 1024;
    pf_finalize_kernel<<<blocks, 1024, 0, st>>>(c->P, *bufs, c->S, c->envs_per_block, c->num_blocks, count_step);
    PF_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int pf_post_physics_step(pf_ctx* c, const pf_env_buffers* bufs, int64_t common_step_counter, void* stream) {
    if (int rc = pf_check_bufs(c, bufs)) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    pf_post_physics_kernel<<<c->num_blocks, PF_THREADS, c->smem_bytes, st>>>(c->P, *bufs, c->S, c->envs_per_block,
                                                                             (uint32_t)common_step_counter);
    PF_CUDA(cudaGetLastError());
    return pf_launch_finalize(c, bufs, st, 1);
}

extern "C" int pf_reset_all(pf_ctx* c, const pf_env_buffers* bufs, int64_t common_step_counter, void* stream) {
    if (int rc = pf_check_bufs(c, bufs)) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    PF_CUDA(cudaMemsetAsync(bufs->reset_buf, 1, (size_t)c->P.num_envs, st));
    pf_reset_all_kernel<<<c->num_blocks, PF_THREADS, c->smem_bytes, st>>>(c->P, *bufs, c->S, c->envs_per_block,
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
    pf_push_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(c->P, *bufs, forces, max_push_force,
                                                             (uint32_t)common_step_counter);
    PF_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int pf_step_host(pf_ctx* c, const pf_env_buffers* bufs, const pf_host_state* hs, const float* actions_host,
                            float* obs_host, float* priv_obs_host, float* rew_host, uint8_t* reset_host,
                            float* torques_host, int64_t common_step_counter, void* stream) {
    if (int rc = pf_check_bufs(c, bufs)) return rc;
    if (!actions_host || !obs_host || !rew_host || !reset_host) return pf_fail(1, "%s", "pf_step_host: null host buffer");
    const pf_params& P = c->P;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t N = (size_t)P.num_envs;
    const cudaMemcpyKind H2D = cudaMemcpyHostToDevice, D2H = cudaMemcpyDeviceToHost;
    if (hs) {
        if (hs->root_states) PF_CUDA(cudaMemcpyAsync(bufs->root_states, hs->root_states, 4 * N * 13, H2D, st));
        if (hs->dof_state) PF_CUDA(cudaMemcpyAsync(bufs->dof_state, hs->dof_state, 4 * N * P.num_dof * 2, H2D, st));
        if (hs->contact_forces)
            PF_CUDA(cudaMemcpyAsync((void*)bufs->contact_forces, hs->contact_forces, 4 * N * P.num_bodies * 3, H2D, st));
        if (hs->rigid_body_states)
            PF_CUDA(cudaMemcpyAsync((void*)bufs->rigid_body_states, hs->rigid_body_states, 4 * N * P.num_bodies * 13, H2D, st));
    }
    PF_CUDA(cudaMemcpyAsync(c->d_actions_in, actions_host, sizeof(float) * N * P.num_dof, H2D, st));
    for (int i = 0; i < P.decimation; ++i)
        if (int rc = pf_pd_torques(c, bufs, i == 0 ? (const float*)c->d_actions_in : nullptr, stream)) return rc;
    if (int rc = pf_post_physics_step(c, bufs, common_step_counter, stream)) return rc;
    PF_CUDA(cudaMemcpyAsync(obs_host, bufs->obs_buf, sizeof(float) * N * P.num_obs, D2H, st));
    if (priv_obs_host && bufs->priv_obs_buf)
        PF_CUDA(cudaMemcpyAsync(priv_obs_host, bufs->priv_obs_buf, sizeof(float) * N * P.num_priv_obs, D2H, st));
    PF_CUDA(cudaMemcpyAsync(rew_host, bufs->rew_buf, sizeof(float) * N, D2H, st));
    PF_CUDA(cudaMemcpyAsync(reset_host, bufs->reset_buf, N, D2H, st));
    if (torques_host) PF_CUDA(cudaMemcpyAsync(torques_host, bufs->torques, sizeof(float) * N * P.num_dof, D2H, st));
    PF_CUDA(cudaStreamSynchronize(st));
    return 0;
}
