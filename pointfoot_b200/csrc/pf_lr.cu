// LeggedRobot base-class step: legged_gym/envs/base/legged_robot.py (LR) -- the env of the a1 /
// anymal / cassie tasks -- as four small kernels behind context-free C-ABI calls (include/pointfoot_b200.h, "LeggedRobot").
//   pf_lr_pd_torques_kernel   LR:353-375          one thread per (env, DOF)
//   pf_lr_scan_kernel         LR:777-813          one WARP per env, lanes = scan points (187 for the LeggedRobotCfg grid)
//   pf_lr_step_kernel         LR:105-136, 138-226 one thread per env: base quantities, callback, termination, rewards, reset_idx,
//                                                 the 48 leading observation columns, tail copies
//   pf_lr_obs_heights_kernel  LR:219-224          one thread per (env, point): height columns (post-reset root z, pre-reset scan)
//   pf_lr_finalize_kernel     LR:128, 175-187     ascending reset ids, count, extras["episode"] (deterministic block reductions)
// Generic in DOF / body / foot counts (<= 12 / 24 / 4) and reward-term set: this is the widening row, written for parity first;
// the pointfoot step (pf_env.cu) stays the tuned one.  Compiled with -fmad=false: thresholded results must round like the
// reference's eager ops; the few fmas ATen itself uses (2- / 3-element norms, cross) are spelled out in pf_env_core.cuh.
#include <cuda_runtime.h>
#include <stdio.h>
#include <string.h>
#include "pf_lr_core.cuh"
#include "pf_launch.h"

void pf_set_error(const char* msg);

// ------------------------------------------------------------------------------------------------ kernels
__global__ void __launch_bounds__(256) pf_lr_pd_torques_kernel(const pf_lr_params P, const pf_lr_buffers B,
                                                               const float* __restrict__ actions_in) {
    pf_grid_dep_launch();
    pf_grid_dep_wait();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < P.num_envs * P.num_dof) lr_pd_torque_elem(P, B, actions_in, i);
}

// LR:777-813 with the PRE-reset pose (the callback runs before reset_idx, LR:122-128): measured_heights[N,P]; one warp per env
__global__ void __launch_bounds__(256) pf_lr_scan_kernel(const pf_lr_params P, const pf_lr_buffers B) {
    pf_grid_dep_launch();
    pf_grid_dep_wait();
    const int lane = threadIdx.x & 31;
    const int e = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    if (e >= P.num_envs) return;
    const float* r = B.root_states + (size_t)e * 13;
    float zn, wn;
    pf_yaw_quat(r + 3, zn, wn);
    const float bx = r[0], by = r[1];
    for (int p = lane; p < P.num_height_points; p += 32) lr_scan_point(P, B, e, p, zn, wn, bx, by);
}

__global__ void __launch_bounds__(128) pf_lr_step_kernel(const pf_lr_params P, const pf_lr_buffers B, const uint32_t step, const int push) {
    pf_grid_dep_launch();
    pf_grid_dep_wait();
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e < P.num_envs) lr_step_env(P, B, e, step, push);
}

__global__ void __launch_bounds__(128) pf_lr_reset_all_kernel(const pf_lr_params P, const pf_lr_buffers B, const uint32_t step) {
    pf_grid_dep_launch();
    pf_grid_dep_wait();
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e < P.num_envs) lr_reset_all_env(P, B, e, step);
}

__global__ void __launch_bounds__(256) pf_lr_obs_heights_kernel(const pf_lr_params P, const pf_lr_buffers B, const uint32_t step) {
    pf_grid_dep_launch();
    pf_grid_dep_wait();
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < (long long)P.num_envs * P.num_height_points) lr_obs_height(P, B, i, step);
}

// ------------------------------------------------------------------------------------------------ finalize
// CTA k < R: extras["episode"]["rew_<term k>"] = mean over the reset envs of episode_sums[k] / max_episode_length_s, then those
// entries are zeroed (LR:175-178); CTA R: ascending reset ids (LR:128 nonzero()), count, terrain level mean, extras["time_outs"].
// Fixed strides and fixed trees: deterministic.
__global__ void __launch_bounds__(1024) pf_lr_finalize_kernel(const pf_lr_params P, const pf_lr_buffers B, const int R) {
    __shared__ double s_sum[32];
    __shared__ int s_cnt[32];
    __shared__ int s_base;
    pf_grid_dep_launch();
    pf_grid_dep_wait();
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, N = P.num_envs, k = blockIdx.x;
    if (k < R) {
        double acc = 0.0; int cnt = 0;
        float* row = B.episode_sums + (size_t)k * N;
        for (int e = tid; e < N; e += 1024)
            if (B.reset_buf[e]) { acc += (double)row[e]; row[e] = 0.0f; ++cnt; }
        for (int o = 16; o > 0; o >>= 1) { acc += __shfl_xor_sync(0xffffffffu, acc, o); cnt += __shfl_xor_sync(0xffffffffu, cnt, o); }
        if (lane == 0) { s_sum[warp] = acc; s_cnt[warp] = cnt; }
        __syncthreads();
        if (tid == 0) {
            double t = 0.0; int c = 0;
            for (int w = 0; w < 32; ++w) { t += s_sum[w]; c += s_cnt[w]; }
            if (c > 0) B.episode_stats[k] = (float)(t / (double)c) / P.max_episode_length_s;
        }
        return;
    }
    // ---- CTA R ----
    if (tid == 0) s_base = 0;
    double lv = 0.0;
    __syncthreads();
    for (int e0 = 0; e0 < N; e0 += 1024) {
        const int e = e0 + tid;
        const bool flag = e < N && B.reset_buf[e] != 0;
        if (e < N && B.terrain_levels) lv += (double)B.terrain_levels[e];
        const unsigned m = __ballot_sync(0xffffffffu, flag);
        if (lane == 0) s_cnt[warp] = __popc(m);
        __syncthreads();
        int off = s_base;
        for (int w = 0; w < warp; ++w) off += s_cnt[w];
        if (flag) B.reset_ids[off + __popc(m & ((1u << lane) - 1u))] = e;
        __syncthreads();
        if (tid == 0) { int t = 0; for (int w = 0; w < 32; ++w) t += s_cnt[w]; s_base += t; }
        __syncthreads();
    }
    const int total = s_base;
    for (int o = 16; o > 0; o >>= 1) lv += __shfl_xor_sync(0xffffffffu, lv, o);
    if (lane == 0) s_sum[warp] = lv;
    __syncthreads();
    if (tid == 0) {
        *B.reset_count = total;
        if (total > 0) {
            double t = 0.0;
            for (int w = 0; w < 32; ++w) t += s_sum[w];
            B.episode_stats[R] = (float)(t / (double)N);
            B.episode_stats[R + 1] = (float)total;
        }
    }
    if (total > 0 && B.extras_time_outs)
        for (int e = tid; e < N; e += 1024) B.extras_time_outs[e] = B.time_out_buf[e];
}

// ================================================================================================ C ABI
static int lr_fail(int code, const char* msg) { pf_set_error(msg); return code; }
#define LR_CUDA(call) do { if ((call) != cudaSuccess) return lr_fail(2, #call " failed"); } while (0)

extern "C" int pf_lr_sizeof_params(void) { return (int)sizeof(pf_lr_params); }
extern "C" int pf_lr_sizeof_buffers(void) { return (int)sizeof(pf_lr_buffers); }

static int lr_check(const pf_lr_params* p, const pf_lr_buffers* b, bool post) {
    if (!p || !b) return lr_fail(1, "pf_lr: null params / buffers");
    if (p->abi_version != PF_ABI_VERSION) return lr_fail(1, "pf_lr: ABI version mismatch");
    if (p->num_envs <= 0 || p->num_dof <= 0 || p->num_dof > PF_LR_MAX_DOF || p->num_bodies <= 0 || p->num_bodies > PF_LR_MAX_BODIES ||
        p->num_feet < 0 || p->num_feet > PF_LR_MAX_FEET || p->num_commands < 3 || p->num_commands > 4 || p->num_reward_terms < 0 ||
        p->num_reward_terms + (p->has_termination_term ? 1 : 0) > PF_MAX_REWARD_TERMS || p->num_height_points < 0 ||
        p->control_type < PF_CONTROL_P || p->control_type > PF_CONTROL_T)
        return lr_fail(1, "pf_lr: parameter out of range");
    if (!b->dof_state || !b->actions || !b->torques || !b->last_dof_vel) return lr_fail(1, "pf_lr_buffers: a required pointer is NULL");
    if (!post) return 0;
    // LR:209-224: the observation width must be what compute_observations builds
    const int want = 12 + 3 * p->num_dof + (p->measure_heights ? p->num_height_points : 0);
    if (p->num_obs != want) return lr_fail(3, "pf_lr: computed observation width does not match num_observations");
    if (!b->root_states || !b->contact_forces || !b->commands || !b->episode_length_buf || !b->base_lin_vel || !b->base_ang_vel ||
        !b->projected_gravity || !b->last_actions || !b->last_root_vel || !b->feet_air_time || !b->last_contacts || !b->episode_sums ||
        !b->rew_buf || !b->reset_buf || !b->time_out_buf || !b->obs_buf || !b->env_origins || !b->reset_ids || !b->reset_count ||
        !b->episode_stats || !b->noise_scale_vec)
        return lr_fail(1, "pf_lr_buffers: a required pointer is NULL");
    if (p->measure_heights && (!b->measured_heights || !b->height_points_xy)) return lr_fail(1, "pf_lr_buffers: height scan buffers are NULL");
    if (p->terrain_kind == PF_TERRAIN_HEIGHTFIELD && (!b->height_samples || !b->terrain_origins || !b->terrain_levels || !b->terrain_types))
        return lr_fail(1, "pf_lr_buffers: terrain pointers are NULL for a heightfield task");
    return 0;
}

extern "C" int pf_lr_pd_torques(const pf_lr_params* p, const pf_lr_buffers* b, const float* actions_in, void* stream) {
    if (int rc = lr_check(p, b, false)) return rc;
    const int n = p->num_envs * p->num_dof;
    LR_CUDA(pf_launch_pdl(pf_lr_pd_torques_kernel, dim3((n + 255) / 256), dim3(256), 0, (cudaStream_t)stream, *p, *b, actions_in));
    return 0;
}

extern "C" int pf_lr_reset_all(const pf_lr_params* p, const pf_lr_buffers* b, int64_t common_step_counter, void* stream) {
    if (int rc = lr_check(p, b, true)) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const int N = p->num_envs, R = p->num_reward_terms + (p->has_termination_term ? 1 : 0);
    LR_CUDA(pf_launch_pdl(pf_lr_reset_all_kernel, dim3((N + 127) / 128), dim3(128), 0, st, *p, *b, (uint32_t)common_step_counter));
    LR_CUDA(pf_launch_pdl(pf_lr_finalize_kernel, dim3(R + 1), dim3(1024), 0, st, *p, *b, R));
    return 0;
}

extern "C" int pf_lr_post_physics_step(const pf_lr_params* p, const pf_lr_buffers* b, int64_t common_step_counter, int32_t push,
                                       void* stream) {
    if (int rc = lr_check(p, b, true)) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    const uint32_t step = (uint32_t)common_step_counter;
    const int N = p->num_envs, R = p->num_reward_terms + (p->has_termination_term ? 1 : 0);
    if (p->measure_heights)
        LR_CUDA(pf_launch_pdl(pf_lr_scan_kernel, dim3((N + 7) / 8), dim3(256), 0, st, *p, *b));
    LR_CUDA(pf_launch_pdl(pf_lr_step_kernel, dim3((N + 127) / 128), dim3(128), 0, st, *p, *b, step, (int)(push != 0)));
    if (p->measure_heights) {
        const long long n = (long long)N * p->num_height_points;
        LR_CUDA(pf_launch_pdl(pf_lr_obs_heights_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, st, *p, *b, step));
    }
    LR_CUDA(pf_launch_pdl(pf_lr_finalize_kernel, dim3(R + 1), dim3(1024), 0, st, *p, *b, R));
    return 0;
}
