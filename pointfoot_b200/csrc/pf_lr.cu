// LeggedRobot base-class step: legged_gym/envs/base/legged_robot.py (LR) -- the env of the a1 /
// anymal / cassie tasks -- as four small kernels behind context-free C-ABI calls (include/pointfoot_b200.h, "LeggedRobot").
//   pf_lr_pd_torques_kernel   LR:353-375          one thread per (env, DOF)
//   pf_lr_scan_kernel         LR:777-813, 219-224 one WARP per env, lanes = scan points (187 for the LeggedRobotCfg grid)
//   pf_lr_step_kernel         LR:105-136, 138-226 one thread per env: base quantities, callback, termination, rewards, reset_idx,
//                                                 the 48 leading observation columns, tail copies
//   (height columns, LR:219-224: written by the scan kernel's lanes; pf_lr_obs_heights_fix_kernel redoes them for the envs that
//    reset -- post-reset root z against the pre-reset scan -- one warp per entry of the reset id list)
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

// LR:777-813 with the PRE-reset pose (the callback runs before reset_idx, LR:122-128): measured_heights[N,P]; one warp per env.
// The same lanes write the height observation columns (LR:219-224) with the root z they have at hand: right for every env
// that does not reset on this step; the few that do are redone by pf_lr_obs_heights_fix_kernel with the post-reset z.
__global__ void __launch_bounds__(256) pf_lr_scan_kernel(const pf_lr_params P, const pf_lr_buffers B, const uint32_t step) {
    pf_grid_dep_launch();
    pf_grid_dep_wait();
    const int lane = threadIdx.x & 31;
    const int e = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    if (e >= P.num_envs) return;
    const float* r = B.root_states + (size_t)e * 13;
    float zn, wn;
    pf_yaw_quat(r + 3, zn, wn);
    const float bx = r[0], by = r[1], rz = r[2];
    const uint32_t nkey = lr_noise_key(P, e, step);
    for (int p = lane; p < P.num_height_points; p += 32) {
        const float h = lr_scan_point(P, B, e, p, zn, wn, bx, by);
        lr_obs_height(P, B, e, p, h, rz, nkey);
    }
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

// height columns of the envs that reset on this step (LR:219-224 runs after reset_idx: post-reset root z against the pre-reset
// scan): one warp per entry of the device-side reset id list
__global__ void __launch_bounds__(256) pf_lr_obs_heights_fix_kernel(const pf_lr_params P, const pf_lr_buffers B, const uint32_t step) {
    pf_grid_dep_launch();
    pf_grid_dep_wait();
    const int lane = threadIdx.x & 31, nw = (int)((gridDim.x * blockDim.x) >> 5);
    const int cnt = min(*B.reset_count, P.num_envs);
    for (int k = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5); k < cnt; k += nw) {
        const int e = B.reset_ids[k];
        const float rz = B.root_states[(size_t)e * 13 + 2];
        const uint32_t nkey = lr_noise_key(P, e, step);
        for (int p = lane; p < P.num_height_points; p += 32)
            lr_obs_height(P, B, e, p, B.measured_heights[(size_t)e * P.num_height_points + p], rz, nkey);
    }
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
        const unsigned char* rb = reinterpret_cast<const unsigned char*>(B.reset_buf);
        // 16 reset flags per 128-bit load, four loads in flight per thread: the flags are sparse, and a byte load consumed at once
        // per 1,024 envs made this loop 256 dependent L2 round trips (256 us at 262,144 envs)
        const int n16 = (reinterpret_cast<uintptr_t>(rb) & 15) == 0 ? N >> 4 : 0;
        for (int i0 = 0; i0 < n16; i0 += 4 * 1024) {
            uint4 f[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) { const int i = i0 + u * 1024 + tid; f[u] = make_uint4(0u, 0u, 0u, 0u); if (i < n16) f[u] = *(reinterpret_cast<const uint4*>(rb) + i); }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if ((f[u].x | f[u].y | f[u].z | f[u].w) == 0u) continue;
                const uint32_t w4[4] = {f[u].x, f[u].y, f[u].z, f[u].w};
                const int e0 = (i0 + u * 1024 + tid) << 4;
#pragma unroll
                for (int q = 0; q < 16; ++q)
                    if ((w4[q >> 2] >> (8 * (q & 3))) & 0xffu) { acc += (double)row[e0 + q]; row[e0 + q] = 0.0f; ++cnt; }
            }
        }
        for (int e = (n16 << 4) + tid; e < N; e += 1024)
            if (rb[e]) { acc += (double)row[e]; row[e] = 0.0f; ++cnt; }
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
    // ---- CTA R: warp w owns the contiguous env range [w * per, (w + 1) * per) (per a multiple of 512): count, one barrier, then
    //      write -- no barrier per 1024 envs (that loop was 362 us at 262,144 envs); 16 flags per lane per 128-bit load ----
    const unsigned char* rb = reinterpret_cast<const unsigned char*>(B.reset_buf);
    const bool al16 = (reinterpret_cast<uintptr_t>(rb) & 15) == 0;
    const int per = ((N + 31) / 32 + 511) & ~511, e_beg = min(N, warp * per), e_end = min(N, e_beg + per);
    // nonzero-byte mask (16 bits) of the lane's 16 envs starting at e (zeros beyond e_end)
    auto flags16 = [&](int e) -> unsigned {
        unsigned m = 0;
        if (al16 && e + 16 <= e_end) {
            const uint4 f = *reinterpret_cast<const uint4*>(rb + e);
            const uint32_t w4[4] = {f.x, f.y, f.z, f.w};
#pragma unroll
            for (int q = 0; q < 16; ++q) m |= ((w4[q >> 2] >> (8 * (q & 3))) & 0xffu) ? (1u << q) : 0u;
        } else {
            for (int q = 0; q < 16; ++q) if (e + q < e_end && rb[e + q]) m |= 1u << q;
        }
        return m;
    };
    int cnt = 0;
    for (int e0 = e_beg; e0 < e_end; e0 += 4 * 512) {
        unsigned m[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) m[u] = flags16(e0 + u * 512 + lane * 16);
#pragma unroll
        for (int u = 0; u < 4; ++u) cnt += __popc(m[u]);
    }
    double lv = 0.0;
    if (B.terrain_levels) {                   // mean level over ALL envs (LR:180): 2 MB at 262,144 envs through one SM -- 16 values in flight per lane
        const long long* tl = reinterpret_cast<const long long*>(B.terrain_levels);
        if ((reinterpret_cast<uintptr_t>(tl) & 15) == 0 && (e_beg & 1) == 0) {
            const longlong2* t2 = reinterpret_cast<const longlong2*>(tl);
            const int p_beg = e_beg >> 1, p_end = e_end >> 1;              // pairs
            long long acc = 0;
            for (int i0 = p_beg; i0 < p_end; i0 += 8 * 32) {
                longlong2 t[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) { const int i = i0 + u * 32 + lane; t[u] = make_longlong2(0, 0); if (i < p_end) t[u] = t2[i]; }
#pragma unroll
                for (int u = 0; u < 8; ++u) acc += t[u].x + t[u].y;
            }
            if ((e_end & 1) && lane == 0) acc += tl[e_end - 1];
            lv = (double)acc;
        } else {
            long long acc = 0;
            for (int e = e_beg + lane; e < e_end; e += 32) acc += tl[e];
            lv = (double)acc;
        }
    }
    for (int o = 16; o > 0; o >>= 1) { lv += __shfl_xor_sync(0xffffffffu, lv, o); cnt += __shfl_xor_sync(0xffffffffu, cnt, o); }
    if (lane == 0) { s_cnt[warp] = cnt; s_sum[warp] = lv; }
    __syncthreads();
    int off = 0, total = 0;
    for (int w = 0; w < 32; ++w) { if (w < warp) off += s_cnt[w]; total += s_cnt[w]; }
    for (int e0 = e_beg; e0 < e_end; e0 += 4 * 512) {
        unsigned m[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) m[u] = flags16(e0 + u * 512 + lane * 16);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int c = __popc(m[u]);
            int incl = c;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
            int o2 = off + incl - c;
            const int e = e0 + u * 512 + lane * 16;
            for (unsigned mm = m[u]; mm; mm &= mm - 1) B.reset_ids[o2++] = e + __ffs(mm) - 1;
            off += __shfl_sync(0xffffffffu, incl, 31);
        }
    }
    if (tid == 0) {
        *B.reset_count = total;
        if (total > 0) {
            double t = 0.0;
            for (int w = 0; w < 32; ++w) t += s_sum[w];
            B.episode_stats[R] = (float)(t / (double)N);
            B.episode_stats[R + 1] = (float)total;
        }
    }
    if (total > 0 && B.extras_time_outs) {    // extras["time_outs"] = time_out_buf (LR:186-187): 128-bit copies, four loads in flight
        const unsigned char* src = reinterpret_cast<const unsigned char*>(B.time_out_buf);
        unsigned char* dst = reinterpret_cast<unsigned char*>(B.extras_time_outs);
        const int n16 = ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0 ? N >> 4 : 0;
        for (int i0 = 0; i0 < n16; i0 += 4 * 1024) {
            uint4 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) { const int i = i0 + u * 1024 + tid; if (i < n16) v[u] = reinterpret_cast<const uint4*>(src)[i]; }
#pragma unroll
            for (int u = 0; u < 4; ++u) { const int i = i0 + u * 1024 + tid; if (i < n16) reinterpret_cast<uint4*>(dst)[i] = v[u]; }
        }
        for (int e = (n16 << 4) + tid; e < N; e += 1024) dst[e] = src[e];
    }
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
        LR_CUDA(pf_launch_pdl(pf_lr_scan_kernel, dim3((N + 7) / 8), dim3(256), 0, st, *p, *b, step));
    LR_CUDA(pf_launch_pdl(pf_lr_step_kernel, dim3((N + 127) / 128), dim3(128), 0, st, *p, *b, step, (int)(push != 0)));
    LR_CUDA(pf_launch_pdl(pf_lr_finalize_kernel, dim3(R + 1), dim3(1024), 0, st, *p, *b, R));
    if (p->measure_heights) {                 // after the id list: only the envs that reset have their height columns redone
        int blocks = (N + 7) / 8;
        if (blocks > 148 * 4) blocks = 148 * 4;
        LR_CUDA(pf_launch_pdl(pf_lr_obs_heights_fix_kernel, dim3(blocks), dim3(256), 0, st, *p, *b, step));
    }
    return 0;
}
