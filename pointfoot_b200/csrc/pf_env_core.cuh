// Per-environment logic of PointFoot.step(), written once as host/device inline functions.
// The CUDA kernels (pf_env.cu) call it with rows staged in shared memory; tests/host_emu compiles the same
// header with g++ and calls it in a serial loop, so the logic can be checked without a GPU.  That emulation
// is test infrastructure: it is not linked into libpointfoot_b200.so.
//
// Floating point: the reference runs torch CPU eager (PF:41,50 hard-code 'cpu'), i.e. one rounding per op and
// no contraction, EXCEPT where ATen itself fuses: norm over 2/3 elements is an fma chain and cross() is
// fma(a1,b2,-(a2*b1)) (measured).  This file is compiled with -fmad=false / -ffp-contract=off
// and spells those fmas out, so integer results that depend on fp32 thresholds (termination, contacts,
// terrain indices, curriculum levels) come out bit-identical.
#pragma once
#include <stdint.h>
#include <math.h>
#include "../../include/pointfoot_b200.h"
#include "pf_rng.cuh"

#ifdef __CUDACC__
#define PF_HD __host__ __device__ __forceinline__
#else
#define PF_HD static inline
#endif

PF_HD float pf_minf(float a, float b) { return a < b ? a : b; }
PF_HD float pf_maxf(float a, float b) { return a > b ? a : b; }
PF_HD float pf_clipf(float x, float lo, float hi) { return pf_minf(pf_maxf(x, lo), hi); }
PF_HD float pf_sq(float x) { return x * x; }

// ATen CPU norm(p=2) over the last dim: 2 and 3 elements = fma chain, 4 elements = plain sequential sum.
PF_HD float pf_norm2(float x, float y) { return sqrtf(fmaf(y, y, x * x)); }
PF_HD float pf_norm3(float x, float y, float z) { return sqrtf(fmaf(z, z, fmaf(y, y, x * x))); }

// torch.cross, ATen CPU: out[i] = fma(a[i+1], b[i+2], -(a[i+2]*b[i+1]))
PF_HD void pf_cross(const float a[3], const float b[3], float o[3]) {
    o[0] = fmaf(a[1], b[2], -(a[2] * b[1]));
    o[1] = fmaf(a[2], b[0], -(a[0] * b[2]));
    o[2] = fmaf(a[0], b[1], -(a[1] * b[0]));
}

// isaacgym.torch_utils.quat_rotate_inverse / quat_rotate; q is xyzw.
PF_HD void pf_quat_rotate(const float q[4], const float v[3], float o[3], bool inverse) {
    const float qw = q[3];
    const float s = 2.0f * (qw * qw) - 1.0f;
    float c[3];
    pf_cross(q, v, c);
    const float d = (q[0] * v[0] + q[1] * v[1]) + q[2] * v[2];     // bmm(q_vec, v)
    for (int i = 0; i < 3; ++i) {
        const float a = v[i] * s;
        const float b = (c[i] * qw) * 2.0f;
        const float cc = (q[i] * d) * 2.0f;
        o[i] = inverse ? (a - b) + cc : (a + b) + cc;
    }
}

// isaacgym.torch_utils.quat_apply
PF_HD void pf_quat_apply(const float q[4], const float b[3], float o[3]) {
    float t[3], u[3];
    pf_cross(q, b, t);
    t[0] *= 2.0f; t[1] *= 2.0f; t[2] *= 2.0f;
    pf_cross(q, t, u);
    for (int i = 0; i < 3; ++i) o[i] = (b[i] + q[3] * t[i]) + u[i];
}

// (x / horizontal_scale) of PF:1102.  CPU eager: true divide.  CUDA eager: x * (1.0f / scale).
PF_HD float pf_index_div(float x, float scale, int mode) {
    return mode == PF_DIV_TRUE ? x / scale : x * (1.0f / scale);
}

// _get_terrain_heights_from_points, PF:1100-1112: min of 3 neighbouring int16 samples (NOT bilinear).
PF_HD int16_t pf_terrain_min3(const pf_params& P, const int16_t* __restrict__ hs, float x, float y) {
    x = x + P.border_size;
    y = y + P.border_size;
    x = pf_index_div(x, P.horizontal_scale, P.index_div_mode);
    y = pf_index_div(y, P.horizontal_scale, P.index_div_mode);
    long long px = (long long)x;       // .long(): truncation toward zero
    long long py = (long long)y;
    const long long mx = P.hf_rows - 2, my = P.hf_cols - 2;
    px = px < 0 ? 0 : (px > mx ? mx : px);
    py = py < 0 ? 0 : (py > my ? my : py);
    const int16_t* r = hs + px * (long long)P.hf_cols + py;
#ifdef __CUDA_ARCH__
    const int16_t h1 = __ldg(r), h2 = __ldg(r + P.hf_cols), h3 = __ldg(r + 1);   // read-only path: may move across stores
#else
    const int16_t h1 = r[0], h2 = r[P.hf_cols], h3 = r[1];
#endif
    const int16_t m = h1 < h2 ? h1 : h2;
    return m < h3 ? m : h3;
}

// Normalised yaw-only quaternion of quat_apply_yaw (math.py:38-42): returns (z, w) of normalize((0,0,qz,qw)).
PF_HD void pf_yaw_quat(const float q[4], float& zn, float& wn) {
    float n = sqrtf(q[2] * q[2] + q[3] * q[3]);     // norm over 4 elements: plain sum; the two zeros are exact
    n = pf_maxf(n, 1e-9f);
    zn = q[2] / n;
    wn = q[3] / n;
}

// One height-scan sample, PF:1070-1075: rotate grid point (gx,gy,0) by the yaw quaternion, add base xyz,
// look the terrain up.  Returns the raw int16 sample.
PF_HD int16_t pf_scan_point(const pf_params& P, const int16_t* __restrict__ hs, float zn, float wn,
                            float gx, float gy, float base_x, float base_y) {
    // quat_apply((0,0,zn,wn), (gx,gy,0)) with the cross() formula above; zero products dropped (exact)
    const float t0 = -(zn * gy) * 2.0f;
    const float t1 = (zn * gx) * 2.0f;
    const float ox = (gx + wn * t0) + (-(zn * t1));
    const float oy = (gy + wn * t1) + (zn * t0);
    return pf_terrain_min3(P, hs, ox + base_x, oy + base_y);
}

// wrap_to_pi, math.py:45-48 (torch remainder = fmod with the divisor's sign)
PF_HD float pf_wrap_to_pi(float a) {
    const float two_pi = 6.283185307179586f;
    float m = fmodf(a, two_pi);
    if (m != 0.0f && m < 0.0f) m += two_pi;
    m -= two_pi * (m > 3.141592653589793f ? 1.0f : 0.0f);
    return m;
}

// ------------------------------------------------------------------------------------------------
// Inputs of one env for one step.  Row pointers may point into shared memory (kernel) or straight into the
// tensors (host emulation).
struct PfEnvIn {
    const float* root;        // 13
    const float* dof;         // D pairs (pos, vel)
    const float* cf;          // B*3 contact forces
    const float* feet_pos[PF_NUM_FEET];   // xyz of rigid_body_states[feet_indices[f]]
    float height_mean;        // mean_p (root_z - measured_heights[p]) over the P scan points: fp32 terms summed in
                              // double, rounded once (ATen's cascade sum order depends on the host ISA; this is
                              // the closest order-free statement of torch.mean, PF:1144-1147)
    bool have_heights;        // measured_heights is a tensor (rough) rather than the int 0 (PF:775)
};

// Everything the step leaves behind for the observation stage.
struct PfEnvOut {
    float ang_vel[3], gravity[3];     // pre-reset (PF:210-211)
    float dof_pos[PF_MAX_DOF], dof_vel[PF_MAX_DOF];   // post-reset
    float root_z;                     // post-reset
    float commands[4];                // after the heading update
    bool reset;
};

// post_physics_step up to (and including) reset_idx: PF:204-222.  `e` = local env index, writes go to B.
PF_HD void pf_env_step(const pf_params& P, const pf_env_buffers& B, int e, const PfEnvIn& in,
                       uint32_t step, float* sums_out /* [R] scaled reward terms of this step or NULL */,
                       PfEnvOut& out) {
    const int D = P.num_dof;
    const float dt = P.dt;
    const int N = P.num_envs;
    // Everything the step reads from env-owned global buffers, loaded up front: the loads are independent of each
    // other, so they are all in flight together.  Left where they are used, each one sits behind an earlier store
    // the compiler must assume aliases it, and the step pays ~60 memory latencies one after another.
    const int R_all = P.num_reward_terms + (P.has_termination_term ? 1 : 0);
    float es[PF_MAX_REWARD_TERMS + 1];
    for (int k = 0; k < R_all; ++k) es[k] = B.episode_sums[(size_t)k * N + e];
    float tq_[PF_MAX_DOF], ldv_[PF_MAX_DOF], la_[PF_MAX_DOF], ac_[PF_MAX_DOF];
    for (int d = 0; d < D; ++d) {
        tq_[d] = B.torques[e * D + d]; ldv_[d] = B.last_dof_vel[e * D + d];
        la_[d] = B.last_actions[e * D + d]; ac_[d] = B.actions[e * D + d];
    }
    bool fc_old_[PF_NUM_FEET], cfilt_old_[PF_NUM_FEET], lc_old_[PF_NUM_FEET];
    float fat_old_[PF_NUM_FEET], lfat_old_[PF_NUM_FEET], lmfh_old_[PF_NUM_FEET], cmfh_old_[PF_NUM_FEET];
    for (int f = 0; f < PF_NUM_FEET; ++f) {
        const int k = e * PF_NUM_FEET + f;
        fc_old_[f] = B.first_contact[k] != 0; cfilt_old_[f] = B.contact_filt[k] != 0; lc_old_[f] = B.last_contacts[k] != 0;
        fat_old_[f] = B.feet_air_time[k]; lfat_old_[f] = B.last_feet_air_time[k];
        lmfh_old_[f] = B.last_max_feet_height[k];
        cmfh_old_[f] = P.include_feet_height_rewards ? B.current_max_feet_height[k] : 0.0f;
    }
    float cmd_old_[4];
    for (int i = 0; i < 4; ++i) cmd_old_[i] = i < P.num_commands ? B.commands[e * P.num_commands + i] : 0.0f;
    // ---- PF:204 ----
    const long long ep_len = B.episode_length_buf[e] + 1;

    // ---- PF:208-211 base quantities ----
    const float* q = in.root + 3;
    float lin_vel[3];
    const float grav[3] = {0.0f, 0.0f, -1.0f};
    pf_quat_rotate(q, in.root + 7, lin_vel, true);
    pf_quat_rotate(q, in.root + 10, out.ang_vel, true);
    pf_quat_rotate(q, grav, out.gravity, true);
    for (int i = 0; i < 3; ++i) {
        B.base_lin_vel[e * 3 + i] = lin_vel[i];
        B.base_ang_vel[e * 3 + i] = out.ang_vel[i];
        B.projected_gravity[e * 3 + i] = out.gravity[i];
    }

    // ---- PF:1114-1129 feet state machine ----
    float fat[PF_NUM_FEET], lfat[PF_NUM_FEET], lmfh[PF_NUM_FEET];
    bool first_contact[PF_NUM_FEET];
    for (int f = 0; f < PF_NUM_FEET; ++f) {
        const int k = e * PF_NUM_FEET + f;
        const bool fc_old = fc_old_[f];
        const bool cfilt_old = cfilt_old_[f];
        float fat_f = fat_old_[f];
        lfat[f] = fc_old ? fat_f : lfat_old_[f];
        fat_f = cfilt_old ? 0.0f : fat_f;
        lmfh[f] = lmfh_old_[f];
        if (P.include_feet_height_rewards) {
            float cm = cmfh_old_[f];
            lmfh[f] = fc_old ? cm : lmfh[f];
            cm = cfilt_old ? 0.0f : cm;
            float below = 0.0f;                                   // PF:1089-1090 plane
            if (P.terrain_kind == PF_TERRAIN_HEIGHTFIELD)
                below = (float)pf_terrain_min3(P, B.height_samples, in.feet_pos[f][0], in.feet_pos[f][1])
                        * P.vertical_scale;
            const float fh = in.feet_pos[f][2] - below;
            cm = pf_maxf(cm, fh);
            B.feet_height[k] = fh;
            B.current_max_feet_height[k] = cm;
            B.last_max_feet_height[k] = lmfh[f];
        }
        const bool contact = in.cf[P.feet_indices[f] * 3 + 2] > 1.0f;
        const bool cfilt = contact || lc_old_[f];
        first_contact[f] = (fat_f > 0.0f) && cfilt;
        fat_f = fat_f + dt;
        fat[f] = fat_f;
        B.contact_filt[k] = cfilt; B.last_contacts[k] = contact; B.first_contact[k] = first_contact[f];
        B.feet_air_time[k] = fat_f; B.last_feet_air_time[k] = lfat[f];
    }

    // ---- PF:479-493 callback: commands ----
    const int C = P.num_commands;
    for (int i = 0; i < 4; ++i) out.commands[i] = cmd_old_[i];
    const uint32_t genv = (uint32_t)(P.env_id_offset + e);
    if (P.resample_commands && (ep_len % P.resample_period) == 0) {       // LR:337-351
        const uint32_t ek = pf_rng_env_key(pf_rng_key(P.seed, step, PF_RNG_RESAMPLE), genv);
        out.commands[0] = (P.cmd_range[0][1] - P.cmd_range[0][0]) * pf_rng_uniform(ek, 0) + P.cmd_range[0][0];
        out.commands[1] = (P.cmd_range[1][1] - P.cmd_range[1][0]) * pf_rng_uniform(ek, 1) + P.cmd_range[1][0];
        if (P.heading_command)
            out.commands[3] = (P.cmd_range[3][1] - P.cmd_range[3][0]) * pf_rng_uniform(ek, 3) + P.cmd_range[3][0];
        else
            out.commands[2] = (P.cmd_range[2][1] - P.cmd_range[2][0]) * pf_rng_uniform(ek, 2) + P.cmd_range[2][0];
        const float keep = pf_norm2(out.commands[0], out.commands[1]) > 0.2f ? 1.0f : 0.0f;
        out.commands[0] *= keep; out.commands[1] *= keep;
    }
    if (P.heading_command) {                                              // PF:487-490
        const float fwd_b[3] = {1.0f, 0.0f, 0.0f};
        float fwd[3];
        pf_quat_apply(q, fwd_b, fwd);
        const float heading = atan2f(fwd[1], fwd[0]);
        out.commands[2] = pf_clipf(0.5f * pf_wrap_to_pi(out.commands[3] - heading), -1.0f, 1.0f);
    }
    for (int i = 0; i < C; ++i) B.commands[e * C + i] = out.commands[i];

    // ---- PF:239-247 termination ----
    bool terminate = false;
    for (int i = 0; i < P.num_termination; ++i) {
        const float* F = in.cf + P.termination_indices[i] * 3;
        terminate = terminate || (pf_norm3(F[0], F[1], F[2]) > 1.0f);
    }
    const bool time_out = (float)ep_len > P.max_episode_length;
    const bool reset = terminate || time_out;
    B.terminate[e] = terminate; B.time_out_buf[e] = time_out; B.reset_buf[e] = reset;
    out.reset = reset;

    // ---- PF:301-318 rewards, summed in cfg (alphabetical) order ----
    float rew = 0.0f;
    for (int k = 0; k < P.num_reward_terms; ++k) {
        float r = 0.0f;
        switch (P.reward_ids[k]) {
        case PF_REW_LIN_VEL_Z: r = pf_sq(lin_vel[2]); break;
        case PF_REW_ANG_VEL_XY: r = pf_sq(out.ang_vel[0]) + pf_sq(out.ang_vel[1]); break;
        case PF_REW_ORIENTATION: r = pf_sq(out.gravity[0]) + pf_sq(out.gravity[1]); break;
        case PF_REW_BASE_HEIGHT: {
            const float bh = in.have_heights ? in.height_mean : in.root[2];
            r = pf_sq(bh - P.base_height_target);
        } break;
        case PF_REW_TORQUES: for (int d = 0; d < D; ++d) r += pf_sq(tq_[d]); break;
        case PF_REW_DOF_VEL: for (int d = 0; d < D; ++d) r += pf_sq(in.dof[2 * d + 1]); break;
        case PF_REW_DOF_ACC:
            for (int d = 0; d < D; ++d) r += pf_sq((ldv_[d] - in.dof[2 * d + 1]) / dt);
            break;
        case PF_REW_ACTION_RATE:
            for (int d = 0; d < D; ++d) r += pf_sq(la_[d] - ac_[d]);
            break;
        case PF_REW_COLLISION:
            for (int i = 0; i < P.num_penalised; ++i) {
                const float* F = in.cf + P.penalised_indices[i] * 3;
                r += (pf_norm3(F[0], F[1], F[2]) > 0.1f) ? 1.0f : 0.0f;
            }
            break;
        case PF_REW_TERMINATION: r = (reset && !time_out) ? 1.0f : 0.0f; break;
        case PF_REW_DOF_POS_LIMITS:
            for (int d = 0; d < D; ++d) {
                const float qd = in.dof[2 * d];
                r += -pf_minf(qd - P.dof_pos_limits[d][0], 0.0f) + pf_maxf(qd - P.dof_pos_limits[d][1], 0.0f);
            }
            break;
        case PF_REW_DOF_VEL_LIMITS:
            for (int d = 0; d < D; ++d)
                r += pf_clipf(fabsf(in.dof[2 * d + 1]) - P.dof_vel_limits[d] * P.soft_dof_vel_limit, 0.0f, 1.0f);
            break;
        case PF_REW_TORQUE_LIMITS:
            for (int d = 0; d < D; ++d)
                r += pf_maxf(fabsf(tq_[d]) - P.torque_limits[d] * P.soft_torque_limit, 0.0f);
            break;
        case PF_REW_TRACKING_LIN_VEL: {
            const float err = pf_sq(out.commands[0] - lin_vel[0]) + pf_sq(out.commands[1] - lin_vel[1]);
            r = expf(-err / P.tracking_sigma);
        } break;
        case PF_REW_TRACKING_ANG_VEL:
            r = expf(-pf_sq(out.commands[2] - out.ang_vel[2]) / P.tracking_sigma);
            break;
        case PF_REW_FEET_AIR_TIME: {
            float lo = 0.0f, hi = 0.0f;
            for (int f = 0; f < PF_NUM_FEET; ++f) {
                const float m = first_contact[f] ? 1.0f : 0.0f;
                lo += pf_minf(fat[f] - P.min_feet_air_time, 0.0f) * m;
                hi += pf_minf(P.max_feet_air_time - fat[f], 0.0f) * m;
            }
            r = lo + hi;
        } break;
        case PF_REW_NO_FLY: {
            int n = 0;
            for (int f = 0; f < PF_NUM_FEET; ++f) n += in.cf[P.feet_indices[f] * 3 + 2] > 0.1f;
            r = n == 1 ? 1.0f : 0.0f;
        } break;
        case PF_REW_UNBALANCE_FEET_AIR_TIME: {     // torch.var over 2 values, unbiased
            const float m = (lfat[0] + lfat[1]) / 2.0f;
            r = pf_sq(lfat[0] - m) + pf_sq(lfat[1] - m);
        } break;
        case PF_REW_UNBALANCE_FEET_HEIGHT: {
            const float m = (lmfh[0] + lmfh[1]) / 2.0f;
            r = pf_sq(lmfh[0] - m) + pf_sq(lmfh[1] - m);
        } break;
        case PF_REW_STUMBLE: {
            bool any = false;
            for (int f = 0; f < PF_NUM_FEET; ++f) {
                const float* F = in.cf + P.feet_indices[f] * 3;
                any = any || (pf_norm2(F[0], F[1]) > 5.0f * fabsf(F[2]));
            }
            r = any ? 1.0f : 0.0f;
        } break;
        case PF_REW_STAND_STILL:                   // signed comparison, PF:1233-1234
            r = (fabsf(lin_vel[0]) * (out.commands[0] < 0.1f ? 1.0f : 0.0f)
                 + fabsf(lin_vel[1]) * (out.commands[1] < 0.1f ? 1.0f : 0.0f))
                + fabsf(out.ang_vel[2]) * (out.commands[2] < 0.1f ? 1.0f : 0.0f);
            break;
        case PF_REW_FEET_CONTACT_FORCES:
            for (int f = 0; f < PF_NUM_FEET; ++f) {
                const float* F = in.cf + P.feet_indices[f] * 3;
                r += pf_maxf(pf_norm3(F[0], F[1], F[2]) - P.max_contact_force, 0.0f);
            }
            break;
        case PF_REW_FEET_DISTANCE: {
            const float d = pf_norm2(in.feet_pos[0][0] - in.feet_pos[1][0], in.feet_pos[0][1] - in.feet_pos[1][1]);
            r = pf_clipf(P.min_feet_distance - d, 0.0f, 1.0f);
        } break;
        case PF_REW_SURVIVAL: r = (reset ? 0.0f : 1.0f) * dt; break;
        default: break;
        }
        r = r * P.reward_scales[k];
        rew += r;
        const float s = es[k] + r;
        B.episode_sums[(size_t)k * N + e] = s;
        if (sums_out) sums_out[k] = s;
    }
    if (P.only_positive_rewards) rew = pf_maxf(rew, 0.0f);
    if (P.has_termination_term) {
        const int k = P.num_reward_terms;
        const float r = ((reset && !time_out) ? 1.0f : 0.0f) * P.termination_scale;
        rew += r;
        const float s = es[k] + r;
        B.episode_sums[(size_t)k * N + e] = s;
        if (sums_out) sums_out[k] = s;
    }
    B.rew_buf[e] = rew;
    B.episode_length_buf[e] = ep_len;

    for (int d = 0; d < D; ++d) { out.dof_pos[d] = in.dof[2 * d]; out.dof_vel[d] = in.dof[2 * d + 1]; }
    out.root_z = in.root[2];
}

// ------------------------------------------------------------------------------------------------
// reset_idx for ONE env (PF:249-299, 549-588, 613-636).  Writes the reset rows in place into root_states /
// dof_state (PhysX-owned), updates the curriculum, zeroes the per-env buffers and the episode sums.
// `sums_out[R]` receives the episode sums before they are zeroed (for extras["episode"], PF:276-280).
// `root_pre` = this env's root row BEFORE the reset (the curriculum measures the distance walked).
PF_HD void pf_env_reset(const pf_params& P, const pf_env_buffers& B, int e, const float* root_pre,
                        const float cmd_xy[2], uint32_t step, bool update_curriculum,
                        float* sums_out, PfEnvOut& out) {
    const int D = P.num_dof, N = P.num_envs;
    const uint32_t genv = (uint32_t)(P.env_id_offset + e);

    // ---- PF:613-636 terrain curriculum ----
    if (P.terrain_curriculum && update_curriculum) {
        const float dist = pf_norm2(root_pre[0] - B.env_origins[e * 3 + 0], root_pre[1] - B.env_origins[e * 3 + 1]);
        const bool up = dist > P.terrain_env_length / 2.0f;
        const bool down = (dist < (pf_norm2(cmd_xy[0], cmd_xy[1]) * P.max_episode_length_s) * 0.5f) && !up;
        long long lvl = B.terrain_levels[e] + (up ? 1 : 0) - (down ? 1 : 0);
        if (lvl >= P.max_terrain_level) {
            const uint32_t ek = pf_rng_env_key(pf_rng_key(P.seed, step, PF_RNG_CURRICULUM), genv);
            lvl = (long long)(pf_rng_u32(ek, 0) % (uint32_t)P.max_terrain_level);
        } else if (lvl < 0) {
            lvl = 0;
        }
        B.terrain_levels[e] = lvl;
        const float* o = B.terrain_origins + ((size_t)lvl * P.num_terrain_types + B.terrain_types[e]) * 3;
        for (int i = 0; i < 3; ++i) B.env_origins[e * 3 + i] = o[i];
    }

    // ---- PF:549-559 dofs: pos = default * U(0.5, 1.5), vel = 0 ----
    {
        const uint32_t ek = pf_rng_env_key(pf_rng_key(P.seed, step, PF_RNG_RESET_DOF), genv);
        for (int d = 0; d < D; ++d) {
            const float u = (1.5f - 0.5f) * pf_rng_uniform(ek, d) + 0.5f;
            const float p = P.default_dof_pos[d] * u;
            B.dof_state[((size_t)e * D + d) * 2 + 0] = p;
            B.dof_state[((size_t)e * D + d) * 2 + 1] = 0.0f;
            out.dof_pos[d] = p; out.dof_vel[d] = 0.0f;
        }
    }
    // ---- PF:566-584 root state ----
    {
        const uint32_t ek = pf_rng_env_key(pf_rng_key(P.seed, step, PF_RNG_RESET_ROOT), genv);
        float r[13];
        for (int i = 0; i < 13; ++i) r[i] = P.base_init_state[i];
        for (int i = 0; i < 3; ++i) r[i] = r[i] + B.env_origins[e * 3 + i];
        if (P.custom_origins)
            for (int i = 0; i < 2; ++i) r[i] = r[i] + ((1.0f - (-1.0f)) * pf_rng_uniform(ek, i) + (-1.0f));
        for (int i = 0; i < 6; ++i) r[7 + i] = (0.5f - (-0.5f)) * pf_rng_uniform(ek, 2 + i) + (-0.5f);
        for (int i = 0; i < 13; ++i) B.root_states[(size_t)e * 13 + i] = r[i];
        out.root_z = r[2];
    }
    // ---- PF:290-299 buffers ----
    for (int d = 0; d < D; ++d) { B.last_actions[e * D + d] = 0.0f; B.last_dof_vel[e * D + d] = 0.0f; }
    for (int f = 0; f < PF_NUM_FEET; ++f) {
        const int k = e * PF_NUM_FEET + f;
        B.feet_air_time[k] = 0.0f; B.last_feet_air_time[k] = 0.0f;
        B.current_max_feet_height[k] = 0.0f; B.last_max_feet_height[k] = 0.0f;
    }
    B.episode_length_buf[e] = 0;
    B.reset_buf[e] = 1;
    // ---- PF:276-280 episode sums ----
    const int R = P.num_reward_terms + (P.has_termination_term ? 1 : 0);
    for (int k = 0; k < R; ++k) {
        if (sums_out) sums_out[k] = B.episode_sums[(size_t)k * N + e];
        B.episode_sums[(size_t)k * N + e] = 0.0f;
    }
}

// Number of leading (non height-scan) observation columns: PF:376-382.
PF_HD int pf_obs_base_width(const pf_params& P) { return 6 + 3 * P.num_dof + 3; }

// Column j (< base width) of the observation row before noise, PF:376-382.
PF_HD float pf_obs_base_col(const pf_params& P, const PfEnvOut& o, const float* actions, int j) {
    const int D = P.num_dof;
    if (j < 3) return o.ang_vel[j] * P.obs_scale_ang_vel;
    if (j < 6) return o.gravity[j - 3];
    j -= 6;
    if (j < D) return (o.dof_pos[j] - P.default_dof_pos[j]) * P.obs_scale_dof_pos;
    j -= D;
    if (j < D) return o.dof_vel[j] * P.obs_scale_dof_vel;
    j -= D;
    if (j < D) return actions[j];
    j -= D;
    return o.commands[j] * P.commands_scale[j];
}

// Height-scan observation column, PF:367-369: clip(root_z - 0.5 - h, -1, 1) * scale
PF_HD float pf_obs_height_col(const pf_params& P, float root_z, int16_t raw) {
    const float h = (float)raw * P.vertical_scale;
    return pf_clipf((root_z - 0.5f) - h, -1.0f, 1.0f) * P.obs_scale_height;
}

// _add_noise_to_obs (PF:328-338) for column `lane` of either buffer, then the clip of PF:189-192.
PF_HD float pf_obs_finish(const pf_params& P, float v, uint32_t noise_env_key, int lane,
                          const float* __restrict__ noise_vec) {
    if (P.add_noise) {
        const float s = noise_vec[lane];
        if (s != 0.0f) v = v + (2.0f * pf_rng_uniform(noise_env_key, lane) - 1.0f) * s;
    }
    return pf_clipf(v, -P.clip_obs, P.clip_obs);
}

// _compute_torques for one DOF, PF:523-547 (LR:353-375)
PF_HD float pf_pd_torque(const pf_params& P, int d, float action, float q, float qd, float last_qd) {
    const float as = action * P.action_scale;
    float t;
    if (P.control_type == PF_CONTROL_P)
        t = P.p_gains[d] * ((as + P.default_dof_pos[d]) - q) - P.d_gains[d] * qd;
    else if (P.control_type == PF_CONTROL_V)
        t = P.p_gains[d] * (as - qd) - (P.d_gains[d] * (qd - last_qd)) / P.sim_dt;
    else
        t = as;
    return pf_clipf(t, -P.torque_limits[d], P.torque_limits[d]);
}

// _push_robots for one env, PF:590-611: force on body 0 = quat_rotate(base_quat, U(-F, F)^3), z * 0.5
PF_HD void pf_push_force(const pf_params& P, int e, const float* quat, float max_push_force, uint32_t step,
                         float out[3]) {
    const uint32_t ek = pf_rng_env_key(pf_rng_key(P.seed, step, PF_RNG_PUSH), (uint32_t)(P.env_id_offset + e));
    float f[3];
    for (int i = 0; i < 3; ++i)
        f[i] = (max_push_force - (-max_push_force)) * pf_rng_uniform(ek, i) + (-max_push_force);
    pf_quat_rotate(quat, f, out, false);
    out[2] = out[2] * 0.5f;
}
