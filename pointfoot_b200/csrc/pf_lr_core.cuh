// Per-env statements of the LeggedRobot step (see pf_lr.cu): host + device, so that tests/host_emu can drive the SAME code
// with serial loops in a container without a GPU.  Arithmetic is spelled in the reference's evaluation order; compile with
// -fmad=false (nvcc) / -ffp-contract=off (g++).
#pragma once
#include "pf_env_core.cuh"

// _get_heights index path, LR:793-810 (min of 3 neighbouring int16 samples); same statement as pf_terrain_min3
PF_HD int16_t lr_terrain_min3(const pf_lr_params& P, const int16_t* __restrict__ hs, float x, float y) {
    x = x + P.border_size;
    y = y + P.border_size;
    x = pf_index_div(x, P.horizontal_scale, P.index_div_mode);
    y = pf_index_div(y, P.horizontal_scale, P.index_div_mode);
    long long px = (long long)x, py = (long long)y;
    const long long mx = P.hf_rows - 2, my = P.hf_cols - 2;
    px = px < 0 ? 0 : (px > mx ? mx : px);
    py = py < 0 ? 0 : (py > my ? my : py);
    const int16_t* r = hs + px * (long long)P.hf_cols + py;
#ifdef __CUDA_ARCH__
    const int16_t h1 = __ldg(r), h2 = __ldg(r + P.hf_cols), h3 = __ldg(r + 1);
#else
    const int16_t h1 = r[0], h2 = r[P.hf_cols], h3 = r[1];
#endif
    const int16_t m = h1 < h2 ? h1 : h2;
    return m < h3 ? m : h3;
}

// ------------------------------------------------------------------------------------------------ torques
PF_HD void lr_pd_torque_elem(const pf_lr_params& P, const pf_lr_buffers& B, const float* __restrict__ actions_in, int i) {
    const int d = i % P.num_dof;
    float a;
    if (actions_in) { a = pf_clipf(actions_in[i], -P.clip_actions, P.clip_actions); B.actions[i] = a; }
    else a = B.actions[i];
    struct { float x, y; } s = {B.dof_state[2 * i], B.dof_state[2 * i + 1]};
    const float as = a * P.action_scale;
    float t;
    if (P.control_type == PF_CONTROL_P) t = P.p_gains[d] * ((as + P.default_dof_pos[d]) - s.x) - P.d_gains[d] * s.y;
    else if (P.control_type == PF_CONTROL_V) t = P.p_gains[d] * (as - s.y) - (P.d_gains[d] * (s.y - B.last_dof_vel[i])) / P.sim_dt;
    else t = as;
    B.torques[i] = pf_clipf(t, -P.torque_limits[d], P.torque_limits[d]);
}

// ------------------------------------------------------------------------------------------------ height scan
// LR:777-813 with the PRE-reset pose (the callback runs before reset_idx, LR:122-128): measured_heights[N,P]
PF_HD float lr_scan_point(const pf_lr_params& P, const pf_lr_buffers& B, int e, int p, float zn, float wn, float bx, float by) {
    float h = 0.0f;                                                   // 'plane': zeros, LR:789-790
    if (P.terrain_kind == PF_TERRAIN_HEIGHTFIELD) {
        const float gx = B.height_points_xy[2 * p], gy = B.height_points_xy[2 * p + 1];
        const float t0 = -(zn * gy) * 2.0f, t1 = (zn * gx) * 2.0f;   // quat_apply_yaw, math.py:38-42 (zero products dropped: exact)
        const float ox = (gx + wn * t0) + (-(zn * t1)), oy = (gy + wn * t1) + (zn * t0);
        h = (float)lr_terrain_min3(P, B.height_samples, ox + bx, oy + by) * P.vertical_scale;
    }
    B.measured_heights[(size_t)e * P.num_height_points + p] = h;
    return h;
}

// ------------------------------------------------------------------------------------------------ rewards
struct LrRewIn {
    const float *lin_vel, *ang_vel, *gravity, *commands;       // [3], [3], [3], [4]
    const float *tq, *ldv, *la, *ac, *dof, *cf;                // rows of this env: [D], [D], [D], [D], [D][2], [B][3]
    float* fat; uint8_t* lc;                                   // feet_air_time / last_contacts rows [F] (updated by feet_air_time)
    float bh;                                                  // base height, LR:828-831
    bool reset, time_out;
};

// One _reward_* term, unscaled (LR:816-906)
PF_HD float lr_reward_term(const pf_lr_params& P, int id, const LrRewIn& c) {
    const int D = P.num_dof;
    float r = 0.0f;
    switch (id) {
    case PF_REW_LIN_VEL_Z: r = pf_sq(c.lin_vel[2]); break;
    case PF_REW_ANG_VEL_XY: r = pf_sq(c.ang_vel[0]) + pf_sq(c.ang_vel[1]); break;
    case PF_REW_ORIENTATION: r = pf_sq(c.gravity[0]) + pf_sq(c.gravity[1]); break;
    case PF_REW_BASE_HEIGHT: r = pf_sq(c.bh - P.base_height_target); break;
    case PF_REW_TORQUES: for (int d = 0; d < D; ++d) r += pf_sq(c.tq[d]); break;
    case PF_REW_DOF_VEL: for (int d = 0; d < D; ++d) r += pf_sq(c.dof[2 * d + 1]); break;
    case PF_REW_DOF_ACC: for (int d = 0; d < D; ++d) r += pf_sq((c.ldv[d] - c.dof[2 * d + 1]) / P.dt); break;
    case PF_REW_ACTION_RATE: for (int d = 0; d < D; ++d) r += pf_sq(c.la[d] - c.ac[d]); break;
    case PF_REW_COLLISION:
        for (int i = 0; i < P.num_penalised; ++i) {
            const float* F = c.cf + P.penalised_indices[i] * 3;
            r += pf_norm3(F[0], F[1], F[2]) > 0.1f ? 1.0f : 0.0f;
        }
        break;
    case PF_REW_TERMINATION: r = (c.reset && !c.time_out) ? 1.0f : 0.0f; break;
    case PF_REW_DOF_POS_LIMITS:
        for (int d = 0; d < D; ++d) {
            const float q = c.dof[2 * d];
            r += -pf_minf(q - P.dof_pos_limits[d][0], 0.0f) + pf_maxf(q - P.dof_pos_limits[d][1], 0.0f);
        }
        break;
    case PF_REW_DOF_VEL_LIMITS:
        for (int d = 0; d < D; ++d) r += pf_clipf(fabsf(c.dof[2 * d + 1]) - P.dof_vel_limits[d] * P.soft_dof_vel_limit, 0.0f, 1.0f);
        break;
    case PF_REW_TORQUE_LIMITS:
        for (int d = 0; d < D; ++d) r += pf_maxf(fabsf(c.tq[d]) - P.torque_limits[d] * P.soft_torque_limit, 0.0f);
        break;
    case PF_REW_TRACKING_LIN_VEL:
        r = expf(-(pf_sq(c.commands[0] - c.lin_vel[0]) + pf_sq(c.commands[1] - c.lin_vel[1])) / P.tracking_sigma);
        break;
    case PF_REW_TRACKING_ANG_VEL: r = expf(-pf_sq(c.commands[2] - c.ang_vel[2]) / P.tracking_sigma); break;
    case PF_REW_FEET_AIR_TIME: {                               // LR:882-893: the ONLY place the feet state is updated
        for (int f = 0; f < P.num_feet; ++f) {
            const bool contact = c.cf[P.feet_indices[f] * 3 + 2] > 1.0f;
            const bool filt = contact || c.lc[f] != 0;
            c.lc[f] = contact;
            const bool first = (c.fat[f] > 0.0f) && filt;
            const float t = c.fat[f] + P.dt;
            r += (t - 0.5f) * (first ? 1.0f : 0.0f);
            c.fat[f] = t * (filt ? 0.0f : 1.0f);
        }
        r *= pf_norm2(c.commands[0], c.commands[1]) > 0.1f ? 1.0f : 0.0f;
    } break;
    case PF_REW_STUMBLE: {
        bool any = false;
        for (int f = 0; f < P.num_feet; ++f) {
            const float* F = c.cf + P.feet_indices[f] * 3;
            any = any || (pf_norm2(F[0], F[1]) > 5.0f * fabsf(F[2]));
        }
        r = any ? 1.0f : 0.0f;
    } break;
    case PF_REW_STAND_STILL:                                   // LR:900-902: joint pose at (near) zero command
        for (int d = 0; d < D; ++d) r += fabsf(c.dof[2 * d] - P.default_dof_pos[d]);
        r *= pf_norm2(c.commands[0], c.commands[1]) < 0.1f ? 1.0f : 0.0f;
        break;
    case PF_REW_FEET_CONTACT_FORCES:
        for (int f = 0; f < P.num_feet; ++f) {
            const float* F = c.cf + P.feet_indices[f] * 3;
            r += pf_maxf(pf_norm3(F[0], F[1], F[2]) - P.max_contact_force, 0.0f);
        }
        break;
    default: break;
    }
    return r;
}

PF_HD void lr_resample(const pf_lr_params& P, uint32_t ek, float* cmd) {       // LR:337-351
    cmd[0] = (P.cmd_range[0][1] - P.cmd_range[0][0]) * pf_rng_uniform(ek, 0) + P.cmd_range[0][0];
    cmd[1] = (P.cmd_range[1][1] - P.cmd_range[1][0]) * pf_rng_uniform(ek, 1) + P.cmd_range[1][0];
    if (P.heading_command) cmd[3] = (P.cmd_range[3][1] - P.cmd_range[3][0]) * pf_rng_uniform(ek, 2) + P.cmd_range[3][0];
    else cmd[2] = (P.cmd_range[2][1] - P.cmd_range[2][0]) * pf_rng_uniform(ek, 2) + P.cmd_range[2][0];
    const float keep = pf_norm2(cmd[0], cmd[1]) > 0.2f ? 1.0f : 0.0f;
    cmd[0] *= keep; cmd[1] *= keep;
}

// LR:145-173 reset_idx for one env (the episode sums are read and zeroed by pf_lr_finalize_kernel); root_pre = the pose before
// the reset, cmd = this env's command row (in registers)
PF_HD void lr_reset_env(const pf_lr_params& P, const pf_lr_buffers& B, int e, uint32_t genv, uint32_t step,
                                             const float* root_pre, float* cmd) {
    const int D = P.num_dof;
    float* root = B.root_states + (size_t)e * 13;
    float* dof = B.dof_state + (size_t)e * D * 2;
    if (P.terrain_curriculum) {                        // LR:421-441
        const float dist = pf_norm2(root_pre[0] - B.env_origins[e * 3 + 0], root_pre[1] - B.env_origins[e * 3 + 1]);
        const bool up = dist > P.terrain_env_length / 2.0f;
        const bool down = (dist < (pf_norm2(cmd[0], cmd[1]) * P.max_episode_length_s) * 0.5f) && !up;
        long long lvl = B.terrain_levels[e] + (up ? 1 : 0) - (down ? 1 : 0);
        if (lvl >= P.max_terrain_level)
            lvl = (long long)(pf_rng_u32(pf_rng_env_key(pf_rng_key(P.seed, step, PF_RNG_CURRICULUM), genv), 0) % (uint32_t)P.max_terrain_level);
        else if (lvl < 0) lvl = 0;
        B.terrain_levels[e] = lvl;
        const float* o = B.terrain_origins + ((size_t)lvl * P.num_terrain_types + B.terrain_types[e]) * 3;
        for (int i = 0; i < 3; ++i) B.env_origins[e * 3 + i] = o[i];
    }
    {                                                  // LR:377-390
        const uint32_t ek = pf_rng_env_key(pf_rng_key(P.seed, step, PF_RNG_RESET_DOF), genv);
        for (int d = 0; d < D; ++d) {
            dof[2 * d] = P.default_dof_pos[d] * ((1.5f - 0.5f) * pf_rng_uniform(ek, d) + 0.5f);
            dof[2 * d + 1] = 0.0f;
        }
    }
    {                                                  // LR:392-412
        const uint32_t ek = pf_rng_env_key(pf_rng_key(P.seed, step, PF_RNG_RESET_ROOT), genv);
        float r[13];
        for (int i = 0; i < 13; ++i) r[i] = P.base_init_state[i];
        for (int i = 0; i < 3; ++i) r[i] = r[i] + B.env_origins[e * 3 + i];
        if (P.custom_origins)
            for (int i = 0; i < 2; ++i) r[i] = r[i] + ((1.0f - (-1.0f)) * pf_rng_uniform(ek, i) + (-1.0f));
        for (int i = 0; i < 6; ++i) r[7 + i] = (0.5f - (-0.5f)) * pf_rng_uniform(ek, 2 + i) + (-0.5f);
        for (int i = 0; i < 13; ++i) root[i] = r[i];
    }
    lr_resample(P, pf_rng_env_key(pf_rng_key(P.seed, step, PF_RNG_RESAMPLE_RESET), genv), cmd);      // LR:166
    for (int d = 0; d < D; ++d) { B.last_actions[(size_t)e * D + d] = 0.0f; B.last_dof_vel[(size_t)e * D + d] = 0.0f; }   // LR:169-173
    for (int f = 0; f < P.num_feet; ++f) B.feet_air_time[(size_t)e * P.num_feet + f] = 0.0f;
}

// ------------------------------------------------------------------------------------------------ the step, one thread per env
PF_HD void lr_step_env(const pf_lr_params& P, const pf_lr_buffers& B, const int e, const uint32_t step, const int push) {
    const int D = P.num_dof, N = P.num_envs, C = P.num_commands;
    const uint32_t genv = (uint32_t)(P.env_id_offset + e);
    float* root = B.root_states + (size_t)e * 13;
    float* dof = B.dof_state + (size_t)e * D * 2;
    const float* cf = B.contact_forces + (size_t)e * P.num_bodies * 3;
    float root_pre[13];
#pragma unroll
    for (int i = 0; i < 13; ++i) root_pre[i] = root[i];

    // ---- LR:112-120 ----
    const long long ep_len = B.episode_length_buf[e] + 1;
    const float* q = root_pre + 3;
    float lin_vel[3], ang_vel[3], gravity[3];
    const float grav[3] = {0.0f, 0.0f, -1.0f};
    pf_quat_rotate(q, root_pre + 7, lin_vel, true);
    pf_quat_rotate(q, root_pre + 10, ang_vel, true);
    pf_quat_rotate(q, grav, gravity, true);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        B.base_lin_vel[e * 3 + i] = lin_vel[i]; B.base_ang_vel[e * 3 + i] = ang_vel[i]; B.projected_gravity[e * 3 + i] = gravity[i];
    }

    // ---- LR:320-335 callback: resample, heading, (heights: pf_lr_scan_kernel), push ----
    float cmd[4] = {0.f, 0.f, 0.f, 0.f};
    for (int i = 0; i < C && i < 4; ++i) cmd[i] = B.commands[(size_t)e * C + i];
    if ((ep_len % P.resample_period) == 0) lr_resample(P, pf_rng_env_key(pf_rng_key(P.seed, step, PF_RNG_RESAMPLE), genv), cmd);
    if (P.heading_command) {
        const float fwd_b[3] = {1.0f, 0.0f, 0.0f};
        float fwd[3];
        pf_quat_apply(q, fwd_b, fwd);
        cmd[2] = pf_clipf(0.5f * pf_wrap_to_pi(cmd[3] - atan2f(fwd[1], fwd[0])), -1.0f, 1.0f);
    }
    if (push) {                                            // LR:414-419: the root xy velocity is overwritten in place
        const uint32_t ek = pf_rng_env_key(pf_rng_key(P.seed, step, PF_RNG_PUSH), genv);
        for (int i = 0; i < 2; ++i) root[7 + i] = (P.max_push_vel_xy - (-P.max_push_vel_xy)) * pf_rng_uniform(ek, i) + (-P.max_push_vel_xy);
    }

    // ---- LR:138-143 termination ----
    bool terminate = false;
    for (int i = 0; i < P.num_termination; ++i) {
        const float* F = cf + P.termination_indices[i] * 3;
        terminate = terminate || pf_norm3(F[0], F[1], F[2]) > 1.0f;
    }
    const bool time_out = (float)ep_len > P.max_episode_length;
    const bool reset = terminate || time_out;
    B.time_out_buf[e] = time_out;
    B.reset_buf[e] = reset;

    // ---- LR:190-207 rewards, summed in cfg (dir()) order ----
    LrRewIn c;
    c.lin_vel = lin_vel; c.ang_vel = ang_vel; c.gravity = gravity; c.commands = cmd;
    c.tq = B.torques + (size_t)e * D; c.ldv = B.last_dof_vel + (size_t)e * D; c.la = B.last_actions + (size_t)e * D;
    c.ac = B.actions + (size_t)e * D; c.dof = dof; c.cf = cf;
    c.fat = B.feet_air_time + (size_t)e * P.num_feet; c.lc = B.last_contacts + (size_t)e * P.num_feet;
    c.reset = reset; c.time_out = time_out;
    c.bh = root_pre[2];                                    // measured_heights is the int 0 when nothing measures (LR:521)
    if (P.measure_heights) {                               // mean over the scan points, fp32 terms summed in double (cf. PfEnvIn::height_mean)
        double acc = 0.0;
        const float* mh = B.measured_heights + (size_t)e * P.num_height_points;
        for (int p = 0; p < P.num_height_points; ++p) acc += (double)(root_pre[2] - mh[p]);
        c.bh = (float)(acc / (double)P.num_height_points);
    }
    float rew = 0.0f;
    for (int k = 0; k < P.num_reward_terms; ++k) {
        const float r = lr_reward_term(P, P.reward_ids[k], c) * P.reward_scales[k];
        rew += r;
        B.episode_sums[(size_t)k * N + e] += r;
    }
    if (P.only_positive_rewards) rew = pf_maxf(rew, 0.0f);
    if (P.has_termination_term) {
        const float r = ((reset && !time_out) ? 1.0f : 0.0f) * P.termination_scale;
        rew += r;
        B.episode_sums[(size_t)P.num_reward_terms * N + e] += r;
    }
    B.rew_buf[e] = rew;

    // ---- LR:145-188 reset_idx (the episode sums are read and zeroed by pf_lr_finalize_kernel) ----
    long long ep_out = ep_len;
    if (reset) { lr_reset_env(P, B, e, genv, step, root_pre, cmd); ep_out = 0; }
    B.episode_length_buf[e] = ep_out;
    for (int i = 0; i < C && i < 4; ++i) B.commands[(size_t)e * C + i] = cmd[i];

    // ---- LR:209-226 the 48 leading observation columns (+ noise LR:225-226, clip LR:99-100) ----
    const uint32_t nkey = pf_rng_env_key(pf_rng_key(P.seed, step, PF_RNG_OBS_NOISE), genv);
    float* obs = B.obs_buf + (size_t)e * P.num_obs;
    int j = 0;
    auto put = [&](float v) {
        if (P.add_noise) {
            const float s = B.noise_scale_vec[j];
            if (s != 0.0f) v = v + (2.0f * pf_rng_uniform(nkey, (uint32_t)j) - 1.0f) * s;
        }
        obs[j++] = pf_clipf(v, -P.clip_obs, P.clip_obs);
    };
    for (int i = 0; i < 3; ++i) put(lin_vel[i] * P.obs_scale_lin_vel);
    for (int i = 0; i < 3; ++i) put(ang_vel[i] * P.obs_scale_ang_vel);
    for (int i = 0; i < 3; ++i) put(gravity[i]);
    for (int i = 0; i < 3; ++i) put(cmd[i] * P.commands_scale[i]);
    for (int d = 0; d < D; ++d) put((dof[2 * d] - P.default_dof_pos[d]) * P.obs_scale_dof_pos);
    for (int d = 0; d < D; ++d) put(dof[2 * d + 1] * P.obs_scale_dof_vel);
    for (int d = 0; d < D; ++d) put(B.actions[(size_t)e * D + d]);

    // ---- LR:130-132 tail copies (post-reset, post-push state) ----
    for (int d = 0; d < D; ++d) {
        B.last_actions[(size_t)e * D + d] = B.actions[(size_t)e * D + d];
        B.last_dof_vel[(size_t)e * D + d] = dof[2 * d + 1];
    }
    for (int i = 0; i < 6; ++i) B.last_root_vel[(size_t)e * 6 + i] = root[7 + i];
}

// base_task.py:105-109 reset(): reset_idx(all envs) (the zero-action step that follows is the host's)
PF_HD void lr_reset_all_env(const pf_lr_params& P, const pf_lr_buffers& B, const int e, const uint32_t step) {
    const int C = P.num_commands;
    float root_pre[13], cmd[4] = {0.f, 0.f, 0.f, 0.f};
    for (int i = 0; i < 13; ++i) root_pre[i] = B.root_states[(size_t)e * 13 + i];
    for (int i = 0; i < C && i < 4; ++i) cmd[i] = B.commands[(size_t)e * C + i];
    lr_reset_env(P, B, e, (uint32_t)(P.env_id_offset + e), step, root_pre, cmd);
    for (int i = 0; i < C && i < 4; ++i) B.commands[(size_t)e * C + i] = cmd[i];
    B.episode_length_buf[e] = 0;
    B.reset_buf[e] = 1;                                    // LR:174
}

// LR:219-224: heights = clip(root_z - 0.5 - measured_heights, -1, 1) * scale; root z AFTER reset_idx, scan from BEFORE it
// `h` = the env's measured height at point p, `root_z` its root z (post-reset for the reference's result), `nkey` the env's
// OBS_NOISE key (computed once per env by the caller)
PF_HD void lr_obs_height(const pf_lr_params& P, const pf_lr_buffers& B, const int e, const int p, const float h, const float root_z,
                         const uint32_t nkey) {
    const int base = P.num_obs - P.num_height_points;
    float v = pf_clipf((root_z - 0.5f) - h, -1.0f, 1.0f) * P.obs_scale_height;
    if (P.add_noise) {
        const float s = B.noise_scale_vec[base + p];
        if (s != 0.0f) v = v + (2.0f * pf_rng_uniform(nkey, (uint32_t)(base + p)) - 1.0f) * s;
    }
    B.obs_buf[(size_t)e * P.num_obs + base + p] = pf_clipf(v, -P.clip_obs, P.clip_obs);
}
PF_HD uint32_t lr_noise_key(const pf_lr_params& P, const int e, const uint32_t step) {
    return pf_rng_env_key(pf_rng_key(P.seed, step, PF_RNG_OBS_NOISE), (uint32_t)(P.env_id_offset + e));
}
