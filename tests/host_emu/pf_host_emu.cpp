// Host emulation of the env kernels: the SAME pf_env_core.cuh the CUDA kernels use, driven by serial loops
// over host memory.  TEST INFRASTRUCTURE ONLY (built by tests/host_emu/build.sh into tests/host_emu/_build/,
// never linked into or loaded by the product).  It lets the per-env logic be checked against the oracle in a
// container without a GPU; the GPU parity tests then only have to catch staging / indexing mistakes.
#include "../../pointfoot_b200/csrc/pf_env_core.cuh"
#include <vector>
#include <stdlib.h>

extern "C" int pf_emu_pd_torques(const pf_params* Pp, const pf_env_buffers* Bp, const float* actions_in) {
    const pf_params& P = *Pp; const pf_env_buffers& B = *Bp;
    const int D = P.num_dof;
    for (int e = 0; e < P.num_envs; ++e)
        for (int d = 0; d < D; ++d) {
            if (actions_in) B.actions[e * D + d] = pf_clipf(actions_in[e * D + d], -P.clip_actions, P.clip_actions);
            B.torques[e * D + d] = pf_pd_torque(P, d, B.actions[e * D + d], B.dof_state[(e * D + d) * 2],
                                                B.dof_state[(e * D + d) * 2 + 1], B.last_dof_vel[e * D + d]);
        }
    return 0;
}

// _push_robots, PF:590-611 (pf_push_kernel's loop, serial)
extern "C" int pf_emu_push_forces(const pf_params* Pp, const pf_env_buffers* Bp, float* forces, float max_push_force,
                                  long long common_step_counter) {
    const pf_params& P = *Pp; const pf_env_buffers& B = *Bp;
    const int row = 3 * P.num_bodies;
    for (int e = 0; e < P.num_envs; ++e) {
        for (int c = 0; c < row; ++c) forces[(size_t)e * row + c] = 0.0f;
        pf_push_force(P, e, B.root_states + (size_t)e * 13 + 3, max_push_force, (uint32_t)common_step_counter,
                      forces + (size_t)e * row);
    }
    return 0;
}

extern "C" int pf_emu_post_physics_step(const pf_params* Pp, const pf_env_buffers* Bp, const float* noise_vec,
                                        const float* height_points_xy, long long common_step_counter) {
    const pf_params& P = *Pp; const pf_env_buffers& B = *Bp;
    const int D = P.num_dof, N = P.num_envs, Pn = P.num_height_points;
    const uint32_t step = (uint32_t)common_step_counter;
    const int R = P.num_reward_terms + (P.has_termination_term ? 1 : 0);
    const int base_w = pf_obs_base_width(P);
    // PF_EMU_GENERIC=1 / PF_EMU_TABLE=1 force the order-agnostic reward loop / the table-driven statement
    // (all three statements must agree bit for bit)
    PfRewTab T;
    const int canonical = pf_step_mode(P, T, getenv("PF_EMU_GENERIC") != nullptr, getenv("PF_EMU_TABLE") != nullptr);
    std::vector<int16_t> raw(Pn > 0 ? Pn : 1);
    std::vector<double> stat(R, 0.0);
    int count = 0;
    for (int e = 0; e < N; ++e) {
        PfEnvIn in; PfEnvOut out;
        float root_pre[13];
        for (int i = 0; i < 13; ++i) root_pre[i] = B.root_states[e * 13 + i];
        in.root = root_pre;
        in.dof = B.dof_state + (size_t)e * D * 2;
        in.cf = B.contact_forces + (size_t)e * P.num_bodies * 3;
        for (int f = 0; f < PF_NUM_FEET; ++f)
            in.feet_pos[f] = B.rigid_body_states + ((size_t)e * P.num_bodies + P.feet_indices[f]) * 13;
        double hsum = 0.0; in.height_mean = 0.0f; in.have_heights = Pn > 0;
        if (Pn > 0) {
            float zn, wn; pf_yaw_quat(root_pre + 3, zn, wn);
            for (int p = 0; p < Pn; ++p) {
                raw[p] = pf_scan_point(P, B.height_samples, zn, wn, height_points_xy[2 * p],
                                       height_points_xy[2 * p + 1], root_pre[0], root_pre[1], canonical ? &T : nullptr);
                const float h = (float)raw[p] * P.vertical_scale;
                if (B.measured_heights) B.measured_heights[(size_t)e * Pn + p] = h;
                hsum += (double)(root_pre[2] - h);
            }
            in.height_mean = (float)(hsum / (double)Pn);
        }
        pf_env_step(P, B, T, e, in, step, nullptr, out, canonical);
        if (out.reset) {
            std::vector<float> sums(R);
            pf_env_reset(P, B, e, root_pre, out.commands, step, true, sums.data(), out);
            for (int k = 0; k < R; ++k) stat[k] += sums[k];
            B.reset_ids[count++] = e;
        }
        // observations, PF:320-382 + clip PF:189-192
        const uint32_t nk = pf_rng_env_key(pf_rng_key(P.seed, step, PF_RNG_OBS_NOISE), (uint32_t)(P.env_id_offset + e));
        const float* act = B.actions + e * D;
        for (int j = 0; j < P.num_obs; ++j) {
            const float v = j < base_w ? pf_obs_base_col(P, out, act, j) : pf_obs_height_col(P, out.root_z, raw[j - base_w]);
            B.obs_buf[(size_t)e * P.num_obs + j] = pf_obs_finish(P, v, nk, j, noise_vec);
        }
        if (B.priv_obs_buf)
            for (int j = 0; j < P.num_priv_obs; ++j) {
                const float v = j < base_w ? pf_obs_base_col(P, out, act, j) : pf_obs_height_col(P, out.root_z, raw[j - base_w]);
                B.priv_obs_buf[(size_t)e * P.num_priv_obs + j] = pf_obs_finish(P, v, nk, j, noise_vec);
            }
        // PF:225-227
        for (int d = 0; d < D; ++d) { B.last_actions[e * D + d] = act[d]; B.last_dof_vel[e * D + d] = out.dof_vel[d]; }
        for (int i = 0; i < 6; ++i) B.last_root_vel[e * 6 + i] = B.root_states[e * 13 + 7 + i];
    }
    *B.reset_count = count;
    if (count > 0 && B.extras_time_outs) for (int e = 0; e < N; ++e) B.extras_time_outs[e] = B.time_out_buf[e];
    if (count > 0) {
        for (int k = 0; k < R; ++k) B.episode_stats[k] = (float)(stat[k] / count) / P.max_episode_length_s;
        double lv = 0; if (B.terrain_levels) for (int e = 0; e < N; ++e) lv += (double)B.terrain_levels[e];
        B.episode_stats[R] = (float)(lv / N);
        B.episode_stats[R + 1] = (float)count;
    }
    return 0;
}

// ------------------------------------------------------------------------------------------------ LeggedRobot (N2)
// The SAME pf_lr_core.cuh the pf_lr_* kernels use, driven serially; the finalize step is restated with serial loops
// (the kernel's fixed-stride tree only changes the rounding of the double-precision sums).
#include "../../pointfoot_b200/csrc/pf_lr_core.cuh"

extern "C" int pf_emu_lr_pd_torques(const pf_lr_params* Pp, const pf_lr_buffers* Bp, const float* actions_in) {
    for (int i = 0; i < Pp->num_envs * Pp->num_dof; ++i) lr_pd_torque_elem(*Pp, *Bp, actions_in, i);
    return 0;
}

static void lr_emu_finalize(const pf_lr_params& P, const pf_lr_buffers& B) {
    const int N = P.num_envs, R = P.num_reward_terms + (P.has_termination_term ? 1 : 0);
    int count = 0;
    for (int e = 0; e < N; ++e) if (B.reset_buf[e]) B.reset_ids[count++] = e;
    *B.reset_count = count;
    for (int k = 0; k < R; ++k) {
        double t = 0.0;
        for (int e = 0; e < N; ++e) if (B.reset_buf[e]) { t += (double)B.episode_sums[(size_t)k * N + e]; B.episode_sums[(size_t)k * N + e] = 0.0f; }
        if (count) B.episode_stats[k] = (float)(t / (double)count) / P.max_episode_length_s;
    }
    if (count) {
        double lv = 0.0;
        if (B.terrain_levels) for (int e = 0; e < N; ++e) lv += (double)B.terrain_levels[e];
        B.episode_stats[R] = (float)(lv / (double)N);
        B.episode_stats[R + 1] = (float)count;
        if (B.extras_time_outs) for (int e = 0; e < N; ++e) B.extras_time_outs[e] = B.time_out_buf[e];
    }
}

extern "C" int pf_emu_lr_post_physics_step(const pf_lr_params* Pp, const pf_lr_buffers* Bp, long long common_step_counter, int push) {
    const pf_lr_params& P = *Pp; const pf_lr_buffers& B = *Bp;
    const uint32_t step = (uint32_t)common_step_counter;
    if (P.measure_heights)
        for (int e = 0; e < P.num_envs; ++e) {
            const float* r = B.root_states + (size_t)e * 13;
            float zn, wn; pf_yaw_quat(r + 3, zn, wn);
            for (int p = 0; p < P.num_height_points; ++p) lr_scan_point(P, B, e, p, zn, wn, r[0], r[1]);
        }
    for (int e = 0; e < P.num_envs; ++e) lr_step_env(P, B, e, step, push);
    if (P.measure_heights)          // LR:219-224 after reset_idx: post-reset root z against the pre-reset scan
        for (int e = 0; e < P.num_envs; ++e) {
            const uint32_t nkey = lr_noise_key(P, e, step);
            for (int p = 0; p < P.num_height_points; ++p)
                lr_obs_height(P, B, e, p, B.measured_heights[(size_t)e * P.num_height_points + p], B.root_states[(size_t)e * 13 + 2], nkey);
        }
    lr_emu_finalize(P, B);
    return 0;
}

extern "C" int pf_emu_lr_reset_all(const pf_lr_params* Pp, const pf_lr_buffers* Bp, long long common_step_counter) {
    for (int e = 0; e < Pp->num_envs; ++e) lr_reset_all_env(*Pp, *Bp, e, (uint32_t)common_step_counter);
    lr_emu_finalize(*Pp, *Bp);
    return 0;
}

// ------------------------------------------------------------------------------------------------ terrain generation (N1)
#include "../../pointfoot_b200/csrc/pf_terrain_core.cuh"

extern "C" int pf_emu_terrain_generate(const pf_terrain_params* Pp, const pf_subterrain* subs, int16_t* hf, double* origins) {
    const pf_terrain_params& P = *Pp;
    for (int r = 0; r < P.tot_rows; ++r)
        for (int c = 0; c < P.tot_cols; ++c) hf[(size_t)r * P.tot_cols + c] = pf_terrain_map_pixel(P, subs, r, c);
    const int W = P.cell_px;
    for (int i = 0; i < P.num_rows; ++i)
        for (int j = 0; j < P.num_cols; ++j) {
            int m = -32768;
            for (int x = P.origin_patch[0]; x < P.origin_patch[1]; ++x)
                for (int y = P.origin_patch[2]; y < P.origin_patch[3]; ++y) {
                    const int v = hf[(size_t)(P.border_px + i * W + x) * P.tot_cols + (P.border_px + j * W + y)];
                    m = v > m ? v : m;
                }
            double* o = origins + (size_t)(i * P.num_cols + j) * 3;
            o[0] = ((double)i + 0.5) * P.env_length; o[1] = ((double)j + 0.5) * P.env_width; o[2] = (double)m * P.vertical_scale;
        }
    return 0;
}

extern "C" int pf_emu_terrain_trimesh(const int16_t* hf, int rows, int cols, double hs, double vs, double thr, float* vertices,
                                      uint32_t* tri) {
    for (int r = 0; r < rows; ++r)
        for (int c = 0; c < cols; ++c) pf_terrain_vertex(hf, rows, cols, hs, vs, thr, r, c, vertices + ((size_t)r * cols + c) * 3);
    for (int r = 0; r < rows - 1; ++r)
        for (int c = 0; c < cols - 1; ++c) {
            const uint32_t i0 = (uint32_t)(r * cols + c), i1 = i0 + 1, i2 = i0 + (uint32_t)cols, i3 = i2 + 1;
            uint32_t* t = tri + ((size_t)r * (cols - 1) + c) * 6;
            t[0] = i0; t[1] = i3; t[2] = i1; t[3] = i0; t[4] = i2; t[5] = i3;
        }
    return 0;
}
