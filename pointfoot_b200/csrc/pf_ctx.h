// The opaque pf_ctx of include/pointfoot_b200.h: constants + small device scratch.  Internal header.
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>
#include "../../include/pointfoot_b200.h"
#include "pf_env_core.cuh"

// Device scratch the step kernel hands to the finalize kernel (per-CTA partials).
struct PfCtxScratch {
    int* blk_count;        // [num_slabs] resets in the warp slab (envs_per_warp consecutive envs)
    long long* blk_level;  // [num_slabs] sum of terrain levels in the slab
    double* blk_sums;      // [num_slabs][PF_MAX_REWARD_TERMS] episode sums of the slab's reset envs (valid iff count > 0)
    double* cta_sums;      // [finalize CTAs][PF_MAX_REWARD_TERMS] per-CTA partials of the episode statistics
    long long* cta_lvl;    // [finalize CTAs]
    int* ticket;           // last-CTA-done counter of the finalize kernel (left at 0)
    float* height_mean;    // [num_envs] mean_p (root_z - measured_heights[p]): pf_height_scan_kernel -> step kernel
    const float* noise_vec;       // [max(num_obs, num_priv_obs)]
    const float* height_points;   // [P][2]
    PfRewTab rew;                 // reward terms by id (row, scale): constant-bank operands of the step kernel
};

typedef void (*pf_step_kernel_t)(const pf_params, const pf_env_buffers, const PfCtxScratch, const int, const uint32_t);

struct pf_ctx {
    pf_params P;
    pf_step_kernel_t step_kernel;   // instantiation of pf_post_physics_kernel picked at pf_ctx_create
    pf_step_kernel_t generic_kernel;   // the order-agnostic instantiation: also taken when a state tensor is not 8-byte aligned
    PfCtxScratch S;
    int device;
    int num_sms;
    int envs_per_warp;     // envs owned by one warp of the step kernel (8, 16 or 32)
    int num_slabs;         // ceil(num_envs / envs_per_warp)
    int threads;           // CTA size of the step kernel (32..256: small batches use small CTAs)
    int canonical;         // statement of the step that applies (pf_step_mode): 0 generic, 1 table-driven, 2 pointfoot term set
    int num_blocks;        // CTAs of the step kernel
    int max_blocks;
    size_t smem_bytes;
    // pinned host staging for pf_step_host
    void* d_actions_in;    // [N,D] device copy of host actions
    // pf_step_host with pinned buffers: the transfers the first torque kernel does not need, and the results' way
    // out, run on io_stream beside the kernels (fork / join by events, so the pattern can be stream-captured)
    cudaStream_t io_stream;
    cudaEvent_t ev_fork, ev_in, ev_post, ev_out, ev_fin;
    int fin_pending;       // a finalize kernel is in flight on io_stream: pf_post_physics_step_join makes a stream wait for it
    int out_pending;       // a scatter is in flight on io_stream: pf_step_host_join makes a stream wait for it
};
