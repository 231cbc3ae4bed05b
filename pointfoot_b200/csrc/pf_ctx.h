// The opaque pf_ctx of include/pointfoot_b200.h: constants + small device scratch.  Internal header.
#pragma once
#include <stdint.h>
#include "../../include/pointfoot_b200.h"

// Device scratch the step kernel hands to the finalize kernel (per-CTA partials).
struct PfCtxScratch {
    int* blk_count;        // [max_blocks] resets in the CTA
    long long* blk_level;  // [max_blocks] sum of terrain levels in the CTA
    double* blk_sums;      // [max_blocks][PF_MAX_REWARD_TERMS] episode sums of the CTA's reset envs
    const float* noise_vec;       // [max(num_obs, num_priv_obs)]
    const float* height_points;   // [P][2]
};

struct pf_ctx {
    pf_params P;
    PfCtxScratch S;
    int device;
    int num_sms;
    int envs_per_block;    // EPB of the step kernel
    int num_blocks;        // ceil(num_envs / EPB)
    int max_blocks;
    size_t smem_bytes;
    // pinned host staging for pf_step_host
    void* d_actions_in;    // [N,D] device copy of host actions
};
