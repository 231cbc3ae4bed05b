// Internal layout of pf_rl_ctx (include/pointfoot_b200.h), shared by pf_rl.cu (fp32 kernels, everything that is
// not a GEMM) and pf_mlp_tc.cu (tcgen05 tiles).
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>
#include "../../include/pointfoot_b200.h"

struct PfNet {                                  // one MLP of ActorCritic (actor_critic.py)
    int layers;
    int dims[PF_MAX_MLP_LAYERS + 1];
    long long w_off[PF_MAX_MLP_LAYERS];         // offsets into the flat fp32 parameter buffer
    long long b_off[PF_MAX_MLP_LAYERS];
    float* act[PF_MAX_MLP_LAYERS];              // saved post-ELU activations [rows_cap, dims[l+1]] (last = output)
    // tensor-core path: packed bf16 weights in UMMA tile order + fp32 biases, see pf_mlp_tc.cu
    void* tc_weights; size_t tc_weight_bytes;
    int tc_plan_id;
};

struct pf_rl_ctx {
    pf_rl_params P;
    int device, num_sms;
    long long n_params, std_off;
    PfNet net[2];                               // 0 actor, 1 critic
    int rows_cap;                               // max(num_envs, minibatch size)
    int mb_size;                                // (N*T) / num_mini_batches
    int* perm;                                  // [N*T] minibatch permutation of this iteration
    float* dbuf[2];                             // [rows_cap, max_width] backward ping-pong (dY / dH)
    float* part;                                // split-K / column-sum partials
    long long part_floats;
    int wgrad_splits;
    float* mu; float* val;                      // [rows_cap, A], [rows_cap] forward outputs
    float* dmu; float* dval;                    // loss gradients
    double* loss_part;                          // [loss_blocks][8 + A] per-CTA partials of the loss kernel
    int loss_blocks;
    double* gae_part;                           // [gae_blocks][2]
    int gae_blocks;
    double* norm_part;                          // [norm_blocks] grad-norm^2 partials
    int norm_blocks;
    void* tc;                                   // opaque state of the tensor-core path (pf_mlp_tc.cu), or NULL
    void* wm;                                   // opaque state of the register-resident small-net path (pf_mlp_wm.cu), or NULL
    void* tc_gemm;                              // opaque state of the tensor-core training GEMM (pf_gemm_tc.cu), or NULL
    // Packed inference weights: were they (re)written by the work most recently queued for them?  Then the next inference
    // launch may be the packing kernel's direct successor and must not touch them before its programmatic-dependency
    // wait; otherwise it stages / streams them under its predecessor's tail (pf_rl_weights_early()).
    int weights_fresh;
    int early_in_graphs;                        // pf_rl_early_weights(): the caller vouches for launches captured into graphs
};
// 1 if an inference launch on `st` may read the packed weights before its dependency wait (clears the fresh mark)
int pf_rl_weights_early(pf_rl_ctx* c, cudaStream_t st);

// ---- tensor-core path entry points (pf_mlp_tc.cu) ----
int pf_tc_create(pf_rl_ctx* c);                                   // allocates packed weights, builds tile plans
void pf_tc_destroy(pf_rl_ctx* c);
int pf_tc_pack_weights(pf_rl_ctx* c, const float* params, cudaStream_t st);
// fused MLP forward of one net on `rows` rows of x[rows, dims[0]] (fp32, row stride ldx):
// out[rows, dims[layers]] fp32.  Optionally copies the input rows to `copy_to` (rollout storage row).
int pf_tc_mlp_forward(pf_rl_ctx* c, int net, const float* params, const float* x, long long ldx, int rows,
                      float* out, float* copy_to, cudaStream_t st);
// PPO.act of both nets + Normal sample / log-prob + rollout-storage writes in ONE launch (when pf_tc_act_fusable).
int pf_tc_act_fusable(const pf_rl_ctx* c);
int pf_tc_act(pf_rl_ctx* c, const float* params, const float* obs, const float* critic_obs, const pf_rl_storage* storage,
              int t_row, int64_t policy_step, float* actions_out, float* values_out, float* logp_out, cudaStream_t st);

// ---- register-resident warp-MMA path for small nets (pf_mlp_wm.cu); same contracts as the pf_tc_* calls above ----
int pf_wm_create(pf_rl_ctx* c);                                   // sets c->wm when both nets qualify (else leaves it NULL, returns 0)
void pf_wm_destroy(pf_rl_ctx* c);
int pf_wm_pack_weights(pf_rl_ctx* c, const float* params, cudaStream_t st);
int pf_wm_mlp_forward(pf_rl_ctx* c, int net, const float* x, long long ldx, int rows, float* out, float* copy_to, cudaStream_t st);
int pf_wm_act(pf_rl_ctx* c, const float* params, const float* obs, const float* critic_obs, const pf_rl_storage* storage,
              int t_row, int64_t policy_step, float* actions_out, float* values_out, float* logp_out, cudaStream_t st);

// ---- GEMM contract of the fp32 kernel (pf_rl.cu) ----
struct GemmArgs {
    const float* A; long long sam, sak; const int* a_rows;   // a_rows gathers the m index (forward)
    const float* B; long long sbk, sbn; const int* b_rows;   // b_rows gathers the k index (wgrad)
    float* C; long long ldc;
    int M, N, K;
    int k_per_split;              // K range of one blockIdx.z slice
    long long split_stride;       // C offset between slices (partials buffer), 0 when unsplit
    const float* bias;            // [N] or NULL
    int elu;                      // 1: C = elu(acc + bias)
    const float* elu_out;         // dgrad: multiply by ELU'(x) recovered from the saved activation y: y > 0 ? 1 : y + 1
    long long ld_elu;
};

// ---- tensor-core PPO.update (pf_train_tc.cu): state hangs off pf_rl_ctx::tc_gemm ----
int pf_tt_create(pf_rl_ctx* c);
void pf_tt_destroy(pf_rl_ctx* c);
int pf_tt_pack_weights(pf_rl_ctx* c, const float* params, cudaStream_t st);
// gather + forward of both nets + `loss` (fills c->dmu / c->dval from c->mu / c->val, and the std gradient) + backward
int pf_tt_minibatch(pf_rl_ctx* c, const pf_rl_storage* storage, const pf_rl_learner* L, const int* idx, cudaStream_t st,
                    int (*loss)(pf_rl_ctx*, const pf_rl_storage*, const pf_rl_learner*, const int*, cudaStream_t));
