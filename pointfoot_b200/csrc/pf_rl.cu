// rsl_rl v1.0.2 side of the hot path (sm_100a): ActorCritic forward, sampling + log-prob, rollout-storage
// writes, GAE reverse scan, the PPO minibatch loss / backward and the fused clip + Adam step.
// This file holds the fp32 kernels (PF_RL_FP32: the parity path) and every non-GEMM kernel; the tcgen05
// tensor-core tiles (PF_RL_BF16_TC) live in pf_mlp_tc.cu and are called from here through pf_rl_internal.h.
// Cited upstream files: rsl_rl/modules/actor_critic.py, algorithms/ppo.py, storage/rollout_storage.py (tag v1.0.2;
// third party).
#include <cuda_runtime.h>
#include <stdio.h>
#include <string.h>
#include <math.h>
#include "pf_rl_internal.h"
#include "pf_rng.cuh"
#include "pf_launch.h"

static thread_local char g_rl_err[512] = "";
extern "C" const char* pf_last_error(void);
void pf_set_error(const char* msg);      // pf_env.cu

static int rl_fail(int code, const char* what, const char* detail) {
    snprintf(g_rl_err, sizeof(g_rl_err), "%s: %s", what, detail ? detail : "");
    pf_set_error(g_rl_err);
    return code;
}
#define RL_CUDA(call)                                                                 \
    do {                                                                              \
        cudaError_t err__ = (call);                                                   \
        if (err__ != cudaSuccess) return rl_fail(2, #call, cudaGetErrorString(err__)); \
    } while (0)

// ================================================================================================ GEMM (fp32)
// C[M,N] = epilogue( sum_k A(m,k) * B(k,n) ) with fully generic element strides, so one kernel serves
//   forward  Y = X W^T      A = X[row(m), k]      B = W[n, k]
//   dgrad    dX = dY W      A = dY[m, k]          B = W[k, n]
//   wgrad    dW = dY^T X    A = dY[k, m]          B = X[row(k), n]     (K = batch, split over blockIdx.z)
// row(): optional gather through the minibatch permutation (mini_batch_generator's indexing), applied to the
// batch index, so minibatches are never materialised.
// Tile 64x64x16, 256 threads, 4x4 outputs per thread; the accumulation order over k is fixed (ascending), and
// split-K partials are reduced in a fixed order by a second kernel: results are run-to-run deterministic.
#define GT 64
#define GK 16
__global__ void __launch_bounds__(256) pf_sgemm_kernel(const GemmArgs g) {
    __shared__ float As[GK][GT + 4];
    __shared__ float Bs[GK][GT + 4];
    const int tid = threadIdx.x;
    const int m0 = blockIdx.y * GT, n0 = blockIdx.x * GT;
    const int kbeg = blockIdx.z * g.k_per_split;
    const int kend = min(g.K, kbeg + g.k_per_split);
    const int tx = tid & 15, ty = tid >> 4;           // thread computes rows ty*4.., cols tx*4..
    float acc[4][4] = {};
    const bool a_kfast = g.sak == 1, b_nfast = g.sbn == 1;
    for (int k0 = kbeg; k0 < kend; k0 += GK) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int i = tid + 256 * j;
            {   // A tile [64 m][16 k]
                const int m = a_kfast ? (i >> 4) : (i & 63), k = a_kfast ? (i & 15) : (i >> 6);
                float v = 0.0f;
                if (m0 + m < g.M && k0 + k < kend) {
                    const long long row = g.a_rows ? (long long)g.a_rows[m0 + m] : (long long)(m0 + m);
                    v = g.A[row * g.sam + (long long)(k0 + k) * g.sak];
                }
                As[k][m] = v;
            }
            {   // B tile [16 k][64 n]
                const int n = b_nfast ? (i & 63) : (i >> 4), k = b_nfast ? (i >> 6) : (i & 15);
                float v = 0.0f;
                if (n0 + n < g.N && k0 + k < kend) {
                    const long long kr = g.b_rows ? (long long)g.b_rows[k0 + k] : (long long)(k0 + k);
                    v = g.B[kr * g.sbk + (long long)(n0 + n) * g.sbn];
                }
                Bs[k][n] = v;
            }
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < GK; ++k) {
            const float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
            const float4 b = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
            const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        __syncthreads();
    }
    float* C = g.C + (long long)blockIdx.z * g.split_stride;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int m = m0 + ty * 4 + i;
        if (m >= g.M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int n = n0 + tx * 4 + j;
            if (n >= g.N) continue;
            float v = acc[i][j];
            if (g.bias) v += g.bias[n];
            if (g.elu) v = v > 0.0f ? v : expm1f(v);
            if (g.elu_out) {
                const float y = g.elu_out[(long long)m * g.ld_elu + n];
                v *= y > 0.0f ? 1.0f : y + 1.0f;
            }
            C[(long long)m * g.ldc + n] = v;
        }
    }
}

// out[i] = sum_s partial[s][i], fixed order (split-K wgrad partials, bias column sums)
__global__ void pf_reduce_partials_kernel(const float* __restrict__ part, int splits, long long stride, float* out,
                                          long long n) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float s = 0.0f;
    for (int k = 0; k < splits; ++k) s += part[(long long)k * stride + i];
    out[i] = s;
}

// column sums of dY[B, n] over a row chunk (bias gradient partials): grid (ceil(n/32), splits), block (32, 8)
__global__ void pf_colsum_kernel(const float* __restrict__ dY, long long ld, int rows, int n, int rows_per_split,
                                 float* part, long long stride) {
    __shared__ float red[8][33];
    const int c = blockIdx.x * 32 + threadIdx.x;
    const int r0 = blockIdx.y * rows_per_split, r1 = min(rows, r0 + rows_per_split);
    float s = 0.0f;
    if (c < n)
        for (int r = r0 + threadIdx.y; r < r1; r += 8) s += dY[(long long)r * ld + c];
    red[threadIdx.y][threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.y == 0 && c < n) {
        float t = 0.0f;
        for (int k = 0; k < 8; ++k) t += red[k][threadIdx.x];
        part[(long long)blockIdx.y * stride + c] = t;
    }
}

static int launch_gemm(pf_rl_ctx* c, const GemmArgs& g, int splits, cudaStream_t st) {
    (void)c;
    dim3 grid((g.N + GT - 1) / GT, (g.M + GT - 1) / GT, splits);
    PF_LAUNCHED(), pf_sgemm_kernel<<<grid, 256, 0, st>>>(g);
    RL_CUDA(cudaGetLastError());
    return 0;
}

// ================================================================================================ ctx
extern "C" int pf_sizeof_rl_params(void) { return (int)sizeof(pf_rl_params); }

static int setup_net(PfNet& n, int layers, const int32_t* dims, long long& off) {
    if (layers < 1 || layers > PF_MAX_MLP_LAYERS) return 1;
    n.layers = layers;
    for (int i = 0; i <= layers; ++i) {
        if (dims[i] <= 0 || dims[i] > 4096) return 1;
        n.dims[i] = dims[i];
    }
    for (int l = 0; l < layers; ++l) {
        n.w_off[l] = off; off += (long long)dims[l] * dims[l + 1];
        n.b_off[l] = off; off += dims[l + 1];
    }
    return 0;
}

static int rl_ctx_init(pf_rl_ctx* c);

extern "C" int pf_rl_create(const pf_rl_params* params, pf_rl_ctx** out) {
    if (!params || !out) return rl_fail(1, "pf_rl_create", "null argument");
    const pf_rl_params& P = *params;
    if (P.abi_version != PF_ABI_VERSION) return rl_fail(1, "pf_rl_create", "ABI version mismatch");
    if (P.num_envs <= 0 || P.num_steps <= 0 || P.num_actions <= 0 || P.num_actions > 32 || P.num_mini_batches <= 0)
        return rl_fail(1, "pf_rl_create", "parameter out of range");
    if (P.actor_dims[0] != P.num_obs || P.critic_dims[0] != P.num_critic_obs ||
        P.actor_dims[P.actor_layers] != P.num_actions || P.critic_dims[P.critic_layers] != 1)
        return rl_fail(1, "pf_rl_create", "MLP dims do not match the observation / action widths");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) return rl_fail(2, "pf_rl_create: no CUDA device; there is no CPU fallback", cudaGetErrorString(e));
    pf_rl_ctx* c = new pf_rl_ctx();      // value-initialised: every pointer starts NULL, so a failed create can be destroyed
    c->P = P;
    if (int rc = rl_ctx_init(c)) { pf_rl_destroy(c); return rc; }
    *out = c;
    return 0;
}

// the part of pf_rl_create that can fail once the context exists (the caller destroys it on a non-zero return)
static int rl_ctx_init(pf_rl_ctx* c) {
    const pf_rl_params& P = c->P;
    RL_CUDA(cudaGetDevice(&c->device));
    RL_CUDA(cudaDeviceGetAttribute(&c->num_sms, cudaDevAttrMultiProcessorCount, c->device));
    long long off = 0;
    c->std_off = off; off += P.num_actions;                       // ActorCritic.parameters(): std first
    if (setup_net(c->net[0], P.actor_layers, P.actor_dims, off) || setup_net(c->net[1], P.critic_layers, P.critic_dims, off))
        return rl_fail(1, "pf_rl_create", "bad MLP description");
    c->n_params = off;
    const long long total = (long long)P.num_envs * P.num_steps;
    c->mb_size = (int)(total / P.num_mini_batches);
    c->rows_cap = c->mb_size > P.num_envs ? c->mb_size : P.num_envs;
    int maxw = 1;
    long long max_wsize = 1;
    for (int k = 0; k < 2; ++k)
        for (int l = 0; l < c->net[k].layers; ++l) {
            const int w = c->net[k].dims[l + 1];
            maxw = w > maxw ? w : maxw;
            if (c->net[k].dims[l] > maxw) maxw = c->net[k].dims[l];
            const long long ws = (long long)c->net[k].dims[l] * w;
            max_wsize = ws > max_wsize ? ws : max_wsize;
            RL_CUDA(cudaMalloc(&c->net[k].act[l], sizeof(float) * (size_t)c->rows_cap * w));
        }
    RL_CUDA(cudaMalloc(&c->perm, sizeof(int) * (size_t)total));
    for (int i = 0; i < 2; ++i) RL_CUDA(cudaMalloc(&c->dbuf[i], sizeof(float) * (size_t)c->rows_cap * maxw));
    c->wgrad_splits = (c->mb_size + 383) / 384;
    if (c->wgrad_splits > 64) c->wgrad_splits = 64;
    if (c->wgrad_splits < 1) c->wgrad_splits = 1;
    c->part_floats = (long long)c->wgrad_splits * max_wsize;
    RL_CUDA(cudaMalloc(&c->part, sizeof(float) * (size_t)c->part_floats));
    RL_CUDA(cudaMalloc(&c->mu, sizeof(float) * (size_t)c->rows_cap * P.num_actions));
    RL_CUDA(cudaMalloc(&c->val, sizeof(float) * (size_t)c->rows_cap));
    RL_CUDA(cudaMalloc(&c->dmu, sizeof(float) * (size_t)c->rows_cap * P.num_actions));
    RL_CUDA(cudaMalloc(&c->dval, sizeof(float) * (size_t)c->rows_cap));
    c->loss_blocks = (c->mb_size + 255) / 256;
    RL_CUDA(cudaMalloc(&c->loss_part, sizeof(double) * (size_t)c->loss_blocks * (8 + 32)));
    c->gae_blocks = (P.num_envs + 255) / 256;
    RL_CUDA(cudaMalloc(&c->gae_part, sizeof(double) * (size_t)c->gae_blocks * 2));
    c->norm_blocks = 256;
    RL_CUDA(cudaMalloc(&c->norm_part, sizeof(double) * (size_t)(c->norm_blocks + 8)));
    c->tc = nullptr;
    c->wm = nullptr;
    if (P.precision == PF_RL_BF16_TC) {
        // inference side: small nets (the flat policy) run register-resident on warp-level MMAs, everything else on tcgen05
        if (int rc = pf_wm_create(c)) return rc;
        if (!c->wm) { if (int rc = pf_tc_create(c)) return rc; }
        if (int rc = pf_tt_create(c)) return rc;
    } else if (P.precision != PF_RL_FP32) {
        return rl_fail(1, "pf_rl_create", "unknown precision");
    }
    return 0;
}

extern "C" int pf_rl_destroy(pf_rl_ctx* c) {
    if (!c) return 0;
    if (c->tc) pf_tc_destroy(c);
    pf_wm_destroy(c);
    for (int k = 0; k < 2; ++k) cudaFree(c->net[k].tc_weights);       // (a pf_tc_create that failed half-way)
    pf_tt_destroy(c);
    for (int k = 0; k < 2; ++k)
        for (int l = 0; l < c->net[k].layers; ++l) cudaFree(c->net[k].act[l]);
    cudaFree(c->perm); cudaFree(c->dbuf[0]); cudaFree(c->dbuf[1]); cudaFree(c->part);
    cudaFree(c->mu); cudaFree(c->val); cudaFree(c->dmu); cudaFree(c->dval);
    cudaFree(c->loss_part); cudaFree(c->gae_part); cudaFree(c->norm_part);
    delete c;
    return 0;
}

extern "C" int pf_rl_inference_path(const pf_rl_ctx* c) { return !c ? -1 : (c->wm ? 2 : (c->tc ? 1 : 0)); }

extern "C" int64_t pf_rl_param_count(const pf_rl_ctx* c) { return c ? c->n_params : -1; }

extern "C" int64_t pf_rl_param_offset(const pf_rl_ctx* c, int net, int layer, int is_bias) {
    if (!c) return -1;
    if (net < 0) return c->std_off;
    if (net > 1 || layer < 0 || layer >= c->net[net].layers) return -1;
    return is_bias ? c->net[net].b_off[layer] : c->net[net].w_off[layer];
}

int pf_rl_weights_early(pf_rl_ctx* c, cudaStream_t st) {
    static const bool off = [] { const char* e = getenv("PF_RL_EARLY_WEIGHTS"); return e && atoi(e) == 0; }();
    if (off) return 0;

    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(st, &cs) != cudaSuccess) { cudaGetLastError(); return 0; }
    // captured: the launch will be replayed in surroundings this context cannot see -- only on the caller's word
    if (cs != cudaStreamCaptureStatusNone) return c->early_in_graphs;
    const int early = !c->weights_fresh;
    c->weights_fresh = 0;
    return early;
}

extern "C" int pf_rl_weights_written(pf_rl_ctx* c) {
    if (!c) return rl_fail(1, "pf_rl_weights_written", "null context");
    c->weights_fresh = 1;
    return 0;
}

extern "C" int pf_rl_early_weights(pf_rl_ctx* c, int on) {
    if (!c) return rl_fail(1, "pf_rl_early_weights", "null context");
    c->early_in_graphs = on ? 1 : 0;
    return 0;
}

extern "C" int pf_rl_params_changed(pf_rl_ctx* c, const float* params, void* stream) {
    if (!c || !params) return rl_fail(1, "pf_rl_params_changed", "null argument");
    c->weights_fresh = 1;
    if (c->tc) { if (int rc = pf_tc_pack_weights(c, params, (cudaStream_t)stream)) return rc; }
    if (c->wm) { if (int rc = pf_wm_pack_weights(c, params, (cudaStream_t)stream)) return rc; }
    if (c->tc_gemm) return pf_tt_pack_weights(c, params, (cudaStream_t)stream);
    return 0;
}

// ================================================================================================ forward
// fp32 forward of one net over `rows` rows of x (row stride ldx; optional gather of the row index through
// `rows_idx`).  Hidden activations are kept in net.act[] for the backward pass; returns the output pointer.
static int forward_fp32(pf_rl_ctx* c, int k, const float* params, const float* x, long long ldx, const int* rows_idx,
                        int rows, cudaStream_t st) {
    PfNet& n = c->net[k];
    const float* in = x;
    long long ld = ldx;
    for (int l = 0; l < n.layers; ++l) {
        GemmArgs g = {};
        g.A = in; g.sam = ld; g.sak = 1; g.a_rows = l == 0 ? rows_idx : nullptr;
        g.B = params + n.w_off[l]; g.sbk = 1; g.sbn = n.dims[l];
        g.C = n.act[l]; g.ldc = n.dims[l + 1];
        g.M = rows; g.N = n.dims[l + 1]; g.K = n.dims[l]; g.k_per_split = g.K; g.split_stride = 0;
        g.bias = params + n.b_off[l];
        g.elu = l < n.layers - 1;
        if (int rc = launch_gemm(c, g, 1, st)) return rc;
        in = n.act[l]; ld = n.dims[l + 1];
    }
    return 0;
}

// net output [rows, out] after a forward (fp32 path: last activation buffer)
static inline float* net_out(pf_rl_ctx* c, int k) { return c->net[k].act[c->net[k].layers - 1]; }

// ================================================================================================ act
// Normal(mean, std).sample() under the RNG contract + log_prob(actions).sum(-1) (actor_critic.py act /
// get_actions_log_prob), written straight into row t of the rollout storage (rollout_storage.py add_transitions).
__global__ void pf_act_sample_kernel(const pf_rl_params P, const float* __restrict__ mean, const float* __restrict__ std,
                                     const float* __restrict__ value, pf_rl_storage S, int has_storage, int t,
                                     uint32_t policy_step, float* actions_out, float* values_out, float* logp_out) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= P.num_envs) return;
    const int A = P.num_actions;
    const uint32_t ek = pf_rng_env_key(pf_rng_key(P.seed, policy_step, PF_RNG_POLICY), (uint32_t)(P.env_id_offset + e));
    const long long row = (long long)t * P.num_envs + e;
    float logp = 0.0f;
    for (int a = 0; a < A; ++a) {
        const float m = mean[(long long)e * A + a], s = std[a];
        const float act = m + s * pf_rng_normal(ek, (uint32_t)a);
        // torch.distributions.Normal.log_prob: -((x - loc)^2) / (2 var) - log(scale) - log(sqrt(2 pi))
        const float d = act - m;
        logp += -(d * d) / (2.0f * (s * s)) - logf(s) - 0.9189385332046727f;
        if (actions_out) actions_out[(long long)e * A + a] = act;
        if (has_storage) {
            S.actions[row * A + a] = act;
            S.mu[row * A + a] = m;
            S.sigma[row * A + a] = s;
        }
    }
    const float v = value[e];
    if (values_out) values_out[e] = v;
    if (logp_out) logp_out[e] = logp;
    if (has_storage) {
        S.values[row] = v;
        S.actions_log_prob[row] = logp;
    }
}

static int check_rl(pf_rl_ctx* c, const void* a, const void* b, const char* what) {
    if (!c || !a || !b) return rl_fail(1, what, "null argument");
    return 0;
}

static int mlp_forward(pf_rl_ctx* c, int k, const float* params, const float* x, int rows, float* out_tc,
                       float* copy_to, cudaStream_t st, float** out) {
    PfNet& n = c->net[k];
    if (c->wm) {
        if (int rc = pf_wm_mlp_forward(c, k, x, n.dims[0], rows, out_tc, copy_to, st)) return rc;
        *out = out_tc;
        return 0;
    }
    if (c->tc) {
        if (int rc = pf_tc_mlp_forward(c, k, params, x, n.dims[0], rows, out_tc, copy_to, st)) return rc;
        *out = out_tc;
        return 0;
    }
    if (int rc = forward_fp32(c, k, params, x, n.dims[0], nullptr, rows, st)) return rc;
    if (copy_to)
        RL_CUDA(cudaMemcpyAsync(copy_to, x, sizeof(float) * (size_t)rows * n.dims[0], cudaMemcpyDeviceToDevice, st));
    *out = net_out(c, k);
    return 0;
}

extern "C" int pf_rl_act(pf_rl_ctx* c, const float* params, const float* obs, const float* critic_obs, int32_t t,
                         const pf_rl_storage* storage, int64_t policy_step, float* actions_out, float* values_out,
                         float* logp_out, void* stream) {
    if (int rc = check_rl(c, params, obs, "pf_rl_act")) return rc;
    const pf_rl_params& P = c->P;
    cudaStream_t st = (cudaStream_t)stream;
    if (storage && (t < 0 || t >= P.num_steps)) return rl_fail(5, "pf_rl_act", "Rollout buffer overflow");
    if (!critic_obs) critic_obs = obs;
    if (c->wm)
        return pf_wm_act(c, params, obs, critic_obs, storage, storage ? t : 0, policy_step, actions_out, values_out, logp_out, st);
    if (c->tc && pf_tc_act_fusable(c))
        return pf_tc_act(c, params, obs, critic_obs, storage, storage ? t : 0, policy_step, actions_out, values_out, logp_out, st);
    const long long row0 = (long long)t * P.num_envs;
    float* mean = nullptr; float* val = nullptr;
    float* copy_obs = storage ? storage->observations + row0 * P.num_obs : nullptr;
    float* copy_priv = (storage && storage->privileged_observations)
                           ? storage->privileged_observations + row0 * P.num_critic_obs : nullptr;
    if (int rc = mlp_forward(c, 0, params, obs, P.num_envs, c->mu, copy_obs, st, &mean)) return rc;
    if (int rc = mlp_forward(c, 1, params, critic_obs, P.num_envs, c->val, copy_priv, st, &val)) return rc;
    pf_rl_storage S = {};
    if (storage) S = *storage;
    PF_LAUNCHED(), pf_act_sample_kernel<<<(P.num_envs + 127) / 128, 128, 0, st>>>(P, mean, params + c->std_off, val, S, storage != nullptr,
                                                                   t, (uint32_t)policy_step, actions_out, values_out, logp_out);
    RL_CUDA(cudaGetLastError());
    return 0;
}

// ActorCritic.act_inference / evaluate accept any batch size upstream (a single observation row in play / export
// checks, a whole rollout in analysis scripts): `rows` rows are processed in chunks of at most rows_cap, the size
// the activation workspace was allocated for.
static int mlp_rows(pf_rl_ctx* c, int k, const float* params, const float* x, int rows, float* out, cudaStream_t st) {
    const int in_w = c->net[k].dims[0], out_w = c->net[k].dims[c->net[k].layers];
    for (int r0 = 0; r0 < rows; r0 += c->rows_cap) {
        const int n = rows - r0 < c->rows_cap ? rows - r0 : c->rows_cap;
        float* dst = out + (size_t)r0 * out_w;
        float* res = nullptr;
        if (int rc = mlp_forward(c, k, params, x + (size_t)r0 * in_w, n, dst, nullptr, st, &res)) return rc;
        if (res != dst)
            RL_CUDA(cudaMemcpyAsync(dst, res, sizeof(float) * (size_t)n * out_w, cudaMemcpyDeviceToDevice, st));
    }
    return 0;
}

extern "C" int pf_rl_act_inference(pf_rl_ctx* c, const float* params, const float* obs, int32_t rows, float* actions_out,
                                   void* stream) {
    if (int rc = check_rl(c, params, obs, "pf_rl_act_inference")) return rc;
    if (!actions_out) return rl_fail(1, "pf_rl_act_inference", "null output");
    if (rows < 0) return rl_fail(1, "pf_rl_act_inference", "negative row count");
    return mlp_rows(c, 0, params, obs, rows, actions_out, (cudaStream_t)stream);
}

extern "C" int pf_rl_evaluate(pf_rl_ctx* c, const float* params, const float* critic_obs, int32_t rows, float* values_out,
                              void* stream) {
    if (int rc = check_rl(c, params, critic_obs, "pf_rl_evaluate")) return rc;
    if (!values_out) return rl_fail(1, "pf_rl_evaluate", "null output");
    if (rows < 0) return rl_fail(1, "pf_rl_evaluate", "negative row count");
    return mlp_rows(c, 1, params, critic_obs, rows, values_out, (cudaStream_t)stream);
}

// PPO.process_env_step (ppo.py): rewards = rew.clone(); rewards += gamma * squeeze(values * time_outs)
__global__ void pf_process_env_step_kernel(const pf_rl_params P, pf_rl_storage S, int t, const float* __restrict__ rew,
                                           const uint8_t* __restrict__ dones, const uint8_t* __restrict__ time_outs) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= P.num_envs) return;
    const long long row = (long long)t * P.num_envs + e;
    float r = rew[e];
    if (time_outs) r = __fadd_rn(r, __fmul_rn(P.gamma, __fmul_rn(S.values[row], time_outs[e] ? 1.0f : 0.0f)));
    S.rewards[row] = r;
    S.dones[row] = dones[e] ? 1 : 0;
}

extern "C" int pf_rl_process_env_step(pf_rl_ctx* c, const pf_rl_storage* storage, int32_t t, const float* rew,
                                      const uint8_t* dones, const uint8_t* time_outs, void* stream) {
    if (int rc = check_rl(c, storage, rew, "pf_rl_process_env_step")) return rc;
    if (!dones) return rl_fail(1, "pf_rl_process_env_step", "null dones");
    if (t < 0 || t >= c->P.num_steps) return rl_fail(5, "pf_rl_process_env_step", "Rollout buffer overflow");
    PF_LAUNCHED(), pf_process_env_step_kernel<<<(c->P.num_envs + 255) / 256, 256, 0, (cudaStream_t)stream>>>(c->P, *storage, t, rew, dones, time_outs);
    RL_CUDA(cudaGetLastError());
    return 0;
}

// ================================================================================================ GAE
// RolloutStorage.compute_returns (rollout_storage.py): reverse scan over T, one thread per env, coalesced over
// the env dimension.  Each op is one fp32 rounding in the reference's order (torch eager), hence the explicit
// _rn intrinsics: returns / raw advantages are bit-identical to the torch loop.
__global__ void __launch_bounds__(256) pf_gae_kernel(const pf_rl_params P, pf_rl_storage S,
                                                     const float* __restrict__ last_values, double* part) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    const int N = P.num_envs, T = P.num_steps;
    double s1 = 0.0, s2 = 0.0;
    if (e < N) {
        float adv = 0.0f;
        float next_v = last_values[e];
        for (int t = T - 1; t >= 0; --t) {
            const long long i = (long long)t * N + e;
            const float v = S.values[i], r = S.rewards[i];
            const float nt = __fsub_rn(1.0f, S.dones[i] ? 1.0f : 0.0f);
            const float ntg = __fmul_rn(nt, P.gamma);
            const float delta = __fsub_rn(__fadd_rn(r, __fmul_rn(ntg, next_v)), v);
            adv = __fadd_rn(delta, __fmul_rn(__fmul_rn(ntg, P.lam), adv));
            const float ret = __fadd_rn(adv, v);
            S.returns[i] = ret;
            const float a = __fsub_rn(ret, v);          // self.advantages = self.returns - self.values
            S.advantages[i] = a;
            s1 += (double)a; s2 += (double)a * (double)a;
            next_v = v;
        }
    }
    __shared__ double r1[8], r2[8];
    for (int o = 16; o > 0; o >>= 1) { s1 += __shfl_xor_sync(0xffffffffu, s1, o); s2 += __shfl_xor_sync(0xffffffffu, s2, o); }
    if ((threadIdx.x & 31) == 0) { r1[threadIdx.x >> 5] = s1; r2[threadIdx.x >> 5] = s2; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0, b = 0.0;
        for (int w = 0; w < 8; ++w) { a += r1[w]; b += r2[w]; }
        part[2 * blockIdx.x] = a; part[2 * blockIdx.x + 1] = b;
    }
}

__global__ void pf_gae_stats_kernel(const double* __restrict__ part, int blocks, double count, double* stats) {
    __shared__ double r1[32], r2[32];
    double a = 0.0, b = 0.0;
    for (int i = threadIdx.x; i < blocks; i += blockDim.x) { a += part[2 * i]; b += part[2 * i + 1]; }
    for (int o = 16; o > 0; o >>= 1) { a += __shfl_xor_sync(0xffffffffu, a, o); b += __shfl_xor_sync(0xffffffffu, b, o); }
    if ((threadIdx.x & 31) == 0) { r1[threadIdx.x >> 5] = a; r2[threadIdx.x >> 5] = b; }
    __syncthreads();
    if (threadIdx.x == 0) {
        a = b = 0.0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) { a += r1[w]; b += r2[w]; }
        stats[0] = a; stats[1] = b; stats[2] = count;
    }
}

// advantages = (advantages - mean) / (std_unbiased + 1e-8), mean / std over all T*N samples (of all devices)
__global__ void pf_norm_adv_kernel(float* adv, long long n, const double* __restrict__ stats) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double cnt = stats[2], mean = stats[0] / cnt;
    double var = (stats[1] - cnt * mean * mean) / (cnt - 1.0);
    var = var > 0.0 ? var : 0.0;
    const float m = (float)mean, sd = (float)sqrt(var);
    adv[i] = __fdiv_rn(__fsub_rn(adv[i], m), __fadd_rn(sd, 1e-8f));
}

extern "C" int pf_rl_gae(pf_rl_ctx* c, const pf_rl_storage* storage, const float* last_values, double* stats, void* stream) {
    if (int rc = check_rl(c, storage, last_values, "pf_rl_gae")) return rc;
    if (!stats) return rl_fail(1, "pf_rl_gae", "null stats");
    cudaStream_t st = (cudaStream_t)stream;
    PF_LAUNCHED(), pf_gae_kernel<<<c->gae_blocks, 256, 0, st>>>(c->P, *storage, last_values, c->gae_part);
    RL_CUDA(cudaGetLastError());
    PF_LAUNCHED(), pf_gae_stats_kernel<<<1, 256, 0, st>>>(c->gae_part, c->gae_blocks, (double)c->P.num_envs * c->P.num_steps, stats);
    RL_CUDA(cudaGetLastError());
    return 0;
}

// ------------------------------------------------------------------------------------------------ GAE without allocation
// pf_rl_gae_noalloc: the same two kernels as pf_rl_gae (pf_gae_kernel + pf_gae_stats_kernel, unchanged), for callers that
// cannot allocate device memory: no pf_rl_ctx (pf_rl_create calls cudaMalloc), the per-CTA partials live in a file-scope
// device array instead of the context's gae_part.  No counterpart in the
// reference.
//   * the array is PF_GAE_NOALLOC_MAX_ENVS / 256 CTAs x 2 doubles = 64 KB; every GPU that loads the library holds it until
//     the process exits;
//   * in HOST code the array's name is only a host-side stand-in for the symbol, not a device address: the device address
//     is looked up once per device (with that device current) by cudaGetSymbolAddress and that pointer is what the
//     kernels are handed;
//   * every call on a device shares the array.  Nothing in it has to start at zero -- every CTA of pf_gae_kernel overwrites
//     its own two partials and pf_gae_stats_kernel reads exactly the CTAs of this call -- so a call does not depend on what
//     the previous one left; but two calls that OVERLAP on the GPU (different streams, no ordering between them) race
//     on it.  Calls on one stream, or ordered by events, are safe.
#define PF_GAE_NOALLOC_MAX_DEVICES 64
static __device__ __align__(256) double g_pf_gae_noalloc_part[2 * (PF_GAE_NOALLOC_MAX_ENVS / 256)];
static double* g_pf_gae_noalloc_ptr[PF_GAE_NOALLOC_MAX_DEVICES] = {nullptr};      // device address of the array, per device

// device address of the scratch array on the CURRENT device (looked up on the first call there); NULL + error on failure
static double* gae_noalloc_scratch(void) {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= PF_GAE_NOALLOC_MAX_DEVICES) {
        rl_fail(2, "pf_rl_gae_noalloc", "no current CUDA device (or device ordinal >= 64)");
        return nullptr;
    }
    double* p = __atomic_load_n(&g_pf_gae_noalloc_ptr[dev], __ATOMIC_ACQUIRE);
    if (p) return p;
    void* sym = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&sym, g_pf_gae_noalloc_part);
    if (e != cudaSuccess || !sym) {
        rl_fail(2, "pf_rl_gae_noalloc: cudaGetSymbolAddress", cudaGetErrorString(e));
        return nullptr;
    }
    // (two threads racing here store the same address: the symbol has one address per device)
    __atomic_store_n(&g_pf_gae_noalloc_ptr[dev], static_cast<double*>(sym), __ATOMIC_RELEASE);
    return static_cast<double*>(sym);
}

extern "C" int pf_rl_gae_noalloc(const float* rewards, const uint8_t* dones, const float* values, const float* last_values,
                                 int32_t num_envs, int32_t num_steps, float gamma, float lam, float* returns,
                                 float* advantages, double* stats, void* stream) {
    if (!rewards || !dones || !values || !last_values || !returns || !advantages || !stats)
        return rl_fail(1, "pf_rl_gae_noalloc", "null argument");
    if (num_envs <= 0 || num_steps <= 0) return rl_fail(1, "pf_rl_gae_noalloc", "num_envs and num_steps must be positive");
    if (num_envs > PF_GAE_NOALLOC_MAX_ENVS) {
        char msg[160];
        snprintf(msg, sizeof(msg), "num_envs = %d exceeds PF_GAE_NOALLOC_MAX_ENVS = %d (the static scratch array holds no more); "
                 "use pf_rl_gae with a pf_rl_ctx for larger batches", (int)num_envs, (int)PF_GAE_NOALLOC_MAX_ENVS);
        return rl_fail(8, "pf_rl_gae_noalloc", msg);
    }
    double* part = gae_noalloc_scratch();
    if (!part) return 2;
    pf_rl_params P;
    memset(&P, 0, sizeof(P));
    P.abi_version = PF_ABI_VERSION; P.num_envs = num_envs; P.num_steps = num_steps; P.gamma = gamma; P.lam = lam;
    pf_rl_storage S;
    memset(&S, 0, sizeof(S));
    S.rewards = const_cast<float*>(rewards); S.dones = const_cast<uint8_t*>(dones); S.values = const_cast<float*>(values);
    S.returns = returns; S.advantages = advantages;
    cudaStream_t st = (cudaStream_t)stream;
    const int blocks = (num_envs + 255) / 256;
    PF_LAUNCHED(), pf_gae_kernel<<<blocks, 256, 0, st>>>(P, S, last_values, part);
    RL_CUDA(cudaGetLastError());
    PF_LAUNCHED(), pf_gae_stats_kernel<<<1, 256, 0, st>>>(part, blocks, (double)num_envs * num_steps, stats);
    RL_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int pf_rl_normalize_advantages(pf_rl_ctx* c, const pf_rl_storage* storage, const double* stats, void* stream) {
    if (int rc = check_rl(c, storage, stats, "pf_rl_normalize_advantages")) return rc;
    const long long n = (long long)c->P.num_envs * c->P.num_steps;
    PF_LAUNCHED(), pf_norm_adv_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(storage->advantages, n, stats);
    RL_CUDA(cudaGetLastError());
    return 0;
}

// ================================================================================================ shuffle
// mini_batch_generator's randperm under the contract of oracle/rsl_rl_ref.py feistel_permutation: 4-round Feistel
// network over an even number of bits with cycle walking -- a bijection computed independently per element.
__global__ void pf_permutation_kernel(int* perm, unsigned n, int half, uint32_t k0, uint32_t k1, uint32_t k2, uint32_t k3) {
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t mask = (1u << half) - 1u;
    const uint32_t keys[4] = {k0, k1, k2, k3};
    uint32_t x = i;
    do {
        uint32_t l = x >> half, r = x & mask;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t f = pf_fmix32(r + keys[j]) & mask;
            const uint32_t nl = r;
            r = l ^ f;
            l = nl;
        }
        x = (l << half) | r;
    } while (x >= n);
    perm[i] = (int)x;
}

extern "C" int pf_rl_make_permutation(pf_rl_ctx* c, int64_t iteration, void* stream) {
    if (!c) return rl_fail(1, "pf_rl_make_permutation", "null ctx");
    const unsigned n = (unsigned)c->P.num_mini_batches * (unsigned)c->mb_size;
    int bits = 2;
    while (bits < 32 && (1ull << bits) < (unsigned long long)n) ++bits;
    if (bits < 2) bits = 2;
    bits += bits & 1;
    const uint32_t ek = pf_rng_env_key(pf_rng_key(c->P.seed, (uint32_t)iteration, 7u), 0u);
    PF_LAUNCHED(), pf_permutation_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(c->perm, n, bits / 2, pf_rng_u32(ek, 0), pf_rng_u32(ek, 1),
                                                                           pf_rng_u32(ek, 2), pf_rng_u32(ek, 3));
    RL_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int pf_rl_get_permutation(pf_rl_ctx* c, int32_t* out, void* stream) {
    if (!c || !out) return rl_fail(1, "pf_rl_get_permutation", "null argument");
    const size_t n = (size_t)c->P.num_mini_batches * (size_t)c->mb_size;
    RL_CUDA(cudaMemcpyAsync(out, c->perm, n * sizeof(int), cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return 0;
}

// ================================================================================================ PPO loss
// ppo.py update(), per sample of the minibatch: log-prob of the stored action under the new policy, the
// adaptive-KL statistic, clipped surrogate, clipped value loss, entropy; and their gradients w.r.t. the network
// outputs (mu, value) and std.  Reductions are per-CTA partials in double, summed in a fixed order afterwards.
// loss = mean(max(s1, s2)) + c_v * mean(max(v1, v2)) - c_e * mean(entropy)
#define LOSS_THREADS 256
#define LOSS_SLOTS (8 + 32)
__global__ void __launch_bounds__(LOSS_THREADS) pf_ppo_loss_kernel(const pf_rl_params P, pf_rl_storage S,
                                                                   const int* __restrict__ perm, int mb_size,
                                                                   const float* __restrict__ mu, const float* __restrict__ val,
                                                                   const float* __restrict__ std, float* dmu, float* dval,
                                                                   double* part) {
    const int b = blockIdx.x * LOSS_THREADS + threadIdx.x;
    const int A = P.num_actions;
    const float inv_n = 1.0f / (float)mb_size;
    double acc[5] = {0, 0, 0, 0, 0};              // surrogate, value loss, entropy, kl, count
    float dstd[32];
#pragma unroll
    for (int a = 0; a < 32; ++a) dstd[a] = 0.0f;
    if (b < mb_size) {
        const long long i = perm[b];
        float logp = 0.0f, ent = 0.0f, kl = 0.0f;
        for (int a = 0; a < A; ++a) {
            const float m = mu[(long long)b * A + a], s = std[a];
            const float x = S.actions[i * A + a];
            const float d = x - m;
            logp += -(d * d) / (2.0f * (s * s)) - logf(s) - 0.9189385332046727f;
            ent += 0.5f + 0.9189385332046727f + logf(s);
            const float om = S.mu[i * A + a], os = S.sigma[i * A + a];
            kl += logf(s / os + 1.e-5f) + (os * os + (om - m) * (om - m)) / (2.0f * (s * s)) - 0.5f;
        }
        const float adv = S.advantages[i];
        const float ratio = expf(logp - S.actions_log_prob[i]);
        const float s1 = -adv * ratio;
        const float s2 = -adv * fminf(fmaxf(ratio, 1.0f - P.clip_param), 1.0f + P.clip_param);
        const float sur = fmaxf(s1, s2);
        const float g_ratio = s1 >= s2 ? -adv : 0.0f;          // torch.max: ties happen only inside the clip range
        const float dlogp = g_ratio * ratio * inv_n;
        for (int a = 0; a < A; ++a) {
            const float m = mu[(long long)b * A + a], s = std[a];
            const float d = S.actions[i * A + a] - m;
            dmu[(long long)b * A + a] = dlogp * d / (s * s);
            // d logp / d sigma = d^2 / s^3 - 1/s ;  entropy: d/d sigma = 1/s
            dstd[a] = dlogp * (d * d / (s * s * s) - 1.0f / s) - P.entropy_coef * inv_n / s;
        }
        const float v = val[b], ret = S.returns[i], tv = S.values[i];
        float vl, gv;
        if (P.use_clipped_value_loss) {
            const float dv = v - tv;
            const float vc = tv + fminf(fmaxf(dv, -P.clip_param), P.clip_param);
            const float l1 = (v - ret) * (v - ret), l2 = (vc - ret) * (vc - ret);
            const float in_range = (dv >= -P.clip_param && dv <= P.clip_param) ? 1.0f : 0.0f;
            const float g1 = 2.0f * (v - ret), g2 = 2.0f * (vc - ret) * in_range;
            vl = fmaxf(l1, l2);
            gv = l1 > l2 ? g1 : (l1 < l2 ? g2 : 0.5f * (g1 + g2));   // torch.max splits the gradient on ties
        } else {
            vl = (ret - v) * (ret - v);
            gv = 2.0f * (v - ret);
        }
        dval[b] = P.value_loss_coef * gv * inv_n;
        acc[0] = sur; acc[1] = vl; acc[2] = ent; acc[3] = kl; acc[4] = 1.0;
    }
    // CTA reduction (fixed tree)
    __shared__ double red[LOSS_THREADS / 32][LOSS_SLOTS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < 5; ++k) {
        double v = acc[k];
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) red[warp][k] = v;
    }
#pragma unroll
    for (int a = 0; a < 32; ++a) {
        if (a < A) {
            double v = (double)dstd[a];
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if (lane == 0) red[warp][8 + a] = v;
        }
    }
    __syncthreads();
    if (threadIdx.x < LOSS_SLOTS) {
        const int k = threadIdx.x;
        double v = 0.0;
        if (k < 5 || (k >= 8 && k < 8 + A))
            for (int w = 0; w < LOSS_THREADS / 32; ++w) v += red[w][k];
        part[(long long)blockIdx.x * LOSS_SLOTS + k] = v;
    }
}

// partials -> loss_stats[0..4] and the gradient of std.  One warp per slot (lanes stride over the CTAs, fixed shuffle
// tree: deterministic), instead of one thread walking every CTA's partial.
__global__ void __launch_bounds__(1024) pf_ppo_loss_final_kernel(const double* __restrict__ part, int blocks, int A, double* loss_stats,
                                                                 float* grad_std) {
    pf_grid_dep_launch();
    pf_grid_dep_wait();
    const int lane = threadIdx.x & 31;
    for (int k = threadIdx.x >> 5; k < LOSS_SLOTS; k += (int)(blockDim.x >> 5)) {
        if (!(k < 5 || (k >= 8 && k < 8 + A))) continue;
        double v = 0.0;
        for (int i = lane; i < blocks; i += 32) v += part[(long long)i * LOSS_SLOTS + k];
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) {
            if (k < 5) loss_stats[k] = v;
            else grad_std[k - 8] = (float)v;
        }
    }
}

// ================================================================================================ backward
static int backward_fp32(pf_rl_ctx* c, int k, const pf_rl_learner* L, const float* x0, long long ldx0, const int* rows_idx,
                         float* dY, int rows, cudaStream_t st) {
    PfNet& n = c->net[k];
    const float* params = L->params;
    int cur = 0;                                  // dY lives in the caller's buffer first, then in dbuf[]
    for (int l = n.layers - 1; l >= 0; --l) {
        const int out = n.dims[l + 1], in = n.dims[l];
        const float* X = l == 0 ? x0 : n.act[l - 1];
        const long long ldx = l == 0 ? ldx0 : in;
        // ---- wgrad: dW[out,in] = dY^T X, split over the batch ----
        int splits = c->wgrad_splits;
        int kps = (rows + splits - 1) / splits;
        const int kq = GK;
        kps = (kps + kq - 1) / kq * kq;
        splits = (rows + kps - 1) / kps;
        GemmArgs g = {};
        g.A = dY; g.sam = 1; g.sak = out; g.a_rows = nullptr;
        g.B = X; g.sbk = ldx; g.sbn = 1; g.b_rows = l == 0 ? rows_idx : nullptr;
        g.C = c->part; g.ldc = in; g.M = out; g.N = in; g.K = rows; g.k_per_split = kps;
        g.split_stride = (long long)out * in;
        if (int rc = launch_gemm(c, g, splits, st)) return rc;
        const long long wn = (long long)out * in;
        PF_LAUNCHED(), pf_reduce_partials_kernel<<<(unsigned)((wn + 255) / 256), 256, 0, st>>>(c->part, splits, wn, L->grads + n.w_off[l], wn);
        RL_CUDA(cudaGetLastError());
        // ---- bias grad: column sums of dY ----
        dim3 cg((out + 31) / 32, splits), cb(32, 8);
        PF_LAUNCHED(), pf_colsum_kernel<<<cg, cb, 0, st>>>(dY, out, rows, out, kps, c->part, out);
        RL_CUDA(cudaGetLastError());
        PF_LAUNCHED(), pf_reduce_partials_kernel<<<(out + 255) / 256, 256, 0, st>>>(c->part, splits, out, L->grads + n.b_off[l], out);
        RL_CUDA(cudaGetLastError());
        // ---- dgrad: dH = dY W, times ELU'(saved activation) ----
        if (l > 0) {
            GemmArgs d = {};
            d.A = dY; d.sam = out; d.sak = 1;
            d.B = params + n.w_off[l]; d.sbk = in; d.sbn = 1;
            d.C = c->dbuf[cur]; d.ldc = in; d.M = rows; d.N = in; d.K = out; d.k_per_split = out;
            d.elu_out = n.act[l - 1]; d.ld_elu = in;
            if (int rc = launch_gemm(c, d, 1, st)) return rc;
            dY = c->dbuf[cur];
            cur ^= 1;
        }
    }
    return 0;
}

// pf_ppo_loss_kernel over the outputs in c->mu / c->val (the tensor-core forward's fp32 results): fills c->dmu / c->dval,
// loss_stats and the gradient of std
static int loss_on_ctx_outputs(pf_rl_ctx* c, const pf_rl_storage* storage, const pf_rl_learner* L, const int* idx, cudaStream_t st) {
    PF_LAUNCHED(), pf_ppo_loss_kernel<<<c->loss_blocks, LOSS_THREADS, 0, st>>>(c->P, *storage, idx, c->mb_size, c->mu, c->val,
                                                                                L->params + c->std_off, c->dmu, c->dval, c->loss_part);
    RL_CUDA(cudaGetLastError());
    PF_LAUNCHED(), pf_ppo_loss_final_kernel<<<1, 1024, 0, st>>>(c->loss_part, c->loss_blocks, c->P.num_actions, L->loss_stats, L->grads + c->std_off);
    RL_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int pf_rl_minibatch_grad(pf_rl_ctx* c, const pf_rl_storage* storage, const pf_rl_learner* L,
                                    int32_t minibatch, void* stream) {
    if (int rc = check_rl(c, storage, L, "pf_rl_minibatch_grad")) return rc;
    if (!L->params || !L->grads || !L->loss_stats) return rl_fail(1, "pf_rl_minibatch_grad", "null learner buffer");
    if (minibatch < 0 || minibatch >= c->P.num_mini_batches) return rl_fail(1, "pf_rl_minibatch_grad", "minibatch index out of range");
    const pf_rl_params& P = c->P;
    cudaStream_t st = (cudaStream_t)stream;
    const int MB = c->mb_size;
    const int* idx = c->perm + (long long)minibatch * MB;
    const float* cobs = storage->privileged_observations ? storage->privileged_observations : storage->observations;
    if (c->tc_gemm) return pf_tt_minibatch(c, storage, L, idx, st, loss_on_ctx_outputs);      // PF_RL_BF16_TC: tcgen05 GEMMs
    // forward (activations kept for the backward pass); the minibatch rows are gathered on the fly
    if (int rc = forward_fp32(c, 0, L->params, storage->observations, P.num_obs, idx, MB, st)) return rc;
    if (int rc = forward_fp32(c, 1, L->params, cobs, P.num_critic_obs, idx, MB, st)) return rc;
    PF_LAUNCHED(), pf_ppo_loss_kernel<<<c->loss_blocks, LOSS_THREADS, 0, st>>>(P, *storage, idx, MB, net_out(c, 0), net_out(c, 1),
                                                                 L->params + c->std_off, c->dmu, c->dval, c->loss_part);
    RL_CUDA(cudaGetLastError());
    PF_LAUNCHED(), pf_ppo_loss_final_kernel<<<1, 1024, 0, st>>>(c->loss_part, c->loss_blocks, P.num_actions, L->loss_stats, L->grads + c->std_off);
    RL_CUDA(cudaGetLastError());
    if (int rc = backward_fp32(c, 0, L, storage->observations, P.num_obs, idx, c->dmu, MB, st)) return rc;
    if (int rc = backward_fp32(c, 1, L, cobs, P.num_critic_obs, idx, c->dval, MB, st)) return rc;
    return 0;
}

// ================================================================================================ optimizer
// ppo.py update(): adaptive-KL learning rate, nn.utils.clip_grad_norm_(params, max_grad_norm), torch.optim.Adam step.
// Three launches: (1) per-CTA partial sums of grad^2 (double), (2) one thread turns them and loss_stats into
// the step's scalars (lr, clip coefficient, bias corrections) entirely on the device -- no host sync in the
// PPO loop, (3) the fused clip + Adam update over the flat parameter buffer.
__global__ void __launch_bounds__(256) pf_gradnorm_kernel(const float* __restrict__ g, long long n, double* part) {
    double s = 0.0;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) {
        const double v = (double)g[i];
        s += v * v;
    }
    __shared__ double red[8];
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < 8; ++w) t += red[w];
        part[blockIdx.x] = t;
    }
}

struct AdamScalars { float clip_coef, w1, beta2, w2, bc2_sqrt, eps, neg_step_size, pad; };

__global__ void __launch_bounds__(256) pf_adam_scalars_kernel(const pf_rl_params P, const double* __restrict__ part, int blocks,
                                                              pf_rl_learner L, AdamScalars* out) {
    // grad-norm^2: the <= 256 CTA partials summed by a fixed tree (one load per thread instead of a serial walk)
    __shared__ double red[8];
    double n2 = (int)threadIdx.x < blocks ? part[threadIdx.x] : 0.0;
    for (int o = 16; o > 0; o >>= 1) n2 += __shfl_xor_sync(0xffffffffu, n2, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = n2;
    __syncthreads();
    if (threadIdx.x != 0) return;
    n2 = 0.0;
    for (int w = 0; w < 8; ++w) n2 += red[w];
    L.loss_stats[5] = n2;
    double lr = *L.lr;
    if (P.adaptive_schedule && P.desired_kl > 0.0) {
        const float kl_mean = (float)(L.loss_stats[3] / L.loss_stats[4]);
        if (kl_mean > (float)(P.desired_kl * 2.0)) lr = fmax(1e-5, lr / 1.5);
        else if (kl_mean < (float)(P.desired_kl / 2.0) && kl_mean > 0.0f) lr = fmin(1e-2, lr * 1.5);
        *L.lr = lr;
    }
    const long long step = *L.adam_step + 1;
    *L.adam_step = step;
    const float total_norm = (float)sqrt(n2);
    float coef = P.max_grad_norm / (total_norm + 1e-6f);
    coef = coef < 1.0f ? coef : 1.0f;
    const double bc1 = 1.0 - pow(P.adam_beta1, (double)step), bc2 = 1.0 - pow(P.adam_beta2, (double)step);
    AdamScalars s;
    s.clip_coef = coef;
    s.w1 = (float)(1.0 - P.adam_beta1);
    s.beta2 = (float)P.adam_beta2;
    s.w2 = (float)(1.0 - P.adam_beta2);
    s.bc2_sqrt = (float)sqrt(bc2);
    s.eps = (float)P.adam_eps;
    s.neg_step_size = (float)(-(lr / bc1));
    s.pad = 0.0f;
    *out = s;
}

__global__ void __launch_bounds__(256) pf_adam_kernel(long long n, pf_rl_learner L, const AdamScalars* __restrict__ sc) {
    const AdamScalars s = *sc;
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < n; i += (long long)gridDim.x * 256) {
        const float g = L.grads[i] * s.clip_coef;
        L.grads[i] = g;                                      // clip_grad_norm_ scales .grad in place
        float m = L.adam_m[i], v = L.adam_v[i];
        m = m + s.w1 * (g - m);                              // exp_avg.lerp_(grad, 1 - beta1)
        v = v * s.beta2 + s.w2 * g * g;                      // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1 - beta2)
        L.adam_m[i] = m; L.adam_v[i] = v;
        const float denom = sqrtf(v) / s.bc2_sqrt + s.eps;
        L.params[i] = L.params[i] + s.neg_step_size * (m / denom);
    }
}

extern "C" int pf_rl_adam_step(pf_rl_ctx* c, const pf_rl_learner* L, void* stream) {
    if (!c || !L || !L->params || !L->grads || !L->adam_m || !L->adam_v || !L->lr || !L->adam_step || !L->loss_stats)
        return rl_fail(1, "pf_rl_adam_step", "null argument");
    cudaStream_t st = (cudaStream_t)stream;
    const long long n = c->n_params;
    int blocks = (int)((n + 255) / 256);
    if (blocks > c->norm_blocks) blocks = c->norm_blocks;
    AdamScalars* sc = reinterpret_cast<AdamScalars*>(c->norm_part + c->norm_blocks);
    PF_LAUNCHED(), pf_gradnorm_kernel<<<blocks, 256, 0, st>>>(L->grads, n, c->norm_part);
    RL_CUDA(cudaGetLastError());
    PF_LAUNCHED(), pf_adam_scalars_kernel<<<1, 256, 0, st>>>(c->P, c->norm_part, blocks, *L, sc);
    RL_CUDA(cudaGetLastError());
    int ublocks = (int)((n + 255) / 256);
    if (ublocks > c->num_sms * 8) ublocks = c->num_sms * 8;
    PF_LAUNCHED(), pf_adam_kernel<<<ublocks, 256, 0, st>>>(n, *L, sc);
    RL_CUDA(cudaGetLastError());
    c->weights_fresh = 1;
    if (c->tc) { if (int rc = pf_tc_pack_weights(c, L->params, st)) return rc; }
    if (c->wm) { if (int rc = pf_wm_pack_weights(c, L->params, st)) return rc; }
    if (c->tc_gemm) return pf_tt_pack_weights(c, L->params, st);
    return 0;
}
