// Actuator-network torques: legged_gym/envs/anymal_c/anymal.py:62-78 Anymal._compute_torques
// with cfg.control.use_actuator_network, i.e. the TorchScript module resources/actuator_nets/anydrive_v3_lstm.pt
// ("LSTMsea"): per (env, DOF) one sample x = (action * action_scale + default_dof_pos - dof_pos, dof_vel) * in_scale ->
// 2-layer LSTM, hidden size 8, sequence length 1, recurrent state (h, c) carried between sub-steps -> Linear(8 -> 1)
// * out_scale = torque.  And the state reset of anymal.py:52-56 (reset_idx).
//
// One thread per (env, DOF): 969 parameters staged once per CTA in shared memory (every thread reads the same weight at
// the same time: broadcast, no bank conflicts), ~600 FMAs and 8 transcendental pairs per thread against 256 B of
// recurrent state read and written -> HBM bound: 8 (q, qd) + 4 (action) + 4 (torque) + 2 x 128 (h, c of both layers, in
// and out) = 272 B per DOF per sub-step.  State layout is the reference's: [2 layers, N * A, 8] fp32 (anymal.py:48-49),
// so a thread's 8 values are two 128-bit accesses.  Context-free: no allocation, nothing retained.
#include <cuda_runtime.h>
#include <stdio.h>
#include "../../include/pointfoot_b200.h"
#include "pf_launch.h"

void pf_set_error(const char* msg);

#define AN_H 8                       // LSTM hidden size
#define AN_G (4 * AN_H)              // gate rows (PyTorch order: input, forget, cell, output)
// offsets inside the flat weight buffer = state_dict order of the module without its two scale buffers
#define AN_WIH0 0                    // lstm.weight_ih_l0 [32, 2]
#define AN_WHH0 (AN_WIH0 + AN_G * 2) // lstm.weight_hh_l0 [32, 8]
#define AN_BIH0 (AN_WHH0 + AN_G * AN_H)
#define AN_BHH0 (AN_BIH0 + AN_G)
#define AN_WIH1 (AN_BHH0 + AN_G)     // lstm.weight_ih_l1 [32, 8]
#define AN_WHH1 (AN_WIH1 + AN_G * AN_H)
#define AN_BIH1 (AN_WHH1 + AN_G * AN_H)
#define AN_BHH1 (AN_BIH1 + AN_G)
#define AN_WLIN (AN_BHH1 + AN_G)     // linear.weight [1, 8]
#define AN_BLIN (AN_WLIN + AN_H)     // linear.bias [1]
#define AN_NW (AN_BLIN + 1)
static_assert(AN_NW == PF_ACTUATOR_NET_WEIGHTS, "weight layout and header constant disagree");

__device__ __forceinline__ float an_sigmoid(float x) { return 1.0f / (1.0f + expf(-x)); }

// one LSTM cell (torch.nn.LSTM, one time step): gates = W_ih x + b_ih + W_hh h + b_hh; c' = f * c + i * g; h' = o * tanh(c')
template <int kIn>
__device__ __forceinline__ void an_cell(const float* __restrict__ w_ih, const float* __restrict__ w_hh,
                                        const float* __restrict__ b_ih, const float* __restrict__ b_hh,
                                        const float (&x)[kIn], float (&h)[AN_H], float (&c)[AN_H]) {
    float hn[AN_H];
#pragma unroll
    for (int j = 0; j < AN_H; ++j) {
        float g[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int r = k * AN_H + j;
            float a = b_ih[r] + b_hh[r];
#pragma unroll
            for (int i = 0; i < kIn; ++i) a = fmaf(w_ih[r * kIn + i], x[i], a);
#pragma unroll
            for (int i = 0; i < AN_H; ++i) a = fmaf(w_hh[r * AN_H + i], h[i], a);
            g[k] = a;
        }
        const float cn = an_sigmoid(g[1]) * c[j] + an_sigmoid(g[0]) * tanhf(g[2]);
        c[j] = cn;
        hn[j] = an_sigmoid(g[3]) * tanhf(cn);
    }
#pragma unroll
    for (int j = 0; j < AN_H; ++j) h[j] = hn[j];
}

__device__ __forceinline__ void an_ld8(const float* p, float (&v)[AN_H]) {
    const float4 a = reinterpret_cast<const float4*>(p)[0], b = reinterpret_cast<const float4*>(p)[1];
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ void an_st8(float* p, const float (&v)[AN_H]) {
    reinterpret_cast<float4*>(p)[0] = make_float4(v[0], v[1], v[2], v[3]);
    reinterpret_cast<float4*>(p)[1] = make_float4(v[4], v[5], v[6], v[7]);
}

__global__ void __launch_bounds__(256) pf_actuator_net_kernel(const pf_actuator_net_params P, const float* __restrict__ weights,
                                                              const float* __restrict__ actions, const float* __restrict__ dof_state,
                                                              float* hidden, float* cell, float* torques) {
    __shared__ float w[AN_NW];
    pf_grid_dep_launch();
    for (int i = threadIdx.x; i < AN_NW; i += blockDim.x) w[i] = weights[i];       // (constants: may load before the wait)
    pf_grid_dep_wait();
    __syncthreads();
    const long long n = (long long)P.num_envs * P.num_dof;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int d = (int)(i % P.num_dof);
    const float2 s = reinterpret_cast<const float2*>(dof_state)[i];                 // (dof_pos, dof_vel), anymal.py:66-67
    float x[2];
    x[0] = ((actions[i] * P.action_scale + P.default_dof_pos[d]) - s.x) * P.in_scale[0];
    x[1] = s.y * P.in_scale[1];
    float h0[AN_H], c0[AN_H], h1[AN_H], c1[AN_H];
    an_ld8(hidden + i * AN_H, h0); an_ld8(cell + i * AN_H, c0);
    an_ld8(hidden + (n + i) * AN_H, h1); an_ld8(cell + (n + i) * AN_H, c1);
    an_cell<2>(w + AN_WIH0, w + AN_WHH0, w + AN_BIH0, w + AN_BHH0, x, h0, c0);
    an_cell<AN_H>(w + AN_WIH1, w + AN_WHH1, w + AN_BIH1, w + AN_BHH1, h0, h1, c1);
    float t = w[AN_BLIN];
#pragma unroll
    for (int j = 0; j < AN_H; ++j) t = fmaf(w[AN_WLIN + j], h1[j], t);
    torques[i] = P.out_scale * t;
    an_st8(hidden + i * AN_H, h0); an_st8(cell + i * AN_H, c0);
    an_st8(hidden + (n + i) * AN_H, h1); an_st8(cell + (n + i) * AN_H, c1);
}

// anymal.py:52-56: sea_hidden_state_per_env[:, env_ids] = 0, sea_cell_state_per_env[:, env_ids] = 0.  One thread per
// (reset env, layer, DOF); the count is read on the device (no host sync): ids / count are what pf_post_physics_step leaves.
__global__ void __launch_bounds__(256) pf_actuator_net_reset_kernel(int num_envs, int num_dof, const int32_t* __restrict__ reset_ids,
                                                                    const int32_t* __restrict__ reset_count, float* hidden, float* cell) {
    pf_grid_dep_launch();
    pf_grid_dep_wait();
    const int cnt = min(*reset_count, num_envs);
    const long long n = (long long)num_envs * num_dof;
    for (long long u = (long long)blockIdx.x * blockDim.x + threadIdx.x; u < (long long)cnt * 2 * num_dof;
         u += (long long)gridDim.x * blockDim.x) {
        const int k = (int)(u / (2 * num_dof)), r = (int)(u % (2 * num_dof)), layer = r / num_dof, d = r % num_dof;
        const int e = reset_ids[k];
        if (e < 0 || e >= num_envs) continue;
        const long long o = (layer * n + (long long)e * num_dof + d) * AN_H;
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        reinterpret_cast<float4*>(hidden + o)[0] = z; reinterpret_cast<float4*>(hidden + o)[1] = z;
        reinterpret_cast<float4*>(cell + o)[0] = z; reinterpret_cast<float4*>(cell + o)[1] = z;
    }
}

static int an_fail(int code, const char* msg) { pf_set_error(msg); return code; }

extern "C" int pf_actuator_net_torques(const pf_actuator_net_params* p, const float* weights, const float* actions,
                                       const float* dof_state, float* hidden, float* cell, float* torques, void* stream) {
    if (!p || !weights || !actions || !dof_state || !hidden || !cell || !torques)
        return an_fail(1, "pf_actuator_net_torques: null argument");
    if (p->num_envs <= 0 || p->num_dof <= 0 || p->num_dof > PF_ACTUATOR_MAX_DOF)
        return an_fail(1, "pf_actuator_net_torques: num_envs / num_dof out of range");
    if (((reinterpret_cast<uintptr_t>(hidden) | reinterpret_cast<uintptr_t>(cell)) & 15) || (reinterpret_cast<uintptr_t>(dof_state) & 7))
        return an_fail(1, "pf_actuator_net_torques: hidden / cell must be 16-byte aligned, dof_state 8-byte aligned");
    const long long n = (long long)p->num_envs * p->num_dof;
    if (pf_launch_pdl(pf_actuator_net_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, (cudaStream_t)stream, *p, weights,
                      actions, dof_state, hidden, cell, torques) != cudaSuccess)
        return an_fail(2, "pf_actuator_net_kernel: launch failed");
    return 0;
}

extern "C" int pf_actuator_net_reset(int32_t num_envs, int32_t num_dof, const int32_t* reset_ids, const int32_t* reset_count,
                                     float* hidden, float* cell, void* stream) {
    if (!reset_ids || !reset_count || !hidden || !cell) return an_fail(1, "pf_actuator_net_reset: null argument");
    if (num_envs <= 0 || num_dof <= 0 || num_dof > PF_ACTUATOR_MAX_DOF) return an_fail(1, "pf_actuator_net_reset: size out of range");
    if (pf_launch_pdl(pf_actuator_net_reset_kernel, dim3(64), dim3(256), 0, (cudaStream_t)stream, (int)num_envs, (int)num_dof,
                      reset_ids, reset_count, hidden, cell) != cudaSuccess)
        return an_fail(2, "pf_actuator_net_reset_kernel: launch failed");
    return 0;
}
