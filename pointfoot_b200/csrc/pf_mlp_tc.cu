// placeholder until the tcgen05 path lands (next commit): PF_RL_BF16_TC contexts fail loudly
#include "pf_rl_internal.h"
void pf_set_error(const char* msg);
int pf_tc_create(pf_rl_ctx*) { pf_set_error("pf_rl_create: tensor-core path not built"); return 6; }
void pf_tc_destroy(pf_rl_ctx*) {}
int pf_tc_pack_weights(pf_rl_ctx*, const float*, cudaStream_t) { return 6; }
int pf_tc_mlp_forward(pf_rl_ctx*, int, const float*, const float*, long long, int, float*, float*, cudaStream_t) { return 6; }
