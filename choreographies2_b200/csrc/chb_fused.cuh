// chb_fused.cuh -- placeholder interface of the fused action+gradient kernel (filled in later).
#pragma once
#include <vector>
#include "chb_core.cuh"

struct ChbFusedPlan {
    int active = 0;
};

// returns 1 if the fused kernel can serve this level, 0 if not, <0 on CUDA error
static inline int chb_fused_setup(ChbFusedPlan &F, const ChbSys &S, const std::vector<ChbLoop> &loops, int G, int device,
                                  size_t smem_limit, int num_sms, std::vector<void *> &level_allocs) {
    (void)S; (void)loops; (void)G; (void)device; (void)smem_limit; (void)num_sms; (void)level_allocs;
    F.active = 0;
    return 0;
}
static inline int chb_fused_action_grad(ChbFusedPlan &F, const ChbSys &S, int64_t batch, const double *params,
                                        double *action, double *grad, cudaStream_t st) {
    (void)F; (void)S; (void)batch; (void)params; (void)action; (void)grad; (void)st;
    return -1;
}
