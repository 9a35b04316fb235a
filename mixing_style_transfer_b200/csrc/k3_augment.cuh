#pragma once
#include "plan.cuh"
namespace mixfe {
inline int run_augment(mixfe_plan*, const float* const*, float* const*, long long, long long, int, int,
                       const mixfe_aug_params*, const float*, int, cudaStream_t) {
    return fail(MIXFE_EUNSUPPORTED, "augmentation kernels not built yet");
}
}
