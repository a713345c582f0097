// Internal interface between exo_policy.cu (the C-ABI of include/exo_policy.h) and exo_policy_fp32.cu (the float32 path).
#ifndef EXO_POLICY_FP32_H
#define EXO_POLICY_FP32_H
#include "../../include/exo_policy.h"

struct ExoPolicyF32;
int exo_policy_f32_create(int device, int n_sm, ExoPolicyF32** out);
void exo_policy_f32_destroy(ExoPolicyF32* p);
int exo_policy_f32_set_weights(ExoPolicyF32* p, const ExoPolicyWeights* w);
int exo_policy_f32_forward(ExoPolicyF32* p, const float* obs_dev, long long n, float max_action, float* actions_dev, float* zs_dev, void* stream);
int exo_policy_f32_forward_noalloc(ExoPolicyF32* p, const float* obs_dev, long long n, float max_action, float* actions_dev, float* zs_dev, void* stream);
#endif
