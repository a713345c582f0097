/* exo_policy.h — C-ABI of the device-resident policy forward pass.
 *
 * Replaces, for a batch of observations that already lives in HBM (the batched env's obs tensor), the network part of
 *   Agent.select_action            Agent/TD7_multi_agent_Pink_noise.py:208-228
 *     zs     = encoder.zs(state)   :101-105   zs1 80->300, ELU, zs2 300->300, ELU, zs3 300->300, AvgL1Norm (:59-60)
 *     action = actor(state, zs)    :79-84     l0 80->300, AvgL1Norm, cat[a, zs], l1 600->300, ReLU, l2 300->300, ReLU,
 *                                             l3 300->7, tanh
 *     action.clip(-1, 1) * max_action   :228
 * (exploration noise, :219-226, stays on the host side of the binding: it is a per-episode table lookup).
 * The dimensions are those of the shipped checkpoints and of Hyperparameters (:39-55): state 80, zs_dim = enc_hdim =
 * actor_hdim = 300, action 7; exo_policy_set_weights rejects anything else.
 *
 * One fused sm_100a kernel (csrc/exo_policy.cu): 128 observations per CTA, every layer a tcgen05.mma (kind::tf32,
 * fp32 accumulators in tensor memory) whose weights stream from L2 through cp.async.bulk stages, activations never
 * leave the SM between layers. Arithmetic: TF32 products (inputs rounded to nearest, 10-bit mantissa), fp32
 * accumulation, fp32 bias/activation - the arithmetic of torch with allow_tf32, not of its default fp32 matmul;
 * the parity tolerance this implies is stated in tests/test_policy.py.
 *
 * Conventions as in exo_env.h: plain C types, device pointers owned by the caller, kernels launch on the caller's
 * stream, int status + exo_last_error(), no host synchronisation inside exo_policy_forward.
 */
#ifndef EXO_POLICY_H
#define EXO_POLICY_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define EXO_POLICY_STATE_DIM 80
#define EXO_POLICY_HIDDEN_DIM 300
#define EXO_POLICY_ACTION_DIM 7

typedef struct ExoPolicy ExoPolicy;

/* torch.nn.Linear parameters as the reference's state dicts hold them (weight [out, in] row-major float32, bias [out]),
 * HOST pointers. Encoder (:87-99): zs1, zs2, zs3. Actor (:66-76): l0, l1, l2, l3. */
typedef struct ExoPolicyWeights {
    const float *zs1_w, *zs1_b;   /* [300, 80],  [300] */
    const float *zs2_w, *zs2_b;   /* [300, 300], [300] */
    const float *zs3_w, *zs3_b;   /* [300, 300], [300] */
    const float *l0_w, *l0_b;     /* [300, 80],  [300] */
    const float *l1_w, *l1_b;     /* [300, 600], [300]  (columns 0..299 multiply AvgL1Norm(l0(state)), 300..599 zs) */
    const float *l2_w, *l2_b;     /* [300, 300], [300] */
    const float *l3_w, *l3_b;     /* [7, 300],   [7]   */
} ExoPolicyWeights;

/* Fails unless `device` is an sm_100 GPU (there is no CPU path). Environment EXO_POLICY_PAIR=1 selects the CTA-pair
 * (cta_group::2) variant of the kernel for this handle; results are bit-identical. */
int exo_policy_create(int32_t device, ExoPolicy** out);
int exo_policy_destroy(ExoPolicy* p);

/* Agent.load (:349-366) for the two networks select_action uses: packs the matrices into the tensor-core operand
 * layout (TF32, round to nearest) and uploads them; synchronous. state_dim/hidden_dim/action_dim must be 80/300/7. */
int exo_policy_set_weights(ExoPolicy* p, const ExoPolicyWeights* w, int32_t state_dim, int32_t hidden_dim, int32_t action_dim);

/* actions[n,7] = clip(actor(obs, encoder.zs(obs)), -1, 1) * max_action for n observations [n,80] (device, float32,
 * row-major). If zs_dev is not NULL it also receives encoder.zs(obs) [n,300] (tests / the learner). */
int exo_policy_forward(ExoPolicy* p, const float* obs_dev, int64_t n, float max_action, float* actions_dev, float* zs_dev,
                       void* stream);

/* The same forward pass in the reference's own arithmetic - float32 products and sums on the FP32 pipe, as torch's
 * default matmul (csrc/exo_policy_fp32.cu) - for runs that must reproduce the reference's actions to summation order.
 * About 8x slower than the tensor-core kernel (1.8 ms per 65 536 observations). Same arguments as exo_policy_forward. */
int exo_policy_forward_fp32(ExoPolicy* p, const float* obs_dev, int64_t n, float max_action, float* actions_dev, float* zs_dev,
                            void* stream);

/* exo_policy_forward_fp32 for callers that cannot allocate device memory (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc). NO COUNTERPART IN THE REFERENCE.
 * Same kernel, arguments and results (bit-identical) as exo_policy_forward_fp32, but the kernel's per-CTA scratch lives in a
 * file-scope `static __device__` array in csrc/exo_policy_fp32.cu - 64 CTAs x 77 824 B = 4.75 MiB that every GPU which loads
 * the library holds until exit - and the call performs no cudaMalloc / cudaFree / allocating library call: it checks its
 * arguments and launches one kernel on `stream`. (The handle - weights, the array's device address - is made beforehand by
 * exo_policy_create / exo_policy_set_weights, which do allocate.)
 *   limit   n <= EXO_POLICY_NOALLOC_MAX_OBS observations per call; larger n fails with EXO_ERR_ARG and a message naming the
 *           limit (nothing is launched).
 *   sharing every call on a device uses the same array, which keeps the previous call's contents. Nothing needs clearing
 *           (the kernel overwrites a CTA's slice before reading it), but calls that OVERLAP on the GPU - two streams, or
 *           two handles on one device - race on it: serialise them (one stream, or events). */
#define EXO_POLICY_NOALLOC_MAX_OBS 4096
int exo_policy_forward_fp32_noalloc(ExoPolicy* p, const float* obs_dev, int64_t n, float max_action, float* actions_dev,
                                    float* zs_dev, void* stream);

/* One forward pass (max_action 1) that also records the SM clock (clock64) of CTA 0's first tile; synchronises.
 * clocks_host[32]: [2*g], [2*g+1] = MMA issuer starts / has issued GEMM phase g (0..7: l0, zs1, zs2, zs3, l1-zs half,
 * l1-a0 half, l2, l3); [16+2*k], [17+2*k] = epilogue k (0 = observation load, 1..7 after each accumulator) begins /
 * ends. A tuning aid: tools/policy_probe.py prints the timeline. */
int exo_policy_profile(ExoPolicy* p, const float* obs_dev, int64_t n, float* actions_dev, int64_t* clocks_host, void* stream);

/* kernels launched so far through this handle */
int64_t exo_policy_launch_count(const ExoPolicy* p);

#ifdef __cplusplus
}
#endif
#endif
