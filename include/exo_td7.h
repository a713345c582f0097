/* exo_td7.h — C-ABI of the device-resident TD7 update (second half).
 *
 * Replaces, for a batch that already lives in HBM (exo_lap_sample's output), one iteration of
 *   Agent.train                    Agent/TD7_multi_agent_Pink_noise.py:230-312
 *     encoder update               :238-246   mse(encoder.zsa(encoder.zs(s), a), encoder.zs(s')) -> Adam
 *     critic update                :251-276   target from (fixed_encoder_target, actor_target, critic_target) with
 *                                             clipped policy noise and value clipping, LAP_huber(|Q - Q_target|) -> Adam
 *     priorities                   :281-282   max_h |td| .clamp(min_priority) ** alpha  (for exo_lap_update_priority)
 *     actor update                 :287-298   every policy_freq iterations: -critic(s, actor(s, zs), ...).mean() -> Adam
 *     roll-over                    :303-312   every target_update_rate iterations: actor_target <- actor,
 *                                             critic_target <- critic, fixed_encoder_target <- fixed_encoder,
 *                                             fixed_encoder <- encoder, max_target/min_target <- max/min
 * and holds the nine networks Agent owns (:151-163) plus the three Adam states (:152,156,160: lr, weight_decay 1e-7,
 * torch's default betas/eps) in device memory. With cfg.offline the actor loss carries the TD3+BC term (:292-293).
 *
 * Arithmetic (csrc/exo_td7.cu): float32 storage, sums and element-wise math; the matrix products run on the tensor
 * cores as 3xTF32 (each operand split into two TF32 parts, three MMAs, float32 accumulation: every product within
 * 2^-21 of the float32 one), or - environment EXO_TD7_FFMA=1 at exo_td7_create - as float32 FFMA on the FP32 pipe, the
 * arithmetic of torch's default matmul. Gradients are hand-derived; weight gradients are summed over the batch with
 * split-K atomics, so two runs agree to summation order, not bit for bit. Tolerances: tests/test_td7.py.
 *
 * Conventions as in exo_env.h: plain C types, batch pointers are device pointers owned by the caller, kernels launch on
 * the caller's stream, int status + exo_last_error(). Layer sizes are arguments (the reference's Hyperparameters), not
 * constants.
 */
#ifndef EXO_TD7_H
#define EXO_TD7_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ExoTD7 ExoTD7;

/* Hyperparameters (:12-55) + the sizes Agent.__init__ passes to the networks. batch_rows = env_num * batch_size, the
 * row count of LAP.sample's tensors. policy_noise_decrease = target_policy_noise / learning_steps (:191). */
typedef struct ExoTD7Config {
    int32_t state_dim, action_dim, zs_dim, enc_hdim, critic_hdim, actor_hdim;
    int32_t batch_rows;
    int32_t target_update_rate, policy_freq;
    float discount, target_policy_noise, noise_clip, policy_noise_decrease;
    float alpha, min_priority;
    float actor_lr, critic_lr, encoder_lr, weight_decay;
    uint64_t seed;                 /* of the on-device normal generator used when exo_td7_train gets randn_dev = NULL */
    int32_t offline;               /* Agent(offline=True): the TD3+BC term of the actor loss (:292-293) */
    float lmbda;                   /* its weight, hp.lmbda */
} ExoTD7Config;

/* network ids */
#define EXO_TD7_ACTOR 0
#define EXO_TD7_ACTOR_TARGET 1
#define EXO_TD7_CHECKPOINT_ACTOR 2
#define EXO_TD7_CRITIC 3
#define EXO_TD7_CRITIC_TARGET 4
#define EXO_TD7_ENCODER 5
#define EXO_TD7_FIXED_ENCODER 6
#define EXO_TD7_FIXED_ENCODER_TARGET 7
#define EXO_TD7_CHECKPOINT_ENCODER 8
#define EXO_TD7_N_NETS 9
/* optimiser ids (the network each one steps: actor, critic, encoder) and the per-parameter vectors they hold */
#define EXO_TD7_OPT_ACTOR 0
#define EXO_TD7_OPT_CRITIC 1
#define EXO_TD7_OPT_ENCODER 2
#define EXO_TD7_EXP_AVG 0
#define EXO_TD7_EXP_AVG_SQ 1
#define EXO_TD7_GRAD 2            /* the gradient of the last backward pass (read-only use: tests) */

/* Fails unless `device` is an sm_100 GPU (there is no CPU path). Networks start zeroed: load them with exo_td7_set_net. */
int exo_td7_create(int32_t device, const ExoTD7Config* cfg, ExoTD7** out);
int exo_td7_destroy(ExoTD7* t);

/* A network as ONE flat float32 vector: its state dict in torch's order, each nn.Linear as weight [out,in] row-major
 * then bias [out]. Actor: l0 l1 l2 l3 (:70-76); critic: q01 q1 q2 q3 q02 q4 q5 q6 (:118-126); encoder: zs1 zs2 zs3 zsa1
 * zsa2 zsa3 (:92-99). exo_td7_net_size gives the length. set/get take HOST pointers and synchronise. */
int64_t exo_td7_net_size(const ExoTD7* t, int32_t net);
int exo_td7_set_net(ExoTD7* t, int32_t net, const float* host, int64_t n);
int exo_td7_get_net(ExoTD7* t, int32_t net, float* host, int64_t n);
/* dst <- src on the device (load_state_dict between two of Agent's networks, e.g. :325-326 checkpoint_actor <- actor). */
int exo_td7_copy_net(ExoTD7* t, int32_t dst, int32_t src);

/* Adam state in the same flat layout (Agent.save/load's *_optimizer files, :349-362). `step` = the optimiser's step count. */
int exo_td7_set_opt(ExoTD7* t, int32_t opt, int32_t which, const float* host, int64_t n, int64_t step);
int exo_td7_get_opt(ExoTD7* t, int32_t opt, int32_t which, float* host, int64_t n, int64_t* step);

/* One Agent.train() iteration on a sampled batch: state / next_state [batch_rows, state_dim], action [batch_rows,
 * action_dim], reward / not_done [batch_rows] (device, float32, row-major). randn_dev [batch_rows, action_dim] are the
 * standard-normal draws behind the target-policy noise (:254) or NULL to draw them on the device. priority_dev
 * [batch_rows] receives the new priorities. *rolled_over (host, may be NULL) is set to 1 when this iteration rolled the
 * target networks over, i.e. when the caller must also call replay_buffer.reset_max_priority() (:309). Asynchronous. The
 * launch sequence of an iteration is captured into a CUDA graph the first time a set of batch pointers is seen (up to 32
 * sets; exo_lap_sample always returns the same buffers) and replayed afterwards; EXO_TD7_GRAPH=0 at exo_td7_create or a
 * gradient hook select plain stream launches. */
int exo_td7_train(ExoTD7* t, const float* state_dev, const float* action_dev, const float* next_state_dev,
                  const float* reward_dev, const float* not_done_dev, const float* randn_dev, float* priority_dev,
                  int32_t* rolled_over, void* stream);

/* Data-parallel learners (one process per GPU, each with its own replay-buffer shard and batch): `hook` is called on the
 * host, inside exo_td7_train, each time the gradient of one optimiser's network is complete on `stream` and before its
 * Adam step is enqueued. The hook enqueues the exchange on that stream - an all-reduce (mean) of grad_dev[n] over the
 * ranks (NCCL) - so that every rank takes the same step: N ranks x batch_rows rows behave as one batch of N x batch_rows
 * (the three losses are means over rows). NULL removes the hook. The value bounds stay per rank; callers that want the
 * single-batch semantics reduce max / min over the ranks when an iteration rolls over (exo_td7_scalars / _set_scalars). */
typedef void (*ExoTD7GradHook)(int32_t opt, float* grad_dev, int64_t n, void* stream, void* user);
int exo_td7_set_grad_hook(ExoTD7* t, ExoTD7GradHook hook, void* user);

/* Synchronises and reads the learner's scalars: out[0..3] = max, min, max_target, min_target (:255-259,311-312),
 * [4] = target_policy_noise, [5] = training_steps, [6..8] = encoder / critic / actor loss of the last iteration that
 * computed them. */
int exo_td7_scalars(ExoTD7* t, double* out9);
/* Agent.load's counterpart for those scalars (resuming a run). */
int exo_td7_set_scalars(ExoTD7* t, const double* in6);

/* The generator exo_td7_train uses when randn_dev is NULL, exposed for tests: fills out_dev[count] with the standard
 * normals of iteration `step` (a counter hash of (seed, step, index) + Box-Muller; not torch's stream). */
int exo_td7_randn(ExoTD7* t, float* out_dev, int64_t count, int64_t step, void* stream);

/* kernels launched so far through this handle */
int64_t exo_td7_launch_count(const ExoTD7* t);

#ifdef __cplusplus
}
#endif
#endif
