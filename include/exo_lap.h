/* exo_lap.h — C-ABI of the B200-resident prioritized replay buffer that consumes the batched env's outputs
 * without leaving HBM. Part of libexo_env.so.
 *
 * Replaces the reference's `LAP` (Agent/TD7_buffer_multi_agent.py:5-150): per-env ("tremor_num") rows of
 * transitions with one shared write pointer, loss-adjusted prioritized sampling (cumsum + searchsorted per row),
 * priority updates and the running max priority. Semantics kept, quirks included:
 *   - add (:49-63): writes slot `ptr` of row `tremor_num`, priority = max_priority; the pointer (and size) advance
 *     only on adds whose running `count` is a multiple of num_envs (:59-61), count increments on every add;
 *   - sample (:65-111): per row, ind = searchsorted(cumsum(priority[row, :size]), u * total) for batch_size
 *     uniforms u in [0,1), gathered into [num_envs*batch_size, dim] float32 arrays, row-major by env;
 *   - update_priority (:113-117): priority[row, ind] = new priority, max_priority = max(old, max(new));
 *   - reset_max_priority (:119-120): max over the WHOLE priority array; reset_buffer (:122-139).
 * Storage is float32 (the reference stores float64 and casts to float32 in sample(); its inputs are float32
 * observations, so the sampled values are identical). ptr / count / size / max_priority live on the device:
 * no call synchronises the stream except exo_lap_counters.
 *
 * Conventions as in exo_env.h: plain C types, device pointers owned by the caller, int status + exo_last_error().
 */
#ifndef EXO_LAP_H
#define EXO_LAP_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ExoLap ExoLap;

/* LAP.__init__ (:6-47). max_action / normalize_actions as in the reference (actions are stored divided by
 * max_action when normalize_actions != 0). Fails unless `device` is an sm_100 GPU. */
int exo_lap_create(int32_t state_dim, int32_t action_dim, int32_t num_envs, int64_t max_size, int32_t batch_size,
                   double max_action, int32_t normalize_actions, int32_t device, ExoLap** out);
int exo_lap_destroy(ExoLap* b);

/* k calls of LAP.add (:49-63) in index order, on the device: entry j is (state[j], action[j], next_state[j],
 * reward[j], done[j]) for row tremor_num[j] (NULL: j % num_envs); entries with skip[j] != 0 are not added (the
 * training loop's `if not done[i]`, Exoskeleton_agent_train.py:139). When several entries of one call land on the
 * same (row, slot) the last one wins, as it would sequentially. reward is float32. */
int exo_lap_add(ExoLap* b, const float* state_dev /*[k,sd]*/, const float* action_dev /*[k,ad]*/,
                const float* next_state_dev /*[k,sd]*/, const float* reward_dev /*[k]*/, const uint8_t* done_dev /*[k]*/,
                const int32_t* tremor_num_dev /*[k] or NULL*/, const uint8_t* skip_dev /*[k] or NULL*/, int32_t k, void* stream);

/* LAP.sample (:65-111), prioritized branch. uniform_dev: num_envs*batch_size uniforms in [0,1), row-major by env
 * (the reference draws torch.rand(batch_size) once per env, in env order). Outputs are float32, [num_envs*batch_size, dim].
 * The sampled indices are kept for exo_lap_update_priority and also written to ind_dev if it is not NULL. */
int exo_lap_sample(ExoLap* b, const float* uniform_dev, float* state_dev, float* action_dev, float* next_state_dev,
                   float* reward_dev, float* not_done_dev, int32_t* ind_dev, void* stream);

/* LAP.update_priority (:113-117): priority_dev float32 [num_envs*batch_size] for the indices of the last sample. */
int exo_lap_update_priority(ExoLap* b, const float* priority_dev, void* stream);
/* LAP.reset_max_priority (:119-120) and LAP.reset_buffer (:122-139). */
int exo_lap_reset_max_priority(ExoLap* b, void* stream);
int exo_lap_reset_buffer(ExoLap* b, void* stream);

/* ptr, count, size (int64) and max_priority (as double) -> host; synchronises `stream`. */
int exo_lap_counters(ExoLap* b, int64_t* ptr, int64_t* count, int64_t* size, double* max_priority, void* stream);
/* Device views for tests / checkpoints: priority float32 [num_envs, max_size]; field 0 state, 1 action, 2 next_state,
 * 3 reward, 4 not_done, 5 priority. Returns the device pointer and the row pitch in elements. */
int exo_lap_storage(ExoLap* b, int32_t field, void** dev_ptr, int64_t* elems_per_slot);
/* kernels launched so far by this buffer */
int64_t exo_lap_launch_count(const ExoLap* b);

#ifdef __cplusplus
}
#endif
#endif
