// exo_launch.h — host launchers, one translation unit per precision so the library builds in parallel.
#pragma once
#include <cuda_runtime.h>
#include "exo_types.cuh"

namespace exo {
template <typename R>
void launch_step(const EnvArrays<R>& A, const ChainConst<R>& C, const StepConst<R>& K, int nax, bool info_on,
                 const float* actions, float* obs, R* reward, uint8_t* done, const InfoPtrs<R>& ip, cudaStream_t s);
// EXO_MOTOR_FULL: the same step with the FULL template flag (targets handed to the motor kernel), then the general
// multibody step itself (exo_motor.cuh)
template <typename R>
void launch_step_full(const EnvArrays<R>& A, const ChainConst<R>& C, const StepConst<R>& K, int nax, bool info_on,
                      const float* actions, float* obs, R* reward, uint8_t* done, const InfoPtrs<R>& ip, cudaStream_t s);
template <typename R>
void launch_motor(const EnvArrays<R>& A, const ChainConst<R>& C, const DynConst<R>& D, int lag, int env0, int env1, cudaStream_t s);
// exo_gen_kernel: draws, tremor table, parameters and propagator of the NEXT episode into the spare slot of the envs of
// work list K.list (>= 0) or selected by mask (NULL: all). Returns a cudaError_t.
template <typename R>
int launch_gen(const EnvArrays<R>& A, const ResetConst& K, int nax, int n_sm, const uint8_t* mask, int expect, cudaStream_t s);
// exo_begin_kernel: start the prefetched episode (slot flip, counts = 2, first observation) of the envs of work list
// K.list (>= 0; `count_hint` bounds the grid) or selected by mask (NULL: all), appending them to work list K.append.
template <typename R>
void launch_begin(const EnvArrays<R>& A, const ChainConst<R>& C, const ResetConst& K, int nax, const uint8_t* mask, float* obs,
                  cudaStream_t s);
template <typename R>
void launch_fk(const R* theta, R* p, int n, R L1, R L2, cudaStream_t s);
// gather (set = 0) / scatter (set = 1) one field between the resident layout and an env-major [n, k] buffer;
// dtype: 0 real, 1 f64, 2 i32, 3 u32
template <typename R>
void launch_field(const EnvArrays<R>& A, int field, int k, int dtype, int naxv, int next, void* buf, int set, cudaStream_t s);
void launch_seed(uint32_t* mt, int n, const uint32_t* seeds, cudaStream_t s);
}  // namespace exo
