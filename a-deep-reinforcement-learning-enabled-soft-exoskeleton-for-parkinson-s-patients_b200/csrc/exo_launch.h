// exo_launch.h — host launchers, one translation unit per precision so the library builds in parallel.
#pragma once
#include <cuda_runtime.h>
#include "exo_types.cuh"

namespace exo {
template <typename R>
void launch_step(const EnvArrays<R>& A, const ChainConst<R>& C, const StepConst<R>& K, int nax, bool info_on,
                 const float* actions, float* obs, R* reward, uint8_t* done, const InfoPtrs<R>& ip, cudaStream_t s);
template <typename R>
void launch_reset(const EnvArrays<R>& A, const ChainConst<R>& C, const ResetConst& K, const StepConst<R>& SK, int nax,
                  const uint8_t* mask, float* obs, cudaStream_t s);   // 2 kernels: reset + propagator
template <typename R>
void launch_fk(const R* theta, R* p, int n, R L1, R L2, cudaStream_t s);
void launch_seed(uint32_t* mt, int n, int stride, const uint32_t* seeds, int ctor_draw, cudaStream_t s);
}  // namespace exo
