#define REAL float
// Instantiates the general multibody step (exo_motor.cuh) and the FULL variant of the step kernel for one arithmetic type.
#include "exo_launch.h"
#include "exo_reset.cuh"       // begin_episode
#include "exo_motor.cuh"

namespace exo {
template <typename R>
void launch_step_full(const EnvArrays<R>& A, const ChainConst<R>& C, const StepConst<R>& K, int nax, bool info_on,
                      const float* actions, float* obs, R* reward, uint8_t* done, const InfoPtrs<R>& ip, cudaStream_t s) {
    const int grid = (K.env1 - K.env0 + EXO_STEP_THREADS - 1) / EXO_STEP_THREADS;
    if (nax == 4) {
        if (info_on) exo_step_kernel<R, true, 4, true><<<grid, EXO_STEP_THREADS, 0, s>>>(A, C, K, actions, obs, reward, done, ip);
        else exo_step_kernel<R, false, 4, true><<<grid, EXO_STEP_THREADS, 0, s>>>(A, C, K, actions, obs, reward, done, ip);
    } else {
        if (info_on) exo_step_kernel<R, true, 7, true><<<grid, EXO_STEP_THREADS, 0, s>>>(A, C, K, actions, obs, reward, done, ip);
        else exo_step_kernel<R, false, 7, true><<<grid, EXO_STEP_THREADS, 0, s>>>(A, C, K, actions, obs, reward, done, ip);
    }
}
template <typename R>
void launch_motor(const EnvArrays<R>& A, const ChainConst<R>& C, const DynConst<R>& D, int lag, int env0, int env1, cudaStream_t s) {
    const int grid = (env1 - env0 + EXO_MOTOR_THREADS - 1) / EXO_MOTOR_THREADS;
    const size_t smem = (size_t)(EXO_NTRI + EXO_NDOF) * EXO_MOTOR_THREADS * sizeof(R);
    static bool configured[64] = {false};
    int dev = 0;
    cudaGetDevice(&dev);
    if (!configured[dev & 63]) {
        cudaFuncSetAttribute(exo_motor_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        configured[dev & 63] = true;
    }
    exo_motor_kernel<R><<<grid, EXO_MOTOR_THREADS, smem, s>>>(A, C, D, lag, env0, env1);
}
template void launch_step_full<REAL>(const EnvArrays<REAL>&, const ChainConst<REAL>&, const StepConst<REAL>&, int, bool, const float*,
                                     float*, REAL*, uint8_t*, const InfoPtrs<REAL>&, cudaStream_t);
template void launch_motor<REAL>(const EnvArrays<REAL>&, const ChainConst<REAL>&, const DynConst<REAL>&, int, int, int, cudaStream_t);
}  // namespace exo
