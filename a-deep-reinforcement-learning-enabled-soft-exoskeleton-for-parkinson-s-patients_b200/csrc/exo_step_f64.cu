#define REAL double
// Instantiates the step kernel for one arithmetic type (see exo_step.cuh).
#include "exo_launch.h"
#include "exo_reset.cuh"       // begin_episode (the auto-reset path of the step kernel) + exo_step.cuh

namespace exo {
template <typename R>
__global__ void exo_fk_kernel(const R* __restrict__ theta, R* __restrict__ p, int n, R L1, R L2) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    R th[7], o[3];
#pragma unroll
    for (int i = 0; i < 7; i++) th[i] = theta[(size_t)e * 7 + i];
    dh_end_effector(th, L1, L2, o);
    p[(size_t)e * 3] = o[0]; p[(size_t)e * 3 + 1] = o[1]; p[(size_t)e * 3 + 2] = o[2];
}

template <typename R>
void launch_step(const EnvArrays<R>& A, const ChainConst<R>& C, const StepConst<R>& K, int nax, bool info_on,
                 const float* actions, float* obs, R* reward, uint8_t* done, const InfoPtrs<R>& ip, cudaStream_t s) {
    const int grid = (K.env1 - K.env0 + EXO_STEP_THREADS - 1) / EXO_STEP_THREADS;
    if (nax == 4) {
        if (info_on) exo_step_kernel<R, true, 4><<<grid, EXO_STEP_THREADS, 0, s>>>(A, C, K, actions, obs, reward, done, ip);
        else exo_step_kernel<R, false, 4><<<grid, EXO_STEP_THREADS, 0, s>>>(A, C, K, actions, obs, reward, done, ip);
    } else {
        if (info_on) exo_step_kernel<R, true, 7><<<grid, EXO_STEP_THREADS, 0, s>>>(A, C, K, actions, obs, reward, done, ip);
        else exo_step_kernel<R, false, 7><<<grid, EXO_STEP_THREADS, 0, s>>>(A, C, K, actions, obs, reward, done, ip);
    }
}
template <typename R>
void launch_fk(const R* theta, R* p, int n, R L1, R L2, cudaStream_t s) {
    exo_fk_kernel<R><<<(n + 127) / 128, 128, 0, s>>>(theta, p, n, L1, L2);
}
template <typename R>
void launch_field(const EnvArrays<R>& A, int field, int k, int dtype, int naxv, int next, void* buf, int set, cudaStream_t s) {
    const size_t total = (size_t)A.n * k;
    const int grid = (int)((total + 255) / 256);
#define EXO_FIELD_GO(T)                                                                                  \
    do {                                                                                                 \
        if (set) exo_field_kernel<R, T, true><<<grid, 256, 0, s>>>(A, field, k, naxv, next, (T*)buf);          \
        else exo_field_kernel<R, T, false><<<grid, 256, 0, s>>>(A, field, k, naxv, next, (T*)buf);             \
    } while (0)
    if (dtype == 0) EXO_FIELD_GO(R);
    else if (dtype == 1) EXO_FIELD_GO(double);
    else if (dtype == 2) EXO_FIELD_GO(int);
    else EXO_FIELD_GO(uint32_t);
#undef EXO_FIELD_GO
}
template void launch_step<REAL>(const EnvArrays<REAL>&, const ChainConst<REAL>&, const StepConst<REAL>&, int, bool, const float*,
                                float*, REAL*, uint8_t*, const InfoPtrs<REAL>&, cudaStream_t);
template void launch_fk<REAL>(const REAL*, REAL*, int, REAL, REAL, cudaStream_t);
template void launch_field<REAL>(const EnvArrays<REAL>&, int, int, int, int, int, void*, int, cudaStream_t);
}  // namespace exo
