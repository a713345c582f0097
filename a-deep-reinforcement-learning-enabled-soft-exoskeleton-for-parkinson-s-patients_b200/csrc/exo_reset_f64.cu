#define REAL double
// Instantiates the reset + propagator kernels for one arithmetic type (see exo_reset.cuh).
#include "exo_launch.h"
#include "exo_reset.cuh"

namespace exo {
template <typename R>
void launch_reset(const EnvArrays<R>& A, const ChainConst<R>& C, const ResetConst& K, const StepConst<R>& SK, int nax,
                  const uint8_t* mask, float* obs, cudaStream_t s) {
    const int n = A.n;
    if (nax == 4) exo_reset_kernel<R, 4><<<(n + 63) / 64, 64, 0, s>>>(A, C, K, SK, mask, obs);
    else exo_reset_kernel<R, 7><<<(n + 63) / 64, 64, 0, s>>>(A, C, K, SK, mask, obs);
    exo_propagator_kernel<R><<<(n + PW_WARPS - 1) / PW_WARPS, PW_WARPS * 32, 0, s>>>(A, K.dt, mask);
}
template void launch_reset<REAL>(const EnvArrays<REAL>&, const ChainConst<REAL>&, const ResetConst&, const StepConst<REAL>&, int,
                                 const uint8_t*, float*, cudaStream_t);
#ifdef EXO_WITH_SEED
void launch_seed(uint32_t* mt, int n, int stride, const uint32_t* seeds, int ctor_draw, cudaStream_t s) {
    exo_seed_kernel<<<(n + 127) / 128, 128, 0, s>>>(mt, n, stride, seeds, ctor_draw);
}
#endif
}  // namespace exo
