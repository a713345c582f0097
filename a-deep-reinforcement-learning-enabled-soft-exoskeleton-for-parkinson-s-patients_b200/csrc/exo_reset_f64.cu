#define REAL double
// Instantiates the reset kernel for one arithmetic type (see exo_reset.cuh).
#include "exo_launch.h"
#include "exo_reset.cuh"

namespace exo {
template <typename R, int NAX>
static int launch_reset_nax(const EnvArrays<R>& A, const ChainConst<R>& C, const ResetConst& K, int n_sm,
                            const uint8_t* mask, float* obs, cudaStream_t s) {
    const size_t smem = RST_WARPS * reset_smem_per_warp(A.table_len);
    static size_t configured[64] = {0};           // per device: the attribute belongs to the device's copy of the kernel
    int dev = 0;
    cudaGetDevice(&dev);
    if (smem > configured[dev & 63]) {
        cudaError_t e = cudaFuncSetAttribute(exo_reset_kernel<R, NAX>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return (int)e;
        configured[dev & 63] = smem;
    }
    // persistent-style grid: as many blocks as fit (shared memory bound), each warp strides over the work list
    int per_sm = (int)((227 * 1024) / (smem + 1024));
    if (per_sm < 1) per_sm = 1;
    if (per_sm > 4) per_sm = 4;
    int grid = n_sm * per_sm;
    if (K.list_parity < 0) { const int need = (A.n + RST_WARPS - 1) / RST_WARPS; if (need < grid) grid = need; }
    // Auto-reset launches follow the step kernel in the same stream: programmatic dependent launch lets this grid be
    // scheduled while the step kernel drains (it signals at its start); the kernel waits with griddepcontrol.wait
    // before it touches the work list, so an empty list costs almost nothing on the device timeline.
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(RST_WARPS * 32); cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = (K.list_parity >= 0 && !K.motor_full) ? 1 : 0;   // FULL: the motor kernel sits between
    return (int)cudaLaunchKernelEx(&cfg, exo_reset_kernel<R, NAX>, A, C, K, mask, obs);
}
template <typename R>
int launch_reset(const EnvArrays<R>& A, const ChainConst<R>& C, const ResetConst& K, int nax, int n_sm,
                 const uint8_t* mask, float* obs, cudaStream_t s) {
    return nax == 4 ? launch_reset_nax<R, 4>(A, C, K, n_sm, mask, obs, s) : launch_reset_nax<R, 7>(A, C, K, n_sm, mask, obs, s);
}
template int launch_reset<REAL>(const EnvArrays<REAL>&, const ChainConst<REAL>&, const ResetConst&, int, int, const uint8_t*,
                                float*, cudaStream_t);
#ifdef EXO_WITH_SEED
void launch_seed(uint32_t* mt, int n, const uint32_t* seeds, cudaStream_t s) {
    exo_seed_kernel<<<(n + 31) / 32, 32, 0, s>>>(mt, n, seeds);
}
#endif
}  // namespace exo
