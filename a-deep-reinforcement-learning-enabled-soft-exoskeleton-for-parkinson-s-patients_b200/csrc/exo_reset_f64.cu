#define REAL double
// Instantiates the reset kernel for one arithmetic type (see exo_reset.cuh).
#include <stdlib.h>
#include "exo_launch.h"
#include "exo_reset.cuh"

namespace exo {
template <typename R, int NAX>
static int launch_gen_nax(const EnvArrays<R>& A, const ResetConst& K, int n_sm, const uint8_t* mask, int expect, cudaStream_t s) {
    const size_t smem = RST_WARPS * reset_smem_per_warp(A.table_len);
    static size_t configured[64] = {0};           // per device: the attribute belongs to the device's copy of the kernel
    int dev = 0;
    cudaGetDevice(&dev);
    if (smem > configured[dev & 63]) {
        cudaError_t e = cudaFuncSetAttribute(exo_gen_kernel<R, NAX>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return (int)e;
        configured[dev & 63] = smem;
    }
    // persistent-style grid: as many blocks as fit (shared memory bound), each warp strides over the work list.
    // Work-list passes run BESIDE the step kernel (side stream): one-warp blocks (4 k registers, 12.5 KB of shared memory)
    // slot into the space the step kernel's 7 blocks per SM leave free, so the step stays a single wave; a four-warp block
    // would take the place of one step block on its SM and push that block into a second wave.
    const bool small_list = K.list >= 0 && expect < 4096;      // `expect`: the host's guess of the list length (a step's finished envs vs reset() of everything)
    int warps = small_list ? 1 : RST_WARPS;
    if (const char* e = getenv("EXO_GEN_WARPS")) { const int w = atoi(e); if (w >= 1 && w <= RST_WARPS) warps = w; }
    int per_sm = (int)((227 * 1024) / (smem + 1024));
    if (per_sm < 1) per_sm = 1;
    if (per_sm > 4) per_sm = 4;
    int grid = n_sm * per_sm * (RST_WARPS / warps);
    if (!small_list) { const int need = (A.n + warps - 1) / warps; if (need < grid) grid = need; }
    else {
        grid = n_sm * 2;                          // a work list holds ~n / 280 envs at steady state; two one-warp blocks per SM walk it
        if (const char* e = getenv("EXO_GEN_GRID")) { const int g = atoi(e); if (g >= 1) grid = g; }
    }
    exo_gen_kernel<R, NAX><<<grid, warps * 32, (size_t)warps * reset_smem_per_warp(A.table_len), s>>>(A, K, mask);
    return (int)cudaGetLastError();
}
template <typename R>
int launch_gen(const EnvArrays<R>& A, const ResetConst& K, int nax, int n_sm, const uint8_t* mask, int expect, cudaStream_t s) {
    return nax == 4 ? launch_gen_nax<R, 4>(A, K, n_sm, mask, expect, s) : launch_gen_nax<R, 7>(A, K, n_sm, mask, expect, s);
}
template <typename R>
void launch_begin(const EnvArrays<R>& A, const ChainConst<R>& C, const ResetConst& K, int nax, const uint8_t* mask, float* obs,
                  cudaStream_t s) {
    const int grid = (A.n + 63) / 64;             // list mode: threads past the list length idle
    if (nax == 4) exo_begin_kernel<R, 4><<<grid, 64, 0, s>>>(A, C, K, mask, obs);
    else exo_begin_kernel<R, 7><<<grid, 64, 0, s>>>(A, C, K, mask, obs);
}
template int launch_gen<REAL>(const EnvArrays<REAL>&, const ResetConst&, int, int, const uint8_t*, int, cudaStream_t);
template void launch_begin<REAL>(const EnvArrays<REAL>&, const ChainConst<REAL>&, const ResetConst&, int, const uint8_t*, float*, cudaStream_t);
#ifdef EXO_WITH_SEED
void launch_seed(uint32_t* mt, int n, const uint32_t* seeds, cudaStream_t s) {
    exo_seed_kernel<<<(n + 31) / 32, 32, 0, s>>>(mt, n, seeds);
}
#endif
}  // namespace exo
