// exo_policy_fp32.cu — the policy forward pass of include/exo_policy.h in the reference's own arithmetic: float32
// products and sums on the FP32 pipe (torch's default matmul), for runs that must reproduce the reference's actions to
// summation order (evaluation replays, parity tests). The fast path is the TF32 tensor-core kernel of exo_policy.cu.
//
//   encoder.zs + actor of Agent/TD7_multi_agent_Pink_noise.py:59-105 as select_action evaluates them (:208-228).
//
// One fused kernel, 64 observations per CTA tile, 512 threads: warp w owns rows 8(w&7)..8(w&7)+7 of the tile and the column
// half w>>3 (0..159 / 160..319), lane l the columns l, l+32, ..., l+128 of that half (40 accumulators per thread; four warps
// per scheduler - the first version, 8 warps with 80 accumulators each, left the FMA pipe at 47 %). Activations ping-pong between two 64 x 304 shared-memory buffers;
// every layer streams its weights (host-transposed to [K][320], zero padded) from L2 in 16-row chunks through a
// cp.async double buffer. cat[a0, zs] * l1' is two accumulating passes; a0 waits in an L2-resident per-CTA scratch.
// FP32-pipe bound: 7 shared-memory loads per 40 FFMA in the inner loop.
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>
#include <mutex>
#include <vector>

#include "../../include/exo_env.h"
#include "../../include/exo_policy.h"
#include "exo_policy_fp32.h"

int exo_set_error(int code, const char* fmt, ...);     // exo_api.cu

namespace {

constexpr int SD = EXO_POLICY_STATE_DIM, HD = EXO_POLICY_HIDDEN_DIM, AD = EXO_POLICY_ACTION_DIM;
constexpr int HP = 304;          // activation row pitch / K of a hidden layer (300 padded to a multiple of the 16-row chunk)
constexpr int WP = 320;          // weight row pitch: 10 columns per lane
constexpr int TM = 64, KC = 16, NT = 512, NC = WP / 64, HALF = WP / 2;
constexpr int BUF_FLOATS = TM * HP;
constexpr int SMEM_BYTES = (2 * BUF_FLOATS + 2 * KC * WP) * 4;      // 196 608

// transposed weight blocks in the order of use, each [K_pad][WP]; biases [7][WP]
constexpr int K_OBS = SD, K_HID = HP;
constexpr size_t OFF_L0 = 0, OFF_ZS1 = OFF_L0 + (size_t)K_OBS * WP, OFF_ZS2 = OFF_ZS1 + (size_t)K_OBS * WP,
                 OFF_ZS3 = OFF_ZS2 + (size_t)K_HID * WP, OFF_L1B = OFF_ZS3 + (size_t)K_HID * WP, OFF_L1A = OFF_L1B + (size_t)K_HID * WP,
                 OFF_L2 = OFF_L1A + (size_t)K_HID * WP, OFF_L3 = OFF_L2 + (size_t)K_HID * WP, WT_FLOATS = OFF_L3 + (size_t)AD * HP;
enum { B_L0 = 0, B_ZS1 = WP, B_ZS2 = 2 * WP, B_ZS3 = 3 * WP, B_L1 = 4 * WP, B_L2 = 5 * WP, B_L3 = 6 * WP, BIAS_FLOATS = 7 * WP };

struct F32Args {
    const float *obs, *wt, *bias;
    float *scratch, *actions, *zs;
    long long n;
    float max_action;
    int n_tiles;
};

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// acc[r][c] += sum_k X[8*rg + r][k] * WT[k][col0 + 32 c], k < K (a multiple of 16); rg = warp & 7, col0 = (warp >> 3)*160 + lane.
__device__ __forceinline__ void gemm_acc(float (&acc)[8][NC], const float* __restrict__ X, const float* __restrict__ wt, int K, float* ws,
                                         int warp, int lane, int tid) {
    const int rg = warp & 7, col0 = (warp >> 3) * HALF + lane;
    const int nchunk = K / KC;
    auto load_chunk = [&](int c, int buf) {
        const float4* src = reinterpret_cast<const float4*>(wt + (size_t)c * KC * WP);
        float4* dst = reinterpret_cast<float4*>(ws + buf * KC * WP);
        for (int i = tid; i < KC * WP / 4; i += NT) cp_async16(dst + i, src + i);
        cp_async_commit();
    };
    load_chunk(0, 0);
    for (int c = 0; c < nchunk; ++c) {
        cp_async_wait_all();
        __syncthreads();                           // chunk c has landed for everyone; chunk c-1's buffer is free
        if (c + 1 < nchunk) load_chunk(c + 1, (c + 1) & 1);
        const float* w = ws + (c & 1) * KC * WP + col0;
        const float* x = X + (size_t)(rg * 8) * HP + c * KC;
#pragma unroll
        for (int k4 = 0; k4 < KC; k4 += 4) {
            float4 xv[8];
#pragma unroll
            for (int r = 0; r < 8; ++r) xv[r] = *reinterpret_cast<const float4*>(x + r * HP + k4);     // warp-wide broadcast
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
                float wv[NC];
#pragma unroll
                for (int cc = 0; cc < NC; ++cc) wv[cc] = w[(k4 + kk) * WP + 32 * cc];
#pragma unroll
                for (int r = 0; r < 8; ++r) {
                    const float xs = kk == 0 ? xv[r].x : kk == 1 ? xv[r].y : kk == 2 ? xv[r].z : xv[r].w;
#pragma unroll
                    for (int cc = 0; cc < NC; ++cc) acc[r][cc] = fmaf(xs, wv[cc], acc[r][cc]);
                }
            }
        }
    }
    __syncthreads();                               // every warp is done with X and with the last chunk
}

__device__ __forceinline__ void zero_acc(float (&acc)[8][NC]) {
#pragma unroll
    for (int r = 0; r < 8; ++r)
#pragma unroll
        for (int c = 0; c < NC; ++c) acc[r][c] = 0.f;
}

// ACT 0 ELU, 1 ReLU, 2 AvgL1Norm (:59-60). Adds the bias, applies ACT, stores the thread's 8 x 10 values at dst[row * pitch + col]
// (columns >= 304 are dropped; 300..303 are exactly 0: zero weights and bias).
template <int ACT>
__device__ __forceinline__ void store_act(float (&acc)[8][NC], const float* __restrict__ bias, float* dst, int pitch, int warp, int lane,
                                          float* zs_out, long long row0, long long n, float* part /* smem [2][TM] */) {
    const int rg = warp & 7, half = warp >> 3, col0 = half * HALF + lane;
    float b[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) b[c] = __ldg(bias + col0 + 32 * c);
    float tot[8];
    if (ACT == 2) {                                  // row sums of |x|: this warp's half, then the other half's through shared memory
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            float s = 0.f;
#pragma unroll
            for (int c = 0; c < NC; ++c) s += fabsf(acc[r][c] + b[c]);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            tot[r] = s;
            if (lane == 0) part[half * TM + rg * 8 + r] = s;
        }
        __syncthreads();
#pragma unroll
        for (int r = 0; r < 8; ++r) tot[r] += part[(half ^ 1) * TM + rg * 8 + r];
        __syncthreads();
    }
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        float v[NC];
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            v[c] = acc[r][c] + b[c];
            if (ACT == 0) v[c] = v[c] > 0.f ? v[c] : expm1f(v[c]);
            if (ACT == 1) v[c] = fmaxf(v[c], 0.f);
        }
        if (ACT == 2) {
            const float denom = fmaxf(tot[r] / (float)HD, 1e-8f);
#pragma unroll
            for (int c = 0; c < NC; ++c) v[c] = v[c] / denom;
        }
        const int row = rg * 8 + r;
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            const int col = col0 + 32 * c;
            if (col < HP) dst[(size_t)row * pitch + col] = v[c];
            if (ACT == 2 && zs_out && col < HD && row0 + row < n) zs_out[(row0 + row) * HD + col] = v[c];
        }
    }
}

__global__ void __launch_bounds__(NT, 1) exo_policy_fp32_kernel(const F32Args a) {
    extern __shared__ __align__(16) float sm[];
    float* bufA = sm;
    float* bufB = sm + BUF_FLOATS;
    float* ws = sm + 2 * BUF_FLOATS;
    __shared__ float part[2 * TM];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    float* scratch = a.scratch + (size_t)blockIdx.x * BUF_FLOATS;
    float acc[8][NC];
    for (int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x) {
        const long long row0 = (long long)tile * TM;
        // observations -> bufA[:, 0:80]
        for (int i = tid; i < TM * SD / 4; i += NT) {
            const int r = i / (SD / 4), c = i % (SD / 4);
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (row0 + r < a.n) v = __ldg(reinterpret_cast<const float4*>(a.obs + (row0 + r) * SD) + c);
            *reinterpret_cast<float4*>(bufA + r * HP + 4 * c) = v;
        }
        __syncthreads();
        // a0 = AvgL1Norm(l0(obs)) -> scratch
        zero_acc(acc);
        gemm_acc(acc, bufA, a.wt + OFF_L0, K_OBS, ws, warp, lane, tid);
        store_act<2>(acc, a.bias + B_L0, scratch, HP, warp, lane, nullptr, row0, a.n, part);
        // zs = AvgL1Norm(zs3(elu(zs2(elu(zs1(obs))))))
        zero_acc(acc);
        gemm_acc(acc, bufA, a.wt + OFF_ZS1, K_OBS, ws, warp, lane, tid);
        store_act<0>(acc, a.bias + B_ZS1, bufB, HP, warp, lane, nullptr, row0, a.n, part);
        __syncthreads();
        zero_acc(acc);
        gemm_acc(acc, bufB, a.wt + OFF_ZS2, K_HID, ws, warp, lane, tid);
        store_act<0>(acc, a.bias + B_ZS2, bufA, HP, warp, lane, nullptr, row0, a.n, part);
        __syncthreads();
        zero_acc(acc);
        gemm_acc(acc, bufA, a.wt + OFF_ZS3, K_HID, ws, warp, lane, tid);
        store_act<2>(acc, a.bias + B_ZS3, bufB, HP, warp, lane, a.zs, row0, a.n, part);
        __syncthreads();
        // l1(cat[a0, zs]) = l1[:, 300:600] zs + l1[:, 0:300] a0 + b: the zs half first, a0 comes back from the scratch meanwhile
        for (int i = tid; i < BUF_FLOATS / 4; i += NT)
            reinterpret_cast<float4*>(bufA)[i] = reinterpret_cast<const float4*>(scratch)[i];     // this CTA's own writes (L2/L1 coherent within the CTA after the barriers)
        zero_acc(acc);
        gemm_acc(acc, bufB, a.wt + OFF_L1B, K_HID, ws, warp, lane, tid);
        gemm_acc(acc, bufA, a.wt + OFF_L1A, K_HID, ws, warp, lane, tid);
        store_act<1>(acc, a.bias + B_L1, bufB, HP, warp, lane, nullptr, row0, a.n, part);
        __syncthreads();
        zero_acc(acc);
        gemm_acc(acc, bufB, a.wt + OFF_L2, K_HID, ws, warp, lane, tid);
        store_act<1>(acc, a.bias + B_L2, bufA, HP, warp, lane, nullptr, row0, a.n, part);
        __syncthreads();
        // l3 + tanh + clip: warp w, rows 4w..4w+3; lane l sums k = l, l+32, ...
        const float* w3 = a.wt + OFF_L3;
#pragma unroll 1
        for (int r = 0; r < TM / (NT / 32); ++r) {
            const int row = warp * (TM / (NT / 32)) + r;
            float s[AD];
#pragma unroll
            for (int j = 0; j < AD; ++j) s[j] = 0.f;
            for (int k = lane; k < HD; k += 32) {
                const float x = bufA[row * HP + k];
#pragma unroll
                for (int j = 0; j < AD; ++j) s[j] = fmaf(x, __ldg(w3 + j * HP + k), s[j]);
            }
#pragma unroll
            for (int j = 0; j < AD; ++j)
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) s[j] += __shfl_xor_sync(0xffffffffu, s[j], o);
            if (lane < AD && row0 + row < a.n) {
                float v = 0.f;
#pragma unroll
                for (int j = 0; j < AD; ++j) if (lane == j) v = s[j];
                v = tanhf(v + __ldg(a.bias + B_L3 + lane));
                a.actions[(row0 + row) * AD + lane] = fminf(fmaxf(v, -1.f), 1.f) * a.max_action;
            }
        }
        __syncthreads();
    }
}

void transpose_into(float* dst, const float* W, int n_rows, int ld, int col0, int k_real) {     // dst[k][WP] = W[n][col0 + k]
    for (int n = 0; n < n_rows; ++n)
        for (int k = 0; k < k_real; ++k) dst[(size_t)k * WP + n] = W[(size_t)n * ld + col0 + k];
}

}  // namespace

// Scratch of exo_policy_forward_fp32_noalloc: one BUF_FLOATS slice (64 rows x 304 floats = 77 824 B) per CTA for
// EXO_POLICY_NOALLOC_MAX_OBS / 64 = 64 CTAs, 4.75 MiB, resident on every GPU that loads the library until the process
// exits. In HOST code this name is a host-side stand-in, not a device address: noalloc_scratch_ptr() asks the runtime for
// the address (cudaGetSymbolAddress, once per device, with that device current) and the launch passes that.
constexpr int NOALLOC_MAX_CTAS = EXO_POLICY_NOALLOC_MAX_OBS / TM;
static_assert(EXO_POLICY_NOALLOC_MAX_OBS % TM == 0, "whole tiles");
static __device__ __align__(256) float g_noalloc_scratch[(size_t)NOALLOC_MAX_CTAS * BUF_FLOATS];

struct ExoPolicyF32 {
    int device, n_sm;
    float *wt, *bias, *scratch;
    float* scratch_static;       // device address of g_noalloc_scratch on `device`
};

namespace {
constexpr int MAX_DEVICES = 64;
std::mutex g_noalloc_mu;
float* g_noalloc_ptr[MAX_DEVICES] = {};
// device address of g_noalloc_scratch on the CURRENT device (= `device`); resolved once per device
cudaError_t noalloc_scratch_ptr(int device, float** out) {
    if (device < 0 || device >= MAX_DEVICES) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (g_noalloc_ptr[device] == nullptr) {
        float* ptr = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        g_noalloc_ptr[device] = ptr;
    }
    *out = g_noalloc_ptr[device];
    return cudaSuccess;
}
}  // namespace

int exo_policy_f32_create(int device, int n_sm, ExoPolicyF32** out) {
    ExoPolicyF32* p = new ExoPolicyF32();
    p->device = device; p->n_sm = n_sm; p->wt = p->bias = p->scratch = nullptr; p->scratch_static = nullptr;
    cudaError_t e = cudaMalloc((void**)&p->wt, WT_FLOATS * sizeof(float));
    if (e == cudaSuccess) e = cudaMalloc((void**)&p->bias, BIAS_FLOATS * sizeof(float));
    if (e == cudaSuccess) e = cudaMalloc((void**)&p->scratch, (size_t)n_sm * BUF_FLOATS * sizeof(float));
    if (e == cudaSuccess) e = cudaFuncSetAttribute(exo_policy_fp32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e == cudaSuccess) e = noalloc_scratch_ptr(device, &p->scratch_static);      // (exo_policy_create made `device` current)
    if (e != cudaSuccess) {
        exo_policy_f32_destroy(p);
        return exo_set_error(EXO_ERR_CUDA, "exo_policy (fp32 path): %s", cudaGetErrorString(e));
    }
    *out = p;
    return EXO_OK;
}

void exo_policy_f32_destroy(ExoPolicyF32* p) {
    if (!p) return;
    cudaFree(p->wt); cudaFree(p->bias); cudaFree(p->scratch);
    delete p;
}

int exo_policy_f32_set_weights(ExoPolicyF32* p, const ExoPolicyWeights* w) {
    std::vector<float> wt(WT_FLOATS, 0.f), bias(BIAS_FLOATS, 0.f);
    transpose_into(&wt[OFF_L0], w->l0_w, HD, SD, 0, SD);
    transpose_into(&wt[OFF_ZS1], w->zs1_w, HD, SD, 0, SD);
    transpose_into(&wt[OFF_ZS2], w->zs2_w, HD, HD, 0, HD);
    transpose_into(&wt[OFF_ZS3], w->zs3_w, HD, HD, 0, HD);
    transpose_into(&wt[OFF_L1B], w->l1_w, HD, 2 * HD, HD, HD);
    transpose_into(&wt[OFF_L1A], w->l1_w, HD, 2 * HD, 0, HD);
    transpose_into(&wt[OFF_L2], w->l2_w, HD, HD, 0, HD);
    for (int j = 0; j < AD; ++j) memcpy(&wt[OFF_L3 + (size_t)j * HP], w->l3_w + (size_t)j * HD, HD * sizeof(float));
    memcpy(&bias[B_L0], w->l0_b, HD * 4);   memcpy(&bias[B_ZS1], w->zs1_b, HD * 4); memcpy(&bias[B_ZS2], w->zs2_b, HD * 4);
    memcpy(&bias[B_ZS3], w->zs3_b, HD * 4); memcpy(&bias[B_L1], w->l1_b, HD * 4);   memcpy(&bias[B_L2], w->l2_b, HD * 4);
    memcpy(&bias[B_L3], w->l3_b, AD * 4);
    cudaError_t e = cudaMemcpy(p->wt, wt.data(), WT_FLOATS * sizeof(float), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(p->bias, bias.data(), BIAS_FLOATS * sizeof(float), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) return exo_set_error(EXO_ERR_CUDA, "exo_policy_set_weights (fp32 path): %s", cudaGetErrorString(e));
    return EXO_OK;
}

int exo_policy_f32_forward(ExoPolicyF32* p, const float* obs_dev, long long n, float max_action, float* actions_dev, float* zs_dev, void* stream) {
    F32Args a;
    a.obs = obs_dev; a.wt = p->wt; a.bias = p->bias; a.scratch = p->scratch; a.actions = actions_dev; a.zs = zs_dev;
    a.n = n; a.max_action = max_action;
    const long long tiles = (n + TM - 1) / TM;
    if (tiles > 0x7FFFFFFF) return exo_set_error(EXO_ERR_ARG, "exo_policy_forward: too many observations");
    a.n_tiles = (int)tiles;
    const int grid = (int)(tiles < p->n_sm ? tiles : p->n_sm);
    exo_policy_fp32_kernel<<<grid, NT, SMEM_BYTES, (cudaStream_t)stream>>>(a);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return exo_set_error(EXO_ERR_CUDA, "exo_policy_forward (fp32 path): %s", cudaGetErrorString(e));
    return EXO_OK;
}

// exo_policy_forward_fp32_noalloc: the same kernel, one CTA per 64-row tile, its scratch slices in g_noalloc_scratch. Nothing
// is allocated or freed here (one kernel launch + cudaGetLastError). The kernel writes every word of a CTA's slice (a0 for all
// 64 rows x 304 columns) before it reads it back, so the array's contents from the previous call need no clearing; two calls
// that overlap on the GPU (different streams, or different handles on one device) would share the slices and race.
int exo_policy_f32_forward_noalloc(ExoPolicyF32* p, const float* obs_dev, long long n, float max_action, float* actions_dev, float* zs_dev,
                                   void* stream) {
    if (n > EXO_POLICY_NOALLOC_MAX_OBS)
        return exo_set_error(EXO_ERR_ARG, "exo_policy_forward_fp32_noalloc: %lld observations exceed the static scratch's maximum of %d "
                             "(EXO_POLICY_NOALLOC_MAX_OBS); split the batch or use exo_policy_forward_fp32", n, EXO_POLICY_NOALLOC_MAX_OBS);
    if (p->scratch_static == nullptr) return exo_set_error(EXO_ERR_STATE, "exo_policy_forward_fp32_noalloc: static scratch not resolved");
    F32Args a;
    a.obs = obs_dev; a.wt = p->wt; a.bias = p->bias; a.scratch = p->scratch_static; a.actions = actions_dev; a.zs = zs_dev;
    a.n = n; a.max_action = max_action;
    a.n_tiles = (int)((n + TM - 1) / TM);            // <= NOALLOC_MAX_CTAS
    exo_policy_fp32_kernel<<<a.n_tiles, NT, SMEM_BYTES, (cudaStream_t)stream>>>(a);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return exo_set_error(EXO_ERR_CUDA, "exo_policy_forward_fp32_noalloc: %s", cudaGetErrorString(e));
    return EXO_OK;
}
