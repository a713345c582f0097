// exo_td7.cu — the TD7 update of include/exo_td7.h: one Agent.train() iteration (Agent/TD7_multi_agent_Pink_noise.py
// :230-312) with every network, gradient and Adam state resident in HBM.
//
// Building blocks:
//   tc5::gemm_tc5_kernel  (exo_td7_tc5.cuh, the DEFAULT) the GEMM contract below on tcgen05: 128 x TN tiles, 3xTF32 products
//                         with the accumulator in tensor memory. EXO_TD7_GEMM=mma selects gemm_kernel instead.
//   gemm_kernel<TA,TB,TC,TN>  C = op(A) op(B) on 64 x TN x 16 tiles (TN = 64 | 80, picked per launch so the tiles fill one
//                         wave of SMs), 256 threads, a 4-stage cp.async ring. TC: the products run on the tensor cores as
//                         3xTF32 (split operands, mma.sync m16n8k8, fp32 accumulation; fp32-class accuracy); !TC
//                         (EXO_TD7_FFMA=1): float32 FFMA on the FP32 pipe, the reference's exact arithmetic. One kernel serves
//                           forward      Y  = act(X W^T + b)                      (TA=0, TB=1, bias + activation epilogue)
//                           input grad   dX = (dY W[:, c0:c1]) * act'(saved out)  (TA=0, TB=0, derivative epilogue)
//                           weight grad  dW += dY^T X                             (TA=1, TB=0, split over the batch,
//                                                                                  atomicAdd epilogue)
//                         torch.cat is never materialised for outputs: producers write into column slices (ld*).
//   row kernels           AvgL1Norm forward / backward (one warp per row), column sums (bias gradients), the loss
//                         heads (mse, LAP_huber + priorities, target value with clipping and the running max / min).
//   adam_kernel           torch.optim.Adam over a network's flat parameter vector (one launch per optimiser).
// The orchestration (exo_td7_train) is a fixed launch sequence on the caller's stream; nothing synchronises.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <initializer_list>
#include <utility>
#include <vector>

#include "../../include/exo_env.h"
#include "../../include/exo_td7.h"

int exo_set_error(int code, const char* fmt, ...);     // exo_api.cu

namespace {

enum { ACT_NONE = 0, ACT_RELU = 1, ACT_ELU = 2, ACT_TANH = 3 };
enum { OUT_STORE = 0, OUT_ADD = 1, OUT_ATOMIC = 2 };

struct GemmArgs {
    const float* A; int lda;
    const float* B; int ldb;
    float* C; int ldc;
    int M, N, K, kchunk;
    const float* bias; int act;          // forward epilogue
    const float* D; int ldd; int dact;   // backward epilogue: multiply by act'(D(m,n)), D = the saved forward OUTPUT
    int out;
};

__device__ __forceinline__ float act_fwd(float v, int act) {
    if (act == ACT_RELU) return fmaxf(v, 0.f);
    if (act == ACT_ELU) return v > 0.f ? v : expm1f(v);
    if (act == ACT_TANH) return tanhf(v);
    return v;
}
__device__ __forceinline__ float act_bwd(float y, int act) {      // derivative from the output y
    if (act == ACT_RELU) return y > 0.f ? 1.f : 0.f;
    if (act == ACT_ELU) return y > 0.f ? 1.f : y + 1.f;
    if (act == ACT_TANH) return 1.f - y * y;
    return 1.f;
}

constexpr int TM = 64, TK = 16, STAGES = 4;

// ---- 3xTF32: the GEMM on the tensor cores with fp32-class accuracy. Every operand x is split in registers into
// hi = x rounded to TF32 (10 explicit mantissa bits, nearest) and lo = x - hi (exact in float32, |lo| <= 2^-11 |x|);
// a b ~ lo_a hi_b + hi_a lo_b + hi_a hi_b: three mma.sync.m16n8k8 (tf32 in, f32 accumulate) per fragment pair, small
// terms first. The tensor core reads only the upper 19 bits of lo (truncation, <= 2^-11 |lo|, sign-symmetric), so each
// product is off by <= 2^-21 |a b| including the dropped lo_a lo_b term - a few float32 roundings.
// cvt.rna.tf32.f32 is a 4-instruction sequence on sm_100a (Inf/NaN guard included) and made the splits 60 % of the
// loop's instructions; the integer form below is 3 instructions per operand (finite inputs).
__device__ __forceinline__ void split_tf32(float x, uint32_t& hi, uint32_t& lo) {
    hi = (__float_as_uint(x) + 0x1000u) & 0xffffe000u;
    lo = __float_as_uint(x - __uint_as_float(hi));
}
__device__ __forceinline__ void mma_tf32(float (&c)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

__device__ __forceinline__ void gemm_store(const GemmArgs& g, int m, int n, float v) {
    if (m >= g.M || n >= g.N) return;
    if (g.bias) v += g.bias[n];
    v = act_fwd(v, g.act);
    if (g.D) v *= act_bwd(g.D[(size_t)m * g.ldd + n], g.dact);
    float* c = g.C + (size_t)m * g.ldc + n;
    if (g.out == OUT_STORE) *c = v;
    else if (g.out == OUT_ADD) *c += v;
    else atomicAdd(c, v);
}

#include "exo_td7_tc5.cuh"      // gemm_tc5_kernel: the same GEMM on tcgen05 + tensor memory (EXO_TD7_GEMM=tc5)

// C tile 64 x TN per CTA (TN = 64 or 80), 256 threads, k in steps of 16 through a 4-stage cp.async ring (4-byte copies:
// any alignment, either operand transposed, zero fill outside the matrix). The host picks TN so that the tiles fill one
// wave of SMs: 1920 x 300 is 30 x 5 = 150 tiles of 64 x 64 - two more than a B200 has SMs, i.e. two waves - but
// 30 x 4 = 120 tiles of 64 x 80. Global pointers and shared offsets of the copies are set up once; per k-step a copy
// costs a bound compare, the cp.async and a pointer increment.
//   TC = false  FP32 pipe: 4 x (TN/16) accumulators per thread, tiles k-major (A pitch 68, float4 reads).
//   TC = true   tensor cores, 3xTF32: 8 warps as 4 (m) x 2 (n), warp tile 16 x TN/2; tiles k-major with pitch = 8 mod 32
//               and the column index XOR-ed with 2 ((k >> 2) & 3): conflict-free for the fragment reads (lanes = 4 k x
//               8 m) and for both staging patterns (lanes along k or along m).
template <bool TA, bool TB, bool TC, int TN>
__global__ void __launch_bounds__(256) gemm_kernel(GemmArgs g) {
    constexpr int PA = TC ? 72 : 68;
    constexpr int PB = TC ? (TN == 64 ? 72 : 104) : (TN == 64 ? 68 : 84);
    constexpr int NB = TN * TK / 256;             // B-tile elements per thread: 4 or 5
    constexpr int NF = TN / 16;                   // TC: n8 fragments per warp; !TC: columns per thread
    __shared__ __align__(16) float As[STAGES][TK][PA];
    __shared__ __align__(16) float Bs[STAGES][TK][PB];
    const int t = threadIdx.x;
    const int m0 = blockIdx.y * TM, n0 = blockIdx.x * TN;
    const int kbeg = blockIdx.z * g.kchunk;
    const int kend = min(g.K, kbeg + g.kchunk);
    auto swz = [](int k) { return TC ? ((k >> 2) & 3) << 1 : 0; };
    // copies: element (mm, kk) of the A tile / (kk, nn) of the B tile, contiguous along the operand's fast axis
    const float* pa[4];
    const float* pb[NB];
    uint32_t da[4], db[NB];
    int ka[4], kb[NB];                            // k offset of the element inside the tile; -1 = outside M / N
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int mm = TA ? (t & 63) : (t >> 4) + 16 * i, kk = TA ? (t >> 6) + 4 * i : (t & 15);
        const int gm = m0 + mm;
        ka[i] = gm < g.M ? kk : -1;
        pa[i] = gm < g.M ? (TA ? g.A + (size_t)(kbeg + kk) * g.lda + gm : g.A + (size_t)gm * g.lda + kbeg + kk) : g.A;
        da[i] = (uint32_t)__cvta_generic_to_shared(&As[0][kk][mm ^ swz(kk)]);
    }
#pragma unroll
    for (int i = 0; i < NB; ++i) {
        const int e = t + 256 * i;
        const int nn = TB ? e / TK : e % TN, kk = TB ? e % TK : e / TN;
        const int gn = n0 + nn;
        kb[i] = gn < g.N ? kk : -1;
        pb[i] = gn < g.N ? (TB ? g.B + (size_t)gn * g.ldb + kbeg + kk : g.B + (size_t)(kbeg + kk) * g.ldb + gn) : g.B;
        db[i] = (uint32_t)__cvta_generic_to_shared(&Bs[0][kk][nn ^ swz(kk)]);
    }
    const size_t stepa = TA ? (size_t)TK * g.lda : TK, stepb = TB ? TK : (size_t)TK * g.ldb;
    int knext = kbeg;                             // k of the next tile to copy
    auto issue = [&](int stage) {
        if (knext < kend) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const bool ok = ka[i] >= 0 && knext + ka[i] < kend;
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(da[i] + stage * (TK * PA * 4)), "l"(ok ? pa[i] : g.A), "r"(ok ? 4 : 0) : "memory");
                pa[i] += stepa;
            }
#pragma unroll
            for (int i = 0; i < NB; ++i) {
                const bool ok = kb[i] >= 0 && knext + kb[i] < kend;
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(db[i] + stage * (TK * PB * 4)), "l"(ok ? pb[i] : g.B), "r"(ok ? 4 : 0) : "memory");
                pb[i] += stepb;
            }
            knext += TK;
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    const int lane = t & 31, warp = t >> 5, gid = lane >> 2, tig = lane & 3;
    const int wm = (warp >> 1) * 16, wn = (warp & 1) * (TN / 2);      // TC
    const int tx = t & 15, ty = t >> 4;                                // !TC
    float acc[NF * 4];
#pragma unroll
    for (int i = 0; i < NF * 4; ++i) acc[i] = 0.f;
#pragma unroll
    for (int s = 0; s < STAGES - 1; ++s) issue(s);
    int stage = 0;
    for (int k0 = kbeg; k0 < kend; k0 += TK) {
        asm volatile("cp.async.wait_group %0;" ::"n"(STAGES - 2) : "memory");
        __syncthreads();                       // tile `stage` has landed for everyone; tile stage-1 is no longer being read
        issue((stage + STAGES - 1) % STAGES);
        if (TC) {
#pragma unroll
            for (int ks = 0; ks < TK; ks += 8) {
                const int s0 = swz(ks + tig), s1 = swz(ks + tig + 4);
                uint32_t ah[4], al[4], bh[NF][2], bl[NF][2];
                const int m = wm + gid;
                split_tf32(As[stage][ks + tig][m ^ s0], ah[0], al[0]);
                split_tf32(As[stage][ks + tig][(m + 8) ^ s0], ah[1], al[1]);
                split_tf32(As[stage][ks + tig + 4][m ^ s1], ah[2], al[2]);
                split_tf32(As[stage][ks + tig + 4][(m + 8) ^ s1], ah[3], al[3]);
#pragma unroll
                for (int j = 0; j < NF; ++j) {
                    const int n = wn + 8 * j + gid;
                    split_tf32(Bs[stage][ks + tig][n ^ s0], bh[j][0], bl[j][0]);
                    split_tf32(Bs[stage][ks + tig + 4][n ^ s1], bh[j][1], bl[j][1]);
                }
#pragma unroll
                for (int j = 0; j < NF; ++j) {
                    float(&c)[4] = *reinterpret_cast<float(*)[4]>(&acc[j * 4]);
                    mma_tf32(c, al, bh[j]);
                    mma_tf32(c, ah, bl[j]);
                    mma_tf32(c, ah, bh[j]);
                }
            }
        } else {
#pragma unroll
            for (int kk = 0; kk < TK; ++kk) {
                const float4 a = *reinterpret_cast<const float4*>(&As[stage][kk][ty * 4]);
                const float av[4] = {a.x, a.y, a.z, a.w};
                float bv[NF];
#pragma unroll
                for (int j = 0; j < NF; ++j) bv[j] = Bs[stage][kk][tx * NF + j];
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < NF; ++j) acc[i * NF + j] = fmaf(av[i], bv[j], acc[i * NF + j]);
            }
        }
        stage = (stage + 1) % STAGES;
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    if (TC) {
#pragma unroll
        for (int j = 0; j < NF; ++j)
#pragma unroll
            for (int r = 0; r < 4; ++r)
                gemm_store(g, m0 + wm + gid + ((r >> 1) << 3), n0 + wn + 8 * j + 2 * tig + (r & 1), acc[j * 4 + r]);
    } else {
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < NF; ++j) gemm_store(g, m0 + ty * 4 + i, n0 + tx * NF + j, acc[i * NF + j]);
    }
}

template <bool TA, bool TB>
void launch_gemm(const GemmArgs& g, bool tc, int tn, dim3 grid, cudaStream_t s) {
    if (tc) {
        if (tn == 64) gemm_kernel<TA, TB, true, 64><<<grid, 256, 0, s>>>(g);
        else gemm_kernel<TA, TB, true, 80><<<grid, 256, 0, s>>>(g);
    } else {
        if (tn == 64) gemm_kernel<TA, TB, false, 64><<<grid, 256, 0, s>>>(g);
        else gemm_kernel<TA, TB, false, 80><<<grid, 256, 0, s>>>(g);
    }
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// y = x / max(mean|x|, 1e-8) per row (:59-60); one warp per row
__global__ void avgl1_fwd_kernel(const float* x, int ldx, float* y, int ldy, int rows, int n) {
    const int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (r >= rows) return;
    const float* xr = x + (size_t)r * ldx;
    float s = 0.f;
    for (int c = lane; c < n; c += 32) s += fabsf(xr[c]);
    const float d = fmaxf(warp_sum(s) / (float)n, 1e-8f);
    float* yr = y + (size_t)r * ldy;
    for (int c = lane; c < n; c += 32) yr[c] = xr[c] / d;
}

// dx = dy / d - sign(x) (sum dy x) / (n d^2), the second term only where mean|x| >= 1e-8 (clamp's gradient)
__global__ void avgl1_bwd_kernel(const float* dy, int lddy, const float* x, int ldx, float* dx, int lddx, int rows, int n) {
    const int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (r >= rows) return;
    const float* xr = x + (size_t)r * ldx;
    const float* gr = dy + (size_t)r * lddy;
    float s = 0.f, p = 0.f;
    for (int c = lane; c < n; c += 32) { s += fabsf(xr[c]); p += gr[c] * xr[c]; }
    const float m = warp_sum(s) / (float)n;
    p = warp_sum(p);
    const float d = fmaxf(m, 1e-8f);
    const float k = m >= 1e-8f ? p / ((float)n * d * d) : 0.f;
    float* o = dx + (size_t)r * lddx;
    for (int c = lane; c < n; c += 32) {
        const float xv = xr[c];
        const float sg = xv > 0.f ? 1.f : (xv < 0.f ? -1.f : 0.f);
        o[c] = gr[c] / d - sg * k;
    }
}

__global__ void copy_cols_kernel(float* dst, int ldd, const float* src, int lds, int rows, int cols) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)rows * cols) return;
    const int r = (int)(i / cols), c = (int)(i % cols);
    dst[(size_t)r * ldd + c] = src[(size_t)r * lds + c];
}

// gb[n] += sum_rows dy[r, n]: block = 32 columns x 8 row lanes over one slice of the rows (gridDim.y slices, atomicAdd)
__global__ void colsum_kernel(const float* dy, int ld, float* gb, int rows, int n, int rows_per_slice) {
    __shared__ float part[8][33];
    const int c = blockIdx.x * 32 + threadIdx.x;
    const int r0 = blockIdx.y * rows_per_slice, r1 = min(rows, r0 + rows_per_slice);
    float s = 0.f;
    if (c < n)
        for (int r = r0 + threadIdx.y; r < r1; r += 8) s += dy[(size_t)r * ld + c];
    part[threadIdx.y][threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.y == 0 && c < n) {
        for (int k = 1; k < 8; ++k) s += part[k][threadIdx.x];
        atomicAdd(&gb[c], s);
    }
}

// device scalars of the learner
enum { SC_MAX = 0, SC_MIN, SC_MAX_TARGET, SC_MIN_TARGET, SC_ENC_LOSS, SC_CRITIC_LOSS, SC_ACTOR_LOSS, SC_ABSQ, SC_BC_MSE, SC_COUNT };

__device__ __forceinline__ float block_sum(float v, float* sh) {
    v = warp_sum(v);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    float r = 0.f;
    if (threadIdx.x < 32) {
        r = threadIdx.x < (blockDim.x >> 5) ? sh[threadIdx.x] : 0.f;
        r = warp_sum(r);
    }
    __syncthreads();
    return r;       // valid in warp 0
}

// F.mse_loss(pred, next_zs) (:243): d = 2 (pred - next) / count, loss accumulated into sc[SC_ENC_LOSS]
__global__ void mse_grad_kernel(const float* pred, const float* next, float* d, size_t count, float* sc) {
    __shared__ float sh[32];
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    float e = 0.f;
    if (i < count) {
        e = pred[i] - next[i];
        d[i] = e * (2.0f / (float)count);
    }
    const float s = block_sum(e * e, sh);
    if (threadIdx.x == 0) atomicAdd(&sc[SC_ENC_LOSS], s / (float)count);
}

// next_action = (actor_target(...) + (randn * sigma).clamp(-clip, clip)).clamp(-1, 1)   (:254-256)
// per-iteration scalars, written by the host before each iteration (so that a captured graph can be replayed unchanged)
enum { DP_SIGMA = 0, DP_ADAM = 1 /* + 2 opt: step size, sqrt(bias correction 2) */, DP_STEP = 8 /* uint64 */, DP_COUNT = 16 };

__global__ void target_action_kernel(float* na, const float* randn, int count, const float* dp, float clip) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const float sigma = dp[DP_SIGMA];
    const float nz = fminf(fmaxf(randn[i] * sigma, -clip), clip);
    na[i] = fminf(fmaxf(na[i] + nz, -1.f), 1.f);
}

// Q_target = reward + not_done * discount * min_h(Q_h).clamp(min_target, max_target); running max / min (:260-264).
// One block: the running extrema are updated without atomics.
__global__ void q_target_kernel(const float* q2, const float* reward, const float* not_done, float* qt, int rows, float discount, float* sc) {
    __shared__ float smax[32], smin[32];
    const float lo = sc[SC_MIN_TARGET], hi = sc[SC_MAX_TARGET];
    float mx = -INFINITY, mn = INFINITY;
    for (int b = threadIdx.x; b < rows; b += blockDim.x) {
        float q = fminf(q2[2 * b], q2[2 * b + 1]);
        q = fminf(fmaxf(q, lo), hi);
        const float v = reward[b] + not_done[b] * discount * q;
        qt[b] = v;
        mx = fmaxf(mx, v); mn = fminf(mn, v);
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) { mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o)); mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o)); }
    if ((threadIdx.x & 31) == 0) { smax[threadIdx.x >> 5] = mx; smin[threadIdx.x >> 5] = mn; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < (int)(blockDim.x >> 5); ++w) { mx = fmaxf(mx, smax[w]); mn = fminf(mn, smin[w]); }
        sc[SC_MAX] = fmaxf(sc[SC_MAX], mx);
        sc[SC_MIN] = fminf(sc[SC_MIN], mn);
    }
}

// td = |Q - Q_target|; LAP_huber (:63-64) and its gradient; priority = max_h td .clamp(min_priority) ** alpha (:281)
__global__ void td_head_kernel(const float* q2, const float* qt, float* dq, float* prio, int rows, float min_priority, float alpha, float* sc) {
    __shared__ float sh[32];
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    float loss = 0.f;
    if (b < rows) {
        float tmax = 0.f;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const float d = q2[2 * b + h] - qt[b];
            const float td = fabsf(d);
            const float sg = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);
            loss += td < min_priority ? 0.5f * td * td : min_priority * td;
            dq[2 * b + h] = (td < min_priority ? td : min_priority) * sg / (float)rows;
            tmax = fmaxf(tmax, td);
        }
        prio[b] = powf(fmaxf(tmax, min_priority), alpha);
    }
    const float s = block_sum(loss, sh);
    if (threadIdx.x == 0) atomicAdd(&sc[SC_CRITIC_LOSS], s / (float)rows);
}

// actor_loss = -Q.mean() over [rows, 2] (:291): dQ = -1 / (2 rows); also |Q|.mean() for the offline term
__global__ void actor_head_kernel(const float* q2, float* dq, int rows, float* sc) {
    __shared__ float sh[32];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    float q = 0.f;
    if (i < 2 * rows) { q = q2[i]; dq[i] = -1.0f / (float)(2 * rows); }
    const float s = block_sum(q, sh);
    const float a = block_sum(fabsf(q), sh);
    if (threadIdx.x == 0) {
        atomicAdd(&sc[SC_ACTOR_LOSS], -s / (float)(2 * rows));
        atomicAdd(&sc[SC_ABSQ], a / (float)(2 * rows));
    }
}

// d = (da + db [+ bc]) * (1 - act^2): the gradient paths into the actor's tanh output - through the critic's sa input,
// through fixed_encoder.zsa, and offline (:292-293) bc = lmbda |Q|.mean() * 2 (act - action) / count
__global__ void tanh_bwd_kernel(float* d, const float* da, const float* db, const float* act, int count, const float* action,
                                float bc_scale, float* sc) {
    __shared__ float sh[32];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    float e = 0.f;
    if (i < count) {
        const float y = act[i];
        float g = da[i] + db[i];
        if (bc_scale != 0.f) {
            e = y - action[i];
            g += bc_scale * sc[SC_ABSQ] * e;
        }
        d[i] = g * (1.f - y * y);
    }
    if (bc_scale != 0.f) {                       // uniform per launch
        const float s = block_sum(e * e, sh);
        if (threadIdx.x == 0) atomicAdd(&sc[SC_BC_MSE], s / (float)count);
    }
}

// torch.optim.Adam, single-tensor arithmetic: g += wd p; m.lerp_(g, 1-b1); v = b2 v + (1-b2) g g;
// p -= (lr / bc1) * m / (sqrt(v) / sqrt(bc2) + eps)
__global__ void adam_kernel(float* p, const float* g, float* m, float* v, size_t n, float wd, const float* dp) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float step_size = dp[0], bc2_sqrt = dp[1];
    const float b1 = 0.9f, b2 = 0.999f, eps = 1e-8f;
    const float pv = p[i];
    const float gv = g[i] + wd * pv;
    const float mv = m[i] + (gv - m[i]) * (1.f - b1);
    const float vv = v[i] * b2 + gv * gv * (1.f - b2);
    m[i] = mv; v[i] = vv;
    const float denom = sqrtf(vv) / bc2_sqrt + eps;
    p[i] = pv - step_size * (mv / denom);
}

// standard normals for the target-policy noise when the caller passes none: splitmix64 counter hash + Box-Muller
__global__ void randn_kernel(float* out, int count, uint64_t seed, const uint64_t* step_ptr) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const uint64_t step = *step_ptr;
    uint64_t z = seed + 0x9E3779B97F4A7C15ull * (step * 0x100000000ull + (uint64_t)i + 1);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z ^= z >> 31;
    const float u1 = ((float)(uint32_t)(z >> 40) + 1.0f) * (1.0f / 16777216.0f);      // (0, 1]
    const float u2 = (float)(uint32_t)((z >> 8) & 0xFFFFFF) * (1.0f / 16777216.0f);
    out[i] = sqrtf(-2.f * logf(u1)) * cospif(2.f * u2);
}

struct Lin { int out, in; size_t w, b; };
struct NetDef { std::vector<Lin> L; size_t size = 0; };

NetDef make_def(std::initializer_list<std::pair<int, int>> layers) {
    NetDef d;
    for (auto& l : layers) {
        Lin x;
        x.out = l.first; x.in = l.second;
        x.w = d.size; d.size += (size_t)x.out * x.in;
        x.b = d.size; d.size += (size_t)x.out;
        d.L.push_back(x);
    }
    return d;
}

inline int kind_of(int net) { return net <= EXO_TD7_CHECKPOINT_ACTOR ? 0 : (net <= EXO_TD7_CRITIC_TARGET ? 1 : 2); }
const int OPT_NET[3] = {EXO_TD7_ACTOR, EXO_TD7_CRITIC, EXO_TD7_ENCODER};

}  // namespace

struct ExoTD7 {
    int device = 0, n_sm = 148;
    ExoTD7GradHook hook = nullptr;
    void* hook_user = nullptr;
    bool ffma = false;                      // EXO_TD7_FFMA=1: the FP32-pipe GEMM instead of the 3xTF32 tensor-core one
    int tc5_cluster = 1;                    // EXO_TD7_TC5_CLUSTER: largest split-k cluster the tcgen05 GEMM may use (1, 2, 4)
    bool tc5 = true;                        // 3xTF32 on tcgen05 / tensor memory (exo_td7_tc5.cuh); EXO_TD7_GEMM=mma: the mma.sync kernel
    ExoTD7Config c;
    int64_t launches = 0, steps = 0;
    double sigma = 0;                       // hp.target_policy_noise, decays by policy_noise_decrease per iteration
    NetDef def[3];                          // actor, critic, encoder
    float* net[EXO_TD7_N_NETS] = {};
    float *grad[3] = {}, *m[3] = {}, *v[3] = {};
    int64_t opt_step[3] = {0, 0, 0};
    float* sc = nullptr;
    std::vector<float*> owned;
    // activations ([batch_rows, *])
    // side streams: 0 the target branch, 1 the fixed-encoder branch (both beside the encoder update), 2 the critic's
    // second head, 3 the weight gradients (off the dX chain, which is the critical path of every backward pass)
    cudaStream_t side[4] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t ev_fork = nullptr, ev_join[2] = {nullptr, nullptr};
    cudaEvent_t evs[32] = {};
    int ev_next = 0;
    // per-iteration scalars: device block + pinned host ring (the copy is enqueued ahead of the iteration)
    float* dparams = nullptr;
    float* hparams = nullptr;
    uint64_t* dstep_scratch = nullptr;        // exo_td7_randn's own step cell
    // CUDA graphs of one iteration, per (batch pointers, actor half): replayed on `gstream` between two events
    struct Graph { const void* key[7]; bool actor; cudaGraphExec_t exec; int64_t kernels; };
    std::vector<Graph> graphs;
    cudaStream_t gstream = nullptr;
    cudaEvent_t g_in = nullptr, g_out = nullptr;
    bool use_graph = true;                    // EXO_TD7_GRAPH=0: plain stream launches
    float *G[6], *G2[4];                                        // per-layer gradient buffers of a backward pass (main chain, head 1)
    float *t_h1, *t_h2, *t_z3, *next_zs;                       // no-grad encoder passes (target branch)
    float *n_h1, *n_h2, *n_z3, *f_h1, *f_h2, *f_z3;            // the same scratch for the other two branches
    float *e_h1, *e_h2, *e_z3, *e_cat, *e_g1, *e_g2, *e_pred;  // encoder update
    float *tzs, *t_cat, *tzsa, *na, *rn;                       // target
    float *a_u, *a_cat, *a_h1, *a_h2, *act;                    // actor (target pass, then the actor update's cache)
    float *fzs, *f_cat, *f_g1, *f_g2, *fzsa;                   // fixed encoder
    float *c_sa, *c_u[2], *c_cat[2], *c_h1[2], *c_h2[2], *q2, *qt, *dq;   // critic
    float *dA, *dB, *d_zsa, *d_act1, *d_act2, *dT;             // gradients in flight
};

namespace {

struct Ctx {
    ExoTD7* t;
    cudaStream_t s;
    int B;

    void gemm(bool ta, bool tb, GemmArgs g, int splits) {
        if (g.M <= 0 || g.N <= 0 || g.K <= 0) return;
        if (t->tc5 && !t->ffma) {
            // 128-row tiles: a 1920-row batch is 15 of them. Atomic outputs (weight gradients) split k over blockIdx.z as
            // before; store / add outputs split k over a cluster of 2 (4 for single-column-tile layers) CTAs that sum
            // their partial tiles through distributed shared memory, while the tile count leaves SMs idle.
            const int mt5 = (g.M + tc5::TM5 - 1) / tc5::TM5;
            auto ctas = [&](int tn) { return mt5 * ((g.N + tn - 1) / tn); };
            const int tn5 = (g.N <= 64 || ctas(64) <= t->n_sm) && (g.N % 80 != 0 || g.N <= 64) ? 64 : 80;
            const int atoms = (g.K + tc5::KA - 1) / tc5::KA;
            int ck = 1, ks5 = 1;
            if (g.out == OUT_ATOMIC) {
                g.kchunk = splits <= 1 ? g.K : (((g.K + splits - 1) / splits + tc5::KA - 1) / tc5::KA) * tc5::KA;
                ks5 = (g.K + g.kchunk - 1) / g.kchunk;
            } else {
                g.kchunk = g.K;
                auto fits = [&](int c) { const int per = (atoms + c - 1) / c; return atoms - per * (c - 1) >= 1 && per >= 2; };
                if (t->tc5_cluster >= 2 && ctas(tn5) * 2 <= t->n_sm + 12 && fits(2)) ck = 2;
                if (t->tc5_cluster >= 4 && ctas(tn5) * 4 <= t->n_sm / 2 && fits(4)) ck = 4;
            }
            dim3 grid5((g.N + tn5 - 1) / tn5, mt5, ck > 1 ? ck : ks5);
            if (!ta && tb) tc5::launch<false, true>(g, tn5, ck, grid5, s);          // (errors surface in cudaGetLastError, as for gemm_kernel)
            else if (!ta && !tb) tc5::launch<false, false>(g, tn5, ck, grid5, s);
            else tc5::launch<true, false>(g, tn5, ck, grid5, s);
            t->launches++;
            return;
        }
        g.kchunk = splits <= 1 ? g.K : (((g.K + splits - 1) / splits + TK - 1) / TK) * TK;
        const int ks = (g.K + g.kchunk - 1) / g.kchunk, mt = (g.M + TM - 1) / TM;
        // tile width: the one with the shorter critical path = waves of CTAs x columns per tile
        auto cost = [&](int tn) { return (long)((mt * ((g.N + tn - 1) / tn) * ks + t->n_sm - 1) / t->n_sm) * tn; };
        const int tn = cost(80) < cost(64) ? 80 : 64;
        dim3 grid((g.N + tn - 1) / tn, mt, ks);
        if (!ta && tb) launch_gemm<false, true>(g, !t->ffma, tn, grid, s);
        else if (!ta && !tb) launch_gemm<false, false>(g, !t->ffma, tn, grid, s);
        else launch_gemm<true, false>(g, !t->ffma, tn, grid, s);
        t->launches++;
    }
    // Y = act(X W^T + b)
    void fwd(const float* net, const Lin& L, const float* X, int ldx, float* Y, int ldy, int act) {
        GemmArgs g = {};
        g.A = X; g.lda = ldx; g.B = net + L.w; g.ldb = L.in; g.C = Y; g.ldc = ldy;
        g.M = B; g.N = L.out; g.K = L.in; g.bias = net + L.b; g.act = act; g.out = OUT_STORE;
        gemm(false, true, g, 1);
    }
    // dX[:, 0:nc] (=|+=) dY W[:, c0:c0+nc] (* act'(D))
    void dx(const float* net, const Lin& L, const float* dY, int lddy, float* dX, int lddx, int c0, int nc,
            const float* D, int ldd, int dact, int out = OUT_STORE) {
        GemmArgs g = {};
        g.A = dY; g.lda = lddy; g.B = net + L.w + c0; g.ldb = L.in; g.C = dX; g.ldc = lddx;
        g.M = B; g.N = nc; g.K = L.out; g.D = D; g.ldd = ldd; g.dact = dact; g.out = out;
        gemm(false, false, g, 1);
    }
    // gW += dY^T X, gb += colsum(dY)   (the gradient vector was zeroed at the start of the backward pass)
    void dw(float* grad, const Lin& L, const float* dY, int lddy, const float* X, int ldx) {
        GemmArgs g = {};
        g.A = dY; g.lda = lddy; g.B = X; g.ldb = ldx; g.C = grad + L.w; g.ldc = L.in;
        g.M = L.out; g.N = L.in; g.K = B; g.out = OUT_ATOMIC;
        const int tiles = ((L.out + TM - 1) / TM) * ((L.in + 63) / 64);
        int splits = (2 * t->n_sm + tiles - 1) / tiles;
        splits = std::max(1, std::min(splits, (B + 63) / 64));
        gemm(true, false, g, splits);
        const int slices = std::max(1, std::min((B + 63) / 64, (2 * t->n_sm) / ((L.out + 31) / 32)));
        const int rps = (B + slices - 1) / slices;
        colsum_kernel<<<dim3((L.out + 31) / 32, (B + rps - 1) / rps), dim3(32, 8), 0, s>>>(dY, lddy, grad + L.b, B, L.out, rps);
        t->launches++;
    }
    // `other` waits for everything enqueued so far on this stream
    void before(cudaStream_t other) {
        cudaEvent_t e = t->evs[t->ev_next++ & 31];
        cudaEventRecord(e, s);
        cudaStreamWaitEvent(other, e, 0);
    }
    // the weight gradient of a layer on the weight-gradient stream, as soon as this stream has produced dY
    void dw_side(float* grad, const Lin& L, const float* dY, int lddy, const float* X, int ldx) {
        Ctx w{t, t->side[3], B};
        before(w.s);
        w.dw(grad, L, dY, lddy, X, ldx);
    }
    void join_dw() {
        Ctx w{t, t->side[3], B};
        w.before(s);
    }
    void avgl1(const float* x, int ldx, float* y, int ldy, int n) {
        avgl1_fwd_kernel<<<(B + 7) / 8, 256, 0, s>>>(x, ldx, y, ldy, B, n);
        t->launches++;
    }
    void avgl1_bwd(const float* dy, int lddy, const float* x, int ldx, float* dxp, int lddx, int n) {
        avgl1_bwd_kernel<<<(B + 7) / 8, 256, 0, s>>>(dy, lddy, x, ldx, dxp, lddx, B, n);
        t->launches++;
    }
    void copy(float* dst, int ldd, const float* src, int lds, int cols) {
        const size_t n = (size_t)B * cols;
        copy_cols_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(dst, ldd, src, lds, B, cols);
        t->launches++;
    }
    void adam(int opt) {
        const int net = OPT_NET[opt];
        const size_t n = t->def[kind_of(net)].size;
        if (t->hook) t->hook(opt, t->grad[opt], (int64_t)n, (void*)s, t->hook_user);      // data-parallel: all-reduce first
        adam_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(t->net[net], t->grad[opt], t->m[opt], t->v[opt], n, t->c.weight_decay,
                                                                t->dparams + DP_ADAM + 2 * opt);
        t->launches++;
    }
    void zero_grad(int opt) { cudaMemsetAsync(t->grad[opt], 0, t->def[kind_of(OPT_NET[opt])].size * sizeof(float), s); }

    // Encoder.zs (:101-105) -> zs (column slice allowed); caches h1, h2 (post-ELU) and z3 (pre-norm)
    void enc_zs(const float* net, const float* X, float* zs, int ldzs, float* h1, float* h2, float* z3) {
        const NetDef& d = t->def[2];
        const int S = t->c.state_dim, H = t->c.enc_hdim, Z = t->c.zs_dim;
        fwd(net, d.L[0], X, S, h1, H, ACT_ELU);
        fwd(net, d.L[1], h1, H, h2, H, ACT_ELU);
        fwd(net, d.L[2], h2, H, z3, Z, ACT_NONE);
        avgl1(z3, Z, zs, ldzs, Z);
    }
    // Encoder.zsa (:107-111) on a filled cat[zs, a]
    void enc_zsa(const float* net, const float* cat, float* g1, float* g2, float* out) {
        const NetDef& d = t->def[2];
        const int H = t->c.enc_hdim, Z = t->c.zs_dim, A = t->c.action_dim;
        fwd(net, d.L[3], cat, Z + A, g1, H, ACT_ELU);
        fwd(net, d.L[4], g1, H, g2, H, ACT_ELU);
        fwd(net, d.L[5], g2, H, out, Z, ACT_NONE);
    }
    // Actor.forward (:79-84); caches u (l0 pre-norm), cat[a0, zs], h1, h2 (post-ReLU)
    void actor_fwd(const float* net, const float* X, const float* zs, float* out) {
        const NetDef& d = t->def[0];
        const int S = t->c.state_dim, Ha = t->c.actor_hdim, Z = t->c.zs_dim, A = t->c.action_dim;
        fwd(net, d.L[0], X, S, t->a_u, Ha, ACT_NONE);
        avgl1(t->a_u, Ha, t->a_cat, Ha + Z, Ha);
        copy(t->a_cat + Ha, Ha + Z, zs, Z, Z);
        fwd(net, d.L[1], t->a_cat, Ha + Z, t->a_h1, Ha, ACT_RELU);
        fwd(net, d.L[2], t->a_h1, Ha, t->a_h2, Ha, ACT_RELU);
        fwd(net, d.L[3], t->a_h2, Ha, out, A, ACT_TANH);
    }
    // Critic.forward (:128-144): sa = cat[s, a], per head cat[AvgL1Norm(q0h(sa)), zsa, zs] -> q2[:, h]; the second head
    // runs on side stream 2
    void critic_fwd(const float* net, const float* X, const float* a, const float* zsa, const float* zs) {
        const NetDef& d = t->def[1];
        const int S = t->c.state_dim, A = t->c.action_dim, C = t->c.critic_hdim, Z = t->c.zs_dim, W = C + 2 * Z;
        copy(t->c_sa, S + A, X, S, S);
        copy(t->c_sa + S, S + A, a, A, A);
        Ctx h1{t, t->side[2], B};
        before(h1.s);
        for (int h = 0; h < 2; ++h) {
            Ctx& x = h ? h1 : *this;
            x.copy(t->c_cat[h] + C, W, zsa, Z, Z);
            x.copy(t->c_cat[h] + C + Z, W, zs, Z, Z);
            const Lin* L = &d.L[4 * h];
            x.fwd(net, L[0], t->c_sa, S + A, t->c_u[h], C, ACT_NONE);
            x.avgl1(t->c_u[h], C, t->c_cat[h], W, C);
            x.fwd(net, L[1], t->c_cat[h], W, t->c_h1[h], C, ACT_ELU);
            x.fwd(net, L[2], t->c_h1[h], C, t->c_h2[h], C, ACT_ELU);
            x.fwd(net, L[3], t->c_h2[h], C, t->q2 + h, 2, ACT_NONE);
        }
        h1.before(s);
    }
};

}  // namespace

#define TCK(call)                                                                                      \
    do {                                                                                               \
        cudaError_t e_ = (call);                                                                       \
        if (e_ != cudaSuccess) return exo_set_error(EXO_ERR_CUDA, "exo_td7: %s", cudaGetErrorString(e_)); \
    } while (0)

// One iteration without its host-side bookkeeping: everything here is stream work that a CUDA graph can replay.
static int td7_iteration(ExoTD7* t, const float* sS, const float* sA, const float* sN, const float* sR, const float* sD,
                         const float* randn_dev, float* prio, bool actor_half, cudaStream_t stream) {
    const ExoTD7Config& c = t->c;
    const int B = c.batch_rows, S = c.state_dim, A = c.action_dim, Z = c.zs_dim, H = c.enc_hdim, C = c.critic_hdim, Ha = c.actor_hdim;
    Ctx x{t, stream, B};
    cudaStream_t s = x.s;
    const NetDef &DA = t->def[0], &DC = t->def[1], &DE = t->def[2];
    float* enc = t->net[EXO_TD7_ENCODER];
    const float* fe = t->net[EXO_TD7_FIXED_ENCODER];
    const float* fet = t->net[EXO_TD7_FIXED_ENCODER_TARGET];
    float* critic = t->net[EXO_TD7_CRITIC];
    float* actor = t->net[EXO_TD7_ACTOR];
    TCK(cudaMemsetAsync(t->sc + SC_ENC_LOSS, 0, 2 * sizeof(float), s));

    // Three independent branches until the critic update needs Q_target and fixed_zsa: the encoder update stays on the
    // caller's stream, the target computation and the fixed-encoder passes run on two side streams (fork / join by events).
    TCK(cudaEventRecord(t->ev_fork, s));
    Ctx xt{t, t->side[0], B}, xf{t, t->side[1], B};
    TCK(cudaStreamWaitEvent(xt.s, t->ev_fork, 0));
    TCK(cudaStreamWaitEvent(xf.s, t->ev_fork, 0));

    // ---- critic target (:251-264), side stream 0
    xt.enc_zs(fet, sN, t->tzs, Z, t->t_h1, t->t_h2, t->t_z3);
    xt.actor_fwd(t->net[EXO_TD7_ACTOR_TARGET], sN, t->tzs, t->na);
    const float* rn = randn_dev;
    if (!rn) {
        randn_kernel<<<(B * A + 255) / 256, 256, 0, xt.s>>>(t->rn, B * A, c.seed, reinterpret_cast<const uint64_t*>(t->dparams + DP_STEP));
        t->launches++;
        rn = t->rn;
    }
    target_action_kernel<<<(B * A + 255) / 256, 256, 0, xt.s>>>(t->na, rn, B * A, t->dparams, c.noise_clip);
    t->launches++;
    xt.copy(t->t_cat, Z + A, t->tzs, Z, Z);
    xt.copy(t->t_cat + Z, Z + A, t->na, A, A);
    xt.enc_zsa(fet, t->t_cat, t->t_h1, t->t_h2, t->tzsa);
    xt.critic_fwd(t->net[EXO_TD7_CRITIC_TARGET], sN, t->na, t->tzsa, t->tzs);
    q_target_kernel<<<1, 1024, 0, xt.s>>>(t->q2, sR, sD, t->qt, B, c.discount, t->sc);
    t->launches++;
    TCK(cudaEventRecord(t->ev_join[0], xt.s));

    // ---- fixed_zs, fixed_zsa (:266-267), side stream 1
    xf.enc_zs(fe, sS, t->fzs, Z, t->f_h1, t->f_h2, t->f_z3);
    xf.copy(t->f_cat, Z + A, t->fzs, Z, Z);
    xf.copy(t->f_cat + Z, Z + A, sA, A, A);
    xf.enc_zsa(fe, t->f_cat, t->f_g1, t->f_g2, t->fzsa);
    TCK(cudaEventRecord(t->ev_join[1], xf.s));

    // ---- encoder update (:238-246), caller's stream
    x.enc_zs(enc, sN, t->next_zs, Z, t->n_h1, t->n_h2, t->n_z3);
    x.enc_zs(enc, sS, t->e_cat, Z + A, t->e_h1, t->e_h2, t->e_z3);
    x.copy(t->e_cat + Z, Z + A, sA, A, A);
    x.enc_zsa(enc, t->e_cat, t->e_g1, t->e_g2, t->e_pred);
    {
        const size_t n = (size_t)B * Z;
        mse_grad_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(t->e_pred, t->next_zs, t->G[0], n, t->sc);
        t->launches++;
    }
    x.zero_grad(EXO_TD7_OPT_ENCODER);
    {
        float* g = t->grad[EXO_TD7_OPT_ENCODER];
        float** G = t->G;
        x.dw_side(g, DE.L[5], G[0], Z, t->e_g2, H);
        x.dx(enc, DE.L[5], G[0], Z, G[1], H, 0, H, t->e_g2, H, ACT_ELU);
        x.dw_side(g, DE.L[4], G[1], H, t->e_g1, H);
        x.dx(enc, DE.L[4], G[1], H, G[2], H, 0, H, t->e_g1, H, ACT_ELU);
        x.dw_side(g, DE.L[3], G[2], H, t->e_cat, Z + A);
        x.dx(enc, DE.L[3], G[2], H, G[3], Z, 0, Z, nullptr, 0, ACT_NONE);
        x.avgl1_bwd(G[3], Z, t->e_z3, Z, G[3], Z, Z);
        x.dw_side(g, DE.L[2], G[3], Z, t->e_h2, H);
        x.dx(enc, DE.L[2], G[3], Z, G[4], H, 0, H, t->e_h2, H, ACT_ELU);
        x.dw_side(g, DE.L[1], G[4], H, t->e_h1, H);
        x.dx(enc, DE.L[1], G[4], H, G[5], H, 0, H, t->e_h1, H, ACT_ELU);
        x.dw_side(g, DE.L[0], G[5], H, sS, S);
        x.join_dw();
    }
    x.adam(EXO_TD7_OPT_ENCODER);

    // ---- join; critic update (:268-276) and priorities (:281)
    TCK(cudaStreamWaitEvent(s, t->ev_join[0], 0));
    TCK(cudaStreamWaitEvent(s, t->ev_join[1], 0));
    x.critic_fwd(critic, sS, sA, t->fzsa, t->fzs);
    td_head_kernel<<<(B + 255) / 256, 256, 0, s>>>(t->q2, t->qt, t->dq, prio, B, c.min_priority, c.alpha, t->sc);
    t->launches++;
    x.zero_grad(EXO_TD7_OPT_CRITIC);
    {
        Ctx h1{t, t->side[2], B};
        x.before(h1.s);
        for (int h = 0; h < 2; ++h) {
            Ctx& y = h ? h1 : x;
            float** G = h ? t->G2 : t->G;
            float* g = t->grad[EXO_TD7_OPT_CRITIC];
            const Lin* L = &DC.L[4 * h];
            const int W = C + 2 * Z;
            y.dw_side(g, L[3], t->dq + h, 2, t->c_h2[h], C);
            y.dx(critic, L[3], t->dq + h, 2, G[0], C, 0, C, t->c_h2[h], C, ACT_ELU);
            y.dw_side(g, L[2], G[0], C, t->c_h1[h], C);
            y.dx(critic, L[2], G[0], C, G[1], C, 0, C, t->c_h1[h], C, ACT_ELU);
            y.dw_side(g, L[1], G[1], C, t->c_cat[h], W);
            y.dx(critic, L[1], G[1], C, G[2], C, 0, C, nullptr, 0, ACT_NONE);
            y.avgl1_bwd(G[2], C, t->c_u[h], C, G[2], C, C);
            y.dw_side(g, L[0], G[2], C, t->c_sa, S + A);
        }
        h1.before(s);
        x.join_dw();
    }
    x.adam(EXO_TD7_OPT_CRITIC);

    // ---- actor update (:287-298)
    if (actor_half) {
        TCK(cudaMemsetAsync(t->sc + SC_ACTOR_LOSS, 0, 3 * sizeof(float), s));
        x.actor_fwd(actor, sS, t->fzs, t->act);
        x.copy(t->f_cat + Z, Z + A, t->act, A, A);
        x.enc_zsa(fe, t->f_cat, t->f_g1, t->f_g2, t->fzsa);
        x.critic_fwd(critic, sS, t->act, t->fzsa, t->fzs);
        actor_head_kernel<<<(2 * B + 255) / 256, 256, 0, s>>>(t->q2, t->dq, B, t->sc);
        t->launches++;
        {                                        // the critic's input gradients: d action (through sa) and d zsa
            Ctx h1{t, t->side[2], B};
            x.before(h1.s);
            for (int h = 0; h < 2; ++h) {
                Ctx& y = h ? h1 : x;
                float** G = h ? t->G2 : t->G;
                const Lin* L = &DC.L[4 * h];
                y.dx(critic, L[3], t->dq + h, 2, G[0], C, 0, C, t->c_h2[h], C, ACT_ELU);
                y.dx(critic, L[2], G[0], C, G[1], C, 0, C, t->c_h1[h], C, ACT_ELU);
                y.dx(critic, L[1], G[1], C, G[2], C, 0, C, nullptr, 0, ACT_NONE);
                y.avgl1_bwd(G[2], C, t->c_u[h], C, G[2], C, C);
                if (h) x.before(h1.s);           // head 0 has stored d_zsa and d_act1: head 1 adds to them
                y.dx(critic, L[1], G[1], C, t->d_zsa, Z, C, Z, nullptr, 0, ACT_NONE, h ? OUT_ADD : OUT_STORE);
                y.dx(critic, L[0], G[2], C, t->d_act1, A, S, A, nullptr, 0, ACT_NONE, h ? OUT_ADD : OUT_STORE);
            }
            h1.before(s);
        }
        x.dx(fe, DE.L[5], t->d_zsa, Z, t->dA, H, 0, H, t->f_g2, H, ACT_ELU);       // fixed_encoder.zsa's d action
        x.dx(fe, DE.L[4], t->dA, H, t->dB, H, 0, H, t->f_g1, H, ACT_ELU);
        x.dx(fe, DE.L[3], t->dB, H, t->d_act2, A, Z, A, nullptr, 0, ACT_NONE);
        tanh_bwd_kernel<<<(B * A + 255) / 256, 256, 0, s>>>(t->dT, t->d_act1, t->d_act2, t->act, B * A, sA,
                                                            c.offline ? c.lmbda * 2.0f / (float)(B * A) : 0.f, t->sc);
        t->launches++;
        x.zero_grad(EXO_TD7_OPT_ACTOR);
        float* g = t->grad[EXO_TD7_OPT_ACTOR];
        float** G = t->G;
        x.dw_side(g, DA.L[3], t->dT, A, t->a_h2, Ha);
        x.dx(actor, DA.L[3], t->dT, A, G[0], Ha, 0, Ha, t->a_h2, Ha, ACT_RELU);
        x.dw_side(g, DA.L[2], G[0], Ha, t->a_h1, Ha);
        x.dx(actor, DA.L[2], G[0], Ha, G[1], Ha, 0, Ha, t->a_h1, Ha, ACT_RELU);
        x.dw_side(g, DA.L[1], G[1], Ha, t->a_cat, Ha + Z);
        x.dx(actor, DA.L[1], G[1], Ha, G[2], Ha, 0, Ha, nullptr, 0, ACT_NONE);
        x.avgl1_bwd(G[2], Ha, t->a_u, Ha, G[2], Ha, Ha);
        x.dw_side(g, DA.L[0], G[2], Ha, sS, S);
        x.join_dw();
        x.adam(EXO_TD7_OPT_ACTOR);
    }

    (void)DA;
    return EXO_OK;
}

extern "C" int exo_td7_train(ExoTD7* t, const float* sS, const float* sA, const float* sN, const float* sR, const float* sD,
                             const float* randn_dev, float* prio, int32_t* rolled_over, void* stream) {
    if (!t || !sS || !sA || !sN || !sR || !sD || !prio) return exo_set_error(EXO_ERR_ARG, "exo_td7_train: null argument");
    TCK(cudaSetDevice(t->device));
    const ExoTD7Config& c = t->c;
    cudaStream_t s = (cudaStream_t)stream;
    t->steps++;
    const bool actor_half = t->steps % c.policy_freq == 0;
    // this iteration's scalars: noise scale, Adam's bias-corrected step sizes, the generator's step
    float* hp = t->hparams + (size_t)(t->steps & 63) * DP_COUNT;
    hp[DP_SIGMA] = (float)t->sigma;
    t->sigma -= (double)c.policy_noise_decrease;
    const float lrs[3] = {c.actor_lr, c.critic_lr, c.encoder_lr};
    for (int o = 0; o < 3; ++o) {
        if (o == EXO_TD7_OPT_ACTOR && !actor_half) continue;
        const int64_t k = ++t->opt_step[o];
        hp[DP_ADAM + 2 * o] = (float)((double)lrs[o] / (1.0 - pow(0.9, (double)k)));
        hp[DP_ADAM + 2 * o + 1] = (float)sqrt(1.0 - pow(0.999, (double)k));
    }
    const uint64_t st = (uint64_t)t->steps;
    memcpy(hp + DP_STEP, &st, sizeof(st));
    TCK(cudaMemcpyAsync(t->dparams, hp, DP_COUNT * sizeof(float), cudaMemcpyHostToDevice, s));

    const void* key[7] = {sS, sA, sN, sR, sD, randn_dev, prio};
    if (t->use_graph && !t->hook) {
        ExoTD7::Graph* g = nullptr;
        for (auto& e : t->graphs)
            if (e.actor == actor_half && !memcmp(e.key, key, sizeof(key))) { g = &e; break; }
        if (!g && t->graphs.size() < 32) {
            // capture on the library's own stream (the caller's may be the legacy default stream, which cannot be captured)
            const int64_t l0 = t->launches;
            TCK(cudaStreamBeginCapture(t->gstream, cudaStreamCaptureModeThreadLocal));
            int rc = td7_iteration(t, sS, sA, sN, sR, sD, randn_dev, prio, actor_half, t->gstream);
            cudaGraph_t graph = nullptr;
            cudaError_t e = cudaStreamEndCapture(t->gstream, &graph);
            const int64_t kernels = t->launches - l0;
            t->launches = l0;
            if (rc != EXO_OK) { if (graph) cudaGraphDestroy(graph); return rc; }
            if (e != cudaSuccess) return exo_set_error(EXO_ERR_CUDA, "exo_td7_train: graph capture: %s", cudaGetErrorString(e));
            cudaGraphExec_t exec = nullptr;
            e = cudaGraphInstantiate(&exec, graph, 0);
            cudaGraphDestroy(graph);
            if (e != cudaSuccess) return exo_set_error(EXO_ERR_CUDA, "exo_td7_train: graph instantiate: %s", cudaGetErrorString(e));
            ExoTD7::Graph ng;
            memcpy(ng.key, key, sizeof(key));
            ng.actor = actor_half; ng.exec = exec; ng.kernels = kernels;
            t->graphs.push_back(ng);
            g = &t->graphs.back();
        }
        if (g) {
            TCK(cudaEventRecord(t->g_in, s));
            TCK(cudaStreamWaitEvent(t->gstream, t->g_in, 0));
            TCK(cudaGraphLaunch(g->exec, t->gstream));
            TCK(cudaEventRecord(t->g_out, t->gstream));
            TCK(cudaStreamWaitEvent(s, t->g_out, 0));
            t->launches += g->kernels;
        } else if (int rc = td7_iteration(t, sS, sA, sN, sR, sD, randn_dev, prio, actor_half, s)) {
            return rc;
        }
    } else if (int rc = td7_iteration(t, sS, sA, sN, sR, sD, randn_dev, prio, actor_half, s)) {
        return rc;
    }

    // ---- roll-over (:303-312)
    const NetDef &DA = t->def[0], &DC = t->def[1], &DE = t->def[2];
    const bool roll = t->steps % c.target_update_rate == 0;
    if (roll) {
        TCK(cudaMemcpyAsync(t->net[EXO_TD7_ACTOR_TARGET], t->net[EXO_TD7_ACTOR], DA.size * sizeof(float), cudaMemcpyDeviceToDevice, s));
        TCK(cudaMemcpyAsync(t->net[EXO_TD7_CRITIC_TARGET], t->net[EXO_TD7_CRITIC], DC.size * sizeof(float), cudaMemcpyDeviceToDevice, s));
        TCK(cudaMemcpyAsync(t->net[EXO_TD7_FIXED_ENCODER_TARGET], t->net[EXO_TD7_FIXED_ENCODER], DE.size * sizeof(float), cudaMemcpyDeviceToDevice, s));
        TCK(cudaMemcpyAsync(t->net[EXO_TD7_FIXED_ENCODER], t->net[EXO_TD7_ENCODER], DE.size * sizeof(float), cudaMemcpyDeviceToDevice, s));
        TCK(cudaMemcpyAsync(t->sc + SC_MAX_TARGET, t->sc + SC_MAX, 2 * sizeof(float), cudaMemcpyDeviceToDevice, s));
    }
    if (rolled_over) *rolled_over = roll ? 1 : 0;
    TCK(cudaGetLastError());
    return EXO_OK;
}

extern "C" int exo_td7_destroy(ExoTD7* t) {
    if (!t) return EXO_OK;
    cudaSetDevice(t->device);
    cudaDeviceSynchronize();
    for (float* p : t->owned) cudaFree(p);
    for (int i = 0; i < 4; ++i)
        if (t->side[i]) cudaStreamDestroy(t->side[i]);
    for (int i = 0; i < 2; ++i)
        if (t->ev_join[i]) cudaEventDestroy(t->ev_join[i]);
    for (int i = 0; i < 32; ++i)
        if (t->evs[i]) cudaEventDestroy(t->evs[i]);
    if (t->ev_fork) cudaEventDestroy(t->ev_fork);
    for (auto& g : t->graphs) cudaGraphExecDestroy(g.exec);
    if (t->gstream) cudaStreamDestroy(t->gstream);
    if (t->g_in) cudaEventDestroy(t->g_in);
    if (t->g_out) cudaEventDestroy(t->g_out);
    if (t->hparams) cudaFreeHost(t->hparams);
    delete t;
    return EXO_OK;
}

extern "C" int exo_td7_create(int32_t device, const ExoTD7Config* cfg, ExoTD7** out) {
    if (!out || !cfg) return exo_set_error(EXO_ERR_ARG, "exo_td7_create: null argument");
    const ExoTD7Config& c = *cfg;
    if (c.state_dim < 1 || c.action_dim < 1 || c.zs_dim < 1 || c.enc_hdim < 1 || c.critic_hdim < 1 || c.actor_hdim < 1 || c.batch_rows < 1 ||
        c.target_update_rate < 1 || c.policy_freq < 1)
        return exo_set_error(EXO_ERR_ARG, "exo_td7_create: sizes and periods must be positive");
    int ndev = 0;
    TCK(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return exo_set_error(EXO_ERR_CUDA, "device %d not present: this library has no CPU path", device);
    cudaDeviceProp prop;
    TCK(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) return exo_set_error(EXO_ERR_CUDA, "device %d is sm_%d%d; libexo_env.so carries sm_100a code only", device, prop.major, prop.minor);
    TCK(cudaSetDevice(device));
    ExoTD7* t = new ExoTD7();
    t->device = device; t->n_sm = prop.multiProcessorCount; t->c = c; t->sigma = c.target_policy_noise;
    const char* fm = getenv("EXO_TD7_FFMA");
    t->ffma = fm && fm[0] != '0';
    const char* gemm_mode = getenv("EXO_TD7_GEMM");
    t->tc5 = !(gemm_mode && !strcmp(gemm_mode, "mma"));       // default: tcgen05; EXO_TD7_GEMM=mma keeps the mma.sync kernel
    if (const char* ck = getenv("EXO_TD7_TC5_CLUSTER")) t->tc5_cluster = atoi(ck);
    const int B = c.batch_rows, S = c.state_dim, A = c.action_dim, Z = c.zs_dim, H = c.enc_hdim, C = c.critic_hdim, Ha = c.actor_hdim;
    t->def[0] = make_def({{Ha, S}, {Ha, Z + Ha}, {Ha, Ha}, {A, Ha}});
    t->def[1] = make_def({{C, S + A}, {C, 2 * Z + C}, {C, C}, {1, C}, {C, S + A}, {C, 2 * Z + C}, {C, C}, {1, C}});
    t->def[2] = make_def({{H, S}, {H, H}, {Z, H}, {H, Z + A}, {H, H}, {Z, H}});
    cudaError_t err = cudaSuccess;
    auto alloc = [&](size_t n) -> float* {
        float* p = nullptr;
        if (err != cudaSuccess) return nullptr;
        err = cudaMalloc((void**)&p, std::max<size_t>(n, 1) * sizeof(float));
        if (err == cudaSuccess) { t->owned.push_back(p); err = cudaMemset(p, 0, std::max<size_t>(n, 1) * sizeof(float)); }
        return p;
    };
    for (int n = 0; n < EXO_TD7_N_NETS; ++n) t->net[n] = alloc(t->def[kind_of(n)].size);
    for (int o = 0; o < 3; ++o) {
        const size_t n = t->def[kind_of(OPT_NET[o])].size;
        t->grad[o] = alloc(n); t->m[o] = alloc(n); t->v[o] = alloc(n);
    }
    t->sc = alloc(SC_COUNT);
    auto rows = [&](int w) { return alloc((size_t)B * w); };
    t->t_h1 = rows(H); t->t_h2 = rows(H); t->t_z3 = rows(Z); t->next_zs = rows(Z);
    t->n_h1 = rows(H); t->n_h2 = rows(H); t->n_z3 = rows(Z); t->f_h1 = rows(H); t->f_h2 = rows(H); t->f_z3 = rows(Z);
    for (int i = 0; i < 4 && err == cudaSuccess; ++i) err = cudaStreamCreateWithFlags(&t->side[i], cudaStreamNonBlocking);
    for (int i = 0; i < 2 && err == cudaSuccess; ++i) err = cudaEventCreateWithFlags(&t->ev_join[i], cudaEventDisableTiming);
    for (int i = 0; i < 32 && err == cudaSuccess; ++i) err = cudaEventCreateWithFlags(&t->evs[i], cudaEventDisableTiming);
    if (err == cudaSuccess) err = cudaEventCreateWithFlags(&t->ev_fork, cudaEventDisableTiming);
    t->dparams = alloc(DP_COUNT + 2);
    t->dstep_scratch = reinterpret_cast<uint64_t*>(t->dparams + DP_COUNT);
    if (err == cudaSuccess) err = cudaMallocHost((void**)&t->hparams, 64 * DP_COUNT * sizeof(float));
    if (err == cudaSuccess) err = cudaStreamCreateWithFlags(&t->gstream, cudaStreamNonBlocking);
    if (err == cudaSuccess) err = cudaEventCreateWithFlags(&t->g_in, cudaEventDisableTiming);
    if (err == cudaSuccess) err = cudaEventCreateWithFlags(&t->g_out, cudaEventDisableTiming);
    const char* gm = getenv("EXO_TD7_GRAPH");
    t->use_graph = !(gm && gm[0] == '0');
    t->e_h1 = rows(H); t->e_h2 = rows(H); t->e_z3 = rows(Z); t->e_cat = rows(Z + A); t->e_g1 = rows(H); t->e_g2 = rows(H); t->e_pred = rows(Z);
    t->tzs = rows(Z); t->t_cat = rows(Z + A); t->tzsa = rows(Z); t->na = rows(A); t->rn = rows(A);
    t->a_u = rows(Ha); t->a_cat = rows(Ha + Z); t->a_h1 = rows(Ha); t->a_h2 = rows(Ha); t->act = rows(A);
    t->fzs = rows(Z); t->f_cat = rows(Z + A); t->f_g1 = rows(H); t->f_g2 = rows(H); t->fzsa = rows(Z);
    t->c_sa = rows(S + A);
    for (int h = 0; h < 2; ++h) { t->c_u[h] = rows(C); t->c_cat[h] = rows(C + 2 * Z); t->c_h1[h] = rows(C); t->c_h2[h] = rows(C); }
    t->q2 = rows(2); t->qt = rows(1); t->dq = rows(2);
    const int wmax = std::max(std::max(H, C), std::max(Ha, Z));
    t->dA = rows(wmax); t->dB = rows(wmax);
    for (int i = 0; i < 6; ++i) t->G[i] = rows(wmax);
    for (int i = 0; i < 4; ++i) t->G2[i] = rows(wmax); t->d_zsa = rows(Z); t->d_act1 = rows(A); t->d_act2 = rows(A); t->dT = rows(A);
    if (err == cudaSuccess) {
        const float init[SC_COUNT] = {-1e8f, 1e8f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};      // Agent.__init__ :179-182
        err = cudaMemcpy(t->sc, init, sizeof(init), cudaMemcpyHostToDevice);
    }
    if (err != cudaSuccess) {
        exo_td7_destroy(t);
        return exo_set_error(EXO_ERR_CUDA, "exo_td7_create: %s", cudaGetErrorString(err));
    }
    *out = t;
    return EXO_OK;
}

extern "C" int64_t exo_td7_net_size(const ExoTD7* t, int32_t net) {
    if (!t || net < 0 || net >= EXO_TD7_N_NETS) return -1;
    return (int64_t)t->def[kind_of(net)].size;
}

static int td7_vec(ExoTD7* t, float* dev, size_t size, const float* src, float* dst, int64_t n, const char* what) {
    if (!t || (!src && !dst)) return exo_set_error(EXO_ERR_ARG, "%s: null argument", what);
    if ((size_t)n != size) return exo_set_error(EXO_ERR_ARG, "%s: vector has %lld floats, the network has %zu", what, (long long)n, size);
    TCK(cudaSetDevice(t->device));
    TCK(cudaDeviceSynchronize());
    if (src) TCK(cudaMemcpy(dev, src, size * sizeof(float), cudaMemcpyHostToDevice));
    else TCK(cudaMemcpy(dst, dev, size * sizeof(float), cudaMemcpyDeviceToHost));
    return EXO_OK;
}

extern "C" int exo_td7_set_net(ExoTD7* t, int32_t net, const float* host, int64_t n) {
    if (!t || net < 0 || net >= EXO_TD7_N_NETS) return exo_set_error(EXO_ERR_ARG, "exo_td7_set_net: bad network id");
    return td7_vec(t, t->net[net], t->def[kind_of(net)].size, host, nullptr, n, "exo_td7_set_net");
}
extern "C" int exo_td7_get_net(ExoTD7* t, int32_t net, float* host, int64_t n) {
    if (!t || net < 0 || net >= EXO_TD7_N_NETS) return exo_set_error(EXO_ERR_ARG, "exo_td7_get_net: bad network id");
    return td7_vec(t, t->net[net], t->def[kind_of(net)].size, nullptr, host, n, "exo_td7_get_net");
}
extern "C" int exo_td7_copy_net(ExoTD7* t, int32_t dst, int32_t src) {
    if (!t || dst < 0 || dst >= EXO_TD7_N_NETS || src < 0 || src >= EXO_TD7_N_NETS || kind_of(dst) != kind_of(src))
        return exo_set_error(EXO_ERR_ARG, "exo_td7_copy_net: the two networks must be of the same kind");
    TCK(cudaSetDevice(t->device));
    TCK(cudaDeviceSynchronize());
    if (dst != src) TCK(cudaMemcpy(t->net[dst], t->net[src], t->def[kind_of(src)].size * sizeof(float), cudaMemcpyDeviceToDevice));
    return EXO_OK;
}

static float* td7_opt_vec(ExoTD7* t, int32_t opt, int32_t which) {
    if (!t || opt < 0 || opt > 2) return nullptr;
    return which == EXO_TD7_EXP_AVG ? t->m[opt] : which == EXO_TD7_EXP_AVG_SQ ? t->v[opt] : which == EXO_TD7_GRAD ? t->grad[opt] : nullptr;
}
extern "C" int exo_td7_set_opt(ExoTD7* t, int32_t opt, int32_t which, const float* host, int64_t n, int64_t step) {
    float* d = td7_opt_vec(t, opt, which);
    if (!d || which == EXO_TD7_GRAD || step < 0) return exo_set_error(EXO_ERR_ARG, "exo_td7_set_opt: bad optimiser / vector id or step");
    if (int rc = td7_vec(t, d, t->def[kind_of(OPT_NET[opt])].size, host, nullptr, n, "exo_td7_set_opt")) return rc;
    t->opt_step[opt] = step;
    return EXO_OK;
}
extern "C" int exo_td7_get_opt(ExoTD7* t, int32_t opt, int32_t which, float* host, int64_t n, int64_t* step) {
    float* d = td7_opt_vec(t, opt, which);
    if (!d) return exo_set_error(EXO_ERR_ARG, "exo_td7_get_opt: bad optimiser / vector id");
    if (step) *step = t->opt_step[opt];
    return td7_vec(t, d, t->def[kind_of(OPT_NET[opt])].size, nullptr, host, n, "exo_td7_get_opt");
}

extern "C" int exo_td7_scalars(ExoTD7* t, double* out) {
    if (!t || !out) return exo_set_error(EXO_ERR_ARG, "exo_td7_scalars: null argument");
    TCK(cudaSetDevice(t->device));
    TCK(cudaDeviceSynchronize());
    float h[SC_COUNT];
    TCK(cudaMemcpy(h, t->sc, sizeof(h), cudaMemcpyDeviceToHost));
    out[0] = h[SC_MAX]; out[1] = h[SC_MIN]; out[2] = h[SC_MAX_TARGET]; out[3] = h[SC_MIN_TARGET];
    out[4] = t->sigma; out[5] = (double)t->steps;
    out[6] = h[SC_ENC_LOSS]; out[7] = h[SC_CRITIC_LOSS];
    out[8] = h[SC_ACTOR_LOSS] + (t->c.offline ? t->c.lmbda * h[SC_ABSQ] * h[SC_BC_MSE] : 0.f);
    return EXO_OK;
}
extern "C" int exo_td7_set_scalars(ExoTD7* t, const double* in) {
    if (!t || !in) return exo_set_error(EXO_ERR_ARG, "exo_td7_set_scalars: null argument");
    TCK(cudaSetDevice(t->device));
    TCK(cudaDeviceSynchronize());
    const float h[4] = {(float)in[0], (float)in[1], (float)in[2], (float)in[3]};
    TCK(cudaMemcpy(t->sc, h, sizeof(h), cudaMemcpyHostToDevice));
    t->sigma = in[4];
    t->steps = (int64_t)in[5];
    return EXO_OK;
}

extern "C" int exo_td7_set_grad_hook(ExoTD7* t, ExoTD7GradHook hook, void* user) {
    if (!t) return exo_set_error(EXO_ERR_ARG, "exo_td7_set_grad_hook: null handle");
    t->hook = hook;
    t->hook_user = user;
    return EXO_OK;
}

extern "C" int exo_td7_randn(ExoTD7* t, float* out_dev, int64_t count, int64_t step, void* stream) {
    if (!t || !out_dev || count < 0 || count > 0x7FFFFFFF) return exo_set_error(EXO_ERR_ARG, "exo_td7_randn: bad argument");
    TCK(cudaSetDevice(t->device));
    const uint64_t st = (uint64_t)step;
    TCK(cudaMemcpyAsync(t->dstep_scratch, &st, sizeof(st), cudaMemcpyHostToDevice, (cudaStream_t)stream));     // pageable source: staged before return
    if (count) randn_kernel<<<(unsigned)((count + 255) / 256), 256, 0, (cudaStream_t)stream>>>(out_dev, (int)count, t->c.seed, t->dstep_scratch);
    t->launches++;
    TCK(cudaGetLastError());
    return EXO_OK;
}

extern "C" int64_t exo_td7_launch_count(const ExoTD7* t) { return t ? t->launches : 0; }
