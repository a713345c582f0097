// tc5_gemm.cu — the TD7 update's tcgen05 GEMM (csrc/exo_td7_tc5.cuh) alone: time per launch for the update's shapes and a
// cycle timeline of CTA (0,0,0) (marks compiled in with -DTC5_TIMELINE).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -DTC5_TIMELINE -o tools/ubench/tc5_gemm tools/ubench/tc5_gemm.cu
//   gpurun -- tools/ubench/tc5_gemm
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <cuda_runtime.h>

namespace {
enum { ACT_NONE = 0, ACT_RELU = 1, ACT_ELU = 2, ACT_TANH = 3 };
enum { OUT_STORE = 0, OUT_ADD = 1, OUT_ATOMIC = 2 };
struct GemmArgs {
    const float* A; int lda;
    const float* B; int ldb;
    float* C; int ldc;
    int M, N, K, kchunk;
    const float* bias; int act;
    const float* D; int ldd; int dact;
    int out;
};
__device__ __forceinline__ float act_fwd(float v, int act) {
    if (act == ACT_RELU) return fmaxf(v, 0.f);
    if (act == ACT_ELU) return v > 0.f ? v : expm1f(v);
    if (act == ACT_TANH) return tanhf(v);
    return v;
}
__device__ __forceinline__ float act_bwd(float y, int act) {
    if (act == ACT_RELU) return y > 0.f ? 1.f : 0.f;
    if (act == ACT_ELU) return y > 0.f ? 1.f : y + 1.f;
    if (act == ACT_TANH) return 1.f - y * y;
    return 1.f;
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
#include "../../a-deep-reinforcement-learning-enabled-soft-exoskeleton-for-parkinson-s-patients_b200/csrc/exo_td7_tc5.cuh"
}  // namespace

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

struct Case { const char* name; bool ta, tb; int M, N, K, lda, ldb, tn, splits; bool bias; int act; bool deriv; int out; };
static int g_ck = 1;

int main(int argc, char** argv) {
    if (argc > 1) g_ck = atoi(argv[1]);
    printf("cluster split-k %d (store outputs)\n", g_ck);
    const int rows = 1920;
    std::vector<Case> cases = {
        {"fwd 1920x300x300 elu tn80", false, true, rows, 300, 300, 300, 300, 80, 1, true, ACT_ELU, false, OUT_STORE},
        {"fwd 1920x300x300 none tn80", false, true, rows, 300, 300, 300, 300, 80, 1, false, ACT_NONE, false, OUT_STORE},
        {"fwd 1920x300x300 elu tn64", false, true, rows, 300, 300, 300, 300, 64, 1, true, ACT_ELU, false, OUT_STORE},
        {"fwd 1920x320x920 elu tn80", false, true, rows, 320, 920, 920, 920, 80, 1, true, ACT_ELU, false, OUT_STORE},
        {"fwd 1920x7x300 tanh tn64", false, true, rows, 7, 300, 300, 300, 64, 1, true, ACT_TANH, false, OUT_STORE},
        {"dx  1920x300x300 elu' tn80", false, false, rows, 300, 300, 300, 300, 80, 1, false, ACT_NONE, true, OUT_STORE},
        {"dw  300x300x1920 split8 tn80", true, false, 300, 300, rows, 300, 300, 80, 8, false, ACT_NONE, false, OUT_ATOMIC},
    };
    float *A, *B, *C, *D, *bias;
    const size_t big = (size_t)rows * 1024;
    CK(cudaMalloc(&A, big * 4)); CK(cudaMalloc(&B, big * 4)); CK(cudaMalloc(&C, big * 4)); CK(cudaMalloc(&D, big * 4)); CK(cudaMalloc(&bias, 4096));
    std::vector<float> h(big);
    for (size_t i = 0; i < big; i++) h[i] = (float)((i * 2654435761u) % 2001) / 1000.f - 1.f;
    CK(cudaMemcpy(A, h.data(), big * 4, cudaMemcpyHostToDevice)); CK(cudaMemcpy(B, h.data() + 7, (big - 7) * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(D, h.data() + 13, (big - 13) * 4, cudaMemcpyHostToDevice)); CK(cudaMemcpy(bias, h.data() + 3, 4096, cudaMemcpyHostToDevice));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for (const Case& c : cases) {
        GemmArgs g = {};
        g.A = A; g.lda = c.lda; g.B = B; g.ldb = c.ldb; g.C = C; g.ldc = c.N; g.M = c.M; g.N = c.N; g.K = c.K;
        g.bias = c.bias ? bias : nullptr; g.act = c.act; g.D = c.deriv ? D : nullptr; g.ldd = c.N; g.dact = ACT_ELU; g.out = c.out;
        g.kchunk = c.splits <= 1 ? c.K : (((c.K + c.splits - 1) / c.splits + tc5::KA - 1) / tc5::KA) * tc5::KA;
        const int ck = c.out == OUT_ATOMIC ? 1 : g_ck;
        dim3 grid((c.N + c.tn - 1) / c.tn, (c.M + 127) / 128, ck > 1 ? ck : (c.K + g.kchunk - 1) / g.kchunk);
        auto go = [&]() {
            if (!c.ta && c.tb) CK((tc5::launch<false, true>(g, c.tn, ck, grid, 0)));
            else if (!c.ta && !c.tb) CK((tc5::launch<false, false>(g, c.tn, ck, grid, 0)));
            else CK((tc5::launch<true, false>(g, c.tn, ck, grid, 0)));
        };
        // sampled check against a float64 product on the host
        CK(cudaMemset(C, 0, big * 4));
        go();
        CK(cudaDeviceSynchronize());
        {
            std::vector<float> hc((size_t)c.M * c.N);
            CK(cudaMemcpy(hc.data(), C, hc.size() * 4, cudaMemcpyDeviceToHost));
            double worst = 0;
            for (int sidx = 0; sidx < 4000; sidx++) {
                const int m = (int)((sidx * 7919u) % (unsigned)c.M), n = (int)((sidx * 104729u) % (unsigned)c.N);
                double acc = 0;
                for (int k = 0; k < c.K; k++) {
                    const double a = c.ta ? h[(size_t)k * c.lda + m] : h[(size_t)m * c.lda + k];
                    const double b = c.tb ? h[7 + (size_t)n * c.ldb + k] : h[7 + (size_t)k * c.ldb + n];
                    acc += a * b;
                }
                if (c.bias) acc += h[3 + n];
                if (c.act == ACT_ELU) acc = acc > 0 ? acc : expm1(acc);
                if (c.act == ACT_TANH) acc = tanh(acc);
                if (c.deriv) { const double y = h[13 + (size_t)m * c.N + n]; acc *= y > 0 ? 1.0 : y + 1.0; }
                worst = fmax(worst, fabs(acc - (double)hc[(size_t)m * c.N + n]) / fmax(1.0, fabs(acc)));
            }
            printf("%-32s max rel err over 4000 sampled outputs %.2e\n", c.name, worst);
        }
        for (int i = 0; i < 5; i++) go();
        CK(cudaDeviceSynchronize());
        const int reps = 50;
        CK(cudaEventRecord(e0));
        for (int i = 0; i < reps; i++) go();
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("%-32s grid (%d,%d,%d)  %.2f us per launch (back to back, warm L2)\n", c.name, grid.x, grid.y, grid.z, ms * 1000.f / reps);
#ifdef TC5_TIMELINE
        long long mk[64];
        CK(cudaMemcpyFromSymbol(mk, tc5::tc5_marks, sizeof(mk)));
        const int na = ((g.kchunk + 31) / 32 + ck - 1) / ck;
        printf("   cycles from kernel entry (warp 0): set-up done %lld | accumulator done %lld, tile in smem %lld, bias loaded %lld, first 4 rows %lld, stored %lld\n",
               mk[1] - mk[0], mk[3] - mk[0], mk[4] - mk[0], mk[6] - mk[0], mk[7] - mk[0], mk[5] - mk[0]);
        for (int a = 0; a < na && 11 + a * 4 < 64; a++)
            printf("   atom %2d: stage free @%lld, packed + arrived +%lld, next loads issued +%lld\n", a, mk[8 + a * 4] - mk[0],
                   mk[9 + a * 4] - mk[8 + a * 4], mk[10 + a * 4] - mk[9 + a * 4]);
#endif
    }
    return 0;
}
