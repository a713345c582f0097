// exo_lap.cu — B200-resident prioritized replay buffer (include/exo_lap.h): the reference's LAP
// (Agent/TD7_buffer_multi_agent.py:5-150) with storage, pointer arithmetic, sampling and priority updates on the
// device, so transitions go from the env's step kernel into the buffer and out to the learner without a host copy.
//
// HBM layout: state / next_state [num_envs][max_size][state_dim] f32 (one transition's 320 B contiguous: a warp
// copies it with coalesced 128-bit accesses), action [num_envs][max_size][action_dim], reward / not_done / priority
// [num_envs][max_size]. Counters (ptr, count, size, max_priority) are device-resident; kernels read them, a final
// one-thread kernel advances them, nothing synchronises the stream.
// All kernels are HBM-bound copies/reductions: add moves 2*sd + ad + 3 words per transition, sample reads the
// priority rows once (4 B per stored transition) plus the gathered rows.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <vector>

#include "../../include/exo_env.h"
#include "../../include/exo_lap.h"

extern "C" const char* exo_last_error(void);
int exo_set_error(int code, const char* fmt, ...);     // exo_api.cu

#define LCK(call)                                                                                          \
    do {                                                                                                   \
        cudaError_t _e = (call);                                                                           \
        if (_e != cudaSuccess) return exo_set_error(EXO_ERR_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(_e)); \
    } while (0)

namespace {

constexpr int LAP_CHUNK = 1024;          // priorities per chunk sum (one 256-thread block, one float4 per thread)

struct LapCounters {
    long long ptr, count, size;          // LAP.ptr, LAP.count, LAP.size
    float max_priority;                  // LAP.max_priority
    int kadd;                            // entries actually added by the call in flight
};

struct LapView {
    float *state, *action, *next_state, *reward, *not_done, *priority;
    LapCounters* ctr;
    int sd, ad, num_envs, batch;
    long long max_size;
    float norm;                          // normalize_actions (actions are stored divided by it, :51)
};

// number of multiples of E in [a, b)
__host__ __device__ inline long long multiples_in(long long a, long long b, long long E) {
    return (b + E - 1) / E - (a + E - 1) / E;
}

// ---- add ------------------------------------------------------------------------------------------------
// rank[j] = number of non-skipped entries before j; one block, ballot scan, k / 1024 rounds
__global__ void __launch_bounds__(1024) lap_rank_kernel(const uint8_t* __restrict__ skip, int k, int* __restrict__ rank,
                                                        LapCounters* ctr) {
    __shared__ int warp_tot[32];
    __shared__ int base_s;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) base_s = 0;
    __syncthreads();
    for (int j0 = 0; j0 < k; j0 += 1024) {
        const int j = j0 + tid;
        const bool keep = j < k && skip[j] == 0;
        const unsigned m = __ballot_sync(0xffffffffu, keep);
        if (lane == 0) warp_tot[wid] = __popc(m);
        __syncthreads();
        int before = 0;
        for (int w = 0; w < wid; w++) before += warp_tot[w];
        if (j < k) rank[j] = keep ? base_s + before + __popc(m & ((1u << lane) - 1u)) : -1;
        __syncthreads();
        if (tid == 0) { int t = 0; for (int w = 0; w < 32; w++) t += warp_tot[w]; base_s += t; }
        __syncthreads();
    }
    if (tid == 0) ctr->kadd = base_s;
}

__device__ __forceinline__ int lap_row_of(const int32_t* tremor_num, int j, int E) { return tremor_num ? tremor_num[j] : j % E; }

// winner[(window, row)] = highest rank that writes that (row, slot): sequential last-writer-wins
__global__ void lap_winner_kernel(LapView B, const int32_t* __restrict__ tremor_num, const int* __restrict__ rank, int k,
                                  int* __restrict__ winner, int nwin) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= k) return;
    const int r = rank ? rank[j] : j;
    if (r < 0) return;
    const long long c0 = B.ctr->count, c = c0 + r;
    const long long win = multiples_in(c0, c, B.num_envs);
    const int row = lap_row_of(tremor_num, j, B.num_envs);
    if (row < 0 || row >= B.num_envs || win >= nwin) return;
    atomicMax(&winner[win * B.num_envs + row], r);
}

// one warp per entry: copy the transition into (row, slot)
__global__ void __launch_bounds__(256) lap_add_kernel(LapView B, const float* __restrict__ state, const float* __restrict__ action,
                                                      const float* __restrict__ next_state, const float* __restrict__ reward,
                                                      const uint8_t* __restrict__ done, const int32_t* __restrict__ tremor_num,
                                                      const int* __restrict__ rank, int k, const int* __restrict__ winner, int nwin) {
    const int j = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (j >= k) return;
    const int sd = B.sd, ad = B.ad;
    const float* s = state + (size_t)j * sd;
    const float* ns = next_state + (size_t)j * sd;
    // the common shape (state_dim <= 128, 16-byte aligned rows): the transition's payload is requested before the
    // counters / rank / winner chain resolves, so the copy pays one memory latency instead of three
    const bool vec = (sd & 3) == 0 && sd <= 128 && ((reinterpret_cast<size_t>(s) | reinterpret_cast<size_t>(ns)) & 15) == 0;
    float4 vs = make_float4(0, 0, 0, 0), vn = vs;
    if (vec && lane < (sd >> 2)) { vs = __ldg(reinterpret_cast<const float4*>(s) + lane); vn = __ldg(reinterpret_cast<const float4*>(ns) + lane); }
    const float va = lane < ad ? action[(size_t)j * ad + lane] : 0.0f;
    const float vr = reward[j];
    const uint8_t vd = done[j];
    const int r = rank ? rank[j] : j;
    if (r < 0) return;
    const long long c0 = B.ctr->count, c = c0 + r;
    const long long win = multiples_in(c0, c, B.num_envs);
    const int row = lap_row_of(tremor_num, j, B.num_envs);
    if (row < 0 || row >= B.num_envs || win >= nwin) return;
    if (winner[win * B.num_envs + row] != r) return;                       // overwritten later in this same call
    const long long slot = (B.ctr->ptr + win) % B.max_size;
    const long long t = (long long)row * B.max_size + slot;
    float* ds = B.state + t * sd;
    float* dns = B.next_state + t * sd;
    if (vec) {
        if (lane < (sd >> 2)) { reinterpret_cast<float4*>(ds)[lane] = vs; reinterpret_cast<float4*>(dns)[lane] = vn; }
    } else {
        for (int i = lane; i < sd; i += 32) { ds[i] = s[i]; dns[i] = ns[i]; }
    }
    if (lane < ad) B.action[t * ad + lane] = va / B.norm;
    for (int i = lane + 32; i < ad; i += 32) B.action[t * ad + i] = action[(size_t)j * ad + i] / B.norm;
    if (lane == 0) {
        B.reward[t] = vr;
        B.not_done[t] = 1.0f - (vd ? 1.0f : 0.0f);
        B.priority[t] = B.ctr->max_priority;
    }
}

__global__ void lap_advance_kernel(LapView B, int k_if_no_skip) {
    LapCounters* c = B.ctr;
    const long long kadd = k_if_no_skip >= 0 ? k_if_no_skip : c->kadd;
    const long long adv = multiples_in(c->count, c->count + kadd, B.num_envs);
    c->ptr = (c->ptr + adv) % B.max_size;
    c->size = (c->size + adv < B.max_size) ? c->size + adv : B.max_size;
    c->count += kadd;
}

}  // namespace

namespace {

// ---- sample ---------------------------------------------------------------------------------------------
// chunk_sum[row][c] = sum of priority[row, c*1024 .. min((c+1)*1024, size)) in fp64 (blocks past `size` exit)
__global__ void __launch_bounds__(256) lap_chunk_sum_kernel(LapView B, double* __restrict__ chunk_sum, int nchunk_max) {
    __shared__ double wsum[8];
    const int row = blockIdx.y, c = blockIdx.x, tid = threadIdx.x;
    const long long size = B.ctr->size;
    const long long lo = (long long)c * LAP_CHUNK;
    if (lo >= size) return;
    const float* p = B.priority + (long long)row * B.max_size + lo;
    const long long i = (long long)tid * 4;
    double acc = 0.0;
    if ((B.max_size & 3) == 0 && lo + i + 4 <= size) {
        const float4 v = *reinterpret_cast<const float4*>(p + i);
        acc = ((double)v.x + (double)v.y) + ((double)v.z + (double)v.w);
    } else {
        for (int u = 0; u < 4; u++) if (lo + i + u < size) acc += (double)p[i + u];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((tid & 31) == 0) wsum[tid >> 5] = acc;
    __syncthreads();
    if (tid == 0) {
        double t = 0.0;
        for (int w = 0; w < 8; w++) t += wsum[w];
        chunk_sum[(long long)row * nchunk_max + c] = t;
    }
}

// Blocks (row, slice): every block rebuilds the exclusive prefix over its row's chunk sums (<= 1 K doubles, cheap) and
// serves one sample per warp, so a batch of 128 per row spreads over 16 blocks per row instead of one SM per row
// (r01g ncu: 250 us with one block per row). Per sample: searchsorted(cumsum, val)
// (first index whose running sum is >= val) by locating the chunk and scanning it 32 priorities at a time, then
// the warp gathers that transition into the output batch.
__global__ void __launch_bounds__(256) lap_sample_kernel(LapView B, const double* __restrict__ chunk_sum, int nchunk_max,
                                                         const float* __restrict__ uniform, float* __restrict__ o_state,
                                                         float* __restrict__ o_action, float* __restrict__ o_next,
                                                         float* __restrict__ o_reward, float* __restrict__ o_not_done,
                                                         int* __restrict__ ind_keep, int* __restrict__ ind_out) {
    extern __shared__ double prefix[];               // [nchunk + 1] exclusive prefix, prefix[nchunk] = total
    const int row = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, nwarp = blockDim.x >> 5;
    const long long size = B.ctr->size;
    const int nchunk = (int)((size + LAP_CHUNK - 1) / LAP_CHUNK);
    if (wid == 0) {                                   // warp scan over the chunk sums, 32 at a time
        double run = 0.0;
        for (int c0 = 0; c0 < nchunk; c0 += 32) {
            const int c = c0 + lane;
            double v = c < nchunk ? chunk_sum[(long long)row * nchunk_max + c] : 0.0, inc = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const double t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
            if (c < nchunk) prefix[c] = run + (inc - v);
            run += __shfl_sync(0xffffffffu, inc, 31);
        }
        if (lane == 0) prefix[nchunk] = run;
    }
    __syncthreads();
    const float total = (float)prefix[nchunk];        // csum[-1] (float32 in the reference)
    const float* prow = B.priority + (long long)row * B.max_size;
    for (int b = blockIdx.y * nwarp + wid; b < B.batch; b += gridDim.y * nwarp) {
        const int o = row * B.batch + b;
        long long idx = 0;
        if (size > 0) {
            const float valf = uniform[o] * total;    // torch.rand(...) * csum[-1], float32
            const double val = (double)valf;
            // chunk: first c with inclusive prefix >= val
            int lo = 0, hi = nchunk - 1;
            while (lo < hi) { const int mid = (lo + hi) >> 1; if (prefix[mid + 1] >= val) hi = mid; else lo = mid + 1; }
            long long found = -1;
            for (int c = lo; c < nchunk && found < 0; c++) {          // normally one chunk; rounding may push it to the next
                double run = prefix[c];
                const long long base = (long long)c * LAP_CHUNK;
                for (int i0 = 0; i0 < LAP_CHUNK && found < 0; i0 += 32) {
                    const long long i = base + i0 + lane;
                    const double v = i < size ? (double)prow[i] : 0.0;
                    double inc = v;
#pragma unroll
                    for (int s = 1; s < 32; s <<= 1) { const double t = __shfl_up_sync(0xffffffffu, inc, s); if (lane >= s) inc += t; }
                    const unsigned hit = __ballot_sync(0xffffffffu, i < size && run + inc >= val);
                    if (hit) found = base + i0 + (__ffs(hit) - 1);
                    run += __shfl_sync(0xffffffffu, inc, 31);
                    if (base + i0 + 32 >= size) break;
                }
            }
            idx = found < 0 ? size - 1 : found;       // val above the last running sum (rounding): last valid index
        }
        if (lane == 0) { ind_keep[o] = (int)idx; if (ind_out) ind_out[o] = (int)idx; }
        const long long t = (long long)row * B.max_size + idx;
        const int sd = B.sd, ad = B.ad;
        const float* s = B.state + t * sd;
        const float* ns = B.next_state + t * sd;
        if ((sd & 3) == 0 && ((reinterpret_cast<size_t>(o_state) | reinterpret_cast<size_t>(o_next)) & 15) == 0) {
            for (int i = lane; i < (sd >> 2); i += 32) {
                reinterpret_cast<float4*>(o_state + (size_t)o * sd)[i] = reinterpret_cast<const float4*>(s)[i];
                reinterpret_cast<float4*>(o_next + (size_t)o * sd)[i] = reinterpret_cast<const float4*>(ns)[i];
            }
        } else {
            for (int i = lane; i < sd; i += 32) { o_state[(size_t)o * sd + i] = s[i]; o_next[(size_t)o * sd + i] = ns[i]; }
        }
        for (int i = lane; i < ad; i += 32) o_action[(size_t)o * ad + i] = B.action[t * ad + i];
        if (lane == 0) { o_reward[o] = B.reward[t]; o_not_done[o] = B.not_done[t]; }
    }
}

// ---- priorities -----------------------------------------------------------------------------------------
__device__ __forceinline__ void atomic_max_nonneg(float* addr, float v) {
    if (v >= 0.0f) atomicMax(reinterpret_cast<int*>(addr), __float_as_int(v));     // non-negative floats order like ints
}
__global__ void lap_update_kernel(LapView B, const float* __restrict__ priority, const int* __restrict__ ind, int total) {
    const int o = blockIdx.x * blockDim.x + threadIdx.x;
    float p = 0.0f;
    if (o < total) {
        p = priority[o];
        // numpy fancy assignment (:115): when an index was sampled more than once in a row the LAST entry wins
        const int row = o / B.batch, me = ind[o];
        bool last = true;
        for (int q = o + 1; q < (row + 1) * B.batch; q++) last = last && ind[q] != me;
        if (last) B.priority[(long long)row * B.max_size + me] = p;
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) p = fmaxf(p, __shfl_xor_sync(0xffffffffu, p, s));
    if ((threadIdx.x & 31) == 0) atomic_max_nonneg(&B.ctr->max_priority, p);
}
__global__ void lap_set_max_kernel(LapView B, float v) { B.ctr->max_priority = v; }
__global__ void lap_max_kernel(LapView B, long long n) {
    float m = 0.0f;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        m = fmaxf(m, B.priority[i]);
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, s));
    if ((threadIdx.x & 31) == 0) atomic_max_nonneg(&B.ctr->max_priority, m);
}
__global__ void lap_reset_counters_kernel(LapView B) {
    B.ctr->ptr = 0; B.ctr->count = 0; B.ctr->size = 0; B.ctr->max_priority = 1.0f; B.ctr->kadd = 0;
}

}  // namespace

// ---- C-ABI ----------------------------------------------------------------------------------------------
struct ExoLap {
    LapView v;
    int device, n_sm;
    int nchunk_max;
    double* chunk_sum;        // [num_envs][nchunk_max]
    int* ind;                 // [num_envs * batch] indices of the last sample
    int* rank;                // [rank_cap] scratch of exo_lap_add
    int* winner;              // [winner_cap]
    long long rank_cap, winner_cap;
    bool sampled;
    int64_t launches;
    std::vector<void*> allocs;
};

static int lap_alloc(ExoLap* b, void** p, size_t bytes) {
    LCK(cudaMalloc(p, bytes ? bytes : 16));
    b->allocs.push_back(*p);
    LCK(cudaMemset(*p, 0, bytes ? bytes : 16));
    return 0;
}

extern "C" int exo_lap_create(int32_t state_dim, int32_t action_dim, int32_t num_envs, int64_t max_size, int32_t batch_size,
                              double max_action, int32_t normalize_actions, int32_t device, ExoLap** out) {
    if (!out || state_dim <= 0 || action_dim <= 0 || num_envs <= 0 || max_size <= 0 || batch_size <= 0 || max_action == 0.0)
        return exo_set_error(EXO_ERR_ARG, "exo_lap_create: bad arguments");
    int ndev = 0;
    LCK(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return exo_set_error(EXO_ERR_CUDA, "device %d not present: this library has no CPU path", device);
    cudaDeviceProp prop;
    LCK(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) return exo_set_error(EXO_ERR_CUDA, "device %d is sm_%d%d; libexo_env.so carries sm_100a code only", device, prop.major, prop.minor);
    LCK(cudaSetDevice(device));
    ExoLap* b = new ExoLap();
    memset(&b->v, 0, sizeof(b->v));
    b->device = device; b->n_sm = prop.multiProcessorCount; b->sampled = false; b->launches = 0;
    b->rank = nullptr; b->winner = nullptr; b->rank_cap = b->winner_cap = 0;
    LapView& v = b->v;
    v.sd = state_dim; v.ad = action_dim; v.num_envs = num_envs; v.batch = batch_size; v.max_size = max_size;
    v.norm = (float)(normalize_actions ? max_action : 1.0);
    const size_t slots = (size_t)num_envs * (size_t)max_size;
    b->nchunk_max = (int)((max_size + LAP_CHUNK - 1) / LAP_CHUNK);
    int rc = lap_alloc(b, (void**)&v.state, slots * state_dim * sizeof(float));
    if (!rc) rc = lap_alloc(b, (void**)&v.next_state, slots * state_dim * sizeof(float));
    if (!rc) rc = lap_alloc(b, (void**)&v.action, slots * action_dim * sizeof(float));
    if (!rc) rc = lap_alloc(b, (void**)&v.reward, slots * sizeof(float));
    if (!rc) rc = lap_alloc(b, (void**)&v.not_done, slots * sizeof(float));
    if (!rc) rc = lap_alloc(b, (void**)&v.priority, slots * sizeof(float));
    if (!rc) rc = lap_alloc(b, (void**)&v.ctr, sizeof(LapCounters));
    if (!rc) rc = lap_alloc(b, (void**)&b->chunk_sum, sizeof(double) * (size_t)num_envs * b->nchunk_max);
    if (!rc) rc = lap_alloc(b, (void**)&b->ind, sizeof(int) * (size_t)num_envs * batch_size);
    if (rc) { exo_lap_destroy(b); return rc; }
    lap_reset_counters_kernel<<<1, 1>>>(v);
    LCK(cudaGetLastError());
    LCK(cudaDeviceSynchronize());
    *out = b;
    return 0;
}

extern "C" int exo_lap_destroy(ExoLap* b) {
    if (!b) return 0;
    cudaSetDevice(b->device);
    for (void* p : b->allocs) cudaFree(p);
    if (b->rank) cudaFree(b->rank);
    if (b->winner) cudaFree(b->winner);
    delete b;
    return 0;
}

static int lap_scratch(ExoLap* b, int** p, long long* cap, long long need) {
    if (need <= *cap) return 0;
    if (*p) LCK(cudaFree(*p));
    *p = nullptr; *cap = 0;
    LCK(cudaMalloc((void**)p, sizeof(int) * (size_t)need));
    *cap = need;
    return 0;
}

extern "C" int exo_lap_add(ExoLap* b, const float* state_dev, const float* action_dev, const float* next_state_dev,
                           const float* reward_dev, const uint8_t* done_dev, const int32_t* tremor_num_dev,
                           const uint8_t* skip_dev, int32_t k, void* stream) {
    if (!b || !state_dev || !action_dev || !next_state_dev || !reward_dev || !done_dev || k < 0)
        return exo_set_error(EXO_ERR_ARG, "exo_lap_add: bad arguments");
    if (k == 0) return 0;
    LCK(cudaSetDevice(b->device));
    cudaStream_t s = (cudaStream_t)stream;
    const LapView& v = b->v;
    // windows (= distinct slots) one call can touch: the count can cross at most k / num_envs + 1 multiples of num_envs
    const long long nwin = (long long)k / v.num_envs + 2;
    if (nwin - 1 > v.max_size) return exo_set_error(EXO_ERR_ARG, "exo_lap_add: %d entries wrap the %lld-slot buffer within one call", k, v.max_size);
    if (int rc = lap_scratch(b, &b->winner, &b->winner_cap, nwin * v.num_envs)) return rc;
    int* rank = nullptr;
    if (skip_dev) {
        if (int rc = lap_scratch(b, &b->rank, &b->rank_cap, k)) return rc;
        rank = b->rank;
        lap_rank_kernel<<<1, 1024, 0, s>>>(skip_dev, k, rank, v.ctr);
        b->launches++;
    }
    LCK(cudaMemsetAsync(b->winner, 0xff, sizeof(int) * (size_t)(nwin * v.num_envs), s));     // -1
    lap_winner_kernel<<<(k + 255) / 256, 256, 0, s>>>(v, tremor_num_dev, rank, k, b->winner, (int)nwin);
    lap_add_kernel<<<(k + 7) / 8, 256, 0, s>>>(v, state_dev, action_dev, next_state_dev, reward_dev, done_dev, tremor_num_dev, rank,
                                               k, b->winner, (int)nwin);
    lap_advance_kernel<<<1, 1, 0, s>>>(v, skip_dev ? -1 : k);
    b->launches += 3;
    LCK(cudaGetLastError());
    return 0;
}

extern "C" int exo_lap_sample(ExoLap* b, const float* uniform_dev, float* state_dev, float* action_dev, float* next_state_dev,
                              float* reward_dev, float* not_done_dev, int32_t* ind_dev, void* stream) {
    if (!b || !uniform_dev || !state_dev || !action_dev || !next_state_dev || !reward_dev || !not_done_dev)
        return exo_set_error(EXO_ERR_ARG, "exo_lap_sample: null buffer");
    LCK(cudaSetDevice(b->device));
    cudaStream_t s = (cudaStream_t)stream;
    const LapView& v = b->v;
    lap_chunk_sum_kernel<<<dim3(b->nchunk_max, v.num_envs), 256, 0, s>>>(v, b->chunk_sum, b->nchunk_max);
    const size_t smem = sizeof(double) * (size_t)(b->nchunk_max + 1);
    if (smem > 200 * 1024) return exo_set_error(EXO_ERR_ARG, "exo_lap_sample: max_size too large for the chunk prefix (%d chunks)", b->nchunk_max);
    if (smem > 48 * 1024) LCK(cudaFuncSetAttribute(lap_sample_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    lap_sample_kernel<<<dim3(v.num_envs, (v.batch + 7) / 8), 256, smem, s>>>(v, b->chunk_sum, b->nchunk_max, uniform_dev, state_dev, action_dev, next_state_dev,
                                                    reward_dev, not_done_dev, b->ind, ind_dev);
    b->launches += 2;
    LCK(cudaGetLastError());
    b->sampled = true;
    return 0;
}

extern "C" int exo_lap_update_priority(ExoLap* b, const float* priority_dev, void* stream) {
    if (!b || !priority_dev) return exo_set_error(EXO_ERR_ARG, "exo_lap_update_priority: bad arguments");
    if (!b->sampled) return exo_set_error(EXO_ERR_STATE, "exo_lap_update_priority: no sample() since the last update (priority_indexes is empty)");
    LCK(cudaSetDevice(b->device));
    const int total = b->v.num_envs * b->v.batch;
    lap_update_kernel<<<(total + 255) / 256, 256, 0, (cudaStream_t)stream>>>(b->v, priority_dev, b->ind, total);
    b->launches++;
    LCK(cudaGetLastError());
    b->sampled = false;                                  // priority_indexes.clear() (:117)
    return 0;
}

extern "C" int exo_lap_reset_max_priority(ExoLap* b, void* stream) {
    if (!b) return exo_set_error(EXO_ERR_ARG, "exo_lap_reset_max_priority: null handle");
    LCK(cudaSetDevice(b->device));
    cudaStream_t s = (cudaStream_t)stream;
    lap_set_max_kernel<<<1, 1, 0, s>>>(b->v, 0.0f);
    const long long n = (long long)b->v.num_envs * b->v.max_size;
    long long blocks = (n + 1023) / 1024;
    if (blocks > (long long)b->n_sm * 16) blocks = (long long)b->n_sm * 16;
    lap_max_kernel<<<(int)blocks, 256, 0, s>>>(b->v, n);
    b->launches += 2;
    LCK(cudaGetLastError());
    return 0;
}

extern "C" int exo_lap_reset_buffer(ExoLap* b, void* stream) {
    if (!b) return exo_set_error(EXO_ERR_ARG, "exo_lap_reset_buffer: null handle");
    LCK(cudaSetDevice(b->device));
    cudaStream_t s = (cudaStream_t)stream;
    const LapView& v = b->v;
    const size_t slots = (size_t)v.num_envs * (size_t)v.max_size;
    LCK(cudaMemsetAsync(v.state, 0, slots * v.sd * sizeof(float), s));
    LCK(cudaMemsetAsync(v.next_state, 0, slots * v.sd * sizeof(float), s));
    LCK(cudaMemsetAsync(v.action, 0, slots * v.ad * sizeof(float), s));
    LCK(cudaMemsetAsync(v.reward, 0, slots * sizeof(float), s));
    LCK(cudaMemsetAsync(v.not_done, 0, slots * sizeof(float), s));
    LCK(cudaMemsetAsync(v.priority, 0, slots * sizeof(float), s));
    lap_reset_counters_kernel<<<1, 1, 0, s>>>(v);
    b->launches++;
    LCK(cudaGetLastError());
    b->sampled = false;
    return 0;
}

extern "C" int exo_lap_counters(ExoLap* b, int64_t* ptr, int64_t* count, int64_t* size, double* max_priority, void* stream) {
    if (!b) return exo_set_error(EXO_ERR_ARG, "exo_lap_counters: null handle");
    LCK(cudaSetDevice(b->device));
    LapCounters c;
    LCK(cudaMemcpyAsync(&c, b->v.ctr, sizeof(c), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    LCK(cudaStreamSynchronize((cudaStream_t)stream));
    if (ptr) *ptr = c.ptr;
    if (count) *count = c.count;
    if (size) *size = c.size;
    if (max_priority) *max_priority = (double)c.max_priority;
    return 0;
}

extern "C" int exo_lap_storage(ExoLap* b, int32_t field, void** dev_ptr, int64_t* elems_per_slot) {
    if (!b || !dev_ptr || !elems_per_slot) return exo_set_error(EXO_ERR_ARG, "exo_lap_storage: bad arguments");
    const LapView& v = b->v;
    switch (field) {
        case 0: *dev_ptr = v.state; *elems_per_slot = v.sd; break;
        case 1: *dev_ptr = v.action; *elems_per_slot = v.ad; break;
        case 2: *dev_ptr = v.next_state; *elems_per_slot = v.sd; break;
        case 3: *dev_ptr = v.reward; *elems_per_slot = 1; break;
        case 4: *dev_ptr = v.not_done; *elems_per_slot = 1; break;
        case 5: *dev_ptr = v.priority; *elems_per_slot = 1; break;
        default: return exo_set_error(EXO_ERR_ARG, "exo_lap_storage: unknown field %d", field);
    }
    return 0;
}

extern "C" int64_t exo_lap_launch_count(const ExoLap* b) { return b ? b->launches : 0; }
