// exo_step.cuh — the fused control-step kernel: one thread per environment, one launch per env.step().
//
// Replaces, for N environments at once, Exoskeleton_env.py:368-471 (step) and what it calls:
//   transform_action :256-266 | get_actuator_positions sim:119-158 | get_force_components sim:207-298
//   get_actuator_pos_vect sim:300-336 | return_reference_dummy_pos sim:338-351 | get_torques :177-187
//   tremor lookup :403-406 | solve_diff_eq x2 :409-414 (as q = G*T, see exo_reset.cuh) | angle composition :421-424
//   set_joint_position + stepSimulation :427-433 (restated Bullet position motors, oracle/minibullet.py)
//   get_reward :341-366 | update_state_vector :487-570 | done :457
//   + the caller-side amplitude metric of Exoskeleton_agent_train.py:158-170 (forward_kinematics x3).
//
// Roofline: HBM-bound (1045 B of compulsory traffic per env-step for ~1.5 kflop). The kernel is a
// single wave (65 536 envs = 13.8 warps per SM), so what matters is bytes in flight per instruction and
// instruction count: all resident state moves as 128-bit vector loads/stores of the vector-major SoA
// (exo_types.cuh), the row-major boundary arrays actions[n,7] / obs[n,80] are transposed per WARP through
// shared memory (no block barrier: STS.128 into a 21-chunk pitch, LDS.128 + STG.128 out), and the fp32 build
// uses MUFU-based reciprocal / rsqrt / exp where the stated tolerances allow.
#pragma once
#include <type_traits>
#include "exo_types.cuh"
#include "exo_rk45.cuh"

namespace exo {

// sin and cos of a joint angle, fp32. Joint angles live inside the URDF limits (|q| < 2.9 rad), so the argument reduction is
// three Cody-Waite steps by pi/2 (exact for |k| < 2^7) and the kernels are the degree-7 / degree-8 minimax polynomials on
// [-pi/4, pi/4] (< 1 ulp each): ~28 instructions for the pair against ~78 for sincosf, whose generality (Payne-Hanek path,
// local-memory table) the 5 evaluations per env-step paid for - 20 % of the step kernel's instructions in the r02e profile.
// There is deliberately NO library fallback for large arguments: with one in the source the compiler predicated sincosf's
// fast path into every call (67 instructions per call in the r02v profile instead of 28). An env whose joint angle has left
// +-100 rad has diverged (non-SPD inertia draw, SURVEY hard part 5); its angle is reduced with fp32 accuracy like any other.
__device__ __forceinline__ void sincos_r(float x, float* s, float* c) {
    const float k = rintf(x * 0.63661977236758134308f);
    float r = fmaf(-k, 1.5707962512969970703125f, x);               // pi/2 split in three: 24 + 24 + 24 bits
    r = fmaf(-k, 7.5497894158615963533520698547e-8f, r);
    r = fmaf(-k, 5.3903029534742383e-15f, r);
    const float r2 = r * r;
    float ps = fmaf(r2, -1.9515295891e-4f, 8.3321608736e-3f);
    ps = fmaf(ps, r2, -1.6666654611e-1f);
    const float sn = fmaf(ps * r2, r, r);
    float pc = fmaf(r2, 2.443315711809948e-5f, -1.388731625493765e-3f);
    pc = fmaf(pc, r2, 4.166664568298827e-2f);
    const float cs = fmaf(pc * r2, r2, fmaf(-0.5f, r2, 1.0f));
    const int q = __float2int_rn(k);
    const float a = (q & 1) ? cs : sn, b = (q & 1) ? sn : cs;
    *s = (q & 2) ? -a : a;
    *c = ((q + 1) & 2) ? -b : b;
}
__device__ __forceinline__ void sincos_r(double x, double* s, double* c) { sincos(x, s, c); }
__device__ __forceinline__ float exp_r(float x) { return __expf(x); }
__device__ __forceinline__ double exp_r(double x) { return exp(x); }
__device__ __forceinline__ float sqrt_r(float x) { return sqrtf(x); }
__device__ __forceinline__ double sqrt_r(double x) { return sqrt(x); }
__device__ __forceinline__ float abs_r(float x) { return fabsf(x); }
__device__ __forceinline__ double abs_r(double x) { return fabs(x); }
__device__ __forceinline__ float div_r(float a, float b) { return __fdividef(a, b); }
__device__ __forceinline__ double div_r(double a, double b) { return a / b; }
// d / sqrt(r2), the "direction cosine" cos(atan2(., .)) of sim:222-230
__device__ __forceinline__ float dircos_r(float d, float r2) { return d * rsqrtf(r2); }
__device__ __forceinline__ double dircos_r(double d, double r2) { return d / sqrt(r2); }

__device__ __forceinline__ void ld4(const V4<float>* p, float* o) {
    const float4 t = *reinterpret_cast<const float4*>(p);
    o[0] = t.x; o[1] = t.y; o[2] = t.z; o[3] = t.w;
}
__device__ __forceinline__ void ld4(const V4<double>* p, double* o) {
    const double2 a = reinterpret_cast<const double2*>(p)[0], b = reinterpret_cast<const double2*>(p)[1];
    o[0] = a.x; o[1] = a.y; o[2] = b.x; o[3] = b.y;
}
// read-only data (per-episode parameters): non-coherent path
__device__ __forceinline__ void ld4_ro(const V4<float>* p, float* o) {
    const float4 t = __ldg(reinterpret_cast<const float4*>(p));
    o[0] = t.x; o[1] = t.y; o[2] = t.z; o[3] = t.w;
}
__device__ __forceinline__ void ld4_ro(const V4<double>* p, double* o) {
    const double2 a = __ldg(reinterpret_cast<const double2*>(p)), b = __ldg(reinterpret_cast<const double2*>(p) + 1);
    o[0] = a.x; o[1] = a.y; o[2] = b.x; o[3] = b.y;
}
// Ordered variants: `asm volatile` keeps these in program order, so the kernel decides which loads are in flight
// before the first dependent instruction (the compiler's own schedule put count-dependent loads in front of the
// independent ones and spilled freshly loaded registers, serialising the memory latency).
__device__ __forceinline__ void ld4_ord(const V4<float>* p, float* o) {
    asm volatile("ld.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(o[0]), "=f"(o[1]), "=f"(o[2]), "=f"(o[3]) : "l"(p));
}
__device__ __forceinline__ void ld4_ord(const V4<double>* p, double* o) {
    asm volatile("ld.global.v2.f64 {%0,%1}, [%2];" : "=d"(o[0]), "=d"(o[1]) : "l"(p));
    asm volatile("ld.global.v2.f64 {%0,%1}, [%2+16];" : "=d"(o[2]), "=d"(o[3]) : "l"(p));
}
__device__ __forceinline__ void ld4_ro_ord(const V4<float>* p, float* o) {
    asm volatile("ld.global.nc.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(o[0]), "=f"(o[1]), "=f"(o[2]), "=f"(o[3]) : "l"(p));
}
__device__ __forceinline__ void ld4_ro_ord(const V4<double>* p, double* o) {
    asm volatile("ld.global.nc.v2.f64 {%0,%1}, [%2];" : "=d"(o[0]), "=d"(o[1]) : "l"(p));
    asm volatile("ld.global.nc.v2.f64 {%0,%1}, [%2+16];" : "=d"(o[2]), "=d"(o[3]) : "l"(p));
}
__device__ __forceinline__ int ld_int_ord(const int* p) {
    int v;
    asm volatile("ld.global.s32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ float ld_ro_ord(const float* p) {
    float v;
    asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ double ld_ro_ord(const double* p) {
    double v;
    asm volatile("ld.global.nc.f64 %0, [%1];" : "=d"(v) : "l"(p));
    return v;
}
// 16-byte asynchronous global -> shared copy (LDGSTS): no register, no scoreboard until cp_async_wait
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"((unsigned)__cvta_generic_to_shared(smem)), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

__device__ __forceinline__ void st4(V4<float>* p, const float* o) {
    *reinterpret_cast<float4*>(p) = make_float4(o[0], o[1], o[2], o[3]);
}
__device__ __forceinline__ void st4(V4<double>* p, const double* o) {
    reinterpret_cast<double2*>(p)[0] = make_double2(o[0], o[1]);
    reinterpret_cast<double2*>(p)[1] = make_double2(o[2], o[3]);
}

// out = A * B * Rz(angle); A,B row-major 3x3
template <typename R>
__device__ __forceinline__ void mul_origin_rz(const R* A, const R* B, R s, R c, R* out) {
    R M[9];
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int j = 0; j < 3; j++) M[i * 3 + j] = A[i * 3] * B[j] + A[i * 3 + 1] * B[3 + j] + A[i * 3 + 2] * B[6 + j];
#pragma unroll
    for (int i = 0; i < 3; i++) {
        out[i * 3 + 0] = c * M[i * 3] + s * M[i * 3 + 1];
        out[i * 3 + 1] = c * M[i * 3 + 1] - s * M[i * 3];
        out[i * 3 + 2] = M[i * 3 + 2];
    }
}
template <typename R>
__device__ __forceinline__ void affine(const R* Rm, const R* p, const R* v, R* o) {
#pragma unroll
    for (int i = 0; i < 3; i++) o[i] = p[i] + (Rm[i * 3] * v[0] + Rm[i * 3 + 1] * v[1] + Rm[i * 3 + 2] * v[2]);
}

// World frames of the humerus (link 2) and forearm (link 4) and the elbow reference point for joint vector qs[5].
template <typename R>
struct ChainPose { R R2[9], p2[3], R4[9], p4[3], elref[3]; };

template <typename R>
__device__ __forceinline__ void chain_fk(const ChainConst<R>& C, const R* qs, ChainPose<R>& P) {
    R s, c, Ra[9], Rb[9], pa[3], pb[3];
    const R eye[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    sincos_r(qs[0], &s, &c);
    mul_origin_rz(eye, C.Rorg[0], s, c, Ra);                       // link 0
    pa[0] = C.p0[0]; pa[1] = C.p0[1]; pa[2] = C.p0[2];
    sincos_r(qs[1], &s, &c);
    affine(Ra, pa, C.porg[1], pb);
    mul_origin_rz(Ra, C.Rorg[1], s, c, Rb);                        // link 1
    sincos_r(qs[2], &s, &c);
    affine(Rb, pb, C.porg[2], P.p2);
    mul_origin_rz(Rb, C.Rorg[2], s, c, P.R2);                      // link 2 (humerus)
    sincos_r(qs[3], &s, &c);
    affine(P.R2, P.p2, C.porg[3], pa);
    mul_origin_rz(P.R2, C.Rorg[3], s, c, Ra);                      // link 3 (auxlink3)
    affine(Ra, pa, C.off_elbow, P.elref);
    sincos_r(qs[4], &s, &c);
    affine(Ra, pa, C.porg[4], P.p4);
    mul_origin_rz(Ra, C.Rorg[4], s, c, P.R4);                      // link 4 (forearm)
}

// End effector of the 7-joint Denavit-Hartenberg arm of calculate_arm_end_effector_points.py:18-50
// (alpha = +-pi/2 taken exactly, a = 0, d = (0,0,L1,0,L2,0,0)).
template <typename R>
__device__ __forceinline__ void dh_end_effector(const R* th, R L1, R L2, R* p) {
    R c0[3] = {1, 0, 0}, c1[3] = {0, 1, 0}, c2[3] = {0, 0, 1};
    p[0] = p[1] = p[2] = 0;
#pragma unroll
    for (int k = 0; k < 7; k++) {
        R s, c;
        sincos_r(th[k], &s, &c);
        const R sa = (k == 2) ? R(-1) : R(1);
        const R d = (k == 2) ? L1 : (k == 4 ? L2 : R(0));
#pragma unroll
        for (int i = 0; i < 3; i++) {
            R n0 = c * c0[i] + s * c1[i];
            R n1 = sa * c2[i];
            R n2 = sa * (s * c0[i] - c * c1[i]);
            p[i] += d * c2[i];
            c0[i] = n0; c1[i] = n1; c2[i] = n2;
        }
    }
}

// defined in exo_reset.cuh: start the prefetched episode of env e (slot flip, counts = 2, position reads, first observation)
template <typename R, int NAX>
__device__ __forceinline__ void begin_episode(const EnvArrays<R>& S, const ChainConst<R>& C, int e, int new_slot,
                                           const R* qs, const R* el_old, const R* qpl, float4* obA, float4* obB,
                                           const float4* st_ep, const float4* st_tt);

// FULL: ExoConfig.motor_model = EXO_MOTOR_FULL - the position targets go to exo_motor_kernel (exo_motor.cuh) instead of the
// closed-form tracker, and the actuator end points follow their prismatic dummies (which the general step lets drift).
template <typename R, bool INFO, int NAX, bool FULL = false>
__global__ void __launch_bounds__(EXO_STEP_THREADS, (sizeof(R) == 4 && !INFO && !FULL) ? EXO_STEP_MIN_BLOCKS : 4)
exo_step_kernel(EnvArrays<R> S, ChainConst<R> C, StepConst<R> K, const float* __restrict__ actions,
                float* __restrict__ obs, R* __restrict__ reward, uint8_t* __restrict__ done, InfoPtrs<R> info) {
    constexpr int WARPS = EXO_STEP_THREADS / 32;
    constexpr int NAXV = (NAX + 3) / 4;                               // vectors per tremor-table row
    // Per-warp staging (no block barrier anywhere): obs chunks 0..10 | per-episode parameter vectors, later reused for
    // obs chunks 11..19 | the warp's 32 action rows. Pitches of 11 and 9 sixteen-byte chunks are odd: the STS.128 / LDS.128 of
    // these two tiles are conflict-free (ncu still counts ~28 conflicting wavefronts per warp in the kernel, 10.9 % of its
    // shared-memory wavefronts, in the LDGSTS copies).
    constexpr int EPB_BYTES = (NV_EP * 32 * (int)sizeof(V4<R>) > 32 * 9 * 16) ? NV_EP * 32 * (int)sizeof(V4<R>) : 32 * 9 * 16;
    __shared__ __align__(16) float4 s_obsA[WARPS][32 * 11];
    __shared__ __align__(16) unsigned char s_epB[WARPS][EPB_BYTES];
    __shared__ __align__(16) float s_act[WARPS][32 * 7];
    // fp32: the first three tremor rows of the NEXT episode of a lane whose episode ends in this step (its 11 parameter
    // vectors go to the lane's row of s_obsA, which that lane does not need: its observation is the next episode's first)
    constexpr bool STAGE_NEXT = sizeof(R) == 4;
    __shared__ __align__(16) float4 s_nxt[STAGE_NEXT ? WARPS : 1][STAGE_NEXT ? 32 * 3 : 1];

    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int wbase = K.env0 + blockIdx.x * EXO_STEP_THREADS + wid * 32;     // first env of this warp
    const int e = wbase + lane;
    const int n = K.env1, st = S.stride;                            // this launch covers envs [K.env0, K.env1)
    if (blockIdx.x == 0 && tid == 0) S.fin_count[(K.list + 1) % EXO_RING] = 0;   // the next call's list (its last
                                                                     // consumer finished: the host waited for that generation pass)
    if (wbase >= n) return;                                          // whole warp past the end
    long long* tl = nullptr;
    if (K.timeline != nullptr) {
        tl = K.timeline + (size_t)((wbase - K.env0) >> 5) * 8;
        if (lane == 0) {
            unsigned long long gt; unsigned smid;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt));
            asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
            tl[0] = (long long)gt; tl[1] = smid; tl[2] = clock64();
        }
    }
#define EXO_MARK(k) do { if (tl != nullptr && lane == 0) tl[k] = clock64(); } while (0)
    if (K.stagger_cycles > 0) {
        // The launch is one wave whose warps would all issue their ~30 vector loads at once; the memory system then serves
        // them roughly first-come first-served ACROSS warps, so every warp's last vector arrives when the whole 34 MB burst
        // has been served (~5 us at the HBM rate) and no warp computes before that. Starting the blocks of later dispatch
        // layers a little later gives the early layers their data first: they compute while the rest of the burst streams.
        const int layer = blockIdx.x / K.n_sm;
        if (layer > 0) {
            const long long t0 = clock64(), wait = (long long)layer * K.stagger_cycles;
            while (clock64() - t0 < wait) __nanosleep(64);
        }
    }

    // ---- every global access of the step is issued up front, in this order and unconditionally (a lane past the end
    //      or an idle env re-reads a valid record and discards it): counts/meta first (the second-level addresses
    //      depend on them: counts carries the step index AND which of the two parameter slots holds the running
    //      episode), the action rows as asynchronous global->shared copies, the carried state as 128-bit register
    //      loads, then - one memory latency later - the per-episode parameters (async copies), G, the three tremor rows
    //      and the motion row.
    const bool valid = e < n;
    const int ec = valid ? e : n - 1;
    const int c_word = ld_int_ord(S.counts + ec);
    const int meta = ld_int_ord(S.meta + ec);
    V4<R>* sp = reinterpret_cast<V4<R>*>(&s_epB[wid][0]);            // [NV_EP][32]
    const int rows_here = min(32, n - wbase);
    const float* asrc = actions + (size_t)wbase * 7;
    const bool act_async = rows_here == 32 && (reinterpret_cast<size_t>(asrc) & 15) == 0;   // warp-uniform
    float araw[7];
    if (act_async) {                                                  // 896 contiguous bytes = 56 chunks
        cp_async16(&s_act[wid][lane * 4], asrc + lane * 4);
        if (lane < 24) cp_async16(&s_act[wid][(lane + 32) * 4], asrc + (lane + 32) * 4);
    } else {
#pragma unroll
        for (int k = 0; k < 7; k++) araw[k] = (lane + 32 * k < rows_here * 7) ? ld_ro_ord(asrc + lane + 32 * k) : 0.0f;
    }
    cp_async_commit();
    R s[4 * NV_STATE];
#pragma unroll
    for (int v = 0; v < NV_STATE; v++) ld4_ord(S.state + (size_t)v * st + ec, s + 4 * v);
    const R ret0 = S.ret[ec];
    // second level (clamped so the addresses are always inside the tables; a live env has 2 <= counts <= L - 2, so
    // the clamp never changes what a live env reads)
    const int c_raw = c_word & EXO_COUNT_MASK, slot = (c_word >> EXO_SLOT_SHIFT) & 1;
    const int mid = (meta >> 8) & 0xff, L = meta >> 16;
    const int c = min(max(c_raw, 1), S.table_len - 2);
    {
        const V4<R>* eps = S.ep + (size_t)slot * S.ep_slot;
#pragma unroll
        for (int v = 0; v < NV_EP; v++) {
            const char* src = reinterpret_cast<const char*>(eps + (size_t)v * st + ec);
            char* dst = reinterpret_cast<char*>(sp + v * 32 + lane);
#pragma unroll
            for (int h = 0; h < (int)sizeof(V4<R>) / 16; h++) cp_async16(dst + 16 * h, src + 16 * h);
        }
        cp_async_commit();
    }
    constexpr int NGV = (INFO || NAX > 4) ? NV_G : 4;
    R g[4 * NGV];
    {
        const V4<R>* Gs = S.G + (size_t)slot * S.G_slot;
#pragma unroll
        for (int v = 0; v < NGV; v++) ld4_ro_ord(Gs + (size_t)v * st + ec, g + 4 * v);
    }
    R tm[4 * NAXV], tc[4 * NAXV], tp[4 * NAXV];                       // tremor rows c-1, c, c+1
    {
        const V4<R>* tts = S.tt + (size_t)slot * S.tt_slot;
#pragma unroll
        for (int h = 0; h < NAXV; h++) {
            ld4_ro_ord(tts + ((size_t)(c - 1) * st + ec) * NAXV + h, tm + 4 * h);
            ld4_ro_ord(tts + ((size_t)c * st + ec) * NAXV + h, tc + 4 * h);
            ld4_ro_ord(tts + ((size_t)(c + 1) * st + ec) * NAXV + h, tp + 4 * h);
        }
    }
    // The envs whose episode ends in this step are known NOW (counts + 1 reaches L - 1): reserve their places on the work list
    // here, so the atomic's round trip to a cold L2 line overlaps the whole step instead of sitting at the end of the warp
    const bool will_fin = valid && c_raw < L - 1 && c_raw + 2 >= L;
    const unsigned wf_mask = __ballot_sync(0xffffffffu, will_fin);
    int fin_base = 0;
    if (K.auto_reset && wf_mask != 0u && lane == 0) fin_base = atomicAdd(S.fin_count + K.list, __popc(wf_mask));
    const bool stage_next = STAGE_NEXT && K.fused_begin && will_fin;
    if (K.fused_begin && will_fin) {
        // this env's episode ends in this step: fetch the first rows of its NEXT episode's parameters (spare slot) now, so the
        // episode-start path at the end of the kernel does not wait for memory: asynchronous copies into shared memory
        // (fp32), an L2 prefetch otherwise
        const V4<R>* epn = S.ep + (size_t)(slot ^ 1) * S.ep_slot;
        const V4<R>* ttn = S.tt + (size_t)(slot ^ 1) * S.tt_slot;
        if (STAGE_NEXT) {
#pragma unroll
            for (int v = 0; v < NV_EP; v++) cp_async16(&s_obsA[wid][lane * 11 + v], epn + (size_t)v * st + ec);
#pragma unroll
            for (int k = 0; k < 3; k++) cp_async16(&s_nxt[STAGE_NEXT ? wid : 0][STAGE_NEXT ? lane * 3 + k : 0], ttn + ((size_t)k * st + ec) * NAXV);
        } else {
#pragma unroll
            for (int v = 0; v < NV_EP; v++) asm volatile("prefetch.global.L2 [%0];" :: "l"(epn + (size_t)v * st + ec));
#pragma unroll
            for (int k = 0; k < 3; k++) asm volatile("prefetch.global.L2 [%0];" :: "l"(ttn + ((size_t)k * st + ec) * NAXV));
        }
    }
    cp_async_commit();
    const R* mo = S.motion + (size_t)mid * 5 * S.max_len;
    R mrow[5], mnext[5];
#pragma unroll
    for (int r = 0; r < 5; r++) {
        mrow[r] = ld_ro_ord(mo + r * S.max_len + c);
        mnext[r] = INFO ? ld_ro_ord(mo + r * S.max_len + c + 1) : R(0);
    }
    EXO_MARK(3);                                                     // every load issued, counts arrived
    if (!act_async) {
#pragma unroll
        for (int k = 0; k < 7; k++) s_act[wid][lane + 32 * k] = araw[k];
    }
    cp_async_wait_all();
    __syncwarp();
    EXO_MARK(4);                                                     // staged parameters + actions landed

    const bool live = valid && c_raw < L - 1;   // an env that already reported done idles (Exoskeleton_agent_train.py:139)
    bool fin = false;
    R b_qs[5], b_el[3];                         // auto-reset: joint vector behind the link-state cache / elbow reference after this step
    int v_cross = 0, v_red = 0, v_len = 0;
    R v_ret = 0, v_amp = 0;
    int v_cge = 0, v_clt = 0, v_ntr = 0, v_nam = 0;      // training-script suppression metrics (INFO mode only)
    R v_str = 0, v_sam = 0;
    if (valid && !live) {
        done[e] = 1; reward[e] = R(0);
        if (FULL) { const R z[4] = {0, 0, 0, 0}; st4(S.tgt + (size_t)st + e, z); }     // live flag off: the motor kernel skips this env
    }
    const unsigned full = 0xffffffffu;
    const unsigned live_mask = __ballot_sync(full, live);
    if (live) {
        float ob[80];
        // ---- transform_action (:256-266)
        R pf[4];
        ld4(sp + (W_FMAX >> 2) * 32 + lane, pf);                      // staged parameter vector 10: shifts 40,41 + force maxima
        const R fm_sh = pf[W_FMAX & 3], fm_el = pf[(W_FMAX + 1) & 3];
        R F[7];
#pragma unroll
        for (int i = 0; i < 7; i++) F[i] = ((R(s_act[wid][lane * 7 + i]) + R(1)) / R(2)) * (i < 2 ? fm_el : fm_sh);
        // ---- observation chunks that only need loaded data go to the staging tile NOW (frees ~44 registers for the
        //      arithmetic below): forces[k-2], forces[k-1] (:488-501), tremor rows k-2..k (:502-513; the obs is built
        //      after counts += 1, so rows c-1, c, c+1), prev_position_vectors (:385)
        float4* so = &s_obsA[wid][lane * 11];
        R fp1[7];
#pragma unroll
        for (int i = 0; i < 7; i++) {
            fp1[i] = s[W_FPREV + i];
            const R f0i = i < 2 ? K.inv_fmax_el0 : K.inv_fmax_sh0, f0 = i < 2 ? K.fmax_el0 : K.fmax_sh0;
            if (sizeof(R) == 8) {
                ob[i] = (c >= 3) ? (float)(fp1[i] / f0) : 0.0f;       // forces[k-2]; the history is zeroed at reset (:195)
                ob[7 + i] = (float)(F[i] / f0);                       // forces[k-1]
            } else {
                ob[i] = (c >= 3) ? (float)(fp1[i] * f0i) : 0.0f;
                ob[7 + i] = (float)(F[i] * f0i);
            }
        }
        {
            const R inrm[4] = {R(0.1), R(0.1), R(0.1), R(0.2)};
            const R nrm[4] = {R(10), R(10), R(10), R(5)};
#pragma unroll
            for (int j = 0; j < 4; j++) {
                if (sizeof(R) == 8) {
                    ob[14 + j] = (float)(tm[j] / nrm[j]); ob[18 + j] = (float)(tc[j] / nrm[j]); ob[22 + j] = (float)(tp[j] / nrm[j]);
                } else {
                    ob[14 + j] = (float)(tm[j] * inrm[j]); ob[18 + j] = (float)(tc[j] * inrm[j]); ob[22 + j] = (float)(tp[j] * inrm[j]);
                }
            }
        }
#pragma unroll
        for (int i = 0; i < 21; i++) ob[26 + i] = (float)s[W_POSVEC + i];
        if (!stage_next) {                       // (a lane whose episode ends keeps the staged next-episode parameters in its row)
#pragma unroll
            for (int ch = 0; ch < 11; ch++) so[ch] = make_float4(ob[4 * ch], ob[4 * ch + 1], ob[4 * ch + 2], ob[4 * ch + 3]);
        }
        // ---- link-state cache -> world points (getLinkState reads, sim:129-142,194-195)
        R qb[5], qs[5];
#pragma unroll
        for (int j = 0; j < 5; j++) { qb[j] = s[W_Q + j]; qs[j] = K.lag ? s[W_QLAG + j] : qb[j]; }
        ChainPose<R> P;
        chain_fk(C, qs, P);
        R el_old[3];
#pragma unroll
        for (int k = 0; k < 3; k++) el_old[k] = s[W_ELREF + k];
        R qpl[12];                                                    // FULL: prismatic positions behind the link-state cache
        if (FULL) {
            const int w0 = K.lag ? W_QPLAG : W_QP;                    // words w0 .. w0+8 of the dyn record
#pragma unroll
            for (int d = 0; d < EXO_NPRIS; d++) qpl[d] = *exo_word(S.dyn, w0 + d, e, st);
        }
        // new pos_vec -> obs[47:68]
        // ---- per actuator: endpoints, force direction, pos_vec, torque (sim:207-336, env:177-187)
        R at0 = 0, at1 = 0, at2 = 0, ty0 = 0, ty1 = 0;
        // one instantiation per actuator (compile-time index: every array subscript below is static - the loop over a
        // run-time order[] table left the shift selects as local-memory accesses), called in the summation order of
        // :395-397 (3,4,5,7,6)
        auto actuator = [&](auto IC) {
            constexpr int i = decltype(IC)::value;
            constexpr int fix_slot = i - 2;                  // fixed slot of act_i1 for the shoulder actuators
            R a1[3], a2[3];
            if (i < 2) {
                if (FULL) {
                    R oh[3], oa[3];
#pragma unroll
                    for (int k = 0; k < 3; k++) { oh[k] = C.off_hum[i][k] + C.pax_hum[i][k] * qpl[i]; oa[k] = C.off_alk[i][k] + C.pax_alk[i][k] * qpl[7 + i]; }
                    affine(P.R2, P.p2, oh, a1);
                    affine(P.R4, P.p4, oa, a2);
                } else {
                    affine(P.R2, P.p2, C.off_hum[i], a1);               // act11 / act21 ride on the humerus
                    affine(P.R4, P.p4, C.off_alk[i], a2);               // act12 / act22 ride on the forearm
                }
            } else {
                a1[0] = C.fixed[fix_slot < 0 ? 0 : fix_slot][0]; a1[1] = C.fixed[fix_slot < 0 ? 0 : fix_slot][1]; a1[2] = C.fixed[fix_slot < 0 ? 0 : fix_slot][2];
                if (FULL) {
                    R oh[3];
#pragma unroll
                    for (int k = 0; k < 3; k++) oh[k] = C.off_hum[i][k] + C.pax_hum[i][k] * qpl[i];
                    affine(P.R2, P.p2, oh, a2);
                } else affine(P.R2, P.p2, C.off_hum[i], a2);
            }
            {   // dummy_shift_coordinates of this actuator's two ends: words 6i..6i+5 of the staged record = two vectors
                constexpr int W0 = W_SHIFT;
                constexpr int w = W0 + 6 * i, va = w >> 2;
                R pa[4], pb[4];
                ld4(sp + va * 32 + lane, pa);
                ld4(sp + (va + 1) * 32 + lane, pb);
#pragma unroll
                for (int k = 0; k < 3; k++) {
                    const int w1 = w + k, w2 = w + 3 + k;
                    a1[k] += (w1 >> 2) == va ? pa[w1 & 3] : pb[w1 & 3];
                    a2[k] += (w2 >> 2) == va ? pa[w2 & 3] : pb[w2 & 3];
                }
            }
            R dx, dy, dz;
            if (sizeof(R) == 8) {   // the reference's +5 shift (sim:212) only matters at fp64 resolution
                dx = (a2[0] + R(5)) - (a1[0] + R(5)); dy = (a2[1] + R(5)) - (a1[1] + R(5)); dz = (a2[2] + R(5)) - (a1[2] + R(5));
            } else { dx = a2[0] - a1[0]; dy = a2[1] - a1[1]; dz = a2[2] - a1[2]; }
            // cos(atan2(dy,dx)), cos(atan2(dx,dy)), cos(atan2(dz,dx))  (not a unit vector; reproduced as is)
            const R rxy2 = dx * dx + dy * dy, rxz2 = dx * dx + dz * dz;
            const R fx = F[i] * (rxy2 > R(0) ? dircos_r(dx, rxy2) : R(1));
            const R fy = F[i] * (rxy2 > R(0) ? dircos_r(dy, rxy2) : R(1));
            const R fz = F[i] * (rxz2 > R(0) ? dircos_r(dx, rxz2) : R(1));
            const R* stale = (i < 2) ? el_old : C.sh_ref;               // refs cached by the previous read (sim:300-336)
            const R* fresh = (i < 2) ? P.elref : C.sh_ref;
            R r[3];
#pragma unroll
            for (int k = 0; k < 3; k++) {
                const R pv = a2[k] - stale[k];
                r[k] = fresh[k] - a2[k];
                s[W_POSVEC + i * 3 + k] = pv;
                ob[47 + i * 3 + k] = (float)pv;
            }
            const R tx = fy * r[2] - fz * r[1], ty = fz * r[0] - fx * r[2], tz = fx * r[1] - fy * r[0];   // cross(F, r) :179
            if (i == 0) ty0 = ty;
            else if (i == 1) ty1 = ty;
            else { at0 += ty; at1 += tx; at2 += tz; }
        };
        actuator(std::integral_constant<int, 0>{}); actuator(std::integral_constant<int, 1>{});
        actuator(std::integral_constant<int, 2>{}); actuator(std::integral_constant<int, 3>{});
        actuator(std::integral_constant<int, 4>{}); actuator(std::integral_constant<int, 6>{});
        actuator(std::integral_constant<int, 5>{});
        const R at[4] = {at0, at1, at2, abs_r(ty0) - abs_r(ty1)};      // :394-400
        EXO_MARK(5);                                                   // kinematics + actuators done
        // ---- tremor lookup (:403-406) and the rows the observation needs (:502-513)
        R ttc[7], tv[7];
#pragma unroll
        for (int j = 0; j < 7; j++) {
            ttc[j] = (j < NAX) ? tc[j] : R(0);
            tv[j] = ttc[j] + (j < 4 ? at[j] : R(0));
        }
        // ---- joint model: q(dt) = G * T from rest (:409-418); rows 0..3 drive the motors
        const R r2d = R(57.29577951308232087679815481410517), d2r = R(0.017453292519943295769236907684886);
        R qa[7], qt[7];
        constexpr int ROWS = INFO ? 7 : 4;
        bool ode_done = false;
        if constexpr (sizeof(R) == 8) {
            if (K.ode_rk45) {
                // verification mode (fp64 handles): the reference's own integrator, scipy's RK45 restated (exo_rk45.cuh), on
                // this episode's I, D, S - twice per step as in Exoskeleton_env.py:409-414
                const double* ids = S.IDS + ((size_t)slot * S.n + e) * EXO_IDS_PITCH;
                double Tv[7], q7[7];
#pragma unroll
                for (int j = 0; j < 7; j++) Tv[j] = tv[j];
                rk45_solve(ids, Tv, 1.0 / 40.0, q7);
#pragma unroll
                for (int i = 0; i < ROWS; i++) qa[i] = q7[i] * r2d;
                if (INFO) {
#pragma unroll
                    for (int j = 0; j < 7; j++) Tv[j] = ttc[j];
                    rk45_solve(ids, Tv, 1.0 / 40.0, q7);
#pragma unroll
                    for (int i = 0; i < 7; i++) qt[i] = q7[i] * r2d;
                }
                ode_done = true;
            }
        }
        if (!ode_done) {
#pragma unroll
            for (int i = 0; i < ROWS; i++) {
                R acc = 0, acc2 = 0;
#pragma unroll
                for (int j = 0; j < 7; j++) {
                    if (j < 4 || j < NAX) {
                        const R gij = g[exo_g_word(i, j)];
                        acc += gij * tv[j];
                        if (INFO) acc2 += gij * ttc[j];
                    }
                }
                qa[i] = acc * r2d;
                if (INFO) qt[i] = acc2 * r2d;
            }
        }
        // ---- angle composition + position motors + one Bullet step (:421-433, oracle/minibullet.py)
        const R tgt[5] = {(mrow[4] + qa[2]) * d2r, (mrow[3] + qa[0]) * d2r, (mrow[2] + qa[1]) * d2r, (mrow[0] + qa[3]) * d2r,
                          mrow[1] * d2r};
        int crossed = 0;
        if (FULL) {                                   // hand the targets to exo_motor_kernel; q itself is its output
            const R t0[4] = {tgt[0], tgt[1], tgt[2], tgt[3]}, t1[4] = {tgt[4], R(1), R(0), R(0)};
            st4(S.tgt + e, t0); st4(S.tgt + (size_t)st + e, t1);
        }
#pragma unroll
        for (int j = 0; j < 5; j++) {
            if (FULL) { if (K.lag) s[W_QLAG + j] = qb[j]; continue; }
            R qn = qb[j] + R(0.025) * div_r(R(0.1) * (tgt[j] - qb[j]), R(0.025));   // q + dt*v*, v* = kp*(target-q)/dt
            if (qn < C.lower[j] || qn > C.upper[j]) {
                crossed++;
                if (K.limits) qn = qn < C.lower[j] ? C.lower[j] : C.upper[j];
            }
            s[W_Q + j] = qn;
            if (K.lag) s[W_QLAG + j] = qb[j];
        }
#pragma unroll
        for (int k = 0; k < 3; k++) { s[W_ELREF + k] = P.elref[k]; b_el[k] = P.elref[k]; }
#pragma unroll
        for (int j = 0; j < 5; j++) b_qs[j] = K.lag ? qb[j] : s[W_Q + j];
        // ---- reward (:268-366)
        const R eps = R(1e-10);
        const R m = fm_el + fm_sh;
        const int nax = __popc(meta & 0x7f);
        const R nf = (R)nax;
        R sum_unw = 0, sum_tor = 0, sumF = 0, smooth = 0;
        int axis_cnt = 0, arm_red = 0;
#pragma unroll
        for (int j = 0; j < 7; j++) {
            const R a = abs_r(tv[j]), b = abs_r(ttc[j]);
            if (j < 4) {
                if ((meta >> j) & 1) sum_tor += div_r(a - b, b) + R(1);
                else sum_unw += a;
            }
            if (j < 4 || j < NAX) {
                const R red = div_r(a - b, b + eps) * R(100);
                if (red < R(0)) { axis_cnt++; if (j < 4) arm_red = 1; }      // nan/inf -> 0 -> not counted (:302-303)
            }   // axes without a table entry: a == b == 0 -> reduction 0 -> not counted
        }
#pragma unroll
        for (int i = 0; i < 7; i++) {
            const R fp2 = s[W_FPREV2 + i];
            const R d = F[i] - R(2) * fp1[i] + fp2;
            smooth += d * d;
            sumF += F[i];
            s[W_FPREV2 + i] = fp1[i];
            s[W_FPREV + i] = F[i];
        }
        const R r_unw = exp_r(-div_r(sum_unw, div_r(m / R(4), nf)) + eps) * R(0.5);
        const R r_tor = exp_r(div_r(-sum_tor + eps, nf)) * R(0.9);
        const R r_axis = (R)axis_cnt * R(0.5);
        const R r_ctl = exp_r(-div_r(sumF, m / R(2)) + eps) * R(0.05);
        const R r_smo = R(0.05) * exp_r(-div_r(div_r(smooth, R(7)), m / R(4)) + eps);
        const R max_reward = nf * R(0.5) + R(0.9) + R(0.05) + R(0.05) + R(0.5);
        const R rew = div_r(r_axis + r_tor + r_smo + r_ctl + r_unw, max_reward);
        reward[e] = rew;
        EXO_MARK(6);                                                   // reward done
        // ---- write the carried state back
#pragma unroll
        for (int v = 0; v < NV_STATE; v++) st4(S.state + (size_t)v * st + e, s + 4 * v);
        // ---- counts, observation head, done (:448-457)
        const int k1 = c + 1;
        S.counts[e] = k1 | (slot << EXO_SLOT_SHIFT);
#pragma unroll
        for (int k = 0; k < 3; k++) {
            ob[68 + k] = (float)C.sh_ref[k]; ob[71 + k] = (float)el_old[k];     // ref_prev (:383)
            ob[74 + k] = (float)C.sh_ref[k]; ob[77 + k] = (float)P.elref[k];    // ref (:387)
        }
        __syncwarp(live_mask);            // every live lane has read its staged parameters: the region becomes obs chunks 11..19
        {
            float4* sob = reinterpret_cast<float4*>(&s_epB[wid][0]) + lane * 9;
#pragma unroll
            for (int ch = 11; ch < 20; ch++) sob[ch - 11] = make_float4(ob[4 * ch], ob[4 * ch + 1], ob[4 * ch + 2], ob[4 * ch + 3]);
        }
        const bool dn = k1 >= L - 1;
        done[e] = dn ? 1 : 0;
        fin = dn;
        // ---- episode statistics (device-side; all-reduced over NCCL by the host)
        const R ret = ret0 + rew;
        S.ret[e] = dn ? R(0) : ret;
        R amp = 0;
        if (INFO) {
            // Exoskeleton_agent_train.py:158-170 with return_original_joint_angles at the incremented counts
            R b[7] = {mnext[2] * d2r, mnext[3] * d2r, mnext[4] * d2r, mnext[0] * d2r, mnext[1] * d2r, R(0), R(0)};
            R th[7], p0[3], p1[3], p2[3];
            dh_end_effector(b, K.L1, K.L2, p0);
#pragma unroll
            for (int i = 0; i < 7; i++) th[i] = qa[i] * d2r + b[i];
            dh_end_effector(th, K.L1, K.L2, p1);
#pragma unroll
            for (int i = 0; i < 7; i++) th[i] = qt[i] * d2r + b[i];
            dh_end_effector(th, K.L1, K.L2, p2);
            const R ds = sqrt_r((p1[0] - p0[0]) * (p1[0] - p0[0]) + (p1[1] - p0[1]) * (p1[1] - p0[1]) + (p1[2] - p0[2]) * (p1[2] - p0[2]));
            const R du = sqrt_r((p2[0] - p0[0]) * (p2[0] - p0[0]) + (p2[1] - p0[1]) * (p2[1] - p0[1]) + (p2[2] - p0[2]) * (p2[2] - p0[2]));
            amp = ((ds - du) / du) * R(100);
#pragma unroll
            for (int j = 0; j < 7; j++) {
                if (info.actuator_torques) info.actuator_torques[(size_t)e * 7 + j] = j < 4 ? at[j] : R(0);
                if (info.torque_val) info.torque_val[(size_t)e * 7 + j] = tv[j];
                if (info.ampl_val) info.ampl_val[(size_t)e * 7 + j] = qa[j];
                if (info.tremor_torque_val) info.tremor_torque_val[(size_t)e * 7 + j] = ttc[j];
                if (info.tremor_ampl_val) info.tremor_ampl_val[(size_t)e * 7 + j] = qt[j];
            }
            if (info.reward_parts) {
                R* rp = info.reward_parts + (size_t)e * 5;
                rp[0] = r_unw; rp[1] = r_tor; rp[2] = r_axis; rp[3] = r_ctl; rp[4] = r_smo;
            }
            if (info.amp_metric) info.amp_metric[e] = amp;
            // Exoskeleton_agent_train.py:150-155,172-173,186-188: per-axis suppression percentages with nan/inf -> 0
            // (a zero tremor value), occurrence counters over the four arm axes, and - with the script's `disregard`
            // masking, positive values dropped - the sums/counts behind its mean-of-non-zero suppression figures
#pragma unroll
            for (int j = 0; j < 7; j++) {
                const R tb = abs_r(ttc[j]), ab = abs_r(qt[j]);
                const R tr = tb > R(0) ? (abs_r(tv[j]) - tb) / tb * R(100) : R(0);
                const R ta = ab > R(0) ? (abs_r(qa[j]) - ab) / ab * R(100) : R(0);
                if (j < 4) { if (tr >= R(0)) v_cge++; else v_clt++; }
                if (tr < R(0)) { v_str += tr; v_ntr++; }
                if (ta < R(0)) { v_sam += ta; v_nam++; }
            }
        }
        v_cross = crossed; v_red = arm_red;
        if (dn) { v_len = k1 - 2; v_ret = ret; }
        if (INFO && amp < R(0)) v_amp = amp;
    }
    // ---- auto-reset: an env whose episode just ended starts its next one HERE - the parameters were generated one episode
    //      ahead into its spare slot (exo_gen_kernel on the side stream), so this is a slot flip, counts = 2, the position
    //      reads of Exoskeleton_env.py:230-250 and the first observation (:252), which replaces the env's row in the staging
    //      tiles. Nothing on the step's critical path waits for a random draw.
    if (K.fused_begin && fin) {
        R qplb[EXO_NPRIS];
        if (FULL) {
#pragma unroll
            for (int d = 0; d < EXO_NPRIS; d++) qplb[d] = *exo_word(S.dyn, W_QP + d, e, st);
        }
        begin_episode<R, NAX>(S, C, e, slot ^ 1, b_qs, b_el, FULL ? qplb : nullptr, &s_obsA[wid][lane * 11],
                              reinterpret_cast<float4*>(&s_epB[wid][0]) + lane * 9,
                              stage_next ? &s_obsA[wid][lane * 11] : nullptr, stage_next ? &s_nxt[STAGE_NEXT ? wid : 0][STAGE_NEXT ? lane * 3 : 0] : nullptr);
    }
    // ---- warp-level epilogue: every lane of the warp arrives here
    const unsigned fin_mask = __ballot_sync(full, fin);
    if (live_mask == 0) return;
    __syncwarp();
    // staging tiles -> row-major [n,80]: the warp's 32 rows are 10 240 contiguous bytes; both tiles are read
    // linearly (conflict-free LDS.128) and written as 128-bit stores, 176 B + 144 B per row
    if (obs != nullptr) {
        float4* dst = reinterpret_cast<float4*>(obs + (size_t)wbase * 80);
        const float4* srcA = &s_obsA[wid][0];
        const float4* srcB = reinterpret_cast<const float4*>(&s_epB[wid][0]);
#pragma unroll
        for (int it = 0; it < 11; it++) {
            const int v = it * 32 + lane;
            const int row = v / 11, ch = v - row * 11;
            if ((live_mask >> row) & 1u) dst[row * 20 + ch] = srcA[v];
        }
#pragma unroll
        for (int it = 0; it < 9; it++) {
            const int v = it * 32 + lane;
            const int row = v / 9, ch = v - row * 9;
            if ((live_mask >> row) & 1u) dst[row * 20 + 11 + ch] = srcB[v];
        }
    }
    // finished envs -> this call's work list (one atomic per warp that has any): their vacated slot gets the episode after next
    if (K.auto_reset && fin_mask) {                                   // fin_mask == wf_mask: the places were reserved at load time
        const int base = __shfl_sync(full, fin_base, 0);
        if (fin) S.fin_list[(size_t)K.list * st + base + __popc(fin_mask & ((1u << lane) - 1u))] = e;
    }
    EXO_MARK(7);                                                     // observation rows written
    // statistics: integer counters by warp-reduce, the real-valued ones only in warps where an episode ended
    {
        const int n_live = __popc(live_mask);
        const int t_cross = __reduce_add_sync(full, v_cross);
        const int t_red = __reduce_add_sync(full, v_red);
        double t_ret = 0.0, t_amp = 0.0;
        int t_len = 0, t_ampn = 0;
        if (fin_mask) {
            t_len = __reduce_add_sync(full, v_len);
            double x = (double)v_ret;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(full, x, o);
            t_ret = x;
        }
        if (INFO) {
            const unsigned am = __ballot_sync(full, v_amp < R(0));
            t_ampn = __popc(am);
            if (am) {
                double x = (double)v_amp;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(full, x, o);
                t_amp = x;
            }
        }
        if (lane == 0) {
            if (fin_mask) {
                atomicAdd(S.stats + 0, (double)__popc(fin_mask)); atomicAdd(S.stats + 1, t_ret); atomicAdd(S.stats + 2, (double)t_len);
            }
            atomicAdd(S.stats + 3, (double)n_live);
            if (t_cross) atomicAdd(S.stats + 4, (double)t_cross);
            if (t_red) atomicAdd(S.stats + 5, (double)t_red);
            if (INFO && t_ampn) { atomicAdd(S.stats + 6, t_amp); atomicAdd(S.stats + 7, (double)t_ampn); }
        }
        if (INFO) {
            const int t_cge = __reduce_add_sync(full, v_cge), t_clt = __reduce_add_sync(full, v_clt);
            const int t_ntr = __reduce_add_sync(full, v_ntr), t_nam = __reduce_add_sync(full, v_nam);
            double xs = (double)v_str, xa = (double)v_sam;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) { xs += __shfl_xor_sync(full, xs, o); xa += __shfl_xor_sync(full, xa, o); }
            if (lane == 0) {
                atomicAdd(S.stats + 8, (double)t_cge); atomicAdd(S.stats + 9, (double)t_clt);
                if (t_ntr) { atomicAdd(S.stats + 10, xs); atomicAdd(S.stats + 11, (double)t_ntr); }
                if (t_nam) { atomicAdd(S.stats + 12, xa); atomicAdd(S.stats + 13, (double)t_nam); }
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// exo_get_field / exo_set_field: gather/scatter between the resident layout and env-major [n, k] buffers.
// ---------------------------------------------------------------------------------------------------------
// `next` != 0 addresses the prefetched (spare) slot of the per-episode fields instead of the one in use.
template <typename R, typename T, bool SET>
__global__ void exo_field_kernel(EnvArrays<R> S, int field, int k, int naxv, int next, T* __restrict__ buf) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (size_t)S.n * k) return;
    const int e = (int)(idx / k), i = (int)(idx - (size_t)e * k), st = S.stride;
    const int slot = ((S.counts[e] >> EXO_SLOT_SHIFT) & 1) ^ (next ? 1 : 0);
    T* a = nullptr;
    switch (field) {
        case F_Q: a = (T*)exo_word(S.state, W_Q + i, e, st); break;
        case F_Q_LAG: a = (T*)exo_word(S.state, W_QLAG + i, e, st); break;
        case F_POS_VEC: a = (T*)exo_word(S.state, W_POSVEC + i, e, st); break;
        case F_ELBOW_REF: a = (T*)exo_word(S.state, W_ELREF + i, e, st); break;
        case F_PREV_ACTION: a = (T*)exo_word(S.state, W_FPREV + i, e, st); break;
        case F_PREV_ACTION2: a = (T*)exo_word(S.state, W_FPREV2 + i, e, st); break;
        case F_FORCE_MAX: a = (T*)exo_word(S.ep + (size_t)slot * S.ep_slot, W_FMAX + i, e, st); break;
        case F_SHIFT: a = (T*)exo_word(S.ep + (size_t)slot * S.ep_slot, W_SHIFT + i, e, st); break;
        case F_G: a = (T*)exo_word(S.G + (size_t)slot * S.G_slot, exo_g_word(i / 7, i % 7), e, st); break;
        case F_IDS: a = (T*)(S.IDS + ((size_t)slot * S.n + e) * EXO_IDS_PITCH + i); break;
        case F_TREMOR_TABLE: {
            const int nax = k / S.table_len, row = i / nax, j = i - row * nax;
            a = (T*)&(S.tt + (size_t)slot * S.tt_slot)[((size_t)row * st + e) * naxv + (j >> 2)].v[j & 3];
            break;
        }
        case F_TREMOR_MAG: a = (T*)(S.mag + (size_t)slot * st + e); break;
        case F_COUNTS:                                                  // the slot bit is private: masked on get, kept on set
            if (SET) S.counts[e] = ((int)buf[idx] & EXO_COUNT_MASK) | (S.counts[e] & ~EXO_COUNT_MASK);
            else buf[idx] = (T)(S.counts[e] & EXO_COUNT_MASK);
            return;
        case F_META: a = (T*)(S.meta + e); break;
        case F_MT_STATE: a = (T*)(S.mt + (size_t)e * EXO_MT_PITCH + i); break;
        case F_EP_RETURN: a = (T*)(S.ret + e); break;
        case F_QD: if (!S.dyn) return; a = (T*)exo_word(S.dyn, W_QD + i, e, st); break;
        case F_Q_PRIS: if (!S.dyn) return; a = (T*)exo_word(S.dyn, W_QP + i, e, st); break;
        case F_Q_PRIS_LAG: if (!S.dyn) return; a = (T*)exo_word(S.dyn, W_QPLAG + i, e, st); break;
        default: return;
    }
    if (SET) *a = buf[idx]; else buf[idx] = *a;
}

}  // namespace exo
