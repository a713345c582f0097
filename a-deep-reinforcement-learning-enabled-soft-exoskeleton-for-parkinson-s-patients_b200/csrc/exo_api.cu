// exo_api.cu — host side of libexo_env.so: the C-ABI declared in include/exo_env.h.
// Owns the resident SoA state in HBM, flattens the URDF tree into the 5-joint chain constants and launches
// the kernels of exo_step.cuh / exo_reset.cuh on the caller's stream. No CPU compute path exists here.
#include <cuda_runtime.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <vector>

#include "../../include/exo_env.h"
#include "exo_launch.h"

using namespace exo;

static thread_local char g_err[512] = "";
static int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}
#define CK(call)                                                                                         \
    do {                                                                                                 \
        cudaError_t _e = (call);                                                                         \
        if (_e != cudaSuccess) return fail(EXO_ERR_CUDA, "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(_e)); \
    } while (0)

// shared with exo_lap.cu (one thread-local error string per library)
int exo_set_error(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}
extern "C" const char* exo_last_error(void) { return g_err; }
extern "C" int exo_abi_version(void) { return 1; }

#define EXO_GEN_STREAMS 4

struct ExoHandle {
    ExoConfig cfg;
    ExoGeometry geo;
    int n, stride, device, nax, max_len, n_motions, n_sm;
    // episode prefetch: generation passes run on s_gen, one work list of the ring per exo_step / exo_reset call
    int64_t tick;                // calls so far; ring slot = tick % EXO_RING
    int last_gen[EXO_GEN_STREAMS];   // per side stream: ring slot of its newest generation pass (-1: none)
    bool gen_ready;              // s_gen / events created
    bool ev_gen_used[EXO_RING];
    int n_gen_streams;
    cudaStream_t s_gen[EXO_GEN_STREAMS];   // passes of consecutive calls touch disjoint envs: they round-robin over these streams and
                                           // overlap, so a pass's ~50 us chain latency does not pace the steps (one in-order stream would)
    cudaEvent_t ev_gen[EXO_RING], ev_main[EXO_RING];
    bool have_motions, have_types, seeded;
    int64_t launches;
    std::vector<void*> allocs;
    EnvArrays<float> Af;
    EnvArrays<double> Ad;
    ChainConst<float> Cf;
    ChainConst<double> Cd;
    long long* timeline;         // debug: exo_debug_timeline
    DynConst<float> Df;          // EXO_MOTOR_FULL: inertial data + solver settings (exo_set_motor_model)
    DynConst<double> Dd;
    bool have_dyn;
    int dummy_link[EXO_NPRIS];   // URDF link index of dummy d (humerus slots 0..6, forearm 7..8)
    // staging for the host-buffer entry points
    float* d_act; float* d_obs; void* d_rew; uint8_t* d_done; uint8_t* d_mask;
    uint32_t* d_seeds;
    // exo_step_host pipeline: copy-in / copy-out streams and per-chunk events (created on first use)
    cudaStream_t s_in, s_out;
    cudaEvent_t ev_start, ev_done, ev_in[8], ev_k[8];
    bool pipe_ready;
    std::vector<int> motion_len;
};

static int dev_alloc(ExoHandle* h, void** p, size_t bytes, bool zero = true) {
    CK(cudaMalloc(p, bytes ? bytes : 16));
    h->allocs.push_back(*p);
    if (zero) CK(cudaMemset(*p, 0, bytes ? bytes : 16));
    return 0;
}

static void mat3mul(const double* A, const double* B, double* C) {
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) C[i * 3 + j] = A[i * 3] * B[j] + A[i * 3 + 1] * B[3 + j] + A[i * 3 + 2] * B[6 + j];
}

// Walk from a link up through prismatic joints (held at 0 by Bullet's default velocity motors) to the revolute
// link (or base) that carries it; returns the carrier index and the point in the carrier's frame.
static int carrier_of(const ExoGeometry* g, int link, double* off) {
    double p[3] = {g->com_xyz[link][0], g->com_xyz[link][1], g->com_xyz[link][2]};
    int l = link;
    while (l >= 0 && g->joint_type[l] == 1) {
        double q[3];
        for (int i = 0; i < 3; i++)
            q[i] = g->origin_xyz[l][i] + g->origin_R[l][i * 3] * p[0] + g->origin_R[l][i * 3 + 1] * p[1] + g->origin_R[l][i * 3 + 2] * p[2];
        memcpy(p, q, sizeof(p));
        l = g->parent[l];
    }
    memcpy(off, p, sizeof(p));
    return l;
}

template <typename R>
static void cast_chain(const ChainConst<double>& s, ChainConst<R>& d) {
    const double* sp = reinterpret_cast<const double*>(&s);
    R* dp = reinterpret_cast<R*>(&d);
    for (size_t i = 0; i < sizeof(ChainConst<double>) / sizeof(double); i++) dp[i] = (R)sp[i];
}

static int build_chain(ExoHandle* h) {
    const ExoGeometry* g = &h->geo;
    ChainConst<double>& C = h->Cd;
    memset(&C, 0, sizeof(C));
    for (int i = 0; i < 5; i++) {
        if (g->joint_type[i] != 0 || g->parent[i] != i - 1)
            return fail(EXO_ERR_TOPOLOGY, "links 0..4 must be a serial chain of revolute joints (link %d)", i);
        if (fabs(g->axis[i][0]) > 1e-12 || fabs(g->axis[i][1]) > 1e-12 || fabs(g->axis[i][2] - 1.0) > 1e-12)
            return fail(EXO_ERR_TOPOLOGY, "revolute joint %d: axis must be local z", i);
        memcpy(C.Rorg[i], g->origin_R[i], 9 * sizeof(double));
        memcpy(C.porg[i], g->origin_xyz[i], 3 * sizeof(double));
        C.lower[i] = g->lower[i];
        C.upper[i] = g->upper[i];
    }
    for (int k = 0; k < 3; k++) C.p0[k] = g->base_pos[k] + g->origin_xyz[0][k];
    // expected carriers, in the order act11,act12,act21,...,act72 : humerus=2, forearm=4, base=-1
    static const int want[14] = {2, 4, 2, 4, -1, 2, -1, 2, -1, 2, -1, 2, -1, 2};
    int hum = 0, alk = 0, fix = 0;
    // slot order inside off_hum: act11, act21, act32, act42, act52, act62, act72
    static const int hum_slot[14] = {0, -1, 1, -1, -1, 2, -1, 3, -1, 4, -1, 5, -1, 6};
    for (int a = 0; a < 14; a++) {
        double off[3];
        const int c = carrier_of(g, g->act_link[a], off);
        if (c != want[a]) return fail(EXO_ERR_TOPOLOGY, "actuator endpoint %d rides on link %d, kernel expects %d", a, c, want[a]);
        // prismatic axis of the dummy in its carrier's frame (the dummy's joint hangs directly off the carrier)
        const int dl = g->act_link[a];
        double pax[3] = {0, 0, 0};
        if (c >= 0) {
            if (g->joint_type[dl] != 1 || g->parent[dl] != c) return fail(EXO_ERR_TOPOLOGY, "actuator endpoint %d: expected a prismatic leaf on link %d", a, c);
            const double an = sqrt(g->axis[dl][0] * g->axis[dl][0] + g->axis[dl][1] * g->axis[dl][1] + g->axis[dl][2] * g->axis[dl][2]);
            for (int i = 0; i < 3; i++)
                pax[i] = (g->origin_R[dl][i * 3] * g->axis[dl][0] + g->origin_R[dl][i * 3 + 1] * g->axis[dl][1] + g->origin_R[dl][i * 3 + 2] * g->axis[dl][2]) / an;
        }
        if (c == 2) { memcpy(C.off_hum[hum_slot[a]], off, sizeof(off)); memcpy(C.pax_hum[hum_slot[a]], pax, sizeof(pax)); h->dummy_link[hum_slot[a]] = dl; hum++; }
        else if (c == 4) { memcpy(C.off_alk[alk], off, sizeof(off)); memcpy(C.pax_alk[alk], pax, sizeof(pax)); h->dummy_link[7 + alk] = dl; alk++; }
        else { for (int k = 0; k < 3; k++) C.fixed[fix][k] = g->base_pos[k] + off[k]; fix++; }
    }
    double off[3];
    if (carrier_of(g, g->shoulder_link, off) != 0 || fabs(off[0]) + fabs(off[1]) + fabs(off[2]) > 0)
        return fail(EXO_ERR_TOPOLOGY, "shoulder reference must be the origin of link 0");
    memcpy(C.sh_ref, C.p0, sizeof(C.p0));
    if (carrier_of(g, g->elbow_link, off) != 3) return fail(EXO_ERR_TOPOLOGY, "elbow reference must ride on link 3");
    memcpy(C.off_elbow, off, sizeof(off));
    cast_chain(C, h->Cf);
    return 0;
}

extern "C" void exo_default_motor_model(ExoMotorModel* m) {
    if (!m) return;
    memset(m, 0, sizeof(*m));
    m->inertia_from_file = 0; m->iterations = 50; m->row_order = 0; m->residual_threshold = 1e-7;
    m->kp = 0.1; m->kd = 1.0; m->max_force = 1.0e5; m->default_motor_impulse = 1.0;
    m->linear_damping = 0.04; m->angular_damping = 0.04;
    m->gravity[0] = 0; m->gravity[1] = 0; m->gravity[2] = -9.81;
    m->erp = 0.2; m->limit_max_impulse = 100.0;
}

// Order in which btAlignedObjectArray::quickSort leaves n elements whose keys all compare equal (the world's multibody
// constraints all carry one island id): the partition step swaps symmetric pairs, then both halves recurse.
static void bullet_quicksort_equal(std::vector<int>& a, int lo, int hi) {
    int i = lo, j = hi;
    do {
        if (i <= j) { std::swap(a[i], a[j]); i++; j--; }
    } while (i <= j);
    if (lo < j) bullet_quicksort_equal(a, lo, j);
    if (i < hi) bullet_quicksort_equal(a, i, hi);
}

// principal inertias of link l in its inertial frame: the URDF's, or (pybullet default) those of the collision box's AABB
static void link_inertia_diag(const ExoLinkDynamics* L, int l, bool from_file, double* diag) {
    if (from_file) { memcpy(diag, L->inertia_diag[l], 3 * sizeof(double)); return; }
    double ext[3];
    for (int i = 0; i < 3; i++) {            // |R_inertial^T box_R| * size
        ext[i] = 0;
        for (int j = 0; j < 3; j++) {
            double rij = 0;
            for (int k = 0; k < 3; k++) rij += L->inertial_R[l][k * 3 + i] * L->box_R[l][k * 3 + j];
            ext[i] += fabs(rij) * L->box_size[l][j];
        }
    }
    const double m = L->mass[l] / 12.0;
    diag[0] = m * (ext[1] * ext[1] + ext[2] * ext[2]); diag[1] = m * (ext[0] * ext[0] + ext[2] * ext[2]); diag[2] = m * (ext[0] * ext[0] + ext[1] * ext[1]);
}

// T = A diag A^T as (xx,xy,xz,yy,yz,zz)
static void sym_from_diag(const double* A, const double* diag, double* out6) {
    double M[9];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) M[i * 3 + j] = A[i * 3] * diag[0] * A[j * 3] + A[i * 3 + 1] * diag[1] * A[j * 3 + 1] + A[i * 3 + 2] * diag[2] * A[j * 3 + 2];
    out6[0] = M[0]; out6[1] = M[1]; out6[2] = M[2]; out6[3] = M[4]; out6[4] = M[5]; out6[5] = M[8];
}

template <typename R>
static void cast_dyn(const DynConst<double>& s, DynConst<R>& d) {
#define CP(f) for (size_t i = 0; i < sizeof(s.f) / sizeof(double); i++) ((R*)&d.f)[i] = (R)((const double*)&s.f)[i]
    CP(m_link); CP(com_link); CP(I_link); CP(m_d); CP(I_d); CP(g);
#undef CP
    d.kp = (R)s.kp; d.kd = (R)s.kd; d.cap_pos = (R)s.cap_pos; d.cap_vel = (R)s.cap_vel; d.k_lin = (R)s.k_lin; d.k_ang = (R)s.k_ang;
    d.dt = (R)s.dt; d.erp = (R)s.erp; d.limit_cap = (R)s.limit_cap; d.threshold = (R)s.threshold;
    d.iterations = s.iterations; d.n_cons = s.n_cons;
    memcpy(d.cons, s.cons, sizeof(s.cons));
    memcpy(d.slot_dof, s.slot_dof, sizeof(s.slot_dof)); memcpy(d.dof_slot, s.dof_slot, sizeof(s.dof_slot));
    memcpy(d.tri, s.tri, sizeof(s.tri)); memcpy(d.etri, s.etri, sizeof(s.etri));
}

extern "C" int exo_set_motor_model(ExoHandle* h, const ExoLinkDynamics* L, const ExoMotorModel* model) {
    if (!h || !L) return fail(EXO_ERR_ARG, "exo_set_motor_model: bad arguments");
    if (h->cfg.motor_model != EXO_MOTOR_FULL) return fail(EXO_ERR_STATE, "exo_set_motor_model: the handle was created with EXO_MOTOR_IDEAL");
    ExoMotorModel mm;
    if (model) mm = *model; else exo_default_motor_model(&mm);
    if (mm.iterations < 1 || mm.iterations > 100000) return fail(EXO_ERR_ARG, "exo_set_motor_model: iterations out of range");
    const ExoGeometry* g = &h->geo;
    DynConst<double>& D = h->Dd;
    memset(&D, 0, sizeof(D));
    for (int i = 0; i < 5; i++) {
        double diag[3];
        link_inertia_diag(L, i, mm.inertia_from_file != 0, diag);
        D.m_link[i] = L->mass[i];
        memcpy(D.com_link[i], g->com_xyz[i], 3 * sizeof(double));
        sym_from_diag(L->inertial_R[i], diag, D.I_link[i]);
    }
    for (int d = 0; d < EXO_NPRIS; d++) {
        const int l = h->dummy_link[d];
        double diag[3], A[9];
        link_inertia_diag(L, l, mm.inertia_from_file != 0, diag);
        mat3mul(g->origin_R[l], L->inertial_R[l], A);          // inertial frame in the carrier frame
        D.m_d[d] = L->mass[l];
        if (!(D.m_d[d] > 0)) return fail(EXO_ERR_ARG, "exo_set_motor_model: dummy link %d has no mass", l);
        sym_from_diag(A, diag, D.I_d[d]);
    }
    D.dt = 1.0 / 40.0;                                           // Exoskeleton_env.py:59,117
    D.kp = mm.kp; D.kd = mm.kd; D.cap_pos = mm.max_force * D.dt; D.cap_vel = mm.default_motor_impulse;
    D.k_lin = mm.linear_damping; D.k_ang = mm.angular_damping;
    memcpy(D.g, mm.gravity, sizeof(D.g));
    D.erp = mm.erp; D.limit_cap = mm.limit_max_impulse; D.threshold = mm.residual_threshold; D.iterations = mm.iterations;
    // world constraint list: joint-limit constraints in link order (URDF import), then createJointMotors' motors in link order
    std::vector<int> cons;                     // link index for a motor, -(link+1) for a limit constraint
    for (int l = 0; l < EXO_NLINKS; l++) if (g->lower[l] <= g->upper[l]) cons.push_back(-(l + 1));
    for (int l = 0; l < EXO_NLINKS; l++) cons.push_back(l);
    if (mm.row_order == 0 && cons.size() > 1) bullet_quicksort_equal(cons, 0, (int)cons.size() - 1);
    int dof_of[EXO_NLINKS];
    for (int l = 0; l < EXO_NLINKS; l++) dof_of[l] = l < 5 ? l : -1;
    for (int d = 0; d < EXO_NPRIS; d++) dof_of[h->dummy_link[d]] = 5 + d;
    D.n_cons = 0;
    for (int c : cons) {
        const int l = c >= 0 ? c : -c - 1, dof = dof_of[l];
        if (dof < 0) continue;                 // the dummies on the fixed base: decoupled, held at rest (exo_motor.cuh header)
        if (c < 0 && dof >= 5) continue;       // prismatic limits (+-0.5 m) are never reached
        if (D.n_cons >= EXO_MAX_ROWS) return fail(EXO_ERR_TOPOLOGY, "exo_set_motor_model: too many constraint rows");
        D.cons[D.n_cons++] = (signed char)(c >= 0 ? dof : -(dof + 1));
    }
    int slot = 0;
    for (int ci = 0; ci < D.n_cons; ci++)
        if (D.cons[ci] >= 0) { D.slot_dof[slot] = D.cons[ci]; D.dof_slot[(int)D.cons[ci]] = (signed char)slot; slot++; }
    if (slot != EXO_NDOF) return fail(EXO_ERR_TOPOLOGY, "exo_set_motor_model: %d motor rows, expected %d", slot, EXO_NDOF);
    for (int a = 0; a < EXO_NDOF; a++)
        for (int b = 0; b < EXO_NDOF; b++) D.tri[a][b] = (unsigned char)(a >= b ? a * (a + 1) / 2 + b : b * (b + 1) / 2 + a);
    memset(D.etri, 0, sizeof(D.etri));
    for (int a = 0; a < D.n_cons && a < 19; a++)
        for (int b = 0; b < D.n_cons && b < 19; b++) {
            const int da = D.cons[a] >= 0 ? D.cons[a] : -D.cons[a] - 1, db = D.cons[b] >= 0 ? D.cons[b] : -D.cons[b] - 1;
            D.etri[a][b] = D.tri[(int)D.dof_slot[da]][(int)D.dof_slot[db]];
        }
    cast_dyn(D, h->Df);
    h->have_dyn = true;
    return 0;
}

template <typename R>
static int alloc_arrays(ExoHandle* h, EnvArrays<R>& A) {
    const size_t st = h->stride, n = h->n;
    memset(&A, 0, sizeof(A));
    A.n = h->n; A.stride = h->stride;
#define AL(field, T, count) if (int rc = dev_alloc(h, (void**)&A.field, sizeof(T) * (size_t)(count))) return rc
    AL(state, V4<R>, NV_STATE * st); AL(ep, V4<R>, 2 * NV_EP * st); AL(G, V4<R>, 2 * NV_G * st);
    A.ep_slot = NV_EP * st; A.G_slot = NV_G * st;
    AL(ret, R, st); AL(counts, int, st); AL(meta, int, st); AL(mag, double, 2 * st);
    AL(IDS, double, 2 * n * EXO_IDS_PITCH); AL(mt, uint32_t, n * EXO_MT_PITCH);
    AL(stats, double, EXO_N_STATS); AL(fin_count, int, EXO_RING); AL(fin_list, int, (size_t)EXO_RING * st);
    if (h->cfg.motor_model == EXO_MOTOR_FULL) { AL(dyn, V4<R>, NV_DYN * st); AL(tgt, V4<R>, NV_TGT * st); }
#undef AL
    return 0;
}

extern "C" int exo_create(const ExoConfig* cfg, const ExoGeometry* geo, int32_t n_envs, int32_t device, ExoHandle** out) {
    if (!cfg || !geo || !out || n_envs <= 0) return fail(EXO_ERR_ARG, "exo_create: bad arguments");
    if (cfg->precision != EXO_F32 && cfg->precision != EXO_F64) return fail(EXO_ERR_ARG, "precision must be EXO_F32 or EXO_F64");
    if (cfg->ode != EXO_ODE_PROPAGATOR && cfg->ode != EXO_ODE_RK45) return fail(EXO_ERR_ARG, "ode must be EXO_ODE_PROPAGATOR or EXO_ODE_RK45");
    if (cfg->ode == EXO_ODE_RK45 && cfg->precision != EXO_F64) return fail(EXO_ERR_ARG, "EXO_ODE_RK45 is an fp64 verification mode: precision must be EXO_F64");
    if (cfg->motor_model != EXO_MOTOR_IDEAL && cfg->motor_model != EXO_MOTOR_FULL) return fail(EXO_ERR_ARG, "motor_model must be EXO_MOTOR_IDEAL or EXO_MOTOR_FULL");
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return fail(EXO_ERR_CUDA, "device %d not present (%d CUDA devices): this library has no CPU path", device, ndev);
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) return fail(EXO_ERR_CUDA, "device %d is sm_%d%d; libexo_env.so carries sm_100a code only", device, prop.major, prop.minor);
    CK(cudaSetDevice(device));
    ExoHandle* h = new ExoHandle();
    h->cfg = *cfg; h->geo = *geo; h->n = n_envs; h->stride = (n_envs + 31) & ~31; h->device = device;
    h->nax = 4; h->max_len = 0; h->n_motions = 0; h->have_motions = h->have_types = h->seeded = false; h->launches = 0;
    h->n_sm = prop.multiProcessorCount; h->pipe_ready = false; h->have_dyn = false;
    h->timeline = nullptr;
    h->tick = 0; h->gen_ready = false; h->n_gen_streams = EXO_GEN_STREAMS;
    if (const char* e = getenv("EXO_GEN_STREAMS")) { const int k = atoi(e); if (k >= 1 && k <= EXO_GEN_STREAMS) h->n_gen_streams = k; }
    for (int i = 0; i < EXO_GEN_STREAMS; i++) h->last_gen[i] = -1;
    for (int i = 0; i < EXO_RING; i++) h->ev_gen_used[i] = false;
    h->d_act = nullptr; h->d_obs = nullptr; h->d_rew = nullptr; h->d_done = nullptr; h->d_mask = nullptr; h->d_seeds = nullptr;
    if (int rc = build_chain(h)) { delete h; return rc; }
    int rc = (cfg->precision == EXO_F32) ? alloc_arrays(h, h->Af) : alloc_arrays(h, h->Ad);
    if (!rc) rc = dev_alloc(h, (void**)&h->d_act, sizeof(float) * 7 * (size_t)n_envs);
    if (!rc) rc = dev_alloc(h, (void**)&h->d_obs, sizeof(float) * 80 * (size_t)n_envs);
    if (!rc) rc = dev_alloc(h, (void**)&h->d_rew, sizeof(double) * (size_t)n_envs);
    if (!rc) rc = dev_alloc(h, (void**)&h->d_done, (size_t)n_envs);
    if (!rc) rc = dev_alloc(h, (void**)&h->d_mask, (size_t)n_envs);
    if (!rc) rc = dev_alloc(h, (void**)&h->d_seeds, sizeof(uint32_t) * (size_t)n_envs);
    if (rc) { exo_destroy(h); return rc; }
    *out = h;
    return 0;
}

extern "C" int exo_destroy(ExoHandle* h) {
    if (!h) return 0;
    cudaSetDevice(h->device);
    for (void* p : h->allocs) cudaFree(p);
    if (h->gen_ready) {
        for (int i = 0; i < EXO_GEN_STREAMS; i++) { cudaStreamSynchronize(h->s_gen[i]); cudaStreamDestroy(h->s_gen[i]); }
        for (int i = 0; i < EXO_RING; i++) { cudaEventDestroy(h->ev_gen[i]); cudaEventDestroy(h->ev_main[i]); }
    }
    if (h->pipe_ready) {
        cudaStreamDestroy(h->s_in); cudaStreamDestroy(h->s_out);
        cudaEventDestroy(h->ev_start); cudaEventDestroy(h->ev_done);
        for (int i = 0; i < 8; i++) { cudaEventDestroy(h->ev_in[i]); cudaEventDestroy(h->ev_k[i]); }
    }
    delete h;
    return 0;
}

template <typename R>
static int load_motions_t(ExoHandle* h, EnvArrays<R>& A, const double* tables, const int32_t* lengths, int nm, int max_len) {
    std::vector<R> host((size_t)nm * 5 * max_len);
    for (size_t i = 0; i < host.size(); i++) host[i] = (R)tables[i];
    R* d; int* dl;
    if (int rc = dev_alloc(h, (void**)&d, sizeof(R) * host.size(), false)) return rc;
    if (int rc = dev_alloc(h, (void**)&dl, sizeof(int) * nm, false)) return rc;
    CK(cudaMemcpy(d, host.data(), sizeof(R) * host.size(), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dl, lengths, sizeof(int) * nm, cudaMemcpyHostToDevice));
    A.motion = d; A.motion_len = dl; A.max_len = max_len; A.n_motions = nm;
    return 0;
}

extern "C" int exo_load_motions(ExoHandle* h, const double* tables, const int32_t* lengths, int32_t n_motions, int32_t max_len) {
    if (!h || !tables || !lengths || n_motions <= 0 || n_motions > 255) return fail(EXO_ERR_ARG, "exo_load_motions: bad arguments");
    if (h->have_motions) return fail(EXO_ERR_STATE, "motions already loaded");
    for (int m = 0; m < n_motions; m++)
        if (lengths[m] < 5 || lengths[m] > max_len || lengths[m] > EXO_MAX_LEN) return fail(EXO_ERR_ARG, "motion %d: length %d out of range", m, lengths[m]);
    CK(cudaSetDevice(h->device));
    int rc = h->cfg.precision == EXO_F32 ? load_motions_t(h, h->Af, tables, lengths, n_motions, max_len)
                                         : load_motions_t(h, h->Ad, tables, lengths, n_motions, max_len);
    if (rc) return rc;
    h->motion_len.assign(lengths, lengths + n_motions);
    h->max_len = max_len; h->n_motions = n_motions; h->have_motions = true;
    return 0;
}

template <typename R>
static int alloc_tables(ExoHandle* h, EnvArrays<R>& A) {
    int tl = 0;
    for (int l : h->motion_len) tl = l > tl ? l : tl;
    A.table_len = tl;
    const size_t naxv = (h->nax + 3) / 4;
    A.tt_slot = (size_t)tl * h->stride * naxv;
    return dev_alloc(h, (void**)&A.tt, sizeof(V4<R>) * 2 * A.tt_slot);
}

extern "C" int exo_set_env_types(ExoHandle* h, const uint8_t* seq, const int32_t* motion_id) {
    if (!h || !seq || !motion_id) return fail(EXO_ERR_ARG, "exo_set_env_types: bad arguments");
    if (!h->have_motions) return fail(EXO_ERR_STATE, "call exo_load_motions first");
    if (h->have_types) return fail(EXO_ERR_STATE, "env types already set");
    CK(cudaSetDevice(h->device));
    std::vector<int> meta(h->stride, 0);
    int nax = 4;
    for (int e = 0; e < h->n; e++) {
        int bits = 0;
        for (int j = 0; j < 7; j++) {
            if (seq[(size_t)e * 7 + j] > 1) return fail(EXO_ERR_ARG, "tremor_sequence entries must be 0 or 1 (env %d)", e);
            bits |= (seq[(size_t)e * 7 + j] & 1) << j;
        }
        if (bits >> 4) nax = 7;
        if (motion_id[e] < 0 || motion_id[e] >= h->n_motions) return fail(EXO_ERR_ARG, "env %d: motion id %d out of range", e, motion_id[e]);
        meta[e] = bits | (motion_id[e] << 8) | (h->motion_len[motion_id[e]] << 16);   // episode length rides along: no dependent load in the kernels
    }
    h->nax = nax;
    int* dmeta = h->cfg.precision == EXO_F32 ? h->Af.meta : h->Ad.meta;
    CK(cudaMemcpy(dmeta, meta.data(), sizeof(int) * h->stride, cudaMemcpyHostToDevice));
    int rc = h->cfg.precision == EXO_F32 ? alloc_tables(h, h->Af) : alloc_tables(h, h->Ad);
    if (rc) return rc;
    h->have_types = true;
    return 0;
}

template <typename R>
static StepConst<R> step_const(const ExoHandle* h) {
    StepConst<R> K;
    K.fmax_sh0 = (R)h->cfg.max_force_shoulder; K.fmax_el0 = (R)h->cfg.max_force_elbow;
    K.inv_fmax_sh0 = R(1) / K.fmax_sh0; K.inv_fmax_el0 = R(1) / K.fmax_el0;
    K.list = 0; K.fused_begin = 0;
    // staggered load issue (exo_step.cuh): one dispatch layer's share of the launch's HBM burst takes ~0.75 us at the measured
    // 6.5 TB/s (34 MB / 7 layers) = ~1470 SM cycles; only single-wave launches (<= 8 blocks per SM) are staggered - in a
    // multi-wave launch successive waves overlap their phases by themselves. EXO_STEP_STAGGER overrides (0 = off).
    static const int stagger = getenv("EXO_STEP_STAGGER") ? atoi(getenv("EXO_STEP_STAGGER")) : 1470;
    const int blocks = (h->n + EXO_STEP_THREADS - 1) / EXO_STEP_THREADS;
    K.stagger_cycles = (blocks > 2 * h->n_sm && blocks <= 8 * h->n_sm) ? stagger : 0; K.n_sm = h->n_sm;
    K.timeline = h->timeline;
    K.L1 = (R)h->cfg.humerus_length; K.L2 = (R)h->cfg.forearm_length;
    K.lag = h->cfg.link_state_lag; K.limits = h->cfg.joint_limits; K.auto_reset = h->cfg.auto_reset; K.metrics = h->cfg.metrics;
    K.motor_full = h->cfg.motor_model == EXO_MOTOR_FULL;
    K.ode_rk45 = h->cfg.ode == EXO_ODE_RK45;
    K.env0 = 0; K.env1 = h->n;
    return K;
}

static ResetConst reset_const(const ExoHandle* h) {
    ResetConst K;
    K.amp_lo = h->cfg.tremor_amplitude_range[0]; K.amp_hi = h->cfg.tremor_amplitude_range[1];
    K.f1_lo = h->cfg.first_harmonics_interval[0]; K.f1_hi = h->cfg.first_harmonics_interval[1];
    K.f2_lo = h->cfg.second_harmonics_interval[0]; K.f2_hi = h->cfg.second_harmonics_interval[1];
    K.shift_range = h->cfg.dr_actuator_end_pos_shift; K.act_range = h->cfg.dr_actuator_range; K.noise_frac = h->cfg.matrix_noise_fraction;
    K.fmax_sh0 = h->cfg.max_force_shoulder; K.fmax_el0 = h->cfg.max_force_elbow;
    K.dt = 1.0 / 40.0;                       // Exoskeleton_env.py:59
    K.lag = h->cfg.link_state_lag;
    K.motor_full = h->cfg.motor_model == EXO_MOTOR_FULL;
    K.ctor_skip = 0; K.list = -1; K.append = -1;
    return K;
}

static int gen_setup(ExoHandle* h) {
    if (h->gen_ready) return 0;
    // lowest priority: when blocks of a step kernel and of a generation pass are both waiting for an SM, the step goes first
    int prio_lo = 0, prio_hi = 0;
    CK(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    if (getenv("EXO_GEN_NOPRIO")) prio_lo = 0;
    for (int i = 0; i < EXO_GEN_STREAMS; i++) CK(cudaStreamCreateWithPriority(&h->s_gen[i], cudaStreamNonBlocking, prio_lo));
    for (int i = 0; i < EXO_RING; i++) {
        CK(cudaEventCreateWithFlags(&h->ev_gen[i], cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&h->ev_main[i], cudaEventDisableTiming));
    }
    h->gen_ready = true;
    return 0;
}

// `s` waits for every generation pass launched so far (they run in order on s_gen)
static int wait_all_gen(ExoHandle* h, cudaStream_t s) {
    if (h->gen_ready)
        for (int i = 0; i < EXO_GEN_STREAMS; i++)
            if (h->last_gen[i] >= 0) CK(cudaStreamWaitEvent(s, h->ev_gen[h->last_gen[i]], 0));
    return 0;
}

// generation pass for work list `list` (filled by kernels already queued on `s`), on the side stream
template <typename R>
static int gen_list_async(ExoHandle* h, EnvArrays<R>& A, int list, cudaStream_t s, int expect) {
    if (int rc = gen_setup(h)) return rc;
    const int si = list % h->n_gen_streams;
    cudaStream_t g = h->s_gen[si];
    CK(cudaEventRecord(h->ev_main[list], s));
    CK(cudaStreamWaitEvent(g, h->ev_main[list], 0));
    ResetConst K = reset_const(h);
    K.list = list;
    if (int e = launch_gen<R>(A, K, h->nax, h->n_sm, nullptr, expect, g))
        return fail(EXO_ERR_CUDA, "exo_gen_kernel: %s", cudaGetErrorString((cudaError_t)e));
    h->launches += 1;
    CK(cudaEventRecord(h->ev_gen[list], g));
    h->ev_gen_used[list] = true;
    h->last_gen[si] = list;
    return 0;
}

// env.reset() for the masked envs: start their prefetched episode (exo_begin_kernel) on `s`, refill the spare slot on s_gen
template <typename R>
static int reset_t(ExoHandle* h, EnvArrays<R>& A, const ChainConst<R>& C, const uint8_t* mask, float* obs, cudaStream_t s) {
    if (int rc = wait_all_gen(h, s)) return rc;        // an env reset here may still have its spare slot in flight
    const int list = (int)(h->tick++ % EXO_RING);
    CK(cudaMemsetAsync(A.fin_count + list, 0, sizeof(int), s));
    ResetConst K = reset_const(h);
    K.append = list;
    launch_begin<R>(A, C, K, h->nax, mask, obs, s);
    h->launches += 1;
    CK(cudaGetLastError());
    return gen_list_async<R>(h, A, list, s, mask == nullptr ? h->n : 0);
}

static int check_ready(ExoHandle* h, const char* who) {
    if (!h) return fail(EXO_ERR_ARG, "%s: null handle", who);
    if (!h->seeded) return fail(EXO_ERR_STATE, "%s: call exo_load_motions, exo_set_env_types and exo_seed first", who);
    if (h->cfg.motor_model == EXO_MOTOR_FULL && !h->have_dyn) return fail(EXO_ERR_STATE, "%s: EXO_MOTOR_FULL needs exo_set_motor_model", who);
    CK(cudaSetDevice(h->device));
    return 0;
}

// After np.random.seed: every env gets the parameters of its first episode into the spare slot (the invariant the step
// kernel's auto-reset and exo_reset rely on). With reference_ctor the constructor's draws come first: one rand()
// (Exoskeleton_env.py:70) and a whole initialize_movement (:172), which also leaves its position vectors in the state.
template <typename R>
static int seed_prime(ExoHandle* h, EnvArrays<R>& A, const ChainConst<R>& C, cudaStream_t s) {
    if (h->gen_ready) { for (int i = 0; i < EXO_GEN_STREAMS; i++) CK(cudaStreamSynchronize(h->s_gen[i])); }
    for (int i = 0; i < EXO_GEN_STREAMS; i++) h->last_gen[i] = -1;
    CK(cudaMemsetAsync(A.counts, 0, sizeof(int) * h->stride, s));
    ResetConst K = reset_const(h);
    if (h->cfg.reference_ctor) {
        K.ctor_skip = 2;
        if (int e = launch_gen<R>(A, K, h->nax, h->n_sm, nullptr, h->n, s)) return fail(EXO_ERR_CUDA, "exo_gen_kernel: %s", cudaGetErrorString((cudaError_t)e));
        K.ctor_skip = 0;
        launch_begin<R>(A, C, K, h->nax, nullptr, nullptr, s);
        h->launches += 2;
    }
    if (int e = launch_gen<R>(A, K, h->nax, h->n_sm, nullptr, h->n, s)) return fail(EXO_ERR_CUDA, "exo_gen_kernel: %s", cudaGetErrorString((cudaError_t)e));
    h->launches += 1;
    CK(cudaGetLastError());
    return 0;
}

extern "C" int exo_seed(ExoHandle* h, const uint32_t* seeds, void* stream) {
    if (!h || !seeds) return fail(EXO_ERR_ARG, "exo_seed: bad arguments");
    if (!h->have_types) return fail(EXO_ERR_STATE, "call exo_load_motions and exo_set_env_types first");
    CK(cudaSetDevice(h->device));
    cudaStream_t s = (cudaStream_t)stream;
    CK(cudaMemcpyAsync(h->d_seeds, seeds, sizeof(uint32_t) * h->n, cudaMemcpyHostToDevice, s));
    uint32_t* mt = h->cfg.precision == EXO_F32 ? h->Af.mt : h->Ad.mt;
    launch_seed(mt, h->n, h->d_seeds, s);
    h->launches += 1;
    CK(cudaGetLastError());
    h->seeded = true;
    int rc = h->cfg.precision == EXO_F32 ? seed_prime(h, h->Af, h->Cf, s) : seed_prime(h, h->Ad, h->Cd, s);
    if (rc) return rc;
    CK(cudaStreamSynchronize(s));            // seeds is a host buffer the caller may free
    return 0;
}

extern "C" int exo_reset(ExoHandle* h, const uint8_t* mask_dev, float* obs_dev, void* stream) {
    if (int rc = check_ready(h, "exo_reset")) return rc;
    if (!obs_dev) return fail(EXO_ERR_ARG, "exo_reset: obs_dev is null");
    cudaStream_t s = (cudaStream_t)stream;
    return h->cfg.precision == EXO_F32 ? reset_t(h, h->Af, h->Cf, mask_dev, obs_dev, s)
                                       : reset_t(h, h->Ad, h->Cd, mask_dev, obs_dev, s);
}

template <typename R> static const DynConst<R>& dyn_const(const ExoHandle* h);
template <> const DynConst<float>& dyn_const<float>(const ExoHandle* h) { return h->Df; }
template <> const DynConst<double>& dyn_const<double>(const ExoHandle* h) { return h->Dd; }

template <typename R>
static int step_t(ExoHandle* h, EnvArrays<R>& A, const ChainConst<R>& C, const float* actions, float* obs, void* reward,
                  uint8_t* done, const ExoInfo* info, cudaStream_t s, int env0 = 0, int env1 = -1) {
    const int list = (int)(h->tick++ % EXO_RING);
    StepConst<R> K = step_const<R>(h);
    K.list = list;
    K.fused_begin = K.auto_reset && !(K.motor_full && !K.lag);      // (general step without link-state lag: the first observation
                                                                    //  needs the joint angles the motor kernel is about to produce)
    static const bool dbg_no_fused = getenv("EXO_DEBUG_NO_FUSED") != nullptr, dbg_no_gen = getenv("EXO_DEBUG_NO_GEN") != nullptr;
    if (dbg_no_fused) K.fused_begin = 0;                            // timing experiments only
    if (env1 >= 0) { K.env0 = env0; K.env1 = env1; }
    // this launch clears the counter of ring slot list+1: the generation pass that last consumed it must have finished
    if (h->gen_ready && h->ev_gen_used[(list + 1) % EXO_RING]) CK(cudaStreamWaitEvent(s, h->ev_gen[(list + 1) % EXO_RING], 0));
    InfoPtrs<R> ip;
    memset(&ip, 0, sizeof(ip));
    const bool want_info = info != nullptr || h->cfg.metrics;
    if (info) {
        ip.actuator_torques = (R*)info->actuator_torques; ip.torque_val = (R*)info->torque_val; ip.ampl_val = (R*)info->ampl_val;
        ip.tremor_torque_val = (R*)info->tremor_torque_val; ip.tremor_ampl_val = (R*)info->tremor_ampl_val;
        ip.reward_parts = (R*)info->reward_parts; ip.amp_metric = (R*)info->amp_metric;
    }
    if (K.motor_full) {
        launch_step_full<R>(A, C, K, h->nax, want_info, actions, obs, (R*)reward, done, ip, s);
        launch_motor<R>(A, C, dyn_const<R>(h), K.lag, K.env0, K.env1, s);
        h->launches += 2;
    } else {
        launch_step<R>(A, C, K, h->nax, want_info, actions, obs, (R*)reward, done, ip, s);
        h->launches += 1;
    }
    CK(cudaGetLastError());
    if (K.auto_reset) {
        if (!K.fused_begin) {                      // start the finished envs' next episode after the motor kernel
            ResetConst RK = reset_const(h);
            RK.list = list;
            launch_begin<R>(A, C, RK, h->nax, nullptr, obs, s);
            h->launches += 1;
            CK(cudaGetLastError());
        }
        if (dbg_no_gen) return 0;                  // timing experiments only: the spare slots go stale
        return gen_list_async<R>(h, A, list, s, 0);   // refill the vacated slots, off the critical path
    }
    return 0;
}

extern "C" int exo_step(ExoHandle* h, const float* actions_dev, float* obs_dev, void* reward_dev, uint8_t* done_dev,
                        const ExoInfo* info, void* stream) {
    if (int rc = check_ready(h, "exo_step")) return rc;
    if (!actions_dev || !obs_dev || !reward_dev || !done_dev) return fail(EXO_ERR_ARG, "exo_step: null buffer");
    cudaStream_t s = (cudaStream_t)stream;
    return h->cfg.precision == EXO_F32 ? step_t(h, h->Af, h->Cf, actions_dev, obs_dev, reward_dev, done_dev, info, s)
                                       : step_t(h, h->Ad, h->Cd, actions_dev, obs_dev, reward_dev, done_dev, info, s);
}

extern "C" int exo_step_host(ExoHandle* h, const float* actions_host, float* obs_host, void* reward_host, uint8_t* done_host, void* stream) {
    if (int rc = check_ready(h, "exo_step_host")) return rc;
    if (!actions_host || !obs_host || !reward_host || !done_host) return fail(EXO_ERR_ARG, "exo_step_host: null buffer");
    cudaStream_t s = (cudaStream_t)stream;
    const size_t n = h->n, rsz = h->cfg.precision == EXO_F32 ? 4 : 8;
    // The 325 B per env that go back over PCIe bound this entry point (21 MB at 65 536 envs ~ 0.39 ms at 55 GB/s), so
    // the population is cut into chunks whose H2D copy, kernels and D2H copies run on three streams: while chunk i's
    // observations cross the bus, chunk i+1 steps. Small populations take one chunk.
    int chunks = n >= 16384 ? 4 : 1;
    if (const char* e = getenv("EXO_HOST_CHUNKS")) { const int c = atoi(e); if (c >= 1 && c <= 8 && n >= 64 * (size_t)c) chunks = c; }
    if (chunks > 1 && !h->pipe_ready) {
        CK(cudaStreamCreateWithFlags(&h->s_in, cudaStreamNonBlocking));
        CK(cudaStreamCreateWithFlags(&h->s_out, cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&h->ev_start, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&h->ev_done, cudaEventDisableTiming));
        for (int i = 0; i < 8; i++) {
            CK(cudaEventCreateWithFlags(&h->ev_in[i], cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&h->ev_k[i], cudaEventDisableTiming));
        }
        h->pipe_ready = true;
    }
    if (chunks == 1) {
        CK(cudaMemcpyAsync(h->d_act, actions_host, n * 7 * sizeof(float), cudaMemcpyHostToDevice, s));
        if (int rc = exo_step(h, h->d_act, h->d_obs, h->d_rew, h->d_done, nullptr, stream)) return rc;
        CK(cudaMemcpyAsync(obs_host, h->d_obs, n * 80 * sizeof(float), cudaMemcpyDeviceToHost, s));
        CK(cudaMemcpyAsync(reward_host, h->d_rew, n * rsz, cudaMemcpyDeviceToHost, s));
        CK(cudaMemcpyAsync(done_host, h->d_done, n, cudaMemcpyDeviceToHost, s));
        CK(cudaStreamSynchronize(s));
        return 0;
    }
    CK(cudaEventRecord(h->ev_start, s));                       // copies start after whatever the caller queued on `stream`
    CK(cudaStreamWaitEvent(h->s_in, h->ev_start, 0));
    CK(cudaStreamWaitEvent(h->s_out, h->ev_start, 0));
    const size_t per = ((n / chunks + 63) / 64) * 64;          // chunk boundaries on block boundaries of the step kernel
    for (int i = 0; i < chunks; i++) {
        const size_t b0 = (size_t)i * per, b1 = (i == chunks - 1) ? n : (b0 + per < n ? b0 + per : n);
        if (b0 >= b1) break;
        CK(cudaMemcpyAsync(h->d_act + b0 * 7, actions_host + b0 * 7, (b1 - b0) * 7 * sizeof(float), cudaMemcpyHostToDevice, h->s_in));
        CK(cudaEventRecord(h->ev_in[i], h->s_in));
        CK(cudaStreamWaitEvent(s, h->ev_in[i], 0));
        int rc = h->cfg.precision == EXO_F32
                     ? step_t(h, h->Af, h->Cf, h->d_act, h->d_obs, h->d_rew, h->d_done, nullptr, s, (int)b0, (int)b1)
                     : step_t(h, h->Ad, h->Cd, h->d_act, h->d_obs, h->d_rew, h->d_done, nullptr, s, (int)b0, (int)b1);
        if (rc) return rc;
        CK(cudaEventRecord(h->ev_k[i], s));
        CK(cudaStreamWaitEvent(h->s_out, h->ev_k[i], 0));
        CK(cudaMemcpyAsync(obs_host + b0 * 80, h->d_obs + b0 * 80, (b1 - b0) * 80 * sizeof(float), cudaMemcpyDeviceToHost, h->s_out));
    }
    // reward and done are small: one copy each behind the last chunk instead of a latency-bound copy per chunk
    CK(cudaMemcpyAsync(reward_host, h->d_rew, n * rsz, cudaMemcpyDeviceToHost, h->s_out));
    CK(cudaMemcpyAsync(done_host, h->d_done, n, cudaMemcpyDeviceToHost, h->s_out));
    CK(cudaEventRecord(h->ev_done, h->s_out));
    CK(cudaStreamWaitEvent(s, h->ev_done, 0));
    CK(cudaStreamSynchronize(s));
    return 0;
}

extern "C" int exo_reset_host(ExoHandle* h, const uint8_t* mask_host, float* obs_host, void* stream) {
    if (int rc = check_ready(h, "exo_reset_host")) return rc;
    if (!obs_host) return fail(EXO_ERR_ARG, "exo_reset_host: null buffer");
    cudaStream_t s = (cudaStream_t)stream;
    const size_t n = h->n;
    if (mask_host) CK(cudaMemcpyAsync(h->d_mask, mask_host, n, cudaMemcpyHostToDevice, s));
    if (int rc = exo_reset(h, mask_host ? h->d_mask : nullptr, h->d_obs, stream)) return rc;
    CK(cudaMemcpyAsync(obs_host, h->d_obs, n * 80 * sizeof(float), cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    return 0;
}

// elements per env and dtype of each EXO_FIELD_*
static int field_shape(const ExoHandle* h, int field, int* k, int* dtype) {
    static const int K0[EXO_N_FIELDS] = {5, 5, 21, 3, 7, 7, 2, 42, 49, 147, 0, 1, 1, 1, 625, 1, EXO_NDOF, EXO_NPRIS, EXO_NPRIS};
    static const int D0[EXO_N_FIELDS] = {0, 0, 0, 0, 0, 0, 0, 0, 0, EXO_DT_F64, 0, EXO_DT_F64, EXO_DT_I32, EXO_DT_I32, EXO_DT_U32, 0, 0, 0, 0};
    if (field < 0 || field >= EXO_N_FIELDS) return -1;
    *k = K0[field]; *dtype = D0[field];
    if (field == EXO_FIELD_TREMOR_TABLE) *k = (h->cfg.precision == EXO_F32 ? h->Af.table_len : h->Ad.table_len) * h->nax;
    return 0;
}

extern "C" int exo_field_info(ExoHandle* h, int32_t field, int32_t* k, int32_t* dtype) {
    if (!h || !k || !dtype) return fail(EXO_ERR_ARG, "exo_field_info: bad arguments");
    if (!h->have_types) return fail(EXO_ERR_STATE, "exo_field_info: env types not set yet");
    int kk, dd;
    if (field_shape(h, field & ~EXO_FIELD_NEXT, &kk, &dd)) return fail(EXO_ERR_ARG, "exo_field_info: unknown field %d", field);
    *k = kk; *dtype = dd;
    return 0;
}

static int field_io(ExoHandle* h, int32_t field, void* buf, int set, void* stream, const char* who) {
    if (!h || !buf) return fail(EXO_ERR_ARG, "%s: bad arguments", who);
    if (!h->have_types) return fail(EXO_ERR_STATE, "%s: env types not set yet", who);
    int k, dt;
    const int next = (field & EXO_FIELD_NEXT) ? 1 : 0;
    field &= ~EXO_FIELD_NEXT;
    if (field_shape(h, field, &k, &dt)) return fail(EXO_ERR_ARG, "%s: unknown field %d", who, field);
    if (field >= EXO_FIELD_QD && field <= EXO_FIELD_Q_PRIS_LAG && h->cfg.motor_model != EXO_MOTOR_FULL)
        return fail(EXO_ERR_STATE, "%s: field %d exists only with EXO_MOTOR_FULL", who, field);
    CK(cudaSetDevice(h->device));
    cudaStream_t s = (cudaStream_t)stream;
    const int naxv = (h->nax + 3) / 4;
    if (int rc = wait_all_gen(h, s)) return rc;          // the spare slots / MT keys may still be being written on the side stream
    if (h->cfg.precision == EXO_F32) launch_field<float>(h->Af, field, k, dt, naxv, next, buf, set, s);
    else launch_field<double>(h->Ad, field, k, dt, naxv, next, buf, set, s);
    h->launches += 1;
    CK(cudaGetLastError());
    return 0;
}

extern "C" int exo_get_field(ExoHandle* h, int32_t field, void* out_dev, void* stream) {
    return field_io(h, field, out_dev, 0, stream, "exo_get_field");
}
extern "C" int exo_set_field(ExoHandle* h, int32_t field, const void* in_dev, void* stream) {
    return field_io(h, field, const_cast<void*>(in_dev), 1, stream, "exo_set_field");
}

extern "C" int exo_stats(ExoHandle* h, double* out_dev, int32_t reset, void* stream) {
    if (int rc = check_ready(h, "exo_stats")) return rc;
    if (!out_dev) return fail(EXO_ERR_ARG, "exo_stats: null buffer");
    cudaStream_t s = (cudaStream_t)stream;
    double* st = h->cfg.precision == EXO_F32 ? h->Af.stats : h->Ad.stats;
    CK(cudaMemcpyAsync(out_dev, st, sizeof(double) * EXO_N_STATS, cudaMemcpyDeviceToDevice, s));
    if (reset) CK(cudaMemsetAsync(st, 0, sizeof(double) * EXO_N_STATS, s));
    return 0;
}

// Debug: per-warp phase timestamps of the next exo_step launches into buf_dev ([ceil(n/32)][8] int64; NULL: off).
// Word 0 globaltimer (ns) at entry, 1 SM id, 2..7 clock64 at: entry, all loads issued, staged data landed, kinematics +
// actuators done, reward done, observation rows written.
extern "C" int exo_debug_timeline(ExoHandle* h, void* buf_dev) {
    if (!h) return fail(EXO_ERR_ARG, "exo_debug_timeline: null handle");
    h->timeline = (long long*)buf_dev;
    return 0;
}

extern "C" int exo_wait_prefetch(ExoHandle* h, void* stream) {
    if (int rc = check_ready(h, "exo_wait_prefetch")) return rc;
    return wait_all_gen(h, (cudaStream_t)stream);
}

extern "C" int exo_set_auto_reset(ExoHandle* h, int32_t on) {
    if (!h) return fail(EXO_ERR_ARG, "exo_set_auto_reset: null handle");
    h->cfg.auto_reset = on ? 1 : 0;
    return 0;
}

extern "C" int64_t exo_launch_count(const ExoHandle* h) { return h ? h->launches : 0; }

extern "C" int exo_forward_kinematics(ExoHandle* h, const void* theta_dev, void* p_dev, int32_t n, double L1, double L2, double L3, void* stream) {
    if (!h || !theta_dev || !p_dev || n <= 0) return fail(EXO_ERR_ARG, "exo_forward_kinematics: bad arguments");
    (void)L3;                                // unused by the reference too (a = 0 in every D-H row)
    CK(cudaSetDevice(h->device));
    cudaStream_t s = (cudaStream_t)stream;
    if (h->cfg.precision == EXO_F32) launch_fk<float>((const float*)theta_dev, (float*)p_dev, n, (float)L1, (float)L2, s);
    else launch_fk<double>((const double*)theta_dev, (double*)p_dev, n, L1, L2, s);
    h->launches += 1;
    CK(cudaGetLastError());
    return 0;
}
