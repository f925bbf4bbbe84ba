// go1_warp.cuh -- device code of one physics substep (the mj_step pipeline the reference runs at
// go1_mujoco_env.py:481), ONE ENVIRONMENT PER WARP, state staged in shared memory.
//
// Why this mapping (profiles/r1_ncu_k_step_lane_per_leg_*): environments take different
// branches (contact sets, cone zones, Newton / line-search trip counts).  With one environment per warp
// all of that is warp-uniform control flow -- no intra-warp divergence between environments -- and the
// code is loops over lanes instead of per-lane unrolled straight-line code, so the hot Newton loop stays
// inside the instruction cache.  Lanes split the work of one environment:
//   bodies (13)        lane b      forward kinematics, inertia about the COM, RNE forces
//   dofs (18)          lane d      joint-space vectors live in registers of lane d (qacc, M*qacc, grad, search...)
//   (leg, component)   lane 6k+e   spatial 6-vectors along the four leg chains (body motion / force passes)
//   geoms (42)         lane g      plane narrowphase, two rounds
//   contacts           lane c      constraint rows of contact c (cone zones, forces, line-search polynomials)
//   6x6 entries (21)   lane e      per-body contact stiffness K_b = sum G' H G, composite pass, Schur complement
// The joint-space matrices (M, the Newton Hessian H = M + J'DJ, M + h*diag(damping)) are block-arrow
// (trunk 6x6 + four 3x3 leg blocks + 6x3 couplings): 117 numbers, assembled CRBA-style from 6x6 spatial
// matrices, so contact Jacobians are never materialised.
#pragma once
#include "go1_model.h"

#ifdef GO1_EMU_COUNT
extern long g_emu_rounds[4];
#endif
namespace go1 {

#define GO1_DEV __device__ __forceinline__
#define GO1_NOINL __device__ __noinline__

constexpr unsigned FULL = 0xffffffffu;
constexpr int NB = 13;      // bodies: 0 trunk, 1 + 3*leg + link
constexpr int NDOF = 18;    // 0..5 trunk (3 lin world, 3 rot body frame), 6 + 3*leg + link
#ifndef GO1_LS_REL
#define GO1_LS_REL 1e-1
#endif
#ifndef GO1_LOCKSTEP
#define GO1_LOCKSTEP 1
#endif
#ifndef GO1_LS_FAST_MAX
#define GO1_LS_FAST_MAX 10
#endif
#ifndef GO1_PHASE_BARRIERS
#define GO1_PHASE_BARRIERS 1  // block barriers at the phase boundaries inside a substep / Newton round (see GO1_PB)
#endif
// The warps of a block run different environments; data-dependent phases (robot-robot narrowphase, pair Hessian, the
// coupled-leg factorisation, line-search trip counts) let them drift apart inside a round, and every drifted warp is its
// own instruction-fetch stream (the kernel is bound by instruction delivery).  A block barrier at
// each phase boundary regroups them.  Every warp of the block executes the same NUMBER of barriers per round / substep.
#ifndef GO1_PB_MASK
#define GO1_PB_MASK 31         // which phase barriers exist: 1 F1 (before the pair collision), 2 F2 (after it), 4 B1 (before the Hessian),
#endif                         // 8 B2 (before the factorisation), 16 B3 (before the line search); timing experiments only
#if defined(GO1_PHASE_TIMING) && !defined(GO1_EMU)
// Diagnostic build only (tools/phase_times.py): clock64 at every block barrier.  Bucket k holds the WORK a warp did since the
// previous barrier (twork[k]) and the time it then WAITED at barrier k (twait[k]); k: 0 F1, 1 F2, 2 B1, 3 B2, 4 B3, 5 the
// per-round barrier at the loop top, 6 the barrier at the start of a substep.
#define GO1_PT_SYNC(k, stmt) do { const long long t0_ = clock64(); out.twork[k] += t0_ - out.tlast; stmt; const long long t1_ = clock64(); \
                                  out.twait[k] += t1_ - t0_; out.tlast = t1_; } while (0)
#define GO1_PT_K(bit) ((bit) == 1 ? 0 : (bit) == 2 ? 1 : (bit) == 4 ? 2 : (bit) == 8 ? 3 : 4)
#define GO1_PBX(bit) do { if ((GO1_PB_MASK & (bit)) && (bit) <= 16) GO1_PT_SYNC(GO1_PT_K(bit), __syncthreads()); } while (0)
#elif GO1_LOCKSTEP && GO1_PHASE_BARRIERS
#define GO1_PBX(bit) do { if (GO1_PB_MASK & (bit)) __syncthreads(); } while (0)
#else
#define GO1_PBX(bit) ((void)0)
#endif
#define GO1_PB3() do { GO1_PBX(4); GO1_PBX(8); GO1_PBX(16); } while (0)
#ifndef GO1_WARM_EXTRAP
#define GO1_WARM_EXTRAP 1     // fp32 build: warm start = 2 a(t-1) - a(t-2) inside a control step (see substep_w)
#endif
#ifndef GO1_GRAD_BOUND_EXIT
#define GO1_GRAD_BOUND_EXIT 1
#endif
#ifndef GO1_SKIP_SMOOTH_CHECK
#define GO1_SKIP_SMOOTH_CHECK 1
#endif
#ifndef GO1_DEC_EXIT
#define GO1_DEC_EXIT 1        // fp32 build: Newton-decrement exit before the line search (see the Newton loop)
#endif
#ifndef GO1_CW
#define GO1_CW 40
#endif
#ifndef GO1_SELF_COLLISION
#define GO1_SELF_COLLISION 1   // robot-robot contacts (0: floor contacts only, the round-1 physics; timing A/B only)
#endif
constexpr int CW = GO1_CW;  // contacts kept per environment (geom order; the rest is counted as overflow)
#ifndef GO1_PW
#define GO1_PW 16
#endif
constexpr int PW = GO1_PW;  // robot-robot contacts kept per environment (stored after the floor contacts; both share the CW slots)
constexpr int NARROW = 117; // block-arrow storage: tt[21] | tk[4][6][3] | kk[4][6]
constexpr int A_TK = 21, A_KK = 93;

template <typename T> struct Num;
template <> struct Num<float> {
    static GO1_DEV float sqrt_(float x) { return sqrtf(x); }
    static GO1_DEV float rsqrt_(float x) { return rsqrtf(x); }
    static GO1_DEV void sincos_(float x, float* s, float* c) { sincosf(x, s, c); }
    static GO1_DEV float pow_(float a, float b) { return powf(a, b); }
    static GO1_DEV float abs_(float a) { return fabsf(a); }
    static GO1_DEV float max_(float a, float b) { return fmaxf(a, b); }
    static GO1_DEV float min_(float a, float b) { return fminf(a, b); }
    static constexpr float MINVAL = 1e-15f;
    static constexpr float GEOM_EPS = 1e-6f;      // below this distance between two centres the direction is rounding noise (fp32)
};
template <> struct Num<double> {
    static GO1_DEV double sqrt_(double x) { return sqrt(x); }
    static GO1_DEV double rsqrt_(double x) { return 1.0 / sqrt(x); }
    static GO1_DEV void sincos_(double x, double* s, double* c) { sincos(x, s, c); }
    static GO1_DEV double pow_(double a, double b) { return pow(a, b); }
    static GO1_DEV double abs_(double a) { return fabs(a); }
    static GO1_DEV double max_(double a, double b) { return fmax(a, b); }
    static GO1_DEV double min_(double a, double b) { return fmin(a, b); }
    static constexpr double MINVAL = 1e-15;
    static constexpr double GEOM_EPS = 1e-15;     // mjMINVAL, as MuJoCo and the oracle
};

// butterfly: every lane gets the bit-identical sum
template <typename T> GO1_DEV T warp_sum(T v) {
    v += __shfl_xor_sync(FULL, v, 16);
    v += __shfl_xor_sync(FULL, v, 8);
    v += __shfl_xor_sync(FULL, v, 4);
    v += __shfl_xor_sync(FULL, v, 2);
    v += __shfl_xor_sync(FULL, v, 1);
    return v;
}

template <typename T> GO1_DEV T dot3(const T* a, const T* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
template <typename T> GO1_DEV void cross3(T* r, const T* a, const T* b) {
    T x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
    r[0] = x; r[1] = y; r[2] = z;
}
template <typename T> GO1_DEV void matvec3(T* r, const T* R, const T* v) {  // R row-major
    T x = R[0] * v[0] + R[1] * v[1] + R[2] * v[2], y = R[3] * v[0] + R[4] * v[1] + R[5] * v[2], z = R[6] * v[0] + R[7] * v[1] + R[8] * v[2];
    r[0] = x; r[1] = y; r[2] = z;
}
template <typename T> GO1_DEV T dot6(const T* a, const T* b) {
    return a[0] * b[0] + a[1] * b[1] + a[2] * b[2] + a[3] * b[3] + a[4] * b[4] + a[5] * b[5];
}
// mju_mulInertVec
template <typename T> GO1_DEV void mul_inert(T* r, const T* i, const T* v) {
    T r0 = i[0] * v[0] + i[3] * v[1] + i[4] * v[2] - i[8] * v[4] + i[7] * v[5];
    T r1 = i[3] * v[0] + i[1] * v[1] + i[5] * v[2] + i[8] * v[3] - i[6] * v[5];
    T r2 = i[4] * v[0] + i[5] * v[1] + i[2] * v[2] - i[7] * v[3] + i[6] * v[4];
    T r3 = i[8] * v[1] - i[7] * v[2] + i[9] * v[3];
    T r4 = i[6] * v[2] - i[8] * v[0] + i[9] * v[4];
    T r5 = i[7] * v[0] - i[6] * v[1] + i[9] * v[5];
    r[0] = r0; r[1] = r1; r[2] = r2; r[3] = r3; r[4] = r4; r[5] = r5;
}
// mju_crossMotion / mju_crossForce
template <typename T> GO1_DEV void cross_motion(T* r, const T* vel, const T* v) {
    T a[3], b[3], c[3];
    cross3(a, vel, v); cross3(b, vel, v + 3); cross3(c, vel + 3, v);
    r[0] = a[0]; r[1] = a[1]; r[2] = a[2]; r[3] = b[0] + c[0]; r[4] = b[1] + c[1]; r[5] = b[2] + c[2];
}
template <typename T> GO1_DEV void cross_force(T* r, const T* vel, const T* f) {
    T a[3], b[3], c[3];
    cross3(a, vel, f); cross3(b, vel + 3, f + 3); cross3(c, vel, f + 3);
    r[0] = a[0] + b[0]; r[1] = a[1] + b[1]; r[2] = a[2] + b[2]; r[3] = c[0]; r[4] = c[1]; r[5] = c[2];
}
// mju_inertCom with a body-frame inertia tensor Ib = (xx,yy,zz,xy,xz,yz): R Ib R' + m(|d|^2 1 - d d'), m d, m
template <typename T> GO1_DEV void cinert_of(T* res, const T* R, const T* Ib, T mass, const T* dif) {
    T t[9];
#pragma unroll
    for (int r = 0; r < 3; r++) {
        t[3 * r + 0] = R[3 * r] * Ib[0] + R[3 * r + 1] * Ib[3] + R[3 * r + 2] * Ib[4];
        t[3 * r + 1] = R[3 * r] * Ib[3] + R[3 * r + 1] * Ib[1] + R[3 * r + 2] * Ib[5];
        t[3 * r + 2] = R[3 * r] * Ib[4] + R[3 * r + 1] * Ib[5] + R[3 * r + 2] * Ib[2];
    }
    res[0] = t[0] * R[0] + t[1] * R[1] + t[2] * R[2] + mass * (dif[1] * dif[1] + dif[2] * dif[2]);
    res[1] = t[3] * R[3] + t[4] * R[4] + t[5] * R[5] + mass * (dif[0] * dif[0] + dif[2] * dif[2]);
    res[2] = t[6] * R[6] + t[7] * R[7] + t[8] * R[8] + mass * (dif[0] * dif[0] + dif[1] * dif[1]);
    res[3] = t[0] * R[3] + t[1] * R[4] + t[2] * R[5] - mass * dif[0] * dif[1];
    res[4] = t[0] * R[6] + t[1] * R[7] + t[2] * R[8] - mass * dif[0] * dif[2];
    res[5] = t[3] * R[6] + t[4] * R[7] + t[5] * R[8] - mass * dif[1] * dif[2];
    res[6] = mass * dif[0]; res[7] = mass * dif[1]; res[8] = mass * dif[2]; res[9] = mass;
}

#define SYM(r, c) ((r) >= (c) ? ((r) * ((r) + 1) / 2 + (c)) : ((c) * ((c) + 1) / 2 + (r)))

template <typename T> GO1_DEV void quat_to_mat(T* R, const T* q) {
    T q00 = q[0] * q[0], q01 = q[0] * q[1], q02 = q[0] * q[2], q03 = q[0] * q[3];
    T q11 = q[1] * q[1], q12 = q[1] * q[2], q13 = q[1] * q[3], q22 = q[2] * q[2], q23 = q[2] * q[3], q33 = q[3] * q[3];
    R[0] = q00 + q11 - q22 - q33; R[4] = q00 - q11 + q22 - q33; R[8] = q00 - q11 - q22 + q33;
    R[1] = 2 * (q12 - q03); R[2] = 2 * (q13 + q02); R[3] = 2 * (q12 + q03);
    R[5] = 2 * (q23 - q01); R[6] = 2 * (q13 - q02); R[7] = 2 * (q23 + q01);
}

// getimpedance (engine_core_constraint.c) with pre-clamped solimp
template <typename T> GO1_DEV T impedance(T d0, T d1, T width, T mid, T power, T pos_minus_margin) {
    typedef Num<T> N;
    if (d0 == d1 || width <= N::MINVAL) return T(0.5) * (d0 + d1);
    T x = N::abs_(pos_minus_margin) / width;
    if (x >= 1) return d1;
    if (x <= 0) return d0;
    T y;
    if (power == 1) y = x;
    else if (x <= mid) {
        T a = (power == 2) ? T(1) / mid : T(1) / N::pow_(mid, power - 1);
        y = a * ((power == 2) ? x * x : N::pow_(x, power));
    } else {
        T b = (power == 2) ? T(1) / (1 - mid) : T(1) / N::pow_(1 - mid, power - 1);
        T omx = 1 - x;
        y = 1 - b * ((power == 2) ? omx * omx : N::pow_(omx, power));
    }
    return d0 + y * (d1 - d0);
}

}  // namespace go1
#include "go1_pair.cuh"
namespace go1 {

// ---- per-environment shared-memory workspace (one per warp)
template <typename T> struct WEnv {
    // persistent state row, same order as the global state slab (S_* offsets in go1sim.cu): 78 values + pad
    T qpos[19], qvel[18], warm[18], last_act[12], air[4], dev[3], vdes[3], ep_ret, pad_[2];
    T ctrl[12];
    int ei[8];                       // step, unhealthy_steps, last_contacts, reset_count, bad_state
    T com[4];                        // robot COM relative to the trunk origin
    T comw[4];                       // world COM at the start of the last substep (subtree_com)
    T cdof[NDOF][6];                 // motion axes [ang; lin] about the COM
    T crb[NB][10];                   // composite spatial inertia about the COM (mju_inertCom layout)
    T Kc[NB][21];                    // per-body contact stiffness -> composite -> + crb.  During kinematics / collision
                                     // the same words hold the body frames (xmat 13x9, xpos 13x3; dead before the Newton loop)
    T M[120], H[120];                // block-arrow matrices; H is factored in place
    T bv[2][NB][6];                  // per-body spatial vectors: 0 = cvel, then A (for qacc); 1 = S (search) or body wrenches
    T xv[20];                        // joint-space vector being broadcast
    T cfrc[NB][6];                   // own contact wrench of every body about the COM (cfrc_ext of the last evaluation)
    T c_r[3][CW], c_D0[CW], c_aref[6][CW];
    T c_s[14][CW];                   // scratch: evaluate -> wrench(6) N T U1..U5 zone ; line search -> polynomial(11)
    T c_n[3][PW];                    // normal (geom1 -> geom2) of the robot-robot contacts [nfl, ncon)
    int c_meta[CW];                  // body2 | (body1 + 1) << 4 (0: floor) | class << 8 | geom2 id << 16 | geom1 id << 22
    int bstart[NB + 3];              // floor contacts of body b: [bstart[b], bstart[b+1])
    int ncon, overflow, nfl, pad2_;  // nfl: floor contacts (sorted by body); [nfl, ncon): robot-robot contacts
    // aliases (lifetimes do not overlap, see DESIGN.md "shared-memory budget")
    GO1_DEV T (*xmat())[9] { return reinterpret_cast<T (*)[9]>(&Kc[0][0]); }
    GO1_DEV T (*xpos())[3] { return reinterpret_cast<T (*)[3]>(&Kc[0][0] + NB * 9); }
    GO1_DEV T (*ubuf())[6] { return reinterpret_cast<T (*)[6]>(&c_s[0][0]); }      // assemble_arrow scratch (c_s is dead then)
    GO1_DEV T (*wrench())[6] { return bv[1]; }
    GO1_DEV T* kd() { return &Kc[0][0]; }              // dense 12x12 leg block of the Hessian (environments with leg-leg contacts)
    GO1_DEV T* jbuf() { return &Kc[0][0] + 144; }      // 6 x 12 Jacobian rows of one robot-robot contact
    GO1_DEV T* g6() { return &Kc[0][0] + 216; }        // packed 6x6 block of one coupled pair of legs (21) + its 3x3 coupling block (9)
};

GO1_DEV int meta_b2(int meta) { return meta & 15; }
GO1_DEV int meta_b1p(int meta) { return (meta >> 4) & 15; }     // body1 + 1; 0 = the floor
GO1_DEV int meta_cls(int meta) { return (meta >> 8) & 15; }

// contact frame (rows: normal, two tangents): the floor's constant frame, or mju_makeFrame of the stored pair normal
template <typename T> GO1_DEV void contact_frame(const DevModel<T>& m, const WEnv<T>& E, int c, int dim, T* F) {
    if (c < E.nfl) {
#pragma unroll
        for (int e = 0; e < 9; e++) F[e] = m.plane_frame[e];
    } else {
        const int p = c - E.nfl;
        F[0] = E.c_n[0][p]; F[1] = E.c_n[1][p]; F[2] = E.c_n[2][p];
        if (dim > 1) make_frame_(F);
        else { F[3] = F[4] = F[5] = F[6] = F[7] = F[8] = 0; }
    }
}
// spatial motion of geom2's body relative to geom1's body (the floor does not move)
template <typename T> GO1_DEV void contact_image(const T (*bv)[6], int meta, T* A) {
    const int b2 = meta_b2(meta), b1p = meta_b1p(meta);
#pragma unroll
    for (int e = 0; e < 6; e++) A[e] = bv[b2][e];
    if (b1p) {
#pragma unroll
        for (int e = 0; e < 6; e++) A[e] -= bv[b1p - 1][e];
    }
}



// ---------------------------------------------------------------------------------------------
// Y_b = sum over the dofs on the path to body b of cdof_d * x_d   (x: 18 joint-space values in shared memory)
// lanes 0..23 = (leg k, component e); trunk written by leg 0's lanes.
template <typename T> GO1_NOINL void body_motion(WEnv<T>& E, const T* x, T (*out)[6], int lane) {
    if (lane < 24) {
        const int k = lane / 6, e = lane - 6 * k;
        T t = 0;
#pragma unroll
        for (int j = 0; j < 6; j++) t += E.cdof[j][e] * x[j];
        if (k == 0) out[0][e] = t;
#pragma unroll
        for (int i = 0; i < 3; i++) {
            t += E.cdof[6 + 3 * k + i][e] * x[6 + 3 * k + i];
            out[1 + 3 * k + i][e] = t;
        }
    }
    __syncwarp();
}

// joint-space force of per-body spatial forces W (own, not composite; destroyed): returns this lane's dof value
template <typename T> GO1_NOINL T force_to_joint(WEnv<T>& E, T (*W)[6], int lane) {
    T w1 = 0;
    if (lane < 24) {
        const int k = lane / 6, e = lane - 6 * k;
        const T w3 = W[3 + 3 * k][e];
        const T w2 = W[2 + 3 * k][e] + w3;
        w1 = W[1 + 3 * k][e] + w2;
        W[2 + 3 * k][e] = w2; W[1 + 3 * k][e] = w1;
    }
    // trunk composite: lanes e < 6 add the four hips' composites
    const T h1 = __shfl_sync(FULL, w1, (lane % 6) + 6), h2 = __shfl_sync(FULL, w1, (lane % 6) + 12), h3 = __shfl_sync(FULL, w1, (lane % 6) + 18);
    if (lane < 6) W[0][lane] = ((W[0][lane] + w1) + h1) + (h2 + h3);
    __syncwarp();
    T r = 0;
    if (lane < NDOF) {
        const int b = lane < 6 ? 0 : lane - 5;
#pragma unroll
        for (int e = 0; e < 6; e++) r += E.cdof[lane][e] * W[b][e];
    }
    return r;
}

// composite pass over the 10-parameter spatial inertias (mj_crb accumulation): lanes e < 10
template <typename T> GO1_DEV void composite10(T (*I)[10], int lane) {
    if (lane < 10) {
        T acc = I[0][lane];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const T k3 = I[3 + 3 * k][lane];
            const T k2 = I[2 + 3 * k][lane] + k3;
            const T k1 = I[1 + 3 * k][lane] + k2;
            acc += k1;
            I[2 + 3 * k][lane] = k2; I[1 + 3 * k][lane] = k1;
        }
        I[0][lane] = acc;
    }
    __syncwarp();
}

// K[b] <- (sum of K over the subtree of b, if `accumulate`) + the composite inertia crb[b] expanded to a packed 6x6
// (lanes e < 21; entry e of the 6x6 is sgn * crb[idx], see store order in DevModel::tab)
template <typename T> GO1_DEV void composite21(const DevModel<T>& m, T (*K)[21], const T (*crb)[10], bool accumulate, int lane) {
    if (lane < 21) {
        const int idx = m.tab.crb_idx[lane];
        const T sg = (T)m.tab.crb_sgn[lane];
        T acc = accumulate ? K[0][lane] : T(0);
#pragma unroll
        for (int k = 0; k < 4; k++) {
            T k3 = 0, k2 = 0, k1 = 0;
            if (accumulate) {
                k3 = K[3 + 3 * k][lane];
                k2 = K[2 + 3 * k][lane] + k3;
                k1 = K[1 + 3 * k][lane] + k2;
                acc += k1;
            }
            K[3 + 3 * k][lane] = k3 + sg * crb[3 + 3 * k][idx]; K[2 + 3 * k][lane] = k2 + sg * crb[2 + 3 * k][idx];
            K[1 + 3 * k][lane] = k1 + sg * crb[1 + 3 * k][idx];
        }
        K[0][lane] = acc + sg * crb[0][idx];
    }
    __syncwarp();
}

// block-arrow matrix of the composite 6x6 matrices K:  out[(u,c)] = cdof_c' K[body(u)] cdof_u   (CRBA form)
template <typename T> GO1_NOINL void assemble_arrow(const DevModel<T>& m, WEnv<T>& E, const T (*K)[21], T* out, int lane) {
#pragma unroll 1
    for (int idx = lane; idx < NDOF * 6; idx += 32) {
        const int d = idx / 6, r = idx - 6 * d;
        const int b = d < 6 ? 0 : d - 5;
        T s = 0;
#pragma unroll
        for (int c = 0; c < 6; c++) s += K[b][SYM(r, c)] * E.cdof[d][c];
        E.ubuf()[d][r] = s;
    }
    __syncwarp();
#pragma unroll 1
    for (int idx = lane; idx < NARROW; idx += 32) {
        const int u = m.tab.ent_u[idx], c = m.tab.ent_c[idx];
        T s = 0;
#pragma unroll
        for (int e = 0; e < 6; e++) s += E.cdof[c][e] * E.ubuf()[u][e];
        out[idx] = s;
    }
    __syncwarp();
}

// y = A x for the block-arrow A; x is read from E.xv (must be written + synced by the caller); returns y_lane
template <typename T> GO1_NOINL T arrow_matvec(const WEnv<T>& E, const T* A, int lane) {
    T s = 0;
    if (lane < 6) {
#pragma unroll
        for (int c = 0; c < 6; c++) s += A[SYM(lane, c)] * E.xv[c];
#pragma unroll
        for (int k = 0; k < 4; k++)
#pragma unroll
            for (int i = 0; i < 3; i++) s += A[A_TK + 18 * k + 3 * lane + i] * E.xv[6 + 3 * k + i];
    } else if (lane < NDOF) {
        const int k = (lane - 6) / 3, i = lane - 6 - 3 * k;
#pragma unroll
        for (int j = 0; j < 6; j++) s += A[A_TK + 18 * k + 3 * j + i] * E.xv[j];
#pragma unroll
        for (int c = 0; c < 3; c++) s += A[A_KK + 6 * k + SYM(i, c)] * E.xv[6 + 3 * k + c];
    }
    return s;
}

// dense 6x6 Cholesky of the packed tt block (diagonals inverted), every lane redundantly in registers; lane 0 stores
template <typename T> GO1_DEV void chol6(T* A, int lane) {
    typedef Num<T> N;
    T p[21];
#pragma unroll
    for (int e = 0; e < 21; e++) p[e] = A[e];
    __syncwarp();
#pragma unroll
    for (int j = 0; j < 6; j++) {
        T d = p[SYM(j, j)];
#pragma unroll
        for (int k = 0; k < j; k++) d -= p[SYM(j, k)] * p[SYM(j, k)];
        const T inv = N::rsqrt_(N::max_(d, N::MINVAL));
        p[SYM(j, j)] = inv;
#pragma unroll
        for (int i = j + 1; i < 6; i++) {
            T s = p[SYM(i, j)];
#pragma unroll
            for (int k = 0; k < j; k++) s -= p[SYM(i, k)] * p[SYM(j, k)];
            p[SYM(i, j)] = s * inv;
        }
    }
    if (lane == 0) {
#pragma unroll
        for (int e = 0; e < 21; e++) A[e] = p[e];
    }
    __syncwarp();
}
// x <- (L L')^-1 x with the factor of chol6
template <typename T> GO1_DEV void solve6(const T* F, T* x) {
#pragma unroll
    for (int i = 0; i < 6; i++) {
        T s = x[i];
#pragma unroll
        for (int k = 0; k < i; k++) s -= F[SYM(i, k)] * x[k];
        x[i] = s * F[SYM(i, i)];
    }
#pragma unroll
    for (int i = 5; i >= 0; i--) {
        T s = x[i];
#pragma unroll
        for (int k = i + 1; k < 6; k++) s -= F[SYM(k, i)] * x[k];
        x[i] = s * F[SYM(i, i)];
    }
}

// in-place Cholesky of the block-arrow A: kk[k] -> L_k (diagonals inverted), tk[k] -> Y_k = C_k L_k^-T,
// tt -> Cholesky factor of the Schur complement (diagonals inverted)
template <typename T> GO1_NOINL void arrow_factor(const DevModel<T>& m, T* A, int lane) {
    typedef Num<T> N;
    if (lane < 4) {
        T* a = A + A_KK + 6 * lane;
        const T i00 = N::rsqrt_(N::max_(a[0], N::MINVAL));
        const T l10 = a[1] * i00, l20 = a[3] * i00;
        const T i11 = N::rsqrt_(N::max_(a[2] - l10 * l10, N::MINVAL));
        const T l21 = (a[4] - l20 * l10) * i11;
        const T i22 = N::rsqrt_(N::max_(a[5] - l20 * l20 - l21 * l21, N::MINVAL));
        a[0] = i00; a[1] = l10; a[2] = i11; a[3] = l20; a[4] = l21; a[5] = i22;
    }
    __syncwarp();
    if (lane < 24) {
        const int k = lane / 6, j = lane - 6 * k;
        const T* l = A + A_KK + 6 * k;
        T* y = A + A_TK + 18 * k + 3 * j;
        const T y0 = y[0] * l[0];
        const T y1 = (y[1] - y0 * l[1]) * l[2];
        const T y2 = (y[2] - y0 * l[3] - y1 * l[4]) * l[5];
        y[0] = y0; y[1] = y1; y[2] = y2;
    }
    __syncwarp();
    if (lane < 21) {
        const int r = m.tab.sym_r[lane], c = m.tab.sym_c[lane];
        T s = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const T* yr = A + A_TK + 18 * k + 3 * r; const T* yc = A + A_TK + 18 * k + 3 * c;
            s += yr[0] * yc[0] + yr[1] * yc[1] + yr[2] * yc[2];
        }
        A[lane] -= s;
    }
    __syncwarp();
    chol6(A, lane);
}

// solves A x = g with the factor F (arrow_factor); g/x are per-dof-lane register values
template <typename T> GO1_NOINL T arrow_solve(WEnv<T>& E, const T* F, T g, int lane) {
    if (lane < NDOF) E.xv[lane] = g;
    __syncwarp();
    if (lane < 4) {
        const T* l = F + A_KK + 6 * lane; T* z = E.xv + 6 + 3 * lane;
        const T z0 = z[0] * l[0];
        const T z1 = (z[1] - l[1] * z0) * l[2];
        const T z2 = (z[2] - l[3] * z0 - l[4] * z1) * l[5];
        z[0] = z0; z[1] = z1; z[2] = z2;
    }
    __syncwarp();
    if (lane < 6) {
        T s = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const T* y = F + A_TK + 18 * k + 3 * lane; const T* z = E.xv + 6 + 3 * k;
            s += y[0] * z[0] + y[1] * z[1] + y[2] * z[2];
        }
        E.xv[lane] -= s;
    }
    __syncwarp();
    T x[6];
#pragma unroll
    for (int i = 0; i < 6; i++) x[i] = E.xv[i];
    solve6(F, x);
    __syncwarp();
    if (lane >= 6 && lane < NDOF) {
        const int k = (lane - 6) / 3, i = lane - 6 - 3 * k;
        T z = E.xv[lane];
#pragma unroll
        for (int j = 0; j < 6; j++) z -= F[A_TK + 18 * k + 3 * j + i] * x[j];
        E.xv[lane] = z;
    }
    if (lane == 0) {
#pragma unroll
        for (int i = 0; i < 6; i++) E.xv[i] = x[i];
    }
    __syncwarp();
    if (lane < 4) {
        const T* l = F + A_KK + 6 * lane; T* z = E.xv + 6 + 3 * lane;
        const T w2 = z[2] * l[5];
        const T w1 = (z[1] - l[4] * w2) * l[2];
        const T w0 = (z[0] - l[1] * w1 - l[3] * w2) * l[0];
        z[0] = w0; z[1] = w1; z[2] = w2;
    }
    __syncwarp();
    const T r = lane < NDOF ? E.xv[lane] : T(0);
    __syncwarp();
    return r;
}


// ---- exactly one pair of legs la < lb coupled by leg-leg contacts: their 6 dofs form one dense block G6 (packed, E.g6(), built
// by pair_hessian), the other two legs keep their 3x3 blocks.  Same elimination as arrow_factor, with a 6x6 Cholesky for the
// pair; the rows of Y that belong to the pair go back into the tk slots of the two legs, so the Schur sum is unchanged.
template <typename T> GO1_NOINL void pairleg_factor(WEnv<T>& E, const DevModel<T>& m, T* A, int la, int lb, int lane) {
    typedef Num<T> N;
    T* G = E.g6();
    if (lane < 4 && lane != la && lane != lb) {
        T* a = A + A_KK + 6 * lane;
        const T i00 = N::rsqrt_(N::max_(a[0], N::MINVAL));
        const T l10 = a[1] * i00, l20 = a[3] * i00;
        const T i11 = N::rsqrt_(N::max_(a[2] - l10 * l10, N::MINVAL));
        const T l21 = (a[4] - l20 * l10) * i11;
        const T i22 = N::rsqrt_(N::max_(a[5] - l20 * l20 - l21 * l21, N::MINVAL));
        a[0] = i00; a[1] = l10; a[2] = i11; a[3] = l20; a[4] = l21; a[5] = i22;
    }
    chol6(G, lane);
    if (lane < 24) {
        const int k = lane / 6, j = lane - 6 * k;
        if (k != la && k != lb) {
            const T* l = A + A_KK + 6 * k;
            T* y = A + A_TK + 18 * k + 3 * j;
            const T y0 = y[0] * l[0];
            const T y1 = (y[1] - y0 * l[1]) * l[2];
            const T y2 = (y[2] - y0 * l[3] - y1 * l[4]) * l[5];
            y[0] = y0; y[1] = y1; y[2] = y2;
        }
    } else if (lane < 30) {          // trunk row lane - 24 of the pair's coupling [C_la C_lb] L6^-T
        const int t = lane - 24;
        T* ya = A + A_TK + 18 * la + 3 * t; T* yb = A + A_TK + 18 * lb + 3 * t;
        T y[6] = {ya[0], ya[1], ya[2], yb[0], yb[1], yb[2]};
#pragma unroll
        for (int c = 0; c < 6; c++) {
            T v = y[c];
#pragma unroll
            for (int p = 0; p < c; p++) v -= y[p] * G[SYM(c, p)];
            y[c] = v * G[SYM(c, c)];
        }
        ya[0] = y[0]; ya[1] = y[1]; ya[2] = y[2]; yb[0] = y[3]; yb[1] = y[4]; yb[2] = y[5];
    }
    __syncwarp();
    if (lane < 21) {
        const int r = m.tab.sym_r[lane], c = m.tab.sym_c[lane];
        T s = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const T* yr = A + A_TK + 18 * k + 3 * r; const T* yc = A + A_TK + 18 * k + 3 * c;
            s += yr[0] * yc[0] + yr[1] * yc[1] + yr[2] * yc[2];
        }
        A[lane] -= s;
    }
    __syncwarp();
    chol6(A, lane);
}

template <typename T> GO1_NOINL T pairleg_solve(WEnv<T>& E, const T* F, T g, int la, int lb, int lane) {
    const T* G = E.g6();
    if (lane < NDOF) E.xv[lane] = g;
    __syncwarp();
    if (lane < 4 && lane != la && lane != lb) {
        const T* l = F + A_KK + 6 * lane; T* z = E.xv + 6 + 3 * lane;
        const T z0 = z[0] * l[0];
        const T z1 = (z[1] - l[1] * z0) * l[2];
        const T z2 = (z[2] - l[3] * z0 - l[4] * z1) * l[5];
        z[0] = z0; z[1] = z1; z[2] = z2;
    } else if (lane == 4) {
        T* za = E.xv + 6 + 3 * la; T* zb = E.xv + 6 + 3 * lb;
        T z[6] = {za[0], za[1], za[2], zb[0], zb[1], zb[2]};
#pragma unroll
        for (int i = 0; i < 6; i++) {
            T s = z[i];
#pragma unroll
            for (int k = 0; k < i; k++) s -= G[SYM(i, k)] * z[k];
            z[i] = s * G[SYM(i, i)];
        }
        za[0] = z[0]; za[1] = z[1]; za[2] = z[2]; zb[0] = z[3]; zb[1] = z[4]; zb[2] = z[5];
    }
    __syncwarp();
    if (lane < 6) {
        T s = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const T* y = F + A_TK + 18 * k + 3 * lane; const T* z = E.xv + 6 + 3 * k;
            s += y[0] * z[0] + y[1] * z[1] + y[2] * z[2];
        }
        E.xv[lane] -= s;
    }
    __syncwarp();
    T x[6];
#pragma unroll
    for (int i = 0; i < 6; i++) x[i] = E.xv[i];
    solve6(F, x);
    __syncwarp();
    if (lane >= 6 && lane < NDOF) {
        const int k = (lane - 6) / 3, i = lane - 6 - 3 * k;
        T z = E.xv[lane];
#pragma unroll
        for (int j = 0; j < 6; j++) z -= F[A_TK + 18 * k + 3 * j + i] * x[j];
        E.xv[lane] = z;
    }
    if (lane == 0) {
#pragma unroll
        for (int i = 0; i < 6; i++) E.xv[i] = x[i];
    }
    __syncwarp();
    if (lane < 4 && lane != la && lane != lb) {
        const T* l = F + A_KK + 6 * lane; T* z = E.xv + 6 + 3 * lane;
        const T w2 = z[2] * l[5];
        const T w1 = (z[1] - l[4] * w2) * l[2];
        const T w0 = (z[0] - l[1] * w1 - l[3] * w2) * l[0];
        z[0] = w0; z[1] = w1; z[2] = w2;
    } else if (lane == 4) {
        T* za = E.xv + 6 + 3 * la; T* zb = E.xv + 6 + 3 * lb;
        T z[6] = {za[0], za[1], za[2], zb[0], zb[1], zb[2]};
#pragma unroll
        for (int i = 5; i >= 0; i--) {
            T s = z[i];
#pragma unroll
            for (int k = i + 1; k < 6; k++) s -= G[SYM(k, i)] * z[k];
            z[i] = s * G[SYM(i, i)];
        }
        za[0] = z[0]; za[1] = z[1]; za[2] = z[2]; zb[0] = z[3]; zb[1] = z[4]; zb[2] = z[5];
    }
    __syncwarp();
    const T r = lane < NDOF ? E.xv[lane] : T(0);
    __syncwarp();
    return r;
}

// ---- environments with a leg-leg contact: the 12 x 12 leg block KD = E.kd() is dense (pair_hessian), the trunk block tt and the
// couplings tk stay in the arrow storage A.  Same elimination order as arrow_factor: KD = L L' (in place, diagonals inverted),
// tk -> Y = C L^-T, tt -> Cholesky factor of the Schur complement tt - Y Y'.
template <typename T> GO1_NOINL void dense_factor(WEnv<T>& E, const DevModel<T>& m, T* A, int lane) {
    typedef Num<T> N;
    T* KD = E.kd();
#pragma unroll 1
    for (int c = 0; c < 12; c++) {
        const T inv = N::rsqrt_(N::max_(KD[c * 12 + c], N::MINVAL));
        T l = 0;
        if (lane > c && lane < 12) l = KD[lane * 12 + c] * inv;
        __syncwarp();
        if (lane > c && lane < 12) KD[lane * 12 + c] = l;
        if (lane == c) KD[c * 12 + c] = inv;
        __syncwarp();
        if (lane > c && lane < 12)
            for (int s2 = c + 1; s2 <= lane; s2++) KD[lane * 12 + s2] -= l * KD[s2 * 12 + c];
        __syncwarp();
    }
    if (lane < 6) {      // row `lane` of Y: forward substitution L y' = c'
        T y[12];
#pragma unroll
        for (int c = 0; c < 12; c++) {
            T v = A[A_TK + 18 * (c / 3) + 3 * lane + (c % 3)];
#pragma unroll
            for (int p = 0; p < c; p++) v -= y[p] * KD[c * 12 + p];
            y[c] = v * KD[c * 12 + c];
        }
#pragma unroll
        for (int c = 0; c < 12; c++) A[A_TK + 18 * (c / 3) + 3 * lane + (c % 3)] = y[c];
    }
    __syncwarp();
    if (lane < 21) {
        const int r = m.tab.sym_r[lane], c = m.tab.sym_c[lane];
        T s = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const T* yr = A + A_TK + 18 * k + 3 * r; const T* yc = A + A_TK + 18 * k + 3 * c;
            s += yr[0] * yc[0] + yr[1] * yc[1] + yr[2] * yc[2];
        }
        A[lane] -= s;
    }
    __syncwarp();
    chol6(A, lane);
}

// solves H x = g with the factor of dense_factor; g / x are per-dof-lane register values
template <typename T> GO1_NOINL T dense_solve(WEnv<T>& E, const T* F, T g, int lane) {
    const T* KD = E.kd();
    const int r = lane - 6;
    const bool leg = lane >= 6 && lane < NDOF;
    T x = g;
#pragma unroll 1
    for (int c = 0; c < 12; c++) {           // z = L^-1 g_legs
        const T zc = __shfl_sync(FULL, x, 6 + c) * KD[c * 12 + c];
        if (leg && r == c) x = zc;
        else if (leg && r > c) x -= KD[r * 12 + c] * zc;
    }
    if (lane < NDOF) E.xv[lane] = x;
    __syncwarp();
    if (lane < 6) {
        T s = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const T* y = F + A_TK + 18 * k + 3 * lane; const T* z = E.xv + 6 + 3 * k;
            s += y[0] * z[0] + y[1] * z[1] + y[2] * z[2];
        }
        E.xv[lane] -= s;
    }
    __syncwarp();
    T x6[6];
#pragma unroll
    for (int i = 0; i < 6; i++) x6[i] = E.xv[i];
    solve6(F, x6);
    if (leg) {
        const int k = r / 3, i = r - 3 * k;
#pragma unroll
        for (int j = 0; j < 6; j++) x -= F[A_TK + 18 * k + 3 * j + i] * x6[j];
    }
#pragma unroll 1
    for (int c = 11; c >= 0; c--) {          // w = L^-T z
        const T wc = __shfl_sync(FULL, x, 6 + c) * KD[c * 12 + c];
        if (leg && r == c) x = wc;
        else if (leg && r < c) x -= KD[c * 12 + r] * wc;
    }
    __syncwarp();
    T res = 0;
    if (lane < 6) res = lane == 0 ? x6[0] : (lane == 1 ? x6[1] : (lane == 2 ? x6[2] : (lane == 3 ? x6[3] : (lane == 4 ? x6[4] : x6[5]))));
    else if (leg) res = x;
    return res;
}


// ---------------------------------------------------------------------------------------------
template <typename T>
struct EnvCfg {
    int frame_skip, max_episode_steps, max_unhealthy_steps, solver_iterations;
    T rand_power, fallen_prob, vdes_min[3], vdes_max[3], reward_w[6], cost_w[11];
    T healthy_z[2], healthy_roll[2], healthy_pitch[2], sigma, obs_clip, tolerance;
    T dt;                      // control period = model timestep x frame_skip (MujocoEnv.dt; air-time bookkeeping, go1_mujoco_env.py:656)
    // bipedal mode (go1_mujoco_env.py:43, :91-101, :967-972); read only when biped != 0
    int biped;
    T biped_w[10], biped_reset_z, biped_reset_qw, biped_reset_qy, foot_site[2][3];
    const T* biped_ready;      // device array [12] (indexed by lane: kept out of the kernel parameters, which must not be indexed dynamically)
    const unsigned char* gp_tab;   // device: candidate geom pairs of the robot-robot collision, a[NGP] then b[NGP] (DevModel::gp_a / gp_b)
};

// kinematic quantities the biped terms read from mjData (world frame, start of the last substep), one row per env
enum { BK_HIP_Z = 0 /* FR FL RR RL */, BK_REAR_HIP_Y = 4 /* RR RL */, BK_FRONT_CALF_Z = 6, BK_FRONT_FOOT_Z = 8, BK_TRUNK_XZ = 10, BK_N = 12 };

// what mjData holds after the last substep, as the env reads it (per-lane registers + E.cfrc)
template <typename T> struct SubOut {
    T qacc, qfa, qfc;          // values of this lane's dof (lanes < 18)
    T qacc_prev, qacc_prev2;   // qacc of the two substeps before the last one (warm-start extrapolation)
    T trunk_z;
    int ncon, iters, iters_total, overflow, bad;
    int nsub;                  // substeps completed in this control step
#if defined(GO1_PHASE_TIMING) && !defined(GO1_EMU)
    long long tlast, twork[7], twait[7];
    int maxmode, rounds;
#endif
};

template <typename T> struct Stage4 { T px[4], py[4], pz[4], d[4]; int n; };
template <typename T> GO1_DEV void stage_emit(Stage4<T>& st, const T* pos, T dist) {
#pragma unroll
    for (int j = 0; j < 4; j++)
        if (st.n == j) { st.px[j] = pos[0]; st.py[j] = pos[1]; st.pz[j] = pos[2]; st.d[j] = dist; }
    st.n++;
}

// plane-vs-geom narrowphase for one geom (engine_collision_primitive.c rules), positions relative to the
// trunk origin; dc = signed distance of the geom centre to the plane
template <typename T>
GO1_DEV void collide_geom(const DevModel<T>& m, const DevGeom<T>& g, const T* Rb, const T* gp, T dc, Stage4<T>& st) {
    typedef Num<T> N;
    const T* n = m.plane_frame;
    const DevClass<T>& cl = m.cls[g.cls];
    const T inc = cl.incmargin, margin = cl.margin;
    T ax[3];
    matvec3(ax, Rb, g.zax);
    if (g.type == G_SPHERE) {
        const T dist = dc - g.size[0];
        if (dist < inc) {
            T pos[3];
#pragma unroll
            for (int k = 0; k < 3; k++) pos[k] = gp[k] - n[k] * (g.size[0] + T(0.5) * dist);
            stage_emit(st, pos, dist);
        }
    } else if (g.type == G_CAPSULE) {
        const T pa = dot3(n, ax) * g.size[1];
#pragma unroll
        for (int s = 0; s < 2; s++) {
            const T sg = s == 0 ? T(1) : T(-1);
            const T dist = dc + sg * pa - g.size[0];
            if (dist < inc) {
                T pos[3];
#pragma unroll
                for (int k = 0; k < 3; k++) pos[k] = gp[k] + sg * ax[k] * g.size[1] - n[k] * (g.size[0] + T(0.5) * dist);
                stage_emit(st, pos, dist);
            }
        }
    } else if (g.type == G_CYLINDER) {
        T prjaxis = dot3(n, ax);
        if (prjaxis > 0) { ax[0] = -ax[0]; ax[1] = -ax[1]; ax[2] = -ax[2]; prjaxis = -prjaxis; }
        T vec[3];
#pragma unroll
        for (int k = 0; k < 3; k++) vec[k] = ax[k] * prjaxis - n[k];
        const T len_sqr = dot3(vec, vec);
        if (len_sqr >= N::MINVAL * N::MINVAL) {
            const T s = g.size[0] / N::sqrt_(len_sqr);
            vec[0] *= s; vec[1] *= s; vec[2] *= s;
        } else {
            T xw[3];
            matvec3(xw, Rb, g.xax);
            vec[0] = xw[0] * g.size[0]; vec[1] = xw[1] * g.size[0]; vec[2] = xw[2] * g.size[0];
        }
        const T prjvec = dot3(vec, n);
        ax[0] *= g.size[1]; ax[1] *= g.size[1]; ax[2] *= g.size[1];
        prjaxis *= g.size[1];
        const T dA = dc + prjaxis + prjvec;
        if (dA <= margin) {
            if (dA < inc) {
                T pos[3];
#pragma unroll
                for (int k = 0; k < 3; k++) pos[k] = gp[k] + vec[k] + ax[k] - n[k] * dA * T(0.5);
                stage_emit(st, pos, dA);
            }
            const T dB = dc - prjaxis + prjvec;
            if (dB <= margin && dB < inc) {
                T pos[3];
#pragma unroll
                for (int k = 0; k < 3; k++) pos[k] = gp[k] + vec[k] - ax[k] - n[k] * dB * T(0.5);
                stage_emit(st, pos, dB);
            }
            const T dT = dc + prjaxis - prjvec * T(0.5);
            if (dT <= margin && dT < inc) {
                T v1[3];
                cross3(v1, vec, ax);
                const T nn = N::sqrt_(dot3(v1, v1));
                if (nn < N::MINVAL) { v1[0] = 1; v1[1] = 0; v1[2] = 0; } else { v1[0] /= nn; v1[1] /= nn; v1[2] /= nn; }
                const T s = g.size[0] * T(0.86602540378443864676);
                v1[0] *= s; v1[1] *= s; v1[2] *= s;
#pragma unroll
                for (int sgi = 0; sgi < 2; sgi++) {
                    const T sg = sgi == 0 ? T(1) : T(-1);
                    T pos[3];
#pragma unroll
                    for (int k = 0; k < 3; k++) pos[k] = gp[k] + sg * v1[k] + ax[k] - vec[k] * T(0.5) - n[k] * dT * T(0.5);
                    stage_emit(st, pos, dT);
                }
            }
        }
    } else {  // box
        T xw[3], yl[3], yw[3];
        matvec3(xw, Rb, g.xax);
        cross3(yl, g.zax, g.xax);
        matvec3(yw, Rb, yl);
        int cnt = 0;
#pragma unroll 1
        for (int i = 0; i < 8 && cnt < 4; i++) {
            const T sx = (i & 1) ? g.size[0] : -g.size[0], sy = (i & 2) ? g.size[1] : -g.size[1], sz = (i & 4) ? g.size[2] : -g.size[2];
            T corner[3];
#pragma unroll
            for (int k = 0; k < 3; k++) corner[k] = xw[k] * sx + yw[k] * sy + ax[k] * sz;
            const T ld = dot3(n, corner);
            if (dc + ld > margin || ld > 0) continue;
            cnt++;
            const T dist = dc + ld;
            if (dist < inc) {
                T pos[3];
#pragma unroll
                for (int k = 0; k < 3; k++) pos[k] = corner[k] + gp[k] - n[k] * dist * T(0.5);
                stage_emit(st, pos, dist);
            }
        }
    }
}


// per-dof-lane constraint rows (friction loss + joint limit of this lane's leg dof)
template <typename T> struct DofRows { T fr_aref, lim_s, lim_D, lim_aref, floss, Rf; };

// Evaluates all constraint rows at the acceleration whose body-space image is E.bv[0] (A) and whose
// dof value on this lane is `a`.  Contacts: lane c (+32..).  Returns the lane-partial cost; qfc = J'f for this
// lane's dof, hdiag = diagonal Hessian term of this lane's friction/limit rows.  Per contact it leaves the
// world wrench and the cone state (N, T, U1..5, zone) in E.c_s for the Hessian pass; E.cfrc gets the per-body sums.
template <typename T>
GO1_DEV T evaluate_rows(const DevModel<T>& m, WEnv<T>& E, int lane, T a, const DofRows<T>& dr, T& qfc, T& hdiag, T* frc_out) {
    typedef Num<T> N;
    T cost = 0;
    const int ncon = E.ncon;
#pragma unroll 1
    for (int c = lane; c < ncon; c += 32) {
        const int meta = E.c_meta[c];
        const DevClass<T>& cl = m.cls[meta_cls(meta)];
        const int dim = cl.dim;
        const T r[3] = {E.c_r[0][c], E.c_r[1][c], E.c_r[2][c]};
        T A[6], F[9];
        contact_image(E.bv[0], meta, A);
        contact_frame(m, E, c, dim, F);
        T g[3][3];
#pragma unroll
        for (int j = 0; j < 3; j++) cross3(g[j], r, F + 3 * j);
        T jar[6], D[6];
        const T D0 = E.c_D0[c];
#pragma unroll
        for (int j = 0; j < 3; j++) {
            jar[j] = dot3(g[j], A) + dot3(F + 3 * j, A + 3) - E.c_aref[j][c];
            D[j] = D0 * cl.drat[j];
        }
        if (dim == 1) { jar[1] = 0; jar[2] = 0; }       // frictionless robot-robot contact: the normal row only (drat[1..] = 0)
        if (dim == 6) {
#pragma unroll
            for (int j = 0; j < 3; j++) { jar[3 + j] = dot3(F + 3 * j, A) - E.c_aref[3 + j][c]; D[3 + j] = D0 * cl.drat[3 + j]; }
        } else {
#pragma unroll
            for (int j = 3; j < 6; j++) { jar[j] = 0; D[j] = 0; }
        }
        const T mu = cl.mu;
        T U[6];
        U[0] = jar[0] * mu;
        T Tsq = 0;
#pragma unroll
        for (int j = 1; j < 6; j++) { U[j] = jar[j] * cl.fri[j - 1]; Tsq += U[j] * U[j]; }   // rows >= dim have jar = 0
        const T Nn = U[0], Tt = N::sqrt_(Tsq);
        T f[6] = {0, 0, 0, 0, 0, 0};
        T zone = 0;
        if ((Nn >= mu * Tt) || (Tt <= 0 && Nn >= 0)) {
            // top zone: no force
        } else if ((mu * Nn + Tt <= 0) || (Tt <= 0 && Nn < 0)) {
            zone = 1;   // bottom zone: quadratic in every row
#pragma unroll
            for (int j = 0; j < 6; j++) { f[j] = -D[j] * jar[j]; cost += T(0.5) * D[j] * jar[j] * jar[j]; }
        } else {
            zone = 2;   // middle zone (cone surface)
            const T Dm = D0 / N::max_(mu * mu * (1 + mu * mu), N::MINVAL);
            const T NT = Nn - mu * Tt;
            cost += T(0.5) * Dm * NT * NT;
            f[0] = -Dm * NT * mu;
            const T fs = -f[0] / Tt;
#pragma unroll
            for (int j = 1; j < 6; j++) f[j] = fs * U[j] * cl.fri[j - 1];
        }
        // world wrench about the COM: force Fw = sum f_j F_j, torque = r x Fw + sum f_{3+j} F_j
        T Fw[3], Tw[3];
#pragma unroll
        for (int e = 0; e < 3; e++) {
            Fw[e] = f[0] * F[e] + f[1] * F[3 + e] + f[2] * F[6 + e];
            Tw[e] = f[3] * F[e] + f[4] * F[3 + e] + f[5] * F[6 + e];
        }
        T rxF[3];
        cross3(rxF, r, Fw);
#pragma unroll
        for (int e = 0; e < 3; e++) { E.c_s[e][c] = Tw[e] + rxF[e]; E.c_s[3 + e][c] = Fw[e]; }
        E.c_s[6][c] = Nn; E.c_s[7][c] = Tt;
#pragma unroll
        for (int j = 1; j < 6; j++) E.c_s[7 + j][c] = U[j];
        E.c_s[13][c] = zone;
        if (frc_out) {
#pragma unroll
            for (int j = 0; j < 6; j++) frc_out[1 + c * 12 + 6 + j] = f[j];
        }
    }
    // friction-loss (Huber) and joint-limit rows of this lane's leg dof
    T fd = 0;
    hdiag = 0;
    if (lane >= 6 && lane < NDOF) {
        const T R = dr.Rf, fl = dr.floss;
        if (fl > 0) {
            const T jar = a - dr.fr_aref;
            const T Rf = R * fl;
            if (jar <= -Rf) { fd += fl; cost += -fl * (T(0.5) * Rf + jar); }
            else if (jar >= Rf) { fd += -fl; cost += -fl * (T(0.5) * Rf - jar); }
            else { const T D = T(1) / R; fd += -D * jar; cost += T(0.5) * D * jar * jar; hdiag += D; }
        }
        if (dr.lim_s != 0) {
            const T jar = dr.lim_s * a - dr.lim_aref;
            if (jar < 0) { const T D = dr.lim_D; fd += dr.lim_s * (-D * jar); cost += T(0.5) * D * jar * jar; hdiag += D; }
        }
    }
    __syncwarp();
    // per-body sums of the contact wrenches (floor contacts are sorted by body; a robot-robot contact pushes geom2's body
    // with +w and geom1's body with -w, mj_rnePostConstraint)
    const int nfl = E.nfl;
#pragma unroll 1
    for (int idx = lane; idx < NB * 6; idx += 32) {
        const int b = idx / 6, e = idx - 6 * b;
        T s = 0;
        for (int c = E.bstart[b]; c < E.bstart[b + 1]; c++) s += E.c_s[e][c];
        for (int c = nfl; c < ncon; c++) {
            const int meta = E.c_meta[c];
            if (meta_b2(meta) == b) s += E.c_s[e][c];
            if (meta_b1p(meta) == b + 1) s -= E.c_s[e][c];
        }
        E.wrench()[b][e] = s; E.cfrc[b][e] = s;
    }
    __syncwarp();
    qfc = force_to_joint(E, E.wrench(), lane) + fd;
    return cost;
}

// One entry gx' Hc gy of a contact's Hessian block (gx / gy: the 6 constraint-row values of two columns).  Hc = diag(D) in
// the bottom zone; in the middle zone the cone Hessian Dm*S*Hu*S (mj_constraintUpdate) is diagonal + rank 2:
// sum_r w_r g_r g_r' + al (p q' + q p') + be q q'.
template <typename T>
GO1_DEV T cone_quad(const DevClass<T>& cl, const WEnv<T>& E, int c, int zone, T D0, const T* gx, const T* gy) {
    typedef Num<T> N;
    T val = 0;
    if (zone == 1) {
#pragma unroll
        for (int j = 0; j < 6; j++) val += D0 * cl.drat[j] * gx[j] * gy[j];     // rows >= dim: g = 0
    } else {
        const T mu = cl.mu;
        const T Nn = E.c_s[6][c], Tt = E.c_s[7][c];
        const T Dm = D0 / N::max_(mu * mu * (1 + mu * mu), N::MINVAL);
        const T a1 = mu * Nn / (Tt * Tt * Tt), a2 = mu * mu - mu * Nn / Tt;
        T qx = 0, qy = 0;
        val = mu * mu * gx[0] * gy[0];
#pragma unroll
        for (int j = 1; j < 6; j++) {
            const T sj = cl.fri[j - 1], Uj = E.c_s[7 + j][c];
            qx += sj * Uj * gx[j]; qy += sj * Uj * gy[j];
            val += a2 * sj * sj * gx[j] * gy[j];
        }
        val += (-mu * mu / Tt) * (gx[0] * qy + qx * gy[0]) + a1 * qx * qy;
        val *= Dm;
    }
    return val;
}

// per-body contact stiffness K_b += G' Hc G for every FLOOR contact in the bottom / middle zone (lanes e < 21 own one
// entry of the packed symmetric 6x6).  Robot-robot contacts couple two chains and are added by pair_hessian.
template <typename T>
GO1_DEV bool contact_stiffness(const DevModel<T>& m, WEnv<T>& E, int lane) {
    const int ncon = E.nfl;
    bool any = false;
    for (int c = lane; c < ncon; c += 32) any |= E.c_s[13][c] != 0;
    if (!__any_sync(FULL, any)) return false;
#pragma unroll 1
    for (int idx = lane; idx < NB * 21; idx += 32) (&E.Kc[0][0])[idx] = 0;
    __syncwarp();
    const T* F = m.plane_frame;
    const int el = lane < 21 ? lane : 0;
    const int x = m.tab.sym_r[el], y = m.tab.sym_c[el];
#pragma unroll 1
    for (int c = 0; c < ncon; c++) {
        const int zone = (int)E.c_s[13][c];
        if (zone == 0) continue;
        const int meta = E.c_meta[c];
        const int b = meta_b2(meta);
        const DevClass<T>& cl = m.cls[meta_cls(meta)];
        const int dim = cl.dim;
        const T D0 = E.c_D0[c];
        const T r[3] = {E.c_r[0][c], E.c_r[1][c], E.c_r[2][c]};
        // G rows restricted to this lane's two indices: translational rows j: (r x F_j ; F_j), rotational: (F_j ; 0)
        T gx[6], gy[6];
#pragma unroll
        for (int j = 0; j < 3; j++) {
            T cr[3];
            cross3(cr, r, F + 3 * j);
            gx[j] = x < 3 ? (x == 0 ? cr[0] : (x == 1 ? cr[1] : cr[2])) : F[3 * j + x - 3];
            gy[j] = y < 3 ? (y == 0 ? cr[0] : (y == 1 ? cr[1] : cr[2])) : F[3 * j + y - 3];
            gx[3 + j] = (dim == 6 && x < 3) ? F[3 * j + x] : T(0);
            gy[3 + j] = (dim == 6 && y < 3) ? F[3 * j + y] : T(0);
        }
        const T val = cone_quad(cl, E, c, zone, D0, gx, gy);
        if (lane < 21) E.Kc[b][lane] += val;
    }
    __syncwarp();
    return true;
}

// Jacobian rows of robot-robot contact c restricted to the 12 leg dofs: JB[j * 12 + u] = J_j[6 + u] (lanes u < 12).  Both bodies ride
// on the trunk, so J = G (S_body2 - S_body1) is zero on the six trunk dofs.
template <typename T>
GO1_DEV void pair_jacobian(const DevModel<T>& m, WEnv<T>& E, int c, int meta, int dim, T* JB, int lane) {
    if (lane < 12) {
        const int b2 = meta_b2(meta), b1 = meta_b1p(meta) - 1;
        const int k = lane / 3, i = lane - 3 * k;          // dof 6 + lane moves body b iff b is link >= i of leg k
        const bool m2 = b2 >= 1 && (b2 - 1) / 3 == k && (b2 - 1) - 3 * k >= i;
        const bool m1 = b1 >= 1 && (b1 - 1) / 3 == k && (b1 - 1) - 3 * k >= i;
        const T sgn = (m2 ? T(1) : T(0)) - (m1 ? T(1) : T(0));
        T F[9], cd[6];
        contact_frame(m, E, c, dim, F);
        const T r[3] = {E.c_r[0][c], E.c_r[1][c], E.c_r[2][c]};
#pragma unroll
        for (int e = 0; e < 6; e++) cd[e] = E.cdof[6 + lane][e];
#pragma unroll
        for (int j = 0; j < 3; j++) {
            T cr[3];
            cross3(cr, r, F + 3 * j);
            JB[j * 12 + lane] = j < dim ? sgn * (dot3(cr, cd) + dot3(F + 3 * j, cd + 3)) : T(0);
            JB[(3 + j) * 12 + lane] = dim == 6 ? sgn * dot3(F + 3 * j, cd) : T(0);
        }
    }
    __syncwarp();
}

// Hessian terms J' Hc J of the robot-robot contacts in the bottom / middle zone; they live in the 12 x 12 leg block.  E.H must
// hold the arrow.  Returns
//   0  no such contact is active
//   1  only contacts inside one leg or between a leg and the trunk: added to the legs' 3x3 blocks of E.H (still block-arrow)
//   2 | (4 la + lb) << 2   exactly one pair of legs la < lb is coupled: their 6 dofs form one dense 6x6 block, packed in E.g6()
//      (pairleg_factor / pairleg_solve); the 3x3 blocks of E.H hold the diagonal parts
//   3  several leg pairs are coupled: complete dense 12 x 12 leg block in E.kd() (dense_factor / dense_solve)
template <typename T>
GO1_NOINL int pair_hessian(const DevModel<T>& m, WEnv<T>& E, int lane) {
    const int nfl = E.nfl, ncon = E.ncon;
    bool any = false;
    int code = -1;                                     // 4 * min(leg) + max(leg) of an active leg-leg contact (PW <= 32: one per lane)
    for (int c = nfl + lane; c < ncon; c += 32) {
        if (E.c_s[13][c] != 0) {
            any = true;
            const int meta = E.c_meta[c];
            const int b2 = meta_b2(meta), b1 = meta_b1p(meta) - 1;
            if (b1 >= 1 && b2 >= 1) {
                const int l1 = (b1 - 1) / 3, l2 = (b2 - 1) / 3;
                if (l1 != l2) code = l1 < l2 ? 4 * l1 + l2 : 4 * l2 + l1;
            }
        }
    }
    if (!__any_sync(FULL, any)) return 0;
    const unsigned cm = __ballot_sync(FULL, code >= 0);
    int first = -1;
    bool multi = false;
    if (cm) {
        first = __shfl_sync(FULL, code, __ffs((int)cm) - 1);
        multi = __any_sync(FULL, code >= 0 && code != first);
    }
    T* JB = E.jbuf();
    if (multi) {
        T* KD = E.kd();
#pragma unroll 1
        for (int idx = lane; idx < 144; idx += 32) KD[idx] = 0;
        __syncwarp();
#pragma unroll 1
        for (int c = nfl; c < ncon; c++) {
            const int zone = (int)E.c_s[13][c];
            if (zone == 0) continue;
            const int meta = E.c_meta[c];
            const DevClass<T>& cl = m.cls[meta_cls(meta)];
            pair_jacobian(m, E, c, meta, cl.dim, JB, lane);
            const T D0 = E.c_D0[c];
#pragma unroll 1
            for (int idx = lane; idx < 144; idx += 32) {
                const int u = idx / 12, v = idx - 12 * u;
                T gx[6], gy[6];
#pragma unroll
                for (int j = 0; j < 6; j++) { gx[j] = JB[j * 12 + u]; gy[j] = JB[j * 12 + v]; }
                KD[idx] += cone_quad(cl, E, c, zone, D0, gx, gy);
            }
            __syncwarp();
        }
        if (lane < 24) {
            const int k = lane / 6, e = lane - 6 * k;
            const int i = m.tab.sym_r[e], cc = m.tab.sym_c[e];
            const T hv = E.H[A_KK + 6 * k + e];
            KD[(3 * k + i) * 12 + 3 * k + cc] += hv;
            if (i != cc) KD[(3 * k + cc) * 12 + 3 * k + i] += hv;
        }
        __syncwarp();
        return 3;
    }
    T* X = E.g6() + 21;                                 // 3x3 coupling block (rows: leg lb, columns: leg la)
    if (first >= 0 && lane < 9) X[lane] = 0;
    __syncwarp();
#pragma unroll 1
    for (int c = nfl; c < ncon; c++) {
        const int zone = (int)E.c_s[13][c];
        if (zone == 0) continue;
        const int meta = E.c_meta[c];
        const DevClass<T>& cl = m.cls[meta_cls(meta)];
        pair_jacobian(m, E, c, meta, cl.dim, JB, lane);
        const int b2 = meta_b2(meta), b1 = meta_b1p(meta) - 1;
        const int l2 = b2 >= 1 ? (b2 - 1) / 3 : -1, l1 = b1 >= 1 ? (b1 - 1) / 3 : -1;
        const bool two = l1 >= 0 && l2 >= 0 && l1 != l2;
        const int ka = two ? (l1 < l2 ? l1 : l2) : (l2 >= 0 ? l2 : l1), kb = two ? (l1 < l2 ? l2 : l1) : -1;
        // lanes 0..5: 3x3 block of leg ka; 6..11: leg kb; 12..20: coupling block (row in kb, column in ka)
        if (lane < 6 || (two && lane < 21)) {
            int u, v;
            T* dst;
            if (lane < 12) {
                const int k = lane < 6 ? ka : kb, e = lane < 6 ? lane : lane - 6;
                u = 3 * k + m.tab.sym_r[e]; v = 3 * k + m.tab.sym_c[e];
                dst = E.H + A_KK + 6 * k + e;
            } else {
                const int e = lane - 12, i = e / 3;
                u = 3 * kb + i; v = 3 * ka + (e - 3 * i);
                dst = X + e;
            }
            T gx[6], gy[6];
#pragma unroll
            for (int j = 0; j < 6; j++) { gx[j] = JB[j * 12 + u]; gy[j] = JB[j * 12 + v]; }
            *dst += cone_quad(cl, E, c, zone, E.c_D0[c], gx, gy);
        }
        __syncwarp();
    }
    if (first < 0) return 1;
    const int la = first >> 2, lb = first & 3;
    if (lane < 21) {
        const int r = m.tab.sym_r[lane], cc = m.tab.sym_c[lane];
        T v;
        if (r < 3) v = E.H[A_KK + 6 * la + SYM(r, cc)];
        else if (cc >= 3) v = E.H[A_KK + 6 * lb + SYM(r - 3, cc - 3)];
        else v = X[(r - 3) * 3 + cc];
        E.g6()[lane] = v;
    }
    __syncwarp();
    return 2 | (first << 2);
}

// line-search polynomials of every contact for the direction whose body-space image is E.bv[1] (S), at the
// point whose image is E.bv[0] (A):  c_s[0..2] quad, 3 U0, 4 V0, 5 UU, 6 UV, 7 VV, 8 T0, 9 zone0, 10 x0
template <typename T>
GO1_DEV void ls_prepare(const DevModel<T>& m, WEnv<T>& E, int lane) {
    typedef Num<T> N;
    const int ncon = E.ncon;
#pragma unroll 1
    for (int c = lane; c < ncon; c += 32) {
        const int meta = E.c_meta[c];
        const DevClass<T>& cl = m.cls[meta_cls(meta)];
        const int dim = cl.dim;
        const T r[3] = {E.c_r[0][c], E.c_r[1][c], E.c_r[2][c]};
        T A[6], S[6], F[9];
        contact_image(E.bv[0], meta, A);
        contact_image(E.bv[1], meta, S);
        contact_frame(m, E, c, dim, F);
        const T D0 = E.c_D0[c], mu = cl.mu;
        T q0 = 0, q1 = 0, q2 = 0, U0 = 0, V0 = 0, UU = 0, UV = 0, VV = 0;
#pragma unroll
        for (int j = 0; j < 6; j++) {
            if (j >= dim) continue;
            T jar, jv;
            if (j < 3) {
                T g[3];
                cross3(g, r, F + 3 * j);
                jar = dot3(g, A) + dot3(F + 3 * j, A + 3) - E.c_aref[j][c];
                jv = dot3(g, S) + dot3(F + 3 * j, S + 3);
            } else {
                jar = dot3(F + 3 * (j - 3), A) - E.c_aref[j][c];
                jv = dot3(F + 3 * (j - 3), S);
            }
            const T D = D0 * cl.drat[j];
            q0 += T(0.5) * D * jar * jar; q1 += D * jar * jv; q2 += T(0.5) * D * jv * jv;
            if (j == 0) { U0 = jar * mu; V0 = jv * mu; }
            else { const T u = jar * cl.fri[j - 1], v = jv * cl.fri[j - 1]; UU += u * u; UV += u * v; VV += v * v; }
        }
        // zone at alpha = 0 (0 top, 1 bottom, 2 middle): reference point of the cost DIFFERENCES
        const T T0 = N::sqrt_(N::max_(UU, T(0)));
        T z0;
        if (UU <= 0) z0 = U0 < 0 ? T(1) : T(0);
        else if (U0 >= mu * T0) z0 = 0;
        else if (mu * U0 + T0 <= 0) z0 = 1;
        else z0 = 2;
        E.c_s[0][c] = q0; E.c_s[1][c] = q1; E.c_s[2][c] = q2; E.c_s[3][c] = U0; E.c_s[4][c] = V0; E.c_s[5][c] = UU;
        E.c_s[6][c] = UV; E.c_s[7][c] = VV; E.c_s[8][c] = T0; E.c_s[9][c] = z0; E.c_s[10][c] = U0 - mu * T0;
    }
    __syncwarp();
}

template <typename T> struct LSPoint { T alpha, cost, d1, d2; };
template <typename T> struct LSRow { T fr_jar, fr_jv, lim_jar, lim_jv; };

// PrimalEval at step alpha.  `cost` is accumulated as the DIFFERENCE to the cost at alpha = 0, row by row: rows
// that stay in the same piece contribute alpha*q1 + alpha^2*q2 (no constant term), so the comparison of two
// candidate steps is meaningful in fp32 even when the total cost is ~1e4 (MuJoCo compares absolute fp64 costs).
template <typename T>
GO1_NOINL LSPoint<T> ls_eval(const DevModel<T>& m, const WEnv<T>& E, int lane, T alpha, T g1, T g2, const DofRows<T>& dr, const LSRow<T>& lr) {
    typedef Num<T> N;
    T cost = 0, d1 = 0, d2 = 0;
    if (lane == 0) { cost = alpha * (g1 + alpha * g2); d1 = 2 * alpha * g2 + g1; d2 = 2 * g2; }   // Gauss term, relative to alpha = 0
    const int ncon = E.ncon;
#pragma unroll 1
    for (int c = lane; c < ncon; c += 32) {
        const DevClass<T>& cl = m.cls[meta_cls(E.c_meta[c])];
        const T mu = cl.mu;
        const T q0 = E.c_s[0][c], q1 = E.c_s[1][c], q2 = E.c_s[2][c], q3 = E.c_s[3][c], q4 = E.c_s[4][c], q5 = E.c_s[5][c];
        const T q6 = E.c_s[6][c], q7 = E.c_s[7][c], T0 = E.c_s[8][c], x0 = E.c_s[10][c];
        const int z0 = (int)E.c_s[9][c];
        const T Dm = E.c_D0[c] / N::max_(mu * mu * (1 + mu * mu), N::MINVAL);
        const T Nn = q3 + alpha * q4, Tsq = q5 + alpha * (2 * q6 + alpha * q7);
        int za;
        T Tt = 0;
        if (Tsq <= 0) za = Nn < 0 ? 1 : 0;
        else {
            Tt = N::sqrt_(Tsq);
            if (Nn >= mu * Tt) za = 0;
            else if (mu * Nn + Tt <= 0) za = 1;
            else za = 2;
        }
        T ca = 0;   // cost at alpha (only needed when the zone changed)
        if (za == 1) {
            d1 += q1 + 2 * alpha * q2; d2 += 2 * q2;
            ca = q0 + alpha * (q1 + alpha * q2);
        } else if (za == 2) {
            const T N1 = q4, T1 = (q6 + alpha * q7) / Tt, T2 = q7 / Tt - (q6 + alpha * q7) * T1 / (Tt * Tt);
            const T NT = Nn - mu * Tt, NT1 = N1 - mu * T1;
            d1 += Dm * NT * NT1;
            d2 += Dm * (NT1 * NT1 + NT * (-mu * T2));
            ca = T(0.5) * Dm * NT * NT;
        }
        if (za == z0) {
            if (za == 1) cost += alpha * (q1 + alpha * q2);
            else if (za == 2) {
                const T dx = alpha * (q4 - mu * (2 * q6 + alpha * q7) / (Tt + T0));
                cost += T(0.5) * Dm * dx * ((Nn - mu * Tt) + x0);
            }
        } else {
            const T c0 = z0 == 1 ? q0 : (z0 == 2 ? T(0.5) * Dm * x0 * x0 : T(0));
            cost += ca - c0;
        }
    }
    if (lane >= 6 && lane < NDOF) {
        const T fl = dr.floss;
        if (fl > 0) {
            const T R = dr.Rf, Rf = R * fl, jar = lr.fr_jar, jv = lr.fr_jv;
            const T x = jar + alpha * jv;
            const int za = (-Rf < x && x < Rf) ? 0 : (x <= -Rf ? 1 : 2);
            const int z0 = (-Rf < jar && jar < Rf) ? 0 : (jar <= -Rf ? 1 : 2);
            const T D = T(1) / R;
            if (za == 0) { d1 += D * jar * jv + alpha * D * jv * jv; d2 += D * jv * jv; }
            else if (za == 1) d1 += -fl * jv;
            else d1 += fl * jv;
            if (za == z0) {
                if (za == 0) cost += alpha * (D * jar * jv + T(0.5) * alpha * D * jv * jv);
                else if (za == 1) cost += -fl * alpha * jv;
                else cost += fl * alpha * jv;
            } else {
                const T ca = za == 0 ? T(0.5) * D * x * x : (za == 1 ? fl * (T(-0.5) * Rf - x) : fl * (T(-0.5) * Rf + x));
                const T c0 = z0 == 0 ? T(0.5) * D * jar * jar : (z0 == 1 ? fl * (T(-0.5) * Rf - jar) : fl * (T(-0.5) * Rf + jar));
                cost += ca - c0;
            }
        }
        if (dr.lim_s != 0) {
            const T jar = lr.lim_jar, jv = lr.lim_jv, D = dr.lim_D;
            const T x = jar + alpha * jv;
            const bool aa = x < 0, a0 = jar < 0;
            if (aa) { d1 += D * jar * jv + alpha * D * jv * jv; d2 += D * jv * jv; }
            if (aa && a0) cost += alpha * (D * jar * jv + T(0.5) * alpha * D * jv * jv);
            else cost += (aa ? T(0.5) * D * x * x : T(0)) - (a0 ? T(0.5) * D * jar * jar : T(0));
        }
    }
    LSPoint<T> p;
    p.alpha = alpha; p.cost = warp_sum(cost); p.d1 = warp_sum(d1); p.d2 = warp_sum(d2);
    if (p.d2 <= 0) p.d2 = N::MINVAL;
    return p;
}


#define GO1_LEGF(field, j) ((&m.field[0][0])[j])

// ---------------------------------------------------------------------------------------------
// one physics substep (mj_step) of the environment owned by this warp
template <typename T>
GO1_NOINL void substep_w(const DevModel<T>& m, const EnvCfg<T>& cfg, WEnv<T>& E, const int lane, SubOut<T>& out,
                         const bool capture, T* con_out, const bool active, T* bip_out = nullptr) {
    typedef Num<T> N;
    const T h = m.timestep;
    const int jl = lane >= 6 && lane < NDOF ? lane - 6 : 0;      // leg-dof index of this lane (0..11)
    const bool legdof = lane >= 6 && lane < NDOF;
    T qfa = 0, qfs = 0, as_ = 0;
    DofRows<T> dr;
    dr.fr_aref = 0; dr.lim_s = 0; dr.lim_D = 0; dr.lim_aref = 0; dr.floss = 0; dr.Rf = 1;
    int ncon = 0;
#if defined(GO1_PHASE_TIMING) && !defined(GO1_EMU)
    GO1_PT_SYNC(6, __syncthreads());
#elif GO1_LOCKSTEP
    __syncthreads();     // the warps of a block run the substep phases together: they share instruction fetches
#endif
    if (active) {
    // ------------------------------------------------------------------ mj_kinematics
    T R0[9];
    {
        T q[4] = {E.qpos[3], E.qpos[4], E.qpos[5], E.qpos[6]};
        const T nq = N::sqrt_(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
        if (nq < N::MINVAL) { q[0] = 1; q[1] = q[2] = q[3] = 0; }
        else { const T inv = T(1) / nq; q[0] *= inv; q[1] *= inv; q[2] *= inv; q[3] *= inv; }
        __syncwarp();
        if (lane == 0) { E.qpos[3] = q[0]; E.qpos[4] = q[1]; E.qpos[5] = q[2]; E.qpos[6] = q[3]; }
        quat_to_mat(R0, q);
    }
    const T plane_d0 = m.plane_frame[0] * (E.qpos[0] - m.plane_pos[0]) + m.plane_frame[1] * (E.qpos[1] - m.plane_pos[1]) +
                       m.plane_frame[2] * (E.qpos[2] - m.plane_pos[2]);
    out.trunk_z = E.qpos[2];
    // body lanes: lane < 12 -> leg body (k, i) = body 1 + lane; lane 12 -> trunk (body 0)
    const int bk = lane < 12 ? lane / 3 : 0, bi = lane < 12 ? lane - 3 * bk : 0;
    const int mybody = lane < 12 ? lane + 1 : 0;
    T Rb[9], pb[3] = {0, 0, 0}, axw[3] = {0, 0, 0};
#pragma unroll
    for (int e = 0; e < 9; e++) Rb[e] = R0[e];
    {
        T sn = 0, cs = 1;
        if (lane < 12) N::sincos_(E.qpos[7 + lane] - GO1_LEGF(qpos0, lane), &sn, &cs);
#pragma unroll
        for (int j = 0; j < 3; j++) {
            const T sj = __shfl_sync(FULL, sn, 3 * bk + j), cj = __shfl_sync(FULL, cs, 3 * bk + j);
            if (lane < 12 && j <= bi) {
                const T* bp = m.body_pos[bk][j];
                const T* u = m.jnt_axis[bk][j];
                T t[3];
                matvec3(t, Rb, bp);
                pb[0] += t[0]; pb[1] += t[1]; pb[2] += t[2];
                matvec3(axw, Rb, u);
                const T oc = 1 - cj;
                const T Rl[9] = {cj + oc * u[0] * u[0], oc * u[0] * u[1] - sj * u[2], oc * u[0] * u[2] + sj * u[1],
                                 oc * u[0] * u[1] + sj * u[2], cj + oc * u[1] * u[1], oc * u[1] * u[2] - sj * u[0],
                                 oc * u[0] * u[2] - sj * u[1], oc * u[1] * u[2] + sj * u[0], cj + oc * u[2] * u[2]};
                T Rn[9];
#pragma unroll
                for (int r = 0; r < 3; r++)
#pragma unroll
                    for (int cc = 0; cc < 3; cc++) Rn[3 * r + cc] = Rb[3 * r] * Rl[cc] + Rb[3 * r + 1] * Rl[3 + cc] + Rb[3 * r + 2] * Rl[6 + cc];
#pragma unroll
                for (int e = 0; e < 9; e++) Rb[e] = Rn[e];
            }
        }
    }
    // ------------------------------------------------------------------ mj_comPos: COM, cinert, cdof
    T cin[10];
    {
        const T* ipos = lane < 12 ? m.ipos[bk][bi] : m.t_ipos;
        const T mass = lane < 12 ? m.mass[bk][bi] : (lane == 12 ? m.t_mass : T(0));
        const T* Ib = lane < 12 ? m.Ib[bk][bi] : m.t_Ib;
        T xip[3];
        matvec3(xip, Rb, ipos);
        xip[0] += pb[0]; xip[1] += pb[1]; xip[2] += pb[2];
        const T inv = T(1) / m.total_mass;
        T com[3];
#pragma unroll
        for (int k = 0; k < 3; k++) com[k] = warp_sum(mass * xip[k]) * inv;
        if (lane < NB) {
#pragma unroll
            for (int e = 0; e < 9; e++) E.xmat()[mybody][e] = Rb[e];
#pragma unroll
            for (int e = 0; e < 3; e++) E.xpos()[mybody][e] = pb[e];
            if (bip_out) {      // biped mode, last substep: data.xpos / site_xpos / xmat as the reward terms will read them
                const T wy = E.qpos[1] + pb[1], wz = E.qpos[2] + pb[2];
                if (lane < 12 && bi == 0) bip_out[BK_HIP_Z + bk] = wz;
                if (lane < 12 && bi == 0 && bk >= 2) bip_out[BK_REAR_HIP_Y + bk - 2] = wy;
                if (lane < 12 && bi == 2 && bk < 2) {
                    const T sx = bk ? cfg.foot_site[1][0] : cfg.foot_site[0][0], sy = bk ? cfg.foot_site[1][1] : cfg.foot_site[0][1],
                            sz = bk ? cfg.foot_site[1][2] : cfg.foot_site[0][2];
                    bip_out[BK_FRONT_CALF_Z + bk] = wz;
                    bip_out[BK_FRONT_FOOT_Z + bk] = wz + Rb[6] * sx + Rb[7] * sy + Rb[8] * sz;
                }
                if (lane == 12) bip_out[BK_TRUNK_XZ] = R0[6];
            }
            const T dif[3] = {xip[0] - com[0], xip[1] - com[1], xip[2] - com[2]};
            cinert_of(cin, Rb, Ib, mass, dif);
#pragma unroll
            for (int e = 0; e < 10; e++) E.crb[mybody][e] = cin[e];
        }
        if (lane < 12) {
            const T off[3] = {com[0] - pb[0], com[1] - pb[1], com[2] - pb[2]};
            T cl[3];
            cross3(cl, axw, off);
#pragma unroll
            for (int e = 0; e < 3; e++) { E.cdof[6 + lane][e] = axw[e]; E.cdof[6 + lane][3 + e] = cl[e]; }
        } else if (lane < NDOF) {
            const int j = lane - 12;
            if (j < 3) {
#pragma unroll
                for (int e = 0; e < 6; e++) E.cdof[j][e] = (e == 3 + j) ? T(1) : T(0);
            } else {
                const T col[3] = {R0[j - 3], R0[3 + j - 3], R0[6 + j - 3]};
                T tl[3];
                cross3(tl, col, com);     // offset = com - trunk origin
#pragma unroll
                for (int e = 0; e < 3; e++) { E.cdof[j][e] = col[e]; E.cdof[j][3 + e] = tl[e]; }
            }
        }
        if (lane < 3) E.com[lane] = com[lane];
    }
    __syncwarp();
    GO1_PBX(128);      // F0 (experiment)
    // ------------------------------------------------------------------ mj_collision (plane vs geoms), contacts in geom order
    {
        int base = 0;
        const T com[3] = {E.com[0], E.com[1], E.com[2]};
#pragma unroll 1
        for (int g0 = 0; g0 < m.nflat; g0 += 32) {
            const int gi = g0 + lane;
            Stage4<T> st;
            st.n = 0;
#pragma unroll
            for (int j = 0; j < 4; j++) { st.px[j] = st.py[j] = st.pz[j] = st.d[j] = 0; }
            int meta = 0;
            if (gi < m.nflat) {
                const DevGeom<T>& g = m.flat[gi];
                const int b = g.slot;
                T Rg[9], gp[3];
#pragma unroll
                for (int e = 0; e < 9; e++) Rg[e] = E.xmat()[b][e];
                matvec3(gp, Rg, g.pos);
                gp[0] += E.xpos()[b][0]; gp[1] += E.xpos()[b][1]; gp[2] += E.xpos()[b][2];
                const T dc = plane_d0 + dot3(m.plane_frame, gp);
                meta = b | (g.cls << 8) | (g.gid << 16);
                if (dc - m.flat_rb[gi] <= m.cls[g.cls].margin) collide_geom(m, g, Rg, gp, dc, st);
            }
            int incl = st.n;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) { const int v = __shfl_up_sync(FULL, incl, d); if (lane >= d) incl += v; }
            const int total = __shfl_sync(FULL, incl, 31);
            const int off = base + incl - st.n;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int c = off + j;
                if (j < st.n && c < CW) {
                    E.c_r[0][c] = st.px[j] - com[0]; E.c_r[1][c] = st.py[j] - com[1]; E.c_r[2][c] = st.pz[j] - com[2];
                    E.c_D0[c] = st.d[j]; E.c_meta[c] = meta;
                }
            }
            base += total;
        }
        __syncwarp();
        const int nfl = base < CW ? base : CW;
        int nover = base > CW ? base - CW : 0;
        // body ranges (floor contacts are sorted by body): bstart[b] = number of floor contacts with body < b
        if (lane <= NB) {
            int lo = 0, hi = nfl;
            while (lo < hi) { const int mid = (lo + hi) >> 1; if (meta_b2(E.c_meta[mid]) < lane) lo = mid + 1; else hi = mid; }
            E.bstart[lane] = lo;
        }
        GO1_PBX(1);    // F1
        // ---- robot-robot pairs (go1_pair.cuh).  Candidate geom pairs come from a table in (body1, body2, geom1, geom2) order --
        // mj_collision's order: bodies that are parent and child, geoms of one body and hip-motor pairs are not in it.  Pass 1:
        // bounding-sphere test of every table entry, the hits are compacted (order kept).  Pass 2: one lane per hit runs the
        // narrowphase.  The body frames are still in the Kc words; geom centres and the hit list use the (dead) contact scratch.
        int nall = nfl;
#if GO1_SELF_COLLISION
        {
            T* gc = &E.c_s[0][0];
            int* hits = reinterpret_cast<int*>(gc + 3 * GFLAT);
#pragma unroll 1
            for (int g = lane; g < m.nflat; g += 32) {
                const DevGeom<T>& G = m.flat[g];
                T cw[3];
                matvec3(cw, E.xmat()[G.slot], G.pos);
#pragma unroll
                for (int e = 0; e < 3; e++) gc[3 * g + e] = cw[e] + E.xpos()[G.slot][e];
            }
            __syncwarp();
            const unsigned char* gpa = cfg.gp_tab;
            const unsigned char* gpb = cfg.gp_tab + NGP;
            const T pmargin = m.pair_margin;
            int nhit = 0, p0 = 0;
#pragma unroll 1
            for (;;) {
                // fill: bounding-sphere test of the next table entries until a full pass of 32 candidates is waiting (hits[] holds < 64)
#pragma unroll 1
                while (p0 < m.ngp && nhit < 32) {
                    const int pi = p0 + lane;
                    bool hit = false;
                    if (pi < m.ngp) {
                        const int a = gpa[pi], b = gpb[pi];
                        const T d[3] = {gc[3 * a] - gc[3 * b], gc[3 * a + 1] - gc[3 * b + 1], gc[3 * a + 2] - gc[3 * b + 2]};
                        const T bound = m.flat_rb[a] + m.flat_rb[b] + pmargin;
                        hit = dot3(d, d) <= bound * bound;
                    }
                    const unsigned mask = __ballot_sync(FULL, hit);
                    if (hit) hits[nhit + __popc(mask & ((1u << lane) - 1u))] = pi;
                    nhit += __popc(mask);
                    p0 += 32;
                }
                __syncwarp();
#if GO1_SELF_COLLISION == 3
                nhit = 0;          // timing experiment only: bounding-sphere scan without the narrowphase
#endif
                if (nhit == 0) break;
                const int take = nhit < 32 ? nhit : 32;
                Stage2<T> st;
                st.cnt = 0;
#pragma unroll
                for (int k = 0; k < 3; k++) { st.p[0][k] = st.p[1][k] = st.n[0][k] = st.n[1][k] = 0; }
                st.d[0] = st.d[1] = 0;
                int meta = 0;
                if (lane < take) {
                    const int pi = hits[lane];
                    int g1 = gpa[pi], g2 = gpb[pi];             // flat indices; geom1 = the lower geom type (mj_collideGeoms)
                    if (m.flat[g1].type > m.flat[g2].type) { const int t_ = g1; g1 = g2; g2 = t_; }
                    GeomW<T> w1, w2;
#pragma unroll
                    for (int q = 0; q < 2; q++) {
                        const int gq = q == 0 ? g1 : g2;
                        const DevGeom<T>& g = m.flat[gq];
                        GeomW<T>& w = q == 0 ? w1 : w2;
                        const T* Rg = E.xmat()[g.slot];
                        w.type = g.type;
                        matvec3(w.z, Rg, g.zax);
                        matvec3(w.x, Rg, g.xax);
#pragma unroll
                        for (int k = 0; k < 3; k++) { w.c[k] = gc[3 * gq + k]; w.s[k] = g.size[k]; }
                    }
                    const DevGeom<T>& G1 = m.flat[g1]; const DevGeom<T>& G2 = m.flat[g2];
                    const int ci = m.pair_cls[G1.kind][G2.kind];
                    pp_collide(w1, w2, m.cls[ci].margin, m.cls[ci].incmargin, st);
                    meta = G2.slot | ((G1.slot + 1) << 4) | (ci << 8) | (G2.gid << 16) | (G1.gid << 22);
                }
                const int cnt = st.cnt < 2 ? st.cnt : 2;
                int incl = cnt;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) { const int v = __shfl_up_sync(FULL, incl, d); if (lane >= d) incl += v; }
                const int total = __shfl_sync(FULL, incl, 31);
                const int off = nall + incl - cnt;
#pragma unroll
                for (int j = 0; j < 2; j++) {
                    const int c = off + j;
                    if (j < cnt && c < CW && c - nfl < PW) {
                        E.c_r[0][c] = st.p[j][0] - com[0]; E.c_r[1][c] = st.p[j][1] - com[1]; E.c_r[2][c] = st.p[j][2] - com[2];
                        E.c_D0[c] = st.d[j]; E.c_meta[c] = meta;
                        E.c_n[0][c - nfl] = st.n[j][0]; E.c_n[1][c - nfl] = st.n[j][1]; E.c_n[2][c - nfl] = st.n[j][2];
                    }
                }
                nall += total;
                // the candidates that did not fit into this pass move to the front
                const int rest = 32 + lane < nhit ? hits[32 + lane] : 0;
                __syncwarp();
                if (32 + lane < nhit) hits[lane] = rest;
                nhit -= take;
                __syncwarp();
            }
            const int cap = nfl + PW < CW ? nfl + PW : CW;
            if (nall > cap) { nover += nall - cap; nall = cap; }
#if GO1_SELF_COLLISION == 2
            nall = nfl;      // timing experiment only: collision phase executed, contacts discarded
#endif
        }
#endif
        __syncwarp();
        if (lane == 0) { E.ncon = nall; E.nfl = nfl; E.overflow = nover; }
        __syncwarp();
    }
    ncon = E.ncon;
    GO1_PBX(2);        // F2
    // ------------------------------------------------------------------ mj_crb -> block-arrow M (+ armature)
    composite10(E.crb, lane);
    composite21(m, E.Kc, E.crb, false, lane);          // (overwrites xmat / xpos: collision is done)
    assemble_arrow(m, E, E.Kc, E.M, lane);
    if (legdof) { const int k = jl / 3, i = jl - 3 * k; E.M[A_KK + 6 * k + SYM(i, i)] += GO1_LEGF(armature, jl); }
    __syncwarp();
    // ------------------------------------------------------------------ mj_comVel + mj_rne (bias) + passive + actuation
    body_motion(E, E.qvel, E.bv[0], lane);
    T bias;
    {
        T angw[3], vx[3];
        const T wb[3] = {E.qvel[3], E.qvel[4], E.qvel[5]}, vl[3] = {E.qvel[0], E.qvel[1], E.qvel[2]};
        matvec3(angw, R0, wb);
        cross3(vx, vl, angw);
        T cacc[6] = {0, 0, 0, -m.gravity[0] + vx[0], -m.gravity[1] + vx[1], -m.gravity[2] + vx[2]};
#pragma unroll
        for (int j = 0; j < 3; j++) {
            if (lane < 12 && j <= bi) {
                const int pbody = j == 0 ? 0 : 3 * bk + j;
                T cvp[6], cd[6], cdd[6];
#pragma unroll
                for (int e = 0; e < 6; e++) { cvp[e] = E.bv[0][pbody][e]; cd[e] = E.cdof[6 + 3 * bk + j][e]; }
                cross_motion(cdd, cvp, cd);
                const T v = E.qvel[6 + 3 * bk + j];
#pragma unroll
                for (int e = 0; e < 6; e++) cacc[e] += cdd[e] * v;
            }
        }
        if (lane < NB) {
            T cv[6], t1[6], t2[6], t3[6];
#pragma unroll
            for (int e = 0; e < 6; e++) cv[e] = E.bv[0][mybody][e];
            mul_inert(t1, cin, cacc);
            mul_inert(t2, cin, cv);
            cross_force(t3, cv, t2);
#pragma unroll
            for (int e = 0; e < 6; e++) E.wrench()[mybody][e] = t1[e] + t3[e];
        }
        __syncwarp();
        bias = force_to_joint(E, E.wrench(), lane);
    }
    qfs = -bias;
    if (legdof) {
        const T ql = E.qpos[7 + jl], vl = E.qvel[6 + jl];
        T u = E.ctrl[jl];
        if (GO1_LEGF(act_climited, jl)) u = N::min_(GO1_LEGF(act_chi, jl), N::max_(GO1_LEGF(act_clo, jl), u));
        const T gear = GO1_LEGF(act_gear, jl);
        T force = GO1_LEGF(act_gain, jl) * u;
        if (GO1_LEGF(act_bias, jl)) force += GO1_LEGF(act_b0, jl) + GO1_LEGF(act_b1, jl) * (gear * ql) + GO1_LEGF(act_b2, jl) * (gear * vl);
        if (GO1_LEGF(act_flimited, jl)) force = N::min_(GO1_LEGF(act_fhi, jl), N::max_(GO1_LEGF(act_flo, jl), force));
        qfa = gear * force;
        qfs = -GO1_LEGF(damping, jl) * vl - bias + qfa;
        // friction-loss and limit rows of this dof (mj_makeConstraint / mj_referenceConstraint)
        dr.floss = GO1_LEGF(floss, jl); dr.Rf = GO1_LEGF(Rf, jl);
        dr.fr_aref = -GO1_LEGF(Bf, jl) * vl;
        if (GO1_LEGF(limited, jl)) {
            const T margin = GO1_LEGF(lim_margin, jl);
            const T dlo = ql - GO1_LEGF(lim_lo, jl), dhi = GO1_LEGF(lim_hi, jl) - ql;
            T pos = 0, sg = 0;
            if (dlo < margin) { pos = dlo; sg = 1; }
            else if (dhi < margin) { pos = dhi; sg = -1; }
            if (sg != 0) {
                const T* si = m.lim_imp[jl / 3][jl % 3];
                const T imp = impedance(si[0], si[1], si[2], si[3], si[4], pos - margin);
                const T R = N::max_(N::MINVAL, (1 - imp) * GO1_LEGF(dof_invw, jl) / imp);
                dr.lim_s = sg; dr.lim_D = T(1) / R;
                dr.lim_aref = -GO1_LEGF(lim_B, jl) * (sg * vl) - GO1_LEGF(lim_K, jl) * imp * (pos - margin);
            }
        }
    }
    // ------------------------------------------------------------------ qacc_smooth = M^-1 qfrc_smooth
#pragma unroll 1
    for (int idx = lane; idx < NARROW; idx += 32) E.H[idx] = E.M[idx];
    __syncwarp();
    arrow_factor(m, E.H, lane);
    as_ = arrow_solve(E, E.H, qfs, lane);
    // ------------------------------------------------------------------ contact rows: impedance, D, reference acceleration
    {
#pragma unroll 1
        for (int c = lane; c < ncon; c += 32) {
            const int meta = E.c_meta[c];
            const int b = meta_b2(meta), b1p = meta_b1p(meta);
            const DevClass<T>& cl = m.cls[meta_cls(meta)];
            const T dist = E.c_D0[c];
            const T pm = dist - cl.incmargin;
            const T imp = impedance(cl.d0, cl.d1, cl.width, cl.mid, cl.power, pm);
            // mj_diagApprox: translational body_invweight0 of both bodies (the world's is 0)
            T dA = (b == 0) ? m.t_invw[0] : (&m.invw[0][0][0])[2 * (b - 1)];
            if (b1p) dA += (b1p == 1) ? m.t_invw[0] : (&m.invw[0][0][0])[2 * (b1p - 2)];
            const T Rr = N::max_(N::MINVAL, (1 - imp) * dA / imp);
            E.c_D0[c] = T(1) / Rr;
            const T r[3] = {E.c_r[0][c], E.c_r[1][c], E.c_r[2][c]};
            T cv[6], vp[3], F[9];
            contact_image(E.bv[0], meta, cv);
            contact_frame(m, E, c, cl.dim, F);
            cross3(vp, cv, r);
            vp[0] += cv[3]; vp[1] += cv[4]; vp[2] += cv[5];
#pragma unroll
            for (int j = 0; j < 3; j++) {
                T ar = -cl.B * dot3(F + 3 * j, vp);
                if (j == 0) ar -= cl.K * imp * pm;
                E.c_aref[j][c] = ar;
                E.c_aref[3 + j][c] = -cl.B * dot3(F + 3 * j, cv);
            }
#ifdef GO1_EMU_TRACE
            std::printf("ROW c %d b2 %d b1p %d cls %d dim %d dist %.6g D0 %.6g imp %.6g dA %.6g aref %.6g %.6g %.6g r %.5f %.5f %.5f n %.5f %.5f %.5f\n", c, b, b1p, meta_cls(meta), cl.dim, (double)dist,
                        (double)E.c_D0[c], (double)imp, (double)dA, (double)E.c_aref[0][c], (double)E.c_aref[1][c], (double)E.c_aref[2][c], (double)r[0], (double)r[1], (double)r[2], (double)F[0], (double)F[1], (double)F[2]);
#endif
            if (capture) {
                T* e = con_out + 1 + c * 12;
                e[0] = (T)(meta >> 16); e[1] = (T)cl.dim; e[2] = dist;
                e[3] = r[0] + E.com[0] + E.qpos[0]; e[4] = r[1] + E.com[1] + E.qpos[1]; e[5] = r[2] + E.com[2] + E.qpos[2];
            }
        }
        if (capture && lane == 0) con_out[0] = (T)ncon;
    }
    __syncwarp();
        GO1_PBX(64);   // F3 (experiment)
    } else { GO1_PBX(128); GO1_PBX(1); GO1_PBX(2); GO1_PBX(64); }   // active
    // ------------------------------------------------------------------ mj_fwdConstraint: warmstart + Newton
    T a = as_, Ma = qfs, qfc = 0, hdiag = 0, grad = 0, cost = 0, oldcost = 0;
    const T scale = T(1) / (m.meaninertia * T(18));
    const T tol = cfg.tolerance;
    const int maxiter = cfg.solver_iterations;
    int iter = 0, phase = 0, nstall = 0;
    T cost_s = 0, improvement = 0, gradient = 0;
    bool first = true;
    T* frc_out = capture ? con_out : (T*)nullptr;
    bool conv = !active;
#if GO1_SKIP_SMOOTH_CHECK
    // MuJoCo compares the cost of the warm start with the cost of qacc_smooth and starts from the better one.  After the
    // first substep of a control step the warm start is the (extrapolated) solution of 2 ms ago and practically always
    // wins; the fp32 build skips that comparison there (one evaluation round per substep).  Same optimum either way.
    if (sizeof(T) == 4 && active && out.nsub >= 1) {
        a = lane < NDOF ? E.warm[lane] : T(0);
        if (GO1_WARM_EXTRAP && out.nsub >= 2) a = T(2) * a - out.qacc_prev;
        if (lane < NDOF) E.xv[lane] = a;
        __syncwarp();
        Ma = arrow_matvec(E, E.M, lane);
        __syncwarp();
        phase = 2;
    }
#endif
#pragma unroll 1
    while (true) {
#if defined(GO1_PHASE_TIMING) && !defined(GO1_EMU)
        bool all_conv_;
        GO1_PT_SYNC(5, all_conv_ = __syncthreads_and(conv));
        if (all_conv_) break;
#elif GO1_LOCKSTEP
        if (__syncthreads_and(conv)) break;       // one Newton round per barrier, until every warp of the block converged
#else
        if (conv) break;
#endif
        if (conv) { GO1_PBX(32); GO1_PB3(); continue; }
        do {
        if (lane < NDOF) E.xv[lane] = a;
        __syncwarp();
        body_motion(E, E.xv, E.bv[0], lane);
        const T rows = evaluate_rows(m, E, lane, a, dr, qfc, hdiag, frc_out);
        GO1_PBX(32);   // B0 (experiment)
        const T gauss = lane < NDOF ? (Ma - qfs) * (a - as_) : T(0);
        const T cnew = warp_sum(rows + T(0.5) * gauss);
        if (phase == 0) {           // cost at qacc_smooth; now try the warmstart
            cost_s = cnew; phase = 1;
            a = lane < NDOF ? E.warm[lane] : T(0);
            // MuJoCo warm-starts from the previous qacc.  The optimum is unique, so the start only decides how many
            // Newton iterations are needed: inside a control step (constant ctrl) the solution moves smoothly and the
            // linear extrapolation of the last two solutions is a closer start (-15 % iterations on the golden trace).
            // The candidate still has to beat qacc_smooth's cost below, exactly like MuJoCo's warm start.
            if (GO1_WARM_EXTRAP == 2 && sizeof(T) == 4 && out.nsub >= 3) a = T(3) * (a - out.qacc_prev) + out.qacc_prev2;
            else if (GO1_WARM_EXTRAP && sizeof(T) == 4 && out.nsub >= 2) a = T(2) * a - out.qacc_prev;
            if (lane < NDOF) E.xv[lane] = a;
            __syncwarp();
            Ma = arrow_matvec(E, E.M, lane);
            __syncwarp();
            GO1_PB3();
            break;
        }
        if (phase == 1) {
            phase = 2;
            if (cnew > cost_s) {    // qacc_smooth is the better start: evaluate it again
                a = as_;
                if (lane < NDOF) E.xv[lane] = a;
                __syncwarp();
                Ma = arrow_matvec(E, E.M, lane);
                __syncwarp();
                GO1_PB3();
                break;
            }
        }
        grad = lane < NDOF ? Ma - qfs - qfc : T(0);
        if (!first) {
            improvement = scale * (oldcost - cnew);
            // fp32 noise floor: a cost of magnitude c cannot resolve improvements below ~eps*c.  Heavily loaded states
            // (tens of deep contacts, cost ~1e6) otherwise never meet the absolute tolerance and run to the iteration cap.
            if (sizeof(T) == 4) nstall = improvement < N::max_(tol, T(2.4e-7) * scale * N::abs_(cnew)) ? nstall + 1 : 0;
            gradient = scale * N::sqrt_(warp_sum(grad * grad));
            iter++;
        }
        first = false;
        cost = cnew;
#ifdef GO1_DBG_FIXED_ITERS
        if (iter >= GO1_DBG_FIXED_ITERS) { conv = true; GO1_PB3(); break; }      // timing experiment only: same trip count everywhere
#else
        if (iter >= maxiter) { conv = true; GO1_PB3(); break; }
        // MuJoCo's exit test (after >= 1 move): gradient small, or improvement small.  The measured improvement is a
        // difference of O(cost) numbers and is pure rounding noise in fp32 once it drops below ~1e-7*cost, so
        // "improvement small" is only trusted when the Newton decrement -g.s/2 (the predicted improvement of the
        // next move, free of cancellation) is small too; see DESIGN.md.
        if (iter > 0 && gradient < tol) { conv = true; GO1_PB3(); break; }
        // fp32: |grad| never reaches MuJoCo's 1e-8 (rounding floor of M a - qfrc), but H >= M >= lambda_lb*I bounds the Newton
        // decrement from the gradient alone: g'H^-1 g / 2 <= |g|^2 / (2 lambda_lb).  If even that bound is below the
        // tolerance the point is converged and the Hessian assembly / factorisation / solve of this round are skipped.
        if (GO1_GRAD_BOUND_EXIT && sizeof(T) == 4 && iter > 0 && m.lambda_lb > 0 &&
            T(0.5) * gradient * gradient < tol * scale * m.lambda_lb) { conv = true; GO1_PB3(); break; }
        if (nstall >= 3) { conv = true; GO1_PB3(); break; }
#endif
        // ---- Hessian H = M + J'DJ (+ cone blocks), block-arrow; search = -H^-1 grad
        GO1_PBX(4);    // B1
        if (contact_stiffness(m, E, lane)) {
            composite21(m, E.Kc, E.crb, true, lane);
            assemble_arrow(m, E, E.Kc, E.H, lane);
            if (legdof) { const int k = jl / 3, i = jl - 3 * k; E.H[A_KK + 6 * k + SYM(i, i)] += GO1_LEGF(armature, jl) + hdiag; }
        } else {
#pragma unroll 1
            for (int idx = lane; idx < NARROW; idx += 32) E.H[idx] = E.M[idx];
            __syncwarp();
            if (legdof) { const int k = jl / 3, i = jl - 3 * k; E.H[A_KK + 6 * k + SYM(i, i)] += hdiag; }
        }
        __syncwarp();
        // robot-robot contacts (rare): inside one leg block, or leg-leg fill -> dense 12 x 12 leg block
        const int pmode = E.ncon > E.nfl ? pair_hessian(m, E, lane) : 0;
#if defined(GO1_PHASE_TIMING) && !defined(GO1_EMU)
        if ((pmode & 3) > out.maxmode) out.maxmode = pmode & 3;
        out.rounds++;
#endif
#ifdef GO1_EMU_COUNT
        if (lane == 0) { __atomic_fetch_add(&g_emu_rounds[pmode & 3], 1, __ATOMIC_RELAXED); if (E.ncon > E.nfl) __atomic_fetch_add(&g_emu_rounds[3], 1, __ATOMIC_RELAXED); }
#endif
        GO1_PBX(8);    // B2
        T s;
        if ((pmode & 3) == 3) {
            dense_factor(E, m, E.H, lane);
            s = -dense_solve(E, E.H, grad, lane);
        } else if ((pmode & 3) == 2) {
            const int la = pmode >> 4, lb = (pmode >> 2) & 3;
            pairleg_factor(E, m, E.H, la, lb, lane);
            s = -pairleg_solve(E, E.H, grad, la, lb, lane);
        } else {
            arrow_factor(m, E.H, lane);
            s = -arrow_solve(E, E.H, grad, lane);
        }
        // Newton decrement: -g.s/2 is the improvement the quadratic model predicts for the move about to be made.  If it
        // is below the tolerance, MuJoCo would make the move, measure "improvement < tolerance" and stop; stopping here
        // reaches the same optimum to within the same tolerance and skips that line search and the extra evaluation
        // round (-19 % iterations on the golden sets).  fp64 check build: only after MuJoCo's own test has fired.
        if ((GO1_DEC_EXIT && sizeof(T) == 4 && (iter > 0 || GO1_DEC_EXIT > 1)) || (iter > 0 && improvement < tol)) {
            const T dec = warp_sum(grad * s);
#ifndef GO1_DBG_FIXED_ITERS
            if (T(-0.5) * scale * dec < tol) { conv = true; GO1_PBX(16); break; }
#endif
        }
        // ---- exact line search (PrimalSearch)
        GO1_PBX(16);   // B3
        if (lane < NDOF) E.xv[lane] = s;
        __syncwarp();
        const T Ms = arrow_matvec(E, E.M, lane);
        body_motion(E, E.xv, E.bv[1], lane);
#ifdef GO1_EMU_TRACE
        int g_trace_it = 0;
#endif
        T alpha = 0;
        {
            const T snorm = N::sqrt_(warp_sum(s * s));
            if (snorm >= N::MINVAL) {
                T gtol = tol * m.ls_tolerance * snorm / scale;
                ls_prepare(m, E, lane);
                LSRow<T> lr;
                lr.fr_jar = a - dr.fr_aref; lr.fr_jv = s;
                lr.lim_jar = dr.lim_s * a - dr.lim_aref; lr.lim_jv = dr.lim_s * s;
                const T g1 = warp_sum(lane < NDOF ? s * (Ma - qfs) : T(0));
                const T g2 = warp_sum(lane < NDOF ? T(0.5) * s * Ms : T(0));
#ifdef GO1_DBG_LSCAP
                const int lsmax = GO1_DBG_LSCAP;      // timing experiment only
#else
                const int lsmax = m.ls_iterations;
#endif
                int it = 0;
#ifdef GO1_EMU_TRACE
#define GO1_TRACE_IT g_trace_it = it
#else
#define GO1_TRACE_IT
#endif
                if (sizeof(T) == 4) {
                    // fp32 product path: safeguarded Newton on the (continuous, increasing) derivative d1(alpha).
                    //  * the point alpha = 0 needs no evaluation: d1(0) = grad.s and, for the exact Hessian, d2(0) = -grad.s,
                    //    so the first trial is the full Newton move alpha = 1;
                    //  * MuJoCo's absolute tolerance (1e-10 |s|) is below the fp32 rounding noise of d1: stop when d1 has
                    //    dropped by GO1_LS_REL.  The Newton exit tests are unchanged, so the optimum is the same;
                    //  * one evaluation per refinement (Newton step inside the bracket, else secant, else bisection) instead
                    //    of MuJoCo's three.
                    const T d10 = warp_sum(grad * s);
                    gtol = N::max_(gtol, T(GO1_LS_REL) * N::abs_(d10));
                    LSPoint<T> lo, hi, p;
                    lo.alpha = 0; lo.cost = 0; lo.d1 = d10; lo.d2 = N::max_(-d10, N::MINVAL);
                    hi = lo;
                    bool have_hi = false;
                    int stall = 0;
                    p = ls_eval(m, E, lane, T(1), g1, g2, dr, lr); it++;
#pragma unroll 1
                    for (;;) {
                        if (N::abs_(p.d1) < gtol) { alpha = p.alpha; break; }
                        if (it >= (lsmax < GO1_LS_FAST_MAX ? lsmax : GO1_LS_FAST_MAX)) {
                            // out of budget (d1 at its rounding-noise floor): the evaluated point of lowest cost; costs are
                            // differences to alpha = 0, so a negative cost is a guaranteed decrease
                            T best = 0;
                            if (p.cost < best) { best = p.cost; alpha = p.alpha; }
                            if (lo.cost < best) { best = lo.cost; alpha = lo.alpha; }
                            if (have_hi && hi.cost < best) { best = hi.cost; alpha = hi.alpha; }
                            break;
                        }
                        if (p.d1 < 0) { stall = (have_hi && N::abs_(p.d1) > T(0.5) * N::abs_(lo.d1)) ? stall + 1 : 0; lo = p; }
                        else { stall = (have_hi && N::abs_(p.d1) > T(0.5) * N::abs_(hi.d1)) ? stall + 1 : 0; hi = p; have_hi = true; }
                        T cand = p.alpha - p.d1 / p.d2;
                        if (have_hi) {
                            const T w = hi.alpha - lo.alpha;
                            if (!(cand > lo.alpha && cand < hi.alpha)) cand = lo.alpha - lo.d1 * w / (hi.d1 - lo.d1);
                            if (stall >= 2 || !(cand > lo.alpha && cand < hi.alpha)) { cand = lo.alpha + T(0.5) * w; stall = 0; }
                            if (!(w > T(1e-3) * hi.alpha)) { alpha = lo.cost <= hi.cost ? lo.alpha : hi.alpha; break; }   // bracket collapsed
                        } else if (!(cand > p.alpha)) cand = T(2) * p.alpha;
                        p = ls_eval(m, E, lane, cand, g1, g2, dr, lr); it++;
#ifdef GO1_EMU_TRACE
                        if (lane == 0 && it > 10 && it < 24) std::printf("LSDBG it %d lo %.6g (%.3g) hi %.6g (%.3g) have %d cand %.8g -> d1 %.4g d2 %.4g cost %.4g gtol %.3g d10 %.3g\n", it, (double)lo.alpha, (double)lo.d1, (double)hi.alpha, (double)hi.d1, (int)have_hi, (double)cand, (double)p.d1, (double)p.d2, (double)p.cost, (double)gtol, (double)d10);
#endif
                    }
                } else {
                LSPoint<T> p0, p1, p2, pmid, p1n, p2n;
                p0 = ls_eval(m, E, lane, T(0), g1, g2, dr, lr); it++;
                // fp32: MuJoCo's absolute tolerance (1e-10 * |search|) is below the rounding noise of d1, the search would
                // run to the iteration cap; stop when the directional derivative has dropped by GO1_LS_REL instead.  The
                // Newton exit tests (gradient / decrement) are unchanged, so the optimum is the same.
                if (sizeof(T) == 4) gtol = N::max_(gtol, T(GO1_LS_REL) * N::abs_(p0.d1));
                p1 = ls_eval(m, E, lane, p0.alpha - p0.d1 / p0.d2, g1, g2, dr, lr); it++;
                if (p0.cost < p1.cost) p1 = p0;
                bool done = false;
                if (N::abs_(p1.d1) < gtol) { alpha = p1.alpha; done = true; }
                if (!done) {
                    const T dir = p1.d1 < 0 ? T(1) : T(-1);
                    bool p2update = false;
                    p2 = p1;
                    while (p1.d1 * dir <= -gtol && it < lsmax) {
                        p2 = p1; p2update = true;
                        p1 = ls_eval(m, E, lane, p1.alpha - p1.d1 / p1.d2, g1, g2, dr, lr); it++;
                        if (N::abs_(p1.d1) < gtol) { alpha = p1.alpha; done = true; break; }
                    }
                    if (!done && (it >= lsmax || !p2update)) { alpha = p1.alpha; done = true; }
                    if (!done) {
                        p2n = p1;
                        p1n = ls_eval(m, E, lane, p1.alpha - p1.d1 / p1.d2, g1, g2, dr, lr); it++;
                        while (it < lsmax && !done) {
                            pmid = ls_eval(m, E, lane, T(0.5) * (p1.alpha + p2.alpha), g1, g2, dr, lr); it++;
                            T bestc = 0; bool have = false; T besta = 0;
                            if (N::abs_(p1n.d1) < gtol) { have = true; bestc = p1n.cost; besta = p1n.alpha; }
                            if (N::abs_(p2n.d1) < gtol && (!have || p2n.cost < bestc)) { have = true; bestc = p2n.cost; besta = p2n.alpha; }
                            if (N::abs_(pmid.d1) < gtol && (!have || pmid.cost < bestc)) { have = true; bestc = pmid.cost; besta = pmid.alpha; }
                            if (have) { alpha = besta; done = true; break; }
                            int b1 = 0, b2 = 0;
                            const LSPoint<T> c0 = p1n, c1 = p2n, c2 = pmid;
#pragma unroll
                            for (int q = 0; q < 3; q++) {
                                const LSPoint<T>& cd = q == 0 ? c0 : (q == 1 ? c1 : c2);
                                if (p1.d1 < 0 && cd.d1 < 0 && p1.d1 < cd.d1) { p1 = cd; b1 = 1; }
                                else if (p1.d1 > 0 && cd.d1 > 0 && p1.d1 > cd.d1) { p1 = cd; b1 = 2; }
                            }
                            if (b1) { p1n = ls_eval(m, E, lane, p1.alpha - p1.d1 / p1.d2, g1, g2, dr, lr); it++; }
#pragma unroll
                            for (int q = 0; q < 3; q++) {
                                const LSPoint<T>& cd = q == 0 ? c0 : (q == 1 ? c1 : c2);
                                if (p2.d1 < 0 && cd.d1 < 0 && p2.d1 < cd.d1) { p2 = cd; b2 = 1; }
                                else if (p2.d1 > 0 && cd.d1 > 0 && p2.d1 > cd.d1) { p2 = cd; b2 = 2; }
                            }
                            if (b2) { p2n = ls_eval(m, E, lane, p2.alpha - p2.d1 / p2.d2, g1, g2, dr, lr); it++; }
                            if (!b1 && !b2) { alpha = pmid.cost < p0.cost ? pmid.alpha : T(0); done = true; }
                        }
                        if (!done) {
                            if (p1.cost <= p2.cost && p1.cost < p0.cost) alpha = p1.alpha;
                            else if (p2.cost <= p1.cost && p2.cost < p0.cost) alpha = p2.alpha;
                            else alpha = 0;
                        }
                    }
                }
                }
                GO1_TRACE_IT;
            }
        }
#ifdef GO1_EMU_TRACE
        if (lane == 0) std::printf("TRACE iter %d ncon %d lsevals %d alpha %.6g cost %.9g impr %.3g grad %.3g\n", iter, ncon, g_trace_it, (double)alpha, (double)cost, (double)improvement, (double)gradient);
#endif
#ifndef GO1_DBG_FIXED_ITERS
        if (alpha == 0) { conv = true; break; }
#endif
        a += alpha * s; Ma += alpha * Ms;
        oldcost = cost;
        } while (0);
    }
    if (!active) return;
    out.iters = iter; out.iters_total += iter;
    out.ncon = ncon;
    out.overflow += E.overflow;
    out.qacc_prev2 = out.qacc_prev; out.qacc_prev = out.qacc; out.nsub++;
    out.qacc = a; out.qfa = qfa; out.qfc = qfc;
    if (lane < 3) E.comw[lane] = E.com[lane] + E.qpos[lane];
    // mj_checkAcc
    if (__any_sync(FULL, lane < NDOF && !(N::abs_(a) <= T(1e10)))) { out.bad = 1; return; }
    // ------------------------------------------------------------------ mj_Euler (implicit joint damping) + mj_advance
    {
#pragma unroll 1
        for (int idx = lane; idx < NARROW; idx += 32) E.H[idx] = E.M[idx];
        __syncwarp();
        const T hd = legdof ? h * GO1_LEGF(damping, jl) : T(0);
        if (legdof) { const int k = jl / 3, i = jl - 3 * k; E.H[A_KK + 6 * k + SYM(i, i)] += hd; }
        __syncwarp();
        arrow_factor(m, E.H, lane);
        // MuJoCo solves (M + hB) a_d = qfrc_smooth + qfrc_constraint.  At the solver's optimum the right-hand side
        // equals M*qacc, hence a_d = qacc - h (M + hB)^-1 (B o qacc): same result in exact arithmetic, but it does
        // not amplify the residual gradient the fp32 Newton leaves along stiff contact directions.
        // The fp64 check build keeps MuJoCo's literal right-hand side (bit-level comparable with the oracle).
        T ad;
        if (sizeof(T) == 8) ad = arrow_solve(E, E.H, qfs + qfc, lane);
        else ad = a - arrow_solve(E, E.H, hd * a, lane);
        if (lane < NDOF) {
            const T v = E.qvel[lane] + h * ad;
            E.qvel[lane] = v;
            E.warm[lane] = a;
            if (lane >= 6) E.qpos[7 + jl] += h * v;
        }
        __syncwarp();
        if (lane == 0) {
#pragma unroll
            for (int c = 0; c < 3; c++) E.qpos[c] += h * E.qvel[c];
            const T w[3] = {E.qvel[3], E.qvel[4], E.qvel[5]};
            const T wn = N::sqrt_(dot3(w, w));
            if (wn >= N::MINVAL) {
                const T ang = h * wn, inv = T(1) / wn;
                T sn, cs;
                N::sincos_(T(0.5) * ang, &sn, &cs);
                const T qr[4] = {cs, w[0] * inv * sn, w[1] * inv * sn, w[2] * inv * sn};
                const T* q = E.qpos + 3;
                const T r0 = q[0] * qr[0] - q[1] * qr[1] - q[2] * qr[2] - q[3] * qr[3];
                const T r1 = q[0] * qr[1] + q[1] * qr[0] + q[2] * qr[3] - q[3] * qr[2];
                const T r2 = q[0] * qr[2] - q[1] * qr[3] + q[2] * qr[0] + q[3] * qr[1];
                const T r3 = q[0] * qr[3] + q[1] * qr[2] - q[2] * qr[1] + q[3] * qr[0];
                const T nn = T(1) / N::sqrt_(r0 * r0 + r1 * r1 + r2 * r2 + r3 * r3);
                E.qpos[3] = r0 * nn; E.qpos[4] = r1 * nn; E.qpos[5] = r2 * nn; E.qpos[6] = r3 * nn;
            }
        }
        __syncwarp();
    }
}

}  // namespace go1
