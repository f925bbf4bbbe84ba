// go1sim.cu -- kernels + C ABI (include/go1sim.h) of the batched Go1 simulator for sm_100a.
//
// One control step of every environment is ONE kernel launch (k_step): load the SoA state, run
// frame_skip physics substeps entirely in registers / local scratch, evaluate the fused
// observation / reward / cost / termination logic (go1_mujoco_env.py:483-532), auto-reset finished
// environments with Philox noise, store the state back.  Thread mapping: 4 lanes per environment
// (one per leg), 8 environments per warp, 16 per 64-thread block.
//
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo [-DGO1_REAL=double] (see build.py)
#ifdef GO1_EMU
#include "cuda_emu.h"   // tests/emu: CPU SIMT shim used only by the test-suite (never by the product)
#define GO1_DYN_SMEM(name) unsigned char* name = g_emu_block.dyn_smem
#define GO1_LAUNCH(kernel, grid, block, smem, stream, ...) emu_launch(kernel, grid, block, __VA_ARGS__)
#else
#include <cuda_runtime.h>
#define GO1_DYN_SMEM(name) extern __shared__ __align__(16) unsigned char name[]
#define GO1_LAUNCH(kernel, grid, block, smem, stream, ...) kernel<<<grid, block, smem, stream>>>(__VA_ARGS__)
#endif
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/go1sim.h"
#include "go1_physics.cuh"

#ifndef GO1_REAL
#define GO1_REAL float
#endif
#ifndef GO1_BLOCK
#define GO1_BLOCK 64
#endif
#ifndef GO1_MINBLOCKS
#define GO1_MINBLOCKS 7
#endif

namespace go1 {

typedef GO1_REAL real;

// ---------------------------------------------------------------------------------------------
// SoA state layout: real fields [field][N], int fields [field][N]
enum {
    S_QPOS = 0, S_QVEL = 19, S_WARM = 37, S_LASTACT = 55, S_AIR = 67, S_DEV = 71, S_VDES = 74, S_EPRET = 77, S_NREAL = 78
};
enum { I_STEP = 0, I_UNHEALTHY = 1, I_LASTCON = 2, I_RESETCNT = 3, I_BAD = 4, I_NINT = 5 };
// diag layout (real, per env, [field][N])
enum { D_QACC = 0, D_QFA = 18, D_QFC = 36, D_CFRC = 54, D_TRUNKZ = 138, D_COM = 139, D_NREAL = 142 };
enum { DI_NCON = 0, DI_ITER = 1, DI_ITERTOT = 2, DI_OVERFLOW = 3, DI_NINT = 4 };
// contact capture: per env, per lane, CMAX entries: gid, dim, dist, pos3, force6
enum { DC_STRIDE = 12 };

struct EnvCfg {
    int frame_skip, max_episode_steps, max_unhealthy_steps, solver_iterations;
    real rand_power, fallen_prob, vdes_min[3], vdes_max[3], reward_w[6], cost_w[11];
    real healthy_z[2], healthy_roll[2], healthy_pitch[2], sigma, obs_clip, tolerance;
};

template <typename T>
struct KParams {
    const DevModel<T>* model;
    EnvCfg cfg;
    int n_envs;
    T* sreal; int* sint;
    const float* actions;          // [N][12] or null
    int action_mode;               // 0 buffer, 1 random, 2 zero
    const uint8_t* reset_mask;     // reset-only launches
    int reset_only;
    float* obs; float* rew; uint8_t* done; uint8_t* trunc; float* term_obs; float* info;
    T* diag; int* diag_i; T* diag_con;   // nullable
    double* stats;                 // nullable: 8 doubles
    unsigned long long seed, env_offset, launch_counter;
};

// ---------------------------------------------------------------------------------------------
// Philox4x32-10
struct U4 { unsigned x, y, z, w; };
__device__ __forceinline__ U4 philox(unsigned long long seed, unsigned c0, unsigned c1, unsigned c2, unsigned c3) {
    unsigned k0 = (unsigned)seed, k1 = (unsigned)(seed >> 32);
#pragma unroll
    for (int i = 0; i < 10; i++) {
        unsigned hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        unsigned hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        unsigned n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    return U4{c0, c1, c2, c3};
}
__device__ __forceinline__ float u01(unsigned u) { return (float)(u >> 8) * (1.0f / 16777216.0f); }          // [0,1)
__device__ __forceinline__ float u01_open(unsigned u) { return ((float)(u >> 8) + 0.5f) * (1.0f / 16777216.0f); }  // (0,1)

// ---------------------------------------------------------------------------------------------
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

template <typename T> struct EnvRegs {
    T qt[7], ql[3], vt[6], vl[3], wt[6], wl[3];
};
template <typename T> struct StepOut {   // what mjData holds after the last substep, as the env reads it
    T qacc_t[6], qacc_k[3], qfa[3], qfc_t[6], qfc_k[3];
    T W[4][6];        // per-body external contact wrench about the COM (own): slot 0 is this lane's share of the trunk's
    T trunk_z, com[3];
    int ncon, iters, iters_total, overflow, bad;
};

// emit an active contact
template <typename T> GO1_DEV void emit(Contacts<T>& C, const T* pos, T dist, int cls, int gid, const T* com) {
    if (C.n >= CMAX) { C.overflow++; return; }
    int c = C.n++;
    C.r[c][0] = pos[0] - com[0]; C.r[c][1] = pos[1] - com[1]; C.r[c][2] = pos[2] - com[2];
    C.dist[c] = dist; C.cls[c] = (short)cls; C.gid[c] = (short)gid;
}

// plane-vs-geom narrowphase for the geoms of one body slot (engine_collision_primitive.c rules).
// Positions are relative to the trunk origin; d0 = signed distance of that origin to the plane.
template <typename T>
GO1_DEV void collide_slot(const DevModel<T>& m, int leg, int slot, const T* pb, const T* Rb, T d0, const T* com, Contacts<T>& C) {
    typedef Num<T> N;
    const T* n = m.plane_frame;
    for (int gi = m.gstart[leg][slot]; gi < m.gstart[leg][slot + 1]; gi++) {
        const DevGeom<T>& g = m.geoms[leg][gi];
        const DevClass<T>& cl = m.cls[g.cls];
        T gp[3], ax[3];
        matvec3(gp, Rb, g.pos);
        gp[0] += pb[0]; gp[1] += pb[1]; gp[2] += pb[2];
        matvec3(ax, Rb, g.zax);
        const T dc = d0 + dot3(n, gp);
        const T inc = cl.incmargin, margin = cl.margin;
        if (g.type == G_SPHERE) {
            const T dist = dc - g.size[0];
            if (dist < inc) {
                T pos[3];
#pragma unroll
                for (int k = 0; k < 3; k++) pos[k] = gp[k] - n[k] * (g.size[0] + T(0.5) * dist);
                emit(C, pos, dist, g.cls, g.gid, com);
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
                    emit(C, pos, dist, g.cls, g.gid, com);
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
                    emit(C, pos, dA, g.cls, g.gid, com);
                }
                const T dB = dc - prjaxis + prjvec;
                if (dB <= margin && dB < inc) {
                    T pos[3];
#pragma unroll
                    for (int k = 0; k < 3; k++) pos[k] = gp[k] + vec[k] - ax[k] - n[k] * dB * T(0.5);
                    emit(C, pos, dB, g.cls, g.gid, com);
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
                        emit(C, pos, dT, g.cls, g.gid, com);
                    }
                }
            }
        } else {  // box
            T xw[3], yl[3], yw[3];
            matvec3(xw, Rb, g.xax);
            cross3(yl, g.zax, g.xax);
            matvec3(yw, Rb, yl);
            int cnt = 0;
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
                    emit(C, pos, dist, g.cls, g.gid, com);
                }
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// one physics substep (mj_step) for this lane's share of the environment
template <typename T>
__device__ __noinline__ void substep(const DevModel<T>& m, const EnvCfg& cfg, int leg, unsigned mask, EnvRegs<T>& S,
                                     const T* ctrl, StepOut<T>& out, bool last, bool capture, Contacts<T>& C) {
    typedef Num<T> N;
    Lane<T> L;
    const T h = m.timestep;
    // ------------------------------------------------------------------ mj_kinematics
    {
        T nq = N::sqrt_(S.qt[3] * S.qt[3] + S.qt[4] * S.qt[4] + S.qt[5] * S.qt[5] + S.qt[6] * S.qt[6]);
        if (nq < N::MINVAL) { S.qt[3] = 1; S.qt[4] = S.qt[5] = S.qt[6] = 0; }
        else { T inv = T(1) / nq; S.qt[3] *= inv; S.qt[4] *= inv; S.qt[5] *= inv; S.qt[6] *= inv; }
    }
    quat_to_mat(L.R0, S.qt + 3);
    L.plane_d0 = m.plane_frame[0] * (S.qt[0] - m.plane_pos[0]) + m.plane_frame[1] * (S.qt[1] - m.plane_pos[1]) +
                 m.plane_frame[2] * (S.qt[2] - m.plane_pos[2]);
    out.trunk_z = S.qt[2];
    T axw[3][3], xip[3][3];
    T msum[3] = {0, 0, 0};
#pragma unroll
    for (int i = 0; i < 3; i++) {
        const T* Rp = i == 0 ? L.R0 : L.R[i - 1];
        T t[3];
        matvec3(t, Rp, m.body_pos[leg][i]);
        if (i == 0) { L.p[0][0] = t[0]; L.p[0][1] = t[1]; L.p[0][2] = t[2]; }
        else { L.p[i][0] = L.p[i - 1][0] + t[0]; L.p[i][1] = L.p[i - 1][1] + t[1]; L.p[i][2] = L.p[i - 1][2] + t[2]; }
        const T* u = m.jnt_axis[leg][i];
        T s, c;
        N::sincos_(S.ql[i] - m.qpos0[leg][i], &s, &c);
        const T oc = 1 - c;
        T Rl[9] = {c + oc * u[0] * u[0], oc * u[0] * u[1] - s * u[2], oc * u[0] * u[2] + s * u[1],
                   oc * u[0] * u[1] + s * u[2], c + oc * u[1] * u[1], oc * u[1] * u[2] - s * u[0],
                   oc * u[0] * u[2] - s * u[1], oc * u[1] * u[2] + s * u[0], c + oc * u[2] * u[2]};
#pragma unroll
        for (int r = 0; r < 3; r++)
#pragma unroll
            for (int cc = 0; cc < 3; cc++) L.R[i][3 * r + cc] = Rp[3 * r] * Rl[cc] + Rp[3 * r + 1] * Rl[3 + cc] + Rp[3 * r + 2] * Rl[6 + cc];
        matvec3(axw[i], Rp, u);
        matvec3(t, L.R[i], m.ipos[leg][i]);
        xip[i][0] = L.p[i][0] + t[0]; xip[i][1] = L.p[i][1] + t[1]; xip[i][2] = L.p[i][2] + t[2];
        msum[0] += m.mass[leg][i] * xip[i][0]; msum[1] += m.mass[leg][i] * xip[i][1]; msum[2] += m.mass[leg][i] * xip[i][2];
    }
    T xip0[3];
    matvec3(xip0, L.R0, m.t_ipos);
    // ------------------------------------------------------------------ mj_comPos
    {
        const T inv = T(1) / m.total_mass;
#pragma unroll
        for (int k = 0; k < 3; k++) L.com[k] = (tile_sum(msum[k], mask) + m.t_mass * xip0[k]) * inv;
    }
    T ci[3][10], ci0[10];
    {
        T dif[3];
#pragma unroll
        for (int i = 0; i < 3; i++) {
            dif[0] = xip[i][0] - L.com[0]; dif[1] = xip[i][1] - L.com[1]; dif[2] = xip[i][2] - L.com[2];
            cinert_of(ci[i], L.R[i], m.Ib[leg][i], m.mass[leg][i], dif);
            T off[3] = {L.com[0] - L.p[i][0], L.com[1] - L.p[i][1], L.com[2] - L.p[i][2]};
            L.ca[i][0] = axw[i][0]; L.ca[i][1] = axw[i][1]; L.ca[i][2] = axw[i][2];
            cross3(L.cl[i], axw[i], off);
        }
        dif[0] = xip0[0] - L.com[0]; dif[1] = xip0[1] - L.com[1]; dif[2] = xip0[2] - L.com[2];
        cinert_of(ci0, L.R0, m.t_Ib, m.t_mass, dif);
#pragma unroll
        for (int j = 0; j < 3; j++) {
            T col[3] = {L.R0[j], L.R0[3 + j], L.R0[6 + j]};
            cross3(L.tl[j], col, L.com);   // offset = com - trunk origin (= 0)
        }
    }
    // ------------------------------------------------------------------ mj_collision (plane vs geoms)
    C.n = 0; C.overflow = 0;
    {
        T p0[3] = {0, 0, 0};
        C.start[0] = 0;
        collide_slot(m, leg, 0, p0, L.R0, L.plane_d0, L.com, C);
        C.start[1] = C.n;
        collide_slot(m, leg, 1, L.p[0], L.R[0], L.plane_d0, L.com, C);
        C.start[2] = C.n;
        collide_slot(m, leg, 2, L.p[1], L.R[1], L.plane_d0, L.com, C);
        C.start[3] = C.n;
        collide_slot(m, leg, 3, L.p[2], L.R[2], L.plane_d0, L.com, C);
        C.start[4] = C.n;
    }
    // ------------------------------------------------------------------ mj_crb -> block-arrow M
    {
        T crb[3][10], crb0[10];
#pragma unroll
        for (int e = 0; e < 10; e++) { crb[2][e] = ci[2][e]; crb[1][e] = ci[1][e] + crb[2][e]; crb[0][e] = ci[0][e] + crb[1][e]; }
#pragma unroll
        for (int e = 0; e < 10; e++) crb0[e] = ci0[e] + tile_sum(crb[0][e], mask);
#pragma unroll
        for (int i = 0; i < 3; i++) {
            T cd[6] = {L.ca[i][0], L.ca[i][1], L.ca[i][2], L.cl[i][0], L.cl[i][1], L.cl[i][2]};
            T buf[6];
            mul_inert(buf, crb[i], cd);
#pragma unroll
            for (int j = 0; j <= i; j++) L.M.kk[SYM(i, j)] = dot3(L.ca[j], buf) + dot3(L.cl[j], buf + 3);
            L.M.kk[SYM(i, i)] += m.armature[leg][i];
#pragma unroll
            for (int j = 0; j < 3; j++) {
                L.M.tk[j][i] = buf[3 + j];
                T col[3] = {L.R0[j], L.R0[3 + j], L.R0[6 + j]};
                L.M.tk[3 + j][i] = dot3(col, buf) + dot3(L.tl[j], buf + 3);
            }
        }
        T bt[6][6];
#pragma unroll
        for (int j = 0; j < 3; j++) {
            T cdt[6] = {0, 0, 0, j == 0 ? T(1) : T(0), j == 1 ? T(1) : T(0), j == 2 ? T(1) : T(0)};
            mul_inert(bt[j], crb0, cdt);
            T cdr[6] = {L.R0[j], L.R0[3 + j], L.R0[6 + j], L.tl[j][0], L.tl[j][1], L.tl[j][2]};
            mul_inert(bt[3 + j], crb0, cdr);
        }
#pragma unroll
        for (int r = 0; r < 6; r++)
#pragma unroll
            for (int c = 0; c <= r; c++) {
                T v;
                if (c < 3) v = bt[r][3 + c];
                else { T col[3] = {L.R0[c - 3], L.R0[3 + c - 3], L.R0[6 + c - 3]}; v = dot3(col, bt[r]) + dot3(L.tl[c - 3], bt[r] + 3); }
                L.M.tt[SYM(r, c)] = v;
            }
    }
    // ------------------------------------------------------------------ mj_comVel + mj_rne (bias) + passive + actuation
    T bias_t[6], bias_k[3];
    {
        T angw[3];
        matvec3(angw, L.R0, S.vt + 3);
#pragma unroll
        for (int c = 0; c < 3; c++) {
            L.cvel[0][c] = angw[c];
            L.cvel[0][3 + c] = S.vt[c] + L.tl[0][c] * S.vt[3] + L.tl[1][c] * S.vt[4] + L.tl[2][c] * S.vt[5];
        }
        T cacc[4][6];
        {
            T vx[3];
            cross3(vx, S.vt, angw);
            cacc[0][0] = cacc[0][1] = cacc[0][2] = 0;
#pragma unroll
            for (int c = 0; c < 3; c++) cacc[0][3 + c] = -m.gravity[c] + vx[c];
        }
#pragma unroll
        for (int i = 0; i < 3; i++) {
            T cd[6] = {L.ca[i][0], L.ca[i][1], L.ca[i][2], L.cl[i][0], L.cl[i][1], L.cl[i][2]};
            T cdd[6];
            cross_motion(cdd, L.cvel[i], cd);
#pragma unroll
            for (int c = 0; c < 6; c++) { L.cvel[i + 1][c] = L.cvel[i][c] + cd[c] * S.vl[i]; cacc[i + 1][c] = cacc[i][c] + cdd[c] * S.vl[i]; }
        }
        T f[4][6];
#pragma unroll
        for (int b = 0; b < 4; b++) {
            const T* I = b == 0 ? ci0 : ci[b - 1];
            T t1[6], t2[6], t3[6];
            mul_inert(t1, I, cacc[b]);
            mul_inert(t2, I, L.cvel[b]);
            cross_force(t3, L.cvel[b], t2);
#pragma unroll
            for (int c = 0; c < 6; c++) f[b][c] = t1[c] + t3[c];
        }
#pragma unroll
        for (int c = 0; c < 6; c++) { f[2][c] += f[3][c]; f[1][c] += f[2][c]; }
#pragma unroll
        for (int c = 0; c < 6; c++) f[0][c] += tile_sum(f[1][c], mask);
#pragma unroll
        for (int i = 0; i < 3; i++) bias_k[i] = dot3(L.ca[i], f[i + 1]) + dot3(L.cl[i], f[i + 1] + 3);
#pragma unroll
        for (int j = 0; j < 3; j++) {
            bias_t[j] = f[0][3 + j];
            T col[3] = {L.R0[j], L.R0[3 + j], L.R0[6 + j]};
            bias_t[3 + j] = dot3(col, f[0]) + dot3(L.tl[j], f[0] + 3);
        }
    }
#pragma unroll
    for (int i = 0; i < 3; i++) {
        T u = ctrl[i];
        if (m.act_climited[leg][i]) u = N::min_(m.act_chi[leg][i], N::max_(m.act_clo[leg][i], u));
        const T gear = m.act_gear[leg][i];
        T force = m.act_gain[leg][i] * u;
        if (m.act_bias[leg][i]) force += m.act_b0[leg][i] + m.act_b1[leg][i] * (gear * S.ql[i]) + m.act_b2[leg][i] * (gear * S.vl[i]);
        if (m.act_flimited[leg][i]) force = N::min_(m.act_fhi[leg][i], N::max_(m.act_flo[leg][i], force));
        L.qfa[i] = gear * force;
        L.qfs_k[i] = -m.damping[leg][i] * S.vl[i] - bias_k[i] + L.qfa[i];
    }
#pragma unroll
    for (int j = 0; j < 6; j++) L.qfs_t[j] = -bias_t[j];
    // ------------------------------------------------------------------ qacc_smooth = M^-1 qfrc_smooth
    ArrowFactor<T> FM;
    arrow_factor(FM, L.M, mask);
#pragma unroll
    for (int j = 0; j < 6; j++) L.as_t[j] = L.qfs_t[j];
#pragma unroll
    for (int i = 0; i < 3; i++) L.as_k[i] = L.qfs_k[i];
    arrow_solve(FM, L.as_t, L.as_k, mask);
    // ------------------------------------------------------------------ mj_makeConstraint / impedance / reference
    const T* F = m.plane_frame;
#pragma unroll
    for (int b = 0; b < 4; b++) {
        for (int c = C.start[b]; c < C.start[b + 1]; c++) {
            const DevClass<T>& cl = m.cls[C.cls[c]];
            const T pm = C.dist[c] - cl.incmargin;
            const T imp = impedance(cl.d0, cl.d1, cl.width, cl.mid, cl.power, pm);
            const T dA = (b == 0) ? m.t_invw[0] : m.invw[leg][b - 1][0];
            const T R0 = N::max_(N::MINVAL, (1 - imp) * dA / imp);
            C.D0[c] = T(1) / R0;
            T r[3] = {C.r[c][0], C.r[c][1], C.r[c][2]};
            T vp[3];
            cross3(vp, L.cvel[b], r);
            vp[0] += L.cvel[b][3]; vp[1] += L.cvel[b][4]; vp[2] += L.cvel[b][5];
#pragma unroll
            for (int j = 0; j < 3; j++) {
                C.aref[c][j] = -cl.B * dot3(F + 3 * j, vp);
                C.aref[c][3 + j] = -cl.B * dot3(F + 3 * j, L.cvel[b]);
            }
            C.aref[c][0] -= cl.K * imp * pm;
        }
    }
#pragma unroll
    for (int i = 0; i < 3; i++) {
        L.fr_aref[i] = -m.Bf[leg][i] * S.vl[i];
        L.lim_s[i] = 0; L.lim_D[i] = 0; L.lim_aref[i] = 0;
        if (m.limited[leg][i]) {
            const T margin = m.lim_margin[leg][i];
            const T dlo = S.ql[i] - m.lim_lo[leg][i], dhi = m.lim_hi[leg][i] - S.ql[i];
            T pos = 0, sg = 0;
            if (dlo < margin) { pos = dlo; sg = 1; }
            else if (dhi < margin) { pos = dhi; sg = -1; }
            if (sg != 0) {
                const T* si = m.lim_imp[leg][i];
                const T imp = impedance(si[0], si[1], si[2], si[3], si[4], pos - margin);
                const T R = N::max_(N::MINVAL, (1 - imp) * m.dof_invw[leg][i] / imp);
                L.lim_s[i] = sg; L.lim_D[i] = T(1) / R;
                L.lim_aref[i] = -m.lim_B[leg][i] * (sg * S.vl[i]) - m.lim_K[leg][i] * imp * (pos - margin);
            }
        }
    }
    const int ncon_tile = tile_sum_i(C.n, mask);
    // ------------------------------------------------------------------ mj_fwdConstraint: warmstart + Newton
    T at[6], ak[3];
    T W[4][6], K[4][21], hd[3], fkd[3];
    {
        T Mt[6], Mk[3];
        arrow_mul(L.M, S.wt, S.wl, Mt, Mk, mask);
        T cw = eval_rows<T, false>(m, L, C, leg, S.wt, S.wl, W, K, hd, fkd, false);
        T gw = 0;
#pragma unroll
        for (int i = 0; i < 3; i++) gw += (Mk[i] - L.qfs_k[i]) * (S.wl[i] - L.as_k[i]);
        if (leg == 0) {
#pragma unroll
            for (int j = 0; j < 6; j++) gw += (Mt[j] - L.qfs_t[j]) * (S.wt[j] - L.as_t[j]);
        }
        cw = tile_sum(cw + T(0.5) * gw, mask);
        T cs = tile_sum(eval_rows<T, false>(m, L, C, leg, L.as_t, L.as_k, W, K, hd, fkd, false), mask);
        const bool use_smooth = cw > cs;
#pragma unroll
        for (int j = 0; j < 6; j++) at[j] = use_smooth ? L.as_t[j] : S.wt[j];
#pragma unroll
        for (int i = 0; i < 3; i++) ak[i] = use_smooth ? L.as_k[i] : S.wl[i];
    }
    T Mat[6], Mak[3], qfc_t[6], qfc_k[3], grad_t[6], grad_k[3];
    T cost;
    const T scale = T(1) / (m.meaninertia * T(18));
    const T tol = cfg.tolerance;
    const int maxiter = cfg.solver_iterations;
    arrow_mul(L.M, at, ak, Mat, Mak, mask);
    auto evaluate = [&](bool final_capture) {
        hd[0] = hd[1] = hd[2] = 0;
        T c = eval_rows<T, true>(m, L, C, leg, at, ak, W, K, hd, fkd, final_capture);
        if (last) {
#pragma unroll
            for (int b = 0; b < 4; b++)
#pragma unroll
                for (int e = 0; e < 6; e++) out.W[b][e] = W[b][e];
        }
        body_force_to_joint(L, W, qfc_t, qfc_k, mask);
        T g = 0;
#pragma unroll
        for (int i = 0; i < 3; i++) { qfc_k[i] += fkd[i]; g += (Mak[i] - L.qfs_k[i]) * (ak[i] - L.as_k[i]); }
        if (leg == 0) {
#pragma unroll
            for (int j = 0; j < 6; j++) g += (Mat[j] - L.qfs_t[j]) * (at[j] - L.as_t[j]);
        }
        cost = tile_sum(c + T(0.5) * g, mask);
#pragma unroll
        for (int j = 0; j < 6; j++) grad_t[j] = Mat[j] - L.qfs_t[j] - qfc_t[j];
#pragma unroll
        for (int i = 0; i < 3; i++) grad_k[i] = Mak[i] - L.qfs_k[i] - qfc_k[i];
    };
    evaluate(capture);
    int iter = 0;
    T improvement = 0, gradient = 0;
    while (iter < maxiter) {
        // MuJoCo's exit test (after >= 1 move): gradient small, or improvement small.  The measured
        // improvement is a difference of O(cost) numbers and is pure rounding noise in fp32 once it drops
        // below ~1e-7*cost, so "improvement small" is only trusted when the Newton decrement -g.s/2 (the
        // predicted improvement of the next move, free of cancellation) is small too; see DESIGN.md.
        if (iter > 0 && gradient < tol) break;
        // ---- Hessian H = M + J'DJ (+ cone blocks), block-arrow; search = -H^-1 grad
        Arrow<T> H = L.M;
#pragma unroll
        for (int i = 0; i < 3; i++) H.kk[SYM(i, i)] += hd[i];
        if (ncon_tile > 0) {
#pragma unroll
            for (int e = 0; e < 21; e++) { K[2][e] += K[3][e]; K[1][e] += K[2][e]; }
#pragma unroll
            for (int e = 0; e < 21; e++) K[0][e] = tile_sum(K[0][e] + K[1][e], mask);
#pragma unroll
            for (int i = 0; i < 3; i++) {
                T cd[6] = {L.ca[i][0], L.ca[i][1], L.ca[i][2], L.cl[i][0], L.cl[i][1], L.cl[i][2]};
                T u[6];
                sym6_mul(u, K[i + 1], cd);
#pragma unroll
                for (int j = 0; j <= i; j++) H.kk[SYM(i, j)] += dot3(L.ca[j], u) + dot3(L.cl[j], u + 3);
#pragma unroll
                for (int j = 0; j < 3; j++) {
                    H.tk[j][i] += u[3 + j];
                    T col[3] = {L.R0[j], L.R0[3 + j], L.R0[6 + j]};
                    H.tk[3 + j][i] += dot3(col, u) + dot3(L.tl[j], u + 3);
                }
            }
#pragma unroll
            for (int c = 0; c < 6; c++) {
                T cd[6];
                if (c < 3) { cd[0] = cd[1] = cd[2] = 0; cd[3] = c == 0; cd[4] = c == 1; cd[5] = c == 2; }
                else { cd[0] = L.R0[c - 3]; cd[1] = L.R0[3 + c - 3]; cd[2] = L.R0[6 + c - 3]; cd[3] = L.tl[c - 3][0]; cd[4] = L.tl[c - 3][1]; cd[5] = L.tl[c - 3][2]; }
                T u[6];
                sym6_mul(u, K[0], cd);
#pragma unroll
                for (int r = c; r < 6; r++) {
                    T v;
                    if (r < 3) v = u[3 + r];
                    else { T col[3] = {L.R0[r - 3], L.R0[3 + r - 3], L.R0[6 + r - 3]}; v = dot3(col, u) + dot3(L.tl[r - 3], u + 3); }
                    H.tt[SYM(r, c)] += v;
                }
            }
        }
        ArrowFactor<T> FH;
        arrow_factor(FH, H, mask);
        T st[6], sk[3];
#pragma unroll
        for (int j = 0; j < 6; j++) st[j] = grad_t[j];
#pragma unroll
        for (int i = 0; i < 3; i++) sk[i] = grad_k[i];
        arrow_solve(FH, st, sk, mask);
#pragma unroll
        for (int j = 0; j < 6; j++) st[j] = -st[j];
#pragma unroll
        for (int i = 0; i < 3; i++) sk[i] = -sk[i];
        if (iter > 0 && improvement < tol) {
            T dec = grad_k[0] * sk[0] + grad_k[1] * sk[1] + grad_k[2] * sk[2];
            dec = tile_sum(dec, mask);
#pragma unroll
            for (int j = 0; j < 6; j++) dec += grad_t[j] * st[j];
#ifdef GO1_EMU_TRACE
            if (leg == 0) printf("   decrement %.3g\n", (double)(T(-0.5) * scale * dec));
#endif
            if (T(-0.5) * scale * dec < tol) break;
        }
        // ---- exact line search (PrimalSearch)
        T Mvt[6], Mvk[3];
        arrow_mul(L.M, st, sk, Mvt, Mvk, mask);
        T alpha = 0;
        {
            T sn = sk[0] * sk[0] + sk[1] * sk[1] + sk[2] * sk[2];
            sn = tile_sum(sn, mask);
#pragma unroll
            for (int j = 0; j < 6; j++) sn += st[j] * st[j];
            const T snorm = N::sqrt_(sn);
            if (snorm >= N::MINVAL) {
                const T gtol = tol * m.ls_tolerance * snorm / scale;
                LSRows<T> rows;
                ls_prepare(m, L, C, leg, at, ak, st, sk, rows);
                // Gauss polynomial, distributed: leg parts on every lane, trunk part on lane 0
                T g0 = 0, g1 = 0, g2 = 0;
#pragma unroll
                for (int i = 0; i < 3; i++) {
                    g0 += T(0.5) * (Mak[i] - L.qfs_k[i]) * (ak[i] - L.as_k[i]);
                    g1 += sk[i] * (Mak[i] - L.qfs_k[i]);
                    g2 += T(0.5) * sk[i] * Mvk[i];
                }
                if (leg == 0) {
#pragma unroll
                    for (int j = 0; j < 6; j++) {
                        g0 += T(0.5) * (Mat[j] - L.qfs_t[j]) * (at[j] - L.as_t[j]);
                        g1 += st[j] * (Mat[j] - L.qfs_t[j]);
                        g2 += T(0.5) * st[j] * Mvt[j];
                    }
                }
                auto ls_eval = [&](LSPoint<T>& p) {
                    T c = p.alpha * (g1 + p.alpha * g2), d1 = 2 * p.alpha * g2 + g1, d2 = 2 * g2;   // cost relative to alpha = 0
                    ls_eval_lane(m, L, C, leg, rows, p.alpha, c, d1, d2);
                    p.cost = tile_sum(c, mask); p.d1 = tile_sum(d1, mask); p.d2 = tile_sum(d2, mask);
                    if (p.d2 <= 0) p.d2 = N::MINVAL;
                };
                const int lsmax = m.ls_iterations;
                int it = 0;
                LSPoint<T> p0, p1, p2, pmid, p1n, p2n;
                p0.alpha = 0; ls_eval(p0); it++;
                p1.alpha = p0.alpha - p0.d1 / p0.d2; ls_eval(p1); it++;
                if (p0.cost < p1.cost) p1 = p0;
                bool done = false;
                if (N::abs_(p1.d1) < gtol) { alpha = p1.alpha; done = true; }
                if (!done) {
                    const T dir = p1.d1 < 0 ? T(1) : T(-1);
                    bool p2update = false;
                    p2 = p1;
                    while (p1.d1 * dir <= -gtol && it < lsmax) {
                        p2 = p1; p2update = true;
                        p1.alpha -= p1.d1 / p1.d2; ls_eval(p1); it++;
                        if (N::abs_(p1.d1) < gtol) { alpha = p1.alpha; done = true; break; }
                    }
                    if (!done && (it >= lsmax || !p2update)) { alpha = p1.alpha; done = true; }
                    if (!done) {
                        p2n = p1;
                        p1n.alpha = p1.alpha - p1.d1 / p1.d2; ls_eval(p1n); it++;
                        while (it < lsmax && !done) {
                            pmid.alpha = T(0.5) * (p1.alpha + p2.alpha); ls_eval(pmid); it++;
                            // candidates: p1n, p2n, pmid
                            T bestc = 0; bool have = false; T besta = 0;
                            if (N::abs_(p1n.d1) < gtol) { have = true; bestc = p1n.cost; besta = p1n.alpha; }
                            if (N::abs_(p2n.d1) < gtol && (!have || p2n.cost < bestc)) { have = true; bestc = p2n.cost; besta = p2n.alpha; }
                            if (N::abs_(pmid.d1) < gtol && (!have || pmid.cost < bestc)) { have = true; bestc = pmid.cost; besta = pmid.alpha; }
                            if (have) { alpha = besta; done = true; break; }
                            // update brackets
                            int b1 = 0, b2 = 0;
                            {
                                const LSPoint<T> cand[3] = {p1n, p2n, pmid};
#pragma unroll
                                for (int q = 0; q < 3; q++) {
                                    if (p1.d1 < 0 && cand[q].d1 < 0 && p1.d1 < cand[q].d1) { p1 = cand[q]; b1 = 1; }
                                    else if (p1.d1 > 0 && cand[q].d1 > 0 && p1.d1 > cand[q].d1) { p1 = cand[q]; b1 = 2; }
                                }
                                if (b1) { p1n.alpha = p1.alpha - p1.d1 / p1.d2; ls_eval(p1n); it++; }
#pragma unroll
                                for (int q = 0; q < 3; q++) {
                                    if (p2.d1 < 0 && cand[q].d1 < 0 && p2.d1 < cand[q].d1) { p2 = cand[q]; b2 = 1; }
                                    else if (p2.d1 > 0 && cand[q].d1 > 0 && p2.d1 > cand[q].d1) { p2 = cand[q]; b2 = 2; }
                                }
                                if (b2) { p2n.alpha = p2.alpha - p2.d1 / p2.d2; ls_eval(p2n); it++; }
                            }
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
        }
#ifdef GO1_EMU_TRACE
        if (leg == 0 && alpha == 0) printf("   alpha==0 exit\n");
#endif
        if (alpha == 0) break;
#pragma unroll
        for (int j = 0; j < 6; j++) { at[j] += alpha * st[j]; Mat[j] += alpha * Mvt[j]; }
#pragma unroll
        for (int i = 0; i < 3; i++) { ak[i] += alpha * sk[i]; Mak[i] += alpha * Mvk[i]; }
        const T oldcost = cost;
        evaluate(capture);
        T gn = grad_k[0] * grad_k[0] + grad_k[1] * grad_k[1] + grad_k[2] * grad_k[2];
        gn = tile_sum(gn, mask);
#pragma unroll
        for (int j = 0; j < 6; j++) gn += grad_t[j] * grad_t[j];
        improvement = scale * (oldcost - cost); gradient = scale * N::sqrt_(gn);
        iter++;
#ifdef GO1_EMU_TRACE
        if (leg == 0) printf("   it %d alpha %.6g cost %.9g impr %.3g grad %.3g\n", iter, (double)alpha, (double)cost, (double)improvement, (double)gradient);
#endif
    }
    out.iters = iter; out.iters_total += iter;
    out.ncon = ncon_tile;
    out.overflow += tile_sum_i(C.overflow, mask);
#pragma unroll
    for (int j = 0; j < 6; j++) { out.qacc_t[j] = at[j]; out.qfc_t[j] = qfc_t[j]; }
#pragma unroll
    for (int i = 0; i < 3; i++) { out.qacc_k[i] = ak[i]; out.qfc_k[i] = qfc_k[i]; out.qfa[i] = L.qfa[i]; }
    out.com[0] = L.com[0] + S.qt[0]; out.com[1] = L.com[1] + S.qt[1]; out.com[2] = L.com[2] + S.qt[2];
    // mj_checkAcc
    {
        int badl = 0;
#pragma unroll
        for (int i = 0; i < 3; i++) badl |= !(N::abs_(ak[i]) <= T(1e10));
#pragma unroll
        for (int j = 0; j < 6; j++) badl |= !(N::abs_(at[j]) <= T(1e10));
        if (tile_sum_i(badl, mask)) { out.bad = 1; return; }
    }
    // ------------------------------------------------------------------ mj_Euler (implicit joint damping) + mj_advance
    {
        Arrow<T> A2 = L.M;
#pragma unroll
        for (int i = 0; i < 3; i++) A2.kk[SYM(i, i)] += h * m.damping[leg][i];
        ArrowFactor<T> F2;
        arrow_factor(F2, A2, mask);
        // MuJoCo solves (M + hB) a_d = qfrc_smooth + qfrc_constraint.  At the solver's optimum the right-hand side
        // equals M*qacc, hence a_d = qacc - h (M + hB)^-1 (B o qacc): same result in exact arithmetic, but it does
        // not amplify the residual gradient the fp32 Newton leaves along stiff contact directions.
        // The fp64 check build keeps MuJoCo's literal right-hand side (bit-level comparable with the oracle).
        T dt_[6], dk[3];
        if (sizeof(T) == 8) {
#pragma unroll
            for (int j = 0; j < 6; j++) dt_[j] = L.qfs_t[j] + qfc_t[j];
#pragma unroll
            for (int i = 0; i < 3; i++) dk[i] = L.qfs_k[i] + qfc_k[i];
            arrow_solve(F2, dt_, dk, mask);
        } else {
#pragma unroll
            for (int j = 0; j < 6; j++) dt_[j] = 0;
#pragma unroll
            for (int i = 0; i < 3; i++) dk[i] = h * m.damping[leg][i] * ak[i];
            arrow_solve(F2, dt_, dk, mask);
#pragma unroll
            for (int j = 0; j < 6; j++) dt_[j] = at[j] - dt_[j];
#pragma unroll
            for (int i = 0; i < 3; i++) dk[i] = ak[i] - dk[i];
        }
#pragma unroll
        for (int j = 0; j < 6; j++) S.vt[j] += h * dt_[j];
#pragma unroll
        for (int i = 0; i < 3; i++) { S.vl[i] += h * dk[i]; S.ql[i] += h * S.vl[i]; }
#pragma unroll
        for (int c = 0; c < 3; c++) S.qt[c] += h * S.vt[c];
        T w[3] = {S.vt[3], S.vt[4], S.vt[5]};
        T wn = N::sqrt_(dot3(w, w));
        if (wn >= N::MINVAL) {
            const T ang = h * wn, inv = T(1) / wn;
            T s, c;
            N::sincos_(T(0.5) * ang, &s, &c);
            const T qr[4] = {c, w[0] * inv * s, w[1] * inv * s, w[2] * inv * s};
            const T* q = S.qt + 3;
            T r0 = q[0] * qr[0] - q[1] * qr[1] - q[2] * qr[2] - q[3] * qr[3];
            T r1 = q[0] * qr[1] + q[1] * qr[0] + q[2] * qr[3] - q[3] * qr[2];
            T r2 = q[0] * qr[2] - q[1] * qr[3] + q[2] * qr[0] + q[3] * qr[1];
            T r3 = q[0] * qr[3] + q[1] * qr[2] - q[2] * qr[1] + q[3] * qr[0];
            const T nn = T(1) / N::sqrt_(r0 * r0 + r1 * r1 + r2 * r2 + r3 * r3);
            S.qt[3] = r0 * nn; S.qt[4] = r1 * nn; S.qt[5] = r2 * nn; S.qt[6] = r3 * nn;
        }
#pragma unroll
        for (int j = 0; j < 6; j++) S.wt[j] = at[j];
#pragma unroll
        for (int i = 0; i < 3; i++) S.wl[i] = ak[i];
    }
}

// ---------------------------------------------------------------------------------------------
// env logic helpers (go1_mujoco_env.py)
template <typename T> GO1_DEV void projected_gravity(const T* q, const T* grav, T* pg) {   // :596-608, :1017-1038
    const T w = q[0], x = q[1], y = q[2], z = q[3];
    T t2 = T(2) * (w * y - z * x);
    t2 = t2 > T(1) ? T(1) : t2; t2 = t2 < T(-1) ? T(-1) : t2;
    T e[3];
    if (sizeof(T) == 4) {
        e[0] = atan2f(T(2) * (w * x + y * z), T(1) - T(2) * (x * x + y * y));
        e[1] = asinf(t2);
        e[2] = atan2f(T(2) * (w * z + x * y), T(1) - T(2) * (y * y + z * z));
    } else {
        e[0] = atan2((double)(T(2) * (w * x + y * z)), (double)(T(1) - T(2) * (x * x + y * y)));
        e[1] = asin((double)t2);
        e[2] = atan2((double)(T(2) * (w * z + x * y)), (double)(T(1) - T(2) * (y * y + z * z)));
    }
    const T d = grav[0] * e[0] + grav[1] * e[1] + grav[2] * e[2];
    T v[3] = {d * e[0], d * e[1], d * e[2]};
    const T n = Num<T>::sqrt_(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
    if (n == 0) { pg[0] = v[0]; pg[1] = v[1]; pg[2] = v[2]; }
    else { pg[0] = v[0] / n; pg[1] = v[1] / n; pg[2] = v[2] / n; }
}
template <typename T> GO1_DEV T clipv(T v, T lim) { return v > lim ? lim : (v < -lim ? -lim : v); }

// writes this lane's share of the 48-d observation (:915-947)
template <typename T>
GO1_DEV void write_obs(const DevModel<T>& m, const EnvCfg& cfg, int leg, const EnvRegs<T>& S, const T* vdes, const T* last_act,
                       const T* pg, float* o) {
    const T lim = cfg.obs_clip;
    if (leg == 0) { for (int c = 0; c < 3; c++) o[c] = (float)clipv(S.vt[c] * T(2), lim); }
    else if (leg == 1) { for (int c = 0; c < 3; c++) o[3 + c] = (float)clipv(S.vt[3 + c] * T(0.25), lim); }
    else if (leg == 2) { for (int c = 0; c < 3; c++) o[6 + c] = (float)clipv(pg[c], lim); }
    else { for (int c = 0; c < 3; c++) o[9 + c] = (float)clipv(vdes[c] * T(2), lim); }
#pragma unroll
    for (int i = 0; i < 3; i++) {
        o[12 + 3 * leg + i] = (float)clipv(S.ql[i] - m.key_qpos[7 + 3 * leg + i], lim);
        o[24 + 3 * leg + i] = (float)clipv(S.vl[i] * T(0.05), lim);
        o[36 + 3 * leg + i] = (float)clipv(last_act[i], lim);
    }
}

template <typename T> struct Episode {
    T last_act[3], air, dev[3], vdes[3], ep_ret;
    int step, unhealthy, lastcon, resetcnt;
};

// reset_model (:949-1001) with Philox noise; replicated values are computed identically on the 4 lanes
template <typename T>
GO1_DEV void reset_env(const DevModel<T>& m, const EnvCfg& cfg, int leg, unsigned long long seed, unsigned long long gid,
                       EnvRegs<T>& S, Episode<T>& E) {
    E.resetcnt += 1;
    const U4 r0 = philox(seed, (unsigned)gid, (unsigned)(gid >> 32), (unsigned)E.resetcnt, 0x52535430u);
    const U4 r1 = philox(seed, (unsigned)gid, (unsigned)(gid >> 32), (unsigned)E.resetcnt, 0x52535431u);
#pragma unroll
    for (int c = 0; c < 7; c++) S.qt[c] = m.key_qpos[c];
#pragma unroll
    for (int i = 0; i < 3; i++) S.ql[i] = m.key_qpos[7 + 3 * leg + i];
    if (u01(r0.x) < cfg.fallen_prob) {
        const float lo = 3.14159265358979f / 2.1f, hi = 3.14159265358979f / 1.5f;
        const float ang = lo + (hi - lo) * u01(r0.y);
        float s, c;
        sincosf(0.5f * ang, &s, &c);
        S.qt[3] = c; S.qt[4] = 0; S.qt[5] = 0; S.qt[6] = 0;
        if (u01(r0.z) < 0.5f) S.qt[4] = s; else S.qt[5] = s;
        S.qt[2] = T(0.1);
    }
    if (cfg.rand_power > 0) {
        const U4 rn = philox(seed, (unsigned)gid, (unsigned)(gid >> 32), (unsigned)E.resetcnt, 0x52535440u + leg);
        const float a0 = sqrtf(-2.0f * logf(u01_open(rn.x))), a1 = sqrtf(-2.0f * logf(u01_open(rn.z)));
        float s0, c0, s1, c1;
        sincospif(2.0f * u01(rn.y), &s0, &c0);
        sincospif(2.0f * u01(rn.w), &s1, &c1);
        const float nz[3] = {a0 * c0, a0 * s0, a1 * c1};
        (void)s1;
#pragma unroll
        for (int i = 0; i < 3; i++) {
            T q = S.ql[i] + T(0.1) * cfg.rand_power * T(nz[i]);
            q = q < m.lim_lo[leg][i] ? m.lim_lo[leg][i] : q;
            q = q > m.lim_hi[leg][i] ? m.lim_hi[leg][i] : q;
            S.ql[i] = q;
        }
    }
#pragma unroll
    for (int j = 0; j < 6; j++) { S.vt[j] = 0; S.wt[j] = 0; }
#pragma unroll
    for (int i = 0; i < 3; i++) { S.vl[i] = 0; S.wl[i] = 0; E.last_act[i] = 0; E.dev[i] = 0; }
    const float uu[3] = {u01(r0.w), u01(r1.x), u01(r1.y)};
#pragma unroll
    for (int c = 0; c < 3; c++) E.vdes[c] = cfg.vdes_min[c] + (cfg.vdes_max[c] - cfg.vdes_min[c]) * T(uu[c]);
    E.air = 0; E.step = 0; E.unhealthy = 0; E.lastcon = 0; E.ep_ret = 0;
}

// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(GO1_BLOCK, GO1_MINBLOCKS) k_step(const KParams<T> P) {
    GO1_DYN_SMEM(smem_raw);
    __shared__ double sstat[8];
    DevModel<T>& m = *reinterpret_cast<DevModel<T>*>(smem_raw);
    if (threadIdx.x < 8) sstat[threadIdx.x] = 0;
    {
        const int* src = reinterpret_cast<const int*>(P.model);
        int* dst = reinterpret_cast<int*>(smem_raw);
        for (int i = threadIdx.x; i < (int)(sizeof(DevModel<T>) / 4); i += blockDim.x) dst[i] = src[i];
    }
    __syncthreads();
    const int env = blockIdx.x * (GO1_BLOCK / NLEG) + (threadIdx.x >> 2);
    const int leg = threadIdx.x & 3;
    const unsigned mask = 0xFu << ((threadIdx.x & 31) & ~3);
    if (env >= P.n_envs) return;
    const EnvCfg& cfg = P.cfg;
    const int N = P.n_envs;
    const unsigned long long gid = P.env_offset + (unsigned long long)env;
    T* sr = P.sreal + env;
    int* si = P.sint + env;
    // ---- load state
    EnvRegs<T> S;
    Episode<T> E;
#pragma unroll
    for (int c = 0; c < 7; c++) S.qt[c] = sr[(size_t)(S_QPOS + c) * N];
#pragma unroll
    for (int j = 0; j < 6; j++) { S.vt[j] = sr[(size_t)(S_QVEL + j) * N]; S.wt[j] = sr[(size_t)(S_WARM + j) * N]; }
#pragma unroll
    for (int i = 0; i < 3; i++) {
        S.ql[i] = sr[(size_t)(S_QPOS + 7 + 3 * leg + i) * N];
        S.vl[i] = sr[(size_t)(S_QVEL + 6 + 3 * leg + i) * N];
        S.wl[i] = sr[(size_t)(S_WARM + 6 + 3 * leg + i) * N];
        E.last_act[i] = sr[(size_t)(S_LASTACT + 3 * leg + i) * N];
        E.dev[i] = sr[(size_t)(S_DEV + i) * N];
        E.vdes[i] = sr[(size_t)(S_VDES + i) * N];
    }
    E.air = sr[(size_t)(S_AIR + leg) * N];
    E.ep_ret = sr[(size_t)S_EPRET * N];
    E.step = si[(size_t)I_STEP * N]; E.unhealthy = si[(size_t)I_UNHEALTHY * N]; E.lastcon = si[(size_t)I_LASTCON * N];
    E.resetcnt = si[(size_t)I_RESETCNT * N];
    int badflag = si[(size_t)I_BAD * N];

    auto store_state = [&]() {
        if (leg == 0) {
#pragma unroll
            for (int c = 0; c < 7; c++) sr[(size_t)(S_QPOS + c) * N] = S.qt[c];
#pragma unroll
            for (int j = 0; j < 6; j++) { sr[(size_t)(S_QVEL + j) * N] = S.vt[j]; sr[(size_t)(S_WARM + j) * N] = S.wt[j]; }
#pragma unroll
            for (int i = 0; i < 3; i++) { sr[(size_t)(S_DEV + i) * N] = E.dev[i]; sr[(size_t)(S_VDES + i) * N] = E.vdes[i]; }
            sr[(size_t)S_EPRET * N] = E.ep_ret;
            si[(size_t)I_STEP * N] = E.step; si[(size_t)I_UNHEALTHY * N] = E.unhealthy; si[(size_t)I_LASTCON * N] = E.lastcon;
            si[(size_t)I_RESETCNT * N] = E.resetcnt; si[(size_t)I_BAD * N] = badflag;
        }
#pragma unroll
        for (int i = 0; i < 3; i++) {
            sr[(size_t)(S_QPOS + 7 + 3 * leg + i) * N] = S.ql[i];
            sr[(size_t)(S_QVEL + 6 + 3 * leg + i) * N] = S.vl[i];
            sr[(size_t)(S_WARM + 6 + 3 * leg + i) * N] = S.wl[i];
            sr[(size_t)(S_LASTACT + 3 * leg + i) * N] = E.last_act[i];
        }
        sr[(size_t)(S_AIR + leg) * N] = E.air;
    };

    if (P.reset_only) {
        if (P.reset_mask && !P.reset_mask[env]) return;
        reset_env(m, cfg, leg, P.seed, gid, S, E);
        badflag = 0;
        if (P.obs) {
            T pg[3];
            projected_gravity(S.qt + 3, m.gravity, pg);
            write_obs(m, cfg, leg, S, E.vdes, E.last_act, pg, P.obs + (size_t)env * GO1_OBS_DIM);
        }
        store_state();
        return;
    }

    // ---- action for this leg's three joints
    T act[3];
    if (P.action_mode == 0) {
#pragma unroll
        for (int i = 0; i < 3; i++) act[i] = (T)P.actions[(size_t)env * GO1_ACT_DIM + 3 * leg + i];
    } else if (P.action_mode == 1) {
        const U4 r = philox(P.seed ^ 0xA5A5A5A55A5A5A5Aull, (unsigned)gid, (unsigned)(gid >> 32), (unsigned)P.launch_counter, 0x41435400u + leg);
        const float u[3] = {u01(r.x), u01(r.y), u01(r.z)};
#pragma unroll
        for (int i = 0; i < 3; i++) act[i] = m.ctrl_lo[3 * leg + i] + (m.ctrl_hi[3 * leg + i] - m.ctrl_lo[3 * leg + i]) * T(u[i]);
    } else {
#pragma unroll
        for (int i = 0; i < 3; i++) act[i] = 0;
    }

    // ---- do_simulation: frame_skip x mj_step, then mj_rnePostConstraint (:481)
    E.step += 1;
    StepOut<T> out;
    out.iters_total = 0; out.overflow = 0; out.bad = 0; out.ncon = 0; out.iters = 0;
    Contacts<T> C;
    const bool capture = P.diag_con != nullptr;
    typedef Num<T> NN;
    {
        // mj_checkPos / mj_checkVel: a non-finite or exploded state is never simulated
        int pre = 0;
#pragma unroll
        for (int i = 0; i < 3; i++) pre |= !(NN::abs_(S.ql[i]) <= T(1e10)) | !(NN::abs_(S.vl[i]) <= T(1e10)) | !(NN::abs_(S.wl[i]) <= T(1e10));
#pragma unroll
        for (int c = 0; c < 7; c++) pre |= !(NN::abs_(S.qt[c]) <= T(1e10));
#pragma unroll
        for (int j = 0; j < 6; j++) pre |= !(NN::abs_(S.vt[j]) <= T(1e10)) | !(NN::abs_(S.wt[j]) <= T(1e10));
        if (tile_sum_i(pre, mask)) out.bad = 1;
    }
    for (int s = 0; s < cfg.frame_skip && !out.bad; s++) {
        substep(m, cfg, leg, mask, S, act, out, s == cfg.frame_skip - 1, capture && s == cfg.frame_skip - 1, C);
    }
    if (out.bad) {   // mjData of a skipped step: nothing the reward code may read is left uninitialised
        C.n = 0;
#pragma unroll
        for (int b = 0; b < 4; b++)
#pragma unroll
            for (int e = 0; e < 6; e++) out.W[b][e] = 0;
#pragma unroll
        for (int i = 0; i < 3; i++) { out.qacc_k[i] = 0; out.qfa[i] = 0; out.qfc_k[i] = 0; }
#pragma unroll
        for (int j = 0; j < 6; j++) { out.qacc_t[j] = 0; out.qfc_t[j] = 0; }
        out.trunk_z = 0; out.com[0] = out.com[1] = out.com[2] = 0;
    }
    // ---- health (:433-471) on the post-step state; non-finite states also raise the sticky bad flag
    int nonfinite = 0;
#pragma unroll
    for (int i = 0; i < 3; i++) nonfinite |= !(NN::abs_(S.ql[i]) <= T(1e10)) | !(NN::abs_(S.vl[i]) <= T(1e10));
    if (leg == 0) {
#pragma unroll
        for (int c = 0; c < 7; c++) nonfinite |= !(NN::abs_(S.qt[c]) <= T(1e10));
#pragma unroll
        for (int j = 0; j < 6; j++) nonfinite |= !(NN::abs_(S.vt[j]) <= T(1e10));
    }
    nonfinite = tile_sum_i(nonfinite, mask) | out.bad;
    int reason = 0;
    if (nonfinite) reason = 1;
    else if (!(cfg.healthy_z[0] <= S.qt[2] && S.qt[2] <= cfg.healthy_z[1])) reason = 2;
    else if (!(cfg.healthy_roll[0] <= S.qt[4] && S.qt[4] <= cfg.healthy_roll[1])) reason = 3;
    else if (!(cfg.healthy_pitch[0] <= S.qt[5] && S.qt[5] <= cfg.healthy_pitch[1])) reason = 4;
    const bool ok = reason == 0;

    // ---- observation (:483, uses a_{t-1})
    T pg[3];
    projected_gravity(S.qt + 3, m.gravity, pg);
    float* obs_row = P.obs + (size_t)env * GO1_OBS_DIM;
    float* tobs_row = P.term_obs ? P.term_obs + (size_t)env * GO1_OBS_DIM : nullptr;

    // ---- rewards (:791-913)
    const T* rw = cfg.reward_w; const T* cw = cfg.cost_w;
    const T ex = E.vdes[0] - S.vt[0], ey = E.vdes[1] - S.vt[1], ez = E.vdes[2] - S.vt[5];
    T lin_r, ang_r;
    if (sizeof(T) == 4) { lin_r = expf(-(ex * ex + ey * ey) / cfg.sigma) * rw[0]; ang_r = expf(-(ez * ez) / cfg.sigma) * rw[1]; }
    else { lin_r = exp((double)(-(ex * ex + ey * ey) / cfg.sigma)) * rw[0]; ang_r = exp((double)(-(ez * ez) / cfg.sigma)) * rw[1]; }
    const T healthy_r = ok ? rw[2] : T(0);
    // feet air time (:633-668): this lane = one foot
    T air_l;
    {
        const T fmag = NN::sqrt_(dot6(out.W[3], out.W[3]));
        const int curr = fmag > T(1);
        const int filt = curr | ((E.lastcon >> leg) & 1);
        const T first = (E.air > 0 && filt) ? T(1) : T(0);
        E.air += T(0.02);
        const T ts = E.air - T(0.2) > 0 ? E.air - T(0.2) : T(0);
        air_l = ts * ts * first;
        if (filt) E.air = 0;
        int bits = curr << leg;
        bits |= __shfl_xor_sync(mask, bits, 1);
        bits |= __shfl_xor_sync(mask, bits, 2);
        E.lastcon = bits;
    }
    T air_r = tile_sum(air_l, mask);
    if (!(NN::sqrt_(E.vdes[0] * E.vdes[0] + E.vdes[1] * E.vdes[1]) > T(0.1))) air_r = 0;
    air_r *= rw[3];
    // recovery (:534-580)
    T dev[3] = {0, 0, 0};
    {
        const T z = S.qt[2], roll = S.qt[4], pitch = S.qt[5];
        if (!(cfg.healthy_z[0] <= z && z <= cfg.healthy_z[1])) dev[0] = NN::min_(NN::abs_(z - cfg.healthy_z[0]), NN::abs_(z - cfg.healthy_z[1]));
        if (!(cfg.healthy_roll[0] <= roll && roll <= cfg.healthy_roll[1])) dev[1] = NN::min_(NN::abs_(roll - cfg.healthy_roll[0]), NN::abs_(roll - cfg.healthy_roll[1]));
        if (!(cfg.healthy_pitch[0] <= pitch && pitch <= cfg.healthy_pitch[1])) dev[2] = NN::min_(NN::abs_(pitch - cfg.healthy_pitch[0]), NN::abs_(pitch - cfg.healthy_pitch[1]));
    }
    T rec = 0;
    if (ok) { E.dev[0] = E.dev[1] = E.dev[2] = 0; }
    else { rec = (E.dev[0] - dev[0]) + (E.dev[1] - dev[1]) + (E.dev[2] - dev[2]); E.dev[0] = dev[0]; E.dev[1] = dev[1]; E.dev[2] = dev[2]; }
    const T rec_r = rec * rw[4];
    T gu = 0;
    if (!ok) {
        T hr = out.trunk_z / cfg.healthy_z[0];
        hr = hr < 0 ? T(0) : (hr > 1 ? T(1) : hr);
        gu = hr * T(1.5) + (T(1) - (pg[0] * pg[0] + pg[1] * pg[1])) * T(0.5);
    }
    const T gu_r = gu * rw[5];
    const T rewards = lin_r + ang_r + healthy_r + air_r + rec_r + gu_r;
    // costs: per-lane partial sums over this leg's joints, reduced across the tile
    T s_tau = 0, s_rate = 0, s_jl = 0, s_jv = 0, s_acc = 0, s_djp = 0, s_coll = 0;
#pragma unroll
    for (int i = 0; i < 3; i++) {
        const int u = 3 * leg + i;
        s_tau += out.qfa[i] * out.qfa[i];
        const T dr = E.last_act[i] - act[i];
        s_rate += dr * dr;
        const T lo = m.soft_lo[u] - S.ql[i], hi = S.ql[i] - m.soft_hi[u];
        s_jl += (lo > 0 ? lo : T(0)) + (hi > 0 ? hi : T(0));
        s_jv += S.vl[i] * S.vl[i];
        s_acc += out.qacc_k[i] * out.qacc_k[i] / (NN::abs_(S.vl[i]) + T(1e-6));
        const T dj = S.ql[i] - m.key_ctrl[u];
        s_djp += dj * dj;
    }
    s_coll = dot6(out.W[1], out.W[1]) + dot6(out.W[2], out.W[2]);
    s_tau = tile_sum(s_tau, mask); s_rate = tile_sum(s_rate, mask); s_jl = tile_sum(s_jl, mask); s_jv = tile_sum(s_jv, mask);
    s_acc = tile_sum(s_acc, mask); s_djp = tile_sum(s_djp, mask); s_coll = tile_sum(s_coll, mask);
    const T disc = ok ? T(1) : T(0.1);
    const T ctrl_c = s_tau * disc * cw[0];
    const T vert_c = S.vt[2] * S.vt[2] * cw[1];
    const T xyang_c = (S.vt[3] * S.vt[3] + S.vt[4] * S.vt[4]) * cw[2];
    const T rate_c = s_rate * disc * cw[3];
    const T jl_c = s_jl * cw[4];
    const T jv_c = s_jv * cw[5];
    const T acc_c = s_acc * disc * cw[6];
    const T orient_c = (pg[0] * pg[0] + pg[1] * pg[1]) * cw[7];
    const T coll_c = (NN::sqrt_(s_coll) > T(0.1) ? T(1) : T(0)) * cw[8];
    const T djp_c = s_djp * cw[9];
    const T unh_c = (ok ? T(0) : T(1)) * cw[10];
    const T costs = ctrl_c + rate_c + vert_c + xyang_c + jl_c + jv_c + acc_c + coll_c + unh_c + orient_c + djp_c;
    T reward = rewards - costs;
    reward = reward > 0 ? reward : T(0);
    if (nonfinite) reward = 0;
    // ---- termination (:487-498)
    E.unhealthy = ok ? 0 : E.unhealthy + 1;
    const bool terminated = E.unhealthy > cfg.max_unhealthy_steps;
    const bool truncated = E.step >= cfg.max_episode_steps;
    const bool done = terminated || truncated || nonfinite;
    E.ep_ret += reward;
    const int ep_len = E.step;
    const T ep_ret = E.ep_ret;
    const float x_pos = (float)S.qt[0], y_pos = (float)S.qt[1];
    // ---- optional capture of mjData (tests)
    if (P.diag) {
        T* dg = P.diag + env;
        if (leg == 0) {
#pragma unroll
            for (int j = 0; j < 6; j++) { dg[(size_t)(D_QACC + j) * N] = out.qacc_t[j]; dg[(size_t)(D_QFA + j) * N] = 0; dg[(size_t)(D_QFC + j) * N] = out.qfc_t[j]; }
            dg[(size_t)D_TRUNKZ * N] = out.trunk_z;
#pragma unroll
            for (int c = 0; c < 3; c++) dg[(size_t)(D_COM + c) * N] = out.com[c];
#pragma unroll
            for (int c = 0; c < 6; c++) dg[(size_t)(D_CFRC + c) * N] = 0;   // world body
            int* di = P.diag_i + env;
            di[(size_t)DI_NCON * N] = out.ncon; di[(size_t)DI_ITER * N] = out.iters; di[(size_t)DI_ITERTOT * N] = out.iters_total;
            di[(size_t)DI_OVERFLOW * N] = out.overflow;
        }
#pragma unroll
        for (int i = 0; i < 3; i++) {
            dg[(size_t)(D_QACC + 6 + 3 * leg + i) * N] = out.qacc_k[i];
            dg[(size_t)(D_QFA + 6 + 3 * leg + i) * N] = out.qfa[i];
            dg[(size_t)(D_QFC + 6 + 3 * leg + i) * N] = out.qfc_k[i];
        }
#pragma unroll
        for (int c = 0; c < 6; c++) {
            const T tw = tile_sum(out.W[0][c], mask);
            if (leg == 0) dg[(size_t)(D_CFRC + 6 + c) * N] = tw;   // trunk = body 1
#pragma unroll
            for (int b = 1; b < 4; b++) dg[(size_t)(D_CFRC + 6 * (2 + 3 * leg + b - 1) + c) * N] = out.W[b][c];
        }
        if (P.diag_con) {
            T* dc = P.diag_con + ((size_t)env * NLEG + leg) * (CMAX * DC_STRIDE + 1);
            dc[0] = (T)C.n;
            for (int c = 0; c < C.n; c++) {
                T* e = dc + 1 + c * DC_STRIDE;
                e[0] = (T)C.gid[c]; e[1] = (T)m.cls[C.cls[c]].dim; e[2] = C.dist[c];
                e[3] = C.r[c][0] + out.com[0]; e[4] = C.r[c][1] + out.com[1]; e[5] = C.r[c][2] + out.com[2];
#pragma unroll
                for (int j = 0; j < 6; j++) e[6 + j] = C.frc[c][j];
            }
        }
    }
    // ---- outputs + auto-reset (SubprocVecEnv worker semantics)
    {
        // obs uses the PREVIOUS action a_{t-1} (:483 runs before :529)
        const T prev[3] = {E.last_act[0], E.last_act[1], E.last_act[2]};
        E.last_act[0] = act[0]; E.last_act[1] = act[1]; E.last_act[2] = act[2];    // :529
        if (done) {
            if (tobs_row) write_obs(m, cfg, leg, S, E.vdes, prev, pg, tobs_row);
        } else {
            write_obs(m, cfg, leg, S, E.vdes, prev, pg, obs_row);
        }
    }
    if (leg == 0) {
        P.rew[env] = (float)reward;
        P.done[env] = done ? 1 : 0;
        P.trunc[env] = (truncated && !terminated && !nonfinite) ? 1 : 0;
        if (P.info) {
            float* inf = P.info + (size_t)env * GO1_INFO_DIM;
            inf[0] = x_pos; inf[1] = y_pos; inf[2] = (float)lin_r; inf[3] = (float)(-ctrl_c); inf[4] = (float)healthy_r;
            inf[5] = (float)rec_r; inf[6] = (float)gu_r; inf[7] = (float)(-unh_c); inf[8] = (float)(-orient_c); inf[9] = (float)(-djp_c);
            inf[10] = (float)ep_ret; inf[11] = (float)ep_len; inf[12] = (float)reason; inf[13] = ok ? 1.f : 0.f;
            inf[14] = (float)out.iters_total; inf[15] = (float)out.ncon;
        }
    }
    if (P.stats) {
        // block-level reduction in shared memory, then at most 8 global atomics per block
        if (leg == 0) {
            atomicAdd(&sstat[0], (double)reward); atomicAdd(&sstat[7], 1.0);
            if (done) { atomicAdd(&sstat[1], 1.0); atomicAdd(&sstat[2], (double)ep_len); atomicAdd(&sstat[3], (double)ep_ret); }
            if (terminated) atomicAdd(&sstat[4], 1.0);
            if (truncated && !terminated) atomicAdd(&sstat[5], 1.0);
            if (nonfinite) atomicAdd(&sstat[6], 1.0);
        }
        __syncthreads();
        if (threadIdx.x < 8 && sstat[threadIdx.x] != 0.0) atomicAdd(P.stats + threadIdx.x, sstat[threadIdx.x]);
    }
    if (done) {
        reset_env(m, cfg, leg, P.seed, gid, S, E);
        badflag = nonfinite ? badflag + 1 : badflag;
        T pg2[3];
        projected_gravity(S.qt + 3, m.gravity, pg2);
        write_obs(m, cfg, leg, S, E.vdes, E.last_act, pg2, obs_row);
    }
    store_state();
}

// 48-64-64-12 tanh MLP + Gaussian sampling + clip: one thread per env (weights in shared memory).
// Round-1 version on the FP32 pipe; the tensor-core (tcgen05) version is tracked in DESIGN.md "next".
__global__ void __launch_bounds__(128) k_policy(const float* __restrict__ obs, float* __restrict__ act, int n, Go1Policy pol,
                                                 const float* lo, const float* hi, unsigned long long seed, unsigned long long counter,
                                                 unsigned long long env_offset) {
    __shared__ float w0[64 * 48], b0[64], w1[64 * 64], b1[64], w2[12 * 64], b2[12], ls[12], slo[12], shi[12];
    for (int i = threadIdx.x; i < 64 * 48; i += blockDim.x) w0[i] = pol.w0[i];
    for (int i = threadIdx.x; i < 64 * 64; i += blockDim.x) w1[i] = pol.w1[i];
    for (int i = threadIdx.x; i < 12 * 64; i += blockDim.x) w2[i] = pol.w2[i];
    for (int i = threadIdx.x; i < 64; i += blockDim.x) { b0[i] = pol.b0[i]; b1[i] = pol.b1[i]; }
    for (int i = threadIdx.x; i < 12; i += blockDim.x) { b2[i] = pol.b2[i]; ls[i] = pol.log_std ? pol.log_std[i] : -1e30f; slo[i] = lo[i]; shi[i] = hi[i]; }
    __syncthreads();
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    float x[48], h0[64], h1[64];
    for (int i = 0; i < 48; i++) x[i] = obs[(size_t)e * 48 + i];
    for (int o = 0; o < 64; o++) { float s = b0[o]; for (int i = 0; i < 48; i++) s += w0[o * 48 + i] * x[i]; h0[o] = tanhf(s); }
    for (int o = 0; o < 64; o++) { float s = b1[o]; for (int i = 0; i < 64; i++) s += w1[o * 64 + i] * h0[i]; h1[o] = tanhf(s); }
    const unsigned long long gid = env_offset + e;
    for (int o = 0; o < 12; o += 4) {
        const U4 r = philox(seed ^ 0x504F4C4943590000ull, (unsigned)gid, (unsigned)(gid >> 32), (unsigned)counter, 0x50000000u + o);
        const float a0 = sqrtf(-2.0f * logf(u01_open(r.x))), a1 = sqrtf(-2.0f * logf(u01_open(r.z)));
        float s0, c0, s1, c1;
        sincospif(2.0f * u01(r.y), &s0, &c0);
        sincospif(2.0f * u01(r.w), &s1, &c1);
        const float nz[4] = {a0 * c0, a0 * s0, a1 * c1, a1 * s1};
        for (int q = 0; q < 4; q++) {
            float s = b2[o + q];
            for (int i = 0; i < 64; i++) s += w2[(o + q) * 64 + i] * h1[i];
            if (pol.log_std) s += expf(ls[o + q]) * nz[q];
            s = fminf(shi[o + q], fmaxf(slo[o + q], s));
            act[(size_t)e * 12 + o + q] = s;
        }
    }
}

}  // namespace go1

// =============================================================================================
// host side / C ABI
using namespace go1;

struct Go1Sim {
    int n_envs = 0, device = 0;
    EnvCfg cfg;
    DevModel<double> hmodel;
    DevModel<real>* dmodel = nullptr;
    real* sreal = nullptr; int* sint = nullptr;
    real* diag = nullptr; int* diag_i = nullptr; real* diag_con = nullptr;
    bool diag_enabled = false;
    double* stats = nullptr;
    float *d_act = nullptr, *d_obs = nullptr, *d_rew = nullptr, *d_tobs = nullptr, *d_info = nullptr;
    uint8_t *d_done = nullptr, *d_trunc = nullptr, *d_mask = nullptr;
    float *h_act = nullptr, *h_obs = nullptr, *h_rew = nullptr, *h_tobs = nullptr, *h_info = nullptr;   // pinned staging
    uint8_t *h_done = nullptr, *h_trunc = nullptr, *h_mask = nullptr;
    float *d_lo = nullptr, *d_hi = nullptr;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    unsigned long long seed = 0, env_offset = 0, launch_counter = 0;
    float last_ms = 0; int last_launches = 0;
    std::string err;
};

static std::string g_create_err;

#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { h->err = std::string(#call) + ": " + cudaGetErrorString(e_); return GO1_ECUDA; } } while (0)

extern "C" {

void go1_default_config(Go1EnvConfig* c) {
    std::memset(c, 0, sizeof(*c));
    c->biped = 0; c->frame_skip = 10; c->max_episode_steps = 6000; c->max_unhealthy_steps = 750;
    c->rand_power = 0.0f; c->fallen_start_prob = 0.2f;
    c->desired_vel_min[0] = -0.5f; c->desired_vel_max[0] = 0.5f;
    const float rw[6] = {2.0f, 1.0f, 1.0f, 5.0f, 10.0f, 20.0f};
    const float cw[11] = {0.0002f, 2.0f, 0.05f, 0.01f, 10.0f, 0.01f, 2.0e-4f, 1.0f, 1.0f, 0.1f, 5.0f};
    std::memcpy(c->reward_w, rw, sizeof(rw)); std::memcpy(c->cost_w, cw, sizeof(cw));
    c->healthy_z[0] = 0.22f; c->healthy_z[1] = 1.8f;
    c->healthy_roll[0] = -1.3962634015954636f; c->healthy_roll[1] = 1.3962634015954636f;
    c->healthy_pitch[0] = -3.14159265358979323846f; c->healthy_pitch[1] = 0.0f;
    c->tracking_sigma = 0.25f; c->obs_clip = 100.0f; c->solver_tolerance = -1.0f; c->solver_iterations = 0;
}

int go1_precision_bits(void) { return (int)sizeof(real) * 8; }
const char* go1_version(void) { return "go1sim 0.1 (sm_100a)"; }
const char* go1_last_error(go1_handle h) { return h ? h->err.c_str() : g_create_err.c_str(); }

static void fill_cfg(EnvCfg& k, const Go1EnvConfig& c, const DevModel<double>& M) {
    k.frame_skip = c.frame_skip; k.max_episode_steps = c.max_episode_steps; k.max_unhealthy_steps = c.max_unhealthy_steps;
    k.solver_iterations = c.solver_iterations > 0 ? c.solver_iterations : M.iterations;
    k.rand_power = c.rand_power; k.fallen_prob = c.fallen_start_prob;
    for (int i = 0; i < 3; i++) { k.vdes_min[i] = c.desired_vel_min[i]; k.vdes_max[i] = c.desired_vel_max[i]; }
    // weights / ranges are the reference's fp64 literals: widen from the float config, but keep exact
    // double values for the defaults (0.0002f etc. are not representable)
    Go1EnvConfig def; go1_default_config(&def);
    const double rwd[6] = {2.0, 1.0, 1.0, 5.0, 10.0, 20.0};
    const double cwd[11] = {0.0002, 2.0, 0.05, 0.01, 10.0, 0.01, 2.0e-4, 1.0, 1.0, 0.1, 5.0};
    for (int i = 0; i < 6; i++) k.reward_w[i] = (real)(c.reward_w[i] == def.reward_w[i] ? rwd[i] : (double)c.reward_w[i]);
    for (int i = 0; i < 11; i++) k.cost_w[i] = (real)(c.cost_w[i] == def.cost_w[i] ? cwd[i] : (double)c.cost_w[i]);
    const double hz[2] = {0.22, 1.8}, hr[2] = {-1.3962634015954636, 1.3962634015954636}, hp[2] = {-3.14159265358979323846, 0.0};
    for (int i = 0; i < 2; i++) {
        k.healthy_z[i] = (real)(c.healthy_z[i] == def.healthy_z[i] ? hz[i] : (double)c.healthy_z[i]);
        k.healthy_roll[i] = (real)(c.healthy_roll[i] == def.healthy_roll[i] ? hr[i] : (double)c.healthy_roll[i]);
        k.healthy_pitch[i] = (real)(c.healthy_pitch[i] == def.healthy_pitch[i] ? hp[i] : (double)c.healthy_pitch[i]);
    }
    k.sigma = (real)c.tracking_sigma; k.obs_clip = (real)c.obs_clip;
    k.tolerance = (real)(c.solver_tolerance >= 0 ? (double)c.solver_tolerance : M.tolerance);
}

int go1_create(const void* model_blob, size_t nbytes, const Go1EnvConfig* cfg_in, int n_envs, int cuda_device, uint64_t seed,
               uint64_t env_id_offset, go1_handle* out) {
    if (!out || !model_blob || n_envs <= 0) { g_create_err = "go1_create: bad argument"; return GO1_EINVAL; }
    *out = nullptr;
    Go1EnvConfig cfg;
    if (cfg_in) cfg = *cfg_in; else go1_default_config(&cfg);
    if (cfg.biped) { g_create_err = "biped=1 is not implemented (DESIGN.md: next)"; return GO1_EUNSUPPORTED; }
    if (cfg.frame_skip <= 0 || cfg.max_episode_steps <= 0) { g_create_err = "bad config"; return GO1_EINVAL; }
    Go1Sim* h = new Go1Sim();
    if (!build_dev_model(model_blob, nbytes, h->hmodel, h->err)) { g_create_err = "model rejected: " + h->err; delete h; return GO1_EMODEL; }
    h->n_envs = n_envs; h->device = cuda_device; h->seed = seed; h->env_offset = env_id_offset;
    fill_cfg(h->cfg, cfg, h->hmodel);
    auto fail = [&](const char* what, cudaError_t e) { g_create_err = std::string(what) + ": " + cudaGetErrorString(e); go1_destroy(h); return GO1_ECUDA; };
    cudaError_t e;
    if ((e = cudaSetDevice(cuda_device)) != cudaSuccess) return fail("cudaSetDevice", e);
    if ((e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking)) != cudaSuccess) return fail("cudaStreamCreate", e);
    cudaEventCreate(&h->ev0); cudaEventCreate(&h->ev1);
    DevModel<real>* hm = new DevModel<real>();
    cast_dev_model(h->hmodel, *hm);
    if ((e = cudaMalloc(&h->dmodel, sizeof(DevModel<real>))) != cudaSuccess) { delete hm; return fail("cudaMalloc model", e); }
    e = cudaMemcpy(h->dmodel, hm, sizeof(DevModel<real>), cudaMemcpyHostToDevice);
    delete hm;
    if (e != cudaSuccess) return fail("cudaMemcpy model", e);
    const size_t N = (size_t)n_envs;
#define ALLOC(p, bytes) if ((e = cudaMalloc((void**)&(p), (bytes))) != cudaSuccess) return fail("cudaMalloc " #p, e); cudaMemset((p), 0, (bytes))
    ALLOC(h->sreal, sizeof(real) * S_NREAL * N); ALLOC(h->sint, sizeof(int) * I_NINT * N);
    ALLOC(h->stats, sizeof(double) * 8);
    ALLOC(h->d_act, sizeof(float) * 12 * N); ALLOC(h->d_obs, sizeof(float) * 48 * N); ALLOC(h->d_rew, sizeof(float) * N);
    ALLOC(h->d_tobs, sizeof(float) * 48 * N); ALLOC(h->d_info, sizeof(float) * GO1_INFO_DIM * N);
    ALLOC(h->d_done, N); ALLOC(h->d_trunc, N); ALLOC(h->d_mask, N);
    ALLOC(h->d_lo, sizeof(float) * 12); ALLOC(h->d_hi, sizeof(float) * 12);
#undef ALLOC
#define HALLOC(p, bytes) if ((e = cudaMallocHost((void**)&(p), (bytes))) != cudaSuccess) return fail("cudaMallocHost " #p, e)
    HALLOC(h->h_act, sizeof(float) * 12 * N); HALLOC(h->h_obs, sizeof(float) * 48 * N); HALLOC(h->h_rew, sizeof(float) * N);
    HALLOC(h->h_tobs, sizeof(float) * 48 * N); HALLOC(h->h_info, sizeof(float) * GO1_INFO_DIM * N);
    HALLOC(h->h_done, N); HALLOC(h->h_trunc, N); HALLOC(h->h_mask, N);
#undef HALLOC
    float lo[12], hi[12];
    for (int i = 0; i < 12; i++) { lo[i] = (float)h->hmodel.ctrl_lo[i]; hi[i] = (float)h->hmodel.ctrl_hi[i]; }
    cudaMemcpy(h->d_lo, lo, sizeof(lo), cudaMemcpyHostToDevice); cudaMemcpy(h->d_hi, hi, sizeof(hi), cudaMemcpyHostToDevice);
    if ((e = cudaFuncSetAttribute(k_step<real>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(DevModel<real>))) != cudaSuccess)
        return fail("cudaFuncSetAttribute", e);
    *out = h;
    int rc = go1_reset(h, nullptr, nullptr, nullptr);
    if (rc != GO1_OK) { g_create_err = h->err; go1_destroy(h); *out = nullptr; return rc; }
    if ((e = cudaStreamSynchronize(h->stream)) != cudaSuccess) { *out = nullptr; return fail("initial reset", e); }
    return GO1_OK;
}

int go1_destroy(go1_handle h) {
    if (!h) return GO1_EINVAL;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    void* dev[] = {h->dmodel, h->sreal, h->sint, h->diag, h->diag_i, h->diag_con, h->stats, h->d_act, h->d_obs, h->d_rew, h->d_tobs,
                   h->d_info, h->d_done, h->d_trunc, h->d_mask, h->d_lo, h->d_hi};
    for (void* p : dev) if (p) cudaFree(p);
    void* host[] = {h->h_act, h->h_obs, h->h_rew, h->h_tobs, h->h_info, h->h_done, h->h_trunc, h->h_mask};
    for (void* p : host) if (p) cudaFreeHost(p);
    if (h->ev0) cudaEventDestroy(h->ev0);
    if (h->ev1) cudaEventDestroy(h->ev1);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
    return GO1_OK;
}

int go1_num_envs(go1_handle h) { return h ? h->n_envs : GO1_EINVAL; }

int go1_action_bounds(go1_handle h, float* low12, float* high12) {
    if (!h || !low12 || !high12) return GO1_EINVAL;
    for (int i = 0; i < 12; i++) { low12[i] = (float)h->hmodel.ctrl_lo[i]; high12[i] = (float)h->hmodel.ctrl_hi[i]; }
    return GO1_OK;
}

static KParams<real> base_params(go1_handle h) {
    KParams<real> P;
    std::memset(&P, 0, sizeof(P));
    P.model = h->dmodel; P.cfg = h->cfg; P.n_envs = h->n_envs; P.sreal = h->sreal; P.sint = h->sint;
    P.seed = h->seed; P.env_offset = h->env_offset; P.launch_counter = h->launch_counter;
    if (h->diag_enabled) { P.diag = h->diag; P.diag_i = h->diag_i; P.diag_con = h->diag_con; }
    return P;
}
static int launch(go1_handle h, const KParams<real>& P, cudaStream_t s) {
    const int envs_per_block = GO1_BLOCK / NLEG;
    const int grid = (h->n_envs + envs_per_block - 1) / envs_per_block;
    GO1_LAUNCH(k_step<real>, grid, GO1_BLOCK, sizeof(DevModel<real>), s, P);
    CK(cudaGetLastError());
    return GO1_OK;
}

int go1_reset(go1_handle h, const uint8_t* mask_dev, float* obs_dev, void* stream) {
    if (!h) return GO1_EINVAL;
    CK(cudaSetDevice(h->device));
    KParams<real> P = base_params(h);
    P.reset_only = 1; P.reset_mask = mask_dev; P.obs = obs_dev;
    return launch(h, P, stream ? (cudaStream_t)stream : h->stream);
}

int go1_step(go1_handle h, const float* actions_dev, float* obs_dev, float* rew_dev, uint8_t* done_dev, uint8_t* trunc_dev,
             float* terminal_obs_dev, float* info_dev, void* stream) {
    if (!h || !actions_dev || !obs_dev || !rew_dev || !done_dev || !trunc_dev) { if (h) h->err = "go1_step: null buffer"; return GO1_EINVAL; }
    CK(cudaSetDevice(h->device));
    KParams<real> P = base_params(h);
    P.actions = actions_dev; P.action_mode = 0; P.obs = obs_dev; P.rew = rew_dev; P.done = done_dev; P.trunc = trunc_dev;
    P.term_obs = terminal_obs_dev; P.info = info_dev;
    h->launch_counter++;
    return launch(h, P, stream ? (cudaStream_t)stream : h->stream);
}

int go1_step_host(go1_handle h, const float* actions, float* obs, float* rew, uint8_t* done, uint8_t* trunc, float* terminal_obs, float* info) {
    if (!h || !actions || !obs || !rew || !done || !trunc) { if (h) h->err = "go1_step_host: null buffer"; return GO1_EINVAL; }
    CK(cudaSetDevice(h->device));
    const size_t N = (size_t)h->n_envs;
    std::memcpy(h->h_act, actions, sizeof(float) * 12 * N);
    CK(cudaMemcpyAsync(h->d_act, h->h_act, sizeof(float) * 12 * N, cudaMemcpyHostToDevice, h->stream));
    int rc = go1_step(h, h->d_act, h->d_obs, h->d_rew, h->d_done, h->d_trunc, h->d_tobs, info ? h->d_info : nullptr, h->stream);
    if (rc != GO1_OK) return rc;
    CK(cudaMemcpyAsync(h->h_obs, h->d_obs, sizeof(float) * 48 * N, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaMemcpyAsync(h->h_rew, h->d_rew, sizeof(float) * N, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaMemcpyAsync(h->h_done, h->d_done, N, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaMemcpyAsync(h->h_trunc, h->d_trunc, N, cudaMemcpyDeviceToHost, h->stream));
    if (terminal_obs) CK(cudaMemcpyAsync(h->h_tobs, h->d_tobs, sizeof(float) * 48 * N, cudaMemcpyDeviceToHost, h->stream));
    if (info) CK(cudaMemcpyAsync(h->h_info, h->d_info, sizeof(float) * GO1_INFO_DIM * N, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    std::memcpy(obs, h->h_obs, sizeof(float) * 48 * N); std::memcpy(rew, h->h_rew, sizeof(float) * N);
    std::memcpy(done, h->h_done, N); std::memcpy(trunc, h->h_trunc, N);
    if (terminal_obs) std::memcpy(terminal_obs, h->h_tobs, sizeof(float) * 48 * N);
    if (info) std::memcpy(info, h->h_info, sizeof(float) * GO1_INFO_DIM * N);
    return GO1_OK;
}

int go1_reset_host(go1_handle h, const uint8_t* mask, float* obs) {
    if (!h) return GO1_EINVAL;
    CK(cudaSetDevice(h->device));
    const size_t N = (size_t)h->n_envs;
    const uint8_t* dmask = nullptr;
    if (mask) {
        std::memcpy(h->h_mask, mask, N);
        CK(cudaMemcpyAsync(h->d_mask, h->h_mask, N, cudaMemcpyHostToDevice, h->stream));
        dmask = h->d_mask;
    }
    int rc = go1_reset(h, dmask, obs ? h->d_obs : nullptr, h->stream);
    if (rc != GO1_OK) return rc;
    if (obs) CK(cudaMemcpyAsync(h->h_obs, h->d_obs, sizeof(float) * 48 * N, cudaMemcpyDeviceToHost, h->stream));
    CK(cudaStreamSynchronize(h->stream));
    if (obs) {
        if (!mask) std::memcpy(obs, h->h_obs, sizeof(float) * 48 * N);
        else for (size_t i = 0; i < N; i++) if (mask[i]) std::memcpy(obs + 48 * i, h->h_obs + 48 * i, sizeof(float) * 48);
    }
    return GO1_OK;
}

int go1_rollout(go1_handle h, int n_steps, int action_mode, const Go1Policy* policy, Go1Stats* out, void* stream) {
    if (!h || n_steps < 0) return GO1_EINVAL;
    if (action_mode == GO1_ACTION_POLICY && !policy) { h->err = "go1_rollout: policy mode needs a policy"; return GO1_EINVAL; }
    CK(cudaSetDevice(h->device));
    cudaStream_t s = stream ? (cudaStream_t)stream : h->stream;
    CK(cudaMemsetAsync(h->stats, 0, sizeof(double) * 8, s));
    int launches = 0;
    CK(cudaEventRecord(h->ev0, s));
    for (int t = 0; t < n_steps; t++) {
        KParams<real> P = base_params(h);
        P.obs = h->d_obs; P.rew = h->d_rew; P.done = h->d_done; P.trunc = h->d_trunc; P.term_obs = nullptr; P.info = nullptr;
        P.stats = h->stats;
        if (action_mode == GO1_ACTION_POLICY) {
#ifdef GO1_EMU
            const int tb = 1;
#else
            const int tb = 128;
#endif
            const int grid = (h->n_envs + tb - 1) / tb;
            GO1_LAUNCH(k_policy, grid, tb, 0, s, h->d_obs, h->d_act, h->n_envs, *policy, h->d_lo, h->d_hi, h->seed, h->launch_counter, h->env_offset);
            launches++;
            P.actions = h->d_act; P.action_mode = 0;
        } else {
            P.action_mode = action_mode == GO1_ACTION_RANDOM ? 1 : 2;
        }
        h->launch_counter++;
        int rc = launch(h, P, s);
        if (rc != GO1_OK) return rc;
        launches++;
    }
    CK(cudaEventRecord(h->ev1, s));
    CK(cudaStreamSynchronize(s));
    CK(cudaEventElapsedTime(&h->last_ms, h->ev0, h->ev1));
    h->last_launches = launches;
    if (out) {
        double st[8];
        CK(cudaMemcpy(st, h->stats, sizeof(st), cudaMemcpyDeviceToHost));
        out->sum_reward = st[0]; out->n_episodes = st[1]; out->sum_episode_length = st[2]; out->sum_episode_return = st[3];
        out->n_terminated = st[4]; out->n_truncated = st[5]; out->n_bad_state = st[6]; out->env_steps = st[7];
    }
    return GO1_OK;
}

int go1_last_rollout_timing(go1_handle h, float* ms, int* launches) {
    if (!h) return GO1_EINVAL;
    if (ms) *ms = h->last_ms;
    if (launches) *launches = h->last_launches;
    return GO1_OK;
}

// state access: strided gathers through a small staging buffer (test/diagnostic path, synchronous)
static int copy_fields(go1_handle h, bool to_device, void* base, size_t elem, int nfields, int env, void* host) {
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    const size_t N = (size_t)h->n_envs;
    if (to_device) CK(cudaMemcpy2D((char*)base + elem * env, elem * N, host, elem, elem, nfields, cudaMemcpyHostToDevice));
    else CK(cudaMemcpy2D(host, elem, (char*)base + elem * env, elem * N, elem, nfields, cudaMemcpyDeviceToHost));
    return GO1_OK;
}

int go1_set_state(go1_handle h, int env, const Go1State* s) {
    if (!h || !s || env < 0 || env >= h->n_envs) return GO1_EINVAL;
    real r[S_NREAL]; int iv[I_NINT];
    for (int i = 0; i < 19; i++) r[S_QPOS + i] = (real)s->qpos[i];
    for (int i = 0; i < 18; i++) { r[S_QVEL + i] = (real)s->qvel[i]; r[S_WARM + i] = (real)s->qacc_warmstart[i]; }
    for (int i = 0; i < 12; i++) r[S_LASTACT + i] = (real)s->last_action[i];
    for (int i = 0; i < 4; i++) r[S_AIR + i] = (real)s->feet_air_time[i];
    for (int i = 0; i < 3; i++) { r[S_DEV + i] = (real)s->last_health_deviation[i]; r[S_VDES + i] = (real)s->desired_velocity[i]; }
    r[S_EPRET] = (real)s->episode_return;
    iv[I_STEP] = s->step; iv[I_UNHEALTHY] = s->unhealthy_steps;
    iv[I_LASTCON] = (s->last_contacts[0] ? 1 : 0) | (s->last_contacts[1] ? 2 : 0) | (s->last_contacts[2] ? 4 : 0) | (s->last_contacts[3] ? 8 : 0);
    iv[I_RESETCNT] = s->reset_count; iv[I_BAD] = s->bad_state;
    int rc = copy_fields(h, true, h->sreal, sizeof(real), S_NREAL, env, r);
    if (rc) return rc;
    return copy_fields(h, true, h->sint, sizeof(int), I_NINT, env, iv);
}

int go1_get_state(go1_handle h, int env, Go1State* s) {
    if (!h || !s || env < 0 || env >= h->n_envs) return GO1_EINVAL;
    real r[S_NREAL]; int iv[I_NINT];
    int rc = copy_fields(h, false, h->sreal, sizeof(real), S_NREAL, env, r);
    if (rc) return rc;
    rc = copy_fields(h, false, h->sint, sizeof(int), I_NINT, env, iv);
    if (rc) return rc;
    for (int i = 0; i < 19; i++) s->qpos[i] = r[S_QPOS + i];
    for (int i = 0; i < 18; i++) { s->qvel[i] = r[S_QVEL + i]; s->qacc_warmstart[i] = r[S_WARM + i]; }
    for (int i = 0; i < 12; i++) s->last_action[i] = r[S_LASTACT + i];
    for (int i = 0; i < 4; i++) { s->feet_air_time[i] = r[S_AIR + i]; s->last_contacts[i] = (iv[I_LASTCON] >> i) & 1; }
    for (int i = 0; i < 3; i++) { s->last_health_deviation[i] = r[S_DEV + i]; s->desired_velocity[i] = r[S_VDES + i]; }
    s->episode_return = r[S_EPRET];
    s->step = iv[I_STEP]; s->unhealthy_steps = iv[I_UNHEALTHY]; s->reset_count = iv[I_RESETCNT]; s->bad_state = iv[I_BAD];
    return GO1_OK;
}

int go1_enable_diag(go1_handle h, int enable) {
    if (!h) return GO1_EINVAL;
    CK(cudaSetDevice(h->device));
    if (enable && !h->diag) {
        const size_t N = (size_t)h->n_envs;
        CK(cudaMalloc((void**)&h->diag, sizeof(real) * D_NREAL * N)); CK(cudaMemset(h->diag, 0, sizeof(real) * D_NREAL * N));
        CK(cudaMalloc((void**)&h->diag_i, sizeof(int) * DI_NINT * N)); CK(cudaMemset(h->diag_i, 0, sizeof(int) * DI_NINT * N));
        const size_t cb = sizeof(real) * N * NLEG * (CMAX * DC_STRIDE + 1);
        CK(cudaMalloc((void**)&h->diag_con, cb)); CK(cudaMemset(h->diag_con, 0, cb));
    }
    h->diag_enabled = enable != 0;
    return GO1_OK;
}

int go1_get_diag(go1_handle h, int env, Go1Diag* d) {
    if (!h || !d || env < 0 || env >= h->n_envs) return GO1_EINVAL;
    if (!h->diag) { h->err = "diag capture not enabled (go1_enable_diag)"; return GO1_EINVAL; }
    real r[D_NREAL]; int iv[DI_NINT];
    int rc = copy_fields(h, false, h->diag, sizeof(real), D_NREAL, env, r);
    if (rc) return rc;
    rc = copy_fields(h, false, h->diag_i, sizeof(int), DI_NINT, env, iv);
    if (rc) return rc;
    for (int i = 0; i < 18; i++) { d->qacc[i] = r[D_QACC + i]; d->qfrc_actuator[i] = r[D_QFA + i]; d->qfrc_constraint[i] = r[D_QFC + i]; }
    for (int b = 0; b < 14; b++) for (int c = 0; c < 6; c++) d->cfrc_ext[b][c] = r[D_CFRC + 6 * b + c];
    d->trunk_xpos_z = r[D_TRUNKZ];
    for (int c = 0; c < 3; c++) d->subtree_com[c] = r[D_COM + c];
    d->ncon = iv[DI_NCON]; d->solver_iter = iv[DI_ITER]; d->solver_iter_total = iv[DI_ITERTOT]; d->contact_overflow = iv[DI_OVERFLOW];
    return GO1_OK;
}

int go1_get_contacts(go1_handle h, int env, Go1Contact* buf, int cap, int* n) {
    if (!h || !buf || !n || env < 0 || env >= h->n_envs) return GO1_EINVAL;
    if (!h->diag_con) { h->err = "diag capture not enabled (go1_enable_diag)"; return GO1_EINVAL; }
    CK(cudaSetDevice(h->device));
    CK(cudaStreamSynchronize(h->stream));
    const size_t per = (size_t)NLEG * (CMAX * DC_STRIDE + 1);
    std::vector<real> tmp(per);
    CK(cudaMemcpy(tmp.data(), h->diag_con + (size_t)env * per, sizeof(real) * per, cudaMemcpyDeviceToHost));
    int cnt = 0;
    for (int k = 0; k < NLEG; k++) {
        const real* p = tmp.data() + (size_t)k * (CMAX * DC_STRIDE + 1);
        int nk = (int)p[0];
        for (int c = 0; c < nk && cnt < cap; c++, cnt++) {
            const real* e = p + 1 + c * DC_STRIDE;
            buf[cnt].geom = (int)e[0]; buf[cnt].dim = (int)e[1]; buf[cnt].dist = e[2];
            for (int a = 0; a < 3; a++) buf[cnt].pos[a] = e[3 + a];
            for (int a = 0; a < 6; a++) buf[cnt].force[a] = e[6 + a];
        }
    }
    *n = cnt;
    return GO1_OK;
}

int go1_set_attr_f32(go1_handle h, const char* name, float v) {
    if (!h || !name) return GO1_EINVAL;
    std::string s(name);
    if (s == "rand_power" || s == "_rand_power") h->cfg.rand_power = v;
    else if (s == "fallen_start_prob") h->cfg.fallen_prob = v;
    else if (s.rfind("desired_vel_min_", 0) == 0 && s.size() == 17 && s[16] >= 'x' && s[16] <= 'z') h->cfg.vdes_min[s[16] - 'x'] = v;
    else if (s.rfind("desired_vel_max_", 0) == 0 && s.size() == 17 && s[16] >= 'x' && s[16] <= 'z') h->cfg.vdes_max[s[16] - 'x'] = v;
    else if (s == "solver_tolerance") h->cfg.tolerance = (real)v;
    else { h->err = "unknown attribute " + s; return GO1_EINVAL; }
    return GO1_OK;
}

int go1_get_attr_f32(go1_handle h, const char* name, float* v) {
    if (!h || !name || !v) return GO1_EINVAL;
    std::string s(name);
    if (s == "rand_power" || s == "_rand_power") *v = (float)h->cfg.rand_power;
    else if (s == "fallen_start_prob") *v = (float)h->cfg.fallen_prob;
    else if (s.rfind("desired_vel_min_", 0) == 0 && s.size() == 17 && s[16] >= 'x' && s[16] <= 'z') *v = (float)h->cfg.vdes_min[s[16] - 'x'];
    else if (s.rfind("desired_vel_max_", 0) == 0 && s.size() == 17 && s[16] >= 'x' && s[16] <= 'z') *v = (float)h->cfg.vdes_max[s[16] - 'x'];
    else if (s == "solver_tolerance") *v = (float)h->cfg.tolerance;
    else { h->err = "unknown attribute " + s; return GO1_EINVAL; }
    return GO1_OK;
}

}  // extern "C"
