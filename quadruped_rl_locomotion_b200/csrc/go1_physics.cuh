// go1_physics.cuh -- device code of one physics substep (the mj_step pipeline the reference runs at
// go1_mujoco_env.py:481) for ONE environment spread over a 4-lane tile, one lane per leg.
//
// Mapping: lane k of the tile owns leg k (3 bodies, 3 dofs, its collision geoms, its
// contacts, its friction-loss / limit rows) and a replica of the trunk quantities.  The joint-space
// matrices (M, the Newton Hessian H = M + J'DJ, M + h*diag(damping)) are block-arrow
//     [ T (6x6)   C_0 .. C_3 ]
//     [ C_k'      L_k (3x3)  ]
// so every factorisation is four independent 3x3 Choleskys (one per lane) + a 6x6 Schur complement
// reduced across the tile with two xor-shuffles.  Contact Jacobians are never materialised: a floor
// contact on body b acts through the spatial motion of b about the robot COM, so J*a, J'*f and J'DJ
// are evaluated through per-body spatial vectors / 6x6 spatial matrices (a CRBA-like accumulation).
#pragma once
#include "go1_model.h"

namespace go1 {

#define GO1_DEV __device__ __forceinline__

template <typename T> struct Num;
template <> struct Num<float> {
    static GO1_DEV float sqrt_(float x) { return sqrtf(x); }
    static GO1_DEV void sincos_(float x, float* s, float* c) { sincosf(x, s, c); }
    static GO1_DEV float pow_(float a, float b) { return powf(a, b); }
    static GO1_DEV float abs_(float a) { return fabsf(a); }
    static GO1_DEV float max_(float a, float b) { return fmaxf(a, b); }
    static GO1_DEV float min_(float a, float b) { return fminf(a, b); }
    static constexpr float MINVAL = 1e-15f;
};
template <> struct Num<double> {
    static GO1_DEV double sqrt_(double x) { return sqrt(x); }
    static GO1_DEV void sincos_(double x, double* s, double* c) { sincos(x, s, c); }
    static GO1_DEV double pow_(double a, double b) { return pow(a, b); }
    static GO1_DEV double abs_(double a) { return fabs(a); }
    static GO1_DEV double max_(double a, double b) { return fmax(a, b); }
    static GO1_DEV double min_(double a, double b) { return fmin(a, b); }
    static constexpr double MINVAL = 1e-15;
};

// ---- tile (4-lane) reductions: butterfly => all four lanes get the bit-identical sum
template <typename T> GO1_DEV T tile_sum(T v, unsigned mask) {
    v += __shfl_xor_sync(mask, v, 1);
    v += __shfl_xor_sync(mask, v, 2);
    return v;
}
GO1_DEV int tile_sum_i(int v, unsigned mask) {
    v += __shfl_xor_sync(mask, v, 1);
    v += __shfl_xor_sync(mask, v, 2);
    return v;
}

// ---- small linear algebra
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

// y = K x for a packed symmetric 6x6
template <typename T> GO1_DEV void sym6_mul(T* y, const T* K, const T* x) {
#pragma unroll
    for (int r = 0; r < 6; r++) {
        T s = 0;
#pragma unroll
        for (int c = 0; c < 6; c++) s += K[SYM(r, c)] * x[c];
        y[r] = s;
    }
}
// K += w * g g'
template <typename T> GO1_DEV void sym6_rank1(T* K, const T* g, T w) {
#pragma unroll
    for (int r = 0; r < 6; r++) {
        T wr = w * g[r];
#pragma unroll
        for (int c = 0; c <= r; c++) K[SYM(r, c)] += wr * g[c];
    }
}

// ---- block-arrow matrix: T (packed sym 6x6, replicated), C (6x3, this lane's coupling), L (packed sym 3x3)
template <typename T> struct Arrow { T tt[21]; T tk[6][3]; T kk[6]; };
template <typename T> struct ArrowFactor { T l[6]; T y[6][3]; T ls[21]; };

template <typename T> GO1_DEV void arrow_factor(ArrowFactor<T>& F, const Arrow<T>& A, unsigned mask) {
    typedef Num<T> N;
    T l00 = N::sqrt_(N::max_(A.kk[0], N::MINVAL));
    T i00 = T(1) / l00;
    T l10 = A.kk[1] * i00, l20 = A.kk[3] * i00;
    T l11 = N::sqrt_(N::max_(A.kk[2] - l10 * l10, N::MINVAL));
    T i11 = T(1) / l11;
    T l21 = (A.kk[4] - l20 * l10) * i11;
    T l22 = N::sqrt_(N::max_(A.kk[5] - l20 * l20 - l21 * l21, N::MINVAL));
    T i22 = T(1) / l22;
    F.l[0] = i00; F.l[1] = l10; F.l[2] = i11; F.l[3] = l20; F.l[4] = l21; F.l[5] = i22;  // diagonals stored inverted
    T p[21];
#pragma unroll
    for (int j = 0; j < 6; j++) {
        T y0 = A.tk[j][0] * i00;
        T y1 = (A.tk[j][1] - y0 * l10) * i11;
        T y2 = (A.tk[j][2] - y0 * l20 - y1 * l21) * i22;
        F.y[j][0] = y0; F.y[j][1] = y1; F.y[j][2] = y2;
    }
#pragma unroll
    for (int r = 0; r < 6; r++)
#pragma unroll
        for (int c = 0; c <= r; c++) p[SYM(r, c)] = F.y[r][0] * F.y[c][0] + F.y[r][1] * F.y[c][1] + F.y[r][2] * F.y[c][2];
#pragma unroll
    for (int e = 0; e < 21; e++) p[e] = A.tt[e] - tile_sum(p[e], mask);
    // dense Cholesky of the 6x6 Schur complement, diagonal stored inverted
#pragma unroll
    for (int j = 0; j < 6; j++) {
        T d = p[SYM(j, j)];
#pragma unroll
        for (int k = 0; k < j; k++) d -= F.ls[SYM(j, k)] * F.ls[SYM(j, k)];
        T inv = T(1) / N::sqrt_(N::max_(d, N::MINVAL));
        F.ls[SYM(j, j)] = inv;
#pragma unroll
        for (int i = j + 1; i < 6; i++) {
            T s = p[SYM(i, j)];
#pragma unroll
            for (int k = 0; k < j; k++) s -= F.ls[SYM(i, k)] * F.ls[SYM(j, k)];
            F.ls[SYM(i, j)] = s * inv;
        }
    }
}
// solves A x = g in place (gt replicated trunk part, gk this lane's leg part)
template <typename T> GO1_DEV void arrow_solve(const ArrowFactor<T>& F, T* gt, T* gk, unsigned mask) {
    T z0 = gk[0] * F.l[0];
    T z1 = (gk[1] - F.l[1] * z0) * F.l[2];
    T z2 = (gk[2] - F.l[3] * z0 - F.l[4] * z1) * F.l[5];
    T x[6];
#pragma unroll
    for (int j = 0; j < 6; j++) x[j] = gt[j] - tile_sum(F.y[j][0] * z0 + F.y[j][1] * z1 + F.y[j][2] * z2, mask);
#pragma unroll
    for (int i = 0; i < 6; i++) {
        T s = x[i];
#pragma unroll
        for (int k = 0; k < i; k++) s -= F.ls[SYM(i, k)] * x[k];
        x[i] = s * F.ls[SYM(i, i)];
    }
#pragma unroll
    for (int i = 5; i >= 0; i--) {
        T s = x[i];
#pragma unroll
        for (int k = i + 1; k < 6; k++) s -= F.ls[SYM(k, i)] * x[k];
        x[i] = s * F.ls[SYM(i, i)];
    }
#pragma unroll
    for (int j = 0; j < 6; j++) {
        gt[j] = x[j];
        z0 -= F.y[j][0] * x[j]; z1 -= F.y[j][1] * x[j]; z2 -= F.y[j][2] * x[j];
    }
    T w2 = z2 * F.l[5];
    T w1 = (z1 - F.l[4] * w2) * F.l[2];
    T w0 = (z0 - F.l[1] * w1 - F.l[3] * w2) * F.l[0];
    gk[0] = w0; gk[1] = w1; gk[2] = w2;
}
// y = A x
template <typename T> GO1_DEV void arrow_mul(const Arrow<T>& A, const T* xt, const T* xk, T* yt, T* yk, unsigned mask) {
#pragma unroll
    for (int r = 0; r < 6; r++) {
        T s = A.tk[r][0] * xk[0] + A.tk[r][1] * xk[1] + A.tk[r][2] * xk[2];
        s = tile_sum(s, mask);
#pragma unroll
        for (int c = 0; c < 6; c++) s += A.tt[SYM(r, c)] * xt[c];
        yt[r] = s;
    }
#pragma unroll
    for (int i = 0; i < 3; i++) {
        T s = 0;
#pragma unroll
        for (int c = 0; c < 3; c++) s += A.kk[SYM(i, c)] * xk[c];
#pragma unroll
        for (int j = 0; j < 6; j++) s += A.tk[j][i] * xt[j];
        yk[i] = s;
    }
}

// ---- per-lane contact storage (local memory; only touched slots ever become cache lines)
template <typename T> struct Contacts {
    int n;               // active contacts of this lane
    int start[5];        // slot boundaries (contacts are emitted in slot order)
    int overflow;
    T r[CMAX][3];        // contact point relative to the COM, world axes
    T dist[CMAX];
    T D0[CMAX];
    T aref[CMAX][6];
    T q[CMAX][12];       // line-search polynomial: quad(3) | U0 V0 UU | UV VV Dm | T0 zone0 x0
    T frc[CMAX][6];      // contact-frame force of the last evaluation (diag only)
    short cls[CMAX];
    short gid[CMAX];
};

// ---- everything a substep keeps live for one lane
template <typename T> struct Lane {
    // kinematics (positions relative to the trunk origin, world axes)
    T R0[9], R[3][9], p[3][3], com[3];
    T ca[3][3], cl[3][3];          // leg cdof: angular / linear (about the COM)
    T tl[3][3];                    // trunk rotational cdof, linear part (angular part = columns of R0)
    T cvel[4][6];                  // spatial velocity of trunk, hip, thigh, calf about the COM
    Arrow<T> M;
    T qfs_t[6], qfs_k[3];          // qfrc_smooth
    T as_t[6], as_k[3];            // qacc_smooth
    T qfa[3];                      // qfrc_actuator (leg part)
    // constraint rows owned by this lane
    T fr_aref[3];                  // friction-loss rows
    T lim_aref[3], lim_D[3], lim_s[3];  // limit rows (lim_s = 0 -> inactive, else Jacobian sign)
    T plane_d0;                    // signed distance of the trunk origin to the floor plane
};

// spatial motion of the four bodies for a joint-space vector (at, ak):  A_b = sum cdof_j x_j
template <typename T> GO1_DEV void body_motion(const Lane<T>& L, const T* at, const T* ak, T A[4][6]) {
#pragma unroll
    for (int c = 0; c < 3; c++) {
        A[0][c] = L.R0[3 * c + 0] * at[3] + L.R0[3 * c + 1] * at[4] + L.R0[3 * c + 2] * at[5];
        A[0][3 + c] = at[c] + L.tl[0][c] * at[3] + L.tl[1][c] * at[4] + L.tl[2][c] * at[5];
    }
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
        for (int c = 0; c < 3; c++) {
            A[i + 1][c] = A[i][c] + L.ca[i][c] * ak[i];
            A[i + 1][3 + c] = A[i][3 + c] + L.cl[i][c] * ak[i];
        }
}
// joint-space force from per-body spatial forces W (own, not yet composite). W is destroyed.
template <typename T> GO1_DEV void body_force_to_joint(const Lane<T>& L, T W[4][6], T* ft, T* fk, unsigned mask) {
#pragma unroll
    for (int c = 0; c < 6; c++) { W[2][c] += W[3][c]; W[1][c] += W[2][c]; }
#pragma unroll
    for (int c = 0; c < 6; c++) W[0][c] = tile_sum(W[0][c] + W[1][c], mask);
#pragma unroll
    for (int i = 0; i < 3; i++) fk[i] = dot3(L.ca[i], W[i + 1]) + dot3(L.cl[i], W[i + 1] + 3);
#pragma unroll
    for (int j = 0; j < 3; j++) {
        ft[j] = W[0][3 + j];
        T col[3] = {L.R0[j], L.R0[3 + j], L.R0[6 + j]};
        ft[3 + j] = dot3(col, W[0]) + dot3(L.tl[j], W[0] + 3);
    }
}

// rows of the contact map G (6 x [ang;lin]) for the floor frame F and lever r:  lin rows (r x f_j ; f_j), ang rows (f_j ; 0)
template <typename T> GO1_DEV void contact_rows(const T* F, const T* r, T g[3][3]) {
#pragma unroll
    for (int j = 0; j < 3; j++) cross3(g[j], r, F + 3 * j);
}

struct EvalOut { };

// Evaluates all constraint rows of this lane at joint-space acceleration (at, ak):
//   cost (lane-local, caller reduces), joint-space constraint force via W, optional Hessian blocks.
// mode bit0: accumulate Hessian (K per body, diagonal terms), bit1: store contact-frame forces
template <typename T, bool HESS>
GO1_DEV T eval_rows(const DevModel<T>& m, const Lane<T>& L, Contacts<T>& C, int leg, const T* at, const T* ak,
                    T W[4][6], T K[4][21], T* hdiag, T* fk_direct, bool store_frc) {
    typedef Num<T> N;
    T A[4][6];
    body_motion(L, at, ak, A);
    T cost = 0;
#pragma unroll
    for (int b = 0; b < 4; b++)
#pragma unroll
        for (int c = 0; c < 6; c++) W[b][c] = 0;
    if (HESS) {
#pragma unroll
        for (int b = 0; b < 4; b++)
#pragma unroll
            for (int e = 0; e < 21; e++) K[b][e] = 0;
    }
    const T* F = m.plane_frame;
#pragma unroll
    for (int b = 0; b < 4; b++) {
        for (int c = C.start[b]; c < C.start[b + 1]; c++) {
            const DevClass<T>& cl = m.cls[C.cls[c]];
            const int dim = cl.dim;
            T r[3] = {C.r[c][0], C.r[c][1], C.r[c][2]};
            T g[3][3];
            contact_rows(F, r, g);
            T jar[6], D[6];
            const T D0 = C.D0[c];
#pragma unroll
            for (int j = 0; j < 3; j++) {
                jar[j] = dot3(g[j], A[b]) + dot3(F + 3 * j, A[b] + 3) - C.aref[c][j];
                D[j] = D0 * cl.drat[j];
            }
            if (dim == 6) {
#pragma unroll
                for (int j = 0; j < 3; j++) { jar[3 + j] = dot3(F + 3 * j, A[b]) - C.aref[c][3 + j]; D[3 + j] = D0 * cl.drat[3 + j]; }
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
            if ((Nn >= mu * Tt) || (Tt <= 0 && Nn >= 0)) {
                // top zone: no force
            } else if ((mu * Nn + Tt <= 0) || (Tt <= 0 && Nn < 0)) {
                // bottom zone: quadratic in every row
#pragma unroll
                for (int j = 0; j < 6; j++) { f[j] = -D[j] * jar[j]; cost += T(0.5) * D[j] * jar[j] * jar[j]; }
                if (HESS) {
#pragma unroll
                    for (int j = 0; j < 3; j++) {
                        T gr[6] = {g[j][0], g[j][1], g[j][2], F[3 * j], F[3 * j + 1], F[3 * j + 2]};
                        sym6_rank1(K[b], gr, D[j]);
                    }
                    if (dim == 6) {
#pragma unroll
                        for (int j = 0; j < 3; j++) {  // ang rows touch only the angular 3x3 block
                            const T w = D[3 + j];
#pragma unroll
                            for (int rr = 0; rr < 3; rr++)
#pragma unroll
                                for (int cc = 0; cc <= rr; cc++) K[b][SYM(rr, cc)] += w * F[3 * j + rr] * F[3 * j + cc];
                        }
                    }
                }
            } else {
                // middle zone (cone surface)
                const T Dm = D0 / N::max_(mu * mu * (1 + mu * mu), N::MINVAL);
                const T NT = Nn - mu * Tt;
                cost += T(0.5) * Dm * NT * NT;
                f[0] = -Dm * NT * mu;
                const T fs = -f[0] / Tt;
#pragma unroll
                for (int j = 1; j < 6; j++) f[j] = fs * U[j] * cl.fri[j - 1];
                if (HESS) {
                    // cone Hessian in jar coordinates: H = S (Dm * Hu) S with S = diag(mu, fri), Hu as in mj_constraintUpdate
                    T sc[6];
                    sc[0] = mu;
#pragma unroll
                    for (int j = 1; j < 6; j++) sc[j] = cl.fri[j - 1];
                    const T a1 = mu * Nn / (Tt * Tt * Tt), a2 = mu * mu - mu * Nn / Tt;
                    // h_r = sum_s H[r][s] g_s  (6-vectors), then K += sum_r g_r h_r'
                    T G[6][6];
#pragma unroll
                    for (int j = 0; j < 3; j++) {
                        G[j][0] = g[j][0]; G[j][1] = g[j][1]; G[j][2] = g[j][2]; G[j][3] = F[3 * j]; G[j][4] = F[3 * j + 1]; G[j][5] = F[3 * j + 2];
                        G[3 + j][0] = F[3 * j]; G[3 + j][1] = F[3 * j + 1]; G[3 + j][2] = F[3 * j + 2]; G[3 + j][3] = 0; G[3 + j][4] = 0; G[3 + j][5] = 0;
                    }
#pragma unroll
                    for (int rr = 0; rr < 6; rr++) {
                        if (rr >= dim) continue;
                        T h[6] = {0, 0, 0, 0, 0, 0};
#pragma unroll
                        for (int ss = 0; ss < 6; ss++) {
                            if (ss >= dim) continue;
                            T hu;
                            if (rr == 0 && ss == 0) hu = 1;
                            else if (rr == 0) hu = -mu * U[ss] / Tt;
                            else if (ss == 0) hu = -mu * U[rr] / Tt;
                            else hu = a1 * U[rr] * U[ss] + (rr == ss ? a2 : T(0));
                            const T w = Dm * sc[rr] * sc[ss] * hu;
#pragma unroll
                            for (int e = 0; e < 6; e++) h[e] += w * G[ss][e];
                        }
#pragma unroll
                        for (int e = 0; e < 6; e++)
#pragma unroll
                            for (int e2 = 0; e2 <= e; e2++) K[b][SYM(e, e2)] += G[rr][e] * h[e2];  // sum over rr is symmetric
                    }
                }
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
            for (int e = 0; e < 3; e++) { W[b][e] += Tw[e] + rxF[e]; W[b][3 + e] += Fw[e]; }
            if (store_frc) {
#pragma unroll
                for (int j = 0; j < 6; j++) C.frc[c][j] = f[j];
            }
        }
    }
    // friction-loss rows (Huber) and joint-limit rows of this leg's dofs
#pragma unroll
    for (int i = 0; i < 3; i++) {
        T fd = 0;
        {
            const T R = m.Rf[leg][i], fl = m.floss[leg][i];
            if (fl > 0) {
                const T jar = ak[i] - L.fr_aref[i];
                const T Rf = R * fl;
                if (jar <= -Rf) { fd += fl; cost += -fl * (T(0.5) * Rf + jar); }
                else if (jar >= Rf) { fd += -fl; cost += -fl * (T(0.5) * Rf - jar); }
                else { const T D = T(1) / R; fd += -D * jar; cost += T(0.5) * D * jar * jar; if (HESS) hdiag[i] += D; }
            }
        }
        if (L.lim_s[i] != 0) {
            const T jar = L.lim_s[i] * ak[i] - L.lim_aref[i];
            if (jar < 0) { const T D = L.lim_D[i]; fd += L.lim_s[i] * (-D * jar); cost += T(0.5) * D * jar * jar; if (HESS) hdiag[i] += D; }
        }
        fk_direct[i] = fd;
    }
    return cost;
}

// Line-search pieces -----------------------------------------------------------------------------
template <typename T> struct LSRows {   // friction/limit rows of this lane at the line-search origin
    T fr_jar[3], fr_jv[3], lim_jar[3], lim_jv[3];
};

// prepares per-contact polynomials (C.q) for the direction (st, sk) at the point (at, ak)
template <typename T>
GO1_DEV void ls_prepare(const DevModel<T>& m, const Lane<T>& L, Contacts<T>& C, int leg, const T* at, const T* ak,
                        const T* st, const T* sk, LSRows<T>& rows) {
    typedef Num<T> N;
    T A[4][6], S[4][6];
    body_motion(L, at, ak, A);
    body_motion(L, st, sk, S);
    const T* F = m.plane_frame;
#pragma unroll
    for (int b = 0; b < 4; b++) {
        for (int c = C.start[b]; c < C.start[b + 1]; c++) {
            const DevClass<T>& cl = m.cls[C.cls[c]];
            const int dim = cl.dim;
            T r[3] = {C.r[c][0], C.r[c][1], C.r[c][2]};
            T g[3][3];
            contact_rows(F, r, g);
            const T D0 = C.D0[c], mu = cl.mu;
            T q0 = 0, q1 = 0, q2 = 0, U0 = 0, V0 = 0, UU = 0, UV = 0, VV = 0;
#pragma unroll
            for (int j = 0; j < 6; j++) {
                if (j >= dim) continue;
                T jar, jv;
                if (j < 3) {
                    jar = dot3(g[j], A[b]) + dot3(F + 3 * j, A[b] + 3) - C.aref[c][j];
                    jv = dot3(g[j], S[b]) + dot3(F + 3 * j, S[b] + 3);
                } else {
                    jar = dot3(F + 3 * (j - 3), A[b]) - C.aref[c][j];
                    jv = dot3(F + 3 * (j - 3), S[b]);
                }
                const T D = D0 * cl.drat[j];
                q0 += T(0.5) * D * jar * jar; q1 += D * jar * jv; q2 += T(0.5) * D * jv * jv;
                if (j == 0) { U0 = jar * mu; V0 = jv * mu; }
                else { const T u = jar * cl.fri[j - 1], v = jv * cl.fri[j - 1]; UU += u * u; UV += u * v; VV += v * v; }
            }
            C.q[c][0] = q0; C.q[c][1] = q1; C.q[c][2] = q2; C.q[c][3] = U0; C.q[c][4] = V0; C.q[c][5] = UU;
            C.q[c][6] = UV; C.q[c][7] = VV; C.q[c][8] = D0 / N::max_(mu * mu * (1 + mu * mu), N::MINVAL);
            // zone at alpha = 0 (0 top, 1 bottom, 2 middle): reference point of the cost DIFFERENCES
            const T T0 = N::sqrt_(N::max_(UU, T(0)));
            T z0;
            if (UU <= 0) z0 = U0 < 0 ? T(1) : T(0);
            else if (U0 >= mu * T0) z0 = 0;
            else if (mu * U0 + T0 <= 0) z0 = 1;
            else z0 = 2;
            C.q[c][9] = T0; C.q[c][10] = z0; C.q[c][11] = U0 - mu * T0;
        }
    }
#pragma unroll
    for (int i = 0; i < 3; i++) {
        rows.fr_jar[i] = ak[i] - L.fr_aref[i]; rows.fr_jv[i] = sk[i];
        rows.lim_jar[i] = L.lim_s[i] * ak[i] - L.lim_aref[i]; rows.lim_jv[i] = L.lim_s[i] * sk[i];
    }
}

// lane-local part of PrimalEval at step alpha: adds into cost / d1 / d2.
// `cost` is accumulated as the DIFFERENCE to the cost at alpha = 0, row by row: rows that stay in the same
// piece contribute alpha*q1 + alpha^2*q2 (no constant term), so the comparison of two candidate steps is
// meaningful in fp32 even when the total cost is ~1e4 (MuJoCo compares absolute fp64 costs).
template <typename T>
GO1_DEV void ls_eval_lane(const DevModel<T>& m, const Lane<T>& L, const Contacts<T>& C, int leg, const LSRows<T>& rows,
                          T alpha, T& cost, T& d1, T& d2) {
    typedef Num<T> N;
    for (int c = 0; c < C.n; c++) {
        const T mu = m.cls[C.cls[c]].mu;
        const T* q = C.q[c];
        const T Nn = q[3] + alpha * q[4], Tsq = q[5] + alpha * (2 * q[6] + alpha * q[7]);
        const T Dm = q[8], T0 = q[9], x0 = q[11];
        const int z0 = (int)q[10];
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
            d1 += q[1] + 2 * alpha * q[2]; d2 += 2 * q[2];
            ca = q[0] + alpha * (q[1] + alpha * q[2]);
        } else if (za == 2) {
            const T N1 = q[4], T1 = (q[6] + alpha * q[7]) / Tt, T2 = q[7] / Tt - (q[6] + alpha * q[7]) * T1 / (Tt * Tt);
            const T NT = Nn - mu * Tt, NT1 = N1 - mu * T1;
            d1 += Dm * NT * NT1;
            d2 += Dm * (NT1 * NT1 + NT * (-mu * T2));
            ca = T(0.5) * Dm * NT * NT;
        }
        if (za == z0) {
            if (za == 1) cost += alpha * (q[1] + alpha * q[2]);
            else if (za == 2) {
                const T dx = alpha * (q[4] - mu * (2 * q[6] + alpha * q[7]) / (Tt + T0));
                cost += T(0.5) * Dm * dx * ((Nn - mu * Tt) + x0);
            }
        } else {
            const T c0 = z0 == 1 ? q[0] : (z0 == 2 ? T(0.5) * Dm * x0 * x0 : T(0));
            cost += ca - c0;
        }
    }
#pragma unroll
    for (int i = 0; i < 3; i++) {
        const T fl = m.floss[leg][i];
        if (fl > 0) {
            const T R = m.Rf[leg][i], Rf = R * fl, jar = rows.fr_jar[i], jv = rows.fr_jv[i];
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
        if (L.lim_s[i] != 0) {
            const T jar = rows.lim_jar[i], jv = rows.lim_jv[i], D = L.lim_D[i];
            const T x = jar + alpha * jv;
            const bool aa = x < 0, a0 = jar < 0;
            if (aa) { d1 += D * jar * jv + alpha * D * jv * jv; d2 += D * jv * jv; }
            if (aa && a0) cost += alpha * (D * jar * jv + T(0.5) * alpha * D * jv * jv);
            else cost += (aa ? T(0.5) * D * x * x : T(0)) - (a0 ? T(0.5) * D * jar * jar : T(0));
        }
    }
}

template <typename T> struct LSPoint { T alpha, cost, d1, d2; };

}  // namespace go1
