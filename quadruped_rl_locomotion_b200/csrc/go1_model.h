// go1_model.h -- device model (quadruped-specialised) and its construction from the flat model blob.
//
// The blob is what quadruped_rl_locomotion_b200/mjcf.py emits (mjModel-style named arrays).  The CUDA
// path is specialised to the topology the reference's MJCF has (go1_torque.xml:83-208): one free
// trunk + NLEG identical 3-hinge chains.  build_dev_model() validates exactly that and refuses
// anything else loudly (GO1_EMODEL) instead of guessing.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

namespace go1 {

constexpr int NLEG = 4;       // lanes per environment: one lane per leg
constexpr int GMAX = 12;      // collision geoms handled by one lane (its leg's + a share of the trunk's)
constexpr int CMAX = 16;      // active contacts kept per lane
constexpr int NCLS = 8;       // distinct contact parameter classes (floor-mixed + robot-robot mixed)
constexpr int NKIND = 4;      // distinct per-geom contact parameter sets
constexpr int NGP = 704;      // candidate geom pairs of the robot-robot collision (bodies not parent and child; cylinder-cylinder: left / right hips only)
constexpr int GFLAT = 44;     // robot collision geoms (flat list, sorted by body)

// index tables of the block-arrow layout tt[21] | tk[4][6][3] | kk[4][6]  (go1_warp.cuh)
// crb_idx / crb_sgn: entry e of the packed 6x6 spatial inertia = crb_sgn[e] * cinert[crb_idx[e]] (mju_mulInertVec)
struct ArrowTab { unsigned char sym_r[21], sym_c[21], pad[2]; unsigned char ent_u[120], ent_c[120]; signed char crb_idx[24], crb_sgn[24]; };

enum { G_SPHERE = 2, G_CAPSULE = 3, G_CYLINDER = 5, G_BOX = 6 };

template <typename T>
struct DevGeom {
    int type, slot, cls, gid;  // slot: body index (0 trunk, 1 + 3*leg + link); cls: class of the contact with the floor; gid: MuJoCo geom id
    int kind, pad;             // kind: index of this geom's own contact parameter set (robot-robot classes: DevModel::pair_cls)
    T pos[3], zax[3], xax[3], size[3];
};

template <typename T>
struct DevClass {
    int dim, pad;
    T mu;           // contact.mu = friction[0]*sqrt(R1/R0) = friction[0]/sqrt(impratio)
    T fri[5];       // unpacked friction (f0,f0,f1,f2,f2), floor-mixed
    T drat[6];      // D_j / D_0  (1, impratio*f_{j-1}^2/f_0^2 ...)
    T K, B;         // reference acceleration constants of the normal row (friction rows: K = 0)
    T d0, d1, width, mid, power;  // clamped solimp
    T margin, incmargin;
};

template <typename T>
struct DevModel {
    T timestep, gravity[3], impratio, tolerance, ls_tolerance, meaninertia, total_mass;
    T lambda_lb;     // rigorous lower bound of the eigenvalues of M (any pose): min(leg armature, trunk mass, trunk principal inertias)
    int iterations, ls_iterations, nclass, pad0;
    T plane_pos[3], plane_frame[9];
    // trunk
    T t_mass, t_ipos[3], t_Ib[6], t_invw[2];
    // legs [leg][link]
    T body_pos[NLEG][3][3], jnt_axis[NLEG][3][3], ipos[NLEG][3][3], mass[NLEG][3], Ib[NLEG][3][6], invw[NLEG][3][2];
    T qpos0[NLEG][3], damping[NLEG][3], armature[NLEG][3], floss[NLEG][3], Rf[NLEG][3], Bf[NLEG][3];
    T lim_lo[NLEG][3], lim_hi[NLEG][3], lim_K[NLEG][3], lim_B[NLEG][3], lim_imp[NLEG][3][5], lim_margin[NLEG][3], dof_invw[NLEG][3];
    int limited[NLEG][3];
    // actuators [leg][link]
    int act_id[NLEG][3], act_climited[NLEG][3], act_flimited[NLEG][3], act_bias[NLEG][3];
    T act_gear[NLEG][3], act_gain[NLEG][3], act_b0[NLEG][3], act_b1[NLEG][3], act_b2[NLEG][3];
    T act_clo[NLEG][3], act_chi[NLEG][3], act_flo[NLEG][3], act_fhi[NLEG][3];
    DevClass<T> cls[NCLS];
    // flat geom list sorted by body (slot = body index: 0 trunk, 1 + 3*leg + link) + bounding radii
    int nflat, pad1;
    DevGeom<T> flat[GFLAT];
    T flat_rb[GFLAT];
    ArrowTab tab;
    // robot-robot collision (go1_pair.cuh): number of candidate geom pairs, class of a contact between geoms of kinds (a, b)
    int ngp, pad2;
    int pair_cls[NKIND][NKIND];
    T pair_margin;
    // env constants
    T key_qpos[19], key_ctrl[12], soft_lo[12], soft_hi[12], ctrl_lo[12], ctrl_hi[12];
    // ---- NOT staged in shared memory (read through EnvCfg::gp_tab): candidate geom pairs (flat indices) in mj_collision's order
    unsigned char gp_a[NGP], gp_b[NGP];
};

// ---------------------------------------------------------------------------------------------
// host side: blob reader
struct Blob {
    const uint8_t* b = nullptr;
    size_t n = 0;
    std::string err;
    const void* find(const char* name, int* count, int want_code) {
        if (n < 16 || std::memcmp(b, "GO1MDL01", 8) != 0) { err = "bad model blob magic"; return nullptr; }
        uint32_t cnt;
        std::memcpy(&cnt, b + 8, 4);
        const uint8_t* p = b + 16;
        for (uint32_t i = 0; i < cnt; i++, p += 88) {
            if ((size_t)(p - b) + 88 > n) break;
            if (std::strncmp((const char*)p, name, 48) == 0) {
                int32_t code; int64_t off, nb;
                std::memcpy(&code, p + 48, 4); std::memcpy(&off, p + 72, 8); std::memcpy(&nb, p + 80, 8);
                if (code != want_code || off < 0 || (size_t)(off + nb) > n) { err = std::string("bad entry ") + name; return nullptr; }
                if (count) *count = (int)(nb / (code == 0 ? 8 : (code == 1 ? 4 : 1)));
                return b + off;
            }
        }
        err = std::string("model blob lacks field ") + name;
        return nullptr;
    }
    const double* d(const char* name, int* count = nullptr) { return (const double*)find(name, count, 0); }
    const int* i(const char* name, int* count = nullptr) { return (const int*)find(name, count, 1); }
};

inline void quat2mat(const double* q, double* R) {
    double q00 = q[0] * q[0], q01 = q[0] * q[1], q02 = q[0] * q[2], q03 = q[0] * q[3];
    double q11 = q[1] * q[1], q12 = q[1] * q[2], q13 = q[1] * q[3], q22 = q[2] * q[2], q23 = q[2] * q[3], q33 = q[3] * q[3];
    R[0] = q00 + q11 - q22 - q33; R[4] = q00 - q11 + q22 - q33; R[8] = q00 - q11 - q22 + q33;
    R[1] = 2 * (q12 - q03); R[2] = 2 * (q13 + q02); R[3] = 2 * (q12 + q03);
    R[5] = 2 * (q23 - q01); R[6] = 2 * (q13 - q02); R[7] = 2 * (q23 + q01);
}

// body-frame inertia tensor (xx,yy,zz,xy,xz,yz) from principal inertia + iquat
inline void body_inertia(const double* iquat, const double* diag, double* out6) {
    double R[9];
    quat2mat(iquat, R);
    double M[9];
    for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++) {
            double s = 0;
            for (int k = 0; k < 3; k++) s += R[3 * r + k] * diag[k] * R[3 * c + k];
            M[3 * r + c] = s;
        }
    out6[0] = M[0]; out6[1] = M[4]; out6[2] = M[8]; out6[3] = M[1]; out6[4] = M[2]; out6[5] = M[5];
}

inline void make_frame(double* f) {  // mju_makeFrame
    double n = std::sqrt(f[0] * f[0] + f[1] * f[1] + f[2] * f[2]);
    for (int k = 0; k < 3; k++) f[k] /= n;
    f[3] = f[4] = f[5] = 0;
    if (f[1] < 0.5 && f[1] > -0.5) f[4] = 1; else f[5] = 1;
    double d = f[0] * f[3] + f[1] * f[4] + f[2] * f[5];
    for (int k = 0; k < 3; k++) f[3 + k] -= f[k] * d;
    n = std::sqrt(f[3] * f[3] + f[4] * f[4] + f[5] * f[5]);
    for (int k = 0; k < 3; k++) f[3 + k] /= n;
    f[6] = f[1] * f[5] - f[2] * f[4]; f[7] = f[2] * f[3] - f[0] * f[5]; f[8] = f[0] * f[4] - f[1] * f[3];
}

#define GO1_REQ(cond, msg) do { if (!(cond)) { err = (msg); return false; } } while (0)
#define GO1_GETD(var, name) const double* var = blob.d(name); if (!var) { err = blob.err; return false; }
#define GO1_GETI(var, name) const int* var = blob.i(name); if (!var) { err = blob.err; return false; }

// Builds the specialised device model (in double; cast to T by the caller). Returns false + err on rejection.
inline bool build_dev_model(const void* data, size_t nbytes, DevModel<double>& M, std::string& err) {
    Blob blob; blob.b = (const uint8_t*)data; blob.n = nbytes;
    std::memset(&M, 0, sizeof(M));
    GO1_GETI(p_nq, "nq") GO1_GETI(p_nv, "nv") GO1_GETI(p_nu, "nu") GO1_GETI(p_nbody, "nbody") GO1_GETI(p_njnt, "njnt")
    GO1_GETI(p_ngeom, "ngeom") GO1_GETI(p_nkey, "nkey")
    int nq = *p_nq, nv = *p_nv, nu = *p_nu, nbody = *p_nbody, njnt = *p_njnt, ngeom = *p_ngeom;
    GO1_REQ(nq == 19 && nv == 18 && nbody == 14 && njnt == 13, "model is not a free trunk + 4 x 3-hinge legs (nq/nv/nbody/njnt)");
    GO1_REQ(nu == 12, "expected 12 actuators");
    GO1_REQ(*p_nkey >= 1, "model needs keyframe 0 (home pose)");
    GO1_GETD(o_ts, "opt_timestep") GO1_GETD(o_g, "opt_gravity") GO1_GETD(o_imp, "opt_impratio") GO1_GETD(o_tol, "opt_tolerance")
    GO1_GETD(o_lst, "opt_ls_tolerance") GO1_GETD(o_mi, "stat_meaninertia")
    GO1_GETI(o_it, "opt_iterations") GO1_GETI(o_lsit, "opt_ls_iterations") GO1_GETI(o_cone, "opt_cone")
    GO1_GETI(o_integ, "opt_integrator") GO1_GETI(o_solver, "opt_solver")
    GO1_REQ(*o_cone == 1 && *o_integ == 0 && *o_solver == 2, "only cone=elliptic, integrator=Euler, solver=Newton are implemented");
    M.timestep = *o_ts; M.impratio = *o_imp; M.tolerance = *o_tol; M.ls_tolerance = *o_lst; M.meaninertia = *o_mi;
    M.iterations = *o_it; M.ls_iterations = *o_lsit;
    for (int k = 0; k < 3; k++) M.gravity[k] = o_g[k];

    GO1_GETI(body_parentid, "body_parentid") GO1_GETI(body_jntnum, "body_jntnum") GO1_GETI(body_jntadr, "body_jntadr")
    GO1_GETD(body_pos, "body_pos") GO1_GETD(body_quat, "body_quat") GO1_GETD(body_ipos, "body_ipos") GO1_GETD(body_iquat, "body_iquat")
    GO1_GETD(body_mass, "body_mass") GO1_GETD(body_inertia_, "body_inertia") GO1_GETD(body_invw, "body_invweight0")
    GO1_GETI(jnt_type, "jnt_type") GO1_GETI(jnt_qposadr, "jnt_qposadr") GO1_GETI(jnt_dofadr, "jnt_dofadr") GO1_GETI(jnt_limited, "jnt_limited")
    GO1_GETD(jnt_pos, "jnt_pos") GO1_GETD(jnt_axis, "jnt_axis") GO1_GETD(jnt_range, "jnt_range") GO1_GETD(jnt_solref, "jnt_solref")
    GO1_GETD(jnt_solimp, "jnt_solimp") GO1_GETD(jnt_margin, "jnt_margin")
    GO1_GETD(dof_armature, "dof_armature") GO1_GETD(dof_damping, "dof_damping") GO1_GETD(dof_floss, "dof_frictionloss")
    GO1_GETD(dof_invw, "dof_invweight0") GO1_GETD(dof_solref, "dof_solref") GO1_GETD(dof_solimp, "dof_solimp")
    GO1_GETD(qpos0, "qpos0")
    GO1_GETI(geom_type, "geom_type") GO1_GETI(geom_contype, "geom_contype") GO1_GETI(geom_conaff, "geom_conaffinity")
    GO1_GETI(geom_condim, "geom_condim") GO1_GETI(geom_bodyid, "geom_bodyid") GO1_GETI(geom_priority, "geom_priority")
    GO1_GETD(geom_size, "geom_size") GO1_GETD(geom_pos, "geom_pos") GO1_GETD(geom_quat, "geom_quat") GO1_GETD(geom_friction, "geom_friction")
    GO1_GETD(geom_margin, "geom_margin") GO1_GETD(geom_gap, "geom_gap") GO1_GETD(geom_solmix, "geom_solmix")
    GO1_GETD(geom_solref, "geom_solref") GO1_GETD(geom_solimp, "geom_solimp")
    GO1_GETI(act_trnid, "actuator_trnid") GO1_GETI(act_cl, "actuator_ctrllimited") GO1_GETI(act_fl, "actuator_forcelimited")
    GO1_GETI(act_bt, "actuator_biastype")
    GO1_GETD(act_gear, "actuator_gear") GO1_GETD(act_gain, "actuator_gainprm") GO1_GETD(act_bias, "actuator_biasprm")
    GO1_GETD(act_cr, "actuator_ctrlrange") GO1_GETD(act_fr, "actuator_forcerange")
    GO1_GETD(key_qpos, "key_qpos") GO1_GETD(key_ctrl, "key_ctrl")

    // ---- topology: body 1 = free trunk (child of world), then 4 chains of 3 single-hinge bodies
    GO1_REQ(body_parentid[1] == 0 && body_jntnum[1] == 1 && jnt_type[body_jntadr[1]] == 0, "body 1 must be the free-floating trunk");
    GO1_REQ(jnt_qposadr[body_jntadr[1]] == 0 && jnt_dofadr[body_jntadr[1]] == 0, "trunk free joint must own qpos[0:7]/qvel[0:6]");
    int legbody[NLEG][3];
    for (int k = 0; k < NLEG; k++)
        for (int i = 0; i < 3; i++) {
            int b = 2 + 3 * k + i;
            legbody[k][i] = b;
            GO1_REQ(body_parentid[b] == (i == 0 ? 1 : b - 1), "legs must be 3-body chains hanging off the trunk, in body order");
            GO1_REQ(body_jntnum[b] == 1, "each leg body needs exactly one joint");
            int j = body_jntadr[b];
            GO1_REQ(jnt_type[j] == 3, "leg joints must be hinges");
            GO1_REQ(jnt_qposadr[j] == 7 + 3 * k + i && jnt_dofadr[j] == 6 + 3 * k + i, "unexpected joint addressing");
            GO1_REQ(std::fabs(jnt_pos[3 * j]) + std::fabs(jnt_pos[3 * j + 1]) + std::fabs(jnt_pos[3 * j + 2]) == 0.0, "leg joint anchors must be at the body origin");
            const double* q = body_quat + 4 * b;
            GO1_REQ(q[0] == 1.0 && q[1] == 0.0 && q[2] == 0.0 && q[3] == 0.0, "leg bodies must have identity body_quat");
        }
    double total = 0;
    for (int b = 1; b < nbody; b++) total += body_mass[b];
    M.total_mass = total;
    M.t_mass = body_mass[1];
    for (int c = 0; c < 3; c++) M.t_ipos[c] = body_ipos[3 + c];
    body_inertia(body_iquat + 4, body_inertia_ + 3, M.t_Ib);
    // M = sum_b J_b' I_b J_b + diag(armature) >= J_trunk' I_trunk J_trunk + diag(armature): block diagonal (trunk spatial inertia,
    // armature on the 12 hinges), so lambda_min(M) >= min(trunk mass, smallest trunk principal inertia, smallest armature)
    M.lambda_lb = std::min(body_mass[1], std::min(body_inertia_[3], std::min(body_inertia_[4], body_inertia_[5])));
    for (int d = 6; d < 18; d++) M.lambda_lb = std::min(M.lambda_lb, dof_armature[d]);
    M.t_invw[0] = body_invw[2]; M.t_invw[1] = body_invw[3];
    for (int k = 0; k < NLEG; k++)
        for (int i = 0; i < 3; i++) {
            int b = legbody[k][i], j = body_jntadr[b], dof = 6 + 3 * k + i;
            for (int c = 0; c < 3; c++) {
                M.body_pos[k][i][c] = body_pos[3 * b + c];
                M.jnt_axis[k][i][c] = jnt_axis[3 * j + c];
                M.ipos[k][i][c] = body_ipos[3 * b + c];
            }
            M.mass[k][i] = body_mass[b];
            body_inertia(body_iquat + 4 * b, body_inertia_ + 3 * b, M.Ib[k][i]);
            M.invw[k][i][0] = body_invw[2 * b]; M.invw[k][i][1] = body_invw[2 * b + 1];
            M.qpos0[k][i] = qpos0[7 + 3 * k + i];
            M.damping[k][i] = dof_damping[dof]; M.armature[k][i] = dof_armature[dof]; M.floss[k][i] = dof_floss[dof];
            M.dof_invw[k][i] = dof_invw[dof];
            // friction-loss row: pos = margin = 0 -> impedance = clamp(solimp[0]); R, B are constants
            {
                const double* si = dof_solimp + 5 * dof; const double* sr = dof_solref + 2 * dof;
                double d0 = std::fmin(0.9999, std::fmax(0.0001, si[0])), d1 = std::fmin(0.9999, std::fmax(0.0001, si[1]));
                double imp = (si[0] == si[1] || si[2] <= 1e-15) ? 0.5 * (d0 + d1) : d0;
                M.Rf[k][i] = std::fmax(1e-15, (1 - imp) * dof_invw[dof] / imp);
                GO1_REQ(sr[0] > 0, "direct (negative) solref is not implemented");
                double tc = std::fmax(sr[0], 2 * M.timestep);
                M.Bf[k][i] = 2 / std::fmax(1e-15, d1 * tc);
            }
            M.limited[k][i] = jnt_limited[j];
            M.lim_lo[k][i] = jnt_range[2 * j]; M.lim_hi[k][i] = jnt_range[2 * j + 1]; M.lim_margin[k][i] = jnt_margin[j];
            {
                const double* si = jnt_solimp + 5 * j; const double* sr = jnt_solref + 2 * j;
                GO1_REQ(sr[0] > 0, "direct (negative) solref is not implemented");
                double d1 = std::fmin(0.9999, std::fmax(0.0001, si[1]));
                double tc = std::fmax(sr[0], 2 * M.timestep);
                M.lim_K[k][i] = 1 / std::fmax(1e-15, d1 * d1 * tc * tc * sr[1] * sr[1]);
                M.lim_B[k][i] = 2 / std::fmax(1e-15, d1 * tc);
                M.lim_imp[k][i][0] = std::fmin(0.9999, std::fmax(0.0001, si[0])); M.lim_imp[k][i][1] = d1;
                M.lim_imp[k][i][2] = std::fmax(0.0, si[2]); M.lim_imp[k][i][3] = std::fmin(0.9999, std::fmax(0.0001, si[3]));
                M.lim_imp[k][i][4] = std::fmax(1.0, si[4]);
            }
            M.act_id[k][i] = -1;
        }
    GO1_REQ(dof_floss[0] == 0 && dof_damping[0] == 0 && dof_armature[0] == 0 && dof_floss[3] == 0 && dof_damping[3] == 0 && dof_armature[3] == 0,
            "free joint must have no damping/armature/frictionloss");
    // ---- actuators: one per leg joint
    for (int u = 0; u < nu; u++) {
        int j = act_trnid[u];
        GO1_REQ(j >= 1 && j <= 12, "actuator must drive a leg hinge");
        int k = (j - 1) / 3, i = (j - 1) % 3;
        GO1_REQ(M.act_id[k][i] < 0, "two actuators on one joint are not supported");
        M.act_id[k][i] = u;
        M.act_gear[k][i] = act_gear[u]; M.act_gain[k][i] = act_gain[3 * u];
        M.act_bias[k][i] = act_bt[u];
        M.act_b0[k][i] = act_bias[3 * u]; M.act_b1[k][i] = act_bias[3 * u + 1]; M.act_b2[k][i] = act_bias[3 * u + 2];
        M.act_climited[k][i] = act_cl[u]; M.act_clo[k][i] = act_cr[2 * u]; M.act_chi[k][i] = act_cr[2 * u + 1];
        M.act_flimited[k][i] = act_fl[u]; M.act_flo[k][i] = act_fr[2 * u]; M.act_fhi[k][i] = act_fr[2 * u + 1];
        M.ctrl_lo[u] = act_cr[2 * u]; M.ctrl_hi[u] = act_cr[2 * u + 1];
        double off = 0.5 * (1 - 0.9) * (act_cr[2 * u + 1] - act_cr[2 * u]);   // go1_mujoco_env.py:134-145
        M.soft_lo[u] = act_cr[2 * u] + off; M.soft_hi[u] = act_cr[2 * u + 1] - off;
        GO1_REQ(u == 3 * k + i, "actuator order must follow joint order (action layout of the env)");
    }
    for (int k = 0; k < NLEG; k++) for (int i = 0; i < 3; i++) GO1_REQ(M.act_id[k][i] >= 0, "every leg joint needs an actuator");
    for (int c = 0; c < 19; c++) M.key_qpos[c] = key_qpos[c];
    for (int c = 0; c < 12; c++) M.key_ctrl[c] = key_ctrl[c];

    // ---- floor plane: exactly one colliding plane on the world body
    int floor = -1;
    for (int g = 0; g < ngeom; g++)
        if (geom_type[g] == 0 && geom_bodyid[g] == 0 && (geom_contype[g] || geom_conaff[g])) { GO1_REQ(floor < 0, "more than one floor plane"); floor = g; }
    GO1_REQ(floor >= 0, "no floor plane");
    {
        double R[9];
        quat2mat(geom_quat + 4 * floor, R);
        for (int c = 0; c < 3; c++) M.plane_pos[c] = geom_pos[3 * floor + c];
        double f[9] = {R[2], R[5], R[8], 0, 0, 0, 0, 0, 0};
        make_frame(f);
        for (int c = 0; c < 9; c++) M.plane_frame[c] = f[c];
    }
    // ---- collision geoms -> lanes; contact parameter classes (mixing with the floor, mj_contactParam rules)
    // contact parameter class of a geom pair (mj_contactParam rules) -> index into M.cls (created on first use)
    std::string class_err;
    auto class_of = [&](int ga, int gb) -> int {
        DevClass<double> C;
        std::memset(&C, 0, sizeof(C));
        double fri[3], solref[2], solimp[5];
        int p1 = geom_priority[ga], p2 = geom_priority[gb];
        if (p1 != p2) {
            int w = p1 > p2 ? ga : gb;
            C.dim = geom_condim[w];
            for (int c = 0; c < 3; c++) fri[c] = geom_friction[3 * w + c];
            for (int c = 0; c < 2; c++) solref[c] = geom_solref[2 * w + c];
            for (int c = 0; c < 5; c++) solimp[c] = geom_solimp[5 * w + c];
        } else {
            C.dim = geom_condim[ga] > geom_condim[gb] ? geom_condim[ga] : geom_condim[gb];
            for (int c = 0; c < 3; c++) fri[c] = std::fmax(geom_friction[3 * ga + c], geom_friction[3 * gb + c]);
            double m1 = geom_solmix[ga], m2 = geom_solmix[gb], mix;
            if (m1 >= 1e-15 && m2 >= 1e-15) mix = m1 / (m1 + m2);
            else if (m1 < 1e-15 && m2 < 1e-15) mix = 0.5;
            else mix = m1 < 1e-15 ? 0.0 : 1.0;
            for (int c = 0; c < 2; c++) solref[c] = mix * geom_solref[2 * ga + c] + (1 - mix) * geom_solref[2 * gb + c];
            for (int c = 0; c < 5; c++) solimp[c] = mix * geom_solimp[5 * ga + c] + (1 - mix) * geom_solimp[5 * gb + c];
        }
        if (!(C.dim == 1 || C.dim == 3 || C.dim == 6)) { class_err = "contacts must have condim 1, 3 or 6 after mixing"; return -1; }
        if (!(solref[0] > 0)) { class_err = "direct (negative) solref is not implemented"; return -1; }
        C.fri[0] = C.fri[1] = std::fmax(1e-5, fri[0]); C.fri[2] = std::fmax(1e-5, fri[1]);
        C.fri[3] = C.fri[4] = std::fmax(1e-5, fri[2]);
        C.mu = C.fri[0] / std::sqrt(std::fmax(1e-15, M.impratio));
        C.drat[0] = 1;
        for (int jj = 1; jj < 6; jj++) C.drat[jj] = jj < C.dim ? M.impratio * C.fri[jj - 1] * C.fri[jj - 1] / (C.fri[0] * C.fri[0]) : 0.0;
        C.d0 = std::fmin(0.9999, std::fmax(0.0001, solimp[0])); C.d1 = std::fmin(0.9999, std::fmax(0.0001, solimp[1]));
        C.width = std::fmax(0.0, solimp[2]); C.mid = std::fmin(0.9999, std::fmax(0.0001, solimp[3])); C.power = std::fmax(1.0, solimp[4]);
        double tc = std::fmax(solref[0], 2 * M.timestep);
        C.K = 1 / std::fmax(1e-15, C.d1 * C.d1 * tc * tc * solref[1] * solref[1]);
        C.B = 2 / std::fmax(1e-15, C.d1 * tc);
        C.margin = std::fmax(geom_margin[ga], geom_margin[gb]);
        C.incmargin = C.margin - std::fmax(geom_gap[ga], geom_gap[gb]);
        for (int c = 0; c < M.nclass; c++) {
            DevClass<double> a = M.cls[c];
            if (std::memcmp(&a, &C, sizeof(C)) == 0) return c;
        }
        if (M.nclass >= NCLS) { class_err = "too many distinct contact parameter classes"; return -1; }
        M.cls[M.nclass] = C;
        return M.nclass++;
    };
    // per-geom contact parameter set ("kind"): geoms with identical own parameters share one
    std::vector<int> kind_rep;
    auto same_params = [&](int a, int b) {
        bool eq = geom_priority[a] == geom_priority[b] && geom_condim[a] == geom_condim[b] && geom_solmix[a] == geom_solmix[b] &&
                  geom_margin[a] == geom_margin[b] && geom_gap[a] == geom_gap[b];
        for (int c = 0; c < 3; c++) eq = eq && geom_friction[3 * a + c] == geom_friction[3 * b + c];
        for (int c = 0; c < 2; c++) eq = eq && geom_solref[2 * a + c] == geom_solref[2 * b + c];
        for (int c = 0; c < 5; c++) eq = eq && geom_solimp[5 * a + c] == geom_solimp[5 * b + c];
        return eq;
    };
    auto kind_of = [&](int g) -> int {
        for (size_t k = 0; k < kind_rep.size(); k++) if (same_params(kind_rep[k], g)) return (int)k;
        if ((int)kind_rep.size() >= NKIND) return -1;
        kind_rep.push_back(g);
        return (int)kind_rep.size() - 1;
    };
    int trunk_share = 0;
    int cnt[NLEG][4];
    std::memset(cnt, 0, sizeof(cnt));
    std::vector<DevGeom<double>> tmp[NLEG][4];
    M.nclass = 0;
    for (int g = 0; g < ngeom; g++) {
        int b = geom_bodyid[g];
        if (b == 0) {
            GO1_REQ(g == floor || !(geom_contype[g] || geom_conaff[g]), "only one static collider (the floor) is supported");
            continue;
        }
        if (!((geom_contype[floor] & geom_conaff[g]) || (geom_contype[g] & geom_conaff[floor]))) continue;  // visual geoms
        int t = geom_type[g];
        GO1_REQ(t == G_SPHERE || t == G_CAPSULE || t == G_CYLINDER || t == G_BOX, "unsupported collision geom type");
        int lane, slot;
        if (b == 1) { lane = trunk_share++ % NLEG; slot = 0; }
        else { lane = (b - 2) / 3; slot = (b - 2) % 3 + 1; }
        DevGeom<double> dg;
        std::memset(&dg, 0, sizeof(dg));
        dg.type = t; dg.slot = slot; dg.gid = g;
        double R[9];
        quat2mat(geom_quat + 4 * g, R);
        for (int c = 0; c < 3; c++) {
            dg.pos[c] = geom_pos[3 * g + c]; dg.size[c] = geom_size[3 * g + c];
            dg.zax[c] = R[3 * c + 2]; dg.xax[c] = R[3 * c + 0];
        }
        dg.kind = kind_of(g);
        GO1_REQ(dg.kind >= 0, "too many distinct per-geom contact parameter sets");
        int ci = class_of(floor, g);
        GO1_REQ(ci >= 0, class_err.c_str());
        GO1_REQ(M.cls[ci].dim == 3 || M.cls[ci].dim == 6, "floor contacts must have condim 3 or 6 after mixing");
        dg.cls = ci;
        tmp[lane][slot].push_back(dg);
    }
    // flat list in body order: trunk geoms (kept in geom-id order), then every leg body
    {
        std::vector<DevGeom<double>> trunk;
        for (int k = 0; k < NLEG; k++) for (auto& dg : tmp[k][0]) trunk.push_back(dg);
        for (size_t a = 0; a < trunk.size(); a++)
            for (size_t b2 = a + 1; b2 < trunk.size(); b2++)
                if (trunk[b2].gid < trunk[a].gid) std::swap(trunk[a], trunk[b2]);
        int n = 0;
        auto push = [&](DevGeom<double> dg, int body) -> bool {
            if (n >= GFLAT) return false;
            dg.slot = body;
            double rb = 0;
            if (dg.type == G_SPHERE) rb = dg.size[0];
            else if (dg.type == G_CAPSULE) rb = dg.size[0] + dg.size[1];
            else if (dg.type == G_CYLINDER) rb = std::sqrt(dg.size[0] * dg.size[0] + dg.size[1] * dg.size[1]);
            else rb = std::sqrt(dg.size[0] * dg.size[0] + dg.size[1] * dg.size[1] + dg.size[2] * dg.size[2]);
            M.flat_rb[n] = rb * (1 + 1e-6) + 1e-9;
            M.flat[n++] = dg;
            return true;
        };
        for (auto& dg : trunk) GO1_REQ(push(dg, 0), "too many collision geoms");
        for (int k = 0; k < NLEG; k++)
            for (int s = 1; s < 4; s++)
                for (auto& dg : tmp[k][s]) GO1_REQ(push(dg, 1 + 3 * k + (s - 1)), "too many collision geoms");
        M.nflat = n;
    }
    // ---- robot-robot collision tables (mj_collision: body pairs that are not parent and child; geoms of one body never collide)
    {
        int body_g0[15];
        for (int b = 0; b <= 13; b++) {
            int g = 0;
            while (g < M.nflat && M.flat[g].slot < b) g++;
            body_g0[b] = g;
        }
        body_g0[14] = M.nflat;
        auto parent = [](int b) { return b == 0 ? -1 : ((b - 1) % 3 == 0 ? 0 : b - 1); };
        M.ngp = 0;
        M.pair_margin = 0;
        for (int a = 0; a < 13; a++)
            for (int b = a + 1; b < 13; b++) {
                if (parent(b) == a || parent(a) == b) continue;
                for (int g1 = body_g0[a]; g1 < body_g0[a + 1]; g1++)
                    for (int g2 = body_g0[b]; g2 < body_g0[b + 1]; g2++) {
                        // cylinder-cylinder: only the left / right hip motors can meet (coplanar pair, go1_pair.cuh:pp_cyl_cyl); the
                        // other cylinder pairs (front vs rear hips) are 0.37 m apart and are left out of the table
                        const bool cc = M.flat[g1].type == G_CYLINDER && M.flat[g2].type == G_CYLINDER;
                        if (cc) {
                            const bool hips = (a - 1) % 3 == 0 && (b - 1) % 3 == 0 && a >= 1;
                            const int la = (a - 1) / 3, lb = (b - 1) / 3;           // legs FR FL RR RL
                            if (!(hips && la / 2 == lb / 2)) continue;
                        }
                        const int lo = std::min(M.flat[g1].type, M.flat[g2].type), hi = std::max(M.flat[g1].type, M.flat[g2].type);
                        GO1_REQ(cc || lo == G_SPHERE || lo == G_CAPSULE, "robot-robot pair without a sphere or capsule is not implemented");
                        (void)hi;
                        GO1_REQ(M.ngp < NGP, "too many candidate geom pairs");
                        M.gp_a[M.ngp] = (unsigned char)g1; M.gp_b[M.ngp] = (unsigned char)g2; M.ngp++;
                    }
            }
        for (int ka = 0; ka < NKIND; ka++)
            for (int kb = 0; kb < NKIND; kb++) {
                M.pair_cls[ka][kb] = 0;
                if (ka < (int)kind_rep.size() && kb < (int)kind_rep.size()) {
                    int ci = class_of(kind_rep[ka], kind_rep[kb]);
                    GO1_REQ(ci >= 0, class_err.c_str());
                    M.pair_cls[ka][kb] = ci;
                    M.pair_margin = std::fmax(M.pair_margin, M.cls[ci].margin);
                }
            }
    }
    // block-arrow index tables
    {
        ArrowTab& t = M.tab;
        std::memset(&t, 0, sizeof(t));
        const int ci[21] = {0, 3, 1, 4, 5, 2, 0, 8, 7, 9, 8, 0, 6, 0, 9, 7, 6, 0, 0, 0, 9};
        const int cs[21] = {1, 1, 1, 1, 1, 1, 0, 1, -1, 1, -1, 0, 1, 0, 1, 1, -1, 0, 0, 0, 1};
        for (int e = 0; e < 21; e++) { t.crb_idx[e] = (signed char)ci[e]; t.crb_sgn[e] = (signed char)cs[e]; }
        for (int r = 0; r < 6; r++)
            for (int c = 0; c <= r; c++) { int e = r * (r + 1) / 2 + c; t.sym_r[e] = r; t.sym_c[e] = c; t.ent_u[e] = r; t.ent_c[e] = c; }
        for (int k = 0; k < NLEG; k++) {
            for (int j = 0; j < 6; j++)
                for (int i = 0; i < 3; i++) { int e = 21 + 18 * k + 3 * j + i; t.ent_u[e] = 6 + 3 * k + i; t.ent_c[e] = j; }
            for (int i = 0; i < 3; i++)
                for (int c = 0; c <= i; c++) { int e = 93 + 6 * k + i * (i + 1) / 2 + c; t.ent_u[e] = 6 + 3 * k + i; t.ent_c[e] = 6 + 3 * k + c; }
        }
    }
    return true;
}

// Foot sites (the first site attached to each calf; go1.xml names them FR, FL, RR, RL): position in the calf frame.
// Only the biped terms read them (go1_mujoco_env.py:360, :409), so they travel in EnvCfg, not in DevModel.
inline bool read_foot_sites(const void* data, size_t nbytes, double out[NLEG][3], std::string& err) {
    Blob blob; blob.b = (const uint8_t*)data; blob.n = nbytes;
    int ns = 0;
    const int* bodyid = blob.i("site_bodyid", &ns);
    const double* pos = blob.d("site_pos");
    if (!bodyid || !pos) { err = blob.err; return false; }
    for (int k = 0; k < NLEG; k++) {
        int found = -1;
        for (int s = 0; s < ns && found < 0; s++)
            if (bodyid[s] == 4 + 3 * k) found = s;      // MuJoCo body ids: world 0, trunk 1, leg k: hip 2+3k, thigh 3+3k, calf 4+3k
        if (found < 0) { err = "model has no foot site on a calf body"; return false; }
        for (int e = 0; e < 3; e++) out[k][e] = pos[3 * found + e];
    }
    return true;
}

template <typename T>
inline void cast_dev_model(const DevModel<double>& S, DevModel<T>& D) {
    // field-wise cast: both structs have identical layout up to the scalar type, walk them as arrays
    std::memset(&D, 0, sizeof(D));
#define CP(f) D.f = (T)S.f
#define CPA(f) do { const double* s_ = (const double*)&S.f; T* d_ = (T*)&D.f; for (size_t i_ = 0; i_ < sizeof(S.f) / sizeof(double); i_++) d_[i_] = (T)s_[i_]; } while (0)
#define CPI(f) std::memcpy(&D.f, &S.f, sizeof(S.f))
    CP(timestep); CPA(gravity); CP(impratio); CP(tolerance); CP(ls_tolerance); CP(meaninertia); CP(total_mass); CP(lambda_lb);
    CP(iterations); CP(ls_iterations); CP(nclass);
    CPA(plane_pos); CPA(plane_frame); CP(t_mass); CPA(t_ipos); CPA(t_Ib); CPA(t_invw);
    CPA(body_pos); CPA(jnt_axis); CPA(ipos); CPA(mass); CPA(Ib); CPA(invw); CPA(qpos0); CPA(damping); CPA(armature);
    CPA(floss); CPA(Rf); CPA(Bf); CPA(lim_lo); CPA(lim_hi); CPA(lim_K); CPA(lim_B); CPA(lim_imp); CPA(lim_margin); CPA(dof_invw);
    CPI(limited); CPI(act_id); CPI(act_climited); CPI(act_flimited); CPI(act_bias);
    CPA(act_gear); CPA(act_gain); CPA(act_b0); CPA(act_b1); CPA(act_b2); CPA(act_clo); CPA(act_chi); CPA(act_flo); CPA(act_fhi);
    D.nflat = S.nflat;
    CPA(flat_rb); CPI(tab);
    D.ngp = S.ngp; CPI(pair_cls); CP(pair_margin); CPI(gp_a); CPI(gp_b);
    for (int g = 0; g < GFLAT; g++) {
        const DevGeom<double>& s = S.flat[g]; DevGeom<T>& d = D.flat[g];
        d.type = s.type; d.slot = s.slot; d.cls = s.cls; d.gid = s.gid; d.kind = s.kind;
        for (int c = 0; c < 3; c++) { d.pos[c] = (T)s.pos[c]; d.zax[c] = (T)s.zax[c]; d.xax[c] = (T)s.xax[c]; d.size[c] = (T)s.size[c]; }
    }
    for (int c = 0; c < NCLS; c++) {
        const DevClass<double>& s = S.cls[c]; DevClass<T>& d = D.cls[c];
        d.dim = s.dim; d.mu = (T)s.mu; d.K = (T)s.K; d.B = (T)s.B; d.d0 = (T)s.d0; d.d1 = (T)s.d1; d.width = (T)s.width;
        d.mid = (T)s.mid; d.power = (T)s.power; d.margin = (T)s.margin; d.incmargin = (T)s.incmargin;
        for (int j = 0; j < 5; j++) d.fri[j] = (T)s.fri[j];
        for (int j = 0; j < 6; j++) d.drat[j] = (T)s.drat[j];
    }
    CPA(key_qpos); CPA(key_ctrl); CPA(soft_lo); CPA(soft_hi); CPA(ctrl_lo); CPA(ctrl_hi);
#undef CP
#undef CPA
#undef CPI
}

}  // namespace go1
