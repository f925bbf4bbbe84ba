/*
 * go1_oracle.c -- TEST INFRASTRUCTURE ONLY (never linked into or called by the product path).
 *
 * fp64 CPU restatement of the physics the reference runs through
 *     go1_mujoco_env.py:481  self.do_simulation(action, self.frame_skip)
 *        -> gymnasium MujocoEnv: data.ctrl[:] = a; mujoco.mj_step(model, data, nstep=10);
 *           mujoco.mj_rnePostConstraint(model, data)
 * i.e. the third-party dependency `mujoco==2.3.0` (reference requirements.txt:81), which is
 * NOT vendored in /root/reference and not installable in this image.  The algorithm below is
 * a restatement of MuJoCo's published pipeline (engine_forward.c, engine_core_smooth.c,
 * engine_collision_primitive.c, engine_core_constraint.c, engine_solver.c, engine_passive.c,
 * engine_support.c, engine_util_spatial.c; function names are kept in comments) for the
 * feature set the Go1 MJCF selects: free + hinge joints, plane-vs-{sphere,capsule,cylinder,
 * box} contacts, elliptic cones (impratio), joint limits, dof friction loss, joint-space
 * motors / position servos, Newton solver, Euler integrator with implicit joint damping.
 *
 * PARITY UNPINNED vs MuJoCo itself: no MuJoCo binary, golden vector or test exists in the
 * reference for this path.  The oracle is pinned instead by (i) the notebook's
 * model constants (tests/test_mjcf.py), (ii) physics invariants (tests/test_oracle.py: mass
 * matrix vs independent Jacobian formula, momentum/energy, KKT residual, cone membership).
 *
 * Robot-robot pairs (every collision geom has contype = conaffinity = 1, go1_torque.xml:9,26) are
 * generated for all bodies that are not parent and child; pairs with a cylinder use the exact
 * closest-feature solution in place of MuJoCo 2.3.0's MPR (see the narrowphase section).
 *
 * Build: see oracle/Makefile (gcc -O2 -shared -fPIC).
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define mjMINVAL 1e-15
#define mjMINIMP 0.0001
#define mjMAXIMP 0.9999
#define mjMINMU 1e-5
#define mjMAXVAL 1e10
#define MAXCON 256
#define MAXEFC 1200

enum { GEOM_PLANE = 0, GEOM_SPHERE = 2, GEOM_CAPSULE = 3, GEOM_CYLINDER = 5, GEOM_BOX = 6, GEOM_MESH = 7 };
enum { JNT_FREE = 0, JNT_BALL = 1, JNT_SLIDE = 2, JNT_HINGE = 3 };
enum { CN_FRICTION_DOF = 1, CN_LIMIT_JOINT = 3, CN_CONTACT_FRICTIONLESS = 5, CN_CONTACT_ELLIPTIC = 7 };
enum { ST_SATISFIED = 0, ST_QUADRATIC = 1, ST_LINEARNEG = 2, ST_LINEARPOS = 3, ST_CONE = 4 };

typedef struct {
    int nq, nv, nu, nbody, njnt, ngeom, nsite, nkey, nM;
    double timestep, gravity[3], impratio, tolerance, ls_tolerance, meaninertia;
    int iterations, ls_iterations, cone;
    const int *body_parentid, *body_rootid, *body_jntnum, *body_jntadr, *body_dofnum, *body_dofadr;
    const int *body_geomnum, *body_geomadr;
    const double *body_pos, *body_quat, *body_ipos, *body_iquat, *body_mass, *body_inertia, *body_invweight0;
    const int *jnt_type, *jnt_qposadr, *jnt_dofadr, *jnt_bodyid, *jnt_limited;
    const double *jnt_pos, *jnt_axis, *jnt_range, *jnt_solref, *jnt_solimp, *jnt_margin;
    const int *dof_bodyid, *dof_jntid, *dof_parentid, *dof_Madr;
    const double *dof_armature, *dof_damping, *dof_frictionloss, *dof_invweight0, *dof_solref, *dof_solimp;
    const double *qpos0;
    const int *geom_type, *geom_contype, *geom_conaffinity, *geom_condim, *geom_bodyid, *geom_priority;
    const double *geom_size, *geom_pos, *geom_quat, *geom_friction, *geom_margin, *geom_gap, *geom_solmix;
    const double *geom_solref, *geom_solimp, *geom_rbound;
    const int *site_bodyid;
    const double *site_pos, *site_quat;
    const int *actuator_trnid, *actuator_ctrllimited, *actuator_forcelimited, *actuator_biastype;
    const double *actuator_gear, *actuator_gainprm, *actuator_biasprm, *actuator_ctrlrange, *actuator_forcerange;
    const double *key_qpos, *key_qvel, *key_ctrl;
    void *blob; /* owned copy */
} O1Model;

typedef struct {
    double dist, pos[3], frame[9], includemargin, friction[5], solref[2], solimp[5], mu, H[36];
    int dim, geom1, geom2, exclude, efc_address;
} O1Contact;

typedef struct {
    const O1Model *m;
    double time;
    double *qpos, *qvel, *ctrl, *qacc, *qacc_warmstart, *qacc_smooth;
    double *qfrc_bias, *qfrc_passive, *qfrc_actuator, *qfrc_smooth, *qfrc_constraint, *actuator_force;
    double *xpos, *xquat, *xmat, *xipos, *ximat, *xanchor, *xaxis, *geom_xpos, *geom_xmat, *site_xpos, *site_xmat;
    double *subtree_com, *cinert, *crb, *cdof, *cdof_dot, *cvel, *cacc, *cfrc_body, *cfrc_ext;
    double *qM, *qLD, *qLDiagInv, *qH, *qHDiagInv; /* qH: factor of M + h*diag(damping) */
    O1Contact *contact;
    int ncon, nefc, ne, nf;
    double *efc_J, *efc_pos, *efc_margin, *efc_frictionloss, *efc_diagApprox, *efc_R, *efc_D, *efc_KBIP;
    double *efc_vel, *efc_aref, *efc_force;
    int *efc_type, *efc_id, *efc_state;
    int solver_iter, ls_total, warn_bad, warn_contact_full, warn_unsupported_pair, self_collision;
    double solver_improvement, solver_gradient;
    /* optional overrides used by tests */
    double tolerance_override; /* <0: use model */
    int flop_count_enabled;
    double flops;
} O1Data;

#define EXPORT __attribute__((visibility("default")))

/* ------------------------------------------------------------------------------------------ */
/* blob parsing (format: quadruped_rl_locomotion_b200/mjcf.py Model.to_blob)                    */
static const void *blob_find(const uint8_t *b, size_t nbytes, const char *name, int *count) {
    uint32_t n;
    if (nbytes < 16 || memcmp(b, "GO1MDL01", 8) != 0) return NULL;
    memcpy(&n, b + 8, 4);
    const uint8_t *p = b + 16;
    for (uint32_t i = 0; i < n; i++, p += 88) {
        if (strncmp((const char *)p, name, 48) == 0) {
            int32_t code, ndim, shape[4];
            int64_t off, nb;
            memcpy(&code, p + 48, 4); memcpy(&ndim, p + 52, 4); memcpy(shape, p + 56, 16);
            memcpy(&off, p + 72, 8); memcpy(&nb, p + 80, 8);
            int isz = code == 0 ? 8 : (code == 1 ? 4 : 1);
            if (count) *count = (int)(nb / isz);
            (void)ndim;
            return b + off;
        }
    }
    return NULL;
}

#define GETD(field) do { m->field = (const double *)blob_find(b, nbytes, #field, NULL); if (!m->field) { fprintf(stderr, "oracle: blob lacks %s\n", #field); goto fail; } } while (0)
#define GETI(field) do { m->field = (const int *)blob_find(b, nbytes, #field, NULL); if (!m->field) { fprintf(stderr, "oracle: blob lacks %s\n", #field); goto fail; } } while (0)
#define GETSI(dst, name) do { const int *p_ = (const int *)blob_find(b, nbytes, name, NULL); if (!p_) { fprintf(stderr, "oracle: blob lacks %s\n", name); goto fail; } dst = *p_; } while (0)
#define GETSD(dst, name) do { const double *p_ = (const double *)blob_find(b, nbytes, name, NULL); if (!p_) { fprintf(stderr, "oracle: blob lacks %s\n", name); goto fail; } dst = *p_; } while (0)

EXPORT O1Model *o1_model_load(const void *blob, size_t nbytes) {
    O1Model *m = (O1Model *)calloc(1, sizeof(O1Model));
    m->blob = malloc(nbytes);
    memcpy(m->blob, blob, nbytes);
    const uint8_t *b = (const uint8_t *)m->blob;
    GETSI(m->nq, "nq"); GETSI(m->nv, "nv"); GETSI(m->nu, "nu"); GETSI(m->nbody, "nbody"); GETSI(m->njnt, "njnt");
    GETSI(m->ngeom, "ngeom"); GETSI(m->nsite, "nsite"); GETSI(m->nkey, "nkey"); GETSI(m->nM, "nM");
    GETSD(m->timestep, "opt_timestep"); GETSD(m->impratio, "opt_impratio"); GETSD(m->tolerance, "opt_tolerance");
    GETSD(m->ls_tolerance, "opt_ls_tolerance"); GETSD(m->meaninertia, "stat_meaninertia");
    GETSI(m->iterations, "opt_iterations"); GETSI(m->ls_iterations, "opt_ls_iterations"); GETSI(m->cone, "opt_cone");
    {
        const double *g = (const double *)blob_find(b, nbytes, "opt_gravity", NULL);
        if (!g) goto fail;
        memcpy(m->gravity, g, 24);
        int integ, solver;
        GETSI(integ, "opt_integrator"); GETSI(solver, "opt_solver");
        if (integ != 0 || solver != 2 || m->cone != 1) {
            fprintf(stderr, "oracle: only Euler + Newton + elliptic is restated (got integrator %d solver %d cone %d)\n", integ, solver, m->cone);
            goto fail;
        }
    }
    GETI(body_parentid); GETI(body_rootid); GETI(body_jntnum); GETI(body_jntadr); GETI(body_dofnum); GETI(body_dofadr);
    GETI(body_geomnum); GETI(body_geomadr);
    GETD(body_pos); GETD(body_quat); GETD(body_ipos); GETD(body_iquat); GETD(body_mass); GETD(body_inertia); GETD(body_invweight0);
    GETI(jnt_type); GETI(jnt_qposadr); GETI(jnt_dofadr); GETI(jnt_bodyid); GETI(jnt_limited);
    GETD(jnt_pos); GETD(jnt_axis); GETD(jnt_range); GETD(jnt_solref); GETD(jnt_solimp); GETD(jnt_margin);
    GETI(dof_bodyid); GETI(dof_jntid); GETI(dof_parentid); GETI(dof_Madr);
    GETD(dof_armature); GETD(dof_damping); GETD(dof_frictionloss); GETD(dof_invweight0); GETD(dof_solref); GETD(dof_solimp);
    GETD(qpos0);
    GETI(geom_type); GETI(geom_contype); GETI(geom_conaffinity); GETI(geom_condim); GETI(geom_bodyid); GETI(geom_priority);
    GETD(geom_size); GETD(geom_pos); GETD(geom_quat); GETD(geom_friction); GETD(geom_margin); GETD(geom_gap); GETD(geom_solmix);
    GETD(geom_solref); GETD(geom_solimp); GETD(geom_rbound);
    GETI(site_bodyid); GETD(site_pos); GETD(site_quat);
    GETI(actuator_trnid); GETI(actuator_ctrllimited); GETI(actuator_forcelimited); GETI(actuator_biastype);
    GETD(actuator_gear); GETD(actuator_gainprm); GETD(actuator_biasprm); GETD(actuator_ctrlrange); GETD(actuator_forcerange);
    GETD(key_qpos); GETD(key_qvel); GETD(key_ctrl);
    return m;
fail:
    free(m->blob);
    free(m);
    return NULL;
}

EXPORT void o1_model_free(O1Model *m) {
    if (!m) return;
    free(m->blob);
    free(m);
}

/* ------------------------------------------------------------------------------------------ */
static double *dalloc(int n) { return (double *)calloc((size_t)(n > 0 ? n : 1), sizeof(double)); }

EXPORT O1Data *o1_data_new(const O1Model *m) {
    O1Data *d = (O1Data *)calloc(1, sizeof(O1Data));
    int nv = m->nv, nb = m->nbody;
    d->m = m;
    d->qpos = dalloc(m->nq); d->qvel = dalloc(nv); d->ctrl = dalloc(m->nu); d->qacc = dalloc(nv);
    d->qacc_warmstart = dalloc(nv); d->qacc_smooth = dalloc(nv);
    d->qfrc_bias = dalloc(nv); d->qfrc_passive = dalloc(nv); d->qfrc_actuator = dalloc(nv); d->qfrc_smooth = dalloc(nv);
    d->qfrc_constraint = dalloc(nv); d->actuator_force = dalloc(m->nu);
    d->xpos = dalloc(3 * nb); d->xquat = dalloc(4 * nb); d->xmat = dalloc(9 * nb); d->xipos = dalloc(3 * nb); d->ximat = dalloc(9 * nb);
    d->xanchor = dalloc(3 * m->njnt); d->xaxis = dalloc(3 * m->njnt);
    d->geom_xpos = dalloc(3 * m->ngeom); d->geom_xmat = dalloc(9 * m->ngeom);
    d->site_xpos = dalloc(3 * m->nsite); d->site_xmat = dalloc(9 * m->nsite);
    d->subtree_com = dalloc(3 * nb); d->cinert = dalloc(10 * nb); d->crb = dalloc(10 * nb);
    d->cdof = dalloc(6 * nv); d->cdof_dot = dalloc(6 * nv); d->cvel = dalloc(6 * nb); d->cacc = dalloc(6 * nb);
    d->cfrc_body = dalloc(6 * nb); d->cfrc_ext = dalloc(6 * nb);
    d->qM = dalloc(m->nM); d->qLD = dalloc(m->nM); d->qLDiagInv = dalloc(nv); d->qH = dalloc(m->nM); d->qHDiagInv = dalloc(nv);
    d->contact = (O1Contact *)calloc(MAXCON, sizeof(O1Contact));
    d->efc_J = dalloc(MAXEFC * nv); d->efc_pos = dalloc(MAXEFC); d->efc_margin = dalloc(MAXEFC);
    d->efc_frictionloss = dalloc(MAXEFC); d->efc_diagApprox = dalloc(MAXEFC); d->efc_R = dalloc(MAXEFC);
    d->efc_D = dalloc(MAXEFC); d->efc_KBIP = dalloc(4 * MAXEFC); d->efc_vel = dalloc(MAXEFC); d->efc_aref = dalloc(MAXEFC);
    d->efc_force = dalloc(MAXEFC);
    d->efc_type = (int *)calloc(MAXEFC, sizeof(int)); d->efc_id = (int *)calloc(MAXEFC, sizeof(int));
    d->efc_state = (int *)calloc(MAXEFC, sizeof(int));
    d->tolerance_override = -1.0;
    d->self_collision = 1;
    return d;
}

EXPORT void o1_data_free(O1Data *d) {
    if (!d) return;
    double **ptrs[] = {&d->qpos, &d->qvel, &d->ctrl, &d->qacc, &d->qacc_warmstart, &d->qacc_smooth, &d->qfrc_bias,
        &d->qfrc_passive, &d->qfrc_actuator, &d->qfrc_smooth, &d->qfrc_constraint, &d->actuator_force, &d->xpos, &d->xquat,
        &d->xmat, &d->xipos, &d->ximat, &d->xanchor, &d->xaxis, &d->geom_xpos, &d->geom_xmat, &d->site_xpos, &d->site_xmat,
        &d->subtree_com, &d->cinert, &d->crb, &d->cdof, &d->cdof_dot, &d->cvel, &d->cacc, &d->cfrc_body, &d->cfrc_ext,
        &d->qM, &d->qLD, &d->qLDiagInv, &d->qH, &d->qHDiagInv, &d->efc_J, &d->efc_pos, &d->efc_margin, &d->efc_frictionloss,
        &d->efc_diagApprox, &d->efc_R, &d->efc_D, &d->efc_KBIP, &d->efc_vel, &d->efc_aref, &d->efc_force};
    for (size_t i = 0; i < sizeof(ptrs) / sizeof(ptrs[0]); i++) free(*ptrs[i]);
    free(d->contact); free(d->efc_type); free(d->efc_id); free(d->efc_state);
    free(d);
}

/* mj_resetData: qpos<-qpos0, everything else zero (gymnasium MujocoEnv.reset -> _reset_simulation). */
EXPORT void o1_reset_data(O1Data *d) {
    const O1Model *m = d->m;
    int nv = m->nv, nb = m->nbody;
    memcpy(d->qpos, m->qpos0, sizeof(double) * m->nq);
#define Z(p, n) memset(d->p, 0, sizeof(double) * (size_t)(n))
    Z(qvel, nv); Z(ctrl, m->nu); Z(qacc, nv); Z(qacc_warmstart, nv); Z(qacc_smooth, nv); Z(qfrc_bias, nv); Z(qfrc_passive, nv);
    Z(qfrc_actuator, nv); Z(qfrc_smooth, nv); Z(qfrc_constraint, nv); Z(actuator_force, m->nu);
    Z(xpos, 3 * nb); Z(xquat, 4 * nb); Z(xmat, 9 * nb); Z(xipos, 3 * nb); Z(ximat, 9 * nb); Z(xanchor, 3 * m->njnt);
    Z(xaxis, 3 * m->njnt); Z(geom_xpos, 3 * m->ngeom); Z(geom_xmat, 9 * m->ngeom); Z(site_xpos, 3 * m->nsite);
    Z(site_xmat, 9 * m->nsite); Z(subtree_com, 3 * nb); Z(cinert, 10 * nb); Z(crb, 10 * nb); Z(cdof, 6 * nv);
    Z(cdof_dot, 6 * nv); Z(cvel, 6 * nb); Z(cacc, 6 * nb); Z(cfrc_body, 6 * nb); Z(cfrc_ext, 6 * nb);
    Z(qM, m->nM); Z(qLD, m->nM); Z(qLDiagInv, nv); Z(qH, m->nM); Z(qHDiagInv, nv);
#undef Z
    d->time = 0;
    d->ncon = d->nefc = d->ne = d->nf = 0;
    d->solver_iter = 0;
}

/* ------------------------------------------------------------------------------------------ */
/* small vector helpers                                                                         */
static inline double dot3(const double *a, const double *b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
static inline void cross3(double *r, const double *a, const double *b) {
    double x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
    r[0] = x; r[1] = y; r[2] = z;
}
static inline double norm3(const double *a) { return sqrt(dot3(a, a)); }
static double normalize3(double *v) { /* mju_normalize3 */
    double n = norm3(v);
    if (n < mjMINVAL) { v[0] = 1; v[1] = 0; v[2] = 0; }
    else { v[0] /= n; v[1] /= n; v[2] /= n; }
    return n;
}
static double normalize4(double *q) { /* mju_normalize4 */
    double n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
    if (n < mjMINVAL) { q[0] = 1; q[1] = q[2] = q[3] = 0; }
    else { q[0] /= n; q[1] /= n; q[2] /= n; q[3] /= n; }
    return n;
}
static void mulQuat(double *r, const double *a, const double *b) {
    double t[4] = {a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3],
                   a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2],
                   a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1],
                   a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0]};
    memcpy(r, t, sizeof t);
}
static void quat2Mat(double *R, const double *q) { /* row-major 3x3 */
    double q00 = q[0] * q[0], q01 = q[0] * q[1], q02 = q[0] * q[2], q03 = q[0] * q[3];
    double q11 = q[1] * q[1], q12 = q[1] * q[2], q13 = q[1] * q[3], q22 = q[2] * q[2], q23 = q[2] * q[3], q33 = q[3] * q[3];
    R[0] = q00 + q11 - q22 - q33; R[4] = q00 - q11 + q22 - q33; R[8] = q00 - q11 - q22 + q33;
    R[1] = 2 * (q12 - q03); R[2] = 2 * (q13 + q02); R[3] = 2 * (q12 + q03);
    R[5] = 2 * (q23 - q01); R[6] = 2 * (q13 - q02); R[7] = 2 * (q23 + q01);
}
static void rotVecQuat(double *r, const double *v, const double *q) {
    double R[9]; quat2Mat(R, q);
    double t[3] = {R[0] * v[0] + R[1] * v[1] + R[2] * v[2], R[3] * v[0] + R[4] * v[1] + R[5] * v[2], R[6] * v[0] + R[7] * v[1] + R[8] * v[2]};
    memcpy(r, t, sizeof t);
}
static void mulMatVec3(double *r, const double *R, const double *v) {
    double t[3] = {R[0] * v[0] + R[1] * v[1] + R[2] * v[2], R[3] * v[0] + R[4] * v[1] + R[5] * v[2], R[6] * v[0] + R[7] * v[1] + R[8] * v[2]};
    memcpy(r, t, sizeof t);
}
static void axisAngle2Quat(double *q, const double *axis, double angle) {
    if (angle == 0) { q[0] = 1; q[1] = q[2] = q[3] = 0; return; }
    double s = sin(angle * 0.5);
    q[0] = cos(angle * 0.5); q[1] = axis[0] * s; q[2] = axis[1] * s; q[3] = axis[2] * s;
}

/* mju_mulInertVec: 10-vector inertia [Ixx Iyy Izz Ixy Ixz Iyz mx my mz m] times spatial motion [w; v] */
static void mulInertVec(double *res, const double *i, const double *v) {
    double r[6];
    r[0] = i[0] * v[0] + i[3] * v[1] + i[4] * v[2] - i[8] * v[4] + i[7] * v[5];
    r[1] = i[3] * v[0] + i[1] * v[1] + i[5] * v[2] + i[8] * v[3] - i[6] * v[5];
    r[2] = i[4] * v[0] + i[5] * v[1] + i[2] * v[2] - i[7] * v[3] + i[6] * v[4];
    r[3] = i[8] * v[1] - i[7] * v[2] + i[9] * v[3];
    r[4] = i[6] * v[2] - i[8] * v[0] + i[9] * v[4];
    r[5] = i[7] * v[0] - i[6] * v[1] + i[9] * v[5];
    memcpy(res, r, sizeof r);
}
/* mju_crossMotion: res = vel x_m v */
static void crossMotion(double *res, const double *vel, const double *v) {
    double r[6];
    cross3(r, vel, v);
    double a[3], b[3];
    cross3(a, vel, v + 3); cross3(b, vel + 3, v);
    r[3] = a[0] + b[0]; r[4] = a[1] + b[1]; r[5] = a[2] + b[2];
    memcpy(res, r, sizeof r);
}
/* mju_crossForce: res = vel x_f f */
static void crossForce(double *res, const double *vel, const double *f) {
    double r[6], a[3], b[3];
    cross3(a, vel, f); cross3(b, vel + 3, f + 3);
    r[0] = a[0] + b[0]; r[1] = a[1] + b[1]; r[2] = a[2] + b[2];
    cross3(r + 3, vel, f + 3);
    memcpy(res, r, sizeof r);
}

/* ------------------------------------------------------------------------------------------ */
/* mj_kinematics                                                                                */
static void o1_kinematics(O1Data *d) {
    const O1Model *m = d->m;
    d->xquat[0] = 1; d->xmat[0] = d->xmat[4] = d->xmat[8] = 1;
    d->ximat[0] = d->ximat[4] = d->ximat[8] = 1;
    for (int i = 1; i < m->nbody; i++) {
        double pos[3], quat[4];
        int jn = m->body_jntnum[i], ja = m->body_jntadr[i], p = m->body_parentid[i];
        if (jn == 1 && m->jnt_type[ja] == JNT_FREE) {
            int qa = m->jnt_qposadr[ja];
            normalize4(d->qpos + qa + 3); /* normalised in place, as mj_kinematics does */
            memcpy(pos, d->qpos + qa, 24); memcpy(quat, d->qpos + qa + 3, 32);
            memcpy(d->xanchor + 3 * ja, pos, 24);
            d->xaxis[3 * ja] = 0; d->xaxis[3 * ja + 1] = 0; d->xaxis[3 * ja + 2] = 1;
        } else {
            mulMatVec3(pos, d->xmat + 9 * p, m->body_pos + 3 * i);
            for (int k = 0; k < 3; k++) pos[k] += d->xpos[3 * p + k];
            mulQuat(quat, d->xquat + 4 * p, m->body_quat + 4 * i);
            for (int j = ja; j < ja + jn; j++) {
                int qa = m->jnt_qposadr[j];
                double v[3];
                rotVecQuat(v, m->jnt_pos + 3 * j, quat);
                for (int k = 0; k < 3; k++) d->xanchor[3 * j + k] = pos[k] + v[k];
                rotVecQuat(d->xaxis + 3 * j, m->jnt_axis + 3 * j, quat);
                if (m->jnt_type[j] == JNT_HINGE) {
                    double qloc[4];
                    axisAngle2Quat(qloc, m->jnt_axis + 3 * j, d->qpos[qa] - m->qpos0[qa]);
                    mulQuat(quat, quat, qloc);
                    rotVecQuat(v, m->jnt_pos + 3 * j, quat);
                    for (int k = 0; k < 3; k++) pos[k] = d->xanchor[3 * j + k] - v[k];
                } else if (m->jnt_type[j] == JNT_SLIDE) {
                    for (int k = 0; k < 3; k++) pos[k] += d->xaxis[3 * j + k] * (d->qpos[qa] - m->qpos0[qa]);
                }
            }
        }
        normalize4(quat);
        memcpy(d->xpos + 3 * i, pos, 24); memcpy(d->xquat + 4 * i, quat, 32);
        quat2Mat(d->xmat + 9 * i, quat);
        double v[3], q[4];
        mulMatVec3(v, d->xmat + 9 * i, m->body_ipos + 3 * i);
        for (int k = 0; k < 3; k++) d->xipos[3 * i + k] = pos[k] + v[k];
        mulQuat(q, quat, m->body_iquat + 4 * i);
        quat2Mat(d->ximat + 9 * i, q);
    }
    for (int g = 0; g < m->ngeom; g++) {
        int b = m->geom_bodyid[g];
        double v[3], q[4];
        mulMatVec3(v, d->xmat + 9 * b, m->geom_pos + 3 * g);
        for (int k = 0; k < 3; k++) d->geom_xpos[3 * g + k] = d->xpos[3 * b + k] + v[k];
        mulQuat(q, d->xquat + 4 * b, m->geom_quat + 4 * g);
        quat2Mat(d->geom_xmat + 9 * g, q);
    }
    for (int s = 0; s < m->nsite; s++) {
        int b = m->site_bodyid[s];
        double v[3], q[4];
        mulMatVec3(v, d->xmat + 9 * b, m->site_pos + 3 * s);
        for (int k = 0; k < 3; k++) d->site_xpos[3 * s + k] = d->xpos[3 * b + k] + v[k];
        mulQuat(q, d->xquat + 4 * b, m->site_quat + 4 * s);
        quat2Mat(d->site_xmat + 9 * s, q);
    }
}

/* mj_comPos: subtree COM, cinert (about the root subtree COM, world axes), cdof */
static void o1_comPos(O1Data *d) {
    const O1Model *m = d->m;
    int nb = m->nbody;
    double *mass_sub = (double *)alloca(sizeof(double) * nb);
    for (int i = 0; i < nb; i++) {
        mass_sub[i] = m->body_mass[i];
        for (int k = 0; k < 3; k++) d->subtree_com[3 * i + k] = m->body_mass[i] * d->xipos[3 * i + k];
    }
    for (int i = nb - 1; i > 0; i--) {
        int p = m->body_parentid[i];
        mass_sub[p] += mass_sub[i];
        for (int k = 0; k < 3; k++) d->subtree_com[3 * p + k] += d->subtree_com[3 * i + k];
    }
    for (int i = 0; i < nb; i++) {
        if (mass_sub[i] < mjMINVAL) memcpy(d->subtree_com + 3 * i, d->xipos + 3 * i, 24);
        else for (int k = 0; k < 3; k++) d->subtree_com[3 * i + k] /= mass_sub[i];
    }
    memset(d->cinert, 0, sizeof(double) * 10);
    for (int i = 1; i < nb; i++) {
        const double *com = d->subtree_com + 3 * m->body_rootid[i];
        const double *R = d->ximat + 9 * i, *I = m->body_inertia + 3 * i;
        double dif[3], mass = m->body_mass[i], *res = d->cinert + 10 * i;
        for (int k = 0; k < 3; k++) dif[k] = d->xipos[3 * i + k] - com[k];
        /* mju_inertCom: R diag(I) R^T + m (|dif|^2 1 - dif dif^T) */
        double t[9];
        for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) t[3 * r + c] = R[3 * r + c] * I[c];
        res[0] = t[0] * R[0] + t[1] * R[1] + t[2] * R[2];
        res[1] = t[3] * R[3] + t[4] * R[4] + t[5] * R[5];
        res[2] = t[6] * R[6] + t[7] * R[7] + t[8] * R[8];
        res[3] = t[0] * R[3] + t[1] * R[4] + t[2] * R[5];
        res[4] = t[0] * R[6] + t[1] * R[7] + t[2] * R[8];
        res[5] = t[3] * R[6] + t[4] * R[7] + t[5] * R[8];
        res[0] += mass * (dif[1] * dif[1] + dif[2] * dif[2]);
        res[1] += mass * (dif[0] * dif[0] + dif[2] * dif[2]);
        res[2] += mass * (dif[0] * dif[0] + dif[1] * dif[1]);
        res[3] -= mass * dif[0] * dif[1];
        res[4] -= mass * dif[0] * dif[2];
        res[5] -= mass * dif[1] * dif[2];
        res[6] = mass * dif[0]; res[7] = mass * dif[1]; res[8] = mass * dif[2];
        res[9] = mass;
    }
    for (int j = 0; j < m->njnt; j++) {
        int b = m->jnt_bodyid[j], da = m->jnt_dofadr[j];
        double off[3];
        for (int k = 0; k < 3; k++) off[k] = d->subtree_com[3 * m->body_rootid[b] + k] - d->xanchor[3 * j + k];
        if (m->jnt_type[j] == JNT_FREE) {
            memset(d->cdof + 6 * da, 0, sizeof(double) * 18);
            for (int k = 0; k < 3; k++) d->cdof[6 * (da + k) + 3 + k] = 1;
            for (int k = 0; k < 3; k++) {
                double ax[3] = {d->xmat[9 * b + k], d->xmat[9 * b + 3 + k], d->xmat[9 * b + 6 + k]};
                double *c = d->cdof + 6 * (da + 3 + k);
                memcpy(c, ax, 24);
                cross3(c + 3, ax, off);
            }
        } else if (m->jnt_type[j] == JNT_HINGE) {
            double *c = d->cdof + 6 * da;
            memcpy(c, d->xaxis + 3 * j, 24);
            cross3(c + 3, d->xaxis + 3 * j, off);
        } else { /* slide */
            double *c = d->cdof + 6 * da;
            c[0] = c[1] = c[2] = 0;
            memcpy(c + 3, d->xaxis + 3 * j, 24);
        }
    }
}

/* mj_crb */
static void o1_crb(O1Data *d) {
    const O1Model *m = d->m;
    int nv = m->nv, nb = m->nbody;
    memcpy(d->crb, d->cinert, sizeof(double) * 10 * nb);
    for (int i = nb - 1; i > 0; i--) {
        int p = m->body_parentid[i];
        if (p > 0) for (int k = 0; k < 10; k++) d->crb[10 * p + k] += d->crb[10 * i + k];
    }
    memset(d->qM, 0, sizeof(double) * m->nM);
    for (int i = 0; i < nv; i++) {
        double buf[6];
        int adr = m->dof_Madr[i];
        mulInertVec(buf, d->crb + 10 * m->dof_bodyid[i], d->cdof + 6 * i);
        d->qM[adr] = m->dof_armature[i];
        for (int j = i; j >= 0; j = m->dof_parentid[j]) {
            double s = 0;
            for (int k = 0; k < 6; k++) s += d->cdof[6 * j + k] * buf[k];
            d->qM[adr++] += s;
        }
    }
}

/* sparse L'DL factorisation (mj_factorM layout) of an arbitrary tree-structured matrix */
static void factor_tree(const O1Model *m, double *LD, double *diaginv) {
    int nv = m->nv;
    for (int k = nv - 1; k >= 0; k--) {
        int kk = m->dof_Madr[k];
        int i = m->dof_parentid[k], ki = kk + 1;
        while (i >= 0) {
            double tmp = LD[ki] / LD[kk];
            int cnt = m->dof_Madr[i + 1] - m->dof_Madr[i];
            for (int p = 0; p < cnt; p++) LD[m->dof_Madr[i] + p] -= tmp * LD[ki + p];
            LD[ki] = tmp;
            i = m->dof_parentid[i];
            ki++;
        }
    }
    for (int k = 0; k < nv; k++) diaginv[k] = 1.0 / LD[m->dof_Madr[k]];
}
static void solve_tree(const O1Model *m, const double *LD, const double *diaginv, double *x) {
    int nv = m->nv;
    for (int i = nv - 1; i >= 0; i--) {
        if (x[i] != 0) {
            int adr = m->dof_Madr[i] + 1;
            for (int j = m->dof_parentid[i]; j >= 0; j = m->dof_parentid[j]) x[j] -= LD[adr++] * x[i];
        }
    }
    for (int i = 0; i < nv; i++) x[i] *= diaginv[i];
    for (int i = 0; i < nv; i++) {
        int adr = m->dof_Madr[i] + 1;
        for (int j = m->dof_parentid[i]; j >= 0; j = m->dof_parentid[j]) x[i] -= LD[adr++] * x[j];
    }
}
/* res = M * vec (sparse) */
static void mulM(const O1Model *m, const double *qM, double *res, const double *vec) {
    int nv = m->nv;
    for (int i = 0; i < nv; i++) res[i] = 0;
    for (int i = 0; i < nv; i++) {
        int adr = m->dof_Madr[i];
        res[i] += qM[adr] * vec[i];
        int a = adr + 1;
        for (int j = m->dof_parentid[i]; j >= 0; j = m->dof_parentid[j], a++) {
            res[i] += qM[a] * vec[j];
            res[j] += qM[a] * vec[i];
        }
    }
}
static void fullM(const O1Model *m, const double *qM, double *dst) {
    int nv = m->nv;
    memset(dst, 0, sizeof(double) * nv * nv);
    for (int i = 0; i < nv; i++) {
        int a = m->dof_Madr[i];
        for (int j = i; j >= 0; j = m->dof_parentid[j], a++) dst[i * nv + j] = dst[j * nv + i] = qM[a];
    }
}

/* ------------------------------------------------------------------------------------------ */
/* collision: plane vs primitives (engine_collision_primitive.c)                               */
typedef struct { double dist, pos[3], frame[9]; } RawCon;

static int planeSphereRaw(const double *ppos, const double *pn, const double *c, double r, double margin, RawCon *con) {
    double t[3] = {c[0] - ppos[0], c[1] - ppos[1], c[2] - ppos[2]};
    double cdist = dot3(t, pn);
    if (cdist > margin + r) return 0;
    con->dist = cdist - r;
    for (int k = 0; k < 3; k++) con->pos[k] = c[k] + pn[k] * (-con->dist * 0.5 - r);
    memset(con->frame, 0, sizeof con->frame);
    memcpy(con->frame, pn, 24);
    return 1;
}
static int mjc_PlaneSphere(const double *ppos, const double *pmat, const double *gpos, const double *gsize, double margin, RawCon *con) {
    double n[3] = {pmat[2], pmat[5], pmat[8]};
    return planeSphereRaw(ppos, n, gpos, gsize[0], margin, con);
}
static int mjc_PlaneCapsule(const double *ppos, const double *pmat, const double *gpos, const double *gmat, const double *gsize, double margin, RawCon *con) {
    double n[3] = {pmat[2], pmat[5], pmat[8]};
    double axis[3] = {gmat[2], gmat[5], gmat[8]};
    double c1[3], c2[3];
    for (int k = 0; k < 3; k++) { c1[k] = gpos[k] + axis[k] * gsize[1]; c2[k] = gpos[k] - axis[k] * gsize[1]; }
    int n1 = planeSphereRaw(ppos, n, c1, gsize[0], margin, con);
    int n2 = planeSphereRaw(ppos, n, c2, gsize[0], margin, con + n1);
    /* align contact frames with capsule axis */
    if (n1) memcpy(con[0].frame + 3, axis, 24);
    if (n2) memcpy(con[n1].frame + 3, axis, 24);
    return n1 + n2;
}
static int mjc_PlaneCylinder(const double *ppos, const double *pmat, const double *gpos, const double *gmat, const double *gsize, double margin, RawCon *con) {
    double n[3] = {pmat[2], pmat[5], pmat[8]};
    double axis[3] = {gmat[2], gmat[5], gmat[8]};
    double prjaxis = dot3(n, axis);
    if (prjaxis > 0) { axis[0] = -axis[0]; axis[1] = -axis[1]; axis[2] = -axis[2]; prjaxis = -prjaxis; }
    double t[3] = {gpos[0] - ppos[0], gpos[1] - ppos[1], gpos[2] - ppos[2]};
    double dist0 = dot3(t, n);
    double vec[3];
    for (int k = 0; k < 3; k++) vec[k] = axis[k] * prjaxis - n[k];
    double len_sqr = dot3(vec, vec);
    if (len_sqr >= mjMINVAL * mjMINVAL) {
        double s = gsize[0] / sqrt(len_sqr);
        for (int k = 0; k < 3; k++) vec[k] *= s;
    } else {
        vec[0] = gmat[0] * gsize[0]; vec[1] = gmat[3] * gsize[0]; vec[2] = gmat[6] * gsize[0];
    }
    double prjvec = dot3(vec, n);
    for (int k = 0; k < 3; k++) axis[k] *= gsize[1];
    prjaxis *= gsize[1];
    int cnt = 0;
    if (dist0 + prjaxis + prjvec <= margin) {
        con[cnt].dist = dist0 + prjaxis + prjvec;
        for (int k = 0; k < 3; k++) con[cnt].pos[k] = gpos[k] + vec[k] + axis[k] - n[k] * con[cnt].dist * 0.5;
        memset(con[cnt].frame, 0, 72); memcpy(con[cnt].frame, n, 24);
        cnt++;
    } else return 0;
    if (dist0 - prjaxis + prjvec <= margin) {
        con[cnt].dist = dist0 - prjaxis + prjvec;
        for (int k = 0; k < 3; k++) con[cnt].pos[k] = gpos[k] + vec[k] - axis[k] - n[k] * con[cnt].dist * 0.5;
        memset(con[cnt].frame, 0, 72); memcpy(con[cnt].frame, n, 24);
        cnt++;
    }
    double prjvec1 = -prjvec * 0.5;
    if (dist0 + prjaxis + prjvec1 <= margin) {
        double vec1[3];
        cross3(vec1, vec, axis);
        normalize3(vec1);
        double s = gsize[0] * sqrt(3.0) * 0.5;
        for (int k = 0; k < 3; k++) vec1[k] *= s;
        double dd = dist0 + prjaxis + prjvec1;
        for (int sgn = 1; sgn >= -1; sgn -= 2) {
            con[cnt].dist = dd;
            for (int k = 0; k < 3; k++) con[cnt].pos[k] = gpos[k] + sgn * vec1[k] + axis[k] - vec[k] * 0.5 - n[k] * dd * 0.5;
            memset(con[cnt].frame, 0, 72); memcpy(con[cnt].frame, n, 24);
            cnt++;
        }
    }
    return cnt;
}
static int mjc_PlaneBox(const double *ppos, const double *pmat, const double *gpos, const double *gmat, const double *gsize, double margin, RawCon *con) {
    double n[3] = {pmat[2], pmat[5], pmat[8]};
    double t[3] = {gpos[0] - ppos[0], gpos[1] - ppos[1], gpos[2] - ppos[2]};
    double dist = dot3(t, n);
    int cnt = 0;
    for (int i = 0; i < 8; i++) {
        double vec[3] = {(i & 1 ? gsize[0] : -gsize[0]), (i & 2 ? gsize[1] : -gsize[1]), (i & 4 ? gsize[2] : -gsize[2])};
        double corner[3];
        mulMatVec3(corner, gmat, vec);
        double ldist = dot3(n, corner);
        if (dist + ldist > margin || ldist > 0) continue;
        con[cnt].dist = dist + ldist;
        for (int k = 0; k < 3; k++) con[cnt].pos[k] = corner[k] + gpos[k] - n[k] * con[cnt].dist * 0.5;
        memset(con[cnt].frame, 0, 72); memcpy(con[cnt].frame, n, 24);
        if (++cnt >= 4) return 4;
    }
    return cnt;
}

/* mju_makeFrame */
static void makeFrame(double *f) {
    normalize3(f);
    if (norm3(f + 3) < 0.5) {
        f[3] = f[4] = f[5] = 0;
        if (f[1] < 0.5 && f[1] > -0.5) f[4] = 1; else f[5] = 1;
    }
    double dd = dot3(f, f + 3);
    for (int k = 0; k < 3; k++) f[3 + k] -= f[k] * dd;
    normalize3(f + 3);
    cross3(f + 6, f, f + 3);
}

/* ------------------------------------------------------------------------------------------ */
/* robot-robot narrowphase (engine_collision_primitive.c / engine_collision_box.c).  The normal points
 * from geom1 to geom2 (type1 <= type2 after mj_collideGeoms' swap).  sphere-sphere, sphere-capsule,
 * capsule-capsule and sphere-box restate MuJoCo's analytic functions [recalled].  MuJoCo 2.3.0 sends every
 * pair with a cylinder other than plane-cylinder through libccd's MPR (mjc_Convex, one contact, tolerance
 * 1e-6), which cannot be reproduced analytically: here those pairs use the exact closest-feature solution
 * that MPR approximates for sphere-cylinder, and for a capsule against a cylinder or a box the centre (box:
 * for long patches both ends) of the CONTACT PATCH of the capsule axis (seg_patch): the closest point alone
 * is ill-conditioned where a capsule lies along a face, the patch is not (MuJoCo's mjc_CapsuleBox also
 * returns up to two contacts; its exact rules are not reproduced).  The CUDA kernel implements exactly the
 * same rules (go1_pair.cuh). */
static int sphereSphereRaw(const double *c1, double r1, const double *c2, double r2, const double *fall1, const double *fall2,
                           double margin, RawCon *con) {
    double dif[3] = {c2[0] - c1[0], c2[1] - c1[1], c2[2] - c1[2]};
    double cd2 = dot3(dif, dif), md = margin + r1 + r2;
    if (cd2 > md * md) return 0;
    double len = sqrt(cd2);
    con->dist = len - r1 - r2;
    memset(con->frame, 0, sizeof con->frame);
    if (len < mjMINVAL) {   /* coincident centres: cross product of the z axes (mjraw_SphereSphere) */
        cross3(con->frame, fall1, fall2);
        normalize3(con->frame);
    } else for (int k = 0; k < 3; k++) con->frame[k] = dif[k] / len;
    for (int k = 0; k < 3; k++) con->pos[k] = c1[k] + con->frame[k] * (r1 + 0.5 * con->dist);
    return 1;
}
static int mjc_SphereCapsule(const double *p1, const double *m1, double r1, const double *p2, const double *m2, const double *s2,
                             double margin, RawCon *con) {
    double ax[3] = {m2[2], m2[5], m2[8]}, v[3] = {p1[0] - p2[0], p1[1] - p2[1], p1[2] - p2[2]};
    double x = fmin(s2[1], fmax(-s2[1], dot3(ax, v)));
    double q[3] = {p2[0] + ax[0] * x, p2[1] + ax[1] * x, p2[2] + ax[2] * x};
    double z1[3] = {m1[2], m1[5], m1[8]};
    return sphereSphereRaw(p1, r1, q, s2[0], z1, ax, margin, con);
}
static int mjc_CapsuleCapsule(const double *p1, const double *m1, const double *s1, const double *p2, const double *m2, const double *s2,
                              double margin, RawCon *con) {
    double z1[3] = {m1[2], m1[5], m1[8]}, z2[3] = {m2[2], m2[5], m2[8]};
    double a1[3] = {z1[0] * s1[1], z1[1] * s1[1], z1[2] * s1[1]}, a2[3] = {z2[0] * s2[1], z2[1] * s2[1], z2[2] * s2[1]};
    double dif[3] = {p1[0] - p2[0], p1[1] - p2[1], p1[2] - p2[2]};
    double ma = dot3(a1, a1), mb = -dot3(a1, a2), mc = dot3(a2, a2), u = -dot3(a1, dif), v = dot3(a2, dif);
    double det = ma * mc - mb * mb;
    double v1[3], v2[3];
    if (fabs(det) >= mjMINVAL) {
        double x1 = (mc * u - mb * v) / det, x2 = (ma * v - mb * u) / det;
        if (x1 > 1) { x1 = 1; x2 = (v - mb) / mc; }
        else if (x1 < -1) { x1 = -1; x2 = (v + mb) / mc; }
        if (x2 > 1) { x2 = 1; x1 = (u - mb) / ma; }
        else if (x2 < -1) { x2 = -1; x1 = (u + mb) / ma; }
        x1 = fmin(1.0, fmax(-1.0, x1));
        for (int k = 0; k < 3; k++) { v1[k] = p1[k] + a1[k] * x1; v2[k] = p2[k] + a2[k] * x2; }
        return sphereSphereRaw(v1, s1[0], v2, s2[0], z1, z2, margin, con);
    }
    /* parallel axes: up to two contacts from the four end-point projections */
    int n = 0;
    for (int e = 0; e < 4 && n < 2; e++) {
        double x1, x2;
        if (e == 0) { x1 = 1; x2 = (v - mb) / mc; x2 = fmin(1.0, fmax(-1.0, x2)); }
        else if (e == 1) { x1 = -1; x2 = (v + mb) / mc; x2 = fmin(1.0, fmax(-1.0, x2)); }
        else if (e == 2) { x2 = 1; x1 = (u - mb) / ma; x1 = fmin(1.0, fmax(-1.0, x1)); }
        else { x2 = -1; x1 = (u + mb) / ma; x1 = fmin(1.0, fmax(-1.0, x1)); }
        for (int k = 0; k < 3; k++) { v1[k] = p1[k] + a1[k] * x1; v2[k] = p2[k] + a2[k] * x2; }
        n += sphereSphereRaw(v1, s1[0], v2, s2[0], z1, z2, margin, con + n);
    }
    return n;
}
/* sphere (centre c, radius r) vs box (p2, m2, half sizes s2): mjc_SphereBox */
static int sphereBoxRaw(const double *c, double r, const double *p2, const double *m2, const double *s2, double margin, RawCon *con) {
    double t[3] = {c[0] - p2[0], c[1] - p2[1], c[2] - p2[2]}, ctr[3], cl[3], dv[3];
    for (int i = 0; i < 3; i++) ctr[i] = m2[i] * t[0] + m2[3 + i] * t[1] + m2[6 + i] * t[2];   /* box frame */
    for (int i = 0; i < 3; i++) { cl[i] = fmin(s2[i], fmax(-s2[i], ctr[i])); dv[i] = cl[i] - ctr[i]; }
    double dist = norm3(dv), nl[3];
    if (dist - r > margin) return 0;
    memset(con->frame, 0, sizeof con->frame);
    double pl[3];
    if (dist <= mjMINVAL) {   /* centre inside the box: leave through the nearest face */
        double closest = 1e30; int k = 0;
        for (int i = 0; i < 6; i++) {
            double fd = fabs((i % 2 ? 1.0 : -1.0) * s2[i / 2] - ctr[i / 2]);
            if (fd < closest) { closest = fd; k = i; }
        }
        nl[0] = nl[1] = nl[2] = 0; nl[k / 2] = (k % 2) ? -1.0 : 1.0;
        con->dist = -closest - r;
        for (int i = 0; i < 3; i++) pl[i] = ctr[i] + nl[i] * (r - closest) * 0.5;
    } else {
        for (int i = 0; i < 3; i++) nl[i] = dv[i] / dist;
        con->dist = dist - r;
        for (int i = 0; i < 3; i++) pl[i] = ctr[i] + nl[i] * (r + 0.5 * con->dist);
    }
    mulMatVec3(con->frame, m2, nl);
    double pw[3]; mulMatVec3(pw, m2, pl);
    for (int k = 0; k < 3; k++) con->pos[k] = pw[k] + p2[k];
    return 1;
}
/* sphere vs solid cylinder (p2, axis = z column of m2, radius s2[0], half length s2[1]) */
static int sphereCylinderRaw(const double *c, double r, const double *p2, const double *m2, const double *s2, double margin, RawCon *con) {
    double ax[3] = {m2[2], m2[5], m2[8]}, t[3] = {c[0] - p2[0], c[1] - p2[1], c[2] - p2[2]};
    double z = dot3(ax, t), w[3] = {t[0] - z * ax[0], t[1] - z * ax[1], t[2] - z * ax[2]};
    double rho = norm3(w), R = s2[0], h = s2[1], e[3];
    if (rho < mjMINVAL) { e[0] = m2[0]; e[1] = m2[3]; e[2] = m2[6]; } else { e[0] = w[0] / rho; e[1] = w[1] / rho; e[2] = w[2] / rho; }
    double dz = fabs(z) - h, dr = rho - R, sz = z < 0 ? -1.0 : 1.0, n[3];
    memset(con->frame, 0, sizeof con->frame);
    if (dz <= 0 && dr <= 0) {     /* centre inside: leave through the nearer of side and cap */
        double closest;
        if (-dr < -dz) { closest = -dr; for (int k = 0; k < 3; k++) n[k] = -e[k]; }
        else { closest = -dz; for (int k = 0; k < 3; k++) n[k] = -sz * ax[k]; }
        if (-closest - r > margin) return 0;
        con->dist = -closest - r;
        for (int k = 0; k < 3; k++) con->pos[k] = c[k] + n[k] * (r - closest) * 0.5;
    } else {
        double zc = fmin(h, fmax(-h, z)), rc = fmin(rho, R), q[3], dv[3];
        for (int k = 0; k < 3; k++) { q[k] = p2[k] + zc * ax[k] + rc * e[k]; dv[k] = q[k] - c[k]; }
        double dist = norm3(dv);
        if (dist - r > margin) return 0;
        if (dist < mjMINVAL) return 0;
        for (int k = 0; k < 3; k++) n[k] = dv[k] / dist;
        con->dist = dist - r;
        for (int k = 0; k < 3; k++) con->pos[k] = c[k] + n[k] * (r + 0.5 * con->dist);
    }
    memcpy(con->frame, n, 24);
    return 1;
}
/* Signed distance of the 2-D point (px, py) to the rectangle |x| <= hx, |y| <= hy and its outward unit normal there. */
static double rectSdf(double px, double py, double hx, double hy, double *gx, double *gy) {
    double qx = fabs(px) - hx, qy = fabs(py) - hy, sx = px < 0 ? -1.0 : 1.0, sy = py < 0 ? -1.0 : 1.0;
    if (qx <= 0 && qy <= 0) {               /* inside: leave through the nearer side */
        if (qx > qy) { *gx = sx; *gy = 0; return qx; }
        *gx = 0; *gy = sy; return qy;
    }
    double ox = fmax(qx, 0.0), oy = fmax(qy, 0.0), d = sqrt(ox * ox + oy * oy);
    *gx = sx * ox / d; *gy = sy * oy / d;
    return d;
}
/* Two solid cylinders whose axes AND centres lie in one plane (the left and right hip motors of this robot: both turn about
 * the trunk's x axis, go1_torque.xml:31-46).  Both shapes are mirror-symmetric about that plane, so a closest pair of points
 * lies in it and the problem is rectangle vs rectangle (half length along the axis, radius across): the closest pair of two
 * convex polygons always contains a vertex, so each rectangle's vertex with the smallest signed distance to the OTHER rectangle
 * is a contact candidate (normal = that rectangle's outward normal there, position half way); each candidate within the margin
 * becomes a contact (one in general, two in near-symmetric poses).  MuJoCo 2.3.0 sends this pair through libccd's MPR (always
 * one contact); DESIGN.md section 1, deviation 1.  Returns -1 when the configuration
 * is not coplanar (no such pair on this robot). */
static int cylinderCylinderCoplanar(const double *p1, const double *m1, const double *s1, const double *p2, const double *m2, const double *s2,
                                    double margin, RawCon *con) {
    double z1[3] = {m1[2], m1[5], m1[8]}, z2[3] = {m2[2], m2[5], m2[8]}, t[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
    double nrm[3], e2[3];
    cross3(nrm, z1, z2);
    if (norm3(nrm) < 1e-6) cross3(nrm, z1, t);                                  /* parallel axes */
    if (norm3(nrm) < 1e-9) { nrm[0] = m1[0]; nrm[1] = m1[3]; nrm[2] = m1[6]; }    /* coaxial: any direction across the axis */
    normalize3(nrm);
    if (fabs(dot3(nrm, t)) > 1e-6 || fabs(dot3(nrm, z2)) > 1e-6) return -1;
    cross3(e2, nrm, z1);
    double cx = dot3(t, z1), cy = dot3(t, e2), ax = dot3(z2, z1), ay = dot3(z2, e2);       /* rectangle 2 in the frame of rectangle 1 */
    /* the deepest vertex of EACH rectangle in the other one gives a contact of its own (no minimum over the two families: in the
     * mirror-symmetric pose, both hips at the same limit, the two candidates tie and rounding would pick the normal) */
    double best[2] = {1e30, 1e30}, bn[2][2] = {{0, 0}, {0, 0}}, bp[2][2] = {{0, 0}, {0, 0}};
    for (int k = 0; k < 4; k++) {           /* vertices of rectangle 2 against rectangle 1 */
        double u = (k & 2) ? -s2[1] : s2[1], v = (k & 1) ? -s2[0] : s2[0];
        double vx = cx + u * ax - v * ay, vy = cy + u * ay + v * ax, gx, gy;
        double d = rectSdf(vx, vy, s1[1], s1[0], &gx, &gy);
        if (d < best[0]) { best[0] = d; bn[0][0] = gx; bn[0][1] = gy; bp[0][0] = vx - 0.5 * d * gx; bp[0][1] = vy - 0.5 * d * gy; }
    }
    for (int k = 0; k < 4; k++) {           /* vertices of rectangle 1 against rectangle 2 (in the frame of rectangle 2) */
        double wx = (k & 2) ? -s1[1] : s1[1], wy = (k & 1) ? -s1[0] : s1[0];
        double rx = wx - cx, ry = wy - cy, lx = rx * ax + ry * ay, ly = -rx * ay + ry * ax, gx, gy;
        double d = rectSdf(lx, ly, s2[1], s2[0], &gx, &gy);
        if (d < best[1]) {
            double nx = gx * ax - gy * ay, ny = gx * ay + gy * ax;                  /* outward normal of rectangle 2, frame 1 */
            best[1] = d; bn[1][0] = -nx; bn[1][1] = -ny; bp[1][0] = wx - 0.5 * d * nx; bp[1][1] = wy - 0.5 * d * ny;
        }
    }
    int num = 0;
    for (int f = 0; f < 2; f++) {
        if (best[f] > margin) continue;
        memset(con[num].frame, 0, sizeof con[num].frame);
        for (int k = 0; k < 3; k++) {
            con[num].frame[k] = bn[f][0] * z1[k] + bn[f][1] * e2[k];
            con[num].pos[k] = p1[k] + bp[f][0] * z1[k] + bp[f][1] * e2[k];
        }
        con[num].dist = best[f];
        num++;
    }
    return num;
}
/* F(t) = squared distance between the capsule-axis point pl + t*al (t in [-1, 1]) and a box / solid cylinder given in ITS frame
 * (cylinder: box == 0, sz = {R, R, h}), g(t) = F'(t) / 2 and g'(t).  F is convex along the line. */
static void seg_eval(const double *pl, const double *al, const double *sz, int box, double t, double *F, double *g, double *dg) {
    double x[3] = {pl[0] + t * al[0], pl[1] + t * al[1], pl[2] + t * al[2]};
    *F = 0; *g = 0; *dg = 0;
    if (box) {
        for (int i = 0; i < 3; i++) {
            double ex = x[i] - fmin(sz[i], fmax(-sz[i], x[i]));
            *F += ex * ex; *g += ex * al[i];
            if (ex != 0) *dg += al[i] * al[i];
        }
        return;
    }
    double ez = x[2] - fmin(sz[2], fmax(-sz[2], x[2]));
    *F = ez * ez; *g = ez * al[2];
    if (ez != 0) *dg = al[2] * al[2];
    double rho = sqrt(x[0] * x[0] + x[1] * x[1]);
    if (rho > sz[0]) {
        double u = (x[0] * al[0] + x[1] * al[1]) / rho, a2 = al[0] * al[0] + al[1] * al[1];
        *F += (rho - sz[0]) * (rho - sz[0]);
        *g += (rho - sz[0]) * u;
        *dg += u * u + (rho - sz[0]) * (a2 - u * u) / rho;
    }
}
/* Contact patch [t1, t2] of the capsule axis: the axis points within `reach` (radius + margin) of the body; 0 when there are none.
 * First a point of minimum distance (root of the monotone g: Newton inside a shrinking bracket); along a face / the cylinder side it
 * is not unique, it only serves as a point inside the patch.  The patch ends are roots of the convex F(t) = reach^2 approached by
 * Newton from outside (monotone).  The kernel runs the identical steps (go1_pair.cuh seg_patch). */
static int seg_patch(const double *pl, const double *al, const double *sz, int box, double reach, double *t1, double *t2) {
    const double eps = 1e-13, r2 = reach * reach;
    double Flo, Fhi, glo, ghi, dg, ts, Fs;
    seg_eval(pl, al, sz, box, -1.0, &Flo, &glo, &dg);
    seg_eval(pl, al, sz, box, 1.0, &Fhi, &ghi, &dg);
    if (glo >= 0) { ts = -1; Fs = Flo; }
    else if (ghi <= 0) { ts = 1; Fs = Fhi; }
    else {
        double lo = -1, hi = 1, gt;
        ts = lo - glo * (hi - lo) / (ghi - glo);
        for (int it = 0; it < 16; it++) {
            seg_eval(pl, al, sz, box, ts, &Fs, &gt, &dg);
            if (gt == 0) break;
            if (gt < 0) lo = ts; else hi = ts;
            double tn = dg > 0 ? ts - gt / dg : 0.5 * (lo + hi);
            if (!(tn > lo && tn < hi)) tn = 0.5 * (lo + hi);
            int done = fabs(tn - ts) <= eps;
            ts = tn;
            if (done) break;
        }
        seg_eval(pl, al, sz, box, ts, &Fs, &gt, &dg);
    }
    if (Fs > r2) return 0;
    *t1 = -1; *t2 = 1;
    if (Flo > r2) {
        double t = -1, F = Flo, g = glo;
        for (int it = 0; it < 12 && F > r2 && g < 0; it++) {
            double tn = t - 0.5 * (F - r2) / g;
            if (tn > ts) tn = ts;
            int done = fabs(tn - t) <= eps;
            t = tn;
            if (done) break;
            seg_eval(pl, al, sz, box, t, &F, &g, &dg);
        }
        *t1 = t;
    }
    if (Fhi > r2) {
        double t = 1, F = Fhi, g = ghi;
        for (int it = 0; it < 12 && F > r2 && g > 0; it++) {
            double tn = t - 0.5 * (F - r2) / g;
            if (tn < ts) tn = ts;
            int done = fabs(tn - t) <= eps;
            t = tn;
            if (done) break;
            seg_eval(pl, al, sz, box, t, &F, &g, &dg);
        }
        *t2 = t;
    }
    return 1;
}
/* capsule (geom1) vs box or cylinder (geom2): sphere-vs-body contacts at the centre of the contact patch; a patch longer than half
 * the capsule against a BOX gives two contacts at its ends (mjc_CapsuleBox also returns up to two), a cylinder pair one (as MPR) */
static int capsuleConvex(const double *p1, const double *m1, const double *s1, const double *p2, const double *m2, const double *s2,
                         int box, double margin, RawCon *con) {
    double ax[3] = {m1[2], m1[5], m1[8]}, t[3] = {p1[0] - p2[0], p1[1] - p2[1], p1[2] - p2[2]}, pl[3], al[3];
    for (int i = 0; i < 3; i++) {
        pl[i] = m2[i] * t[0] + m2[3 + i] * t[1] + m2[6 + i] * t[2];
        al[i] = (m2[i] * ax[0] + m2[3 + i] * ax[1] + m2[6 + i] * ax[2]) * s1[1];
    }
    double sz[3] = {s2[0], box ? s2[1] : s2[0], box ? s2[2] : s2[1]};
    /* separating-axis prefilter on the face normals of the (bounding) box */
    for (int i = 0; i < 3; i++) {
        double lim = sz[i] + s1[0] + margin, a = pl[i] - al[i], b = pl[i] + al[i];
        if ((a > lim && b > lim) || (a < -lim && b < -lim)) return 0;
    }
    double t1, t2;
    if (!seg_patch(pl, al, sz, box, s1[0] + margin, &t1, &t2)) return 0;
    int two = box && (t2 - t1 > 1.0), n = 0;
    for (int e = 0; e < (two ? 2 : 1); e++) {
        double tt = two ? (e == 0 ? t1 : t2) : 0.5 * (t1 + t2);
        double c[3] = {p1[0] + ax[0] * s1[1] * tt, p1[1] + ax[1] * s1[1] * tt, p1[2] + ax[2] * s1[1] * tt};
        n += box ? sphereBoxRaw(c, s1[0], p2, m2, s2, margin, con + n) : sphereCylinderRaw(c, s1[0], p2, m2, s2, margin, con + n);
    }
    return n;
}

/* contact parameter mixing (mj_contactParam semantics) + add contacts for one geom pair */
static void collide_pair(O1Data *d, int g1, int g2) {
    const O1Model *m = d->m;
    /* contype / conaffinity filter */
    if (!((m->geom_contype[g1] & m->geom_conaffinity[g2]) || (m->geom_contype[g2] & m->geom_conaffinity[g1]))) return;
    double margin = fmax(m->geom_margin[g1], m->geom_margin[g2]);
    double gap = fmax(m->geom_gap[g1], m->geom_gap[g2]);
    /* mj_collideGeoms: the lower geom type is geom1 */
    if (m->geom_type[g1] > m->geom_type[g2]) { int t_ = g1; g1 = g2; g2 = t_; }
    RawCon raw[8];
    int num = 0;
    const double *p1 = d->geom_xpos + 3 * g1, *m1 = d->geom_xmat + 9 * g1, *s1 = m->geom_size + 3 * g1;
    const double *p2 = d->geom_xpos + 3 * g2, *m2 = d->geom_xmat + 9 * g2, *s2 = m->geom_size + 3 * g2;
    if (m->geom_type[g1] == GEOM_PLANE) {
        /* bounding-sphere vs plane filter */
        double n[3] = {m1[2], m1[5], m1[8]}, t[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
        if (dot3(t, n) > margin + m->geom_rbound[g2]) return;
        switch (m->geom_type[g2]) {
            case GEOM_SPHERE: num = mjc_PlaneSphere(p1, m1, p2, s2, margin, raw); break;
            case GEOM_CAPSULE: num = mjc_PlaneCapsule(p1, m1, p2, m2, s2, margin, raw); break;
            case GEOM_CYLINDER: num = mjc_PlaneCylinder(p1, m1, p2, m2, s2, margin, raw); break;
            case GEOM_BOX: num = mjc_PlaneBox(p1, m1, p2, m2, s2, margin, raw); break;
            default: return;
        }
    } else {
        /* bounding-sphere filter */
        double t[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]}, bound = m->geom_rbound[g1] + m->geom_rbound[g2] + margin;
        if (dot3(t, t) > bound * bound) return;
        int t1 = m->geom_type[g1], t2 = m->geom_type[g2];
        double z1[3] = {m1[2], m1[5], m1[8]}, z2[3] = {m2[2], m2[5], m2[8]};
        if (t1 == GEOM_SPHERE && t2 == GEOM_SPHERE) num = sphereSphereRaw(p1, s1[0], p2, s2[0], z1, z2, margin, raw);
        else if (t1 == GEOM_SPHERE && t2 == GEOM_CAPSULE) num = mjc_SphereCapsule(p1, m1, s1[0], p2, m2, s2, margin, raw);
        else if (t1 == GEOM_SPHERE && t2 == GEOM_CYLINDER) num = sphereCylinderRaw(p1, s1[0], p2, m2, s2, margin, raw);
        else if (t1 == GEOM_SPHERE && t2 == GEOM_BOX) num = sphereBoxRaw(p1, s1[0], p2, m2, s2, margin, raw);
        else if (t1 == GEOM_CAPSULE && t2 == GEOM_CAPSULE) num = mjc_CapsuleCapsule(p1, m1, s1, p2, m2, s2, margin, raw);
        else if (t1 == GEOM_CAPSULE && t2 == GEOM_CYLINDER) num = capsuleConvex(p1, m1, s1, p2, m2, s2, 0, margin, raw);
        else if (t1 == GEOM_CAPSULE && t2 == GEOM_BOX) num = capsuleConvex(p1, m1, s1, p2, m2, s2, 1, margin, raw);
        else if (t1 == GEOM_CYLINDER && t2 == GEOM_CYLINDER) {
            /* hip motor vs hip motor (MuJoCo: MPR): exact for this robot's coplanar pairs, anything else is COUNTED, not collided */
            num = cylinderCylinderCoplanar(p1, m1, s1, p2, m2, s2, margin, raw);
            if (num < 0) { d->warn_unsupported_pair++; return; }
        }
        else { d->warn_unsupported_pair++; return; }   /* cylinder-box / box-box: no such non-adjacent pair on this robot */
    }
    if (!num) return;
    /* mix parameters */
    int condim;
    double fri[3], solref[2], solimp[5];
    int pr1 = m->geom_priority[g1], pr2 = m->geom_priority[g2];
    if (pr1 != pr2) {
        int g = pr1 > pr2 ? g1 : g2;
        condim = m->geom_condim[g];
        memcpy(fri, m->geom_friction + 3 * g, 24);
        memcpy(solref, m->geom_solref + 2 * g, 16);
        memcpy(solimp, m->geom_solimp + 5 * g, 40);
    } else {
        condim = m->geom_condim[g1] > m->geom_condim[g2] ? m->geom_condim[g1] : m->geom_condim[g2];
        for (int k = 0; k < 3; k++) fri[k] = fmax(m->geom_friction[3 * g1 + k], m->geom_friction[3 * g2 + k]);
        double mix1 = m->geom_solmix[g1], mix2 = m->geom_solmix[g2], mix;
        if (mix1 >= mjMINVAL && mix2 >= mjMINVAL) mix = mix1 / (mix1 + mix2);
        else if (mix1 < mjMINVAL && mix2 < mjMINVAL) mix = 0.5;
        else if (mix1 < mjMINVAL) mix = 0.0;
        else mix = 1.0;
        if (m->geom_solref[2 * g1] > 0 && m->geom_solref[2 * g2] > 0)
            for (int k = 0; k < 2; k++) solref[k] = mix * m->geom_solref[2 * g1 + k] + (1 - mix) * m->geom_solref[2 * g2 + k];
        else
            for (int k = 0; k < 2; k++) solref[k] = fmin(m->geom_solref[2 * g1 + k], m->geom_solref[2 * g2 + k]);
        for (int k = 0; k < 5; k++) solimp[k] = mix * m->geom_solimp[5 * g1 + k] + (1 - mix) * m->geom_solimp[5 * g2 + k];
    }
    for (int i = 0; i < num; i++) {
        if (d->ncon >= MAXCON) { d->warn_contact_full++; return; }
        O1Contact *c = d->contact + d->ncon;
        memset(c, 0, sizeof *c);
        c->dist = raw[i].dist;
        memcpy(c->pos, raw[i].pos, 24);
        memcpy(c->frame, raw[i].frame, 72);
        makeFrame(c->frame);
        c->includemargin = margin - gap;
        c->dim = condim;
        c->geom1 = g1; c->geom2 = g2;
        c->friction[0] = fmax(mjMINMU, fri[0]); c->friction[1] = fmax(mjMINMU, fri[0]);
        c->friction[2] = fmax(mjMINMU, fri[1]);
        c->friction[3] = fmax(mjMINMU, fri[2]); c->friction[4] = fmax(mjMINMU, fri[2]);
        memcpy(c->solref, solref, 16); memcpy(c->solimp, solimp, 40);
        c->exclude = (c->dist >= c->includemargin) ? 1 : 0;
        c->efc_address = -1;
        d->ncon++;
    }
}

/* mj_collision: body pairs in (body1, body2) order (mj_broadphase's sorted list), filterBodyPair (same body, parent and
 * child unless the parent is the world), then every geom pair of the two bodies through collide_pair.  With
 * self_collision == 0 only the world-vs-robot pairs are visited (round-1 behaviour, kept for A/B tests). */
static void o1_collision(O1Data *d) {
    const O1Model *m = d->m;
    d->ncon = 0;
    for (int b1 = 0; b1 < m->nbody; b1++) {
        for (int b2 = b1 + 1; b2 < m->nbody; b2++) {
            if (b1 != 0 && !d->self_collision) continue;
            if (b1 != 0 && (m->body_parentid[b2] == b1 || m->body_parentid[b1] == b2)) continue;
            for (int i1 = 0; i1 < m->body_geomnum[b1]; i1++) {
                int g1 = m->body_geomadr[b1] + i1;
                if (m->geom_type[g1] == GEOM_MESH) continue;
                if (b1 == 0 && m->geom_type[g1] != GEOM_PLANE) continue;
                for (int i2 = 0; i2 < m->body_geomnum[b2]; i2++) {
                    int g2 = m->body_geomadr[b2] + i2;
                    if (m->geom_type[g2] == GEOM_PLANE || m->geom_type[g2] == GEOM_MESH) continue;
                    collide_pair(d, g1, g2);
                }
            }
        }
    }
}

/* ------------------------------------------------------------------------------------------ */
/* mj_jac: jacp/jacr (3 x nv, row-major) of `point` on `body` */
static void o1_jac(const O1Data *d, double *jacp, double *jacr, const double *point, int body) {
    const O1Model *m = d->m;
    int nv = m->nv;
    if (jacp) memset(jacp, 0, sizeof(double) * 3 * nv);
    if (jacr) memset(jacr, 0, sizeof(double) * 3 * nv);
    if (body <= 0) return;
    double off[3];
    for (int k = 0; k < 3; k++) off[k] = point[k] - d->subtree_com[3 * m->body_rootid[body] + k];
    /* last dof of the nearest body with dofs */
    int b = body;
    while (b > 0 && m->body_dofnum[b] == 0) b = m->body_parentid[b];
    if (b <= 0) return;
    int i = m->body_dofadr[b] + m->body_dofnum[b] - 1;
    for (; i >= 0; i = m->dof_parentid[i]) {
        const double *c = d->cdof + 6 * i;
        if (jacr) { jacr[i] = c[0]; jacr[nv + i] = c[1]; jacr[2 * nv + i] = c[2]; }
        if (jacp) {
            double t[3];
            cross3(t, c, off);
            jacp[i] = c[3] + t[0]; jacp[nv + i] = c[4] + t[1]; jacp[2 * nv + i] = c[5] + t[2];
        }
    }
}

static int add_row(O1Data *d, const double *jac, double pos, double margin, double frictionloss, int type, int id) {
    int nv = d->m->nv;
    if (d->nefc >= MAXEFC) { d->warn_contact_full++; return -1; }
    int r = d->nefc++;
    memcpy(d->efc_J + (size_t)r * nv, jac, sizeof(double) * nv);
    d->efc_pos[r] = pos; d->efc_margin[r] = margin; d->efc_frictionloss[r] = frictionloss;
    d->efc_type[r] = type; d->efc_id[r] = id;
    return r;
}

/* getimpedance (engine_core_constraint.c) */
static void get_impedance(const double *solimp_in, double pos, double margin, double *imp, double *impP) {
    double s[5];
    s[0] = fmin(mjMAXIMP, fmax(mjMINIMP, solimp_in[0]));
    s[1] = fmin(mjMAXIMP, fmax(mjMINIMP, solimp_in[1]));
    s[2] = fmax(0.0, solimp_in[2]);
    s[3] = fmin(mjMAXIMP, fmax(mjMINIMP, solimp_in[3]));
    s[4] = fmax(1.0, solimp_in[4]);
    if (s[0] == s[1] || s[2] <= mjMINVAL) { *imp = 0.5 * (s[0] + s[1]); *impP = 0; return; }
    double x = (pos - margin) / s[2], sgn = 1;
    if (x < 0) { x = -x; sgn = -1; }
    if (x >= 1 || x <= 0) { *imp = (x >= 1 ? s[1] : s[0]); *impP = 0; return; }
    double y, yP;
    if (s[4] == 1) { y = x; yP = 1; }
    else if (x <= s[3]) {
        double a = 1 / pow(s[3], s[4] - 1);
        y = a * pow(x, s[4]); yP = s[4] * a * pow(x, s[4] - 1);
    } else {
        double b = 1 / pow(1 - s[3], s[4] - 1);
        y = 1 - b * pow(1 - x, s[4]); yP = s[4] * b * pow(1 - x, s[4] - 1);
    }
    *imp = s[0] + y * (s[1] - s[0]);
    *impP = yP * sgn * (s[1] - s[0]) / s[2];
}

/* mj_makeConstraint + mj_diagApprox + mj_makeImpedance */
static void o1_makeConstraint(O1Data *d) {
    const O1Model *m = d->m;
    int nv = m->nv;
    double *jac = (double *)alloca(sizeof(double) * nv);
    d->nefc = d->ne = d->nf = 0;
    /* friction loss (mj_instantiateFriction) */
    for (int i = 0; i < nv; i++) {
        if (m->dof_frictionloss[i] > 0) {
            memset(jac, 0, sizeof(double) * nv);
            jac[i] = 1;
            add_row(d, jac, 0, 0, m->dof_frictionloss[i], CN_FRICTION_DOF, i);
        }
    }
    d->nf = d->nefc;
    /* joint limits (mj_instantiateLimit) */
    for (int j = 0; j < m->njnt; j++) {
        if (!m->jnt_limited[j]) continue;
        if (m->jnt_type[j] != JNT_HINGE && m->jnt_type[j] != JNT_SLIDE) continue;
        double value = d->qpos[m->jnt_qposadr[j]], margin = m->jnt_margin[j];
        for (int side = -1; side <= 1; side += 2) {
            double dist = side * (m->jnt_range[2 * j + (side + 1) / 2] - value);
            if (dist < margin) {
                memset(jac, 0, sizeof(double) * nv);
                jac[m->jnt_dofadr[j]] = -side;
                add_row(d, jac, dist, margin, 0, CN_LIMIT_JOINT, j);
            }
        }
    }
    /* contacts (mj_instantiateContact, elliptic / frictionless) */
    double *jp1 = (double *)alloca(sizeof(double) * 3 * nv), *jr1 = (double *)alloca(sizeof(double) * 3 * nv);
    double *jp2 = (double *)alloca(sizeof(double) * 3 * nv), *jr2 = (double *)alloca(sizeof(double) * 3 * nv);
    for (int c = 0; c < d->ncon; c++) {
        O1Contact *con = d->contact + c;
        if (con->exclude) continue;
        int b1 = m->geom_bodyid[con->geom1], b2 = m->geom_bodyid[con->geom2];
        o1_jac(d, jp1, jr1, con->pos, b1);
        o1_jac(d, jp2, jr2, con->pos, b2);
        for (int k = 0; k < 3 * nv; k++) { jp2[k] -= jp1[k]; jr2[k] -= jr1[k]; }
        int dim = con->dim;
        con->efc_address = d->nefc;
        for (int r = 0; r < dim; r++) {
            const double *ax = con->frame + 3 * (r < 3 ? r : r - 3);
            const double *src = r < 3 ? jp2 : jr2;
            for (int k = 0; k < nv; k++) jac[k] = ax[0] * src[k] + ax[1] * src[nv + k] + ax[2] * src[2 * nv + k];
            int row = add_row(d, jac, r == 0 ? con->dist : 0, r == 0 ? con->includemargin : 0, 0,
                              dim == 1 ? CN_CONTACT_FRICTIONLESS : CN_CONTACT_ELLIPTIC, c);
            if (row < 0) { con->efc_address = -1; con->exclude = 3; d->nefc -= r; break; }
        }
    }
    /* diagApprox, impedance, R, KBIP */
    for (int i = 0; i < d->nefc; i++) {
        const double *solref, *solimp;
        double dA;
        int is_friction_row = 0;
        switch (d->efc_type[i]) {
            case CN_FRICTION_DOF:
                dA = m->dof_invweight0[d->efc_id[i]];
                solref = m->dof_solref + 2 * d->efc_id[i]; solimp = m->dof_solimp + 5 * d->efc_id[i];
                is_friction_row = 1;
                break;
            case CN_LIMIT_JOINT:
                dA = m->dof_invweight0[m->jnt_dofadr[d->efc_id[i]]];
                solref = m->jnt_solref + 2 * d->efc_id[i]; solimp = m->jnt_solimp + 5 * d->efc_id[i];
                break;
            default: {
                const O1Contact *con = d->contact + d->efc_id[i];
                int b1 = m->geom_bodyid[con->geom1], b2 = m->geom_bodyid[con->geom2];
                int j = i - con->efc_address;
                double tran = m->body_invweight0[2 * b1] + m->body_invweight0[2 * b2];
                double rot = m->body_invweight0[2 * b1 + 1] + m->body_invweight0[2 * b2 + 1];
                dA = j < 3 ? tran : rot;
                solref = con->solref; solimp = con->solimp;
                if (j > 0) is_friction_row = 1;
            }
        }
        d->efc_diagApprox[i] = dA;
        double imp, impP;
        get_impedance(solimp, d->efc_pos[i], d->efc_margin[i], &imp, &impP);
        d->efc_R[i] = fmax(mjMINVAL, (1 - imp) * dA / imp);
        double K, B, dmax = fmin(mjMAXIMP, fmax(mjMINIMP, solimp[1]));
        if (solref[0] > 0) {
            double timeconst = fmax(solref[0], 2 * m->timestep), dampratio = solref[1];
            K = 1 / fmax(mjMINVAL, dmax * dmax * timeconst * timeconst * dampratio * dampratio);
            B = 2 / fmax(mjMINVAL, dmax * timeconst);
        } else {
            K = -solref[0] / fmax(mjMINVAL, dmax * dmax);
            B = -solref[1] / fmax(mjMINVAL, dmax);
        }
        if (is_friction_row) K = 0;
        d->efc_KBIP[4 * i] = K; d->efc_KBIP[4 * i + 1] = B; d->efc_KBIP[4 * i + 2] = imp; d->efc_KBIP[4 * i + 3] = impP;
    }
    /* elliptic friction rows: R scaled by impratio and friction coefficients; contact mu */
    for (int c = 0; c < d->ncon; c++) {
        O1Contact *con = d->contact + c;
        if (con->efc_address < 0 || con->dim == 1) continue;
        int i = con->efc_address;
        d->efc_R[i + 1] = d->efc_R[i] / fmax(mjMINVAL, m->impratio);
        for (int j = 1; j < con->dim; j++)
            d->efc_R[i + j] = d->efc_R[i + 1] * con->friction[0] * con->friction[0] / (con->friction[j - 1] * con->friction[j - 1]);
        con->mu = con->friction[0] * sqrt(d->efc_R[i + 1] / d->efc_R[i]);
    }
    for (int i = 0; i < d->nefc; i++) d->efc_D[i] = 1 / d->efc_R[i];
}

/* ------------------------------------------------------------------------------------------ */
/* mj_comVel + mj_rne(flg_acc=0) -> qfrc_bias                                                  */
static void o1_comVel(O1Data *d) {
    const O1Model *m = d->m;
    memset(d->cvel, 0, sizeof(double) * 6);
    for (int i = 1; i < m->nbody; i++) {
        double cvel[6];
        memcpy(cvel, d->cvel + 6 * m->body_parentid[i], 48);
        int da = m->body_dofadr[i];
        for (int jj = 0; jj < m->body_jntnum[i]; jj++) {
            int j = m->body_jntadr[i] + jj;
            int a = m->jnt_dofadr[j];
            if (m->jnt_type[j] == JNT_FREE) {
                memset(d->cdof_dot + 6 * a, 0, sizeof(double) * 18);
                for (int k = 0; k < 3; k++) for (int c = 0; c < 6; c++) cvel[c] += d->cdof[6 * (a + k) + c] * d->qvel[a + k];
                for (int k = 3; k < 6; k++) crossMotion(d->cdof_dot + 6 * (a + k), cvel, d->cdof + 6 * (a + k));
                for (int k = 3; k < 6; k++) for (int c = 0; c < 6; c++) cvel[c] += d->cdof[6 * (a + k) + c] * d->qvel[a + k];
            } else {
                crossMotion(d->cdof_dot + 6 * a, cvel, d->cdof + 6 * a);
                for (int c = 0; c < 6; c++) cvel[c] += d->cdof[6 * a + c] * d->qvel[a];
            }
        }
        (void)da;
        memcpy(d->cvel + 6 * i, cvel, 48);
    }
}
static void o1_rne_bias(O1Data *d) {
    const O1Model *m = d->m;
    int nb = m->nbody, nv = m->nv;
    memset(d->cacc, 0, 48);
    for (int k = 0; k < 3; k++) d->cacc[3 + k] = -m->gravity[k];
    for (int i = 1; i < nb; i++) {
        double *cacc = d->cacc + 6 * i;
        memcpy(cacc, d->cacc + 6 * m->body_parentid[i], 48);
        int da = m->body_dofadr[i];
        for (int k = 0; k < m->body_dofnum[i]; k++)
            for (int c = 0; c < 6; c++) cacc[c] += d->cdof_dot[6 * (da + k) + c] * d->qvel[da + k];
        double t1[6], t2[6], t3[6];
        mulInertVec(t1, d->cinert + 10 * i, cacc);
        mulInertVec(t2, d->cinert + 10 * i, d->cvel + 6 * i);
        crossForce(t3, d->cvel + 6 * i, t2);
        for (int c = 0; c < 6; c++) d->cfrc_body[6 * i + c] = t1[c] + t3[c];
    }
    memset(d->cfrc_body, 0, 48);
    for (int i = nb - 1; i > 0; i--) {
        int p = m->body_parentid[i];
        if (p > 0) for (int c = 0; c < 6; c++) d->cfrc_body[6 * p + c] += d->cfrc_body[6 * i + c];
    }
    for (int i = 0; i < nv; i++) {
        double s = 0;
        for (int c = 0; c < 6; c++) s += d->cdof[6 * i + c] * d->cfrc_body[6 * m->dof_bodyid[i] + c];
        d->qfrc_bias[i] = s;
    }
}

/* ------------------------------------------------------------------------------------------ */
/* constraint update (mj_constraintUpdate): forces, states, cost; optional cone Hessians       */
static double constraint_update(O1Data *d, const double *jar, int flg_cone_hessian) {
    const O1Model *m = d->m;
    (void)m;
    double cost = 0;
    for (int i = 0; i < d->nefc; i++) {
        double D = d->efc_D[i], R = d->efc_R[i];
        int tp = d->efc_type[i];
        if (tp == CN_FRICTION_DOF) {
            double f = d->efc_frictionloss[i];
            if (jar[i] <= -R * f) { d->efc_force[i] = f; d->efc_state[i] = ST_LINEARNEG; cost += -f * (0.5 * R * f + jar[i]); }
            else if (jar[i] >= R * f) { d->efc_force[i] = -f; d->efc_state[i] = ST_LINEARPOS; cost += -f * (0.5 * R * f - jar[i]); }
            else { d->efc_force[i] = -D * jar[i]; d->efc_state[i] = ST_QUADRATIC; cost += 0.5 * D * jar[i] * jar[i]; }
        } else if (tp == CN_CONTACT_ELLIPTIC) {
            O1Contact *con = d->contact + d->efc_id[i];
            int dim = con->dim;
            double mu = con->mu, U[6], N, T = 0;
            U[0] = jar[i] * mu;
            for (int j = 1; j < dim; j++) { U[j] = jar[i + j] * con->friction[j - 1]; T += U[j] * U[j]; }
            N = U[0]; T = sqrt(T);
            if ((N >= mu * T) || (T <= 0 && N >= 0)) {
                for (int j = 0; j < dim; j++) { d->efc_force[i + j] = 0; d->efc_state[i + j] = ST_SATISFIED; }
            } else if ((mu * N + T <= 0) || (T <= 0 && N < 0)) {
                for (int j = 0; j < dim; j++) {
                    d->efc_force[i + j] = -d->efc_D[i + j] * jar[i + j];
                    d->efc_state[i + j] = ST_QUADRATIC;
                    cost += 0.5 * d->efc_D[i + j] * jar[i + j] * jar[i + j];
                }
            } else {
                double Dm = D / fmax(mu * mu * (1 + mu * mu), mjMINVAL);
                double NT = N - mu * T;
                cost += 0.5 * Dm * NT * NT;
                d->efc_force[i] = -Dm * NT * mu;
                for (int j = 1; j < dim; j++) d->efc_force[i + j] = -d->efc_force[i] / T * U[j] * con->friction[j - 1];
                for (int j = 0; j < dim; j++) d->efc_state[i + j] = ST_CONE;
                if (flg_cone_hessian) {
                    double *H = con->H;
                    memset(H, 0, sizeof(double) * 36);
                    H[0] = 1;
                    for (int j = 1; j < dim; j++) H[j] = -mu * U[j] / T;
                    double scl = mu * N / (T * T * T);
                    for (int k = 1; k < dim; k++) for (int j = k; j < dim; j++) H[k * dim + j] = scl * U[j] * U[k];
                    scl = mu * mu - mu * N / T;
                    for (int j = 1; j < dim; j++) H[j * (dim + 1)] += scl;
                    for (int k = 0; k < dim; k++) {
                        double sk = Dm * (k == 0 ? mu : con->friction[k - 1]);
                        for (int j = k; j < dim; j++) H[k * dim + j] *= sk * (j == 0 ? mu : con->friction[j - 1]);
                    }
                    for (int k = 0; k < dim; k++) for (int j = k + 1; j < dim; j++) H[j * dim + k] = H[k * dim + j];
                }
            }
            i += dim - 1;
        } else { /* limit, frictionless contact */
            if (jar[i] < 0) { d->efc_force[i] = -D * jar[i]; d->efc_state[i] = ST_QUADRATIC; cost += 0.5 * D * jar[i] * jar[i]; }
            else { d->efc_force[i] = 0; d->efc_state[i] = ST_SATISFIED; }
        }
    }
    return cost;
}

typedef struct { double alpha, cost, deriv[2]; } LSPoint;

typedef struct {
    int nv, nefc;
    double *jar, *Jv, *Ma, *Mv, *grad, *Mgrad, *search, *quad, *H;
    double cost, gauss, quadGauss[3];
} Ctx;

static void ls_eval(const O1Data *d, const Ctx *c, LSPoint *p) {
    double cost = 0, alpha = p->alpha, deriv[2] = {0, 0};
    double qt[3] = {c->quadGauss[0], c->quadGauss[1], c->quadGauss[2]};
    for (int i = 0; i < d->nefc; i++) {
        const double *q = c->quad + 3 * i;
        int tp = d->efc_type[i];
        if (tp == CN_FRICTION_DOF) {
            double x = c->jar[i] + alpha * c->Jv[i], f = d->efc_frictionloss[i], Rf = d->efc_R[i] * f;
            if (-Rf < x && x < Rf) { qt[0] += q[0]; qt[1] += q[1]; qt[2] += q[2]; }
            else if (x <= -Rf) { qt[0] += f * (-0.5 * Rf - c->jar[i]); qt[1] += -f * c->Jv[i]; }
            else { qt[0] += f * (-0.5 * Rf + c->jar[i]); qt[1] += f * c->Jv[i]; }
        } else if (tp == CN_CONTACT_ELLIPTIC) {
            const O1Contact *con = d->contact + d->efc_id[i];
            double mu = con->mu;
            const double *q1 = q + 3, *q2 = q + 6;
            double U0 = q1[0], V0 = q1[1], UU = q1[2], UV = q2[0], VV = q2[1], Dm = q2[2];
            double N = U0 + alpha * V0, Tsqr = UU + alpha * (2 * UV + alpha * VV);
            if (Tsqr <= 0) {
                if (N < 0) { qt[0] += q[0]; qt[1] += q[1]; qt[2] += q[2]; }
            } else {
                double T = sqrt(Tsqr);
                if (N >= mu * T) { /* top zone */ }
                else if (mu * N + T <= 0) { qt[0] += q[0]; qt[1] += q[1]; qt[2] += q[2]; }
                else {
                    double N1 = V0, T1 = (UV + alpha * VV) / T, T2 = VV / T - (UV + alpha * VV) * T1 / (T * T);
                    cost += 0.5 * Dm * (N - mu * T) * (N - mu * T);
                    deriv[0] += Dm * (N - mu * T) * (N1 - mu * T1);
                    deriv[1] += Dm * ((N1 - mu * T1) * (N1 - mu * T1) + (N - mu * T) * (-mu * T2));
                }
            }
            i += con->dim - 1;
        } else {
            double x = c->jar[i] + alpha * c->Jv[i];
            if (x < 0) { qt[0] += q[0]; qt[1] += q[1]; qt[2] += q[2]; }
        }
    }
    cost += alpha * alpha * qt[2] + alpha * qt[1] + qt[0];
    deriv[0] += 2 * alpha * qt[2] + qt[1];
    deriv[1] += 2 * qt[2];
    if (deriv[1] <= 0) deriv[1] = mjMINVAL;
    p->cost = cost; p->deriv[0] = deriv[0]; p->deriv[1] = deriv[1];
}

/* update one side of the bracket from the candidate list; returns 1 if updated */
static int update_bracket(const O1Data *d, const Ctx *c, LSPoint *p, const LSPoint cand[3], LSPoint *pnext) {
    int flag = 0;
    for (int i = 0; i < 3; i++) {
        if (p->deriv[0] < 0 && cand[i].deriv[0] < 0 && p->deriv[0] < cand[i].deriv[0]) { *p = cand[i]; flag = 1; }
        else if (p->deriv[0] > 0 && cand[i].deriv[0] > 0 && p->deriv[0] > cand[i].deriv[0]) { *p = cand[i]; flag = 2; }
    }
    if (flag) {
        pnext->alpha = p->alpha - p->deriv[0] / p->deriv[1];
        ls_eval(d, c, pnext);
    }
    return flag;
}

/* PrimalSearch: exact line search on the piecewise cost along `search` */
static double primal_search(O1Data *d, Ctx *c, double tolerance) {
    const O1Model *m = d->m;
    int nv = m->nv, nefc = d->nefc;
    double snorm = 0;
    for (int i = 0; i < nv; i++) snorm += c->search[i] * c->search[i];
    snorm = sqrt(snorm);
    double scale = 1 / (m->meaninertia * (nv > 1 ? nv : 1));
    double gtol = tolerance * m->ls_tolerance * snorm / scale;
    if (snorm < mjMINVAL) return 0;
    mulM(m, d->qM, c->Mv, c->search);
    for (int i = 0; i < nefc; i++) {
        double s = 0;
        const double *J = d->efc_J + (size_t)i * nv;
        for (int k = 0; k < nv; k++) s += J[k] * c->search[k];
        c->Jv[i] = s;
    }
    c->quadGauss[0] = c->gauss;
    c->quadGauss[1] = 0; c->quadGauss[2] = 0;
    for (int i = 0; i < nv; i++) {
        c->quadGauss[1] += c->search[i] * (c->Ma[i] - d->qfrc_smooth[i]);
        c->quadGauss[2] += 0.5 * c->search[i] * c->Mv[i];
    }
    for (int i = 0; i < nefc; i++) {
        double D = d->efc_D[i], *q = c->quad + 3 * i;
        q[0] = 0.5 * D * c->jar[i] * c->jar[i]; q[1] = D * c->jar[i] * c->Jv[i]; q[2] = 0.5 * D * c->Jv[i] * c->Jv[i];
        if (d->efc_type[i] == CN_CONTACT_ELLIPTIC) {
            const O1Contact *con = d->contact + d->efc_id[i];
            int dim = con->dim;
            double mu = con->mu, U[6], V[6];
            U[0] = c->jar[i] * mu; V[0] = c->Jv[i] * mu;
            double UU = 0, UV = 0, VV = 0;
            for (int j = 1; j < dim; j++) {
                double Dj = d->efc_D[i + j];
                U[j] = c->jar[i + j] * con->friction[j - 1]; V[j] = c->Jv[i + j] * con->friction[j - 1];
                q[0] += 0.5 * Dj * c->jar[i + j] * c->jar[i + j];
                q[1] += Dj * c->jar[i + j] * c->Jv[i + j];
                q[2] += 0.5 * Dj * c->Jv[i + j] * c->Jv[i + j];
                UU += U[j] * U[j]; UV += U[j] * V[j]; VV += V[j] * V[j];
            }
            q[3] = U[0]; q[4] = V[0]; q[5] = UU;
            q[6] = UV; q[7] = VV; q[8] = D / fmax(mu * mu * (1 + mu * mu), mjMINVAL);
            i += dim - 1;
        }
    }
    LSPoint p0, p1, p2, pmid, p1next, p2next;
    int iter = 0, maxiter = m->ls_iterations;
    p0.alpha = 0; ls_eval(d, c, &p0); iter++;
    p1.alpha = p0.alpha - p0.deriv[0] / p0.deriv[1]; ls_eval(d, c, &p1); iter++;
    if (p0.cost < p1.cost) p1 = p0;
    if (fabs(p1.deriv[0]) < gtol) { d->ls_total += iter; return p1.alpha; }
    int dir = p1.deriv[0] < 0 ? 1 : -1;
    int p2update = 0;
    p2 = p1;
    while (p1.deriv[0] * dir <= -gtol && iter < maxiter) {
        p2 = p1; p2update = 1;
        p1.alpha -= p1.deriv[0] / p1.deriv[1]; ls_eval(d, c, &p1); iter++;
        if (fabs(p1.deriv[0]) < gtol) { d->ls_total += iter; return p1.alpha; }
    }
    if (iter >= maxiter || !p2update) { d->ls_total += iter; return p1.alpha; }
    p2next = p1;
    p1next.alpha = p1.alpha - p1.deriv[0] / p1.deriv[1]; ls_eval(d, c, &p1next); iter++;
    while (iter < maxiter) {
        pmid.alpha = 0.5 * (p1.alpha + p2.alpha); ls_eval(d, c, &pmid); iter++;
        LSPoint cand[3] = {p1next, p2next, pmid};
        int best = -1;
        for (int i = 0; i < 3; i++)
            if (fabs(cand[i].deriv[0]) < gtol && (best < 0 || cand[i].cost < cand[best].cost)) best = i;
        if (best >= 0) { d->ls_total += iter; return cand[best].alpha; }
        int b1 = update_bracket(d, c, &p1, cand, &p1next); if (b1) iter++;
        int b2 = update_bracket(d, c, &p2, cand, &p2next); if (b2) iter++;
        if (!b1 && !b2) { d->ls_total += iter; return pmid.cost < p0.cost ? pmid.alpha : 0.0; }
    }
    d->ls_total += iter;
    if (p1.cost <= p2.cost && p1.cost < p0.cost) return p1.alpha;
    if (p2.cost <= p1.cost && p2.cost < p0.cost) return p2.alpha;
    return 0;
}

/* dense Cholesky helpers (mju_cholFactor / mju_cholSolve) */
static int chol_factor(double *A, int n) {
    int rank = n;
    for (int j = 0; j < n; j++) {
        double t = A[j * n + j];
        for (int k = 0; k < j; k++) t -= A[j * n + k] * A[j * n + k];
        if (t < mjMINVAL) { t = mjMINVAL; rank--; }
        A[j * n + j] = sqrt(t);
        double inv = 1 / A[j * n + j];
        for (int i = j + 1; i < n; i++) {
            double s = A[i * n + j];
            for (int k = 0; k < j; k++) s -= A[i * n + k] * A[j * n + k];
            A[i * n + j] = s * inv;
        }
    }
    return rank;
}
static void chol_solve(const double *L, int n, double *x) {
    for (int i = 0; i < n; i++) {
        double s = x[i];
        for (int k = 0; k < i; k++) s -= L[i * n + k] * x[k];
        x[i] = s / L[i * n + i];
    }
    for (int i = n - 1; i >= 0; i--) {
        double s = x[i];
        for (int k = i + 1; k < n; k++) s -= L[k * n + i] * x[k];
        x[i] = s / L[i * n + i];
    }
}

/* PrimalUpdateConstraint */
static void primal_update_constraint(O1Data *d, Ctx *c) {
    const O1Model *m = d->m;
    int nv = m->nv;
    c->cost = constraint_update(d, c->jar, 1);
    for (int k = 0; k < nv; k++) d->qfrc_constraint[k] = 0;
    for (int i = 0; i < d->nefc; i++) {
        double f = d->efc_force[i];
        if (f == 0) continue;
        const double *J = d->efc_J + (size_t)i * nv;
        for (int k = 0; k < nv; k++) d->qfrc_constraint[k] += J[k] * f;
    }
    c->gauss = 0;
    for (int k = 0; k < nv; k++) c->gauss += 0.5 * (c->Ma[k] - d->qfrc_smooth[k]) * (d->qacc[k] - d->qacc_smooth[k]);
    c->cost += c->gauss;
}
/* PrimalUpdateGradient (Newton): grad, H = M + J'DJ + cone blocks, Mgrad = H \ grad */
static void primal_update_gradient(O1Data *d, Ctx *c) {
    const O1Model *m = d->m;
    int nv = m->nv;
    for (int k = 0; k < nv; k++) c->grad[k] = c->Ma[k] - d->qfrc_smooth[k] - d->qfrc_constraint[k];
    fullM(m, d->qM, c->H);
    for (int i = 0; i < d->nefc; i++) {
        if (d->efc_state[i] != ST_QUADRATIC) continue;
        const double *J = d->efc_J + (size_t)i * nv;
        double D = d->efc_D[i];
        for (int a = 0; a < nv; a++) {
            if (J[a] == 0) continue;
            double s = D * J[a];
            for (int b = 0; b < nv; b++) c->H[a * nv + b] += s * J[b];
        }
    }
    for (int ci = 0; ci < d->ncon; ci++) {
        const O1Contact *con = d->contact + ci;
        if (con->efc_address < 0 || con->dim == 1) continue;
        int i = con->efc_address, dim = con->dim;
        if (d->efc_state[i] != ST_CONE) continue;
        for (int r = 0; r < dim; r++) for (int s = 0; s < dim; s++) {
            double h = con->H[r * dim + s];
            if (h == 0) continue;
            const double *Jr = d->efc_J + (size_t)(i + r) * nv, *Js = d->efc_J + (size_t)(i + s) * nv;
            for (int a = 0; a < nv; a++) {
                if (Jr[a] == 0) continue;
                double t = h * Jr[a];
                for (int b = 0; b < nv; b++) c->H[a * nv + b] += t * Js[b];
            }
        }
    }
    chol_factor(c->H, nv);
    memcpy(c->Mgrad, c->grad, sizeof(double) * nv);
    chol_solve(c->H, nv, c->Mgrad);
}

/* mj_solNewton */
static void o1_solNewton(O1Data *d, int maxiter, double tolerance) {
    const O1Model *m = d->m;
    int nv = m->nv, nefc = d->nefc;
    Ctx c;
    c.nv = nv; c.nefc = nefc;
    c.jar = dalloc(nefc); c.Jv = dalloc(nefc); c.Ma = dalloc(nv); c.Mv = dalloc(nv); c.grad = dalloc(nv);
    c.Mgrad = dalloc(nv); c.search = dalloc(nv); c.quad = dalloc(3 * nefc + 9); c.H = dalloc(nv * nv);
    for (int i = 0; i < nefc; i++) {
        double s = 0;
        const double *J = d->efc_J + (size_t)i * nv;
        for (int k = 0; k < nv; k++) s += J[k] * d->qacc[k];
        c.jar[i] = s - d->efc_aref[i];
    }
    mulM(m, d->qM, c.Ma, d->qacc);
    primal_update_constraint(d, &c);
    primal_update_gradient(d, &c);
    for (int k = 0; k < nv; k++) c.search[k] = -c.Mgrad[k];
    double scale = 1 / (m->meaninertia * (nv > 1 ? nv : 1));
    int iter = 0;
    d->solver_improvement = d->solver_gradient = 0;
    while (iter < maxiter) {
        double alpha = primal_search(d, &c, tolerance);
        if (alpha == 0) break;
        for (int k = 0; k < nv; k++) { d->qacc[k] += alpha * c.search[k]; c.Ma[k] += alpha * c.Mv[k]; }
        for (int i = 0; i < nefc; i++) c.jar[i] += alpha * c.Jv[i];
        double oldcost = c.cost;
        primal_update_constraint(d, &c);
        primal_update_gradient(d, &c);
        for (int k = 0; k < nv; k++) c.search[k] = -c.Mgrad[k];
        double improvement = scale * (oldcost - c.cost), gn = 0;
        for (int k = 0; k < nv; k++) gn += c.grad[k] * c.grad[k];
        double gradient = scale * sqrt(gn);
        d->solver_improvement = improvement; d->solver_gradient = gradient;
        iter++;
        if (improvement < tolerance || gradient < tolerance) break;
    }
    d->solver_iter = iter;
    free(c.jar); free(c.Jv); free(c.Ma); free(c.Mv); free(c.grad); free(c.Mgrad); free(c.search); free(c.quad); free(c.H);
}

/* ------------------------------------------------------------------------------------------ */
static int bad(const double *x, int n) {
    for (int i = 0; i < n; i++) if (!(x[i] == x[i]) || x[i] > mjMAXVAL || x[i] < -mjMAXVAL) return 1;
    return 0;
}

/* mj_forward */
EXPORT void o1_forward(O1Data *d) {
    const O1Model *m = d->m;
    int nv = m->nv;
    /* --- fwdPosition --- */
    o1_kinematics(d);
    o1_comPos(d);
    o1_crb(d);
    memcpy(d->qLD, d->qM, sizeof(double) * m->nM);
    factor_tree(m, d->qLD, d->qLDiagInv);
    o1_collision(d);
    o1_makeConstraint(d);
    /* --- fwdVelocity --- */
    o1_comVel(d);
    for (int i = 0; i < nv; i++) d->qfrc_passive[i] = -m->dof_damping[i] * d->qvel[i];
    for (int i = 0; i < d->nefc; i++) { /* mj_referenceConstraint */
        double s = 0;
        const double *J = d->efc_J + (size_t)i * nv;
        for (int k = 0; k < nv; k++) s += J[k] * d->qvel[k];
        d->efc_vel[i] = s;
        const double *KBIP = d->efc_KBIP + 4 * i;
        d->efc_aref[i] = -KBIP[1] * s - KBIP[0] * KBIP[2] * (d->efc_pos[i] - d->efc_margin[i]);
    }
    o1_rne_bias(d);
    /* --- fwdActuation --- */
    memset(d->qfrc_actuator, 0, sizeof(double) * nv);
    for (int u = 0; u < m->nu; u++) {
        double ctrl = d->ctrl[u];
        if (m->actuator_ctrllimited[u]) ctrl = fmin(m->actuator_ctrlrange[2 * u + 1], fmax(m->actuator_ctrlrange[2 * u], ctrl));
        int j = m->actuator_trnid[u];
        double gear = m->actuator_gear[u];
        double length = gear * d->qpos[m->jnt_qposadr[j]], velocity = gear * d->qvel[m->jnt_dofadr[j]];
        double force = m->actuator_gainprm[3 * u] * ctrl;
        if (m->actuator_biastype[u] == 1)
            force += m->actuator_biasprm[3 * u] + m->actuator_biasprm[3 * u + 1] * length + m->actuator_biasprm[3 * u + 2] * velocity;
        if (m->actuator_forcelimited[u]) force = fmin(m->actuator_forcerange[2 * u + 1], fmax(m->actuator_forcerange[2 * u], force));
        d->actuator_force[u] = force;
        d->qfrc_actuator[m->jnt_dofadr[j]] += gear * force;
    }
    /* --- fwdAcceleration --- */
    for (int i = 0; i < nv; i++) {
        d->qfrc_smooth[i] = d->qfrc_passive[i] - d->qfrc_bias[i] + d->qfrc_actuator[i];
        d->qacc_smooth[i] = d->qfrc_smooth[i];
    }
    solve_tree(m, d->qLD, d->qLDiagInv, d->qacc_smooth);
    /* --- fwdConstraint --- */
    if (d->nefc == 0) {
        memcpy(d->qacc, d->qacc_smooth, sizeof(double) * nv);
        memset(d->qfrc_constraint, 0, sizeof(double) * nv);
        d->solver_iter = 0;
        return;
    }
    {
        double tol = d->tolerance_override >= 0 ? d->tolerance_override : m->tolerance;
        double *jar = (double *)alloca(sizeof(double) * d->nefc), *Ma = (double *)alloca(sizeof(double) * nv);
        /* warmstart: compare cost at qacc_warmstart and at qacc_smooth */
        for (int i = 0; i < d->nefc; i++) {
            double s = 0;
            const double *J = d->efc_J + (size_t)i * nv;
            for (int k = 0; k < nv; k++) s += J[k] * d->qacc_warmstart[k];
            jar[i] = s - d->efc_aref[i];
        }
        double cost_warm = constraint_update(d, jar, 0);
        mulM(m, d->qM, Ma, d->qacc_warmstart);
        for (int k = 0; k < nv; k++) cost_warm += 0.5 * (Ma[k] - d->qfrc_smooth[k]) * (d->qacc_warmstart[k] - d->qacc_smooth[k]);
        for (int i = 0; i < d->nefc; i++) {
            double s = 0;
            const double *J = d->efc_J + (size_t)i * nv;
            for (int k = 0; k < nv; k++) s += J[k] * d->qacc_smooth[k];
            jar[i] = s - d->efc_aref[i];
        }
        double cost_smooth = constraint_update(d, jar, 0);
        if (cost_warm > cost_smooth) memcpy(d->qacc, d->qacc_smooth, sizeof(double) * nv);
        else memcpy(d->qacc, d->qacc_warmstart, sizeof(double) * nv);
        o1_solNewton(d, m->iterations, tol);
    }
}

/* mj_Euler with implicit joint damping + mj_advance */
static void o1_euler(O1Data *d) {
    const O1Model *m = d->m;
    int nv = m->nv;
    double h = m->timestep;
    double *qacc = (double *)alloca(sizeof(double) * nv);
    int damped = 0;
    for (int i = 0; i < nv; i++) if (m->dof_damping[i] > 0) damped = 1;
    if (damped) {
        memcpy(d->qH, d->qM, sizeof(double) * m->nM);
        for (int i = 0; i < nv; i++) d->qH[m->dof_Madr[i]] += h * m->dof_damping[i];
        factor_tree(m, d->qH, d->qHDiagInv);
        for (int i = 0; i < nv; i++) qacc[i] = d->qfrc_smooth[i] + d->qfrc_constraint[i];
        solve_tree(m, d->qH, d->qHDiagInv, qacc);
    } else memcpy(qacc, d->qacc, sizeof(double) * nv);
    for (int i = 0; i < nv; i++) d->qvel[i] += h * qacc[i];
    /* mj_integratePos */
    for (int j = 0; j < m->njnt; j++) {
        int qa = m->jnt_qposadr[j], da = m->jnt_dofadr[j];
        if (m->jnt_type[j] == JNT_FREE) {
            for (int k = 0; k < 3; k++) d->qpos[qa + k] += h * d->qvel[da + k];
            double ax[3] = {d->qvel[da + 3], d->qvel[da + 4], d->qvel[da + 5]}, qrot[4];
            double angle = h * normalize3(ax);
            axisAngle2Quat(qrot, ax, angle);
            normalize4(d->qpos + qa + 3);
            mulQuat(d->qpos + qa + 3, d->qpos + qa + 3, qrot);
            normalize4(d->qpos + qa + 3);
        } else d->qpos[qa] += h * d->qvel[da];
    }
    d->time += h;
    memcpy(d->qacc_warmstart, d->qacc, sizeof(double) * nv);
}

/* mj_step */
EXPORT void o1_step(O1Data *d) {
    const O1Model *m = d->m;
    if (bad(d->qpos, m->nq) || bad(d->qvel, m->nv)) { d->warn_bad++; o1_reset_data(d); }
    o1_forward(d);
    if (bad(d->qacc, m->nv)) { d->warn_bad++; o1_reset_data(d); o1_forward(d); }
    o1_euler(d);
}
EXPORT void o1_step_n(O1Data *d, int n) { for (int i = 0; i < n; i++) o1_step(d); }

/* mj_rnePostConstraint, the part the env reads: cfrc_ext from contacts */
EXPORT void o1_rne_post_constraint(O1Data *d) {
    const O1Model *m = d->m;
    memset(d->cfrc_ext, 0, sizeof(double) * 6 * m->nbody);
    for (int c = 0; c < d->ncon; c++) {
        const O1Contact *con = d->contact + c;
        if (con->efc_address < 0) continue;
        double lfrc[6] = {0, 0, 0, 0, 0, 0}, cfrc[6];
        for (int j = 0; j < con->dim; j++) lfrc[j] = d->efc_force[con->efc_address + j];
        /* rotVecMatT: world = frame^T * local */
        for (int k = 0; k < 3; k++) {
            cfrc[3 + k] = con->frame[k] * lfrc[0] + con->frame[3 + k] * lfrc[1] + con->frame[6 + k] * lfrc[2];
            cfrc[k] = con->frame[k] * lfrc[3] + con->frame[3 + k] * lfrc[4] + con->frame[6 + k] * lfrc[5];
        }
        int bs[2] = {m->geom_bodyid[con->geom1], m->geom_bodyid[con->geom2]};
        for (int s = 0; s < 2; s++) {
            int k = bs[s];
            if (!k) continue;
            const double *com = d->subtree_com + 3 * m->body_rootid[k];
            double r[3] = {con->pos[0] - com[0], con->pos[1] - com[1], con->pos[2] - com[2]}, t[3];
            cross3(t, r, cfrc + 3);
            double sign = s == 0 ? -1.0 : 1.0;
            for (int a = 0; a < 3; a++) {
                d->cfrc_ext[6 * k + a] += sign * (cfrc[a] + t[a]);
                d->cfrc_ext[6 * k + 3 + a] += sign * cfrc[3 + a];
            }
        }
    }
}

/* ------------------------------------------------------------------------------------------ */
/* accessors for ctypes                                                                         */
EXPORT double *o1_field(O1Data *d, const char *name, int *n) {
    const O1Model *m = d->m;
    int nv = m->nv, nb = m->nbody;
#define F(nm, cnt) if (strcmp(name, #nm) == 0) { if (n) *n = (cnt); return d->nm; }
    F(qpos, m->nq) F(qvel, nv) F(ctrl, m->nu) F(qacc, nv) F(qacc_warmstart, nv) F(qacc_smooth, nv)
    F(qfrc_bias, nv) F(qfrc_passive, nv) F(qfrc_actuator, nv) F(qfrc_smooth, nv) F(qfrc_constraint, nv) F(actuator_force, m->nu)
    F(xpos, 3 * nb) F(xquat, 4 * nb) F(xmat, 9 * nb) F(xipos, 3 * nb) F(ximat, 9 * nb) F(xanchor, 3 * m->njnt) F(xaxis, 3 * m->njnt)
    F(geom_xpos, 3 * m->ngeom) F(geom_xmat, 9 * m->ngeom) F(site_xpos, 3 * m->nsite) F(site_xmat, 9 * m->nsite)
    F(subtree_com, 3 * nb) F(cinert, 10 * nb) F(crb, 10 * nb) F(cdof, 6 * nv) F(cdof_dot, 6 * nv) F(cvel, 6 * nb) F(cacc, 6 * nb)
    F(cfrc_ext, 6 * nb) F(qM, m->nM) F(qLD, m->nM) F(qLDiagInv, nv)
    F(efc_J, d->nefc * nv) F(efc_pos, d->nefc) F(efc_margin, d->nefc) F(efc_frictionloss, d->nefc) F(efc_diagApprox, d->nefc)
    F(efc_R, d->nefc) F(efc_D, d->nefc) F(efc_KBIP, 4 * d->nefc) F(efc_vel, d->nefc) F(efc_aref, d->nefc) F(efc_force, d->nefc)
#undef F
    if (strcmp(name, "time") == 0) { if (n) *n = 1; return &d->time; }
    if (strcmp(name, "tolerance_override") == 0) { if (n) *n = 1; return &d->tolerance_override; }
    if (strcmp(name, "solver_improvement") == 0) { if (n) *n = 1; return &d->solver_improvement; }
    if (strcmp(name, "solver_gradient") == 0) { if (n) *n = 1; return &d->solver_gradient; }
    return NULL;
}
EXPORT int *o1_ifield(O1Data *d, const char *name, int *n) {
#define F(nm, cnt) if (strcmp(name, #nm) == 0) { if (n) *n = (cnt); return &d->nm; }
#define FP(nm, cnt) if (strcmp(name, #nm) == 0) { if (n) *n = (cnt); return d->nm; }
    F(ncon, 1) F(nefc, 1) F(nf, 1) F(solver_iter, 1) F(ls_total, 1) F(warn_bad, 1) F(warn_contact_full, 1) F(warn_unsupported_pair, 1) F(self_collision, 1)
    FP(efc_type, d->nefc) FP(efc_id, d->nefc) FP(efc_state, d->nefc)
#undef F
#undef FP
    return NULL;
}
/* contact i -> flat record: dist,pos3,frame9,includemargin,friction5,solref2,solimp5,mu,dim,geom1,geom2,exclude,efc_address (32 doubles) */
EXPORT int o1_contact(const O1Data *d, int i, double *out) {
    if (i < 0 || i >= d->ncon) return -1;
    const O1Contact *c = d->contact + i;
    int k = 0;
    out[k++] = c->dist;
    for (int a = 0; a < 3; a++) out[k++] = c->pos[a];
    for (int a = 0; a < 9; a++) out[k++] = c->frame[a];
    out[k++] = c->includemargin;
    for (int a = 0; a < 5; a++) out[k++] = c->friction[a];
    for (int a = 0; a < 2; a++) out[k++] = c->solref[a];
    for (int a = 0; a < 5; a++) out[k++] = c->solimp[a];
    out[k++] = c->mu; out[k++] = c->dim; out[k++] = c->geom1; out[k++] = c->geom2; out[k++] = c->exclude; out[k++] = c->efc_address;
    return k;
}
EXPORT int o1_nq(const O1Model *m) { return m->nq; }
EXPORT int o1_nv(const O1Model *m) { return m->nv; }
EXPORT int o1_nu(const O1Model *m) { return m->nu; }
EXPORT int o1_nbody(const O1Model *m) { return m->nbody; }
